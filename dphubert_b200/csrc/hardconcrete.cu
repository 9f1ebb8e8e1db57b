// HardConcrete (L0) gates: batched sampling, expected-L0 norms, the differentiable expected model size and the
// Lagrangian sparsity term — a handful of launches instead of the ~200 scalar kernels per step of the reference.
//
// Replaces wav2vec2/hardconcrete.py:76-99 (l0_norm, train-mode forward), the get_num_params tree
// (components.py:122-134,187-196,276-277,438-449,750-761,859-865,929-934,1005-1009; model.py:109-113) and
// lightning.py:267-271 (expected sparsity + Lagrangian), including their autograd backward.
#include "common.cuh"
#include "../../include/dphubert_b200.h"

namespace dph {
void set_error(const char* fmt, ...);

// u ~ U(eps, 1-eps) from the counter-based hash (matches the distribution of Tensor.uniform_(eps, 1-eps))
DPH_DEVICE float hc_uniform(unsigned long long seed, unsigned long long idx, float eps) {
  const uint32_t r = hash_u32(seed, idx);
  const float u01 = ((float)(r >> 8) + 0.5f) * (1.0f / 16777216.0f);
  return eps + u01 * (1.0f - 2.0f * eps);
}

__global__ void hardconcrete_fwd_kernel(const float* __restrict__ la, const float* __restrict__ u_in,
                                        float* __restrict__ u_out, float* __restrict__ mask, long long n, float beta,
                                        float l, float r, float eps, unsigned long long seed) {
  pdl_wait();
  pdl_trigger();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float u = u_in ? u_in[i] : hc_uniform(seed, (unsigned long long)i, eps);
  if (u_out) u_out[i] = u;
  const float s = sigmoidf_((logf(u / (1.f - u)) + la[i]) / beta);
  const float m = s * (r - l) + l;
  mask[i] = fminf(fmaxf(m, 0.f), 1.f);
}

// dla += dmask * d mask / d la
__global__ void hardconcrete_bwd_kernel(const float* __restrict__ la, const float* __restrict__ u_in,
                                        const float* __restrict__ dmask, float* __restrict__ dla, long long n,
                                        float beta, float l, float r) {
  pdl_wait();
  pdl_trigger();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float u = u_in[i];
  const float s = sigmoidf_((logf(u / (1.f - u)) + la[i]) / beta);
  const float m = s * (r - l) + l;
  float g = 0.f;
  if (m > 0.f && m < 1.f) g = dmask[i] * (r - l) * s * (1.f - s) / beta;
  dla[i] += g;
}

// norms[s] = sum_{j in segment s} sigmoid(la[j] + bias)      one block per segment
__global__ void __launch_bounds__(256)
    l0_norms_kernel(const float* __restrict__ la_base, const long long* __restrict__ seg_off,
                    const int* __restrict__ seg_len, float bias, float* __restrict__ norms) {
  __shared__ float red[32];
  const int s = blockIdx.x;
  const float* la = la_base + seg_off[s];
  const int n = seg_len[s];
  float acc = 0.f;
  for (int j = threadIdx.x; j < n; j += blockDim.x) acc += sigmoidf_(la[j] + bias);
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) norms[s] = acc;
}

// Expected parameter count and its derivative wrt every segment norm. `prog` layout (int32):
//   [0] n_conv; then per conv layer 5 ints: seg (-1 = no gate), kernel, has_bias, has_norm, static_out
//   then: proj_out, static_extra (pos-conv numel + 2*D), D, n_layers
//   then per layer 9 ints: att_present, heads_seg, heads_static, head_dim, attgate_seg,
//                           ff_present, ff_seg, ff_static, ffgate_seg
// Single thread: the whole formula is ~10 L scalar operations.
__global__ void num_params_kernel(const int* __restrict__ prog, const float* __restrict__ norms, int nseg,
                                  float* __restrict__ out_n, float* __restrict__ dnorm) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  for (int s = 0; s < nseg; ++s) dnorm[s] = 0.f;
  int p = 0;
  const int n_conv = prog[p++];
  double total = 0.0;
  // conv chain: params_i = in * out_i * k (+ out_i) (+ 2 out_i); in_{i+1} = out_i.  d/d out_i couples layer i and i+1.
  double in_ch = 1.0;
  int prev_seg = -1;
  double prev_coef = 0.0;  // d(total)/d(in_ch) contributed by the current layer is out*k; handled when visiting it
  const int conv_base = p;
  for (int i = 0; i < n_conv; ++i) {
    const int seg = prog[p], k = prog[p + 1], hb = prog[p + 2], hn = prog[p + 3], so = prog[p + 4];
    p += 5;
    const double out = seg >= 0 ? (double)norms[seg] : (double)so;
    total += in_ch * out * k + (hb ? out : 0.0) + (hn ? 2.0 * out : 0.0);
    if (seg >= 0) dnorm[seg] += (float)(in_ch * k + (hb ? 1.0 : 0.0) + (hn ? 2.0 : 0.0));
    if (prev_seg >= 0) dnorm[prev_seg] += (float)(out * k);  // this layer's dependence on its input width
    (void)prev_coef;
    prev_seg = seg;
    in_ch = out;
  }
  (void)conv_base;
  const int proj_out = prog[p++], static_extra = prog[p++], D = prog[p++], n_layers = prog[p++];
  // dummy weight + feature projection: in + in*2 + (in+1)*proj_out
  total += in_ch + in_ch * 2.0 + (in_ch + 1.0) * proj_out;
  if (prev_seg >= 0) dnorm[prev_seg] += (float)(3.0 + proj_out);
  total += (double)static_extra;
  for (int l = 0; l < n_layers; ++l) {
    const int att = prog[p], hseg = prog[p + 1], hst = prog[p + 2], hd = prog[p + 3], ags = prog[p + 4];
    const int ff = prog[p + 5], fseg = prog[p + 6], fst = prog[p + 7], fgs = prog[p + 8];
    p += 9;
    total += 4.0 * D;
    if (att) {
      const double h = hseg >= 0 ? (double)norms[hseg] : (double)hst;
      const double a = (D + 1.0) * h * hd * 3.0 + (h * hd + 1.0) * D;
      const double da_dh = (D + 1.0) * hd * 3.0 + (double)hd * D;
      const double g = ags >= 0 ? (double)norms[ags] : 1.0;
      total += a * g;
      if (hseg >= 0) dnorm[hseg] += (float)(da_dh * g);
      if (ags >= 0) dnorm[ags] += (float)a;
    }
    if (ff) {
      const double f = fseg >= 0 ? (double)norms[fseg] : (double)fst;
      const double a = (D + 1.0) * f + (f + 1.0) * D;
      const double da_df = (D + 1.0) + D;
      const double g = fgs >= 0 ? (double)norms[fgs] : 1.0;
      total += a * g;
      if (fseg >= 0) dnorm[fseg] += (float)(da_df * g);
      if (fgs >= 0) dnorm[fgs] += (float)a;
    }
  }
  out_n[0] = (float)total;
}

// Lagrangian: s = 1 - N/N0; loss = l1 (s - t) + l2 (s - t)^2.  out: [loss, s, dloss/dN, dloss/dl1, dloss/dl2]
__global__ void lagrangian_kernel(const float* __restrict__ n_hat, const float* __restrict__ lambdas, float n0,
                                  const float* __restrict__ target, float* __restrict__ out) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const float s = 1.f - n_hat[0] / n0;
  const float t = target[0];
  const float l1 = lambdas[0], l2 = lambdas[1];
  const float d = s - t;
  out[0] = l1 * d + l2 * d * d;
  out[1] = s;
  out[2] = -(l1 + 2.f * l2 * d) / n0;
  out[3] = d;
  out[4] = d * d;
}

// dla[j] += upstream * coef[0] * dnorm[seg] * sigmoid'(la_j + bias)      one block per segment
__global__ void __launch_bounds__(256)
    l0_norms_bwd_kernel(const float* __restrict__ la_base, float* __restrict__ dla_base,
                        const long long* __restrict__ seg_off, const int* __restrict__ seg_len, float bias,
                        const float* __restrict__ dnorm, const float* __restrict__ upstream,
                        const float* __restrict__ coef) {
  const int s = blockIdx.x;
  const float* la = la_base + seg_off[s];
  float* dla = dla_base + seg_off[s];
  const int n = seg_len[s];
  float g = dnorm[s];
  if (upstream) g *= upstream[0];
  if (coef) g *= coef[0];
  for (int j = threadIdx.x; j < n; j += blockDim.x) {
    const float sg = sigmoidf_(la[j] + bias);
    dla[j] += g * sg * (1.f - sg);
  }
}

}  // namespace dph

using namespace dph;

extern "C" int dph_hardconcrete_fwd(const float* log_alpha, const float* u_in, float* u_out, float* mask, int64_t n,
                                    float beta, float limit_l, float limit_r, float eps, uint64_t seed, void* stream) {
  if (!log_alpha || !mask || n <= 0) {
    set_error("hardconcrete_fwd: bad arguments");
    return DPH_ERR_ARG;
  }
  launch_pdl(hardconcrete_fwd_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream),
             log_alpha, u_in, u_out, mask, n, beta, limit_l, limit_r, eps, seed);
  dph_count_launch();
  return dph_check_last("hardconcrete_fwd");
}

extern "C" int dph_hardconcrete_bwd(const float* log_alpha, const float* u, const float* dmask, float* dlog_alpha,
                                    int64_t n, float beta, float limit_l, float limit_r, void* stream) {
  if (!log_alpha || !u || !dmask || !dlog_alpha || n <= 0) {
    set_error("hardconcrete_bwd: bad arguments");
    return DPH_ERR_ARG;
  }
  hardconcrete_bwd_kernel<<<(unsigned)((n + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      log_alpha, u, dmask, dlog_alpha, n, beta, limit_l, limit_r);
  dph_count_launch();
  return dph_check_last("hardconcrete_bwd");
}

extern "C" int dph_l0_norms(const float* la_base, const int64_t* seg_off, const int* seg_len, int nseg, float bias,
                            float* norms, void* stream) {
  if (!la_base || !seg_off || !seg_len || !norms || nseg <= 0) {
    set_error("l0_norms: bad arguments");
    return DPH_ERR_ARG;
  }
  l0_norms_kernel<<<nseg, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(la_base, (const long long*)seg_off, seg_len,
                                                                           bias, norms);
  dph_count_launch();
  return dph_check_last("l0_norms");
}

extern "C" int dph_num_params(const int* prog, const float* norms, int nseg, float* out_n, float* dnorm, void* stream) {
  if (!prog || !out_n || !dnorm) {
    set_error("num_params: bad arguments");
    return DPH_ERR_ARG;
  }
  num_params_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(prog, norms, nseg, out_n, dnorm);
  dph_count_launch();
  return dph_check_last("num_params");
}

extern "C" int dph_lagrangian(const float* n_hat, const float* lambdas, float n0, const float* target, float* out,
                              void* stream) {
  if (!n_hat || !lambdas || !target || !out) {
    set_error("lagrangian: bad arguments");
    return DPH_ERR_ARG;
  }
  lagrangian_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(n_hat, lambdas, n0, target, out);
  dph_count_launch();
  return dph_check_last("lagrangian");
}

extern "C" int dph_l0_norms_bwd(const float* la_base, float* dla_base, const int64_t* seg_off, const int* seg_len,
                                int nseg, float bias, const float* dnorm, const float* upstream, const float* coef,
                                void* stream) {
  if (!la_base || !dla_base || !seg_off || !seg_len || !dnorm || nseg <= 0) {
    set_error("l0_norms_bwd: bad arguments");
    return DPH_ERR_ARG;
  }
  l0_norms_bwd_kernel<<<nseg, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      la_base, dla_base, (const long long*)seg_off, seg_len, bias, dnorm, upstream, coef);
  dph_count_launch();
  return dph_check_last("l0_norms_bwd");
}
