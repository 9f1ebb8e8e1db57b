// Weight preparation for the implicit-GEMM convolutions (tiny, once per optimizer step):
//  * strided 512->512 convs (components.py:81-87): (Cout, Cin, k) master weight -> tap-major GEMM operands
//      wf[co][tap][ci]            forward  B operand  (K index = tap * Cin + ci)
//      wd[perm(tap)][ci][co]      dgrad    B operand  (taps re-ordered so the even/odd output phases are contiguous)
//  * positional conv (components.py:298-307): weight_norm(dim=2) reparameterisation w = g * v / ||v||_(out,in)
//      wn[d][ci][tap]  fp32 normalised weight (for the backward), plus per-group tap-major operands
//      wp[g][co][tap][cip]  forward,   wdp[g][ci][K-1-tap][cop]  dgrad (flipped taps), channels padded to Cgp
//    and the backward of the reparameterisation (dv, dg from dW).
#include "common.cuh"
#include "../../include/dphubert_b200.h"

namespace dph {
void set_error(const char* fmt, ...);

struct TapPerm {
  int p[16];
};

template <typename T>
__global__ void conv_weight_prep_kernel(const float* __restrict__ w, T* __restrict__ wf, T* __restrict__ wd, int Cout,
                                        int Cin, int k, TapPerm perm) {
  const long long total = (long long)Cout * Cin * k;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int tap = (int)(i % k);
    const long long r = i / k;
    const int ci = (int)(r % Cin), co = (int)(r / Cin);
    const float v = w[i];
    if (wf) wf[((long long)co * k + tap) * Cin + ci] = from_f<T>(v);
    if (wd) wd[((long long)perm.p[tap] * Cin + ci) * Cout + co] = from_f<T>(v);
  }
}

// dW[co][ci][tap] += src[co][tap][ci]
__global__ void conv_wgrad_unpermute_kernel(const float* __restrict__ src, float* __restrict__ dw, int Cout, int Cin,
                                            int k) {
  const long long total = (long long)Cout * Cin * k;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int tap = (int)(i % k);
    const long long r = i / k;
    const int ci = (int)(r % Cin), co = (int)(r / Cin);
    dw[i] += src[((long long)co * k + tap) * Cin + ci];
  }
}

// norm[tap] = || v[:, :, tap] ||_2     one block per tap
__global__ void __launch_bounds__(256)
    posconv_norm_kernel(const float* __restrict__ v, float* __restrict__ norm, int D, int Cg, int K) {
  __shared__ float red[32];
  const int tap = blockIdx.x;
  float acc = 0.f;
  const int n = D * Cg;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float x = v[(long long)i * K + tap];
    acc += x * x;
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) norm[tap] = sqrtf(acc);
}

template <typename T>
__global__ void posconv_weight_prep_kernel(const float* __restrict__ v, const float* __restrict__ g,
                                           const float* __restrict__ norm, float* __restrict__ wn, T* __restrict__ wp,
                                           T* __restrict__ wdp, int D, int G, int Cg, int Cgp, int K) {
  const long long total = (long long)D * Cg * K;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int tap = (int)(i % K);
    const long long r = i / K;
    const int ci = (int)(r % Cg), d = (int)(r / Cg);
    const int grp = d / Cg, co = d - grp * Cg;
    const float val = g[tap] * v[i] / norm[tap];
    if (wn) wn[i] = val;
    if (wp) wp[(((long long)grp * Cg + co) * K + tap) * Cgp + ci] = from_f<T>(val);
    if (wdp) wdp[(((long long)grp * Cg + ci) * K + (K - 1 - tap)) * Cgp + co] = from_f<T>(val);
  }
}

// pass 1 of the weight-norm backward: s[tap] = sum_{o,i} dW[o,i,tap] * v[o,i,tap], where dW is read from the
// per-group tap-major fp32 layout the wgrad GEMM produced: dwp[g][co (ld Cop)][tap][cip]
__global__ void __launch_bounds__(256)
    posconv_wbwd_dot_kernel(const float* __restrict__ dwp, const float* __restrict__ v, float* __restrict__ s, int D,
                            int G, int Cg, int Cop, int Cgp, int K) {
  __shared__ float red[32];
  const int tap = blockIdx.x;
  float acc = 0.f;
  const int n = D * Cg;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int ci = i % Cg, d = i / Cg;
    const int grp = d / Cg, co = d - grp * Cg;
    acc += dwp[(((long long)grp * Cop + co) * K + tap) * Cgp + ci] * v[(long long)i * K + tap];
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) s[tap] = acc;
}

// pass 2: dg[tap] += s/norm ;  dv += g/norm * (dW - v * s / norm^2)
__global__ void posconv_wbwd_apply_kernel(const float* __restrict__ dwp, const float* __restrict__ v,
                                          const float* __restrict__ g, const float* __restrict__ norm,
                                          const float* __restrict__ s, float* __restrict__ dv, float* __restrict__ dg,
                                          int D, int G, int Cg, int Cop, int Cgp, int K) {
  const long long total = (long long)D * Cg * K;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int tap = (int)(i % K);
    const long long r = i / K;
    const int ci = (int)(r % Cg), d = (int)(r / Cg);
    const int grp = d / Cg, co = d - grp * Cg;
    const float dw = dwp[(((long long)grp * Cop + co) * K + tap) * Cgp + ci];
    const float nr = norm[tap];
    dv[i] += g[tap] / nr * (dw - v[i] * s[tap] / (nr * nr));
    if (i < K) dg[i] += s[i] / norm[i];
  }
}

}  // namespace dph

using namespace dph;

static int grid_for(long long n) {
  long long g = (n + 255) / 256;
  return (int)(g < 148 * 8 ? (g < 1 ? 1 : g) : 148 * 8);
}

extern "C" int dph_conv_weight_prep(const float* w, void* wf, void* wd, int Cout, int Cin, int k, const int* tap_perm,
                                    int dtype, void* stream) {
  if (!w || k <= 0 || k > 16) {
    set_error("conv_weight_prep: bad arguments");
    return DPH_ERR_ARG;
  }
  TapPerm perm;
  for (int i = 0; i < 16; ++i) perm.p[i] = (tap_perm && i < k) ? tap_perm[i] : i;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const long long total = (long long)Cout * Cin * k;
  if (dtype == DPH_DTYPE_BF16)
    conv_weight_prep_kernel<bf16><<<grid_for(total), 256, 0, s>>>(w, (bf16*)wf, (bf16*)wd, Cout, Cin, k, perm);
  else
    conv_weight_prep_kernel<float><<<grid_for(total), 256, 0, s>>>(w, (float*)wf, (float*)wd, Cout, Cin, k, perm);
  dph_count_launch();
  return dph_check_last("conv_weight_prep");
}

extern "C" int dph_conv_wgrad_unpermute(const float* src, float* dw, int Cout, int Cin, int k, void* stream) {
  const long long total = (long long)Cout * Cin * k;
  conv_wgrad_unpermute_kernel<<<grid_for(total), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(src, dw, Cout, Cin, k);
  dph_count_launch();
  return dph_check_last("conv_wgrad_unpermute");
}

extern "C" int dph_posconv_weight_prep(const float* v, const float* g, float* norm, float* wn, void* wp, void* wdp, int D,
                                       int G, int Cg, int Cgp, int K, int dtype, void* stream) {
  if (!v || !g || !norm || D != G * Cg) {
    set_error("posconv_weight_prep: bad arguments");
    return DPH_ERR_ARG;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  posconv_norm_kernel<<<K, 256, 0, s>>>(v, norm, D, Cg, K);
  const long long total = (long long)D * Cg * K;
  if (dtype == DPH_DTYPE_BF16)
    posconv_weight_prep_kernel<bf16><<<grid_for(total), 256, 0, s>>>(v, g, norm, wn, (bf16*)wp, (bf16*)wdp, D, G, Cg, Cgp, K);
  else
    posconv_weight_prep_kernel<float><<<grid_for(total), 256, 0, s>>>(v, g, norm, wn, (float*)wp, (float*)wdp, D, G, Cg, Cgp,
                                                                      K);
  dph_count_launch(2);
  return dph_check_last("posconv_weight_prep");
}

// scratch: K floats
extern "C" int dph_posconv_weight_bwd(const float* dwp, const float* v, const float* g, const float* norm, float* scratch,
                                      float* dv, float* dg, int D, int G, int Cg, int Cop, int Cgp, int K, void* stream) {
  if (!dwp || !v || !g || !norm || !scratch || !dv || !dg) {
    set_error("posconv_weight_bwd: bad arguments");
    return DPH_ERR_ARG;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  posconv_wbwd_dot_kernel<<<K, 256, 0, s>>>(dwp, v, scratch, D, G, Cg, Cop, Cgp, K);
  const long long total = (long long)D * Cg * K;
  posconv_wbwd_apply_kernel<<<grid_for(total), 256, 0, s>>>(dwp, v, g, norm, scratch, dv, dg, D, G, Cg, Cop, Cgp, K);
  dph_count_launch(2);
  return dph_check_last("posconv_weight_bwd");
}
