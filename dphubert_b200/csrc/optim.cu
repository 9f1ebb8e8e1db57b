// Optimizer step over flat parameter arenas: global grad-norm, clip, AdamW (3 param groups incl. the negative-lr
// Lagrange multipliers), refresh of the bf16 GEMM weight copy and grad zeroing — one pass over the state.
//
// Replaces lightning.py:200-238 (torch.optim.AdamW with groups main / log_alpha / lambda) and distill.py:48
// (gradient_clip_val -> clip_grad_norm_): ~6 passes over 95.6 M elements become 2 (sum of squares, fused update).
// Hyper-parameters that change every step (lr per group, step count, clip norm) live in device memory so the
// whole step can sit inside a CUDA graph.
#include "common.cuh"
#include "../../include/dphubert_b200.h"

namespace dph {
void set_error(const char* fmt, ...);

__global__ void __launch_bounds__(256) sumsq_kernel(const float* __restrict__ x, long long n, double* __restrict__ out) {
  pdl_wait();
  pdl_trigger();
  __shared__ double sh[8];
  double acc = 0.0;
  const long long stride = (long long)gridDim.x * blockDim.x * 4;
  const long long n4 = n & ~3LL;
  for (long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < n4; i += stride) {
    const float4 v = *reinterpret_cast<const float4*>(x + i);
    acc += (double)(v.x * v.x + v.y * v.y) + (double)(v.z * v.z + v.w * v.w);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (long long i = n4; i < n; ++i) acc += (double)x[i] * (double)x[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int q = 0; q < 8; ++q) t += sh[q];
    atomicAdd(out, t);
  }
}

__global__ void counter_inc_kernel(int* c) {
  pdl_wait();
  pdl_trigger();
  if (threadIdx.x == 0 && blockIdx.x == 0) c[0] += 1;
}

// state[0] (double) = sum of squared grads over all arenas (0 => no clipping info); lr_dev[0] = group lr.
template <bool WRITE_BF16>
__global__ void __launch_bounds__(256)
    adamw_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, long long n,
                 const float* __restrict__ lr_dev, float beta1, float beta2, float eps, float wd,
                 const int* __restrict__ step_dev, const double* __restrict__ gnorm_sq, float max_norm,
                 bf16* __restrict__ p_bf16, int zero_grad, int vec_ok) {
  pdl_wait();
  pdl_trigger();
  const float lr = lr_dev[0];
  const int step = step_dev[0];
  float clip = 1.f;
  if (gnorm_sq != nullptr && max_norm > 0.f) {
    const float total = (float)sqrt(gnorm_sq[0]);
    clip = fminf(1.f, max_norm / (total + 1e-6f));
  }
  const float bc1 = 1.f - powf(beta1, (float)step);
  const float bc2 = 1.f - powf(beta2, (float)step);
  const float step_size = lr / bc1;
  const float inv_sqrt_bc2 = rsqrtf(bc2);
  const float decay = 1.f - lr * wd;
  auto upd = [&](float& pi, float gi, float& mi, float& vi) {
    const float gr = gi * clip;
    pi *= decay;
    mi = beta1 * mi + (1.f - beta1) * gr;
    vi = beta2 * vi + (1.f - beta2) * gr * gr;
    const float denom = sqrtf(vi) * inv_sqrt_bc2 + eps;
    pi -= step_size * mi / denom;
  };
  // 16-byte main loop (all four arrays 16-byte aligned: checked by the launcher), scalar tail
  const long long n4 = vec_ok ? n / 4 : 0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    // master weights and moments are touched once per step by this kernel only: streaming (evict-first) accesses keep
    // them from displacing the bf16 weight mirror and the activations the next step's GEMMs find in L2
    float4 p4 = __ldcs(reinterpret_cast<const float4*>(p) + i);
    const float4 g4 = reinterpret_cast<const float4*>(g)[i];
    float4 m4 = __ldcs(reinterpret_cast<const float4*>(m) + i);
    float4 v4 = __ldcs(reinterpret_cast<const float4*>(v) + i);
    upd(p4.x, g4.x, m4.x, v4.x);
    upd(p4.y, g4.y, m4.y, v4.y);
    upd(p4.z, g4.z, m4.z, v4.z);
    upd(p4.w, g4.w, m4.w, v4.w);
    __stcs(reinterpret_cast<float4*>(p) + i, p4);
    __stcs(reinterpret_cast<float4*>(m) + i, m4);
    __stcs(reinterpret_cast<float4*>(v) + i, v4);
    if (WRITE_BF16) {
      uint2 u;
      u.x = pack_bf16x2(p4.x, p4.y);
      u.y = pack_bf16x2(p4.z, p4.w);
      reinterpret_cast<uint2*>(p_bf16)[i] = u;
    }
    if (zero_grad) reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  for (long long i = n4 * 4 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    float pi = p[i], mi = m[i], vi = v[i];
    upd(pi, g[i], mi, vi);
    p[i] = pi;
    m[i] = mi;
    v[i] = vi;
    if (WRITE_BF16) p_bf16[i] = __float2bfloat16_rn(pi);
    if (zero_grad) g[i] = 0.f;
  }
}

// g *= 1/world (after a SUM all-reduce)  -- only needed when the reduction is not already an average
__global__ void scale_kernel(float* __restrict__ x, long long n, float a) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    x[i] *= a;
}

}  // namespace dph

using namespace dph;

extern "C" int dph_sumsq(const float* x, int64_t n, double* out, void* stream) {
  if (!x || !out || n <= 0 || (reinterpret_cast<uintptr_t>(x) & 15) != 0) {
    set_error("sumsq: bad arguments (pointer must be 16-byte aligned)");
    return DPH_ERR_ARG;
  }
  long long blocks = (n / 4 + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  launch_pdl(sumsq_kernel, dim3((unsigned)blocks), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), x, n, out);
  dph_count_launch();
  return dph_check_last("sumsq");
}

extern "C" int dph_counter_inc(int* counter, void* stream) {
  launch_pdl(counter_inc_kernel, dim3(1), dim3(32), 0, reinterpret_cast<cudaStream_t>(stream), counter);
  dph_count_launch();
  return dph_check_last("counter_inc");
}

extern "C" int dph_adamw(float* p, float* g, float* m, float* v, int64_t n, const float* lr_dev, float beta1, float beta2,
                         float eps, float weight_decay, const int* step_dev, const double* gnorm_sq, float max_norm,
                         void* p_bf16, int zero_grad, void* stream) {
  if (!p || !g || !m || !v || !lr_dev || !step_dev || n <= 0) {
    set_error("adamw: bad arguments");
    return DPH_ERR_ARG;
  }
  auto al = [](const void* q, uintptr_t a) { return (reinterpret_cast<uintptr_t>(q) & (a - 1)) == 0; };
  const int vec_ok = al(p, 16) && al(g, 16) && al(m, 16) && al(v, 16) && (!p_bf16 || al(p_bf16, 8));
  long long blocks = ((vec_ok ? n / 4 : n) + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (p_bf16)
    launch_pdl(adamw_kernel<true>, dim3((unsigned)blocks), dim3(256), 0, s, p, g, m, v, n, lr_dev, beta1, beta2, eps,
               weight_decay, step_dev, gnorm_sq, max_norm, (bf16*)p_bf16, zero_grad, vec_ok);
  else
    launch_pdl(adamw_kernel<false>, dim3((unsigned)blocks), dim3(256), 0, s, p, g, m, v, n, lr_dev, beta1, beta2, eps,
               weight_decay, step_dev, gnorm_sq, max_norm, (bf16*)nullptr, zero_grad, vec_ok);
  dph_count_launch();
  return dph_check_last("adamw");
}

extern "C" int dph_scale(float* x, int64_t n, float a, void* stream) {
  long long blocks = (n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  scale_kernel<<<(unsigned)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(x, n, a);
  dph_count_launch();
  return dph_check_last("scale");
}
