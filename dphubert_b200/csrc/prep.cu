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

// ---- coalesced variants (used when the geometry allows; the kernels above remain the general fallback) -------------
// sum of squares per tap: consecutive threads read consecutive taps (v is [row][tap]); requires 256 % K == 0, so a
// thread always meets the same tap.  normsq must be zero on entry.
__global__ void __launch_bounds__(256)
    posconv_normsq_kernel(const float* __restrict__ v, float* __restrict__ normsq, long long total, int K) {
  __shared__ float red[256];
  const long long chunk = ((total + gridDim.x - 1) / gridDim.x + 255) / 256 * 256;
  const long long e0 = (long long)blockIdx.x * chunk, e1 = e0 + chunk < total ? e0 + chunk : total;
  float acc = 0.f;
  for (long long e = e0 + threadIdx.x; e < e1; e += 256) {
    const float x = v[e];
    acc = fmaf(x, x, acc);
  }
  red[threadIdx.x] = acc;
  __syncthreads();
  if ((int)threadIdx.x < K) {
    float t = 0.f;
    for (int q = threadIdx.x; q < 256; q += K) t += red[q];
    atomicAdd(normsq + threadIdx.x, t);
  }
}
__global__ void posconv_sqrt_kernel(float* __restrict__ x, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) x[i] = sqrtf(x[i]);
}

// One block per (group, chunk of TC taps): stages v[g Cg + co][ci][tap0 .. tap0 + TC) in shared memory and writes both
// tap-major operand copies with their contiguous dimension (ci for wp, co for wdp) across the threads.
template <typename T, int TC>
__global__ void __launch_bounds__(256)
    posconv_weight_prep_tiled_kernel(const float* __restrict__ v, const float* __restrict__ g,
                                     const float* __restrict__ norm, float* __restrict__ wn, T* __restrict__ wp,
                                     T* __restrict__ wdp, int G, int Cg, int Cgp, int K) {
  extern __shared__ float tile[];  // [co][ci][TC]
  const int grp = blockIdx.y, tap0 = blockIdx.x * TC;
  const int n = Cg * Cg;
  float sc[TC];
#pragma unroll
  for (int q = 0; q < TC; ++q) sc[q] = (tap0 + q < K) ? g[tap0 + q] / norm[tap0 + q] : 0.f;
  for (int e = threadIdx.x; e < n; e += 256) {
    const int co = e / Cg, ci = e - co * Cg;
    const long long base = ((long long)(grp * Cg + co) * Cg + ci) * K + tap0;
#pragma unroll
    for (int q = 0; q < TC; ++q) {
      float val = 0.f;
      if (tap0 + q < K) {
        val = sc[q] * v[base + q];
        if (wn) wn[base + q] = val;
      }
      tile[e * TC + q] = val;
    }
  }
  __syncthreads();
  for (int e = threadIdx.x; e < n * TC; e += 256) {
    // wp: (co, tap, ci) with ci fastest
    const int ci = e % Cg, r = e / Cg, q = r % TC, co = r / TC;
    if (tap0 + q < K) {
      const float val = tile[(co * Cg + ci) * TC + q];
      if (wp) wp[(((long long)grp * Cg + co) * K + tap0 + q) * Cgp + ci] = from_f<T>(val);
    }
  }
  for (int e = threadIdx.x; e < n * TC; e += 256) {
    // wdp: (ci, K-1-tap, co) with co fastest
    const int co = e % Cg, r = e / Cg, q = r % TC, ci = r / TC;
    if (tap0 + q < K) {
      const float val = tile[(co * Cg + ci) * TC + q];
      if (wdp) wdp[(((long long)grp * Cg + ci) * K + (K - 1 - tap0 - q)) * Cgp + co] = from_f<T>(val);
    }
  }
}

// s[tap] += sum_ci dwp[g][co][tap][ci] * v[g Cg + co][ci][tap] for one (g, co) per block: v's slab is staged in shared
// memory (row pitch K + 1), dwp is read with ci across the lanes.  s must be zero on entry.
__global__ void __launch_bounds__(256)
    posconv_wbwd_dot_tiled_kernel(const float* __restrict__ dwp, const float* __restrict__ v, float* __restrict__ s,
                                  int Cg, int Cop, int Cgp, int K) {
  extern __shared__ float vs[];  // [ci][K + 1]
  const int d = blockIdx.x;      // g * Cg + co
  const int grp = d / Cg, co = d - grp * Cg;
  const float* vsrc = v + (long long)d * Cg * K;
  for (int e = threadIdx.x; e < Cg * K; e += 256) vs[(e / K) * (K + 1) + (e % K)] = vsrc[e];
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const float* src = dwp + ((long long)grp * Cop + co) * K * Cgp;
  for (int tap = w; tap < K; tap += 8) {
    float acc = 0.f;
    for (int ci = lane; ci < Cg; ci += 32) acc = fmaf(src[(long long)tap * Cgp + ci], vs[ci * (K + 1) + tap], acc);
    acc = warp_sum(acc);
    if (lane == 0) atomicAdd(s + tap, acc);
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
  const long long total = (long long)D * Cg * K;
  int launches = 2;
  if (256 % K == 0) {
    cudaMemsetAsync(norm, 0, sizeof(float) * K, s);
    posconv_normsq_kernel<<<148 * 2, 256, 0, s>>>(v, norm, total, K);
    posconv_sqrt_kernel<<<1, 256, 0, s>>>(norm, K);
    launches = 3;
  } else {
    posconv_norm_kernel<<<K, 256, 0, s>>>(v, norm, D, Cg, K);
  }
  constexpr int TC = 4;
  const size_t tile_bytes = sizeof(float) * Cg * Cg * TC;
  if (tile_bytes <= 48 * 1024) {
    dim3 grid((K + TC - 1) / TC, G);
    if (dtype == DPH_DTYPE_BF16)
      posconv_weight_prep_tiled_kernel<bf16, TC><<<grid, 256, tile_bytes, s>>>(v, g, norm, wn, (bf16*)wp, (bf16*)wdp, G,
                                                                               Cg, Cgp, K);
    else
      posconv_weight_prep_tiled_kernel<float, TC><<<grid, 256, tile_bytes, s>>>(v, g, norm, wn, (float*)wp, (float*)wdp,
                                                                                G, Cg, Cgp, K);
  } else if (dtype == DPH_DTYPE_BF16) {
    posconv_weight_prep_kernel<bf16><<<grid_for(total), 256, 0, s>>>(v, g, norm, wn, (bf16*)wp, (bf16*)wdp, D, G, Cg, Cgp, K);
  } else {
    posconv_weight_prep_kernel<float><<<grid_for(total), 256, 0, s>>>(v, g, norm, wn, (float*)wp, (float*)wdp, D, G, Cg, Cgp,
                                                                      K);
  }
  dph_count_launch(launches);
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
  const size_t vs_bytes = sizeof(float) * Cg * (K + 1);
  if (vs_bytes <= 48 * 1024) {
    cudaMemsetAsync(scratch, 0, sizeof(float) * K, s);
    posconv_wbwd_dot_tiled_kernel<<<D, 256, vs_bytes, s>>>(dwp, v, scratch, Cg, Cop, Cgp, K);
  } else {
    posconv_wbwd_dot_kernel<<<K, 256, 0, s>>>(dwp, v, scratch, D, G, Cg, Cop, Cgp, K);
  }
  const long long total = (long long)D * Cg * K;
  posconv_wbwd_apply_kernel<<<grid_for(total), 256, 0, s>>>(dwp, v, g, norm, scratch, dv, dg, D, G, Cg, Cop, Cgp, K);
  dph_count_launch(2);
  return dph_check_last("posconv_weight_bwd");
}
