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

// Tiled variant (Cout % 32 == 0, Cin % 32 == 0, k <= 4): a block moves a 32 x 32 (co, ci) tile with all its taps
// through shared memory, so that the reads (ci, tap contiguous) and both writes (ci contiguous for wf, co contiguous
// for wd) are 64 .. 128-byte runs instead of 2-byte scatters (12.4 -> ~3 us per 512 x 512 x 3 layer).
template <typename T>
__global__ void __launch_bounds__(256)
    conv_weight_prep_tiled_kernel(const float* __restrict__ w, T* __restrict__ wf, T* __restrict__ wd, int Cout, int Cin,
                                  int k, TapPerm perm) {
  pdl_wait();
  pdl_trigger();
  __shared__ float tile[32][32 * 4 + 1];
  const int ci0 = blockIdx.x * 32, co0 = blockIdx.y * 32;
  const int rowlen = 32 * k;
  for (int e = threadIdx.x; e < 32 * rowlen; e += 256) {
    const int col = e / rowlen, rem = e - col * rowlen;   // rem = ci_l * k + tap
    tile[col][rem] = w[((long long)(co0 + col) * Cin + ci0) * k + rem];
  }
  __syncthreads();
  if (wf) {
    for (int e = threadIdx.x; e < 32 * rowlen; e += 256) {
      const int cil = e & 31, r = e >> 5, tap = r % k, col = r / k;
      wf[((long long)(co0 + col) * k + tap) * Cin + ci0 + cil] = from_f<T>(tile[col][cil * k + tap]);
    }
  }
  if (wd) {
    for (int e = threadIdx.x; e < 32 * rowlen; e += 256) {
      const int col = e & 31, r = e >> 5, cil = r & 31, tap = r >> 5;
      wd[((long long)perm.p[tap] * Cin + ci0 + cil) * Cout + co0 + col] = from_f<T>(tile[col][cil * k + tap]);
    }
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

// tiled variant (same envelope and tile as conv_weight_prep_tiled_kernel): 128-byte reads, contiguous 32 k float
// read-modify-writes
__global__ void __launch_bounds__(256)
    conv_wgrad_unpermute_tiled_kernel(const float* __restrict__ src, float* __restrict__ dw, int Cout, int Cin, int k) {
  pdl_wait();
  pdl_trigger();
  __shared__ float tile[32][32 * 4 + 1];
  const int ci0 = blockIdx.x * 32, co0 = blockIdx.y * 32;
  const int rowlen = 32 * k;
  for (int e = threadIdx.x; e < 32 * rowlen; e += 256) {
    const int cil = e & 31, r = e >> 5, tap = r % k, col = r / k;
    tile[col][cil * k + tap] = src[((long long)(co0 + col) * k + tap) * Cin + ci0 + cil];
  }
  __syncthreads();
  for (int e = threadIdx.x; e < 32 * rowlen; e += 256) {
    const int col = e / rowlen, rem = e - col * rowlen;
    dw[((long long)(co0 + col) * Cin + ci0) * k + rem] += tile[col][rem];
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
  pdl_wait();
  pdl_trigger();
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
  pdl_wait();
  pdl_trigger();
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
  pdl_wait();
  pdl_trigger();
  extern __shared__ float tile[];  // [q][co][Cg + 1]
  const int grp = blockIdx.y, tap0 = blockIdx.x * TC;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int pitch = Cg + 1, plane = Cg * pitch;
  float sc[TC];
#pragma unroll
  for (int q = 0; q < TC; ++q) sc[q] = (tap0 + q < K) ? g[tap0 + q] / norm[tap0 + q] : 0.f;
  const bool full = (tap0 + TC <= K) && (K % 4 == 0) &&   // 16-byte pieces of v / wn
                    ((reinterpret_cast<uintptr_t>(v) | reinterpret_cast<uintptr_t>(wn)) & 15) == 0;
  // a warp per output channel, the input channels across its lanes: TC consecutive taps (TC * 4 bytes) per element
  for (int co = warp; co < Cg; co += 8) {
    for (int ci = lane; ci < Cg; ci += 32) {
      const long long base = ((long long)(grp * Cg + co) * Cg + ci) * K + tap0;
      float val[TC];
      if (full) {
#pragma unroll
        for (int q = 0; q < TC; q += 4) {
          const float4 x = *reinterpret_cast<const float4*>(v + base + q);
          val[q] = sc[q] * x.x; val[q + 1] = sc[q + 1] * x.y; val[q + 2] = sc[q + 2] * x.z; val[q + 3] = sc[q + 3] * x.w;
          if (wn) *reinterpret_cast<float4*>(wn + base + q) = make_float4(val[q], val[q + 1], val[q + 2], val[q + 3]);
        }
      } else {
#pragma unroll
        for (int q = 0; q < TC; ++q) {
          val[q] = 0.f;
          if (tap0 + q < K) {
            val[q] = sc[q] * v[base + q];
            if (wn) wn[base + q] = val[q];
          }
        }
      }
#pragma unroll
      for (int q = 0; q < TC; ++q) tile[q * plane + co * pitch + ci] = val[q];
    }
  }
  __syncthreads();
  // wp: (co, tap, ci) with ci fastest;  wdp: (ci, K-1-tap, co) with co fastest — a warp per (row, tap), lanes along the row
  for (int r = warp; r < Cg * TC; r += 8) {
    const int q = r % TC, a = r / TC;
    if (tap0 + q >= K) continue;
    for (int l = lane; l < Cg; l += 32) {
      if (wp) wp[(((long long)grp * Cg + a) * K + tap0 + q) * Cgp + l] = from_f<T>(tile[q * plane + a * pitch + l]);
      if (wdp) wdp[(((long long)grp * Cg + a) * K + (K - 1 - tap0 - q)) * Cgp + l] = from_f<T>(tile[q * plane + l * pitch + a]);
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

// Weight-norm backward of the positional conv, tiled (one block per output channel d = g Cg + co, 256 % K == 0):
// the block's slab of dW — dwp[g][co][tap][ci], ci fastest — is staged in shared memory with coalesced reads, after
// which consecutive threads walk consecutive TAPS, the fastest index of v / dv.  (The kernels above read one of the
// two operands with a stride of Cgp or K floats: 103 + 50 us for 4.7 M elements.)
//   PASS 1:  s[tap] += sum_ci dW[d, ci, tap] v[d, ci, tap]          (s zero on entry)
//   PASS 2:  dv += g / norm * (dW - v s / norm^2),   dg += s / norm  (block 0)
template <int PASS>
__global__ void __launch_bounds__(256)
    posconv_wbwd_tiled_kernel(const float* __restrict__ dwp, const float* __restrict__ v, const float* __restrict__ g,
                              const float* __restrict__ norm, float* __restrict__ s, float* __restrict__ dv,
                              float* __restrict__ dg, int Cg, int Cop, int Cgp, int K) {
  pdl_wait();
  pdl_trigger();
  extern __shared__ float dws[];   // [tap][Cg + 1]
  __shared__ float red[256];
  const int d = blockIdx.x;
  const int grp = d / Cg, co = d - grp * Cg;
  const float* src = dwp + ((long long)grp * Cop + co) * K * Cgp;
  for (int e = threadIdx.x; e < K * Cg; e += 256) {
    const int tap = e / Cg, ci = e - tap * Cg;
    dws[tap * (Cg + 1) + ci] = src[(long long)tap * Cgp + ci];
  }
  __syncthreads();
  const int tap = threadIdx.x % K;          // 256 % K == 0: a thread always meets the same tap
  const long long base = (long long)d * Cg * K;
  if (PASS == 1) {
    float acc = 0.f;
#pragma unroll 4
    for (int ci = threadIdx.x / K; ci < Cg; ci += 256 / K) acc = fmaf(dws[tap * (Cg + 1) + ci], v[base + (long long)ci * K + tap], acc);
    red[threadIdx.x] = acc;
    __syncthreads();
    if ((int)threadIdx.x < K) {
      float t = 0.f;
      for (int q = threadIdx.x; q < 256; q += K) t += red[q];
      atomicAdd(s + threadIdx.x, t);
    }
  } else {
    const float nr = norm[tap], st = s[tap];
    const float a = g[tap] / nr, c = st / (nr * nr);
#pragma unroll 4
    for (int ci = threadIdx.x / K; ci < Cg; ci += 256 / K) {
      const long long i = base + (long long)ci * K + tap;
      dv[i] += a * (dws[tap * (Cg + 1) + ci] - v[i] * c);
    }
    if (d == 0 && (int)threadIdx.x < K) dg[threadIdx.x] += st / nr;
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
  if (k <= 4 && Cout % 32 == 0 && Cin % 32 == 0) {
    const dim3 grid(Cin / 32, Cout / 32);
    if (dtype == DPH_DTYPE_BF16)
      launch_pdl(conv_weight_prep_tiled_kernel<bf16>, grid, dim3(256), 0, s, w, (bf16*)wf, (bf16*)wd, Cout, Cin, k, perm);
    else
      launch_pdl(conv_weight_prep_tiled_kernel<float>, grid, dim3(256), 0, s, w, (float*)wf, (float*)wd, Cout, Cin, k, perm);
    dph_count_launch();
    return dph_check_last("conv_weight_prep");
  }
  if (dtype == DPH_DTYPE_BF16)
    conv_weight_prep_kernel<bf16><<<grid_for(total), 256, 0, s>>>(w, (bf16*)wf, (bf16*)wd, Cout, Cin, k, perm);
  else
    conv_weight_prep_kernel<float><<<grid_for(total), 256, 0, s>>>(w, (float*)wf, (float*)wd, Cout, Cin, k, perm);
  dph_count_launch();
  return dph_check_last("conv_weight_prep");
}

extern "C" int dph_conv_wgrad_unpermute(const float* src, float* dw, int Cout, int Cin, int k, void* stream) {
  const long long total = (long long)Cout * Cin * k;
  if (k <= 4 && Cout % 32 == 0 && Cin % 32 == 0)
    launch_pdl(conv_wgrad_unpermute_tiled_kernel, dim3(Cin / 32, Cout / 32), dim3(256), 0,
               reinterpret_cast<cudaStream_t>(stream), src, dw, Cout, Cin, k);
  else
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
    launch_pdl(posconv_normsq_kernel, dim3(148 * 2), dim3(256), 0, s, v, norm, total, K);
    launch_pdl(posconv_sqrt_kernel, dim3(1), dim3(256), 0, s, norm, K);
    launches = 3;
  } else {
    posconv_norm_kernel<<<K, 256, 0, s>>>(v, norm, D, Cg, K);
  }
  constexpr int TC = 8;   // 8 taps = one full 32-byte sector of v per (co, ci)
  const size_t tile_bytes = sizeof(float) * Cg * (Cg + 1) * TC;
  if (tile_bytes <= 96 * 1024) {
    static bool attr = false;
    if (!attr) {
      cudaFuncSetAttribute(posconv_weight_prep_tiled_kernel<bf16, TC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
      cudaFuncSetAttribute(posconv_weight_prep_tiled_kernel<float, TC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
      attr = true;
    }
    dim3 grid((K + TC - 1) / TC, G);
    if (dtype == DPH_DTYPE_BF16)
      launch_pdl(posconv_weight_prep_tiled_kernel<bf16, TC>, grid, dim3(256), tile_bytes, s, v, g, (const float*)norm, wn,
                 (bf16*)wp, (bf16*)wdp, G, Cg, Cgp, K);
    else
      launch_pdl(posconv_weight_prep_tiled_kernel<float, TC>, grid, dim3(256), tile_bytes, s, v, g, (const float*)norm, wn,
                 (float*)wp, (float*)wdp, G, Cg, Cgp, K);
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
  const size_t dws_bytes = sizeof(float) * K * (Cg + 1);
  if (K <= 256 && 256 % K == 0 && dws_bytes <= 46 * 1024) {
    cudaMemsetAsync(scratch, 0, sizeof(float) * K, s);
    launch_pdl(posconv_wbwd_tiled_kernel<1>, dim3(D), dim3(256), dws_bytes, s, dwp, v, g, norm, scratch, dv, dg, Cg, Cop, Cgp, K);
    launch_pdl(posconv_wbwd_tiled_kernel<2>, dim3(D), dim3(256), dws_bytes, s, dwp, v, g, norm, scratch, dv, dg, Cg, Cop, Cgp, K);
    dph_count_launch(2);
    return dph_check_last("posconv_weight_bwd");
  }
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
