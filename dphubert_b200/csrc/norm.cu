// Row LayerNorm forward / backward (memory-bound; one warp per row, 128-bit accesses where aligned).
//
// Replaces nn.LayerNorm / the transposed channel LayerNorm of the reference:
//   wav2vec2/components.py:54-61 (conv-stack channel LN, Large), :271 (feature projection), :808-811,:839,:850-856
//   (encoder layers), :879,:888-889,:905-906 (transformer).  Activations are channels-last here, so every one of
//   these is a plain LayerNorm over the last dimension.
// Optional fusions: GELU + per-column mask after the norm (conv blocks of the Large extractor, components.py:108-114),
// dropout after the norm (components.py:891), second upstream gradient added in the backward (hidden states that
// feed both the next layer and a distillation head).
#include "common.cuh"
#include "vec.cuh"
#include "../../include/dphubert_b200.h"

namespace dph {
void set_error(const char* fmt, ...);

constexpr int LN_MAX_PER_LANE = 40;  // supports D <= 32 * 40 = 1280

template <typename T, int NPL>
__global__ void __launch_bounds__(256)
    layernorm_fwd_kernel(const T* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                         T* __restrict__ y, T* __restrict__ y_pre, float* __restrict__ mean_out,
                         float* __restrict__ rstd_out, const float* __restrict__ colmask, int rows, int D, int Dv,
                         float eps, int act, float drop_p, unsigned long long seed) {
  // Dv <= D: only the first Dv columns of a row take part (zero-padded channel layout of ragged pruned models);
  // the padding columns are written as zeros.
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const T* xr = x + (long long)warp * D;
  float v[NPL];
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < NPL; ++k) {
    const int i = lane + 32 * k;
    v[k] = (i < Dv) ? to_f(xr[i]) : 0.f;
    s += v[k];
  }
  s = warp_sum(s);
  const float mean = s / Dv;
  float q = 0.f;
#pragma unroll
  for (int k = 0; k < NPL; ++k) {
    const int i = lane + 32 * k;
    float d = (i < Dv) ? v[k] - mean : 0.f;
    q += d * d;
  }
  q = warp_sum(q);
  const float rstd = rsqrtf(q / Dv + eps);
  if (lane == 0) {
    if (mean_out) mean_out[warp] = mean;
    if (rstd_out) rstd_out[warp] = rstd;
  }
  const uint32_t thresh = (uint32_t)(drop_p * 4294967296.0);
  const float inv_keep = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
  T* yr = y + (long long)warp * D;
#pragma unroll
  for (int k = 0; k < NPL; ++k) {
    const int i = lane + 32 * k;
    if (i >= D) break;
    if (i >= Dv) {
      if (y_pre) y_pre[(long long)warp * D + i] = from_f<T>(0.f);
      yr[i] = from_f<T>(0.f);
      continue;
    }
    float o = (v[k] - mean) * rstd * __ldg(w + i) + __ldg(b + i);
    if (y_pre) y_pre[(long long)warp * D + i] = from_f<T>(o);
    if (act == 1) o = gelu_erf(o);
    if (colmask) o *= __ldg(colmask + i);
    if (drop_p > 0.f) o *= dropout_scale(seed, (unsigned long long)warp * D + i, thresh, inv_keep);
    yr[i] = from_f<T>(o);
  }
}

// Backward. dy_total = (dy + dy2) passed through dropout / colmask / GELU' first if those were fused in forward.
// Per-CTA partial sums of dw / db / dmask are reduced in shared memory, then atomically added to the fp32 outputs.
template <typename T, int NPL>
__global__ void __launch_bounds__(256)
    layernorm_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ dy2, const T* __restrict__ x,
                         const float* __restrict__ w, const float* __restrict__ b, const float* __restrict__ mean_in,
                         const float* __restrict__ rstd_in, T* __restrict__ dx, float* __restrict__ dw,
                         float* __restrict__ db, const float* __restrict__ colmask, float* __restrict__ dmask, int rows,
                         int D, int Dv, int act, float drop_p, unsigned long long seed, int rows_per_warp,
                         int rows_per_batch, long long out_bs, int out_row_off) {
  extern __shared__ float sh[];  // [3][D] : dw, db, dmask partials
  float* s_dw = sh;
  float* s_db = sh + D;
  float* s_dm = sh + 2 * D;
  for (int i = threadIdx.x; i < 3 * D; i += blockDim.x) sh[i] = 0.f;
  __syncthreads();
  const int warp_in_blk = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int warps_per_blk = blockDim.x >> 5;
  const int gwarp = blockIdx.x * warps_per_blk + warp_in_blk;
  const uint32_t thresh = (uint32_t)(drop_p * 4294967296.0);
  const float inv_keep = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
  float a_dw[NPL], a_db[NPL], a_dm[NPL];
#pragma unroll
  for (int k = 0; k < NPL; ++k) a_dw[k] = a_db[k] = a_dm[k] = 0.f;

  for (int rr = 0; rr < rows_per_warp; ++rr) {
    const int row = gwarp * rows_per_warp + rr;
    if (row >= rows) break;
    const long long off = (long long)row * D;
    const float mean = mean_in[row], rstd = rstd_in[row];
    float g[NPL], xh[NPL];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int k = 0; k < NPL; ++k) {
      const int i = lane + 32 * k;
      g[k] = 0.f;
      xh[k] = 0.f;
      if (i >= Dv) continue;
      float d = to_f(dy[off + i]);
      if (dy2) d += to_f(dy2[off + i]);
      float xhat = (to_f(x[off + i]) - mean) * rstd;
      if (drop_p > 0.f) d *= dropout_scale(seed, (unsigned long long)off + i, thresh, inv_keep);
      if (act == 1 || colmask) {
        float u = xhat * __ldg(w + i) + __ldg(b + i);
        float m = colmask ? __ldg(colmask + i) : 1.f;
        if (act == 1) {
          if (colmask) a_dm[k] += d * gelu_erf(u);
          d = d * m * gelu_erf_grad(u);
        } else {
          a_dm[k] += d * u;
          d = d * m;
        }
      }
      a_dw[k] += d * xhat;
      a_db[k] += d;
      float gg = d * __ldg(w + i);
      g[k] = gg;
      xh[k] = xhat;
      s1 += gg;
      s2 += gg * xhat;
    }
    s1 = warp_sum(s1) / Dv;
    s2 = warp_sum(s2) / Dv;
    const int ob = row / rows_per_batch;
    const long long ooff = (long long)ob * out_bs + (long long)(row - ob * rows_per_batch + out_row_off) * D;
#pragma unroll
    for (int k = 0; k < NPL; ++k) {
      const int i = lane + 32 * k;
      if (i < D) dx[ooff + i] = from_f<T>(i < Dv ? rstd * (g[k] - s1 - xh[k] * s2) : 0.f);
    }
  }
#pragma unroll
  for (int k = 0; k < NPL; ++k) {
    const int i = lane + 32 * k;
    if (i >= Dv) break;
    atomicAdd(&s_dw[i], a_dw[k]);
    atomicAdd(&s_db[i], a_db[k]);
    if (dmask) atomicAdd(&s_dm[i], a_dm[k]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < D; i += blockDim.x) {
    if (dw) atomicAdd(dw + i, s_dw[i]);
    if (db) atomicAdd(db + i, s_db[i]);
    if (dmask) atomicAdd(dmask + i, s_dm[i]);
  }
}


// ---------------------------------------------------------------------------------------------------------------
// Vectorised variants for the hot shapes (D a multiple of 32 * VEC, plain LayerNorm with optional dropout):
// 128-bit accesses, one warp per row, persistent grid; the backward reduces dw / db across the warps of a CTA in
// shared memory (no shared-memory atomics) and issues one coalesced RED per column per CTA.
// ---------------------------------------------------------------------------------------------------------------
template <typename T, int Q>
__global__ void __launch_bounds__(256)
    layernorm_fwd_vec_kernel(const T* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                             T* __restrict__ y, float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows,
                             int D, float eps, float drop_p, unsigned long long seed) {
  constexpr int VEC = VecOf<T>::N;
  const int lane = threadIdx.x & 31;
  const int warps_per_blk = blockDim.x >> 5;
  const uint32_t thresh = (uint32_t)(drop_p * 4294967296.0);
  const float inv_keep = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
  pdl_wait();
  pdl_trigger();
  // scale / shift of this lane's columns: row-invariant, requested before the row data so that their latency hides
  // behind the statistics instead of following them
  float wv[Q][VEC], bv[Q][VEC];
#pragma unroll
  for (int q = 0; q < Q; ++q) {
    load_vec_f32(w + (lane + 32 * q) * VEC, wv[q]);
    load_vec_f32(b + (lane + 32 * q) * VEC, bv[q]);
  }
  for (int row = blockIdx.x * warps_per_blk + (threadIdx.x >> 5); row < rows; row += gridDim.x * warps_per_blk) {
    const T* xr = x + (long long)row * D;
    float v[Q][VEC];
    float s = 0.f;
#pragma unroll
    for (int q = 0; q < Q; ++q) {
      load_vec(xr + (lane + 32 * q) * VEC, v[q]);
#pragma unroll
      for (int j = 0; j < VEC; ++j) s += v[q][j];
    }
    const float mean = warp_sum(s) / D;
    float qs = 0.f;
#pragma unroll
    for (int q = 0; q < Q; ++q)
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        const float d = v[q][j] - mean;
        qs += d * d;
      }
    const float rstd = rsqrtf(warp_sum(qs) / D + eps);
    if (lane == 0) {
      if (mean_out) mean_out[row] = mean;
      if (rstd_out) rstd_out[row] = rstd;
    }
    T* yr = y + (long long)row * D;
#pragma unroll
    for (int q = 0; q < Q; ++q) {
      const int c0 = (lane + 32 * q) * VEC;
      float o[VEC];
#pragma unroll
      for (int j = 0; j < VEC; ++j) o[j] = (v[q][j] - mean) * rstd * wv[q][j] + bv[q][j];
      if (drop_p > 0.f) dropout_apply_vec<VEC>(seed, (unsigned long long)row * D + c0, thresh, inv_keep, o);
      store_vec(yr + c0, o);
    }
  }
}

constexpr int LNB_WARPS = 16;

template <typename T, int Q>
__global__ void __launch_bounds__(LNB_WARPS * 32, 1)
    layernorm_bwd_vec_kernel(const T* __restrict__ dy, const T* __restrict__ dy2, const T* __restrict__ x,
                             const float* __restrict__ w, const float* __restrict__ mean_in,
                             const float* __restrict__ rstd_in, T* __restrict__ dx, float* __restrict__ dw,
                             float* __restrict__ db, int rows, int D, float drop_p, unsigned long long seed) {
  constexpr int VEC = VecOf<T>::N;
  extern __shared__ float sh[];  // [LNB_WARPS][D]
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const uint32_t thresh = (uint32_t)(drop_p * 4294967296.0);
  const float inv_keep = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
  float a_dw[Q][VEC], a_db[Q][VEC], wv[Q][VEC];
  pdl_wait();
  pdl_trigger();
#pragma unroll
  for (int q = 0; q < Q; ++q) {
    load_vec_f32(w + (lane + 32 * q) * VEC, wv[q]);
#pragma unroll
    for (int j = 0; j < VEC; ++j) a_dw[q][j] = a_db[q][j] = 0.f;
  }
  for (int row = blockIdx.x * LNB_WARPS + wid; row < rows; row += gridDim.x * LNB_WARPS) {
    const long long off = (long long)row * D;
    const float mean = mean_in[row], rstd = rstd_in[row];
    float g[Q][VEC], xh[Q][VEC];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int q = 0; q < Q; ++q) {
      const int c0 = (lane + 32 * q) * VEC;
      load_vec(dy + off + c0, g[q]);
      load_vec(x + off + c0, xh[q]);
      if (dy2) {
        float t[VEC];
        load_vec(dy2 + off + c0, t);
#pragma unroll
        for (int j = 0; j < VEC; ++j) g[q][j] += t[j];
      }
      if (drop_p > 0.f) dropout_apply_vec<VEC>(seed, (unsigned long long)off + c0, thresh, inv_keep, g[q]);
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        const float d = g[q][j];
        const float xhat = (xh[q][j] - mean) * rstd;
        a_dw[q][j] += d * xhat;
        a_db[q][j] += d;
        const float gg = d * wv[q][j];
        g[q][j] = gg;
        xh[q][j] = xhat;
        s1 += gg;
        s2 += gg * xhat;
      }
    }
    s1 = warp_sum(s1) / D;
    s2 = warp_sum(s2) / D;
#pragma unroll
    for (int q = 0; q < Q; ++q) {
      float o[VEC];
#pragma unroll
      for (int j = 0; j < VEC; ++j) o[j] = rstd * (g[q][j] - s1 - xh[q][j] * s2);
      store_vec(dx + off + (lane + 32 * q) * VEC, o);
    }
  }
  // cross-warp reduction of the parameter gradients, one quantity at a time through [warps][D] shared memory
#pragma unroll
  for (int which = 0; which < 2; ++which) {
    float* outp = which == 0 ? dw : db;
    __syncthreads();
#pragma unroll
    for (int q = 0; q < Q; ++q)
#pragma unroll
      for (int j = 0; j < VEC; ++j) sh[wid * D + (lane + 32 * q) * VEC + j] = which == 0 ? a_dw[q][j] : a_db[q][j];
    __syncthreads();
    if (outp) {
      for (int c = threadIdx.x; c < D; c += blockDim.x) {
        float t = 0.f;
#pragma unroll
        for (int q = 0; q < LNB_WARPS; ++q) t += sh[q * D + c];
        atomicAdd(outp + c, t);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Vectorised variants WITH the conv-block fusions of the Large extractor (components.py:54-61, 108-114):
//   forward   y = colmask * GELU(LN(x))
//   backward  through mask / GELU / LN in one pass, mask gradient reduced like dw / db, output rows optionally
//             re-mapped into the zero-padded per-utterance layout the strided-conv dgrad GEMM reads.
// These tensors are the largest of the whole Large step (8 x 47999 x 512 in layer 0); the generic one-scalar-per-lane
// kernels ran them at a fifth of the memory bandwidth.  bf16 uses the tanh-form GELU of the other bf16 kernels
// (error below bf16 resolution; one MUFU per element keeps the kernel memory-bound), fp32 the erf form.
// ---------------------------------------------------------------------------------------------------------------
template <typename T>
DPH_DEVICE void gelu_pair_fwd(float& a, float& b) {
  if constexpr (sizeof(T) == 2) {
    upk2(gelu_tanh2(pk2(a, b)), a, b);
  } else {
    a = gelu_erf(a);
    b = gelu_erf(b);
  }
}
template <typename T>
DPH_DEVICE void gelu_pair_bwd(float ua, float ub, float& ga, float& gb, float& da, float& db) {
  if constexpr (sizeof(T) == 2) {
    f32x2 g2, d2;
    gelu_and_grad_tanh2(pk2(ua, ub), g2, d2);
    upk2(g2, ga, gb);
    upk2(d2, da, db);
  } else {
    ga = gelu_erf(ua);
    gb = gelu_erf(ub);
    da = gelu_erf_grad(ua);
    db = gelu_erf_grad(ub);
  }
}

template <typename T, int Q>
__global__ void __launch_bounds__(256)
    layernorm_fwd_vecf_kernel(const T* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                              T* __restrict__ y, float* __restrict__ mean_out, float* __restrict__ rstd_out,
                              const float* __restrict__ colmask, int rows, int D, float eps, int act) {
  constexpr int VEC = VecOf<T>::N;
  const int lane = threadIdx.x & 31;
  const int warps_per_blk = blockDim.x >> 5;
  pdl_wait();
  pdl_trigger();
  float wv[Q][VEC], bv[Q][VEC], mv[Q][VEC];
#pragma unroll
  for (int q = 0; q < Q; ++q) {
    load_vec_f32(w + (lane + 32 * q) * VEC, wv[q]);
    load_vec_f32(b + (lane + 32 * q) * VEC, bv[q]);
    if (colmask) {
      load_vec_f32(colmask + (lane + 32 * q) * VEC, mv[q]);
    } else {
#pragma unroll
      for (int j = 0; j < VEC; ++j) mv[q][j] = 1.f;
    }
  }
  for (int row = blockIdx.x * warps_per_blk + (threadIdx.x >> 5); row < rows; row += gridDim.x * warps_per_blk) {
    const T* xr = x + (long long)row * D;
    float v[Q][VEC];
    float s = 0.f;
#pragma unroll
    for (int q = 0; q < Q; ++q) {
      load_vec(xr + (lane + 32 * q) * VEC, v[q]);
#pragma unroll
      for (int j = 0; j < VEC; ++j) s += v[q][j];
    }
    const float mean = warp_sum(s) / D;
    float qs = 0.f;
#pragma unroll
    for (int q = 0; q < Q; ++q)
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        const float d = v[q][j] - mean;
        qs += d * d;
      }
    const float rstd = rsqrtf(warp_sum(qs) / D + eps);
    if (lane == 0) {
      if (mean_out) mean_out[row] = mean;
      if (rstd_out) rstd_out[row] = rstd;
    }
    T* yr = y + (long long)row * D;
#pragma unroll
    for (int q = 0; q < Q; ++q) {
      float o[VEC];
#pragma unroll
      for (int j = 0; j < VEC; ++j) o[j] = (v[q][j] - mean) * rstd * wv[q][j] + bv[q][j];
      if (act == 1) {
#pragma unroll
        for (int j = 0; j < VEC; j += 2) gelu_pair_fwd<T>(o[j], o[j + 1]);
      }
#pragma unroll
      for (int j = 0; j < VEC; ++j) o[j] *= mv[q][j];
      store_vec(yr + (lane + 32 * q) * VEC, o);
    }
  }
}

template <typename T, int Q>
__global__ void __launch_bounds__(LNB_WARPS * 32, 1)
    layernorm_bwd_vecf_kernel(const T* __restrict__ dy, const T* __restrict__ x, const float* __restrict__ w,
                              const float* __restrict__ b, const float* __restrict__ mean_in,
                              const float* __restrict__ rstd_in, T* __restrict__ dx, float* __restrict__ dw,
                              float* __restrict__ db, const float* __restrict__ colmask, float* __restrict__ dmask,
                              int rows, int D, int act, int rows_per_batch, long long out_bs, int out_row_off) {
  constexpr int VEC = VecOf<T>::N;
  extern __shared__ float sh[];  // [LNB_WARPS][D]
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  // (512 threads -> 128 registers: scale, shift and mask are re-read per row through L1 instead of living in registers)
  float a_dw[Q][VEC], a_db[Q][VEC], a_dm[Q][VEC];
  pdl_wait();
  pdl_trigger();
#pragma unroll
  for (int q = 0; q < Q; ++q) {
#pragma unroll
    for (int j = 0; j < VEC; ++j) a_dw[q][j] = a_db[q][j] = a_dm[q][j] = 0.f;
  }
  for (int row = blockIdx.x * LNB_WARPS + wid; row < rows; row += gridDim.x * LNB_WARPS) {
    const long long off = (long long)row * D;
    const float mean = mean_in[row], rstd = rstd_in[row];
    float g[Q][VEC], xh[Q][VEC];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int q = 0; q < Q; ++q) {
      const int c0 = (lane + 32 * q) * VEC;
      load_vec(dy + off + c0, g[q]);
      load_vec(x + off + c0, xh[q]);
      float wvq[VEC], bvq[VEC], mvq[VEC];
      load_vec_f32(w + c0, wvq);
      load_vec_f32(b + c0, bvq);
      if (colmask) {
        load_vec_f32(colmask + c0, mvq);
      } else {
#pragma unroll
        for (int j = 0; j < VEC; ++j) mvq[j] = 1.f;
      }
#pragma unroll
      for (int j = 0; j < VEC; j += 2) {
        const float xa = (xh[q][j] - mean) * rstd, xb = (xh[q][j + 1] - mean) * rstd;
        const float ua = xa * wvq[j] + bvq[j], ub = xb * wvq[j + 1] + bvq[j + 1];
        float ga = ua, gb = ub, da = 1.f, dbb = 1.f;
        if (act == 1) gelu_pair_bwd<T>(ua, ub, ga, gb, da, dbb);
        float d0 = g[q][j], d1 = g[q][j + 1];
        a_dm[q][j] += d0 * ga;            // d mask = sum dy * act(u)
        a_dm[q][j + 1] += d1 * gb;
        d0 *= mvq[j] * da;                // through the mask and the activation
        d1 *= mvq[j + 1] * dbb;
        a_dw[q][j] += d0 * xa;
        a_dw[q][j + 1] += d1 * xb;
        a_db[q][j] += d0;
        a_db[q][j + 1] += d1;
        const float g0 = d0 * wvq[j], g1 = d1 * wvq[j + 1];
        g[q][j] = g0;
        g[q][j + 1] = g1;
        xh[q][j] = xa;
        xh[q][j + 1] = xb;
        s1 += g0 + g1;
        s2 += g0 * xa + g1 * xb;
      }
    }
    s1 = warp_sum(s1) / D;
    s2 = warp_sum(s2) / D;
    const int ob = row / rows_per_batch;
    const long long ooff = (long long)ob * out_bs + (long long)(row - ob * rows_per_batch + out_row_off) * D;
#pragma unroll
    for (int q = 0; q < Q; ++q) {
      float o[VEC];
#pragma unroll
      for (int j = 0; j < VEC; ++j) o[j] = rstd * (g[q][j] - s1 - xh[q][j] * s2);
      store_vec(dx + ooff + (lane + 32 * q) * VEC, o);
    }
  }
#pragma unroll
  for (int which = 0; which < 3; ++which) {
    float* outp = which == 0 ? dw : (which == 1 ? db : dmask);
    __syncthreads();
#pragma unroll
    for (int q = 0; q < Q; ++q)
#pragma unroll
      for (int j = 0; j < VEC; ++j)
        sh[wid * D + (lane + 32 * q) * VEC + j] = which == 0 ? a_dw[q][j] : (which == 1 ? a_db[q][j] : a_dm[q][j]);
    __syncthreads();
    if (outp) {
      for (int c = threadIdx.x; c < D; c += blockDim.x) {
        float t = 0.f;
#pragma unroll
        for (int q = 0; q < LNB_WARPS; ++q) t += sh[q * D + c];
        atomicAdd(outp + c, t);
      }
    }
  }
}

}  // namespace dph

using namespace dph;

extern "C" int dph_layernorm_fwd(const void* x, const float* w, const float* b, void* y, void* y_pre, float* mean,
                                 float* rstd, const float* colmask, int rows, int D, int d_valid, float eps, int act,
                                 float drop_p, uint64_t seed, int dtype, void* stream) {
  if (d_valid <= 0) d_valid = D;
  if (!x || !w || !b || !y || rows <= 0 || D <= 0 || D > 32 * LN_MAX_PER_LANE || d_valid > D) {
    set_error("layernorm_fwd: bad arguments (D=%d)", D);
    return DPH_ERR_ARG;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int warps_per_blk = 8;
  {
    // vectorised fast path
    const int vec = dtype == DPH_DTYPE_BF16 ? 8 : 4;
    const bool al = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(w) |
                      reinterpret_cast<uintptr_t>(b)) & 15) == 0;
    const int q = D / (32 * vec);
    if (al && !y_pre && !colmask && act == 0 && d_valid == D && D % (32 * vec) == 0 && q >= 1 && q <= 4) {
      int g = (rows + warps_per_blk - 1) / warps_per_blk;
      if (g > 148 * 8) g = 148 * 8;
#define LN_FV(T, Q)                                                                                                  \
  launch_pdl(layernorm_fwd_vec_kernel<T, Q>, dim3(g), dim3(256), 0, s, (const T*)x, w, b, (T*)y, mean, rstd, rows, D,  \
             eps, drop_p, (unsigned long long)seed)
      if (dtype == DPH_DTYPE_BF16) {
        if (q == 1) LN_FV(bf16, 1); else if (q == 2) LN_FV(bf16, 2); else if (q == 3) LN_FV(bf16, 3); else LN_FV(bf16, 4);
      } else {
        if (q == 1) LN_FV(float, 1); else if (q == 2) LN_FV(float, 2); else if (q == 3) LN_FV(float, 3); else LN_FV(float, 4);
      }
#undef LN_FV
      dph_count_launch();
      return dph_check_last("layernorm_fwd_vec");
    }
    // the same with GELU / column mask fused (conv blocks of the Large extractor)
    const bool alm = !colmask || (reinterpret_cast<uintptr_t>(colmask) & 15) == 0;
    if (al && alm && !y_pre && (colmask || act == 1) && drop_p == 0.f && d_valid == D && D % (32 * vec) == 0 &&
        q >= 1 && q <= 4) {
      int g = (rows + warps_per_blk - 1) / warps_per_blk;
      if (g > 148 * 8) g = 148 * 8;
#define LN_FF(T, Q)                                                                                                   \
  launch_pdl(layernorm_fwd_vecf_kernel<T, Q>, dim3(g), dim3(256), 0, s, (const T*)x, w, b, (T*)y, mean, rstd, colmask, \
             rows, D, eps, act)
      if (dtype == DPH_DTYPE_BF16) {
        if (q == 1) LN_FF(bf16, 1); else if (q == 2) LN_FF(bf16, 2); else if (q == 3) LN_FF(bf16, 3); else LN_FF(bf16, 4);
      } else {
        if (q == 1) LN_FF(float, 1); else if (q == 2) LN_FF(float, 2); else if (q == 3) LN_FF(float, 3); else LN_FF(float, 4);
      }
#undef LN_FF
      dph_count_launch();
      return dph_check_last("layernorm_fwd_vecf");
    }
  }
  const int grid = (rows + warps_per_blk - 1) / warps_per_blk;
#define LN_FWD(NPL)                                                                                              \
  do {                                                                                                           \
    if (dtype == DPH_DTYPE_BF16)                                                                                 \
      layernorm_fwd_kernel<bf16, NPL><<<grid, 256, 0, s>>>((const bf16*)x, w, b, (bf16*)y, (bf16*)y_pre, mean,   \
                                                           rstd, colmask, rows, D, d_valid, eps, act, drop_p,    \
                                                           seed);                                                \
    else                                                                                                         \
      layernorm_fwd_kernel<float, NPL><<<grid, 256, 0, s>>>((const float*)x, w, b, (float*)y, (float*)y_pre,     \
                                                            mean, rstd, colmask, rows, D, d_valid, eps, act,       \
                                                            drop_p, seed);                                       \
  } while (0)
  if (D <= 512) LN_FWD(16);
  else if (D <= 768) LN_FWD(24);
  else if (D <= 1024) LN_FWD(32);
  else LN_FWD(LN_MAX_PER_LANE);
#undef LN_FWD
  dph_count_launch();
  return dph_check_last("layernorm_fwd");
}

extern "C" int dph_layernorm_bwd(const void* dy, const void* dy2, const void* x, const float* w, const float* b,
                                 const float* mean, const float* rstd, void* dx, float* dw, float* db,
                                 const float* colmask, float* dmask, int rows, int D, int d_valid, int act,
                                 float drop_p, uint64_t seed, int rows_per_batch, int64_t out_bs, int out_row_off,
                                 int dtype, void* stream) {
  if (d_valid <= 0) d_valid = D;
  if (rows_per_batch <= 0) {
    rows_per_batch = rows;
    out_bs = 0;
    out_row_off = 0;
  }
  if (!dy || !x || !w || !b || !mean || !rstd || !dx || rows <= 0 || D <= 0 || D > 32 * LN_MAX_PER_LANE ||
      d_valid > D) {
    set_error("layernorm_bwd: bad arguments (D=%d)", D);
    return DPH_ERR_ARG;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int warps_per_blk = 8;
  {
    const int vec = dtype == DPH_DTYPE_BF16 ? 8 : 4;
    const bool al = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(dx) |
                      reinterpret_cast<uintptr_t>(dy2) | reinterpret_cast<uintptr_t>(w)) & 15) == 0;
    const int q = D / (32 * vec);
    const bool fp32_ok = dtype == DPH_DTYPE_BF16 || q <= 4;
    if (al && !colmask && !dmask && act == 0 && d_valid == D && rows_per_batch == rows && out_row_off == 0 &&
        D % (32 * vec) == 0 &&
        q >= 1 && q <= (dtype == DPH_DTYPE_BF16 ? 4 : 8) && fp32_ok) {
      int g = (rows + LNB_WARPS - 1) / LNB_WARPS;
      if (g > 148) g = 148;
      const size_t smem = (size_t)LNB_WARPS * D * sizeof(float);
#define LN_BV(T, Q)                                                                                                   \
  do {                                                                                                                \
    static bool attr = false;                                                                                         \
    if (!attr) {                                                                                                      \
      cudaFuncSetAttribute(layernorm_bwd_vec_kernel<T, Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);   \
      attr = true;                                                                                                    \
    }                                                                                                                 \
    launch_pdl(layernorm_bwd_vec_kernel<T, Q>, dim3(g), dim3(LNB_WARPS * 32), smem, s, (const T*)dy, (const T*)dy2,   \
               (const T*)x, w, mean, rstd, (T*)dx, dw, db, rows, D, drop_p, (unsigned long long)seed);               \
  } while (0)
      if (dtype == DPH_DTYPE_BF16) {
        if (q == 1) LN_BV(bf16, 1); else if (q == 2) LN_BV(bf16, 2); else if (q == 3) LN_BV(bf16, 3); else LN_BV(bf16, 4);
      } else {
        if (q == 1) LN_BV(float, 1); else if (q == 2) LN_BV(float, 2); else if (q == 3) LN_BV(float, 3);
        else if (q == 4) LN_BV(float, 4); else if (q == 6) LN_BV(float, 6); else LN_BV(float, 8);
      }
#undef LN_BV
      dph_count_launch();
      return dph_check_last("layernorm_bwd_vec");
    }
    // fused variant: mask / GELU backward, mask gradient, re-mapped output rows (no dropout, no second gradient)
    const bool alm = !colmask || (reinterpret_cast<uintptr_t>(colmask) & 15) == 0;
    const bool rows_al = (out_bs % vec == 0);
    if (al && alm && rows_al && !dy2 && drop_p == 0.f && d_valid == D && D % (32 * vec) == 0 && q >= 1 &&
        q * vec <= 16 /* <= 16 elements per lane: D <= 512 */ && (size_t)LNB_WARPS * D * sizeof(float) <= 96 * 1024) {
      int g = (rows + LNB_WARPS - 1) / LNB_WARPS;
      if (g > 148) g = 148;
      const size_t smem = (size_t)LNB_WARPS * D * sizeof(float);
#define LN_BF(T, Q)                                                                                                   \
  do {                                                                                                                \
    static bool attr = false;                                                                                         \
    if (!attr) {                                                                                                      \
      cudaFuncSetAttribute(layernorm_bwd_vecf_kernel<T, Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);  \
      attr = true;                                                                                                    \
    }                                                                                                                 \
    launch_pdl(layernorm_bwd_vecf_kernel<T, Q>, dim3(g), dim3(LNB_WARPS * 32), smem, s, (const T*)dy, (const T*)x, w, \
               b, mean, rstd, (T*)dx, dw, db, colmask, dmask, rows, D, act, rows_per_batch, (long long)out_bs,        \
               out_row_off);                                                                                          \
  } while (0)
      if (dtype == DPH_DTYPE_BF16) {
        if (q == 1) LN_BF(bf16, 1); else if (q == 2) LN_BF(bf16, 2); else if (q == 3) LN_BF(bf16, 3); else LN_BF(bf16, 4);
      } else {
        if (q == 1) LN_BF(float, 1); else if (q == 2) LN_BF(float, 2); else if (q == 3) LN_BF(float, 3); else LN_BF(float, 4);
      }
#undef LN_BF
      dph_count_launch();
      return dph_check_last("layernorm_bwd_vecf");
    }
  }
  // ~4 CTAs per SM worth of work; each warp walks rows_per_warp consecutive rows
  int target_warps = 148 * 4 * warps_per_blk;
  int rows_per_warp = (rows + target_warps - 1) / target_warps;
  if (rows_per_warp < 1) rows_per_warp = 1;
  const int nwarps = (rows + rows_per_warp - 1) / rows_per_warp;
  const int grid = (nwarps + warps_per_blk - 1) / warps_per_blk;
  const size_t smem = 3 * (size_t)D * sizeof(float);
#define LN_BWD(NPL)                                                                                              \
  do {                                                                                                           \
    if (dtype == DPH_DTYPE_BF16)                                                                                 \
      layernorm_bwd_kernel<bf16, NPL><<<grid, 256, smem, s>>>((const bf16*)dy, (const bf16*)dy2, (const bf16*)x, \
                                                              w, b, mean, rstd, (bf16*)dx, dw, db, colmask,      \
                                                              dmask, rows, D, d_valid, act, drop_p, seed,       \
                                                              rows_per_warp,                                     \
                                                              rows_per_batch, out_bs, out_row_off);              \
    else                                                                                                         \
      layernorm_bwd_kernel<float, NPL><<<grid, 256, smem, s>>>((const float*)dy, (const float*)dy2,              \
                                                               (const float*)x, w, b, mean, rstd, (float*)dx,    \
                                                               dw, db, colmask, dmask, rows, D, d_valid, act,   \
                                                               drop_p, seed, rows_per_warp, rows_per_batch,      \
                                                               out_bs, out_row_off);                             \
  } while (0)
  if (D <= 512) LN_BWD(16);
  else if (D <= 768) LN_BWD(24);
  else if (D <= 1024) LN_BWD(32);
  else LN_BWD(LN_MAX_PER_LANE);
#undef LN_BWD
  dph_count_launch();
  return dph_check_last("layernorm_bwd");
}
