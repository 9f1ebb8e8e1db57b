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
#include "../../include/dphubert_b200.h"

namespace dph {
void set_error(const char* fmt, ...);

constexpr int LN_MAX_PER_LANE = 40;  // supports D <= 32 * 40 = 1280

template <typename T, int NPL>
__global__ void __launch_bounds__(256)
    layernorm_fwd_kernel(const T* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                         T* __restrict__ y, T* __restrict__ y_pre, float* __restrict__ mean_out,
                         float* __restrict__ rstd_out, const float* __restrict__ colmask, int rows, int D, float eps,
                         int act, float drop_p, unsigned long long seed) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const T* xr = x + (long long)warp * D;
  float v[NPL];
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < NPL; ++k) {
    const int i = lane + 32 * k;
    v[k] = (i < D) ? to_f(xr[i]) : 0.f;
    s += v[k];
  }
  s = warp_sum(s);
  const float mean = s / D;
  float q = 0.f;
#pragma unroll
  for (int k = 0; k < NPL; ++k) {
    const int i = lane + 32 * k;
    float d = (i < D) ? v[k] - mean : 0.f;
    q += d * d;
  }
  q = warp_sum(q);
  const float rstd = rsqrtf(q / D + eps);
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
                         int D, int act, float drop_p, unsigned long long seed, int rows_per_warp,
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
      if (i >= D) continue;
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
    s1 = warp_sum(s1) / D;
    s2 = warp_sum(s2) / D;
    const int ob = row / rows_per_batch;
    const long long ooff = (long long)ob * out_bs + (long long)(row - ob * rows_per_batch + out_row_off) * D;
#pragma unroll
    for (int k = 0; k < NPL; ++k) {
      const int i = lane + 32 * k;
      if (i < D) dx[ooff + i] = from_f<T>(rstd * (g[k] - s1 - xh[k] * s2));
    }
  }
#pragma unroll
  for (int k = 0; k < NPL; ++k) {
    const int i = lane + 32 * k;
    if (i >= D) break;
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

}  // namespace dph

using namespace dph;

extern "C" int dph_layernorm_fwd(const void* x, const float* w, const float* b, void* y, void* y_pre, float* mean,
                                 float* rstd, const float* colmask, int rows, int D, float eps, int act, float drop_p,
                                 uint64_t seed, int dtype, void* stream) {
  if (!x || !w || !b || !y || rows <= 0 || D <= 0 || D > 32 * LN_MAX_PER_LANE) {
    set_error("layernorm_fwd: bad arguments (D=%d)", D);
    return DPH_ERR_ARG;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int warps_per_blk = 8;
  const int grid = (rows + warps_per_blk - 1) / warps_per_blk;
#define LN_FWD(NPL)                                                                                              \
  do {                                                                                                           \
    if (dtype == DPH_DTYPE_BF16)                                                                                 \
      layernorm_fwd_kernel<bf16, NPL><<<grid, 256, 0, s>>>((const bf16*)x, w, b, (bf16*)y, (bf16*)y_pre, mean,   \
                                                           rstd, colmask, rows, D, eps, act, drop_p, seed);      \
    else                                                                                                         \
      layernorm_fwd_kernel<float, NPL><<<grid, 256, 0, s>>>((const float*)x, w, b, (float*)y, (float*)y_pre,     \
                                                            mean, rstd, colmask, rows, D, eps, act, drop_p, seed); \
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
                                 const float* colmask, float* dmask, int rows, int D, int act, float drop_p,
                                 uint64_t seed, int rows_per_batch, int64_t out_bs, int out_row_off, int dtype,
                                 void* stream) {
  if (rows_per_batch <= 0) {
    rows_per_batch = rows;
    out_bs = 0;
    out_row_off = 0;
  }
  if (!dy || !x || !w || !b || !mean || !rstd || !dx || rows <= 0 || D <= 0 || D > 32 * LN_MAX_PER_LANE) {
    set_error("layernorm_bwd: bad arguments (D=%d)", D);
    return DPH_ERR_ARG;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int warps_per_blk = 8;
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
                                                              dmask, rows, D, act, drop_p, seed, rows_per_warp, \
                                                              rows_per_batch, out_bs, out_row_off);              \
    else                                                                                                         \
      layernorm_bwd_kernel<float, NPL><<<grid, 256, smem, s>>>((const float*)dy, (const float*)dy2,              \
                                                               (const float*)x, w, b, mean, rstd, (float*)dx,    \
                                                               dw, db, colmask, dmask, rows, D, act, drop_p,     \
                                                               seed, rows_per_warp, rows_per_batch, out_bs,      \
                                                               out_row_off);                                     \
  } while (0)
  if (D <= 512) LN_BWD(16);
  else if (D <= 768) LN_BWD(24);
  else if (D <= 1024) LN_BWD(32);
  else LN_BWD(LN_MAX_PER_LANE);
#undef LN_BWD
  dph_count_launch();
  return dph_check_last("layernorm_bwd");
}
