// Layer-wise distillation loss (MSE + L1 + cosine) as two fused reduction kernels.
//
// Replaces lightning.py:116-139 (DistillLoss.forward: nn.MSELoss, nn.L1Loss, nn.CosineSimilarity(dim=-1) with
// eps 1e-8, "raw" or "log_sig" cosine variant) and its autograd backward: ~8 elementwise passes over
// 2 x (B, L, T, D) become one read of both tensors forward and one read + one write backward.
//
// One warp per feature row (D <= 1280).  Forward writes per-row (dot, |a|, |b|) so the backward does not have to
// re-reduce, and accumulates the three loss components into acc[0..2] (fp32 atomics, one per CTA).
#include "common.cuh"
#include "vec.cuh"
#include "../../include/dphubert_b200.h"

namespace dph {
void set_error(const char* fmt, ...);

constexpr int DL_NPL = 40;

template <typename T>
__global__ void __launch_bounds__(256)
    distill_loss_fwd_kernel(const T* __restrict__ a, const T* __restrict__ b, float* __restrict__ rowstat,
                            float* __restrict__ acc, long long rows, int D, int cos_type, float eps, int vec) {
  pdl_wait();
  pdl_trigger();
  __shared__ float red[3][8];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const long long row = (long long)blockIdx.x * 8 + w;
  float l2 = 0.f, l1 = 0.f, cosv = 0.f;
  if (row < rows) {
    const T* ar = a + row * D;
    const T* br = b + row * D;
    float dot = 0.f, na = 0.f, nb = 0.f;
    constexpr int VEC = VecOf<T>::N;
    if (vec) {  // 16-byte loads (D a multiple of the vector width, 16-byte aligned rows)
      for (int c0 = lane * VEC; c0 < D; c0 += 32 * VEC) {
        float xa[VEC], xb[VEC];
        load_vec(ar + c0, xa);
        load_vec(br + c0, xb);
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
          const float d = xa[j] - xb[j];
          l2 = fmaf(d, d, l2);
          l1 += fabsf(d);
          dot = fmaf(xa[j], xb[j], dot);
          na = fmaf(xa[j], xa[j], na);
          nb = fmaf(xb[j], xb[j], nb);
        }
      }
    } else {
      for (int i = lane; i < D; i += 32) {
        const float x = to_f(ar[i]), y = to_f(br[i]);
        const float d = x - y;
        l2 += d * d;
        l1 += fabsf(d);
        dot += x * y;
        na += x * x;
        nb += y * y;
      }
    }
    l2 = warp_sum(l2);
    l1 = warp_sum(l1);
    dot = warp_sum(dot);
    na = sqrtf(warp_sum(na));
    nb = sqrtf(warp_sum(nb));
    const float c = dot / (fmaxf(na, eps) * fmaxf(nb, eps));
    if (lane == 0) {
      rowstat[row * 3 + 0] = dot;
      rowstat[row * 3 + 1] = na;
      rowstat[row * 3 + 2] = nb;
    }
    // log_sig: -log(sigmoid(c)) = softplus(-c)
    cosv = (cos_type == 0) ? c : -log1pf(expf(-c));
  }
  if (lane == 0) {
    red[0][w] = l2;
    red[1][w] = l1;
    red[2][w] = cosv;
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    float s = 0.f;
    for (int q = 0; q < 8; ++q) s += red[threadIdx.x][q];
    atomicAdd(acc + threadIdx.x, s);
  }
}

// out[0] = total, out[1] = mse, out[2] = l1, out[3] = cos loss   (means + weights applied here)
__global__ void distill_loss_finalize_kernel(const float* __restrict__ acc, float* __restrict__ out, float n_elem,
                                             float n_rows, float w2, float w1, float wc) {
  pdl_wait();
  pdl_trigger();
  if (threadIdx.x != 0) return;
  const float mse = (w2 != 0.f) ? acc[0] / n_elem : 0.f;
  const float l1 = (w1 != 0.f) ? acc[1] / n_elem : 0.f;
  const float lc = (wc != 0.f) ? -acc[2] / n_rows : 0.f;
  out[0] = w2 * mse + w1 * l1 + wc * lc;
  out[1] = mse;
  out[2] = l1;
  out[3] = lc;
}

// da = upstream * ( w2 * 2 (a-b)/N + w1 * sign(a-b)/N + wc * d(-mean cos)/da )
template <typename T>
__global__ void __launch_bounds__(256)
    distill_loss_bwd_kernel(const T* __restrict__ a, const T* __restrict__ b, const float* __restrict__ rowstat,
                            const float* __restrict__ upstream, T* __restrict__ da, long long rows, int D, float w2,
                            float w1, float wc, int cos_type, float eps, float n_elem, float n_rows, int vec) {
  pdl_wait();
  pdl_trigger();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const long long row = (long long)blockIdx.x * 8 + w;
  if (row >= rows) return;
  const float g = upstream ? upstream[0] : 1.f;
  const float dot = rowstat[row * 3 + 0], na = rowstat[row * 3 + 1], nb = rowstat[row * 3 + 2];
  const float na_c = fmaxf(na, eps), nb_c = fmaxf(nb, eps);
  const float c = dot / (na_c * nb_c);
  // d cos / d a_i = b_i / (na nb) - cos * a_i / na^2   (for na > eps)
  float k = -wc / n_rows;  // loss_cos = -mean(f(cos))
  if (cos_type != 0) k *= 1.f / (1.f + expf(c));  // d log(sigmoid(c)) / dc = sigmoid(-c)
  const float kb = k / (na_c * nb_c);
  const float ka = (na > eps) ? k * c / (na * na) : 0.f;
  const float c2 = 2.f * w2 / n_elem, c1 = w1 / n_elem;
  const T* ar = a + row * D;
  const T* br = b + row * D;
  T* dr = da + row * D;
  constexpr int VEC = VecOf<T>::N;
  if (vec) {
    for (int c0 = lane * VEC; c0 < D; c0 += 32 * VEC) {
      float xa[VEC], xb[VEC], o[VEC];
      load_vec(ar + c0, xa);
      load_vec(br + c0, xb);
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        const float d = xa[j] - xb[j];
        const float gr = c2 * d + c1 * ((d > 0.f) ? 1.f : ((d < 0.f) ? -1.f : 0.f)) + kb * xb[j] - ka * xa[j];
        o[j] = g * gr;
      }
      store_vec(dr + c0, o);
    }
    return;
  }
  for (int i = lane; i < D; i += 32) {
    const float x = to_f(ar[i]), y = to_f(br[i]);
    const float d = x - y;
    float gr = c2 * d + c1 * ((d > 0.f) ? 1.f : ((d < 0.f) ? -1.f : 0.f)) + kb * y - ka * x;
    dr[i] = from_f<T>(g * gr);
  }
}

}  // namespace dph

using namespace dph;

// acc: 3 floats of scratch; rowstat: rows*3 floats; out: 4 floats.  The loss over L stacked layers is taken by
// calling this once per layer (no torch.stack copy): `first` zeroes acc, `last` finalises with `total_rows`.
extern "C" int dph_distill_loss_fwd(const void* a, const void* b, float* rowstat, float* acc, float* out, int64_t rows,
                                    int D, float w2, float w1, float wc, int cos_type, int first, int last,
                                    int64_t total_rows, int dtype, void* stream) {
  if (!a || !b || !rowstat || !acc || !out || rows <= 0 || D <= 0) {
    set_error("distill_loss_fwd: bad arguments");
    return DPH_ERR_ARG;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (first) cudaMemsetAsync(acc, 0, 3 * sizeof(float), s);
  const unsigned grid = (unsigned)((rows + 7) / 8);
  const int vw = dtype == DPH_DTYPE_BF16 ? 8 : 4;
  const int vec = (D % vw == 0) && (((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) == 0);
  if (dtype == DPH_DTYPE_BF16)
    launch_pdl(distill_loss_fwd_kernel<bf16>, dim3(grid), dim3(256), 0, s, (const bf16*)a, (const bf16*)b, rowstat, acc, rows,
               D, cos_type, 1e-8f, vec);
  else
    launch_pdl(distill_loss_fwd_kernel<float>, dim3(grid), dim3(256), 0, s, (const float*)a, (const float*)b, rowstat, acc,
               rows, D, cos_type, 1e-8f, vec);
  if (last)
    launch_pdl(distill_loss_finalize_kernel, dim3(1), dim3(32), 0, s, (const float*)acc, out, (float)total_rows * (float)D,
               (float)total_rows, w2, w1, wc);
  dph_count_launch(last ? 2 : 1);
  return dph_check_last("distill_loss_fwd");
}

extern "C" int dph_distill_loss_bwd(const void* a, const void* b, const float* rowstat, const float* upstream, void* da,
                                    int64_t rows, int64_t total_rows, int D, float w2, float w1, float wc, int cos_type,
                                    int dtype, void* stream) {
  if (!a || !b || !rowstat || !da || rows <= 0 || D <= 0) {
    set_error("distill_loss_bwd: bad arguments");
    return DPH_ERR_ARG;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const unsigned grid = (unsigned)((rows + 7) / 8);
  const float ne = (float)total_rows * (float)D, nr = (float)total_rows;
  const int vw = dtype == DPH_DTYPE_BF16 ? 8 : 4;
  const int vec = (D % vw == 0) && (((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) |
                                      reinterpret_cast<uintptr_t>(da)) & 15) == 0);
  if (dtype == DPH_DTYPE_BF16)
    launch_pdl(distill_loss_bwd_kernel<bf16>, dim3(grid), dim3(256), 0, s, (const bf16*)a, (const bf16*)b, rowstat, upstream,
               (bf16*)da, rows, D, w2, w1, wc, cos_type, 1e-8f, ne, nr, vec);
  else
    launch_pdl(distill_loss_bwd_kernel<float>, dim3(grid), dim3(256), 0, s, (const float*)a, (const float*)b, rowstat,
               upstream, (float*)da, rows, D, w2, w1, wc, cos_type, 1e-8f, ne, nr, vec);
  dph_count_launch();
  return dph_check_last("distill_loss_bwd");
}
