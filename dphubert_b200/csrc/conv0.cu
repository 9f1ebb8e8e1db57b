// Layer 0 of the waveform feature extractor: Conv1d(1 -> C, k, stride, bias?) -> GroupNorm(C groups) -> GELU -> mask,
// fused into one pass that reads the waveform once and writes the (B, T1, C) channels-last activation exactly once.
//
// Replaces wav2vec2/components.py:107-114 for conv_layers[0] (k=10, s=5, C=512) and its autograd backward.
//
// GroupNorm statistics are obtained WITHOUT touching the big activation: z[t,c] = sum_k w[c,k] x[s t + k], hence
//   sum_t z      = w[c,:] . S          S[k]    = sum_t x[s t + k]
//   sum_t z^2    = w[c,:] R w[c,:]^T   R[k,k'] = sum_t x[s t + k] x[s t + k']
// i.e. a k + k(k+1)/2 element autocorrelation of the waveform per utterance (accumulated in fp64) gives the exact
// per-(utterance, channel) mean / variance.  The backward recomputes z from the waveform instead of saving it.
#include <mutex>

#include "common.cuh"
#include "vec.cuh"
#include "../../include/dphubert_b200.h"

namespace dph {
void set_error(const char* fmt, ...);

constexpr int C0_MAXK = 10;
constexpr int C0_TT = 128;       // output frames per CTA
constexpr int C0_THREADS = 256;  // 64 channel groups (8 channels each) x 4 frame lanes
constexpr int C0_NR = C0_MAXK + C0_MAXK * (C0_MAXK + 1) / 2;  // 65 moments

// ------------------------------------------------------------------------------------------------------------
// waveform moments  (B, 65) fp64
// ------------------------------------------------------------------------------------------------------------
// One thread per moment: the block stages MOM_TT frames of the waveform in shared memory, thread (m, grp) walks the
// frames grp, grp + 3, ... for its moment m (the first k: sums, the rest: products x_j x_l, j <= l) in fp64, the three
// frame groups are folded in shared memory and the block issues one fp64 atomic per moment.
constexpr int MOM_TT = 1024;
__global__ void __launch_bounds__(256)
    conv0_moments_kernel(const float* __restrict__ wave, long long wave_bs, int T1, int k, int stride,
                         double* __restrict__ R) {
  pdl_wait();
  pdl_trigger();
  extern __shared__ float mx[];  // (MOM_TT - 1) * stride + k samples
  __shared__ double part[3][C0_NR];
  const int b = blockIdx.y;
  const int t0 = blockIdx.x * MOM_TT;
  const int nt = min(MOM_TT, T1 - t0);
  const float* x = wave + (long long)b * wave_bs + (long long)t0 * stride;
  const int nsamp = (nt - 1) * stride + k;
  for (int i = threadIdx.x; i < nsamp; i += blockDim.x) mx[i] = x[i];
  __syncthreads();
  const int m = threadIdx.x % C0_NR, grp = threadIdx.x / C0_NR;
  if (grp < 3) {
    // moment index -> (j, l): m < MAXK is the plain sum of tap j = m; the products follow row by row (j <= l)
    int j = m, l = -1;
    if (m >= C0_MAXK) {
      int r = m - C0_MAXK;
      j = 0;
      while (r >= C0_MAXK - j) {
        r -= C0_MAXK - j;
        ++j;
      }
      l = j + r;
    }
    double acc = 0.0;
    if (j < k && l < k) {
      if (l < 0) {
        for (int t = grp; t < nt; t += 3) acc += (double)mx[t * stride + j];
      } else {
        for (int t = grp; t < nt; t += 3) acc += (double)mx[t * stride + j] * (double)mx[t * stride + l];
      }
    }
    part[grp][m] = acc;
  }
  __syncthreads();
  if (threadIdx.x < C0_NR)
    atomicAdd(R + (long long)b * C0_NR + threadIdx.x, part[0][threadIdx.x] + part[1][threadIdx.x] + part[2][threadIdx.x]);
}

// mean / rstd per (b, c) from the moments
__global__ void conv0_stats_kernel(const double* __restrict__ R, const float* __restrict__ w,
                                   const float* __restrict__ bias, float* __restrict__ mean, float* __restrict__ rstd,
                                   int B, int C, int k, int T1, float eps) {
  pdl_wait();
  pdl_trigger();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * C) return;
  const int b = i / C, c = i - b * C;
  const double* r = R + (long long)b * C0_NR;
  double wk[C0_MAXK];
  for (int j = 0; j < C0_MAXK; ++j) wk[j] = j < k ? (double)w[c * k + j] : 0.0;
  double s = 0.0, q = 0.0;
  int idx = C0_MAXK;
  for (int j = 0; j < C0_MAXK; ++j) {
    s += wk[j] * r[j];
    for (int l = j; l < C0_MAXK; ++l) {
      double term = wk[j] * wk[l] * r[idx++];
      q += (l == j) ? term : 2.0 * term;
    }
  }
  const double m0 = s / T1;  // mean without bias
  double var = q / T1 - m0 * m0;
  if (var < 0.0) var = 0.0;
  mean[i] = (float)(m0 + (bias ? (double)bias[c] : 0.0));
  rstd[i] = (float)(1.0 / sqrt(var + (double)eps));
}

// ------------------------------------------------------------------------------------------------------------
// forward apply
// ------------------------------------------------------------------------------------------------------------
// Two adjacent channels per thread, every frame of the tile in turn: a warp writes 64 consecutive channels (128 B of
// bf16) per frame, the 2 x k weights stay in registers, the waveform window is broadcast from shared memory.
constexpr int C0A_TT = 256;  // frames per CTA

template <typename T>
__global__ void __launch_bounds__(C0_THREADS, 3)
    conv0_apply_kernel(const float* __restrict__ wave, long long wave_bs, const float* __restrict__ w,
                       const float* __restrict__ bias, const float* __restrict__ gamma, const float* __restrict__ beta,
                       const float* __restrict__ mean, const float* __restrict__ rstd, const float* __restrict__ mask,
                       T* __restrict__ y, int C, int k, int stride, int T1, int norm, int act, int pair_vec) {
  __shared__ float sx[C0A_TT * 8 + C0_MAXK + 8];
  const int b = blockIdx.y;
  const int t0 = blockIdx.x * C0A_TT;
  const int nt = min(C0A_TT, T1 - t0);
  const float* x = wave + (long long)b * wave_bs + (long long)t0 * stride;
  const int nsamp = (nt - 1) * stride + k;
  for (int i = threadIdx.x; i < nsamp; i += blockDim.x) sx[i] = x[i];
  __syncthreads();
  for (int cbase = 0; cbase < C; cbase += 2 * C0_THREADS) {
    const int c0 = cbase + 2 * threadIdx.x;
    if (c0 >= C) continue;
    float wr[2][C0_MAXK], scale[2], shift[2], mk[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int c = c0 + j;
      const bool ok = c < C;
#pragma unroll
      for (int q = 0; q < C0_MAXK; ++q) wr[j][q] = (ok && q < k) ? __ldg(w + c * k + q) : 0.f;
      const float bb = (ok && bias) ? __ldg(bias + c) : 0.f;
      if (norm && ok) {
        const float rs = rstd[b * C + c], mu = mean[b * C + c], g = __ldg(gamma + c);
        scale[j] = rs * g;
        shift[j] = __ldg(beta + c) + (bb - mu) * rs * g;
      } else {
        scale[j] = 1.f;
        shift[j] = bb;
      }
      mk[j] = (ok && mask) ? __ldg(mask + c) : 1.f;
    }
    const bool two = c0 + 1 < C;
    T* dst = y + ((long long)b * T1 + t0) * C + c0;
#pragma unroll 4
    for (int tt = 0; tt < nt; ++tt, dst += C) {
      float xs[C0_MAXK];
#pragma unroll
      for (int q = 0; q < C0_MAXK; ++q) xs[q] = (q < k) ? sx[tt * stride + q] : 0.f;
      float o[2];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        float z = 0.f;
#pragma unroll
        for (int q = 0; q < C0_MAXK; ++q) z = fmaf(wr[j][q], xs[q], z);
        float u = fmaf(z, scale[j], shift[j]);
        if (act) u = gelu_erf_fast(u);
        o[j] = u * mk[j];
      }
      if constexpr (sizeof(T) == 2) {
        if (pair_vec && two) {
          *reinterpret_cast<uint32_t*>(dst) = pack_bf16x2(o[0], o[1]);
        } else {
          dst[0] = from_f<T>(o[0]);
          if (two) dst[1] = from_f<T>(o[1]);
        }
      } else {
        dst[0] = from_f<T>(o[0]);
        if (two) dst[1] = from_f<T>(o[1]);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------------------
// Single-pass backward (used whenever the waveform moments of the forward pass are available).
//   du = dy * mask * act'(u),  yn = (w.x + bias - mean) * rstd,  u = yn * gamma + beta
// GroupNorm's backward  dz = rstd * gamma * (du - A/T - yn * Bq/T)  needs the full-utterance sums A = sum du and
// Bq = sum du * yn before the weight gradient  dW[c,q] = sum_t dz * x[s t + q]  can be formed -- but dW is linear in
// them:  dW = rstd * gamma * (S1 - (A/T) Sx - (Bq/T) S3)  with  S1[c,q] = sum_t du x_q  (accumulated here, in the
// same pass as A and Bq),  Sx[q] = sum_t x_q  and  S3[c,q] = sum_t yn x_q = rstd * (sum_p w_p R[p,q] + (bias - mean)
// Sx[q]),  both of which follow from the 65 fp64 waveform moments R of the forward pass.  dy is read once.
// scratch layout: per (b, c) 16 floats  [A, Bq, S1[0..9], unused x4].
// ------------------------------------------------------------------------------------------------------------
constexpr int C0F_TT = 256;  // frames per CTA

template <typename T>
__global__ void __launch_bounds__(C0_THREADS, 2)
    conv0_bwd_fused_kernel(const float* __restrict__ wave, long long wave_bs, const float* __restrict__ w,
                           const float* __restrict__ bias, const float* __restrict__ gamma,
                           const float* __restrict__ beta, const float* __restrict__ mean,
                           const float* __restrict__ rstd, const float* __restrict__ mask, const T* __restrict__ dy,
                           float* __restrict__ scratch, float* __restrict__ dmask, int C, int k, int stride, int T1,
                           int norm, int act, int pair_vec) {
  __shared__ float sx[C0F_TT * 8 + C0_MAXK + 8];
  const int b = blockIdx.y;
  const int t0 = blockIdx.x * C0F_TT;
  const int nt = min(C0F_TT, T1 - t0);
  const float* x = wave + (long long)b * wave_bs + (long long)t0 * stride;
  const int nsamp = (nt - 1) * stride + k;
  for (int i = threadIdx.x; i < nsamp; i += blockDim.x) sx[i] = x[i];
  __syncthreads();
  // 2 adjacent channels per thread: a warp reads 64 consecutive channels (128 B of bf16) per frame
  for (int cbase = 0; cbase < C; cbase += 2 * C0_THREADS) {
    const int c0 = cbase + 2 * threadIdx.x;
    if (c0 >= C) continue;
    float wr[2][C0_MAXK], s1[2][C0_MAXK], nsc[2], nsh[2], gm[2], bt[2], mk[2], aA[2], aB[2], aM[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int c = c0 + j;
      const bool ok = c < C;
#pragma unroll
      for (int q = 0; q < C0_MAXK; ++q) {
        wr[j][q] = (ok && q < k) ? __ldg(w + c * k + q) : 0.f;
        s1[j][q] = 0.f;
      }
      const float bb = (ok && bias) ? __ldg(bias + c) : 0.f;
      nsc[j] = 1.f; nsh[j] = bb; gm[j] = 1.f; bt[j] = 0.f;
      if (norm && ok) {
        const float rs = rstd[b * C + c], mu = mean[b * C + c];
        nsc[j] = rs;
        nsh[j] = (bb - mu) * rs;
        gm[j] = __ldg(gamma + c);
        bt[j] = __ldg(beta + c);
      }
      mk[j] = (ok && mask) ? __ldg(mask + c) : 1.f;
      aA[j] = aB[j] = aM[j] = 0.f;
    }
    const bool two = c0 + 1 < C;
    const T* src = dy + ((long long)b * T1 + t0) * C + c0;
    bool done = false;
    if constexpr (sizeof(T) == 2) {
      if (pair_vec && two) {
        // bf16 path: the channel pair of this thread runs on packed fp32x2 arithmetic (FFMA2: half the issue slots
        // for the taps and the weight-gradient sums) with the tanh-form GELU / GELU' of the tensor-core forward
        f32x2 wr2[C0_MAXK], s12[C0_MAXK];
#pragma unroll
        for (int q = 0; q < C0_MAXK; ++q) {
          wr2[q] = pk2(wr[0][q], wr[1][q]);
          s12[q] = pk2(0.f, 0.f);
        }
        const f32x2 nsc2 = pk2(nsc[0], nsc[1]), nsh2 = pk2(nsh[0], nsh[1]), gm2 = pk2(gm[0], gm[1]),
                    bt2 = pk2(bt[0], bt[1]), mk2 = pk2(mk[0], mk[1]);
        f32x2 aA2 = pk2(0.f, 0.f), aB2 = aA2, aM2 = aA2;
#pragma unroll 2
        for (int tt = 0; tt < nt; ++tt, src += C) {
          float xs[C0_MAXK];
#pragma unroll
          for (int q = 0; q < C0_MAXK; ++q) xs[q] = (q < k) ? sx[tt * stride + q] : 0.f;
          const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(src));
          const f32x2 d2 = pk2(f.x, f.y);
          f32x2 z2 = pk2(0.f, 0.f);
#pragma unroll
          for (int q = 0; q < C0_MAXK; ++q) z2 = fma2(wr2[q], bc2(xs[q]), z2);
          const f32x2 yn2 = fma2(z2, nsc2, nsh2);
          const f32x2 u2 = fma2(yn2, gm2, bt2);
          f32x2 au2 = u2, gu2 = pk2(1.f, 1.f);
          if (act) gelu_and_grad_tanh2(u2, au2, gu2);
          aM2 = fma2(d2, au2, aM2);
          const f32x2 du2 = mul2(mul2(d2, mk2), gu2);
          aA2 = add2(aA2, du2);
          aB2 = fma2(du2, yn2, aB2);
#pragma unroll
          for (int q = 0; q < C0_MAXK; ++q) s12[q] = fma2(du2, bc2(xs[q]), s12[q]);
        }
        upk2(aA2, aA[0], aA[1]);
        upk2(aB2, aB[0], aB[1]);
        upk2(aM2, aM[0], aM[1]);
#pragma unroll
        for (int q = 0; q < C0_MAXK; ++q) upk2(s12[q], s1[0][q], s1[1][q]);
        done = true;
      }
    }
#pragma unroll 2
    for (int tt = 0; tt < (done ? 0 : nt); ++tt, src += C) {
      float xs[C0_MAXK];
#pragma unroll
      for (int q = 0; q < C0_MAXK; ++q) xs[q] = (q < k) ? sx[tt * stride + q] : 0.f;
      float d[2];
      if constexpr (sizeof(T) == 2) {
        if (pair_vec && two) {
          const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(src));
          d[0] = f.x;
          d[1] = f.y;
        } else {
          d[0] = to_f(src[0]);
          d[1] = two ? to_f(src[1]) : 0.f;
        }
      } else {
        d[0] = to_f(src[0]);
        d[1] = two ? to_f(src[1]) : 0.f;
      }
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        float z = 0.f;
#pragma unroll
        for (int q = 0; q < C0_MAXK; ++q) z = fmaf(wr[j][q], xs[q], z);
        const float yn = fmaf(z, nsc[j], nsh[j]);
        const float u = fmaf(yn, gm[j], bt[j]);
        float au = u, gu = 1.f;
        if (act) gelu_and_grad_fast(u, au, gu);
        aM[j] = fmaf(d[j], au, aM[j]);
        const float du = d[j] * mk[j] * gu;
        aA[j] += du;
        aB[j] = fmaf(du, yn, aB[j]);
#pragma unroll
        for (int q = 0; q < C0_MAXK; ++q) s1[j][q] = fmaf(du, xs[q], s1[j][q]);
      }
    }
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int c = c0 + j;
      if (c < C) {
        float* o = scratch + ((long long)b * C + c) * 16;
        red_add_v4(o, aA[j], aB[j], s1[j][0], s1[j][1]);
        red_add_v4(o + 4, s1[j][2], s1[j][3], s1[j][4], s1[j][5]);
        red_add_v4(o + 8, s1[j][6], s1[j][7], s1[j][8], s1[j][9]);
        if (dmask) atomicAdd(dmask + c, aM[j]);
      }
    }
  }
}

// One thread per (utterance, channel): fold the per-utterance sums into dW, dbias, dgamma, dbeta (fp64 arithmetic: the
// three terms of dW cancel to leading order).
__global__ void conv0_bwd_finalize_kernel(const float* __restrict__ scratch, const double* __restrict__ R,
                                          const float* __restrict__ w, const float* __restrict__ bias,
                                          const float* __restrict__ gamma, const float* __restrict__ mean,
                                          const float* __restrict__ rstd, float* __restrict__ dW,
                                          float* __restrict__ dbias, float* __restrict__ dgamma,
                                          float* __restrict__ dbeta, int B, int C, int k, int T1, int norm) {
  pdl_wait();
  pdl_trigger();
  // one thread per (utterance, channel): the per-utterance terms are independent and meet in the final atomics
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * C) return;
  const int c = i % C, b0 = i / C;
  double dw[C0_MAXK];
  for (int q = 0; q < C0_MAXK; ++q) dw[q] = 0.0;
  double db = 0.0, dg = 0.0, dbt = 0.0;
  const double invT = 1.0 / (double)T1;
  for (int b = b0; b < b0 + 1; ++b) {
    const float* s = scratch + ((long long)b * C + c) * 16;
    const double A = s[0], Bq = s[1];
    if (!norm) {
      for (int q = 0; q < k; ++q) dw[q] += (double)s[2 + q];
      db += A;
      continue;
    }
    const double* r = R + (long long)b * C0_NR;
    const double rs = rstd[b * C + c], g = gamma[c];
    const double shift = (bias ? (double)bias[c] : 0.0) - (double)mean[b * C + c];
    // second moments as a full symmetric matrix lookup: R2(p, q), p <= q stored row by row after the 10 first moments
    auto r2 = [&](int p_, int q_) {
      const int lo = p_ < q_ ? p_ : q_, hi = p_ < q_ ? q_ : p_;
      // offset of row lo in the packed upper triangle of a C0_MAXK x C0_MAXK matrix
      const int off = lo * C0_MAXK - lo * (lo - 1) / 2;
      return r[C0_MAXK + off + (hi - lo)];
    };
    double syn = 0.0;  // sum_t yn (zero up to rounding)
    for (int p_ = 0; p_ < k; ++p_) syn += (double)w[c * k + p_] * r[p_];
    syn = rs * (syn + shift * (double)T1);
    for (int q = 0; q < k; ++q) {
      double wr2 = 0.0;
      for (int p_ = 0; p_ < k; ++p_) wr2 += (double)w[c * k + p_] * r2(p_, q);
      const double s3 = rs * (wr2 + shift * r[q]);
      dw[q] += rs * g * ((double)s[2 + q] - A * invT * r[q] - Bq * invT * s3);
    }
    db += rs * g * (A - A - Bq * invT * syn);
    dg += Bq;
    dbt += A;
  }
  for (int q = 0; q < k; ++q) atomicAdd(dW + c * k + q, (float)dw[q]);
  if (dbias) atomicAdd(dbias + c, (float)db);
  if (norm && dgamma) atomicAdd(dgamma + c, (float)dg);
  if (norm && dbeta) atomicAdd(dbeta + c, (float)dbt);
}


// Per-utterance waveform normalisation (model.py:96-103: F.layer_norm over the valid samples, zero padding after).
// One CTA per utterance: at most ~250 k samples, read three times from L2.
__global__ void __launch_bounds__(1024)
    wave_norm_kernel(const float* __restrict__ in, long long in_bs, float* __restrict__ out, long long out_bs,
                     const long long* __restrict__ lengths, int T0, float eps) {
  __shared__ double shd[32];
  __shared__ double s_bcast;
  const int b = blockIdx.x;
  const float* x = in + (long long)b * in_bs;
  float* y = out + (long long)b * out_bs;
  long long len = lengths ? lengths[b] : T0;
  if (len > T0) len = T0;
  if (len < 0) len = 0;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  auto block_sum_d = [&](double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if (lane == 0) shd[w] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = 0.0;
      for (int q = 0; q < (int)(blockDim.x >> 5); ++q) t += shd[q];
      s_bcast = t;
    }
    __syncthreads();
    return s_bcast;
  };
  double acc = 0.0;
  for (long long i = threadIdx.x; i < len; i += blockDim.x) acc += (double)x[i];
  const double mean = len > 0 ? block_sum_d(acc) / (double)len : 0.0;
  acc = 0.0;
  for (long long i = threadIdx.x; i < len; i += blockDim.x) {
    const double d = (double)x[i] - mean;
    acc += d * d;
  }
  const double var = len > 0 ? block_sum_d(acc) / (double)len : 0.0;
  const float rstd = (float)(1.0 / sqrt(var + (double)eps));
  const float mu = (float)mean;
  for (long long i = threadIdx.x; i < T0; i += blockDim.x) y[i] = i < len ? (x[i] - mu) * rstd : 0.f;
}

}  // namespace dph

using namespace dph;

namespace dph {
int conv0_bwd_tc(const float* wave, long long wave_bs, const float* w, const float* bias, const float* gamma,
                 const float* beta, const float* mean, const float* rstd, const float* mask, const bf16* dy,
                 float* scratch, float* dmask, int B, int C, int k, int stride, int T1, int norm, int act,
                 cudaStream_t s);
// DPH_CONV0_BWD_TC=0 keeps the CUDA-core backward (debugging / A-B timing)
static bool conv0_bwd_tc_enabled() {
  static const bool on = [] {
    const char* e = getenv("DPH_CONV0_BWD_TC");
    return !(e != nullptr && e[0] == '0');
  }();
  return on;
}
int conv0_fwd_tc(const float* wave, long long wave_bs, const float* w, const float* bias, const float* gamma,
                 const float* beta, const float* mean, const float* rstd, const float* mask, bf16* y, int B, int C, int k,
                 int stride, int T1, int norm, int act, cudaStream_t s);
}

static int conv0_check(int C, int k, int stride, int T1) {
  if (C <= 0 || k <= 0 || k > C0_MAXK || stride <= 0 || stride > 8 || T1 <= 0) {
    set_error("conv0: unsupported shape C=%d k=%d stride=%d T1=%d (k <= 10, stride <= 8)", C, k, stride, T1);
    return DPH_ERR_UNSUPPORTED;
  }
  return DPH_OK;
}

extern "C" int dph_conv0_stats(const float* wave, int64_t wave_bs, const float* w, const float* bias, double* moments,
                               float* mean, float* rstd, int B, int C, int k, int stride, int T1, float eps,
                               void* stream) {
  int rc = conv0_check(C, k, stride, T1);
  if (rc) return rc;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  cudaMemsetAsync(moments, 0, sizeof(double) * B * C0_NR, s);
  const int gx = (T1 + MOM_TT - 1) / MOM_TT;
  const size_t mom_smem = sizeof(float) * ((size_t)(MOM_TT - 1) * stride + k);
  if (mom_smem > 48 * 1024) {
    set_error("conv0_stats: stride %d too large for the moment kernel's staging buffer", stride);
    return DPH_ERR_UNSUPPORTED;
  }
  launch_pdl(conv0_moments_kernel, dim3(gx, B), dim3(256), mom_smem, s, wave, (long long)wave_bs, T1, k, stride, moments);
  launch_pdl(conv0_stats_kernel, dim3((B * C + 127) / 128), dim3(128), 0, s, (const double*)moments, w, bias, mean, rstd, B, C,
             k, T1, eps);
  dph_count_launch(2);
  return dph_check_last("conv0_stats");
}

extern "C" int dph_conv0_fwd(const float* wave, int64_t wave_bs, const float* w, const float* bias, const float* gamma,
                             const float* beta, const float* mean, const float* rstd, const float* mask, void* y, int B,
                             int C, int k, int stride, int T1, int norm, int act, int dtype, void* stream) {
  int rc = conv0_check(C, k, stride, T1);
  if (rc) return rc;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  dim3 grid((T1 + C0A_TT - 1) / C0A_TT, B);
  const int pair_vec = (C % 2 == 0) && ((reinterpret_cast<uintptr_t>(y) & 3) == 0);
  if (dtype == DPH_DTYPE_BF16) {
    // tensor-core path (conv0_tc.cu) for the usual shapes; anything outside its envelope takes the CUDA-core kernel
    rc = conv0_fwd_tc(wave, wave_bs, w, bias, gamma, beta, mean, rstd, mask, (bf16*)y, B, C, k, stride, T1, norm, act, s);
    if (rc != DPH_ERR_UNSUPPORTED) {
      if (rc == DPH_OK) dph_count_launch();
      return rc ? rc : dph_check_last("conv0_fwd_tc");
    }
  }
  if (dtype == DPH_DTYPE_BF16)
    conv0_apply_kernel<bf16><<<grid, C0_THREADS, 0, s>>>(wave, wave_bs, w, bias, gamma, beta, mean, rstd, mask,
                                                         (bf16*)y, C, k, stride, T1, norm, act, pair_vec);
  else
    conv0_apply_kernel<float><<<grid, C0_THREADS, 0, s>>>(wave, wave_bs, w, bias, gamma, beta, mean, rstd, mask,
                                                          (float*)y, C, k, stride, T1, norm, act, pair_vec);
  dph_count_launch();
  return dph_check_last("conv0_fwd");
}

// scratch: 16 * B * C floats (per (b, c): A, Bq, S1[10], pad), zeroed here. `moments` = the (B, 65) fp64 waveform
// moments dph_conv0_stats produced for the forward pass (required when norm != 0).
extern "C" int dph_conv0_bwd(const float* wave, int64_t wave_bs, const float* w, const float* bias, const float* gamma,
                             const float* beta, const float* mean, const float* rstd, const double* moments,
                             const float* mask, const void* dy, float* scratch, float* dW, float* dbias, float* dgamma,
                             float* dbeta, float* dmask, int B, int C, int k, int stride, int T1, int norm, int act,
                             int dtype, void* stream) {
  int rc = conv0_check(C, k, stride, T1);
  if (rc) return rc;
  if (norm && !moments) {
    set_error("conv0_bwd: the GroupNorm backward needs the waveform moments of the forward pass");
    return DPH_ERR_ARG;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  dim3 grid((T1 + C0F_TT - 1) / C0F_TT, B);
  const int pair_vec = (C % 2 == 0) && ((reinterpret_cast<uintptr_t>(dy) & 3) == 0);
  cudaMemsetAsync(scratch, 0, sizeof(float) * 16 * (size_t)B * C, s);
  bool on_tc = false;
  if (dtype == DPH_DTYPE_BF16 && conv0_bwd_tc_enabled()) {
    // tensor-core formulation of the same per-(b, c) sums (conv0_bwd_tc.cu); shapes outside its envelope fall through
    rc = conv0_bwd_tc(wave, wave_bs, w, bias, gamma, beta, mean, rstd, mask, (const bf16*)dy, scratch, dmask, B, C, k,
                      stride, T1, norm, act, s);
    if (rc == DPH_OK) on_tc = true;
    else if (rc != DPH_ERR_UNSUPPORTED) return rc;
  }
  if (on_tc) {
  } else if (dtype == DPH_DTYPE_BF16)
    conv0_bwd_fused_kernel<bf16><<<grid, C0_THREADS, 0, s>>>(wave, wave_bs, w, bias, gamma, beta, mean, rstd, mask,
                                                             (const bf16*)dy, scratch, dmask, C, k, stride, T1, norm,
                                                             act, pair_vec);
  else
    conv0_bwd_fused_kernel<float><<<grid, C0_THREADS, 0, s>>>(wave, wave_bs, w, bias, gamma, beta, mean, rstd, mask,
                                                              (const float*)dy, scratch, dmask, C, k, stride, T1, norm,
                                                              act, pair_vec);
  launch_pdl(conv0_bwd_finalize_kernel, dim3((B * C + 127) / 128), dim3(128), 0, s, (const float*)scratch, moments, w, bias,
             gamma, mean, rstd, dW, dbias, dgamma, dbeta, B, C, k, T1, norm);
  dph_count_launch(2);
  return dph_check_last("conv0_bwd");
}

// ---------------------------------------------------------------------------------------------------------------
// dph_conv0_bwd_static: dph_conv0_bwd with the 16 * B * C float scratch taken from a library-owned, file-scope
// device array instead of a caller buffer, for callers that cannot allocate (CUDA graph capture, serving loops,
// harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Limits: B * C <= DPH_CONV0_STATIC_MAX_BC (16384, e.g. 32 utterances x 512 channels; the array is 1 MiB and every
// GPU that loads the library holds it until exit).  All calls share the one array, which keeps the previous call's
// contents: it is cleared here (stream-ordered memset, as dph_conv0_bwd clears its caller's buffer), and two calls
// that overlap on the GPU (different streams without an ordering between them) race on it.
// In host code `g_conv0_static_scratch` is only a host-side stand-in for the symbol; the device address is looked
// up once per device with cudaGetSymbolAddress while that device is current, and that pointer goes to the kernels.
// ---------------------------------------------------------------------------------------------------------------
#define DPH_CONV0_STATIC_MAX_BC 16384
static __device__ __align__(256) float g_conv0_static_scratch[16 * DPH_CONV0_STATIC_MAX_BC];

static int conv0_static_scratch_ptr(float** out) {
  constexpr int kMaxDev = 64;
  static float* ptrs[kMaxDev] = {};
  static std::mutex mu;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDev) {
    set_error("conv0_bwd_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
    return DPH_ERR_CUDA;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_conv0_static_scratch);  // address of the array on the CURRENT device
    if (e != cudaSuccess || p == nullptr) {
      set_error("conv0_bwd_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return DPH_ERR_CUDA;
    }
    ptrs[dev] = reinterpret_cast<float*>(p);
  }
  *out = ptrs[dev];
  return DPH_OK;
}

extern "C" int dph_conv0_bwd_static(const float* wave, int64_t wave_bs, const float* w, const float* bias,
                                    const float* gamma, const float* beta, const float* mean, const float* rstd,
                                    const double* moments, const float* mask, const void* dy, float* dW, float* dbias,
                                    float* dgamma, float* dbeta, float* dmask, int B, int C, int k, int stride, int T1,
                                    int norm, int act, int dtype, void* stream) {
  if (B <= 0 || C <= 0) {
    set_error("conv0_bwd_static: bad sizes B=%d C=%d", B, C);
    return DPH_ERR_ARG;
  }
  if ((long long)B * C > DPH_CONV0_STATIC_MAX_BC) {
    set_error("conv0_bwd_static: B * C = %lld exceeds the static scratch limit of %d (B=%d, C=%d); use dph_conv0_bwd "
              "with a caller-provided scratch buffer", (long long)B * C, DPH_CONV0_STATIC_MAX_BC, B, C);
    return DPH_ERR_ARG;
  }
  float* scratch = nullptr;
  int rc = conv0_static_scratch_ptr(&scratch);
  if (rc) return rc;
  // same launch sequence as dph_conv0_bwd (which starts by clearing the scratch on `stream`)
  return dph_conv0_bwd(wave, wave_bs, w, bias, gamma, beta, mean, rstd, moments, mask, dy, scratch, dW, dbias, dgamma,
                       dbeta, dmask, B, C, k, stride, T1, norm, act, dtype, stream);
}

extern "C" int dph_wave_norm(const float* in, int64_t in_bs, float* out, int64_t out_bs, const int64_t* lengths, int B,
                             int T0, float eps, void* stream) {
  if (!in || !out || B <= 0 || T0 <= 0) {
    set_error("wave_norm: bad arguments");
    return DPH_ERR_ARG;
  }
  wave_norm_kernel<<<B, 1024, 0, reinterpret_cast<cudaStream_t>(stream)>>>(in, in_bs, out, out_bs,
                                                                           (const long long*)lengths, T0, eps);
  dph_count_launch();
  return dph_check_last("wave_norm");
}
