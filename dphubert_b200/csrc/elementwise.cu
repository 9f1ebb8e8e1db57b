// Memory-bound elementwise / column-reduction kernels around the GEMMs (all 128-bit vectorised when aligned).
//
//  * dph_epilogue_bwd : backward of the fused GEMM epilogue  o = dropout(scalar * colmask * act(z))  -> dz plus the
//    column reductions the reference gets from autograd: bias grad, HardConcrete column-mask grad, layer-gate grad
//    (components.py:110-114 conv blocks, :424-434 attention, :733-746 FFN).
//  * dph_pad_rows_zero: `x[mask] = 0` for padded frames (components.py:979-980).
//  * dph_posconv_pad  : scatter (B,T,D) into the zero-padded, group-padded staging buffer of the positional conv.
//  * casts / small vector ops used by the weight cache and the mask plumbing.
#include "common.cuh"
#include "vec.cuh"
#include "../../include/dphubert_b200.h"

namespace dph {
void set_error(const char* fmt, ...);

// Thread block = 32 column groups (VEC columns each) x 8 row lanes; a block owns a strip of 32*VEC columns and a
// contiguous chunk of rows.  Column partial sums are reduced across the 8 row lanes in shared memory and leave the
// block as ONE line-wide RED per column (same-address float atomics serialise in L2: keep them few and coalesced).
// ZP = false: no pre-activation involved (plain dropout / gate scaling and the bias column sums), the cheap variant.
template <typename T, int VEC, bool ZP>
__global__ void __launch_bounds__(256, 2)
    epilogue_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ z, const float* __restrict__ colmask,
                        const float* __restrict__ scalar, T* __restrict__ dout, float* __restrict__ dmask,
                        float* __restrict__ dbias, float* __restrict__ dscalar, int rows, int C, int rows_per_blk,
                        int act, float drop_p, unsigned long long seed, int rows_per_batch, long long out_bs,
                        int out_row_off, long long in_ld, long long z_ld) {
  __shared__ float red[2][8][32 * VEC + 1];
  __shared__ float red_s[32];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c0 = (blockIdx.x * 32 + tx) * VEC;
  const int r0 = blockIdx.y * rows_per_blk;
  const int r1 = min(rows, r0 + rows_per_blk);
  const bool active = c0 < C;
  float cm[VEC], a_mask[VEC], a_bias[VEC];
  pdl_wait();
  pdl_trigger();
#pragma unroll
  for (int j = 0; j < VEC; ++j) {
    cm[j] = (active && colmask && c0 + j < C) ? __ldg(colmask + c0 + j) : 1.f;
    a_mask[j] = 0.f;
    a_bias[j] = 0.f;
  }
  const float sc = scalar ? __ldg(scalar) : 1.f;
  float a_scalar = 0.f;
  const uint32_t thresh = (uint32_t)(drop_p * 4294967296.0);
  const float inv_keep = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
  constexpr int RB = ZP ? 2 : 4;  // rows in flight per thread (independent loads issued back to back)
  // bf16 x 8 path: the raw 16-byte words of the NEXT row batch are requested before the current batch is processed,
  // so the loads of one batch fly while the arithmetic of the previous one runs (the kernel is bound by bytes in flight)
  constexpr bool PF = (VEC == 8) && (sizeof(T) == 2);
  [[maybe_unused]] uint4 nd[RB], nz[RB];
  [[maybe_unused]] auto fetch = [&](int rbx) {
#pragma unroll
    for (int q = 0; q < RB; ++q) {
      const int r = rbx + 8 * q;
      if (r < r1) {
        nd[q] = *reinterpret_cast<const uint4*>(dy + (long long)r * in_ld + c0);
        if constexpr (ZP) nz[q] = *reinterpret_cast<const uint4*>(z + (long long)r * z_ld + c0);
      }
    }
  };
  [[maybe_unused]] auto unpack8 = [](const uint4& u, float* v) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = __bfloat1622float2(h[e]);
      v[2 * e] = f.x;
      v[2 * e + 1] = f.y;
    }
  };
  if (active) {
    if constexpr (PF) fetch(r0 + ty);
    for (int rb = r0 + ty; rb < r1; rb += 8 * RB) {
      float d[RB][VEC], zz[RB][VEC];
      if constexpr (PF) {
#pragma unroll
        for (int q = 0; q < RB; ++q) {
          if (rb + 8 * q < r1) {
            unpack8(nd[q], d[q]);
            if constexpr (ZP) unpack8(nz[q], zz[q]);
          }
        }
        if (rb + 8 * RB < r1) fetch(rb + 8 * RB);
      }
#pragma unroll
      for (int q = 0; q < (PF ? 0 : RB); ++q) {
        const int r = rb + 8 * q;
        if (r < r1) {
          const long long ioff = (long long)r * in_ld + c0;
          if constexpr (VEC > 1) {
            load_vec(dy + ioff, d[q]);
            if constexpr (ZP) load_vec(z + (long long)r * z_ld + c0, zz[q]);
          } else {
            d[q][0] = to_f(dy[ioff]);
            if constexpr (ZP) zz[q][0] = to_f(z[(long long)r * z_ld + c0]);
          }
        }
      }
#pragma unroll
      for (int q = 0; q < RB; ++q) {
        const int r = rb + 8 * q;
        if (r < r1) {
          if (drop_p > 0.f) dropout_apply_vec<VEC>(seed, (unsigned long long)r * C + c0, thresh, inv_keep, d[q]);
#pragma unroll
          for (int j = 0; j < VEC; ++j) {
            const float ds = d[q][j];
            const float dm = ds * sc;
            float dv = dm * cm[j];
            if constexpr (ZP) {
              float az = zz[q][j], gz = 1.f;
              if (act == 1) gelu_and_grad_fast(zz[q][j], az, gz);
              a_scalar += ds * az * cm[j];
              a_mask[j] += dm * az;
              dv *= gz;
            }
            a_bias[j] += dv;
            d[q][j] = dv;
          }
          if (dout) {
            const int b = r / rows_per_batch, t = r - b * rows_per_batch;
            const long long ooff = (long long)b * out_bs + (long long)(t + out_row_off) * C + c0;
            if constexpr (VEC > 1) {
              store_vec(dout + ooff, d[q]);
            } else {
              dout[ooff] = from_f<T>(d[q][0]);
            }
          }
        }
      }
    }
  }
  if (dmask || dbias) {
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      red[0][ty][tx * VEC + j] = a_mask[j];
      red[1][ty][tx * VEC + j] = a_bias[j];
    }
    __syncthreads();
    for (int c = threadIdx.x; c < 32 * VEC; c += 256) {
      const int col = blockIdx.x * 32 * VEC + c;
      if (col < C) {
        float m = 0.f, bsum = 0.f;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          m += red[0][q][c];
          bsum += red[1][q][c];
        }
        if (dmask) atomicAdd(dmask + col, m);
        if (dbias) atomicAdd(dbias + col, bsum);
      }
    }
  }
  if (dscalar) {
    float tot = block_sum(a_scalar, red_s);
    if (threadIdx.x == 0) atomicAdd(dscalar, tot);
  }
}

template <typename T>
__global__ void pad_rows_zero_kernel(T* __restrict__ x, const long long* __restrict__ lengths, int B, int Tn, int D) {
  pdl_wait();
  pdl_trigger();
  // one block per (b, t) candidate row in the padded region; grid = B * Tn blocks is wasteful, so walk rows per block
  const int b = blockIdx.y;
  long long len = lengths[b];
  if (len < 0) len = 0;
  for (int t = (int)len + blockIdx.x; t < Tn; t += gridDim.x) {
    T* row = x + ((long long)b * Tn + t) * D;
    for (int i = threadIdx.x; i < D; i += blockDim.x) row[i] = from_f<T>(0.f);
  }
}

// (B, T, G*Cg) -> staging buffer rows [front + b*(T+pad) + t], each row G * Cgp wide (channels of a group padded
// with zeros from Cg to Cgp). Pad rows / pad channels are NOT written here (the buffer is zeroed once at allocation).
template <typename T>
__global__ void posconv_pad_kernel(const T* __restrict__ x, T* __restrict__ xp, int B, int Tn, int G, int Cg, int Cgp,
                                   int front, int pad) {
  const long long total = (long long)B * Tn * G * Cgp;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    int c = (int)(i % Cgp);
    long long r = i / Cgp;
    int g = (int)(r % G);
    r /= G;
    int t = (int)(r % Tn);
    int b = (int)(r / Tn);
    T v = from_f<T>(0.f);
    if (c < Cg) v = x[((long long)b * Tn + t) * (G * Cg) + g * Cg + c];
    xp[((long long)front + (long long)b * (Tn + pad) + t) * (G * Cgp) + g * Cgp + c] = v;
  }
}

// bf16, Cg % 8 == 0, Cgp % 8 == 0, 16-byte aligned buffers: one thread per 16-byte piece of an output row (the scalar
// kernel above spends a 64-bit division chain per 2-byte element: 28.6 us for 3992 x 1024, this one moves the same
// 14 MB at copy speed).  blockIdx.y = row, pieces across the threads.
__global__ void __launch_bounds__(128)
    posconv_pad_vec_kernel(const bf16* __restrict__ x, bf16* __restrict__ xp, int Tn, int G, int Cg, int Cgp, int front,
                           int pad) {
  pdl_wait();
  pdl_trigger();
  const int row = blockIdx.x;                 // b * Tn + t
  const int b = row / Tn, t = row - b * Tn;
  const int ppg = Cgp >> 3;                   // pieces per group in the padded row
  const uint4* src = reinterpret_cast<const uint4*>(x + (long long)row * (G * Cg));
  uint4* dst = reinterpret_cast<uint4*>(xp + ((long long)front + (long long)b * (Tn + pad) + t) * (G * Cgp));
  for (int pc = threadIdx.x; pc < G * ppg; pc += blockDim.x) {
    const int g = pc / ppg, j = pc - g * ppg;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (j * 8 < Cg) v = __ldg(src + ((g * Cg) >> 3) + j);
    dst[pc] = v;
  }
}

// inverse gather: staging rows -> dense (B,T,G*Cg), optionally accumulating into the destination
template <typename T>
__global__ void posconv_unpad_kernel(const T* __restrict__ xp, T* __restrict__ x, int B, int Tn, int G, int Cg, int Cgp,
                                     int front, int pad, int accumulate) {
  const long long total = (long long)B * Tn * G * Cg;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    int c = (int)(i % Cg);
    long long r = i / Cg;
    int g = (int)(r % G);
    r /= G;
    int t = (int)(r % Tn);
    int b = (int)(r / Tn);
    float v = to_f(xp[((long long)front + (long long)b * (Tn + pad) + t) * (G * Cgp) + g * Cgp + c]);
    if (accumulate) v += to_f(x[i]);
    x[i] = from_f<T>(v);
  }
}

template <typename TI, typename TO>
__global__ void cast_kernel(const TI* __restrict__ in, TO* __restrict__ out, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    out[i] = from_f<TO>(to_f(in[i]));
}

// out[i] = a[i] * (b ? b[i] : 1) * alpha + (beta ? beta * out[i] : 0)     (fp32 vectors: masks, gates)
__global__ void vec_mul_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ out,
                               float alpha, float beta, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    float v = a[i] * (b ? b[i] : 1.f) * alpha;
    if (beta != 0.f) v += beta * out[i];
    out[i] = v;
  }
}

// (B, C, T) <-> (B, T, C) tile transpose (API-compat path of ConvLayerBlock.forward, not on the hot path)
template <typename TI, typename TO>
__global__ void transpose_kernel(const TI* __restrict__ in, TO* __restrict__ out, int R, int Ccols) {
  // in: (B, R, Ccols) -> out: (B, Ccols, R)
  __shared__ float tile[32][33];
  const int b = blockIdx.z;
  const TI* ib = in + (long long)b * R * Ccols;
  TO* ob = out + (long long)b * R * Ccols;
  int c = blockIdx.x * 32 + threadIdx.x;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    int r = blockIdx.y * 32 + j;
    tile[j][threadIdx.x] = (r < R && c < Ccols) ? to_f(ib[(long long)r * Ccols + c]) : 0.f;
  }
  __syncthreads();
  int r = blockIdx.y * 32 + threadIdx.x;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    int cc = blockIdx.x * 32 + j;
    if (r < R && cc < Ccols) ob[(long long)cc * R + r] = from_f<TO>(tile[threadIdx.x][j]);
  }
}


// dst[i] += src[i]
template <typename T, int VEC>
__global__ void axpy_kernel(T* __restrict__ dst, const T* __restrict__ src, long long n) {
  pdl_wait();
  pdl_trigger();
  const long long nv = n / VEC;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += (long long)gridDim.x * blockDim.x) {
    if constexpr (VEC == 1) {
      dst[i] = from_f<T>(to_f(dst[i]) + to_f(src[i]));
    } else {
      float a[VEC], b[VEC];
      load_vec(dst + i * VEC, a);
      load_vec(src + i * VEC, b);
#pragma unroll
      for (int j = 0; j < VEC; ++j) a[j] += b[j];
      store_vec(dst + i * VEC, a);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (long long i = nv * VEC; i < n; ++i) dst[i] = from_f<T>(to_f(dst[i]) + to_f(src[i]));
}

// Backward of a per-group column mask (attention-head HardConcrete mask, components.py:424-426):
//   dout[r, c] = dy[r, c] * mask[c / group];   dmask[c / group] += sum_r dy[r, c] * z[r, c]
template <typename T>
__global__ void __launch_bounds__(128)
    group_mask_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ z, const float* __restrict__ mask,
                          T* __restrict__ dout, float* __restrict__ dmask, int rows, int C, int group, int rows_per_blk) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r0 = blockIdx.y * rows_per_blk;
  const int r1 = min(rows, r0 + rows_per_blk);
  float acc = 0.f;
  if (c < C) {
    const float m = __ldg(mask + c / group);
    for (int r = r0; r < r1; ++r) {
      const long long o = (long long)r * C + c;
      const float d = to_f(dy[o]);
      acc += d * to_f(z[o]);
      dout[o] = from_f<T>(d * m);
    }
  }
  // threads of a warp cover 32 consecutive columns; when group is a multiple of 32 they share one gate
  if (group % 32 == 0) {
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0 && c < C) atomicAdd(dmask + c / group, acc);
  } else if (c < C) {
    atomicAdd(dmask + c / group, acc);
  }
}

}  // namespace dph

using namespace dph;

extern "C" int dph_epilogue_bwd(const void* dy, const void* z, const float* colmask, const float* scalar, void* dout,
                                float* dmask, float* dbias, float* dscalar, int rows, int C, int act, float drop_p,
                                uint64_t seed, int rows_per_batch, int64_t out_bs, int out_row_off, int64_t in_ld,
                                int64_t z_ld, int dtype, void* stream) {
  if (!dy || rows <= 0 || C <= 0) {
    set_error("epilogue_bwd: bad arguments");
    return DPH_ERR_ARG;
  }
  if ((act == 1 || dmask || dscalar) && !z) {
    set_error("epilogue_bwd: z required for act / dmask / dscalar");
    return DPH_ERR_ARG;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (rows_per_batch <= 0) {
    rows_per_batch = rows;
    out_bs = 0;
  }
  if (in_ld <= 0) in_ld = C;
  if (z_ld <= 0) z_ld = C;
  const int vecw = dtype == DPH_DTYPE_BF16 ? 8 : 4;
  auto al = [&](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  const bool vec_ok = (C % vecw == 0) && (in_ld % vecw == 0) && (z_ld % vecw == 0) && al(dy) && al(z) && al(dout) &&
                      (out_bs % vecw == 0);
  const int VECW = vec_ok ? vecw : 1;
  const int cols_per_blk = 32 * VECW;
  const int gx = (C + cols_per_blk - 1) / cols_per_blk;
  const bool zp = (z != nullptr);
  int gy = (148 * 2) / gx;                       // two resident CTAs of 256 threads per SM, one wave
  (void)zp;
  const int max_gy = (rows + 31) / 32;           // at least 32 rows (one 4-deep load batch per row lane) per CTA
  if (gy > max_gy) gy = max_gy;
  if (gy < 1) gy = 1;
  if (gy > 65535) gy = 65535;
  const int rows_per_blk = (rows + gy - 1) / gy;
  gy = (rows + rows_per_blk - 1) / rows_per_blk;
  dim3 grid(gx, gy);
#define EB_LAUNCH(T, V)                                                                                             \
  do {                                                                                                              \
    if (zp)                                                                                                         \
      launch_pdl(epilogue_bwd_kernel<T, V, true>, grid, dim3(256), 0, s, (const T*)dy, (const T*)z, colmask, scalar, \
                 (T*)dout, dmask, dbias, dscalar, rows, C, rows_per_blk, act, drop_p, (unsigned long long)seed,      \
                 rows_per_batch, (long long)out_bs, out_row_off, (long long)in_ld, (long long)z_ld);                 \
    else                                                                                                            \
      launch_pdl(epilogue_bwd_kernel<T, V, false>, grid, dim3(256), 0, s, (const T*)dy, (const T*)z, colmask,       \
                 scalar, (T*)dout, dmask, dbias, dscalar, rows, C, rows_per_blk, act, drop_p,                       \
                 (unsigned long long)seed, rows_per_batch, (long long)out_bs, out_row_off, (long long)in_ld,        \
                 (long long)z_ld);                                                                                  \
  } while (0)
  if (dtype == DPH_DTYPE_BF16) {
    if (vec_ok) EB_LAUNCH(bf16, 8);
    else EB_LAUNCH(bf16, 1);
  } else {
    if (vec_ok) EB_LAUNCH(float, 4);
    else EB_LAUNCH(float, 1);
  }
#undef EB_LAUNCH
  dph_count_launch();
  return dph_check_last("epilogue_bwd");
}

extern "C" int dph_pad_rows_zero(void* x, const int64_t* lengths, int B, int T, int D, int dtype, void* stream) {
  if (!x || !lengths || B <= 0 || T <= 0 || D <= 0) {
    set_error("pad_rows_zero: bad arguments");
    return DPH_ERR_ARG;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  dim3 grid(8, B);
  if (dtype == DPH_DTYPE_BF16)
    launch_pdl(pad_rows_zero_kernel<bf16>, dim3(grid), dim3(128), 0, s, (bf16*)x, (const long long*)lengths, B, T, D);
  else
    launch_pdl(pad_rows_zero_kernel<float>, dim3(grid), dim3(128), 0, s, (float*)x, (const long long*)lengths, B, T, D);
  dph_count_launch();
  return dph_check_last("pad_rows_zero");
}

extern "C" int dph_posconv_pad(const void* x, void* xp, int B, int T, int G, int Cg, int Cgp, int front, int pad,
                               int dtype, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  long long total = (long long)B * T * G * Cgp;
  int grid = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  if (dtype == DPH_DTYPE_BF16 && Cg % 8 == 0 && Cgp % 8 == 0 &&
      ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(xp)) & 15) == 0 && (long long)B * T < (1LL << 31)) {
    launch_pdl(posconv_pad_vec_kernel, dim3(B * T), dim3(128), 0, s, (const bf16*)x, (bf16*)xp, T, G, Cg, Cgp, front, pad);
    dph_count_launch();
    return dph_check_last("posconv_pad");
  }
  if (dtype == DPH_DTYPE_BF16)
    posconv_pad_kernel<bf16><<<grid, 256, 0, s>>>((const bf16*)x, (bf16*)xp, B, T, G, Cg, Cgp, front, pad);
  else
    posconv_pad_kernel<float><<<grid, 256, 0, s>>>((const float*)x, (float*)xp, B, T, G, Cg, Cgp, front, pad);
  dph_count_launch();
  return dph_check_last("posconv_pad");
}

extern "C" int dph_posconv_unpad(const void* xp, void* x, int B, int T, int G, int Cg, int Cgp, int front, int pad,
                                 int accumulate, int dtype, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  long long total = (long long)B * T * G * Cg;
  int grid = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  if (dtype == DPH_DTYPE_BF16)
    posconv_unpad_kernel<bf16><<<grid, 256, 0, s>>>((const bf16*)xp, (bf16*)x, B, T, G, Cg, Cgp, front, pad, accumulate);
  else
    posconv_unpad_kernel<float><<<grid, 256, 0, s>>>((const float*)xp, (float*)x, B, T, G, Cg, Cgp, front, pad,
                                                     accumulate);
  dph_count_launch();
  return dph_check_last("posconv_unpad");
}

extern "C" int dph_cast(const void* in, int in_dtype, void* out, int out_dtype, int64_t n, void* stream) {
  if (!in || !out || n <= 0) {
    set_error("cast: bad arguments");
    return DPH_ERR_ARG;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  int grid = (int)((n + 255) / 256 < 148 * 16 ? (n + 255) / 256 : 148 * 16);
  if (in_dtype == DPH_DTYPE_F32 && out_dtype == DPH_DTYPE_BF16)
    cast_kernel<float, bf16><<<grid, 256, 0, s>>>((const float*)in, (bf16*)out, n);
  else if (in_dtype == DPH_DTYPE_BF16 && out_dtype == DPH_DTYPE_F32)
    cast_kernel<bf16, float><<<grid, 256, 0, s>>>((const bf16*)in, (float*)out, n);
  else if (in_dtype == DPH_DTYPE_F32)
    cast_kernel<float, float><<<grid, 256, 0, s>>>((const float*)in, (float*)out, n);
  else
    cast_kernel<bf16, bf16><<<grid, 256, 0, s>>>((const bf16*)in, (bf16*)out, n);
  dph_count_launch();
  return dph_check_last("cast");
}

extern "C" int dph_vec_mul(const float* a, const float* b, float* out, float alpha, float beta, int n, void* stream) {
  if (!a || !out || n <= 0) {
    set_error("vec_mul: bad arguments");
    return DPH_ERR_ARG;
  }
  vec_mul_kernel<<<(n + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a, b, out, alpha, beta, n);
  dph_count_launch();
  return dph_check_last("vec_mul");
}

extern "C" int dph_transpose(const void* in, int in_dtype, void* out, int out_dtype, int B, int R, int C, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  dim3 grid((C + 31) / 32, (R + 31) / 32, B), blk(32, 8);
  if (in_dtype == DPH_DTYPE_F32 && out_dtype == DPH_DTYPE_F32)
    transpose_kernel<float, float><<<grid, blk, 0, s>>>((const float*)in, (float*)out, R, C);
  else if (in_dtype == DPH_DTYPE_F32)
    transpose_kernel<float, bf16><<<grid, blk, 0, s>>>((const float*)in, (bf16*)out, R, C);
  else if (out_dtype == DPH_DTYPE_F32)
    transpose_kernel<bf16, float><<<grid, blk, 0, s>>>((const bf16*)in, (float*)out, R, C);
  else
    transpose_kernel<bf16, bf16><<<grid, blk, 0, s>>>((const bf16*)in, (bf16*)out, R, C);
  dph_count_launch();
  return dph_check_last("transpose");
}

extern "C" int dph_memset(void* p, int value, int64_t bytes, void* stream) {
  cudaError_t e = cudaMemsetAsync(p, value, (size_t)bytes, reinterpret_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) {
    set_error("memset: %s", cudaGetErrorString(e));
    return DPH_ERR_CUDA;
  }
  return DPH_OK;
}

extern "C" int dph_memset2d(void* p, int64_t pitch, int value, int64_t width, int64_t height, void* stream) {
  cudaError_t e = cudaMemset2DAsync(p, (size_t)pitch, value, (size_t)width, (size_t)height,
                                    reinterpret_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) {
    set_error("memset2d: %s", cudaGetErrorString(e));
    return DPH_ERR_CUDA;
  }
  return DPH_OK;
}

extern "C" int dph_axpy(void* dst, const void* src, int64_t n, int dtype, void* stream) {
  if (!dst || !src || n <= 0) {
    set_error("axpy: bad arguments");
    return DPH_ERR_ARG;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const bool al = ((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 15) == 0;
  long long nv = n / (dtype == DPH_DTYPE_BF16 ? 8 : 4);
  int grid = (int)((nv + 255) / 256 < 148 * 16 ? (nv + 255) / 256 : 148 * 16);
  if (grid < 1) grid = 1;
  if (dtype == DPH_DTYPE_BF16) {
    if (al) launch_pdl(axpy_kernel<bf16, 8>, dim3(grid), dim3(256), 0, s, (bf16*)dst, (const bf16*)src, n);
    else launch_pdl(axpy_kernel<bf16, 1>, dim3(148 * 8), dim3(256), 0, s, (bf16*)dst, (const bf16*)src, n);
  } else {
    if (al) launch_pdl(axpy_kernel<float, 4>, dim3(grid), dim3(256), 0, s, (float*)dst, (const float*)src, n);
    else launch_pdl(axpy_kernel<float, 1>, dim3(148 * 8), dim3(256), 0, s, (float*)dst, (const float*)src, n);
  }
  dph_count_launch();
  return dph_check_last("axpy");
}

extern "C" int dph_group_mask_bwd(const void* dy, const void* z, const float* mask, void* dout, float* dmask, int rows,
                                  int C, int group, int dtype, void* stream) {
  if (!dy || !z || !mask || !dout || !dmask || rows <= 0 || C <= 0 || group <= 0) {
    set_error("group_mask_bwd: bad arguments");
    return DPH_ERR_ARG;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int gx = (C + 127) / 128;
  int target_y = (148 * 8 + gx - 1) / gx;
  int rows_per_blk = (rows + target_y - 1) / target_y;
  if (rows_per_blk < 4) rows_per_blk = 4;
  dim3 grid(gx, (rows + rows_per_blk - 1) / rows_per_blk);
  if (dtype == DPH_DTYPE_BF16)
    group_mask_bwd_kernel<bf16><<<grid, 128, 0, s>>>((const bf16*)dy, (const bf16*)z, mask, (bf16*)dout, dmask, rows, C,
                                                     group, rows_per_blk);
  else
    group_mask_bwd_kernel<float><<<grid, 128, 0, s>>>((const float*)dy, (const float*)z, mask, (float*)dout, dmask, rows,
                                                      C, group, rows_per_blk);
  dph_count_launch();
  return dph_check_last("group_mask_bwd");
}


// Frame counts after the conv stack (ConvLayerBlock.forward, components.py:116-119: floor((L - k) / s) + 1, clamped
// at 0, layer after layer) for all utterances in one launch.
namespace dph {
struct ConvGeom {
  int n;
  int k[16];
  int s[16];
};
__global__ void conv_out_lengths_kernel(const long long* __restrict__ in, long long* __restrict__ out, int B,
                                        ConvGeom g) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  long long L = in[b];
  for (int i = 0; i < g.n; ++i) {
    const long long num = L - g.k[i];
    long long q = num / g.s[i];
    if ((num % g.s[i] != 0) && ((num < 0) != (g.s[i] < 0))) --q;  // floor division
    L = q + 1;
    if (L < 0) L = 0;
  }
  out[b] = L;
}
}  // namespace dph

extern "C" int dph_conv_out_lengths(const int64_t* lengths, int64_t* out, int B, int nlayers, const int* ksizes,
                                    const int* strides, void* stream) {
  if (!lengths || !out || B <= 0 || nlayers < 0 || nlayers > 16 || (nlayers > 0 && (!ksizes || !strides))) {
    set_error("conv_out_lengths: bad arguments");
    return DPH_ERR_ARG;
  }
  dph::ConvGeom g;
  g.n = nlayers;
  for (int i = 0; i < 16; ++i) {
    g.k[i] = i < nlayers ? ksizes[i] : 1;
    g.s[i] = i < nlayers ? strides[i] : 1;
    if (g.s[i] <= 0) {
      set_error("conv_out_lengths: stride must be positive");
      return DPH_ERR_ARG;
    }
  }
  dph::conv_out_lengths_kernel<<<(B + 127) / 128, 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const long long*>(lengths), reinterpret_cast<long long*>(out), B, g);
  dph_count_launch();
  return dph_check_last("conv_out_lengths");
}
