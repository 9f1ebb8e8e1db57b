// fp32 CUDA-core GEMM for the fp32 parity mode ("TF32 off"): tcgen05 has no true-fp32 MMA, so the
// 1e-4 fp32 parity configuration runs on FFMA. Fully strided operand views (same im2col-free trick as
// the tensor-core path), two batch dimensions, same fused epilogue.
#include "common.cuh"
#include "../../include/dphubert_b200.h"

namespace dph {
void set_error(const char* fmt, ...);

constexpr int SBM = 64, SBN = 64, SBK = 16;

struct SimtParams {
  dph_operand_f32_t a, b;
  int M, N, K, nb0, nb1;
  float* C;
  float* C2;
  long long ldc, c_sb0, c_sb1;
  int accumulate;
  float alpha;
  const float* bias;
  int act;
  const float* colmask;
  const float* scalar_mask;
  const float* residual;
  long long ldr, r_sb0, r_sb1;
  float drop_p;
  unsigned long long seed;
  long long vec_sb0, scalar_sb0;
};

DPH_DEVICE long long opnd_off(const dph_operand_f32_t& o, int row, int k) {
  int kx = k / o.kin;
  int ki = k - kx * o.kin;
  return (long long)row * o.rs + (long long)kx * o.kxs + (long long)ki * o.cs;
}

__global__ void __launch_bounds__(256) gemm_f32_kernel(const SimtParams p) {
  __shared__ float sA[SBK][SBM + 4];
  __shared__ float sB[SBK][SBN + 4];
  const int bz = blockIdx.z;
  const int b0 = bz % p.nb0, b1 = bz / p.nb0;
  const float* A = p.a.ptr + (long long)b0 * p.a.bs0 + (long long)b1 * p.a.bs1;
  const float* B = p.b.ptr + (long long)b0 * p.b.bs0 + (long long)b1 * p.b.bs1;
  const int m0 = blockIdx.y * SBM, n0 = blockIdx.x * SBN;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;

  for (int k0 = 0; k0 < p.K; k0 += SBK) {
    // cooperative load: 64x16 elements each for A and B, 256 threads -> 4 elements per thread per operand
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      int e = threadIdx.x + it * 256;
      int kk, rr;
      if (p.a.cs == 1) {  // k contiguous: let consecutive threads walk k
        kk = e & 15;
        rr = e >> 4;
      } else {
        rr = e & 63;
        kk = e >> 6;
      }
      int gm = m0 + rr, gk = k0 + kk;
      sA[kk][rr] = (gm < p.M && gk < p.K) ? A[opnd_off(p.a, gm, gk)] : 0.0f;
      if (p.b.cs == 1) {
        kk = e & 15;
        rr = e >> 4;
      } else {
        rr = e & 63;
        kk = e >> 6;
      }
      int gn = n0 + rr;
      gk = k0 + kk;
      sB[kk][rr] = (gn < p.N && gk < p.K) ? B[opnd_off(p.b, gn, gk)] : 0.0f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < SBK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = sA[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = sB[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

  const float smask = p.scalar_mask ? __ldg(p.scalar_mask + b0 * p.scalar_sb0) : 1.0f;
  const uint32_t thresh = (uint32_t)(p.drop_p * 4294967296.0);
  const float inv_keep = p.drop_p > 0.0f ? 1.0f / (1.0f - p.drop_p) : 1.0f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int m = m0 + ty * 4 + i;
    if (m >= p.M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n = n0 + tx * 4 + j;
      if (n >= p.N) continue;
      float v = acc[i][j] * p.alpha;
      if (p.bias) v += __ldg(p.bias + b0 * p.vec_sb0 + n);
      long long coff = (long long)b0 * p.c_sb0 + (long long)b1 * p.c_sb1 + (long long)m * p.ldc + n;
      if (p.C2) p.C2[coff] = v;
      if (p.act == 1) v = gelu_erf(v);
      if (p.colmask) v *= __ldg(p.colmask + b0 * p.vec_sb0 + n);
      v *= smask;
      if (p.drop_p > 0.0f) {
        unsigned long long idx =
            (((unsigned long long)b1 * p.nb0 + b0) * (unsigned long long)p.M + m) * (unsigned long long)p.N + n;
        v *= dropout_scale(p.seed, idx, thresh, inv_keep);
      }
      if (p.residual)
        v += p.residual[(long long)b0 * p.r_sb0 + (long long)b1 * p.r_sb1 + (long long)m * p.ldr + n];
      if (p.accumulate == 2) {
        atomicAdd(&p.C[coff], v);
      } else {
        if (p.accumulate) v += p.C[coff];
        p.C[coff] = v;
      }
    }
  }
}

}  // namespace dph

extern "C" int dph_gemm_f32(const dph_gemm_f32_desc_t* d, void* stream) {
  using namespace dph;
  if (!d || !d->a.ptr || !d->b.ptr || !d->c || d->M <= 0 || d->N <= 0 || d->K <= 0 || d->nb0 <= 0 || d->nb1 <= 0) {
    set_error("gemm_f32: bad arguments");
    return DPH_ERR_ARG;
  }
  SimtParams p;
  p.a = d->a;
  p.b = d->b;
  if (p.a.kin <= 0) p.a.kin = d->K;
  if (p.b.kin <= 0) p.b.kin = d->K;
  p.M = d->M;
  p.N = d->N;
  p.K = d->K;
  p.nb0 = d->nb0;
  p.nb1 = d->nb1;
  p.C = d->c;
  p.C2 = d->c2;
  p.ldc = d->ldc;
  p.c_sb0 = d->c_sb0;
  p.c_sb1 = d->c_sb1;
  p.accumulate = d->accumulate;
  p.alpha = d->alpha;
  p.bias = d->bias;
  p.act = d->act;
  p.colmask = d->colmask;
  p.scalar_mask = d->scalar_mask;
  p.residual = d->residual;
  p.ldr = d->ldr;
  p.r_sb0 = d->r_sb0;
  p.r_sb1 = d->r_sb1;
  p.drop_p = d->drop_p;
  p.seed = d->seed;
  p.vec_sb0 = d->vec_sb0;
  p.scalar_sb0 = d->scalar_sb0;
  long long nz = (long long)d->nb0 * d->nb1;
  if (nz > 65535) {
    set_error("gemm_f32: too many batches");
    return DPH_ERR_ARG;
  }
  dim3 grid((d->N + SBN - 1) / SBN, (d->M + SBM - 1) / SBM, (unsigned)nz);
  if (grid.y > 65535) {
    set_error("gemm_f32: M too large for grid.y");
    return DPH_ERR_ARG;
  }
  gemm_f32_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(p);
  dph_count_launch();
  return dph_check_last("gemm_f32_kernel");
}
