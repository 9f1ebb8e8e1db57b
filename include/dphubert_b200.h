/*
 * dphubert_b200 — C ABI of the B200-native kernels behind DPHuBERT's joint distill-and-prune step.
 *
 * The reference (seas2nada/DPHuBERT) has no native layer: every operator below replaces a chain of eager
 * PyTorch calls in the reference's Python modules; the reference file:line each entry point stands in for is
 * cited next to it (paths relative to the reference checkout).  All entry points
 *   - take raw device pointers, plain sizes and a cudaStream_t (passed as void*),
 *   - return 0 on success or a negative DPH_ERR_* code, never throw, never allocate, never synchronise,
 *   - launch on the given stream only (safe to capture in a CUDA graph).
 * dtype codes: 0 = bf16, 1 = fp32.
 */
#ifndef DPHUBERT_B200_H
#define DPHUBERT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DPH_DTYPE_BF16 0
#define DPH_DTYPE_F32 1

/* ------------------------------------------------------------------------------------------------------------
 * Library introspection
 * ---------------------------------------------------------------------------------------------------------- */
/* Number of kernels launched by this library since load (or since dph_launch_count_reset). */
long long dph_launch_count(void);
void dph_launch_count_reset(void);
/* ABI version, bumped on any signature change. */
int dph_abi_version(void);
/* Last error string of the calling thread (static storage). */
const char* dph_last_error(void);

/* ------------------------------------------------------------------------------------------------------------
 * Generic bf16 tensor-core GEMM (tcgen05.mma + TMEM accumulators, operands staged by TMA).
 *   C[b][m, n] = epilogue( alpha * sum_k A[b][m, k] * B[b][n, k] )
 * Operands are described as up-to-5-D strided views so that strided 1-D convolutions (overlapping-row
 * views of a channels-last activation), grouped convolutions and per-head attention batches are the same
 * kernel with a different tensor map — no im2col buffer is ever materialised.
 *
 * Replaces: every nn.Linear / `@` / nn.Conv1d(512,512,k,s=2) / grouped pos-conv call on the path —
 *   wav2vec2/components.py:107 (conv), :272 (projection), :328 (pos conv), :406-408,:411,:422,:430 (attention),
 *   :733,:741 (FFN); lightning.py:258 (distill heads) — and their autograd backward (dgrad / wgrad).
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct {
  const void* ptr;     /* bf16 device pointer, 16-byte aligned */
  int major_mn;        /* 0: K-major view, dims = (k_in, row, kx, b0, b1); k index = kx * dims[0] + k_in      */
                       /* 1: MN-major view, dims = (mn, k, b0, b1, 1)  (reduction index is the slow dimension)  */
  uint64_t dims[5];    /* extent of each dimension in elements, dims[0] is contiguous                           */
  uint64_t strides[4]; /* element strides of dims[1..4]; each a multiple of 8                                    */
  int b0_mul, b1_mul;  /* 1 if the operand varies along batch dim 0 / 1, else 0 (broadcast)                     */
} dph_operand_t;

typedef struct {
  dph_operand_t a, b;
  int M, N, K;        /* per-batch problem size (rows of A, rows of B, reduction length)                        */
  int nb0, nb1;       /* batch extents (>= 1)                                                                  */
  int block_n;        /* 0 = auto, else one of 64 / 128 / 192 / 256                                            */
  int ksplit;         /* >1: split the reduction over ksplit CTAs (requires c_mode == 1)                       */
  /* output */
  void* c;            /* (rows, N) row-major with leading dimension ldc; batch offsets c_sb0 / c_sb1 (elements)*/
  void* c2;           /* optional second output holding the pre-activation value (same layout / dtype as c)    */
  int64_t ldc, c_sb0, c_sb1;
  int c_dtype;        /* DPH_DTYPE_*                                                                           */
  int c_mode;         /* 0 store, 1 atomic add (fp32), 2 read-add-store                                        */
  /* epilogue, applied in this order: *alpha, +bias[n], (c2 store), act, *colmask[n], *scalar_mask[0],
   * dropout, +residual */
  float alpha;
  const float* bias;        /* fp32 (N) or NULL */
  int act;                  /* 0 none, 1 exact GELU */
  const float* colmask;     /* fp32 (N) or NULL — HardConcrete column mask */
  const float* scalar_mask; /* fp32 (1) or NULL — HardConcrete whole-layer mask */
  const void* residual;     /* or NULL */
  int64_t ldr, r_sb0, r_sb1;
  int r_dtype;
  float drop_p;             /* 0 disables dropout */
  uint64_t seed;
  int64_t vec_sb0;          /* element offset added to bias / colmask per batch index b0 (grouped conv, heads) */
  int64_t scalar_sb0;       /* element offset added to scalar_mask per batch index b0 (per-head HardConcrete mask) */
} dph_gemm_desc_t;

int dph_gemm_bf16(const dph_gemm_desc_t* desc, void* stream);

/* fp32 CUDA-core GEMM used by the fp32 parity mode (TF32 off; tcgen05 has no true-fp32 MMA):
 *   C[b][m,n] (+)= epilogue( alpha * sum_k A[b](m,k) * B[b](n,k) )
 * element (row, k) of an operand lives at ptr + b0*bs0 + b1*bs1 + row*rs + (k / kin)*kxs + (k % kin)*cs,
 * which expresses plain, transposed, overlapping-row (strided conv) and tap-major (pos conv) views. */
typedef struct {
  const float* ptr;
  int64_t rs, cs, kxs;
  int kin;            /* <= 0 means kin = K (no split of the reduction index) */
  int64_t bs0, bs1;
} dph_operand_f32_t;

typedef struct {
  dph_operand_f32_t a, b;
  int M, N, K, nb0, nb1;
  float* c;
  float* c2;
  int64_t ldc, c_sb0, c_sb1;
  int accumulate;     /* 1: C += result; 2: atomicAdd (batches may alias the same C) */
  float alpha;
  const float* bias;
  int act;
  const float* colmask;
  const float* scalar_mask;
  const float* residual;
  int64_t ldr, r_sb0, r_sb1;
  float drop_p;
  uint64_t seed;
  int64_t vec_sb0, scalar_sb0;
} dph_gemm_f32_desc_t;

int dph_gemm_f32(const dph_gemm_f32_desc_t* desc, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DPHUBERT_B200_H */
