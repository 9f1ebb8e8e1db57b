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
#define DPH_DTYPE_I16 2 /* 16-bit PCM, dph_collate_crop's sample pool only */

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
  int ksplit;         /* >1: split the reduction over ksplit CTAs (requires c_mode == 1); <0: library chooses    */
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
  int cta_group;            /* 0 = auto; 1 = one CTA per 128-row tile; 2 = CTA pair (tcgen05 cta_group::2) per 256-row tile */
  /* Fused BACKWARD of an activation epilogue (act_bwd != 0; bf16 output, N % 32 == 0, 32-byte aligned rows): the product
   * is d loss / d o for a forward epilogue o = dropout(colmask[n] * act(z)); the kernel reads the saved pre-activation
   * z[m, n] (same dtype as c, own strides), writes c = acc * dropout * colmask[n] * act'(z) and adds the column sums
   * dbias[n] += sum_m c and dmask[n] += sum_m acc * dropout * act(z) (fp32, either may be NULL; indexed like colmask,
   * i.e. + b0 * vec_sb0).  `act`, `colmask`, `drop_p`, `seed` describe the FORWARD epilogue being differentiated.
   * Replaces autograd through components.py:110-114 (conv blocks) and :733-737 (FFN hidden) without a separate
   * elementwise pass over the (rows, N) gradient. */
  int act_bwd;
  const void* z;
  int64_t ldz, z_sb0, z_sb1;
  float* dbias;
  float* dmask;
  /* b_nsplit > 0 (MN-major B only, a multiple of 64 dividing N, nb1 == 1): column n of the product reads B "row"
   * n % b_nsplit of tap n / b_nsplit, the tap being b.dims[3] / b.strides[2] — an overlapping, row-shifted view of the
   * same activation.  Used for the weight gradient of the grouped positional conv (components.py:319-333 backward)
   * with all 128 taps folded into N instead of 128 separate 64 x 64 products. */
  int b_nsplit;
} dph_gemm_desc_t;

int dph_gemm_bf16(const dph_gemm_desc_t* desc, void* stream);

/* Debug hook, not part of the drop-in surface: when `buf` is non-NULL every following dph_gemm_bf16 launch writes a
 * per-CTA event timeline (64 uint64 %globaltimer stamps per CTA, see csrc/gemm_tcgen05.cu); NULL switches it off. */
int dph_gemm_debug_timeline(void* buf);

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


/* ------------------------------------------------------------------------------------------------------------
 * Kernel RNG epoch (dropout / HardConcrete noise): a device-side counter mixed into every counter-based draw, so
 * that a training step captured in a CUDA graph still draws fresh masks on every replay.
 * Replaces torch's global CUDA generator state used by nn.Dropout / Tensor.uniform_ (hardconcrete.py:96).
 * ---------------------------------------------------------------------------------------------------------- */
int dph_rng_set(unsigned long long value, void* stream);
int dph_rng_advance(void* stream);
/* writes the current epoch to *out_dev (device uint64): checkpoint / resume of the noise stream (distill.py:141-144) */
int dph_rng_get(void* out_dev, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Waveform feature extractor, layer 0 (memory-bound): Conv1d(1->C, k<=10, stride) -> GroupNorm(C groups) -> GELU ->
 * HardConcrete channel mask, channels-last (B, T1, C) output written exactly once; statistics come from a 65-term
 * autocorrelation of the waveform (fp64), the backward recomputes the conv from the waveform.
 * Replaces wav2vec2/components.py:107-114 (ConvLayerBlock.forward, layer 0) + autograd; dph_wave_norm replaces
 * wav2vec2/model.py:96-103 (per-utterance F.layer_norm of the waveform).
 * ---------------------------------------------------------------------------------------------------------- */
int dph_conv0_stats(const float* wave, int64_t wave_bs, const float* w, const float* bias, double* moments /* B*65 */,
                    float* mean /* B*C */, float* rstd /* B*C */, int B, int C, int k, int stride, int T1, float eps,
                    void* stream);
int dph_conv0_fwd(const float* wave, int64_t wave_bs, const float* w, const float* bias, const float* gamma,
                  const float* beta, const float* mean, const float* rstd, const float* mask, void* y, int B, int C, int k,
                  int stride, int T1, int norm, int act, int dtype, void* stream);
/* backward in ONE pass over dy: per-(b, c) sums A, Bq, S1[q] = sum_t du x[s t + q] are accumulated together and the
 * GroupNorm-corrected weight gradient is assembled from them and the forward pass's waveform moments. */
int dph_conv0_bwd(const float* wave, int64_t wave_bs, const float* w, const float* bias, const float* gamma,
                  const float* beta, const float* mean, const float* rstd, const double* moments /* B*65, norm only */,
                  const float* mask, const void* dy, float* scratch /* 16*B*C */, float* dW, float* dbias,
                  float* dgamma, float* dbeta, float* dmask, int B, int C, int k, int stride, int T1, int norm, int act,
                  int dtype, void* stream);
/* dph_conv0_bwd without a caller-provided scratch buffer: the 16*B*C floats live in a file-scope device array inside
 * the library (1 MiB per GPU that loads it, held until exit), so the call allocates nothing.  NO COUNTERPART IN THE
 * REFERENCE: added for callers that cannot call cudaMalloc (CUDA graph capture,
 * serving loops, harnesses that forbid it).  Limit: B * C <= 16384 (DPH_ERR_ARG with a message beyond it).  Every call
 * shares the array (cleared at the start of each call, stream-ordered): calls that overlap on the GPU race. */
int dph_conv0_bwd_static(const float* wave, int64_t wave_bs, const float* w, const float* bias, const float* gamma,
                         const float* beta, const float* mean, const float* rstd, const double* moments, const float* mask,
                         const void* dy, float* dW, float* dbias, float* dgamma, float* dbeta, float* dmask, int B, int C,
                         int k, int stride, int T1, int norm, int act, int dtype, void* stream);
int dph_wave_norm(const float* in, int64_t in_bs, float* out, int64_t out_bs, const int64_t* lengths, int B, int T0,
                  float eps, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Row LayerNorm (+ optional GELU, column mask, dropout) forward / backward.
 * Replaces nn.LayerNorm at components.py:54-61 (channel LN of the Large extractor), :271, :808-811, :839, :850-856,
 * :879, :888-889, :905-906 and the dropout at :891.
 * Rows have D columns of which the first d_valid (0 = all) are normalised; the rest is zero padding of a ragged
 * (pruned, prune.py:33-59) channel count and is written back as zeros.
 * ---------------------------------------------------------------------------------------------------------- */
int dph_layernorm_fwd(const void* x, const float* w, const float* b, void* y, void* y_pre, float* mean, float* rstd,
                      const float* colmask, int rows, int D, int d_valid, float eps, int act, float drop_p, uint64_t seed,
                      int dtype, void* stream);
int dph_layernorm_bwd(const void* dy, const void* dy2, const void* x, const float* w, const float* b, const float* mean,
                      const float* rstd, void* dx, float* dw, float* db, const float* colmask, float* dmask, int rows, int D,
                      int d_valid, int act, float drop_p, uint64_t seed, int rows_per_batch, int64_t out_bs, int out_row_off,
                      int dtype, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Elementwise / column-reduction helpers around the GEMMs.
 *  dph_epilogue_bwd: backward of  o = dropout(scalar * colmask * act(z))  -> dz, bias grad, HardConcrete column-mask
 *    grad, layer-gate grad (components.py:110-114, :424-434, :733-746 via autograd).
 *  dph_group_mask_bwd: backward of the per-head mask `output * head_mask[:, None, None]` (components.py:424-426).
 *  dph_pad_rows_zero: `x[mask] = 0` (components.py:979-980).   dph_posconv_pad/unpad: staging layout of the
 *    positional conv (components.py:319-333).   dph_axpy: residual / multi-consumer gradient accumulation.
 * ---------------------------------------------------------------------------------------------------------- */
int dph_epilogue_bwd(const void* dy, const void* z, const float* colmask, const float* scalar, void* dout, float* dmask,
                     float* dbias, float* dscalar, int rows, int C, int act, float drop_p, uint64_t seed,
                     int rows_per_batch, int64_t out_bs, int out_row_off, int64_t in_ld, int64_t z_ld, int dtype,
                     void* stream);
int dph_group_mask_bwd(const void* dy, const void* z, const float* mask, void* dout, float* dmask, int rows, int C,
                       int group, int dtype, void* stream);
int dph_pad_rows_zero(void* x, const int64_t* lengths, int B, int T, int D, int dtype, void* stream);
int dph_posconv_pad(const void* x, void* xp, int B, int T, int G, int Cg, int Cgp, int front, int pad, int dtype,
                    void* stream);
int dph_posconv_unpad(const void* xp, void* x, int B, int T, int G, int Cg, int Cgp, int front, int pad, int accumulate,
                      int dtype, void* stream);
/* Grouped positional convolution as a tap-stationary tensor-core implicit GEMM (bf16; components.py:319-333 +
 * :885-892): out = [residual +] act(conv(xp, wp) + bias), optional pre-activation copy c2.  xp = the staging copy of
 * dph_posconv_pad (rows_total x G*64), wp = (G*Cg, K*64) tap-major weights of dph_posconv_weight_prep (forward) or
 * its tap-flipped transposed copy (input gradient).  rows_b = staging rows per utterance, row0 = staging row that
 * (b = 0, t = 0, tap = 0) reads.  Cgp must be 64, Cg a multiple of 16 (<= 64), K a multiple of 8 in [64, 128]. */
int dph_posconv_tc(const void* xp, long long rows_total, const void* wp, const float* bias, const void* residual,
                   void* out, void* c2, int B, int T, int D, int G, int Cg, int Cgp, int K, int rows_b, int row0,
                   int act, void* stream);
/* Frame counts after the conv stack: L <- max(floor((L - k_i) / s_i) + 1, 0) for i < nlayers (components.py:116-119);
 * ksizes / strides are HOST arrays of nlayers (<= 16) ints. */
int dph_conv_out_lengths(const int64_t* lengths, int64_t* out, int B, int nlayers, const int* ksizes, const int* strides,
                         void* stream);
int dph_axpy(void* dst, const void* src, int64_t n, int dtype, void* stream);
int dph_cast(const void* in, int in_dtype, void* out, int out_dtype, int64_t n, void* stream);
int dph_vec_mul(const float* a, const float* b, float* out, float alpha, float beta, int n, void* stream);
int dph_transpose(const void* in, int in_dtype, void* out, int out_dtype, int B, int R, int C, void* stream);
int dph_memset(void* p, int value, int64_t bytes, void* stream);
int dph_memset2d(void* p, int64_t pitch, int value, int64_t width, int64_t height, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Attention probabilities: masked / biased softmax (+dropout) and WavLM's gated relative position bias.
 * Replaces components.py:411-420 and :546-600, :629-647 (+ autograd); the additive padding mask of
 * components.py:979-983 is rebuilt from `lengths`, no (B,1,T,T) / (B*H,T,T) tensor is materialised.
 * ---------------------------------------------------------------------------------------------------------- */
int dph_softmax_fwd(const float* S, int64_t lds, void* P, void* Pd, int64_t ldp, const int64_t* lengths, int B, int H,
                    int T, float drop_p, uint64_t seed, const int* rel_table, const float* emb, const float* gate,
                    const int* head_index, int Htot, int dtype, void* stream);
int dph_softmax_bwd(const void* P, const float* dPd, int64_t lds, void* dS, int64_t ldp, int B, int H, int T, float drop_p,
                    uint64_t seed, const int* rel_table, const float* emb, const float* gate, const int* head_index,
                    int Htot, float* dgate, float* demb, int num_buckets, int dtype, void* stream);
int dph_wavlm_gate_fwd(const void* x, const float* Wg, const float* bg, const float* cst, float* gate, float* ga, float* gb,
                       int B, int T, int Htot, int Dh, int dtype, void* stream);
int dph_wavlm_gate_bwd(const void* x, const float* Wg, const float* cst, const float* ga, const float* gb,
                       const float* dgate, void* dx, float* dWg, float* dbg, float* dcst, int B, int T, int Htot, int Dh,
                       int dtype, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Fused multi-head self-attention for head_dim = 64, T <= 1024, bf16 (tcgen05 + TMEM + TMA; csrc/attention.cu).
 * Replaces wav2vec2/components.py:406-426 (SelfAttention.forward: q k^T, padding mask, softmax, dropout, p v, head
 * mask) and :629-647 (WavLM gated relative-position bias) together with their autograd backward, without ever
 * materialising a (B, H, T, T) tensor.
 *   qkv   : (B*T, 3*H*64) bf16, the fused projection output laid out [k | v | q] per row
 *   ctx   : (B*T, H*64) bf16 context (head-masked); ctx_un: the same before the head mask (saved for backward; may
 *           be NULL when nothing is saved);  lse: (B, H, T) fp32 log-sum-exp of the scaled scores (may be NULL)
 *   lengths (B) or NULL: keys >= length get the reference's additive -10000 mask
 *   rel_table (2T-1) / emb (num_buckets, Htot) / gate (B, Htot, T) / head_index (H): WavLM bias, all NULL for HuBERT
 *   backward: dctx (B*T, H*64) bf16 -> dqkv (B*T, 3*H*64) bf16; delta (B, H, T) fp32 scratch; dq32 (B*T, H*64) fp32
 *           accumulator that must be zero on entry and is left zeroed; dmask_h (H) fp32 head-mask gradient (+=);
 *           dgate (B, Htot, T) / demb (num_buckets, Htot) fp32 WavLM gradients (+=) or NULL.
 * ---------------------------------------------------------------------------------------------------------- */
int dph_attn_fwd(const void* qkv, void* ctx, void* ctx_un, float* lse, int B, int H, int T, float scale,
                 const int64_t* lengths, const int* rel_table, const float* emb, const float* gate,
                 const int* head_index, int Htot, float drop_p, uint64_t seed, const float* head_mask, void* stream);
int dph_attn_bwd(const void* qkv, const void* dctx, const void* o_un, const float* lse, float* delta, float* dq32,
                 void* dqkv, float* dmask_h, int B, int H, int T, float scale, const int64_t* lengths,
                 const int* rel_table, const float* emb, const float* gate, const int* head_index, int Htot,
                 float drop_p, uint64_t seed, const float* head_mask, float* dgate, float* demb, int num_buckets,
                 void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * HardConcrete gates, expected model size, Lagrangian.
 * Replaces wav2vec2/hardconcrete.py:76-99, the get_num_params tree (components.py:122-134, 187-196, 276-277, 438-449,
 * 750-761, 859-865, 929-934, 1005-1009; model.py:109-113) and lightning.py:267-271 (+ autograd).
 * ---------------------------------------------------------------------------------------------------------- */
int dph_hardconcrete_fwd(const float* log_alpha, const float* u_in, float* u_out, float* mask, int64_t n, float beta,
                         float limit_l, float limit_r, float eps, uint64_t seed, void* stream);
int dph_hardconcrete_bwd(const float* log_alpha, const float* u, const float* dmask, float* dlog_alpha, int64_t n,
                         float beta, float limit_l, float limit_r, void* stream);
int dph_l0_norms(const float* la_base, const int64_t* seg_off, const int* seg_len, int nseg, float bias, float* norms,
                 void* stream);
int dph_num_params(const int* prog, const float* norms, int nseg, float* out_n, float* dnorm, void* stream);
int dph_lagrangian(const float* n_hat, const float* lambdas, float n0, const float* target, float* out /* 5 */,
                   void* stream);
int dph_l0_norms_bwd(const float* la_base, float* dla_base, const int64_t* seg_off, const int* seg_len, int nseg, float bias,
                     const float* dnorm, const float* upstream, const float* coef, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Layer-wise distillation loss (MSE + L1 + cosine), fused reductions.  Replaces lightning.py:116-139 (+ autograd).
 * ---------------------------------------------------------------------------------------------------------- */
int dph_distill_loss_fwd(const void* a, const void* b, float* rowstat, float* acc, float* out /* 4 */, int64_t rows, int D,
                         float w2, float w1, float wc, int cos_type, int first, int last, int64_t total_rows, int dtype,
                         void* stream);
int dph_distill_loss_bwd(const void* a, const void* b, const float* rowstat, const float* upstream, void* da, int64_t rows,
                         int64_t total_rows, int D, float w2, float w1, float wc, int cos_type, int dtype, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Derived GEMM operands: tap-major conv weights (components.py:81-87), weight_norm(dim=2) reparameterisation of the
 * positional conv (components.py:298-307) and its backward.
 * ---------------------------------------------------------------------------------------------------------- */
int dph_conv_weight_prep(const float* w, void* wf, void* wd, int Cout, int Cin, int k, const int* tap_perm, int dtype,
                         void* stream);
int dph_conv_wgrad_unpermute(const float* src, float* dw, int Cout, int Cin, int k, void* stream);
int dph_posconv_weight_prep(const float* v, const float* g, float* norm, float* wn, void* wp, void* wdp, int D, int G,
                            int Cg, int Cgp, int K, int dtype, void* stream);
int dph_posconv_weight_bwd(const float* dwp, const float* v, const float* g, const float* norm, float* scratch, float* dv,
                           float* dg, int D, int G, int Cg, int Cop, int Cgp, int K, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Optimizer over flat arenas: global grad norm, clip, AdamW (lr / step in device memory), bf16 weight refresh,
 * grad zeroing.  Replaces lightning.py:200-238 (torch.optim.AdamW, 3 groups incl. the negative-lr lambda group) and
 * distill.py:48 (gradient_clip_val -> clip_grad_norm_).
 * ---------------------------------------------------------------------------------------------------------- */
int dph_sumsq(const float* x, int64_t n, double* out, void* stream);
int dph_counter_inc(int* counter, void* stream);
int dph_adamw(float* p, float* g, float* m, float* v, int64_t n, const float* lr_dev, float beta1, float beta2, float eps,
              float weight_decay, const int* step_dev, const double* gnorm_sq, float max_norm, void* p_bf16, int zero_grad,
              void* stream);
int dph_scale(float* x, int64_t n, float a, void* stream);
/* In-place all-reduce (x = scale * sum over ranks) of n floats at `mc_ptr`, the MULTICAST address of a slice of a
 * symmetric-memory allocation (torch.distributed._symmetric_memory: same offset on every rank); rank r reduces the
 * r-th 1/world of the slice inside the NVSwitch (multimem.ld_reduce) and stores the result to all ranks (multimem.st).
 * Replaces DDP's NCCL all-reduce of a gradient bucket (distill.py:41).  The caller orders the ranks with a
 * symmetric-memory barrier on `stream` before and after the call.  n % 4 == 0, 16-byte aligned; ctas <= 0: 16. */
int dph_multimem_allreduce(void* mc_ptr, int64_t n, int rank, int world, float scale, int ctas, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Batch assembly (memory-bound gather): out[b, t] = pool[start[b] + t] for t < len[b], zero beyond; lengths_out[b] =
 * min(len[b], T).  `pool` = all pre-decoded utterances back to back (fp32 or int16 PCM / 32768) in HBM or mapped
 * pinned host memory.  Replaces the CPU crop / pad collate at dataset/audio_dataset.py:285-363 (`_crop_audio`,
 * `CollateFnAudio.__call__`) and the per-batch host-to-device copy; offsets are drawn by the host with the reference's
 * RNG calls (audio_dataset.py:309).
 * ---------------------------------------------------------------------------------------------------------- */
int dph_collate_crop(const void* pool, int pool_dtype, const int64_t* start, const int* len, float* out,
                     int64_t* lengths_out, int B, int T, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DPHUBERT_B200_H */
