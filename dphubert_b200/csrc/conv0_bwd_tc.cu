// Layer-0 feature-extractor block, BACKWARD, on the tensor cores (bf16 activation path).
//   u = sc[b,c] * z + sh[b,c],  z[b,t,c] = sum_q w[c,q] x[b, s t + q]       (GroupNorm folded into sc / sh)
//   du = dy * act'(u)            (the HardConcrete mask m[c] is applied once per channel when the sums are flushed)
//   per (b, c):  A = sum_t du,  S1[q] = sum_t du x_q,  Bq = sum_t du yn = rstd * sum_q w_q S1[q] + nsh * A,
//                d mask[c] += sum_t dy * act(u)
// i.e. the same per-(utterance, channel) sums `conv0_bwd_fused_kernel` (conv0.cu) hands to
// `conv0_bwd_finalize_kernel`, which folds them with the fp64 waveform moments into dW / dgamma / dbeta
// (wav2vec2/components.py:107-114 backward for conv_layers[0]).
//
// The CUDA-core kernel spends ~45 issue slots per output (10 FMAs to recompute the conv, 10 for the weight-gradient
// sums, GELU and GELU') and runs at 0.11 of the HBM roofline although it reads dy exactly once.  Here both
// contractions run on tcgen05:
//   MMA 1 (recompute, as the forward kernel conv0_tc.cu):  Z[128 frames, 128 ch] = Win[128, 32] . Wt[128 ch, 32]^T with
//         the hi / lo bf16 split  Win row = [x_hi | x_lo | x_hi | 1 0],  Wt row = [w_hi | w_hi | w_lo | 0 0];
//   epilogue (16 warps, TMEM lane = frame): u, act(u), act'(u) with packed fp32x2 arithmetic, dy read from global,
//         du and g = dy * act(u) written as bf16 tiles [frame][channel] into shared memory;
//   MMA 2 (the gradient sums):  D[128 ch, 32] += DU^T[128 ch, 128 frames] . Win[128 frames, 32]  — the SAME window tile
//         as the MN-major B operand, so columns 0..9 / 10..19 of D are sum du x_hi / sum du x_lo, column 30 (the
//         ones column) is A — plus  G^T . E  with a constant tile E whose column 31 is ones: column 31 = d mask.
// D stays in tensor memory (4 channel blocks x 32 columns) across all frame tiles a CTA processes for one utterance
// and is flushed with vector REDs when the utterance changes.  du is rounded to bf16 for MMA 2; A, Bq and S1 are all
// formed from the SAME rounded values, so GroupNorm's cancelling terms stay consistent (the result is the exact
// gradient for an upstream gradient perturbed by 2^-9 relative, the resolution dy itself is stored at).
#include "common.cuh"
#include "../../include/dphubert_b200.h"

namespace dph {
void set_error(const char* fmt, ...);

namespace c0b {
constexpr int TT = 128;                 // frames per tile (= TMEM lanes of MMA 1, K of MMA 2)
constexpr int NB = 128;                 // channels per unit
constexpr int MAXC = 512;
constexpr int MAXK = 10;
constexpr int MAXBLK = MAXC / NB;
constexpr int PROD_WARPS = 2;           // window builders (warps 0, 1), two frames per thread
constexpr int ISSUER = 2;               // MMA issuer warp (warp 3 idles: 20 warps keep 96 registers per thread)
constexpr int EPI0 = 4;                 // first epilogue warp (a multiple of 4: TMEM lane quadrant = warp % 4)
constexpr int EPI_WARPS = 16;           // 4 TMEM lane quadrants x 4 column slices of 32
constexpr int THREADS = 32 * (EPI0 + EPI_WARPS);
constexpr int S_B = 0;                            // weights [512][64] bf16, 128B swizzle                   64 KB
constexpr int S_A = S_B + MAXC * 128;             // windows 2 x [128][64] bf16                             32 KB
constexpr int S_E = S_A + 2 * TT * 128;           // constant tile: column 31 = 1                           16 KB
constexpr int S_H = S_E + TT * 128;               // du tile: 2 sub-tiles [128 frames][64 ch] bf16          32 KB
constexpr int S_G = S_H + 2 * TT * 128;           // dy * act(u) tile, same layout                          32 KB
constexpr int S_TAB = S_G + 2 * TT * 128;         // per channel pair {sc0, sc1, sh0, sh1}                   4 KB
constexpr int S_DY = S_TAB + (MAXC / 2) * 16;    // dy staging, 2 KB per epilogue warp (coalesced -> row per thread) 32 KB
constexpr int S_BAR = S_DY + EPI_WARPS * 2048;
constexpr int SMEM = S_BAR + 256 + 1024;
constexpr int D2_COL = 2 * NB;                    // TMEM: two Z accumulators (2 x 128 columns), then 4 x 32 columns of D

DPH_DEVICE uint32_t sw128(int r, int c) { return (uint32_t)(r * 128 + ((c ^ (r & 7)) << 4)); }

// 32 lanes x 16 columns of 32-bit: thread i of the warp receives row (lane base + i), columns c..c+15
DPH_DEVICE void tmem_ld_32x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

DPH_DEVICE void split_bf16(float v, unsigned short& hi, unsigned short& lo) {
  const bf16 h = __float2bfloat16_rn(v);
  const bf16 l = __float2bfloat16_rn(v - __bfloat162float(h));
  hi = __bfloat16_as_ushort(h);
  lo = __bfloat16_as_ushort(l);
}

struct Params {
  const float* wave;
  long long wave_bs;
  const float* w;
  const float* bias;
  const float* gamma;
  const float* beta;
  const float* mean;
  const float* rstd;
  const float* mask;
  const bf16* dy;
  float* scratch;   // per (b, c) 16 floats [A, Bq, S1[0..9], unused x 4]
  float* dmask;
  int B, C, k, stride, T1, norm;
  int ntile, nblk;
};

template <bool ACT>
__global__ void __launch_bounds__(THREADS, 1) conv0_bwd_tc_kernel(const __grid_constant__ Params p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (sbase - smem_u32(smem_raw));
  const uint32_t sB = sbase + S_B, sA = sbase + S_A, sE = sbase + S_E, sH = sbase + S_H, sG = sbase + S_G;
  float* tab = reinterpret_cast<float*>(gbase + S_TAB);
  const uint32_t bar0 = sbase + S_BAR;
  auto a_full = [&](int i) { return bar0 + 8u * i; };          // [2] window tile i written
  auto a_free = [&](int i) { return bar0 + 16u + 8u * i; };    // [2] every MMA that reads window tile i is done
  auto t_full = [&](int i) { return bar0 + 32u + 8u * i; };    // [2] Z accumulator i complete
  auto t_empty = [&](int i) { return bar0 + 48u + 8u * i; };   // [2] Z accumulator i read by all epilogue warps
  const uint32_t h_full = bar0 + 64u;                           // du / g tiles of the current unit written
  const uint32_t h_free = bar0 + 72u;                           // MMA 2 of the previous unit has read them
  const uint32_t d2_full = bar0 + 80u;                          // every MMA 2 of the utterance is complete
  const uint32_t d2_empty = bar0 + 88u;                         // ... and D has been flushed
  const uint32_t tmem_slot = bar0 + 96u;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(a_full(i), PROD_WARPS * 32);
      mbar_init(a_free(i), 1);
      mbar_init(t_full(i), 1);
      mbar_init(t_empty(i), EPI_WARPS);
    }
    mbar_init(h_full, EPI_WARPS);
    mbar_init(h_free, 1);
    mbar_init(d2_full, 1);
    mbar_init(d2_empty, 4);
    mbar_fence_init();
  }
  if (warp == 0) {
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  pdl_wait();
  // ---- weights -> swizzled bf16 tile: row c = [w_hi(10) | w_hi(10) | w_lo(10) | 0 0]  (K-major B operand of MMA 1)
  for (int c = tid; c < MAXC; c += THREADS) {
    unsigned short row[32];
#pragma unroll
    for (int q = 0; q < 32; ++q) row[q] = 0;
    if (c < p.C) {
#pragma unroll
      for (int q = 0; q < MAXK; ++q) {
        const float v = q < p.k ? __ldg(p.w + c * p.k + q) : 0.f;
        unsigned short hi, lo;
        split_bf16(v, hi, lo);
        row[q] = hi;
        row[MAXK + q] = hi;
        row[2 * MAXK + q] = lo;
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint4 u;
      u.x = row[8 * j] | ((uint32_t)row[8 * j + 1] << 16);
      u.y = row[8 * j + 2] | ((uint32_t)row[8 * j + 3] << 16);
      u.z = row[8 * j + 4] | ((uint32_t)row[8 * j + 5] << 16);
      u.w = row[8 * j + 6] | ((uint32_t)row[8 * j + 7] << 16);
      *reinterpret_cast<uint4*>(gbase + S_B + sw128(c, j)) = u;
    }
  }
  // ---- constant tile E (column 31 = 1.0) and cleared du / g tiles (columns of a partial channel block stay finite)
  for (int i = tid; i < TT * 8; i += THREADS) {
    const int r = i >> 3, j = i & 7;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (j == 3) v.w = 0x3F800000u;  // elements 24..31 of the row: the high half of the last word is element 31
    *reinterpret_cast<uint4*>(gbase + S_E + sw128(r, j)) = v;
  }
  for (int i = tid; i < 4 * TT * 8; i += THREADS)
    *reinterpret_cast<uint4*>(gbase + S_H + i * 16) = make_uint4(0u, 0u, 0u, 0u);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  // contiguous unit range of this CTA; unit = ((b * ntile + tile) * nblk + blk)
  // (32-bit index arithmetic throughout: the host checks that the unit count fits, and 64-bit divisions are calls)
  const int total = p.B * p.ntile * p.nblk;
  const int u0 = (int)((long long)total * blockIdx.x / gridDim.x), u1 = (int)((long long)total * (blockIdx.x + 1) / gridDim.x);
  const int bt0 = u0 / p.nblk;
  auto utt_of = [&](int u) { return (u / p.nblk) / p.ntile; };
  // the sums in D belong to one utterance: flush after unit u when the next unit is another utterance's (or none)
  auto flush_after = [&](int u) { return u + 1 == u1 || utt_of(u + 1) != utt_of(u); };

  if (warp < PROD_WARPS) {
    // ======================= window builders (128 threads, one frame each) =======================
    const int bt1 = (u1 + p.nblk - 1) / p.nblk;
    int nA = 0;
    for (int bt = bt0; bt < bt1 && u0 < u1; ++bt, ++nA) {
      const int abuf = nA & 1;
      const int b = bt / p.ntile, t0 = (bt % p.ntile) * TT;
      if (nA >= 2) mbar_wait(a_free(abuf), (uint32_t)(((nA >> 1) - 1) & 1));
      const int nt = min(TT, p.T1 - t0);
#pragma unroll 1
      for (int r = tid; r < TT; r += PROD_WARPS * 32) {
        unsigned short row[32];
#pragma unroll
        for (int q = 0; q < 32; ++q) row[q] = 0;
        if (r < nt) {
          const float* x = p.wave + (long long)b * p.wave_bs + (long long)(t0 + r) * p.stride;
          float xv[MAXK];
#pragma unroll
          for (int q = 0; q < MAXK; ++q) xv[q] = q < p.k ? __ldg(x + q) : 0.f;
#pragma unroll
          for (int q = 0; q < MAXK; ++q) {
            unsigned short hi, lo;
            split_bf16(xv[q], hi, lo);
            row[q] = hi;
            row[MAXK + q] = lo;
            row[2 * MAXK + q] = hi;
          }
          row[30] = 0x3F80;  // ones column: MMA 1 multiplies it with a zero weight, MMA 2 turns it into sum_t du
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 v4;
          v4.x = row[8 * j] | ((uint32_t)row[8 * j + 1] << 16);
          v4.y = row[8 * j + 2] | ((uint32_t)row[8 * j + 3] << 16);
          v4.z = row[8 * j + 4] | ((uint32_t)row[8 * j + 5] << 16);
          v4.w = row[8 * j + 6] | ((uint32_t)row[8 * j + 7] << 16);
          *reinterpret_cast<uint4*>(gbase + S_A + abuf * (TT * 128) + sw128(r, j)) = v4;
        }
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      mbar_arrive(a_full(abuf));
    }
  } else if (warp == ISSUER) {
    // ======================= MMA issuer =======================
    const uint32_t idesc1 = make_idesc_bf16(TT, NB, false, false);   // Z[frames, ch]   = Win . Wt^T   (K-major both)
    const uint32_t idesc2 = make_idesc_bf16(NB, 32, true, true);     // D[ch, 32]      += DU^T . Win   (MN-major both)
    uint32_t fresh = 0xFu;   // channel blocks whose D accumulator has not been written since the last flush
    uint32_t nflush = 0;
    // MMA 2 of unit `up` (local index ip): needs the du / g tiles the epilogue wrote for it
    auto mma2 = [&](int up, int ip) {
      const int btp = up / p.nblk;
      const int blkp = up % p.nblk, abufp = (btp - bt0) & 1;
      mbar_wait(h_full, (uint32_t)(ip & 1));
      tc_fence_after();
      if (elect_one()) {
        const uint32_t d = tmem_base + (uint32_t)(D2_COL + blkp * 32);
        const uint32_t win = sA + (uint32_t)abufp * (TT * 128);
#pragma unroll
        for (int kk = 0; kk < TT / 16; ++kk)
          umma_f16_ss(d, make_smem_desc(sH + kk * 2048, TT * 128, 1024, 2), make_smem_desc(win + kk * 2048, TT * 128, 1024, 2),
                      idesc2, (kk > 0 || !((fresh >> blkp) & 1u)) ? 1u : 0u);
#pragma unroll
        for (int kk = 0; kk < TT / 16; ++kk)
          umma_f16_ss(d, make_smem_desc(sG + kk * 2048, TT * 128, 1024, 2), make_smem_desc(sE + kk * 2048, TT * 128, 1024, 2),
                      idesc2, 1u);
        umma_commit(h_free);
        const int tile_end = (btp + 1) * p.nblk < u1 ? (btp + 1) * p.nblk : u1;
        if (up + 1 == tile_end) umma_commit(a_free(abufp));
        if (flush_after(up)) umma_commit(d2_full);
      }
      __syncwarp();
      fresh &= ~(1u << blkp);
      if (flush_after(up)) {
        mbar_wait(d2_empty, nflush & 1u);
        tc_fence_after();
        ++nflush;
        fresh = 0xFu;
      }
    };
    int i = 0;
    for (int u = u0; u < u1; ++u, ++i) {
      const int bt = u / p.nblk;
      const int n = bt - bt0, abuf = n & 1, blk = u % p.nblk, acc = i & 1;
      const int tile_begin = bt * p.nblk > u0 ? bt * p.nblk : u0;
      if (u == tile_begin) mbar_wait(a_full(abuf), (uint32_t)((n >> 1) & 1));
      mbar_wait(t_empty(acc), (uint32_t)(((i >> 1) & 1) ^ 1));
      tc_fence_after();
      if (elect_one()) {
        const uint32_t a_addr = sA + (uint32_t)abuf * (TT * 128), b_addr = sB + (uint32_t)blk * (NB * 128);
#pragma unroll
        for (int kk = 0; kk < 2; ++kk)
          umma_f16_ss(tmem_base + (uint32_t)(acc * NB), make_smem_desc(a_addr + kk * 32, 0, 1024, 2),
                      make_smem_desc(b_addr + kk * 32, 0, 1024, 2), idesc1, kk > 0 ? 1u : 0u);
        umma_commit(t_full(acc));
      }
      __syncwarp();
      if (i >= 1) mma2(u - 1, i - 1);
    }
    if (u1 > u0) mma2(u1 - 1, i - 1);
    pdl_trigger();
  } else if (warp >= EPI0) {
    // ======================= epilogue warps =======================
    const int ew = warp - EPI0;
    const int quad = warp & 3;           // TMEM lane quadrant this warp may read
    const int cslice = ew >> 2;          // which 32-column slice of the unit's 128 channels
    const int etid = tid - EPI0 * 32;
    const int frow = quad * 32 + lane;   // frame of this thread inside the tile
    const int cl = cslice * 32;
    int cur_b = -1;
    uint32_t touched = 0, nflush = 0;
    // position of a unit, advanced incrementally (no divisions in the unit loop)
    // (kernel parameters the unit loop needs, held in registers: a constant-bank load per use is a scoreboard stall)
    const int nblk = p.nblk, ntile = p.ntile, T1 = p.T1, C = p.C;
    const bf16* const dyp = p.dy;
    uint32_t act_mask = 0;   // bit blk: this warp's 32-column slice lies inside channel block blk
    for (int q = 0; q < nblk; ++q) act_mask |= (cl < min(NB, C - q * NB) ? 1u : 0u) << q;
    struct Pos { int b, tile, blk; };
    auto advance = [&](Pos q) {
      if (++q.blk == nblk) {
        q.blk = 0;
        if (++q.tile == ntile) { q.tile = 0; ++q.b; }
      }
      return q;
    };
    // dy of a unit for this warp: 32 frames x 32 channels (64 B per frame).  Loaded COALESCED — request e of lane l is the
    // 16-byte piece l % 4 of frame 8 e + l / 4, so one warp request covers 8 rows x 64 B (a row-per-thread load touches
    // 32 cache lines per request and saturates the L1 tag stage: ncu l1tex 71 %, the first use of dy 19 % of all
    // stall samples) — one unit ahead, and transposed to row-per-thread through a 2 KB shared-memory tile per warp.
    auto load_dy = [&](bool valid, Pos q, uint4 (&v)[4]) {
#pragma unroll
      for (int e = 0; e < 4; ++e) v[e] = make_uint4(0u, 0u, 0u, 0u);
      if (valid && ((act_mask >> q.blk) & 1u)) {
        const int tq = q.tile * TT + quad * 32 + (lane >> 2);
        const bf16* src = dyp + ((long long)q.b * T1 + tq) * C + q.blk * NB + cl + (lane & 3) * 8;
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (tq + 8 * e < T1) v[e] = __ldg(reinterpret_cast<const uint4*>(src + (long long)(8 * e) * C));
      }
    };
    const uint32_t dstage = sbase + S_DY + (uint32_t)ew * 2048u;
    // staging address of (row, 16-byte piece): pieces XOR-swizzled by row pair, conflict-free for both access patterns
    auto dy_addr = [&](int row, int piece) { return dstage + (uint32_t)(row * 64 + ((piece ^ ((row >> 1) & 3)) << 4)); };
    Pos cur;
    cur.blk = u0 % p.nblk;
    cur.tile = (u0 / p.nblk) % p.ntile;
    cur.b = (u0 / p.nblk) / p.ntile;
    Pos nxt = advance(cur);
    uint4 dyn[4];
    load_dy(u0 < u1, cur, dyn);
    int i = 0;
    for (int u = u0; u < u1; ++u, ++i, cur = nxt, nxt = advance(nxt)) {
      const int blk = cur.blk, acc = i & 1;
      const int b = cur.b, t0 = cur.tile * TT;
      const bool flush = u + 1 == u1 || nxt.b != cur.b;   // the sums in D belong to one utterance
      const bool active = (act_mask >> blk) & 1u;   // C is a multiple of 32: a warp's 32 columns are all in or all out
      const int c0 = blk * NB + cl;
      // this thread's 32 upstream gradients (64 bytes) were requested one unit ahead; request the next unit's now
      uint4 dyv[4];
#pragma unroll
      for (int e = 0; e < 4; ++e)
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dy_addr(8 * e + (lane >> 2), lane & 3)), "r"(dyn[e].x),
                     "r"(dyn[e].y), "r"(dyn[e].z), "r"(dyn[e].w)
                     : "memory");
      __syncwarp();
#pragma unroll
      for (int e = 0; e < 4; ++e)
        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(dyv[e].x), "=r"(dyv[e].y), "=r"(dyv[e].z), "=r"(dyv[e].w)
                     : "r"(dy_addr(lane, e))
                     : "memory");
      __syncwarp();
      load_dy(u + 1 < u1, nxt, dyn);
      if (b != cur_b) {
        // per-channel table of this utterance, interleaved per channel PAIR: {sc0, sc1, sh0, sh1}, u = sc z + sh
        asm volatile("bar.sync 1, %0;" ::"n"(EPI_WARPS * 32));
        for (int c = etid; c < MAXC; c += EPI_WARPS * 32) {
          float sc = 1.f, sh = 0.f;
          if (c < p.C) {
            const float bb = p.bias ? __ldg(p.bias + c) : 0.f;
            if (p.norm) {
              const float rs = p.rstd[b * p.C + c], mu = p.mean[b * p.C + c], g = __ldg(p.gamma + c);
              sc = rs * g;
              sh = __ldg(p.beta + c) + (bb - mu) * rs * g;
            } else {
              sh = bb;
            }
          }
          float* t = tab + (c >> 1) * 4 + (c & 1);
          t[0] = sc;
          t[2] = sh;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(EPI_WARPS * 32));
        cur_b = b;
      }
      mbar_wait(t_full(acc), (uint32_t)((i >> 1) & 1));
      tc_fence_after();
      const uint32_t dyw[16] = {dyv[0].x, dyv[0].y, dyv[0].z, dyv[0].w, dyv[1].x, dyv[1].y, dyv[1].z, dyv[1].w,
                                dyv[2].x, dyv[2].y, dyv[2].z, dyv[2].w, dyv[3].x, dyv[3].y, dyv[3].z, dyv[3].w};
      // [frame][channel] bf16 tiles, 64-channel sub-tiles of [128 frames][128 B], 128B swizzle (MN-major A of MMA 2)
      const uint32_t sub = (uint32_t)(cl >> 6) * (TT * 128);
      const int j0 = (cl & 63) >> 3;
      // two halves of 16 columns: keeps the accumulator slice, the products and their packed results in registers
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        uint32_t r[16];
        if (active) {
          tmem_ld_32x16(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * NB + cl + hf * 16), r);
          tmem_ld_wait();
        }
        if (hf == 1) {  // all TMEM reads of this accumulator are done: hand it back
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(t_empty(acc));
        }
        uint32_t o1[8], o2[8];
        if (active) {
          const float4* tp = reinterpret_cast<const float4*>(tab + (c0 >> 1) * 4) + hf * 8;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 t = tp[j];
            const f32x2 z = pk2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
            const f32x2 uu = fma2(z, pk2(t.x, t.y), pk2(t.z, t.w));
            f32x2 au = uu, gu = bc2(1.f);
            if constexpr (ACT) gelu_and_grad_tanh2(uu, au, gu);
            const uint32_t w2 = dyw[hf * 8 + j];
            const f32x2 d2 = pk2(__uint_as_float(w2 << 16), __uint_as_float(w2 & 0xFFFF0000u));
            float a0, a1, g0, g1;
            upk2(mul2(d2, gu), a0, a1);
            upk2(mul2(d2, au), g0, g1);
            o1[j] = pack_bf16x2(a0, a1);
            o2[j] = pack_bf16x2(g0, g1);
          }
        }
        // the tiles are single-buffered: MMA 2 of the previous unit must have read them
        if (hf == 0 && i >= 1) mbar_wait(h_free, (uint32_t)((i - 1) & 1));
        if (active) {
#pragma unroll
          for (int q = 0; q < 2; ++q) {
            const uint32_t off = sub + sw128(frow, j0 + hf * 2 + q);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sH + off), "r"(o1[4 * q]), "r"(o1[4 * q + 1]),
                         "r"(o1[4 * q + 2]), "r"(o1[4 * q + 3])
                         : "memory");
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sG + off), "r"(o2[4 * q]), "r"(o2[4 * q + 1]),
                         "r"(o2[4 * q + 2]), "r"(o2[4 * q + 3])
                         : "memory");
          }
        }
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(h_full);
      touched |= 1u << blk;
      if (flush) {
        if (ew < 4) {
          // ---- flush D: thread = channel (TMEM lane), 32 columns per channel block
          mbar_wait(d2_full, nflush & 1u);
          tc_fence_after();
          for (int fb = 0; fb < p.nblk; ++fb) {
            if (!((touched >> fb) & 1u)) continue;
            uint32_t d[32];
            tmem_ld_32x32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(D2_COL + fb * 32), d);
            tmem_ld_wait();
            const int c = fb * NB + frow;
            if (c < p.C) {
              const float m = p.mask ? __ldg(p.mask + c) : 1.f;
              float s1[MAXK];
              float dot = 0.f;
#pragma unroll
              for (int q = 0; q < MAXK; ++q) {
                s1[q] = (__uint_as_float(d[q]) + __uint_as_float(d[MAXK + q])) * m;
                if (q < p.k) dot = fmaf(__ldg(p.w + c * p.k + q), s1[q], dot);
              }
              const float A = __uint_as_float(d[30]) * m;
              float nsc = 1.f, nsh = p.bias ? __ldg(p.bias + c) : 0.f;
              if (p.norm) {
                const float rs = p.rstd[b * p.C + c], mu = p.mean[b * p.C + c];
                nsc = rs;
                nsh = (nsh - mu) * rs;
              }
              const float Bq = fmaf(nsc, dot, nsh * A);   // sum du yn,  yn = nsc z + nsh,  sum du z = sum_q w_q S1[q]
              float* o = p.scratch + ((long long)b * p.C + c) * 16;
              red_add_v4(o, A, Bq, s1[0], s1[1]);
              red_add_v4(o + 4, s1[2], s1[3], s1[4], s1[5]);
              red_add_v4(o + 8, s1[6], s1[7], s1[8], s1[9]);
              if (p.dmask) atomicAdd(p.dmask + c, __uint_as_float(d[31]));
            }
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(d2_empty);
        }
        touched = 0;
        ++nflush;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace c0b

// Accumulates the per-(b, c) sums into `scratch` (16 * B * C floats, cleared by the caller) and d mask.  Returns
// DPH_ERR_UNSUPPORTED (without setting an error) when the shape is outside this kernel's envelope.
int conv0_bwd_tc(const float* wave, long long wave_bs, const float* w, const float* bias, const float* gamma,
                 const float* beta, const float* mean, const float* rstd, const float* mask, const bf16* dy,
                 float* scratch, float* dmask, int B, int C, int k, int stride, int T1, int norm, int act,
                 cudaStream_t s) {
  using namespace c0b;
  if (C % 32 != 0 || C > MAXC || k > MAXK || (reinterpret_cast<uintptr_t>(dy) & 15) != 0) return DPH_ERR_UNSUPPORTED;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(conv0_bwd_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(conv0_bwd_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) {
      set_error("conv0_bwd_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return DPH_ERR_CUDA;
    }
    attr = true;
  }
  Params p;
  p.wave = wave; p.wave_bs = wave_bs; p.w = w; p.bias = bias; p.gamma = gamma; p.beta = beta; p.mean = mean;
  p.rstd = rstd; p.mask = mask; p.dy = dy; p.scratch = scratch; p.dmask = dmask; p.B = B; p.C = C; p.k = k;
  p.stride = stride; p.T1 = T1; p.norm = norm;
  p.ntile = (T1 + TT - 1) / TT;
  p.nblk = (C + NB - 1) / NB;
  const long long total = (long long)B * p.ntile * p.nblk;
  if (total >= (1LL << 30)) return DPH_ERR_UNSUPPORTED;   // the kernel indexes units with 32-bit integers
  int dev = 0, sms = kNumSMsDefault;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = (int)(total < sms ? total : sms);
  cudaError_t e;
  if (act) e = launch_pdl(conv0_bwd_tc_kernel<true>, dim3(grid), dim3(THREADS), SMEM, s, p);
  else e = launch_pdl(conv0_bwd_tc_kernel<false>, dim3(grid), dim3(THREADS), SMEM, s, p);
  if (e != cudaSuccess) {
    set_error("conv0_bwd_tc launch: %s", cudaGetErrorString(e));
    return DPH_ERR_CUDA;
  }
  return DPH_OK;
}

}  // namespace dph
