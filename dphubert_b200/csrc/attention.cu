// Fused multi-head self-attention for head_dim = 64 on the 5th-generation tensor cores (tcgen05 + TMEM + TMA):
//   forward : S = scale * Q K^T (+ WavLM gated relative-position bias, + additive padding mask) -> online softmax
//             -> dropout -> O = P V -> head mask, with S and P living only in TMEM / shared memory;
//   backward: recomputes S / P per tile from Q, K and the saved log-sum-exp, and produces dQ, dK, dV (and the WavLM
//             gate / embedding gradients) without ever materialising a (B, H, T, T) tensor in HBM.
//
// Replaces wav2vec2/components.py:406-426 (SelfAttention.forward: the two batched matmuls, mask add, row-max
// subtraction, softmax, dropout, head mask) and :629-647 (WavLM bias) plus their autograd backward.  The reference
// materialises (B, H, T, T) fp32 scores (95.6 MB per copy at B=8, T=499) and makes ~6 elementwise passes over them.
//
// Layout: q/k/v are read straight out of the fused projection buffer qkv[(b T + t)][k | v | q] (row stride 3 * H * 64)
// by TMA; the context is written as ctx[(b T + t)][h * 64 + e].
#include "common.cuh"
#include <stdlib.h>
#include <type_traits>

#include "../../include/dphubert_b200.h"

namespace dph {
void set_error(const char* fmt, ...);

constexpr int ATT_D = 64;        // head dim
constexpr int ATT_BQ = 128;      // queries per CTA
constexpr int ATT_BK = 128;      // keys per tile
constexpr float kLog2e = 1.4426950408889634f;
// The relative-position bias table lives in shared memory as tb[j - i + T - 1], pre-multiplied by log2(e).  Query rows
// beyond T (last query tile) index a few entries below 0: TB_PAD zeroed floats in front keep those reads in bounds.
constexpr int TB_PAD = 128;

struct AttnParams {
  int B, H, T, inner;            // inner = H * 64
  float scale;
  const long long* lengths;      // (B) or null
  // WavLM relative position bias (null emb = none)
  const int* rel_table;          // (2T-1)
  const float* emb;              // (num_buckets, Htot)
  const float* gate;             // (B, Htot, T) or null
  const int* head_index;         // (H) or null
  int Htot;
  float drop_p;
  unsigned long long seed;
  const float* head_mask;        // (H) or null
};

DPH_DEVICE void tma_load_2d(uint32_t smem_dst, const CUtensorMap* m, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_dst),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
DPH_DEVICE void tma_load_3d(uint32_t smem_dst, const CUtensorMap* m, uint32_t bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
          smem_dst),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
// raw MUFU.EX2 (no range fix-ups): arguments here are <= ~0, underflow flushes to 0
DPH_DEVICE float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
DPH_DEVICE void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// byte offset of 16-byte chunk `c` (0..7) of row `r` inside a [rows][64 bf16] 128B-swizzled K-major tile
DPH_DEVICE uint32_t sw128(int r, int c) { return (uint32_t)(r * 128 + ((c ^ (r & 7)) << 4)); }

// shared memory map of the forward kernel (all tiles 1024-byte aligned)
// BK = keys per tile.  128: two CTAs per SM (88.5 KB, 256 TMEM columns).  64: 56.5 KB and 128 TMEM columns, THREE CTAs
// per SM — the kernel is latency-bound at four warps per CTA (ncu: issue slots 36 % busy), and 384 query tiles on
// 148 x 2 slots are 1.3 waves whose second round runs on a third of the machine; at three per SM they are one wave.
template <int BK>
struct FwdL {
  static constexpr int SQ = 0;                               // Q   [128][64]                    16 KB
  static constexpr int SK = SQ + ATT_BQ * ATT_D * 2;         // K   [BK][64]                     16 / 8 KB
  static constexpr int SV = SK + BK * ATT_D * 2;             // V   [BK][64]                     16 / 8 KB
  static constexpr int SP = SV + BK * ATT_D * 2;             // P   (BK / 64) x [128][64]        32 / 16 KB
  static constexpr int TB = SP + ATT_BQ * BK * 2;            // bias table, TB_PAD + up to 2*1024 floats 8.5 KB
  static constexpr int BAR = TB + (2048 + TB_PAD) * 4;       // mbarriers + tmem slot
  static constexpr int SMEM = BAR + 64 + 1024;               // + alignment slack
  static constexpr int TMEM = BK + ATT_D <= 128 ? 128 : 256; // S (BK columns) | O tile (64 columns)
};

template <bool BIAS, bool DROP, int BK>
__global__ void __launch_bounds__(128, BK == 64 ? 3 : 2)
    attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmKV, const AttnParams p,
                    bf16* __restrict__ ctx, bf16* __restrict__ ctx_un, float* __restrict__ lse) {
  constexpr int FWD_SQ = FwdL<BK>::SQ, FWD_SK = FwdL<BK>::SK, FWD_SV = FwdL<BK>::SV, FWD_SP = FwdL<BK>::SP,
                FWD_TB = FwdL<BK>::TB, FWD_BAR = FwdL<BK>::BAR;
  constexpr int ATT_BK = BK;   // (shadows the file-scope tile size of the backward kernel)
  extern __shared__ uint8_t smem_raw[];
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (sbase - smem_u32(smem_raw));
  const uint32_t sQ = sbase + FWD_SQ, sK = sbase + FWD_SK, sV = sbase + FWD_SV, sP = sbase + FWD_SP;
  float* tb = reinterpret_cast<float*>(gbase + FWD_TB) + TB_PAD;
  const uint32_t bar_q = sbase + FWD_BAR, bar_k = bar_q + 8, bar_v = bar_q + 16, bar_mma = bar_q + 24,
                 tmem_slot = bar_q + 32, bar_o = bar_q + 40;
  const int tid = threadIdx.x, warp = tid >> 5;
  const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int T = p.T;
  const int q0 = qt * ATT_BQ;
  const int ntiles = (T + ATT_BK - 1) / ATT_BK;
  const int row_q = b * T + q0;

  if (tid == 0) {
    mbar_init(bar_q, 1);
    mbar_init(bar_k, 1);
    mbar_init(bar_v, 1);
    mbar_init(bar_mma, 1);
    mbar_init(bar_o, 1);
    mbar_fence_init();
    tma_prefetch_desc(&tmQKV);
    tma_prefetch_desc(&tmKV);
  }
  if (warp == 0) {
    __syncwarp();
    tmem_alloc(tmem_slot, FwdL<BK>::TMEM);
    tmem_relinquish();
  }
  pdl_wait();  // PDL: barrier init / TMEM allocation above overlap the previous kernel's tail
  int htot = h;
  float gate_i = 1.f;
  const int i = q0 + tid;  // query row of this thread
  if constexpr (BIAS) {
    htot = p.head_index ? p.head_index[h] : h;
    for (int r = tid; r < 2 * T - 1; r += 128)
      tb[r] = __ldg(p.emb + (long long)p.rel_table[r] * p.Htot + htot) * kLog2e;
    for (int r = tid; r < TB_PAD; r += 128) tb[-1 - r] = 0.f;
    if (p.gate && i < T) gate_i = p.gate[((long long)b * p.Htot + htot) * T + i];
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
  const uint32_t tS = tmem_base, tO = tmem_base + BK;
  const uint32_t lane_off = ((uint32_t)(warp * 32)) << 16;

  if (warp == 0 && elect_one()) {
    mbar_expect_tx(bar_q, ATT_BQ * ATT_D * 2);
    tma_load_2d(sQ, &tmQKV, bar_q, 2 * p.inner + h * ATT_D, row_q);
    mbar_expect_tx(bar_k, ATT_BK * ATT_D * 2);
    tma_load_2d(sK, &tmKV, bar_k, h * ATT_D, b * T);
    mbar_expect_tx(bar_v, ATT_BK * ATT_D * 2);
    tma_load_2d(sV, &tmKV, bar_v, p.inner + h * ATT_D, b * T);
  }
  const int len = p.lengths ? (int)p.lengths[b] : T;
  const uint32_t thresh = (uint32_t)(p.drop_p * 4294967296.0);
  const float inv_keep = p.drop_p > 0.f ? 1.f / (1.f - p.drop_p) : 1.f;
  const uint32_t dkey = DROP ? dropout_key(p.seed) : 0u;  // (after pdl_wait: the epoch is written earlier in the step)
  // dropout: the row (b, h, i) folds into the key once, a key pair (2q, 2q+1) then shares one 32-bit hash
  const uint32_t drop_rkey =
      dropout_row_key(dkey, (unsigned long long)(b * p.H + h) * T + (unsigned long long)i);
  const float sc2 = p.scale * kLog2e;

  float m2 = -INFINITY, l = 0.f;  // running max (log2 domain) and sum
  float o[ATT_D];
#pragma unroll
  for (int e = 0; e < ATT_D; ++e) o[e] = 0.f;
  uint32_t mma_phase = 0;
  constexpr uint32_t idesc_s = make_idesc_bf16(ATT_BQ, ATT_BK, false, false);
  constexpr uint32_t idesc_o = make_idesc_bf16(ATT_BQ, ATT_D, false, true);

  // S = Q K^T of tile `tt` (issuing lane only): the scores of tile t + 1 are issued right behind P V of tile t, so the
  // tensor core computes them while the warps fold O (bar_mma: scores ready, bar_o: P V done)
  auto issue_scores = [&](int tt) {
    if (tt == 0) mbar_wait(bar_q, 0);
    mbar_wait(bar_k, (uint32_t)(tt & 1));
    tc_fence_after();
#pragma unroll
    for (int k = 0; k < ATT_D / 16; ++k)
      umma_f16_ss(tS, make_smem_desc(sQ + k * 32, 0, 1024, 2), make_smem_desc(sK + k * 32, 0, 1024, 2), idesc_s,
                  k > 0 ? 1u : 0u);
    umma_commit(bar_mma);
  };
  if (warp == 0 && elect_one()) issue_scores(0);
  uint32_t o_phase = 0;

  for (int t = 0; t < ntiles; ++t) {
    const int k0 = t * ATT_BK;
    mbar_wait(bar_mma, mma_phase);
    mma_phase ^= 1u;
    tc_fence_after();
    if (warp == 0 && elect_one() && t + 1 < ntiles) {  // K tile is free again: fetch the next one while the softmax runs
      mbar_expect_tx(bar_k, ATT_BK * ATT_D * 2);
      tma_load_2d(sK, &tmKV, bar_k, h * ATT_D, b * T + k0 + ATT_BK);
    }

    // Per 32-key chunk: chunks whose keys are all real (below T and below the padding length) take a lean path —
    // without a position bias the row maximum is taken on the raw scores (the scale is positive) and a probability
    // costs one FFMA + one MUFU.EX2; with a bias one shared-memory load and one more FFMA.  Only the chunk that
    // straddles the end of the utterance pays for the general logit (clamped table index, key-validity tests, padding
    // mask).  Warp-uniform choice (in fact CTA-uniform).
    const int lim = min(T, len);
    // general logit (log2 domain) of key j for this thread's query
    auto logit = [&](uint32_t raw, int j) {
      float x = __uint_as_float(raw) * sc2;
      if constexpr (BIAS) x += gate_i * tb[min(max(j - i + T - 1, -TB_PAD), 2 * T - 2)];
      if (j >= len) x += -10000.0f * kLog2e;
      return x;
    };
    // ---- pass A: row maximum of this tile (log2 domain)
    float mx = m2, mr = -INFINITY;
#pragma unroll 1
    for (int c = 0; c < ATT_BK / 32; ++c) {
      uint32_t r[32];
      tmem_ld_32x32(tS + lane_off + c * 32, r);
      tmem_ld_wait();
      const int j0 = k0 + c * 32;
      if (j0 + 32 > lim) {
#pragma unroll
        for (int jj = 0; jj < 32; ++jj) {
          const float x = logit(r[jj], j0 + jj);
          if (j0 + jj < T) mx = fmaxf(mx, x);
        }
      } else if constexpr (BIAS) {
        const float* tp = tb + (j0 - i + T - 1);
#pragma unroll
        for (int jj = 0; jj < 32; ++jj) mx = fmaxf(mx, fmaf(__uint_as_float(r[jj]), sc2, gate_i * tp[jj]));
      } else {
#pragma unroll
        for (int jj = 0; jj < 32; ++jj) mr = fmaxf(mr, __uint_as_float(r[jj]));
      }
    }
    mx = fmaxf(mx, mr * sc2);
    const float alpha = fast_exp2(m2 - mx);  // m2 = -inf on the first tile -> 0; applied to O when the tile is added
    m2 = mx;
    l *= alpha;
    const float nm2 = -m2;
    // ---- pass B: probabilities -> bf16 P tile in shared memory (K-major, 128B swizzle, two 64-key blocks)
    float lsum = 0.f;
    auto pass_b = [&](auto edge_tag, int c) {
      constexpr bool EDGE = decltype(edge_tag)::value;
      uint32_t r[32];
      tmem_ld_32x32(tS + lane_off + c * 32, r);
      tmem_ld_wait();
      uint32_t pk[16];
      [[maybe_unused]] const float* tp = tb + (k0 + c * 32 - i + T - 1);
#pragma unroll
      for (int jj = 0; jj < 32; jj += 2) {
        const int j = k0 + c * 32 + jj;
        float p0, p1;
        if constexpr (!EDGE) {
          float e0, e1;
          if constexpr (BIAS) {
            e0 = fmaf(__uint_as_float(r[jj]), sc2, fmaf(gate_i, tp[jj], nm2));
            e1 = fmaf(__uint_as_float(r[jj + 1]), sc2, fmaf(gate_i, tp[jj + 1], nm2));
          } else {
            upk2(fma2(pk2(__uint_as_float(r[jj]), __uint_as_float(r[jj + 1])), bc2(sc2), bc2(nm2)), e0, e1);
          }
          p0 = fast_exp2(e0);
          p1 = fast_exp2(e1);
        } else {
          p0 = (j < T) ? fast_exp2(logit(r[jj], j) + nm2) : 0.f;
          p1 = (j + 1 < T) ? fast_exp2(logit(r[jj + 1], j + 1) + nm2) : 0.f;
        }
        lsum += p0 + p1;
        if constexpr (DROP) {
          float s0, s1;
          dropout_row_pair(drop_rkey, (uint32_t)j >> 1, thresh, inv_keep, s0, s1);
          p0 *= s0;
          p1 *= s1;
        }
        pk[jj >> 1] = pack_bf16x2(p0, p1);
      }
      // 32 keys = 4 chunks of 16 bytes; key block (c / 2), chunk index within the 64-key block ((c & 1) * 4 + q)
      const uint32_t blk = sP + (uint32_t)(c >> 1) * (ATT_BQ * 128);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint32_t addr = blk + sw128(tid, (c & 1) * 4 + q);
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(pk[4 * q]), "r"(pk[4 * q + 1]),
                     "r"(pk[4 * q + 2]), "r"(pk[4 * q + 3])
                     : "memory");
      }
    };
#pragma unroll 1
    for (int c = 0; c < ATT_BK / 32; ++c) {
      if (k0 + c * 32 + 32 > lim) pass_b(std::true_type{}, c);
      else pass_b(std::false_type{}, c);
    }
    l += lsum;
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    // ---- O_tile = P V
    if (warp == 0 && elect_one()) {
      mbar_wait(bar_v, (uint32_t)(t & 1));
      tc_fence_after();
      const uint32_t vb = sV;
#pragma unroll
      for (int k = 0; k < ATT_BK / 16; ++k)
        umma_f16_ss(tO, make_smem_desc(sP + (k >> 2) * (ATT_BQ * 128) + (k & 3) * 32, 0, 1024, 2),
                    make_smem_desc(vb + k * 2048, ATT_BK * 128, 1024, 2), idesc_o, k > 0 ? 1u : 0u);
      umma_commit(bar_o);
      if (t + 1 < ntiles) issue_scores(t + 1);  // every thread is past its reads of S (the barrier above)
    }
    mbar_wait(bar_o, o_phase);
    o_phase ^= 1u;
    tc_fence_after();
    if (warp == 0 && elect_one() && t + 1 < ntiles) {
      mbar_expect_tx(bar_v, ATT_BK * ATT_D * 2);
      tma_load_2d(sV, &tmKV, bar_v, p.inner + h * ATT_D, b * T + k0 + ATT_BK);
    }
#pragma unroll
    for (int c = 0; c < ATT_D / 32; ++c) {
      uint32_t r[32];
      tmem_ld_32x32(tO + lane_off + c * 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int jj = 0; jj < 32; ++jj) o[c * 32 + jj] = fmaf(o[c * 32 + jj], alpha, __uint_as_float(r[jj]));
    }
    // no barrier here: P (shared memory) is free because P V has completed, and the next P V, which overwrites the O
    // tile, is only issued behind the next tile's block barrier
    tc_fence_before();
  }

  pdl_trigger();  // only the context store is left: the next kernel's prologue may start on this SM
  if (i < T) {
    const float inv_l = 1.f / l;
    const float hm = p.head_mask ? __ldg(p.head_mask + h) : 1.f;
    if (lse) lse[((long long)(b * p.H + h)) * T + i] = (m2 + log2f(l)) * (1.f / kLog2e);
    bf16* dst = ctx + ((long long)(b * T + i)) * p.inner + h * ATT_D;
    bf16* dst_un = ctx_un ? ctx_un + ((long long)(b * T + i)) * p.inner + h * ATT_D : nullptr;
#pragma unroll
    for (int e = 0; e < ATT_D; e += 8) {
      float v[8], vm[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        v[u] = o[e + u] * inv_l;
        vm[u] = v[u] * hm;
      }
      uint4 a, c4;
      a.x = pack_bf16x2(vm[0], vm[1]); a.y = pack_bf16x2(vm[2], vm[3]);
      a.z = pack_bf16x2(vm[4], vm[5]); a.w = pack_bf16x2(vm[6], vm[7]);
      *reinterpret_cast<uint4*>(dst + e) = a;
      if (dst_un) {
        c4.x = pack_bf16x2(v[0], v[1]); c4.y = pack_bf16x2(v[2], v[3]);
        c4.z = pack_bf16x2(v[4], v[5]); c4.w = pack_bf16x2(v[6], v[7]);
        *reinterpret_cast<uint4*>(dst_un + e) = c4;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, FwdL<BK>::TMEM);
  }
}


// ------------------------------------------------------------------------------------------------------------
// backward, step 1: delta'[b,h,i] = sum_e dctx[b,i,h,e] * O_un[b,i,h,e]  (and the head-mask gradient = its sum)
// ------------------------------------------------------------------------------------------------------------
// One warp per row, 16-byte loads: lane l of pass q holds columns (q * 32 + l) * 8 .. + 8, i.e. head (q * 32 + l) / 8;
// the eight lanes of a head fold with three shuffles.  Every load of a row (up to 4 passes x 2 tensors) is issued
// before the first use — the kernel moves 18 MB and is pure latency: with four rows per warp walked one after the
// other it took 13 us at 125 blocks; one row per warp and all loads in flight at once: 499 blocks.
constexpr int PREP_RPW = 1;
__global__ void __launch_bounds__(256)
    attn_bwd_prep_kernel(const bf16* __restrict__ dctx, const bf16* __restrict__ o_un, float* __restrict__ delta,
                         float* __restrict__ dmask_h, int B, int H, int T, int inner, bf16* __restrict__ dq_zero) {
  __shared__ float red[8][32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const long long M = (long long)B * T;
  const int nchunk = inner >> 3;            // 16-byte chunks per row (8 per head)
  const int npass = (nchunk + 31) >> 5;
  pdl_wait();
  pdl_trigger();
  const long long row = (long long)blockIdx.x * 8 + w;
  const bool row_ok = row < M;
  const int b = row_ok ? (int)(row / T) : 0, i = row_ok ? (int)(row - (long long)b * T) : 0;
  for (int q0 = 0; q0 < npass; q0 += 4) {   // up to 4 passes (= 16 heads) per sweep
    uint4 a[4], c[4];
#pragma unroll
    for (int qq = 0; qq < 4; ++qq) {
      const int chunk = (q0 + qq) * 32 + lane;
      a[qq] = make_uint4(0u, 0u, 0u, 0u);
      c[qq] = a[qq];
      if (row_ok && chunk < nchunk) {
        a[qq] = __ldg(reinterpret_cast<const uint4*>(dctx + row * inner + chunk * 8));
        c[qq] = __ldg(reinterpret_cast<const uint4*>(o_un + row * inner + chunk * 8));
        if (dq_zero)  // Q slot of dqkv: the backward kernel accumulates dQ into it with bf16x2 reductions
          *reinterpret_cast<uint4*>(dq_zero + row * (3LL * inner) + 2 * inner + chunk * 8) = make_uint4(0u, 0u, 0u, 0u);
      }
    }
    float hsum[4];
#pragma unroll
    for (int qq = 0; qq < 4; ++qq) {
      const int chunk = (q0 + qq) * 32 + lane;
      const uint32_t av[4] = {a[qq].x, a[qq].y, a[qq].z, a[qq].w}, cv[4] = {c[qq].x, c[qq].y, c[qq].z, c[qq].w};
      float d = 0.f;
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 fa = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&av[e]));
        const float2 fc = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&cv[e]));
        d = fmaf(fa.x, fc.x, d);
        d = fmaf(fa.y, fc.y, d);
      }
      d += __shfl_xor_sync(0xffffffffu, d, 1);
      d += __shfl_xor_sync(0xffffffffu, d, 2);
      d += __shfl_xor_sync(0xffffffffu, d, 4);
      hsum[qq] = 0.f;
      if (row_ok && (lane & 7) == 0 && chunk < nchunk) {
        delta[((long long)(b * H + (chunk >> 3))) * T + i] = d;
        hsum[qq] = d;
      }
    }
    if (dmask_h) {
#pragma unroll
      for (int qq = 0; qq < 4; ++qq) {
        red[w][lane] = hsum[qq];
        __syncthreads();
        if (w == 0 && (lane & 7) == 0) {
          const int chunk = (q0 + qq) * 32 + lane;
          if (chunk < nchunk) {
            float t = 0.f;
            for (int k = 0; k < 8; ++k) t += red[k][lane];
            atomicAdd(dmask_h + (chunk >> 3), t);
          }
        }
        __syncthreads();
      }
    }
  }
}

// dq32 (M, inner) fp32 accumulator -> bf16 Q slot of dqkv, and clear the accumulator for the next layer.
// One thread per 8 consecutive columns (32-byte load, 32-byte zero store, 16-byte bf16 store); blockIdx.y walks rows.
__global__ void __launch_bounds__(256)
    attn_dq_convert_kernel(float* __restrict__ dq32, bf16* __restrict__ dqkv, long long M, int inner) {
  pdl_wait();
  pdl_trigger();
  const int per_row = inner >> 3;  // groups of 8 columns per row
  const long long total = M * per_row;
  const uint32_t zero[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
  for (long long it = (long long)blockIdx.x * blockDim.x + threadIdx.x; it < total;
       it += (long long)gridDim.x * blockDim.x) {
    const long long row = it / per_row;
    const int c8 = (int)(it - row * per_row);
    float* src = dq32 + row * inner + c8 * 8;
    uint32_t v[8];
    asm volatile("ld.global.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "l"(src)
                 : "memory");
    stg_256(src, zero);
    uint4 u;
    u.x = pack_bf16x2(__uint_as_float(v[0]), __uint_as_float(v[1]));
    u.y = pack_bf16x2(__uint_as_float(v[2]), __uint_as_float(v[3]));
    u.z = pack_bf16x2(__uint_as_float(v[4]), __uint_as_float(v[5]));
    u.w = pack_bf16x2(__uint_as_float(v[6]), __uint_as_float(v[7]));
    *reinterpret_cast<uint4*>(dqkv + row * (3LL * inner) + 2 * inner + c8 * 8) = u;
  }
}

// ------------------------------------------------------------------------------------------------------------
// backward, step 2: one CTA per (key tile, head, utterance); loops over the query tiles
// ------------------------------------------------------------------------------------------------------------
constexpr int BWD_THREADS = 256;                         // 8 warps: two per TMEM lane quadrant, each half the columns
constexpr int BWD_SK = 0;                                // K    [128 j][64]            16 KB
constexpr int BWD_SV = BWD_SK + ATT_BK * ATT_D * 2;      // V    [128 j][64]            16 KB
constexpr int BWD_SQ = BWD_SV + ATT_BK * ATT_D * 2;      // Q    2 x [128 i][64]        32 KB (double buffered)
constexpr int BWD_SDO = BWD_SQ + 2 * ATT_BQ * ATT_D * 2; // dO   2 x [128 i][64]        32 KB
constexpr int BWD_SP = BWD_SDO + 2 * ATT_BQ * ATT_D * 2; // Pd   2 x [128 i][64 j]      32 KB
constexpr int BWD_SDS = BWD_SP + ATT_BQ * ATT_BK * 2;    // dS   2 x [128 i][64 j]      32 KB
constexpr int BWD_TB = BWD_SDS + ATT_BQ * ATT_BK * 2;    // bias table (TB_PAD + 2T-1 floats)     8.5 KB
constexpr int BWD_DIAG = BWD_TB + (2048 + TB_PAD) * 4;   // embedding gradient per diagonal j - i (TB_PAD + 2T-1) 8.5 KB
constexpr int BWD_HIST = BWD_DIAG + (2048 + TB_PAD) * 4; // ... folded into buckets at the end    4 KB
constexpr int BWD_WDIAG = BWD_HIST + 1024 * 4;           // per-warp tile-local diagonal sums, 8 x 256 floats   8 KB
constexpr int BWD_BAR = BWD_WDIAG + 8 * 256 * 4;
constexpr int BWD_SMEM = BWD_BAR + 64 + 1024;

// Pipeline per query tile qt (tensor core work is issued by thread 0, everything else by all 8 warps):
//   S = Q K^T, dP = dO V^T (TMEM)  ->  P, dS (registers -> bf16 tiles in smem)  ->  dV += P^T dO, dK += dS^T Q,
//   dQ = dS K (TMEM)  ->  dQ accumulated into the fp32 buffer with 16-byte vector reductions.
// S / dP of tile qt+1 are issued right behind the gradient MMAs of tile qt (Q / dO are double buffered), so the
// tensor core works on them while the warps drain dQ.
template <bool BIAS, bool DROP>
__global__ void __launch_bounds__(BWD_THREADS, 1)
    attn_bwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
                    const __grid_constant__ AttnParams p, const float* __restrict__ lse,
                    const float* __restrict__ delta, float* __restrict__ dq32, bf16* __restrict__ dqkv,
                    float* __restrict__ dgate, float* __restrict__ demb, int num_buckets, int dq_bf16) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (sbase - smem_u32(smem_raw));
  const uint32_t sK = sbase + BWD_SK, sV = sbase + BWD_SV, sQ = sbase + BWD_SQ, sDO = sbase + BWD_SDO,
                 sP = sbase + BWD_SP, sDS = sbase + BWD_SDS;
  constexpr uint32_t QBUF = ATT_BQ * ATT_D * 2;  // bytes of one Q / dO buffer
  float* tb = reinterpret_cast<float*>(gbase + BWD_TB) + TB_PAD;
  float* sdiag = reinterpret_cast<float*>(gbase + BWD_DIAG) + TB_PAD;
  float* hist = reinterpret_cast<float*>(gbase + BWD_HIST);
  float* wdiag_all = reinterpret_cast<float*>(gbase + BWD_WDIAG);
  float* wdiag = wdiag_all + (threadIdx.x >> 5) * 256;  // this warp's private sums, index (j_local - i_local) + 127
  const uint32_t bar_kv = sbase + BWD_BAR, bar_qdo = bar_kv + 8 /* [2] */, bar_s = bar_kv + 24, bar_g = bar_kv + 32,
                 tmem_slot = bar_kv + 40;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int quad = warp & 3, half = warp >> 2;
  const int rit = quad * 32 + lane;  // row of this thread inside a 128-row tile
  const int kt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int T = p.T;
  const int k0 = kt * ATT_BK;
  const int nq = (T + ATT_BQ - 1) / ATT_BQ;

  if (tid == 0) {
    mbar_init(bar_kv, 1);
    mbar_init(bar_qdo, 1);
    mbar_init(bar_qdo + 8, 1);
    mbar_init(bar_s, 1);
    mbar_init(bar_g, 1);
    mbar_fence_init();
    tma_prefetch_desc(&tmQKV);
    tma_prefetch_desc(&tmDO);
  }
  if (warp == 0) {
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  pdl_wait();
  int htot = h;
  if constexpr (BIAS) {
    htot = p.head_index ? p.head_index[h] : h;
    for (int r = tid; r < 2 * T - 1; r += BWD_THREADS)
      tb[r] = __ldg(p.emb + (long long)p.rel_table[r] * p.Htot + htot) * kLog2e;
    for (int r = tid; r < TB_PAD; r += BWD_THREADS) tb[-1 - r] = 0.f;
    if (demb) {
      for (int r = tid; r < num_buckets; r += BWD_THREADS) hist[r] = 0.f;
      for (int r = tid - TB_PAD; r < 2048; r += BWD_THREADS) sdiag[r] = 0.f;
      for (int r = tid; r < 8 * 256; r += BWD_THREADS) wdiag_all[r] = 0.f;
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
  const uint32_t tS = tmem_base, tP = tmem_base + 128, tdV = tmem_base + 256, tdK = tmem_base + 320,
                 tdQ = tmem_base + 384;
  const uint32_t lane_off = ((uint32_t)(quad * 32)) << 16;
  constexpr uint32_t idesc_s = make_idesc_bf16(ATT_BQ, ATT_BK, false, false);   // S, dP : K-major x K-major
  constexpr uint32_t idesc_kv = make_idesc_bf16(ATT_BK, ATT_D, true, true);     // dV, dK: A^T (MN-major) x MN-major
  constexpr uint32_t idesc_q = make_idesc_bf16(ATT_BQ, ATT_D, false, true);     // dQ    : K-major x MN-major

  auto load_qdo = [&](int qt) {  // thread 0 only
    const uint32_t buf = (uint32_t)(qt & 1);
    mbar_expect_tx(bar_qdo + 8 * buf, 2 * ATT_BQ * ATT_D * 2);
    // 3-D maps (column, frame, utterance): frames >= T are zero-filled, never the next utterance's rows
    tma_load_3d(sQ + buf * QBUF, &tmQKV, bar_qdo + 8 * buf, 2 * p.inner + h * ATT_D, qt * ATT_BQ, b);
    tma_load_3d(sDO + buf * QBUF, &tmDO, bar_qdo + 8 * buf, h * ATT_D, qt * ATT_BQ, b);
  };
  auto issue_scores = [&](int qt) {  // thread 0 only: S and dP of query tile qt
    const uint32_t buf = (uint32_t)(qt & 1);
    mbar_wait(bar_qdo + 8 * buf, (uint32_t)((qt >> 1) & 1));
    tc_fence_after();
#pragma unroll
    for (int k = 0; k < ATT_D / 16; ++k)
      umma_f16_ss(tS, make_smem_desc(sQ + buf * QBUF + k * 32, 0, 1024, 2), make_smem_desc(sK + k * 32, 0, 1024, 2),
                  idesc_s, k > 0 ? 1u : 0u);
#pragma unroll
    for (int k = 0; k < ATT_D / 16; ++k)
      umma_f16_ss(tP, make_smem_desc(sDO + buf * QBUF + k * 32, 0, 1024, 2), make_smem_desc(sV + k * 32, 0, 1024, 2),
                  idesc_s, k > 0 ? 1u : 0u);
    umma_commit(bar_s);
  };

  if (warp == 0 && elect_one()) {
    mbar_expect_tx(bar_kv, 2 * ATT_BK * ATT_D * 2);
    tma_load_3d(sK, &tmQKV, bar_kv, h * ATT_D, k0, b);
    tma_load_3d(sV, &tmQKV, bar_kv, p.inner + h * ATT_D, k0, b);
    load_qdo(0);
    if (nq > 1) load_qdo(1);
    mbar_wait(bar_kv, 0);
    issue_scores(0);
  }
  const int len = p.lengths ? (int)p.lengths[b] : T;
  const uint32_t thresh = (uint32_t)(p.drop_p * 4294967296.0);
  const float inv_keep = p.drop_p > 0.f ? 1.f / (1.f - p.drop_p) : 1.f;
  const uint32_t dkey = DROP ? dropout_key(p.seed) : 0u;  // (after pdl_wait: the epoch is written earlier in the step)
  const float sc2 = p.scale * kLog2e;
  const float hm = p.head_mask ? __ldg(p.head_mask + h) : 1.f;

  for (int qt = 0; qt < nq; ++qt) {
    const uint32_t buf = (uint32_t)(qt & 1);
    const uint32_t ph = (uint32_t)(qt & 1);
    const int q0 = qt * ATT_BQ;
    const int i = q0 + rit;
    float L2 = 0.f, dl = 0.f, gate_i = 1.f;
    const bool row_ok = i < T;
    if (row_ok) {
      const long long ro = ((long long)(b * p.H + h)) * T + i;
      L2 = lse[ro] * kLog2e;
      dl = delta[ro] * hm;
      if constexpr (BIAS) {
        if (p.gate) gate_i = p.gate[((long long)b * p.Htot + htot) * T + i];
      }
    }
    const uint32_t drop_rkey =
        dropout_row_key(dkey, (unsigned long long)(b * p.H + h) * T + (unsigned long long)i);
    mbar_wait(bar_s, ph);
    tc_fence_after();
    float dg = 0.f;
    // Interior tile pairs (all 128 queries and keys valid, no padding mask) without a position bias take the lean
    // path: 5 instructions per score (FFMA, EX2, FFMA, FMUL, half a pack pair).  CTA-uniform choice.
    // Frames beyond T arrive as zeros (3-D tensor maps), which makes partial tiles harmless on the lean path: a
    // zero dO row adds nothing to dV, its dP and (L2 = delta = 0) dS vanish, and a zero K / V row adds nothing to dQ
    // while its own dK / dV rows are never stored.  Only a real padding mask (len < T) needs the general path.
    // (decided per 32-key chunk; with a position bias the garbage probabilities of keys beyond T would also leak into
    // the gate / embedding gradients, so there the chunk that straddles T takes the general path as well)
    const int lim = BIAS ? min(T, len) : (len < T ? len : T + ATT_BK);
    const float nL2 = -L2, ndl = -dl;
    auto do_chunk = [&](auto edge_tag, int cc) {
      constexpr bool EDGE = decltype(edge_tag)::value;
      const int c = half * 2 + cc;  // 32-key chunk of this warp
      uint32_t rs[32], rp[32];
      tmem_ld_32x32(tS + lane_off + c * 32, rs);
      tmem_ld_32x32(tP + lane_off + c * 32, rp);
      tmem_ld_wait();
      uint32_t pk[16], dk[16];
      [[maybe_unused]] const float* tp = tb + (k0 + c * 32 - i + T - 1);      // bias (pre-scaled by log2 e)
      [[maybe_unused]] float* wd = wdiag + (c * 32 - rit + 127);               // tile-local diagonal of key c*32 + 0
#pragma unroll
      for (int jj = 0; jj < 32; jj += 2) {
        float pv[2], dv[2];
        float dsc[2] = {1.f, 1.f};  // dropout scales of this key pair (one hash, same mask as the forward)
        if constexpr (DROP)
          dropout_row_pair(drop_rkey, (uint32_t)(k0 + c * 32 + jj) >> 1, thresh, inv_keep, dsc[0], dsc[1]);
        if constexpr (!EDGE) {
          // interior chunk: the key pair on packed fp32x2 lanes (FFMA2 / FMUL2), two MUFU.EX2
          float e0, e1;
          [[maybe_unused]] float b0 = 0.f, b1 = 0.f;
          if constexpr (BIAS) {
            b0 = tp[jj];
            b1 = tp[jj + 1];
            e0 = fmaf(__uint_as_float(rs[jj]), sc2, fmaf(gate_i, b0, nL2));
            e1 = fmaf(__uint_as_float(rs[jj + 1]), sc2, fmaf(gate_i, b1, nL2));
          } else {
            upk2(fma2(pk2(__uint_as_float(rs[jj]), __uint_as_float(rs[jj + 1])), bc2(sc2), bc2(nL2)), e0, e1);
          }
          const f32x2 pe2 = pk2(fast_exp2(e0), fast_exp2(e1));
          const f32x2 rp2 = pk2(__uint_as_float(rp[jj]), __uint_as_float(rp[jj + 1]));
          f32x2 t2, pv2 = pe2;
          if constexpr (DROP) {
            const f32x2 d2 = pk2(dsc[0], dsc[1]);
            t2 = fma2(rp2, mul2(d2, bc2(hm)), bc2(ndl));
            pv2 = mul2(pe2, d2);
          } else {
            t2 = fma2(rp2, bc2(hm), bc2(ndl));
          }
          upk2(pv2, pv[0], pv[1]);
          upk2(mul2(pe2, t2), dv[0], dv[1]);
          if constexpr (BIAS) {
            // rows beyond T carry dS = 0 (their dO, delta are zero), keys are all real in an interior chunk
            dg = fmaf(dv[0], b0, fmaf(dv[1], b1, dg));
            if (demb) {
              // embedding gradient: gate_i * dS_ij summed per diagonal j - i.  The 32 lanes of a warp hold 32
              // consecutive rows at the same key, i.e. 32 distinct diagonals per instruction: plain read-modify-write
              // on the warp's private array (shared-memory float atomics are compare-and-swap loops); successive keys
              // alias across lanes, hence the __syncwarp between them
              wd[jj] = fmaf(dv[0], gate_i, wd[jj]);
              __syncwarp();
              wd[jj + 1] = fmaf(dv[1], gate_i, wd[jj + 1]);
              __syncwarp();
            }
          }
        }
#pragma unroll
        for (int u = 0; u < (EDGE ? 2 : 0); ++u) {
          const int j = k0 + c * 32 + jj + u;
          float pe;
          [[maybe_unused]] float tbv = 0.f;
          [[maybe_unused]] bool ok = true;
          if constexpr (!EDGE) {
            pe = fast_exp2(fmaf(__uint_as_float(rs[jj + u]), sc2, nL2));
          } else {
            float x = fmaf(__uint_as_float(rs[jj + u]), sc2, nL2);
            if constexpr (BIAS) {
              tbv = tb[min(max(j - i + T - 1, -TB_PAD), 2 * T - 2)];
              x += gate_i * tbv;
            }
            if (j >= len) x += -10000.0f * kLog2e;
            ok = row_ok && (j < T);
            pe = ok ? fast_exp2(x) : 0.f;
          }
          // d loss / d p_ij = dP * head mask (* dropout scale); dS = P (dP - delta)
          float t;
          if constexpr (DROP) t = fmaf(__uint_as_float(rp[jj + u]) * hm, dsc[u], ndl);
          else t = fmaf(__uint_as_float(rp[jj + u]), hm, ndl);
          const float ds = pe * t;
          pv[u] = DROP ? pe * dsc[u] : pe;
          dv[u] = ds;
          if constexpr (BIAS) {
            if (ok) dg += ds * tbv;
            if (demb) {  // all lanes take part (warp-uniform condition): masked rows / keys add zero
              wd[jj + u] = fmaf(ok ? ds : 0.f, gate_i, wd[jj + u]);
              __syncwarp();
            }
          }
        }
        pk[jj >> 1] = pack_bf16x2(pv[0], pv[1]);
        dk[jj >> 1] = pack_bf16x2(dv[0], dv[1]);
      }
      const uint32_t off = (uint32_t)(c >> 1) * (ATT_BQ * 128);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint32_t so = off + sw128(rit, (c & 1) * 4 + q);
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sP + so), "r"(pk[4 * q]), "r"(pk[4 * q + 1]),
                     "r"(pk[4 * q + 2]), "r"(pk[4 * q + 3])
                     : "memory");
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sDS + so), "r"(dk[4 * q]), "r"(dk[4 * q + 1]),
                     "r"(dk[4 * q + 2]), "r"(dk[4 * q + 3])
                     : "memory");
      }
    };
#pragma unroll 1
    for (int cc = 0; cc < 2; ++cc) {
      if (k0 + (half * 2 + cc) * 32 + 32 > lim) do_chunk(std::true_type{}, cc);
      else do_chunk(std::false_type{}, cc);
    }
    // (the table is pre-scaled by log2 e: undo it for the gate gradient)
    if (BIAS && dgate && p.gate && row_ok)
      atomicAdd(dgate + ((long long)b * p.Htot + htot) * T + i, dg * (1.f / kLog2e));
    fence_async_smem();
    tc_fence_before();
    __syncthreads();  // P / dS complete, S / dP consumed, dQ of the previous tile drained
    if (warp == 0 && elect_one()) {
      tc_fence_after();
#pragma unroll
      for (int k = 0; k < ATT_BQ / 16; ++k) {  // reduction over the 128 queries of this tile
        const uint32_t acc = (qt > 0 || k > 0) ? 1u : 0u;
        // dV[j, e] += sum_i Pd[i, j] dO[i, e]
        umma_f16_ss(tdV, make_smem_desc(sP + k * 2048, ATT_BQ * 128, 1024, 2),
                    make_smem_desc(sDO + buf * QBUF + k * 2048, ATT_BQ * 128, 1024, 2), idesc_kv, acc);
        // dK[j, e] += sum_i dS[i, j] Q[i, e]
        umma_f16_ss(tdK, make_smem_desc(sDS + k * 2048, ATT_BQ * 128, 1024, 2),
                    make_smem_desc(sQ + buf * QBUF + k * 2048, ATT_BQ * 128, 1024, 2), idesc_kv, acc);
      }
#pragma unroll
      for (int k = 0; k < ATT_BK / 16; ++k)  // dQ[i, e] = sum_j dS[i, j] K[j, e]
        umma_f16_ss(tdQ, make_smem_desc(sDS + (k >> 2) * (ATT_BQ * 128) + (k & 3) * 32, 0, 1024, 2),
                    make_smem_desc(sK + k * 2048, ATT_BK * 128, 1024, 2), idesc_q, k > 0 ? 1u : 0u);
      umma_commit(bar_g);
      if (qt + 1 < nq) issue_scores(qt + 1);  // runs on the tensor core while the warps drain dQ below
    }
    if constexpr (BIAS) {
      if (demb) {
        // fold the eight warps' tile-local diagonal sums into the utterance-level array (one thread per diagonal, each
        // owning its slot: no atomics) and clear them for the next query tile
        if (tid < ATT_BQ + ATT_BK - 1) {
          float acc = 0.f;
#pragma unroll
          for (int w8 = 0; w8 < 8; ++w8) {
            acc += wdiag_all[w8 * 256 + tid];
            wdiag_all[w8 * 256 + tid] = 0.f;
          }
          sdiag[(k0 - q0) + (tid - (ATT_BQ - 1)) + T - 1] += acc;
        }
        __syncthreads();  // the warp arrays are clear before the next query tile adds to them
      }
    }
    mbar_wait(bar_g, ph);
    tc_fence_after();
    if (warp == 0 && elect_one() && qt + 2 < nq) load_qdo(qt + 2);  // this tile's Q / dO buffer is free again
    {
      uint32_t r[32];
      tmem_ld_32x32(tdQ + lane_off + half * 32, r);
      tmem_ld_wait();
      if (row_ok && dq_bf16) {
        // bf16x2 vector reductions straight into the (pre-zeroed) Q slot of dqkv: 8 values per 16-byte RED, a quarter
        // of the L2 atomic-ALU work of the fp32 path (which bounds this kernel), and no conversion pass afterwards
        bf16* dst = dqkv + ((long long)(b * T + i)) * (3LL * p.inner) + 2 * p.inner + h * ATT_D + half * 32;
#pragma unroll
        for (int jj = 0; jj < 32; jj += 8) {
          const uint32_t u0 = pack_bf16x2(__uint_as_float(r[jj]) * p.scale, __uint_as_float(r[jj + 1]) * p.scale);
          const uint32_t u1 = pack_bf16x2(__uint_as_float(r[jj + 2]) * p.scale, __uint_as_float(r[jj + 3]) * p.scale);
          const uint32_t u2 = pack_bf16x2(__uint_as_float(r[jj + 4]) * p.scale, __uint_as_float(r[jj + 5]) * p.scale);
          const uint32_t u3 = pack_bf16x2(__uint_as_float(r[jj + 6]) * p.scale, __uint_as_float(r[jj + 7]) * p.scale);
          asm volatile("red.global.add.noftz.v4.bf16x2 [%0], {%1, %2, %3, %4};" ::"l"(dst + jj), "r"(u0), "r"(u1),
                       "r"(u2), "r"(u3)
                       : "memory");
        }
      } else if (row_ok) {
        float* dst = dq32 + ((long long)(b * T + i)) * p.inner + h * ATT_D + half * 32;
#pragma unroll
        for (int jj = 0; jj < 32; jj += 4)
          red_add_v4(dst + jj, __uint_as_float(r[jj]) * p.scale, __uint_as_float(r[jj + 1]) * p.scale,
                     __uint_as_float(r[jj + 2]) * p.scale, __uint_as_float(r[jj + 3]) * p.scale);
      }
    }
    tc_fence_before();  // orders this thread's TMEM reads before the next tile's __syncthreads / MMAs
  }
  // ---- epilogue: dV (x head mask, warps 0-3) and dK (x scale, warps 4-7) of this key tile
  pdl_trigger();
  {
    const int j = k0 + rit;
    const uint32_t ta = half == 0 ? tdV : tdK;
    const float mul = half == 0 ? hm : p.scale;
    bf16* dst = dqkv + ((long long)(b * T + j)) * (3LL * p.inner) + (half == 0 ? p.inner : 0) + h * ATT_D;
#pragma unroll
    for (int c = 0; c < ATT_D / 32; ++c) {
      uint32_t r[32];
      tmem_ld_32x32(ta + lane_off + c * 32, r);
      tmem_ld_wait();
      if (j < T) {
#pragma unroll
        for (int e = 0; e < 32; e += 8) {
          uint4 u;
          u.x = pack_bf16x2(__uint_as_float(r[e]) * mul, __uint_as_float(r[e + 1]) * mul);
          u.y = pack_bf16x2(__uint_as_float(r[e + 2]) * mul, __uint_as_float(r[e + 3]) * mul);
          u.z = pack_bf16x2(__uint_as_float(r[e + 4]) * mul, __uint_as_float(r[e + 5]) * mul);
          u.w = pack_bf16x2(__uint_as_float(r[e + 6]) * mul, __uint_as_float(r[e + 7]) * mul);
          *reinterpret_cast<uint4*>(dst + c * 32 + e) = u;
        }
      }
    }
  }
  if (demb && p.emb) {
    __syncthreads();
    for (int r = tid; r < 2 * T - 1; r += BWD_THREADS) {
      const float v = sdiag[r];
      if (v != 0.f) atomicAdd(&hist[p.rel_table[r]], v);
    }
    __syncthreads();
    for (int r = tid; r < num_buckets; r += BWD_THREADS) {
      const float v = hist[r];
      if (v != 0.f) atomicAdd(demb + (long long)r * p.Htot + htot, v);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace dph

using namespace dph;

namespace dph {
int make_tmap_2d(CUtensorMap* out, const void* ptr, uint64_t cols, uint64_t rows, uint64_t row_stride_elems,
                 uint32_t box_cols, uint32_t box_rows);
int make_tmap_nd_bf16(CUtensorMap* out, const void* ptr, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box, int swizzle);
}

static int attn_check(int B, int H, int T, int num_buckets) {
  if (B <= 0 || H <= 0 || T <= 0 || T > 1024 || num_buckets > 1024 || B > 65535 || H > 65535) {
    set_error("fused attention: unsupported shape B=%d H=%d T=%d (head_dim must be 64, T <= 1024)", B, H, T);
    return DPH_ERR_UNSUPPORTED;
  }
  return DPH_OK;
}

static AttnParams attn_params(int B, int H, int T, float scale, const int64_t* lengths, const int* rel_table,
                              const float* emb, const float* gate, const int* head_index, int Htot, float drop_p,
                              uint64_t seed, const float* head_mask) {
  AttnParams p;
  p.B = B;
  p.H = H;
  p.T = T;
  p.inner = H * ATT_D;
  p.scale = scale;
  p.lengths = reinterpret_cast<const long long*>(lengths);
  p.rel_table = rel_table;
  p.emb = emb;
  p.gate = gate;
  p.head_index = head_index;
  p.Htot = Htot;
  p.drop_p = drop_p;
  p.seed = seed;
  p.head_mask = head_mask;
  return p;
}

extern "C" int dph_attn_fwd(const void* qkv, void* ctx, void* ctx_un, float* lse, int B, int H, int T, float scale,
                            const int64_t* lengths, const int* rel_table, const float* emb, const float* gate,
                            const int* head_index, int Htot, float drop_p, uint64_t seed, const float* head_mask,
                            void* stream) {
  int rc = attn_check(B, H, T, 0);
  if (rc) return rc;
  if (!qkv || !ctx) {
    set_error("attn_fwd: null pointer");
    return DPH_ERR_ARG;
  }
  const int inner = H * ATT_D;
  // keys per tile: 64 (three CTAs per SM) unless DPH_ATT_BK=128 asks for the two-per-SM variant (A / B timing)
  static const int fwd_bk = [] {
    const char* e = getenv("DPH_ATT_BK");
    return (e != nullptr && atoi(e) == 128) ? 128 : 64;
  }();
  CUtensorMap tm, tmkv;
  rc = make_tmap_2d(&tm, qkv, 3ull * inner, (uint64_t)B * T, 3ull * inner, ATT_D, ATT_BQ);
  if (rc) return rc;
  rc = make_tmap_2d(&tmkv, qkv, 3ull * inner, (uint64_t)B * T, 3ull * inner, ATT_D, fwd_bk);
  if (rc) return rc;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaSuccess;
#define ATT_ATTR(BI, DR)                                                                                                   \
  if (e == cudaSuccess) e = cudaFuncSetAttribute(attn_fwd_kernel<BI, DR, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, FwdL<64>::SMEM); \
  if (e == cudaSuccess) e = cudaFuncSetAttribute(attn_fwd_kernel<BI, DR, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize, FwdL<128>::SMEM);
    ATT_ATTR(false, false) ATT_ATTR(false, true) ATT_ATTR(true, false) ATT_ATTR(true, true)
#undef ATT_ATTR
    if (e != cudaSuccess) {
      set_error("attn_fwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return DPH_ERR_CUDA;
    }
    attr = true;
  }
  AttnParams p = attn_params(B, H, T, scale, lengths, rel_table, emb, gate, head_index, Htot, drop_p, seed, head_mask);
  dim3 grid((T + ATT_BQ - 1) / ATT_BQ, H, B);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const bool bias = emb != nullptr, drop = drop_p > 0.f;
#define ATT_FWD(BI, DR)                                                                                              \
  do {                                                                                                               \
    if (fwd_bk == 64)                                                                                                \
      launch_pdl(attn_fwd_kernel<BI, DR, 64>, grid, dim3(128), FwdL<64>::SMEM, s, tm, tmkv, p, (bf16*)ctx,           \
                 (bf16*)ctx_un, lse);                                                                                \
    else                                                                                                             \
      launch_pdl(attn_fwd_kernel<BI, DR, 128>, grid, dim3(128), FwdL<128>::SMEM, s, tm, tmkv, p, (bf16*)ctx,         \
                 (bf16*)ctx_un, lse);                                                                                \
  } while (0)
  if (bias) {
    if (drop) ATT_FWD(true, true); else ATT_FWD(true, false);
  } else {
    if (drop) ATT_FWD(false, true); else ATT_FWD(false, false);
  }
#undef ATT_FWD
  dph_count_launch();
  return dph_check_last("attn_fwd");
}

// delta: (B, H, T) fp32 scratch; dq32: (B*T, H*64) fp32 accumulator that must be zero on entry (it is left zeroed).
extern "C" int dph_attn_bwd(const void* qkv, const void* dctx, const void* o_un, const float* lse, float* delta,
                            float* dq32, void* dqkv, float* dmask_h, int B, int H, int T, float scale,
                            const int64_t* lengths, const int* rel_table, const float* emb, const float* gate,
                            const int* head_index, int Htot, float drop_p, uint64_t seed, const float* head_mask,
                            float* dgate, float* demb, int num_buckets, void* stream) {
  int rc = attn_check(B, H, T, num_buckets);
  if (rc) return rc;
  if (!qkv || !dctx || !o_un || !lse || !delta || !dq32 || !dqkv) {
    set_error("attn_bwd: null pointer");
    return DPH_ERR_ARG;
  }
  const int inner = H * ATT_D;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  CUtensorMap tmq, tmd;
  {
    const uint64_t qd[3] = {3ull * inner, (uint64_t)T, (uint64_t)B};
    const uint64_t qs[2] = {3ull * inner * 2, (uint64_t)T * 3ull * inner * 2};
    const uint32_t box[3] = {ATT_D, ATT_BK, 1};
    rc = make_tmap_nd_bf16(&tmq, qkv, 3, qd, qs, box, 3);
    if (rc) return rc;
    const uint64_t dd[3] = {(uint64_t)inner, (uint64_t)T, (uint64_t)B};
    const uint64_t ds[2] = {(uint64_t)inner * 2, (uint64_t)T * inner * 2};
    rc = make_tmap_nd_bf16(&tmd, dctx, 3, dd, ds, box, 3);
  }
  if (rc) return rc;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(attn_bwd_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(attn_bwd_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(attn_bwd_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(attn_bwd_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM);
    if (e != cudaSuccess) {
      set_error("attn_bwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return DPH_ERR_CUDA;
    }
    attr = true;
  }
  const long long M = (long long)B * T;
  static const int dq_bf16 = [] {
    const char* e = getenv("DPH_DQ_BF16");
    return (e != nullptr && e[0] == '0') ? 0 : 1;
  }();
  launch_pdl(attn_bwd_prep_kernel, dim3((unsigned)((M + 8 * PREP_RPW - 1) / (8 * PREP_RPW))), dim3(256), 0, s,
             (const bf16*)dctx, (const bf16*)o_un, delta, dmask_h, B, H, T, inner,
             dq_bf16 ? (bf16*)dqkv : (bf16*)nullptr);
  AttnParams p = attn_params(B, H, T, scale, lengths, rel_table, emb, gate, head_index, Htot, drop_p, seed, head_mask);
  dim3 grid((T + ATT_BK - 1) / ATT_BK, H, B);
  const bool bias = emb != nullptr, drop = drop_p > 0.f;
#define ATT_BWD(BI, DR)                                                                                             \
  launch_pdl(attn_bwd_kernel<BI, DR>, grid, dim3(BWD_THREADS), BWD_SMEM, s, tmq, tmd, p, lse, delta, dq32,          \
             (bf16*)dqkv, dgate, demb, num_buckets, dq_bf16)
  if (bias) {
    if (drop) ATT_BWD(true, true); else ATT_BWD(true, false);
  } else {
    if (drop) ATT_BWD(false, true); else ATT_BWD(false, false);
  }
#undef ATT_BWD
  if (!dq_bf16) {
    const long long items = M * (inner / 8);
    long long nb = (items + 255) / 256;
    if (nb > 148LL * 16) nb = 148LL * 16;
    launch_pdl(attn_dq_convert_kernel, dim3((unsigned)nb), dim3(256), 0, s, dq32, (bf16*)dqkv, M, inner);
  }
  dph_count_launch(dq_bf16 ? 2 : 3);
  return dph_check_last("attn_bwd");
}
