// bf16 GEMM on the 5th-generation tensor cores (tcgen05.mma, fp32 accumulators in TMEM, TMA-fed smem ring).
//
// Persistent kernel, one CTA per SM, 10 warps:
//   warp 0      TMA producer   (one lane issues cp.async.bulk.tensor into a STAGES-deep smem ring)
//   warp 1      MMA issuer     (one lane issues tcgen05.mma; owns the TMEM allocation)
//   warps 2..9  epilogue       (tcgen05.ld TMEM -> registers -> fused epilogue -> global; two warps per TMEM lane
//                               quadrant, each taking half of the tile's columns)
// Three mbarrier pipelines: smem full/empty (TMA <-> MMA), TMEM full/empty (MMA <-> epilogue, two accumulator
// buffers so the epilogue of tile i overlaps the mainloop of tile i+1), and the static persistent tile walk.
//
// CG = 1: tile 128 x BN x 64 per CTA.
// CG = 2: the two CTAs of a cluster (one TPC) work on one 256 x BN tile with tcgen05.mma.cta_group::2: each CTA
//         stages its own 128 rows of A and only BN/2 rows of B, the leader CTA issues the MMAs for both, every CTA
//         runs the epilogue of its own 128 rows.  Halves the B traffic from L2 and the smem reads per flop.
// Operands may be K-major or MN-major (the latter is what weight-gradient products need: their reduction index is
// the slow memory dimension).
#include "common.cuh"
#include "gemm.h"

namespace dph {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int kEpiWarps = 8;
constexpr int kGemmThreads = 64 + 32 * kEpiWarps;

template <int BN, int CG>
struct GemmCfg {
  static constexpr int B_ROWS = BN / CG;  // rows of B staged by one CTA
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = B_ROWS * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES_RAW = (200 * 1024) / STAGE_BYTES;
  static constexpr int STAGES = STAGES_RAW > 8 ? 8 : STAGES_RAW;
  static constexpr int TMEM_COLS = (2 * BN <= 128) ? 128 : ((2 * BN <= 256) ? 256 : 512);
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
};

struct TileCoord {
  int m0, n0, b0, b1, ks;
};

// m0 is the first row of the (128 * CG)-row tile; n0 is the N-tile index (multiplied by BN by the caller).
template <int CG>
DPH_DEVICE TileCoord decode_tile(const GemmParams& p, int t) {
  TileCoord c;
  c.ks = t % p.ksplit;
  t /= p.ksplit;
  int nb = t % p.ntiles;
  t /= p.ntiles;
  int mb = t % p.mtiles;
  t /= p.mtiles;
  c.b0 = t % p.nb0;
  c.b1 = t / p.nb0;
  c.m0 = mb * (BM * CG);
  c.n0 = nb;
  return c;
}

// Generic (ragged / unaligned) epilogue of one 32-column chunk: scalar accesses, any nvalid. Out of line so that its
// dynamically indexed arrays do not force the fast path's registers onto the stack.
__device__ __noinline__ void epilogue_chunk_generic(const GemmParams& p, const TileCoord& tc, int row, int col0,
                                                    const uint32_t (&r)[32]) {
  // row: row index inside the batch (already checked < M); col0: first column of this 32-wide chunk.
  const int nvalid = min(32, p.N - col0);
  if (nvalid <= 0) return;
  const long long boff = (long long)tc.b0 * p.c_sb0 + (long long)tc.b1 * p.c_sb1;
  const long long coff = boff + (long long)row * p.ldc + col0;
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * p.alpha;

  if (p.bias != nullptr) {
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (j < nvalid) v[j] += __ldg(p.bias + tc.b0 * p.vec_sb0 + col0 + j);
  }
  const bool full = (nvalid == 32) && p.vec_ok;
  if (p.C2 != nullptr) {
    if (p.c_dtype == 0) {
      bf16* o = reinterpret_cast<bf16*>(p.C2) + coff;
      if (full) {
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
          uint4 u;
          u.x = pack_bf16x2(v[j], v[j + 1]);
          u.y = pack_bf16x2(v[j + 2], v[j + 3]);
          u.z = pack_bf16x2(v[j + 4], v[j + 5]);
          u.w = pack_bf16x2(v[j + 6], v[j + 7]);
          *reinterpret_cast<uint4*>(o + j) = u;
        }
      } else {
        for (int j = 0; j < nvalid; ++j) o[j] = __float2bfloat16_rn(v[j]);
      }
    } else {
      float* o = reinterpret_cast<float*>(p.C2) + coff;
      for (int j = 0; j < nvalid; ++j) o[j] = v[j];
    }
  }
  if (p.act == 1) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = gelu_erf_fast(v[j]);
  }
  if (p.colmask != nullptr) {
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (j < nvalid) v[j] *= __ldg(p.colmask + tc.b0 * p.vec_sb0 + col0 + j);
  }
  if (p.scalar_mask != nullptr) {
    const float s = __ldg(p.scalar_mask + tc.b0 * p.scalar_sb0);
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] *= s;
  }
  if (p.drop_p > 0.0f) {
    const uint32_t thresh = (uint32_t)(p.drop_p * 4294967296.0);
    const float inv_keep = 1.0f / (1.0f - p.drop_p);
    const unsigned long long base =
        (((unsigned long long)tc.b1 * p.nb0 + tc.b0) * (unsigned long long)p.M + row) * (unsigned long long)p.N + col0;
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] *= dropout_scale(p.seed, base + j, thresh, inv_keep);
  }
  if (p.residual != nullptr) {
    const long long roff = (long long)tc.b0 * p.r_sb0 + (long long)tc.b1 * p.r_sb1 + (long long)row * p.ldr + col0;
    if (p.r_dtype == 0) {
      const bf16* rp = reinterpret_cast<const bf16*>(p.residual) + roff;
      if (full) {
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
          uint4 u = *reinterpret_cast<const uint4*>(rp + j);
          const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float2 f = __bfloat1622float2(h[q]);
            v[j + 2 * q] += f.x;
            v[j + 2 * q + 1] += f.y;
          }
        }
      } else {
        for (int j = 0; j < nvalid; ++j) v[j] += __bfloat162float(rp[j]);
      }
    } else {
      const float* rp = reinterpret_cast<const float*>(p.residual) + roff;
      for (int j = 0; j < nvalid; ++j) v[j] += rp[j];
    }
  }
  // ---- store ----
  if (p.c_dtype == 0) {
    bf16* o = reinterpret_cast<bf16*>(p.C) + coff;
    if (p.c_mode == 2) {
      for (int j = 0; j < nvalid; ++j) v[j] += __bfloat162float(o[j]);
    }
    if (full) {
#pragma unroll
      for (int j = 0; j < 32; j += 8) {
        uint4 u;
        u.x = pack_bf16x2(v[j], v[j + 1]);
        u.y = pack_bf16x2(v[j + 2], v[j + 3]);
        u.z = pack_bf16x2(v[j + 4], v[j + 5]);
        u.w = pack_bf16x2(v[j + 6], v[j + 7]);
        *reinterpret_cast<uint4*>(o + j) = u;
      }
    } else {
      for (int j = 0; j < nvalid; ++j) o[j] = __float2bfloat16_rn(v[j]);
    }
  } else {
    float* o = reinterpret_cast<float*>(p.C) + coff;
    if (p.c_mode == 1) {
      if (full) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) red_add_v4(o + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
      } else {
        for (int j = 0; j < nvalid; ++j) atomicAdd(o + j, v[j]);
      }
    } else if (full) {
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        float4 f = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        if (p.c_mode == 2) {
          float4 old = *reinterpret_cast<const float4*>(o + j);
          f.x += old.x;
          f.y += old.y;
          f.z += old.z;
          f.w += old.w;
        }
        *reinterpret_cast<float4*>(o + j) = f;
      }
    } else {
      for (int j = 0; j < nvalid; ++j) o[j] = (p.c_mode == 2 ? o[j] : 0.0f) + v[j];
    }
  }
}


DPH_DEVICE void unpack_bf16x8(const uint4& u, float* f) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    float2 t = __bfloat1622float2(h[q]);
    f[2 * q] = t.x;
    f[2 * q + 1] = t.y;
  }
}

// Fast path: all 32 columns valid and every pointer 16-byte aligned (p.vec_ok). Every global load of the chunk
// (bias, column mask, residual; all read-only in this kernel) is issued up front as an independent 16-byte load so
// that their latencies overlap; no dynamically indexed arrays.
DPH_DEVICE void epilogue_chunk(const GemmParams& p, const TileCoord& tc, int row, int col0, uint32_t (&r)[32]) {
  if (!(p.vec_ok && col0 + 32 <= p.N)) {
    if (col0 < p.N) epilogue_chunk_generic(p, tc, row, col0, r);
    return;
  }
  const long long coff = (long long)tc.b0 * p.c_sb0 + (long long)tc.b1 * p.c_sb1 + (long long)row * p.ldc + col0;
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * p.alpha;

  uint4 res[4];
  const bool has_res16 = (p.residual != nullptr) && (p.r_dtype == 0);
  if (has_res16) {
    const long long roff = (long long)tc.b0 * p.r_sb0 + (long long)tc.b1 * p.r_sb1 + (long long)row * p.ldr + col0;
    const uint4* rp = reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(p.residual) + roff);
#pragma unroll
    for (int j = 0; j < 4; ++j) res[j] = __ldg(rp + j);
  }
  if (p.bias != nullptr) {
    const float4* bp = reinterpret_cast<const float4*>(p.bias + tc.b0 * p.vec_sb0 + col0);
    float4 bv[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) bv[j] = __ldg(bp + j);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      v[4 * j] += bv[j].x;
      v[4 * j + 1] += bv[j].y;
      v[4 * j + 2] += bv[j].z;
      v[4 * j + 3] += bv[j].w;
    }
  }
  if (p.C2 != nullptr) {
    if (p.c_dtype == 0) {
      uint4* o = reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(p.C2) + coff);
#pragma unroll
      for (int j = 0; j < 4; ++j)
        o[j] = make_uint4(pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                          pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
    } else {
      float4* o = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.C2) + coff);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    }
  }
  if (p.act == 1) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = gelu_erf_fast(v[j]);
  }
  if (p.colmask != nullptr) {
    const float4* mp = reinterpret_cast<const float4*>(p.colmask + tc.b0 * p.vec_sb0 + col0);
    float4 mv[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) mv[j] = __ldg(mp + j);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      v[4 * j] *= mv[j].x;
      v[4 * j + 1] *= mv[j].y;
      v[4 * j + 2] *= mv[j].z;
      v[4 * j + 3] *= mv[j].w;
    }
  }
  if (p.scalar_mask != nullptr) {
    const float s = __ldg(p.scalar_mask + tc.b0 * p.scalar_sb0);
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] *= s;
  }
  if (p.drop_p > 0.0f) {
    const uint32_t thresh = (uint32_t)(p.drop_p * 4294967296.0);
    const float inv_keep = 1.0f / (1.0f - p.drop_p);
    const unsigned long long base =
        (((unsigned long long)tc.b1 * p.nb0 + tc.b0) * (unsigned long long)p.M + row) * (unsigned long long)p.N + col0;
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] *= dropout_scale(p.seed, base + j, thresh, inv_keep);
  }
  if (has_res16) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float f[8];
      unpack_bf16x8(res[j], f);
#pragma unroll
      for (int q = 0; q < 8; ++q) v[8 * j + q] += f[q];
    }
  } else if (p.residual != nullptr) {
    const long long roff = (long long)tc.b0 * p.r_sb0 + (long long)tc.b1 * p.r_sb1 + (long long)row * p.ldr + col0;
    const float4* rp = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.residual) + roff);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float4 f = __ldg(rp + j);
      v[4 * j] += f.x;
      v[4 * j + 1] += f.y;
      v[4 * j + 2] += f.z;
      v[4 * j + 3] += f.w;
    }
  }
  // ---- store ----
  if (p.c_dtype == 0) {
    uint4* o = reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(p.C) + coff);
    if (p.c_mode == 2) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float f[8];
        const uint4 old = o[j];
        unpack_bf16x8(old, f);
#pragma unroll
        for (int q = 0; q < 8; ++q) v[8 * j + q] += f[q];
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
      o[j] = make_uint4(pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                        pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
  } else {
    float* o = reinterpret_cast<float*>(p.C) + coff;
    if (p.c_mode == 1) {
#pragma unroll
      for (int j = 0; j < 32; j += 4) red_add_v4(o + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
    } else {
      float4* o4 = reinterpret_cast<float4*>(o);
      if (p.c_mode == 2) {
        float4 old[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) old[j] = o4[j];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          v[4 * j] += old[j].x;
          v[4 * j + 1] += old[j].y;
          v[4 * j + 2] += old[j].z;
          v[4 * j + 3] += old[j].w;
        }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) o4[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    }
  }
}

template <int BN, bool A_MN, bool B_MN, int CG>
__global__ void __launch_bounds__(kGemmThreads, 1)
    gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                     const GemmParams p) {
  using Cfg = GemmCfg<BN, CG>;
  constexpr int STAGES = Cfg::STAGES;
  static_assert(CG == 1 || CG == 2, "CTA group is 1 or 2");
  static_assert(!(B_MN && CG == 2) || (BN % 128 == 0), "MN-major B halves must be whole 64-column boxes");
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_base = smem_base + STAGES * Cfg::STAGE_BYTES;
  // barrier layout (8 bytes each): full[STAGES], empty[STAGES], tmem_full[2], tmem_empty[2], then tmem ptr slot
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + 2 + a); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int rank = (CG == 2) ? (int)cluster_ctarank() : 0;  // 0 = leader (issues the MMAs)
  const int unit = blockIdx.x / CG;                         // persistent work unit = CTA or CTA pair
  const int nunits = gridDim.x / CG;
  const int total_tiles = p.nb1 * p.nb0 * p.mtiles * p.ntiles * p.ksplit;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < STAGES; ++s) {
        mbar_init(full_bar(s), 1);   // leader: its own arrive.expect_tx (transaction bytes of both CTAs)
        mbar_init(empty_bar(s), 1);  // one (multicast) tcgen05.commit
      }
      for (int a = 0; a < 2; ++a) {
        mbar_init(tfull_bar(a), 1);
        mbar_init(tempty_bar(a), kEpiWarps * CG);  // leader's barrier collects the epilogue warps of both CTAs
      }
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc_cg<CG>(tmem_slot, Cfg::TMEM_COLS);
    tmem_relinquish_cg<CG>();
  }
  tc_fence_before();
  if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp == 0) {
    // ======================= TMA producer (both CTAs of a pair) =======================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = unit; t < total_tiles; t += nunits) {
        TileCoord tc = decode_tile<CG>(p, t);
        const int m0 = tc.m0 + rank * BM;                    // this CTA's rows of A
        const int n0 = tc.n0 * BN + rank * Cfg::B_ROWS;      // this CTA's share of the B tile
        const int kb_begin = tc.ks * p.kb_per_split;
        const int kb_end = min(p.num_kb, kb_begin + p.kb_per_split);
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t sA = smem_base + stage * Cfg::STAGE_BYTES;
          const uint32_t sB = sA + Cfg::A_BYTES;
          uint32_t fb = full_bar(stage);
          if constexpr (CG == 2) {
            if (rank == 0) mbar_expect_tx(fb, 2 * Cfg::STAGE_BYTES);
            fb = mapa_shared(fb, 0);
          } else {
            mbar_expect_tx(fb, Cfg::STAGE_BYTES);
          }
          auto load = [&](uint32_t dst, const CUtensorMap* m, int c0, int c1, int c2, int c3, int c4) {
            if constexpr (CG == 2) tma_load_5d_2sm(dst, m, fb, c0, c1, c2, c3, c4);
            else tma_load_5d(dst, m, fb, c0, c1, c2, c3, c4);
          };
          if constexpr (!A_MN) {
            load(sA, &tmA, (kb % p.a_kdiv) * BK, m0, kb / p.a_kdiv, tc.b0 * p.a_b0, tc.b1 * p.a_b1);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j)
              load(sA + j * (BK * 128), &tmA, m0 + 64 * j, kb * BK, tc.b0 * p.a_b0, tc.b1 * p.a_b1, 0);
          }
          if constexpr (!B_MN) {
            load(sB, &tmB, (kb % p.b_kdiv) * BK, n0, kb / p.b_kdiv, tc.b0 * p.b_b0, tc.b1 * p.b_b1);
          } else {
#pragma unroll
            for (int j = 0; j < Cfg::B_ROWS / 64; ++j)
              load(sB + j * (BK * 128), &tmB, n0 + 64 * j, kb * BK, tc.b0 * p.b_b0, tc.b1 * p.b_b1, 0);
          }
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ======================= MMA issuer (leader CTA only) =======================
    if (rank == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM * CG, BN, A_MN, B_MN);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = unit; t < total_tiles; t += nunits) {
        TileCoord tc = decode_tile<CG>(p, t);
        const int kb_begin = tc.ks * p.kb_per_split;
        const int kb_end = min(p.num_kb, kb_begin + p.kb_per_split);
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BN);
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          if (lane == 0) {
            const uint32_t sA = smem_base + stage * Cfg::STAGE_BYTES;
            const uint32_t sB = sA + Cfg::A_BYTES;
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              const uint64_t da = A_MN ? make_smem_desc(sA + k * 2048, BK * 128, 1024, 2)
                                       : make_smem_desc(sA + k * 32, 0, 1024, 2);
              const uint64_t db = B_MN ? make_smem_desc(sB + k * 2048, BK * 128, 1024, 2)
                                       : make_smem_desc(sB + k * 32, 0, 1024, 2);
              const uint32_t accum = (kb > kb_begin || k > 0) ? 1u : 0u;
              if constexpr (CG == 2) umma_f16_ss_2sm(tmem_d, da, db, idesc, accum);
              else umma_f16_ss(tmem_d, da, db, idesc, accum);
            }
            if constexpr (CG == 2) {
              umma_commit_2sm(empty_bar(stage), 3);
              if (kb == kb_end - 1) umma_commit_2sm(tfull_bar(acc), 3);
            } else {
              umma_commit(empty_bar(stage));
              if (kb == kb_end - 1) umma_commit(tfull_bar(acc));
            }
          }
          __syncwarp();
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
        if (kb_end <= kb_begin && lane == 0) {
          // empty split (cannot happen with the host-side split computation, kept for safety)
          if constexpr (CG == 2) umma_commit_2sm(tfull_bar(acc), 3);
          else umma_commit(tfull_bar(acc));
        }
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1u;
      }
    }
  } else {
    // ======================= epilogue warps (both CTAs: own 128 rows, all BN columns) =======================
    const int q = warp & 3;           // TMEM lane quadrant this warp may access
    const int half = (warp - 2) >> 2; // which half of the tile's columns
    constexpr int CHUNKS = BN / 64;   // 32-column chunks per warp
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int t = unit; t < total_tiles; t += nunits) {
      TileCoord tc = decode_tile<CG>(p, t);
      const int n0 = tc.n0 * BN + half * (BN / 2);
      const int kb_begin = tc.ks * p.kb_per_split;
      const int kb_end = min(p.num_kb, kb_begin + p.kb_per_split);
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const int row = tc.m0 + rank * BM + q * 32 + lane;
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN + half * (BN / 2));
      const bool row_ok = (row < p.M) && (kb_end > kb_begin);
#pragma unroll 1
      for (int c = 0; c < CHUNKS; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(taddr + c * 32, r);
        tmem_ld_wait();
        if (c == CHUNKS - 1) {
          // all TMEM reads of this accumulator are done: hand the buffer back to the (leader's) MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if constexpr (CG == 2) mbar_arrive_cluster(mapa_shared(tempty_bar(acc), 0));
            else mbar_arrive(tempty_bar(acc));
          }
        }
        if (row_ok) epilogue_chunk(p, tc, row, n0 + c * 32, r);
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1u;
    }
  }

  tc_fence_before();
  if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_cg<CG>(tmem_base, Cfg::TMEM_COLS);
  }
}

}  // namespace dph

// =================================================================================================
// Host side: tensor-map construction and dispatch
// =================================================================================================
#include <mutex>
#include <stdio.h>

namespace dph {

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, []() {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
    if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess) fn = reinterpret_cast<PFN_encodeTiled>(p);
  });
  return fn;
}

void set_error(const char* fmt, ...);

static int make_tmap(CUtensorMap* out, const dph_operand_t& op, int box_rows_kmajor) {
  PFN_encodeTiled enc = get_encode_fn();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled unavailable");
    return DPH_ERR_CUDA;
  }
  cuuint64_t dims[5];
  cuuint64_t strides[4];
  cuuint32_t box[5] = {64, 1, 1, 1, 1};
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  for (int i = 0; i < 5; ++i) dims[i] = op.dims[i] ? op.dims[i] : 1;
  for (int i = 0; i < 4; ++i) {
    strides[i] = op.strides[i] * 2;  // bytes
    if (strides[i] % 16 != 0 || strides[i] == 0) {
      // size-1 dims may carry any stride; give them a legal one
      if (dims[i + 1] == 1) {
        strides[i] = 16;
      } else {
        set_error("operand stride %d (=%llu elements) is not a multiple of 8", i, (unsigned long long)op.strides[i]);
        return DPH_ERR_ARG;
      }
    }
  }
  if ((reinterpret_cast<uintptr_t>(op.ptr) & 15) != 0) {
    set_error("operand pointer not 16-byte aligned");
    return DPH_ERR_ARG;
  }
  box[1] = op.major_mn ? 64 : (cuuint32_t)box_rows_kmajor;
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(op.ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed: %d (dims %llu %llu %llu %llu %llu)", (int)r, (unsigned long long)dims[0],
              (unsigned long long)dims[1], (unsigned long long)dims[2], (unsigned long long)dims[3],
              (unsigned long long)dims[4]);
    return DPH_ERR_CUDA;
  }
  return DPH_OK;
}

// 2-D bf16 tensor map (128B swizzle) over a row-major matrix: used by the fused attention kernels.
int make_tmap_2d(CUtensorMap* out, const void* ptr, uint64_t cols, uint64_t rows, uint64_t row_stride_elems,
                 uint32_t box_cols, uint32_t box_rows) {
  PFN_encodeTiled enc = get_encode_fn();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled unavailable");
    return DPH_ERR_CUDA;
  }
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0 || (row_stride_elems * 2) % 16 != 0) {
    set_error("tensor map: pointer / row stride not 16-byte aligned");
    return DPH_ERR_ARG;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {row_stride_elems * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(2d) failed: %d", (int)r);
    return DPH_ERR_CUDA;
  }
  return DPH_OK;
}

static int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = kNumSMsDefault;
  }
  return n;
}

template <int BN, bool A_MN, bool B_MN, int CG>
static int launch_inst(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid,
                       cudaStream_t stream) {
  using Cfg = GemmCfg<BN, CG>;
  static bool attr_set = false;
  auto kfn = gemm_bf16_kernel<BN, A_MN, B_MN, CG>;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return DPH_ERR_CUDA;
    }
    attr_set = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = (CG > 1) ? 1 : 0;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kfn, ta, tb, p);
  if (e != cudaSuccess) {
    set_error("gemm_bf16_kernel launch: %s", cudaGetErrorString(e));
    return DPH_ERR_CUDA;
  }
  dph_count_launch();
  return dph_check_last("gemm_bf16_kernel");
}

template <int BN, int CG>
static int launch_bn(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid, bool a_mn,
                     bool b_mn, cudaStream_t s) {
  if (!a_mn && !b_mn) return launch_inst<BN, false, false, CG>(ta, tb, p, grid, s);
  if (a_mn && !b_mn) return launch_inst<BN, true, false, CG>(ta, tb, p, grid, s);
  if constexpr (CG == 1 || BN % 128 == 0) {
    if (!a_mn && b_mn) return launch_inst<BN, false, true, CG>(ta, tb, p, grid, s);
    return launch_inst<BN, true, true, CG>(ta, tb, p, grid, s);
  } else {
    set_error("gemm: MN-major B with a CTA pair needs block_n 128 or 256");
    return DPH_ERR_ARG;
  }
}

// A CTA pair cannot split an MN-major B tile of 64 or 192 columns into whole 64-column TMA boxes.
static bool bn_legal(int bn, int cg, bool b_mn) { return !(cg == 2 && b_mn && (bn % 128) != 0); }

static int pick_block_n(int M, int N, int nbatch, int units, int cg, bool b_mn) {
  // minimise (waves * tile cost) -- cost of a tile is proportional to BN
  const int cands[4] = {256, 192, 128, 64};
  int best = 128;
  double best_cost = 1e30;
  const long long mt = (M + BM * cg - 1) / (BM * cg);
  for (int i = 0; i < 4; ++i) {
    int bn = cands[i];
    if (!bn_legal(bn, cg, b_mn)) continue;
    if (bn > 128 && N <= bn / 2) continue;                          // do not waste >50% of a tile on padding
    if (bn == 128 && N <= 64 && bn_legal(64, cg, b_mn)) continue;
    long long tiles = mt * ((N + bn - 1) / bn) * nbatch;
    long long waves = (tiles + units - 1) / units;
    double cost = (double)waves * (bn + 24);  // +24: fixed per-tile overhead in "columns"
    if (cost < best_cost - 1e-9) {
      best_cost = cost;
      best = bn;
    }
  }
  return best;
}

int gemm_bf16_launch(const dph_gemm_desc_t* d, cudaStream_t stream) {
  if (!d || !d->a.ptr || !d->b.ptr || !d->c) {
    set_error("gemm: null pointer");
    return DPH_ERR_ARG;
  }
  if (d->M <= 0 || d->N <= 0 || d->K <= 0 || d->nb0 <= 0 || d->nb1 <= 0) {
    set_error("gemm: bad sizes M=%d N=%d K=%d nb=%d,%d", d->M, d->N, d->K, d->nb0, d->nb1);
    return DPH_ERR_ARG;
  }
  if (d->c_mode == 1 && d->c_dtype != DPH_DTYPE_F32) {
    set_error("gemm: atomic accumulation requires fp32 output");
    return DPH_ERR_ARG;
  }
  const int sms = num_sms();
  const bool a_mn = d->a.major_mn != 0, b_mn = d->b.major_mn != 0;
  // CTA pairs (tcgen05 cta_group::2) whenever there is more than one 128-row tile to pair up.
  int cg = d->cta_group;
  if (cg == 0) cg = (d->M > BM && sms >= 2 && !(b_mn && d->block_n != 0 && d->block_n % 128 != 0)) ? 2 : 1;
  if (cg != 1 && cg != 2) {
    set_error("gemm: cta_group must be 0 (auto), 1 or 2");
    return DPH_ERR_ARG;
  }
  const int units = cg == 2 ? sms / 2 : sms;  // persistent work units (CTAs or CTA pairs)
  const bool auto_split = d->ksplit < 0;
  if (auto_split && d->c_mode != 1) {
    set_error("gemm: automatic split-K requires atomic output mode");
    return DPH_ERR_ARG;
  }
  int bn = d->block_n;
  if (bn == 0) {
    // long reductions into a small output: biggest tile (least operand traffic), parallelism comes from split-K
    int big = d->N > 128 ? 256 : (d->N > 64 ? 128 : 64);
    if (!bn_legal(big, cg, b_mn)) big = 128;
    const long long tiles_big =
        (long long)((d->M + BM * cg - 1) / (BM * cg)) * ((d->N + big - 1) / big) * d->nb0 * d->nb1;
    if (auto_split && tiles_big * 2 <= units) bn = big;
    else bn = pick_block_n(d->M, d->N, d->nb0 * d->nb1, units, cg, b_mn);
  }
  if ((bn != 64 && bn != 128 && bn != 192 && bn != 256) || !bn_legal(bn, cg, b_mn)) {
    set_error("gemm: unsupported block_n %d (cta_group %d)", bn, cg);
    return DPH_ERR_ARG;
  }
  GemmParams p;
  p.M = d->M;
  p.N = d->N;
  p.K = d->K;
  p.nb0 = d->nb0;
  p.nb1 = d->nb1;
  p.mtiles = (d->M + BM * cg - 1) / (BM * cg);
  p.ntiles = (d->N + bn - 1) / bn;
  p.num_kb = (d->K + BK - 1) / BK;
  int ksplit = d->ksplit > 1 ? d->ksplit : 1;
  if (auto_split) {
    const long long tiles = (long long)p.mtiles * p.ntiles * d->nb0 * d->nb1;
    long long ks = units / (tiles > 0 ? tiles : 1);
    if (ks > p.num_kb / 4) ks = p.num_kb / 4;  // at least 4 k-blocks of work per split
    ksplit = ks < 1 ? 1 : (int)ks;
  }
  if (ksplit > 1 && d->c_mode != 1) {
    set_error("gemm: split-K requires atomic output mode");
    return DPH_ERR_ARG;
  }
  if (ksplit > p.num_kb) ksplit = p.num_kb;
  p.kb_per_split = (p.num_kb + ksplit - 1) / ksplit;
  ksplit = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;  // no empty splits
  p.ksplit = ksplit;
  auto kdiv_of = [](const dph_operand_t& op) {
    if (op.major_mn) return 1;
    long long kd = ((long long)op.dims[0] + BK - 1) / BK;
    return (int)(kd < 1 ? 1 : kd);
  };
  p.a_kdiv = kdiv_of(d->a);
  p.b_kdiv = kdiv_of(d->b);
  if (!d->a.major_mn && d->a.dims[2] > 1 && d->a.dims[0] % BK != 0) {
    set_error("gemm: A split-K view needs dims[0] %% 64 == 0");
    return DPH_ERR_ARG;
  }
  if (!d->b.major_mn && d->b.dims[2] > 1 && d->b.dims[0] % BK != 0) {
    set_error("gemm: B split-K view needs dims[0] %% 64 == 0");
    return DPH_ERR_ARG;
  }
  p.a_b0 = d->a.b0_mul;
  p.a_b1 = d->a.b1_mul;
  p.b_b0 = d->b.b0_mul;
  p.b_b1 = d->b.b1_mul;
  p.C = d->c;
  p.C2 = d->c2;
  p.ldc = d->ldc;
  p.c_sb0 = d->c_sb0;
  p.c_sb1 = d->c_sb1;
  p.c_dtype = d->c_dtype;
  p.c_mode = d->c_mode;
  // A single split whose batches do not alias one output needs no atomics: read-add-store with vector accesses.
  if (p.c_mode == 1 && ksplit == 1 && (d->nb0 == 1 || d->c_sb0 != 0) && (d->nb1 == 1 || d->c_sb1 != 0)) p.c_mode = 2;
  p.alpha = d->alpha;
  p.bias = d->bias;
  p.act = d->act;
  p.colmask = d->colmask;
  p.scalar_mask = d->scalar_mask;
  p.residual = d->residual;
  p.ldr = d->ldr;
  p.r_sb0 = d->r_sb0;
  p.r_sb1 = d->r_sb1;
  p.r_dtype = d->r_dtype;
  p.drop_p = d->drop_p;
  p.seed = d->seed;
  p.vec_sb0 = d->vec_sb0;
  p.scalar_sb0 = d->scalar_sb0;
  // vector access legality
  const int celt = d->c_dtype == 0 ? 8 : 4;
  bool vec = (d->ldc % celt == 0) && (d->c_sb0 % celt == 0) && (d->c_sb1 % celt == 0) &&
             ((reinterpret_cast<uintptr_t>(d->c) & 15) == 0);
  if (d->c2) vec = vec && ((reinterpret_cast<uintptr_t>(d->c2) & 15) == 0);
  if (d->residual) {
    const int relt = d->r_dtype == 0 ? 8 : 4;
    vec = vec && (d->ldr % relt == 0) && (d->r_sb0 % relt == 0) && (d->r_sb1 % relt == 0) &&
          ((reinterpret_cast<uintptr_t>(d->residual) & 15) == 0);
  }
  if (d->bias) vec = vec && ((reinterpret_cast<uintptr_t>(d->bias) & 15) == 0) && (d->vec_sb0 % 4 == 0);
  if (d->colmask) vec = vec && ((reinterpret_cast<uintptr_t>(d->colmask) & 15) == 0) && (d->vec_sb0 % 4 == 0);
  p.vec_ok = vec ? 1 : 0;

  CUtensorMap ta, tb;
  int rc = make_tmap(&ta, d->a, BM);
  if (rc != DPH_OK) return rc;
  rc = make_tmap(&tb, d->b, bn / cg);
  if (rc != DPH_OK) return rc;

  const long long total = (long long)p.nb1 * p.nb0 * p.mtiles * p.ntiles * p.ksplit;
  const int grid = (int)(total < units ? total : units) * cg;
  if (cg == 2) {
    switch (bn) {
      case 64: return launch_bn<64, 2>(ta, tb, p, grid, a_mn, b_mn, stream);
      case 128: return launch_bn<128, 2>(ta, tb, p, grid, a_mn, b_mn, stream);
      case 192: return launch_bn<192, 2>(ta, tb, p, grid, a_mn, b_mn, stream);
      default: return launch_bn<256, 2>(ta, tb, p, grid, a_mn, b_mn, stream);
    }
  }
  switch (bn) {
    case 64: return launch_bn<64, 1>(ta, tb, p, grid, a_mn, b_mn, stream);
    case 128: return launch_bn<128, 1>(ta, tb, p, grid, a_mn, b_mn, stream);
    case 192: return launch_bn<192, 1>(ta, tb, p, grid, a_mn, b_mn, stream);
    default: return launch_bn<256, 1>(ta, tb, p, grid, a_mn, b_mn, stream);
  }
}

}  // namespace dph

extern "C" int dph_gemm_bf16(const dph_gemm_desc_t* desc, void* stream) {
  return dph::gemm_bf16_launch(desc, reinterpret_cast<cudaStream_t>(stream));
}
