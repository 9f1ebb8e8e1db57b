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
//
// The fused epilogue is compiled per *feature class* (template parameter F, a mask of EpiFeature bits): the epilogue
// warps are latency-bound, and a kernel that carries the unrolled code of every feature (GELU, dropout hash, ...)
// runs its epilogue ~2x slower than one that only contains what the launch uses (instruction-cache misses).  The
// host picks the smallest class that covers the descriptor (gemm_tcgen05.cu); each class is instantiated in its own
// translation unit (gemm_inst_*.cu).
#pragma once
#include "common.cuh"
#include "gemm.h"

namespace dph {

constexpr int BM = 128;
constexpr int BK = 64;
// Eight epilogue warps: measured best for the forward classes. Sixteen (one per 64-column slice of a 256-wide tile)
// cap the kernel at 96 registers per thread and lose more to spills and barrier traffic than the extra latency hiding
// returns.  The fused-backward class (F_ACTBWD = 512) is the exception: its epilogue waits on the pre-activation
// loads (ncu: issue slots 32 % busy, long-scoreboard stalls on z, 16 us per tile against a 3.3 us main loop), so it
// runs sixteen warps over 16-column half chunks that fit the 96-register budget.
__host__ __device__ constexpr int epi_warps(int bn, int f) { return ((f & 512) != 0 && bn % 128 == 0) ? 16 : 8; }
__host__ __device__ constexpr int gemm_threads(int bn, int f) { return 64 + 32 * epi_warps(bn, f); }

enum EpiFeature : int {
  F_BIAS = 1,      // + bias[n]
  F_GELU = 2,      // erf-GELU
  F_COLMASK = 4,   // * colmask[n]      (HardConcrete channel / FFN masks)
  F_SMASK = 8,     // * scalar_mask[b0] (head / layer gates)
  F_DROP = 16,     // counter-based dropout
  F_RES = 32,      // + residual[m, n]
  F_C2 = 64,       // second output: pre-activation copy
  F_ACCUM = 128,   // c_mode 1 (atomic) / 2 (read-add-store)
  F_ALPHA = 256,   // * alpha
  F_ACTBWD = 512,  // fused BACKWARD of an activation epilogue (see epilogue_chunk_bwd)
};
constexpr int EPI_LIN = F_BIAS | F_COLMASK | F_SMASK | F_RES;
constexpr int EPI_ACT = F_BIAS | F_GELU | F_COLMASK | F_C2 | F_RES;
constexpr int EPI_DROP = F_BIAS | F_COLMASK | F_SMASK | F_DROP | F_RES;
constexpr int EPI_ACC = F_ALPHA | F_ACCUM;
constexpr int EPI_ALL = 511;
constexpr int EPI_BWD = F_ACTBWD | F_COLMASK | F_DROP;

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

// Optional per-CTA event timeline (ns, %globaltimer) for performance debugging: 64 slots per CTA (gemm_tcgen05.cu).
extern __device__ unsigned long long* g_gemm_timeline;
DPH_DEVICE void stamp(unsigned long long* tl, int slot) {
  if (tl != nullptr && slot < 64) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    tl[slot] = t;
  }
}

// Ragged / unaligned chunks (edge tiles, odd leading dimensions): reads its 32 accumulator columns from TMEM itself
// and walks them one by one with every feature checked at run time.  Warp-collective (all lanes must call it).
// Out of line, one copy per translation unit: its dynamically indexed array must not pull the fast path's registers
// onto the stack, and being in the kernel's own module keeps it inside the kernel's register budget.
static __device__ __noinline__ void epilogue_chunk_generic(const GemmParams& p, int b0, int b1, int row, bool row_ok, int col0,
                                                    uint32_t taddr) {
  uint32_t r[32];
  tmem_ld_32x32(taddr, r);
  tmem_ld_wait();
  const int nvalid = min(32, p.N - col0);
  if (!row_ok || nvalid <= 0) return;
  float acc[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) acc[j] = __uint_as_float(r[j]);  // lives in local memory: indexed dynamically below
  const long long coff = (long long)b0 * p.c_sb0 + (long long)b1 * p.c_sb1 + (long long)row * p.ldc + col0;
  const long long roff = (long long)b0 * p.r_sb0 + (long long)b1 * p.r_sb1 + (long long)row * p.ldr + col0;
  const uint32_t thresh = (uint32_t)(p.drop_p * 4294967296.0);
  const float inv_keep = 1.0f / (1.0f - p.drop_p);
  const unsigned long long dbase =
      (((unsigned long long)b1 * p.nb0 + b0) * (unsigned long long)p.M + row) * (unsigned long long)p.N + col0;
  const float smask = p.scalar_mask != nullptr ? __ldg(p.scalar_mask + b0 * p.scalar_sb0) : 1.0f;
#pragma unroll 1
  for (int j = 0; j < nvalid; ++j) {
    float v = acc[j] * p.alpha;
    if (p.bias != nullptr) v += __ldg(p.bias + b0 * p.vec_sb0 + col0 + j);
    if (p.C2 != nullptr) {
      if (p.c_dtype == 0) reinterpret_cast<bf16*>(p.C2)[coff + j] = __float2bfloat16_rn(v);
      else reinterpret_cast<float*>(p.C2)[coff + j] = v;
    }
    if (p.act == 1) v = gelu_erf_fast(v);
    if (p.colmask != nullptr) v *= __ldg(p.colmask + b0 * p.vec_sb0 + col0 + j);
    v *= smask;
    if (p.drop_p > 0.0f) v *= dropout_scale(p.seed, dbase + j, thresh, inv_keep);
    if (p.residual != nullptr) {
      v += p.r_dtype == 0 ? __bfloat162float(reinterpret_cast<const bf16*>(p.residual)[roff + j])
                          : reinterpret_cast<const float*>(p.residual)[roff + j];
    }
    if (p.c_dtype == 0) {
      bf16* o = reinterpret_cast<bf16*>(p.C) + coff + j;
      if (p.c_mode == 2) v += __bfloat162float(*o);
      *o = __float2bfloat16_rn(v);
    } else {
      float* o = reinterpret_cast<float*>(p.C) + coff + j;
      if (p.c_mode == 1) atomicAdd(o, v);
      else *o = (p.c_mode == 2 ? *o : 0.0f) + v;
    }
  }
}


// 32 consecutive outputs of one row as full 32-byte sectors (256-bit stores)
DPH_DEVICE void store_bf16x32(bf16* o, const float (&v)[32]) {
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    uint32_t u[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) u[q] = pack_bf16x2(v[16 * j + 2 * q], v[16 * j + 2 * q + 1]);
    stg_256(o + 16 * j, u);
  }
}
DPH_DEVICE void store_f32x32(float* o, const float (&v)[32]) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint32_t u[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) u[q] = __float_as_uint(v[8 * j + q]);
    stg_256(o + 8 * j, u);
  }
}
DPH_DEVICE void add_bf16x16(float* v, const uint32_t (&u)[8]) {
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u[q]));
    v[2 * q] += t.x;
    v[2 * q + 1] += t.y;
  }
}

// Fast path of one 32-column chunk: all 32 columns valid and every pointer / stride 32-byte aligned (p.vec_ok).
// Every global load of the chunk (bias, column mask, residual: read-only in this kernel) is issued up front as an
// independent wide load so that the latencies overlap; outputs leave as 256-bit stores (whole L2 sectors).
template <int F>
DPH_DEVICE void epilogue_chunk(const GemmParams& p, const TileCoord& tc, int row, int col0, uint32_t (&r)[32]) {
  const long long coff = (long long)tc.b0 * p.c_sb0 + (long long)tc.b1 * p.c_sb1 + (long long)row * p.ldc + col0;
  float v[32];
  if constexpr ((F & F_ALPHA) != 0) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * p.alpha;
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
  }

  uint32_t res[2][8];
  bool has_res16 = false;
  if constexpr ((F & F_RES) != 0) {
    has_res16 = (p.residual != nullptr) && (p.r_dtype == 0);
    if (has_res16) {
      const long long roff = (long long)tc.b0 * p.r_sb0 + (long long)tc.b1 * p.r_sb1 + (long long)row * p.ldr + col0;
      const bf16* rp = reinterpret_cast<const bf16*>(p.residual) + roff;
      ldg_nc_256(rp, res[0]);
      ldg_nc_256(rp + 16, res[1]);
    }
  }
  if constexpr ((F & F_BIAS) != 0) {
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
  }
  if constexpr ((F & F_C2) != 0) {
    if (p.C2 != nullptr) {
      if (p.c_dtype == 0) store_bf16x32(reinterpret_cast<bf16*>(p.C2) + coff, v);
      else store_f32x32(reinterpret_cast<float*>(p.C2) + coff, v);
    }
  }
  if constexpr ((F & F_GELU) != 0) {
    if (p.act == 1) {
      if (p.c_dtype == 0) {
        // bf16 output: tanh-form GELU on packed fp32x2 lanes (one MUFU and ~4.5 issue slots per element, error below
        // bf16 resolution); an fp32 output keeps the erf form
#pragma unroll
        for (int j = 0; j < 32; j += 2) upk2(gelu_tanh2(pk2(v[j], v[j + 1])), v[j], v[j + 1]);
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = gelu_erf_fast(v[j]);
      }
    }
  }
  if constexpr ((F & F_COLMASK) != 0) {
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
  }
  if constexpr ((F & F_SMASK) != 0) {
    if (p.scalar_mask != nullptr) {
      const float s = __ldg(p.scalar_mask + tc.b0 * p.scalar_sb0);
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] *= s;
    }
  }
  if constexpr ((F & F_DROP) != 0) {
    if (p.drop_p > 0.0f) {
      const uint32_t thresh = (uint32_t)(p.drop_p * 4294967296.0);
      const float inv_keep = 1.0f / (1.0f - p.drop_p);
      const unsigned long long base =
          (((unsigned long long)tc.b1 * p.nb0 + tc.b0) * (unsigned long long)p.M + row) * (unsigned long long)p.N +
          col0;
      dropout_apply_vec<32>(p.seed, base, thresh, inv_keep, v);
    }
  }
  if constexpr ((F & F_RES) != 0) {
    if (has_res16) {
      add_bf16x16(v, res[0]);
      add_bf16x16(v + 16, res[1]);
    } else if (p.residual != nullptr) {
      const long long roff = (long long)tc.b0 * p.r_sb0 + (long long)tc.b1 * p.r_sb1 + (long long)row * p.ldr + col0;
      const float* rp = reinterpret_cast<const float*>(p.residual) + roff;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint32_t u[8];
        ldg_nc_256(rp + 8 * j, u);
#pragma unroll
        for (int q = 0; q < 8; ++q) v[8 * j + q] += __uint_as_float(u[q]);
      }
    }
  }
  // ---- store ----
  if (p.c_dtype == 0) {
    bf16* o = reinterpret_cast<bf16*>(p.C) + coff;
    if constexpr ((F & F_ACCUM) != 0) {
      if (p.c_mode == 2) {
        const uint4* o4 = reinterpret_cast<const uint4*>(o);
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const uint4 a = o4[2 * j], b = o4[2 * j + 1];
          const uint32_t u[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
          add_bf16x16(v + 16 * j, u);
        }
      }
    }
    store_bf16x32(o, v);
  } else {
    float* o = reinterpret_cast<float*>(p.C) + coff;
    if constexpr ((F & F_ACCUM) != 0) {
      if (p.c_mode == 1) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) red_add_v4(o + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
        return;
      }
      if (p.c_mode == 2) {
        const float4* o4 = reinterpret_cast<const float4*>(o);
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
    }
    store_f32x32(o, v);
  }
}

// Column sums over the 32 lanes of a warp of 32 per-lane values: a reduce-scatter butterfly (16 + 8 + 4 + 2 + 1
// exchanges instead of 32 full reductions).  Returns, in lane l, the sum over all lanes of v[l].  Destroys v.
DPH_DEVICE float warp_colsum32(float (&v)[32], int lane) {
#pragma unroll
  for (int w = 16; w >= 1; w >>= 1) {
    const bool hi = (lane & w) != 0;
#pragma unroll
    for (int j = 0; j < w; ++j) {
      const float send = hi ? v[j] : v[j + w];
      const float keep = hi ? v[j + w] : v[j];
      v[j] = keep + __shfl_xor_sync(0xffffffffu, send, w);
    }
  }
  return v[0];
}

// 16 per-lane values: after the butterfly lanes l and l + 16 both hold the sum over all lanes of v[l % 16]. Destroys v.
DPH_DEVICE float warp_colsum16(float (&v)[16], int lane) {
#pragma unroll
  for (int w = 8; w >= 1; w >>= 1) {
    const bool hi = (lane & w) != 0;
#pragma unroll
    for (int j = 0; j < w; ++j) {
      const float send = hi ? v[j] : v[j + w];
      const float keep = hi ? v[j + w] : v[j];
      v[j] = keep + __shfl_xor_sync(0xffffffffu, send, w);
    }
  }
  return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 16);
}

// 32 lanes x 16 columns of 32-bit from tensor memory
DPH_DEVICE void tmem_ld_32x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

// Fused backward of the activation epilogue  o = dropout(colmask[n] * act(z[m, n]))  (FFN hidden, conv blocks:
// components.py:110-114, 733-737): the accumulator of this GEMM is d loss / d o (it is a dgrad product), and instead
// of writing it out for a separate elementwise pass the epilogue
//   * reads the saved pre-activation tile z (bf16, its own strides),
//   * writes  c = acc * dropout * colmask[n] * act'(z)   (the gradient wrt the pre-activation), and
//   * adds the column sums  dbias[n] += sum_m c,  dmask[n] += sum_m acc * dropout * act(z)  with one coalesced RED per
//     warp, chunk and quantity (the 32 lanes of a warp hold 32 rows: reduce-scatter, then lane l owns column l).
// Warp-collective: every lane calls it; rows beyond M contribute zeros and store nothing.  Fast path only (32 valid
// columns, 32-byte aligned rows) — the host refuses other shapes.
// 32 bf16 of the saved pre-activation tile for (row, col0 .. col0 + 31): issued one chunk ahead of their use (and,
// for the first chunk of a tile, before the wait for the accumulator) so that the global-load latency hides behind
// the previous chunk's arithmetic / the main loop
DPH_DEVICE void bwd_fetch_z(const GemmParams& p, const TileCoord& tc, int row, bool row_ok, int col0, uint32_t (&zr)[16]) {
  if (row_ok) {
    const long long zoff = (long long)tc.b0 * p.z_sb0 + (long long)tc.b1 * p.z_sb1 + (long long)row * p.ldz + col0;
    const bf16* zp = reinterpret_cast<const bf16*>(p.Z) + zoff;
    uint32_t a[8], b[8];
    ldg_nc_256(zp, a);
    ldg_nc_256(zp + 16, b);
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      zr[q] = a[q];
      zr[8 + q] = b[q];
    }
  } else {
#pragma unroll
    for (int q = 0; q < 16; ++q) zr[q] = 0u;
  }
}

template <int F>
DPH_DEVICE void epilogue_chunk_bwd(const GemmParams& p, const TileCoord& tc, int row, bool row_ok, int col0,
                                   uint32_t (&r)[32], const uint32_t (&zr)[16]) {
  const int lane = threadIdx.x & 31;
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
  if (__any_sync(0xffffffffu, !row_ok)) {  // warp-uniform: only the last M tile has rows beyond M
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = row_ok ? v[j] : 0.f;
  }
  if constexpr ((F & F_DROP) != 0) {
    if (p.drop_p > 0.0f) {
      const uint32_t thresh = (uint32_t)(p.drop_p * 4294967296.0);
      const float inv_keep = 1.0f / (1.0f - p.drop_p);
      const unsigned long long base =
          (((unsigned long long)tc.b1 * p.nb0 + tc.b0) * (unsigned long long)p.M + row) * (unsigned long long)p.N +
          col0;
      dropout_apply_vec<32>(p.seed, base, thresh, inv_keep, v);
    }
  }
  float dm[32];  // acc * dropout * act(z): the column-mask gradient's summand
#pragma unroll
  for (int j = 0; j < 32; j += 2) {
    const float2 zz = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&zr[j >> 1]));
    const f32x2 z2 = pk2(zz.x, zz.y), v2 = pk2(v[j], v[j + 1]);
    f32x2 g2 = z2, dg2 = bc2(1.f);
    if (p.act == 1) gelu_and_grad_tanh2(z2, g2, dg2);
    upk2(mul2(v2, g2), dm[j], dm[j + 1]);
    upk2(mul2(v2, dg2), v[j], v[j + 1]);
  }
  if constexpr ((F & F_COLMASK) != 0) {
    if (p.colmask != nullptr) {
      const float4* mp = reinterpret_cast<const float4*>(p.colmask + tc.b0 * p.vec_sb0 + col0);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 m = __ldg(mp + j);
        upk2(mul2(pk2(v[4 * j], v[4 * j + 1]), pk2(m.x, m.y)), v[4 * j], v[4 * j + 1]);
        upk2(mul2(pk2(v[4 * j + 2], v[4 * j + 3]), pk2(m.z, m.w)), v[4 * j + 2], v[4 * j + 3]);
      }
    }
  }
  if (row_ok) {
    const long long coff = (long long)tc.b0 * p.c_sb0 + (long long)tc.b1 * p.c_sb1 + (long long)row * p.ldc + col0;
    store_bf16x32(reinterpret_cast<bf16*>(p.C) + coff, v);
  }
  if (p.dbias != nullptr) {
    const float sum = warp_colsum32(v, lane);
    atomicAdd(p.dbias + tc.b0 * p.vec_sb0 + col0 + lane, sum);
  }
  if (p.dmask != nullptr) {
    const float sum = warp_colsum32(dm, lane);
    atomicAdd(p.dmask + tc.b0 * p.vec_sb0 + col0 + lane, sum);
  }
}

// ---- the same on a 16-column half chunk (sixteen-warp variant of the class: 96 registers per thread) ----
DPH_DEVICE void bwd_fetch_z16(const GemmParams& p, const TileCoord& tc, int row, bool row_ok, int col0, uint32_t (&zr)[8]) {
  if (row_ok) {
    const long long zoff = (long long)tc.b0 * p.z_sb0 + (long long)tc.b1 * p.z_sb1 + (long long)row * p.ldz + col0;
    ldg_nc_256(reinterpret_cast<const bf16*>(p.Z) + zoff, zr);
  } else {
#pragma unroll
    for (int q = 0; q < 8; ++q) zr[q] = 0u;
  }
}

template <int F>
DPH_DEVICE void epilogue_half_bwd(const GemmParams& p, const TileCoord& tc, int row, bool row_ok, int col0,
                                  uint32_t (&r)[16], const uint32_t (&zr)[8]) {
  const int lane = threadIdx.x & 31;
  float v[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) v[j] = row_ok ? __uint_as_float(r[j]) : 0.f;
  if constexpr ((F & F_DROP) != 0) {
    if (p.drop_p > 0.0f) {
      const uint32_t thresh = (uint32_t)(p.drop_p * 4294967296.0);
      const float inv_keep = 1.0f / (1.0f - p.drop_p);
      const unsigned long long base =
          (((unsigned long long)tc.b1 * p.nb0 + tc.b0) * (unsigned long long)p.M + row) * (unsigned long long)p.N +
          col0;
      dropout_apply_vec<16>(p.seed, base, thresh, inv_keep, v);
    }
  }
  float dm[16];
#pragma unroll
  for (int j = 0; j < 16; j += 2) {
    const float2 zz = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&zr[j >> 1]));
    const f32x2 z2 = pk2(zz.x, zz.y), v2 = pk2(v[j], v[j + 1]);
    f32x2 g2 = z2, dg2 = bc2(1.f);
    if (p.act == 1) gelu_and_grad_tanh2(z2, g2, dg2);
    upk2(mul2(v2, g2), dm[j], dm[j + 1]);
    upk2(mul2(v2, dg2), v[j], v[j + 1]);
  }
  if constexpr ((F & F_COLMASK) != 0) {
    if (p.colmask != nullptr) {
      const float4* mp = reinterpret_cast<const float4*>(p.colmask + tc.b0 * p.vec_sb0 + col0);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float4 m = __ldg(mp + j);
        upk2(mul2(pk2(v[4 * j], v[4 * j + 1]), pk2(m.x, m.y)), v[4 * j], v[4 * j + 1]);
        upk2(mul2(pk2(v[4 * j + 2], v[4 * j + 3]), pk2(m.z, m.w)), v[4 * j + 2], v[4 * j + 3]);
      }
    }
  }
  if (row_ok) {
    const long long coff = (long long)tc.b0 * p.c_sb0 + (long long)tc.b1 * p.c_sb1 + (long long)row * p.ldc + col0;
    uint32_t u[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) u[q] = pack_bf16x2(v[2 * q], v[2 * q + 1]);
    stg_256(reinterpret_cast<bf16*>(p.C) + coff, u);
  }
  if (p.dbias != nullptr) {
    const float sum = warp_colsum16(v, lane);
    if (lane < 16) atomicAdd(p.dbias + tc.b0 * p.vec_sb0 + col0 + lane, sum);
  }
  if (p.dmask != nullptr) {
    const float sum = warp_colsum16(dm, lane);
    if (lane < 16) atomicAdd(p.dmask + tc.b0 * p.vec_sb0 + col0 + lane, sum);
  }
}

template <int BN, bool A_MN, bool B_MN, int CG, int F>
__global__ void __launch_bounds__(gemm_threads(BN, F), 1)
    gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                     const __grid_constant__ GemmParams p) {
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
  unsigned long long* tl = g_gemm_timeline;
  if (tl != nullptr) tl += 64 * blockIdx.x;
  if (threadIdx.x == 0) stamp(tl, 0);

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
        mbar_init(tempty_bar(a), epi_warps(BN, F) * CG);  // leader's barrier collects the epilogue warps of both CTAs
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
  // everything above touched only shared / tensor memory: it may overlap the tail of the previous kernel (PDL)
  pdl_wait();
#ifdef DPH_PDL_EARLY
  pdl_trigger();
#endif
  if (threadIdx.x == 0) stamp(tl, 1);

  if (warp == 0) {
    // ======================= TMA producer (both CTAs of a pair) =======================
    // elect.sync instead of `lane == 0`: ptxas then knows that one lane is active and keeps coordinates / descriptors
    // in uniform registers without per-operand waterfall loops
    if (elect_one()) {
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
          if (t == unit && kb == kb_begin) stamp(tl, 2);
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
            for (int j = 0; j < Cfg::B_ROWS / 64; ++j) {
              const int nn = n0 + 64 * j;
              // b_nsplit > 0: the N index is (tap, channel) = (nn / b_nsplit, nn % b_nsplit) and the tap is the 4th
              // tensor-map dimension (an overlapping, row-shifted view of the same activation: grouped-conv weight
              // gradient with the taps folded into N)
              if (p.b_nsplit > 0) load(sB + j * (BK * 128), &tmB, nn % p.b_nsplit, kb * BK, tc.b0 * p.b_b0, nn / p.b_nsplit, 0);
              else load(sB + j * (BK * 128), &tmB, nn, kb * BK, tc.b0 * p.b_b0, tc.b1 * p.b_b1, 0);
            }
          }
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
      stamp(tl, 3);
    }
    // all operand loads of this CTA are in flight: the next kernel may start its prologue on SMs that drain
    pdl_trigger();
  } else if (warp == 1) {
    // ======================= MMA issuer (leader CTA only) =======================
    if (rank == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM * CG, BN, A_MN, B_MN);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      int ti = 0;
      for (int t = unit; t < total_tiles; t += nunits, ++ti) {
        TileCoord tc = decode_tile<CG>(p, t);
        const int kb_begin = tc.ks * p.kb_per_split;
        const int kb_end = min(p.num_kb, kb_begin + p.kb_per_split);
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BN);
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          if (lane == 0 && kb == kb_begin) stamp(tl, 8 + 2 * ti);
          if (lane == 0 && kb == kb_end - 1) stamp(tl, 9 + 2 * ti);
          if (elect_one()) {
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
    constexpr int NSL = epi_warps(BN, F) / 4;   // column slices of the tile (one per group of four warps)
    constexpr int SLW = BN / NSL;               // columns per slice
    const int half = (warp - 2) >> 2; // which slice of the tile's columns
    constexpr int CHUNKS = SLW / 32;  // 32-column chunks per warp
    int acc = 0;
    uint32_t acc_phase = 0;
    int ti = 0;
    for (int t = unit; t < total_tiles; t += nunits, ++ti) {
      TileCoord tc = decode_tile<CG>(p, t);
      const int n0 = tc.n0 * BN + half * SLW;
      const int kb_begin = tc.ks * p.kb_per_split;
      const int kb_end = min(p.num_kb, kb_begin + p.kb_per_split);
      const int row = tc.m0 + rank * BM + q * 32 + lane;
      const bool row_ok = (row < p.M) && (kb_end > kb_begin);
      constexpr bool W16 = (F & F_ACTBWD) != 0 && NSL == 4;   // sixteen-warp fused-backward variant
      [[maybe_unused]] uint32_t zr[W16 ? 1 : 16], zn[W16 ? 1 : 16];
      [[maybe_unused]] uint32_t hz[8], hn[8];
      // (the host admits the fused backward only with 32-byte aligned rows and N % 32 == 0: a 16-column half chunk is
      // either inside N or outside it, never split)
      if constexpr (W16) {
        bwd_fetch_z16(p, tc, row, row_ok && n0 + 16 <= p.N, n0, hn);   // does not depend on the MMAs
      } else if constexpr ((F & F_ACTBWD) != 0) {
        bwd_fetch_z(p, tc, row, row_ok, n0, zn);
      }
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      if (warp == 2 && lane == 0) stamp(tl, 32 + 2 * ti);
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN + half * SLW);
      if constexpr (W16) {
        {
          constexpr int HALVES = SLW / 16;
#pragma unroll 1
          for (int h = 0; h < HALVES; ++h) {
            const int col0 = n0 + h * 16;
            uint32_t r[16];
            tmem_ld_32x16(taddr + h * 16, r);
#pragma unroll
            for (int q2 = 0; q2 < 8; ++q2) hz[q2] = hn[q2];
            if (h + 1 < HALVES) bwd_fetch_z16(p, tc, row, row_ok && col0 + 32 <= p.N, col0 + 16, hn);
            tmem_ld_wait();
            if (h == HALVES - 1) {   // all TMEM reads of this accumulator are done: hand it back
              tc_fence_before();
              __syncwarp();
              if (lane == 0) {
                if constexpr (CG == 2) mbar_arrive_cluster(mapa_shared(tempty_bar(acc), 0));
                else mbar_arrive(tempty_bar(acc));
              }
            }
            if (col0 + 16 <= p.N) epilogue_half_bwd<F>(p, tc, row, row_ok, col0, r, hz);   // warp-collective, uniform
          }
          if (warp == 2 && lane == 0) stamp(tl, 33 + 2 * ti);
          acc ^= 1;
          if (acc == 0) acc_phase ^= 1u;
          continue;
        }
      }
      if constexpr (!W16) {
#pragma unroll 1
      for (int c = 0; c < CHUNKS; ++c) {
        const int col0 = n0 + c * 32;
        const bool fast = p.vec_ok && (col0 + 32 <= p.N);  // warp-uniform
        // once all TMEM reads of this accumulator are done, hand the buffer back to the (leader's) MMA warp
        auto release_acc = [&]() {
          if (c == CHUNKS - 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
              if constexpr (CG == 2) mbar_arrive_cluster(mapa_shared(tempty_bar(acc), 0));
              else mbar_arrive(tempty_bar(acc));
            }
          }
        };
        if (fast) {
          uint32_t r[32];
          tmem_ld_32x32(taddr + c * 32, r);
          if constexpr ((F & F_ACTBWD) != 0) {
#pragma unroll
            for (int q2 = 0; q2 < 16; ++q2) zr[q2] = zn[q2];
            if (c + 1 < CHUNKS) bwd_fetch_z(p, tc, row, row_ok, col0 + 32, zn);
          }
          tmem_ld_wait();
          release_acc();
          if constexpr ((F & F_ACTBWD) != 0) {
            epilogue_chunk_bwd<F>(p, tc, row, row_ok, col0, r, zr);   // warp-collective (column sums)
          } else {
            if (row_ok) epilogue_chunk<F>(p, tc, row, col0, r);
          }
        } else {
          // kept in its own branch: the accumulator registers of the fast path must not be live across this call
          epilogue_chunk_generic(p, tc.b0, tc.b1, row, row_ok, col0, taddr + c * 32);
          release_acc();
        }
      }
      }
      if (warp == 2 && lane == 0) stamp(tl, 33 + 2 * ti);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1u;
    }
  }

  if (threadIdx.x == 0) stamp(tl, 4);
  tc_fence_before();
  if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();
  if (threadIdx.x == 0) stamp(tl, 5);
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_cg<CG>(tmem_base, Cfg::TMEM_COLS);
  }
}


void set_error(const char* fmt, ...);

template <int BN, bool A_MN, bool B_MN, int CG, int F>
static int launch_inst(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid,
                       cudaStream_t stream) {
  using Cfg = GemmCfg<BN, CG>;
  static bool attr_set = false;
  auto kfn = gemm_bf16_kernel<BN, A_MN, B_MN, CG, F>;
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
  cfg.blockDim = dim3(gemm_threads(BN, F));
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (CG > 1) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = CG;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  if (dph_pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kfn, ta, tb, p);
  if (e != cudaSuccess) {
    set_error("gemm_bf16_kernel launch: %s", cudaGetErrorString(e));
    return DPH_ERR_CUDA;
  }
  dph_count_launch();
  return dph_check_last("gemm_bf16_kernel");
}

// A CTA pair cannot split an MN-major B tile of 64 or 192 columns into whole 64-column TMA boxes.
inline bool bn_legal(int bn, int cg, bool b_mn) { return !(cg == 2 && b_mn && (bn % 128) != 0); }

// ALL_MAJORS = false: only K-major x K-major operands are instantiated (forward-only epilogue classes).
template <int BN, int CG, int F, bool ALL_MAJORS>
static int launch_bn(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid, bool a_mn,
                     bool b_mn, cudaStream_t s) {
  if (!a_mn && !b_mn) return launch_inst<BN, false, false, CG, F>(ta, tb, p, grid, s);
  if constexpr (ALL_MAJORS) {
    if (a_mn && !b_mn) return launch_inst<BN, true, false, CG, F>(ta, tb, p, grid, s);
    if constexpr (CG == 1 || BN % 128 == 0) {
      if (!a_mn && b_mn) return launch_inst<BN, false, true, CG, F>(ta, tb, p, grid, s);
      return launch_inst<BN, true, true, CG, F>(ta, tb, p, grid, s);
    }
  }
  set_error("gemm: no kernel for this operand-major / tile combination in epilogue class %d", F);
  return DPH_ERR_UNSUPPORTED;
}

template <int F, bool ALL_MAJORS>
static int launch_class(int bn, int cg, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid,
                        bool a_mn, bool b_mn, cudaStream_t s) {
  if (cg == 2) {
    switch (bn) {
      case 64: return launch_bn<64, 2, F, ALL_MAJORS>(ta, tb, p, grid, a_mn, b_mn, s);
      case 128: return launch_bn<128, 2, F, ALL_MAJORS>(ta, tb, p, grid, a_mn, b_mn, s);
      case 192: return launch_bn<192, 2, F, ALL_MAJORS>(ta, tb, p, grid, a_mn, b_mn, s);
      default: return launch_bn<256, 2, F, ALL_MAJORS>(ta, tb, p, grid, a_mn, b_mn, s);
    }
  }
  switch (bn) {
    case 64: return launch_bn<64, 1, F, ALL_MAJORS>(ta, tb, p, grid, a_mn, b_mn, s);
    case 128: return launch_bn<128, 1, F, ALL_MAJORS>(ta, tb, p, grid, a_mn, b_mn, s);
    case 192: return launch_bn<192, 1, F, ALL_MAJORS>(ta, tb, p, grid, a_mn, b_mn, s);
    default: return launch_bn<256, 1, F, ALL_MAJORS>(ta, tb, p, grid, a_mn, b_mn, s);
  }
}

// One launcher per epilogue class, each defined in its own translation unit (gemm_inst_*.cu).
int gemm_launch_lin(int bn, int cg, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid,
                    bool a_mn, bool b_mn, cudaStream_t s);
int gemm_launch_act(int bn, int cg, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid,
                    bool a_mn, bool b_mn, cudaStream_t s);
int gemm_launch_drop(int bn, int cg, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid,
                     bool a_mn, bool b_mn, cudaStream_t s);
int gemm_launch_acc(int bn, int cg, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid,
                    bool a_mn, bool b_mn, cudaStream_t s);
int gemm_launch_all(int bn, int cg, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid,
                    bool a_mn, bool b_mn, cudaStream_t s);
int gemm_launch_bwd(int bn, int cg, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid,
                    bool a_mn, bool b_mn, cudaStream_t s);

}  // namespace dph
