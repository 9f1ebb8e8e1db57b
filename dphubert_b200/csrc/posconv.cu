// Grouped positional convolution (ConvolutionalPositionalEmbedding, wav2vec2/components.py:319-333: Conv1d(D, D, k=128,
// padding=64, groups=16) -> drop last frame -> GELU, plus the residual add of Transformer._preprocess :885-892) as a
// tap-stationary implicit GEMM on the 5th-generation tensor cores, and the same kernel with the tap-flipped weights
// for the input gradient.
//
//   out[b, t, g Cg + n] = epi( sum_{tap < K} sum_{ci < Cgp} xp[row0 + b rows_b + t + tap][g Cgp + ci] * wp[g Cg + n][tap][ci] )
//
// xp is the zero-padded, group-padded staging copy dph_posconv_pad writes; wp the tap-major weights of
// dph_posconv_weight_prep.  The generic GEMM expresses this with an overlapping-row operand view and re-loads a
// 128 x 64 activation tile for every one of the 128 taps -- ~2.5 MB of L2 -> SMEM traffic per output tile, which is
// what bounds it (238 TFLOP/s).  Here the Toeplitz structure is used on chip instead: the rows needed by taps
// {kk, kk+8, kk+16, ...} are ONE shared-memory slab (rows t0+kk .. t0+kk+376), and tap kk+8j is that slab advanced
// by j swizzle atoms (8 rows = 1024 bytes), so the MMA descriptor simply moves and the slab is loaded 8 times per
// tile instead of 128 times.  One CTA produces 256 frames x one group (two M=128 accumulators share every weight tile).
#include "common.cuh"
#include "../../include/dphubert_b200.h"

namespace dph {
void set_error(const char* fmt, ...);
int make_tmap_nd_bf16(CUtensorMap* out, const void* ptr, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box, int swizzle);

namespace pcv {
constexpr int MT = 256;                      // frames per CTA (two 128-row accumulators)
constexpr int CI = 64;                       // padded input channels per group (one 128-byte swizzle row)
constexpr int SLAB_ROWS = 384;               // MT + 128 (taps kk + 8 j, j < 16 reach 120 rows further)
constexpr int SLAB_BYTES = SLAB_ROWS * 128;  // 48 KB
constexpr int BSTAGES = 14;
constexpr int BSTAGE_BYTES = 8192;           // up to 64 output channels x 128 bytes
constexpr int S_A = 0;                                   // 2 slabs
constexpr int S_B = S_A + 2 * SLAB_BYTES;                // weight ring
constexpr int S_BAR = S_B + BSTAGES * BSTAGE_BYTES;
constexpr int SMEM = S_BAR + 256 + 1024;
constexpr int EPI_WARPS = 8;
constexpr int THREADS = 32 * (2 + EPI_WARPS);

DPH_DEVICE void tma_load_2d(uint32_t smem_dst, const CUtensorMap* m, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_dst),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}

struct Params {
  int B, T, D, G, Cg, K;   // K taps (multiple of 8, <= 128), Cg output channels per group (multiple of 16, <= 64)
  int rows_b;              // staging rows per utterance (T + pad)
  int row0;                // staging row of (b = 0, t = 0, tap = 0)
  int act;                 // 1 = exact GELU
  const float* bias;       // (D) or null
  const bf16* residual;    // (B T, D) or null
  bf16* out;               // (B T, D)
  bf16* c2;                // pre-activation copy (B T, D) or null
};

__global__ void __launch_bounds__(THREADS, 1)
    posconv_tc_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
                      const __grid_constant__ Params p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sA = sbase + S_A, sB = sbase + S_B, bar0 = sbase + S_BAR;
  auto a_full = [&](int i) { return bar0 + 8u * i; };
  auto a_empty = [&](int i) { return bar0 + 16u + 8u * i; };
  auto b_full = [&](int i) { return bar0 + 32u + 8u * i; };
  auto b_empty = [&](int i) { return bar0 + 32u + 8u * (BSTAGES + i); };
  const uint32_t acc_full = bar0 + 32u + 16u * BSTAGES, tmem_slot = acc_full + 8u;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int m0 = blockIdx.x * MT, g = blockIdx.y, b = blockIdx.z;
  const int ncopy = 8, ntap_per = p.K / 8;

  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(a_full(i), 1);
      mbar_init(a_empty(i), 1);
    }
    for (int i = 0; i < BSTAGES; ++i) {
      mbar_init(b_full(i), 1);
      mbar_init(b_empty(i), 1);
    }
    mbar_init(acc_full, 1);
    mbar_fence_init();
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmW);
  }
  if (warp == 1) {
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
  pdl_wait();

  if (warp == 0) {
    // ======================= TMA producer: weight tiles (one per tap) =======================
    if (elect_one()) {
      int it = 0;
      for (int kk = 0; kk < ncopy; ++kk) {
        for (int j = 0; j < ntap_per; ++j, ++it) {
          const int s = it % BSTAGES;
          if (it >= BSTAGES) mbar_wait(b_empty(s), (uint32_t)(((it / BSTAGES) - 1) & 1));
          mbar_expect_tx(b_full(s), (uint32_t)(p.Cg * 128));
          tma_load_2d(sB + s * BSTAGE_BYTES, &tmW, b_full(s), (kk + 8 * j) * CI, g * p.Cg);
        }
      }
    }
  } else if (warp == 1) {
    // ======================= MMA issuer =======================
    const uint32_t idesc = make_idesc_bf16(128, p.Cg, false, false);
    const int kslices = (p.Cg + 15) / 16;
    int it = 0;
    for (int kk = 0; kk < ncopy; ++kk) {
      const int buf = kk & 1;
      mbar_wait(a_full(buf), (uint32_t)((kk >> 1) & 1));
      for (int j = 0; j < ntap_per; ++j, ++it) {
        const int s = it % BSTAGES;
        mbar_wait(b_full(s), (uint32_t)((it / BSTAGES) & 1));
        tc_fence_after();
        if (elect_one()) {  // elect.sync: ptxas then knows the descriptors are warp-uniform (no waterfall loops)
          const uint32_t bb = sB + s * BSTAGE_BYTES;
          // eight independent accumulation chains (2 row halves x 4 K slices of the 64 input channels): MMAs that
          // accumulate into the same TMEM tile serialise on the MMA latency, and these MMAs are short (N = Cg)
          // input channels Cg .. 63 of a group are zero padding in both operands: their K slices are skipped
#pragma unroll
          for (int k16 = 0; k16 < CI / 16; ++k16) {
            if (k16 >= kslices) break;
#pragma unroll
            for (int half = 0; half < 2; ++half) {
              const uint32_t aa = sA + buf * SLAB_BYTES + (uint32_t)(half * 128 + 8 * j) * 128u;
              umma_f16_ss(tmem_base + (uint32_t)((half * 4 + k16) * 64), make_smem_desc(aa + k16 * 32, 0, 1024, 2),
                          make_smem_desc(bb + k16 * 32, 0, 1024, 2), idesc, it > 0 ? 1u : 0u);
            }
          }
          umma_commit(b_empty(s));
          if (j == ntap_per - 1) umma_commit(a_empty(buf));
          if (it == ncopy * ntap_per - 1) umma_commit(acc_full);
        }
        __syncwarp();
      }
    }
    pdl_trigger();
  } else {
    // ======================= epilogue: 8 warps, (TMEM lane quadrant) x (accumulator half) =======================
    const int quad = warp & 3, half = (warp - 2) >> 2;
    if (warp == 2) {
      // activation slabs (8 per tile, double buffered) are fetched by an otherwise idle epilogue warp, so that they
      // never queue behind the weight-tile ring
      if (elect_one()) {
        const int xrow = p.row0 + b * p.rows_b + m0;
        for (int kk = 0; kk < ncopy; ++kk) {
          const int buf = kk & 1;
          if (kk >= 2) mbar_wait(a_empty(buf), (uint32_t)(((kk >> 1) - 1) & 1));
          mbar_expect_tx(a_full(buf), SLAB_BYTES);
#pragma unroll
          for (int q = 0; q < SLAB_ROWS / 128; ++q)
            tma_load_2d(sA + buf * SLAB_BYTES + q * (128 * 128), &tmX, a_full(buf), g * CI, xrow + kk + q * 128);
        }
      }
      __syncwarp();
    }
    mbar_wait(acc_full, 0);
    tc_fence_after();
    const int t = m0 + half * 128 + quad * 32 + lane;
    const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(half * 4 * 64);
    const bool ok = t < p.T;
    const long long off = ((long long)b * p.T + t) * p.D + g * p.Cg;
#pragma unroll 1
    for (int c = 0; c < p.Cg; c += 16) {
      float v[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = 0.f;
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {  // sum the K-slice accumulators that were used
        if (ks >= (p.Cg + 15) / 16) break;
        uint32_t r[16];
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, "
            "%15}, [%16];"
            : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
              "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
            : "r"(taddr + ks * 64 + c));
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] += __uint_as_float(r[j]);
      }
      if (!ok) continue;
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] += (p.bias ? __ldg(p.bias + g * p.Cg + c + j) : 0.f);
      if (p.c2) {
        uint4 u0, u1;
        u0.x = pack_bf16x2(v[0], v[1]); u0.y = pack_bf16x2(v[2], v[3]);
        u0.z = pack_bf16x2(v[4], v[5]); u0.w = pack_bf16x2(v[6], v[7]);
        u1.x = pack_bf16x2(v[8], v[9]); u1.y = pack_bf16x2(v[10], v[11]);
        u1.z = pack_bf16x2(v[12], v[13]); u1.w = pack_bf16x2(v[14], v[15]);
        uint4* d = reinterpret_cast<uint4*>(p.c2 + off + c);
        d[0] = u0;
        d[1] = u1;
      }
      if (p.act == 1) {
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = gelu_erf_fast(v[j]);
      }
      if (p.residual) {
        const uint4* rp = reinterpret_cast<const uint4*>(p.residual + off + c);
        const uint4 a0 = rp[0], a1 = rp[1];
        const uint32_t w[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[j]));
          v[2 * j] += f.x;
          v[2 * j + 1] += f.y;
        }
      }
      uint4 u0, u1;
      u0.x = pack_bf16x2(v[0], v[1]); u0.y = pack_bf16x2(v[2], v[3]);
      u0.z = pack_bf16x2(v[4], v[5]); u0.w = pack_bf16x2(v[6], v[7]);
      u1.x = pack_bf16x2(v[8], v[9]); u1.y = pack_bf16x2(v[10], v[11]);
      u1.z = pack_bf16x2(v[12], v[13]); u1.w = pack_bf16x2(v[14], v[15]);
      uint4* d = reinterpret_cast<uint4*>(p.out + off + c);
      d[0] = u0;
      d[1] = u1;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace pcv
}  // namespace dph

using namespace dph;

// xp: (rows_total, G * 64) bf16 staging copy; wp: (G * Cg, K * 64) bf16 tap-major weights (forward) or the
// tap-flipped / transposed copy (input gradient); row0 = staging row read by (b = 0, t = 0, tap = 0).
extern "C" int dph_posconv_tc(const void* xp, long long rows_total, const void* wp, const float* bias,
                              const void* residual, void* out, void* c2, int B, int T, int D, int G, int Cg, int Cgp,
                              int K, int rows_b, int row0, int act, void* stream) {
  using namespace pcv;
  if (!xp || !wp || !out || B <= 0 || T <= 0) {
    set_error("posconv_tc: bad arguments");
    return DPH_ERR_ARG;
  }
  if (Cgp != CI || Cg % 16 != 0 || Cg > 64 || Cg < 16 || K % 8 != 0 || K > 128 || K < 64 || D != G * Cg || row0 < 0 ||
      (D * 2) % 16 != 0 || (Cg * 2) % 16 != 0 || B > 65535 || G > 65535) {
    set_error("posconv_tc: unsupported geometry (Cg=%d Cgp=%d K=%d)", Cg, Cgp, K);
    return DPH_ERR_UNSUPPORTED;
  }
  CUtensorMap tmX, tmW;
  {
    const uint64_t dims[2] = {(uint64_t)G * CI, (uint64_t)rows_total};
    const uint64_t strides[1] = {(uint64_t)G * CI * 2};
    const uint32_t box[2] = {CI, 128};
    int rc = make_tmap_nd_bf16(&tmX, xp, 2, dims, strides, box, 3);
    if (rc) return rc;
    const uint64_t wd[2] = {(uint64_t)K * CI, (uint64_t)G * Cg};
    const uint64_t ws[1] = {(uint64_t)K * CI * 2};
    const uint32_t wb[2] = {CI, (uint32_t)Cg};
    rc = make_tmap_nd_bf16(&tmW, wp, 2, wd, ws, wb, 3);
    if (rc) return rc;
  }
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(posconv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) {
      set_error("posconv_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return DPH_ERR_CUDA;
    }
    attr = true;
  }
  Params p;
  p.B = B; p.T = T; p.D = D; p.G = G; p.Cg = Cg; p.K = K; p.rows_b = rows_b; p.row0 = row0; p.act = act;
  p.bias = bias; p.residual = (const bf16*)residual; p.out = (bf16*)out; p.c2 = (bf16*)c2;
  dim3 grid((T + MT - 1) / MT, G, B);
  cudaError_t e = launch_pdl(posconv_tc_kernel, grid, dim3(THREADS), SMEM, reinterpret_cast<cudaStream_t>(stream), tmX,
                             tmW, p);
  if (e != cudaSuccess) {
    set_error("posconv_tc launch: %s", cudaGetErrorString(e));
    return DPH_ERR_CUDA;
  }
  dph_count_launch();
  return dph_check_last("posconv_tc");
}
