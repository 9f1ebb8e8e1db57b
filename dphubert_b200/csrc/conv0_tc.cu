// Layer-0 feature-extractor block on the tensor cores (bf16 activation path):
//   y[b, t, c] = mask[c] * GELU( scale[b,c] * (sum_q w[c,q] x[b, s t + q]) + shift[b,c] )
// (wav2vec2/components.py:107-114 for conv_layers[0]; GroupNorm statistics come from dph_conv0_stats).
//
// The layer is bound by HBM (it writes 2 * C * T1 bytes per utterance and reads almost nothing), but a CUDA-core
// version needs ~30 instructions per output (10 FMAs for the taps, an erf GELU with two MUFU operations) and ends up
// bound by instruction issue at ~1/5 of the HBM roofline.  Here
//   * the k <= 10 taps run as ONE tcgen05 MMA pair per 128 x 256 output tile: the waveform window of every frame is
//     written to shared memory as a K = 32 bf16 row [x_hi | x_lo | x_hi | 0] and multiplied with [w_hi | w_hi | w_lo | 0]
//     (hi = bf16(v), lo = bf16(v - hi): x w ~ x_hi w_hi + x_lo w_hi + x_hi w_lo, 16 mantissa bits, fp32 accumulation
//     in tensor memory) -- no precision is given up against the fp32 FMA chain at bf16 output;
//   * the epilogue reads the accumulator from TMEM and spends ~6 issue slots per output: packed fp32x2 FMAs
//     (FFMA2) for the affine part and for GELU(u) = u * (0.5 + 0.5 tanh(u (a + b u^2 + c u^4))) (|Phi error| < 2e-5
//     before MUFU.TANH's own 2^-11, i.e. below bf16 resolution), one MUFU per output;
//   * persistent CTAs (one per SM) walk contiguous (utterance, frame tile, channel block) units; the waveform window,
//     the MMA and the epilogue of neighbouring units overlap through two A buffers and two TMEM accumulators.
#include "common.cuh"
#include "../../include/dphubert_b200.h"

namespace dph {
void set_error(const char* fmt, ...);

int make_tmap_nd_bf16(CUtensorMap* out, const void* ptr, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box, int swizzle);

namespace c0tc {
constexpr int TT = 128;                 // frames per tile (= TMEM lanes)
constexpr int NB = 256;                 // channels per MMA / accumulator
constexpr int MAXC = 512;
constexpr int MAXK = 10;
constexpr int EPI_WARPS = 16;                   // four per TMEM lane quadrant, 64 accumulator columns each
constexpr int EPI_COLS = NB / (EPI_WARPS / 4);  // columns per epilogue warp
constexpr int EPI_CHUNKS = EPI_COLS / 32;
constexpr int PROD_WARPS = 4;           // one window row per producer thread
constexpr int THREADS = 32 * (PROD_WARPS + EPI_WARPS);
constexpr int S_B = 0;                                       // weights  [512][64] bf16, 128B swizzle   64 KB
constexpr int S_A = S_B + MAXC * 128;                        // windows  2 x [128][64] bf16             32 KB
constexpr int S_TAB = S_A + 2 * TT * 128;                    // per-channel {sc, sh, sc m, sh m} pairs  8 KB
constexpr int S_OUT = S_TAB + MAXC * 16;                     // output staging, 4 KB per epilogue warp          64 KB
constexpr int S_BAR = S_OUT + EPI_WARPS * 4096;              // barriers
constexpr int SMEM = S_BAR + 128 + 1024;

DPH_DEVICE uint32_t sw128(int r, int c) { return (uint32_t)(r * 128 + ((c ^ (r & 7)) << 4)); }

// hi / lo bf16 split of a float
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
  bf16* y;
  int B, C, k, stride, T1, norm;
  int ntile, nblk;
};

template <bool ACT>
__global__ void __launch_bounds__(THREADS, 1)
    conv0_fwd_tc_kernel(const __grid_constant__ Params p, const __grid_constant__ CUtensorMap tmY) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (sbase - smem_u32(smem_raw));
  const uint32_t sB = sbase + S_B, sA = sbase + S_A;
  float* tab = reinterpret_cast<float*>(gbase + S_TAB);
  const uint32_t bar0 = sbase + S_BAR;
  auto a_free = [&](int i) { return bar0 + 8u * i; };        // [2] MMAs that read A buffer i have completed
  auto t_full = [&](int i) { return bar0 + 16u + 8u * i; };  // [2] accumulator i is complete
  auto t_empty = [&](int i) { return bar0 + 32u + 8u * i; }; // [2] accumulator i has been read by all epilogue warps
  const uint32_t tmem_slot = bar0 + 48u;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(a_free(i), 1);
      mbar_init(t_full(i), 1);
      mbar_init(t_empty(i), EPI_WARPS);
    }
    mbar_fence_init();
  }
  if (warp == 0) {
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  pdl_wait();
  // ---- weights -> swizzled bf16 B tile: row c = [w_hi(10) | w_hi(10) | w_lo(10) | 0 0], K chunks 4..7 are never read
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
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  // contiguous unit range of this CTA; unit = ((b * ntile + tile) * nblk + blk)
  const long long total = (long long)p.B * p.ntile * p.nblk;
  const long long u0 = total * blockIdx.x / gridDim.x, u1 = total * (blockIdx.x + 1) / gridDim.x;

  if (warp < PROD_WARPS) {
    // ======================= window builders (128 threads, one frame each) + MMA issuer (thread 0) ===============
    int acc = 0, nA = 0;
    uint32_t acc_ph = 0;
    const long long bt0 = u0 / p.nblk, bt1 = (u1 + p.nblk - 1) / p.nblk;
    for (long long bt = bt0; bt < bt1 && u0 < u1; ++bt, ++nA) {
      const int abuf = nA & 1;
      const int b = (int)(bt / p.ntile), t0 = (int)(bt % p.ntile) * TT;
      // wait until the MMAs that read this buffer two tiles ago are done
      if (nA >= 2) mbar_wait(a_free(abuf), (uint32_t)(((nA >> 1) - 1) & 1));
      {
        const int r = tid;  // frame of this thread inside the tile
        const int nt = min(TT, p.T1 - t0);
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
      asm volatile("bar.sync 2, %0;" ::"n"(PROD_WARPS * 32));
      if (warp == 0) {
        const long long ua = bt * p.nblk > u0 ? bt * p.nblk : u0;
        const long long ub = (bt + 1) * p.nblk < u1 ? (bt + 1) * p.nblk : u1;
        for (long long u = ua; u < ub; ++u) {
          const int blk = (int)(u % p.nblk);
          mbar_wait(t_empty(acc), acc_ph ^ 1u);  // accumulator drained by the epilogue
          tc_fence_after();
          if (elect_one()) {
            const int nb = min(NB, p.C - blk * NB);
            const uint32_t idesc = make_idesc_bf16(TT, nb, false, false);
            const uint32_t a_addr = sA + (uint32_t)abuf * (TT * 128), b_addr = sB + (uint32_t)blk * (NB * 128);
#pragma unroll
            for (int kk = 0; kk < 2; ++kk)
              umma_f16_ss(tmem_base + (uint32_t)(acc * NB), make_smem_desc(a_addr + kk * 32, 0, 1024, 2),
                          make_smem_desc(b_addr + kk * 32, 0, 1024, 2), idesc, kk > 0 ? 1u : 0u);
            umma_commit(t_full(acc));
            if (u + 1 == ub) umma_commit(a_free(abuf));
          }
          __syncwarp();
          acc ^= 1;
          if (acc == 0) acc_ph ^= 1u;
        }
      }
    }
    pdl_trigger();
  } else {
    // ======================= epilogue warps =======================
    const int ew = warp - PROD_WARPS;
    const int quad = warp & 3;           // TMEM lane quadrant this warp may read
    const int chalf = ew >> 2;           // which EPI_COLS-wide slice of the accumulator's 256 columns
    const int etid = tid - PROD_WARPS * 32;
    const uint32_t stage = sbase + S_OUT + (uint32_t)ew * 4096u;
    int acc = 0, cur_b = -1;
    uint32_t acc_ph = 0;
    constexpr float GA = 0.7977176713187712f, GB = 0.03679851046109153f, GC = -0.00031586434495541917f;
    for (long long u = u0; u < u1; ++u) {
      const int blk = (int)(u % p.nblk);
      const long long bt = u / p.nblk;
      const int b = (int)(bt / p.ntile), t0 = (int)(bt % p.ntile) * TT;
      if (b != cur_b) {
        // per-channel table of this utterance: {sc, sh} and {sc * m, sh * m}, interleaved per channel PAIR so that
        // one 16-byte shared load yields an fp32x2 operand pair
        asm volatile("bar.sync 1, %0;" ::"n"(EPI_WARPS * 32));
        for (int c = etid; c < MAXC; c += EPI_WARPS * 32) {
          float sc = 1.f, sh = 0.f, m = 1.f;
          if (c < p.C) {
            const float bb = p.bias ? __ldg(p.bias + c) : 0.f;
            if (p.norm) {
              const float rs = p.rstd[b * p.C + c], mu = p.mean[b * p.C + c], g = __ldg(p.gamma + c);
              sc = rs * g;
              sh = __ldg(p.beta + c) + (bb - mu) * rs * g;
            } else {
              sh = bb;
            }
            m = p.mask ? __ldg(p.mask + c) : 1.f;
          }
          // layout per pair P = c / 2, e = c & 1: tab[P * 8 + {0,1}] = sc, [2,3] = sh, [4,5] = sc m, [6,7] = sh m
          float* t = tab + (c >> 1) * 8 + (c & 1);
          t[0] = sc;
          t[2] = sh;
          t[4] = sc * m;
          t[6] = sh * m;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(EPI_WARPS * 32));
        cur_b = b;
      }
      mbar_wait(t_full(acc), acc_ph);
      tc_fence_after();
      const int nb = min(NB, p.C - blk * NB);
      const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * NB + chalf * EPI_COLS);
      const bool active = chalf * EPI_COLS < nb;  // C is a multiple of 64: a warp's 64 columns are all in or all out
      if (active) {
        // the TMA store of the previous unit must have finished reading this warp's staging tile
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncwarp();
      }
#pragma unroll 1
      for (int ch = 0; ch < EPI_CHUNKS; ++ch) {
        const int cl = chalf * EPI_COLS + ch * 32;  // column inside the accumulator (warp-uniform)
        uint32_t r[32];
        if (active) {
          tmem_ld_32x32(taddr + ch * 32, r);
          tmem_ld_wait();
        }
        if (ch == EPI_CHUNKS - 1) {  // all TMEM reads of this accumulator are done: hand it back
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(t_empty(acc));
        }
        if (!active) continue;
        const int c0 = blk * NB + cl;
        const float4* tp = reinterpret_cast<const float4*>(tab + (c0 >> 1) * 8);
        uint32_t o[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float4 t0v = tp[2 * j], t1v = tp[2 * j + 1];  // {sc0, sc1, sh0, sh1}, {scm0, scm1, shm0, shm1}
          const unsigned long long z = pk2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
          unsigned long long um = fma2(z, pk2(t1v.x, t1v.y), pk2(t1v.z, t1v.w));  // (z sc + sh) m
          float y0, y1;
          if constexpr (ACT) {
            const unsigned long long uu = fma2(z, pk2(t0v.x, t0v.y), pk2(t0v.z, t0v.w));  // u = z sc + sh
            float s0, s1;
            upk2(mul2(uu, uu), s0, s1);
            const unsigned long long s = pk2(fminf(s0, 64.f), fminf(s1, 64.f));
            unsigned long long q = fma2(s, pk2(GC, GC), pk2(GB, GB));
            q = fma2(q, s, pk2(GA, GA));
            float a0, a1;
            upk2(mul2(q, uu), a0, a1);
            const unsigned long long th = pk2(tanh_approx(a0), tanh_approx(a1));
            const unsigned long long h = fma2(th, pk2(0.5f, 0.5f), pk2(0.5f, 0.5f));
            upk2(mul2(um, h), y0, y1);
          } else {
            upk2(um, y0, y1);
          }
          o[j] = pack_bf16x2(y0, y1);
        }
        // stage this thread's 64 bytes (frame = lane, channels ch * 32 ..) in the warp's 128B-swizzled 32 x 64 tile
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint32_t addr = stage + sw128(lane, ch * 4 + q);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(o[4 * q]), "r"(o[4 * q + 1]),
                       "r"(o[4 * q + 2]), "r"(o[4 * q + 3])
                       : "memory");
        }
      }
      if (active) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
          // one coalesced bulk store of the warp's 32 frames x 64 channels; frames >= T1 are clipped by the map
          asm volatile(
              "cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%1, %2, %3}], [%4];" ::"l"(
                  reinterpret_cast<uint64_t>(&tmY)),
              "r"(blk * NB + chalf * EPI_COLS), "r"(t0 + quad * 32), "r"(b), "r"(stage)
              : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
      acc ^= 1;
      if (acc == 0) acc_ph ^= 1u;
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace c0tc

// returns DPH_ERR_UNSUPPORTED (without setting an error) when the shape is outside this kernel's envelope
int conv0_fwd_tc(const float* wave, long long wave_bs, const float* w, const float* bias, const float* gamma,
                 const float* beta, const float* mean, const float* rstd, const float* mask, bf16* y, int B, int C, int k,
                 int stride, int T1, int norm, int act, cudaStream_t s) {
  using namespace c0tc;
  if (C % 64 != 0 || C > MAXC || k > MAXK || (reinterpret_cast<uintptr_t>(y) & 31) != 0)
    return DPH_ERR_UNSUPPORTED;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(conv0_fwd_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(conv0_fwd_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) {
      set_error("conv0_fwd_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return DPH_ERR_CUDA;
    }
    attr = true;
  }
  Params p;
  p.wave = wave; p.wave_bs = wave_bs; p.w = w; p.bias = bias; p.gamma = gamma; p.beta = beta; p.mean = mean;
  p.rstd = rstd; p.mask = mask; p.y = y; p.B = B; p.C = C; p.k = k; p.stride = stride; p.T1 = T1; p.norm = norm;
  p.ntile = (T1 + TT - 1) / TT;
  p.nblk = (C + NB - 1) / NB;
  const long long total = (long long)B * p.ntile * p.nblk;
  int dev = 0, sms = kNumSMsDefault;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = (int)(total < sms ? total : sms);
  CUtensorMap tm;
  {
    const uint64_t dims[3] = {(uint64_t)C, (uint64_t)T1, (uint64_t)B};
    const uint64_t strides[2] = {(uint64_t)C * 2, (uint64_t)T1 * C * 2};
    const uint32_t box[3] = {64, 32, 1};
    const int rc = make_tmap_nd_bf16(&tm, y, 3, dims, strides, box, 3);
    if (rc) return rc;
  }
  cudaError_t e;
  if (act) e = launch_pdl(conv0_fwd_tc_kernel<true>, dim3(grid), dim3(THREADS), SMEM, s, p, tm);
  else e = launch_pdl(conv0_fwd_tc_kernel<false>, dim3(grid), dim3(THREADS), SMEM, s, p, tm);
  if (e != cudaSuccess) {
    set_error("conv0_fwd_tc launch: %s", cudaGetErrorString(e));
    return DPH_ERR_CUDA;
  }
  return DPH_OK;
}

}  // namespace dph
