// Attention-probability kernels: masked / biased softmax forward and backward, WavLM gated relative position bias.
//
// Replaces wav2vec2/components.py:411-420 (scores += mask; -rowmax; softmax; dropout) and, for WavLM,
// :629-647 (bucketed relative-position embedding x per-(batch, head, query) GRU gate added to the scores) plus
// their autograd backward.  The additive padding mask (components.py:979-983) is rebuilt from `lengths` on the fly:
// no (B,1,T,T) tensor and no (B*H,T,T) bias tensor is ever materialised.
//
// One warp owns one score row (T <= 1024 keys, i.e. ~20 s of audio; the reference caps utterances at 15.6 s).
#include "common.cuh"
#include "vec.cuh"
#include "../../include/dphubert_b200.h"

namespace dph {
DPH_DEVICE void load_pair(const float* p, float& a, float& b) {
  const float2 f = *reinterpret_cast<const float2*>(p);
  a = f.x;
  b = f.y;
}
DPH_DEVICE void load_pair(const bf16* p, float& a, float& b) {
  const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p));
  a = f.x;
  b = f.y;
}
DPH_DEVICE void store_pair(float* p, float a, float b) { *reinterpret_cast<float2*>(p) = make_float2(a, b); }
DPH_DEVICE void store_pair(bf16* p, float a, float b) {
  *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(a, b);
}
void set_error(const char* fmt, ...);

constexpr int SM_NPL = 32;  // keys per lane -> T <= 1024

struct WavlmBias {
  const int* rel_table;   // (2T-1) bucket index of (j - i) + T - 1, built on the host exactly like the reference
  const float* emb;       // (num_buckets, Htot)
  const float* gate;      // (B, Htot, T) or null (ungated)
  const int* head_index;  // (H) remaining head -> index in [0, Htot)
  int Htot;
};

template <typename T>
__global__ void __launch_bounds__(256)
    softmax_fwd_kernel(const float* __restrict__ S, long long lds, T* __restrict__ P, T* __restrict__ Pd, long long ldp,
                       const long long* __restrict__ lengths, int B, int H, int Tn, float drop_p,
                       unsigned long long seed, WavlmBias wb) {
  const long long row = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const long long nrows = (long long)B * H * Tn;
  if (row >= nrows) return;
  const int i = (int)(row % Tn);
  const int bh = (int)(row / Tn);
  const int h = bh % H, b = bh / H;
  const float* s = S + row * lds;
  const int len = lengths ? (int)lengths[b] : Tn;
  float gate = 1.f;
  int htot = h;
  if (wb.emb) {
    htot = wb.head_index ? wb.head_index[h] : h;
    if (wb.gate) gate = wb.gate[((long long)b * wb.Htot + htot) * Tn + i];
  }
  float v[SM_NPL];
  float mx = -INFINITY;
#pragma unroll
  for (int k = 0; k < SM_NPL; ++k) {
    const int j = lane + 32 * k;
    float x = -INFINITY;
    if (j < Tn) {
      x = s[j];
      if (wb.emb) x += gate * __ldg(wb.emb + (long long)wb.rel_table[j - i + Tn - 1] * wb.Htot + htot);
      if (j >= len) x += -10000.0f;
    }
    v[k] = x;
    mx = fmaxf(mx, x);
  }
  mx = warp_max(mx);
  float sum = 0.f;
#pragma unroll
  for (int k = 0; k < SM_NPL; ++k) {
    const int j = lane + 32 * k;
    float e = (j < Tn) ? expf(v[k] - mx) : 0.f;
    v[k] = e;
    sum += e;
  }
  sum = warp_sum(sum);
  const float inv = 1.f / sum;
  const uint32_t thresh = (uint32_t)(drop_p * 4294967296.0);
  const float inv_keep = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
#pragma unroll
  for (int k = 0; k < SM_NPL; ++k) {
    const int j = lane + 32 * k;
    if (j < Tn) {
      const float p = v[k] * inv;
      P[row * ldp + j] = from_f<T>(p);
      if (Pd) Pd[row * ldp + j] = from_f<T>(p * dropout_row_elem(dropout_row_key(dropout_key(seed), (unsigned long long)row), (uint32_t)j, thresh, inv_keep));
    } else if (j < ldp) {
      // keep the row padding (ldp >= Tn, multiple of 8) finite and zero: it is read as K-tail by the P*V GEMM
      P[row * ldp + j] = from_f<T>(0.f);
      if (Pd) Pd[row * ldp + j] = from_f<T>(0.f);
    }
  }
}

// dS = P * (dP - sum_j dP_j P_j), dP = dPd * dropout mask.  Optionally accumulates the WavLM bias gradients.
template <typename T>
__global__ void __launch_bounds__(256)
    softmax_bwd_kernel(const T* __restrict__ P, const float* __restrict__ dPd, long long lds, T* __restrict__ dS,
                       long long ldp, int B, int H, int Tn, float drop_p, unsigned long long seed, WavlmBias wb,
                       float* __restrict__ dgate, float* __restrict__ demb, int num_buckets, int rows_per_blk) {
  extern __shared__ float s_emb[];  // [num_buckets] when demb != null
  const int warps = blockDim.x >> 5, w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // grid.y = (b, h); grid.x walks the Tn query rows of that head in chunks of rows_per_blk
  const int i0 = blockIdx.x * rows_per_blk;
  const int nloc = min(rows_per_blk, Tn - i0);
  const long long row0 = (long long)blockIdx.y * Tn + i0;
  if (demb) {
    for (int q = threadIdx.x; q < num_buckets; q += blockDim.x) s_emb[q] = 0.f;
    __syncthreads();
  }
  const uint32_t thresh = (uint32_t)(drop_p * 4294967296.0);
  const float inv_keep = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
  int htot_blk = 0;
  for (int rr = w; rr < nloc; rr += warps) {
    const long long row = row0 + rr;
    const int i = (int)(row % Tn);
    const int bh = (int)(row / Tn);
    const int h = bh % H, b = bh / H;
    float p[SM_NPL], d[SM_NPL];
    float dot = 0.f;
#pragma unroll
    for (int k = 0; k < SM_NPL; ++k) {
      const int j = lane + 32 * k;
      p[k] = 0.f;
      d[k] = 0.f;
      if (j < Tn) {
        p[k] = to_f(P[row * ldp + j]);
        float g = dPd[row * lds + j];
        if (drop_p > 0.f) g *= dropout_row_elem(dropout_row_key(dropout_key(seed), (unsigned long long)row), (uint32_t)j, thresh, inv_keep);
        d[k] = g;
        dot += g * p[k];
      }
    }
    dot = warp_sum(dot);
    float gate = 1.f, dg = 0.f;
    int htot = h;
    if (wb.emb) {
      htot = wb.head_index ? wb.head_index[h] : h;
      htot_blk = htot;
      if (wb.gate) gate = wb.gate[((long long)b * wb.Htot + htot) * Tn + i];
    }
#pragma unroll
    for (int k = 0; k < SM_NPL; ++k) {
      const int j = lane + 32 * k;
      if (j < Tn) {
        const float ds = p[k] * (d[k] - dot);
        dS[row * ldp + j] = from_f<T>(ds);
        if (wb.emb) {
          const int bucket = wb.rel_table[j - i + Tn - 1];
          if (dgate) dg += ds * __ldg(wb.emb + (long long)bucket * wb.Htot + htot);
          if (demb) atomicAdd(&s_emb[bucket], ds * gate);
        }
      } else if (j < ldp) {
        dS[row * ldp + j] = from_f<T>(0.f);
      }
    }
    if (dgate) {
      dg = warp_sum(dg);
      if (lane == 0) dgate[((long long)b * wb.Htot + htot) * Tn + i] = dg;
    }
  }
  if (demb) {
    __syncthreads();
    // all rows of this block share (b, h): flush the per-block histogram
    const int bh = (int)(row0 / Tn);
    const int h = bh % H;
    const int htot = wb.head_index ? wb.head_index[h] : h;
    (void)htot_blk;
    for (int q = threadIdx.x; q < num_buckets; q += blockDim.x) {
      const float val = s_emb[q];
      if (val != 0.f) atomicAdd(demb + (long long)q * wb.Htot + htot, val);
    }
  }
}

// WavLM gate (components.py:636-644):  g8 = Linear(64->8)(x[b,t,h,:]);  ga = sigmoid(sum g8[0:4]); gb = sigmoid(sum g8[4:8])
//   gate = ga * (gb * const[h] - 1) + 2.   x is the layer input viewed as (B, T, Htot, head_dim).
// One thread per (b, t, h): the four outputs of each half are only ever summed, so the 8 x Dh weight collapses to two
// rows WA = sum W[0:4], WB = sum W[4:8] (shared memory); the Dh inputs of a thread are contiguous (128-bit loads, and
// consecutive threads read consecutive 2 * Dh-byte segments).  Memory-bound: reads x once (6 MB at B = 8 x 10 s).
template <typename T>
__global__ void __launch_bounds__(256)
    wavlm_gate_fwd_kernel(const T* __restrict__ x, const float* __restrict__ Wg, const float* __restrict__ bg,
                          const float* __restrict__ cst, float* __restrict__ gate, float* __restrict__ ga_out,
                          float* __restrict__ gb_out, int B, int Tn, int Htot, int Dh) {
  extern __shared__ float sw[];  // WA[Dh], WB[Dh]
  float* WA = sw;
  float* WB = sw + Dh;
  for (int d = threadIdx.x; d < Dh; d += blockDim.x) {
    WA[d] = (__ldg(Wg + d) + __ldg(Wg + Dh + d)) + (__ldg(Wg + 2 * Dh + d) + __ldg(Wg + 3 * Dh + d));
    WB[d] = (__ldg(Wg + 4 * Dh + d) + __ldg(Wg + 5 * Dh + d)) + (__ldg(Wg + 6 * Dh + d) + __ldg(Wg + 7 * Dh + d));
  }
  __syncthreads();
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // over (b, t, h)
  const long long total = (long long)B * Tn * Htot;
  if (idx >= total) return;
  const int h = (int)(idx % Htot);
  const long long bt = idx / Htot;
  const int t = (int)(bt % Tn), b = (int)(bt / Tn);
  const T* xv = x + bt * ((long long)Htot * Dh) + (long long)h * Dh;
  constexpr int VEC = VecOf<T>::N;
  float a = (bg[0] + bg[1]) + (bg[2] + bg[3]), c = (bg[4] + bg[5]) + (bg[6] + bg[7]);
  if ((Dh % VEC) == 0 && ((reinterpret_cast<uintptr_t>(xv) & 15) == 0)) {
    for (int d0 = 0; d0 < Dh; d0 += VEC) {
      float v[VEC];
      load_vec(xv + d0, v);
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        a = fmaf(v[j], WA[d0 + j], a);
        c = fmaf(v[j], WB[d0 + j], c);
      }
    }
  } else {
    for (int d = 0; d < Dh; ++d) {
      const float xd = to_f(xv[d]);
      a = fmaf(xd, WA[d], a);
      c = fmaf(xd, WB[d], c);
    }
  }
  const float ga = sigmoidf_(a);
  const float gb = sigmoidf_(c);
  const long long o = ((long long)b * Htot + h) * Tn + t;
  gate[o] = ga * (gb * cst[h] - 1.f) + 2.f;
  if (ga_out) ga_out[o] = ga;
  if (gb_out) gb_out[o] = gb;
}

// backward of the gate: accumulates dWg (8,Dh), dbg (8), dconst (Htot) and ADDS the input gradient into dx.
// One warp per (b, t, h) row, lane l owns head-dim elements l, l+32, ...: the pre-sigmoid gradient takes only two
// distinct values per row (outputs 0-3 share dpa, 4-7 share dpb), so dWg needs two register accumulators per owned
// element and the input gradient two precomputed column sums of Wg; partial sums meet in shared memory once per warp.
template <typename T>
__global__ void __launch_bounds__(256)
    wavlm_gate_bwd_kernel(const T* __restrict__ x, const float* __restrict__ Wg, const float* __restrict__ cst,
                          const float* __restrict__ ga_in, const float* __restrict__ gb_in,
                          const float* __restrict__ dgate, T* __restrict__ dx, float* __restrict__ dWg,
                          float* __restrict__ dbg, float* __restrict__ dcst, int B, int Tn, int Htot, int Dh) {
  // One warp per (b, t, h) row; lane l owns the ADJACENT head-dim elements 2l, 2l+1 (+64 per extra pass), so a row is
  // one coalesced 4-byte-per-lane access for x and for the dx read-modify-write.  Per-warp partial sums live in
  // registers (dWg: two accumulators per owned element, since outputs 0-3 share dpa and 4-7 share dpb) and in a
  // per-warp slice of shared memory (dconst[h], written by lane 0 only: plain adds, no atomics); they meet once at
  // the end of the block.
  constexpr int NP = 2;           // passes of 64 elements (Dh <= 128)
  extern __shared__ float sh[];   // s_A[Dh] s_B[Dh] s_db[2] | per warp: s_dc[Htot]
  float* s_A = sh;
  float* s_B = sh + Dh;
  float* s_db = sh + 2 * Dh;
  float* s_dc_all = s_db + 2;
  const int nwarp = blockDim.x >> 5;
  for (int q = threadIdx.x; q < 2 * Dh + 2 + nwarp * Htot; q += blockDim.x) sh[q] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* s_dc = s_dc_all + warp * Htot;
  float wA[NP][2], wB[NP][2], accA[NP][2], accB[NP][2];
#pragma unroll
  for (int q = 0; q < NP; ++q)
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int d = 64 * q + 2 * lane + e;
      wA[q][e] = wB[q][e] = accA[q][e] = accB[q][e] = 0.f;
      if (d < Dh) {
#pragma unroll
        for (int o8 = 0; o8 < 4; ++o8) {
          wA[q][e] += __ldg(Wg + o8 * Dh + d);
          wB[q][e] += __ldg(Wg + (o8 + 4) * Dh + d);
        }
      }
    }
  float dbA = 0.f, dbB = 0.f;
  const long long total = (long long)B * Tn * Htot;
  const bool even = (Dh % 2) == 0;
  for (long long idx = (long long)blockIdx.x * nwarp + warp; idx < total; idx += (long long)gridDim.x * nwarp) {
    const int h = (int)(idx % Htot);
    const long long bt = idx / Htot;
    const int t = (int)(bt % Tn), b = (int)(bt / Tn);
    const long long o = ((long long)b * Htot + h) * Tn + t;
    const float dg = dgate[o], ga = ga_in[o], gb = gb_in[o], c = cst[h];
    // gate = ga * (gb * c - 1) + 2
    const float dga = dg * (gb * c - 1.f), dgb = dg * ga * c;
    const float dpa = dga * ga * (1.f - ga), dpb = dgb * gb * (1.f - gb);  // wrt the pre-sigmoid sums
    if (lane == 0) s_dc[h] += dg * ga * gb;
    dbA += dpa;
    dbB += dpb;
    const long long xo = bt * ((long long)Htot * Dh) + (long long)h * Dh;
#pragma unroll
    for (int q = 0; q < NP; ++q) {
      const int d = 64 * q + 2 * lane;
      if (d >= Dh) continue;
      float x0, x1 = 0.f;
      if (even) {
        load_pair(x + xo + d, x0, x1);
      } else {
        x0 = to_f(x[xo + d]);
        if (d + 1 < Dh) x1 = to_f(x[xo + d + 1]);
      }
      accA[q][0] = fmaf(dpa, x0, accA[q][0]);
      accB[q][0] = fmaf(dpb, x0, accB[q][0]);
      accA[q][1] = fmaf(dpa, x1, accA[q][1]);
      accB[q][1] = fmaf(dpb, x1, accB[q][1]);
      if (dx) {
        if (even) {
          float g0, g1;
          load_pair(dx + xo + d, g0, g1);
          store_pair(dx + xo + d, g0 + dpa * wA[q][0] + dpb * wB[q][0], g1 + dpa * wA[q][1] + dpb * wB[q][1]);
        } else {
          dx[xo + d] = from_f<T>(to_f(dx[xo + d]) + dpa * wA[q][0] + dpb * wB[q][0]);
          if (d + 1 < Dh) dx[xo + d + 1] = from_f<T>(to_f(dx[xo + d + 1]) + dpa * wA[q][1] + dpb * wB[q][1]);
        }
      }
    }
  }
#pragma unroll
  for (int q = 0; q < NP; ++q)
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int d = 64 * q + 2 * lane + e;
      if (d < Dh) {
        atomicAdd(&s_A[d], accA[q][e]);   // once per warp per element
        atomicAdd(&s_B[d], accB[q][e]);
      }
    }
  if (lane == 0) {
    atomicAdd(&s_db[0], dbA);
    atomicAdd(&s_db[1], dbB);
  }
  __syncthreads();
  for (int q = threadIdx.x; q < 8 * Dh; q += blockDim.x) atomicAdd(dWg + q, (q / Dh) < 4 ? s_A[q % Dh] : s_B[q % Dh]);
  for (int q = threadIdx.x; q < 8; q += blockDim.x) atomicAdd(dbg + q, q < 4 ? s_db[0] : s_db[1]);
  for (int q = threadIdx.x; q < Htot; q += blockDim.x) {
    float tsum = 0.f;
    for (int w8 = 0; w8 < nwarp; ++w8) tsum += s_dc_all[w8 * Htot + q];
    atomicAdd(dcst + q, tsum);
  }
}

// Fast variant for Dh % VEC == 0 and (Dh / VEC) a divisor of 32 (head dim 64: 8 lanes per row in bf16): a warp-pass
// covers 32 / LPR rows with 16-byte accesses, so a warp walks 4x fewer dependent (scalar -> row) load chains than with
// one row per pass; per-lane accumulators for the VEC owned elements, lane groups combined by shuffles at the end.
template <typename T>
__global__ void __launch_bounds__(256)
    wavlm_gate_bwd_vec_kernel(const T* __restrict__ x, const float* __restrict__ Wg, const float* __restrict__ cst,
                              const float* __restrict__ ga_in, const float* __restrict__ gb_in,
                              const float* __restrict__ dgate, T* __restrict__ dx, float* __restrict__ dWg,
                              float* __restrict__ dbg, float* __restrict__ dcst, int B, int Tn, int Htot, int Dh) {
  constexpr int VEC = VecOf<T>::N;
  extern __shared__ float sh[];   // s_A[Dh] s_B[Dh] s_db[2] | per warp: s_dc[Htot]
  float* s_A = sh;
  float* s_B = sh + Dh;
  float* s_db = sh + 2 * Dh;
  float* s_dc_all = s_db + 2;
  const int nwarp = blockDim.x >> 5;
  for (int q = threadIdx.x; q < 2 * Dh + 2 + nwarp * Htot; q += blockDim.x) sh[q] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int lpr = Dh / VEC, rpw = 32 / lpr;      // lanes per row, rows per warp-pass
  const int sub = lane % lpr, grp = lane / lpr;
  const int d0 = sub * VEC;
  float* s_dc = s_dc_all + warp * Htot;
  float wA[VEC], wB[VEC], accA[VEC], accB[VEC];
#pragma unroll
  for (int e = 0; e < VEC; ++e) {
    wA[e] = wB[e] = accA[e] = accB[e] = 0.f;
#pragma unroll
    for (int o8 = 0; o8 < 4; ++o8) {
      wA[e] += __ldg(Wg + o8 * Dh + d0 + e);
      wB[e] += __ldg(Wg + (o8 + 4) * Dh + d0 + e);
    }
  }
  float dbA = 0.f, dbB = 0.f;
  const long long total = (long long)B * Tn * Htot;
  const long long wid = (long long)blockIdx.x * nwarp + warp, nw = (long long)gridDim.x * nwarp;
  for (long long base = wid * rpw; base < total; base += nw * rpw) {
    const long long idx = base + grp;
    if (idx < total) {
      const int h = (int)(idx % Htot);
      const long long bt = idx / Htot;
      const int t = (int)(bt % Tn), b = (int)(bt / Tn);
      const long long o = ((long long)b * Htot + h) * Tn + t;
      const float dg = dgate[o], ga = ga_in[o], gb = gb_in[o], c = cst[h];
      const float dga = dg * (gb * c - 1.f), dgb = dg * ga * c;   // gate = ga * (gb * c - 1) + 2
      const float dpa = dga * ga * (1.f - ga), dpb = dgb * gb * (1.f - gb);
      if (sub == 0) {
        s_dc[h] += dg * ga * gb;    // the row groups of one pass hold distinct heads (host check): no race
        dbA += dpa;
        dbB += dpb;
      }
      const long long xo = bt * ((long long)Htot * Dh) + (long long)h * Dh + d0;
      float xv[VEC];
      load_vec(x + xo, xv);
#pragma unroll
      for (int e = 0; e < VEC; ++e) {
        accA[e] = fmaf(dpa, xv[e], accA[e]);
        accB[e] = fmaf(dpb, xv[e], accB[e]);
      }
      if (dx) {
        float gv[VEC];
        load_vec(dx + xo, gv);
#pragma unroll
        for (int e = 0; e < VEC; ++e) gv[e] += dpa * wA[e] + dpb * wB[e];
        store_vec(dx + xo, gv);
      }
    }
    __syncwarp();
  }
  // combine the row groups of the warp (lanes with equal `sub`), then one shared-memory add per element and warp
  for (int off = lpr; off < 32; off <<= 1) {
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
      accA[e] += __shfl_xor_sync(0xffffffffu, accA[e], off);
      accB[e] += __shfl_xor_sync(0xffffffffu, accB[e], off);
    }
    dbA += __shfl_xor_sync(0xffffffffu, dbA, off);
    dbB += __shfl_xor_sync(0xffffffffu, dbB, off);
  }
  if (grp == 0) {
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
      atomicAdd(&s_A[d0 + e], accA[e]);
      atomicAdd(&s_B[d0 + e], accB[e]);
    }
    if (sub == 0) {
      atomicAdd(&s_db[0], dbA);
      atomicAdd(&s_db[1], dbB);
    }
  }
  __syncthreads();
  for (int q = threadIdx.x; q < 8 * Dh; q += blockDim.x) atomicAdd(dWg + q, (q / Dh) < 4 ? s_A[q % Dh] : s_B[q % Dh]);
  for (int q = threadIdx.x; q < 8; q += blockDim.x) atomicAdd(dbg + q, q < 4 ? s_db[0] : s_db[1]);
  for (int q = threadIdx.x; q < Htot; q += blockDim.x) {
    float tsum = 0.f;
    for (int w8 = 0; w8 < nwarp; ++w8) tsum += s_dc_all[w8 * Htot + q];
    atomicAdd(dcst + q, tsum);
  }
}

}  // namespace dph

using namespace dph;

extern "C" int dph_softmax_fwd(const float* S, int64_t lds, void* P, void* Pd, int64_t ldp, const int64_t* lengths, int B,
                               int H, int T, float drop_p, uint64_t seed, const int* rel_table, const float* emb,
                               const float* gate, const int* head_index, int Htot, int dtype, void* stream) {
  if (!S || !P || B <= 0 || H <= 0 || T <= 0 || T > 32 * SM_NPL || ldp < T || lds < T) {
    set_error("softmax_fwd: bad arguments (T=%d, max %d)", T, 32 * SM_NPL);
    return DPH_ERR_ARG;
  }
  WavlmBias wb{rel_table, emb, gate, head_index, Htot};
  const long long nrows = (long long)B * H * T;
  const int grid = (int)((nrows + 7) / 8);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (dtype == DPH_DTYPE_BF16)
    softmax_fwd_kernel<bf16><<<grid, 256, 0, s>>>(S, lds, (bf16*)P, (bf16*)Pd, ldp, (const long long*)lengths, B, H, T,
                                                  drop_p, seed, wb);
  else
    softmax_fwd_kernel<float><<<grid, 256, 0, s>>>(S, lds, (float*)P, (float*)Pd, ldp, (const long long*)lengths, B, H,
                                                   T, drop_p, seed, wb);
  dph_count_launch();
  return dph_check_last("softmax_fwd");
}

extern "C" int dph_softmax_bwd(const void* P, const float* dPd, int64_t lds, void* dS, int64_t ldp, int B, int H, int T,
                               float drop_p, uint64_t seed, const int* rel_table, const float* emb, const float* gate,
                               const int* head_index, int Htot, float* dgate, float* demb, int num_buckets, int dtype,
                               void* stream) {
  if (!P || !dPd || !dS || B <= 0 || H <= 0 || T <= 0 || T > 32 * SM_NPL) {
    set_error("softmax_bwd: bad arguments");
    return DPH_ERR_ARG;
  }
  WavlmBias wb{rel_table, emb, gate, head_index, Htot};
  // 32 query rows of one (b, h) per block (the bias-embedding histogram is flushed once per block)
  const int rpb = 32;
  const dim3 grid((T + rpb - 1) / rpb, B * H);
  const size_t smem = demb ? sizeof(float) * num_buckets : 0;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (dtype == DPH_DTYPE_BF16)
    softmax_bwd_kernel<bf16><<<grid, 256, smem, s>>>((const bf16*)P, dPd, lds, (bf16*)dS, ldp, B, H, T, drop_p, seed, wb,
                                                     dgate, demb, num_buckets, rpb);
  else
    softmax_bwd_kernel<float><<<grid, 256, smem, s>>>((const float*)P, dPd, lds, (float*)dS, ldp, B, H, T, drop_p, seed,
                                                      wb, dgate, demb, num_buckets, rpb);
  dph_count_launch();
  return dph_check_last("softmax_bwd");
}

extern "C" int dph_wavlm_gate_fwd(const void* x, const float* Wg, const float* bg, const float* cst, float* gate,
                                  float* ga, float* gb, int B, int T, int Htot, int Dh, int dtype, void* stream) {
  const long long total = (long long)B * T * Htot;
  const int grid = (int)((total + 255) / 256);
  const size_t smem = sizeof(float) * 2 * (size_t)Dh;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (dtype == DPH_DTYPE_BF16)
    wavlm_gate_fwd_kernel<bf16><<<grid, 256, smem, s>>>((const bf16*)x, Wg, bg, cst, gate, ga, gb, B, T, Htot, Dh);
  else
    wavlm_gate_fwd_kernel<float><<<grid, 256, smem, s>>>((const float*)x, Wg, bg, cst, gate, ga, gb, B, T, Htot, Dh);
  dph_count_launch();
  return dph_check_last("wavlm_gate_fwd");
}

extern "C" int dph_wavlm_gate_bwd(const void* x, const float* Wg, const float* cst, const float* ga, const float* gb,
                                  const float* dgate, void* dx, float* dWg, float* dbg, float* dcst, int B, int T,
                                  int Htot, int Dh, int dtype, void* stream) {
  if (Dh <= 0 || Dh > 128) {
    set_error("wavlm_gate_bwd: head dim %d not supported (<= 128)", Dh);
    return DPH_ERR_UNSUPPORTED;
  }
  const size_t smem = sizeof(float) * (2 * Dh + 2 + 8 * Htot);   // 8 warps per block, one dconst slice each
  const long long rows = (long long)B * T * Htot;
  const int grid = (int)(rows / 8 < 148 * 4 ? (rows + 7) / 8 : 148 * 4);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int vec = dtype == DPH_DTYPE_BF16 ? 8 : 4;
  // (the rows of one warp-pass are consecutive (b, t, h): distinct heads as long as Htot >= rows per pass)
  const bool fast = (Dh % vec == 0) && (32 % (Dh / vec) == 0) && (Dh / vec <= 32) && (Htot >= 32 / (Dh / vec)) &&
                    ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dx)) & 15) == 0;
  if (fast) {
    const int rpw = 32 / (Dh / vec);
    const long long passes = (rows + rpw - 1) / rpw;
    // two blocks per SM: every block ends with ~8 * Dh global atomics onto the same addresses, keep them few
    const int g2 = (int)(passes / 8 < 148 * 2 ? (passes + 7) / 8 : 148 * 2);
    if (dtype == DPH_DTYPE_BF16)
      wavlm_gate_bwd_vec_kernel<bf16><<<g2, 256, smem, s>>>((const bf16*)x, Wg, cst, ga, gb, dgate, (bf16*)dx, dWg, dbg,
                                                            dcst, B, T, Htot, Dh);
    else
      wavlm_gate_bwd_vec_kernel<float><<<g2, 256, smem, s>>>((const float*)x, Wg, cst, ga, gb, dgate, (float*)dx, dWg,
                                                             dbg, dcst, B, T, Htot, Dh);
  } else if (dtype == DPH_DTYPE_BF16)
    wavlm_gate_bwd_kernel<bf16><<<grid, 256, smem, s>>>((const bf16*)x, Wg, cst, ga, gb, dgate, (bf16*)dx, dWg, dbg, dcst,
                                                        B, T, Htot, Dh);
  else
    wavlm_gate_bwd_kernel<float><<<grid, 256, smem, s>>>((const float*)x, Wg, cst, ga, gb, dgate, (float*)dx, dWg, dbg,
                                                         dcst, B, T, Htot, Dh);
  dph_count_launch();
  return dph_check_last("wavlm_gate_bwd");
}
