// Common device helpers for the dphubert_b200 kernel library (sm_100a only).
// Thin inline-PTX wrappers for mbarrier / TMA / tcgen05 / TMEM plus small math utilities.
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#define DPH_OK 0
#define DPH_ERR_ARG (-1)
#define DPH_ERR_CUDA (-2)
#define DPH_ERR_UNSUPPORTED (-3)

#define DPH_DEVICE __device__ __forceinline__

typedef __nv_bfloat16 bf16;

// Launch counter (host side, bumped by every C-ABI entry that launches a kernel).
extern "C" long long dph_launch_count(void);
void dph_count_launch(int n = 1);
int dph_check_last(const char* what);

extern __device__ unsigned long long dph_rng_epoch;

// Programmatic dependent launch (PDL): 1 unless the environment says DPH_PDL=0 (read once).
bool dph_pdl_enabled();

namespace dph {

// ---------------------------------------------------------------------------------------------
// Programmatic dependent launch.  A kernel launched through launch_pdl() may become resident while its stream
// predecessor is still running (as soon as every CTA of the predecessor has executed pdl_trigger() or exited); it
// must execute pdl_wait() before it touches global memory that any earlier kernel reads or writes.  pdl_wait()
// returns when the predecessor grid has completed and its writes are visible, so chains stay transitively ordered.
// Kernels that use tensor memory trigger only after their own allocation succeeded (a dependent spinning in
// pdl_wait() while holding TMEM can then never starve the grid it waits for).
// ---------------------------------------------------------------------------------------------
DPH_DEVICE void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
DPH_DEVICE void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = dph_pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

constexpr int kNumSMsDefault = 148;

// ---------------------------------------------------------------------------------------------
// Math
// ---------------------------------------------------------------------------------------------
DPH_DEVICE float gelu_erf(float x) {
  // exact (erf) GELU, matches torch.nn.functional.gelu default
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}
DPH_DEVICE float gelu_erf_grad(float x) {
  const float kInvSqrt2Pi = 0.39894228040143267794f;
  float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752440f));
  float pdf = kInvSqrt2Pi * __expf(-0.5f * x * x);
  return cdf + x * pdf;
}
// Branch-free erf-GELU for the tensor-core GEMM epilogue (bf16 outputs): Abramowitz-Stegun 7.1.26 (|erf error| <=
// 1.5e-7) arranged so that neither tail cancels: with z = |x|/sqrt(2), t = 1/(1 + p z) and
// q = poly(t) exp(-z^2) = 1 - erf(z):   gelu(x) = max(x, 0) - 0.5 |x| q = max(x, 0) - z * (poly(t)/sqrt(2)) * exp(-z^2).
// Raw MUFU approximations (rcp / ex2, ~1 ulp): the arguments are >= 1 resp. <= 0, no range fix-ups needed.
// 16 instructions per element, 2 of them MUFU.
DPH_DEVICE float gelu_erf_fast(float x) {
  const float z = fabsf(x) * 0.70710678118654752440f;
  float t, e;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, z, 1.0f)));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(z * (-1.4426950408889634f * z)));
  // A&S coefficients divided by sqrt(2)
  float poly = fmaf(0.75052688f, t, -1.02753344f);
  poly = fmaf(poly, t, 1.00509115f);
  poly = fmaf(poly, t, -0.20116957f);
  poly = fmaf(poly, t, 0.18019173f);
  poly *= t;
  return fmaxf(x, 0.0f) - z * (poly * e);
}
// gelu(x) and d gelu / dx from one rcp + one ex2 (same A&S erfc polynomial as gelu_erf_fast; exp(-x^2/2) is shared
// between the erfc tail and the normal density).  |error| ~3e-7 on either output.
DPH_DEVICE void gelu_and_grad_fast(float x, float& g, float& dg) {
  const float z = fabsf(x) * 0.70710678118654752440f;
  float t, e;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, z, 1.0f)));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(z * (-1.4426950408889634f * z)));
  float poly = fmaf(1.061405429f, t, -1.453152027f);
  poly = fmaf(poly, t, 1.421413741f);
  poly = fmaf(poly, t, -0.284496736f);
  poly = fmaf(poly, t, 0.254829592f);
  const float hq = 0.5f * poly * t * e;            // 0.5 * erfc(z) = Phi(-|x|)
  const float cdf = x >= 0.0f ? 1.0f - hq : hq;    // Phi(x)
  g = fmaxf(x, 0.0f) - fabsf(x) * hq;
  dg = fmaf(x * 0.39894228040143267794f, e, cdf);
}
DPH_DEVICE float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

// ---- packed fp32x2 arithmetic (Blackwell FFMA2 / FMUL2 / FADD2: two fp32 lanes per issue slot) -----------------
typedef unsigned long long f32x2;
DPH_DEVICE f32x2 pk2(float a, float b) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
DPH_DEVICE f32x2 bc2(float a) { return pk2(a, a); }
DPH_DEVICE void upk2(f32x2 r, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(r)); }
DPH_DEVICE f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
DPH_DEVICE f32x2 mul2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
DPH_DEVICE f32x2 add2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
DPH_DEVICE float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// GELU(u) = u Phi(u) with Phi(u) ~ 0.5 + 0.5 tanh(u (a + b u^2 + c u^4)), u^2 clamped at 64 (|Phi error| < 2e-5 before
// MUFU.TANH's own 2^-11: below bf16 resolution).  Used on the bf16 activation path only; fp32 keeps the erf forms.
constexpr float kGeluTA = 0.7977176713187712f, kGeluTB = 0.03679851046109153f, kGeluTC = -0.00031586434495541917f;
// two lanes at once: g = GELU(u), dg = GELU'(u) = Phi + u Phi',  Phi' = 0.5 (1 - tanh^2)(a + 3 b u^2 + 5 c u^4)
DPH_DEVICE void gelu_and_grad_tanh2(f32x2 u, f32x2& g, f32x2& dg) {
  float s0, s1;
  upk2(mul2(u, u), s0, s1);
  const f32x2 s = pk2(fminf(s0, 64.f), fminf(s1, 64.f));
  f32x2 q = fma2(s, bc2(kGeluTC), bc2(kGeluTB));
  q = fma2(q, s, bc2(kGeluTA));
  float a0, a1;
  upk2(mul2(q, u), a0, a1);
  const f32x2 th = pk2(tanh_approx(a0), tanh_approx(a1));
  const f32x2 h = fma2(th, bc2(0.5f), bc2(0.5f));
  g = mul2(u, h);
  f32x2 qp = fma2(s, bc2(5.f * kGeluTC), bc2(3.f * kGeluTB));
  qp = fma2(qp, s, bc2(kGeluTA));
  const f32x2 t2 = fma2(mul2(th, th), bc2(-1.f), bc2(1.f));  // 1 - tanh^2
  const f32x2 w = mul2(mul2(t2, qp), u);
  dg = fma2(w, bc2(0.5f), h);
}
DPH_DEVICE f32x2 gelu_tanh2(f32x2 u) {
  float s0, s1;
  upk2(mul2(u, u), s0, s1);
  const f32x2 s = pk2(fminf(s0, 64.f), fminf(s1, 64.f));
  f32x2 q = fma2(s, bc2(kGeluTC), bc2(kGeluTB));
  q = fma2(q, s, bc2(kGeluTA));
  float a0, a1;
  upk2(mul2(q, u), a0, a1);
  const f32x2 th = pk2(tanh_approx(a0), tanh_approx(a1));
  return mul2(u, fma2(th, bc2(0.5f), bc2(0.5f)));
}

DPH_DEVICE float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
DPH_DEVICE float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum; `red` must hold >= 32 floats of shared memory. All threads get the result.
DPH_DEVICE float block_sum(float v, float* red) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  int nw = (blockDim.x + 31) >> 5;
  float r = (lane < nw) ? red[lane] : 0.0f;
  r = warp_sum(r);
  return r;
}

// Counter-based RNG for dropout: one 32-bit hash per (seed, element index). Stateless so the
// backward pass can regenerate the identical keep mask.
// `dph_rng_epoch` lives in device memory and is advanced by a one-thread kernel (dph_rng_advance) once per
// training step, so a step captured in a CUDA graph (kernel arguments frozen) still draws fresh masks every replay.
// 32-bit arithmetic only (64-bit multiplies cost ~4 instructions each on the integer pipe): the seed, the epoch and
// the high half of the index fold into a key, the low half goes through the murmur3 finaliser.
// The (seed, epoch) part of the key: hoist it out of inner loops (it costs a global load and four multiplies).
DPH_DEVICE uint32_t dropout_key(uint64_t seed) {
  const uint64_t ep = dph_rng_epoch;
  return (uint32_t)seed ^ ((uint32_t)(seed >> 32) * 0x85EBCA6Bu) ^ ((uint32_t)ep * 0xC2B2AE35u) ^
         ((uint32_t)(ep >> 32) * 0x27D4EB2Fu);
}
DPH_DEVICE uint32_t hash_u32_keyed(uint32_t key, uint64_t idx) {
  key ^= (uint32_t)(idx >> 32) * 0x165667B1u;
  uint32_t x = (uint32_t)idx * 0x9E3779B1u + key;
  x ^= x >> 16;
  x *= 0x85EBCA6Bu;
  x ^= x >> 13;
  x *= 0xC2B2AE35u;
  x ^= x >> 16;
  return x;
}
DPH_DEVICE uint32_t hash_u32(uint64_t seed, uint64_t idx) { return hash_u32_keyed(dropout_key(seed), idx); }
// keep-probability test: returns scale (1/(1-p)) or 0.  One 32-bit hash decides TWO neighbouring elements (index
// pair 2q, 2q+1: low / high 16 bits against the top 16 bits of the threshold, i.e. p quantised to 1/65536), so
// vectorised callers whose base index is even pay for half a hash per element (dropout_scale2).
DPH_DEVICE float dropout_scale(uint64_t seed, uint64_t idx, uint32_t thresh, float inv_keep) {
  const uint32_t h = hash_u32(seed, idx >> 1);
  const uint32_t bits = (idx & 1ull) ? (h >> 16) : (h & 0xFFFFu);
  return (bits >= (thresh >> 16)) ? inv_keep : 0.0f;
}
// idx_even must be even: scales of elements idx_even and idx_even + 1.
DPH_DEVICE void dropout_scale2_keyed(uint32_t key, uint64_t idx_even, uint32_t thresh, float inv_keep, float& s0,
                                     float& s1) {
  const uint32_t h = hash_u32_keyed(key, idx_even >> 1);
  const uint32_t t16 = thresh >> 16;
  s0 = ((h & 0xFFFFu) >= t16) ? inv_keep : 0.0f;
  s1 = ((h >> 16) >= t16) ? inv_keep : 0.0f;
}
DPH_DEVICE void dropout_scale2(uint64_t seed, uint64_t idx_even, uint32_t thresh, float inv_keep, float& s0,
                               float& s1) {
  dropout_scale2_keyed(dropout_key(seed), idx_even, thresh, inv_keep, s0, s1);
}

// Attention-probability dropout uses a 2-D index (row = (b H + h) T + i, key j): the row folds into the key once per
// thread and a key PAIR (2 jp, 2 jp + 1) then costs one 32-bit multiply-add plus the finaliser (no 64-bit arithmetic
// in the inner loop).  The fused attention kernels and the unfused softmax kernels share these two functions.
DPH_DEVICE uint32_t dropout_row_key(uint32_t key, uint64_t row) {
  return key ^ ((uint32_t)row * 0x165667B1u) ^ ((uint32_t)(row >> 32) * 0x27D4EB2Fu);
}
DPH_DEVICE uint32_t hash_row_pair(uint32_t row_key, uint32_t jp) {
  uint32_t x = jp * 0x9E3779B1u + row_key;
  x ^= x >> 16;
  x *= 0x85EBCA6Bu;
  x ^= x >> 13;
  x *= 0xC2B2AE35u;
  x ^= x >> 16;
  return x;
}
// scales of keys 2 jp and 2 jp + 1 of a row
DPH_DEVICE void dropout_row_pair(uint32_t row_key, uint32_t jp, uint32_t thresh, float inv_keep, float& s0, float& s1) {
  const uint32_t h = hash_row_pair(row_key, jp);
  const uint32_t t16 = thresh >> 16;
  s0 = ((h & 0xFFFFu) >= t16) ? inv_keep : 0.0f;
  s1 = ((h >> 16) >= t16) ? inv_keep : 0.0f;
}
DPH_DEVICE float dropout_row_elem(uint32_t row_key, uint32_t j, uint32_t thresh, float inv_keep) {
  const uint32_t h = hash_row_pair(row_key, j >> 1);
  const uint32_t bits = (j & 1u) ? (h >> 16) : (h & 0xFFFFu);
  return (bits >= (thresh >> 16)) ? inv_keep : 0.0f;
}

// v[0..N) *= dropout scale of the N consecutive elements starting at idx.  Aligned pairs share one hash when idx is
// even (the usual case: even row strides); otherwise one hash per element.  Same decisions as dropout_scale().
template <int N>
DPH_DEVICE void dropout_apply_vec(uint64_t seed, uint64_t idx, uint32_t thresh, float inv_keep, float (&v)[N]) {
  const uint32_t key = dropout_key(seed);
  const uint32_t t16 = thresh >> 16;
  if ((N % 2) == 0 && (idx & 1ull) == 0ull) {
#pragma unroll
    for (int j = 0; j + 1 < N; j += 2) {
      const uint32_t h = hash_u32_keyed(key, (idx + j) >> 1);
      v[j] *= ((h & 0xFFFFu) >= t16) ? inv_keep : 0.0f;
      v[j + 1] *= ((h >> 16) >= t16) ? inv_keep : 0.0f;
    }
  } else {
#pragma unroll
    for (int j = 0; j < N; ++j) {
      const uint64_t e = idx + j;
      const uint32_t h = hash_u32_keyed(key, e >> 1);
      const uint32_t bits = (e & 1ull) ? (h >> 16) : (h & 0xFFFFu);
      v[j] *= (bits >= t16) ? inv_keep : 0.0f;
    }
  }
}

template <typename T> DPH_DEVICE float to_f(T v);
template <> DPH_DEVICE float to_f<float>(float v) { return v; }
template <> DPH_DEVICE float to_f<bf16>(bf16 v) { return __bfloat162float(v); }
template <typename T> DPH_DEVICE T from_f(float v);
template <> DPH_DEVICE float from_f<float>(float v) { return v; }
template <> DPH_DEVICE bf16 from_f<bf16>(float v) { return __float2bfloat16_rn(v); }

// 16-byte vector reduction (sm_90+): one L2 atomic transaction for four consecutive fp32 (16-byte aligned).
DPH_DEVICE void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

// 32-byte (256-bit, sm_100+) global accesses: one full L2 sector per thread. Address must be 32-byte aligned.
DPH_DEVICE void stg_256(void* addr, const uint32_t (&v)[8]) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(addr), "r"(v[0]), "r"(v[1]), "r"(v[2]),
               "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
               : "memory");
}
DPH_DEVICE void ldg_nc_256(const void* addr, uint32_t (&v)[8]) {
  asm volatile("ld.global.nc.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "l"(addr));
}

DPH_DEVICE uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

// ---------------------------------------------------------------------------------------------
// Shared-memory address / mbarrier
// ---------------------------------------------------------------------------------------------
DPH_DEVICE uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

DPH_DEVICE void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
DPH_DEVICE void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
DPH_DEVICE void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
DPH_DEVICE void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
DPH_DEVICE void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra WAIT_DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "WAIT_DONE:\n\t"
      "}" ::"r"(bar),
      "r"(parity)
      : "memory");
}

// ---- thread-block cluster helpers (CTA pairs for tcgen05 cta_group::2) ----
DPH_DEVICE uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
DPH_DEVICE void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of the same shared-memory offset in CTA `rank` of this cluster
DPH_DEVICE uint32_t mapa_shared(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
DPH_DEVICE void mbar_arrive_cluster(uint32_t cluster_addr) {
  // default semantics (release at CTA scope): a cluster-scope release would drain every outstanding global store
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}

DPH_DEVICE bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "elect.sync _|P1, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, P1;\n\t"
      "}"
      : "=r"(pred));
  return pred != 0;
}

// ---------------------------------------------------------------------------------------------
// TMA (cp.async.bulk.tensor) — 5-D tiled loads completing on an mbarrier
// ---------------------------------------------------------------------------------------------
DPH_DEVICE void tma_prefetch_desc(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
DPH_DEVICE void tma_load_5d(uint32_t smem_dst, const CUtensorMap* m, uint32_t bar, int c0, int c1, int c2, int c3,
                            int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(smem_dst),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}

// CTA-pair variant: the data lands in this CTA's shared memory, the transaction bytes complete on `bar_cluster`
// (a shared::cluster address, normally the leader CTA's barrier).
DPH_DEVICE void tma_load_5d_2sm(uint32_t smem_dst, const CUtensorMap* m, uint32_t bar_cluster, int c0, int c1, int c2,
                                int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(smem_dst),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}

// ---------------------------------------------------------------------------------------------
// tcgen05 / TMEM
// ---------------------------------------------------------------------------------------------
template <int CG> DPH_DEVICE void tmem_alloc_cg(uint32_t smem_dst, uint32_t ncols) {
  if constexpr (CG == 1)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols)
                 : "memory");
  else
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols)
                 : "memory");
}
template <int CG> DPH_DEVICE void tmem_relinquish_cg() {
  if constexpr (CG == 1)
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  else
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <int CG> DPH_DEVICE void tmem_dealloc_cg(uint32_t taddr, uint32_t ncols) {
  if constexpr (CG == 1)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
  else
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// cta_group::2 MMA: issued by the leader CTA of a pair; A/B descriptors address the same shared-memory offsets in
// both CTAs (A: 128 rows each, B: N/2 rows each), D is the same TMEM address in both.
DPH_DEVICE void umma_f16_ss_2sm(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrive (once) on the barrier at the same offset in every CTA of `cta_mask` when the MMAs issued so far complete.
DPH_DEVICE void umma_commit_2sm(uint32_t bar, uint16_t cta_mask) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
      "h"(cta_mask)
      : "memory");
}

DPH_DEVICE void tmem_alloc(uint32_t smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols)
               : "memory");
}
DPH_DEVICE void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
DPH_DEVICE void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
DPH_DEVICE void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
DPH_DEVICE void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem desc] * B[smem desc], kind::f16 (bf16/fp16 inputs, fp32 accumulate)
DPH_DEVICE void umma_f16_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// A operand from TMEM (used by fused attention: P stays on chip)
DPH_DEVICE void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrive on an mbarrier when all previously issued tcgen05.mma of this thread have completed.
DPH_DEVICE void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// 32 lanes x 32 columns of 32-bit: thread i of the warp receives row (lane base + i), columns c..c+31
DPH_DEVICE void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
DPH_DEVICE void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// 32 lanes x 32 columns store (registers -> TMEM), used by fused attention to stage P.
DPH_DEVICE void tmem_st_32x32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
DPH_DEVICE void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// Shared-memory matrix descriptor for tcgen05.mma (sm_100 format, version = 1).
//   bits [0,14)  start address >> 4        bits [16,30) leading byte offset >> 4
//   bits [32,46) stride byte offset >> 4   bits [46,48) version (1)
//   bits [61,64) layout type: 0 none, 2 = 128B swizzle, 4 = 64B, 6 = 32B
DPH_DEVICE uint64_t make_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout_type) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFFu);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)(layout_type & 7u) << 61;
  return d;
}

// Instruction descriptor for kind::f16 with bf16 A/B and fp32 accumulate.
//   bits [4,6) c format (1 = f32); [7,10) a format (1 = bf16); [10,13) b format; bit 15 a major (1 = MN);
//   bit 16 b major; [17,23) N >> 3; [24,29) M >> 4.
__host__ __device__ constexpr uint32_t make_idesc_bf16(int M, int N, bool a_mn, bool b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

}  // namespace dph
