// In-place gradient all-reduce (mean) over NVSwitch multicast memory — the data-parallel exchange of the training
// step (reference: distill.py:41 strategy="ddp", i.e. DDP's NCCL all-reduce of every gradient bucket).
//
// The gradient arena lives in symmetric memory (the same allocation at the same offset on every rank, bound to one
// multicast object).  Rank r owns the r-th 1/world of a slice: one `multimem.ld_reduce` returns the SUM over all ranks
// of 16 bytes — the reduction happens inside the switch, each rank pulls its share once — and one `multimem.st` writes
// the mean back to every rank.  Per rank the slice crosses NVLink once in each direction and the kernel is a plain
// streaming loop, so a handful of CTAs saturate the links; NCCL's ring / tree kernels hold 16-32 CTAs for the same
// transfer while the dgrad GEMMs of the backward pass want all 148 SMs.
//
// Ordering across ranks (every rank's slice complete before anyone reduces, every mean stored before anyone reads) is
// the caller's: a symmetric-memory barrier on the same stream before and after this launch.
#include "common.cuh"
#include "../../include/dphubert_b200.h"

namespace dph {
void set_error(const char* fmt, ...);

DPH_DEVICE float4 multimem_ld_reduce_add(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
DPH_DEVICE void multimem_st(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}

// mc: multicast address of the slice (16-byte aligned); n4: its length in float4; [q0, q1): this rank's share
__global__ void __launch_bounds__(512) multimem_allreduce_kernel(float* __restrict__ mc, long long q0, long long q1,
                                                                  float scale) {
  pdl_wait();
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long i = q0 + (long long)blockIdx.x * blockDim.x + threadIdx.x;
  // eight independent 16-byte reductions in flight per thread: a switch-side reduction is a ~2 us round trip, so the
  // bytes in flight (CTAs x 512 threads x 128 B) set the bandwidth — measured on 2 GPUs with four in flight: 16 CTAs
  // reached 360 GB/s where NCCL's copy kernels did 520
  constexpr int U = 8;
  for (; i + (U - 1) * stride < q1; i += U * stride) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = multimem_ld_reduce_add(mc + 4 * (i + u * stride));
#pragma unroll
    for (int u = 0; u < U; ++u) {
      v[u].x *= scale; v[u].y *= scale; v[u].z *= scale; v[u].w *= scale;
      multimem_st(mc + 4 * (i + u * stride), v[u]);
    }
  }
  for (; i < q1; i += stride) {
    float4 v = multimem_ld_reduce_add(mc + 4 * i);
    v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
    multimem_st(mc + 4 * i, v);
  }
}

}  // namespace dph

using namespace dph;

extern "C" int dph_multimem_allreduce(void* mc_ptr, int64_t n, int rank, int world, float scale, int ctas, void* stream) {
  if (!mc_ptr || n <= 0 || world <= 0 || rank < 0 || rank >= world || (n & 3) != 0 ||
      (reinterpret_cast<uintptr_t>(mc_ptr) & 15) != 0) {
    set_error("multimem_allreduce: bad arguments (n=%lld must be a multiple of 4 floats, the address 16-byte aligned)",
              (long long)n);
    return DPH_ERR_ARG;
  }
  const long long n4 = n / 4;
  const long long per = (n4 + world - 1) / world;
  const long long q0 = per * rank < n4 ? per * rank : n4;
  const long long q1 = q0 + per < n4 ? q0 + per : n4;
  if (q1 <= q0) return DPH_OK;   // (every rank still takes part in the caller's barriers)
  if (ctas <= 0) ctas = 16;
  long long need = (q1 - q0 + 511) / 512;
  if (need < ctas) ctas = (int)need;
  cudaError_t e = launch_pdl(multimem_allreduce_kernel, dim3(ctas), dim3(512), 0, reinterpret_cast<cudaStream_t>(stream),
                             reinterpret_cast<float*>(mc_ptr), q0, q1, scale);
  if (e != cudaSuccess) {
    set_error("multimem_allreduce launch: %s", cudaGetErrorString(e));
    return DPH_ERR_CUDA;
  }
  dph_count_launch();
  return dph_check_last("multimem_allreduce");
}
