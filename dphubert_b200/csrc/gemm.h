// Internal C++ interface of the tcgen05 GEMM (shared by the C-ABI wrappers in abi_*.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/dphubert_b200.h"

namespace dph {

// Device-side parameter block (passed by value as a kernel argument).
struct GemmParams {
  int M, N, K;  // per-batch valid rows / cols / reduction length
  int nb0, nb1;
  int mtiles, ntiles, ksplit, num_kb, kb_per_split;
  int a_kdiv, b_kdiv;
  int a_b0, a_b1, b_b0, b_b1;
  // output
  void* C;
  void* C2;
  long long ldc, c_sb0, c_sb1;
  int c_dtype;  // 0 bf16, 1 fp32
  int c_mode;   // 0 store, 1 atomic add (fp32 only), 2 read-add-store
  int vec_ok;   // 16-byte vector stores/loads allowed
  float alpha;
  const float* bias;
  int act;
  const float* colmask;
  const float* scalar_mask;
  const void* residual;
  long long ldr, r_sb0, r_sb1;
  int r_dtype;
  float drop_p;
  unsigned long long seed;
  long long vec_sb0, scalar_sb0;
  // fused activation backward (EPI_BWD): saved pre-activation tile + column-sum outputs
  const void* Z;
  long long ldz, z_sb0, z_sb1;
  float* dbias;
  float* dmask;
  int b_nsplit;  // > 0: MN-major B whose N index is (tap, channel): see the producer warp in gemm_kernel.cuh
};

int gemm_bf16_launch(const dph_gemm_desc_t* d, cudaStream_t stream);

}  // namespace dph
