// tcgen05 GEMM kernels of epilogue class EPI_ALL (see gemm_kernel.cuh): one translation unit per class so that the
// classes compile in parallel.
#include "gemm_kernel.cuh"

namespace dph {

int gemm_launch_all(int bn, int cg, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid,
                   bool a_mn, bool b_mn, cudaStream_t s) {
  return launch_class<EPI_ALL, true>(bn, cg, ta, tb, p, grid, a_mn, b_mn, s);
}

}  // namespace dph
