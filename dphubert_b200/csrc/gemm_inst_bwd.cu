// tcgen05 GEMM kernels of epilogue class EPI_BWD (fused activation backward, see epilogue_chunk_bwd in
// gemm_kernel.cuh): dgrad products whose epilogue applies act'(z) * column mask * dropout and reduces the bias /
// mask gradients, replacing a separate elementwise pass over the (rows, N) gradient.
#include "gemm_kernel.cuh"

namespace dph {

int gemm_launch_bwd(int bn, int cg, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid,
                    bool a_mn, bool b_mn, cudaStream_t s) {
  return launch_class<EPI_BWD, true>(bn, cg, ta, tb, p, grid, a_mn, b_mn, s);
}

}  // namespace dph
