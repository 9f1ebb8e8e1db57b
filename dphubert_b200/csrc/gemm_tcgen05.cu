// tcgen05 GEMM, host side: tensor-map construction, tile / split / CTA-group choice, epilogue-class dispatch; plus the
// out-of-line generic epilogue shared by every kernel instance. The kernel itself lives in gemm_kernel.cuh.
#include "gemm_kernel.cuh"

namespace dph {

__device__ unsigned long long* g_gemm_timeline = nullptr;

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

// General bf16 tensor map (rank <= 5; strides in bytes for dims 1..rank-1; swizzle: 0 none, 1 32B, 2 64B, 3 128B).
int make_tmap_nd_bf16(CUtensorMap* out, const void* ptr, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box, int swizzle) {
  PFN_encodeTiled enc = get_encode_fn();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled unavailable");
    return DPH_ERR_CUDA;
  }
  cuuint64_t d[5], st[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) {
    d[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
    if (i > 0) st[i - 1] = strides_bytes[i - 1];
  }
  const CUtensorMapSwizzle sw = swizzle == 3 ? CU_TENSOR_MAP_SWIZZLE_128B
                                : swizzle == 2 ? CU_TENSOR_MAP_SWIZZLE_64B
                                : swizzle == 1 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(ptr), d, st, bx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(rank %d) failed: %d", rank, (int)r);
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
  // the fast epilogue uses 32-byte accesses: leading dimensions / batch strides / pointers must keep that alignment
  const int celt = d->c_dtype == 0 ? 16 : 8;
  bool vec = (d->ldc % celt == 0) && (d->c_sb0 % celt == 0) && (d->c_sb1 % celt == 0) &&
             ((reinterpret_cast<uintptr_t>(d->c) & 31) == 0);
  if (d->c2) vec = vec && ((reinterpret_cast<uintptr_t>(d->c2) & 31) == 0);
  if (d->residual) {
    const int relt = d->r_dtype == 0 ? 16 : 8;
    vec = vec && (d->ldr % relt == 0) && (d->r_sb0 % relt == 0) && (d->r_sb1 % relt == 0) &&
          ((reinterpret_cast<uintptr_t>(d->residual) & 31) == 0);
  }
  if (d->bias) vec = vec && ((reinterpret_cast<uintptr_t>(d->bias) & 15) == 0) && (d->vec_sb0 % 4 == 0);
  if (d->colmask) vec = vec && ((reinterpret_cast<uintptr_t>(d->colmask) & 15) == 0) && (d->vec_sb0 % 4 == 0);
  p.vec_ok = vec ? 1 : 0;
  p.Z = d->z;
  p.ldz = d->ldz;
  p.z_sb0 = d->z_sb0;
  p.z_sb1 = d->z_sb1;
  p.dbias = d->dbias;
  p.dmask = d->dmask;
  p.b_nsplit = d->b_nsplit;
  if (d->b_nsplit != 0) {
    if (!b_mn || d->b_nsplit < 64 || d->b_nsplit % 64 != 0 || d->N % d->b_nsplit != 0 || d->nb1 != 1 ||
        (long long)d->b.dims[0] != d->b_nsplit || (long long)d->b.dims[3] * d->b_nsplit != d->N) {
      set_error("gemm: b_nsplit needs an MN-major B with dims[0] == b_nsplit (a multiple of 64), dims[3] == N / "
                "b_nsplit taps and nb1 == 1");
      return DPH_ERR_ARG;
    }
  }
  if (d->act_bwd) {
    // fused activation backward: fast epilogue path only, nothing else may be fused into the same launch
    const bool z_ok = d->z != nullptr && (d->ldz % 16 == 0) && (d->z_sb0 % 16 == 0) && (d->z_sb1 % 16 == 0) &&
                      ((reinterpret_cast<uintptr_t>(d->z) & 31) == 0);
    if (d->c_dtype != DPH_DTYPE_BF16 || !vec || !z_ok || d->N % 32 != 0 || d->c_mode != 0 || ksplit != 1 || d->bias ||
        d->residual || d->c2 || d->scalar_mask || d->alpha != 1.0f || (d->act != 0 && d->act != 1)) {
      set_error("gemm: act_bwd needs a bf16 output, N %% 32 == 0 (N=%d), 32-byte aligned c / z rows, a pre-activation "
                "tensor, and no bias / residual / c2 / scalar mask / alpha / accumulation in the same launch", d->N);
      return DPH_ERR_UNSUPPORTED;
    }
  }

  CUtensorMap ta, tb;
  int rc = make_tmap(&ta, d->a, BM);
  if (rc != DPH_OK) return rc;
  rc = make_tmap(&tb, d->b, bn / cg);
  if (rc != DPH_OK) return rc;

  const long long total = (long long)p.nb1 * p.nb0 * p.mtiles * p.ntiles * p.ksplit;
  const int grid = (int)(total < units ? total : units) * cg;

  // smallest epilogue class that covers what this launch asks for
  int need = 0;
  if (d->bias) need |= F_BIAS;
  if (d->act != 0) need |= F_GELU;
  if (d->colmask) need |= F_COLMASK;
  if (d->scalar_mask) need |= F_SMASK;
  if (d->drop_p > 0.0f) need |= F_DROP;
  if (d->residual) need |= F_RES;
  if (d->c2) need |= F_C2;
  if (p.c_mode != 0) need |= F_ACCUM;
  if (d->alpha != 1.0f) need |= F_ALPHA;
  const bool kk = !a_mn && !b_mn;
  if (d->act_bwd) return gemm_launch_bwd(bn, cg, ta, tb, p, grid, a_mn, b_mn, stream);
  if ((need & ~EPI_LIN) == 0) return gemm_launch_lin(bn, cg, ta, tb, p, grid, a_mn, b_mn, stream);
  if ((need & ~EPI_ACC) == 0) return gemm_launch_acc(bn, cg, ta, tb, p, grid, a_mn, b_mn, stream);
  if (kk && (need & ~EPI_ACT) == 0) return gemm_launch_act(bn, cg, ta, tb, p, grid, a_mn, b_mn, stream);
  if (kk && (need & ~EPI_DROP) == 0) return gemm_launch_drop(bn, cg, ta, tb, p, grid, a_mn, b_mn, stream);
  return gemm_launch_all(bn, cg, ta, tb, p, grid, a_mn, b_mn, stream);
}

}  // namespace dph

// Debug hook (not part of the drop-in surface): per-CTA event timeline of subsequent GEMM launches, 64 u64 per CTA.
extern "C" int dph_gemm_debug_timeline(void* buf) {
  unsigned long long* pbuf = reinterpret_cast<unsigned long long*>(buf);
  cudaError_t e = cudaMemcpyToSymbol(dph::g_gemm_timeline, &pbuf, sizeof(pbuf));
  return e == cudaSuccess ? DPH_OK : DPH_ERR_CUDA;
}

extern "C" int dph_gemm_bf16(const dph_gemm_desc_t* desc, void* stream) {
  return dph::gemm_bf16_launch(desc, reinterpret_cast<cudaStream_t>(stream));
}
