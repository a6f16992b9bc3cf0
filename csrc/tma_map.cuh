// Host-side helper shared by the TMA-fed grouped GEMMs (fedb_gemm_tma.cu, fedb_gemm_tma2.cu, fedb_gemm_tma2p.cu): the 3-D bf16
// tensor map {inner, rows, groups} with box {64, box_rows, 1} and the 128-byte swizzle the UMMA panels use.
// cuTensorMapEncodeTiled is resolved at run time (cudaGetDriverEntryPoint): the library does not link libcuda.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace fedb {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn tma_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

inline int make_tma_map_bf16(CUtensorMap* m, const void* base, int64_t inner, int64_t rows, int64_t groups, int64_t ld,
                             int64_t group_stride, int box_rows) {
  EncodeTiledFn enc = tma_encode_fn();
  if (!enc) return set_err(FEDB_E_UNSUPPORTED, "cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t dims[3] = {(cuuint64_t)inner, (cuuint64_t)rows, (cuuint64_t)groups};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)(groups > 1 ? group_stride : ld * rows) * 2};
  cuuint32_t box[3] = {64, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    snprintf(g_err, sizeof(g_err), "cuTensorMapEncodeTiled failed with CUresult %d (inner %lld rows %lld ld %lld)", (int)r, (long long)inner,
             (long long)rows, (long long)ld);
    return FEDB_E_CUDA;
  }
  return FEDB_OK;
}

// the two operand maps of a FedbGemm: K-major sources are boxed {64 k, 128 rows}, MN-major ones {64 rows-of-MN, 64 k}
template <typename Gemm>
inline int make_gemm_maps(const Gemm& p, int n_groups, CUtensorMap* tmA, CUtensorMap* tmB) {
  if (int rc = p.a_mn ? make_tma_map_bf16(tmA, p.A, p.M, p.K, n_groups, p.lda, p.a_group_stride, 64)
                      : make_tma_map_bf16(tmA, p.A, p.K, p.M, n_groups, p.lda, p.a_group_stride, 128)) return rc;
  if (int rc = p.b_mn ? make_tma_map_bf16(tmB, p.B, p.N, p.K, n_groups, p.ldb, p.b_group_stride, 64)
                      : make_tma_map_bf16(tmB, p.B, p.K, p.N, n_groups, p.ldb, p.b_group_stride, 128)) return rc;
  return FEDB_OK;
}

}  // namespace fedb
