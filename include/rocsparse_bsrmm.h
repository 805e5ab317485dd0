/*
 * rocsparse_bsrmm.h -- drop-in replacement for the reference header of the
 * same name (xuyifangreeneyes/spmm-denseblock rocsparse_bsrmm.h:102-256).
 *
 * rocsparse_bsrmm_template<T>() keeps the reference's signature, status codes
 * and check order (handle, descr, operations, sizes, quick return, pointers,
 * leading dimensions: rocsparse_bsrmm.h:110-176) and forwards to
 * spmmdb_bsrmm_f32 in libspmmdb.so.  C = alpha * A_bsr * op(B) + beta * C with
 * column-major C; op(B) = N: column-major B, op(B) = T: row-major B.
 *
 * Differences on purpose:
 *   - the result is the mathematical product; the reference kernels derive the
 *     output column from blockDim.y (rocsparse_bsrmm_impl.h:27,326,413) and leave
 *     C unwritten for n <= 256 -- that bug is not reproduced;
 *   - n == 1 and block_dim == 1 work (the reference asserts, :182-188);
 *   - SPMMDB_BSRMM_MATH (environment) or spmmdb_set_bsrmm_math() selects the
 *     arithmetic: fp32 CUDA cores (default, <= 1e-5), tf32 or bf16 tcgen05
 *     tensor cores (<= 1e-2).
 */
#ifndef SPMMDB_DROPIN_ROCSPARSE_BSRMM_H
#define SPMMDB_DROPIN_ROCSPARSE_BSRMM_H

#include <cusparse.h>

#include <type_traits>

#include "spmmdb.h"

inline cusparseStatus_t bsrmm_status_to_cusparse(int rc) {
  switch (rc) {
    case SPMMDB_OK: return CUSPARSE_STATUS_SUCCESS;
    case SPMMDB_ERR_INVALID_VALUE: return CUSPARSE_STATUS_INVALID_VALUE;
    case SPMMDB_ERR_NOT_SUPPORTED: return CUSPARSE_STATUS_MATRIX_TYPE_NOT_SUPPORTED;
    case SPMMDB_ERR_ALLOC: return CUSPARSE_STATUS_ALLOC_FAILED;
    case SPMMDB_ERR_ARCH: return CUSPARSE_STATUS_ARCH_MISMATCH;
    default: return CUSPARSE_STATUS_EXECUTION_FAILED;
  }
}

template <typename T>
cusparseStatus_t rocsparse_bsrmm_template(cusparseHandle_t handle, cusparseDirection_t dir,
                                          cusparseOperation_t trans_A, cusparseOperation_t trans_B, int mb, int n,
                                          int kb, int nnzb, T alpha, const cusparseMatDescr_t descr,
                                          const T* bsr_val, const int* bsr_row_ptr, const int* bsr_col_ind,
                                          int block_dim, T* B, int ldb, T beta, T* C, int ldc) {
  static_assert(std::is_same<T, float>::value,
                "spmmdb: only T = float is provided (the only instantiation the reference uses, run_bsrmm.cu:156)");
  if (handle == nullptr) return CUSPARSE_STATUS_NOT_INITIALIZED;
  if (descr == nullptr) return CUSPARSE_STATUS_INVALID_VALUE;
  if (trans_A != CUSPARSE_OPERATION_NON_TRANSPOSE) return CUSPARSE_STATUS_MATRIX_TYPE_NOT_SUPPORTED;
  if (trans_B != CUSPARSE_OPERATION_NON_TRANSPOSE && trans_B != CUSPARSE_OPERATION_TRANSPOSE)
    return CUSPARSE_STATUS_MATRIX_TYPE_NOT_SUPPORTED;
  /* the handle's stream is honoured (the reference launches on stream 0; a fresh handle is on stream 0 too) */
  cudaStream_t stream = nullptr;
  if (cusparseGetStream(handle, &stream) != CUSPARSE_STATUS_SUCCESS) stream = nullptr;
  const int rc = spmmdb_bsrmm_f32(dir == CUSPARSE_DIRECTION_ROW ? SPMMDB_DIR_ROW : SPMMDB_DIR_COLUMN,
                                  trans_B == CUSPARSE_OPERATION_TRANSPOSE ? 1 : 0, mb, n, kb, nnzb, block_dim, alpha,
                                  bsr_val, bsr_row_ptr, bsr_col_ind, B, ldb, beta, C, ldc, SPMMDB_COL_MAJOR,
                                  spmmdb_get_bsrmm_math(), stream);
  return bsrmm_status_to_cusparse(rc);
}

#endif /* SPMMDB_DROPIN_ROCSPARSE_BSRMM_H */
