// ref_gpu_wrap.cu -- C entry points around the REFERENCE's own GPU kernels, compiled for sm_100a from the
// headers where they lie under /root/reference (gespmm_csrmm.h, rocsparse_bsrmm.h + rocsparse_bsrmm_impl.h),
// into baseline/_ref/libref_gpu.so (git-ignored; travels to the GPU box).  Comparator only: tests/ and
// baseline/compare.py run these kernels next to libspmmdb.so on the same inputs ("the reference
// spmm_test2 kernel compiled for sm_100a ... kept out of the product tree").  Nothing in the product links it.
#include <cuda_runtime.h>
#include <cusparse.h>

#include "gespmm_csrmm.h"      // /root/reference/gespmm_csrmm.h
#include "rocsparse_bsrmm.h"   // /root/reference/rocsparse_bsrmm.h

extern "C" {

// gespmm_csrmm<float> (gespmm_csrmm.h:422-426): row-major B and C, stream 0.
int ref_gespmm_csrmm_f32(int m, int K, int* rowptr, int* colind, float* val, float* B, float* C) {
  gespmm_csrmm<float>(m, K, rowptr, colind, val, B, C);
  return (int)cudaGetLastError();
}

// spmmWrapper<float>(method, tile_row, ...) (gespmm_csrmm.h:379-420).
int ref_spmm_wrapper_f32(int method, int tile_row, int m, int K, int* rowptr, int* colind, float* val, float* B,
                         float* C) {
  spmmWrapper<float>(method, tile_row, m, K, rowptr, colind, val, B, C);
  return (int)cudaGetLastError();
}

// rocsparse_bsrmm_template<float> (rocsparse_bsrmm.h:102-256) with a real handle / descriptor.
int ref_rocsparse_bsrmm_f32(int dir, int trans_b, int mb, int n, int kb, int nnzb, float alpha, const float* bsr_val,
                            const int* bsr_rowptr, const int* bsr_colind, int block_dim, float* B, int ldb, float beta,
                            float* C, int ldc) {
  static cusparseHandle_t handle = nullptr;
  static cusparseMatDescr_t descr = nullptr;
  if (handle == nullptr) {
    if (cusparseCreate(&handle) != CUSPARSE_STATUS_SUCCESS) return -100;
    if (cusparseCreateMatDescr(&descr) != CUSPARSE_STATUS_SUCCESS) return -101;
  }
  cusparseStatus_t st = rocsparse_bsrmm_template<float>(
      handle, dir ? CUSPARSE_DIRECTION_COLUMN : CUSPARSE_DIRECTION_ROW, CUSPARSE_OPERATION_NON_TRANSPOSE,
      trans_b ? CUSPARSE_OPERATION_TRANSPOSE : CUSPARSE_OPERATION_NON_TRANSPOSE, mb, n, kb, nnzb, alpha, descr, bsr_val,
      bsr_rowptr, bsr_colind, block_dim, B, ldb, beta, C, ldc);
  if (st != CUSPARSE_STATUS_SUCCESS) return (int)st;
  return (int)cudaGetLastError();
}

}  // extern "C"
