// vendor_cmp.cu -- the LIBRARY calls the reference makes on this path, behind plain C entry points, so that
// tests/ and baseline/compare.py can run them on the same device arrays as libspmmdb.so:
//   cusparseXcsr2bsrNnz + cusparseScsr2bsr   run_bsrmm.cu:121-142, check_result.cu:172-193, csr2bsr.cu:176-192
//   cusparseSbsr2csr                         bsr2csr.cu:186-188
//   cusparseSbsrmm                           run_bsrmm.cu:161, check_result.cu:223-228, bsrmm.cu:141-144
//   cusparseSpMM (CSR)                       successor of cusparseScsrmm / csrmm2 (run_csrmm.cu:133-142), which
//                                            left cuSPARSE after CUDA 10.2
//   cublasSgemm per non-zero block           block_cublas.cu:123-136
// Comparator / test infrastructure only: built into baseline/_ref/libvendor_cmp.so, never linked by the product.
// Every function returns 0 or a negative code; all pointers are device pointers unless named h_*.
#include <cublas_v2.h>
#include <cuda_runtime.h>
#include <cusparse.h>
#include <stdint.h>
#include <stdio.h>

namespace {
cusparseHandle_t sp() {
  static cusparseHandle_t h = nullptr;
  if (h == nullptr) cusparseCreate(&h);
  return h;
}
cusparseMatDescr_t descr() {
  static cusparseMatDescr_t d = nullptr;
  if (d == nullptr) {
    cusparseCreateMatDescr(&d);
    cusparseSetMatType(d, CUSPARSE_MATRIX_TYPE_GENERAL);
    cusparseSetMatIndexBase(d, CUSPARSE_INDEX_BASE_ZERO);
  }
  return d;
}
cublasHandle_t bl() {
  static cublasHandle_t h = nullptr;
  if (h == nullptr) cublasCreate(&h);
  return h;
}
#define SP_OK(x) do { cusparseStatus_t _s = (x); if (_s != CUSPARSE_STATUS_SUCCESS) return -1000 - (int)_s; } while (0)
#define CU_OK(x) do { cudaError_t _e = (x); if (_e != cudaSuccess) return -2000 - (int)_e; } while (0)
}  // namespace

extern "C" {

int vendor_csr2bsr_nnz(int dir, int m, int n, const int* rowptr, const int* colind, int bs, int* bsr_rowptr,
                       int* h_nnzb) {
  const int mb = (m + bs - 1) / bs;
  (void)mb;
  SP_OK(cusparseXcsr2bsrNnz(sp(), dir ? CUSPARSE_DIRECTION_COLUMN : CUSPARSE_DIRECTION_ROW, m, n, descr(), rowptr,
                            colind, bs, descr(), bsr_rowptr, h_nnzb));
  CU_OK(cudaDeviceSynchronize());
  return 0;
}

int vendor_csr2bsr(int dir, int m, int n, const int* rowptr, const int* colind, const float* val, int bs,
                   int* bsr_rowptr, int* bsr_colind, float* bsr_val) {
  SP_OK(cusparseScsr2bsr(sp(), dir ? CUSPARSE_DIRECTION_COLUMN : CUSPARSE_DIRECTION_ROW, m, n, descr(), val, rowptr,
                         colind, bs, descr(), bsr_val, bsr_rowptr, bsr_colind));
  return 0;
}

int vendor_bsr2csr(int dir, int mb, int nb, const int* bsr_rowptr, const int* bsr_colind, const float* bsr_val, int bs,
                   int* csr_rowptr, int* csr_colind, float* csr_val) {
  SP_OK(cusparseSbsr2csr(sp(), dir ? CUSPARSE_DIRECTION_COLUMN : CUSPARSE_DIRECTION_ROW, mb, nb, descr(), bsr_val,
                         bsr_rowptr, bsr_colind, bs, descr(), csr_val, csr_rowptr, csr_colind));
  return 0;
}

// C(col-major, ldc) = alpha * A_bsr * op(B) + beta * C; trans_b 0: B col-major (ldb >= kb*bs), 1: B row-major [k][ldb]
int vendor_bsrmm(int dir, int trans_b, int mb, int n, int kb, int nnzb, float alpha, const float* bsr_val,
                 const int* bsr_rowptr, const int* bsr_colind, int bs, const float* B, int ldb, float beta, float* C,
                 int ldc) {
  SP_OK(cusparseSbsrmm(sp(), dir ? CUSPARSE_DIRECTION_COLUMN : CUSPARSE_DIRECTION_ROW, CUSPARSE_OPERATION_NON_TRANSPOSE,
                       trans_b ? CUSPARSE_OPERATION_TRANSPOSE : CUSPARSE_OPERATION_NON_TRANSPOSE, mb, n, kb, nnzb,
                       &alpha, descr(), bsr_val, bsr_rowptr, bsr_colind, bs, B, ldb, &beta, C, ldc));
  return 0;
}

// cusparseSpMM, CSR x dense.  Layouts as spmmdb_csrmm_f32 (0 row-major, 1 column-major).  alg: 0 default,
// 2 / 3 = CUSPARSE_SPMM_CSR_ALG2 / ALG3.  The descriptors, the workspace and (alg >= 0, `preprocess`) the
// preprocessing are made once per distinct call signature and kept, so a timed loop measures the SpMM only.
struct SpmmState {
  const void* key_rp = nullptr; const void* key_b = nullptr; const void* key_c = nullptr;
  int m = 0, k = 0, K = 0, alg = 0;
  cusparseSpMatDescr_t A = nullptr; cusparseDnMatDescr_t B = nullptr, C = nullptr;
  void* ws = nullptr; size_t ws_bytes = 0;
};
static SpmmState g_spmm;

int vendor_spmm_csr(int m, int k, int K, int64_t nnz, const int* rowptr, const int* colind, const float* val,
                    const float* B, int64_t ldb, int b_layout, float* C, int64_t ldc, int c_layout, float alpha,
                    float beta, int alg) {
  SpmmState& s = g_spmm;
  const cusparseSpMMAlg_t a = alg == 2 ? CUSPARSE_SPMM_CSR_ALG2 : (alg == 3 ? CUSPARSE_SPMM_CSR_ALG3
                                       : (alg == 1 ? CUSPARSE_SPMM_CSR_ALG1 : CUSPARSE_SPMM_ALG_DEFAULT));
  if (s.key_rp != rowptr || s.key_b != B || s.key_c != C || s.m != m || s.k != k || s.K != K || s.alg != alg) {
    if (s.A) { cusparseDestroySpMat(s.A); cusparseDestroyDnMat(s.B); cusparseDestroyDnMat(s.C); s.A = nullptr; }
    SP_OK(cusparseCreateCsr(&s.A, m, k, nnz, (void*)rowptr, (void*)colind, (void*)val, CUSPARSE_INDEX_32I,
                            CUSPARSE_INDEX_32I, CUSPARSE_INDEX_BASE_ZERO, CUDA_R_32F));
    SP_OK(cusparseCreateDnMat(&s.B, k, K, ldb, (void*)B, CUDA_R_32F, b_layout ? CUSPARSE_ORDER_COL : CUSPARSE_ORDER_ROW));
    SP_OK(cusparseCreateDnMat(&s.C, m, K, ldc, (void*)C, CUDA_R_32F, c_layout ? CUSPARSE_ORDER_COL : CUSPARSE_ORDER_ROW));
    size_t need = 0;
    SP_OK(cusparseSpMM_bufferSize(sp(), CUSPARSE_OPERATION_NON_TRANSPOSE, CUSPARSE_OPERATION_NON_TRANSPOSE, &alpha, s.A,
                                  s.B, &beta, s.C, CUDA_R_32F, a, &need));
    if (need > s.ws_bytes) {
      if (s.ws) cudaFree(s.ws);
      CU_OK(cudaMalloc(&s.ws, need + 256));
      s.ws_bytes = need;
    }
    SP_OK(cusparseSpMM_preprocess(sp(), CUSPARSE_OPERATION_NON_TRANSPOSE, CUSPARSE_OPERATION_NON_TRANSPOSE, &alpha, s.A,
                                  s.B, &beta, s.C, CUDA_R_32F, a, s.ws));
    s.key_rp = rowptr; s.key_b = B; s.key_c = C; s.m = m; s.k = k; s.K = K; s.alg = alg;
  }
  SP_OK(cusparseSpMM(sp(), CUSPARSE_OPERATION_NON_TRANSPOSE, CUSPARSE_OPERATION_NON_TRANSPOSE, &alpha, s.A, s.B, &beta,
                     s.C, CUDA_R_32F, a, s.ws));
  return 0;
}

// block_cublas.cu:123-136: for every non-zero block, z(col-major, ld m) += A_blk(col-major bs x bs) * Y[col*bs.., :]
// (row-major, ld dim) as one cublasSgemm(N, T, bs, dim, bs); index arrays on the HOST like the reference.
int vendor_block_gemm_loop(int mb, int dim, int bs, const int* h_bsr_rowptr, const int* h_bsr_colind,
                           const float* bsr_val, const float* Y, float* Z) {
  const float one = 1.0f;
  const int m = mb * bs;
  for (int i = 0; i < mb; ++i)
    for (int p = h_bsr_rowptr[i]; p < h_bsr_rowptr[i + 1]; ++p) {
      const int col = h_bsr_colind[p];
      cublasStatus_t st = cublasSgemm(bl(), CUBLAS_OP_N, CUBLAS_OP_T, bs, dim, bs, &one, bsr_val + (size_t)p * bs * bs, bs,
                                      Y + (size_t)col * bs * dim, dim, &one, Z + (size_t)i * bs, m);
      if (st != CUBLAS_STATUS_SUCCESS) return -3000 - (int)st;
    }
  return 0;
}

}  // extern "C"
