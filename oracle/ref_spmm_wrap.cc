// ref_spmm_wrap.cc -- C entry points around the reference's OWN CPU SpMM,
// compiled from /root/reference/spmm.cc where it lies (never copied).
// TEST INFRASTRUCTURE: built by oracle/Makefile into oracle/_ref/.
#define main ref_spmm_cc_main
#include "spmm.cc"   // found through -I$(REF); csr_spmm spmm.cc:7-25, coo_spmm :27-43
#undef main
#include <omp.h>

extern "C" {

void ref_csr_spmm(int64_t rows, int64_t cols, int64_t K, int64_t* indptr,
                  int64_t* indices, double* dense, double* out) {
  CSRMatrix csr(rows, cols, indptr, indices);
  DenseMatrix d(cols, K, dense);
  DenseMatrix o(rows, K, out);
  csr_spmm(csr, d, o);
}

void ref_coo_spmm(int64_t rows, int64_t cols, int64_t nnz, int64_t K,
                  int64_t* row, int64_t* col, double* dense, double* out) {
  COOMatrix coo(rows, cols, nnz, row, col);
  DenseMatrix d(cols, K, dense);
  DenseMatrix o(rows, K, out);
  coo_spmm(coo, d, o);
}

int ref_omp_max_threads() { return omp_get_max_threads(); }
void ref_omp_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }

}  // extern "C"
