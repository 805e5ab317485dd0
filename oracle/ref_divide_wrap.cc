// ref_divide_wrap.cc -- C entry point around the reference's OWN divide_matrix
// (divide.cu:52-127), compiled from /root/reference/divide.cu where it lies (nvcc, main renamed,
// shim/cusparse_legacy.h force-included for the removed cusparseScsrmm2).  TEST INFRASTRUCTURE.
#include <cstring>
#include <vector>

void divide_matrix(const std::vector<std::vector<int>>& edges, std::vector<int>& csr_row_ptr,
                   std::vector<int>& csr_col_ind, std::vector<int>& bsr_row_ptr, std::vector<int>& bsr_col_ind,
                   std::vector<float>& bsr_val, int n, int bsize, float density);  // divide.cu:52-56

extern "C" {

// Two calls: with out pointers NULL returns sizes (nnzb; *csr_nnz_out), then copies.
long long ref_divide_matrix(int n, const int* rowptr, const int* colind, int bs, float density, int* csr_rowptr,
                            int* csr_colind, int* bsr_rowptr, int* bsr_colind, float* bsr_val,
                            long long* csr_nnz_out) {
  std::vector<std::vector<int>> e(n);
  for (int i = 0; i < n; ++i) e[i].assign(colind + rowptr[i], colind + rowptr[i + 1]);
  std::vector<int> crp, cci, brp, bci;
  std::vector<float> bv;
  divide_matrix(e, crp, cci, brp, bci, bv, n, bs, density);
  if (csr_nnz_out) *csr_nnz_out = (long long)cci.size();
  if (csr_rowptr) {
    std::memcpy(csr_rowptr, crp.data(), sizeof(int) * crp.size());
    if (!cci.empty()) std::memcpy(csr_colind, cci.data(), sizeof(int) * cci.size());
    std::memcpy(bsr_rowptr, brp.data(), sizeof(int) * brp.size());
    if (!bci.empty()) std::memcpy(bsr_colind, bci.data(), sizeof(int) * bci.size());
    if (!bv.empty()) std::memcpy(bsr_val, bv.data(), sizeof(float) * bv.size());
  }
  return (long long)bci.size();
}

}  // extern "C"
