// hybrid.cu -- the hybrid dense-block product as ONE object: the reference's `divide` program (divide.cu:205-391)
// splits the adjacency matrix once (divide_matrix, :52-127) and then multiplies the two parts into one C
// (:348-373).  spmmdb_hybrid_create does the split on the device and keeps both parts; spmmdb_hybrid_spmm_f32 is
//   C = alpha * A_csr * B + beta * C            (CSR kernel: the entries outside dense blocks)
//   C += alpha * A_bsr * B                      (tcgen05 BSR kernel: the blocks whose fill reaches `density`)
// This is the BSR entry to use on graph adjacency matrices: plain BSR multiplies mostly zeros at the 1-3 % block
// fill a reordering leaves (6-7 ms against 0.45 ms for the hybrid on the Gorder-reordered
// proteins-shaped graph, 1.0 ms for CSR alone).  m and n need not be multiples of the block size: B and C keep
// their own shapes (the BSR kernels clip, RowClip), unlike the reference drivers, which pad both (run_bsrmm.cu:86-94).
#include "common.cuh"

using namespace spmmdb;

struct spmmdb_hybrid {
  int m = 0, n = 0, bs = 0, device = 0;
  int64_t nnzb = 0, csr_nnz = 0;
  int *bsr_rowptr = nullptr, *bsr_colind = nullptr, *csr_rowptr = nullptr, *csr_colind = nullptr;
  float *bsr_val = nullptr, *csr_val = nullptr;
};

extern "C" int spmmdb_hybrid_destroy(spmmdb_hybrid* h) {
  if (h == nullptr) return SPMMDB_OK;
  cudaFree(h->bsr_rowptr); cudaFree(h->bsr_colind); cudaFree(h->bsr_val);
  cudaFree(h->csr_rowptr); cudaFree(h->csr_colind); cudaFree(h->csr_val);
  delete h;
  return SPMMDB_OK;
}

extern "C" int spmmdb_hybrid_create(int m, int n, const int* rowptr, const int* colind, const float* val, int block_dim,
                                    float density, spmmdb_hybrid** out, void* stream) {
  if (out == nullptr || m < 0 || n < 0 || block_dim <= 0 || block_dim > 128) return SPMMDB_ERR_INVALID_VALUE;
  if (m > 0 && rowptr == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  spmmdb_hybrid* h = new spmmdb_hybrid();
  h->m = m; h->n = n; h->bs = block_dim;
  cudaGetDevice(&h->device);
  const int mb = (m + block_dim - 1) / block_dim;
  int rc = SPMMDB_OK;
  auto fail = [&](int code) { spmmdb_hybrid_destroy(h); return code; };
  if (cudaMalloc(&h->bsr_rowptr, sizeof(int) * ((size_t)mb + 1)) != cudaSuccess ||
      cudaMalloc(&h->csr_rowptr, sizeof(int) * ((size_t)m + 1)) != cudaSuccess)
    return fail(SPMMDB_ERR_ALLOC);
  int nnzb = 0, cnnz = 0;
  rc = spmmdb_divide_nnz(m, n, rowptr, colind, block_dim, density, h->bsr_rowptr, h->csr_rowptr, &nnzb, &cnnz, stream);
  if (rc != SPMMDB_OK) return fail(rc);
  h->nnzb = nnzb; h->csr_nnz = cnnz;
  const size_t bvals = (size_t)nnzb * block_dim * block_dim;
  if (cudaMalloc(&h->bsr_colind, sizeof(int) * (size_t)(nnzb > 0 ? nnzb : 1)) != cudaSuccess ||
      cudaMalloc(&h->bsr_val, sizeof(float) * (bvals > 0 ? bvals : 1)) != cudaSuccess ||
      cudaMalloc(&h->csr_colind, sizeof(int) * (size_t)(cnnz > 0 ? cnnz : 1)) != cudaSuccess ||
      cudaMalloc(&h->csr_val, sizeof(float) * (size_t)(cnnz > 0 ? cnnz : 1)) != cudaSuccess)
    return fail(SPMMDB_ERR_ALLOC);
  rc = spmmdb_divide_f32(SPMMDB_DIR_ROW, m, n, rowptr, colind, val, block_dim, density, h->bsr_rowptr, h->bsr_colind,
                         h->bsr_val, h->csr_rowptr, h->csr_colind, h->csr_val, stream);
  if (rc != SPMMDB_OK) return fail(rc);
  if (cudaStreamSynchronize((cudaStream_t)stream) != cudaSuccess) return fail(SPMMDB_ERR_CUDA);
  *out = h;
  return SPMMDB_OK;
}

extern "C" int spmmdb_hybrid_info(const spmmdb_hybrid* h, int64_t* nnzb, int64_t* csr_nnz) {
  if (h == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  if (nnzb) *nnzb = h->nnzb;
  if (csr_nnz) *csr_nnz = h->csr_nnz;
  return SPMMDB_OK;
}

extern "C" int spmmdb_hybrid_spmm_f32(const spmmdb_hybrid* h, int K, const float* B, int64_t ldb, float* C, int64_t ldc,
                                      float alpha, float beta, int math, void* stream) {
  if (h == nullptr || K < 0) return SPMMDB_ERR_INVALID_VALUE;
  if (h->m == 0 || K == 0) return SPMMDB_OK;
  if (B == nullptr || C == nullptr || ldb < K || ldc < K) return SPMMDB_ERR_INVALID_VALUE;
  // the remainder first: it writes every row of C (empty rows included), the blocks then accumulate
  int rc = spmmdb_csrmm_f32(h->m, h->n, K, h->csr_nnz, h->csr_rowptr, h->csr_colind, h->csr_val, B, ldb, SPMMDB_ROW_MAJOR, C,
                            ldc, SPMMDB_ROW_MAJOR, alpha, beta, stream);
  if (rc != SPMMDB_OK || h->nnzb == 0) return rc;
  const int mb = (h->m + h->bs - 1) / h->bs, kb = (h->n + h->bs - 1) / h->bs;
  ScopedRowClip clip(h->n, h->m);  // B has n rows, C has m rows: no padding to whole blocks asked of the caller
  return spmmdb_bsrmm_f32(SPMMDB_DIR_ROW, /*trans_b=*/1, mb, K, kb, h->nnzb, h->bs, alpha, h->bsr_val, h->bsr_rowptr,
                          h->bsr_colind, B, ldb, 1.0f, C, ldc, SPMMDB_ROW_MAJOR, math, stream);
}
