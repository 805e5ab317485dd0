// api.cu -- C-ABI entry points for the SpMM calls: argument validation in the
// reference's order (rocsparse_bsrmm.h:110-188), layout adaptation, dispatch.
#include "common.cuh"

namespace spmmdb {

int csrmm_rowmajor(int m, int K, int64_t nnz, const int* rowptr, const int* colind, const float* val,
                   const float* B, int64_t ldb, float* C, int64_t ldc, float alpha, float beta, cudaStream_t st,
                   int b_bf16 = 0, const unsigned* hot = nullptr);
int bsrmm_simt(int dir, int trans_b, int mb, int n, int bs, float alpha, const float* val, const int* rowptr,
               const int* colind, const float* B, int64_t ldb, float beta, float* C, int64_t ldc, int c_layout,
               cudaStream_t st);
int bsrmm_tc(int dir, int trans_b, int mb, int n, int kb, int64_t nnzb, int bs, float alpha, const float* val,
             const int* rowptr, const int* colind, const float* B, int64_t ldb, float beta, float* C, int64_t ldc,
             int c_layout, int math, cudaStream_t st);
bool bsrmm_tc_supported(int trans_b, int n, int bs, int64_t ldb, const float* B, const float* val, int c_layout,
                        int64_t ldc, const float* C);
int bsr_pack_bf16(int64_t elems, const float* in, void* out, cudaStream_t st);
int bsr_transpose_blocks(int64_t nnzb, int bs, const float* in, float* out, cudaStream_t st);

namespace {

// out[r * ldo + c] = in[c * ldi + r]   (rows x cols result), 32x32 smem tiles.
// mode 0: plain transpose.  mode 1: out = in^T + beta * out (C epilogue).
__global__ void __launch_bounds__(256) transpose_kernel(const float* __restrict__ in, int64_t ldi, float* out,
                                                        int64_t ldo, int rows, int cols, float beta) {
  __shared__ float tile[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  // 1-D grid over the tiles (either dimension may exceed the 65535 limit of grid.y: C5 has 76 533 row tiles)
  const int64_t tiles_r = ((int64_t)rows + 31) / 32;
  const int64_t r0 = ((int64_t)blockIdx.x % tiles_r) * 32, c0 = ((int64_t)blockIdx.x / tiles_r) * 32;
  for (int i = ty; i < 32; i += 8) {
    const int64_t c = c0 + i, r = r0 + tx;
    tile[i][tx] = (c < cols && r < rows) ? in[c * ldi + r] : 0.f;
  }
  __syncthreads();
  for (int i = ty; i < 32; i += 8) {
    const int64_t r = r0 + i, c = c0 + tx;
    if (r < rows && c < cols) {
      float v = tile[tx][i];
      float* o = out + r * ldo + c;
      if (beta != 0.f) v = fmaf(beta, *o, v);
      *o = v;
    }
  }
}

int transpose(const float* in, int64_t ldi, float* out, int64_t ldo, int rows, int cols, float beta,
              cudaStream_t st) {
  if (rows == 0 || cols == 0) return SPMMDB_OK;
  const int64_t tiles = (((int64_t)rows + 31) / 32) * (((int64_t)cols + 31) / 32);
  if (tiles > 0x7fffffffLL) return SPMMDB_ERR_INVALID_VALUE;
  transpose_kernel<<<(unsigned)tiles, 256, 0, st>>>(in, ldi, out, ldo, rows, cols, beta);
  SPMMDB_CUDA_OK(cudaGetLastError());
  return SPMMDB_OK;
}

// nnz = rowptr[m] when the caller does not know it (gespmm_csrmm's signature has no nnz): one 4-byte
// read-back per call, which synchronises `st`.  (A cache keyed by the rowptr address served a stale nnz when
// an allocator handed the same address to another graph, and nnz decides the row segmentation.)  Callers
// that know nnz pass it and stay asynchronous.
int resolve_nnz(const int* rowptr, int m, cudaStream_t st, int64_t* nnz) {
  int h = 0;
  SPMMDB_CUDA_OK(cudaMemcpyAsync(&h, rowptr + m, sizeof(int), cudaMemcpyDeviceToHost, st));
  SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
  if (h < 0) return SPMMDB_ERR_INVALID_VALUE;
  *nnz = h;
  return SPMMDB_OK;
}

}  // namespace
}  // namespace spmmdb

using namespace spmmdb;

extern "C" int spmmdb_csrmm_f32(int m, int k, int K, int64_t nnz, const int* rowptr, const int* colind,
                                const float* val, const float* B, int64_t ldb, int b_layout, float* C, int64_t ldc,
                                int c_layout, float alpha, float beta, void* stream) {
  if (m < 0 || K < 0 || k < 0) return SPMMDB_ERR_INVALID_VALUE;
  if ((b_layout != SPMMDB_ROW_MAJOR && b_layout != SPMMDB_COL_MAJOR) ||
      (c_layout != SPMMDB_ROW_MAJOR && c_layout != SPMMDB_COL_MAJOR))
    return SPMMDB_ERR_INVALID_VALUE;
  if (m == 0 || K == 0) return SPMMDB_OK;
  if (rowptr == nullptr || B == nullptr || C == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  if (b_layout == SPMMDB_ROW_MAJOR ? ldb < K : ldb < k) return SPMMDB_ERR_INVALID_VALUE;
  if (c_layout == SPMMDB_ROW_MAJOR ? ldc < K : ldc < m) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  cudaStream_t st = (cudaStream_t)stream;
  if (nnz < 0) {
    int rc = resolve_nnz(rowptr, m, st, &nnz);
    if (rc != SPMMDB_OK) return rc;
  }
  if (nnz > 0 && (colind == nullptr || val == nullptr)) return SPMMDB_ERR_INVALID_VALUE;

  const int Kpad = (K + 3) & ~3;
  const float* Bk = B;
  int64_t ldbk = ldb;
  if (b_layout == SPMMDB_COL_MAJOR) {
    // column-major B (cusparseScsrmm, run_csrmm.cu:135): re-lay as row-major once
    float* Bt = (float*)scratch(kScratchLayoutB, sizeof(float) * (size_t)k * Kpad, st);
    if (Bt == nullptr) return SPMMDB_ERR_ALLOC;
    int rc = transpose(B, ldb, Bt, Kpad, k, K, 0.f, st);
    if (rc != SPMMDB_OK) return rc;
    Bk = Bt;
    ldbk = Kpad;
  }
  if (c_layout == SPMMDB_ROW_MAJOR)
    return csrmm_rowmajor(m, K, nnz, rowptr, colind, val, Bk, ldbk, C, ldc, alpha, beta, st);
  // column-major C (every cuSPARSE call site): compute row-major, fold beta*C in the transpose
  float* Ct = (float*)scratch(kScratchLayoutC, sizeof(float) * (size_t)m * Kpad, st);
  if (Ct == nullptr) return SPMMDB_ERR_ALLOC;
  int rc = csrmm_rowmajor(m, K, nnz, rowptr, colind, val, Bk, ldbk, Ct, Kpad, alpha, 0.f, st);
  if (rc != SPMMDB_OK) return rc;
  // C[c * ldc + r] = Ct[r * Kpad + c] (+ beta*C): "in" is Ct seen as column-major K x m with ld Kpad
  return transpose(Ct, Kpad, C, ldc, K, m, beta, st);
}

extern "C" int spmmdb_bsrmm_f32(int dir, int trans_b, int mb, int n, int kb, int64_t nnzb, int block_dim,
                                float alpha, const float* bsr_val, const int* bsr_rowptr, const int* bsr_colind,
                                const float* B, int64_t ldb, float beta, float* C, int64_t ldc, int c_layout,
                                int math, void* stream) {
  // same order of checks as rocsparse_bsrmm.h:110-176
  if (dir != SPMMDB_DIR_ROW && dir != SPMMDB_DIR_COLUMN) return SPMMDB_ERR_INVALID_VALUE;
  if (trans_b != 0 && trans_b != 1) return SPMMDB_ERR_NOT_SUPPORTED;
  if (mb < 0 || n < 0 || kb < 0 || nnzb < 0 || block_dim <= 0) return SPMMDB_ERR_INVALID_VALUE;
  if (mb == 0 || n == 0 || kb == 0 || nnzb == 0) return SPMMDB_OK;  // quick return (:151-153)
  if (bsr_val == nullptr || bsr_rowptr == nullptr || bsr_colind == nullptr || B == nullptr || C == nullptr)
    return SPMMDB_ERR_INVALID_VALUE;
  const int64_t m = (int64_t)mb * block_dim;
  if (trans_b == 0 ? ldb < (int64_t)kb * block_dim : ldb < n) return SPMMDB_ERR_INVALID_VALUE;
  if (c_layout == SPMMDB_COL_MAJOR ? ldc < m : ldc < n) return SPMMDB_ERR_INVALID_VALUE;
  if (c_layout != SPMMDB_COL_MAJOR && c_layout != SPMMDB_ROW_MAJOR) return SPMMDB_ERR_INVALID_VALUE;
  if (math != SPMMDB_MATH_FP32 && math != SPMMDB_MATH_TF32 && math != SPMMDB_MATH_BF16 &&
      math != SPMMDB_MATH_BF16_PACKED)
    return SPMMDB_ERR_INVALID_VALUE;
  if (alpha == 0.f && beta == 1.f) return SPMMDB_OK;  // rocsparse_bsrmm.h:76-78
  const int sm = device_sm();
  if (sm < 0) return SPMMDB_ERR_CUDA;
  cudaStream_t st = (cudaStream_t)stream;
  if (math != SPMMDB_MATH_FP32 && bsrmm_tc_supported(trans_b, n, block_dim, ldb, B, bsr_val, c_layout, ldc, C)) {
    if (sm / 10 != 10) return SPMMDB_ERR_ARCH;
    const int rc = bsrmm_tc(dir, trans_b, mb, n, kb, nnzb, block_dim, alpha, bsr_val, bsr_rowptr, bsr_colind, B, ldb,
                            beta, C, ldc, c_layout, math, st);
    // a shape the tensor-core kernel does not lay out (tf32 with 64-wide blocks) runs the fp32 kernel
    if (rc != SPMMDB_ERR_NOT_SUPPORTED) return rc;
  }
  if (math == SPMMDB_MATH_BF16_PACKED) return SPMMDB_ERR_NOT_SUPPORTED;  // bf16 blocks: tensor-core shapes only
  return bsrmm_simt(dir, trans_b, mb, n, block_dim, alpha, bsr_val, bsr_rowptr, bsr_colind, B, ldb, beta, C, ldc,
                    c_layout, st);
}

extern "C" int spmmdb_block_gemm_f32(int mb, int dim, int64_t nnzb, int block_dim, const int* bsr_rowptr,
                                     const int* bsr_colind, const float* bsr_val, const float* Y, float* Z,
                                     int math, void* stream) {
  // block_cublas.cu:132: sgemm(N, T, bs, dim, bs, 1, A (col-major block), bs, Y + col*bs*dim, dim, 1, Z + i*bs, m)
  if (mb < 0 || dim < 0 || nnzb < 0 || block_dim <= 0) return SPMMDB_ERR_INVALID_VALUE;
  const int64_t m = (int64_t)mb * block_dim;
  return spmmdb_bsrmm_f32(SPMMDB_DIR_COLUMN, /*trans_b=*/1, mb, dim, /*kb=*/mb > 0 ? mb : 1, nnzb, block_dim, 1.0f,
                          bsr_val, bsr_rowptr, bsr_colind, Y, dim, 1.0f, Z, m, SPMMDB_COL_MAJOR, math, stream);
}

extern "C" int spmmdb_bsr_pack_bf16(int64_t nnzb, int block_dim, const float* bsr_val, void* out_bf16, void* stream) {
  if (nnzb < 0 || block_dim <= 0 || (block_dim * block_dim) % 4 != 0) return SPMMDB_ERR_INVALID_VALUE;
  if (nnzb == 0) return SPMMDB_OK;
  if (bsr_val == nullptr || out_bf16 == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  if (((uintptr_t)bsr_val | (uintptr_t)out_bf16) % 16 != 0) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  return bsr_pack_bf16(nnzb * block_dim * block_dim, bsr_val, out_bf16, (cudaStream_t)stream);
}

extern "C" int spmmdb_bsr_transpose_blocks_f32(int64_t nnzb, int block_dim, const float* bsr_val, float* out, void* stream) {
  if (nnzb < 0 || block_dim <= 0) return SPMMDB_ERR_INVALID_VALUE;
  if (nnzb == 0) return SPMMDB_OK;
  if (bsr_val == nullptr || out == nullptr || bsr_val == out) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  return bsr_transpose_blocks(nnzb, block_dim, bsr_val, out, (cudaStream_t)stream);
}
