// host.cu -- host-buffer entry points: what a caller that keeps its matrices in
// host memory (the reference drivers' state before their cudaMemcpy block,
// run_csrmm.cu:94-116) calls.  Copies in, runs the device path, copies out.
#include "common.cuh"

using namespace spmmdb;

extern "C" int spmmdb_csrmm_host_f32(int m, int k, int K, int64_t nnz, const int* h_rowptr, const int* h_colind,
                                     const float* h_val, const float* h_B, float* h_C, float alpha, float beta,
                                     void* stream) {
  if (m < 0 || k < 0 || K < 0 || nnz < 0) return SPMMDB_ERR_INVALID_VALUE;
  if (m == 0 || K == 0) return SPMMDB_OK;
  if (h_rowptr == nullptr || h_B == nullptr || h_C == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  if (nnz > 0 && (h_colind == nullptr || h_val == nullptr)) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t a_bytes = sizeof(int) * ((size_t)m + 1) + (sizeof(int) + sizeof(float)) * (size_t)nnz + 64;
  const size_t b_bytes = sizeof(float) * (size_t)k * K;
  const size_t c_bytes = sizeof(float) * (size_t)m * K;
  char* dA = (char*)scratch(kScratchHostA, a_bytes, st);
  float* dB = (float*)scratch(kScratchHostB, b_bytes, st);
  float* dC = (float*)scratch(kScratchHostC, c_bytes, st);
  if (dA == nullptr || dB == nullptr || dC == nullptr) return SPMMDB_ERR_ALLOC;
  int* d_rowptr = (int*)dA;
  const size_t off_ci = ((sizeof(int) * ((size_t)m + 1) + 15) / 16) * 16;
  int* d_colind = (int*)(dA + off_ci);
  float* d_val = (float*)(dA + off_ci + ((sizeof(int) * (size_t)nnz + 15) / 16) * 16);
  SPMMDB_CUDA_OK(cudaMemcpyAsync(d_rowptr, h_rowptr, sizeof(int) * ((size_t)m + 1), cudaMemcpyHostToDevice, st));
  if (nnz > 0) {
    SPMMDB_CUDA_OK(cudaMemcpyAsync(d_colind, h_colind, sizeof(int) * (size_t)nnz, cudaMemcpyHostToDevice, st));
    SPMMDB_CUDA_OK(cudaMemcpyAsync(d_val, h_val, sizeof(float) * (size_t)nnz, cudaMemcpyHostToDevice, st));
  }
  SPMMDB_CUDA_OK(cudaMemcpyAsync(dB, h_B, b_bytes, cudaMemcpyHostToDevice, st));
  if (beta != 0.f) SPMMDB_CUDA_OK(cudaMemcpyAsync(dC, h_C, c_bytes, cudaMemcpyHostToDevice, st));
  int rc = spmmdb_csrmm_f32(m, k, K, nnz, d_rowptr, d_colind, d_val, dB, K, SPMMDB_ROW_MAJOR, dC, K,
                            SPMMDB_ROW_MAJOR, alpha, beta, st);
  if (rc != SPMMDB_OK) return rc;
  SPMMDB_CUDA_OK(cudaMemcpyAsync(h_C, dC, c_bytes, cudaMemcpyDeviceToHost, st));
  SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
  return SPMMDB_OK;
}

extern "C" int spmmdb_shard_plan(const int* h_rowptr, int m, int parts, int align, int* bounds) {
  if (h_rowptr == nullptr || bounds == nullptr || m < 0 || parts < 1 || align < 1) return SPMMDB_ERR_INVALID_VALUE;
  const int64_t nnz = h_rowptr[m];
  bounds[0] = 0;
  for (int g = 1; g < parts; ++g) {
    const int64_t target = (int64_t)g * nnz / parts;
    // first row with rowptr[row] >= target (lower bound)
    int lo = 0, hi = m;
    while (lo < hi) {
      const int mid = lo + (hi - lo) / 2;
      if (h_rowptr[mid] >= target) hi = mid; else lo = mid + 1;
    }
    int b = (lo / align) * align;
    if (b < bounds[g - 1]) b = bounds[g - 1];
    bounds[g] = b;
  }
  bounds[parts] = m;
  return SPMMDB_OK;
}
