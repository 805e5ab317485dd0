// csrplan.cu -- optional per-matrix analysis and operand packing for the CSR path (extensions; the drop-in
// entry points do not need them):
//   spmmdb_csr_hot_columns   which B rows are worth keeping in L2.  A B row is re-read once per non-zero of its
//                            column; with ~6.5 TB/s of misses flowing through a 126 MB L2 a line lives ~20 us,
//                            so under plain LRU only columns hit more often than that survive (ogbn-products
//                            shape: the top 12 K columns, 26 % of the gathers).  Marking the top columns that FIT
//                            the cache evict_last and the rest evict_first keeps ~100 K rows resident instead.
//   spmmdb_dense_pack_bf16   fp32 -> bf16 copy of a row-major dense matrix (B for spmmdb_csrmm_ex_f32 with
//                            SPMMDB_DTYPE_BF16: half the gather bytes, fp32 accumulation).
#include <cuda_bf16.h>

#include "common.cuh"

namespace spmmdb {
int csrmm_rowmajor(int m, int K, int64_t nnz, const int* rowptr, const int* colind, const float* val,
                   const float* B, int64_t ldb, float* C, int64_t ldc, float alpha, float beta, cudaStream_t st,
                   int b_bf16, const unsigned* hot);
namespace {

constexpr int kBins = 4096;

__global__ void __launch_bounds__(256) col_hist_kernel(const int* __restrict__ colind, int64_t nnz, int k,
                                                       int* __restrict__ counts) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += stride) {
    const int c = __ldg(colind + i);
    if ((unsigned)c < (unsigned)k) atomicAdd(counts + c, 1);
  }
}

// bins[min(count, kBins - 1)] += 1 per column (shared-memory histogram per CTA, merged with global atomics)
__global__ void __launch_bounds__(256) count_hist_kernel(const int* __restrict__ counts, int k, int* __restrict__ bins) {
  __shared__ int sh[kBins];
  for (int i = threadIdx.x; i < kBins; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const int stride = gridDim.x * blockDim.x;
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < k; c += stride) atomicAdd(sh + min(counts[c], kBins - 1), 1);
  __syncthreads();
  for (int i = threadIdx.x; i < kBins; i += blockDim.x)
    if (sh[i] != 0) atomicAdd(bins + i, sh[i]);
}

// thr[0] = smallest t >= 1 with #{columns : count >= t} <= budget; thr[1] = that number of columns.
// Columns at the saturated last bin are indistinguishable: if even they exceed the budget nothing is marked.
__global__ void threshold_kernel(const int* __restrict__ bins, int64_t budget, int* __restrict__ thr) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  int64_t above = 0;  // columns with count >= t, for t walking down
  int t = kBins;      // nothing qualifies yet
  int64_t picked = 0;
  for (int b = kBins - 1; b >= 1; --b) {
    if (above + bins[b] > budget) break;
    above += bins[b];
    t = b;
    picked = above;
  }
  thr[0] = t;
  thr[1] = (int)min(picked, (int64_t)0x7fffffff);
}

__global__ void __launch_bounds__(256) hot_bitmap_kernel(const int* __restrict__ counts, int k, const int* __restrict__ thr,
                                                         unsigned* __restrict__ bitmap) {
  const int t = thr[0];
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const bool hot = c < k && counts[c] >= t;
  const unsigned word = __ballot_sync(0xffffffffu, hot);
  if ((threadIdx.x & 31) == 0 && c < k) bitmap[c >> 5] = word;
}

__global__ void __launch_bounds__(256) pack_bf16_2d_kernel(const float* __restrict__ in, int64_t ldi, __nv_bfloat16* __restrict__ out,
                                                           int64_t ldo, int64_t rows, int K) {
  // K % 4 == 0 and 16-byte aligned rows: one float4 -> 4 bf16 per thread step
  const int kv = K >> 2;
  const int64_t total = rows * kv;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int64_t r = i / kv;
    const int j = (int)(i - r * kv);
    const float4 x = __ldcs(reinterpret_cast<const float4*>(in + r * ldi) + j);
    __nv_bfloat162 lo = __floats2bfloat162_rn(x.x, x.y), hi = __floats2bfloat162_rn(x.z, x.w);
    uint2 w;
    w.x = *reinterpret_cast<unsigned*>(&lo);
    w.y = *reinterpret_cast<unsigned*>(&hi);
    *(reinterpret_cast<uint2*>(out + r * ldo) + j) = w;
  }
}

__global__ void __launch_bounds__(256) pack_bf16_scalar_kernel(const float* __restrict__ in, int64_t ldi, __nv_bfloat16* __restrict__ out,
                                                               int64_t ldo, int64_t rows, int K) {
  const int64_t total = rows * K;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int64_t r = i / K;
    const int j = (int)(i - r * K);
    out[r * ldo + j] = __float2bfloat16_rn(in[r * ldi + j]);
  }
}

}  // namespace
}  // namespace spmmdb

using namespace spmmdb;

extern "C" int spmmdb_csr_hot_columns(int k, int64_t nnz, const int* colind, int64_t max_hot_columns,
                                      unsigned* hot_bitmap, int* n_hot_host, void* stream) {
  if (k < 0 || nnz < 0 || max_hot_columns < 0) return SPMMDB_ERR_INVALID_VALUE;
  if (k == 0) { if (n_hot_host) *n_hot_host = 0; return SPMMDB_OK; }
  if (hot_bitmap == nullptr || (nnz > 0 && colind == nullptr)) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t ints = (size_t)k + kBins + 8;
  int* ws = (int*)scratch(kScratchConvert, sizeof(int) * ints, st);
  if (ws == nullptr) return SPMMDB_ERR_ALLOC;
  int* counts = ws;
  int* bins = ws + k;
  int* thr = bins + kBins;
  SPMMDB_CUDA_OK(cudaMemsetAsync(ws, 0, sizeof(int) * ints, st));
  const int nsm = device_num_sms() > 0 ? device_num_sms() : 148;
  if (nnz > 0) col_hist_kernel<<<nsm * 8, 256, 0, st>>>(colind, nnz, k, counts);
  count_hist_kernel<<<nsm, 256, 0, st>>>(counts, k, bins);
  threshold_kernel<<<1, 32, 0, st>>>(bins, max_hot_columns, thr);
  hot_bitmap_kernel<<<(k + 255) / 256, 256, 0, st>>>(counts, k, thr, hot_bitmap);
  SPMMDB_CUDA_OK(cudaGetLastError());
  if (n_hot_host != nullptr) {
    SPMMDB_CUDA_OK(cudaMemcpyAsync(n_hot_host, thr + 1, sizeof(int), cudaMemcpyDeviceToHost, st));
    SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
  }
  return SPMMDB_OK;
}

extern "C" int spmmdb_dense_pack_bf16(int64_t rows, int K, const float* in, int64_t ldi, void* out_bf16, int64_t ldo,
                                      void* stream) {
  if (rows < 0 || K < 0 || ldi < K || ldo < K) return SPMMDB_ERR_INVALID_VALUE;
  if (rows == 0 || K == 0) return SPMMDB_OK;
  if (in == nullptr || out_bf16 == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  cudaStream_t st = (cudaStream_t)stream;
  const int nsm = device_num_sms() > 0 ? device_num_sms() : 148;
  const bool vec = (K % 4 == 0) && (ldi % 4 == 0) && (ldo % 4 == 0) && ((uintptr_t)in % 16 == 0) &&
                   ((uintptr_t)out_bf16 % 8 == 0);
  if (vec) pack_bf16_2d_kernel<<<nsm * 16, 256, 0, st>>>(in, ldi, (__nv_bfloat16*)out_bf16, ldo, rows, K);
  else pack_bf16_scalar_kernel<<<nsm * 16, 256, 0, st>>>(in, ldi, (__nv_bfloat16*)out_bf16, ldo, rows, K);
  SPMMDB_CUDA_OK(cudaGetLastError());
  return SPMMDB_OK;
}

extern "C" int spmmdb_csrmm_ex_f32(int m, int k, int K, int64_t nnz, const int* rowptr, const int* colind, const float* val,
                                   const void* B, int b_dtype, int64_t ldb, float* C, int64_t ldc, float alpha, float beta,
                                   const unsigned* hot_bitmap, void* stream) {
  if (m < 0 || K < 0 || k < 0 || nnz < 0) return SPMMDB_ERR_INVALID_VALUE;
  if (b_dtype != SPMMDB_DTYPE_F32 && b_dtype != SPMMDB_DTYPE_BF16) return SPMMDB_ERR_INVALID_VALUE;
  if (m == 0 || K == 0) return SPMMDB_OK;
  if (rowptr == nullptr || B == nullptr || C == nullptr || ldb < K || ldc < K) return SPMMDB_ERR_INVALID_VALUE;
  if (nnz > 0 && (colind == nullptr || val == nullptr)) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  return csrmm_rowmajor(m, K, nnz, rowptr, colind, val, (const float*)B, ldb, C, ldc, alpha, beta, (cudaStream_t)stream,
                        b_dtype == SPMMDB_DTYPE_BF16 ? 1 : 0, hot_bitmap);
}
