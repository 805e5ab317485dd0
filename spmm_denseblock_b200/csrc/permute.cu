// permute.cu -- apply a node permutation on the device: A' = P A P^T on a CSR matrix (relabel the
// columns, move row i to position old2new[i], re-sort every row ascending), B' = P B on the rows
// of a dense matrix, and the block heat map the reordering study looks at.  Integer work,
// bit-exact.  Replaces the CPU steps between "a reordering produced old2new" and "the SpMM is
// timed": permutate (reorder_strategy.cc:26-39, sortNeighbors :15-24), getHeatmap
// (utility.cc:71-88).  The reordering algorithms themselves stay on the CPU (north_star).
//
// Rows are sorted as 64-bit keys (new column << 32 | position in the old row), so a value array
// follows its entries with one gather: rows of <= 32 entries by a warp (bitonic network over
// shuffles), longer rows by a CTA in shared memory (two size classes), rows beyond 16384 entries
// by a CTA in a global staging area.
#include "common.cuh"

namespace spmmdb {
namespace {

constexpr int kSmallRow = 32;
constexpr int kMediumRow = 1024;
constexpr int kLargeRow = 16384;

// new_deg[old2new[i]] = deg(i); rows longer than a warp's worth go to one of three lists
__global__ void __launch_bounds__(256) perm_degree_kernel(int n, const int* __restrict__ rowptr,
                                                          const int* __restrict__ old2new, int* __restrict__ new_deg,
                                                          int* __restrict__ counters, int* __restrict__ list_m,
                                                          int* __restrict__ list_l, int* __restrict__ list_h) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int d = __ldg(rowptr + i + 1) - __ldg(rowptr + i);
  new_deg[__ldg(old2new + i)] = d;
  if (d > kLargeRow) { list_h[atomicAdd(counters + 2, 1)] = i; atomicMax(counters + 3, d); }
  else if (d > kMediumRow) list_l[atomicAdd(counters + 1, 1)] = i;
  else if (d > kSmallRow) list_m[atomicAdd(counters + 0, 1)] = i;
}

__device__ __forceinline__ unsigned long long shfl_xor_u64(unsigned long long v, int m) {
  const unsigned lo = __shfl_xor_sync(0xffffffffu, (unsigned)v, m);
  const unsigned hi = __shfl_xor_sync(0xffffffffu, (unsigned)(v >> 32), m);
  return ((unsigned long long)hi << 32) | lo;
}

// one warp per old row of at most 32 entries
__global__ void __launch_bounds__(256) perm_small_kernel(int n, const int* __restrict__ rowptr,
                                                         const int* __restrict__ colind, const float* __restrict__ val,
                                                         const int* __restrict__ old2new,
                                                         const int* __restrict__ new_rowptr,
                                                         int* __restrict__ new_colind, float* __restrict__ new_val) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t i = warp; i < n; i += nwarps) {
    const int s = __ldg(rowptr + i), d = __ldg(rowptr + i + 1) - s;
    if (d == 0 || d > kSmallRow) continue;
    unsigned long long key = ~0ull;
    if (lane < d) key = ((unsigned long long)(unsigned)__ldg(old2new + __ldg(colind + s + lane)) << 32) | (unsigned)lane;
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
      for (int j = k >> 1; j > 0; j >>= 1) {
        const unsigned long long other = shfl_xor_u64(key, j);
        const bool up = ((lane & k) == 0);
        const bool lower = ((lane & j) == 0);
        const bool take_min = (up == lower);
        key = take_min ? (key < other ? key : other) : (key > other ? key : other);
      }
    }
    if (lane < d) {
      const int64_t o = (int64_t)__ldg(new_rowptr + __ldg(old2new + i)) + lane;
      new_colind[o] = (int)(key >> 32);
      if (val != nullptr) new_val[o] = __ldg(val + s + (int)(unsigned)key);
    }
  }
}

// bitonic sort of `np2` keys by the whole CTA (shared or global memory)
__device__ __forceinline__ void cta_bitonic(unsigned long long* keys, int np2) {
  for (int k = 2; k <= np2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = threadIdx.x; t < np2 / 2; t += blockDim.x) {
        const int lo = ((t / j) * 2 * j) + (t % j);  // index of the lower partner
        const int hi = lo + j;
        const bool up = ((lo & k) == 0);
        const unsigned long long a = keys[lo], b = keys[hi];
        if ((a > b) == up) { keys[lo] = b; keys[hi] = a; }
      }
      __syncthreads();
    }
  }
}

// one CTA per listed row; keys staged in shared memory (cap > 0) or in `gstage` (cap == 0: the row's region
// starts at 2 * rowptr[i], pow2(d) < 2 d entries, disjoint between rows)
__global__ void __launch_bounds__(256) perm_cta_kernel(const int* __restrict__ list, const int* __restrict__ count,
                                                       int cap, const int* __restrict__ rowptr,
                                                       const int* __restrict__ colind, const float* __restrict__ val,
                                                       const int* __restrict__ old2new,
                                                       const int* __restrict__ new_rowptr, int* __restrict__ new_colind,
                                                       float* __restrict__ new_val, unsigned long long* gstage) {
  extern __shared__ unsigned long long skeys[];
  const int nlist = *count;
  for (int r = blockIdx.x; r < nlist; r += gridDim.x) {
    const int i = list[r];
    const int s = __ldg(rowptr + i), d = __ldg(rowptr + i + 1) - s;
    int np2 = 64;
    while (np2 < d) np2 <<= 1;
    unsigned long long* keys = cap > 0 ? skeys : gstage + 2 * (int64_t)s;
    for (int t = threadIdx.x; t < np2; t += blockDim.x)
      keys[t] = t < d ? (((unsigned long long)(unsigned)__ldg(old2new + __ldg(colind + s + t)) << 32) | (unsigned)t) : ~0ull;
    __syncthreads();
    cta_bitonic(keys, np2);
    const int64_t o = __ldg(new_rowptr + __ldg(old2new + i));
    for (int t = threadIdx.x; t < d; t += blockDim.x) {
      const unsigned long long k = keys[t];
      new_colind[o + t] = (int)(k >> 32);
      if (val != nullptr) new_val[o + t] = __ldg(val + s + (int)(unsigned)k);
    }
    __syncthreads();
  }
}

// out[old2new[r]][:] = in[r][:]
__global__ void __launch_bounds__(256) perm_rows_kernel(int rows, int K, const float* __restrict__ in, int64_t ldi,
                                                        const int* __restrict__ old2new, float* __restrict__ out,
                                                        int64_t ldo, int vec4) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < rows; r += nwarps) {
    const float* __restrict__ src = in + r * ldi;
    float* __restrict__ dst = out + (int64_t)__ldg(old2new + r) * ldo;
    if (vec4) {
      for (int c = lane; c < K / 4; c += 32) __stcs(reinterpret_cast<float4*>(dst) + c, __ldcs(reinterpret_cast<const float4*>(src) + c));
    } else {
      for (int c = lane; c < K; c += 32) dst[c] = src[c];
    }
  }
}

// heat[(row / bs) * nb + col / bs] += 1 (utility.cc:71-88)
__global__ void __launch_bounds__(256) heatmap_kernel(int n, int bs, int nb, const int* __restrict__ rowptr,
                                                      const int* __restrict__ colind, int* __restrict__ heat) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t i = warp; i < n; i += nwarps) {
    const int s = __ldg(rowptr + i), e = __ldg(rowptr + i + 1);
    int* __restrict__ hrow = heat + (int64_t)(i / bs) * nb;
    for (int p = s + lane; p < e; p += 32) atomicAdd(hrow + __ldg(colind + p) / bs, 1);
  }
}

}  // namespace

int exclusive_scan_i32(const int* counts, int n, int* sums, int* out, cudaStream_t st);  // convert.cu

}  // namespace spmmdb

using namespace spmmdb;

extern "C" int spmmdb_permute_csr_f32(int n, const int* rowptr, const int* colind, const float* val,
                                      const int* old2new, int* new_rowptr, int* new_colind, float* new_val,
                                      void* stream) {
  if (n < 0) return SPMMDB_ERR_INVALID_VALUE;
  if (new_rowptr == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) { SPMMDB_CUDA_OK(cudaMemsetAsync(new_rowptr, 0, sizeof(int), st)); return SPMMDB_OK; }
  if (rowptr == nullptr || old2new == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  if (val != nullptr && new_val == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  // scratch: counters[8] | new_deg[n] | scan sums | three row lists of n entries
  const int ntiles = (n + 2047) / 2048 + 1;
  const size_t ints = 8 + (size_t)n + ntiles + 8 + 3 * (size_t)n;
  int* ws = (int*)scratch(kScratchConvert, sizeof(int) * ints, st);
  if (ws == nullptr) return SPMMDB_ERR_ALLOC;
  int* counters = ws;
  int* new_deg = ws + 8;
  int* sums = new_deg + n;
  int* list_m = sums + ntiles + 8;
  int* list_l = list_m + n;
  int* list_h = list_l + n;
  SPMMDB_CUDA_OK(cudaMemsetAsync(counters, 0, 8 * sizeof(int), st));
  perm_degree_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, rowptr, old2new, new_deg, counters, list_m, list_l, list_h);
  SPMMDB_CUDA_OK(cudaGetLastError());
  int rc = exclusive_scan_i32(new_deg, n, sums, new_rowptr, st);
  if (rc != SPMMDB_OK) return rc;
  int h[4] = {0, 0, 0, 0};
  SPMMDB_CUDA_OK(cudaMemcpyAsync(h, counters, sizeof(h), cudaMemcpyDeviceToHost, st));
  SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
  const int nsm = device_num_sms() > 0 ? device_num_sms() : 148;
  perm_small_kernel<<<nsm * 8, 256, 0, st>>>(n, rowptr, colind, val, old2new, new_rowptr, new_colind, new_val);
  SPMMDB_CUDA_OK(cudaGetLastError());
  if (h[0] > 0) {
    const int grid = h[0] < nsm * 24 ? h[0] : nsm * 24;
    // 64-thread CTAs: these rows have 33..1024 entries, and four times as many of them are resident per SM
    perm_cta_kernel<<<grid, 64, kMediumRow * 8, st>>>(list_m, counters + 0, kMediumRow, rowptr, colind, val, old2new,
                                                      new_rowptr, new_colind, new_val, nullptr);
    SPMMDB_CUDA_OK(cudaGetLastError());
  }
  if (h[1] > 0) {
    SPMMDB_CUDA_OK(cudaFuncSetAttribute(perm_cta_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kLargeRow * 8));
    const int grid = h[1] < nsm ? h[1] : nsm;
    perm_cta_kernel<<<grid, 256, kLargeRow * 8, st>>>(list_l, counters + 1, kLargeRow, rowptr, colind, val, old2new,
                                                     new_rowptr, new_colind, new_val, nullptr);
    SPMMDB_CUDA_OK(cudaGetLastError());
  }
  if (h[2] > 0) {
    // rows beyond 16384 entries: stage in global memory, 2 * nnz keys (read nnz back: rowptr[n])
    int nnz = 0;
    SPMMDB_CUDA_OK(cudaMemcpyAsync(&nnz, rowptr + n, sizeof(int), cudaMemcpyDeviceToHost, st));
    SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
    unsigned long long* stage = (unsigned long long*)scratch(kScratchLayoutB, sizeof(unsigned long long) * 2 * (size_t)nnz, st);
    if (stage == nullptr) return SPMMDB_ERR_ALLOC;
    const int grid = h[2] < nsm * 2 ? h[2] : nsm * 2;
    perm_cta_kernel<<<grid, 256, 0, st>>>(list_h, counters + 2, 0, rowptr, colind, val, old2new, new_rowptr, new_colind,
                                          new_val, stage);
    SPMMDB_CUDA_OK(cudaGetLastError());
  }
  return SPMMDB_OK;
}

extern "C" int spmmdb_permute_rows_f32(int rows, int K, const float* in, int64_t ldi, const int* old2new, float* out,
                                       int64_t ldo, void* stream) {
  if (rows < 0 || K < 0 || ldi < K || ldo < K) return SPMMDB_ERR_INVALID_VALUE;
  if (rows == 0 || K == 0) return SPMMDB_OK;
  if (in == nullptr || out == nullptr || old2new == nullptr || in == out) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  const int nsm = device_num_sms() > 0 ? device_num_sms() : 148;
  const int vec4 = (K % 4 == 0) && (ldi % 4 == 0) && (ldo % 4 == 0) && (((uintptr_t)in | (uintptr_t)out) % 16 == 0);
  perm_rows_kernel<<<nsm * 16, 256, 0, (cudaStream_t)stream>>>(rows, K, in, ldi, old2new, out, ldo, vec4);
  SPMMDB_CUDA_OK(cudaGetLastError());
  return SPMMDB_OK;
}

extern "C" int spmmdb_heatmap(int n, const int* rowptr, const int* colind, int block_dim, int* heat, void* stream) {
  if (n < 0 || block_dim <= 0) return SPMMDB_ERR_INVALID_VALUE;
  if (n == 0) return SPMMDB_OK;
  if (rowptr == nullptr || heat == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  cudaStream_t st = (cudaStream_t)stream;
  const int nb = (n + block_dim - 1) / block_dim;
  SPMMDB_CUDA_OK(cudaMemsetAsync(heat, 0, sizeof(int) * (size_t)nb * nb, st));
  const int nsm = device_num_sms() > 0 ? device_num_sms() : 148;
  heatmap_kernel<<<nsm * 16, 256, 0, st>>>(n, block_dim, nb, rowptr, colind, heat);
  SPMMDB_CUDA_OK(cudaGetLastError());
  return SPMMDB_OK;
}
