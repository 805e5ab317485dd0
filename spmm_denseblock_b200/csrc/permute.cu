// permute.cu -- apply a node permutation on the device: A' = P A P^T on a CSR matrix (relabel the
// columns, move row i to position old2new[i], re-sort every row ascending), B' = P B on the rows
// of a dense matrix, and the block heat map the reordering study looks at.  Integer work,
// bit-exact.  Replaces the CPU steps between "a reordering produced old2new" and "the SpMM is
// timed": permutate (reorder_strategy.cc:26-39, sortNeighbors :15-24), getHeatmap
// (utility.cc:71-88).  The reordering algorithms themselves stay on the CPU (north_star).
//
// Rows are sorted as keys (new column << bits | position in the old row), so a value array follows
// its entries with one gather: rows of <= 32 / 128 / 512 entries by a warp with 1 / 4 / 16 keys per
// lane in registers (bitonic network, shuffles below distance 32; 32-bit keys where column and
// position fit), longer rows by a CTA whose warps sort 512-key chunks in registers and merge through
// shared memory (two size classes), rows beyond 16384 entries by a CTA in a global staging area.
#include "common.cuh"

namespace spmmdb {
namespace {

constexpr int kWarpRow1 = 32;     // one key per lane
constexpr int kWarpRow4 = 128;    // four keys per lane
constexpr int kWarpRow16 = 512;   // sixteen keys per lane
constexpr int kSmallCtaRow = 1024; // CTA of 2 warps, 8 KB of keys in shared memory
constexpr int kMediumRow = 4096;  // CTA of 8 warps, 32 KB
constexpr int kLargeRow = 16384;  // CTA of 16 warps, 128 KB; longer rows are staged in global memory

// new_deg[old2new[i]] = deg(i); rows longer than a warp's worth go to one of six lists
__global__ void __launch_bounds__(256) perm_degree_kernel(int n, const int* __restrict__ rowptr,
                                                          const int* __restrict__ old2new, int* __restrict__ new_deg,
                                                          int* __restrict__ counters, int* __restrict__ list4,
                                                          int* __restrict__ list16, int* __restrict__ list_s,
                                                          int* __restrict__ list_m, int* __restrict__ list_l,
                                                          int* __restrict__ list_h, int rows_stay) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int d = __ldg(rowptr + i + 1) - __ldg(rowptr + i);
  if (!rows_stay) new_deg[__ldg(old2new + i)] = d;
  if (d > kLargeRow) list_h[atomicAdd(counters + 5, 1)] = i;
  else if (d > kMediumRow) list_l[atomicAdd(counters + 4, 1)] = i;
  else if (d > kSmallCtaRow) list_m[atomicAdd(counters + 3, 1)] = i;
  else if (d > kWarpRow16) list_s[atomicAdd(counters + 2, 1)] = i;
  else if (d > kWarpRow4) list16[atomicAdd(counters + 1, 1)] = i;
  else if (d > kWarpRow1) list4[atomicAdd(counters + 0, 1)] = i;
}

// Sort keys: (new column << pbits) | position in the old row.  64-bit keys (pbits = 32) always work; when the
// column and the position fit 32 bits together (n <= 2^(32 - pbits)) the warp kernels sort 32-bit keys, at a third
// of the instructions per compare-exchange.
__device__ __forceinline__ unsigned shfl_xor_key(unsigned v, int m) { return __shfl_xor_sync(0xffffffffu, v, m); }
__device__ __forceinline__ unsigned long long shfl_xor_key(unsigned long long v, int m) {
  const unsigned lo = __shfl_xor_sync(0xffffffffu, (unsigned)v, m);
  const unsigned hi = __shfl_xor_sync(0xffffffffu, (unsigned)(v >> 32), m);
  return ((unsigned long long)hi << 32) | lo;
}

// One compare-exchange pass at distance 2^LJ of bitonic stage k over 32 * E keys held by a warp, key (e, lane) being
// element base + 32 e + lane of the sequence: ascending where (element & k) == 0.  Distances below 32 are shuffles,
// the others stay inside a lane.
template <int E, typename K, int LJ>
__device__ __forceinline__ void bitonic_pass(K (&key)[E], int lane, unsigned k, unsigned base) {
  constexpr int j = 1 << LJ;
  if (j >= 32) {
    constexpr int je = j >> 5;
#pragma unroll
    for (int e = 0; e < E; ++e) {
      if ((e & je) == 0) {
        const bool up = (((base | (unsigned)(e << 5)) & k) == 0);
        const K a = key[e], b = key[e | je];
        const K mn = a < b ? a : b, mx = a < b ? b : a;
        key[e] = up ? mn : mx;
        key[e | je] = up ? mx : mn;
      }
    }
  } else {
    const bool lower = (lane & j) == 0;
#pragma unroll
    for (int e = 0; e < E; ++e) {
      const bool up = (((base | (unsigned)(e << 5) | (unsigned)lane) & k) == 0);
      const K mine = key[e], other = shfl_xor_key(mine, j);
      const K mn = mine < other ? mine : other, mx = mine < other ? other : mine;
      key[e] = (up == lower) ? mn : mx;
    }
  }
}

// passes at distances 2^LJ, 2^(LJ-1), ..., 1 of stage k
template <int E, typename K, int LJ>
struct BitonicTail {
  static __device__ __forceinline__ void run(K (&key)[E], int lane, unsigned k, unsigned base) {
    bitonic_pass<E, K, LJ>(key, lane, k, base);
    BitonicTail<E, K, LJ - 1>::run(key, lane, k, base);
  }
};
template <int E, typename K>
struct BitonicTail<E, K, -1> {
  static __device__ __forceinline__ void run(K (&)[E], int, unsigned, unsigned) {}
};

// stages k = 2, 4, ..., 2^LK: a full sort of the warp's 32 * E keys when LK = log2(32 E).  `base` only matters for
// the last stage (the direction of this chunk inside a longer sequence).
template <int E, typename K, int LK>
struct BitonicSort {
  static __device__ __forceinline__ void run(K (&key)[E], int lane, unsigned base) {
    BitonicSort<E, K, LK - 1>::run(key, lane, base);
    constexpr bool last = (1 << LK) == 32 * E;
    BitonicTail<E, K, LK - 1>::run(key, lane, 1u << LK, last ? base : 0u);
  }
};
template <int E, typename K>
struct BitonicSort<E, K, 0> {
  static __device__ __forceinline__ void run(K (&)[E], int, unsigned) {}
};
template <int E> struct Log2Keys;
template <> struct Log2Keys<4> { static constexpr int v = 7; };
template <> struct Log2Keys<16> { static constexpr int v = 9; };

// One warp per listed old row of at most 32 * E entries (E = 4, 16).
template <int E, typename K>
__global__ void __launch_bounds__(128) perm_warp_kernel(int n, const int* __restrict__ list, const int* __restrict__ count,
                                                        const int* __restrict__ rowptr, const int* __restrict__ colind,
                                                        const float* __restrict__ val, const int* __restrict__ old2new,
                                                        const int* __restrict__ new_rowptr, int* __restrict__ new_colind,
                                                        float* __restrict__ new_val, int pbits, int rows_stay) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int total = *count;
  const K pos_mask = (((K)1) << pbits) - 1;
  for (int64_t r = warp; r < total; r += nwarps) {
    const int i = __ldg(list + r);
    const int s = __ldg(rowptr + i), d = __ldg(rowptr + i + 1) - s;
    K key[E];
#pragma unroll
    for (int e = 0; e < E; ++e) {
      const int t = (e << 5) | lane;
      key[e] = ~(K)0;
      if (t < d) key[e] = ((K)(unsigned)__ldg(old2new + __ldg(colind + s + t)) << pbits) | (K)(unsigned)t;
    }
    BitonicSort<E, K, Log2Keys<E>::v>::run(key, lane, 0u);
    const int64_t o = __ldg(new_rowptr + (rows_stay ? i : __ldg(old2new + i)));
#pragma unroll
    for (int e = 0; e < E; ++e) {
      const int t = (e << 5) | lane;
      if (t < d) {
        new_colind[o + t] = (int)(key[e] >> pbits);
        if (val != nullptr) new_val[o + t] = __ldg(val + s + (int)(key[e] & pos_mask));
      }
    }
  }
}

// Rows of at most 32 entries, R consecutive old rows per warp at a time: the loads of the R rows (row bounds, column
// ids, their new ids, the new row start) are independent and overlap, which is what this latency-bound kernel needs.
template <typename K, int R>
__global__ void __launch_bounds__(128) perm_warp1_kernel(int n, const int* __restrict__ rowptr, const int* __restrict__ colind,
                                                         const float* __restrict__ val, const int* __restrict__ old2new,
                                                         const int* __restrict__ new_rowptr, int* __restrict__ new_colind,
                                                         float* __restrict__ new_val, int pbits, int rows_stay) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const K pos_mask = (((K)1) << pbits) - 1;
  for (int64_t r0 = warp * R; r0 < n; r0 += nwarps * R) {
    int s[R], d[R];
    int64_t o[R];
    K key[R][1];
#pragma unroll
    for (int u = 0; u < R; ++u) {
      const int64_t i = r0 + u;
      s[u] = 0; d[u] = 0; o[u] = 0;
      if (i < n) {
        s[u] = __ldg(rowptr + i);
        d[u] = __ldg(rowptr + i + 1) - s[u];
        if (d[u] > kWarpRow1) d[u] = 0;
      }
    }
#pragma unroll
    for (int u = 0; u < R; ++u) {
      key[u][0] = ~(K)0;
      if (lane < d[u]) key[u][0] = ((K)(unsigned)__ldg(old2new + __ldg(colind + s[u] + lane)) << pbits) | (K)(unsigned)lane;
      if (d[u] > 0) o[u] = __ldg(new_rowptr + (rows_stay ? (int)(r0 + u) : __ldg(old2new + r0 + u)));
    }
#pragma unroll
    for (int u = 0; u < R; ++u) BitonicSort<1, K, 5>::run(key[u], lane, 0u);
#pragma unroll
    for (int u = 0; u < R; ++u) {
      if (lane < d[u]) {
        new_colind[o[u] + lane] = (int)(key[u][0] >> pbits);
        if (val != nullptr) new_val[o[u] + lane] = __ldg(val + s[u] + (int)(key[u][0] & pos_mask));
      }
    }
  }
}

// One CTA (2 to 16 warps) per listed row of more than 512 entries: 64-bit keys staged in shared memory (cap > 0) or in
// `gstage` (cap == 0: the row's region starts at 2 * rowptr[i]; pow2(d) < 2 d entries, disjoint between rows).
// Chunks of 512 keys are sorted in registers by a warp; every later stage runs its distances >= 512 in the staging
// memory, one pass each, and its distances below 512 in registers again, one round trip per stage.
__global__ void __launch_bounds__(512) perm_cta_kernel(const int* __restrict__ list, const int* __restrict__ count,
                                                       int cap, const int* __restrict__ rowptr,
                                                       const int* __restrict__ colind, const float* __restrict__ val,
                                                       const int* __restrict__ old2new,
                                                       const int* __restrict__ new_rowptr, int* __restrict__ new_colind,
                                                       float* __restrict__ new_val, unsigned long long* gstage,
                                                       int rows_stay) {
  typedef unsigned long long K;
  constexpr int E = 16, kChunk = 32 * E;
  extern __shared__ unsigned long long skeys[];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int nlist = *count;
  for (int r = blockIdx.x; r < nlist; r += gridDim.x) {
    const int i = list[r];
    const int s = __ldg(rowptr + i), d = __ldg(rowptr + i + 1) - s;
    int np2 = 2 * kChunk;
    while (np2 < d) np2 <<= 1;
    K* keys = cap > 0 ? skeys : gstage + 2 * (int64_t)s;
    K key[E];
    for (int base = wid * kChunk; base < np2; base += nw * kChunk) {
#pragma unroll
      for (int e = 0; e < E; ++e) {
        const int t = base | (e << 5) | lane;
        key[e] = ~(K)0;
        if (t < d) key[e] = ((K)(unsigned)__ldg(old2new + __ldg(colind + s + t)) << 32) | (K)(unsigned)t;
      }
      BitonicSort<E, K, 9>::run(key, lane, (unsigned)base);
#pragma unroll
      for (int e = 0; e < E; ++e) keys[base | (e << 5) | lane] = key[e];
    }
    __syncthreads();
    for (int lk = 10; (1 << lk) <= np2; ++lk) {
      const unsigned k = 1u << lk;
      for (int lj = lk - 1; lj >= 9; --lj) {
        const int j = 1 << lj;
        for (int t = threadIdx.x; t < np2 / 2; t += blockDim.x) {
          const int lo = ((t >> lj) << (lj + 1)) | (t & (j - 1));
          const bool up = ((lo & k) == 0);
          const K a = keys[lo], b = keys[lo + j];
          if ((a > b) == up) { keys[lo] = b; keys[lo + j] = a; }
        }
        __syncthreads();
      }
      for (int base = wid * kChunk; base < np2; base += nw * kChunk) {
#pragma unroll
        for (int e = 0; e < E; ++e) key[e] = keys[base | (e << 5) | lane];
        BitonicTail<E, K, 8>::run(key, lane, k, (unsigned)base);
#pragma unroll
        for (int e = 0; e < E; ++e) keys[base | (e << 5) | lane] = key[e];
      }
      __syncthreads();
    }
    const int64_t o = __ldg(new_rowptr + (rows_stay ? i : __ldg(old2new + i)));
    for (int t = threadIdx.x; t < d; t += blockDim.x) {
      const K k = keys[t];
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

namespace spmmdb {
// The relabel-and-sort pass behind spmmdb_permute_csr_f32.  rows_stay == 0: A' = P A P^T (rows move to old2new[i],
// new_rowptr is written).  rows_stay == 1: rows keep their place (new_rowptr must be rowptr) and only the columns are
// relabelled and sorted -- new_colind[t] = old2new[column], ascending within the row: every adjacency list ordered
// by a per-vertex key (spmmdb_order_rcm: the key is the degree rank).
int permute_csr_impl(int n, const int* rowptr, const int* colind, const float* val, const int* old2new, int rows_stay,
                     int* new_rowptr, int* new_colind, float* new_val, cudaStream_t st) {
  // scratch: counters[8] | new_deg[n] | scan sums | six row lists of n entries
  const int ntiles = (n + 2047) / 2048 + 1;
  const size_t ints = 8 + (size_t)n + ntiles + 8 + 6 * (size_t)n;
  int* ws = (int*)scratch(kScratchConvert, sizeof(int) * ints, st);
  if (ws == nullptr) return SPMMDB_ERR_ALLOC;
  int* counters = ws;
  int* new_deg = ws + 8;
  int* sums = new_deg + n;
  int* list4 = sums + ntiles + 8;
  int* list16 = list4 + n;
  int* list_s = list16 + n;
  int* list_m = list_s + n;
  int* list_l = list_m + n;
  int* list_h = list_l + n;
  SPMMDB_CUDA_OK(cudaMemsetAsync(counters, 0, 8 * sizeof(int), st));
  perm_degree_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, rowptr, old2new, new_deg, counters, list4, list16, list_s, list_m,
                                                      list_l, list_h, rows_stay);
  SPMMDB_CUDA_OK(cudaGetLastError());
  if (!rows_stay) {
    int rc = exclusive_scan_i32(new_deg, n, sums, new_rowptr, st);
    if (rc != SPMMDB_OK) return rc;
  }
  int h[6] = {0, 0, 0, 0, 0, 0};
  SPMMDB_CUDA_OK(cudaMemcpyAsync(h, counters, sizeof(h), cudaMemcpyDeviceToHost, st));
  SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
  const int nsm = device_num_sms() > 0 ? device_num_sms() : 148;
  // 32-bit keys where (column, position) fit: n < 2^(32 - log2(keys per warp))
  const bool wide = SPMMDB_TUNE("PERM_KEY64", 0) != 0;
  auto narrow = [&](int pbits) { return !wide && (int64_t)n < ((int64_t)1 << (32 - pbits)); };
  auto warp_grid = [&](int64_t rows) { return (int)(rows < (int64_t)nsm * 64 ? (rows + 3) / 4 : (int64_t)nsm * 16); };
#define SPMMDB_PERM_WARP(E, PBITS, LIST, COUNT, ROWS)                                                              \
  do {                                                                                                             \
    if (narrow(PBITS))                                                                                             \
      perm_warp_kernel<E, unsigned><<<warp_grid(ROWS), 128, 0, st>>>(n, LIST, COUNT, rowptr, colind, val, old2new, \
                                                                     new_rowptr, new_colind, new_val, PBITS,       \
                                                                     rows_stay);                                   \
    else                                                                                                           \
      perm_warp_kernel<E, unsigned long long><<<warp_grid(ROWS), 128, 0, st>>>(                                    \
          n, LIST, COUNT, rowptr, colind, val, old2new, new_rowptr, new_colind, new_val, 32, rows_stay);           \
    SPMMDB_CUDA_OK(cudaGetLastError());                                                                            \
  } while (0)
  {
    const int64_t want = ((int64_t)n + 15) / 16;  // 4 warps x 4 rows per CTA and iteration
    const int grid = (int)(want < (int64_t)nsm * 16 ? want : (int64_t)nsm * 16);
    if (narrow(5))
      perm_warp1_kernel<unsigned, 4><<<grid, 128, 0, st>>>(n, rowptr, colind, val, old2new, new_rowptr, new_colind, new_val, 5,
                                                           rows_stay);
    else
      perm_warp1_kernel<unsigned long long, 4><<<grid, 128, 0, st>>>(n, rowptr, colind, val, old2new, new_rowptr, new_colind,
                                                                     new_val, 32, rows_stay);
    SPMMDB_CUDA_OK(cudaGetLastError());
  }
  if (h[0] > 0) SPMMDB_PERM_WARP(4, 7, list4, counters + 0, h[0]);
  if (h[1] > 0) SPMMDB_PERM_WARP(16, 9, list16, counters + 1, h[1]);
#undef SPMMDB_PERM_WARP
  if (h[2] > 0) {
    const int grid = h[2] < nsm * 12 ? h[2] : nsm * 12;
    perm_cta_kernel<<<grid, 64, kSmallCtaRow * 8, st>>>(list_s, counters + 2, kSmallCtaRow, rowptr, colind, val, old2new,
                                                       new_rowptr, new_colind, new_val, nullptr, rows_stay);
    SPMMDB_CUDA_OK(cudaGetLastError());
  }
  if (h[3] > 0) {
    const int grid = h[3] < nsm * 3 ? h[3] : nsm * 3;
    perm_cta_kernel<<<grid, 256, kMediumRow * 8, st>>>(list_m, counters + 3, kMediumRow, rowptr, colind, val, old2new,
                                                       new_rowptr, new_colind, new_val, nullptr, rows_stay);
    SPMMDB_CUDA_OK(cudaGetLastError());
  }
  if (h[4] > 0) {
    SPMMDB_CUDA_OK(cudaFuncSetAttribute(perm_cta_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kLargeRow * 8));
    const int grid = h[4] < nsm ? h[4] : nsm;
    perm_cta_kernel<<<grid, 512, kLargeRow * 8, st>>>(list_l, counters + 4, kLargeRow, rowptr, colind, val, old2new,
                                                     new_rowptr, new_colind, new_val, nullptr, rows_stay);
    SPMMDB_CUDA_OK(cudaGetLastError());
  }
  if (h[5] > 0) {
    // rows beyond 16384 entries: stage in global memory, 2 * nnz keys (read nnz back: rowptr[n])
    int nnz = 0;
    SPMMDB_CUDA_OK(cudaMemcpyAsync(&nnz, rowptr + n, sizeof(int), cudaMemcpyDeviceToHost, st));
    SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
    unsigned long long* stage = (unsigned long long*)scratch(kScratchLayoutB, sizeof(unsigned long long) * 2 * (size_t)nnz, st);
    if (stage == nullptr) return SPMMDB_ERR_ALLOC;
    const int grid = h[5] < nsm * 2 ? h[5] : nsm * 2;
    perm_cta_kernel<<<grid, 512, 0, st>>>(list_h, counters + 5, 0, rowptr, colind, val, old2new, new_rowptr, new_colind,
                                          new_val, stage, rows_stay);
    SPMMDB_CUDA_OK(cudaGetLastError());
  }
  return SPMMDB_OK;
}
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
  return permute_csr_impl(n, rowptr, colind, val, old2new, 0, new_rowptr, new_colind, new_val, st);
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
