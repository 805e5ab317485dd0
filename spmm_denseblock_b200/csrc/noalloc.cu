// noalloc.cu -- spmmdb_csrmm_noalloc_f32: C = alpha * A * B + beta * C (CSR x dense, row-major B and C) for callers
// that must not see a cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  No counterpart
// in the reference.  It is a variant of the row / long-row / reduce
// kernels of csrmm.cu whose scratch (long-row list, segment descriptors, per-segment partial sums) is a file-scope
// __device__ array of fixed size instead of the grow-only cudaMalloc'ed scratch of common.cu.  The entry point calls no
// cudaMalloc / cudaFree and no allocating library routine; every GPU that loads libspmmdb.so holds the array
// (kNaBytes, about 8.4 MB) until the process exits.
//
// Every call on a device shares the one array and finds the previous call's contents in it: the two counters are
// cleared at the start of each call, everything else is written before it is read.  Calls that overlap on the GPU
// (two streams, two host threads) race on it; calls on one stream are ordered by the stream.
//
// Sums: a row of at most kNaLongRow non-zeros is one fp32 FMA chain in CSR order (the reference kernel's order,
// gespmm_csrmm.h:95-166); a longer row is summed per segment of kNaSeg non-zeros and the segment sums are added in
// order, whether its segments found room in the array or (array full) the row was summed in place: the result does
// not depend on which of the two happened.
#include <mutex>

#include "common.cuh"

namespace spmmdb {
namespace {

constexpr int kNaMaxK = SPMMDB_NOALLOC_MAX_K;
constexpr int kNaSeg = 512;             // non-zeros per segment of a long row
constexpr int kNaLongRow = 2 * kNaSeg;  // rows beyond this are split
constexpr int kNaMaxLongRows = 16384;
constexpr int kNaMaxSegs = 16384;
constexpr int64_t kNaPartialFloats = (int64_t)1 << 21;  // segment capacity = min(kNaMaxSegs, kNaPartialFloats / K)

// layout of the array (bytes)
constexpr size_t kNaCountersOff = 0;                                                  // 64 ints
constexpr size_t kNaLongOff = 256;                                                    // int4 {row, base, nseg, 0}
constexpr size_t kNaSegOff = kNaLongOff + sizeof(int4) * (size_t)kNaMaxLongRows;      // int2 {row, segment}
constexpr size_t kNaPartialOff = kNaSegOff + sizeof(int2) * (size_t)kNaMaxSegs;       // floats
constexpr size_t kNaBytes = kNaPartialOff + sizeof(float) * (size_t)kNaPartialFloats;
static_assert(kNaPartialOff % 256 == 0, "partial sums start on a 256-byte boundary");
static_assert(kNaBytes == SPMMDB_NOALLOC_SCRATCH_BYTES, "spmmdb.h states the size of the array");

// The scratch.  In HOST code this name is a stand-in, not a device address: the kernels get the pointer that
// cudaGetSymbolAddress returns (na_scratch_ptr below).
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNaBytes];

struct NaArgs {
  int m, K;
  const int* rowptr;
  const int* colind;
  const float* val;
  const float* B;
  int64_t ldb;
  float* C;
  int64_t ldc;
  float alpha, beta;
  int seg_cap;  // segments that fit the partial-sum area at this K
  int* counters;
  int4* long_rows;
  int2* seg_desc;
  float* partial;
};

// V4: a lane owns 4 consecutive columns (tile of 128), else one column (tile of 32)
template <bool V4> struct NaAcc;
template <> struct NaAcc<true> {
  float4 v;
  static constexpr int kTile = 128;
  __device__ __forceinline__ void zero() { v = make_float4(0.f, 0.f, 0.f, 0.f); }
  __device__ __forceinline__ void fma(float w, const float* p) {
    const float4 b = __ldg(reinterpret_cast<const float4*>(p));
    v.x = fmaf(w, b.x, v.x); v.y = fmaf(w, b.y, v.y); v.z = fmaf(w, b.z, v.z); v.w = fmaf(w, b.w, v.w);
  }
  __device__ __forceinline__ void add(const NaAcc& o) { v.x += o.v.x; v.y += o.v.y; v.z += o.v.z; v.w += o.v.w; }
  __device__ __forceinline__ void add(const float* p) {
    const float4 b = *reinterpret_cast<const float4*>(p);
    v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w;
  }
  __device__ __forceinline__ void store(float* p) const { *reinterpret_cast<float4*>(p) = v; }
  __device__ __forceinline__ void store_c(float* p, float alpha, float beta) const {
    float4 o = make_float4(alpha * v.x, alpha * v.y, alpha * v.z, alpha * v.w);
    if (beta != 0.f) {
      const float4 c = *reinterpret_cast<const float4*>(p);
      o.x = fmaf(beta, c.x, o.x); o.y = fmaf(beta, c.y, o.y); o.z = fmaf(beta, c.z, o.z); o.w = fmaf(beta, c.w, o.w);
    }
    *reinterpret_cast<float4*>(p) = o;
  }
  static __device__ __forceinline__ int col(int c0, int lane) { return c0 + 4 * lane; }
};
template <> struct NaAcc<false> {
  float v;
  static constexpr int kTile = 32;
  __device__ __forceinline__ void zero() { v = 0.f; }
  __device__ __forceinline__ void fma(float w, const float* p) { v = fmaf(w, __ldg(p), v); }
  __device__ __forceinline__ void add(const NaAcc& o) { v += o.v; }
  __device__ __forceinline__ void add(const float* p) { v += *p; }
  __device__ __forceinline__ void store(float* p) const { *p = v; }
  __device__ __forceinline__ void store_c(float* p, float alpha, float beta) const {
    float o = alpha * v;
    if (beta != 0.f) o = fmaf(beta, *p, o);
    *p = o;
  }
  static __device__ __forceinline__ int col(int c0, int lane) { return c0 + lane; }
};

// acc += sum over non-zeros [s, e) of val * B[colind][col ...], one FMA chain in CSR order; the warp reads 32
// (column index, value) pairs at a time and hands them round by shuffle
template <bool V4>
__device__ __forceinline__ void na_accumulate(const NaArgs& a, int s, int e, int col, bool colok, int lane, NaAcc<V4>& acc) {
  for (int p = s; p < e; p += 32) {
    const int cnt = min(32, e - p);
    int ci = 0;
    float w = 0.f;
    if (lane < cnt) { ci = __ldg(a.colind + p + lane); w = __ldg(a.val + p + lane); }
    for (int j = 0; j < cnt; ++j) {
      const int c = __shfl_sync(0xffffffffu, ci, j);
      const float x = __shfl_sync(0xffffffffu, w, j);
      if (colok) acc.fma(x, a.B + (int64_t)c * a.ldb + col);
    }
  }
}

// One warp per row.  Rows of at most kNaLongRow non-zeros are finished here; a longer row reserves an entry of the
// long-row list and a run of segment slots (lane 0, two atomics) and leaves its sums to the next two kernels -- or,
// when either is full, is summed here, segment by segment.
template <bool V4>
__global__ void __launch_bounds__(256) na_rows_kernel(NaArgs a) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < a.m; r += nwarps) {
    const int s = __ldg(a.rowptr + r), e = __ldg(a.rowptr + r + 1);
    const int d = e - s;
    float* crow = a.C + r * a.ldc;
    if (d <= kNaLongRow) {
      for (int c0 = 0; c0 < a.K; c0 += NaAcc<V4>::kTile) {
        const int col = NaAcc<V4>::col(c0, lane);
        const bool colok = col < a.K;
        NaAcc<V4> acc;
        acc.zero();
        na_accumulate<V4>(a, s, e, col, colok, lane, acc);
        if (colok) acc.store_c(crow + col, a.alpha, a.beta);
      }
      continue;
    }
    const int nseg = (d + kNaSeg - 1) / kNaSeg;
    int idx = 0, base = 0;
    if (lane == 0) {
      idx = atomicAdd(a.counters + 0, 1);
      base = atomicAdd(a.counters + 1, nseg);
    }
    idx = __shfl_sync(0xffffffffu, idx, 0);
    base = __shfl_sync(0xffffffffu, base, 0);
    const bool fits = idx < kNaMaxLongRows && (int64_t)base + nseg <= a.seg_cap;
    if (idx < kNaMaxLongRows && lane == 0) a.long_rows[idx] = make_int4((int)r, base, fits ? nseg : 0, 0);
    // slots of this run inside the capacity: the row's segments, or "nothing here" when the run does not fit
    for (int g = lane; g < nseg && base + g < a.seg_cap; g += 32) a.seg_desc[base + g] = make_int2(fits ? (int)r : -1, g);
    if (fits) continue;
    for (int c0 = 0; c0 < a.K; c0 += NaAcc<V4>::kTile) {
      const int col = NaAcc<V4>::col(c0, lane);
      const bool colok = col < a.K;
      NaAcc<V4> total;
      total.zero();
      for (int g = 0; g < nseg; ++g) {
        NaAcc<V4> acc;
        acc.zero();
        na_accumulate<V4>(a, s + g * kNaSeg, min(e, s + (g + 1) * kNaSeg), col, colok, lane, acc);
        total.add(acc);
      }
      if (colok) total.store_c(crow + col, a.alpha, a.beta);
    }
  }
}

// one warp per reserved segment slot: partial[slot][0..K) = the segment's sums
template <bool V4>
__global__ void __launch_bounds__(256) na_segments_kernel(NaArgs a) {
  const int lane = threadIdx.x & 31;
  const int warp = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  const int nwarps = (int)(((int64_t)gridDim.x * blockDim.x) >> 5);
  const int nslots = min(a.counters[1], a.seg_cap);
  for (int slot = warp; slot < nslots; slot += nwarps) {
    const int2 sd = a.seg_desc[slot];
    if (sd.x < 0) continue;
    const int rs = __ldg(a.rowptr + sd.x), re = __ldg(a.rowptr + sd.x + 1);
    const int s = rs + sd.y * kNaSeg, e = min(re, s + kNaSeg);
    float* prow = a.partial + (int64_t)slot * a.K;
    for (int c0 = 0; c0 < a.K; c0 += NaAcc<V4>::kTile) {
      const int col = NaAcc<V4>::col(c0, lane);
      const bool colok = col < a.K;
      NaAcc<V4> acc;
      acc.zero();
      na_accumulate<V4>(a, s, e, col, colok, lane, acc);
      if (colok) acc.store(prow + col);
    }
  }
}

// one warp per listed long row: segment sums added in segment order
template <bool V4>
__global__ void __launch_bounds__(256) na_reduce_kernel(NaArgs a) {
  const int lane = threadIdx.x & 31;
  const int warp = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  const int nwarps = (int)(((int64_t)gridDim.x * blockDim.x) >> 5);
  const int nlong = min(a.counters[0], kNaMaxLongRows);
  for (int i = warp; i < nlong; i += nwarps) {
    const int4 lr = a.long_rows[i];
    if (lr.z == 0) continue;  // summed in place by na_rows_kernel
    float* crow = a.C + (int64_t)lr.x * a.ldc;
    for (int c0 = 0; c0 < a.K; c0 += NaAcc<V4>::kTile) {
      const int col = NaAcc<V4>::col(c0, lane);
      if (col >= a.K) continue;
      NaAcc<V4> total;
      total.zero();
      for (int g = 0; g < lr.z; ++g) total.add(a.partial + (int64_t)(lr.y + g) * a.K + col);
      total.store_c(crow + col, a.alpha, a.beta);
    }
  }
}

// Device address of the array, looked up once per device with that device current.
std::mutex g_na_mu;
unsigned char* g_na_ptr[16];

int na_scratch_ptr(unsigned char** out) {
  int dev = -1;
  SPMMDB_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 16) return SPMMDB_ERR_NOT_SUPPORTED;
  std::lock_guard<std::mutex> lk(g_na_mu);
  if (g_na_ptr[dev] == nullptr) {
    unsigned char* p = nullptr;
    SPMMDB_CUDA_OK(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
    if (p == nullptr) return SPMMDB_ERR_CUDA;
    g_na_ptr[dev] = p;
  }
  *out = g_na_ptr[dev];
  return SPMMDB_OK;
}

}  // namespace
}  // namespace spmmdb

using namespace spmmdb;

extern "C" int spmmdb_csrmm_noalloc_f32(int m, int k, int K, const int* rowptr, const int* colind, const float* val,
                                        const float* B, int64_t ldb, float* C, int64_t ldc, float alpha, float beta,
                                        void* stream) {
  if (m < 0 || k < 0 || K < 0) return SPMMDB_ERR_INVALID_VALUE;
  if (K > kNaMaxK) {
    char msg[160];
    snprintf(msg, sizeof(msg), "spmmdb_csrmm_noalloc_f32: K = %d exceeds SPMMDB_NOALLOC_MAX_K = %d (fixed scratch)", K, kNaMaxK);
    set_last_error(msg, cudaErrorInvalidValue, __FILE__, __LINE__);
    return SPMMDB_ERR_NOT_SUPPORTED;
  }
  if (m == 0 || K == 0) return SPMMDB_OK;
  if (rowptr == nullptr || colind == nullptr || val == nullptr || B == nullptr || C == nullptr)
    return SPMMDB_ERR_INVALID_VALUE;
  if (ldb < K || ldc < K) return SPMMDB_ERR_INVALID_VALUE;
  const int nsm = device_num_sms();
  if (nsm <= 0) return SPMMDB_ERR_CUDA;
  unsigned char* base = nullptr;
  int rc = na_scratch_ptr(&base);
  if (rc != SPMMDB_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;

  NaArgs a;
  a.m = m; a.K = K;
  a.rowptr = rowptr; a.colind = colind; a.val = val;
  a.B = B; a.ldb = ldb; a.C = C; a.ldc = ldc;
  a.alpha = alpha; a.beta = beta;
  const int64_t cap = kNaPartialFloats / K;
  a.seg_cap = (int)(cap < kNaMaxSegs ? cap : kNaMaxSegs);
  a.counters = reinterpret_cast<int*>(base + kNaCountersOff);
  a.long_rows = reinterpret_cast<int4*>(base + kNaLongOff);
  a.seg_desc = reinterpret_cast<int2*>(base + kNaSegOff);
  a.partial = reinterpret_cast<float*>(base + kNaPartialOff);
  // the array keeps the last call's contents: the counters must start at zero
  SPMMDB_CUDA_OK(cudaMemsetAsync(a.counters, 0, 256, st));
  const bool v4 = (K % 4 == 0) && (ldb % 4 == 0) && (ldc % 4 == 0) && (((uintptr_t)B | (uintptr_t)C) % 16 == 0);
  const int64_t want = ((int64_t)m + 7) / 8;
  const int grid_rows = (int)(want < (int64_t)nsm * 8 ? want : (int64_t)nsm * 8);
  if (v4) {
    na_rows_kernel<true><<<grid_rows, 256, 0, st>>>(a);
    na_segments_kernel<true><<<nsm * 4, 256, 0, st>>>(a);
    na_reduce_kernel<true><<<nsm, 256, 0, st>>>(a);
  } else {
    na_rows_kernel<false><<<grid_rows, 256, 0, st>>>(a);
    na_segments_kernel<false><<<nsm * 4, 256, 0, st>>>(a);
    na_reduce_kernel<false><<<nsm, 256, 0, st>>>(a);
  }
  SPMMDB_CUDA_OK(cudaGetLastError());
  return SPMMDB_OK;
}
