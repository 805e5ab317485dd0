// csrmm.cu -- CSR x dense (fp32) for sm_100a.  HBM/L2-bandwidth-bound gather:
// no tensor cores.  Replaces gespmm_csrmm.h:95-166 (spmm_test2) and the
// cuSPARSE csrmm / csrmm2 call sites (run_csrmm.cu:133-142).
//
// Mapping: one "group" of LPR lanes owns one row segment; every lane owns VEC
// vectors (float4 when 16-byte aligned, else float) of the output row, so a
// B-row gather is one fully coalesced LPR*16-byte request and the (col, val)
// pairs are read once, coalesced, and broadcast with warp shuffles.  U*VEC
// independent 16-byte loads are in flight per lane before the FMAs retire them
// in CSR order, so each output element is the same fp32 FMA chain as the
// reference kernel (bit-exact) while the memory system sees deep MLP.
//
// Rows longer than `seg` non-zeros are cut into row-relative segments of `seg`
// non-zeros: the main kernel queues them, the long-row kernel computes one
// partial per segment and the group that finishes a row's last segment adds
// the partials in segment order.  The split depends only on the row, so the
// result is deterministic and independent of any row sharding.
#include <stdlib.h>

#include <atomic>
#include <map>
#include <mutex>
#include <utility>

#include <cuda.h>

#include "common.cuh"

namespace spmmdb {
bool make_map_rows(CUtensorMap* map, bool bf16, const void* base, uint64_t cols, uint64_t rows, uint64_t row_bytes);  // bsrmm_tc.cu
namespace {

struct LongRow { int row; int seg_base; int nseg; int epoch; };

struct CsrArgs {
  int m, K;
  const int* rowptr;
  const int* colind;
  const float* val;
  const float* B;       // fp32 rows, or (b_bf16) bf16 rows behind the same pointer
  int64_t ldb;          // in elements of B
  int b_rows;           // rows of B (the TMA-gather kernel's tensor map needs the extent)
  int b_bf16;           // stream kernel only: B holds bf16 (8 columns per 16-byte vector)
  const unsigned* hot;  // stream kernel only: bitmap of the columns whose B rows should stay in L2 (or null)
  int hot_policy;       // L2 eviction class of the hot rows: 0 evict_last, 1 evict_normal
  int cold_policy;      // L2 eviction class of the other rows: 0 normal, 1 evict_first, 2 evict_unchanged
  float* C;
  int64_t ldc;
  float alpha, beta;
  int seg;              // row segment length (power of two)
  // long-row queue (scratch)
  int* epoch_dev;       // device word holding the call number (see csrmm_begin_kernel)
  int* counters;        // [0] = long rows, [1] = segments, [2] = overflow flag, [3] = work items taken,
                        // [4] = segments taken (stream path)
  LongRow* long_rows;   // capacity cap_rows
  int* seg_owner;       // capacity cap_segs: segment -> long row slot
  unsigned long long* seg_entry;  // stream path: (epoch << 32 | slot), published AFTER long_rows[slot]
  int* done;            // capacity cap_rows, zero between calls
  float* partial;       // capacity cap_segs * Kpad floats
  int cap_rows, cap_segs;
  int Kpad;             // K rounded up to 4
  int epoch;            // call counter: tells this call's queue entries from stale ones.  Filled in ON THE DEVICE from
                        // counters[7] (csrmm_begin_kernel increments it once per call), not by the host
  int rpi;              // stream kernel: rows per work item (<= 32)
  int inline_segs;      // stream kernel: warps take queued segments between items (else: long-row kernel after)
  int prequeued;        // the long-row queue was built by csrmm_scan_long_kernel: row kernels only skip long rows
};

// First launch of every call: clears the queue counters and advances the call number (a scratch word of its own that
// is never re-allocated while the other queue buffers live, so numbers never repeat against their stale entries).  The
// number is device state so that a call captured into a CUDA graph gets a fresh one on every replay (stale queue
// entries of the previous replay must not look current).
__global__ void csrmm_begin_kernel(int* counters, int* epoch_dev) {
  const int l = threadIdx.x;
  if (l < 8) counters[l] = 0;
  if (l == 8) {
    int e = *epoch_dev + 1;
    if (e <= 0) e = 1;
    *epoch_dev = e;
  }
}

template <typename V> struct VecOps;
template <> struct VecOps<float4> {
  static constexpr int W = 4;
  __device__ static float4 zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }
  __device__ static void fma(float4& a, float v, const float4& b) {
    a.x = fmaf(v, b.x, a.x); a.y = fmaf(v, b.y, a.y); a.z = fmaf(v, b.z, a.z); a.w = fmaf(v, b.w, a.w);
  }
  __device__ static void add(float4& a, const float4& b) { a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
  __device__ static float4 axpby(float al, const float4& a, float be, const float4& c) {
    return make_float4(fmaf(be, c.x, al * a.x), fmaf(be, c.y, al * a.y), fmaf(be, c.z, al * a.z), fmaf(be, c.w, al * a.w));
  }
  __device__ static float4 scale(float al, const float4& a) { return make_float4(al * a.x, al * a.y, al * a.z, al * a.w); }
};
template <> struct VecOps<float> {
  static constexpr int W = 1;
  __device__ static float zero() { return 0.f; }
  __device__ static void fma(float& a, float v, const float& b) { a = fmaf(v, b, a); }
  __device__ static void add(float& a, const float& b) { a += b; }
  __device__ static float axpby(float al, const float& a, float be, const float& c) { return fmaf(be, c, al * a); }
  __device__ static float scale(float al, const float& a) { return al * a; }
};

// Accumulate one row segment [lb, hb) into acc[VEC].  Must be called by all 32
// lanes of the warp (shuffles); groups whose range is empty idle.
// nvec = number of V-vectors per row of B that this column tile covers,
// vbase = first vector index of the tile.
// `c`, `v` hold the group's first chunk (entry lb + sub, or 0 past hb).  PF: the next chunk is loaded before
// the current one is consumed, so only the first chunk of a row is on the latency chain (two more live
// registers: used where rows / segments span several chunks).
template <typename V, int LPR, int VEC, int U, bool PF, typename Args>
__device__ __forceinline__ void accumulate_preloaded(const Args& a, int lb, int hb, int sub, int grp,
                                                     int vbase, int nvec, V (&acc)[VEC], int c, float v) {
  using Ops = VecOps<V>;
  const V* __restrict__ Bv = reinterpret_cast<const V*>(a.B);
  const int64_t ldbv = a.ldb / Ops::W;
  bool colok[VEC];
#pragma unroll
  for (int i = 0; i < VEC; ++i) colok[i] = (sub + i * LPR) < nvec;

  static_assert(LPR % U == 0, "a gather batch must not straddle two chunks");
  for (int base = lb; __any_sync(0xffffffffu, base < hb); base += LPR) {
    const int nidx = base + LPR + sub;
    int cn = 0;
    float vn = 0.f;
    if (PF && nidx < hb) {
      cn = __ldg(a.colind + nidx);
      vn = __ldg(a.val + nidx);
    }
    const int cnt = hb - base;  // entries of this group's chunk (may be <= 0 or > LPR)
#pragma unroll 1
    for (int kk = 0; kk < LPR; kk += U) {
      if (__all_sync(0xffffffffu, kk >= cnt)) break;
      V b[U][VEC];
      float vv[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int src = grp * LPR + kk + u;
        const int cc = __shfl_sync(0xffffffffu, c, src);
        vv[u] = __shfl_sync(0xffffffffu, v, src);
        const V* __restrict__ brow = Bv + (int64_t)cc * ldbv + vbase + sub;
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
          if (kk + u < cnt && colok[i]) b[u][i] = __ldg(brow + i * LPR);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (kk + u < cnt) {
#pragma unroll
          for (int i = 0; i < VEC; ++i) {
            if (colok[i]) Ops::fma(acc[i], vv[u], b[u][i]);
          }
        }
      }
    }
    if (PF) {
      c = cn;
      v = vn;
    } else if (nidx < hb) {
      c = __ldg(a.colind + nidx);
      v = __ldg(a.val + nidx);
    }
  }
}

template <typename V, int LPR, int VEC, int U, bool PF = false, typename Args>
__device__ __forceinline__ void accumulate_segment(const Args& a, int lb, int hb, int sub, int grp,
                                                   int vbase, int nvec, V (&acc)[VEC]) {
  int c = 0;
  float v = 0.f;
  if (lb + sub < hb) {
    c = __ldg(a.colind + lb + sub);
    v = __ldg(a.val + lb + sub);
  }
  accumulate_preloaded<V, LPR, VEC, U, PF>(a, lb, hb, sub, grp, vbase, nvec, acc, c, v);
}

template <typename V, int LPR, int VEC, typename Args>
__device__ __forceinline__ void store_row(const Args& a, int row, int sub, int vbase, int nvec,
                                          const V (&acc)[VEC]) {
  using Ops = VecOps<V>;
  V* __restrict__ Cv = reinterpret_cast<V*>(a.C) + (int64_t)row * (a.ldc / Ops::W) + vbase + sub;
#pragma unroll
  for (int i = 0; i < VEC; ++i) {
    if (sub + i * LPR < nvec) {
      V out;
      if (a.beta != 0.f) out = Ops::axpby(a.alpha, acc[i], a.beta, Cv[i * LPR]);
      else if (a.alpha != 1.f) out = Ops::scale(a.alpha, acc[i]);
      else out = acc[i];
      __stcs(Cv + i * LPR, out);
    }
  }
}

// Main kernel: one group (LPR lanes) per row.  K > 32 lanes * VEC vectors is
// covered by looping over column tiles inside the kernel.
template <typename V, int LPR, int VEC, int U>
__global__ void __launch_bounds__(256) csrmm_rows_kernel(const CsrArgs a_in) {
  CsrArgs a = a_in;
  a.epoch = __ldcg(a_in.epoch_dev);  // this call's number lives on the device: a captured CUDA graph replays correctly
  constexpr int GPW = 32 / LPR;
  using Ops = VecOps<V>;
  const int lane = threadIdx.x & 31;
  const int sub = lane % LPR, grp = lane / LPR;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t row64 = warp * GPW + grp;
  const int nvec_total = (a.K + Ops::W - 1) / Ops::W;

  const bool valid = row64 < a.m;
  const int row = valid ? (int)row64 : 0;
  int lb = 0, hb = 0;
  if (valid) {
    lb = __ldg(a.rowptr + row);
    hb = __ldg(a.rowptr + row + 1);
  }
  // Long row: hand it to the long-row kernel.  If the queue is full (caller
  // passed a too small nnz) the row is processed here, slowly but correctly.
  const bool is_long = (hb - lb) > a.seg;
  const int nseg = is_long ? (hb - lb + a.seg - 1) / a.seg : 0;
  int slot = -1, seg_base = 0;
  if (is_long && sub == 0) {
    slot = atomicAdd(a.counters + 0, 1);
    seg_base = atomicAdd(a.counters + 1, nseg);
    if (slot >= a.cap_rows || seg_base + nseg > a.cap_segs) {
      atomicExch(a.counters + 2, 1);
      slot = -1;
    } else {
      a.long_rows[slot] = LongRow{row, seg_base, nseg, a.epoch};
    }
  }
  slot = __shfl_sync(0xffffffffu, slot, grp * LPR);
  seg_base = __shfl_sync(0xffffffffu, seg_base, grp * LPR);
  const bool queued = is_long && slot >= 0;
  if (queued) {
    for (int s = sub; s < nseg; s += LPR) a.seg_owner[seg_base + s] = slot;
    hb = lb;
  }

  for (int vbase = 0; vbase < nvec_total; vbase += LPR * VEC) {
    const int nvec = min(LPR * VEC, nvec_total - vbase);
    V acc[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) acc[i] = Ops::zero();
    accumulate_segment<V, LPR, VEC, U>(a, lb, hb, sub, grp, vbase, nvec, acc);
    if (valid && !queued) store_row<V, LPR, VEC>(a, row, sub, vbase, nvec, acc);
  }
}

// Rows kernel for small short-row matrices whose dense width fits one column tile (K <= 4 * LPR * VEC):
// persistent groups walk rows gid, gid + G, gid + 2G, ... and load the NEXT row's first chunk of
// (column, value) pairs and the row pointers of the row after it before gathering for the current
// row, so the rowptr -> colind -> B chain of a row overlaps the gathers of the row before it.  Same
// arithmetic per row as csrmm_rows_kernel (bit-identical output), same long-row queue.
template <typename V, int LPR, int VEC, int U>
__global__ void __launch_bounds__(256, 3) csrmm_rows_pipe_kernel(const CsrArgs a_in) {
  CsrArgs a = a_in;
  a.epoch = __ldcg(a_in.epoch_dev);  // this call's number lives on the device: a captured CUDA graph replays correctly
  constexpr int GPW = 32 / LPR;
  using Ops = VecOps<V>;
  const int lane = threadIdx.x & 31;
  const int sub = lane % LPR, grp = lane / LPR;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t G = (((int64_t)gridDim.x * blockDim.x) >> 5) * GPW;
  const int nvec = (a.K + Ops::W - 1) / Ops::W;

  int64_t row = warp * GPW + grp;
  int lb = 0, hb = 0;
  if (row < a.m) {
    lb = __ldg(a.rowptr + row);
    hb = __ldg(a.rowptr + row + 1);
  }
  int c = 0;
  float v = 0.f;
  if (lb + sub < hb && hb - lb <= a.seg) {
    c = __ldg(a.colind + lb + sub);
    v = __ldg(a.val + lb + sub);
  }
  int nlb = 0, nhb = 0;
  if (row + G < a.m) {
    nlb = __ldg(a.rowptr + row + G);
    nhb = __ldg(a.rowptr + row + G + 1);
  }
  while (__any_sync(0xffffffffu, row < a.m)) {
    int cn = 0;
    float vn = 0.f;
    if (nlb + sub < nhb && nhb - nlb <= a.seg) {
      cn = __ldg(a.colind + nlb + sub);
      vn = __ldg(a.val + nlb + sub);
    }
    int nnlb = 0, nnhb = 0;
    if (row + 2 * G < a.m) {
      nnlb = __ldg(a.rowptr + row + 2 * G);
      nnhb = __ldg(a.rowptr + row + 2 * G + 1);
    }
    const bool valid = row < a.m;
    const bool is_long = (hb - lb) > a.seg;
    bool queued = false;
    if (a.prequeued) {
      queued = is_long;
      if (queued) hb = lb;
    } else if (__any_sync(0xffffffffu, is_long)) {
      const int nseg = is_long ? (hb - lb + a.seg - 1) / a.seg : 0;
      int slot = -1, seg_base = 0;
      if (is_long && sub == 0) {
        slot = atomicAdd(a.counters + 0, 1);
        seg_base = atomicAdd(a.counters + 1, nseg);
        if (slot >= a.cap_rows || seg_base + nseg > a.cap_segs) {
          atomicExch(a.counters + 2, 1);
          slot = -1;
        } else {
          a.long_rows[slot] = LongRow{(int)row, seg_base, nseg, a.epoch};
        }
      }
      slot = __shfl_sync(0xffffffffu, slot, grp * LPR);
      seg_base = __shfl_sync(0xffffffffu, seg_base, grp * LPR);
      queued = is_long && slot >= 0;
      if (queued) {
        for (int s = sub; s < nseg; s += LPR) a.seg_owner[seg_base + s] = slot;
        hb = lb;
      } else if (is_long && lb + sub < hb) {  // queue full: done here, the first chunk was not preloaded
        c = __ldg(a.colind + lb + sub);
        v = __ldg(a.val + lb + sub);
      }
    }
    V acc[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) acc[i] = Ops::zero();
    accumulate_preloaded<V, LPR, VEC, U, false>(a, lb, hb, sub, grp, 0, nvec, acc, c, v);
    if (valid && !queued) store_row<V, LPR, VEC>(a, (int)row, sub, 0, nvec, acc);
    row += G;
    lb = nlb; hb = nhb; c = cn; v = vn;
    nlb = nnlb; nhb = nnhb;
  }
}

// Queue a row longer than `seg` for the long-row kernel (one lane; what csrmm_rows_kernel does with a group).
// Returns false when the queue is full: the caller then processes the row itself.
__device__ __forceinline__ bool queue_long_row(const CsrArgs& a, int row, int len) {
  const int nseg = (len + a.seg - 1) / a.seg;
  const int slot = atomicAdd(a.counters + 0, 1);
  const int seg_base = atomicAdd(a.counters + 1, nseg);
  if (slot >= a.cap_rows || seg_base + nseg > a.cap_segs) {
    atomicExch(a.counters + 2, 1);
    return false;
  }
  a.long_rows[slot] = LongRow{row, seg_base, nseg, a.epoch};
  for (int s = 0; s < nseg; ++s) a.seg_owner[seg_base + s] = slot;
  return true;
}

// Builds the long-row queue ahead of the row kernels (overlapped small-matrix path): a lane per row, rows longer
// than `seg` are queued.  With the queue complete before anything else starts, the long-row kernels can run on
// a second stream next to the rows kernel instead of after it.  A row that does not fit the queue (the caller's
// nnz was too small) is computed here by the whole warp, one float4 / float column per lane, as the sequential
// chain over the whole row -- what the rows kernels do with such a row.
template <typename V>
__global__ void __launch_bounds__(256) csrmm_scan_long_kernel(const CsrArgs a_in) {
  CsrArgs a = a_in;
  a.epoch = __ldcg(a_in.epoch_dev);  // this call's number lives on the device: a captured CUDA graph replays correctly
  using Ops = VecOps<V>;
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
  const int nvec = (a.K + Ops::W - 1) / Ops::W;
  const V* __restrict__ Bv = reinterpret_cast<const V*>(a.B);
  const int64_t ldbv = a.ldb / Ops::W;
  const int64_t rows_padded = ((int64_t)a.m + 31) & ~(int64_t)31;
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < rows_padded; row += nthreads) {
    int lb = 0, hb = 0;
    if (row < a.m) { lb = __ldg(a.rowptr + row); hb = __ldg(a.rowptr + row + 1); }
    bool failed = false;
    if (hb - lb > a.seg) failed = !queue_long_row(a, (int)row, hb - lb);
    unsigned todo = __ballot_sync(FULL, failed);
    while (todo) {
      const int src = __ffs(todo) - 1;
      todo &= todo - 1;
      const int r = (int)(row - lane + src);
      const int l0 = __shfl_sync(FULL, lb, src), h0 = __shfl_sync(FULL, hb, src);
      for (int j = lane; j < nvec; j += 32) {
        V acc[1] = {Ops::zero()};
        for (int q = l0; q < h0; ++q) Ops::fma(acc[0], __ldg(a.val + q), __ldg(Bv + (int64_t)__ldg(a.colind + q) * ldbv + j));
        store_row<V, 32, 1>(a, r, 0, j, nvec - j, acc);
      }
    }
  }
}

// Second half of the long-row path: one warp per queued row adds that row's partial rows strictly in segment
// order (the order every path uses, so results do not depend on which kernel ran) and writes the row of C.
// One vector column per lane, RB partial rows in flight per lane: a hub row with hundreds of segments was the
// critical path when the last warp to finish a segment did this inside the long-row kernel.
template <typename V>
__global__ void __launch_bounds__(256) csrmm_reduce_kernel(const CsrArgs a_in) {
  CsrArgs a = a_in;
  a.epoch = __ldcg(a_in.epoch_dev);  // this call's number lives on the device: a captured CUDA graph replays correctly
  using Ops = VecOps<V>;
  constexpr int RB = 16;
  const int lane = threadIdx.x & 31;
  const int warp = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  const int nwarps = (int)(((int64_t)gridDim.x * blockDim.x) >> 5);
  const int nrows = min(*(volatile int*)(a.counters + 0), a.cap_rows);
  const int nvec_total = (a.K + Ops::W - 1) / Ops::W;
  const int kpadv = a.Kpad / Ops::W;
  const V* __restrict__ Pv = reinterpret_cast<const V*>(a.partial);
  for (int slot = warp; slot < nrows; slot += nwarps) {
    const LongRow lr = a.long_rows[slot];
    if (lr.epoch != a.epoch || lr.seg_base < 0 || (int64_t)lr.seg_base + lr.nseg > a.cap_segs) continue;
    for (int j = lane; j < nvec_total; j += 32) {
      const V* __restrict__ p = Pv + (int64_t)lr.seg_base * kpadv + j;
      V acc = Ops::zero();
      for (int s0 = 0; s0 < lr.nseg; s0 += RB) {
        V t[RB];
#pragma unroll
        for (int jj = 0; jj < RB; ++jj)
          if (s0 + jj < lr.nseg) t[jj] = __ldcg(p + (int64_t)(s0 + jj) * kpadv);
#pragma unroll
        for (int jj = 0; jj < RB; ++jj)
          if (s0 + jj < lr.nseg) {
            if (s0 + jj == 0) acc = t[jj]; else Ops::add(acc, t[jj]);
          }
      }
      V* __restrict__ Cv = reinterpret_cast<V*>(a.C) + (int64_t)lr.row * (a.ldc / Ops::W) + j;
      V out;
      if (a.beta != 0.f) out = Ops::axpby(a.alpha, acc, a.beta, *Cv);
      else if (a.alpha != 1.f) out = Ops::scale(a.alpha, acc);
      else out = acc;
      __stcs(Cv, out);
    }
  }
}

// Long-row kernel: one group per queued segment (all column tiles), partial
// sums to scratch; csrmm_reduce_kernel adds them afterwards.
template <typename V, int LPR, int VEC, int U>
__global__ void __launch_bounds__(256) csrmm_long_kernel(const CsrArgs a_in) {
  CsrArgs a = a_in;
  a.epoch = __ldcg(a_in.epoch_dev);  // this call's number lives on the device: a captured CUDA graph replays correctly
  constexpr int GPW = 32 / LPR;
  using Ops = VecOps<V>;
  const int lane = threadIdx.x & 31;
  const int sub = lane % LPR, grp = lane / LPR;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int nseg_total = min(*(volatile int*)(a.counters + 1), a.cap_segs);
  const int nvec_total = (a.K + Ops::W - 1) / Ops::W;
  const int kpadv = a.Kpad / Ops::W;
  V* __restrict__ Pv = reinterpret_cast<V*>(a.partial);

  for (int64_t item0 = warp * GPW; item0 < nseg_total; item0 += nwarps * GPW) {
    const int64_t item = item0 + grp;
    bool valid = item < nseg_total;
    int slot = 0;
    LongRow lr = LongRow{0, 0, 0, 0};
    int lb = 0, hb = 0;
    if (valid) {
      slot = a.seg_owner[item];
      valid = slot >= 0 && slot < a.cap_rows;
      if (valid) {
        lr = a.long_rows[slot];
        // entries left over from an overflowed / earlier queue fail this test
        valid = lr.epoch == a.epoch && item >= lr.seg_base && item < (int64_t)lr.seg_base + lr.nseg;
      }
    }
    if (valid) {
      const int s = (int)item - lr.seg_base;
      const int row_lb = __ldg(a.rowptr + lr.row);
      const int row_hb = __ldg(a.rowptr + lr.row + 1);
      lb = row_lb + s * a.seg;
      hb = min(row_hb, lb + a.seg);
    }
    for (int vbase = 0; vbase < nvec_total; vbase += LPR * VEC) {
      const int nvec = min(LPR * VEC, nvec_total - vbase);
      V acc[VEC];
#pragma unroll
      for (int i = 0; i < VEC; ++i) acc[i] = Ops::zero();
      accumulate_segment<V, LPR, VEC, U, true>(a, lb, hb, sub, grp, vbase, nvec, acc);
      if (valid) {
#pragma unroll
        for (int i = 0; i < VEC; ++i)
          if (sub + i * LPR < nvec) Pv[item * kpadv + vbase + sub + i * LPR] = acc[i];
      }
    }
  }
}

// ---------------------------------------------------------------------------
// Segment queue of the stream path.  Producers (any lane that meets a row longer than `seg`) take a
// slot and a range of tickets, write the row record, fence, publish one 64-bit entry per ticket
// (epoch << 32 | slot) and release that many units of a counting semaphore.  Consumers take a unit
// of the semaphore (so the number of claims never passes the number of published tickets), then
// the next ticket index, and wait for that entry if another producer is still between its
// allocation and its publication (it publishes without waiting on anybody).  A CAS on the ticket
// index instead of the semaphore serialised the whole GPU on one round trip per claim.  A row record that did not fit the
// queue publishes void entries and is processed by its producer instead.
__device__ __forceinline__ int publish_long_row(const CsrArgs& a, int row, int len) {
  const int nseg = (len + a.seg - 1) / a.seg;
  const int slot = atomicAdd(a.counters + 0, 1);
  const int seg_base = atomicAdd(a.counters + 1, nseg);
  const bool ok = slot < a.cap_rows && seg_base + nseg <= a.cap_segs;
  if (ok) {
    a.long_rows[slot] = LongRow{row, seg_base, nseg, a.epoch};
    __threadfence();
  } else {
    atomicExch(a.counters + 2, 1);  // queue full (caller's nnz was too small): the row is done in place
  }
  if (!a.inline_segs) {
    if (ok) for (int t = 0; t < nseg; ++t) a.seg_owner[seg_base + t] = slot;
    return ok ? 1 : 0;
  }
  const unsigned long long ent = ((unsigned long long)(unsigned)a.epoch << 32) | (ok ? (unsigned)slot : 0xffffffffu);
  int pub = 0;
  for (int t = 0; t < nseg; ++t)
    if (seg_base + t < a.cap_segs) { *(volatile unsigned long long*)(a.seg_entry + seg_base + t) = ent; ++pub; }
  __threadfence();
  atomicAdd(a.counters + 5, pub);  // semaphore: tickets that may be claimed
  return ok ? 1 : 0;
}

// ---------------------------------------------------------------------------
// Stream kernel (K % 4 == 0, K <= 512): persistent warps pull work items of
// `rpi` consecutive rows from a global counter and walk the item's non-zero
// stream [rowptr[r0], rowptr[r0 + rpi]) in coalesced chunks of 32 (col, val)
// pairs, the next chunk prefetched while the current one is consumed.  B-row
// gathers are issued U rows at a time ACROSS row boundaries, so short rows keep
// the same memory-level parallelism as long ones; the FMA phase then walks the
// batch in CSR order and stores a row whenever the stream crosses its end.  All
// control flow is warp-uniform (every lane handles the same non-zero, VEC
// float4 columns of it), and each output element is still the sequential fp32
// FMA chain of the reference kernel.
//
// Rows longer than `seg` are cut into row-relative segments and queued; a queued
// segment is just a one-row work item whose output is a partial-sum row in
// scratch, so the SAME loop body serves both.  Warps take queued segments before
// the next item: a hub row is then gathered while the rows around it are being
// streamed and its neighbourhood of B is in L2 (a separate launch afterwards
// re-read it from DRAM: 9 GB per call on the products-shaped graph).  The warp
// that finishes a row's last segment adds the partials in segment order, so the
// result does not depend on who computed what.  The kernel is launched twice:
// the second launch finds the item counter exhausted and only sweeps segments
// queued after a warp's last look at the queue.
// How the stream kernel reads a 16-byte vector of a B row and multiplies it in.
//   fp32 B: 4 columns per vector (one float4 of accumulators), bf16 B: 8 columns (two float4).
// HINT: the load carries an L2 eviction policy (createpolicy): evict_last for the rows of "hot" columns
// (those the plan marked to stay resident), the caller's choice for the rest.
template <bool BF16> struct BVec;
template <> struct BVec<false> {
  using Raw = float4;
  static constexpr int OV = 1;  // float4 accumulators per raw vector
  __device__ static void fma(float4 (&acc)[1], float v, const Raw& b) { VecOps<float4>::fma(acc[0], v, b); }
};
template <> struct BVec<true> {
  using Raw = uint4;
  static constexpr int OV = 2;
  __device__ static void fma(float4 (&acc)[2], float v, const Raw& b) {
    // bf16 -> fp32 is a 16-bit shift: element 2j in the low half of word j, element 2j + 1 in the high half
    acc[0].x = fmaf(v, __uint_as_float(b.x << 16), acc[0].x);
    acc[0].y = fmaf(v, __uint_as_float(b.x & 0xffff0000u), acc[0].y);
    acc[0].z = fmaf(v, __uint_as_float(b.y << 16), acc[0].z);
    acc[0].w = fmaf(v, __uint_as_float(b.y & 0xffff0000u), acc[0].w);
    acc[1].x = fmaf(v, __uint_as_float(b.z << 16), acc[1].x);
    acc[1].y = fmaf(v, __uint_as_float(b.z & 0xffff0000u), acc[1].y);
    acc[1].z = fmaf(v, __uint_as_float(b.w << 16), acc[1].z);
    acc[1].w = fmaf(v, __uint_as_float(b.w & 0xffff0000u), acc[1].w);
  }
};
__device__ __forceinline__ float4 ldg_raw(const float4* p) { return __ldg(p); }
__device__ __forceinline__ uint4 ldg_raw(const uint4* p) { return __ldg(p); }
__device__ __forceinline__ float4 ldg_policy(const float4* p, uint64_t pol) {
  float4 r;
  asm volatile("ld.global.nc.L2::cache_hint.v4.f32 {%0, %1, %2, %3}, [%4], %5;"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p), "l"(pol));
  return r;
}
__device__ __forceinline__ uint4 ldg_policy(const uint4* p, uint64_t pol) {
  uint4 r;
  asm volatile("ld.global.nc.L2::cache_hint.v4.u32 {%0, %1, %2, %3}, [%4], %5;"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p), "l"(pol));
  return r;
}
__device__ __forceinline__ uint64_t policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 0f3F800000;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 0f3F800000;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t policy_evict_normal() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 0f3F800000;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t policy_evict_unchanged() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_unchanged.b64 %0, 0f3F800000;" : "=l"(p));
  return p;
}

// ---- what the persistent kernels (stream, lean, TMA gather) share: how a warp picks its next unit of work and how a
// queued long-row segment is finished.
struct WorkUnit {
  int t;          // >= 0: ticket of a queued segment; -1: a work item of consecutive rows
  int next_item;  // lane 0: the item claimed for the round after this one
  int nrows;      // rows of the unit (1 for a segment)
  int s, e;       // lane r < nrows: non-zero range of row r (segment: lane 0)
  int lng;        // lane r: row r is long and was queued (not computed here)
  int ldo;        // output row stride in float4 (0: the unit's single partial-sum row)
  float4* obase;  // output row 0
};
enum { kWorkRun = 0, kWorkVoid = 1, kWorkNone = 2 };

// A queued segment if there is one (counting semaphore counters[5], ticket index counters[4]), else the item already
// claimed.  kWorkVoid: the ticket's row did not fit the queue (nothing to do, look again); kWorkNone: out of items.
__device__ __forceinline__ int pick_work(const CsrArgs& a, int lane, int item, int nitems, int kpadv, int& pending, WorkUnit& w) {
  constexpr unsigned FULL = 0xffffffffu;
  int t = -1;
  if (pending) {
    if (lane == 0 && *(volatile int*)(a.counters + 5) > 0) {
      if (atomicSub(a.counters + 5, 1) > 0) t = atomicAdd(a.counters + 4, 1);
      else atomicAdd(a.counters + 5, 1);
    }
    t = __shfl_sync(FULL, t, 0);
  }
  pending = 0;
  w.t = t;
  w.next_item = item;
  w.nrows = 1; w.s = 0; w.e = 0; w.lng = 0;
  if (t >= 0) {
    volatile unsigned long long* pe = a.seg_entry + t;
    unsigned long long ent;
    do { ent = *pe; } while ((unsigned)(ent >> 32) != (unsigned)a.epoch);
    __threadfence();
    pending = 1;  // keep draining the queue
    if ((unsigned)ent == 0xffffffffu) return kWorkVoid;
    const int4 raw = __ldcg(reinterpret_cast<const int4*>(a.long_rows + (unsigned)ent));  // written during this kernel
    if (lane == 0) {
      w.s = __ldg(a.rowptr + raw.x) + (t - raw.y) * a.seg;
      w.e = min(__ldg(a.rowptr + raw.x + 1), w.s + a.seg);
    }
    // partial sums are stored raw; alpha / beta are applied by the final add
    w.obase = reinterpret_cast<float4*>(a.partial) + (int64_t)t * kpadv;
    w.ldo = 0;
    return kWorkRun;
  }
  if (item >= nitems) return kWorkNone;
  if (lane == 0) {
    w.next_item = atomicAdd(a.counters + 3, 1);  // consumed at the bottom of the caller's loop
    // queue look, also consumed at the bottom (a little stale is fine: the second launch sweeps up)
    pending = a.inline_segs && *(volatile int*)(a.counters + 5) > 0;
  }
  const int r0 = item * a.rpi;
  w.nrows = min(a.rpi, a.m - r0);
  if (lane < w.nrows) {
    w.s = __ldg(a.rowptr + r0 + lane);
    w.e = __ldg(a.rowptr + r0 + lane + 1);
  }
  w.lng = (w.e - w.s) > a.seg;
  if (__any_sync(FULL, w.lng)) {
    pending = a.inline_segs;
    if (w.lng) w.lng = publish_long_row(a, r0 + lane, w.e - w.s);
  }
  w.ldo = (int)(a.ldc >> 2);
  w.obase = reinterpret_cast<float4*>(a.C) + (int64_t)r0 * w.ldo;
  return kWorkRun;
}

// The partial sum of ticket t is written: publish it; the warp that finishes the row's last segment adds the partials
// in segment order (alpha / beta applied here) and restores the row's `done` counter.  `acc` is scratch of the caller
// (the stream kernel hands over its accumulators, which are free at this point).
template <int VEC, int OV>
__device__ __forceinline__ void finish_queued_segment(const CsrArgs& a, int t, int lane, int kpadv, const bool (&colok)[VEC],
                                                      float4 (&acc)[VEC][OV]) {
  using Ops = VecOps<float4>;
  constexpr unsigned FULL = 0xffffffffu;
  __threadfence();
  __syncwarp();
  const int seg_slot = (int)(unsigned)*(volatile unsigned long long*)(a.seg_entry + t);
  const int4 raw = __ldcg(reinterpret_cast<const int4*>(a.long_rows + seg_slot));
  const int r0 = raw.x, seg_base = raw.y, seg_nseg = raw.z;
  int last = 0;
  if (lane == 0) last = (atomicAdd(a.done + seg_slot, 1) == seg_nseg - 1);
  last = __shfl_sync(FULL, last, 0);
  if (!last) return;
  __threadfence();
  if (lane == 0) a.done[seg_slot] = 0;
  const float4* __restrict__ p = reinterpret_cast<const float4*>(a.partial) + (int64_t)seg_base * kpadv;
  for (int q = 0; q < seg_nseg; ++q, p += kpadv) {
#pragma unroll
    for (int i = 0; i < VEC; ++i)
      if (colok[i]) {
#pragma unroll
        for (int h = 0; h < OV; ++h) {
          const float4 x = __ldcg(p + OV * (lane + 32 * i) + h);
          if (q == 0) acc[i][h] = x; else Ops::add(acc[i][h], x);
        }
      }
  }
  float4* __restrict__ Cv = reinterpret_cast<float4*>(a.C) + (int64_t)r0 * (a.ldc >> 2);
#pragma unroll
  for (int i = 0; i < VEC; ++i)
    if (colok[i]) {
#pragma unroll
      for (int h = 0; h < OV; ++h) {
        float4* o = Cv + OV * (lane + 32 * i) + h;
        float4 out;
        if (a.beta != 0.f) out = Ops::axpby(a.alpha, acc[i][h], a.beta, *o);
        else if (a.alpha != 1.f) out = Ops::scale(a.alpha, acc[i][h]);
        else out = acc[i][h];
        __stcs(o, out);
      }
    }
}

template <int VEC, int U, int MINB, bool BF16, bool HINT>
__global__ void __launch_bounds__(256, MINB) csrmm_stream_kernel(const CsrArgs a_in) {
  CsrArgs a = a_in;
  a.epoch = __ldcg(a_in.epoch_dev);  // this call's number lives on the device: a captured CUDA graph replays correctly
  using Ops = VecOps<float4>;
  using BV = BVec<BF16>;
  using Raw = typename BV::Raw;
  constexpr int OV = BV::OV;
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int nvec = a.K / (4 * OV);  // 16-byte vectors per row of B
  const int nitems = (a.m + a.rpi - 1) / a.rpi;
  const Raw* __restrict__ Bv = reinterpret_cast<const Raw*>(a.B) + lane;
  const int64_t ldbv = a.ldb / (4 * OV);
  const int kpadv = a.Kpad >> 2;
  bool colok[VEC];
#pragma unroll
  for (int i = 0; i < VEC; ++i) colok[i] = (lane + 32 * i) < nvec;
  uint64_t pol_hot = 0, pol_cold = 0;
  if (HINT) {
    pol_hot = a.hot_policy == 1 ? policy_evict_normal() : policy_evict_last();
    pol_cold = a.cold_policy == 1 ? policy_evict_first() : (a.cold_policy == 2 ? policy_evict_unchanged() : policy_evict_normal());
  }
  // (column, value) chunk loads; with HINT the column's bit of the hot bitmap rides in the sign bit
  auto load_cv = [&](int idx, int& c, float& v) {
    c = __ldg(a.colind + idx);
    v = __ldg(a.val + idx);
    if (HINT) c |= (int)(((__ldg(a.hot + (c >> 5)) >> (c & 31)) & 1u) << 31);
  };

  int item = 0;
  if (lane == 0) item = atomicAdd(a.counters + 3, 1);
  item = __shfl_sync(FULL, item, 0);
  int pending = a.inline_segs;  // look at the queue before the first item: the second launch has only the queue left
  for (;;) {
    // ---- pick work: a queued segment if there is one, else the item already claimed
    WorkUnit w;
    const int what = pick_work(a, lane, item, nitems, kpadv, pending, w);
    if (what == kWorkNone) break;
    if (what == kWorkVoid) continue;
    const int t = w.t, next_item = w.next_item, nrows = w.nrows, s = w.s, e = w.e, lng = w.lng, ldo = w.ldo;
    float4* __restrict__ obase = w.obase;  // output row 0 of this work unit
    const int S = __shfl_sync(FULL, s, 0);
    const int E = __shfl_sync(FULL, e, nrows - 1);

    // current row state (warp-uniform)
    int cr = 0;
    int ce = __shfl_sync(FULL, e, 0);
    int cl = __shfl_sync(FULL, lng, 0);
    float4 acc[VEC][OV];
#pragma unroll
    for (int i = 0; i < VEC; ++i)
#pragma unroll
      for (int h = 0; h < OV; ++h) acc[i][h] = Ops::zero();

    auto flush_advance = [&]() {
      if (!cl) {
        float4* __restrict__ Cv = obase + (int64_t)cr * ldo;
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
          if (colok[i]) {
#pragma unroll
            for (int h = 0; h < OV; ++h) {
              float4* o = Cv + OV * (lane + 32 * i) + h;
              float4 out;
              if (t >= 0) out = acc[i][h];
              else if (a.beta != 0.f) out = Ops::axpby(a.alpha, acc[i][h], a.beta, *o);
              else if (a.alpha != 1.f) out = Ops::scale(a.alpha, acc[i][h]);
              else out = acc[i][h];
              __stcs(o, out);
            }
          }
        }
      }
#pragma unroll
      for (int i = 0; i < VEC; ++i)
#pragma unroll
        for (int h = 0; h < OV; ++h) acc[i][h] = Ops::zero();
      ++cr;
      const int src = min(cr, 31);
      ce = __shfl_sync(FULL, e, src);
      cl = __shfl_sync(FULL, lng, src);
      if (cr >= nrows) { ce = 0x7fffffff; cl = 1; }
    };

    int p0 = S & ~31;  // chunk base (128-byte aligned loads of colind / val)
    int qn = S;        // next non-zero to consume
    int c = 0;
    float v = 0.f;
    {
      const int idx = p0 + lane;
      if (idx >= S && idx < E) load_cv(idx, c, v);
    }
    while (p0 < E) {
      int c_n = 0;
      float v_n = 0.f;
      {
        const int idx = p0 + 32 + lane;
        if (idx < E) load_cv(idx, c_n, v_n);
      }
      const int qhi = min(p0 + 32, E);
#pragma unroll 1
      for (int kk = (qn - p0) / U * U; p0 + kk < qhi; kk += U) {
        Raw b[U][VEC];
        float vv[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          int cc = __shfl_sync(FULL, c, kk + u);
          vv[u] = __shfl_sync(FULL, v, kk + u);
          const int q = p0 + kk + u;
          if (q >= qn && q < qhi) {
            if (HINT) {
              const uint64_t pol = cc < 0 ? pol_hot : pol_cold;
              cc &= 0x7fffffff;
              const Raw* __restrict__ brow = Bv + (int64_t)cc * ldbv;
#pragma unroll
              for (int i = 0; i < VEC; ++i)
                if (colok[i]) b[u][i] = ldg_policy(brow + 32 * i, pol);
            } else {
              const Raw* __restrict__ brow = Bv + (int64_t)cc * ldbv;
#pragma unroll
              for (int i = 0; i < VEC; ++i)
                if (colok[i]) b[u][i] = ldg_raw(brow + 32 * i);
            }
          }
        }
        const int q0 = p0 + kk;
        if (q0 >= qn && q0 + U <= qhi && q0 + U <= ce && !cl) {
          // whole batch inside the current row
#pragma unroll
          for (int u = 0; u < U; ++u)
#pragma unroll
            for (int i = 0; i < VEC; ++i)
              if (colok[i]) BV::fma(acc[i], vv[u], b[u][i]);
        } else {
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int q = q0 + u;
            if (q >= qn && q < qhi) {
              while (q >= ce) flush_advance();
              if (!cl) {
#pragma unroll
                for (int i = 0; i < VEC; ++i)
                  if (colok[i]) BV::fma(acc[i], vv[u], b[u][i]);
              }
            }
          }
        }
      }
      qn = qhi;
      p0 += 32;
      c = c_n;
      v = v_n;
      if (qn < E) {
        while (qn >= ce) flush_advance();
        if (cl) {
          // inside a queued long row: jump to its end
          qn = min(ce, E);
          const int np0 = qn & ~31;
          if (np0 != p0) {
            p0 = np0;
            const int idx = p0 + lane;
            c = 0; v = 0.f;
            if (idx >= qn && idx < E) load_cv(idx, c, v);
          }
        }
      }
    }
    while (cr < nrows) flush_advance();  // last row and trailing empty rows

    if (t >= 0) {
      finish_queued_segment<VEC, OV>(a, t, lane, kpadv, colok, acc);
    } else {
      item = __shfl_sync(FULL, next_item, 0);
    }
    pending = __shfl_sync(FULL, pending, 0);
  }
}




// ---------------------------------------------------------------------------
// Lean stream kernel: csrmm_stream_kernel's work distribution (items of `rpi` rows from a global counter, queued
// long-row segments taken between items, partial sums added in segment order by the last finisher) and its
// register gather, with the inner loop rebuilt around what ncu showed (profiles/r2c_*, r2d_*): the stream kernel
// executes 44 (fp32) / 54 (bf16 B) warp instructions per non-zero for its 8 FMAs -- per-element range and row-end
// tests with divergence scaffolding (BRA + BSSY + BSYNC = 9-11 per non-zero) -- and is issue-bound wherever DRAM
// is not the limit (products-shaped graph with locality, bf16 B).  Here
//   * batches are 8 non-zeros aligned to 8 ABSOLUTE indices, so a batch never straddles a 32-entry
//     (column, value) chunk and its shuffles use constant lane offsets;
//   * the row ends that fall inside a batch are ONE warp reduction (an 8-bit mask in a uniform register).  No
//     row end (the common case): 8 x (value, FMAs) with no test at all.  Otherwise the batch is swept once per
//     row piece with uniform branches around whole non-zeros, and the row store appears once in the code.
// Each output element is still the sequential fp32 FMA chain in CSR order (bit-identical to csrmm_stream_kernel).
// (A shared-memory ring fed by cp.async / cp.async.bulk instead of the register gather was built and measured in
// r2b / r2d: correct, but LDGSTS costs 4 issue slots with its three padding instructions on this chip, UBLKCP ~10
// per row through ELECT / R2UR, and the kernel is issue-bound: 5.0-6.4 ms where this one takes 3.1-4.6 ms: profiles/r2b, r2d, r2e.)
template <int VEC, int MINB, bool BF16>
__global__ void __launch_bounds__(256, MINB) csrmm_lean_kernel(const CsrArgs a_in) {
  CsrArgs a = a_in;
  a.epoch = __ldcg(a_in.epoch_dev);  // this call's number lives on the device: a captured CUDA graph replays correctly
  constexpr int G = 8;
  using Ops = VecOps<float4>;
  using BV = BVec<BF16>;
  using Raw = typename BV::Raw;
  constexpr int OV = BV::OV;
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int nvec = a.K / (4 * OV);                 // 16-byte vectors per row of B
  const Raw* __restrict__ Bv = reinterpret_cast<const Raw*>(a.B) + lane;
  const uint32_t ldbv = (uint32_t)(a.ldb / (4 * OV));  // row pitch in 16-byte vectors (< 2^32: checked by the host)
  const int nitems = (a.m + a.rpi - 1) / a.rpi;
  const int kpadv = a.Kpad >> 2;
  bool colok[VEC];
#pragma unroll
  for (int i = 0; i < VEC; ++i) colok[i] = (lane + 32 * i) < nvec;

  int item = 0;
  if (lane == 0) item = atomicAdd(a.counters + 3, 1);
  item = __shfl_sync(FULL, item, 0);
  int pending = a.inline_segs;
  for (;;) {
    // ---- pick work: a queued segment if there is one, else the item already claimed
    WorkUnit w;
    const int what = pick_work(a, lane, item, nitems, kpadv, pending, w);
    if (what == kWorkNone) break;
    if (what == kWorkVoid) continue;
    const int t = w.t, next_item = w.next_item, nrows = w.nrows, s = w.s, e = w.e, lng = w.lng, ldo = w.ldo;
    float4* __restrict__ obase = w.obase;  // output row 0 of this work unit
    const unsigned longmask = __ballot_sync(FULL, lng != 0);

    // ---- runs of consecutive rows that are computed here (a queued long row ends a run)
    int rb = 0;
    while (rb < nrows) {
      const unsigned lm = longmask >> rb;
      const int re = lm ? rb + (__ffs(lm) - 1) : nrows;
      if (re > rb) {
        const int S = __shfl_sync(FULL, s, rb);
        const int E = __shfl_sync(FULL, e, re - 1);
        int cr = rb;
        float4 acc[VEC][OV];
#pragma unroll
        for (int i = 0; i < VEC; ++i)
#pragma unroll
          for (int h = 0; h < OV; ++h) acc[i][h] = Ops::zero();
        auto store_row = [&]() {  // row cr <- acc, acc <- 0
          float4* __restrict__ Cv = obase + (int64_t)cr * ldo;
#pragma unroll
          for (int i = 0; i < VEC; ++i) {
            if (colok[i]) {
#pragma unroll
              for (int h = 0; h < OV; ++h) {
                float4* o = Cv + OV * (lane + 32 * i) + h;
                float4 out;
                if (t >= 0) out = acc[i][h];
                else if (a.beta != 0.f) out = Ops::axpby(a.alpha, acc[i][h], a.beta, *o);
                else if (a.alpha != 1.f) out = Ops::scale(a.alpha, acc[i][h]);
                else out = acc[i][h];
                __stcs(o, out);
              }
            }
#pragma unroll
            for (int h = 0; h < OV; ++h) acc[i][h] = Ops::zero();
          }
          ++cr;
        };
        // rows that end at non-zero index `pos` with nothing in them: written as zeros (acc is zero here)
        auto skip_empty = [&](int pos) {
          while (cr < re && __shfl_sync(FULL, e, cr) == pos) store_row();
        };
        skip_empty(S);

        // (column, value) chunks of 32: [p0, p0+32) holds the batch being gathered, the next one is the prefetch
        int p0 = S & ~31;
        int c0 = 0, c1 = 0;
        float v0 = 0.f, v1 = 0.f;
        {
          const int i0 = p0 + lane, i1 = i0 + 32;
          if (i0 >= S && i0 < E) { c0 = __ldg(a.colind + i0); v0 = __ldg(a.val + i0); }
          if (i1 < E) { c1 = __ldg(a.colind + i1); v1 = __ldg(a.val + i1); }
        }
        const int k0 = S >> 3, k1 = (E + 7) >> 3;  // batches [k0, k1): batch k = non-zeros [8k, 8k + 8)
        // ---- gather of batch k: 8 B rows, VEC 16-byte vectors per lane each, and the 8 values
        auto gather = [&](int k, Raw (&b)[G][VEC], float (&vv)[G]) {
          const int q0 = k << 3;
          if (q0 - p0 >= 32) {  // the batch starts the next chunk: shift, fetch the one after it
            p0 += 32;
            c0 = c1; v0 = v1;
            const int i1 = p0 + 32 + lane;
            c1 = 0; v1 = 0.f;
            if (i1 < E) { c1 = __ldg(a.colind + i1); v1 = __ldg(a.val + i1); }
          }
          const int l0 = q0 & 31;  // the batch's first lane of the chunk registers
          if (q0 >= S && q0 + G <= E) {
#pragma unroll
            for (int j = 0; j < G; ++j) {
              const unsigned cc = (unsigned)__shfl_sync(FULL, c0, l0 + j);
              vv[j] = __shfl_sync(FULL, v0, l0 + j);
              const Raw* __restrict__ brow = Bv + (uint64_t)cc * ldbv;
#pragma unroll
              for (int i = 0; i < VEC; ++i)
                if (colok[i]) b[j][i] = ldg_raw(brow + 32 * i);
            }
          } else {
            const int jlo = max(S - q0, 0), jhi = min(E - q0, G);
#pragma unroll
            for (int j = 0; j < G; ++j) {
              const unsigned cc = (unsigned)__shfl_sync(FULL, c0, l0 + j);
              vv[j] = __shfl_sync(FULL, v0, l0 + j);
              if (j >= jlo && j < jhi) {
                const Raw* __restrict__ brow = Bv + (uint64_t)cc * ldbv;
#pragma unroll
                for (int i = 0; i < VEC; ++i)
                  if (colok[i]) b[j][i] = ldg_raw(brow + 32 * i);
              }
            }
          }
        };
        // ---- multiply batch k in
        auto consume = [&](int k, const Raw (&b)[G][VEC], const float (&vv)[G]) {
          const int q0 = k << 3;
          const int jlo = max(S - q0, 0), jhi = min(E - q0, G);
          // rows whose last non-zero lies in this batch: bit j = "a row ends after non-zero q0 + j"
          const unsigned epos = (unsigned)(e - 1 - q0);
          const unsigned endmask = __reduce_or_sync(FULL, (lane < nrows && epos < (unsigned)G) ? (1u << epos) : 0u);
          if (endmask == 0u && jlo == 0 && jhi == G) {
            // whole batch inside one row (the common case): 8 x FMAs, no test
#pragma unroll
            for (int j = 0; j < G; ++j)
#pragma unroll
              for (int i = 0; i < VEC; ++i)
                if (colok[i]) BV::fma(acc[i], vv[j], b[j][i]);
          } else {
            // one sweep per row piece [ja, jb): uniform branches around whole non-zeros, one copy of the row store
            int ja = jlo;
            while (ja < jhi) {
              const unsigned em = endmask >> ja;
              const bool ends = em != 0u && (ja + __ffs(em)) <= jhi;
              const int jb = ends ? ja + __ffs(em) : jhi;
#pragma unroll
              for (int j = 0; j < G; ++j) {
                if (j >= ja && j < jb) {
#pragma unroll
                  for (int i = 0; i < VEC; ++i)
                    if (colok[i]) BV::fma(acc[i], vv[j], b[j][i]);
                }
              }
              if (ends) {
                store_row();
                skip_empty(q0 + jb);
              }
              ja = jb;
            }
          }
        };
        // (Gathering batch k + 1 into a second register buffer while batch k is multiplied was measured in r2k: at the
        // 2 CTAs / SM its 126 registers allow it loses to 3 CTAs / SM without it, 5.1 vs 3.95 ms on C5 with bf16 B.)
        for (int k = k0; k < k1; ++k) {
          Raw b[G][VEC];
          float vv[G];
          gather(k, b, vv);
          consume(k, b, vv);
        }
        while (cr < re) store_row();  // trailing rows (only reached when the run holds no non-zero at all)
      }
      rb = re + 1;
    }

    if (t >= 0) {
      float4 sums[VEC][OV];
      finish_queued_segment<VEC, OV>(a, t, lane, kpadv, colok, sums);
    } else {
      item = __shfl_sync(FULL, next_item, 0);
    }
    pending = __shfl_sync(FULL, pending, 0);
  }
}

template <int VEC, int MINB, bool BF16>
int launch_lean(const CsrArgs& a, int nsm, cudaStream_t st) {
  auto kernel = csrmm_lean_kernel<VEC, MINB, BF16>;
  kernel<<<nsm * MINB, 256, 0, st>>>(a);
  SPMMDB_CUDA_OK(cudaGetLastError());
  kernel<<<nsm * MINB, 256, 0, st>>>(a);  // sweep: segments queued late
  SPMMDB_CUDA_OK(cudaGetLastError());
  return SPMMDB_OK;
}

// ---------------------------------------------------------------------------
// TMA-gather kernel: the lean kernel's work distribution and multiply loop, with the B-row gather handed to the
// tensor memory accelerator.  Blackwell's tile::gather4 mode of cp.async.bulk.tensor copies FOUR rows of a 2-D
// tensor, named by four independent row coordinates, into consecutive shared-memory rows with one instruction:
// exactly a CSR gather with the column indices as coordinates.  Per warp a ring of NS stages x 8 rows (two gather4
// per stage, issued by one lane, completing on the stage's mbarrier); all lanes read the rows back as conflict-free
// 16-byte shared loads.  ~3 issue slots per non-zero for the gather instead of the register kernel's LDG + address
// arithmetic, nothing held in registers while in flight (NS * 8 rows per warp x 12 warps = 192 KB per SM), and the
// multiply phase never waits on a long scoreboard.  Same sequential FMA chain per output element (bit-identical,
// tested).  MEASURED (profiles/r2n): C5 spec'd fp32 6.89 ms vs 6.98 for the lean kernel (both DRAM-bound), but C5
// local fp32 5.1 vs 4.15 ms and bf16 B 3.7-4.4 vs 3.1-4.3 ms: the TMA unit feeds an SM no faster than its LDG path
// (the L2 -> SM port is the ceiling either way) and 12 warps multiply slower than 16-24.  Opt-in (CSR_TMA = 1 / 2).
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_gather4(uint32_t dst, const CUtensorMap* map, int col, int r0, int r1, int r2, int r3,
                                            uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
      ::"r"(dst), "l"(map), "r"(col), "r"(r0), "r"(r1), "r"(r2), "r"(r3), "r"(bar) : "memory");
}

template <int VEC, int NS, int WARPS, bool BF16>
__global__ void __launch_bounds__(WARPS * 32, 1) csrmm_tma_kernel(const CsrArgs a_in, const __grid_constant__ CUtensorMap mapB) {
  static_assert(NS == 2 || NS == 4, "the issue pointer must stay within two (column, value) chunks of the consume pointer");
  CsrArgs a = a_in;
  a.epoch = __ldcg(a_in.epoch_dev);  // this call's number lives on the device: a captured CUDA graph replays correctly
  constexpr int G = 8;
  using Ops = VecOps<float4>;
  using BV = BVec<BF16>;
  using Raw = typename BV::Raw;
  constexpr int OV = BV::OV;
  constexpr unsigned FULL = 0xffffffffu;
  extern __shared__ __align__(128) unsigned char tma_smem[];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int nvec = a.K / (4 * OV);                 // 16-byte vectors per row of B
  const uint32_t rowb = (uint32_t)nvec * 16u;      // bytes per gathered row = the tensor map's box
  // per warp: a ring of NS stages x 8 rows, then NS mbarriers (one per stage)
  const uint32_t warp_bytes = (uint32_t)(NS * G) * rowb + 128u;
  unsigned char* my = tma_smem + (size_t)wid * warp_bytes;
  const uint32_t ring_s = (uint32_t)__cvta_generic_to_shared(my);
  const uint32_t bars_s = ring_s + (uint32_t)(NS * G) * rowb;
  const Raw* ring = reinterpret_cast<const Raw*>(my) + lane;
  if (lane < NS) mbar_init(bars_s + 8u * lane, 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncwarp();
  unsigned gk = 0;  // stages this warp has issued and consumed so far: stage s of the ring is in its (gk / NS)-th use
  const int nitems = (a.m + a.rpi - 1) / a.rpi;
  const int kpadv = a.Kpad >> 2;
  bool colok[VEC];
#pragma unroll
  for (int i = 0; i < VEC; ++i) colok[i] = (lane + 32 * i) < nvec;

  int item = 0;
  if (lane == 0) item = atomicAdd(a.counters + 3, 1);
  item = __shfl_sync(FULL, item, 0);
  int pending = a.inline_segs;
  for (;;) {
    // ---- pick work: a queued segment if there is one, else the item already claimed
    WorkUnit w;
    const int what = pick_work(a, lane, item, nitems, kpadv, pending, w);
    if (what == kWorkNone) break;
    if (what == kWorkVoid) continue;
    const int t = w.t, next_item = w.next_item, nrows = w.nrows, s = w.s, e = w.e, lng = w.lng, ldo = w.ldo;
    float4* __restrict__ obase = w.obase;  // output row 0 of this work unit
    const unsigned longmask = __ballot_sync(FULL, lng != 0);

    // ---- runs of consecutive rows that are computed here (a queued long row ends a run)
    int rb = 0;
    while (rb < nrows) {
      const unsigned lm = longmask >> rb;
      const int re = lm ? rb + (__ffs(lm) - 1) : nrows;
      if (re > rb) {
        const int S = __shfl_sync(FULL, s, rb);
        const int E = __shfl_sync(FULL, e, re - 1);
        int cr = rb;
        float4 acc[VEC][OV];
#pragma unroll
        for (int i = 0; i < VEC; ++i)
#pragma unroll
          for (int h = 0; h < OV; ++h) acc[i][h] = Ops::zero();
        auto store_row = [&]() {  // row cr <- acc, acc <- 0
          float4* __restrict__ Cv = obase + (int64_t)cr * ldo;
#pragma unroll
          for (int i = 0; i < VEC; ++i) {
            if (colok[i]) {
#pragma unroll
              for (int h = 0; h < OV; ++h) {
                float4* o = Cv + OV * (lane + 32 * i) + h;
                float4 out;
                if (t >= 0) out = acc[i][h];
                else if (a.beta != 0.f) out = Ops::axpby(a.alpha, acc[i][h], a.beta, *o);
                else if (a.alpha != 1.f) out = Ops::scale(a.alpha, acc[i][h]);
                else out = acc[i][h];
                __stcs(o, out);
              }
            }
#pragma unroll
            for (int h = 0; h < OV; ++h) acc[i][h] = Ops::zero();
          }
          ++cr;
        };
        // rows that end at non-zero index `pos` with nothing in them: written as zeros (acc is zero here)
        auto skip_empty = [&](int pos) {
          while (cr < re && __shfl_sync(FULL, e, cr) == pos) store_row();
        };
        skip_empty(S);

        // (column, value) chunks of 32: [p0, p0+32) holds the batch being multiplied, the gathers run at most NS batches
        // ahead (inside the first two chunks), the third chunk is the prefetch in flight
        int p0 = S & ~31;
        int c0 = 0, c1 = 0, c2 = 0;
        float v0 = 0.f, v1 = 0.f, v2 = 0.f;
        {
          const int i0 = p0 + lane, i1 = i0 + 32, i2 = i0 + 64;
          if (i0 >= S && i0 < E) { c0 = __ldg(a.colind + i0); v0 = __ldg(a.val + i0); }
          if (i1 < E) { c1 = __ldg(a.colind + i1); v1 = __ldg(a.val + i1); }
          if (i2 < E) { c2 = __ldg(a.colind + i2); v2 = __ldg(a.val + i2); }
        }
        const int k0 = S >> 3, k1 = (E + 7) >> 3;  // batches [k0, k1): batch k = non-zeros [8k, 8k + 8)
        // ---- gather of batch k: two tile::gather4 TMA copies (4 rows of B each, row ids straight from the column
        // indices) into stage k % NS of the ring, completing on the stage's mbarrier.  Positions outside the run gather
        // row 0 (never read back).
        auto issue = [&](int k) {
          if (k >= k1) return;
          const int q0 = k << 3;
          const int csel = (q0 - p0) >= 32 ? c1 : c0;
          const int l0 = q0 & 31;
          int r[G];
#pragma unroll
          for (int j = 0; j < G; ++j) {
            const int cc = __shfl_sync(FULL, csel, l0 + j);
            r[j] = (q0 + j >= S && q0 + j < E) ? cc : 0;
          }
          const unsigned slot = (gk + (unsigned)(k - k0)) % NS;
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // earlier reads of the slot vs the copies into it
          if (lane == 0) {
            const uint32_t bar = bars_s + 8u * slot;
            const uint32_t dst = ring_s + (slot * G) * rowb;
            mbar_expect_tx(bar, (uint32_t)G * rowb);
            tma_gather4(dst, &mapB, 0, r[0], r[1], r[2], r[3], bar);
            tma_gather4(dst + 4u * rowb, &mapB, 0, r[4], r[5], r[6], r[7], bar);
          }
        };
        // ---- multiply batch k in
        auto consume = [&](int k) {
          const int q0 = k << 3;
          if (q0 - p0 >= 32) {  // consumed a chunk: shift, fetch the one after the next two
            p0 += 32;
            c0 = c1; v0 = v1; c1 = c2; v1 = v2;
            const int i2 = p0 + 64 + lane;
            c2 = 0; v2 = 0.f;
            if (i2 < E) { c2 = __ldg(a.colind + i2); v2 = __ldg(a.val + i2); }
          }
          const unsigned g = gk + (unsigned)(k - k0);
          const Raw* st = ring + (size_t)((g % NS) * G) * nvec;
          float vv[G];
#pragma unroll
          for (int j = 0; j < G; ++j) vv[j] = __shfl_sync(FULL, v0, (q0 & 31) + j);
          mbar_wait(bars_s + 8u * (g % NS), (g / NS) & 1u);
          const int jlo = max(S - q0, 0), jhi = min(E - q0, G);
          // rows whose last non-zero lies in this batch: bit j = "a row ends after non-zero q0 + j"
          const unsigned epos = (unsigned)(e - 1 - q0);
          const unsigned endmask = __reduce_or_sync(FULL, (lane < nrows && epos < (unsigned)G) ? (1u << epos) : 0u);
          if (endmask == 0u && jlo == 0 && jhi == G) {
            // whole batch inside one row (the common case): 8 x FMAs, no test
#pragma unroll
            for (int j = 0; j < G; ++j)
#pragma unroll
              for (int i = 0; i < VEC; ++i)
                if (colok[i]) BV::fma(acc[i], vv[j], st[j * nvec + 32 * i]);
          } else {
            // one sweep per row piece [ja, jb): uniform branches around whole non-zeros, one copy of the row store
            int ja = jlo;
            while (ja < jhi) {
              const unsigned em = endmask >> ja;
              const bool ends = em != 0u && (ja + __ffs(em)) <= jhi;
              const int jb = ends ? ja + __ffs(em) : jhi;
#pragma unroll
              for (int j = 0; j < G; ++j) {
                if (j >= ja && j < jb) {
#pragma unroll
                  for (int i = 0; i < VEC; ++i)
                    if (colok[i]) BV::fma(acc[i], vv[j], st[j * nvec + 32 * i]);
                }
              }
              if (ends) {
                store_row();
                skip_empty(q0 + jb);
              }
              ja = jb;
            }
          }
        };
        for (int k = 0; k < NS; ++k) issue(k0 + k);
        for (int k = k0; k < k1; ++k) {
          consume(k);
          __syncwarp();  // every lane is done reading the stage before it is refilled
          issue(k + NS);
        }
        gk += (unsigned)(k1 - k0);
        while (cr < re) store_row();  // trailing rows (only reached when the run holds no non-zero at all)
      }
      rb = re + 1;
    }

    if (t >= 0) {
      float4 sums[VEC][OV];
      finish_queued_segment<VEC, OV>(a, t, lane, kpadv, colok, sums);
    } else {
      item = __shfl_sync(FULL, next_item, 0);
    }
    pending = __shfl_sync(FULL, pending, 0);
  }
}



template <int VEC, int NS, int WARPS, bool BF16>
int launch_tma(const CsrArgs& a, int nsm, cudaStream_t st) {
  CUtensorMap map;
  // one box = one row of B (K elements); tile::gather4 fetches four of them per instruction
  if (!make_map_rows(&map, BF16, a.B, (uint64_t)a.K, (uint64_t)a.b_rows, (uint64_t)a.ldb * (BF16 ? 2 : 4))) return SPMMDB_ERR_NOT_SUPPORTED;
  const size_t rowb = (size_t)(a.K / (BF16 ? 8 : 4)) * 16;
  const size_t smem = (size_t)WARPS * ((size_t)NS * 8 * rowb + 128);
  auto kernel = csrmm_tma_kernel<VEC, NS, WARPS, BF16>;
  SPMMDB_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kernel<<<nsm, WARPS * 32, smem, st>>>(a, map);
  SPMMDB_CUDA_OK(cudaGetLastError());
  kernel<<<nsm, WARPS * 32, smem, st>>>(a, map);  // sweep: segments queued late
  SPMMDB_CUDA_OK(cudaGetLastError());
  return SPMMDB_OK;
}

template <typename V, int LPR, int VEC, int U>
int launch_long(const CsrArgs& a, int long_ctas, cudaStream_t st) {
  csrmm_long_kernel<V, LPR, VEC, U><<<long_ctas, 256, 0, st>>>(a);
  SPMMDB_CUDA_OK(cudaGetLastError());
  const int nsm = device_num_sms() > 0 ? device_num_sms() : 148;
  csrmm_reduce_kernel<V><<<nsm * 2, 256, 0, st>>>(a);
  SPMMDB_CUDA_OK(cudaGetLastError());
  return SPMMDB_OK;
}

template <int VEC, int U, int MINB, bool BF16 = false>
int launch_stream(const CsrArgs& a, int nsm, int long_ctas, cudaStream_t st) {
  auto kernel = a.hot != nullptr ? csrmm_stream_kernel<VEC, U, MINB, BF16, true> : csrmm_stream_kernel<VEC, U, MINB, BF16, false>;
  kernel<<<nsm * MINB, 256, 0, st>>>(a);
  SPMMDB_CUDA_OK(cudaGetLastError());
  if (!a.inline_segs) {
    if (BF16) {  // the long-row kernel reads fp32 B only: queued segments are always taken inline with bf16 B
      set_last_error("csrmm bf16-B without inline segments", cudaErrorInvalidValue, __FILE__, __LINE__);
      return SPMMDB_ERR_CUDA;
    }
    return launch_long<float4, 32, VEC, (VEC >= 3 ? 4 : 8)>(a, long_ctas, st);
  }
  kernel<<<nsm * MINB, 256, 0, st>>>(a);  // sweep: segments queued late
  SPMMDB_CUDA_OK(cudaGetLastError());
  return SPMMDB_OK;
}

template <typename V, int LPR, int VEC, int U>
int launch_variant(const CsrArgs& a, int long_ctas, cudaStream_t st) {
  constexpr int GPW = 32 / LPR;
  const int rows_per_cta = 8 * GPW;
  dim3 grid((unsigned)((a.m + rows_per_cta - 1) / rows_per_cta));
  csrmm_rows_kernel<V, LPR, VEC, U><<<grid, 256, 0, st>>>(a);
  SPMMDB_CUDA_OK(cudaGetLastError());
  return launch_long<V, LPR, VEC, U>(a, long_ctas, st);
}

// Second stream + fork / join events of the overlapped small-matrix path (created once per process).
struct Overlap {
  cudaStream_t side = nullptr;
  cudaEvent_t fork = nullptr, join = nullptr;
  bool ok = false;
  Overlap() {
    ok = cudaStreamCreateWithFlags(&side, cudaStreamNonBlocking) == cudaSuccess &&
         cudaEventCreateWithFlags(&fork, cudaEventDisableTiming) == cudaSuccess &&
         cudaEventCreateWithFlags(&join, cudaEventDisableTiming) == cudaSuccess;
  }
};

template <typename V, int LPR, int VEC, int U, int LLPR, int LVEC, int LU>
int launch_pipe(const CsrArgs& a, int long_ctas, cudaStream_t st) {
  const int per_sm = SPMMDB_TUNE("CSR_PIPE_CTAS", 3);
  const int overlap_env = SPMMDB_TUNE("CSR_OVERLAP", 1);
  const int nsm = device_num_sms() > 0 ? device_num_sms() : 148;
  if (overlap_env) {
    // queue first, then the rows kernel on the caller's stream next to the long-row kernels on a side stream:
    // neither fills the GPU alone (both are latency bound at ~30 % occupancy, profiles/r1o)
    // one side stream + event pair per (device, caller stream), created lazily: two caller streams in flight
    // together must not share the fork / join events
    static std::map<std::pair<int, cudaStream_t>, Overlap*> per_stream;
    static std::mutex mu;
    int devid = 0;
    if (cudaGetDevice(&devid) != cudaSuccess) devid = -1;
    Overlap* ovp = nullptr;
    if (devid >= 0) {
      std::lock_guard<std::mutex> lk(mu);
      Overlap*& slot = per_stream[std::make_pair(devid, st)];
      if (slot == nullptr && per_stream.size() <= 64) slot = new Overlap();
      ovp = slot;
    }
    if (ovp != nullptr && ovp->ok) {
      Overlap& ov = *ovp;
      CsrArgs b = a;
      b.prequeued = 1;
      csrmm_scan_long_kernel<V><<<nsm * 2, 256, 0, st>>>(b);
      SPMMDB_CUDA_OK(cudaGetLastError());
      SPMMDB_CUDA_OK(cudaEventRecord(ov.fork, st));
      SPMMDB_CUDA_OK(cudaStreamWaitEvent(ov.side, ov.fork, 0));
      const int rc = launch_long<V, LLPR, LVEC, LU>(b, long_ctas, ov.side);
      if (rc != SPMMDB_OK) return rc;
      csrmm_rows_pipe_kernel<V, LPR, VEC, U><<<nsm * per_sm, 256, 0, st>>>(b);
      SPMMDB_CUDA_OK(cudaGetLastError());
      SPMMDB_CUDA_OK(cudaEventRecord(ov.join, ov.side));
      SPMMDB_CUDA_OK(cudaStreamWaitEvent(st, ov.join, 0));
      return SPMMDB_OK;
    }
  }
  csrmm_rows_pipe_kernel<V, LPR, VEC, U><<<nsm * per_sm, 256, 0, st>>>(a);
  SPMMDB_CUDA_OK(cudaGetLastError());
  return launch_long<V, LLPR, LVEC, LU>(a, long_ctas, st);
}

template <typename V>
int dispatch(const CsrArgs& a, int long_ctas, bool two_rows, cudaStream_t st) {
  const int nvec = (a.K + VecOps<V>::W - 1) / VecOps<V>::W;
  if (nvec <= 4) return launch_variant<V, 4, 1, 4>(a, long_ctas, st);
  if (nvec <= 8) return launch_variant<V, 8, 1, 8>(a, long_ctas, st);
  // Small short-row matrices (`two_rows`: under 4 M non-zeros, 32 < K <= 128) are latency / issue bound, not
  // bandwidth bound (profiles/r1o): two rows per warp (16 lanes each), the pipelined rows kernel, and the
  // long-row kernels on a side stream.  C2: 0.154 (stream kernel) -> 0.141 (two rows per warp) -> 0.108 ms;
  // one row per warp 0.190, four rows per warp lose to imbalance inside the warp.
  const int pipe = SPMMDB_TUNE("CSR_PIPE", 1);
  if (nvec <= 16 && two_rows && pipe) return launch_pipe<V, 16, 1, 8, 16, 1, 8>(a, long_ctas, st);
  if (nvec <= 16) return launch_variant<V, 16, 1, 8>(a, long_ctas, st);
  if (nvec <= 32 && two_rows) {
    if (pipe == 1) return launch_pipe<V, 16, 2, 4, 16, 2, 4>(a, long_ctas, st);
    if (pipe == 2) return launch_pipe<V, 16, 2, 4, 32, 1, 8>(a, long_ctas, st);
    return launch_variant<V, 16, 2, 4>(a, long_ctas, st);
  }
  if (nvec <= 32) return launch_variant<V, 32, 1, 8>(a, long_ctas, st);
  if (nvec <= 64 && two_rows && pipe) return launch_pipe<V, 32, 2, 4, 32, 2, 4>(a, long_ctas, st);
  if (nvec <= 64) return launch_variant<V, 32, 2, 8>(a, long_ctas, st);
  if (nvec <= 96) return launch_variant<V, 32, 3, 4>(a, long_ctas, st);
  return launch_variant<V, 32, 4, 4>(a, long_ctas, st);
}

}  // namespace

// Core entry: row-major B and C.
int csrmm_rowmajor(int m, int K, int64_t nnz, const int* rowptr, const int* colind, const float* val,
                   const float* B, int64_t ldb, float* C, int64_t ldc, float alpha, float beta,
                   cudaStream_t st, int b_bf16, const unsigned* hot) {
  CsrArgs a;
  a.m = m; a.K = K; a.rowptr = rowptr; a.colind = colind; a.val = val;
  a.B = B; a.ldb = ldb; a.C = C; a.ldc = ldc; a.alpha = alpha; a.beta = beta;
  a.b_bf16 = b_bf16; a.hot = hot;
  a.b_rows = 0x7fffffff;  // the callee never learns k from gespmm_csrmm's signature; the map only bounds, never reads
  a.cold_policy = SPMMDB_TUNE("CSR_COLD_POLICY", 1);
  a.hot_policy = SPMMDB_TUNE("CSR_HOT_POLICY", 0);
  const int seg_override = current_segment_override();
  a.seg = seg_override ? seg_override : spmmdb_csrmm_segment_for(m, nnz);
  // Small matrices (C2: 1.1 M nnz, 64-entry segments) are latency-bound: a queued segment as a work unit costs
  // ~10 dependent L2 round trips, the grid-stride long-row kernel half of that (0.155 vs 0.194 ms on C2).
  const int inline_force = SPMMDB_TUNE("CSR_INLINE", -1);  // -1: automatic
  a.inline_segs = inline_force >= 0 ? (inline_force != 0) : (a.seg >= 256);
  if (b_bf16) a.inline_segs = 1;  // the separate long-row kernel reads fp32 B only
  a.prequeued = 0;
  a.Kpad = (K + 3) & ~3;
  a.epoch = 0;  // set on the device
  // worst case: every segment queued.  segments <= nnz/seg + long rows, long rows <= nnz/seg
  const int64_t max_long = nnz / a.seg + 1;
  const int64_t max_segs = 2 * max_long;
  a.cap_rows = (int)max_long;
  a.cap_segs = (int)max_segs;
  const size_t hdr = 64;
  const size_t bytes_q = hdr + sizeof(LongRow) * max_long + sizeof(int) * max_segs;
  char* q = (char*)scratch(kScratchCsrLong, bytes_q, st);
  float* part = (float*)scratch(kScratchCsrPartial, sizeof(float) * (size_t)max_segs * a.Kpad, st);
  // `done` and `seg_entry` live in buffers of their own: the layout of the queue buffer moves with nnz, and
  // both are read before this call writes them -- `done` must be all zero (the last finisher of a row restores
  // it), a `seg_entry` must never carry this call's epoch before its producer published it.  In a shared buffer
  // a later call with another nnz found an earlier call's row records under them.
  int* done = (int*)scratch(kScratchCsrDone, sizeof(int) * (size_t)max_long, st);
  unsigned long long* entry =
      (unsigned long long*)scratch(kScratchCsrEntry, sizeof(unsigned long long) * (size_t)max_segs, st);
  a.epoch_dev = (int*)scratch(kScratchCsrEpoch, 64, st);  // fixed size: allocated once per (device, stream)
  if (q == nullptr || part == nullptr || done == nullptr || entry == nullptr || a.epoch_dev == nullptr)
    return SPMMDB_ERR_ALLOC;
  a.counters = (int*)q;
  a.long_rows = (LongRow*)(q + hdr);
  a.seg_owner = (int*)(q + hdr + sizeof(LongRow) * max_long);
  a.done = done;
  a.seg_entry = entry;
  a.partial = part;
  csrmm_begin_kernel<<<1, 32, 0, st>>>(a.counters, a.epoch_dev);  // zero the queue counters, next call number
  SPMMDB_CUDA_OK(cudaGetLastError());
  const int nsm = device_num_sms() > 0 ? device_num_sms() : 148;
  const int lc_env = SPMMDB_TUNE("CSR_LONG_CTAS", 0);
  const int long_ctas = nsm * (lc_env > 0 ? lc_env : 4);  // C2 sweep: 2 / 4 / 8 -> 0.166 / 0.160 / 0.156 ms
  const bool vec4 = (K % 4 == 0) && (ldb % 4 == 0) && (ldc % 4 == 0) &&
                    (((uintptr_t)B | (uintptr_t)C) % 16 == 0);
  const int stream_min_k = SPMMDB_TUNE("CSR_STREAM_MIN_K", 32);
  // Long average rows (C3: 271 nnz/row) keep one warp per row: the rows kernel reaches the L2 gather
  // ceiling there (0.94 ms vs 1.67 ms for the stream kernel, same profile file).
  const bool dense_rows = m > 0 && nnz / m >= 64;
  // small matrices with a 128-wide dense side: the two-rows-per-warp rows kernel (see dispatch)
  const int small_env = SPMMDB_TUNE("CSR_SMALL", 1);
  const bool small_k128 = small_env && vec4 && nnz < ((int64_t)4 << 20) && K > 32 && K <= 256 && !dense_rows;
  if (b_bf16) {
    // bf16-stored B (fp32 accumulate): the stream kernel with 8 columns per 16-byte vector, whatever the shape
    const bool ok = (K % 8 == 0) && K <= 1024 && (ldb % 8 == 0) && (ldc % 4 == 0) && (((uintptr_t)B | (uintptr_t)C) % 16 == 0);
    if (!ok) return SPMMDB_ERR_NOT_SUPPORTED;
    const int64_t avg = m > 0 ? (nnz + m - 1) / m : 1;
    int rpi = 8;
    while (rpi > 1 && rpi * avg > 512) rpi >>= 1;
    const int rpi_env = SPMMDB_TUNE("CSR_RPI", 0);
    if (rpi_env > 0) rpi = rpi_env < 32 ? rpi_env : 32;
    a.rpi = rpi;
    const int vec = (K / 8 + 31) / 32;
    const int variant = SPMMDB_TUNE("CSR_VARIANT", 0);
    const int bulk = SPMMDB_TUNE("CSR_LEAN", 1);  // 0: csrmm_stream_kernel, 1 / 2: csrmm_lean_kernel (two occupancies)
    const int tma = SPMMDB_TUNE("CSR_TMA", 0);  // tile::gather4 ring kernel (rows of at most 256 elements)
    if (tma && K <= 256) {
      if (tma == 2) return launch_tma<1, 2, 16, true>(a, nsm, st);
      return launch_tma<1, 4, 12, true>(a, nsm, st);
    }
    if (bulk) {
      if (vec == 1) return bulk == 2 ? launch_lean<1, 4, true>(a, nsm, st) : launch_lean<1, 3, true>(a, nsm, st);
      if (vec == 2) return launch_lean<2, 2, true>(a, nsm, st);
    }  // wider rows: the stream kernel below (8 x 3 or 4 vectors per lane do not fit the register file)
    if (vec == 1) {
      if (variant == 1) return launch_stream<1, 8, 3, true>(a, nsm, long_ctas, st);
      if (variant == 2) return launch_stream<1, 16, 3, true>(a, nsm, long_ctas, st);
      if (variant == 3) return launch_stream<1, 8, 4, true>(a, nsm, long_ctas, st);
      return launch_stream<1, 16, 2, true>(a, nsm, long_ctas, st);
    }
    if (vec == 2) return launch_stream<2, 8, 2, true>(a, nsm, long_ctas, st);
    if (vec == 3) return launch_stream<3, 3, 3, true>(a, nsm, long_ctas, st);
    return launch_stream<4, 2, 3, true>(a, nsm, long_ctas, st);
  }
  if (hot != nullptr && !(vec4 && K >= 32 && K <= 512)) a.hot = nullptr;  // the hint is a stream-kernel feature
  const bool force_stream = a.hot != nullptr;  // a caller that supplies a plan wants the kernel that honours it
  if (vec4 && K >= stream_min_k && K <= 512 && ((!dense_rows && !small_k128) || force_stream)) {
    // rows per item: measured on B200 (profiles/r1b_csr_rpi_sweep.txt): 8 rows is the best or within 3% of it
    // for avg degree 2 (C4), 6.5 (C2) and 24 (C5); fewer rows per item for denser rows.
    const int64_t avg = m > 0 ? (nnz + m - 1) / m : 1;
    int rpi = 8;
    while (rpi > 1 && rpi * avg > 512) rpi >>= 1;
    const int rpi_env = SPMMDB_TUNE("CSR_RPI", 0);
    if (rpi_env > 0) rpi = rpi_env < 32 ? rpi_env : 32;
    a.rpi = rpi;
    const int vec = (K / 4 + 31) / 32;
    const int bulk = SPMMDB_TUNE("CSR_LEAN", 1);  // 0: csrmm_stream_kernel, 1 / 2: csrmm_lean_kernel (two occupancies)
    const int tma = SPMMDB_TUNE("CSR_TMA", 0);  // tile::gather4 ring kernel (rows of at most 256 elements)
    if (tma && a.inline_segs && a.hot == nullptr && K <= 256) {
      if (vec == 1) return tma == 2 ? launch_tma<1, 2, 16, false>(a, nsm, st) : launch_tma<1, 4, 12, false>(a, nsm, st);
      return tma == 2 ? launch_tma<2, 4, 6, false>(a, nsm, st) : launch_tma<2, 2, 12, false>(a, nsm, st);
    }
    if (bulk && a.inline_segs && a.hot == nullptr && vec <= 2) {
      if (vec == 1) return bulk == 2 ? launch_lean<1, 4, false>(a, nsm, st) : launch_lean<1, 3, false>(a, nsm, st);
      if (vec == 2) return launch_lean<2, 2, false>(a, nsm, st);
    }
    const int variant = SPMMDB_TUNE("CSR_VARIANT", 0);
    if (vec == 1) {
      if (variant == 1) return launch_stream<1, 16, 2>(a, nsm, long_ctas, st);
      if (variant == 3) return launch_stream<1, 8, 4>(a, nsm, long_ctas, st);
      return launch_stream<1, 8, 3>(a, nsm, long_ctas, st);
    }
    if (vec == 2) {
      if (variant == 2) return launch_stream<2, 4, 3>(a, nsm, long_ctas, st);
      if (variant == 3) return launch_stream<2, 4, 4>(a, nsm, long_ctas, st);
      return launch_stream<2, 8, 2>(a, nsm, long_ctas, st);
    }
    if (vec == 3) return launch_stream<3, 3, 3>(a, nsm, long_ctas, st);
    return launch_stream<4, 2, 3>(a, nsm, long_ctas, st);
  }
  a.rpi = 0;
  return vec4 ? dispatch<float4>(a, long_ctas, small_k128, st) : dispatch<float>(a, long_ctas, false, st);
}

}  // namespace spmmdb

extern "C" int spmmdb_csrmm_segment(int64_t nnz) {
  const int seg_env = SPMMDB_TUNE("CSR_SEG", 0);
  if (seg_env >= 32 && (seg_env & (seg_env - 1)) == 0) return seg_env;  // tuning knob (power of two)
  if (nnz < (int64_t)4 << 20) return 64;
  if (nnz < (int64_t)32 << 20) return 256;
  return 512;  // B200 sweep on the products-shaped graph: 256 / 512 / 1024 / 2048 -> 4.75 / 4.74 / 4.97 / 5.21 ms
}

// With long average rows one warp per row is the fast path and a segment shorter than a few average rows
// would push most of the matrix through the partial-sum path (C3 remainder of a hybrid split: 207 nnz per
// row, 256-entry segments, 2.4 ms instead of 0.9): at least the power of two above 4 average rows.
extern "C" int spmmdb_csrmm_segment_for(int m, int64_t nnz) {
  int seg = spmmdb_csrmm_segment(nnz);
  if (m > 0 && nnz > 0) {
    const int64_t want = 4 * ((nnz + m - 1) / m);
    while (seg < want && seg < (1 << 20)) seg <<= 1;
  }
  return seg;
}

namespace spmmdb {
namespace {
std::atomic<int> g_segment_override{0};
thread_local int tl_segment_override = 0;  // set by a library entry point around its own inner calls
}  // namespace
int current_segment_override() { return tl_segment_override ? tl_segment_override : g_segment_override.load(); }
ScopedSegment::ScopedSegment(int seg) : before_(tl_segment_override) { tl_segment_override = seg; }
ScopedSegment::~ScopedSegment() { tl_segment_override = before_; }
}  // namespace spmmdb
extern "C" int spmmdb_set_csrmm_segment(int seg) {
  if (seg != 0 && (seg < 32 || (seg & (seg - 1)) != 0)) return SPMMDB_ERR_INVALID_VALUE;
  spmmdb::g_segment_override.store(seg);
  return SPMMDB_OK;
}
extern "C" int spmmdb_get_csrmm_segment(void) { return spmmdb::g_segment_override.load(); }
