// convert_br.cu -- csr2bsr and the hybrid split with ONE CTA PER BLOCK ROW (the fast path of
// convert.cu's entry points; same outputs bit for bit).  Replaces the same reference calls:
// cusparseXcsr2bsrNnz + cusparseScsr2bsr (run_bsrmm.cu:121-142) and divide_matrix
// (divide.cu:52-127).
//
// The bs rows of a block row are one contiguous piece of the CSR arrays, so the CTA reads them
// coalesced instead of one lane walking one row.  The set of column blocks the block row touches
// is a bitmap in shared memory (atomicOr per entry whose bit is not set yet); a prefix sum of the per-word popcounts turns
// the bitmap into ranks: the t-th set bit is the t-th non-zero block, ascending, with no sort and
// no merge.  Values are scattered into the block row's piece of bsr_val right after the same CTA
// zero-filled it (the lines are still in L2, DRAM sees each line once).  The split additionally
// counts the entries per touched block in shared-memory counters and keeps a second bitmap of
// the blocks that reach the threshold.
#include <limits.h>
#include <stdlib.h>

#include "common.cuh"

namespace spmmdb {
namespace {

constexpr int kThreads = 256;
constexpr int kCntCap = 4096;  // block counters per pass (split only)

// column block of a column; clamped to the last block so that an out-of-range column cannot leave the bitmap
__device__ __forceinline__ int blk_of(int col, int bs, int shift, int last) {
  return min(shift >= 0 ? col >> shift : col / bs, last);
}

// exclusive prefix sum of v over the CTA (256 threads); *total = CTA sum
__device__ __forceinline__ int cta_exclusive_scan(int v, int* total, int* warp_sums) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) warp_sums[wid] = inc;
  __syncthreads();
  if (wid == 0) {
    int w = (lane < 8) ? warp_sums[lane] : 0;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += t;
    }
    if (lane < 8) warp_sums[lane] = w;
  }
  __syncthreads();
  const int woff = (wid == 0) ? 0 : warp_sums[wid - 1];
  *total = warp_sums[7];
  __syncthreads();
  return woff + inc - v;
}

// wpre[w] = number of set bits of bits[wlo..w); returns the number of set bits in [wlo, whi]
__device__ __forceinline__ int rank_words(const unsigned* bits, int* wpre, int wlo, int whi, int* warp_sums) {
  int carry = 0;
  for (int base = wlo; base <= whi; base += kThreads) {
    const int w = base + threadIdx.x;
    const int v = (w <= whi) ? __popc(bits[w]) : 0;
    int total;
    const int ex = cta_exclusive_scan(v, &total, warp_sums);
    if (w <= whi) wpre[w] = carry + ex;
    carry += total;
  }
  return carry;
}

__device__ __forceinline__ int rank_of(const unsigned* bits, const int* wpre, int blk) {
  return wpre[blk >> 5] + __popc(bits[blk >> 5] & ((1u << (blk & 31)) - 1u));
}

// zero floats [lo, hi) of p with the whole CTA
__device__ __forceinline__ void cta_zero(float* p, int64_t lo, int64_t hi) {
  if (((((uintptr_t)p) | (uintptr_t)(lo * 4) | (uintptr_t)(hi * 4)) & 15) == 0) {
    float4* p4 = reinterpret_cast<float4*>(p + lo);
    const int64_t n4 = (hi - lo) >> 2;
    for (int64_t i = threadIdx.x; i < n4; i += kThreads) p4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  } else {
    for (int64_t i = lo + threadIdx.x; i < hi; i += kThreads) p[i] = 0.f;
  }
}

struct BrArgs {
  int dir, m, mb, bs, shift, nw, thr, last;
  const int* rowptr; const int* colind; const float* val;
  int* bsr_counts; int* csr_counts;                 // count passes
  const int* bsr_rowptr; int* bsr_colind; float* bsr_val;
  const int* csr_rowptr; int* csr_colind; float* csr_val;
};

// FILL = false: counts[ib] = distinct column blocks.  FILL = true: bsr_colind + bsr_val of the block row.
template <bool FILL>
__global__ void __launch_bounds__(kThreads) csr2bsr_br_kernel(const BrArgs a) {
  extern __shared__ unsigned br_smem[];
  unsigned* bm = br_smem;
  int* wpre = reinterpret_cast<int*>(br_smem + a.nw);
  __shared__ int warp_sums[8];
  __shared__ int s_lo, s_hi;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int bs = a.bs, bs2 = bs * bs;
  for (int w = threadIdx.x; w < a.nw; w += kThreads) bm[w] = 0u;
  if (threadIdx.x == 0) { s_lo = INT_MAX; s_hi = -1; }
  __syncthreads();
  for (int ib = blockIdx.x; ib < a.mb; ib += gridDim.x) {
    const int row0 = ib * bs, row1 = min(row0 + bs, a.m);
    const int s0 = __ldg(a.rowptr + row0), s1 = __ldg(a.rowptr + row1);
    int lo = INT_MAX, hi = -1;
    for (int idx = s0 + threadIdx.x; idx < s1; idx += kThreads) {
      const int blk = blk_of(__ldg(a.colind + idx), bs, a.shift, a.last);
      // most entries of a dense block row find their block's bit already set: look before the atomic, or every
      // thread of the CTA serialises on the same one or two bitmap words
      const unsigned bit = 1u << (blk & 31);
      if (!(*(volatile unsigned*)(bm + (blk >> 5)) & bit)) atomicOr(bm + (blk >> 5), bit);
      lo = min(lo, blk >> 5); hi = max(hi, blk >> 5);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
      hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if (lane == 0 && hi >= 0) { atomicMin(&s_lo, lo); atomicMax(&s_hi, hi); }
    __syncthreads();
    const int wlo = s_lo, whi = s_hi;  // whi < 0: empty block row
    int total = 0;
    if (whi >= 0) total = rank_words(bm, wpre, wlo, whi, warp_sums);
    if (!FILL) {
      if (threadIdx.x == 0) a.bsr_counts[ib] = total;
    } else if (whi >= 0) {
      const int64_t base = __ldg(a.bsr_rowptr + ib);
      for (int w = wlo + threadIdx.x; w <= whi; w += kThreads) {
        unsigned word = bm[w];
        int64_t o = base + wpre[w];
        while (word) {
          a.bsr_colind[o++] = w * 32 + (__ffs(word) - 1);
          word &= word - 1u;
        }
      }
      cta_zero(a.bsr_val, base * bs2, (base + total) * bs2);
      __syncthreads();
      for (int r = wid; row0 + r < row1; r += kThreads / 32) {
        const int p0 = __ldg(a.rowptr + row0 + r), p1 = __ldg(a.rowptr + row0 + r + 1);
        for (int p = p0 + lane; p < p1; p += 32) {
          const int col = __ldg(a.colind + p);
          const int blk = blk_of(col, bs, a.shift, a.last);
          const int c = col - blk * bs;
          const int64_t slot = base + rank_of(bm, wpre, blk);
          a.bsr_val[slot * bs2 + (a.dir == SPMMDB_DIR_ROW ? r * bs + c : c * bs + r)] = __ldg(a.val + p);
        }
      }
    }
    __syncthreads();
    if (whi >= 0)
      for (int w = wlo + threadIdx.x; w <= whi; w += kThreads) bm[w] = 0u;
    if (threadIdx.x == 0) { s_lo = INT_MAX; s_hi = -1; }
    __syncthreads();
  }
}

// Hybrid split (divide.cu:52-127).  FILL = false: bsr_counts[ib] = blocks with >= thr entries,
// csr_counts[row] = entries of the row outside those blocks.  FILL = true: both parts.
template <bool FILL>
__global__ void __launch_bounds__(kThreads) divide_br_kernel(const BrArgs a) {
  extern __shared__ unsigned br_smem[];
  unsigned* bm = br_smem;                                   // blocks the block row touches
  unsigned* dbm = br_smem + a.nw;                           // blocks that reach the threshold
  int* wpre = reinterpret_cast<int*>(br_smem + 2 * a.nw);   // ranks (of bm, later of dbm)
  int* cnt = wpre + a.nw;                                   // kCntCap entry counters
  __shared__ int warp_sums[8];
  __shared__ int s_lo, s_hi;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int bs = a.bs, bs2 = bs * bs;
  for (int w = threadIdx.x; w < a.nw; w += kThreads) { bm[w] = 0u; dbm[w] = 0u; }
  if (threadIdx.x == 0) { s_lo = INT_MAX; s_hi = -1; }
  __syncthreads();
  for (int ib = blockIdx.x; ib < a.mb; ib += gridDim.x) {
    const int row0 = ib * bs, row1 = min(row0 + bs, a.m);
    const int s0 = __ldg(a.rowptr + row0), s1 = __ldg(a.rowptr + row1);
    int lo = INT_MAX, hi = -1;
    for (int idx = s0 + threadIdx.x; idx < s1; idx += kThreads) {
      const int blk = blk_of(__ldg(a.colind + idx), bs, a.shift, a.last);
      // most entries of a dense block row find their block's bit already set: look before the atomic, or every
      // thread of the CTA serialises on the same one or two bitmap words
      const unsigned bit = 1u << (blk & 31);
      if (!(*(volatile unsigned*)(bm + (blk >> 5)) & bit)) atomicOr(bm + (blk >> 5), bit);
      lo = min(lo, blk >> 5); hi = max(hi, blk >> 5);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
      hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if (lane == 0 && hi >= 0) { atomicMin(&s_lo, lo); atomicMax(&s_hi, hi); }
    __syncthreads();
    const int wlo = s_lo, whi = s_hi;
    int ndense = 0;
    if (whi >= 0) {
      const int touched = rank_words(bm, wpre, wlo, whi, warp_sums);
      // entries per touched block, kCntCap blocks per pass; blocks that reach thr go into dbm
      for (int c0 = 0; c0 < touched; c0 += kCntCap) {
        for (int i = threadIdx.x; i < kCntCap; i += kThreads) cnt[i] = 0;
        __syncthreads();
        for (int idx = s0 + threadIdx.x; idx < s1; idx += kThreads) {
          const int slot = rank_of(bm, wpre, blk_of(__ldg(a.colind + idx), bs, a.shift, a.last)) - c0;
          if (slot >= 0 && slot < kCntCap) atomicAdd(cnt + slot, 1);
        }
        __syncthreads();
        for (int w = wlo + threadIdx.x; w <= whi; w += kThreads) {  // word w always belongs to the same thread
          unsigned word = bm[w], dword = 0u;
          int slot = wpre[w] - c0;
          while (word) {
            const int b = __ffs(word) - 1;
            if (slot >= 0 && slot < kCntCap && cnt[slot] >= a.thr) dword |= 1u << b;
            ++slot;
            word &= word - 1u;
          }
          if (dword) dbm[w] |= dword;
        }
        __syncthreads();
      }
      ndense = rank_words(dbm, wpre, wlo, whi, warp_sums);  // wpre now ranks the dense blocks
    }
    if (!FILL) {
      if (threadIdx.x == 0) a.bsr_counts[ib] = ndense;
      for (int r = wid; row0 + r < row1; r += kThreads / 32) {
        const int p0 = __ldg(a.rowptr + row0 + r), p1 = __ldg(a.rowptr + row0 + r + 1);
        int rem = 0;
        for (int p = p0 + lane; p < p1; p += 32) {
          const int blk = blk_of(__ldg(a.colind + p), bs, a.shift, a.last);
          rem += ((dbm[blk >> 5] >> (blk & 31)) & 1u) ? 0 : 1;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) rem += __shfl_xor_sync(0xffffffffu, rem, o);
        if (lane == 0) a.csr_counts[row0 + r] = rem;
      }
    } else {
      const int64_t base = __ldg(a.bsr_rowptr + ib);
      if (whi >= 0) {
        for (int w = wlo + threadIdx.x; w <= whi; w += kThreads) {
          unsigned word = dbm[w];
          int64_t o = base + wpre[w];
          while (word) {
            a.bsr_colind[o++] = w * 32 + (__ffs(word) - 1);
            word &= word - 1u;
          }
        }
        cta_zero(a.bsr_val, base * bs2, (base + ndense) * bs2);
      }
      __syncthreads();
      for (int r = wid; row0 + r < row1; r += kThreads / 32) {
        const int p0 = __ldg(a.rowptr + row0 + r), p1 = __ldg(a.rowptr + row0 + r + 1);
        int64_t out = __ldg(a.csr_rowptr + row0 + r);  // next remainder slot of this row
        for (int pb = p0; pb < p1; pb += 32) {
          const int p = pb + lane;
          const bool live = p < p1;
          int col = 0, blk = 0;
          float v = 1.0f;
          bool dense = false;
          if (live) {
            col = __ldg(a.colind + p);
            if (a.val != nullptr) v = __ldg(a.val + p);
            blk = blk_of(col, bs, a.shift, a.last);
            dense = (dbm[blk >> 5] >> (blk & 31)) & 1u;
          }
          if (live && dense) {
            const int c = col - blk * bs;
            const int64_t slot = base + rank_of(dbm, wpre, blk);
            a.bsr_val[slot * bs2 + (a.dir == SPMMDB_DIR_ROW ? r * bs + c : c * bs + r)] = v;
          }
          const unsigned keep = __ballot_sync(0xffffffffu, live && !dense);  // remainder keeps the row's order
          if (live && !dense) {
            const int64_t o = out + __popc(keep & ((1u << lane) - 1u));
            a.csr_colind[o] = col;
            a.csr_val[o] = v;
          }
          out += __popc(keep);
        }
      }
    }
    __syncthreads();
    if (whi >= 0)
      for (int w = wlo + threadIdx.x; w <= whi; w += kThreads) { bm[w] = 0u; dbm[w] = 0u; }
    if (threadIdx.x == 0) { s_lo = INT_MAX; s_hi = -1; }
    __syncthreads();
  }
}

constexpr size_t kBrSmemMax = 160 * 1024;

template <typename K>
int launch_br(K kernel, const BrArgs& a, size_t smem, cudaStream_t st) {
  if (smem > 48 * 1024) SPMMDB_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int nsm = device_num_sms() > 0 ? device_num_sms() : 148;
  int per_sm = (int)((200 * 1024) / (smem + 1024));
  per_sm = per_sm < 1 ? 1 : (per_sm > 8 ? 8 : per_sm);
  const int cap_env = SPMMDB_TUNE("CONVERT_BR_CTAS", 0);
  if (cap_env > 0 && per_sm > cap_env) per_sm = cap_env;
  const int grid = a.mb < nsm * per_sm ? a.mb : nsm * per_sm;
  kernel<<<grid, kThreads, smem, st>>>(a);
  SPMMDB_CUDA_OK(cudaGetLastError());
  return SPMMDB_OK;
}

BrArgs br_args(int dir, int m, int n, int bs, const int* rowptr, const int* colind, const float* val) {
  BrArgs a = {};
  a.dir = dir; a.m = m; a.bs = bs; a.mb = (m + bs - 1) / bs;
  a.shift = (bs & (bs - 1)) == 0 ? __builtin_ctz((unsigned)bs) : -1;
  const int64_t nb = ((int64_t)n + bs - 1) / bs;
  a.nw = (int)((nb + 31) / 32);
  a.last = (int)(nb > 0 ? nb - 1 : 0);
  a.rowptr = rowptr; a.colind = colind; a.val = val;
  return a;
}

}  // namespace

// Whether the block-row kernels take this shape (else convert.cu's warp-merge kernels do): the bitmaps and
// ranks of a block row must fit shared memory.  n = number of columns of the matrix.
bool convert_br_ok(int n, int bs, bool split) {
  if (bs <= 0 || bs > 128) return false;
  const int64_t nw = (((int64_t)n + bs - 1) / bs + 31) / 32;
  const size_t smem = split ? (size_t)nw * 12 + sizeof(int) * kCntCap : (size_t)nw * 8;
  return smem <= kBrSmemMax;
}

int csr2bsr_br(bool fill, int dir, int m, int n, const int* rowptr, const int* colind, const float* val, int bs,
               int* counts, const int* bsr_rowptr, int* bsr_colind, float* bsr_val, cudaStream_t st) {
  BrArgs a = br_args(dir, m, n, bs, rowptr, colind, val);
  a.bsr_counts = counts; a.bsr_rowptr = bsr_rowptr; a.bsr_colind = bsr_colind; a.bsr_val = bsr_val;
  const size_t smem = (size_t)a.nw * 8;
  return fill ? launch_br(csr2bsr_br_kernel<true>, a, smem, st) : launch_br(csr2bsr_br_kernel<false>, a, smem, st);
}

int divide_br(bool fill, int dir, int m, int n, const int* rowptr, const int* colind, const float* val, int bs, int thr,
              int* bsr_counts, int* csr_counts, const int* bsr_rowptr, int* bsr_colind, float* bsr_val,
              const int* csr_rowptr, int* csr_colind, float* csr_val, cudaStream_t st) {
  BrArgs a = br_args(dir, m, n, bs, rowptr, colind, val);
  a.thr = thr;
  a.bsr_counts = bsr_counts; a.csr_counts = csr_counts;
  a.bsr_rowptr = bsr_rowptr; a.bsr_colind = bsr_colind; a.bsr_val = bsr_val;
  a.csr_rowptr = csr_rowptr; a.csr_colind = csr_colind; a.csr_val = csr_val;
  const size_t smem = (size_t)a.nw * 12 + sizeof(int) * kCntCap;
  return fill ? launch_br(divide_br_kernel<true>, a, smem, st) : launch_br(divide_br_kernel<false>, a, smem, st);
}

}  // namespace spmmdb
