// convert.cu -- csr2bsr (count + fill) and bsr2csr for sm_100a.  Integer /
// byte-moving work, HBM-bound, bit-exact.  Replaces the cuSPARSE calls at
// run_bsrmm.cu:121-142 (cusparseXcsr2bsrNnz + cusparseScsr2bsr) and
// bsr2csr.cu:186-188 (cusparseSbsr2csr); CPU twins in the reference:
// utility.cc:47-69 (calculateNnzb), divide.cu:52-127, my_run_bsr.cu:62-115.
//
// csr2bsr: one warp per block row runs a bs-way merge over the (ascending)
// rows of the block row.  Lane l walks row l (and l+32, ... when bs > 32); the
// warp minimum of the lanes' current column-block ids is the next non-zero
// block, emitted in ascending order with no sort, no hash and no bound on the
// number of column blocks.  The count pass and the fill pass are the same
// walk; the fill pass assembles each bs x bs block in shared memory and writes
// it out once, fully coalesced (zeros included), so bsr_val is written exactly
// once and never pre-cleared.
#include <limits.h>
#include <stdlib.h>

#include <mutex>

#include "common.cuh"

namespace spmmdb {
namespace {

constexpr int kWarpsPerCta = 8;

__device__ __forceinline__ int warp_min(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// MODE 0: count blocks per block row -> counts[ib]
// MODE 1: fill through a shared-memory tile (bs*bs floats per warp)
// MODE 2: fill by direct scatter into pre-zeroed bsr_val (very large blocks)
template <int R, int MODE>
__global__ void __launch_bounds__(kWarpsPerCta * 32) csr2bsr_kernel(
    int dir, int m, int mb, int bs, const int* __restrict__ rowptr, const int* __restrict__ colind,
    const float* __restrict__ val, int* __restrict__ counts, const int* __restrict__ bsr_rowptr,
    int* __restrict__ bsr_colind, float* __restrict__ bsr_val, int warps_per_cta) {
  extern __shared__ float smem[];
  const int lane = threadIdx.x & 31;
  const int wid = threadIdx.x >> 5;
  if (wid >= warps_per_cta) return;
  const int bs2 = bs * bs;
  const int shift = (bs & (bs - 1)) == 0 ? __ffs(bs) - 1 : -1;  // power-of-two block size: shift instead of divide
  float* tile = smem + (size_t)wid * bs2;
  if (MODE == 1) {
    for (int i = lane; i < bs2; i += 32) tile[i] = 0.f;
    __syncwarp();
  }
  for (int ib = blockIdx.x * warps_per_cta + wid; ib < mb; ib += gridDim.x * warps_per_cta) {
    int p[R], e[R], col[R], head[R];
#pragma unroll
    for (int t = 0; t < R; ++t) {
      const int r = lane + 32 * t;
      const int row = ib * bs + r;
      p[t] = 0; e[t] = 0;
      if (r < bs && row < m) { p[t] = __ldg(rowptr + row); e[t] = __ldg(rowptr + row + 1); }
      col[t] = (p[t] < e[t]) ? __ldg(colind + p[t]) : -1;
      head[t] = (col[t] >= 0) ? (shift >= 0 ? col[t] >> shift : col[t] / bs) : INT_MAX;
    }
    const int64_t base = (MODE == 0) ? 0 : __ldg(bsr_rowptr + ib);
    int step = 0;
    while (true) {
      int mymin = head[0];
#pragma unroll
      for (int t = 1; t < R; ++t) mymin = min(mymin, head[t]);
      const int bmin = warp_min(mymin);
      if (bmin == INT_MAX) break;
      const int col0 = bmin * bs;
      float* out = (MODE == 0) ? nullptr : bsr_val + (base + step) * bs2;
#pragma unroll
      for (int t = 0; t < R; ++t) {
        const int r = lane + 32 * t;
        while (head[t] == bmin) {
          if (MODE != 0) {
            const int c = col[t] - col0;
            const float v = __ldg(val + p[t]);
            const int off = (dir == SPMMDB_DIR_ROW) ? r * bs + c : c * bs + r;
            if (MODE == 1) tile[off] = v; else out[off] = v;
          }
          ++p[t];
          col[t] = (p[t] < e[t]) ? __ldg(colind + p[t]) : -1;
          head[t] = (col[t] >= 0) ? (shift >= 0 ? col[t] >> shift : col[t] / bs) : INT_MAX;
        }
      }
      if (MODE == 1) {
        __syncwarp();
        if ((bs2 & 3) == 0) {
          float4* o4 = reinterpret_cast<float4*>(out);
          float4* t4 = reinterpret_cast<float4*>(tile);
          for (int i = lane; i < bs2 / 4; i += 32) {
            __stcs(o4 + i, t4[i]);
            t4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
          }
        } else {
          for (int i = lane; i < bs2; i += 32) { out[i] = tile[i]; tile[i] = 0.f; }
        }
        __syncwarp();
      }
      if (MODE != 0 && lane == 0) bsr_colind[base + step] = bmin;
      ++step;
    }
    if (MODE == 0 && lane == 0) counts[ib] = step;
  }
}

// ---- exclusive scan of counts[n] -> out[n+1] (three small kernels) ----------
constexpr int kScanTile = 2048;  // elements per CTA (256 threads x 8)

__device__ __forceinline__ int block_exclusive_scan_256(int v, int* total) {
  __shared__ int warp_sums[8];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) warp_sums[wid] = inc;
  __syncthreads();
  if (wid == 0) {
    int w = (lane < 8) ? warp_sums[lane] : 0;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, w, o);
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

__global__ void __launch_bounds__(256) scan_tile_sums(const int* __restrict__ in, int n, int* __restrict__ sums) {
  const int base = blockIdx.x * kScanTile + threadIdx.x * 8;
  int s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) if (base + i < n) s += in[base + i];
  int total;
  block_exclusive_scan_256(s, &total);
  if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(256) scan_sums_inplace(int* sums, int ntiles) {
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < ntiles; base += 256) {
    const int i = base + threadIdx.x;
    const int v = (i < ntiles) ? sums[i] : 0;
    int total;
    const int ex = block_exclusive_scan_256(v, &total);
    if (i < ntiles) sums[i] = carry + ex;
    __syncthreads();
    if (threadIdx.x == 0) carry += total;
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256) scan_apply(const int* __restrict__ in, int n, const int* __restrict__ sums,
                                                  int* __restrict__ out) {
  const int base = blockIdx.x * kScanTile + threadIdx.x * 8;
  int v[8], s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) { v[i] = (base + i < n) ? in[base + i] : 0; s += v[i]; }
  int total;
  int ex = block_exclusive_scan_256(s, &total) + sums[blockIdx.x];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    if (base + i < n) out[base + i] = ex;
    ex += v[i];
    if (base + i == n - 1) out[n] = ex;
  }
}

// ---- bsr2csr ------------------------------------------------------------------
// Closed form of cusparseSbsr2csr (every slot of every block becomes a CSR entry, explicit zeros included):
// block p of block row ib (q-th of its cnt blocks) puts row r of the block at b0*bs^2 + r*cnt*bs + q*bs.
//
// bsr2csr_vec_kernel (bs a multiple of 4 that divides evenly into float4 pieces, row-major blocks, 16-byte
// aligned arrays): a warp owns a run of CH consecutive blocks, finds the block row of the first one with a
// single binary search and walks bsr_rowptr from there; every block is read as float4s (U blocks in flight
// per lane) and each float4 goes out with its int4 of column ids.  bsr2csr_kernel below is the scalar
// any-size / column-major path (one warp per block).
template <int BS, int U>
__global__ void __launch_bounds__(256) bsr2csr_vec_kernel(int mb, const int* __restrict__ bsr_rowptr,
                                                          const int* __restrict__ bsr_colind,
                                                          const float* __restrict__ bsr_val,
                                                          int* __restrict__ csr_rowptr, int* __restrict__ csr_colind,
                                                          float* __restrict__ csr_val, int ch) {
  constexpr int BS2 = BS * BS;
  constexpr int Q = BS / 4;                       // float4 pieces per block row
  constexpr int PIECES = BS2 / 4;                 // per block
  constexpr int V = (PIECES + 31) / 32;           // per lane per block
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int nnzb = __ldg(bsr_rowptr + mb);
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row <= (int64_t)mb * BS;
       row += (int64_t)gridDim.x * blockDim.x) {
    if (row == (int64_t)mb * BS) { csr_rowptr[row] = nnzb * BS2; break; }
    const int ib = (int)(row / BS), r = (int)(row % BS);
    const int b0 = __ldg(bsr_rowptr + ib), cnt = __ldg(bsr_rowptr + ib + 1) - b0;
    csr_rowptr[row] = b0 * BS2 + r * cnt * BS;
  }
  const int64_t nchunks = ((int64_t)nnzb + ch - 1) / ch;
  for (int64_t chunk = warp; chunk < nchunks; chunk += nwarps) {
    const int p0 = (int)(chunk * ch), p1 = min(nnzb, p0 + ch);
    int ib = 0;
    if (lane == 0) {
      int lo = 0, hi = mb;  // largest ib with bsr_rowptr[ib] <= p0
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(bsr_rowptr + mid) <= p0) lo = mid; else hi = mid;
      }
      ib = lo;
    }
    ib = __shfl_sync(0xffffffffu, ib, 0);
    int row_end = __ldg(bsr_rowptr + ib + 1);
    for (int p = p0; p < p1; p += U) {
      float4 v[U][V];
      int64_t out0[U], stride[U];
      int col0[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int pp = p + u;
        if (pp < p1) {
          while (pp >= row_end) row_end = __ldg(bsr_rowptr + (++ib) + 1);  // skips empty block rows
          const int64_t b0 = __ldg(bsr_rowptr + ib);
          stride[u] = (int64_t)(row_end - b0) * BS;
          out0[u] = b0 * BS2 + (int64_t)(pp - b0) * BS;
          col0[u] = __ldg(bsr_colind + pp) * BS;
          const float4* __restrict__ blk = reinterpret_cast<const float4*>(bsr_val + (int64_t)pp * BS2);
#pragma unroll
          for (int j = 0; j < V; ++j) {
            const int f = lane + 32 * j;
            if (PIECES % 32 == 0 || f < PIECES) v[u][j] = __ldcs(blk + f);
          }
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (p + u < p1) {
#pragma unroll
          for (int j = 0; j < V; ++j) {
            const int f = lane + 32 * j;
            if (PIECES % 32 == 0 || f < PIECES) {
              const int r = f / Q, c = (f % Q) * 4;
              const int64_t o = out0[u] + r * stride[u] + c;
              const int cc = col0[u] + c;
              __stcs(reinterpret_cast<int4*>(csr_colind + o), make_int4(cc, cc + 1, cc + 2, cc + 3));
              __stcs(reinterpret_cast<float4*>(csr_val + o), v[u][j]);
            }
          }
        }
      }
    }
  }
}

// One warp per non-zero block.  Lane 0 finds the block row with a binary search
// over bsr_rowptr; the warp then reads the block contiguously and writes each of
// its bs row pieces to its closed-form position.
__global__ void __launch_bounds__(256) bsr2csr_kernel(int dir, int mb, int bs, const int* __restrict__ bsr_rowptr,
                                                      const int* __restrict__ bsr_colind,
                                                      const float* __restrict__ bsr_val, int* __restrict__ csr_rowptr,
                                                      int* __restrict__ csr_colind, float* __restrict__ csr_val) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int nnzb = __ldg(bsr_rowptr + mb);
  const int bs2 = bs * bs;
  // row pointer: closed form, one thread per CSR row
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row <= (int64_t)mb * bs;
       row += (int64_t)gridDim.x * blockDim.x) {
    if (row == (int64_t)mb * bs) { csr_rowptr[row] = nnzb * bs2; break; }
    const int ib = (int)(row / bs), r = (int)(row % bs);
    const int b0 = __ldg(bsr_rowptr + ib), cnt = __ldg(bsr_rowptr + ib + 1) - b0;
    csr_rowptr[row] = b0 * bs2 + r * cnt * bs;
  }
  for (int64_t p = warp; p < nnzb; p += nwarps) {
    int ib = 0;
    if (lane == 0) {
      int lo = 0, hi = mb;  // largest ib with bsr_rowptr[ib] <= p
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(bsr_rowptr + mid) <= p) lo = mid; else hi = mid;
      }
      ib = lo;
    }
    ib = __shfl_sync(0xffffffffu, ib, 0);
    const int64_t b0 = __ldg(bsr_rowptr + ib), cnt = __ldg(bsr_rowptr + ib + 1) - b0;
    const int64_t q = p - b0;
    const int col0 = __ldg(bsr_colind + p) * bs;
    const float* __restrict__ blk = bsr_val + p * bs2;
    const int64_t out0 = b0 * bs2 + q * bs;
    for (int e = lane; e < bs2; e += 32) {
      const int r = e / bs, c = e - r * bs;
      const float v = (dir == SPMMDB_DIR_ROW) ? __ldcs(blk + e) : __ldcs(blk + c * bs + r);
      const int64_t o = out0 + (int64_t)r * cnt * bs + c;
      __stcs(csr_colind + o, col0 + c);
      __stcs(csr_val + o, v);
    }
  }
}

// ---- hybrid dense-block split (divide.cu:52-127 divide_matrix) -------------------
// Same bs-way merge as csr2bsr, one warp per block row, but every column block is
// first COUNTED (the rows are ascending, so a lane's entries of block b are the run
// below column (b + 1) * bs) and only blocks with count >= thr are emitted as BSR
// blocks; the entries of the other blocks go, in their original order, to the
// remainder CSR (lane l owns row l of the block row, so a row's remainder is written
// by one lane, front to back).  thr is the smallest count whose fill
// (float)(count * 1.0 / bs^2) reaches the caller's density, found on the host with
// the reference's own expression (divide.cu:89-93), so the dense / sparse decision is
// the reference's bit for bit.
// MODE 0: counts (blocks per block row, remainder entries per row).  MODE 1: fill.
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <int R, int MODE>
__global__ void __launch_bounds__(kWarpsPerCta * 32) divide_kernel(
    int dir, int m, int mb, int bs, int thr, const int* __restrict__ rowptr, const int* __restrict__ colind,
    const float* __restrict__ val, int* __restrict__ bsr_counts, int* __restrict__ csr_counts,
    const int* __restrict__ bsr_rowptr, int* __restrict__ bsr_colind, float* __restrict__ bsr_val,
    const int* __restrict__ csr_rowptr, int* __restrict__ csr_colind, float* __restrict__ csr_val,
    int warps_per_cta) {
  extern __shared__ float smem[];
  const int lane = threadIdx.x & 31;
  const int wid = threadIdx.x >> 5;
  if (wid >= warps_per_cta) return;
  const int bs2 = bs * bs;
  const int shift = (bs & (bs - 1)) == 0 ? __ffs(bs) - 1 : -1;
  float* tile = smem + (size_t)wid * bs2;
  if (MODE == 1) {
    for (int i = lane; i < bs2; i += 32) tile[i] = 0.f;
    __syncwarp();
  }
  for (int ib = blockIdx.x * warps_per_cta + wid; ib < mb; ib += gridDim.x * warps_per_cta) {
    int p[R], e[R], head[R], rem[R];
#pragma unroll
    for (int t = 0; t < R; ++t) {
      const int r = lane + 32 * t;
      const int row = ib * bs + r;
      p[t] = 0; e[t] = 0;
      if (r < bs && row < m) { p[t] = __ldg(rowptr + row); e[t] = __ldg(rowptr + row + 1); }
      head[t] = (p[t] < e[t]) ? (shift >= 0 ? __ldg(colind + p[t]) >> shift : __ldg(colind + p[t]) / bs) : INT_MAX;
      rem[t] = (MODE == 1 && r < bs && row < m) ? __ldg(csr_rowptr + row) : 0;  // MODE 1: next output slot
    }
    const int64_t base = (MODE == 0) ? 0 : __ldg(bsr_rowptr + ib);
    int step = 0;
    while (true) {
      int mymin = head[0];
#pragma unroll
      for (int t = 1; t < R; ++t) mymin = min(mymin, head[t]);
      const int bmin = warp_min(mymin);
      if (bmin == INT_MAX) break;
      const int col0 = bmin * bs;
      const int64_t lim = (int64_t)col0 + bs;  // first column past the block
      int q[R], mycnt = 0;
#pragma unroll
      for (int t = 0; t < R; ++t) {
        q[t] = p[t];
        if (head[t] == bmin) {
          while (q[t] < e[t] && __ldg(colind + q[t]) < lim) ++q[t];
          mycnt += q[t] - p[t];
        }
      }
      const bool dense = warp_sum(mycnt) >= thr;
      if (MODE == 0) {
        if (dense) ++step;
        else {
#pragma unroll
          for (int t = 0; t < R; ++t) rem[t] += q[t] - p[t];
        }
      } else if (dense) {
        float* out = bsr_val + (base + step) * bs2;
#pragma unroll
        for (int t = 0; t < R; ++t) {
          const int r = lane + 32 * t;
          for (int i = p[t]; i < q[t]; ++i) {
            const int c = __ldg(colind + i) - col0;
            tile[(dir == SPMMDB_DIR_ROW) ? r * bs + c : c * bs + r] = val ? __ldg(val + i) : 1.0f;
          }
        }
        __syncwarp();
        if ((bs2 & 3) == 0) {
          float4* o4 = reinterpret_cast<float4*>(out);
          float4* t4 = reinterpret_cast<float4*>(tile);
          for (int i = lane; i < bs2 / 4; i += 32) {
            __stcs(o4 + i, t4[i]);
            t4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
          }
        } else {
          for (int i = lane; i < bs2; i += 32) { out[i] = tile[i]; tile[i] = 0.f; }
        }
        __syncwarp();
        if (lane == 0) bsr_colind[base + step] = bmin;
        ++step;
      } else {
#pragma unroll
        for (int t = 0; t < R; ++t) {
          for (int i = p[t]; i < q[t]; ++i, ++rem[t]) {
            csr_colind[rem[t]] = __ldg(colind + i);
            csr_val[rem[t]] = val ? __ldg(val + i) : 1.0f;
          }
        }
      }
#pragma unroll
      for (int t = 0; t < R; ++t) {
        if (head[t] == bmin) {
          p[t] = q[t];
          head[t] = (p[t] < e[t]) ? (shift >= 0 ? __ldg(colind + p[t]) >> shift : __ldg(colind + p[t]) / bs) : INT_MAX;
        }
      }
    }
    if (MODE == 0) {
      if (lane == 0) bsr_counts[ib] = step;
#pragma unroll
      for (int t = 0; t < R; ++t) {
        const int r = lane + 32 * t;
        if (r < bs && ib * bs + r < m) csr_counts[ib * bs + r] = rem[t];
      }
    }
  }
}

template <int MODE>
int launch_divide(int dir, int m, int mb, int bs, int thr, const int* rowptr, const int* colind, const float* val,
                  int* bsr_counts, int* csr_counts, const int* bsr_rowptr, int* bsr_colind, float* bsr_val,
                  const int* csr_rowptr, int* csr_colind, float* csr_val, cudaStream_t st) {
  const int R = (bs + 31) / 32;
  int wpc = kWarpsPerCta;
  size_t smem = 0;
  if (MODE == 1) {
    const size_t per_warp = sizeof(float) * (size_t)bs * bs;
    while (wpc > 1 && per_warp * wpc > 200 * 1024) wpc >>= 1;
    smem = per_warp * wpc;
  }
  const int nsm = device_num_sms() > 0 ? device_num_sms() : 148;
  const int64_t want = ((int64_t)mb + wpc - 1) / wpc;
  const int grid = (int)(want < (int64_t)nsm * 16 ? (want > 0 ? want : 1) : (int64_t)nsm * 16);
#define SPMMDB_LAUNCH_DIV(RR)                                                                                    \
  do {                                                                                                           \
    if (smem > 48 * 1024)                                                                                        \
      SPMMDB_CUDA_OK(cudaFuncSetAttribute(divide_kernel<RR, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,   \
                                          (int)smem));                                                           \
    divide_kernel<RR, MODE><<<grid, kWarpsPerCta * 32, smem, st>>>(dir, m, mb, bs, thr, rowptr, colind, val,     \
                                                                   bsr_counts, csr_counts, bsr_rowptr, bsr_colind, \
                                                                   bsr_val, csr_rowptr, csr_colind, csr_val, wpc); \
  } while (0)
  if (R <= 1) SPMMDB_LAUNCH_DIV(1);
  else if (R <= 2) SPMMDB_LAUNCH_DIV(2);
  else SPMMDB_LAUNCH_DIV(4);
#undef SPMMDB_LAUNCH_DIV
  SPMMDB_CUDA_OK(cudaGetLastError());
  return SPMMDB_OK;
}

int exclusive_scan(const int* counts, int n, int* sums, int* out, cudaStream_t st) {
  const int ntiles = (n + kScanTile - 1) / kScanTile;
  scan_tile_sums<<<ntiles, 256, 0, st>>>(counts, n, sums);
  scan_sums_inplace<<<1, 256, 0, st>>>(sums, ntiles);
  scan_apply<<<ntiles, 256, 0, st>>>(counts, n, sums, out);
  SPMMDB_CUDA_OK(cudaGetLastError());
  return SPMMDB_OK;
}

template <int MODE>
int launch_csr2bsr(int dir, int m, int mb, int bs, const int* rowptr, const int* colind, const float* val,
                   int* counts, const int* bsr_rowptr, int* bsr_colind, float* bsr_val, cudaStream_t st) {
  const int R = (bs + 31) / 32;
  int wpc = kWarpsPerCta;
  size_t smem = 0;
  if (MODE == 1) {
    const size_t per_warp = sizeof(float) * (size_t)bs * bs;
    while (wpc > 1 && per_warp * wpc > 200 * 1024) wpc >>= 1;
    smem = per_warp * wpc;
  }
  const int nsm = device_num_sms() > 0 ? device_num_sms() : 148;
  const int64_t want = ((int64_t)mb + wpc - 1) / wpc;
  const int grid = (int)(want < (int64_t)nsm * 16 ? (want > 0 ? want : 1) : (int64_t)nsm * 16);
#define SPMMDB_LAUNCH_C2B(RR)                                                                                   \
  do {                                                                                                          \
    if (smem > 48 * 1024)                                                                                       \
      SPMMDB_CUDA_OK(cudaFuncSetAttribute(csr2bsr_kernel<RR, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                          (int)smem));                                                          \
    csr2bsr_kernel<RR, MODE><<<grid, kWarpsPerCta * 32, smem, st>>>(dir, m, mb, bs, rowptr, colind, val, counts, \
                                                                    bsr_rowptr, bsr_colind, bsr_val, wpc);      \
  } while (0)
  if (R <= 1) SPMMDB_LAUNCH_C2B(1);
  else if (R <= 2) SPMMDB_LAUNCH_C2B(2);
  else if (R <= 4) SPMMDB_LAUNCH_C2B(4);
  else SPMMDB_LAUNCH_C2B(8);
#undef SPMMDB_LAUNCH_C2B
  SPMMDB_CUDA_OK(cudaGetLastError());
  return SPMMDB_OK;
}

}  // namespace

// convert_br.cu: one CTA per block row (coalesced reads, shared-memory bitmap ranks)
bool convert_br_ok(int n, int bs, bool split);
int csr2bsr_br(bool fill, int dir, int m, int n, const int* rowptr, const int* colind, const float* val, int bs,
               int* counts, const int* bsr_rowptr, int* bsr_colind, float* bsr_val, cudaStream_t st);
int divide_br(bool fill, int dir, int m, int n, const int* rowptr, const int* colind, const float* val, int bs, int thr,
              int* bsr_counts, int* csr_counts, const int* bsr_rowptr, int* bsr_colind, float* bsr_val,
              const int* csr_rowptr, int* csr_colind, float* csr_val, cudaStream_t st);
// Which path: the block-row kernels win when a block row holds enough entries to keep a CTA busy (C3: 4-9 K entries
// per block row, csr2bsr 1.64 -> 0.70 ms, split 2.0-4.6 -> 0.5-0.9 ms) and lose on tiny block rows (C4: 34 entries,
// 0.15 -> 0.41 ms); with blocks under ~8 % fill the fill pass is bound by writing zeros and the warp-merge kernel's
// full-line tile stores are the faster way (C3 at 3 % fill: 1.57 vs 1.85 ms).  Phase 1 reads nnz = rowptr[m] back
// (it synchronises anyway) and leaves it, with nnzb, for phase 2 in a small cache keyed by (rowptr, m, block size).
// SPMMDB_CONVERT_BR = 0 / 1 forces the warp-merge / block-row kernels (read per call: the tests exercise both).
struct ConvNote { const int* rowptr; int m, bs, dev; int64_t nnz, nnzb; };
static ConvNote g_conv_notes[8];
static int g_conv_next = 0;
static std::mutex g_conv_mu;  // the notes are a hint between two equivalent kernels: a stale one costs time only
static void conv_note(const int* rowptr, int m, int bs, int64_t nnz, int64_t nnzb) {
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> lk(g_conv_mu);
  for (auto& e : g_conv_notes)
    if (e.rowptr == rowptr && e.m == m && e.bs == bs && e.dev == dev) { e.nnz = nnz; e.nnzb = nnzb; return; }
  g_conv_notes[g_conv_next] = ConvNote{rowptr, m, bs, dev, nnz, nnzb};
  g_conv_next = (g_conv_next + 1) % 8;
}
static bool conv_lookup(const int* rowptr, int m, int bs, int64_t* nnz, int64_t* nnzb) {
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> lk(g_conv_mu);
  for (auto& e : g_conv_notes)
    if (e.rowptr == rowptr && e.rowptr != nullptr && e.m == m && e.bs == bs && e.dev == dev) {
      *nnz = e.nnz; *nnzb = e.nnzb;
      return true;
    }
  return false;
}
// nnz < 0: unknown.  nnzb < 0: unknown / not relevant (count passes, split).
static bool use_br(int m, int n, int bs, bool split, int64_t nnz, int64_t nnzb) {
  if (!convert_br_ok(n, bs, split)) return false;
  const char* e = getenv("SPMMDB_CONVERT_BR");
  if (e != nullptr && e[0] == '0') return false;
  if (e != nullptr && e[0] == '1') return true;
  const int64_t mb = ((int64_t)m + bs - 1) / bs;
  if (nnz >= 0 && mb > 0 && nnz / mb < 512) return false;
  if (nnz >= 0 && nnzb > 0 && (double)nnz < 0.08 * (double)nnzb * bs * bs) return false;
  return true;
}
static int read_nnz(const int* rowptr, int m, cudaStream_t st, int64_t* nnz) {
  int h = 0;
  SPMMDB_CUDA_OK(cudaMemcpyAsync(&h, rowptr + m, sizeof(int), cudaMemcpyDeviceToHost, st));
  SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
  *nnz = h;
  return SPMMDB_OK;
}

int exclusive_scan_i32(const int* counts, int n, int* sums, int* out, cudaStream_t st) {
  return exclusive_scan(counts, n, sums, out, st);
}
}  // namespace spmmdb

using namespace spmmdb;

extern "C" int spmmdb_csr2bsr_nnz(int m, int n, const int* rowptr, const int* colind, int block_dim,
                                  int* bsr_rowptr, int* nnzb_host, void* stream) {
  if (m < 0 || n < 0 || block_dim <= 0 || block_dim > 256) return SPMMDB_ERR_INVALID_VALUE;
  if (bsr_rowptr == nullptr || nnzb_host == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  cudaStream_t st = (cudaStream_t)stream;
  const int mb = (m + block_dim - 1) / block_dim;
  if (mb == 0) {
    SPMMDB_CUDA_OK(cudaMemsetAsync(bsr_rowptr, 0, sizeof(int), st));
    SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
    *nnzb_host = 0;
    return SPMMDB_OK;
  }
  if (rowptr == nullptr) return SPMMDB_ERR_INVALID_VALUE;  // colind may be null when nnz == 0
  const int ntiles = (mb + kScanTile - 1) / kScanTile;
  int* ws = (int*)scratch(kScratchConvert, sizeof(int) * ((size_t)mb + ntiles + 8), st);
  if (ws == nullptr) return SPMMDB_ERR_ALLOC;
  int* counts = ws;
  int* sums = ws + mb;
  int64_t nnz = -1;
  int rc = read_nnz(rowptr, m, st, &nnz);
  if (rc != SPMMDB_OK) return rc;
  rc = use_br(m, n, block_dim, false, nnz, -1)
               ? csr2bsr_br(false, SPMMDB_DIR_ROW, m, n, rowptr, colind, nullptr, block_dim, counts, nullptr, nullptr,
                            nullptr, st)
               : launch_csr2bsr<0>(SPMMDB_DIR_ROW, m, mb, block_dim, rowptr, colind, nullptr, counts, nullptr, nullptr,
                                   nullptr, st);
  if (rc != SPMMDB_OK) return rc;
  scan_tile_sums<<<ntiles, 256, 0, st>>>(counts, mb, sums);
  scan_sums_inplace<<<1, 256, 0, st>>>(sums, ntiles);
  scan_apply<<<ntiles, 256, 0, st>>>(counts, mb, sums, bsr_rowptr);
  SPMMDB_CUDA_OK(cudaGetLastError());
  SPMMDB_CUDA_OK(cudaMemcpyAsync(nnzb_host, bsr_rowptr + mb, sizeof(int), cudaMemcpyDeviceToHost, st));
  SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
  conv_note(rowptr, m, block_dim, nnz, *nnzb_host);
  return SPMMDB_OK;
}

extern "C" int spmmdb_csr2bsr_f32(int dir, int m, int n, const int* rowptr, const int* colind, const float* val,
                                  int block_dim, const int* bsr_rowptr, int* bsr_colind, float* bsr_val,
                                  void* stream) {
  if (m < 0 || n < 0 || block_dim <= 0 || block_dim > 256) return SPMMDB_ERR_INVALID_VALUE;
  if (dir != SPMMDB_DIR_ROW && dir != SPMMDB_DIR_COLUMN) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  const int mb = (m + block_dim - 1) / block_dim;
  if (mb == 0) return SPMMDB_OK;
  if (rowptr == nullptr || bsr_rowptr == nullptr) return SPMMDB_ERR_INVALID_VALUE;  // others may be null when nnz == 0
  cudaStream_t st = (cudaStream_t)stream;
  int64_t nnz = -1, nnzb = -1;
  conv_lookup(rowptr, m, block_dim, &nnz, &nnzb);
  if (use_br(m, n, block_dim, false, nnz, nnzb) && bsr_colind != nullptr && bsr_val != nullptr && val != nullptr)
    return csr2bsr_br(true, dir, m, n, rowptr, colind, val, block_dim, nullptr, bsr_rowptr, bsr_colind, bsr_val, st);
  if (block_dim <= 128)
    return launch_csr2bsr<1>(dir, m, mb, block_dim, rowptr, colind, val, nullptr, bsr_rowptr, bsr_colind, bsr_val, st);
  // very large blocks: clear + scatter (needs nnzb: read it back)
  int nnzb_big = 0;
  SPMMDB_CUDA_OK(cudaMemcpyAsync(&nnzb_big, bsr_rowptr + mb, sizeof(int), cudaMemcpyDeviceToHost, st));
  SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
  SPMMDB_CUDA_OK(cudaMemsetAsync(bsr_val, 0, sizeof(float) * (size_t)nnzb_big * block_dim * block_dim, st));
  return launch_csr2bsr<2>(dir, m, mb, block_dim, rowptr, colind, val, nullptr, bsr_rowptr, bsr_colind, bsr_val, st);
}

extern "C" int spmmdb_bsr2csr_f32(int dir, int mb, int nb, const int* bsr_rowptr, const int* bsr_colind,
                                  const float* bsr_val, int block_dim, int* csr_rowptr, int* csr_colind,
                                  float* csr_val, void* stream) {
  if (mb < 0 || nb < 0 || block_dim <= 0) return SPMMDB_ERR_INVALID_VALUE;
  if (dir != SPMMDB_DIR_ROW && dir != SPMMDB_DIR_COLUMN) return SPMMDB_ERR_INVALID_VALUE;
  if (bsr_rowptr == nullptr || csr_rowptr == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  cudaStream_t st = (cudaStream_t)stream;
  const int nsm = device_num_sms() > 0 ? device_num_sms() : 148;
  const bool aligned = ((((uintptr_t)bsr_val) | ((uintptr_t)csr_val) | ((uintptr_t)csr_colind)) & 15) == 0;
  const int force_scalar = SPMMDB_TUNE("BSR2CSR_SCALAR", 0);
  if (dir == SPMMDB_DIR_ROW && aligned && !force_scalar &&
      (block_dim == 8 || block_dim == 16 || block_dim == 32 || block_dim == 64)) {
    // a warp's run of blocks: ~16 KB of block values, so the one binary search per run is noise
    const int ch = max(1, 4096 / (block_dim * block_dim));
    const int grid = nsm * 8;
#define SPMMDB_B2C(BS, U) \
  bsr2csr_vec_kernel<BS, U><<<grid, 256, 0, st>>>(mb, bsr_rowptr, bsr_colind, bsr_val, csr_rowptr, csr_colind, csr_val, ch)
    if (block_dim == 8) SPMMDB_B2C(8, 8);
    else if (block_dim == 16) SPMMDB_B2C(16, 4);
    else if (block_dim == 32) SPMMDB_B2C(32, 1);
    else SPMMDB_B2C(64, 1);
#undef SPMMDB_B2C
  } else {
    bsr2csr_kernel<<<nsm * 8, 256, 0, st>>>(dir, mb, block_dim, bsr_rowptr, bsr_colind, bsr_val, csr_rowptr,
                                            csr_colind, csr_val);
  }
  SPMMDB_CUDA_OK(cudaGetLastError());
  return SPMMDB_OK;
}

// Smallest block count that the reference's test `float occupy = (count * 1.0) / bnum; occupy >= density`
// (divide.cu:89-93) accepts; at least 1 (the reference would also emit EMPTY blocks for density <= 0).
extern "C" int spmmdb_divide_threshold(int block_dim, float density) {
  if (block_dim <= 0) return SPMMDB_ERR_INVALID_VALUE;
  const int bnum = block_dim * block_dim;
  auto ok = [&](int64_t c) { float occupy = (float)((c * 1.0) / bnum); return occupy >= density; };
  if (!(density == density)) return INT_MAX;  // NaN: nothing passes
  int64_t lo = 1, hi = INT_MAX;               // ok() is monotone in c
  if (ok(lo)) return 1;
  if (!ok(hi)) return INT_MAX;
  while (hi - lo > 1) {
    const int64_t mid = lo + (hi - lo) / 2;
    if (ok(mid)) hi = mid; else lo = mid;
  }
  return (int)hi;
}

extern "C" int spmmdb_divide_nnz(int m, int n, const int* rowptr, const int* colind, int block_dim, float density,
                                 int* bsr_rowptr, int* csr_rowptr, int* nnzb_host, int* csr_nnz_host, void* stream) {
  if (m < 0 || n < 0 || block_dim <= 0 || block_dim > 128) return SPMMDB_ERR_INVALID_VALUE;
  if (bsr_rowptr == nullptr || csr_rowptr == nullptr || nnzb_host == nullptr || csr_nnz_host == nullptr)
    return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  cudaStream_t st = (cudaStream_t)stream;
  const int mb = (m + block_dim - 1) / block_dim;
  if (mb == 0) {
    SPMMDB_CUDA_OK(cudaMemsetAsync(bsr_rowptr, 0, sizeof(int), st));
    SPMMDB_CUDA_OK(cudaMemsetAsync(csr_rowptr, 0, sizeof(int), st));
    SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
    *nnzb_host = 0; *csr_nnz_host = 0;
    return SPMMDB_OK;
  }
  if (rowptr == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  const int ntiles = (m + kScanTile - 1) / kScanTile + 1;
  int* ws = (int*)scratch(kScratchConvert, sizeof(int) * ((size_t)mb + (size_t)m + ntiles + 8), st);
  if (ws == nullptr) return SPMMDB_ERR_ALLOC;
  int* bsr_counts = ws;
  int* csr_counts = ws + mb;
  int* sums = ws + mb + m;
  const int thr = spmmdb_divide_threshold(block_dim, density);
  int64_t nnz = -1;
  int rc = read_nnz(rowptr, m, st, &nnz);
  if (rc != SPMMDB_OK) return rc;
  conv_note(rowptr, m, -block_dim, nnz, -1);  // negative block size: the split's own note
  rc = use_br(m, n, block_dim, true, nnz, -1)
               ? divide_br(false, SPMMDB_DIR_ROW, m, n, rowptr, colind, nullptr, block_dim, thr, bsr_counts, csr_counts,
                           nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, st)
               : launch_divide<0>(SPMMDB_DIR_ROW, m, mb, block_dim, thr, rowptr, colind, nullptr, bsr_counts,
                                  csr_counts, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, st);
  if (rc != SPMMDB_OK) return rc;
  if ((rc = exclusive_scan(bsr_counts, mb, sums, bsr_rowptr, st)) != SPMMDB_OK) return rc;
  if ((rc = exclusive_scan(csr_counts, m, sums, csr_rowptr, st)) != SPMMDB_OK) return rc;
  SPMMDB_CUDA_OK(cudaMemcpyAsync(nnzb_host, bsr_rowptr + mb, sizeof(int), cudaMemcpyDeviceToHost, st));
  SPMMDB_CUDA_OK(cudaMemcpyAsync(csr_nnz_host, csr_rowptr + m, sizeof(int), cudaMemcpyDeviceToHost, st));
  SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
  return SPMMDB_OK;
}

// Fill-decile histogram of the non-empty bs x bs blocks (calculate_block_density_dist, block_density_dist.cpp:47-86):
// dist[d] = number of blocks whose fill lands in decile d, with the reference's own expression
//   float occupy = (num * 1.0) / (bsize * bsize); int ith = (int)(occupy * 10); if (ith >= 10) ith = 9;
// evaluated on the host to find, per decile, the smallest count that reaches it; the device then counts the blocks
// with at least that many entries (the split's count pass) and the histogram is the difference of neighbours.
extern "C" int spmmdb_block_fill_histogram(int m, int n, const int* rowptr, const int* colind, int block_dim,
                                           int64_t* dist10_host, void* stream) {
  if (m < 0 || n < 0 || block_dim <= 0 || block_dim > 128 || dist10_host == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  for (int d = 0; d < 10; ++d) dist10_host[d] = 0;
  const int mb = (m + block_dim - 1) / block_dim;
  if (mb == 0) return SPMMDB_OK;
  if (rowptr == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  cudaStream_t st = (cudaStream_t)stream;
  const int bnum = block_dim * block_dim;
  auto decile = [&](int num) {
    float occupy = (num * 1.0) / bnum;
    int ith = (int)(occupy * 10);
    return ith >= 10 ? 9 : ith;
  };
  int thr[10];
  for (int d = 0; d < 10; ++d) {
    thr[d] = INT_MAX;
    for (int c = 1; c <= bnum; ++c)
      if (decile(c) >= d) { thr[d] = c; break; }  // decile() is monotone in the count
  }
  const int ntiles = (m + kScanTile - 1) / kScanTile + 1;
  int* ws = (int*)scratch(kScratchConvert, sizeof(int) * (2 * (size_t)mb + (size_t)m + ntiles + 16), st);
  if (ws == nullptr) return SPMMDB_ERR_ALLOC;
  int* bsr_counts = ws;
  int* csr_counts = ws + mb;
  int* sums = ws + mb + m;
  int* tmp_rowptr = sums + ntiles + 4;  // mb + 1 entries
  int64_t nnz = -1;
  int rc = read_nnz(rowptr, m, st, &nnz);
  if (rc != SPMMDB_OK) return rc;
  int at_least[11] = {0};
  for (int d = 0; d < 10; ++d) {
    if (thr[d] == INT_MAX) { at_least[d] = 0; continue; }
    if (d > 0 && thr[d] == thr[d - 1]) { at_least[d] = -1; continue; }  // same class as the decile below: filled in later
    rc = use_br(m, n, block_dim, true, nnz, -1)
             ? divide_br(false, SPMMDB_DIR_ROW, m, n, rowptr, colind, nullptr, block_dim, thr[d], bsr_counts, csr_counts,
                         nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, st)
             : launch_divide<0>(SPMMDB_DIR_ROW, m, mb, block_dim, thr[d], rowptr, colind, nullptr, bsr_counts, csr_counts,
                                nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, st);
    if (rc != SPMMDB_OK) return rc;
    if ((rc = exclusive_scan(bsr_counts, mb, sums, tmp_rowptr, st)) != SPMMDB_OK) return rc;
    SPMMDB_CUDA_OK(cudaMemcpyAsync(&at_least[d], tmp_rowptr + mb, sizeof(int), cudaMemcpyDeviceToHost, st));
    SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
  }
  for (int d = 1; d < 10; ++d)
    if (at_least[d] < 0) at_least[d] = at_least[d - 1];
  for (int d = 0; d < 10; ++d) dist10_host[d] = (int64_t)at_least[d] - at_least[d + 1];
  return SPMMDB_OK;
}

extern "C" int spmmdb_divide_f32(int dir, int m, int n, const int* rowptr, const int* colind, const float* val,
                                 int block_dim, float density, const int* bsr_rowptr, int* bsr_colind,
                                 float* bsr_val, const int* csr_rowptr, int* csr_colind, float* csr_val,
                                 void* stream) {
  if (m < 0 || n < 0 || block_dim <= 0 || block_dim > 128) return SPMMDB_ERR_INVALID_VALUE;
  if (dir != SPMMDB_DIR_ROW && dir != SPMMDB_DIR_COLUMN) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  const int mb = (m + block_dim - 1) / block_dim;
  if (mb == 0) return SPMMDB_OK;
  if (rowptr == nullptr || bsr_rowptr == nullptr || csr_rowptr == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  const int thr = spmmdb_divide_threshold(block_dim, density);
  int64_t nnz = -1, nnzb = -1;
  conv_lookup(rowptr, m, -block_dim, &nnz, &nnzb);
  if (use_br(m, n, block_dim, true, nnz, -1))
    return divide_br(true, dir, m, n, rowptr, colind, val, block_dim, thr, nullptr, nullptr, bsr_rowptr, bsr_colind,
                     bsr_val, csr_rowptr, csr_colind, csr_val, (cudaStream_t)stream);
  return launch_divide<1>(dir, m, mb, block_dim, thr, rowptr, colind, val, nullptr, nullptr, bsr_rowptr, bsr_colind,
                          bsr_val, csr_rowptr, csr_colind, csr_val, (cudaStream_t)stream);
}
