// bsrmm_simt.cu -- BSR x dense on CUDA cores (fp32 FMA, fp32 accumulate): the
// exact-arithmetic path (SPMMDB_MATH_FP32) and the path for block sizes the
// tensor-core kernel does not take.  Replaces bsrmm_large_blockdim_device /
// bsrmm_general_blockdim_device / bsrmmn{n,t}_small_blockdim_device
// (rocsparse_bsrmm_impl.h:14-472) with ONE kernel whose output column comes from
// the block index (the reference uses blockDim.y there, impl:27/:326/:413).
//
// One CTA = one block row x one BS-row slice x TN output columns; thread t owns
// column t of the tile and BS accumulators.  Per non-zero block the A sub-block
// is staged transposed in shared memory and read back as broadcast float4s
// (BS/4 LDS.128 per B element), B is read coalesced (row-major B) or as
// contiguous runs (column-major B).
#include "common.cuh"

namespace spmmdb {
namespace {

struct BsrArgs {
  int dir, trans_b, mb, n, bs;
  float alpha, beta;
  const float* val;
  const int* rowptr;
  const int* colind;
  const float* B;
  int64_t ldb;
  float* C;
  int64_t ldc;
  int c_layout;
  int64_t b_rows, c_rows;  // rows of B that exist / rows of C that may be written (RowClip)
};

template <int BS, int TN>
__global__ void __launch_bounds__(TN) bsrmm_simt_kernel(const BsrArgs a) {
  __shared__ __align__(16) float As[BS * BS];  // As[j * BS + r] = A_sub[r][j]
  const int ib = blockIdx.x;
  const int rt = blockIdx.z;                 // row slice inside the block (bs > BS)
  const int c = blockIdx.y * TN + threadIdx.x;
  const bool cok = c < a.n;
  const int bs = a.bs;
  const int r0 = rt * BS;
  const int rows = min(BS, bs - r0);
  const int jt_n = (bs + BS - 1) / BS;
  const int64_t bs2 = (int64_t)bs * bs;

  float acc[BS];
#pragma unroll
  for (int r = 0; r < BS; ++r) acc[r] = 0.f;

  const int pb = __ldg(a.rowptr + ib), pe = __ldg(a.rowptr + ib + 1);
  for (int p = pb; p < pe; ++p) {
    const int64_t col0 = (int64_t)__ldg(a.colind + p) * bs;
    const float* __restrict__ blk = a.val + (int64_t)p * bs2;
    for (int jt = 0; jt < jt_n; ++jt) {
      const int j0 = jt * BS;
      const int cols = min(BS, bs - j0);
      __syncthreads();
      for (int i = threadIdx.x; i < BS * BS; i += TN) {
        // read order follows the block's storage order for coalescing
        int r, j;
        if (a.dir == SPMMDB_DIR_ROW) { r = i / BS; j = i % BS; } else { j = i / BS; r = i % BS; }
        float v = 0.f;
        if (r < rows && j < cols)
          v = (a.dir == SPMMDB_DIR_ROW) ? __ldg(blk + (int64_t)(r0 + r) * bs + j0 + j)
                                        : __ldg(blk + (int64_t)(j0 + j) * bs + r0 + r);
        As[j * BS + r] = v;
      }
      __syncthreads();
      if (cok) {
#pragma unroll 4
        for (int j = 0; j < BS; ++j) {
          if (j < cols) {
            const int64_t brow = col0 + j0 + j;
            const float b = brow >= a.b_rows ? 0.f
                            : (a.trans_b ? __ldg(a.B + brow * a.ldb + c) : __ldg(a.B + (int64_t)c * a.ldb + brow));
            if (BS % 4 == 0) {
              const float4* __restrict__ a4 = reinterpret_cast<const float4*>(As + j * BS);
#pragma unroll
              for (int r4 = 0; r4 < BS / 4; ++r4) {
                const float4 av = a4[r4];
                acc[4 * r4 + 0] = fmaf(av.x, b, acc[4 * r4 + 0]);
                acc[4 * r4 + 1] = fmaf(av.y, b, acc[4 * r4 + 1]);
                acc[4 * r4 + 2] = fmaf(av.z, b, acc[4 * r4 + 2]);
                acc[4 * r4 + 3] = fmaf(av.w, b, acc[4 * r4 + 3]);
              }
            } else {
#pragma unroll
              for (int r = 0; r < BS; ++r) acc[r] = fmaf(As[j * BS + r], b, acc[r]);
            }
          }
        }
      }
    }
  }

  if (!cok) return;
  const int64_t row0 = (int64_t)ib * bs + r0;
#pragma unroll
  for (int r = 0; r < BS; ++r) {
    if (r < rows && row0 + r < a.c_rows) {
      float* dst = (a.c_layout == SPMMDB_COL_MAJOR) ? a.C + (int64_t)c * a.ldc + row0 + r
                                                    : a.C + (row0 + r) * a.ldc + c;
      // same epilogue arithmetic as rocsparse_bsrmm_impl.h:382-388
      *dst = (a.beta == 0.f) ? a.alpha * acc[r] : fmaf(a.beta, *dst, a.alpha * acc[r]);
    }
  }
}

template <int BS>
int launch(const BsrArgs& a, cudaStream_t st) {
  constexpr int TN = 128;
  dim3 grid((unsigned)a.mb, (unsigned)((a.n + TN - 1) / TN), (unsigned)((a.bs + BS - 1) / BS));
  bsrmm_simt_kernel<BS, TN><<<grid, TN, 0, st>>>(a);
  SPMMDB_CUDA_OK(cudaGetLastError());
  return SPMMDB_OK;
}

}  // namespace

int bsrmm_simt(int dir, int trans_b, int mb, int n, int bs, float alpha, const float* val, const int* rowptr,
               const int* colind, const float* B, int64_t ldb, float beta, float* C, int64_t ldc, int c_layout,
               cudaStream_t st) {
  const RowClip clip = current_row_clip();
  BsrArgs a{dir, trans_b, mb, n, bs, alpha, beta, val, rowptr, colind, B, ldb, C, ldc, c_layout, clip.b_rows, clip.c_rows};
  if (bs <= 4) return launch<4>(a, st);
  if (bs <= 8) return launch<8>(a, st);
  if (bs <= 16) return launch<16>(a, st);
  return launch<32>(a, st);
}

}  // namespace spmmdb
