// bsrmm_tc.cu -- BSR x dense on the sm_100a tensor cores (tcgen05.mma, kind::tf32 or
// kind::f16 with bf16 inputs, fp32 accumulators in TMEM), operands staged by TMA.
// Replaces bsrmm_large_blockdim_device / bsrmm_general_blockdim_device
// (rocsparse_bsrmm_impl.h:315-472) and the per-block cublasSgemm loop of
// block_cublas.cu:123-136 for block sizes 16 / 32 / 64.
//
// Formulation (transposed product, so that the gathered operand is the wide one):
//   for a "super-row" of G = NT / bs consecutive block rows (NT rows of A) and every
//   distinct column block j that occurs in it
//       D[f, r] += sum_c  Bslab_j[c, f] * Astrip_j[r, c]        f < 128 features, r < NT, c < bs
//   i.e. one tcgen05.mma with M = 128 (features), N = run * bs (a run of consecutive
//   block rows that hold block j), K = bs.  The B slab (bs rows of B, all features) is
//   fetched ONCE per super-row and column block and shared by all G block rows, which
//   divides the L2 -> SM gather traffic of a plain per-block kernel by the number of
//   blocks per strip.  D lives in TMEM: lane = feature, column = row of A.
//
//   operand "A" of the MMA = B slab:   row-major B[k][ldb] read by TMA boxes of
//       {128 bytes of features, bs rows} -> MN-major (feature-major) canonical layout
//       (bf16: SWIZZLE_128B; tf32: SWIZZLE_128B with 32-byte atoms, the only MN-major 32-bit layout)
//   operand "B" of the MMA = A strip:  each bs x bs block is one TMA box {bs, bs} of
//       bsr_val viewed as a [nnzb*bs][bs] matrix -> K-major (row-major blocks, DIR_ROW) or
//       MN-major (column-major blocks, DIR_COLUMN) canonical layout, swizzle = row pitch
//
// A pre-pass (bsr_plan_kernel, one warp per super-row) does the G-way merge of the block rows'
// sorted column lists and leaves one 8-byte record {column block, presence mask} per step.
//
// Warp roles ((P + 5) warps; two CTAs per SM with 256 TMEM columns each when the accumulator fits,
// so that one CTA's barrier round trips hide behind the other's MMAs):
//   warps 0..P-1  producers: walk the plan, warp k % P fills step slot k: arms the stage's mbarrier
//                 and issues the TMA loads (one 3-D box for the B slab, one box per present A block)
//   warp P        MMA issuer (lane 0 issues): run detection, tcgen05.mma, tcgen05.commit
//   warps P+1..   epilogue (4 warps = 128 TMEM lanes): tcgen05.ld, alpha/beta, stores to C
// Two accumulator buffers in TMEM let the epilogue of super-row s overlap the MMAs of s+1.
// Measured on B200 (profiles/r1c_bsr_tc_notes.md): the kernel is bound by the per-step latency of
// the MMA-issuing warp (mbarrier wait + ~25 instructions per tcgen05.mma), not by HBM or the tensor pipe.
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdlib.h>

#include "common.cuh"

namespace spmmdb {
namespace {

constexpr int kMaxStages = 8;
constexpr uint32_t kEndFlag = 0x80000000u;  // step meta: end of a super-row

// ---------------------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra WAIT_DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "WAIT_DONE:\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t* slot, int cols) {  // cols: 256 (two CTAs per SM) or 512
  if (cols == 256)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(slot)) : "memory");
  else
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(slot)) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, int cols) {
  if (cols == 256) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(addr) : "memory");
  else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(addr) : "memory");
}
// elect.sync: exactly one lane of the (converged) warp gets true, and ptxas KNOWS the guarded region runs on
// one lane -- with `lane == 0` it does not, and wraps every tcgen05.mma in an ELECT / R2UR.BROADCAST /
// BRA.U.ANY loop over the distinct operand values it fears the lanes might hold.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
template <bool BF16>
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  if constexpr (BF16) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}
// 32 lanes x 32 consecutive columns -> 32 registers per thread (thread = TMEM lane)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// Shared-memory matrix descriptors follow cute::UMMA::SmemDescriptor: bits 0-13 start address >> 4,
// 16-29 leading byte offset >> 4, 32-45 stride byte offset >> 4, 46-47 version (1 = Blackwell),
// 61-63 swizzle mode.
constexpr uint32_t kLayoutSw128 = 2, kLayoutSw64 = 4, kLayoutSw32 = 6, kLayoutSw128B32 = 1;

// Developer timing experiments (skip loads / MMAs, cycle counters, device printf) exist only in builds made with
// -DSPMMDB_BSR_DEBUG; in the release kernel the flag word is the constant 0 and the branches fold away.
#ifdef SPMMDB_BSR_DEBUG
#define SPMMDB_BSR_DBG(a) ((a).dbg)
#else
#define SPMMDB_BSR_DBG(a) 0
#endif

struct TcArgs {
  int mb, n, bs, G, NT, tiles, nchunks, stages;
  int dir, c_layout;
  int num_super;
  int64_t m_valid;        // rows of C that may be written (mb * bs unless a RowClip is set)
  int tmem_cols;          // 512 (one CTA per SM) or 256 (two CTAs per SM)
  int dbg;                // developer timing experiments: 1 = no A loads, 2 = no B loads, 4 = no MMAs
  int b_one;              // B slab fetched by one 3-D TMA box (n is a multiple of the chunk width)
  const int* rowptr;
  const int* colind;
  const uint2* plan;      // step records {column block, presence mask}, super-row s starts at rowptr[s*G]
  const int* plan_count;  // steps per super-row
  float* C;
  int64_t ldc;
  float alpha, beta;
  int* counter;
  uint32_t chunk_bytes;   // bs * 128: one {128-byte feature chunk, bs rows} TMA box of B
  uint32_t block_bytes;   // bs * bs * esize: one block of A
  uint32_t stage_b_bytes; // tiles * chunks_per_tile * chunk_bytes
  uint32_t stage_bytes;   // stage_b_bytes + NT * bs * esize, multiple of 1024
};

struct StepMeta { uint32_t mask; int s; };
struct AccInfo { int s; uint32_t touched; };

struct __align__(8) TcSmemCtl {
  uint64_t full[kMaxStages];
  uint64_t empty[kMaxStages];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  StepMeta meta[kMaxStages];
  AccInfo acc[2];
  int srow[2];
  uint32_t tmem_base;
};

// ---------------------------------------------------------------------------------------
// Plan: the G-way merge of a super-row's sorted column-block lists, done once per call by
// one warp per super-row (thousands of warps in flight hide the index-load latency; inside
// the MMA kernel the same merge would serialise on one producer warp).  Step t of super-row
// s = the t-th distinct column block j of its block rows and the mask of block rows that
// hold (i, j).  Records live at plan[rowptr[s*G] + t]: steps(s) <= blocks(s), no scan needed.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) bsr_plan_kernel(int mb, int G, int num_super, const int* __restrict__ rowptr,
                                                       const int* __restrict__ colind, uint2* __restrict__ plan,
                                                       int* __restrict__ plan_count) {
  constexpr int W = 8;
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int s = warp; s < num_super; s += nwarps) {
    const int br = s * G + lane;
    int cur = 0, end = 0;
    if (lane < G && br < mb) { cur = __ldg(rowptr + br); end = __ldg(rowptr + br + 1); }
    const int base = __shfl_sync(0xffffffffu, cur, 0);
    int q[W];
    bool refill = true;
    int t = 0;
    while (true) {
      if (refill) {
#pragma unroll
        for (int i = 0; i < W; ++i) q[i] = (cur + i < end) ? __ldg(colind + cur + i) : 0x7fffffff;
      }
      const int col = q[0];
      int jmin = col;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) jmin = min(jmin, __shfl_xor_sync(0xffffffffu, jmin, o));
      if (jmin == 0x7fffffff) break;
      const bool mine = (col == jmin);
      const uint32_t present = __ballot_sync(0xffffffffu, mine);
      if (lane == 0) plan[base + t] = make_uint2((uint32_t)jmin, present);
      ++t;
      if (mine) {
        ++cur;
#pragma unroll
        for (int i = 0; i + 1 < W; ++i) q[i] = q[i + 1];
        q[W - 1] = 0x7fffffff;
      }
      // a lane that advanced W times has an empty window: refill everybody together
      refill = __any_sync(0xffffffffu, mine && (cur < end) && (q[0] == 0x7fffffff));
    }
    if (lane == 0) plan_count[s] = t;
  }
}

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void tma_load_2d_addr(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
      "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void tma_load_3d_addr(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2,
                                                 uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
          dst),
      "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
      : "memory");
}

template <bool BF16, int P>
__global__ void __launch_bounds__((P + 5) * 32, 2)
bsrmm_tc_kernel(const TcArgs a, const __grid_constant__ CUtensorMap mapB, const __grid_constant__ CUtensorMap mapA) {
  constexpr int ESIZE = BF16 ? 2 : 4;
  constexpr int FC = 128 / ESIZE;           // features per 128-byte chunk
  constexpr int CPT = 128 / FC;             // chunks per 128-feature tile
  constexpr int KI = 32 / ESIZE;            // K of one tcgen05.mma
  constexpr uint32_t FMT = BF16 ? 1u : 2u;  // F16F32Format: BF16 = 1, TF32 = 2
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // stage buffers first (1024-byte aligned for the 128B swizzle), control block after them
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  TcSmemCtl* ctl = (TcSmemCtl*)(smem + (size_t)a.stages * a.stage_bytes);
  // column-major C: one 32 x 33 float transpose tile per epilogue warp, after the control block
  float* tbuf_all = (float*)(((uintptr_t)(ctl + 1) + 15) & ~(uintptr_t)15);
  // warp index broadcast from lane 0: the role dispatch below becomes a warp-uniform branch for the compiler,
  // which lets the MMA / TMA operands live in uniform registers
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < a.stages; ++i) { mbar_init(&ctl->full[i], 1); mbar_init(&ctl->empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&ctl->tmem_full[i], 2); mbar_init(&ctl->tmem_empty[i], 4); }
    fence_barrier_init();
    tma_prefetch_desc(&mapB);
    tma_prefetch_desc(&mapA);
  }
  if (warp == P + 1) tmem_alloc(&ctl->tmem_base, a.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = ctl->tmem_base;
  const int acc_cols = a.tiles * a.NT;  // <= 256

  if (warp < P) {
    // ===================== producers =====================
    // All P warps walk the same step sequence; step slot k (counted over the CTA's whole life,
    // end-of-super-row markers included) is filled by warp k % P into stage k % stages.
    int stage = 0, own = 0;
    uint32_t phase = 0;
    int it = 0, s_next = 0;
    if (warp == 0 && lane == 0) s_next = atomicAdd(a.counter, 1);
    const uint32_t smem_base = smem_u32(smem);
    long long dbg_wait = 0, dbg_steps = 0;
    const long long dbg_t0 = SPMMDB_BSR_DBG(a) ? clock64() : 0;
    while (true) {
      if (warp == 0 && lane == 0) ctl->srow[it & 1] = s_next;
      if (P > 1) named_bar_sync(1, 32 * P); else __syncwarp();
      const int s = *(volatile int*)&ctl->srow[it & 1];
      ++it;
      const bool term = s >= a.num_super;
      if (warp == 0 && lane == 0 && !term) s_next = atomicAdd(a.counter, 1);
      int cnt = 0, base = 0, cur = 0;
      if (!term) {
        base = __ldg(a.rowptr + s * a.G);
        cnt = __ldg(a.plan_count + s);
        const int br = s * a.G + lane;
        if (lane < a.G && br < a.mb) cur = __ldg(a.rowptr + br);
      }
      for (int t0 = 0; t0 < cnt; t0 += 32) {
        uint2 rec = make_uint2(0u, 0u);
        if (t0 + lane < cnt) rec = __ldg(a.plan + base + t0 + lane);
        const int nt = min(32, cnt - t0);
        for (int i = 0; i < nt; ++i) {
          const int j = (int)__shfl_sync(0xffffffffu, rec.x, i);
          const uint32_t mask = __shfl_sync(0xffffffffu, rec.y, i);
          const bool mine = (mask >> lane) & 1u;
          if (own == warp) {
            const long long tw0 = (SPMMDB_BSR_DBG(a) & 32) ? clock64() : 0;
            mbar_wait(&ctl->empty[stage], phase ^ 1);
            if (SPMMDB_BSR_DBG(a) & 32) { dbg_wait += clock64() - tw0; ++dbg_steps; }
            uint64_t* fb = &ctl->full[stage];
            const uint32_t sB = smem_base + (uint32_t)stage * a.stage_bytes;
            if (lane == 0) {
              ctl->meta[stage].mask = mask;
              mbar_arrive_expect_tx(fb, ((SPMMDB_BSR_DBG(a) & 2) ? 0u : a.nchunks * a.chunk_bytes) +
                                            ((SPMMDB_BSR_DBG(a) & 1) ? 0u : __popc(mask) * a.block_bytes));
            }
            __syncwarp();
            if (SPMMDB_BSR_DBG(a) & 2) {
            } else if (a.b_one) {
              if (lane == 0) tma_load_3d_addr(sB, &mapB, 0, j * a.bs, 0, fb);
            } else if (lane < a.nchunks) {
              tma_load_2d_addr(sB + (uint32_t)lane * a.chunk_bytes, &mapB, lane * FC, j * a.bs, fb);
            }
            if (mine && !(SPMMDB_BSR_DBG(a) & 1)) tma_load_2d_addr(sB + a.stage_b_bytes + (uint32_t)lane * a.block_bytes, &mapA, 0, cur * a.bs, fb);
          }
          cur += mine ? 1 : 0;
          if (++own == P) own = 0;
          if (++stage == a.stages) { stage = 0; phase ^= 1; }
        }
      }
      // end of the super-row (s = -1: no more work)
      if (own == warp) {
        mbar_wait(&ctl->empty[stage], phase ^ 1);
        if (lane == 0) {
          ctl->meta[stage].mask = kEndFlag;
          ctl->meta[stage].s = term ? -1 : s;
          mbar_arrive(&ctl->full[stage]);
        }
      }
      if (++own == P) own = 0;
      if (++stage == a.stages) { stage = 0; phase ^= 1; }
      if (term) break;
    }
    if ((SPMMDB_BSR_DBG(a) & 32) && blockIdx.x == 0 && lane == 0)
      printf("producer %d: total %lld cyc, own steps %lld, wait(empty) %lld cyc\n", warp, clock64() - dbg_t0, dbg_steps,
             dbg_wait);
  } else if (warp == P) {
    // ===================== MMA issuer =====================
    // The whole warp walks the loop so that addresses and descriptors stay warp-uniform (uniform
    // registers feed tcgen05.mma without per-operand R2UR moves); lane 0 alone issues.
    {
      const bool leader = lane == 0;
      const uint32_t tmem_base = __shfl_sync(0xffffffffu, ctl->tmem_base, 0);  // warp-uniform copy
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase[2] = {0, 0};
      const uint32_t rp = (uint32_t)a.bs * ESIZE;  // row pitch of a block in shared memory
      const uint32_t a_layout = rp == 128 ? kLayoutSw128 : (rp == 64 ? kLayoutSw64 : kLayoutSw32);
      // dbg & 64: timing experiment only (wrong results) -- describe the slab as a K-major operand
      const bool kmaj_exp = (SPMMDB_BSR_DBG(a) & 64) != 0;
      const uint32_t idesc_base = (1u << 4) | (FMT << 7) | (FMT << 10) | ((kmaj_exp ? 0u : 1u) << 15) |
                                  ((a.dir == SPMMDB_DIR_COLUMN ? 1u : 0u) << 16) | ((128u >> 4) << 24);
      const int ksteps = a.bs / KI;
      const int max_run = 256 / a.bs;
      // descriptor halves that never change (hi: SBO, version 1, swizzle mode; lo bits 16+: LBO)
      // MMA operand A = B slab, MN-major: 128-byte feature chunks LBO apart, k groups SBO apart
      //   bf16: SWIZZLE_128B, groups of 8 k rows (1024 B); tf32: 32-bit MN-major operands exist only in
      //   the 128B-swizzle-with-32B-atoms layout, groups of 4 k rows (512 B)
      const uint32_t a_hi = kmaj_exp ? ((1024u >> 4) | (1u << 14) | (kLayoutSw128 << 29))
                                     : (((BF16 ? 1024u : 512u) >> 4) | (1u << 14) | ((BF16 ? kLayoutSw128 : kLayoutSw128B32) << 29));
      const uint32_t a_lo_fixed = kmaj_exp ? (1u << 16) : (((a.chunk_bytes >> 4) & 0x3fffu) << 16);
      const uint32_t a_kstep = kmaj_exp ? 2u : ((uint32_t)(KI * 128) >> 4);  // KI rows of 128 B per k-step
      // MMA operand B = A strip.  DIR_ROW: K-major, advance 32 B inside the swizzle atom per k-step,
      //   8-row groups SBO = 8 * rp apart.  DIR_COLUMN: MN-major, blocks LBO apart, k groups 8 * rp apart.
      const uint32_t b_hi = ((8u * rp) >> 4) | (1u << 14) | (a_layout << 29);
      const uint32_t b_lo_fixed = (a.dir == SPMMDB_DIR_ROW ? 1u : ((a.block_bytes >> 4) & 0x3fffu)) << 16;
      const uint32_t b_kstep = (a.dir == SPMMDB_DIR_ROW ? 32u : (uint32_t)KI * rp) >> 4;
      const uint32_t smem_base = smem_u32(smem);
      bool done = false;
      long long dbg_wait = 0, dbg_steps = 0, dbg_wait_acc = 0;
      const long long dbg_t0 = SPMMDB_BSR_DBG(a) ? clock64() : 0;
      while (!done) {
        const long long ta0 = (SPMMDB_BSR_DBG(a) & 32) ? clock64() : 0;
        mbar_wait(&ctl->tmem_empty[acc], acc_phase[acc] ^ 1);
        if (SPMMDB_BSR_DBG(a) & 32) dbg_wait_acc += clock64() - ta0;
        tc_fence_after();
        uint32_t touched = 0;
        while (true) {
          const long long tw0 = (SPMMDB_BSR_DBG(a) & 32) ? clock64() : 0;
          mbar_wait(&ctl->full[stage], phase);
          if (SPMMDB_BSR_DBG(a) & 32) { dbg_wait += clock64() - tw0; ++dbg_steps; }
          tc_fence_after();
          // broadcast from lane 0: tells the compiler the value (and the control flow below) is warp-uniform
          const uint32_t mask = __shfl_sync(0xffffffffu, ctl->meta[stage].mask, 0);
          if (mask & kEndFlag) {
            const int s = __shfl_sync(0xffffffffu, ctl->meta[stage].s, 0);
            if (leader) {
              ctl->acc[acc].s = s;
              ctl->acc[acc].touched = touched;
              mbar_arrive(&ctl->tmem_full[acc]);  // release: publishes acc[] to the epilogue
              umma_commit(&ctl->tmem_full[acc]);  // arrives when every MMA issued so far has completed
              mbar_arrive(&ctl->empty[stage]);
            }
            __syncwarp();
            if (++stage == a.stages) { stage = 0; phase ^= 1; }
            if (s < 0) done = true;
            break;
          }
          const uint32_t sB = smem_base + (uint32_t)stage * a.stage_bytes;
          const uint32_t sA16 = (sB + a.stage_b_bytes) >> 4;
          const uint32_t dcol = tmem_base + (uint32_t)(acc * acc_cols);
          uint32_t m = (SPMMDB_BSR_DBG(a) & 4) ? 0u : mask;
          while (m) {
            const int l0 = __ffs(m) - 1;
            const uint32_t tb = (touched >> l0) & 1u;
            const int ones = __ffs(~(m >> l0)) - 1;                                   // present run length
            const uint32_t diff = (tb ? ~touched : touched) >> l0;                    // touched status changes
            const int same = diff ? __ffs(diff) - 1 : 32;
            const int len = min(min(ones, same), max_run);
            m &= ~((len >= 32 ? 0xffffffffu : ((1u << len) - 1u)) << l0);
            const uint32_t idesc = idesc_base | (((uint32_t)(len * a.bs) >> 3) << 17);
            const uint32_t b_lo0 = ((sA16 + (((uint32_t)l0 * a.block_bytes) >> 4)) & 0x3fffu) | b_lo_fixed;
            // tcgen05.mma has per-thread semantics: exactly one lane may issue it
            if (elect_one()) {
              for (int t = 0; t < a.tiles; ++t) {
                const uint32_t d = dcol + (uint32_t)(t * a.NT + l0 * a.bs);
                const uint32_t a_lo0 = (((sB + (uint32_t)(t * CPT) * a.chunk_bytes) >> 4) & 0x3fffu) | a_lo_fixed;
                for (int ks = 0; ks < ksteps; ++ks) {
                  const uint64_t adesc = ((uint64_t)a_hi << 32) | (a_lo0 + (uint32_t)ks * a_kstep);
                  const uint64_t bdesc = ((uint64_t)b_hi << 32) | (b_lo0 + (uint32_t)ks * b_kstep);
                  umma<BF16>(d, adesc, bdesc, idesc, (tb | (ks > 0)) ? 1u : 0u);
                }
              }
            }
            __syncwarp();
          }
          touched |= mask;
          if (elect_one()) {
            if (SPMMDB_BSR_DBG(a) & 8) mbar_arrive(&ctl->empty[stage]);
            else umma_commit(&ctl->empty[stage]);  // frees the stage when these MMAs have read it
          }
          __syncwarp();
          if (++stage == a.stages) { stage = 0; phase ^= 1; }
        }
        acc_phase[acc] ^= 1;
        acc ^= 1;
      }
      if ((SPMMDB_BSR_DBG(a) & 32) && blockIdx.x == 0 && lane == 0)
        printf("mma: total %lld cyc, steps %lld, wait(full) %lld cyc, wait(tmem_empty) %lld cyc\n", clock64() - dbg_t0,
               dbg_steps, dbg_wait, dbg_wait_acc);
    }
  } else {
    // ===================== epilogue (4 warps = 128 TMEM lanes) =====================
    const int qd = warp & 3;  // TMEM lane quarter this warp may access
    int acc = 0;
    uint32_t full_phase[2] = {0, 0};
    const int64_t m_rows = a.m_valid;
    const int cpb = 32 / a.bs;  // block rows per 32 accumulator columns (0 when bs = 64)
    while (true) {
      mbar_wait(&ctl->tmem_full[acc], full_phase[acc]);
      tc_fence_after();
      const int s = ctl->acc[acc].s;
      const uint32_t touched = ctl->acc[acc].touched;
      if (s < 0) break;
      for (int t = 0; t < a.tiles; ++t) {
        const int f = t * 128 + qd * 32 + lane;
        const bool fok = f < a.n;
        for (int c0 = 0; c0 < a.NT; c0 += 32) {
          const int64_t row0 = (int64_t)s * a.NT + c0;
          if (row0 >= m_rows) break;
          uint32_t v[32];
          // block rows covered by these 32 columns: 2 (bs 16), 1 (bs 32) or half of one (bs 64)
          uint32_t tmask;  // bit j set = column j belongs to a block row that received an MMA
          if (cpb == 2) {
            const uint32_t tt = (touched >> (c0 / 16)) & 3u;
            tmask = ((tt & 1u) ? 0x0000ffffu : 0u) | ((tt & 2u) ? 0xffff0000u : 0u);
          } else if (cpb == 1) {
            tmask = ((touched >> (c0 / 32)) & 1u) ? 0xffffffffu : 0u;
          } else {
            tmask = ((touched >> (c0 / 64)) & 1u) ? 0xffffffffu : 0u;
          }
          if (tmask) tmem_ld32(tmem_base + ((uint32_t)(qd * 32) << 16) + (uint32_t)(acc * acc_cols + t * a.NT + c0), v);
          if (tmask != 0xffffffffu) {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (!((tmask >> j) & 1u)) v[j] = 0u;
          }
          if (SPMMDB_BSR_DBG(a) & 16) continue;
          const int nrow = (int)min((int64_t)32, m_rows - row0);
          if (!fok && a.c_layout == SPMMDB_ROW_MAJOR) continue;  // (the column-major path needs the whole warp)
          float* __restrict__ cp;
          int64_t cs;  // stride between consecutive rows of A in C
          if (a.c_layout != SPMMDB_ROW_MAJOR) {
            // Column-major C (the reference's layout, rocsparse_bsrmm.h:102-108): a thread holds 32 consecutive rows of
            // ONE feature, i.e. 128 contiguous bytes of C -- but the 32 lanes sit ldc floats apart.  Transpose the
            // 32 x 32 tile through shared memory so that a warp store covers 32 consecutive rows of one feature
            // (one full 128-byte line) instead of 32 lines 4 bytes at a time (C4: 5.0 -> see profiles/r2).
            float* tb = tbuf_all + (warp & 3) * (32 * 33);
#pragma unroll
            for (int j = 0; j < 32; ++j) tb[lane * 33 + j] = a.alpha * __uint_as_float(v[j]);
            __syncwarp();
            const int f0 = t * 128 + qd * 32;
            const int nf = min(32, a.n - f0);
            if (lane < nrow) {
              float* __restrict__ cq = a.C + (int64_t)f0 * a.ldc + row0 + lane;
              if (a.beta != 0.f) {
                // all 32 reads of C in flight before the first write (one dependent round trip per feature otherwise:
                // the grouped block GEMM, beta = 1, took 10 ms on C4 instead of 0.9)
                float old[32];
#pragma unroll
                for (int ff = 0; ff < 32; ++ff) old[ff] = ff < nf ? cq[(int64_t)ff * a.ldc] : 0.f;
#pragma unroll
                for (int ff = 0; ff < 32; ++ff)
                  if (ff < nf) cq[(int64_t)ff * a.ldc] = fmaf(a.beta, old[ff], tb[ff * 33 + lane]);
              } else {
#pragma unroll 8
                for (int ff = 0; ff < nf; ++ff) cq[(int64_t)ff * a.ldc] = tb[ff * 33 + lane];
              }
            }
            __syncwarp();
            continue;
          }
          cp = a.C + row0 * a.ldc + f;
          cs = a.ldc;
          if (nrow == 32 && a.beta == 0.f) {
#pragma unroll
            for (int j = 0; j < 32; ++j) cp[(int64_t)j * cs] = a.alpha * __uint_as_float(v[j]);
          } else {
            // beta != 0 (the hybrid split's C += A_bsr * B) or a ragged last tile: every read of C is issued before
            // the first write, or each row would cost a dependent round trip
            float old[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) old[j] = (a.beta != 0.f && j < nrow) ? cp[(int64_t)j * cs] : 0.f;
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (j < nrow) cp[(int64_t)j * cs] = fmaf(a.beta, old[j], a.alpha * __uint_as_float(v[j]));
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&ctl->tmem_empty[acc]);
      full_phase[acc] ^= 1;
      acc ^= 1;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == P + 1) tmem_dealloc(tmem_base, a.tmem_cols);
}

// ---------------------------------------------------------------------------------------
// Pre-passes (only when the caller's operands cannot be read by TMA as they are)
// ---------------------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ T cvt_out(float v);
template <> __device__ __forceinline__ float cvt_out<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 cvt_out<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// out[r * ldo + c] = B(r, c) for r < rows, c < cols; B row-major (ldb >= cols) or column-major (ldb >= rows)
template <typename T>
__global__ void __launch_bounds__(256) prep_b_kernel(const float* __restrict__ B, int64_t ldb, int col_major, T* out,
                                                     int64_t ldo, int rows, int cols) {
  __shared__ float tile[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t r0 = (int64_t)blockIdx.x * 32, c0 = (int64_t)blockIdx.y * 32;
  if (col_major) {
    for (int i = ty; i < 32; i += 8) {
      const int64_t c = c0 + i, r = r0 + tx;
      tile[i][tx] = (c < cols && r < rows) ? B[c * ldb + r] : 0.f;
    }
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
      const int64_t r = r0 + i, c = c0 + tx;
      if (r < rows && c < cols) out[r * ldo + c] = cvt_out<T>(tile[tx][i]);
    }
  } else {
    for (int i = ty; i < 32; i += 8) {
      const int64_t r = r0 + i, c = c0 + tx;
      if (r < rows && c < cols) out[r * ldo + c] = cvt_out<T>(B[r * ldb + c]);
    }
  }
}

__global__ void __launch_bounds__(256) cvt_bf16_kernel(const float4* __restrict__ in, uint2* __restrict__ out, int64_t n4) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 v = __ldcs(in + i);
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
    uint2 o;
    o.x = *reinterpret_cast<uint32_t*>(&lo);
    o.y = *reinterpret_cast<uint32_t*>(&hi);
    out[i] = o;
  }
}

// out block [r][c] = in block [c][r]: column-major blocks -> row-major blocks (tf32 path only: 32-bit
// MN-major operands need 128-byte rows, which a 16-wide block does not have)
__global__ void __launch_bounds__(256) block_transpose_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                              int bs, int64_t nnzb) {
  const int64_t bs2 = (int64_t)bs * bs;
  const int64_t total = nnzb * bs2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t p = i / bs2;
    const int e = (int)(i - p * bs2);
    const int r = e / bs, c = e - r * bs;
    out[i] = __ldg(in + p * bs2 + (int64_t)c * bs + r);
  }
}

// ---------------------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess) {
      cudaGetLastError();
      p = nullptr;
    }
    return (EncodeTiledFn)p;
  }();
  return fn;
}

bool make_map_2d(CUtensorMap* map, bool bf16, const void* base, uint64_t inner, uint64_t outer, uint64_t row_bytes,
                 uint32_t box_inner, uint32_t box_outer, uint32_t swizzle_bytes) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) return false;
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {row_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMapSwizzle sw = swizzle_bytes == 129 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B
                          : swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                          : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                                : CU_TENSOR_MAP_SWIZZLE_32B;
  CUresult r = fn(map, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, dims,
                  strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// B slab as ONE box: view row-major B[krows][ld] as (fc features, krows rows, n / fc chunks) with the chunk
// dimension 128 bytes apart; box = {fc, bs, chunks} lands chunk-major in shared memory.
bool make_map_b3d(CUtensorMap* map, bool bf16, const void* base, uint64_t fc, uint64_t krows, uint64_t chunks,
                  uint64_t row_bytes, uint32_t bs) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) return false;
  cuuint64_t dims[3] = {fc, krows, chunks};
  cuuint64_t strides[2] = {row_bytes, 128};
  cuuint32_t box[3] = {(cuuint32_t)fc, bs, (cuuint32_t)chunks};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(map, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)base, dims,
                  strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  bf16 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

template <bool BF16, int P>
int launch_tc(const TcArgs& a, const CUtensorMap& mapB, const CUtensorMap& mapA, int grid, size_t smem_bytes,
              cudaStream_t st) {
  SPMMDB_CUDA_OK(cudaFuncSetAttribute(bsrmm_tc_kernel<BF16, P>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)smem_bytes));
  bsrmm_tc_kernel<BF16, P><<<grid, (P + 5) * 32, smem_bytes, st>>>(a, mapB, mapA);
  SPMMDB_CUDA_OK(cudaGetLastError());
  return SPMMDB_OK;
}

}  // namespace

// Tensor map of a row-major matrix whose box is ONE whole row (cols elements, no swizzle): what the CSR TMA-gather
// kernel's tile::gather4 copies address, four rows per instruction (csrmm.cu).
bool make_map_rows(CUtensorMap* map, bool bf16, const void* base, uint64_t cols, uint64_t rows, uint64_t row_bytes) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr || cols == 0 || cols > 256 || ((uintptr_t)base % 16) != 0 || row_bytes % 16 != 0) return false;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {row_bytes};
  cuuint32_t box[2] = {(cuuint32_t)cols, 1};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, dims,
                  strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// Shapes the tensor-core kernel takes; everything else runs the fp32 CUDA-core kernel.
bool bsrmm_tc_supported(int trans_b, int n, int bs, int64_t ldb, const float* B, const float* val, int c_layout,
                        int64_t ldc, const float* C) {
  (void)trans_b; (void)ldb; (void)B; (void)c_layout; (void)ldc; (void)C;
  if (bs != 16 && bs != 32 && bs != 64) return false;
  if (n < 1) return false;
  if (((uintptr_t)val) % 16 != 0) return false;
  return encode_fn() != nullptr;
}

int bsrmm_tc(int dir, int trans_b, int mb, int n, int kb, int64_t nnzb, int bs, float alpha, const float* val,
             const int* rowptr, const int* colind, const float* B, int64_t ldb, float beta, float* C, int64_t ldc,
             int c_layout, int math, cudaStream_t st) {
  // n > 512 features: column slabs of 512 (TMEM holds 512 accumulator columns)
  if (n > 512) {
    for (int c0 = 0; c0 < n; c0 += 512) {
      const int nn = n - c0 < 512 ? n - c0 : 512;
      const float* Bs = trans_b ? B + c0 : B + (int64_t)c0 * ldb;
      float* Cs = c_layout == SPMMDB_ROW_MAJOR ? C + c0 : C + (int64_t)c0 * ldc;
      int rc = bsrmm_tc(dir, trans_b, mb, nn, kb, nnzb, bs, alpha, val, rowptr, colind, Bs, ldb, beta, Cs, ldc,
                        c_layout, math, st);
      if (rc != SPMMDB_OK) return rc;
    }
    return SPMMDB_OK;
  }
  const bool packed = (math == SPMMDB_MATH_BF16_PACKED);  // `val` already points at bf16 blocks
  bool bf16 = (math == SPMMDB_MATH_BF16) || packed;
  if (!bf16 && bs == 64) bf16 = false;  // tf32 with 256-byte block rows: see below
  const int esize = bf16 ? 2 : 4;
  if (bs * esize > 128) {
    // a block row wider than the 128-byte swizzle span (tf32, bs = 64) is not laid out by one TMA box
    return SPMMDB_ERR_NOT_SUPPORTED;
  }
  const RowClip clip = current_row_clip();
  // rows of B the TMA maps cover: boxes that reach past them are zero-filled by the hardware
  const int64_t krows = (int64_t)kb * bs < clip.b_rows ? (int64_t)kb * bs : clip.b_rows;
  const int fc = 128 / esize;
  const int tiles = (n + 127) / 128;
  const int nchunks = (n + fc - 1) / fc;

  // ---- operand B: TMA needs row-major [krows][ld] with a 16-byte aligned base and row pitch ----
  const void* Bt = B;
  int64_t ldbt = ldb;
  const bool b_direct = !bf16 && trans_b == 1 && (ldb % 4 == 0) && ((uintptr_t)B % 16 == 0);
  if (!b_direct) {
    const int64_t ldo = ((int64_t)n + 7) & ~(int64_t)7;
    void* sb = scratch(kScratchBsrB, (size_t)krows * ldo * esize, st);
    if (sb == nullptr) return SPMMDB_ERR_ALLOC;
    dim3 grid((unsigned)((krows + 31) / 32), (unsigned)((n + 31) / 32));
    if (bf16)
      prep_b_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(B, ldb, trans_b == 0, (__nv_bfloat16*)sb, ldo, (int)krows, n);
    else
      prep_b_kernel<float><<<grid, 256, 0, st>>>(B, ldb, trans_b == 0, (float*)sb, ldo, (int)krows, n);
    SPMMDB_CUDA_OK(cudaGetLastError());
    Bt = sb;
    ldbt = ldo;
  }
  // ---- operand A: fp32 blocks are read as tf32 in place; bf16 needs a converted copy ----
  const void* At = val;
  if (bf16 && !packed) {
    const int64_t elems = nnzb * bs * bs;
    void* sa = scratch(kScratchBsrA, (size_t)elems * 2, st);
    if (sa == nullptr) return SPMMDB_ERR_ALLOC;
    const int64_t n4 = elems / 4;  // bs*bs is a multiple of 4
    int blocks = (int)((n4 + 255) / 256 < 148 * 16 ? (n4 + 255) / 256 : 148 * 16);
    cvt_bf16_kernel<<<blocks, 256, 0, st>>>((const float4*)val, (uint2*)sa, n4);
    SPMMDB_CUDA_OK(cudaGetLastError());
    At = sa;
  }

  if (!bf16 && dir == SPMMDB_DIR_COLUMN) {
    const int64_t elems = nnzb * bs * bs;
    void* sa = scratch(kScratchBsrA, (size_t)elems * 4, st);
    if (sa == nullptr) return SPMMDB_ERR_ALLOC;
    int blocks = (int)((elems + 255) / 256 < 148 * 16 ? (elems + 255) / 256 : 148 * 16);
    block_transpose_kernel<<<blocks, 256, 0, st>>>(val, (float*)sa, bs, nnzb);
    SPMMDB_CUDA_OK(cudaGetLastError());
    At = sa;
    dir = SPMMDB_DIR_ROW;
  }

  TcArgs a;
  a.mb = mb; a.n = n; a.bs = bs; a.tiles = tiles; a.nchunks = nchunks;
  const int nt_env = SPMMDB_TUNE("BSR_NT", 0);
  // Two CTAs per SM (256 TMEM columns and half the shared memory each) hide the per-step latency of
  // the single MMA-issuing warp; they need tiles * NT <= 128 per accumulator buffer.
  const int cps_env = SPMMDB_TUNE("BSR_CTAS_PER_SM", 0);
  int cps = (cps_env == 1 || cps_env == 2) ? cps_env : (tiles == 1 ? 2 : 1);
  if (tiles > 2 || (128 / tiles) < bs) cps = 1;
  const int acc_max = cps == 2 ? 128 : 256;  // accumulator columns per buffer (two buffers)
  int NT = 256;
  if (nt_env == 64 || nt_env == 128 || nt_env == 256) NT = nt_env;
  while (NT * tiles > acc_max) NT >>= 1;
  if (NT < bs) return SPMMDB_ERR_NOT_SUPPORTED;  // bs = 64 with > 4 tiles cannot happen (n <= 512)
  a.tmem_cols = cps == 2 ? 256 : 512;
  a.NT = NT;
  a.G = NT / bs;
#ifdef SPMMDB_BSR_DEBUG
  a.dbg = SPMMDB_TUNE("BSR_DEBUG_SKIP", 0);
#else
  a.dbg = 0;
#endif
  a.dir = dir; a.c_layout = c_layout;
  a.num_super = (mb + a.G - 1) / a.G;
  a.m_valid = (int64_t)mb * bs < clip.c_rows ? (int64_t)mb * bs : clip.c_rows;
  a.rowptr = rowptr; a.colind = colind; a.C = C; a.ldc = ldc; a.alpha = alpha; a.beta = beta;
  a.chunk_bytes = (uint32_t)bs * 128u;
  a.block_bytes = (uint32_t)(bs * bs * esize);
  const int cpt = 128 / fc;
  a.stage_b_bytes = (uint32_t)(tiles * cpt) * a.chunk_bytes;
  a.stage_bytes = (a.stage_b_bytes + (uint32_t)(NT * bs * esize) + 1023u) & ~1023u;
  // control block (+ the four 32 x 33 float transpose tiles of the column-major-C epilogue)
  const size_t ctl_bytes = sizeof(TcSmemCtl) + 64 + (c_layout == SPMMDB_ROW_MAJOR ? 0 : 4 * 32 * 33 * sizeof(float) + 16);
  const size_t smem_max = cps == 2 ? 113 * 1024 : 227 * 1024;
  int stages = (int)((smem_max - 1024 - ctl_bytes) / a.stage_bytes);
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages < 2) return SPMMDB_ERR_NOT_SUPPORTED;
  a.stages = stages;
  const size_t smem_bytes = 1024 + (size_t)stages * a.stage_bytes + ctl_bytes;

  // scratch: [0] super-row counter | plan_count[num_super] | plan[nnzb] (8-byte records)
  const size_t off_count = 256;
  const size_t off_plan = (off_count + sizeof(int) * (size_t)a.num_super + 255) & ~(size_t)255;
  char* sc = (char*)scratch(kScratchBsr, off_plan + sizeof(uint2) * (size_t)nnzb, st);
  if (sc == nullptr) return SPMMDB_ERR_ALLOC;
  SPMMDB_CUDA_OK(cudaMemsetAsync(sc, 0, 4, st));
  a.counter = (int*)sc;
  a.plan_count = (const int*)(sc + off_count);
  a.plan = (const uint2*)(sc + off_plan);
  {
    const int warps = a.num_super;
    const int blocks = (warps + 7) / 8;
    bsr_plan_kernel<<<blocks, 256, 0, st>>>(mb, a.G, a.num_super, rowptr, colind, (uint2*)(sc + off_plan),
                                            (int*)(sc + off_count));
    SPMMDB_CUDA_OK(cudaGetLastError());
  }

  CUtensorMap mapB, mapA;
  a.b_one = (n % fc == 0) ? 1 : 0;
  const int b3d_env = SPMMDB_TUNE("BSR_B3D", 1);
  if (!b3d_env) a.b_one = 0;
  if (a.b_one && !make_map_b3d(&mapB, bf16, Bt, (uint64_t)fc, (uint64_t)krows, (uint64_t)nchunks,
                               (uint64_t)ldbt * esize, (uint32_t)bs))
    a.b_one = 0;  // driver refused the overlapping view: one box per 128-byte chunk instead
  if (!a.b_one && !make_map_2d(&mapB, bf16, Bt, (uint64_t)n, (uint64_t)krows, (uint64_t)ldbt * esize, (uint32_t)fc,
                          (uint32_t)bs, bf16 ? 128 : 129))
    return SPMMDB_ERR_INVALID_VALUE;
  if (!make_map_2d(&mapA, bf16, At, (uint64_t)bs, (uint64_t)nnzb * bs, (uint64_t)bs * esize, (uint32_t)bs, (uint32_t)bs,
                   (uint32_t)(bs * esize)))
    return SPMMDB_ERR_INVALID_VALUE;

  const int nsm = device_num_sms() > 0 ? device_num_sms() : 148;
  const int grid = a.num_super < nsm * cps ? a.num_super : nsm * cps;
  const int p_env = SPMMDB_TUNE("BSR_PRODUCERS", 0);
  // A producer refills stage (its step) % stages and waits on that stage's `empty` parity: with more producers than
  // stages two producers are a whole ring apart on the SAME barrier and the later one can pass a phase early (ABA) --
  // the hang seen with 4 producers on 3 stages in r1l.  So P <= stages, always.
  int P = p_env >= 1 && p_env <= 3 ? p_env : 3;
  if (P > stages) P = stages;
  if (bf16) {
    if (P == 1) return launch_tc<true, 1>(a, mapB, mapA, grid, smem_bytes, st);
    if (P == 2) return launch_tc<true, 2>(a, mapB, mapA, grid, smem_bytes, st);
    return launch_tc<true, 3>(a, mapB, mapA, grid, smem_bytes, st);
  }
  if (P == 1) return launch_tc<false, 1>(a, mapB, mapA, grid, smem_bytes, st);
  if (P == 2) return launch_tc<false, 2>(a, mapB, mapA, grid, smem_bytes, st);
  return launch_tc<false, 3>(a, mapB, mapA, grid, smem_bytes, st);
}

// column-major blocks <-> row-major blocks, once per matrix (spmmdb_bsr_transpose_blocks_f32)
int bsr_transpose_blocks(int64_t nnzb, int bs, const float* in, float* out, cudaStream_t st) {
  const int64_t elems = nnzb * bs * bs;
  if (elems > 0) {
    int blocks = (int)((elems + 255) / 256 < 148 * 16 ? (elems + 255) / 256 : 148 * 16);
    block_transpose_kernel<<<blocks, 256, 0, st>>>(in, out, bs, nnzb);
    SPMMDB_CUDA_OK(cudaGetLastError());
  }
  return SPMMDB_OK;
}

int bsr_pack_bf16(int64_t elems, const float* in, void* out, cudaStream_t st) {
  const int64_t n4 = elems / 4;
  if (n4 > 0) {
    int blocks = (int)((n4 + 255) / 256 < 148 * 16 ? (n4 + 255) / 256 : 148 * 16);
    cvt_bf16_kernel<<<blocks, 256, 0, st>>>((const float4*)in, (uint2*)out, n4);
    SPMMDB_CUDA_OK(cudaGetLastError());
  }
  return SPMMDB_OK;
}

}  // namespace spmmdb
