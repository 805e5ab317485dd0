// common.cuh -- shared helpers of libspmmdb (error capture, per-device scratch).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "spmmdb.h"

namespace spmmdb {

// Last CUDA error text, returned by spmmdb_last_error().
void set_last_error(const char* what, cudaError_t err, const char* file, int line);

#define SPMMDB_CUDA_OK(expr)                                              \
  do {                                                                    \
    cudaError_t _e = (expr);                                              \
    if (_e != cudaSuccess) {                                              \
      ::spmmdb::set_last_error(#expr, _e, __FILE__, __LINE__);            \
      return SPMMDB_ERR_CUDA;                                             \
    }                                                                     \
  } while (0)

// Grow-only scratch buffer per (device, stream, slot).  The library owns it; it is
// never visible to the caller.  Allocation happens on first use / growth only
// (cudaMalloc synchronises, steady-state calls stay asynchronous).  Newly
// allocated memory is zero-filled.  Calls on one stream are ordered by the stream and
// share a buffer; calls on different streams (or devices) never share one, so they
// may be in flight together.  At most kMaxScratchStreams streams per device hold
// scratch at a time: the least recently used one is synchronised and released.
enum ScratchSlot { kScratchCsrLong = 0, kScratchCsrPartial, kScratchCsrDone, kScratchCsrEntry, kScratchCsrEpoch, kScratchLayoutB, kScratchLayoutC,
                   kScratchConvert, kScratchBsr, kScratchBsrA, kScratchBsrB, kScratchHostA, kScratchHostB, kScratchHostC,
                   kScratchSlots };
void* scratch(int slot, size_t bytes, cudaStream_t stream);

// Tuning knobs: value of spmmdb_set_tuning(name, .) if set, else the environment variable
// SPMMDB_<name>, else `dflt`.  SPMMDB_TUNE caches the lookup per call site and re-reads it
// only after a spmmdb_set_tuning call (a generation counter), so the steady state is one
// atomic load.
struct TuneSlot { unsigned long long packed = ~0ull; };  // (generation << 32 | value), read / written atomically
int tune_lookup(TuneSlot& slot, const char* name, int dflt);
#define SPMMDB_TUNE(name, dflt) ([&]() -> int { static ::spmmdb::TuneSlot _s; return ::spmmdb::tune_lookup(_s, name, dflt); }())

// Row-segment length override (spmmdb_set_csrmm_segment, process-wide) or the calling thread's scoped one.
int current_segment_override();
struct ScopedSegment {  // pins the segment length for the calls this thread makes while it lives
  explicit ScopedSegment(int seg);
  ~ScopedSegment();
  int before_;
};

// Row limits for the BSR kernels, set by library entry points whose matrix size is not a multiple of the block size
// (spmmdb_hybrid_spmm_f32): rows of B at or beyond b_rows read as zero, rows of C at or beyond c_rows are not written.
// The reference's contract (B and C padded to whole blocks, run_bsrmm.cu:86-94) is the default: no limit.
struct RowClip { int64_t b_rows, c_rows; };
RowClip current_row_clip();
struct ScopedRowClip {
  ScopedRowClip(int64_t b_rows, int64_t c_rows);
  ~ScopedRowClip();
  RowClip before_;
};

int next_epoch();      // process-wide call counter (atomic), never 0

int device_sm();       // compute capability major*10+minor of the current device, cached
int device_num_sms();  // SM count of the current device, cached

__host__ __device__ inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

}  // namespace spmmdb
