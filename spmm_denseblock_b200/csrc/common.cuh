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

// Grow-only scratch buffer per (device, slot).  The library owns it; it is
// never visible to the caller.  Allocation happens on first use / growth only
// (cudaMalloc synchronises, steady-state calls stay asynchronous).  Newly
// allocated memory is zero-filled.
enum ScratchSlot { kScratchCsrLong = 0, kScratchCsrPartial, kScratchCsrDone, kScratchCsrEntry, kScratchLayoutB, kScratchLayoutC,
                   kScratchConvert, kScratchBsr, kScratchBsrA, kScratchBsrB, kScratchHostA, kScratchHostB, kScratchHostC,
                   kScratchSlots };
void* scratch(int slot, size_t bytes, cudaStream_t stream);

int device_sm();       // compute capability major*10+minor of the current device, cached
int device_num_sms();  // SM count of the current device, cached

__host__ __device__ inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

}  // namespace spmmdb
