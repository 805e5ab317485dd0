// host.cu -- host-buffer entry points: what a caller that keeps its matrices in
// host memory (the reference drivers' state before their cudaMemcpy block,
// run_csrmm.cu:94-116) calls.  Copies in, runs the device path, copies out.
#include <stdlib.h>

#include "common.cuh"

using namespace spmmdb;

// Column widths of the feature slabs of one call (at most 16, each a multiple of 16 columns); returns their number.
static int plan_slabs(int K, int slabs_knob, int tail, int* widths) {
  int main_w = 64;
  if (slabs_knob > 0) main_w = (K % slabs_knob == 0 && (K / slabs_knob) % 16 == 0) ? K / slabs_knob : K;
  if (K % 16 != 0 || main_w >= K) { widths[0] = K; return 1; }
  if (tail < 16 || tail % 16 != 0) tail = main_w;
  int n = 0, left = K;
  while (left > main_w && n < 12) { widths[n++] = main_w; left -= main_w; }               // full-width slabs
  while (n < 15 && left >= 2 * tail && (left / 2) % 16 == 0) { widths[n++] = left / 2; left -= left / 2; }  // halves
  widths[n++] = left;
  return n;
}

extern "C" int spmmdb_csrmm_host_f32(int m, int k, int K, int64_t nnz, const int* h_rowptr, const int* h_colind,
                                     const float* h_val, const float* h_B, float* h_C, float alpha, float beta,
                                     void* stream) {
  if (m < 0 || k < 0 || K < 0 || nnz < 0) return SPMMDB_ERR_INVALID_VALUE;
  if (m == 0 || K == 0) return SPMMDB_OK;
  if (h_rowptr == nullptr || h_B == nullptr || h_C == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  if (nnz > 0 && (h_colind == nullptr || h_val == nullptr)) return SPMMDB_ERR_INVALID_VALUE;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t a_bytes = sizeof(int) * ((size_t)m + 1) + (sizeof(int) + sizeof(float)) * (size_t)nnz + 64;
  const size_t b_bytes = sizeof(float) * (size_t)k * K;
  const size_t c_bytes = sizeof(float) * (size_t)m * K;
  char* dA = (char*)scratch(kScratchHostA, a_bytes, st);
  float* dB = (float*)scratch(kScratchHostB, b_bytes, st);
  float* dC = (float*)scratch(kScratchHostC, c_bytes, st);
  if (dA == nullptr || dB == nullptr || dC == nullptr) return SPMMDB_ERR_ALLOC;
  int* d_rowptr = (int*)dA;
  const size_t off_ci = ((sizeof(int) * ((size_t)m + 1) + 15) / 16) * 16;
  int* d_colind = (int*)(dA + off_ci);
  float* d_val = (float*)(dA + off_ci + ((sizeof(int) * (size_t)nnz + 15) / 16) * 16);
  SPMMDB_CUDA_OK(cudaMemcpyAsync(d_rowptr, h_rowptr, sizeof(int) * ((size_t)m + 1), cudaMemcpyHostToDevice, st));
  if (nnz > 0) {
    SPMMDB_CUDA_OK(cudaMemcpyAsync(d_colind, h_colind, sizeof(int) * (size_t)nnz, cudaMemcpyHostToDevice, st));
    SPMMDB_CUDA_OK(cudaMemcpyAsync(d_val, h_val, sizeof(float) * (size_t)nnz, cudaMemcpyHostToDevice, st));
  }
  // Feature slabs: columns [c0, c0 + Ks) of C need only the same columns of B, so the host<->device copies of one
  // slab overlap the kernel and the opposite-direction copy of its neighbours (PCIe is full duplex): the call costs
  // about H2D(A) + H2D(B) + D2H(one slab) instead of H2D(A) + H2D(B) + kernel + D2H(C).
  // The call ends with the D2H copy of the LAST slab alone on the wire.  Slabs are 64 columns wide (HOST_SLABS > 0:
  // K / HOST_SLABS columns): narrower last slabs (HOST_TAIL = 32 / 16: the rest of K halved down to that width) shorten
  // that tail but lose more in the strided copies than they gain -- rows under 256 bytes run the DMA engines well
  // below the PCIe rate (products-shaped graph, K = 256: 74.1 ms with 64 64 64 64, 79.1 ms with 64 64 64 32 32,
  // 89.8 ms with 64 64 64 32 16 16; profiles/r2q_host_slabs.jsonl).
  int widths[16];
  const int slabs = plan_slabs(K, SPMMDB_TUNE("HOST_SLABS", 0), SPMMDB_TUNE("HOST_TAIL", 64), widths);
  if (slabs <= 1) {
    SPMMDB_CUDA_OK(cudaMemcpyAsync(dB, h_B, b_bytes, cudaMemcpyHostToDevice, st));
    if (beta != 0.f) SPMMDB_CUDA_OK(cudaMemcpyAsync(dC, h_C, c_bytes, cudaMemcpyHostToDevice, st));
    int rc = spmmdb_csrmm_f32(m, k, K, nnz, d_rowptr, d_colind, d_val, dB, K, SPMMDB_ROW_MAJOR, dC, K,
                              SPMMDB_ROW_MAJOR, alpha, beta, st);
    if (rc != SPMMDB_OK) return rc;
    SPMMDB_CUDA_OK(cudaMemcpyAsync(h_C, dC, c_bytes, cudaMemcpyDeviceToHost, st));
    SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
    return SPMMDB_OK;
  }
  // `st` carries the H2D copies (A is already queued on it), a second stream the kernels, a third the D2H copies
  static thread_local cudaStream_t s_run = nullptr, s_out = nullptr;
  static thread_local cudaEvent_t ev_in[16], ev_run[16];
  static thread_local bool ev_made = false;
  if (s_run == nullptr) {
    SPMMDB_CUDA_OK(cudaStreamCreateWithFlags(&s_run, cudaStreamNonBlocking));
    SPMMDB_CUDA_OK(cudaStreamCreateWithFlags(&s_out, cudaStreamNonBlocking));
  }
  if (!ev_made) {
    for (int i = 0; i < 16; ++i) {
      SPMMDB_CUDA_OK(cudaEventCreateWithFlags(&ev_in[i], cudaEventDisableTiming));
      SPMMDB_CUDA_OK(cudaEventCreateWithFlags(&ev_run[i], cudaEventDisableTiming));
    }
    ev_made = true;
  }
  const size_t pitch = sizeof(float) * (size_t)K;
  // the kernels of all slabs share the whole call's segment length (same bits as the one-shot call)
  const int seg_now = current_segment_override();
  ScopedSegment pin(seg_now ? seg_now : spmmdb_csrmm_segment_for(m, nnz));
  int rc = SPMMDB_OK;
  size_t off = 0;
  for (int sl = 0; sl < slabs && rc == SPMMDB_OK; off += widths[sl], ++sl) {
    const int Ks = widths[sl];
    const size_t width = sizeof(float) * (size_t)Ks;
    cudaError_t e = cudaMemcpy2DAsync(dB + off, pitch, h_B + off, pitch, width, (size_t)k, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess && beta != 0.f)
      e = cudaMemcpy2DAsync(dC + off, pitch, h_C + off, pitch, width, (size_t)m, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaEventRecord(ev_in[sl], st);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(s_run, ev_in[sl], 0);
    if (e != cudaSuccess) { set_last_error("host slab copy", e, __FILE__, __LINE__); rc = SPMMDB_ERR_CUDA; break; }
    rc = spmmdb_csrmm_f32(m, k, Ks, nnz, d_rowptr, d_colind, d_val, dB + off, K, SPMMDB_ROW_MAJOR, dC + off, K,
                          SPMMDB_ROW_MAJOR, alpha, beta, s_run);
    if (rc != SPMMDB_OK) break;
    e = cudaEventRecord(ev_run[sl], s_run);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(s_out, ev_run[sl], 0);
    if (e == cudaSuccess)
      e = cudaMemcpy2DAsync(h_C + off, pitch, dC + off, pitch, width, (size_t)m, cudaMemcpyDeviceToHost, s_out);
    if (e != cudaSuccess) { set_last_error("host slab copy", e, __FILE__, __LINE__); rc = SPMMDB_ERR_CUDA; }
  }
  cudaStreamSynchronize(st);
  cudaStreamSynchronize(s_run);
  cudaError_t e = cudaStreamSynchronize(s_out);
  if (rc == SPMMDB_OK && e != cudaSuccess) { set_last_error("host slab sync", e, __FILE__, __LINE__); rc = SPMMDB_ERR_CUDA; }
  return rc;
}

extern "C" int spmmdb_shard_plan(const int* h_rowptr, int m, int parts, int align, int* bounds) {
  if (h_rowptr == nullptr || bounds == nullptr || m < 0 || parts < 1 || align < 1) return SPMMDB_ERR_INVALID_VALUE;
  const int64_t nnz = h_rowptr[m];
  bounds[0] = 0;
  for (int g = 1; g < parts; ++g) {
    const int64_t target = (int64_t)g * nnz / parts;
    // first row with rowptr[row] >= target (lower bound)
    int lo = 0, hi = m;
    while (lo < hi) {
      const int mid = lo + (hi - lo) / 2;
      if (h_rowptr[mid] >= target) hi = mid; else lo = mid + 1;
    }
    int b = (lo / align) * align;
    if (b < bounds[g - 1]) b = bounds[g - 1];
    bounds[g] = b;
  }
  bounds[parts] = m;
  return SPMMDB_OK;
}
