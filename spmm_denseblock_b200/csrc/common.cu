// common.cu -- error capture, device queries and grow-only scratch for libspmmdb.
#include <mutex>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace spmmdb {

static thread_local char g_last_error[512] = "";

void set_last_error(const char* what, cudaError_t err, const char* file, int line) {
  snprintf(g_last_error, sizeof(g_last_error), "%s: %s (%s:%d)", what, cudaGetErrorString(err), file, line);
}

namespace {
constexpr int kMaxDevices = 16;
struct Scratch { void* ptr = nullptr; size_t bytes = 0; };
Scratch g_scratch[kMaxDevices][kScratchSlots];
int g_sm[kMaxDevices];
int g_nsm[kMaxDevices];
bool g_dev_init[kMaxDevices];
std::mutex g_mu;

int current_device() {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return -1; }
  return (dev >= 0 && dev < kMaxDevices) ? dev : -1;
}

bool init_device(int dev) {
  if (g_dev_init[dev]) return true;
  cudaDeviceProp p;
  if (cudaGetDeviceProperties(&p, dev) != cudaSuccess) { cudaGetLastError(); return false; }
  g_sm[dev] = p.major * 10 + p.minor;
  g_nsm[dev] = p.multiProcessorCount;
  g_dev_init[dev] = true;
  return true;
}
}  // namespace

int device_sm() {
  std::lock_guard<std::mutex> lk(g_mu);
  int dev = current_device();
  if (dev < 0 || !init_device(dev)) return -1;
  return g_sm[dev];
}

int device_num_sms() {
  std::lock_guard<std::mutex> lk(g_mu);
  int dev = current_device();
  if (dev < 0 || !init_device(dev)) return -1;
  return g_nsm[dev];
}

void* scratch(int slot, size_t bytes, cudaStream_t stream) {
  std::lock_guard<std::mutex> lk(g_mu);
  int dev = current_device();
  if (dev < 0) return nullptr;
  Scratch& s = g_scratch[dev][slot];
  if (s.bytes >= bytes && s.ptr != nullptr) return s.ptr;
  // Growth: work already queued on `stream` may still use the old buffer.
  if (s.ptr != nullptr) {
    cudaStreamSynchronize(stream);
    cudaFree(s.ptr);
    s.ptr = nullptr;
    s.bytes = 0;
  }
  size_t want = bytes + bytes / 4 + 256;
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, want);
  if (e != cudaSuccess) {
    set_last_error("cudaMalloc(scratch)", e, __FILE__, __LINE__);
    cudaGetLastError();
    return nullptr;
  }
  e = cudaMemsetAsync(p, 0, want, stream);
  if (e != cudaSuccess) {
    set_last_error("cudaMemsetAsync(scratch)", e, __FILE__, __LINE__);
    cudaFree(p);
    return nullptr;
  }
  s.ptr = p;
  s.bytes = want;
  return p;
}

}  // namespace spmmdb

extern "C" const char* spmmdb_version(void) { return "spmmdb 0.1 (sm_100a)"; }
extern "C" const char* spmmdb_last_error(void) { return spmmdb::g_last_error; }
extern "C" int spmmdb_device_sm(void) { return spmmdb::device_sm(); }

namespace {
int g_bsrmm_math = -1;  // -1: not read from the environment yet
int env_bsrmm_math() {
  const char* e = getenv("SPMMDB_BSRMM_MATH");
  if (e == nullptr) return SPMMDB_MATH_FP32;
  if (strcmp(e, "tf32") == 0 || strcmp(e, "1") == 0) return SPMMDB_MATH_TF32;
  if (strcmp(e, "bf16") == 0 || strcmp(e, "2") == 0) return SPMMDB_MATH_BF16;
  return SPMMDB_MATH_FP32;
}
}  // namespace
extern "C" int spmmdb_set_bsrmm_math(int math) {
  if (math != SPMMDB_MATH_FP32 && math != SPMMDB_MATH_TF32 && math != SPMMDB_MATH_BF16)
    return SPMMDB_ERR_INVALID_VALUE;
  g_bsrmm_math = math;
  return SPMMDB_OK;
}
extern "C" int spmmdb_get_bsrmm_math(void) {
  if (g_bsrmm_math < 0) g_bsrmm_math = env_bsrmm_math();
  return g_bsrmm_math;
}
