// common.cu -- error capture, device queries, tuning knobs and grow-only scratch for libspmmdb.
#include <limits.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <map>
#include <mutex>
#include <string>

#include "common.cuh"

namespace spmmdb {

static thread_local char g_last_error[512] = "";

void set_last_error(const char* what, cudaError_t err, const char* file, int line) {
  snprintf(g_last_error, sizeof(g_last_error), "%s: %s (%s:%d)", what, cudaGetErrorString(err), file, line);
}

namespace {
constexpr int kMaxDevices = 16;
constexpr int kMaxScratchStreams = 8;
struct Scratch { void* ptr = nullptr; size_t bytes = 0; };
struct StreamScratch {
  cudaStream_t stream = nullptr;
  bool used = false;
  uint64_t last_use = 0;
  Scratch slot[kScratchSlots];
};
StreamScratch g_scratch[kMaxDevices][kMaxScratchStreams];
uint64_t g_use_clock = 0;
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

// The scratch set of (dev, stream); takes over the least recently used set when all are taken.
StreamScratch* stream_scratch(int dev, cudaStream_t stream) {
  StreamScratch* free_set = nullptr;
  StreamScratch* lru = nullptr;
  for (auto& s : g_scratch[dev]) {
    if (s.used && s.stream == stream) { s.last_use = ++g_use_clock; return &s; }
    if (!s.used) { if (free_set == nullptr) free_set = &s; }
    else if (lru == nullptr || s.last_use < lru->last_use) lru = &s;
  }
  StreamScratch* s = free_set;
  if (s == nullptr) {
    s = lru;
    // work queued on the evicted stream may still use its buffers
    if (cudaStreamSynchronize(s->stream) != cudaSuccess) { cudaGetLastError(); cudaDeviceSynchronize(); }
    for (auto& b : s->slot) {
      if (b.ptr != nullptr) cudaFree(b.ptr);
      b = Scratch{};
    }
  }
  s->used = true;
  s->stream = stream;
  s->last_use = ++g_use_clock;
  return s;
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
  Scratch& s = stream_scratch(dev, stream)->slot[slot];
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

// ---- tuning knobs ---------------------------------------------------------------------------
namespace {
std::atomic<int> g_tune_gen{0};
std::mutex g_tune_mu;
std::map<std::string, int>& tune_overrides() { static std::map<std::string, int> m; return m; }
std::atomic<int> g_epoch{0};
}  // namespace

int tune_lookup(TuneSlot& slot, const char* name, int dflt) {
  const unsigned gen = (unsigned)g_tune_gen.load(std::memory_order_acquire);
  const unsigned long long seen = __atomic_load_n(&slot.packed, __ATOMIC_ACQUIRE);
  if ((unsigned)(seen >> 32) == gen) return (int)(unsigned)seen;
  std::lock_guard<std::mutex> lk(g_tune_mu);
  int v = dflt;
  auto it = tune_overrides().find(name);
  if (it != tune_overrides().end()) {
    v = it->second;
  } else {
    const std::string env = std::string("SPMMDB_") + name;
    const char* e = getenv(env.c_str());
    if (e != nullptr && *e != '\0') v = atoi(e);
  }
  __atomic_store_n(&slot.packed, ((unsigned long long)gen << 32) | (unsigned)v, __ATOMIC_RELEASE);
  return v;
}

namespace { thread_local RowClip tl_clip = {INT64_MAX, INT64_MAX}; }
RowClip current_row_clip() { return tl_clip; }
ScopedRowClip::ScopedRowClip(int64_t b_rows, int64_t c_rows) : before_(tl_clip) { tl_clip = RowClip{b_rows, c_rows}; }
ScopedRowClip::~ScopedRowClip() { tl_clip = before_; }

int next_epoch() {
  int e = g_epoch.fetch_add(1, std::memory_order_relaxed) + 1;
  if (e == 0) e = g_epoch.fetch_add(1, std::memory_order_relaxed) + 1;
  return e;
}

}  // namespace spmmdb

extern "C" const char* spmmdb_version(void) { return "spmmdb 0.2 (sm_100a)"; }
extern "C" const char* spmmdb_last_error(void) { return spmmdb::g_last_error; }
extern "C" int spmmdb_device_sm(void) { return spmmdb::device_sm(); }

extern "C" int spmmdb_set_tuning(const char* name, int value) {
  if (name == nullptr || *name == '\0') return SPMMDB_ERR_INVALID_VALUE;
  std::lock_guard<std::mutex> lk(spmmdb::g_tune_mu);
  if (value == INT_MIN) spmmdb::tune_overrides().erase(name);
  else spmmdb::tune_overrides()[name] = value;
  spmmdb::g_tune_gen.fetch_add(1, std::memory_order_release);
  return SPMMDB_OK;
}

extern "C" int spmmdb_get_tuning(const char* name, int dflt) {
  if (name == nullptr) return dflt;
  spmmdb::TuneSlot s;
  return spmmdb::tune_lookup(s, name, dflt);
}

namespace {
std::atomic<int> g_bsrmm_math{-1};  // -1: not read from the environment yet
int env_bsrmm_math() {
  const char* e = getenv("SPMMDB_BSRMM_MATH");
  if (e == nullptr) return SPMMDB_MATH_TF32;
  if (strcmp(e, "fp32") == 0 || strcmp(e, "0") == 0) return SPMMDB_MATH_FP32;
  if (strcmp(e, "bf16") == 0 || strcmp(e, "2") == 0) return SPMMDB_MATH_BF16;
  return SPMMDB_MATH_TF32;
}
}  // namespace
extern "C" int spmmdb_set_bsrmm_math(int math) {
  if (math != SPMMDB_MATH_FP32 && math != SPMMDB_MATH_TF32 && math != SPMMDB_MATH_BF16)
    return SPMMDB_ERR_INVALID_VALUE;
  g_bsrmm_math.store(math);
  return SPMMDB_OK;
}
extern "C" int spmmdb_get_bsrmm_math(void) {
  int m = g_bsrmm_math.load();
  if (m < 0) { m = env_bsrmm_math(); g_bsrmm_math.store(m); }
  return m;
}
