// drivers/common.h -- host-side helpers shared by the drop-in drivers: the reference's text graph
// formats (load_data.cc:125-184), its seeded random inputs (load_data.cc:12,29-40; the Bernoulli
// pattern generators are O(nnz) here instead of O(m*n)), RAII device buffers, CUDA-event timing and
// an fp64 host loop used only by check_result.  No kernel lives in a driver: everything on the
// device goes through libspmmdb.so (include/spmmdb.h) or the drop-in headers.
#pragma once
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <random>
#include <stdexcept>
#include <string>
#include <vector>

#include "spmmdb.h"

namespace drv {

// "<msg> in <file> at line<N>" + exit(-1): the reference's checkError convention (utility.cc:15-21)
inline void cuda_check(cudaError_t e, const char* file, int line) {
  if (e != cudaSuccess) {
    std::printf("%s in %s at line%d\n", cudaGetErrorString(e), file, line);
    std::exit(-1);
  }
}
#define DRV_CUDA(x) ::drv::cuda_check((x), __FILE__, __LINE__)
inline void spmmdb_check(int rc, const char* what, const char* file, int line) {
  if (rc != SPMMDB_OK) {
    std::printf("%s: spmmdb status %d (%s) in %s at line%d\n", what, rc, spmmdb_last_error(), file, line);
    std::exit(-1);
  }
}
#define DRV_SPMMDB(x) ::drv::spmmdb_check((x), #x, __FILE__, __LINE__)

template <typename T>
class DeviceArray {
 public:
  DeviceArray() = default;
  explicit DeviceArray(size_t n) { resize(n); }
  DeviceArray(const std::vector<T>& h) { resize(h.size()); upload(h); }
  DeviceArray(const DeviceArray&) = delete;
  DeviceArray& operator=(const DeviceArray&) = delete;
  ~DeviceArray() { if (p_) cudaFree(p_); }
  void resize(size_t n) {
    if (p_) cudaFree(p_);
    p_ = nullptr;
    n_ = n;
    if (n) DRV_CUDA(cudaMalloc(&p_, n * sizeof(T)));
  }
  void upload(const std::vector<T>& h) { if (n_) DRV_CUDA(cudaMemcpy(p_, h.data(), n_ * sizeof(T), cudaMemcpyHostToDevice)); }
  std::vector<T> download() const {
    std::vector<T> h(n_);
    if (n_) DRV_CUDA(cudaMemcpy(h.data(), p_, n_ * sizeof(T), cudaMemcpyDeviceToHost));
    return h;
  }
  void zero() { if (n_) DRV_CUDA(cudaMemset(p_, 0, n_ * sizeof(T))); }
  T* get() const { return p_; }
  size_t size() const { return n_; }

 private:
  T* p_ = nullptr;
  size_t n_ = 0;
};

class GpuTimer {
 public:
  GpuTimer() { DRV_CUDA(cudaEventCreate(&a_)); DRV_CUDA(cudaEventCreate(&b_)); }
  ~GpuTimer() { cudaEventDestroy(a_); cudaEventDestroy(b_); }
  void start() { DRV_CUDA(cudaEventRecord(a_, 0)); }
  float stop_ms() {
    DRV_CUDA(cudaEventRecord(b_, 0));
    DRV_CUDA(cudaEventSynchronize(b_));
    float ms = 0.f;
    DRV_CUDA(cudaEventElapsedTime(&ms, a_, b_));
    return ms;
  }

 private:
  cudaEvent_t a_, b_;
};

struct HostCsr {
  int m = 0, n = 0;
  std::vector<int> rowptr, colind;
  std::vector<float> val;
  int nnz() const { return rowptr.empty() ? 0 : rowptr.back(); }
};

// tmp/<graph>_indptr.txt = "<n+1>\n" then n+1 ints; tmp/<graph>_indices.txt = "<nnz>\n" then nnz ints
// (what dumpCSRToFile writes and loadCSRFromFile reads, load_data.cc:125-165), or tmp/<graph>.csrbin when a
// binary copy exists (drivers/graph2bin makes one): the library's multi-threaded readers, not ifstream >>.
inline HostCsr load_csr_text(const std::string& prefix, float fill = 1.0f) {
  HostCsr a;
  int n = 0;
  int64_t nnz = 0;
  int *rp = nullptr, *ci = nullptr;
  int rc = spmmdb_graph_read(prefix.c_str(), SPMMDB_GRAPH_AUTO, &n, &nnz, &rp, &ci);
  if (rc == SPMMDB_ERR_IO) throw std::runtime_error("cannot open or short read in " + prefix + "_indptr.txt / _indices.txt");
  if (rc != SPMMDB_OK) throw std::runtime_error("malformed graph files " + prefix);
  a.rowptr.assign(rp, rp + n + 1);
  a.colind.assign(ci, ci + nnz);
  spmmdb_graph_free(rp);
  spmmdb_graph_free(ci);
  a.m = a.n = n;
  a.val.assign(a.colind.size(), fill);  // unit adjacency weights (run_csrmm.cu:66)
  return a;
}
inline void dump_csr_text(const std::string& prefix, const HostCsr& a) {
  std::ofstream p(prefix + "_indptr.txt"), i(prefix + "_indices.txt");
  p << a.rowptr.size() << "\n";
  for (int x : a.rowptr) p << x << " ";
  i << a.colind.size() << "\n";
  for (int x : a.colind) i << x << " ";
}

inline std::mt19937_64& rng() {  // the reference's generator and seed (load_data.cc:12)
  static std::mt19937_64 g(1234);
  return g;
}
inline std::vector<float> random_dense(size_t rows, size_t cols) {  // U(-1,1), load_data.cc:29-40
  std::uniform_real_distribution<float> d(-1.0f, 1.0f);
  std::vector<float> v(rows * cols);
  for (auto& x : v) x = d(rng());
  return v;
}
// Bernoulli(p) pattern with U(-1,1) values (load_data.cc:42-69), drawn by geometric gaps: O(nnz)
inline HostCsr random_csr(int m, int n, double p) {
  HostCsr a;
  a.m = m; a.n = n;
  a.rowptr.assign((size_t)m + 1, 0);
  std::uniform_real_distribution<double> u(0.0, 1.0);
  std::uniform_real_distribution<float> dv(-1.0f, 1.0f);
  const double lq = std::log1p(-p);
  long long pos = -1;
  const long long total = (long long)m * n;
  while (true) {
    const double x = std::max(u(rng()), 1e-300);
    pos += (long long)std::floor(std::log(x) / lq) + 1;
    if (pos >= total) break;
    a.rowptr[(size_t)(pos / n) + 1]++;
    a.colind.push_back((int)(pos % n));
    a.val.push_back(dv(rng()));
  }
  for (int i = 0; i < m; ++i) a.rowptr[i + 1] += a.rowptr[i];
  return a;
}
// Bernoulli(p) BLOCK pattern with dense U(-1,1) blocks (load_data.cc:81-113)
struct HostBsr {
  int mb = 0, nb = 0, bs = 0;
  std::vector<int> rowptr, colind;
  std::vector<float> val;
  int nnzb() const { return rowptr.empty() ? 0 : rowptr.back(); }
};
inline HostBsr random_bsr(int mb, int nb, int bs, double p) {
  HostCsr pat = random_csr(mb, nb, p);
  HostBsr b;
  b.mb = mb; b.nb = nb; b.bs = bs;
  b.rowptr = pat.rowptr;
  b.colind = pat.colind;
  b.val = random_dense((size_t)b.nnzb(), (size_t)bs * bs);
  return b;
}

// fp64 host loop: C[m][K] row-major = A * B (B row-major [n][K]); check_result's yardstick
inline std::vector<double> host_csrmm(const HostCsr& a, const std::vector<float>& B, int K) {
  std::vector<double> c((size_t)a.m * K, 0.0);
  for (int i = 0; i < a.m; ++i)
    for (int p = a.rowptr[i]; p < a.rowptr[i + 1]; ++p) {
      const double v = a.val[p];
      const float* b = &B[(size_t)a.colind[p] * K];
      double* o = &c[(size_t)i * K];
      for (int k = 0; k < K; ++k) o[k] += v * b[k];
    }
  return c;
}

inline double gflops(double nnz, double K, double ms) { return nnz * K / (ms * 1e-3) / 1e9; }  // the reference's nnz*K/t

}  // namespace drv
