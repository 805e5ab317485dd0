// bsr2csr <p> <blockDim>      (reference: bsr2csr.cu:90-312)
// Random BSR over 4096 x 4096 -> bsr2csr on the GPU (nnz = nnzb * bs^2, :177), bsrmm vs csrmm within 0.05
// (:296-304), prints both times and their ratio like the original.
#include <cmath>

#include "common.h"

int main(int argc, char* argv[]) {
  const double p = argc > 1 ? std::stod(argv[1]) : 0.05;
  const int bs = argc > 2 ? std::stoi(argv[2]) : 4;
  const int m = 4096, dim = 64;
  const int mb = m / bs;
  drv::HostBsr b = drv::random_bsr(mb, mb, bs, p);
  const int nnzb = b.nnzb();
  const long long nnz = (long long)nnzb * bs * bs;
  std::printf("p = %f blockDim = %d nnzb = %d nnz = %lld\n", p, bs, nnzb, nnz);
  std::vector<float> hy = drv::random_dense((size_t)m, (size_t)dim);  // column-major m x dim
  drv::DeviceArray<int> brp(b.rowptr), bci(b.colind), crp((size_t)m + 1), cci((size_t)nnz);
  drv::DeviceArray<float> bv(b.val), cv((size_t)nnz), y(hy), z1((size_t)m * dim), z2((size_t)m * dim);
  DRV_SPMMDB(spmmdb_bsr2csr_f32(SPMMDB_DIR_ROW, mb, mb, brp.get(), bci.get(), bv.get(), bs, crp.get(), cci.get(), cv.get(),
                                nullptr));
  z1.zero(); z2.zero();
  drv::GpuTimer t;
  t.start();
  DRV_SPMMDB(spmmdb_bsrmm_f32(SPMMDB_DIR_ROW, 0, mb, dim, mb, nnzb, bs, 1.0f, bv.get(), brp.get(), bci.get(), y.get(), m,
                              0.0f, z1.get(), m, SPMMDB_COL_MAJOR, SPMMDB_MATH_FP32, nullptr));
  const float t_bsr = t.stop_ms();
  t.start();
  DRV_SPMMDB(spmmdb_csrmm_f32(m, m, dim, nnz, crp.get(), cci.get(), cv.get(), y.get(), m, SPMMDB_COL_MAJOR, z2.get(), m,
                              SPMMDB_COL_MAJOR, 1.0f, 0.0f, nullptr));
  const float t_csr = t.stop_ms();
  std::vector<float> h1 = z1.download(), h2 = z2.download();
  double d = 0;
  for (size_t i = 0; i < h1.size(); ++i) d = std::max(d, (double)std::fabs(h1[i] - h2[i]));
  std::printf("bsrmm cost time: %f ms, csrmm cost time: %f ms, ratio %f\n", t_bsr, t_csr, t_csr / t_bsr);
  std::printf("max |bsrmm - csrmm| = %g\n", d);
  std::printf(d < 0.05 ? "bsr2csr test PASSED\n" : "bsr2csr test FAILED\n");
  return d < 0.05 ? 0 : 1;
}
