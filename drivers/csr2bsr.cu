// csr2bsr <p>       (reference: csr2bsr.cu:87-313)
// Random 1000 x 1200 CSR -> csr2bsr (bs = 2) on the GPU, then csrmm vs bsrmm must agree within 0.01 (:297-305).
#include <cmath>

#include "common.h"

int main(int argc, char* argv[]) {
  const double p = argc > 1 ? std::stod(argv[1]) : 0.1;
  const int m = 1000, n = 1200, bs = 2, dim = 5;
  drv::HostCsr a = drv::random_csr(m, n, p);
  const int nnz = a.nnz(), mb = (m + bs - 1) / bs, nb = (n + bs - 1) / bs;
  std::printf("p = %f m = %d n = %d nnz = %d\n", p, m, n, nnz);
  std::vector<float> hy = drv::random_dense((size_t)nb * bs, (size_t)dim);  // column-major (nb*bs) x dim
  drv::DeviceArray<int> rowptr(a.rowptr), colind(a.colind), brp((size_t)mb + 1);
  drv::DeviceArray<float> val(a.val), y(hy), z1((size_t)m * dim), z2((size_t)mb * bs * dim);
  int nnzb = 0;
  DRV_SPMMDB(spmmdb_csr2bsr_nnz(m, n, rowptr.get(), colind.get(), bs, brp.get(), &nnzb, nullptr));
  std::printf("nnzb = %d\n", nnzb);
  drv::DeviceArray<int> bci((size_t)nnzb);
  drv::DeviceArray<float> bv((size_t)nnzb * bs * bs);
  DRV_SPMMDB(spmmdb_csr2bsr_f32(SPMMDB_DIR_ROW, m, n, rowptr.get(), colind.get(), val.get(), bs, brp.get(), bci.get(),
                                bv.get(), nullptr));
  z1.zero(); z2.zero();
  DRV_SPMMDB(spmmdb_csrmm_f32(m, n, dim, nnz, rowptr.get(), colind.get(), val.get(), y.get(), nb * bs, SPMMDB_COL_MAJOR,
                              z1.get(), m, SPMMDB_COL_MAJOR, 1.0f, 0.0f, nullptr));
  DRV_SPMMDB(spmmdb_bsrmm_f32(SPMMDB_DIR_ROW, 0, mb, dim, nb, nnzb, bs, 1.0f, bv.get(), brp.get(), bci.get(), y.get(),
                              nb * bs, 0.0f, z2.get(), mb * bs, SPMMDB_COL_MAJOR, SPMMDB_MATH_FP32, nullptr));
  std::vector<float> h1 = z1.download(), h2 = z2.download();
  double d = 0;
  for (int c = 0; c < dim; ++c)
    for (int r = 0; r < m; ++r) d = std::max(d, (double)std::fabs(h1[(size_t)c * m + r] - h2[(size_t)c * mb * bs + r]));
  std::printf("max |csrmm - bsrmm| = %g\n", d);
  std::printf(d < 0.01 ? "csr2bsr test PASSED\n" : "csr2bsr test FAILED\n");
  return d < 0.01 ? 0 : 1;
}
