// block_cublas                       -> the 4 x 4 known-answer case of block_cublas.cu:49-56,87-90
// block_cublas <p> <bsize> <dim>     -> test_block_cublas.cu:149-304: random 131072^2 BSR, timed
// "BSR SpMM as one GEMM per non-zero block": the reference launches nnzb cublasSgemm(N, T, bs, dim, bs) calls
// on one stream per block row (block_cublas.cu:123-136); here it is ONE grouped launch, spmmdb_block_gemm_f32
// (column-major blocks, row-major Y with ld = dim, column-major Z with ld = m, alpha = beta = 1 on zeroed Z).
// SPMMDB_BSRMM_MATH = fp32 | tf32 | bf16 selects CUDA cores or the tcgen05 tensor cores.
#include "common.h"

int main(int argc, char* argv[]) {
  const int math = spmmdb_get_bsrmm_math();
  if (argc < 4) {
    const int bs = 2, mb = 2, dim = 3, m = mb * bs;
    std::vector<int> rp = {0, 2, 3}, ci = {0, 1, 1};
    std::vector<float> v = {0, 3, 1, 2, 4, 2, 0, 0, 0, 5, 0, 8};
    std::vector<float> Y = {6, 0, 0, 0, 7, 5, 4, 3, 0, 0, 0, 7};
    drv::DeviceArray<int> d_rp(rp), d_ci(ci);
    drv::DeviceArray<float> d_v(v), d_y(Y), d_z((size_t)m * dim);
    d_z.zero();
    DRV_SPMMDB(spmmdb_block_gemm_f32(mb, dim, 3, bs, d_rp.get(), d_ci.get(), d_v.get(), d_y.get(), d_z.get(), math, nullptr));
    std::vector<float> z = d_z.download();
    const float want[12] = {16, 26, 0, 20, 19, 20, 0, 15, 5, 10, 0, 56};
    bool ok = true;
    for (int i = 0; i < 12; ++i) { std::printf("%g ", z[i]); ok = ok && z[i] == want[i]; }
    std::printf("\n%s\n", ok ? "block_cublas known answer PASSED" : "block_cublas known answer FAILED");
    return ok ? 0 : 1;
  }
  const double p = std::stod(argv[1]);
  const int bs = std::stoi(argv[2]), dim = std::stoi(argv[3]);
  const int m = 2 << 16, mb = m / bs;
  std::printf("p = %f bsize = %d dim = %d\n", p, bs, dim);
  drv::HostBsr a = drv::random_bsr(mb, mb, bs, p);
  const int nnzb = a.nnzb();
  std::printf("nnzb = %d\n", nnzb);
  std::vector<float> hy = drv::random_dense((size_t)m, (size_t)dim);
  drv::DeviceArray<int> rp(a.rowptr), ci(a.colind);
  drv::DeviceArray<float> v(a.val), y(hy), z((size_t)m * dim);
  z.zero();
  DRV_SPMMDB(spmmdb_block_gemm_f32(mb, dim, nnzb, bs, rp.get(), ci.get(), v.get(), y.get(), z.get(), math, nullptr));  // warm-up
  z.zero();
  drv::GpuTimer t;
  t.start();
  DRV_SPMMDB(spmmdb_block_gemm_f32(mb, dim, nnzb, bs, rp.get(), ci.get(), v.get(), y.get(), z.get(), math, nullptr));
  const float ms = t.stop_ms();
  std::printf("block gemm cost time: %6.10f ms\nGFLOPs: %6.10f\n", ms, (nnzb / 1.0e6) * ((double)bs * bs * dim) / ms);
  std::vector<float> hz = z.download();
  double sum = 0;
  for (float x : hz) sum += x;
  std::printf("checksum: %.6e\n", sum);
  return 0;
}
