// test_bsrmm <p> <blockDim> <dim> <rocsparse|cusparse> <transB 0|1>
// Micro-benchmark of the reference (test_bsrmm.cu:46-182): Bernoulli(p) block pattern over a 131072 x 131072
// matrix with dense U(-1,1) blocks, alpha = 1, beta = 0, column-major C; B column-major (transB = 0) or
// row-major (transB = 1).  "bsrmm cost time / GFLOPs" with the reference's convention
// GFLOPs = nnzb * bs^2 * dim / t.  SPMMDB_BSRMM_MATH = fp32 | tf32 | bf16 selects the engine's arithmetic.
#include <cusparse.h>

#include "common.h"
#include "rocsparse_bsrmm.h"

int main(int argc, char* argv[]) {
  if (argc < 6) { std::printf("usage: test_bsrmm <p> <blockDim> <dim> <rocsparse|cusparse> <transB>\n"); return 1; }
  const float p = std::stof(argv[1]);
  const int bs = std::stoi(argv[2]), dim = std::stoi(argv[3]), transB = std::stoi(argv[5]);
  const std::string impl = argv[4];
  std::printf("p = %f blockDim = %d dim = %d bsrmmImpl = %s transposeB = %d\n", p, bs, dim, impl.c_str(), transB);
  const int m = 2 << 16, n = m;
  const int mb = (m + bs - 1) / bs, nb = (n + bs - 1) / bs;
  if (mb * bs != m || nb * bs != n) { std::printf("blockDim must divide %d\n", m); return 1; }
  drv::HostBsr a = drv::random_bsr(mb, nb, bs, p);
  const int nnzb = a.nnzb();
  std::printf("nnzb = %d\n", nnzb);
  std::vector<float> hy = drv::random_dense((size_t)n, (size_t)dim);
  drv::DeviceArray<int> rowptr(a.rowptr), colind(a.colind);
  drv::DeviceArray<float> val(a.val), y(hy), z((size_t)m * dim);
  z.zero();
  const float alpha = 1.0f, beta = 0.0f;
  cusparseHandle_t handle = nullptr;
  cusparseMatDescr_t descr = nullptr;
  cusparseCreate(&handle);
  cusparseCreateMatDescr(&descr);
  cusparseSetMatType(descr, CUSPARSE_MATRIX_TYPE_GENERAL);
  cusparseSetMatIndexBase(descr, CUSPARSE_INDEX_BASE_ZERO);
  const cusparseOperation_t opB = transB ? CUSPARSE_OPERATION_TRANSPOSE : CUSPARSE_OPERATION_NON_TRANSPOSE;
  const int ldb = transB ? dim : n;
  auto run = [&]() -> int {
    if (impl == "rocsparse") {
      cusparseStatus_t st = rocsparse_bsrmm_template<float>(handle, CUSPARSE_DIRECTION_ROW, CUSPARSE_OPERATION_NON_TRANSPOSE,
                                                            opB, mb, dim, nb, nnzb, alpha, descr, val.get(), rowptr.get(),
                                                            colind.get(), bs, y.get(), ldb, beta, z.get(), m);
      return st == CUSPARSE_STATUS_SUCCESS ? 0 : (int)st;
    }
    if (impl == "cusparse") {
      cusparseStatus_t st = cusparseSbsrmm(handle, CUSPARSE_DIRECTION_ROW, CUSPARSE_OPERATION_NON_TRANSPOSE, opB, mb, dim,
                                           nb, nnzb, &alpha, descr, val.get(), rowptr.get(), colind.get(), bs, y.get(),
                                           ldb, &beta, z.get(), m);
      return st == CUSPARSE_STATUS_SUCCESS ? 0 : (int)st;
    }
    return -99;
  };
  int rc = run();  // warm-up
  if (rc != 0) { std::printf("bsrmm failed with status %d\n", rc); return 1; }
  DRV_CUDA(cudaDeviceSynchronize());
  drv::GpuTimer t;
  t.start();
  run();
  const float ms = t.stop_ms();
  std::printf("bsrmm cost time: %6.10f ms\nGFLOPs: %6.10f\n", ms, (nnzb / 1.0e6) * ((double)bs * bs * dim) / ms);
  std::vector<float> hz = z.download();
  double sum = 0;
  for (float v : hz) sum += v;
  std::printf("checksum: %.6e\n", sum);
  cusparseDestroyMatDescr(descr);
  cusparseDestroy(handle);
  return 0;
}
