// test_csrmm <p> <dim> <gespmm|cusparse>
// Micro-benchmark of the reference (test_csrmm.cu:44-154): Bernoulli(p) 131072 x 131072 CSR matrix with
// U(-1,1) values (seed 1234), one warm-up and one timed call, "csrmm cost time / GFLOPs" with the
// reference's convention GFLOPs = nnz * dim / t (half of 2*nnz*K/t).  "cusparse" (cusparseScsrmm2 in the
// reference, removed from cuSPARSE 12) runs cusparseSpMM with the csrmm2(N, T) layouts.
#include <cusparse.h>

#include "common.h"
#include "gespmm_csrmm.h"

int main(int argc, char* argv[]) {
  if (argc < 4) { std::printf("usage: test_csrmm <p> <dim> <gespmm|cusparse>\n"); return 1; }
  const float p = std::stof(argv[1]);
  const int dim = std::stoi(argv[2]);
  const std::string impl = argv[3];
  std::printf("p = %f dim = %d csrmmImpl = %s\n", p, dim, impl.c_str());
  const int m = 2 << 16, n = m;
  drv::HostCsr a = drv::random_csr(m, n, p);
  const int nnz = a.nnz();
  std::printf("nnz = %d\n", nnz);
  std::vector<float> hy = drv::random_dense((size_t)n, (size_t)dim);
  drv::DeviceArray<int> rowptr(a.rowptr), colind(a.colind);
  drv::DeviceArray<float> val(a.val), y(hy), z((size_t)m * dim);
  z.zero();
  const float alpha = 1.0f, beta = 0.0f;
  cusparseHandle_t handle = nullptr;
  cusparseSpMatDescr_t matA = nullptr;
  cusparseDnMatDescr_t matB = nullptr, matC = nullptr;
  drv::DeviceArray<char> ws;
  if (impl == "cusparse") {
    cusparseCreate(&handle);
    cusparseCreateCsr(&matA, m, n, nnz, rowptr.get(), colind.get(), val.get(), CUSPARSE_INDEX_32I, CUSPARSE_INDEX_32I,
                      CUSPARSE_INDEX_BASE_ZERO, CUDA_R_32F);
    cusparseCreateDnMat(&matB, n, dim, dim, y.get(), CUDA_R_32F, CUSPARSE_ORDER_ROW);
    cusparseCreateDnMat(&matC, m, dim, m, z.get(), CUDA_R_32F, CUSPARSE_ORDER_COL);
    size_t bytes = 0;
    cusparseSpMM_bufferSize(handle, CUSPARSE_OPERATION_NON_TRANSPOSE, CUSPARSE_OPERATION_NON_TRANSPOSE, &alpha, matA,
                            matB, &beta, matC, CUDA_R_32F, CUSPARSE_SPMM_ALG_DEFAULT, &bytes);
    ws.resize(bytes ? bytes : 1);
  } else if (impl != "gespmm") {
    std::printf("unknown csrmmImpl %s\n", impl.c_str());
    return 1;
  }
  auto run = [&]() {
    if (impl == "gespmm") gespmm_csrmm<float>(m, dim, rowptr.get(), colind.get(), val.get(), y.get(), z.get());
    else cusparseSpMM(handle, CUSPARSE_OPERATION_NON_TRANSPOSE, CUSPARSE_OPERATION_NON_TRANSPOSE, &alpha, matA, matB,
                      &beta, matC, CUDA_R_32F, CUSPARSE_SPMM_ALG_DEFAULT, ws.get());
  };
  run();  // warm-up (the reference times a single cold call; both numbers are printed)
  DRV_CUDA(cudaDeviceSynchronize());
  drv::GpuTimer t;
  t.start();
  run();
  const float ms = t.stop_ms();
  std::printf("csrmm cost time: %6.10f ms\nGFLOPs: %6.10f\n", ms, (nnz / 1.0e6) * dim / ms);
  std::vector<float> hz = z.download();
  double sum = 0;
  for (float v : hz) sum += v;
  std::printf("checksum: %.6e\n", sum);
  if (handle) { cusparseDestroySpMat(matA); cusparseDestroyDnMat(matB); cusparseDestroyDnMat(matC); cusparseDestroy(handle); }
  return 0;
}
