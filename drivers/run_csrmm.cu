// run_csrmm <graph> <dim> <cusparseScsrmm|cusparseScsrmm2|gespmm|cusparseSpMM> <transB 0|1>
// Drop-in for the reference driver of the same name (run_csrmm.cu:46-172): same positional arguments,
// same graph files (tmp/<graph>_indptr.txt / _indices.txt), unit weights, U(-1,1) B, 10 timed epochs,
// same printed lines.  "gespmm" calls the drop-in gespmm_csrmm<float> (row-major B and C);
// "cusparseScsrmm" / "cusparseScsrmm2" no longer exist in cuSPARSE 12 -- those names run the ENGINE with
// the layouts of those calls (column-major C; B column-major for transB = 0, row-major for transB = 1,
// run_csrmm.cu:133-142); "cusparseSpMM" runs the vendor library on the same layouts, for comparison.
#include <cusparse.h>

#include <iostream>

#include "common.h"
#include "gespmm_csrmm.h"

int main(int argc, char* argv[]) {
  if (argc < 5) {
    std::printf("usage: run_csrmm <graph> <dim> <cusparseScsrmm|cusparseScsrmm2|gespmm|cusparseSpMM> <transB>\n");
    return 1;
  }
  const std::string graph = argv[1], impl = argv[3];
  const int dim = std::stoi(argv[2]), transB = std::stoi(argv[4]);
  std::printf("graph = %s dim = %d csrmmImpl = %s transposeB = %d\n", graph.c_str(), dim, impl.c_str(), transB);
  if (impl == "cusparseScsrmm" && transB != 0) {
    std::printf("cusparseScsrmm takes op(B) = N only\n");
    return 1;
  }

  std::printf("load CSR matrix...\n");
  drv::HostCsr a = drv::load_csr_text("tmp/" + graph);
  const int n = a.m, nnz = a.nnz();
  std::cout << "n=" << n << " nnz=" << nnz << std::endl;

  std::vector<float> hy = drv::random_dense((size_t)n, (size_t)dim);
  std::printf("gpu memory malloc and memcpy...\n");
  drv::DeviceArray<int> rowptr(a.rowptr), colind(a.colind);
  drv::DeviceArray<float> val(a.val), y(hy), z((size_t)n * dim);
  z.zero();

  // gespmm: y is [n][dim] row-major.  cuSPARSE layouts: transB = 0 -> y column-major n x dim (ldb = n);
  // transB = 1 -> y holds B^T column-major = B row-major (ldb = dim).  C column-major, ldc = n.
  const bool vendor = impl == "cusparseSpMM";
  cusparseHandle_t handle = nullptr;
  cusparseSpMatDescr_t matA = nullptr;
  cusparseDnMatDescr_t matB = nullptr, matC = nullptr;
  drv::DeviceArray<char> workspace;
  const float alpha = 1.0f, beta = 0.0f;
  if (vendor) {
    cusparseCreate(&handle);
    cusparseCreateCsr(&matA, n, n, nnz, rowptr.get(), colind.get(), val.get(), CUSPARSE_INDEX_32I, CUSPARSE_INDEX_32I,
                      CUSPARSE_INDEX_BASE_ZERO, CUDA_R_32F);
    if (transB) cusparseCreateDnMat(&matB, n, dim, dim, y.get(), CUDA_R_32F, CUSPARSE_ORDER_ROW);
    else cusparseCreateDnMat(&matB, n, dim, n, y.get(), CUDA_R_32F, CUSPARSE_ORDER_COL);
    cusparseCreateDnMat(&matC, n, dim, n, z.get(), CUDA_R_32F, CUSPARSE_ORDER_COL);
    size_t ws = 0;
    cusparseSpMM_bufferSize(handle, CUSPARSE_OPERATION_NON_TRANSPOSE, CUSPARSE_OPERATION_NON_TRANSPOSE, &alpha, matA,
                            matB, &beta, matC, CUDA_R_32F, CUSPARSE_SPMM_ALG_DEFAULT, &ws);
    workspace.resize(ws ? ws : 1);
  }

  std::printf("%s...\n", impl.c_str());
  const int epoch = 10;
  float total = 0.f;
  drv::GpuTimer timer;
  for (int i = 0; i < epoch; ++i) {
    timer.start();
    if (impl == "gespmm") {
      gespmm_csrmm<float>(n, dim, rowptr.get(), colind.get(), val.get(), y.get(), z.get());
    } else if (impl == "cusparseScsrmm" || impl == "cusparseScsrmm2") {
      DRV_SPMMDB(spmmdb_csrmm_f32(n, n, dim, nnz, rowptr.get(), colind.get(), val.get(), y.get(), transB ? dim : n,
                                  transB ? SPMMDB_ROW_MAJOR : SPMMDB_COL_MAJOR, z.get(), n, SPMMDB_COL_MAJOR, alpha,
                                  beta, nullptr));
    } else if (vendor) {
      cusparseSpMM(handle, CUSPARSE_OPERATION_NON_TRANSPOSE, CUSPARSE_OPERATION_NON_TRANSPOSE, &alpha, matA, matB, &beta,
                   matC, CUDA_R_32F, CUSPARSE_SPMM_ALG_DEFAULT, workspace.get());
    } else {
      std::printf("unknown csrmmImpl %s\n", impl.c_str());
      return 1;
    }
    const float ms = timer.stop_ms();
    std::printf("csrmm cost time:  %3.10f ms \n", ms);
    total += ms;
  }
  std::printf("average csrmm cost time: %3.10f ms\n", total / epoch);
  std::vector<float> hz = z.download();
  double sum = 0;
  for (float v : hz) sum += v;
  std::printf("checksum: %.6e\n", sum);
  if (vendor) {
    cusparseDestroySpMat(matA);
    cusparseDestroyDnMat(matB);
    cusparseDestroyDnMat(matC);
    cusparseDestroy(handle);
  }
  std::printf("end\n");
  return 0;
}
