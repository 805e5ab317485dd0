// run_bsrmm <graph> <blockDim> <dim> <rocsparse|cusparse>
// Drop-in for the reference driver (run_bsrmm.cu:47-187): load the CSR graph, csr2bsr on the GPU (here
// spmmdb_csr2bsr_nnz + spmmdb_csr2bsr_f32 instead of cusparseXcsr2bsrNnz + cusparseScsr2bsr), then time ONE
// bsrmm call with alpha = 1, beta = 1 on zeroed C, op(B) = N (column-major B and C, ld = nb * blockDim).
// "rocsparse" = the drop-in rocsparse_bsrmm_template<float> (arithmetic selected by SPMMDB_BSRMM_MATH =
// fp32 | tf32 | bf16), "cusparse" = the vendor cusparseSbsrmm for comparison.  Same printed lines.
#include <cusparse.h>

#include <iostream>

#include "common.h"
#include "rocsparse_bsrmm.h"

int main(int argc, char* argv[]) {
  if (argc < 5) {
    std::printf("usage: run_bsrmm <graph> <blockDim> <dim> <rocsparse|cusparse>\n");
    return 1;
  }
  const std::string prefix = "tmp/" + std::string(argv[1]), impl = argv[4];
  std::cout << prefix << std::endl;
  const int bs = std::stoi(argv[2]), dim = std::stoi(argv[3]);

  std::printf("load CSR matrix...\n");
  drv::HostCsr a = drv::load_csr_text(prefix);
  const int n = a.m, nnz = a.nnz();
  std::cout << "n=" << n << " nnz=" << nnz << std::endl;
  const int nb = (n + bs - 1) / bs;
  const float alpha = 1.0f, beta = 1.0f;

  std::vector<float> hy = drv::random_dense((size_t)nb * bs, (size_t)dim);
  drv::DeviceArray<float> y(hy), z((size_t)nb * bs * dim);
  z.zero();
  std::printf("gpu memory malloc and memcpy...\n");
  drv::DeviceArray<int> rowptr(a.rowptr), colind(a.colind);
  drv::DeviceArray<float> val(a.val);

  cusparseHandle_t handle = nullptr;
  cusparseMatDescr_t descr = nullptr;
  cusparseCreate(&handle);
  cusparseCreateMatDescr(&descr);
  cusparseSetMatType(descr, CUSPARSE_MATRIX_TYPE_GENERAL);
  cusparseSetMatIndexBase(descr, CUSPARSE_INDEX_BASE_ZERO);

  drv::DeviceArray<int> bsr_rowptr((size_t)nb + 1);
  int nnzb = 0;
  drv::GpuTimer conv;
  conv.start();
  DRV_SPMMDB(spmmdb_csr2bsr_nnz(n, n, rowptr.get(), colind.get(), bs, bsr_rowptr.get(), &nnzb, nullptr));
  const long long numVal = (long long)nnzb * (long long)(bs * bs) * (long long)sizeof(float);
  std::printf("numVal = %lld\n", numVal);
  drv::DeviceArray<int> bsr_colind((size_t)nnzb);
  drv::DeviceArray<float> bsr_val((size_t)nnzb * bs * bs);
  DRV_SPMMDB(spmmdb_csr2bsr_f32(SPMMDB_DIR_ROW, n, n, rowptr.get(), colind.get(), val.get(), bs, bsr_rowptr.get(),
                                bsr_colind.get(), bsr_val.get(), nullptr));
  const float conv_ms = conv.stop_ms();
  std::printf("density:  %3.10f \n", (1.0 * nnzb) / ((nb * 1.0) * (nb * 1.0)));
  std::printf("csr2bsr cost time (incl. allocation):  %3.10f ms, fill %3.6f\n", conv_ms,
              nnzb ? (double)nnz / ((double)nnzb * bs * bs) : 0.0);

  std::printf("%s bsrmm...\n", impl.c_str());
  drv::GpuTimer timer;
  timer.start();
  if (impl == "rocsparse") {
    cusparseStatus_t st = rocsparse_bsrmm_template<float>(
        handle, CUSPARSE_DIRECTION_ROW, CUSPARSE_OPERATION_NON_TRANSPOSE, CUSPARSE_OPERATION_NON_TRANSPOSE, nb, dim, nb,
        nnzb, alpha, descr, bsr_val.get(), bsr_rowptr.get(), bsr_colind.get(), bs, y.get(), nb * bs, beta, z.get(),
        nb * bs);
    if (st != CUSPARSE_STATUS_SUCCESS) {
      std::printf("%s in %s at line%d\n", cusparseGetErrorString(st), __FILE__, __LINE__);
      return -1;
    }
  } else if (impl == "cusparse") {
    cusparseSbsrmm(handle, CUSPARSE_DIRECTION_ROW, CUSPARSE_OPERATION_NON_TRANSPOSE, CUSPARSE_OPERATION_NON_TRANSPOSE, nb,
                   dim, nb, nnzb, &alpha, descr, bsr_val.get(), bsr_rowptr.get(), bsr_colind.get(), bs, y.get(), nb * bs,
                   &beta, z.get(), nb * bs);
  } else {
    std::printf("unknown bsrmmImpl %s\n", impl.c_str());
    return 1;
  }
  const float ms = timer.stop_ms();
  std::printf("bsrmm cost time:  %3.10f ms \n", ms);
  std::vector<float> hz = z.download();
  double sum = 0;
  for (float v : hz) sum += v;
  std::printf("checksum: %.6e\n", sum);
  cusparseDestroyMatDescr(descr);
  cusparseDestroy(handle);
  std::printf("end\n");
  return 0;
}
