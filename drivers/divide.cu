// divide <graph> <bsize> <dim> <rocsparse|cusparse> <transB 0|1> <density>
// Drop-in for the reference's hybrid dense-block driver (divide.cu:205-391): split the graph into a BSR part
// (blocks with fill >= density) and a remainder CSR, then C = A_csr * op(B) and C += A_bsr * op(B), both with
// alpha = beta = 1 into one zeroed column-major C (ld = nb * bsize), each timed on its own, same printed
// lines.  The split runs on the GPU here (spmmdb_divide_nnz + spmmdb_divide_f32) instead of the reference's CPU
// divide_matrix (divide.cu:52-127); the CSR product is spmmdb_csrmm_f32 with the cusparseScsrmm2 layouts
// (divide.cu:348-352; that entry point left cuSPARSE in CUDA 11).  "rocsparse" = the drop-in
// rocsparse_bsrmm_template<float> (arithmetic by SPMMDB_BSRMM_MATH = fp32 | tf32 | bf16), "cusparse" = the vendor
// cusparseSbsrmm.  Extra line at the end: the same product by the plain CSR path, and the largest difference.
#include <cusparse.h>

#include <iostream>

#include "common.h"
#include "rocsparse_bsrmm.h"

int main(int argc, char* argv[]) {
  if (argc < 7) {
    std::printf("usage: divide <graph> <bsize> <dim> <rocsparse|cusparse> <transB 0|1> <density>\n");
    return 1;
  }
  const std::string prefix = "tmp/" + std::string(argv[1]);
  std::cout << prefix << std::endl;
  const int bsize = std::stoi(argv[2]), dim = std::stoi(argv[3]);
  const std::string impl(argv[4]);
  const int transposeB = std::stoi(argv[5]);
  const float density = std::stof(argv[6]);
  if (transposeB != 0 && transposeB != 1) { std::printf("transB must be 0 or 1\n"); return 1; }

  std::cout << "csr to adj..." << std::endl;
  drv::HostCsr a = drv::load_csr_text(prefix);
  const int n = a.m, nnz = a.nnz();
  const int nb = (n + bsize - 1) / bsize, n1 = nb * bsize;
  const float alpha = 1.0f, beta = 1.0f;
  const cusparseOperation_t transB = transposeB ? CUSPARSE_OPERATION_TRANSPOSE : CUSPARSE_OPERATION_NON_TRANSPOSE;
  // as the reference: ldb = n for op(B) = N (B column-major; bsrmm needs ldb >= nb * bsize, so pad to n1), dim for T
  const int ldb = transposeB ? dim : n1;

  std::cout << "divide matrix..." << std::endl;
  drv::DeviceArray<int> rowptr(a.rowptr), colind(a.colind);
  drv::DeviceArray<int> csr_rowptr((size_t)n + 1), bsr_rowptr((size_t)nb + 1);
  int csrNnz = 0, bsrNnzb = 0;
  drv::GpuTimer split;
  split.start();
  DRV_SPMMDB(spmmdb_divide_nnz(n, n, rowptr.get(), colind.get(), bsize, density, bsr_rowptr.get(), csr_rowptr.get(),
                               &bsrNnzb, &csrNnz, nullptr));
  std::printf("csr nnz = %d    bsr nnzb = %d\n", csrNnz, bsrNnzb);
  std::cout << "vec2ptr..." << std::endl;
  std::cout << "gpu memory malloc and memcpy..." << std::endl;
  drv::DeviceArray<int> csr_colind((size_t)csrNnz), bsr_colind((size_t)bsrNnzb);
  drv::DeviceArray<float> csr_val((size_t)csrNnz), bsr_val((size_t)bsrNnzb * bsize * bsize);
  DRV_SPMMDB(spmmdb_divide_f32(SPMMDB_DIR_ROW, n, n, rowptr.get(), colind.get(), /*val=*/nullptr, bsize, density,
                               bsr_rowptr.get(), bsr_colind.get(), bsr_val.get(), csr_rowptr.get(), csr_colind.get(),
                               csr_val.get(), nullptr));
  const float split_ms = split.stop_ms();
  std::printf("divide cost time (GPU, incl. allocation):  %3.10f ms, dense part %3.6f of nnz\n", split_ms,
              nnz ? 1.0 - (double)csrNnz / nnz : 0.0);

  std::cout << "prepare y and z..." << std::endl;
  std::vector<float> hy = drv::random_dense((size_t)n1, (size_t)dim);
  drv::DeviceArray<float> y(hy), z((size_t)n1 * dim);
  z.zero();

  cusparseHandle_t handle = nullptr;
  cusparseMatDescr_t bsrDescr = nullptr;
  cusparseCreate(&handle);
  cusparseCreateMatDescr(&bsrDescr);
  cusparseSetMatType(bsrDescr, CUSPARSE_MATRIX_TYPE_GENERAL);
  cusparseSetMatIndexBase(bsrDescr, CUSPARSE_INDEX_BASE_ZERO);

  drv::GpuTimer t1, t2;
  t1.start();
  DRV_SPMMDB(spmmdb_csrmm_f32(n, n, dim, csrNnz, csr_rowptr.get(), csr_colind.get(), csr_val.get(), y.get(), ldb,
                              transposeB ? SPMMDB_ROW_MAJOR : SPMMDB_COL_MAJOR, z.get(), n1, SPMMDB_COL_MAJOR, alpha,
                              beta, nullptr));
  const float time1 = t1.stop_ms();
  t2.start();
  if (impl == "rocsparse") {
    cusparseStatus_t st = rocsparse_bsrmm_template<float>(
        handle, CUSPARSE_DIRECTION_ROW, CUSPARSE_OPERATION_NON_TRANSPOSE, transB, nb, dim, nb, bsrNnzb, alpha, bsrDescr,
        bsr_val.get(), bsr_rowptr.get(), bsr_colind.get(), bsize, y.get(), ldb, beta, z.get(), n1);
    if (st != CUSPARSE_STATUS_SUCCESS) {
      std::printf("error occurred in %s at line %d\n", __FILE__, __LINE__);
      return -1;
    }
  } else if (impl == "cusparse") {
    if (bsrNnzb > 0)
      cusparseSbsrmm(handle, CUSPARSE_DIRECTION_ROW, CUSPARSE_OPERATION_NON_TRANSPOSE, transB, nb, dim, nb, bsrNnzb, &alpha,
                     bsrDescr, bsr_val.get(), bsr_rowptr.get(), bsr_colind.get(), bsize, y.get(), ldb, &beta, z.get(), n1);
  } else {
    std::printf("unknown bsrmmImpl %s\n", impl.c_str());
    return 1;
  }
  const float time2 = t2.stop_ms();
  std::printf("csrmm cost time:  %3.10f ms \n", time1);
  std::printf("bsrmm cost time:  %3.10f ms \n", time2);
  std::printf("total cost time:  %3.10f ms \n", time1 + time2);
  std::printf("%3.5f+%3.5f=%3.5f\n", time1, time2, time1 + time2);
  std::vector<float> hz = z.download();

  // not in the reference: the same product by the plain CSR path, to show the split changes nothing
  drv::DeviceArray<float> val(a.val), zc((size_t)n1 * dim);
  zc.zero();
  drv::GpuTimer t3;
  t3.start();
  DRV_SPMMDB(spmmdb_csrmm_f32(n, n, dim, nnz, rowptr.get(), colind.get(), val.get(), y.get(), ldb,
                              transposeB ? SPMMDB_ROW_MAJOR : SPMMDB_COL_MAJOR, zc.get(), n1, SPMMDB_COL_MAJOR, 1.0f,
                              0.0f, nullptr));
  const float time3 = t3.stop_ms();
  std::vector<float> hc = zc.download();
  double maxdiff = 0, sum = 0;
  for (size_t i = 0; i < hz.size(); ++i) {
    maxdiff = std::max(maxdiff, (double)std::fabs(hz[i] - hc[i]));
    sum += hz[i];
  }
  std::printf("plain csrmm cost time:  %3.10f ms, max |hybrid - csr| = %.3e, checksum %.6e\n", time3, maxdiff, sum);

  cusparseDestroyMatDescr(bsrDescr);
  cusparseDestroy(handle);
  std::printf("end\n");
  return 0;
}
