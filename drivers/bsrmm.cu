// bsrmm (no arguments) -- the reference's known-answer program bsrmm.cu:35-177: block size 2, mb = 2, kb = 3,
// bsrRowPtr {0,2,4}, bsrColInd {0,2,1,2}, row-major blocks {0,4,2,7 | 1,8,2,0 | 9,0,0,2 | 0,6,7,0}, B = 1..12
// column-major 6 x 2, through the drop-in rocsparse_bsrmm_template<float> (the call the reference makes at
// bsrmm.cu:141-144 goes to cusparseSbsrmm; both run here).  Prints the m * n results with "%f " like
// bsrmm.cu:171-175; exit code 0 only for 61 26 63 43 139 92 153 97 (SURVEY 8c golden G2) from both.
#include <cusparse.h>

#include "common.h"
#include "rocsparse_bsrmm.h"

int main() {
  const int bs = 2, mb = 2, kb = 3, n = 2, nnzb = 4, m = mb * bs, k = kb * bs;
  const std::vector<int> rowptr = {0, 2, 4}, colind = {0, 2, 1, 2};
  const std::vector<float> val = {0, 4, 2, 7, 1, 8, 2, 0, 9, 0, 0, 2, 0, 6, 7, 0};
  const std::vector<float> y = {1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12};
  drv::DeviceArray<int> d_rp(rowptr), d_ci(colind);
  drv::DeviceArray<float> d_v(val), d_y(y), d_z(m * n);
  cusparseHandle_t handle = nullptr;
  cusparseMatDescr_t descr = nullptr;
  if (cusparseCreate(&handle) != CUSPARSE_STATUS_SUCCESS || cusparseCreateMatDescr(&descr) != CUSPARSE_STATUS_SUCCESS) {
    std::printf("CUSPARSE Library initialization failed\n");
    return 1;
  }
  const float want[8] = {61, 26, 63, 43, 139, 92, 153, 97};
  bool ok = true;
  for (int impl = 0; impl < 2; ++impl) {
    DRV_CUDA(cudaMemset(d_z.get(), 0, sizeof(float) * m * n));
    const float one = 1.0f, zero = 0.0f;
    cusparseStatus_t st =
        impl == 0 ? rocsparse_bsrmm_template<float>(handle, CUSPARSE_DIRECTION_ROW, CUSPARSE_OPERATION_NON_TRANSPOSE,
                                                    CUSPARSE_OPERATION_NON_TRANSPOSE, mb, n, kb, nnzb, one, descr, d_v.get(),
                                                    d_rp.get(), d_ci.get(), bs, d_y.get(), k, zero, d_z.get(), m)
                  : cusparseSbsrmm(handle, CUSPARSE_DIRECTION_ROW, CUSPARSE_OPERATION_NON_TRANSPOSE,
                                   CUSPARSE_OPERATION_NON_TRANSPOSE, mb, n, kb, nnzb, &one, descr, d_v.get(), d_rp.get(),
                                   d_ci.get(), bs, d_y.get(), k, &zero, d_z.get(), m);
    if (st != CUSPARSE_STATUS_SUCCESS) { std::printf("bsrmm failed: status %d\n", (int)st); return 1; }
    DRV_CUDA(cudaDeviceSynchronize());
    const std::vector<float> z = d_z.download();
    for (int i = 0; i < m * n; ++i) std::printf("%f ", z[i]);
    std::printf("\n");
    for (int i = 0; i < 8; ++i) ok = ok && z[i] == want[i];
  }
  cusparseDestroyMatDescr(descr);
  cusparseDestroy(handle);
  std::printf("end\n");
  return ok ? 0 : 1;
}
