// csrmm (no arguments) -- the reference's known-answer program csrmm.cu:26-218: the 4 x 4 example matrix
// A = [[1,0,2,3],[0,4,0,0],[5,0,6,7],[0,8,0,9]] times the column-major 4 x 2 matrix (10,20,30,40 | 50,60,70,80)
// through the cuSPARSE-layout entry (column-major B and C, alpha = 1, beta = 0, csrmm.cu:176-186).  Prints
// "testing examples", the 2 * n results with "%f " like csrmm.cu:212-216, and "end"; exit code 0 only when
// the values are 190 80 510 520 430 240 1230 1200 (SURVEY 8c golden G1).
#include "common.h"

int main() {
  std::printf("testing examples\n");
  const int n = 4, nnz = 9;
  const std::vector<int> rowptr = {0, 3, 4, 7, 9};
  const std::vector<int> colind = {0, 2, 3, 1, 0, 2, 3, 1, 3};
  const std::vector<float> val = {1, 2, 3, 4, 5, 6, 7, 8, 9};
  const std::vector<float> y = {10, 20, 30, 40, 50, 60, 70, 80};
  drv::DeviceArray<int> d_rp(rowptr), d_ci(colind);
  drv::DeviceArray<float> d_v(val), d_y(y), d_z(2 * n);
  DRV_CUDA(cudaMemset(d_z.get(), 0, sizeof(float) * 2 * n));
  DRV_SPMMDB(spmmdb_csrmm_f32(n, n, 2, nnz, d_rp.get(), d_ci.get(), d_v.get(), d_y.get(), n, SPMMDB_COL_MAJOR, d_z.get(), n,
                              SPMMDB_COL_MAJOR, 1.0f, 0.0f, nullptr));
  DRV_CUDA(cudaDeviceSynchronize());
  const std::vector<float> z = d_z.download();
  for (int i = 0; i < 2 * n; ++i) std::printf("%f ", z[i]);
  std::printf("\n");
  const float want[8] = {190, 80, 510, 520, 430, 240, 1230, 1200};
  bool ok = true;
  for (int i = 0; i < 8; ++i) ok = ok && z[i] == want[i];
  std::printf("end\n");
  return ok ? 0 : 1;
}
