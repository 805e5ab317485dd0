// bsrmm_tc.cu -- BSR x dense on tcgen05 tensor cores (placeholder until the
// kernel lands: reports "unsupported" so every call takes the fp32 kernel).
#include "common.cuh"

namespace spmmdb {

bool bsrmm_tc_supported(int, int, int, int64_t, const float*, const float*, int, int64_t, const float*) {
  return false;
}

int bsrmm_tc(int, int, int, int, int, int64_t, int, float, const float*, const int*, const int*, const float*,
             int64_t, float, float*, int64_t, int, int, cudaStream_t) {
  return SPMMDB_ERR_NOT_SUPPORTED;
}

}  // namespace spmmdb
