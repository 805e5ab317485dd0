// cusparse_legacy.h -- TEST INFRASTRUCTURE (oracle/_ref build only).
// The reference's divide.cu calls cusparseScsrmm2 from its main() (divide.cu:348-352); that symbol
// left cuSPARSE in CUDA 11, so the file does not compile against the 12.x toolkit.  The oracle only
// wants the file's CPU function divide_matrix (divide.cu:52-127); this forced-include declares the
// missing entry point as a stub so that the translation unit compiles unchanged where it lies.
// The stub is never executed (main is renamed and never called).
#pragma once
#include <cusparse.h>
static inline cusparseStatus_t cusparseScsrmm2(cusparseHandle_t, cusparseOperation_t, cusparseOperation_t, int, int,
                                               int, int, const float*, const cusparseMatDescr_t, const float*,
                                               const int*, const int*, const float*, int, const float*, float*,
                                               int) {
  return CUSPARSE_STATUS_NOT_SUPPORTED;
}
