// lamm_gemm.cu -- prefill path (N > gemv_max_n).  PLACEHOLDER for this commit:
// routes to the column-grouped GEMV (correct, re-reads W once per 8 columns).
// The tcgen05 int8 kernel replaces this file's body.
#include "lamm_common.cuh"

namespace lamm {

int launch_gemm(int w_type, const void *w, int64_t w_stride, const void *xq, int64_t xq_stride,
                float *c, int64_t ldc, int64_t M, int64_t N, int64_t K, cudaStream_t st) {
  return launch_gemv(w_type, w, w_stride, xq, xq_stride, c, ldc, M, N, K, st);
}

}  // namespace lamm
