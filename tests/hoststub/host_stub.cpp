// host_stub.cpp -- TEST INFRASTRUCTURE ONLY.  A stand-in for the device half of the C-ABI so that the host operator
// (la-llama.cpp_b200/csrc/lamm_host.cpp: gates, batch-plane loop, broadcast factors, mul_mat_id grouping, gather and
// scatter) can be exercised on a machine without a GPU: the handful of lammcu_* functions the operator calls are
// implemented here on the CPU oracle (oracle/lamm_oracle.c).  tests/test_host_logic.py links this file, lamm_host.cpp
// and the oracle into tests/hoststub/libhoststub.so.  Nothing of this is built into, or reachable from, the product.
#include "../../include/lammcu.h"
#include "../../oracle/lamm_oracle.h"

#include <cstdint>
#include <vector>

struct lammcu_weight {
  const void *host;
  int type;
  int64_t M, K, stride, row_begin, row_end;
};

static std::vector<lammcu_weight *> g_weights;
static long g_calls = 0, g_rows = 0;

extern "C" {

int lammcu_init(int) { return LAMMCU_OK; }
const char *lammcu_last_error(void) { return "host stub"; }
int lammcu_blck_size(int t) { return orc_blck_size(t); }
size_t lammcu_type_size(int t) { return orc_type_size(t); }
size_t lammcu_row_size(int t, int64_t k) { return orc_row_size(t, k); }
int lammcu_vec_dot_type(int t) { return orc_vec_dot_type(t); }
int lammcu_type_supported(int t) {                         // the product's table (tests/test_abi.py::test_type_table)
  return t == 0 || t == 1 || t == 2 || t == 3 || t == 6 || t == 7 || t == 8 || (t >= 10 && t <= 14);
}
int lammcu_weight_register(const void *w_host, int w_type, int64_t M, int64_t K, int64_t stride, int64_t row_begin,
                           int64_t row_end, lammcu_weight_t *out) {
  lammcu_weight *w = new lammcu_weight{w_host, w_type, M, K, stride, row_begin, row_end};
  g_weights.push_back(w);
  *out = w;
  return LAMMCU_OK;
}
long g_refresh = 0;
int lammcu_weight_for_op(const void *w_host, int w_type, int64_t M, int64_t K, int64_t stride, int refresh,
                         lammcu_weight_t *out) {
  g_refresh += refresh ? 1 : 0;
  return lammcu_weight_register(w_host, w_type, M, K, stride, 0, M, out);
}
long stub_refreshes(void) { return g_refresh; }
int lammcu_mul_mat_host(lammcu_weight_t w, const float *x, int64_t N, int64_t ldx, float *c, int64_t ldc, void *xq_out) {
  g_calls++;
  g_rows += N;
  const int64_t rows = w->row_end - w->row_begin;
  return orc_mul_mat(w->type, (const char *)w->host + w->row_begin * w->stride, w->stride, x, ldx, c + w->row_begin, ldc,
                     rows, N, w->K, xq_out) == 0 ? LAMMCU_OK : LAMMCU_ERR_SHAPE;
}
int lammcu_mul_mat_q_host(lammcu_weight_t w, const void *xq, int64_t xq_stride, int64_t N, float *c, int64_t ldc) {
  g_calls++;
  g_rows += N;
  const int64_t rows = w->row_end - w->row_begin;
  return orc_mul_mat_q(w->type, (const char *)w->host + w->row_begin * w->stride, w->stride, xq, xq_stride,
                       c + w->row_begin, ldc, rows, N, w->K) == 0 ? LAMMCU_OK : LAMMCU_ERR_SHAPE;
}
// test hooks
long stub_calls(void) { return g_calls; }
long stub_rows(void) { return g_rows; }
void stub_reset(void) {
  g_calls = g_rows = g_refresh = 0;
  for (lammcu_weight *w : g_weights) delete w;
  g_weights.clear();
}

}  // extern "C"
