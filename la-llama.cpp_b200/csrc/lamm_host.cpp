// lamm_host.cpp -- the drop-in operator: lamm_can_mul_mat / lamm_mul_mat /
// lamm_get_opt_level with the reference's signatures and gate semantics
// (src/loongarch_matmul.h:18-24, src/loongarch_matmul.cpp:10-145), implemented on
// the C-ABI of include/lammcu.h.  Host C++ only; all device work is behind lammcu_*.
//
// Differences from the reference that stay ABI-compatible:
//   * the work is done once, by thread ith == 0; other threads return at once and
//     see dst after ggml's post-COMPUTE barrier (ggml.c:18443-18450);
//   * the activation quantize is part of the operator (fused on the GPU from the
//     f32 src1->data) instead of ggml's INIT phase; params->wdata is only read when
//     the caller asks for it (lamm_set_writeback_wdata) or src1 is pre-quantized;
//   * all M rows are computed (the reference drops M % nth rows);
//   * the gate also refuses what the CUDA path cannot take, so ggml falls back.
#include "../../include/lamm_b200.h"
#include "../../include/lammcu.h"

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#ifndef LAMM_OPT_LEVEL
#define LAMM_OPT_LEVEL 3        // src/lamm_common.h:96-100
#endif

namespace {

std::atomic<int> g_writeback{0};

int opt_level() {
  static const int lvl = [] {
    const char *e = std::getenv("LAMM_OPT_LEVEL");
    return e ? std::atoi(e) : LAMM_OPT_LEVEL;
  }();
  return lvl;
}

bool device_ok() {
  static const bool ok = (lammcu_init(-1) == LAMMCU_OK);
  return ok;
}

// ggml_is_contiguous (ggml.c:2616-2624)
bool is_contiguous(const ggml_tensor *t) {
  const size_t ts = lammcu_type_size(t->type);
  const int bs = lammcu_blck_size(t->type);
  if (!ts || !bs) return false;
  return t->nb[0] == ts && t->nb[1] == (t->nb[0] * t->ne[0]) / bs &&
         t->nb[2] == t->nb[1] * t->ne[1] && t->nb[3] == t->nb[2] * t->ne[2];
}

// Only immutable model state may be served from the resident device copy.  ggml also hands the hook a
// mutable src0 -- a view of the K/V cache (-ctk q8_0 ...), an F32 activation or KV tensor as src0, a
// trainable weight -- whose address and shape do not change between calls: those are uploaded every call.
bool is_leaf_weight(const ggml_tensor *t) {
  return t->op == 0 /* GGML_OP_NONE */ && t->view_src == nullptr && t->grad == nullptr &&
         (t->flags & (1 | 4)) == 0 /* not GGML_TENSOR_FLAG_INPUT / _PARAM (ggml.h:532-534) */;
}

[[noreturn]] void die(const char *what) {
  std::fprintf(stderr, "lamm_mul_mat: %s: %s\n", what, lammcu_last_error());
  std::abort();
}

}  // namespace

extern "C" {

int lamm_get_opt_level(void) { return opt_level(); }

void lamm_set_writeback_wdata(int enable) { g_writeback.store(enable); }

bool lamm_can_mul_mat(const struct ggml_compute_params *params, const struct ggml_tensor *dst) {
  if (opt_level() == 0) return false;                                   // :12
  if (params->type != GGML_TASK_TYPE_COMPUTE) return false;             // :15
  const ggml_tensor *src0 = dst->src[0], *src1 = dst->src[1];
  if (!src0 || !src1) return false;
  if (!lammcu_type_supported(src0->type)) return false;                 // :37-59 (never asserts)
  const int vec_dot = lammcu_vec_dot_type(src0->type);
  if ((int)src1->type == vec_dot && !is_contiguous(src1)) return false; // :26
  if (src1->nb[0] != lammcu_type_size(src1->type)) return false;        // :29
  if (dst->type != GGML_TYPE_F32) return false;                         // :34
  // src1 must be F32 (quantized here) or already in the vec-dot type
  if (src1->type != GGML_TYPE_F32 && (int)src1->type != vec_dot) return false;
  // --- what the CUDA path additionally needs ---
  const int64_t K = src0->ne[0];
  const int kmul = (src0->type == GGML_TYPE_F32 || src0->type == GGML_TYPE_F16) ? 8 : lammcu_blck_size(src0->type);
  if (K <= 0 || K % kmul) return false;                                 // lamm_impl.hpp:123-126 aborts instead
  if (src0->nb[0] != lammcu_type_size(src0->type)) return false;        // rows are block-contiguous
  if (src0->nb[1] < lammcu_row_size(src0->type, K)) return false;
  if (src0->ne[2] <= 0 || src0->ne[3] <= 0 || src1->ne[2] % src0->ne[2] || src1->ne[3] % src0->ne[3])
    return false;
  if (dst->nb[0] != sizeof(float)) return false;
  return device_ok();                                                   // no GPU: stock ggml path
}

void lamm_mul_mat(const struct ggml_compute_params *params, struct ggml_tensor *dst) {
  if (params->ith != 0) return;            // one launch per node; rows are not thread-split
  const ggml_tensor *src0 = dst->src[0], *src1 = dst->src[1];
  // GGML_TENSOR_BINARY_OP_LOCALS (ggml.h:306-312), the names the reference uses
  const int64_t ne00 = src0->ne[0], ne01 = src0->ne[1], ne02 = src0->ne[2], ne03 = src0->ne[3];
  const size_t nb01 = src0->nb[1], nb02 = src0->nb[2], nb03 = src0->nb[3];
  const int64_t ne11 = src1->ne[1], ne12 = src1->ne[2], ne13 = src1->ne[3];
  const size_t nb11 = src1->nb[1], nb12 = src1->nb[2], nb13 = src1->nb[3];
  const size_t nb1 = dst->nb[1], nb2 = dst->nb[2], nb3 = dst->nb[3];

  const int64_t r2 = ne12 / ne02, r3 = ne13 / ne03;                     // :75-76 broadcast factors
  const int vec_dot = lammcu_vec_dot_type(src0->type);
  const size_t row_size = lammcu_row_size(vec_dot, ne00);               // :80
  const bool src1_quantized = ((int)src1->type == vec_dot) && src0->type != GGML_TYPE_F32;
  const int64_t ldc = (int64_t)(nb1 / sizeof(float));                   // :99

  for (int64_t i13 = 0; i13 < ne13; i13++) {                            // :130-142
    for (int64_t i12 = 0; i12 < ne12; i12++) {
      const char *a = (const char *)src0->data + (i12 / r2) * nb02 + (i13 / r3) * nb03;
      float *c = (float *)((char *)dst->data + i12 * nb2 + i13 * nb3);
      lammcu_weight_t w;
      if (lammcu_weight_for_op(a, src0->type, ne01, ne00, (int64_t)nb01, is_leaf_weight(src0) ? 0 : 1, &w) != LAMMCU_OK)
        die("weight upload");
      const char *b = (const char *)src1->data + i12 * nb12 + i13 * nb13;
      if (src1_quantized) {
        if (lammcu_mul_mat_q_host(w, b, (int64_t)nb11, ne11, c, ldc) != LAMMCU_OK) die("mul_mat_q");
      } else {
        void *xq_out = nullptr;
        if (g_writeback.load() && params->wdata && src0->type != GGML_TYPE_F32)
          xq_out = (char *)params->wdata + (i12 * ne11 + i13 * ne12 * ne11) * row_size;   // :137-138
        if (lammcu_mul_mat_host(w, (const float *)b, ne11, (int64_t)(nb11 / sizeof(float)), c, ldc,
                                xq_out) != LAMMCU_OK)
          die("mul_mat");
      }
    }
  }
}

// ---- GGML_OP_MUL_MAT_ID (no counterpart in the reference) --------------------------

int lammcu_mul_mat_id_plan(const int32_t *ids, int64_t ids_stride, int id, int64_t n_tokens, int n_as,
                           int64_t *counts, int64_t *order) {
  if (!ids || !counts || !order || n_tokens < 0 || n_as < 1 || id < 0 || ids_stride <= id) return LAMMCU_ERR_ARG;
  for (int e = 0; e < n_as; e++) counts[e] = 0;
  for (int64_t j = 0; j < n_tokens; j++) {                 // ggml.c:11070-11077
    const int32_t e = ids[j * ids_stride + id];
    if (e < 0 || e >= n_as) return LAMMCU_ERR_ARG;
    counts[e]++;
  }
  std::vector<int64_t> at((size_t)n_as);
  int64_t run = 0;
  for (int e = 0; e < n_as; e++) { at[(size_t)e] = run;  run += counts[e]; }
  for (int64_t j = 0; j < n_tokens; j++) order[at[(size_t)ids[j * ids_stride + id]]++] = j;
  return LAMMCU_OK;
}

bool lamm_can_mul_mat_id(const struct ggml_compute_params *params, const struct ggml_tensor *dst) {
  if (opt_level() == 0) return false;
  if (params->type != GGML_TASK_TYPE_COMPUTE) return false;
  const ggml_tensor *ids = dst->src[0], *src1 = dst->src[1];
  const int id = dst->op_params[0], n_as = dst->op_params[1];
  if (!ids || !src1 || n_as < 1 || n_as > GGML_MAX_SRC - 2) return false;
  if ((int)ids->type != 26 /* GGML_TYPE_I32 */ || ids->nb[0] != 4 || ids->nb[1] % 4) return false;
  if (id < 0 || id >= ids->ne[0] || ids->ne[1] != src1->ne[1]) return false;
  if (src1->type != GGML_TYPE_F32 || src1->nb[0] != 4 || src1->nb[1] % 4) return false;   // quantized here
  if (src1->ne[2] != 1 || src1->ne[3] != 1 || ids->ne[2] != 1 || ids->ne[3] != 1) return false;
  if (dst->type != GGML_TYPE_F32 || dst->nb[0] != sizeof(float) || dst->nb[1] % 4) return false;
  const ggml_tensor *a0 = dst->src[2];
  if (!a0 || !lammcu_type_supported(a0->type)) return false;
  const int64_t K = a0->ne[0];
  const int kmul = (a0->type == GGML_TYPE_F32 || a0->type == GGML_TYPE_F16) ? 8 : lammcu_blck_size(a0->type);
  if (K <= 0 || K % kmul || K != src1->ne[0]) return false;
  for (int e = 0; e < n_as; e++) {
    const ggml_tensor *a = dst->src[2 + e];
    if (!a || a->type != a0->type || a->ne[0] != K || a->ne[1] != a0->ne[1] || a->ne[2] != 1 || a->ne[3] != 1)
      return false;
    if (a->nb[0] != lammcu_type_size(a->type) || a->nb[1] < lammcu_row_size(a->type, K)) return false;
  }
  if (dst->ne[0] != a0->ne[1] || dst->ne[1] != src1->ne[1]) return false;
  return device_ok();
}

void lamm_mul_mat_id(const struct ggml_compute_params *params, struct ggml_tensor *dst) {
  if (params->ith != 0) return;
  const ggml_tensor *ids = dst->src[0], *src1 = dst->src[1], *a0 = dst->src[2];
  const int id = dst->op_params[0], n_as = dst->op_params[1];
  const int64_t N = src1->ne[1], K = a0->ne[0], M = a0->ne[1];
  const int64_t ldx = (int64_t)(src1->nb[1] / sizeof(float)), ldc = (int64_t)(dst->nb[1] / sizeof(float));
  std::vector<int64_t> counts((size_t)n_as), order((size_t)(N > 0 ? N : 1));
  if (lammcu_mul_mat_id_plan((const int32_t *)ids->data, (int64_t)(ids->nb[1] / 4), id, N, n_as, counts.data(),
                             order.data()) != LAMMCU_OK) {
    std::fprintf(stderr, "lamm_mul_mat_id: expert id outside [0, %d)\n", n_as);      // ggml.c:11074 asserts
    std::abort();
  }
  const float *x = (const float *)src1->data;
  float *c = (float *)dst->data;
  std::vector<float> xg, cg;
  int64_t at = 0;
  for (int e = 0; e < n_as; e++) {
    const int64_t cnt = counts[(size_t)e];
    if (!cnt) continue;
    const ggml_tensor *a = dst->src[2 + e];
    lammcu_weight_t w;
    if (lammcu_weight_for_op(a->data, a->type, M, K, (int64_t)a->nb[1], is_leaf_weight(a) ? 0 : 1, &w) != LAMMCU_OK)
      die("weight upload");
    const int64_t *rows = order.data() + at;
    at += cnt;
    // the expert's tokens are consecutive rows of src1 / dst: no gather
    if (rows[cnt - 1] - rows[0] == cnt - 1) {
      if (lammcu_mul_mat_host(w, x + rows[0] * ldx, cnt, ldx, c + rows[0] * ldc, ldc, nullptr) != LAMMCU_OK)
        die("mul_mat_id");
      continue;
    }
    xg.resize((size_t)(cnt * K));
    cg.resize((size_t)(cnt * M));
    for (int64_t r = 0; r < cnt; r++) std::memcpy(xg.data() + r * K, x + rows[r] * ldx, (size_t)K * 4);
    if (lammcu_mul_mat_host(w, xg.data(), cnt, K, cg.data(), M, nullptr) != LAMMCU_OK) die("mul_mat_id");
    for (int64_t r = 0; r < cnt; r++) std::memcpy(c + rows[r] * ldc, cg.data() + r * M, (size_t)M * 4);
  }
}

}  // extern "C"
