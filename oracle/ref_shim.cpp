// ref_shim.cpp -- thin C entry points over the UNMODIFIED reference, compiled
// from /root/reference where it lies into oracle/_ref/libref.so by
// oracle/Makefile.  TEST INFRASTRUCTURE ONLY (checker + timed CPU baseline).
//
// Nothing here restates reference code: it only calls the reference's public
// functions (ggml.h, ggml-quants.h, loongarch_matmul.h) so tests can compare
// the oracle restatement and the CUDA path against the real thing:
//   - scalar activation quantizers  quantize_row_q8_{0,1,K}_reference
//   - weight quantizers             ggml_quantize_chunk
//   - the operator itself           lamm_can_mul_mat / lamm_mul_mat, both called
//                                   directly and through ggml_graph_compute
//                                   (the hook at ggml.c:10858-10863)
//   - ggml's own vec_dot            second opinion
//   - a shape-configurable timing loop following src/la-benchmark-matmult.cpp
//     :322-390 (the reference binary hard-codes K=11008 M=4096 N=128).
#include "ggml.h"
#include "ggml-quants.h"
#include "loongarch_matmul.h"

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

extern "C" {

static bool g_inited = false;

// ggml_init fills ggml_table_f32_f16 (needed by GGML_FP16_TO_FP32).
void ref_init(void) {
  if (g_inited) return;
  struct ggml_init_params ip = {1024 * 1024, nullptr, false};
  struct ggml_context *ctx = ggml_init(ip);
  ggml_free(ctx);
  g_inited = true;
}

int ref_opt_level(void) { return lamm_get_opt_level(); }

int ref_vec_dot_type(int type) {
  return (int)ggml_internal_get_type_traits((enum ggml_type)type).vec_dot_type;
}
size_t ref_row_size(int type, int64_t k) { return ggml_row_size((enum ggml_type)type, k); }
size_t ref_type_size(int type) { return ggml_type_size((enum ggml_type)type); }
int ref_blck_size(int type) { return ggml_blck_size((enum ggml_type)type); }

// scalar "_reference" quantizers: the bit-exact target
int ref_quantize_act(int vec_dot_type, const float *x, int64_t ldx, void *y,
                     int64_t n_rows, int64_t k) {
  ref_init();
  const size_t rs = ggml_row_size((enum ggml_type)vec_dot_type, k);
  for (int64_t r = 0; r < n_rows; r++) {
    const float *xr = x + r * ldx;
    char *yr = (char *)y + r * rs;
    switch (vec_dot_type) {
    case GGML_TYPE_Q8_0: quantize_row_q8_0_reference(xr, (block_q8_0 *)yr, (int)k); break;
    case GGML_TYPE_Q8_1: quantize_row_q8_1_reference(xr, (block_q8_1 *)yr, (int)k); break;
    case GGML_TYPE_Q8_K: quantize_row_q8_K_reference(xr, (block_q8_K *)yr, (int)k); break;
    default: return -1;
    }
  }
  return 0;
}

// what ggml's INIT phase really runs on this host (AVX2 variant on x86)
int ref_quantize_act_native(int vec_dot_type, const float *x, int64_t ldx,
                            void *y, int64_t n_rows, int64_t k) {
  ref_init();
  ggml_type_traits_t tt = ggml_internal_get_type_traits((enum ggml_type)vec_dot_type);
  if (!tt.from_float) return -1;
  const size_t rs = ggml_row_size((enum ggml_type)vec_dot_type, k);
  for (int64_t r = 0; r < n_rows; r++)
    tt.from_float(x + r * ldx, (char *)y + r * rs, (int)k);
  return 0;
}

int ref_quantize_weights(int type, const float *x, void *y, int64_t n_rows, int64_t k) {
  ref_init();
  if (type == GGML_TYPE_F32) { memcpy(y, x, (size_t)(n_rows * k) * 4); return 0; }
  ggml_quantize_chunk((enum ggml_type)type, x, y, 0, (int)n_rows, (int)k, nullptr);
  return 0;
}

int ref_dequantize_row(int type, const void *x, float *y, int64_t k) {
  ref_init();
  ggml_type_traits_t tt = ggml_internal_get_type_traits((enum ggml_type)type);
  if (!tt.to_float) return -1;
  tt.to_float(x, y, (int)k);
  return 0;
}

// ggml's own dot product for one (W row, Xq row) pair
float ref_vec_dot(int w_type, const void *wrow, const void *xqrow, int64_t k) {
  ref_init();
  ggml_type_traits_t tt = ggml_internal_get_type_traits((enum ggml_type)w_type);
  float s = 0.0f;
  tt.vec_dot((int)k, &s, 0, wrow, 0, xqrow, 0, 1);
  return s;
}

static void make_2d(struct ggml_tensor *t, enum ggml_type type, int64_t ne0,
                    int64_t ne1, size_t nb1, void *data) {
  memset(t, 0, sizeof(*t));
  t->type = type;
  t->ne[0] = ne0; t->ne[1] = ne1; t->ne[2] = 1; t->ne[3] = 1;
  t->nb[0] = ggml_type_size(type);
  t->nb[1] = nb1;
  t->nb[2] = t->nb[1] * ne1;
  t->nb[3] = t->nb[2];
  t->data = data;
}

// Direct call of the reference operator on caller-provided, already quantized
// activations (exactly what the COMPUTE phase sees): src1 is F32 by type,
// wdata holds the packed vec-dot rows.  nth threads are run one after another
// (the operator is stateless; rows are disjoint), nth must divide M (9-3).
// Returns 0, or 1 if lamm_can_mul_mat said no.
int ref_lamm_mul_mat(int w_type, const void *w, int64_t w_stride,
                     const float *x_f32, const void *xq, float *c, int64_t ldc,
                     int64_t M, int64_t N, int64_t K, int nth) {
  ref_init();
  struct ggml_tensor src0, src1, dst;
  make_2d(&src0, (enum ggml_type)w_type, K, M, (size_t)w_stride, (void *)w);
  make_2d(&src1, GGML_TYPE_F32, K, N, (size_t)K * 4, (void *)x_f32);
  make_2d(&dst, GGML_TYPE_F32, M, N, (size_t)ldc * 4, c);
  dst.op = GGML_OP_MUL_MAT;
  dst.src[0] = &src0;
  dst.src[1] = &src1;
  const enum ggml_type vt = ggml_internal_get_type_traits((enum ggml_type)w_type).vec_dot_type;
  struct ggml_compute_params p;
  p.type = GGML_TASK_TYPE_COMPUTE;
  p.nth = nth;
  p.wsize = ggml_row_size(vt, K) * (size_t)N;
  p.wdata = (void *)xq;
  for (int ith = 0; ith < nth; ith++) {
    p.ith = ith;
    if (!lamm_can_mul_mat(&p, &dst)) return 1;
    lamm_mul_mat(&p, &dst);
  }
  return 0;
}

struct ref_graph {
  struct ggml_context *ctx;
  struct ggml_cgraph *g;
  struct ggml_tensor *w, *x, *c;
  std::vector<uint8_t> work;
};

static ref_graph *graph_new(int w_type, const void *w, const float *x, int64_t M,
                            int64_t N, int64_t K) {
  ref_init();
  ref_graph *rg = new ref_graph();
  size_t ctx_size = ggml_row_size((enum ggml_type)w_type, K) * (size_t)M +
                    (size_t)K * N * 4 + (size_t)M * N * 4 + (16u << 20);
  struct ggml_init_params ip = {ctx_size, nullptr, false};
  rg->ctx = ggml_init(ip);
  if (!rg->ctx) { delete rg; return nullptr; }
  rg->w = ggml_new_tensor_2d(rg->ctx, (enum ggml_type)w_type, K, M);
  rg->x = ggml_new_tensor_2d(rg->ctx, GGML_TYPE_F32, K, N);
  if (w) memcpy(rg->w->data, w, ggml_nbytes(rg->w));
  if (x) memcpy(rg->x->data, x, ggml_nbytes(rg->x));
  rg->c = ggml_mul_mat(rg->ctx, rg->w, rg->x);
  rg->g = ggml_new_graph(rg->ctx);
  ggml_build_forward_expand(rg->g, rg->c);
  return rg;
}

// When this shim is linked against liblammcu.so instead of the reference's loongarch_matmul.o (make ref_b200:
// the real ggml.c -DLA_LLAMA driving the B200 operator through its own hook), every graph copies its weights
// into a fresh ggml context, so a freed context's addresses come back with other bytes: tell the operator's
// weight registry when a "model" goes away, as an application does in llama_free_model.  Absent (weak, NULL)
// in the plain reference build.
extern int lammcu_weight_forget_all(void) __attribute__((weak));
extern void lamm_set_writeback_wdata(int) __attribute__((weak));
static void model_freed(void) { if (lammcu_weight_forget_all) lammcu_weight_forget_all(); }
int ref_is_b200(void) { return lammcu_weight_forget_all ? 1 : 0; }

static void graph_compute(ref_graph *rg, int nth) {
  struct ggml_cplan plan = ggml_graph_plan(rg->g, nth);
  if (plan.work_size > 0) {
    rg->work.resize(plan.work_size);
    plan.work_data = rg->work.data();
  }
  ggml_graph_compute(rg->g, &plan);
}

// The reference's whole path: ggml_mul_mat node -> ggml_graph_compute ->
// INIT quantize (ggml.c:10865-10887) -> hook -> lamm_mul_mat.  W contiguous.
int ref_graph_mul_mat(int w_type, const void *w, const float *x, float *c,
                      int64_t M, int64_t N, int64_t K, int nth) {
  ref_graph *rg = graph_new(w_type, w, x, M, N, K);
  if (!rg) return -1;
  graph_compute(rg, nth);
  memcpy(c, rg->c->data, (size_t)M * N * 4);
  model_freed();
  ggml_free(rg->ctx);
  delete rg;
  return 0;
}

// The same graph, also returning what is in the plan's work buffer afterwards (params->wdata: the quantized
// activations of ggml.c:10865-10887, packed at row_size stride).  With the B200 operator and its write-back
// hook enabled these are the GPU's Xq bytes; `reps` > 1 recomputes the same graph (resident weights).
int ref_graph_mul_mat_wdata(int w_type, const void *w, const float *x, float *c, void *wdata_out, size_t wdata_cap,
                            int64_t M, int64_t N, int64_t K, int nth, int reps, int writeback) {
  ref_graph *rg = graph_new(w_type, w, x, M, N, K);
  if (!rg) return -1;
  if (lamm_set_writeback_wdata) lamm_set_writeback_wdata(writeback);
  for (int r = 0; r < (reps < 1 ? 1 : reps); r++) graph_compute(rg, nth);
  if (lamm_set_writeback_wdata) lamm_set_writeback_wdata(0);
  memcpy(c, rg->c->data, (size_t)M * N * 4);
  if (wdata_out) memcpy(wdata_out, rg->work.data(), rg->work.size() < wdata_cap ? rg->work.size() : wdata_cap);
  model_freed();
  ggml_free(rg->ctx);
  delete rg;
  return 0;
}

// ggml_mul_mat_id (ggml.c:4616-4656, forward ggml.c:10992-11170): n_as expert matrices of one
// type/shape packed back to back in `ws`, ids = I32 [n_ids, N] (row per token), the node takes
// column `id`.  SURVEY 8f-2: the other caller of the same inner product (lamm's hook is not in it).
int ref_graph_mul_mat_id(int w_type, const void *ws, int n_as, const int32_t *ids, int n_ids, int id,
                         const float *x, float *c, int64_t M, int64_t N, int64_t K, int nth) {
  ref_init();
  if (n_as < 1 || n_as > GGML_MAX_SRC - 2 || id < 0 || id >= n_ids) return -2;
  const size_t wbytes = ggml_row_size((enum ggml_type)w_type, K) * (size_t)M;
  size_t ctx_size = wbytes * (size_t)n_as + (size_t)K * N * 4 + (size_t)M * N * 4 +
                    (size_t)n_ids * N * 4 + (16u << 20);
  struct ggml_init_params ip = {ctx_size, nullptr, false};
  struct ggml_context *ctx = ggml_init(ip);
  if (!ctx) return -1;
  struct ggml_tensor *as[GGML_MAX_SRC];
  for (int e = 0; e < n_as; e++) {
    as[e] = ggml_new_tensor_2d(ctx, (enum ggml_type)w_type, K, M);
    memcpy(as[e]->data, (const char *)ws + (size_t)e * wbytes, wbytes);
  }
  struct ggml_tensor *tid = ggml_new_tensor_2d(ctx, GGML_TYPE_I32, n_ids, N);
  memcpy(tid->data, ids, (size_t)n_ids * N * 4);
  struct ggml_tensor *tx = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, K, N);
  memcpy(tx->data, x, (size_t)K * N * 4);
  struct ggml_tensor *tc = ggml_mul_mat_id(ctx, as, n_as, tid, id, tx);
  struct ggml_cgraph *g = ggml_new_graph(ctx);
  ggml_build_forward_expand(g, tc);
  struct ggml_cplan plan = ggml_graph_plan(g, nth);
  std::vector<uint8_t> work(plan.work_size + 16);
  plan.work_data = work.data();
  ggml_graph_compute(g, &plan);
  memcpy(c, tc->data, (size_t)M * N * 4);
  model_freed();
  ggml_free(ctx);
  return 0;
}

// the ops either side of the path as llama.cpp's graph builds them.
//   kind 1: llm_build_norm (llama.cpp:5171-5201): ggml_rms_norm(x, eps) then ggml_mul(., norm_w)   [aux = norm_w, K floats]
//   kind 2: llm_build_ffn's SiLU-gated branch (llama.cpp:5203-5290): ggml_silu(gate) then ggml_mul(., up)
//           [x = gate, aux = up, both N x K]
// followed by ggml_mul_mat(w, .).  y_out (may be NULL) receives the f32 activation the matmul consumed.
int ref_graph_xop_mul_mat(int kind, int w_type, const void *w, const float *x, const float *aux, float eps, float *c,
                          float *y_out, int64_t M, int64_t N, int64_t K, int nth) {
  ref_init();
  if (kind != 1 && kind != 2) return -2;
  const size_t wbytes = ggml_row_size((enum ggml_type)w_type, K) * (size_t)M;
  size_t ctx_size = wbytes + (size_t)K * N * 4 * 5 + (size_t)M * N * 4 + (16u << 20);
  struct ggml_init_params ip = {ctx_size, nullptr, false};
  struct ggml_context *ctx = ggml_init(ip);
  if (!ctx) return -1;
  struct ggml_tensor *tw = ggml_new_tensor_2d(ctx, (enum ggml_type)w_type, K, M);
  memcpy(tw->data, w, wbytes);
  struct ggml_tensor *tx = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, K, N);
  memcpy(tx->data, x, (size_t)K * N * 4);
  struct ggml_tensor *ty;
  if (kind == 1) {
    ty = ggml_rms_norm(ctx, tx, eps);
    if (aux) {
      struct ggml_tensor *tn = ggml_new_tensor_1d(ctx, GGML_TYPE_F32, K);
      memcpy(tn->data, aux, (size_t)K * 4);
      ty = ggml_mul(ctx, ty, tn);
    }
  } else {
    struct ggml_tensor *tu = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, K, N);
    memcpy(tu->data, aux, (size_t)K * N * 4);
    ty = ggml_mul(ctx, ggml_silu(ctx, tx), tu);
  }
  struct ggml_tensor *tc = ggml_mul_mat(ctx, tw, ty);
  struct ggml_cgraph *g = ggml_new_graph(ctx);
  ggml_build_forward_expand(g, tc);
  struct ggml_cplan plan = ggml_graph_plan(g, nth);
  std::vector<uint8_t> work(plan.work_size + 16);
  plan.work_data = work.data();
  ggml_graph_compute(g, &plan);
  memcpy(c, tc->data, (size_t)M * N * 4);
  if (y_out) memcpy(y_out, ty->data, (size_t)K * N * 4);
  model_freed();
  ggml_free(ctx);
  return 0;
}

// A GGUF file written by the reference's own writer (gguf_add_tensor + gguf_write_to_file,
// ggml.c:21160-21420), with key/values of every scalar kind, a string array and a numeric array
// ahead of the tensors: the fixture our reader (csrc/lamm_gguf.cpp) is pinned against.
// ne: [n][2] (ne1 == 0: a 1-D tensor); data[i]: the tensor's bytes; alignment 0 = default (32).
int ref_gguf_write(const char *path, uint32_t alignment, int n, const char *const *names, const int *types,
                   const int64_t *ne, const void *const *data) {
  ref_init();
  size_t total = 1u << 20;
  for (int i = 0; i < n; i++)
    total += ggml_row_size((enum ggml_type)types[i], ne[2 * i]) * (size_t)(ne[2 * i + 1] ? ne[2 * i + 1] : 1) + 1024;
  struct ggml_init_params ip = {total, nullptr, false};
  struct ggml_context *ctx = ggml_init(ip);
  if (!ctx) return -1;
  struct gguf_context *g = gguf_init_empty();
  gguf_set_val_str(g, "general.architecture", "llama");
  if (alignment) gguf_set_val_u32(g, "general.alignment", alignment);
  gguf_set_val_u32(g, "llama.block_count", 32);
  gguf_set_val_f32(g, "llama.rope.freq_base", 10000.5f);
  gguf_set_val_bool(g, "test.flag", true);
  gguf_set_val_u64(g, "test.u64", 0x123456789abcdefull);
  gguf_set_val_f64(g, "test.f64", 2.5);
  gguf_set_val_i8(g, "test.i8", -3);
  gguf_set_val_u16(g, "test.u16", 65535);
  const int32_t arr[5] = {1, -2, 3, -4, 5};
  gguf_set_arr_data(g, "test.arr_i32", GGUF_TYPE_INT32, arr, 5);
  const char *toks[4] = {"<s>", "</s>", "", "hello world"};
  gguf_set_arr_str(g, "tokenizer.ggml.tokens", toks, 4);
  gguf_set_val_str(g, "general.name", "lamm gguf fixture");
  for (int i = 0; i < n; i++) {
    struct ggml_tensor *t = ne[2 * i + 1]
        ? ggml_new_tensor_2d(ctx, (enum ggml_type)types[i], ne[2 * i], ne[2 * i + 1])
        : ggml_new_tensor_1d(ctx, (enum ggml_type)types[i], ne[2 * i]);
    ggml_set_name(t, names[i]);
    memcpy(t->data, data[i], ggml_nbytes(t));
    gguf_add_tensor(g, t);
  }
  gguf_write_to_file(g, path, false);
  gguf_free(g);
  ggml_free(ctx);
  return 0;
}

// Timing loop after src/la-benchmark-matmult.cpp:322-390: each iteration times
// one full ggml_graph_compute (plan + INIT quantize + COMPUTE) of graph 1,
// then runs a second, untimed graph on other weights to disturb the caches.
// usec_out[iters] receives per-iteration microseconds.  w may be NULL
// (weights are then quantized from a constant fill like the reference's).
int ref_bench_mul_mat(int w_type, const void *w, const float *x, int64_t M,
                      int64_t N, int64_t K, int nth, int iters, double *usec_out) {
  std::vector<float> xf;
  if (!x) { xf.assign((size_t)K * N, 2.0f); x = xf.data(); }
  std::vector<uint8_t> wq;
  if (!w) {
    std::vector<float> wf((size_t)K * M, 1.0f);
    wq.resize(ggml_row_size((enum ggml_type)w_type, K) * (size_t)M);
    ref_quantize_weights(w_type, wf.data(), wq.data(), M, K);
    w = wq.data();
  }
  ref_graph *g1 = graph_new(w_type, w, x, M, N, K);
  ref_graph *g2 = graph_new(w_type, w, x, M, N, K);
  if (!g1 || !g2) return -1;
  for (int i = 0; i < iters; i++) {
    int64_t t0 = ggml_time_us();
    graph_compute(g1, nth);
    usec_out[i] = (double)(ggml_time_us() - t0);
    graph_compute(g2, nth);
  }
  model_freed();
  ggml_free(g1->ctx); ggml_free(g2->ctx);
  delete g1; delete g2;
  return 0;
}

} // extern "C"
