/*
 * lamm_oracle.h -- CPU oracle for the LAMM quantized mul_mat hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it, and only as the checker.  The product
 * path (la-llama.cpp_b200/) never links or calls it.
 *
 * It is a plain-C restatement (not a copy) of the reference algorithm:
 *   - block layouts ............ llama.cpp-b2430/ggml-common.h:144-225,316-321
 *   - fp16 <-> fp32 ............ llama.cpp-b2430/ggml-impl.h:179-226 (IEEE, RNE)
 *   - activation quantizers .... llama.cpp-b2430/ggml-quants.c:1182-1205 (q8_0),
 *                                :1396-1429 (q8_1), :3981-4018 (q8_K)
 *   - weight quantizers ........ ggml-quants.c:1002,1044,1085,1133 (q4_0..q5_1),
 *                                :1182 (q8_0)   [synthetic tensor generation]
 *   - per-format dot formulas .. src/lamm_kernel_{f32,q2_k,q4_0,q4_1,q5_0,q5_1,
 *                                q8_0}.hpp (naive kernels)
 *   - operator driver .......... src/loongarch_matmul.cpp:64-143
 *
 * Parity pin: checked against the real reference compiled into
 * oracle/_ref/libref.so (tests/test_oracle_vs_ref.py) and against the golden
 * fixtures under tests/golden/ that were generated from that library
 * (oracle/gen_golden.py).
 */
#ifndef LAMM_ORACLE_H
#define LAMM_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ggml type ids (llama.cpp-b2430/ggml.h:341-369) */
enum {
  ORC_TYPE_F32 = 0,
  ORC_TYPE_Q4_0 = 2,
  ORC_TYPE_Q4_1 = 3,
  ORC_TYPE_Q5_0 = 6,
  ORC_TYPE_Q5_1 = 7,
  ORC_TYPE_Q8_0 = 8,
  ORC_TYPE_Q8_1 = 9,
  ORC_TYPE_Q2_K = 10,
  ORC_TYPE_Q8_K = 15,
  /* SURVEY 8f-1 K-quants: outside the reference's lamm table, so the behaviour
   * restated is ggml's own ggml_vec_dot_q{3,4,5,6}_K_q8_K (scalar forms,
   * ggml-quants.c:6583-6646, 7301-7358, 7968-8029, 8695-8737)              */
  ORC_TYPE_Q3_K = 11,
  ORC_TYPE_Q4_K = 12,
  ORC_TYPE_Q5_K = 13,
  ORC_TYPE_Q6_K = 14,
  /* SURVEY 8f-2 (oracle only so far; no CUDA kernel yet): F16 weights.  Not in the reference's
   * lamm table either: ggml converts src1 with ggml_fp32_to_fp16_row (ggml.c:354-372, RNE) and
   * runs ggml_vec_dot_f16 (ggml.c:1589-1630: f32 products of the two halves, f32 SIMD partial
   * sums)                                                                               */
  ORC_TYPE_F16 = 1
};

#pragma pack(push, 1)
typedef struct { uint16_t d; uint8_t qs[16]; } orc_q4_0;                  /* 18 B / 32 */
typedef struct { uint16_t d, m; uint8_t qs[16]; } orc_q4_1;               /* 20 B */
typedef struct { uint16_t d; uint8_t qh[4]; uint8_t qs[16]; } orc_q5_0;   /* 22 B */
typedef struct { uint16_t d, m; uint8_t qh[4]; uint8_t qs[16]; } orc_q5_1;/* 24 B */
typedef struct { uint16_t d; int8_t qs[32]; } orc_q8_0;                   /* 34 B */
typedef struct { uint16_t d, s; int8_t qs[32]; } orc_q8_1;                /* 36 B */
typedef struct { uint8_t scales[16]; uint8_t qs[64]; uint16_t d, dmin; } orc_q2_K; /* 84 B / 256 */
typedef struct { float d; int8_t qs[256]; int16_t bsums[16]; } orc_q8_K;  /* 292 B / 256 */
/* ggml-common.h:228-313 */
typedef struct { uint8_t hmask[32]; uint8_t qs[64]; uint8_t scales[12]; uint16_t d; } orc_q3_K;          /* 110 B */
typedef struct { uint16_t d, dmin; uint8_t scales[12]; uint8_t qs[128]; } orc_q4_K;                      /* 144 B */
typedef struct { uint16_t d, dmin; uint8_t scales[12]; uint8_t qh[32]; uint8_t qs[128]; } orc_q5_K;      /* 176 B */
typedef struct { uint8_t ql[128]; uint8_t qh[64]; int8_t scales[16]; uint16_t d; } orc_q6_K;             /* 210 B */
#pragma pack(pop)

/* ---- type table ------------------------------------------------------- */
int    orc_blck_size(int type);          /* elements per block, 0 if unknown  */
size_t orc_type_size(int type);          /* bytes per block                   */
int    orc_vec_dot_type(int type);       /* activation type paired with type  */
size_t orc_row_size(int type, int64_t k);/* bytes of one row of k elements    */

/* ---- fp16 ------------------------------------------------------------- */
float    orc_fp16_to_fp32(uint16_t h);
uint16_t orc_fp32_to_fp16(float f);

/* ---- activation quantizers (bit-exact targets) ------------------------ */
void orc_quantize_row_q8_0(const float *x, void *y, int64_t k);
void orc_quantize_row_q8_1(const float *x, void *y, int64_t k);
/* bsums of an all-zero 256-block are left untouched, as in the reference */
void orc_quantize_row_q8_K(const float *x, void *y, int64_t k);
/* dispatch on the vec-dot type; returns 0 on success */
int  orc_quantize_act(int vec_dot_type, const float *x, int64_t ldx, void *y,
                      int64_t n_rows, int64_t k);

/* ---- weight quantizers (synthetic tensors) ---------------------------- */
/* returns 0 on success; Q2_K has no restated quantizer (use orc_random_q2_K
 * or the reference's own through oracle/_ref). */
int  orc_quantize_weights(int type, const float *x, void *y, int64_t n_rows,
                          int64_t k);
void orc_random_q2_K(void *y, int64_t n_blocks, uint64_t seed);
/* same for any super-block weight type (Q2_K .. Q6_K): uniform random bit
 * patterns in every quant / scale field, d and dmin ~ fp16(U(0.001, 0.1))   */
int  orc_random_kquant(int type, void *y, int64_t n_blocks, uint64_t seed);
/* dequantize one row of k elements to f32 (ground truth helper) */
int  orc_dequantize_row(int type, const void *x, float *y, int64_t k);

/* ---- the operator ------------------------------------------------------ */
/* C[j*ldc + i] = dot(W row i, Xq row j) with the per-format kernel formula,
 * every block term evaluated in double.  W: M rows, stride w_stride bytes.
 * Xq: N rows of the vec-dot type, stride xq_stride bytes.                   */
int orc_mul_mat_q(int w_type, const void *w, int64_t w_stride, const void *xq,
                  int64_t xq_stride, float *c, int64_t ldc, int64_t M,
                  int64_t N, int64_t K);
/* same formula, f32 running sum in the reference's naive order
 * (src/lamm_kernel_*.hpp lamm_naive_kernel, with 9-1 fixed for Q5_0) */
int orc_mul_mat_q_f32(int w_type, const void *w, int64_t w_stride,
                      const void *xq, int64_t xq_stride, float *c, int64_t ldc,
                      int64_t M, int64_t N, int64_t K);
/* full path: quantize X (f32, N rows, ldx elements) then orc_mul_mat_q.
 * xq_out (optional) receives the quantized rows packed at row_size stride.  */
int orc_mul_mat(int w_type, const void *w, int64_t w_stride, const float *x,
                int64_t ldx, float *c, int64_t ldc, int64_t M, int64_t N,
                int64_t K, void *xq_out);
/* SURVEY 8f-2 (oracle only so far): ggml_mul_mat_id -- token j is multiplied by expert matrix
 * ws[ids[j * ids_stride + id]] (ggml.c:4616-4656; forward ggml.c:10992-11170 groups the rows
 * per expert and runs the same vec_dot).  ws: n_as matrices of one type and shape.
 * Returns -4 when an id is outside [0, n_as).                                          */
int orc_mul_mat_id(int w_type, const void *const *ws, int n_as, int64_t w_stride,
                   const int32_t *ids, int64_t ids_stride, int id, const float *x,
                   int64_t ldx, float *c, int64_t ldc, int64_t M, int64_t N, int64_t K);
/* multi-threaded (pthreads over M) f32-order variant: the timed "port" CPU
 * baseline.  nth threads, rows split like src/lamm_impl.hpp:40-45 but with the
 * remainder rows given to the last thread (defect 9-3 not reproduced).      */
/* llm_build_norm (RMS norm, then the norm weight) and llm_build_ffn's silu(gate) * up, in f32 as ggml
 * computes them; the matmul that follows quantizes the result like any other activation (orc_mul_mat).             */
int orc_rms_norm_mul(const float *x, int64_t ldx, const float *norm_w, float eps, float *y, int64_t ldy,
                     int64_t n_rows, int64_t k);
float orc_silu(float x);
int orc_silu_mul(const float *gate, int64_t ldg, const float *up, int64_t ldu, float *y, int64_t ldy, int64_t n_rows,
                 int64_t k);
int orc_mul_mat_mt(int w_type, const void *w, int64_t w_stride, const float *x,
                   int64_t ldx, float *c, int64_t ldc, int64_t M, int64_t N,
                   int64_t K, int nth, void *xq_scratch);

#ifdef __cplusplus
}
#endif
#endif /* LAMM_ORACLE_H */
