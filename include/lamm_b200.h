/*
 * lamm_b200.h -- the drop-in operator boundary: the same three C symbols that
 * the reference exports from src/loongarch_matmul.h:18-24 and that ggml calls
 * at llama.cpp-b2430/ggml.c:10858-10863.
 *
 *   bool lamm_can_mul_mat(const struct ggml_compute_params*, const struct ggml_tensor* dst);
 *   void lamm_mul_mat    (const struct ggml_compute_params*, struct ggml_tensor* dst);
 *   int  lamm_get_opt_level(void);
 *
 * Implemented in la-llama.cpp_b200/csrc/lamm_host.cpp on top of the C-ABI of
 * lammcu.h.  A ggml build links liblammcu.so instead of loongarch_matmul.o and
 * keeps including the reference's own loongarch_matmul.h.
 *
 * When ggml.h is not on the include path (our tests, the Python binding) the
 * two argument structs are declared here with the b2430 layout
 * (llama.cpp-b2430/ggml.h:552-590 ggml_tensor, :668-677 ggml_compute_params,
 * :227-234 limits).  Define LAMM_HAVE_GGML_H to use the real header instead;
 * tests/test_abi_layout.py checks both agree field by field.
 */
#ifndef LAMM_B200_H
#define LAMM_B200_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef LAMM_HAVE_GGML_H
#include "ggml.h"
#else

#define GGML_MAX_DIMS 4
#define GGML_MAX_SRC 10
#define GGML_MAX_NAME 64
#define GGML_MAX_OP_PARAMS 64

#ifdef __cplusplus
extern "C" {
#endif

enum ggml_type {
  GGML_TYPE_F32 = 0, GGML_TYPE_F16 = 1, GGML_TYPE_Q4_0 = 2, GGML_TYPE_Q4_1 = 3,
  GGML_TYPE_Q5_0 = 6, GGML_TYPE_Q5_1 = 7, GGML_TYPE_Q8_0 = 8, GGML_TYPE_Q8_1 = 9,
  GGML_TYPE_Q2_K = 10, GGML_TYPE_Q3_K = 11, GGML_TYPE_Q4_K = 12,
  GGML_TYPE_Q5_K = 13, GGML_TYPE_Q6_K = 14, GGML_TYPE_Q8_K = 15,
  GGML_TYPE_COUNT = 27
};

enum ggml_task_type {
  GGML_TASK_TYPE_INIT = 0,
  GGML_TASK_TYPE_COMPUTE,
  GGML_TASK_TYPE_FINALIZE
};

struct ggml_compute_params {
  enum ggml_task_type type;
  int ith, nth;          /* thread index / thread count */
  size_t wsize;          /* work buffer shared by all threads */
  void *wdata;
};

struct ggml_tensor {
  enum ggml_type type;
  int backend;                         /* enum ggml_backend_type */
  void *buffer;                        /* struct ggml_backend_buffer* */
  int64_t ne[GGML_MAX_DIMS];           /* elements per dim */
  size_t nb[GGML_MAX_DIMS];            /* byte strides */
  int op;                              /* enum ggml_op */
  int32_t op_params[GGML_MAX_OP_PARAMS / sizeof(int32_t)];
  int32_t flags;
  struct ggml_tensor *grad;
  struct ggml_tensor *src[GGML_MAX_SRC];
  int perf_runs;
  int64_t perf_cycles;
  int64_t perf_time_us;
  struct ggml_tensor *view_src;
  size_t view_offs;
  void *data;
  char name[GGML_MAX_NAME];
  void *extra;
  char padding[8];
};

#ifdef __cplusplus
}
#endif
#endif /* LAMM_HAVE_GGML_H */

#ifdef __cplusplus
extern "C" {
#endif

/* src/loongarch_matmul.cpp:10-62.  True only in the COMPUTE phase, for dst F32
 * and (src0, vec_dot) in {F32/F32, Q2_K/Q8_K, Q4_0/Q8_0, Q4_1/Q8_1, Q5_0/Q8_0,
 * Q5_1/Q8_1, Q8_0/Q8_0}; additionally false (ggml falls through to its own
 * path) when there is no CUDA device, K is not a multiple of the block size,
 * or src0 rows are not block-contiguous.  Never aborts.                       */
bool lamm_can_mul_mat(const struct ggml_compute_params *params,
                      const struct ggml_tensor *dst);

/* src/loongarch_matmul.cpp:64-143.  Thread ith==0 runs the whole operator on
 * the GPU (leaf weights uploaded once per src0->data; a src0 that is a view, an
 * op result, an INPUT or a PARAM is uploaded on every call); other threads return at
 * once -- ggml's barrier after COMPUTE (ggml.c:18443-18450) publishes dst.
 * Activations come from src1->data (f32, quantized on the GPU) or, when src1
 * already has the vec-dot type, are uploaded as they are.  A CUDA failure
 * prints the message and aborts, like the reference's own error paths
 * (src/lamm_impl.hpp:100-103).                                                */
void lamm_mul_mat(const struct ggml_compute_params *params,
                  struct ggml_tensor *dst);

/* src/loongarch_matmul.cpp:145.  LAMM_OPT_LEVEL semantics are kept: 0 disables
 * the hook, >=1 enables the CUDA path.  Compile-time default 3, overridable at
 * run time with the environment variable LAMM_OPT_LEVEL.                      */
int lamm_get_opt_level(void);

/* SURVEY 8f-2, an extension with no counterpart in the reference (its hook sits in
 * ggml_compute_forward_mul_mat only): the same pair for GGML_OP_MUL_MAT_ID, to be called at the top
 * of ggml_compute_forward_mul_mat_id (ggml.c:10992) exactly like the mul_mat hook
 * (ggml.c:10858-10863).  dst->src[0] = ids (I32 [n_ids, N]), src[1] = b (F32 [K, N]),
 * src[2 .. 2+n_as) = the expert matrices (one type and shape), op_params[0] = id,
 * op_params[1] = n_as (ggml.c:4616-4656).  Token j is multiplied by expert ids[j][id]; the tokens
 * of one expert are gathered and run as one mul_mat through the same kernels as lamm_mul_mat
 * (a single token, the decode case, goes straight to the decode path).  2-D operands only.
 * Host logic verified on CPU (tests/test_host_logic.py), the device path by
 * tests/test_gpu_parity.py::test_operator_mul_mat_id.                                             */
bool lamm_can_mul_mat_id(const struct ggml_compute_params *params,
                         const struct ggml_tensor *dst);
void lamm_mul_mat_id(const struct ggml_compute_params *params,
                     struct ggml_tensor *dst);

/* Test hook: after lamm_mul_mat, copy the GPU's quantized activations back
 * into params->wdata (what ggml's INIT phase would have written there).       */
void lamm_set_writeback_wdata(int enable);

#ifdef __cplusplus
}
#endif
#endif /* LAMM_B200_H */
