/*
 * lammcu.h -- C-ABI of the B200 (sm_100a) quantized mul_mat device library
 * (liblammcu.so).  Plain pointers and sizes only; no C++ or torch types.
 *
 * This is the boundary the reference's host code binds to.  It replaces the
 * CPU kernels behind src/loongarch_matmul.cpp:64-143 (lamm_mul_mat) and the
 * activation quantize of llama.cpp-b2430/ggml.c:10865-10887:
 *
 *   reference (CPU)                                  here (CUDA)
 *   ------------------------------------------------ ---------------------------
 *   from_float_to_vec_dot / quantize_row_q8_{0,1,K}   lammcu_quantize_act_dev
 *     ggml-quants.c:1207,1431,4031
 *   LAMMImpl<T>::matmul  src/lamm_impl.hpp:20-29      lammcu_mul_mat_q_dev
 *     + lamm_kernel_*.hpp                              lammcu_mul_mat_dev (fused)
 *   ggml type_traits rows  ggml.c:458-775             lammcu_{blck,type,row}_size,
 *   ggml_type_trait<T>  src/lamm_ggml_type_trait.h     lammcu_vec_dot_type
 *   lamm_mul_mat on host tensors                      lammcu_weight_register +
 *     src/loongarch_matmul.cpp:64-143                  lammcu_mul_mat_host
 *
 * Data layouts are ggml's, unchanged (llama.cpp-b2430/ggml-common.h:144-225,
 * 316-321): W is M rows of K/QK packed blocks, row stride in bytes; X is N rows
 * of K f32; Xq is N rows of vec-dot blocks packed at lammcu_row_size(); C is N
 * rows of M f32, C[j*ldc + i] (src/lamm_kernel_q4_1.hpp:24).
 *
 * All functions return LAMMCU_OK (0) or a negative lammcu_status; the message
 * of the last failure on the calling thread is lammcu_last_error().  There is
 * no CPU fallback: without a CUDA device every compute entry point fails with
 * LAMMCU_ERR_NO_DEVICE.
 *
 * `stream` arguments are a cudaStream_t passed as void* (NULL = the library's
 * own stream for the current device).  *_dev entry points are asynchronous on
 * that stream; *_host entry points return after the result is in host memory.
 */
#ifndef LAMMCU_H
#define LAMMCU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LAMMCU_MAX_RANKS 8

typedef enum {
  LAMMCU_OK = 0,
  LAMMCU_ERR_NO_DEVICE = -1,   /* no CUDA device / driver                      */
  LAMMCU_ERR_CUDA = -2,        /* a CUDA runtime call failed                   */
  LAMMCU_ERR_TYPE = -3,        /* ggml type not on the path                    */
  LAMMCU_ERR_SHAPE = -4,       /* K not a multiple of the block size, M/N<=0.. */
  LAMMCU_ERR_ALIGN = -5,       /* pointer / stride alignment not supported     */
  LAMMCU_ERR_ARG = -6,         /* NULL pointer, bad handle                     */
  LAMMCU_ERR_NOMEM = -7,
  LAMMCU_ERR_TIMEOUT = -8      /* a peer never published what a sharded call polled for */
} lammcu_status;

/* ggml type ids accepted (llama.cpp-b2430/ggml.h:341-369) */
enum {
  LAMMCU_TYPE_F32 = 0, LAMMCU_TYPE_F16 = 1 /* fp16 weights, activations rounded to fp16 (ggml.c:486-496) */,
  LAMMCU_TYPE_Q4_0 = 2, LAMMCU_TYPE_Q4_1 = 3,
  LAMMCU_TYPE_Q5_0 = 6, LAMMCU_TYPE_Q5_1 = 7, LAMMCU_TYPE_Q8_0 = 8,
  LAMMCU_TYPE_Q8_1 = 9, LAMMCU_TYPE_Q2_K = 10, LAMMCU_TYPE_Q8_K = 15,
  /* the K-quants the reference lists as TODO (dev.md:86-95; the reference's gate
   * refuses them, so ggml's own ggml_vec_dot_q{3,4,5,6}_K_q8_K is the behaviour matched here).
   * Same Q8_K activation kernel as Q2_K.                                                      */
  LAMMCU_TYPE_Q3_K = 11, LAMMCU_TYPE_Q4_K = 12, LAMMCU_TYPE_Q5_K = 13, LAMMCU_TYPE_Q6_K = 14
};

/* ---- library / device ------------------------------------------------- */
const char *lammcu_version(void);
const char *lammcu_last_error(void);
int lammcu_device_count(void);              /* >=0, or negative status         */
int lammcu_init(int device);                /* select device, create stream    */
int lammcu_shutdown(void);                  /* free weights, workspaces        */
int lammcu_sm_count(void);
/* number of kernels this library has launched since load (bench's
 * gpu_launches); lammcu_launch_count_reset() zeroes it                      */
uint64_t lammcu_launch_count(void);
void lammcu_launch_count_reset(void);

/* ---- type table (pure host functions, no device needed) ---------------- */
int    lammcu_type_supported(int w_type);   /* 1 if lamm_can_mul_mat's table has it */
int    lammcu_blck_size(int type);
size_t lammcu_type_size(int type);
int    lammcu_vec_dot_type(int w_type);
size_t lammcu_row_size(int type, int64_t k);

/* ---- device-resident operands ------------------------------------------ */
/* X f32 [n_rows, k] (row stride ldx elements) -> Xq packed at row_size.
 * Bit-exact with quantize_row_q8_{0,1,K}_reference; Q8_K bsums of an all-zero
 * super-block are written as 0 (the reference leaves them untouched).        */
int lammcu_quantize_act_dev(int vec_dot_type, const float *x_dev, int64_t ldx,
                            void *xq_dev, int64_t n_rows, int64_t k, void *stream);

/* C = W . Xq^T on already quantized activations (what COMPUTE sees).
 * For LAMMCU_TYPE_F32 weights xq_dev is the f32 X itself.                    */
int lammcu_mul_mat_q_dev(int w_type, const void *w_dev, int64_t w_stride_bytes,
                         const void *xq_dev, int64_t xq_stride_bytes,
                         float *c_dev, int64_t ldc, int64_t M, int64_t N,
                         int64_t K, void *stream);

/* The whole path on device operands: activation quantize + mul_mat.
 * xq_ws_dev: caller workspace of >= N*lammcu_row_size(vec_dot,K) bytes, or
 * NULL to use the library's own (grown on demand, per device).               */
int lammcu_mul_mat_dev(int w_type, const void *w_dev, int64_t w_stride_bytes,
                       const float *x_dev, int64_t ldx, float *c_dev,
                       int64_t ldc, int64_t M, int64_t N, int64_t K,
                       void *xq_ws_dev, void *stream);

/* No counterpart in the reference: added for callers that cannot allocate
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The same result as
 * lammcu_mul_mat_dev for 1 <= N <= 8 quantized-weight matmuls (activation quantize + GEMV), but the quantized-
 * activation scratch lives in a static device array of LAMMCU_NOALLOC_MAX_SCRATCH bytes that every GPU loading the
 * library holds until exit: the call performs no cudaMalloc / cudaFree and uses no library workspace.
 *   - needs N * lammcu_row_size(lammcu_vec_dot_type(w_type), K) <= LAMMCU_NOALLOC_MAX_SCRATCH (e.g. N = 8 up to
 *     K = 57344); larger requests fail with LAMMCU_ERR_SHAPE and a message; F32 weights (no scratch) are refused;
 *   - lammcu_init must have been called before (it owns the library's only allocations); the first call per
 *     process looks the array's device address up (cudaGetSymbolAddress, no allocation): make it outside capture;
 *   - every call shares the one array and leaves its quantized rows there: calls that overlap on the GPU (two
 *     streams, two host threads) race -- serialise them on one stream.  Nothing has to be cleared between calls.
 * Asynchronous on `stream` (NULL = the library's stream).                                                        */
#define LAMMCU_NOALLOC_MAX_SCRATCH (512 * 1024)
int lammcu_mul_mat_noalloc_dev(int w_type, const void *w_dev, int64_t w_stride_bytes, const float *x_dev, int64_t ldx,
                               float *c_dev, int64_t ldc, int64_t M, int64_t N, int64_t K, void *stream);

/* N == 1, sibling matmuls: `count` weight matrices of the same type, K and row stride
 * that multiply the SAME activation row (wq/wk/wv, gate/up in llama.cpp's build_llama,
 * llama.cpp:5726-5800) -- C_j = W_j . x for every j.  Up to 4 of them run as ONE launch of
 * the decode kernel (one activation quantize, one dependency wait); anything the kernel
 * cannot take falls back to one lammcu_mul_mat_dev call per matrix.  Results are identical
 * to separate calls.                                                                     */
int lammcu_mul_mat_batch_dev(int w_type, int count, const void *const *w_dev,
                             int64_t w_stride_bytes, const int64_t *M, const float *x_dev,
                             float *const *c_dev, int64_t K, void *xq_ws_dev, void *stream);

/* ---- the ops either side of the path, fused into the activation quantizer ---------------------
 * llama.cpp's graph (build_llama, llama.cpp:5726-5800) never hands a matmul a raw tensor: src1 is either
 *   LAMMCU_XOP_RMS_NORM_MUL  llm_build_norm (llama.cpp:5171-5201): ggml_rms_norm(x, eps) (ggml.c:10355-10405) then
 *                            ggml_mul by the norm weight -- in front of wq/wk/wv, gate/up and the output head;
 *                            aux = norm weight [K] (NULL: no weight), eps > 0
 *   LAMMCU_XOP_SILU_MUL      llm_build_ffn (llama.cpp:5203-5290): ggml_silu(gate) (ggml.c:1894-1902, the fp16 table)
 *                            times up -- in front of the down projection; x = gate, aux = up [N][ld_aux]
 * and ggml then quantizes that f32 tensor in the matmul's INIT phase (ggml.c:10865-10887).  These entry points replace
 * the three nodes (op, mul, quantize) + the matmul: the f32 intermediate never exists in memory, and for N == 1 the
 * op is applied inside the decode kernel while it stages the row (ONE launch for norm + quantize + matmul).  The
 * arithmetic is ggml's operation for operation (same IEEE roundings; the SiLU table is built by the host libm exactly
 * as ggml_init builds it), so the quantized activations are the ones ggml would have produced.
 * Rows of x and aux must be 16-byte aligned.  xq_ws_dev (may be NULL) receives the quantized rows.               */
enum { LAMMCU_XOP_RMS_NORM_MUL = 1, LAMMCU_XOP_SILU_MUL = 2 };
typedef struct lammcu_xop {
  int kind;
  const float *aux;       /* device pointer */
  int64_t ld_aux;         /* SILU_MUL: floats between rows of aux */
  float eps;              /* RMS_NORM_MUL */
} lammcu_xop;
/* op + quantize only: xq_dev = quantize(vt, op(x)), N rows packed at lammcu_row_size(vt, K); vt may also be
 * LAMMCU_TYPE_F16 / LAMMCU_TYPE_F32 (plain conversion / the f32 result itself, what F16 / F32 weights consume)  */
int lammcu_quantize_act_xop_dev(const lammcu_xop *op, int vt, const float *x_dev, int64_t ldx, void *xq_dev,
                                int64_t N, int64_t K, void *stream);
/* C = W . op(x): lammcu_mul_mat_dev with the op in front */
int lammcu_mul_mat_xop_dev(const lammcu_xop *op, int w_type, const void *w_dev, int64_t w_stride_bytes,
                           const float *x_dev, int64_t ldx, float *c_dev, int64_t ldc, int64_t M, int64_t N, int64_t K,
                           void *xq_ws_dev, void *stream);
/* N == 1 siblings sharing op(x) (wq/wk/wv after attn_norm, gate/up after ffn_norm): lammcu_mul_mat_batch_dev with the
 * op in front -- one launch for the norm, the quantize and up to 4 matrices                                      */
int lammcu_mul_mat_batch_xop_dev(const lammcu_xop *op, int w_type, int count, const void *const *w_dev,
                                 int64_t w_stride_bytes, const int64_t *M, const float *x_dev, float *const *c_dev,
                                 int64_t K, void *xq_ws_dev, void *stream);

/* N == 1, a CHAIN of dependent matmuls as ONE persistent launch (the same extension class as the sibling batch above;
 * in llama.cpp's build_llama, llama.cpp:5726-5800, a decoded token is such a chain: q,k,v = W x; o = Wo v; g,u = W o;
 * x' = Wd u; ...).  Op i multiplies 1..4 sibling matrices (one type, K and row stride) with the activation row x_dev
 * and writes c_dev[j]; ops run in order and an op may read what earlier ops wrote -- a grid-wide barrier separates
 * them inside the kernel.  Every warp's weight stream runs ahead across op boundaries (weights are immutable), so
 * HBM does not idle while an op waits for its activations.  Results are identical to one lammcu_mul_mat_dev /
 * lammcu_mul_mat_batch_dev call per op (same per-row arithmetic).
 *   lammcu_chain_create   validates and uploads the op table (allocates: do it once, outside graph capture);
 *   lammcu_chain_run_dev  one memset node + one cooperative launch on `stream` (grid = SM count: the GPU must be able
 *                         to hold one CTA per SM at once; a resident decode server on the library's stream is stopped);
 *   lammcu_chain_status   *timed_out = 1 if a barrier poll gave up after ~2 s (synchronous D2H read).
 * The pointers in the table must stay valid until lammcu_chain_destroy.                                            */
typedef struct {
  int count;                  /* 1..4 matrices sharing the activation row */
  const void *w_dev[4];
  int64_t M[4];
  float *c_dev[4];
  const float *x_dev;         /* K f32 */
  int64_t K;
  int64_t w_stride_bytes;
} lammcu_chain_op;
typedef struct lammcu_chain *lammcu_chain_t;
int lammcu_chain_create(int w_type, int n_ops, const lammcu_chain_op *ops, lammcu_chain_t *out);
int lammcu_chain_run_dev(lammcu_chain_t chain, void *stream);
int lammcu_chain_status(lammcu_chain_t chain, int *timed_out);
int lammcu_chain_destroy(lammcu_chain_t chain);

/* Path selection (diagnostics / benchmarking): N <= threshold runs the
 * memory-bound GEMV kernels, larger N the tensor-core GEMM.  Default 8.      */
int lammcu_set_gemv_max_n(int n);
int lammcu_get_gemv_max_n(void);
/* Prefill (N > gemv_max_n) tensor-core kernels.  Q2_K always runs on tcgen05 kind::i8 (scale-folded u8 x s8, one
 * drain per 256-super-block, the min term as one extra int8 MMA: exact integer sums).  The 32-element formats:
 *   0 (default)  scaled-operand mainloop (scales folded into fp16 hi/lo operand pairs, per-row power-of-two lifts on
 *                both operands, one accumulator over the whole K loop; error <= 2^-22 per product, inside the
 *                path's 1e-5*sqrt(K) tolerance);
 *   1            exact per-32-block integer sums (bit-identical to the reference's sumi), scaled on the CUDA cores.
 * Environment override at load time: LAMM_GEMM_EXACT.                                                          */
int lammcu_set_gemm_exact(int on);
int lammcu_get_gemm_exact(void);
/* N == 1: 1 (default) = the single-launch decode kernel (fused activation
 * quantize, programmatic dependent launch); 0 = quantize kernel + generic GEMV */
int lammcu_set_decode_kernel(int enable);
int lammcu_get_decode_kernel(void);
/* Diagnostics: every following decode launch writes 8 x u64 %globaltimer stamps per
 * CTA (start, weights issued, dependency wait done, activations staged, done) into
 * consecutive slots of dev_buf (n_u64 entries); NULL switches tracing off.          */
int lammcu_set_trace(void *dev_buf, int64_t n_u64);

/* ---- row-sharded multi-GPU decode (one process per GPU) ------------------------
 * Weight rows (M) are split contiguously across ranks the way the reference splits
 * them across threads (src/lamm_impl.hpp:107-112) and upstream CUDA across devices
 * (llama.cpp-b2430/ggml-cuda.cu:8480-8492).  Every rank computes its row slice and
 * the kernel's epilogue stores it straight into the full output vector of EVERY rank
 * through peer-mapped memory (NVLink): there is no separate all-gather and no
 * collective call.
 *
 * Output vectors of sharded launches are in "LL" format so that no fence, flag or
 * kernel boundary is needed: M pairs {f32 value, u32 tag}, each pair written with ONE
 * 64-bit store.  The tag is the launch's sequence number on the communicator, drawn on
 * the device from an arrival-ticket counter (all ranks must issue the same sequence of
 * sharded calls; every sharded launch uses the same grid size).  A consumer -- a later
 * sharded launch with x_back >= 1, or lammcu_ll_unpack_dev -- polls each pair until its
 * tag equals (own sequence number - x_back), so an element is usable the moment it
 * lands and the consumer does not wait for the producer GRID to complete (it never
 * executes griddepcontrol.wait): the dependency is carried by the data itself.
 * x_back is how many sharded launches ago (on this communicator, in issue order) the
 * vector was produced: 1 = by the previous sharded launch.  Producer and consumer must
 * be issued in the same eager sequence or captured into the same CUDA graph.
 * Buffer reuse: a vector may be overwritten by a sharded launch once an earlier-issued
 * sharded launch has consumed output produced AFTER the vector's last reader (two
 * vectors per role, ping-pong, always satisfy this in a layer chain).
 * A poll gives up after ~2 s (lost peer) and sets the communicator's status word.
 *
 * lammcu_ipc_alloc / _open give peer-mappable device buffers (cudaIpc*): the owner
 * allocates (zero-filled) and ships the 64-byte handle to the other ranks by any host
 * channel (torch.distributed in the Python layer), who open it.                      */
typedef struct lammcu_comm *lammcu_comm_t;
int lammcu_ipc_alloc(size_t bytes, void **dev_ptr, void *handle64);
int lammcu_ipc_open(const void *handle64, void **dev_ptr);
int lammcu_ipc_close(void *dev_ptr);
int lammcu_ipc_free(void *dev_ptr);
/* cudaMemcpyAsync(cudaMemcpyDefault) on the given stream: moves data between caller
 * tensors and peer-mappable buffers without another CUDA binding                     */
int lammcu_memcpy_async(void *dst, const void *src, size_t bytes, void *stream);
int lammcu_comm_create(int rank, int world, lammcu_comm_t *out);
int lammcu_comm_destroy(lammcu_comm_t comm);
/* *timed_out = 1 if any poll on this communicator gave up (synchronous D2H read)    */
int lammcu_comm_status(lammcu_comm_t comm, int *timed_out);
/* the synchronisation point of a sharded chain: waits for the stream, then LAMMCU_ERR_TIMEOUT (and a message) if any
 * poll of any earlier sharded call on this communicator gave up -- the results since then are not to be used.       */
int lammcu_comm_sync(lammcu_comm_t comm, void *stream);
/* bytes of an LL vector of M elements (8 per element); allocate with lammcu_ipc_alloc
 * (zero-filled: tag 0 is never used)                                                  */
size_t lammcu_ll_bytes(int64_t M);
/* N == 1.  w_shard: this rank's rows [row_begin, row_begin + rows) of W.  x: the
 * activation row on THIS rank -- K plain f32 (x_back = 0; the launch then waits for the
 * previous kernel in the stream as usual) or an LL vector (x_back >= 1, see above).
 * c_ll_all[g]: the output LL vector (M elements) on rank g as mapped in this process.
 * world == 1 is allowed (a single GPU chaining its own launches through LL vectors).
 * Asynchronous on `stream`.                                                           */
int lammcu_mul_mat_sharded_dev(lammcu_comm_t comm, int w_type, const void *w_shard,
                               int64_t w_stride_bytes, const void *x, int x_back,
                               void *const *c_ll_all, int64_t row_begin, int64_t rows,
                               int64_t K, void *stream);
/* Sibling matrices that share the activation row (q/k/v, gate/up) as ONE sharded launch (one
 * sequence number for all of them): matrix j has its own LL output vector; c_ll_all is the
 * [count][world] array of their per-rank buffers (c_ll_all[j * world + g]); row_begin[j] / rows[j]
 * describe this rank's shard of matrix j.  2 <= count <= 4 (1 forwards to the call above).      */
int lammcu_mul_mat_sharded_batch_dev(lammcu_comm_t comm, int w_type, int count,
                                     const void *const *w_shards, int64_t w_stride_bytes,
                                     const void *x, int x_back, void *const *c_ll_all,
                                     const int64_t *row_begin, const int64_t *rows, int64_t K,
                                     void *stream);
/* Row-sharded PREFILL (N > 1): this rank multiplies its rows [row_begin, row_begin + rows) of W
 * with the replicated f32 activations x [N][ldx]; the GEMM epilogue stores the result straight into
 * the full f32 result C [N][ldc] (ggml layout, ldc >= M) of EVERY rank: c_all[g] is rank g's C,
 * lammcu_ipc_alloc'ed and opened on every rank.  flags_all[g] is rank g's flag array (>= 64 bytes,
 * zero-initialised, from lammcu_ipc_alloc): when the call's work has run on the stream, every rank's
 * rows are complete in the LOCAL C (a publish + wait launch follow the GEMM; all ranks must issue
 * the same sequence of these calls).  Shapes the tensor-core path does not take still work (the
 * slab is then copied to the peers).                                                           */
int lammcu_mul_mat_sharded_gemm_dev(lammcu_comm_t comm, int w_type, const void *w_shard,
                                    int64_t w_stride_bytes, const float *x, int64_t ldx,
                                    void *const *c_all, int64_t ldc, void *const *flags_all,
                                    int64_t row_begin, int64_t rows, int64_t N, int64_t K,
                                    void *stream);
/* LL vector (local, produced `back` sharded launches ago, >= 1) -> plain f32 out[M],
 * waiting for every element.  An ordinary kernel: it starts after all earlier work on
 * the stream has completed.                                                           */
int lammcu_ll_unpack_dev(lammcu_comm_t comm, const void *c_ll, int back, float *out,
                         int64_t M, void *stream);

/* ---- host operands (what lamm_mul_mat passes) --------------------------- */
typedef struct lammcu_weight *lammcu_weight_t;

/* Upload W once (weights are immutable); rows [row_begin,row_end) only, so a
 * rank can hold its shard.  Keyed by (host pointer, shard) in a registry so a
 * second registration of the same tensor returns the same handle (one more
 * reference: pair every register with a lammcu_weight_release).  Registering a
 * DIFFERENT tensor at the same address evicts the old entry; a holder's handle
 * to it stays valid until that holder releases it.                           */
int lammcu_weight_register(const void *w_host, int w_type, int64_t M, int64_t K,
                           int64_t w_stride_bytes, int64_t row_begin,
                           int64_t row_end, lammcu_weight_t *out);
/* The operator's own way in (lamm_mul_mat / lamm_mul_mat_id): the registry keeps the entry, the caller takes no
 * reference and never releases.  refresh = 0: src0 is immutable model state (a leaf tensor: op == GGML_OP_NONE,
 * no view_src, not a PARAM) -- a hit serves the resident device copy.  refresh != 0: src0 can change between
 * calls while its address and shape stay the same (a K/V-cache view, an activation used as src0, a trainable
 * weight; the reference reads src0->data on every call, src/loongarch_matmul.cpp:131-135) -- the bytes are
 * uploaded again into the existing device buffer on every call.                                            */
int lammcu_weight_for_op(const void *w_host, int w_type, int64_t M, int64_t K, int64_t w_stride_bytes,
                         int refresh, lammcu_weight_t *out);
int lammcu_weight_lookup(const void *w_host, lammcu_weight_t *out); /* LAMMCU_ERR_ARG if absent */
int lammcu_weight_release(lammcu_weight_t w);
/* The registry is keyed by the HOST address of src0->data (weights are immutable model state).  When the
 * application frees a model it must say so, or a later tensor allocated at the same address with the same
 * type and shape would hit the stale device copy: forget one tensor, or everything (e.g. from
 * llama_free_model).  Handles to forgotten entries become invalid.                                      */
int lammcu_weight_forget(const void *w_host);
int lammcu_weight_forget_all(void);

/* ---- GGUF weights straight into (sharded) device memory ---------------------------
 * Replaces gguf_init_from_file + the tensor-data reads of llama.cpp's model loader
 * (llama.cpp-b2430/ggml.c:20655-20850) for the weights of the path: the file is memory-mapped, the
 * container parsed (versions 2 and 3, every read bounds-checked), and a 2-D tensor's rows
 * [row_begin, row_end) are registered from the mapping as they are -- the on-disk block layout is
 * the device layout.  Errors: LAMMCU_ERR_ARG with the reason in lammcu_last_error().
 * Strings returned by lammcu_gguf_tensor stay valid until lammcu_gguf_close; weights registered from
 * a file are keyed by their address in the mapping, so lammcu_weight_forget_all() (or _forget per
 * tensor) belongs before lammcu_gguf_close.                                                      */
typedef struct lammcu_gguf *lammcu_gguf_t;
int      lammcu_gguf_open(const char *path, lammcu_gguf_t *out);
int      lammcu_gguf_close(lammcu_gguf_t g);
int64_t  lammcu_gguf_tensor_count(lammcu_gguf_t g);
uint32_t lammcu_gguf_alignment(lammcu_gguf_t g);          /* general.alignment, default 32 (ggml.h:250) */
uint64_t lammcu_gguf_data_offset(lammcu_gguf_t g);        /* gguf_get_data_offset */
/* any out pointer may be NULL.  ne4: 4 entries, unused dims = 1.  nbytes = 0 for a ggml type this
 * library has no size for (the data pointer is still valid).                                     */
int      lammcu_gguf_tensor(lammcu_gguf_t g, int64_t i, const char **name, int *type, int *n_dims,
                            int64_t *ne4, const void **data, uint64_t *nbytes);
int64_t  lammcu_gguf_find(lammcu_gguf_t g, const char *name);          /* index, or -1 */
int      lammcu_gguf_get_u32(lammcu_gguf_t g, const char *key, uint32_t *out);
int      lammcu_gguf_get_f32(lammcu_gguf_t g, const char *key, float *out);
int64_t  lammcu_gguf_get_str(lammcu_gguf_t g, const char *key, char *buf, int64_t cap); /* length, or -1 */
/* the bytes of rows [row_begin, row_end) of 1-D / 2-D tensor i in the mapping: one contiguous range */
int      lammcu_gguf_rows(lammcu_gguf_t g, int64_t i, int64_t row_begin, int64_t row_end,
                          const void **data, uint64_t *nbytes);
/* lammcu_weight_register on rows [row_begin, row_end) of 2-D tensor i (a rank's shard, or all rows) */
int      lammcu_gguf_upload_rows(lammcu_gguf_t g, int64_t i, int64_t row_begin, int64_t row_end,
                                 lammcu_weight_t *out);

/* Expert selection of ggml_mul_mat_id (ggml.c:11066-11078 builds the same lists as
 * matrix_row_counts / matrix_rows in its INIT phase): token j goes to expert
 * ids[j * ids_stride + id] (ids_stride in int32 elements).  counts[n_as] receives the number of
 * tokens per expert, order[n_tokens] the token indices grouped by expert (expert 0's tokens first,
 * ascending within an expert).  Pure host function: needs no device and takes no lock.
 * LAMMCU_ERR_ARG for a bad argument or an id outside [0, n_as) (ggml asserts there).            */
int lammcu_mul_mat_id_plan(const int32_t *ids, int64_t ids_stride, int id, int64_t n_tokens,
                           int n_as, int64_t *counts, int64_t *order);
const void *lammcu_weight_dev_ptr(lammcu_weight_t w);
int64_t lammcu_weight_rows(lammcu_weight_t w);
int64_t lammcu_weight_bytes(lammcu_weight_t w);   /* device bytes held for it: rows x row_size, a rank's shard only */

/* X (host f32, N rows, ldx elements) -> H2D -> quantize -> mul_mat -> D2H into
 * C (host f32, ldc elements; only the handle's rows [row_begin,row_end) of
 * every output row are written).  xq_out_host (optional) receives the GPU's
 * quantized activations, for the bit-exact check.  Synchronous.              */
int lammcu_mul_mat_host(lammcu_weight_t w, const float *x_host, int64_t N,
                        int64_t ldx, float *c_host, int64_t ldc,
                        void *xq_out_host);
/* Diagnostics: accumulated host-path phase times in microseconds since the last reset:
 * [0] x memcpy to pinned, [1] H2D enqueue, [2] kernel enqueue, [3] wait for completion,
 * [4] C memcpy from pinned, [5] number of calls, [6] calls whose activation row was not sent
 * because it equalled the previous call's (host mode 1), [7] calls answered by a job that was
 * posted ahead of them (sibling speculation, LAMM_SERVER_SPEC).                           */
void lammcu_host_profile(double *out8, int reset);
/* How lammcu_mul_mat_host (and lamm_mul_mat above it) runs a decode-sized (N == 1) call:
 *   0  copy-engine H2D of X, one kernel launch, zero-copy f32 result in pinned memory,
 *      cudaStreamSynchronize (~30 us per call, most of it launch + completion latency);
 *   1  (default) the call is a JOB for a resident decode server: one long-lived launch (a CTA per
 *      SM) that polls a mailbox in pinned host memory, pulls X over PCIe, runs the same decode
 *      body and copies C back; the calling thread only writes X + 16 tagged words and polls one
 *      word.  The server exits by itself after LAMM_SERVER_IDLE_US (default 200) without a job and
 *      is relaunched on demand; every other entry point that needs the library's stream stops it
 *      first.  lammcu_host_quiesce() stops it explicitly (e.g. before timing unrelated GPU work).
 *      Consecutive calls with a bit-identical activation row (ggml gives q/k/v and gate/up the same
 *      src1) send it once: the later jobs reuse the quantized row the resident instance still holds
 *      (memcmp against the pinned copy; LAMM_SERVER_X_REUSE=0 turns this off).  The path also
 *      learns which weight was called next with the same row and posts that weight's job behind
 *      the current one (up to two followers; mailbox and result buffers are rings of 4): the
 *      following call is answered from it only if it is for that weight and its row memcmp-equals
 *      the job's; anything else is posted normally (LAMM_SERVER_SPEC=0 turns this off).
 * Results are bit-identical in both modes.  Environment override at load time: LAMM_HOST_MODE.  */
int lammcu_set_host_mode(int mode);
int lammcu_get_host_mode(void);
int lammcu_host_quiesce(void);
/* The server leaves the GPU after idle_us without a job (default 200, LAMM_SERVER_IDLE_US) and at
 * the first job boundary life_us after its launch (default 5000, LAMM_SERVER_LIFE_US), so that a
 * caller that never pauses cannot starve kernels queued on other streams.  It is relaunched by the
 * next call; both paths are transparent to the caller.                                          */
int lammcu_set_server_timeouts(unsigned int idle_us, unsigned int life_us);
/* Diagnostics (LAMM_SERVER_TRACE=1 at load time): device time stamps (ns) of the last 64 server
 * jobs, out[64][8] = {job seen by CTA 0, X republished in HBM, X quantized, CTA 0's rows done,
 * CTA 0's chunk written to host memory, 0, 0, 0}.  Stops the server.                            */
int lammcu_server_trace(unsigned long long *out512);
/* same, activations already quantized on the host (src1->type == vec_dot_type) */
int lammcu_mul_mat_q_host(lammcu_weight_t w, const void *xq_host,
                          int64_t xq_stride_bytes, int64_t N, float *c_host,
                          int64_t ldc);

#ifdef __cplusplus
}
#endif
#endif /* LAMMCU_H */
