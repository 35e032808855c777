// lamm_common.cuh -- shared device/host definitions for the sm_100a LAMM path.
//
// Block layouts are ggml's, byte for byte (llama.cpp-b2430/ggml-common.h:144-225,
// 316-321); the type table mirrors src/lamm_ggml_type_trait.h:8-56 and the
// (src0, vec_dot) pairs of src/loongarch_matmul.cpp:37-42.
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/lammcu.h"

namespace lamm {

// ---------------------------------------------------------------- type table
struct TypeInfo {
  int blck;       // elements per block
  int size;       // bytes per block
  int vec_dot;    // activation type it is multiplied with
};

__host__ __device__ constexpr TypeInfo type_info(int t) {
  switch (t) {
  case LAMMCU_TYPE_F32:  return {1, 4, LAMMCU_TYPE_F32};
  case LAMMCU_TYPE_F16:  return {1, 2, LAMMCU_TYPE_F16};   // vec_dot_type F16 (ggml.c:486-496)
  case LAMMCU_TYPE_Q4_0: return {32, 18, LAMMCU_TYPE_Q8_0};
  case LAMMCU_TYPE_Q4_1: return {32, 20, LAMMCU_TYPE_Q8_1};
  case LAMMCU_TYPE_Q5_0: return {32, 22, LAMMCU_TYPE_Q8_0};
  case LAMMCU_TYPE_Q5_1: return {32, 24, LAMMCU_TYPE_Q8_1};
  case LAMMCU_TYPE_Q8_0: return {32, 34, LAMMCU_TYPE_Q8_0};
  case LAMMCU_TYPE_Q8_1: return {32, 36, LAMMCU_TYPE_Q8_1};
  case LAMMCU_TYPE_Q2_K: return {256, 84, LAMMCU_TYPE_Q8_K};
  case LAMMCU_TYPE_Q8_K: return {256, 292, LAMMCU_TYPE_Q8_K};
  // K-quants beyond the reference's table (ggml-common.h:228-313)
  case LAMMCU_TYPE_Q3_K: return {256, 110, LAMMCU_TYPE_Q8_K};
  case LAMMCU_TYPE_Q4_K: return {256, 144, LAMMCU_TYPE_Q8_K};
  case LAMMCU_TYPE_Q5_K: return {256, 176, LAMMCU_TYPE_Q8_K};
  case LAMMCU_TYPE_Q6_K: return {256, 210, LAMMCU_TYPE_Q8_K};
  default:               return {0, 0, -1};
  }
}

// K-quants served by the generic GEMV (N <= 8) and the dequantize-once GEMM only: they are not in the
// reference's table (src/loongarch_matmul.cpp:37-42), the single-launch decode kernel has no unit for them
__host__ __device__ constexpr bool is_kquant_ext(int t) {
  return t == LAMMCU_TYPE_Q3_K || t == LAMMCU_TYPE_Q4_K || t == LAMMCU_TYPE_Q5_K || t == LAMMCU_TYPE_Q6_K;
}

// byte offsets inside the activation blocks
constexpr int Q8_0_QS = 2;     // {f16 d; i8 qs[32]}
constexpr int Q8_1_QS = 4;     // {f16 d, s; i8 qs[32]}
constexpr int Q8_K_QS = 4;     // {f32 d; i8 qs[256]; i16 bsums[16]}
constexpr int Q8_K_BSUMS = 260;

// ---------------------------------------------------------------- launch book-keeping
extern unsigned long long g_launch_count;   // lammcu_launch_count()

// ---------------------------------------------------------------- small device helpers
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ float half_bits_to_float(uint32_t h16) {
  return __half2float(__ushort_as_half((unsigned short)h16));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---------------------------------------------------------------- mbarrier / bulk copy (TMA 1-D)
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// Hand a shared-memory buffer back to its TMA producer after ld.shared into raw[]: the arrive must not become visible
// while loads of the warp are still queued.  SYNCS.ARRIVE is NOT held back by in-flight LDS (nothing it needs depends
// on them), and a bank-conflicted burst of loads can sit in the queue longer than a TMA refill from L2 takes: Q5_1 rows
// (192-byte stride, 4-way conflicts, ~1500 cycles per chunk) read the NEXT chunk's bytes in the tail of their row once
// the producer ran a whole chunk ahead (profiles/r2bf; found by tests/test_gpu_gemm.py::test_gemm_every_form_on_ragged_shapes).
// So the barrier address is made data-dependent on one word of every 16-byte load; `always_zero` is a run-time value
// that is 0 by contract (e.g. K & 1) but that the compiler cannot fold, so the dependency survives and is harmless.
template <int NWR>
__device__ __forceinline__ void mbar_arrive_after_loads(uint64_t *bar, const uint32_t (&raw)[NWR], uint32_t always_zero) {
  uint32_t dep = 0;
#pragma unroll
  for (int i = 3; i < NWR; i += 4) dep |= raw[i];
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar) + (dep & always_zero)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "LAMM_WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra LAMM_DONE_%=;\n\t"
      "bra LAMM_WAIT_%=;\n\t"
      "LAMM_DONE_%=:\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// global -> shared bulk copy (SASS: UBLKCP).  dst, src 16 B aligned; bytes % 16 == 0.
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gmem_src, uint32_t bytes,
                                         uint64_t *bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(smem_dst)),
      "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// same with an L2 evict-first policy: weights are streamed once
__device__ __forceinline__ void bulk_g2s_stream(void *smem_dst, const void *gmem_src,
                                                uint32_t bytes, uint64_t *bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1], %2, [%3], %4;" ::"r"(smem_u32(smem_dst)),
      "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}

// ---------------------------------------------------------------- host-side launchers (defined in the .cu files)
int launch_quantize_act(int vec_dot_type, const float *x, int64_t ldx, void *xq, int64_t n_rows,
                        int64_t k, cudaStream_t st);
int launch_gemv(int w_type, const void *w, int64_t w_stride, const void *xq, int64_t xq_stride,
                float *c, int64_t ldc, int64_t M, int64_t N, int64_t K, cudaStream_t st);
int launch_gemm(int w_type, const void *w, int64_t w_stride, const void *xq, int64_t xq_stride,
                float *c, int64_t ldc, int64_t M, int64_t N, int64_t K, cudaStream_t st);
// single-launch decode (N == 1): fused activation quantize (x != NULL) or pre-quantized row (xq)
bool decode_supported(int w_type, const void *w, int64_t w_stride, const void *x_or_xq, int64_t M, int64_t K);
// row-sharded launch: where the results go and how completion is published (lamm_decode.cu)
struct ShardInfo {
  unsigned long long *ll_peer[LAMMCU_MAX_RANKS];   // LL pairs of the output vector on every rank
  const unsigned long long *x_ll;                  // activation row in LL format (or NULL: plain f32 x)
  int x_back;                                      // ... produced this many sharded launches ago
  unsigned int *arrive_ctr, *status;
  int n_peer, my_rank, row_begin;
  // with a BatchInfo: the LL output vector of sibling j on every rank and this rank's first row of it
  unsigned long long *ll_batch[4][LAMMCU_MAX_RANKS];
  int64_t row_begin_b[4];
};
// resident decode server (lamm_decode.cu): jobs arrive through a mailbox in pinned host memory
constexpr int kSrvWords = 16;      // 32-bit words of a job, each posted as a {word, tag} pair
// The mailbox and the result buffer are RINGS of kSrvSlots entries indexed by (job tag % kSrvSlots): the host may post
// the jobs it expects next (the siblings of a q/k/v or gate/up group, which reuse the staged activation row) together
// with the one it is waiting for; CTA 0 picks all of them up in ONE PCIe read and the grid runs them back to back.
constexpr int kSrvSlots = 4;
struct ServerParams {
  const unsigned long long *mbox_host;   // pinned: kSrvSlots x kSrvWords pairs written by the CPU
  unsigned long long *mbox_dev;          // device: the same pairs, republished by CTA 0
  const float *x_host;                   // pinned: activation row of the current job
  unsigned long long *x_ll;              // device: the row as {value, tag} pairs
  unsigned long long *c_ll_host;         // pinned: kSrvSlots result buffers of {value, tag} pairs
  unsigned int c_slot_pairs;             // pairs per result buffer
  volatile unsigned int *done_host;      // pinned: [1] tag the server gave up waiting for
  unsigned int *status;                  // device: set on a poll time-out
  unsigned int seq0;                     // first job tag this launch serves
  unsigned int idle_ns;                  // exit after this long without a job
  unsigned int life_ns;                  // ... or this long after the launch (other streams' work gets the SMs)
  unsigned long long *trace;             // optional: [64 jobs][8] globaltimer stamps (diagnostics)
};
int decode_server_job(int w_type, const void *w, int64_t w_stride, int64_t M, int64_t K, unsigned int *words);
int launch_decode_server(int w_type, const ServerParams &sp, cudaStream_t st);
// sibling batch: `count` (2..4) matrices of one type / K / row stride that share the activation row
struct BatchInfo {
  int count;
  const void *w[4];
  float *c[4];
  int64_t M[4];
};
struct XOp;                                  // lamm_xop.cuh: the op fused in front of the matmul
int launch_decode(int w_type, const void *w, int64_t w_stride, const float *x, const void *xq, void *xq_out,
                  float *c, int64_t M, int64_t K, cudaStream_t st, const ShardInfo *sh = nullptr,
                  const BatchInfo *bt = nullptr, const XOp *xop = nullptr);
// persistent decode chain (lamm_decode.cu): a list of N == 1 mul_mats, each possibly a sibling batch, run by ONE launch
struct ChainOp {
  int count;                 // 1..4 matrices that share the activation row
  const void *w[4];
  int64_t M[4];
  float *c[4];
  const float *x;
  int64_t K, w_stride;
};
struct ChainPlan;            // opaque: device-side stage table + launch geometry
int chain_create(int w_type, int n_ops, const ChainOp *ops, ChainPlan **out);
int chain_run(ChainPlan *plan, cudaStream_t st);
void chain_destroy(ChainPlan *plan);
int chain_status(ChainPlan *plan, int *timed_out);
int launch_gemm_prep(int vt, int perm, const void *xq, int64_t xq_stride, void *x16, float *dxT, float *sxT,
                     int64_t N, int64_t Npad, int64_t K, cudaStream_t st);
int launch_gemm_prep_so(int vt, int perm, const void *xq, int64_t xq_stride, void *xh, void *xl, void *sh, float *colscale,
                        int64_t N, int64_t Npad, int64_t K, cudaStream_t st);
// row-sharded prefill: where the GEMM epilogue stores C (c[g] = rank g's full result + this rank's first column)
struct GemmPeers {
  float *c[LAMMCU_MAX_RANKS];
  int n;
};
void gemm_set_peers(const GemmPeers *peers);     // thread-local; consumed by the next scaled-operand / direct GEMM launch
bool gemm_peers_used();
const GemmPeers *gemm_peers_take();                // the pending peer set (marks it used), or NULL
// CTA-pair (cta_group::2) kernels, lamm_gemm_pair.cu
bool gemm_f16_pair_supported(int64_t M, int64_t N, int64_t K);
int launch_gemm_f16_pair(const void *w, int64_t w_stride, const void *xq, int64_t xq_stride, float *c, int64_t ldc,
                         int64_t M, int64_t N, int64_t K, cudaStream_t st);
size_t gemm_so_ws_bytes(int64_t M, int64_t N, int64_t K);
int launch_wrow_scale(int w_type, const void *w, int64_t w_stride, float *rowscale, int64_t M, int64_t Mpad, int64_t K,
                      cudaStream_t st);
int launch_gemm_so(int w_type, const void *w, int64_t w_stride, const void *xq, int64_t xq_stride, float *c,
                   int64_t ldc, int64_t M, int64_t N, int64_t K, void *ws, cudaStream_t st, const float *x = nullptr,
                   int64_t ldx = 0);
int launch_quant_prep_so(int vt, int perm, const float *x, int64_t ldx, void *xq_out, int64_t xq_stride, void *xh,
                         void *xl, void *sh, float *colscale, int64_t N, int64_t Npad, int64_t K, cudaStream_t st);
bool gemm_tc_supported(int w_type, const void *w, int64_t w_stride, int64_t M, int64_t N, int64_t K);
size_t gemm_direct_ws_bytes(int64_t M, int64_t N, int64_t K);
int launch_gemm_direct(int w_type, const void *w, int64_t w_stride, const void *xq, int64_t xq_stride, float *c,
                       int64_t ldc, int64_t M, int64_t N, int64_t K, void *ws, cudaStream_t st);
int launch_dequant_hilo(int w_type, const void *w, int64_t w_stride, void *ah, void *al, float *rowscale, int64_t M,
                        int64_t Mpad, int64_t K, cudaStream_t st);
int launch_prep_direct(int vt, const void *xq, int64_t xq_stride, void *xh, void *xl, float *colscale, int64_t N,
                       int64_t Npad, int64_t K, cudaStream_t st);
// Q2_K prefill on tcgen05 kind::i8 (lamm_gemm_x.cu): exact integer sums, one drain per 256-super-block
bool gemm_x_supported(int w_type, const void *w, int64_t w_stride, int64_t M, int64_t N, int64_t K);
size_t gemm_x_ws_bytes(int64_t N, int64_t K);
int launch_gemm_x(int w_type, const void *w, int64_t w_stride, const void *xq, int64_t xq_stride, float *c, int64_t ldc,
                  int64_t M, int64_t N, int64_t K, void *ws, cudaStream_t st);
bool gemm_f16_supported(const void *w, int64_t w_stride, const void *xq, int64_t xq_stride, int64_t M, int64_t N, int64_t K);
int launch_gemm_f16(const void *w, int64_t w_stride, const void *xq, int64_t xq_stride, float *c, int64_t ldc, int64_t M,
                    int64_t N, int64_t K, cudaStream_t st);
void set_gemm_exact(int on);
int get_gemm_exact();
void *gemm_workspace(size_t bytes);      // library-owned scratch for the prefill GEMM (grown on demand)
void set_error(const char *fmt, ...);
int sm_count_cached();
// scratch for the dynamic tile scheduler: returns a zeroed pair of counters
unsigned int *next_sched_slot();
// diagnostics: 8 x u64 per CTA of the next decode launch, or NULL (lammcu_set_trace)
unsigned long long *next_trace_slot(int grid);

}  // namespace lamm
