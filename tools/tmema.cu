// tmema.cu -- the TS form of tcgen05.mma (A operand in TENSOR MEMORY, B through a shared-memory descriptor): known-answer
// check of the operand layout and the issue rate, single CTA and CTA pair (diagnostics for the prefill GEMM's
// "unpack straight into TMEM" variant).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/tmema tools/tmema.cu && /tmp/tmema
//
//  (1) known answer: thread = row writes its 64 fp16 of A (small integers) into 32 TMEM columns of its lane with
//      tcgen05.st.32x32b.x16 (word c = elements 2c, 2c + 1), B (N rows x 64 fp16) sits in shared memory in the K-major
//      SWIZZLE_128B layout; 4 x tcgen05.mma kind::f16 K16 with A = [tmem + 8 kk]; D drained with tcgen05.ld, compared
//      exactly.  PAIR: cluster of 2, cta_group::2, M256 x N256 -- each CTA writes ITS 128 rows of A into its own TMEM
//      at the same column and holds 128 of B's 256 rows.
//  (2) rate: `iters` x 4 MMAs per CTA (pair) with A resident in TMEM, as in tools/i8peak.cu.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::"r"(
          smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ uint64_t sw128_desc(uint32_t addr) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__host__ __device__ constexpr uint32_t make_idesc(int M, int N) {   // f32 accumulate, fp16 x fp16, both K-major
  return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__host__ __device__ inline uint32_t sw128_off(int r, int c) {
  return (uint32_t)(r >> 3) * 1024u + (uint32_t)(r & 7) * 128u + ((((uint32_t)c >> 4) ^ (uint32_t)(r & 7)) << 4) + ((uint32_t)c & 15u);
}
template <bool PAIR>
__device__ __forceinline__ void umma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b, uint32_t idesc, uint32_t acc) {
  if (PAIR)
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
                 "r"(a_tmem), "l"(b), "r"(idesc), "r"(acc)
                 : "memory");
  else
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
                 "r"(a_tmem), "l"(b), "r"(idesc), "r"(acc)
                 : "memory");
}
template <bool PAIR>
__device__ __forceinline__ void commit_all(uint64_t *bar) {
  if (PAIR)
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
                 "h"((uint16_t)3)
                 : "memory");
  else
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
template <bool PAIR>
__device__ __forceinline__ uint32_t tmem_setup(uint32_t *slot, int warp) {
  if (warp == 0) {
    if (PAIR) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(512u) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(512u) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  return *slot;
}
template <bool PAIR>
__device__ __forceinline__ void tmem_free(uint32_t tmem, int warp) {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (PAIR) cluster_sync();
  if (warp == 0) {
    if (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
  }
}
// 16 words of this thread's lane at column `col`
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t *v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
      "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}

constexpr uint32_t kACol = 256;   // A lives behind a 256-column accumulator

// ---- (1) known answer ----------------------------------------------------------------------------
template <bool PAIR>
__global__ void __launch_bounds__(128, 1) kat_ts_kernel(const __half *a_rows, const __half *b_rows, float *d_out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t *sb = smem;
  uint64_t *bar = reinterpret_cast<uint64_t *>(smem + 16384);
  uint32_t *slot = reinterpret_cast<uint32_t *>(smem + 16384 + 16);
  const int tid = threadIdx.x, warp = tid >> 5;
  constexpr int N = PAIR ? 256 : 128;
  uint32_t rank = 0;
  if (PAIR) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  const int row = (int)rank * 128 + tid;
  for (int c = 0; c < 128; c += 16)
    *reinterpret_cast<uint4 *>(sb + sw128_off(tid, c)) = *reinterpret_cast<const uint4 *>(reinterpret_cast<const uint8_t *>(b_rows) + row * 128 + c);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (PAIR) {
    __syncthreads();
    cluster_sync();
  }
  const uint32_t tmem = tmem_setup<PAIR>(slot, warp);
  const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
  uint32_t av[32];
  for (int i = 0; i < 32; i++) av[i] = reinterpret_cast<const uint32_t *>(a_rows)[row * 32 + i];
  tmem_st16(lane_base + kACol, av);
  tmem_st16(lane_base + kACol + 16, av + 16);
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (PAIR) cluster_sync();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (tid == 0 && rank == 0) {
    const uint64_t bd = sw128_desc(smem_u32(sb));
    const uint32_t idesc = make_idesc(PAIR ? 256 : 128, N);
#pragma unroll
    for (int kk = 0; kk < 4; kk++) umma_ts<PAIR>(tmem, tmem + kACol + (uint32_t)(8 * kk), bd + (uint64_t)(2 * kk), idesc, kk ? 1u : 0u);
    commit_all<PAIR>(bar);
  }
  mbar_wait(bar, 0u);
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  for (int c16 = 0; c16 < N / 16; c16++) {
    uint32_t v[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(lane_base + (uint32_t)(c16 * 16))
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int i = 0; i < 16; i++) d_out[row * N + c16 * 16 + i] = __uint_as_float(v[i]);
  }
  tmem_free<PAIR>(tmem, warp);
}

// ---- (2) issue rate with A resident in TMEM -------------------------------------------------------
template <bool PAIR>
__global__ void __launch_bounds__(128, 1) rate_ts_kernel(int iters, unsigned long long *cycles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  constexpr int kBBytes = PAIR ? 16384 : 32768;           // N = 256: 128 rows per CTA of a pair, 256 alone
  uint8_t *sb = smem;
  uint64_t *bar = reinterpret_cast<uint64_t *>(smem + kBBytes);
  uint32_t *slot = reinterpret_cast<uint32_t *>(smem + kBBytes + 16);
  const int tid = threadIdx.x, warp = tid >> 5;
  uint32_t rank = 0;
  if (PAIR) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  for (int i = tid; i < kBBytes / 4; i += 128) reinterpret_cast<uint32_t *>(smem)[i] = 0x3c003c00u;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (PAIR) {
    __syncthreads();
    cluster_sync();
  }
  const uint32_t tmem = tmem_setup<PAIR>(slot, warp);
  uint32_t av[16];
  for (int i = 0; i < 16; i++) av[i] = 0x3c003c00u;
  const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
  tmem_st16(lane_base + kACol, av);
  tmem_st16(lane_base + kACol + 16, av);
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (PAIR) cluster_sync();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const unsigned long long t0 = clock64();
  if (tid == 0 && rank == 0) {
    const uint64_t bd = sw128_desc(smem_u32(sb));
    const uint32_t idesc = make_idesc(PAIR ? 256 : 128, 256);
    for (int it = 0; it < iters; it++) {
#pragma unroll
      for (int kk = 0; kk < 4; kk++) umma_ts<PAIR>(tmem, tmem + kACol + (uint32_t)(8 * kk), bd + (uint64_t)(2 * kk), idesc, (it || kk) ? 1u : 0u);
    }
    commit_all<PAIR>(bar);
  }
  if (tid == 0) {
    mbar_wait(bar, 0u);
    cycles[blockIdx.x] = clock64() - t0;
  }
  tmem_free<PAIR>(tmem, warp);
}

template <bool PAIR>
static void launch(void (*k)(int, unsigned long long *), int ctas, int smem, int iters, unsigned long long *cyc) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(ctas);
  cfg.blockDim = dim3(128);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = PAIR ? 2 : 1;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  CK(cudaLaunchKernelEx(&cfg, k, iters, cyc));
}

template <bool PAIR>
static void run_rate(int sms, int iters) {
  const int smem = 32768 + 64 + 1024;
  const int ctas = PAIR ? (sms & ~1) : sms;
  CK(cudaFuncSetAttribute(rate_ts_kernel<PAIR>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  unsigned long long *cyc;
  CK(cudaMalloc(&cyc, ctas * 8));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  double best_ms = 1e30;
  for (int rep = 0; rep < 6; rep++) {
    CK(cudaEventRecord(e0));
    launch<PAIR>(rate_ts_kernel<PAIR>, ctas, smem, iters, cyc);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    if (rep && ms < best_ms) best_ms = ms;
  }
  std::vector<unsigned long long> h(ctas);
  CK(cudaMemcpy(h.data(), cyc, ctas * 8, cudaMemcpyDeviceToHost));
  unsigned long long cmax = 0;
  for (auto v : h) cmax = v > cmax ? v : cmax;
  const double ops = 2.0 * 128 * 256 * 16 * 4.0 * iters * ctas;
  printf("{\"kind\": \"f16\", \"form\": \"TS (A in TMEM)\", \"tile\": \"M%dxN256xK16\", \"cta_group\": %d, \"ctas\": %d, \"iters\": %d, \"ms\": %.4f, "
         "\"tops\": %.1f, \"cycles_per_mma\": %.2f}\n",
         PAIR ? 256 : 128, PAIR ? 2 : 1, ctas, iters, best_ms, ops / (best_ms * 1e-3) / 1e12, (double)cmax / (4.0 * iters));
  CK(cudaFree(cyc));
}

template <bool PAIR>
static int run_kat(const char *name) {
  const int M = PAIR ? 256 : 128, N = M;
  std::vector<__half> a(M * 64), b(N * 64);
  srand(11 + PAIR);
  for (auto &v : a) v = __float2half((float)(rand() % 31 - 15));
  for (auto &v : b) v = __float2half((float)(rand() % 255 - 127));
  __half *da, *db;
  float *dd;
  CK(cudaMalloc(&da, a.size() * 2));
  CK(cudaMalloc(&db, b.size() * 2));
  CK(cudaMalloc(&dd, M * N * 4));
  CK(cudaMemset(dd, 0xff, M * N * 4));
  CK(cudaMemcpy(da, a.data(), a.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(db, b.data(), b.size() * 2, cudaMemcpyHostToDevice));
  const int smem = 16384 + 64 + 1024;
  CK(cudaFuncSetAttribute(kat_ts_kernel<PAIR>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(PAIR ? 2 : 1);
  cfg.blockDim = dim3(128);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = PAIR ? 2 : 1;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  CK(cudaLaunchKernelEx(&cfg, kat_ts_kernel<PAIR>, (const __half *)da, (const __half *)db, dd));
  CK(cudaDeviceSynchronize());
  std::vector<float> d(M * N);
  CK(cudaMemcpy(d.data(), dd, d.size() * 4, cudaMemcpyDeviceToHost));
  int bad = 0;
  for (int m = 0; m < M && bad < 8; m++)
    for (int n = 0; n < N && bad < 8; n++) {
      double want = 0;
      for (int k = 0; k < 64; k++) want += (double)__half2float(a[m * 64 + k]) * (double)__half2float(b[n * 64 + k]);
      if ((double)d[m * N + n] != want) { printf("%s: D[%d][%d] = %.1f, want %.1f\n", name, m, n, d[m * N + n], want); bad++; }
    }
  printf("known answer %s: %s\n", name, bad ? "FAILED" : "ok");
  cudaFree(da); cudaFree(db); cudaFree(dd);
  return bad;
}

int main(int argc, char **argv) {
  int iters = argc > 1 ? atoi(argv[1]) : 4000;
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  printf("%s, %d SMs\n", prop.name, prop.multiProcessorCount);
  int bad = 0;
  bad += run_kat<false>("TS M128xN128 cta_group::1");
  bad += run_kat<true>("TS M256xN256 cta_group::2");
  run_rate<false>(prop.multiProcessorCount, iters);
  run_rate<true>(prop.multiProcessorCount, iters);
  return bad ? 1 : 0;
}
