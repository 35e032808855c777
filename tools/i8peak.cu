// i8peak.cu -- tensor-pipe peaks of this box, measured the way the prefill GEMMs use the pipe, plus a
// known-answer check of the tcgen05 kind::i8 path (diagnostics; SURVEY 8d asks for a MEASURED int8 peak).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/i8peak tools/i8peak.cu && /tmp/i8peak
//
//  (1) known answer: D[128 x 128] (s32, TMEM) = A (u8, 0..45) x B^T (s8), K = 128, both operands written by
//      threads in the K-major SWIZZLE_128B layout the UMMA descriptors describe, 4 x tcgen05.mma kind::i8
//      (K = 32 each, start address advanced by 32 bytes inside the swizzle atom), drained with tcgen05.ld and
//      compared on the host.  The same for s8 x s8 and for kind::f16 with exact small integers.
//  (2) peak: every SM issues `iters` x 4 MMAs (M128 x N256 x K32 for i8, x K16 for f16) on operands that stay
//      resident in shared memory -- no loads, no drain -- timed with CUDA events over the whole grid:
//      TOPS = 2 * 128 * 256 * K * 4 * iters * CTAs / time.  Printed as one JSON line per kind.
//  (3) `i8peak <iters> <seconds>`: the same stream held for seconds -- the SUSTAINED peak under the board's power cap.
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
template <int KIND>   // 0: f16, 1: i8
__device__ __forceinline__ void umma(uint32_t d_tmem, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  if (KIND == 1)
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
                 "l"(a), "l"(b), "r"(idesc), "r"(acc)
                 : "memory");
  else
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
                 "l"(a), "l"(b), "r"(idesc), "r"(acc)
                 : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// idesc: c_format [4,6) (1 = F32, 2 = S32); a_format [7,10), b_format [10,13) (i8: 0 = u8, 1 = s8; f16: 0 = F16);
// N >> 3 at [17,23); M >> 4 at [24,29)
__host__ __device__ constexpr uint32_t make_idesc(int cfmt, int afmt, int bfmt, int M, int N) {
  return ((uint32_t)cfmt << 4) | ((uint32_t)afmt << 7) | ((uint32_t)bfmt << 10) | ((uint32_t)(N >> 3) << 17) |
         ((uint32_t)(M >> 4) << 24);
}

// byte offset of (row r, byte column c) in a K-major SWIZZLE_128B tile with 128-byte rows
__host__ __device__ inline uint32_t sw128_off(int r, int c) {
  return (uint32_t)(r >> 3) * 1024u + (uint32_t)(r & 7) * 128u + ((((uint32_t)c >> 4) ^ (uint32_t)(r & 7)) << 4) + ((uint32_t)c & 15u);
}

// ---- (1) known answer ----------------------------------------------------------------------------
// KIND 1: A bytes / B bytes are 8-bit integers (a_signed selects s8 for A); KIND 0: A / B are fp16 (64 K per row)
template <int KIND>
__global__ void __launch_bounds__(128, 1) kat_kernel(const uint8_t *a_rows, const uint8_t *b_rows, int32_t *d_out, int a_signed) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t *sa = smem, *sb = smem + 16384;
  uint64_t *bar = reinterpret_cast<uint64_t *>(smem + 32768);
  uint32_t *slot = reinterpret_cast<uint32_t *>(smem + 32768 + 16);
  const int tid = threadIdx.x, warp = tid >> 5;
  // thread = row: 128 bytes of A and of B into the swizzled layout
  for (int c = 0; c < 128; c += 16) {
    *reinterpret_cast<uint4 *>(sa + sw128_off(tid, c)) = *reinterpret_cast<const uint4 *>(a_rows + tid * 128 + c);
    *reinterpret_cast<uint4 *>(sb + sw128_off(tid, c)) = *reinterpret_cast<const uint4 *>(b_rows + tid * 128 + c);
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *slot;
  if (tid == 0) {
    const uint64_t ad = sw128_desc(smem_u32(sa)), bd = sw128_desc(smem_u32(sb));
    const uint32_t idesc = KIND == 1 ? make_idesc(2, a_signed ? 1 : 0, 1, 128, 128) : make_idesc(1, 0, 0, 128, 128);
#pragma unroll
    for (int kk = 0; kk < 4; kk++) umma<KIND>(tmem, ad + (uint64_t)(2 * kk), bd + (uint64_t)(2 * kk), idesc, kk ? 1u : 0u);
    umma_commit(bar);
  }
  mbar_wait(bar, 0u);
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tbase = tmem + ((uint32_t)(warp * 32) << 16);
  for (int c16 = 0; c16 < 8; c16++) {
    uint32_t v[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(tbase + (uint32_t)(c16 * 16))
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int i = 0; i < 16; i++) d_out[tid * 128 + c16 * 16 + i] = (int32_t)v[i];
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128u) : "memory");
}

// ---- (2) peak ------------------------------------------------------------------------------------
template <int KIND>
__global__ void __launch_bounds__(128, 1) peak_kernel(int iters, unsigned long long *cycles, int N) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t *sa = smem, *sb = smem + 16384;                 // A 128 rows, B 256 rows, 128 bytes per row
  uint64_t *bar = reinterpret_cast<uint64_t *>(smem + 16384 + 32768);
  uint32_t *slot = reinterpret_cast<uint32_t *>(smem + 16384 + 32768 + 16);
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < (16384 + 32768) / 4; i += 128)     // small integers / small fp16 values: no NaN patterns
    reinterpret_cast<uint32_t *>(smem)[i] = KIND == 1 ? 0x01020301u : 0x3c003c00u;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *slot;
  if (tid == 0) {
    const uint64_t ad = sw128_desc(smem_u32(sa)), bd = sw128_desc(smem_u32(sb));
    const uint32_t idesc = KIND == 1 ? make_idesc(2, 0, 1, 128, N) : make_idesc(1, 0, 0, 128, N);
    const unsigned long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
      // two accumulators (columns 0..255 / 256..511) alternate, as a double-buffered mainloop would
      const uint32_t d = tmem + (uint32_t)((it & 1) * 256);
#pragma unroll
      for (int kk = 0; kk < 4; kk++) umma<KIND>(d, ad + (uint64_t)(2 * kk), bd + (uint64_t)(2 * kk), idesc, (it > 1 || kk) ? 1u : 0u);
    }
    umma_commit(bar);
    mbar_wait(bar, 0u);
    cycles[blockIdx.x] = clock64() - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}

// ---- (2b) the CTA-pair form: tcgen05.mma.cta_group::2, M256 x N256 x K16 (kind::f16), operands resident ----------
// Each CTA of the cluster of 2 holds 128 rows of A and 128 of the 256 rows of B; the leader issues; both time.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) peak_pair_kernel(int iters, unsigned long long *cycles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t *sa = smem, *sb = smem + 16384;                 // A 128 rows, B 128 rows (this CTA's half), 128 bytes per row
  uint64_t *bar = reinterpret_cast<uint64_t *>(smem + 32768);
  uint32_t *slot = reinterpret_cast<uint32_t *>(smem + 32768 + 16);
  const int tid = threadIdx.x, warp = tid >> 5;
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  for (int i = tid; i < 32768 / 4; i += 128) reinterpret_cast<uint32_t *>(smem)[i] = 0x3c003c00u;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *slot;
  const unsigned long long t0 = clock64();
  if (tid == 0 && rank == 0) {
    const uint64_t ad = sw128_desc(smem_u32(sa)), bd = sw128_desc(smem_u32(sb));
    const uint32_t idesc = make_idesc(1, 0, 0, 256, 256);
    for (int it = 0; it < iters; it++) {
      const uint32_t d = tmem + (uint32_t)((it & 1) * 256);
#pragma unroll
      for (int kk = 0; kk < 4; kk++)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
                     "l"(ad + (uint64_t)(2 * kk)), "l"(bd + (uint64_t)(2 * kk)), "r"(idesc), "r"((it > 1 || kk) ? 1u : 0u)
                     : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
                 "h"((uint16_t)3)
                 : "memory");
  }
  if (tid == 0) {
    mbar_wait(bar, 0u);
    cycles[blockIdx.x] = clock64() - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}

static void run_peak_pair(int sms, int iters) {
  const int smem = 32768 + 64 + 1024;
  const int ctas = sms & ~1;
  CK(cudaFuncSetAttribute(peak_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  unsigned long long *cyc;
  CK(cudaMalloc(&cyc, ctas * 8));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  double best_ms = 1e30;
  for (int rep = 0; rep < 6; rep++) {
    CK(cudaEventRecord(e0));
    peak_pair_kernel<<<ctas, 128, smem>>>(iters, cyc);
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
  const double ops = 2.0 * 256 * 256 * 16 * 4.0 * iters * (ctas / 2);
  printf("{\"kind\": \"f16\", \"tile\": \"M256xN256xK16\", \"cta_group\": 2, \"ctas\": %d, \"iters\": %d, \"ms\": %.4f, \"tops\": %.1f, "
         "\"cycles_per_mma\": %.2f}\n",
         ctas, iters, best_ms, ops / (best_ms * 1e-3) / 1e12, (double)cmax / (4.0 * iters));
  CK(cudaFree(cyc));
}

template <int KIND>
static void run_peak(const char *name, int sms, int iters, int N) {
  const int smem = 16384 + 32768 + 64 + 1024;
  CK(cudaFuncSetAttribute(peak_kernel<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  unsigned long long *cyc;
  CK(cudaMalloc(&cyc, sms * 8));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  double best_ms = 1e30;
  for (int rep = 0; rep < 6; rep++) {
    CK(cudaEventRecord(e0));
    peak_kernel<KIND><<<sms, 128, smem>>>(iters, cyc, N);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    if (rep && ms < best_ms) best_ms = ms;
  }
  std::vector<unsigned long long> h(sms);
  CK(cudaMemcpy(h.data(), cyc, sms * 8, cudaMemcpyDeviceToHost));
  unsigned long long cmax = 0;
  for (auto v : h) cmax = v > cmax ? v : cmax;
  const int K = KIND == 1 ? 32 : 16;
  const double ops = 2.0 * 128 * N * K * 4.0 * iters * sms;
  printf("{\"kind\": \"%s\", \"tile\": \"M128xN%dxK%d\", \"cta_group\": 1, \"ctas\": %d, \"iters\": %d, \"ms\": %.4f, \"tops\": %.1f, "
         "\"cycles_per_mma\": %.2f}\n",
         name, N, K, sms, iters, best_ms, ops / (best_ms * 1e-3) / 1e12, (double)cmax / (4.0 * iters));
  CK(cudaFree(cyc));
}

template <int KIND>
static int run_kat(const char *name, int a_signed) {
  std::vector<uint8_t> a(128 * 128), b(128 * 128);
  srand(7 + a_signed + 3 * KIND);
  if (KIND == 1) {
    for (auto &v : a) v = a_signed ? (uint8_t)(int8_t)(rand() % 255 - 127) : (uint8_t)(rand() % 46);
    for (auto &v : b) v = (uint8_t)(int8_t)(rand() % 255 - 127);
  } else {
    __half *ha = reinterpret_cast<__half *>(a.data()), *hb = reinterpret_cast<__half *>(b.data());
    for (int i = 0; i < 128 * 64; i++) { ha[i] = __float2half((float)(rand() % 31 - 15)); hb[i] = __float2half((float)(rand() % 255 - 127)); }
  }
  uint8_t *da, *db;
  int32_t *dd;
  CK(cudaMalloc(&da, a.size()));
  CK(cudaMalloc(&db, b.size()));
  CK(cudaMalloc(&dd, 128 * 128 * 4));
  CK(cudaMemcpy(da, a.data(), a.size(), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(db, b.data(), b.size(), cudaMemcpyHostToDevice));
  const int smem = 32768 + 64 + 1024;
  CK(cudaFuncSetAttribute(kat_kernel<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  kat_kernel<KIND><<<1, 128, smem>>>(da, db, dd, a_signed);
  CK(cudaDeviceSynchronize());
  std::vector<int32_t> d(128 * 128);
  CK(cudaMemcpy(d.data(), dd, d.size() * 4, cudaMemcpyDeviceToHost));
  int bad = 0;
  for (int m = 0; m < 128 && bad < 5; m++)
    for (int n = 0; n < 128 && bad < 5; n++) {
      double want = 0;
      if (KIND == 1) {
        for (int k = 0; k < 128; k++)
          want += (a_signed ? (double)(int8_t)a[m * 128 + k] : (double)a[m * 128 + k]) * (double)(int8_t)b[n * 128 + k];
      } else {
        const __half *ha = reinterpret_cast<const __half *>(a.data()), *hb = reinterpret_cast<const __half *>(b.data());
        for (int k = 0; k < 64; k++) want += (double)__half2float(ha[m * 64 + k]) * (double)__half2float(hb[n * 64 + k]);
      }
      double got;
      if (KIND == 1) got = (double)d[m * 128 + n];
      else { float f; memcpy(&f, &d[m * 128 + n], 4); got = f; }
      if (got != want) { printf("%s: D[%d][%d] = %.1f, want %.1f\n", name, m, n, got, want); bad++; }
    }
  printf("known answer %s: %s\n", name, bad ? "FAILED" : "ok");
  cudaFree(da); cudaFree(db); cudaFree(dd);
  return bad;
}

// ---- (3) sustained: the same resident-operand MMA stream for `seconds`, the clock the SMs actually ran at -------------
// A 1 ms burst runs at the maximum clock; under the board's power cap a long tensor-bound kernel does not.  Launches of
// ~20 ms back to back; the second half of the run is timed; MHz = clock64 cycles of the slowest CTA / launch time.
template <int KIND>
static void run_sustained(const char *name, int sms, double seconds) {
  const int smem = 16384 + 32768 + 64 + 1024, iters = 80000;
  CK(cudaFuncSetAttribute(peak_kernel<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  unsigned long long *cyc;
  CK(cudaMalloc(&cyc, sms * 8));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  peak_kernel<KIND><<<sms, 128, smem>>>(iters, cyc, 256);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  peak_kernel<KIND><<<sms, 128, smem>>>(iters, cyc, 256);
  CK(cudaEventRecord(e1));
  CK(cudaEventSynchronize(e1));
  float ms1;
  CK(cudaEventElapsedTime(&ms1, e0, e1));
  const int n = (int)(seconds * 1e3 / ms1 / 2) + 1;
  for (int i = 0; i < n; i++) peak_kernel<KIND><<<sms, 128, smem>>>(iters, cyc, 256);
  CK(cudaEventRecord(e0));
  for (int i = 0; i < n; i++) peak_kernel<KIND><<<sms, 128, smem>>>(iters, cyc, 256);
  CK(cudaEventRecord(e1));
  CK(cudaEventSynchronize(e1));
  float ms;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  std::vector<unsigned long long> h(sms);
  CK(cudaMemcpy(h.data(), cyc, sms * 8, cudaMemcpyDeviceToHost));
  unsigned long long cmax = 0;
  for (auto v : h) cmax = v > cmax ? v : cmax;
  const int K = KIND == 1 ? 32 : 16;
  const double ops = 2.0 * 128 * 256 * K * 4.0 * iters * sms * n;
  printf("{\"kind\": \"%s\", \"sustained_s\": %.2f, \"launches\": %d, \"first_launch_ms\": %.3f, \"ms_per_launch\": %.3f, \"tops\": %.1f, "
         "\"cycles_per_mma\": %.2f, \"sm_mhz_effective\": %.0f}\n",
         name, ms * 1e-3, n, ms1, ms / n, ops / (ms * 1e-3) / 1e12, (double)cmax / (4.0 * iters), (double)cmax / (ms / n * 1e3));
  CK(cudaFree(cyc));
}

int main(int argc, char **argv) {
  int iters = argc > 1 ? atoi(argv[1]) : 4000;
  if (argc > 2) {                                  // i8peak <iters> <seconds>: sustained legs only
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    run_sustained<0>("f16", prop.multiProcessorCount, atof(argv[2]));
    run_sustained<1>("i8", prop.multiProcessorCount, atof(argv[2]));
    return 0;
  }
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  printf("%s, %d SMs, max SM clock %d MHz\n", prop.name, prop.multiProcessorCount, prop.clockRate / 1000);
  int bad = 0;
  bad += run_kat<1>("i8 u8 x s8", 0);
  bad += run_kat<1>("i8 s8 x s8", 1);
  bad += run_kat<0>("f16 small integers", 0);
  run_peak<0>("f16", prop.multiProcessorCount, iters, 256);
  run_peak<1>("i8", prop.multiProcessorCount, iters, 256);
  // the 128-wide tile the cta_group::1 prefill kernels issue: the pipe's cost per MMA does not halve with N
  run_peak<0>("f16", prop.multiProcessorCount, iters, 128);
  run_peak<1>("i8", prop.multiProcessorCount, iters, 128);
  run_peak<1>("i8", prop.multiProcessorCount, iters, 64);
  run_peak_pair(prop.multiProcessorCount, iters);
  return bad ? 1 : 0;
}
