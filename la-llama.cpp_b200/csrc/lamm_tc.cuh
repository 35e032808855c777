// lamm_tc.cuh -- tcgen05 / TMEM / TMA PTX wrappers shared by the prefill GEMM kernels
// (lamm_gemm.cu: exact per-block drain; lamm_gemm_so.cu: scaled-operand mainloop).
#pragma once
#include "lamm_common.cuh"

#include <cuda.h>

namespace lamm {

constexpr int GM = 128, GN = 128, GK = 64;     // CTA tile; K elements per pipeline stage

// ---- tcgen05 / TMA PTX ------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t *smem_dst, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, one thread issues for the CTA
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// the mbarrier gets one arrival when all previously issued MMAs of this thread have completed
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// 32 lanes x 32 consecutive columns (thread i <- lane base + i)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ void tma_load_2d(void *smem_dst, const CUtensorMap *map, int x, int y, uint64_t *bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(x), "r"(y)
      : "memory");
}
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap *map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

// ---- CTA pairs (cta_group::2): two CTAs of a cluster, one MMA stream issued by the leader (rank 0) -----------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_rank(uint32_t local_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t *smem_dst, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma_f16_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                              uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// TS form: the A operand is read from TENSOR MEMORY (each CTA of the pair from its own, at the same column; lane = row,
// word c of a K16 step = elements 2c, 2c + 1), B through a shared-memory descriptor (layout and rate: tools/tmema.cu)
__device__ __forceinline__ void umma_f16_pair_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                                 uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t"
      "}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// 16 consecutive columns of the calling thread's TMEM lane (warp w owns lanes 32 (w % 4) ..)
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(
          taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
      "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// one arrival on the barrier at this offset in BOTH CTAs of the pair when all MMAs issued so far have completed
__device__ __forceinline__ void umma_commit_pair(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"((uint16_t)3)
               : "memory");
}
// data into THIS CTA's shared memory, completion bytes onto the barrier at cluster address `bar_cluster`
__device__ __forceinline__ void tma_load_2d_pair(void *smem_dst, const CUtensorMap *map, int x, int y, uint32_t bar_cluster) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(map), "r"(bar_cluster), "r"(x), "r"(y)
      : "memory");
}

// arrive on a barrier of another CTA of the cluster (address from mapa_rank).  Default semantics (release at CTA scope):
// what the remote consumer reads afterwards is this CTA's shared memory through the async proxy (the pair's MMA), which
// the caller's fence.proxy.async has already ordered; a cluster-scope release costs a MEMBAR per arrive (measured:
// 17 % of the peer's unpack time, profiles/r2v).
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar_cluster) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(bar_cluster) : "memory");
}
// wait that also acquires what threads of the peer CTA released before arriving
__device__ __forceinline__ void mbar_wait_cluster(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "LAMM_CWAIT_%=:\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra LAMM_CDONE_%=;\n\t"
      "bra LAMM_CWAIT_%=;\n\t"
      "LAMM_CDONE_%=:\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// K-major SWIZZLE_128B operand tile: 8-row groups 1024 B apart, rows 128 B
__device__ __forceinline__ uint64_t make_sw128_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);        // start address
  d |= (uint64_t)1 << 16;                              // leading byte offset (unused for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;                    // stride byte offset
  d |= (uint64_t)1 << 46;                              // descriptor version (sm_100)
  d |= (uint64_t)2 << 61;                              // SWIZZLE_128B
  return d;
}
// kind::f16, A/B = F16 K-major, D = F32, M = 128, N = 128
constexpr uint32_t kIdescF16 = (1u << 4) | ((uint32_t)(GN >> 3) << 17) | ((uint32_t)(GM >> 4) << 24);
// CTA pair: M = 256 (128 rows per CTA), N = 256
constexpr uint32_t kIdescF16Pair = (1u << 4) | ((uint32_t)(256 >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
// same with A/B = BF16 (a_format = b_format = 1, bits [7,10) and [10,13))
constexpr uint32_t kIdescBF16 = kIdescF16 | (1u << 7) | (1u << 10);

__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  uint64_t ra, rb, rc, rd;
  asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(rc) : "f"(c.x), "f"(c.y));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
  float2 d;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
  return d;
}
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
  uint64_t ra, rb, rd;
  asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
  float2 d;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
  return d;
}

// explicit shared-space loads (the carved-up dynamic shared memory is reached through generic
// pointers otherwise, and generic loads are far slower than LDS)
__device__ __forceinline__ float lds_f32(uint32_t addr) {
  float v;
  asm("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
  float4 v;
  asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}


// two fp16 holding the small integers in the low bits of each half of p: (1024 + v) - bias
__device__ __forceinline__ uint32_t magic_pair(uint32_t p, uint32_t bias_h2) {
  const uint32_t m = p | 0x64006400u;
  const __half2 r = __hsub2(*reinterpret_cast<const __half2 *>(&m), *reinterpret_cast<const __half2 *>(&bias_h2));
  return *reinterpret_cast<const uint32_t *>(&r);
}

// ---- weight unpack: one 32-block of the thread's row -> 4 x 16 B chunks of fp16 ------------
template <int OFF, int NWR>
__device__ __forceinline__ uint32_t raw_word(const uint32_t (&w)[NWR]) {
  if constexpr (OFF % 4 == 0) return w[OFF / 4];
  else return __byte_perm(w[OFF / 4], w[OFF / 4 + 1], 0x5432);
}
template <int OFF, int NWR>
__device__ __forceinline__ float raw_half(const uint32_t (&w)[NWR]) {
  if constexpr (OFF % 4 == 0) return half_bits_to_float(w[OFF / 4] & 0xffffu);
  else return half_bits_to_float(w[OFF / 4] >> 16);
}

// JB: block index inside the thread's 8-block raw chunk.  out[w] = chunk w (8 fp16) in the
// element order documented in lamm_gemm_prep.cu; d / m = f32 of the block's scale / min.
template <int WT, int JB, int NWR>
__device__ __forceinline__ void unpack_block(const uint32_t (&raw)[NWR], uint4 (&out)[4], float &d, float &m) {
  constexpr int BS = type_info(WT).size;
  constexpr int O = JB * BS;
  d = raw_half<O, NWR>(raw);
  m = 0.f;
  if constexpr (WT == LAMMCU_TYPE_Q4_0 || WT == LAMMCU_TYPE_Q4_1) {
    constexpr int QS = O + (WT == LAMMCU_TYPE_Q4_0 ? 2 : 4);
    constexpr uint32_t bias = (WT == LAMMCU_TYPE_Q4_0) ? 0x64086408u : 0x64006400u;     // (q - 8) or q
    if constexpr (WT == LAMMCU_TYPE_Q4_1) m = raw_half<O + 2, NWR>(raw);
    const uint32_t q[4] = {raw_word<QS, NWR>(raw), raw_word<QS + 4, NWR>(raw), raw_word<QS + 8, NWR>(raw),
                           raw_word<QS + 12, NWR>(raw)};
#pragma unroll
    for (int w = 0; w < 4; w++) {
      out[w].x = magic_pair(q[w] & 0x000F000Fu, bias);
      out[w].y = magic_pair((q[w] >> 4) & 0x000F000Fu, bias);
      out[w].z = magic_pair((q[w] >> 8) & 0x000F000Fu, bias);
      out[w].w = magic_pair((q[w] >> 12) & 0x000F000Fu, bias);
    }
  } else if constexpr (WT == LAMMCU_TYPE_Q5_0 || WT == LAMMCU_TYPE_Q5_1) {
    constexpr int QH = O + (WT == LAMMCU_TYPE_Q5_0 ? 2 : 4);
    constexpr int QS = QH + 4;
    constexpr uint32_t bias = (WT == LAMMCU_TYPE_Q5_0) ? 0x64106410u : 0x64006400u;     // (q - 16) or q
    if constexpr (WT == LAMMCU_TYPE_Q5_1) m = raw_half<O + 2, NWR>(raw);
    const uint32_t qh = raw_word<QH, NWR>(raw);
    const uint32_t q[4] = {raw_word<QS, NWR>(raw), raw_word<QS + 4, NWR>(raw), raw_word<QS + 8, NWR>(raw),
                           raw_word<QS + 12, NWR>(raw)};
#pragma unroll
    for (int w = 0; w < 4; w++) {
      const uint32_t h = qh >> (4 * w), g = qh >> (16 + 4 * w);
      out[w].x = magic_pair((q[w] & 0x000F000Fu) | ((h & 1u) << 4) | ((h & 4u) << 18), bias);
      out[w].y = magic_pair(((q[w] >> 4) & 0x000F000Fu) | ((g & 1u) << 4) | ((g & 4u) << 18), bias);
      out[w].z = magic_pair(((q[w] >> 8) & 0x000F000Fu) | ((h & 2u) << 3) | ((h & 8u) << 17), bias);
      out[w].w = magic_pair(((q[w] >> 12) & 0x000F000Fu) | ((g & 2u) << 3) | ((g & 8u) << 17), bias);
    }
  } else {   // Q8_0: natural order, (b ^ 0x80) | 0x6400 = 1024 + 128 + b
    constexpr int QS = O + 2;
    const uint32_t q[8] = {raw_word<QS, NWR>(raw),      raw_word<QS + 4, NWR>(raw),  raw_word<QS + 8, NWR>(raw),
                           raw_word<QS + 12, NWR>(raw), raw_word<QS + 16, NWR>(raw), raw_word<QS + 20, NWR>(raw),
                           raw_word<QS + 24, NWR>(raw), raw_word<QS + 28, NWR>(raw)};
    constexpr uint32_t bias = 0x64806480u;
#pragma unroll
    for (int w = 0; w < 4; w++) {
      const uint32_t a = q[2 * w] ^ 0x80808080u, b = q[2 * w + 1] ^ 0x80808080u;
      const uint32_t a0 = __byte_perm(a, 0x64646464u, 0x5140), a1 = __byte_perm(a, 0x64646464u, 0x7362);
      const uint32_t b0 = __byte_perm(b, 0x64646464u, 0x5140), b1 = __byte_perm(b, 0x64646464u, 0x7362);
      const __half2 bh = *reinterpret_cast<const __half2 *>(&bias);
      __half2 r0 = __hsub2(*reinterpret_cast<const __half2 *>(&a0), bh), r1 = __hsub2(*reinterpret_cast<const __half2 *>(&a1), bh);
      __half2 r2 = __hsub2(*reinterpret_cast<const __half2 *>(&b0), bh), r3 = __hsub2(*reinterpret_cast<const __half2 *>(&b1), bh);
      out[w].x = *reinterpret_cast<uint32_t *>(&r0);
      out[w].y = *reinterpret_cast<uint32_t *>(&r1);
      out[w].z = *reinterpret_cast<uint32_t *>(&r2);
      out[w].w = *reinterpret_cast<uint32_t *>(&r3);
    }
  }
}


// host: cuTensorMapEncodeTiled through the runtime's driver entry point (lamm_gemm.cu)
int make_map_2d(CUtensorMap *map, CUtensorMapDataType dt, int elem_bytes, const void *base, uint64_t dim0,
                uint64_t dim1, uint64_t stride1_bytes, uint32_t box0, uint32_t box1, CUtensorMapSwizzle sw);

}  // namespace lamm
