// lamm_gemm_x.cu -- Q2_K prefill GEMM on tcgen05 kind::i8: every per-super-block INTEGER sum of the reference
// (isum, summs) is computed exactly on the int8 tensor pipe and scaled once per 256-element super-block.
//
// Replaces LAMMImpl<Q2_K>::matmul_simd_block + the Q2_K block kernel (src/lamm_impl.hpp:90-147,
// src/lamm_kernel_q2_k.hpp:163-307) for many activation rows -- the fold of SURVEY 7.3c-ii:
//
//   isum  = sum_k W'[m,k] q8[n,k],  W'[m,k] = (scales[g(k)] & 15) * q2[m,k]  in [0, 45]  (u8 x s8 -> s32):
//           8 MMAs (M128 N128 K32) per super-block into ONE TMEM accumulator;
//   summs = sum_g (scales[g] >> 4) * bsums[n,g]: with bsums = 16 hi + lo (lo in [0,15], hi = bsums >> 4 in
//           [-127, 127]) this is ONE more K = 32 MMA per super-block, A'' = [min_0..15, 16 min_0..15] (u8),
//           B'' = [lo_0..15, hi_0..15] (s8), into a second accumulator;
//   C[n,m] += x.d * (f(d) * isum - f(dmin) * summs)      once per super-block (src/lamm_kernel_q2_k.hpp:68-74).
// No weight dequantize pass, no fp16 rounding anywhere: the integers are the reference's, the scales only meet
// in f32 (so any fp16 d / dmin is safe).
//
// One CTA = 128 weight rows (TMEM lanes) x 128 activation rows (TMEM columns), 28 warps:
//   warp 0       TMA producer: int8 activation tiles (SWIZZLE_128B), x.d rows, the min-term B'' tiles
//   warp 1       MMA issuer (one thread); tcgen05.commit hands accumulators to the drain and stages back
//   warp 2       TMEM allocation (512 columns: a ring of 2 x {isum, summs})
//   warps 4-11   weight unpack, thread = (row, half of the super-block): packed 2-bit quants * 4-bit scales -> u8
//                bytes in the UMMA K-major SWIZZLE_128B layout; the packed rows are staged by coalesced cp.async
//   warps 12-27  drain, thread = TMEM lane (weight row) x 32 columns: tcgen05.ld, s32 -> f32, scale, accumulate
//
// Measured and removed (round 2): the same skeleton for the 32-element formats on kind::f16 with exact small
// integers and a drain per 32-block: 195-273 TOPS at N = 512, below the scaled-operand kernel's 274-348 -- the
// drain costs ~110-136 instructions per warp per block against a 128-cycle MMA (DESIGN d3).
#include "lamm_tc.cuh"

#include <cstdlib>

namespace lamm {

constexpr int kXThreads = 896;             // 4 control + 8 unpack + 16 drain warps
constexpr int kXDrainWarps = 16;

struct XParams {
  const float *dxT;           // [K / 256][Npad]: x.d of super-block u of activation row n
  float *c;
  int64_t ldc;
  const uint8_t *w;           // packed Q2_K rows (read by the unpack threads, any 4-byte alignment)
  int64_t w_stride;
  int M, N, Npad, K;
  float *cp[LAMMCU_MAX_RANKS];   // row-sharded prefill: every rank's full C (already at this rank's first column)
  int n_peer;
};

template <bool PEERS>
__device__ __forceinline__ void x_store(const XParams &p, int64_t idx, float v) {
  if constexpr (!PEERS) {
    p.c[idx] = v;
  } else {
#pragma unroll
    for (int g = 0; g < LAMMCU_MAX_RANKS; g++)
      if (g < p.n_peer) p.cp[g][idx] = v;
  }
}

__device__ __forceinline__ void umma_i8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// kind::i8: D = S32 (c_format 2), A = u8 (0), B = s8 (1), K-major, M = 128, N = 128
constexpr uint32_t kIdescI8 = (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t)(GN >> 3) << 17) | ((uint32_t)(GM >> 4) << 24);

__device__ __forceinline__ void mbar_wait_addr(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "LAMM_XW_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra LAMM_XD_%=;\n\t"
      "bra LAMM_XW_%=;\n\t"
      "LAMM_XD_%=:\n\t"
      "}" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_addr(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// packed f32x2 math with one scalar operand
__device__ __forceinline__ float2 fmul2s(float a, float2 b) { return fmul2(make_float2(a, a), b); }
__device__ __forceinline__ float2 ffma2s(float a, float2 b, float2 c) { return ffma2(make_float2(a, a), b, c); }

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

// ---- shared-memory map (offsets from the 1024-aligned base) --------------------------------------------
constexpr int kXStages = 4;                      // 128-K operand stages (A 16 KB + B 16 KB each)
constexpr int kXTRing = 2;                       // TMEM accumulator ring: 2 x {isum 128 columns, summs 128 columns}
constexpr int kXSRing = 8;                       // super-blocks of scales kept in shared memory
constexpr int kXTile = 16384;
constexpr int kXAOff = 0;
constexpr int kXBOff = kXAOff + kXStages * kXTile;
constexpr int kXA2Off = kXBOff + kXStages * kXTile;        // min-term operand tiles: 4 super-blocks per 128-byte row, x 2
constexpr int kXB2Off = kXA2Off + 2 * kXTile;
constexpr int kXDxOff = kXB2Off + 2 * kXTile;              // x.d      [ring][128 activation rows]
constexpr int kXDwOff = kXDxOff + kXSRing * GN * 4;        // f(d)     [ring][128 weight rows]
constexpr int kXMwOff = kXDwOff + kXSRing * GM * 4;        // f(dmin)  [ring][128 weight rows]
constexpr int kXRawOff = kXMwOff + kXSRing * GM * 4;        // raw packed super-blocks [2][128 rows][84 bytes]
constexpr int kXRawUnit = GM * 84;
constexpr int kXBarOff = kXRawOff + 2 * kXRawUnit;
constexpr int kXTotal = kXBarOff + 512;

// ---- Q2_K: one half (128 elements) of a super-block of the thread's row -> int8 operand bytes -----------
// Element k of the super-block (src/lamm_kernel_q2_k.hpp:49-71): half n = k / 128, plane j = (k % 128) / 32,
// l = k % 32: q2 = (qs[32 n + l] >> 2 j) & 3, sub-block g = k / 16.  The 16-byte chunk c (0..7) of half n's
// 128-byte operand row is sub-block 8 n + c: bytes l0 .. l0 + 15 of plane c / 2 (l0 = 16 (c & 1)) times
// (scales[8 n + c] & 15) -- a byte-wise product that cannot carry (3 * 15 = 45).
// qs: the half's 8 words (qs[32 n ..]); sc: the half's 8 scale bytes (2 words).
__device__ __forceinline__ void q2k_half_to_i8(const uint32_t (&qs)[8], const uint32_t (&sc)[2], uint32_t row_base, uint32_t r7) {
#pragma unroll
  for (int c = 0; c < 8; c++) {
    const uint32_t s = (sc[c >> 2] >> (8 * (c & 3))) & 15u;
    uint32_t o[4];
#pragma unroll
    for (int e = 0; e < 4; e++) o[e] = ((qs[4 * (c & 1) + e] >> (2 * (c >> 1))) & 0x03030303u) * s;
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(row_base + (((uint32_t)c ^ r7) << 4)), "r"(o[0]), "r"(o[1]),
                 "r"(o[2]), "r"(o[3])
                 : "memory");
  }
}

template <bool PEERS>
__global__ void __launch_bounds__(kXThreads, 1)
gemm_q2k_i8_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_b2, const XParams p) {
  constexpr int S = kXStages, T = kXTRing, R = kXSRing;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + kXBarOff);
  uint64_t *full_b = bars, *full_a = bars + 4, *empty = bars + 8, *t_full = bars + 12, *t_empty = bars + 14,
           *mb_full = bars + 16, *m_empty = bars + 18;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 24);
  float *dxs = reinterpret_cast<float *>(smem + kXDxOff);
  float *dws = reinterpret_cast<float *>(smem + kXDwOff), *mws = reinterpret_cast<float *>(smem + kXMwOff);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tiles = (p.N + GN - 1) / GN;
  const int m0 = (blockIdx.x / n_tiles) * GM, n0 = (blockIdx.x % n_tiles) * GN;
  const int n_units = p.K / 256, n_stages = 2 * n_units;

  if (threadIdx.x == 0) {
    for (int i = 0; i < S; i++) {
      mbar_init(&full_b[i], 1);
      mbar_init(&full_a[i], 4);          // the four unpack warps of the stage's half
      mbar_init(&empty[i], 1);           // MMA commit
    }
    for (int i = 0; i < 2; i++) {
      mbar_init(&mb_full[i], 1);
      mbar_init(&m_empty[i], 1);
    }
    for (int i = 0; i < T; i++) {
      mbar_init(&t_full[i], 1);
      mbar_init(&t_empty[i], kXDrainWarps);
    }
    fence_barrier_init();
    prefetch_tmap(&tm_x);
    prefetch_tmap(&tm_b2);
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // Registers: 896 threads start with 72 each.  The control warps give 32 back, the drain warps take 8 more.
  // setmaxnreg sits at the top of each warpgroup's own branch (ptxas budgets registers per dominated region) and
  // the increase is served from the CTA's OWN pool -- what its other warpgroups have released, not the SM's
  // unallocated registers: 128 x (72 - 40) released == 512 x (80 - 72) requested (asking for more HANGS: the
  // first version of this kernel asked for 96 from a pool of 5120 and never returned).
  if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
    if (warp == 0) {
      // ===================== TMA producer =====================
      if (lane == 0) {
        for (int st = 0; st < n_stages; st++) {
          if ((st & 7) == 0) {                                    // min-term B'' tile: 4 super-blocks
            const int g = st >> 3, mt = g & 1;
            mbar_wait(&m_empty[mt], ((uint32_t)(g >> 1) & 1u) ^ 1u);
            mbar_arrive_expect_tx(&mb_full[mt], (uint32_t)kXTile);
            tma_load_2d(smem + kXB2Off + mt * kXTile, &tm_b2, g * 128, n0, &mb_full[mt]);
          }
          const int s = st % S;
          mbar_wait(&empty[s], ((uint32_t)(st / S) & 1u) ^ 1u);
          const bool first = (st & 1) == 0;                       // the unit's x.d row rides on its first stage
          mbar_arrive_expect_tx(&full_b[s], (uint32_t)(kXTile + (first ? GN * 4 : 0)));
          tma_load_2d(smem + kXBOff + s * kXTile, &tm_x, st * 128, n0, &full_b[s]);
          if (first) {
            const int u = st >> 1;
            bulk_g2s(dxs + (u % R) * GN, p.dxT + (int64_t)u * p.Npad + n0, GN * 4, &full_b[s]);
          }
        }
      }
    } else if (warp == 1) {
      // ===================== MMA issuer =====================
      if (lane == 0) {
        for (int st = 0; st < n_stages; st++) {
          const int s = st % S;
          const uint32_t ph = (uint32_t)(st / S) & 1u;
          mbar_wait(&full_a[s], ph);
          mbar_wait(&full_b[s], ph);
          tc_fence_after();
          const uint64_t adesc = make_sw128_desc(smem_u32(smem + kXAOff + s * kXTile));
          const uint64_t bdesc = make_sw128_desc(smem_u32(smem + kXBOff + s * kXTile));
          const int u = st >> 1, tb = u % T, h = st & 1;
          if (h == 0) {
            mbar_wait(&t_empty[tb], ((uint32_t)(u / T) & 1u) ^ 1u);
            tc_fence_after();
          }
          const uint32_t d = tmem_base + (uint32_t)(tb * 2 * GN);
#pragma unroll
          for (int kk = 0; kk < 4; kk++) umma_i8(d, adesc + (uint64_t)(2 * kk), bdesc + (uint64_t)(2 * kk), kIdescI8, (h | kk) ? 1u : 0u);
          if (h == 1) {
            const int g = u >> 2, mt = g & 1, q = u & 3;
            if (q == 0) {
              mbar_wait(&mb_full[mt], (uint32_t)(g >> 1) & 1u);
              tc_fence_after();
            }
            const uint64_t a2 = make_sw128_desc(smem_u32(smem + kXA2Off + mt * kXTile));
            const uint64_t b2 = make_sw128_desc(smem_u32(smem + kXB2Off + mt * kXTile));
            umma_i8(d + (uint32_t)GN, a2 + (uint64_t)(2 * q), b2 + (uint64_t)(2 * q), kIdescI8, 0u);
            umma_commit(&t_full[tb]);
            if (q == 3 || u == n_units - 1) umma_commit(&m_empty[mt]);
          }
          umma_commit(&empty[s]);
        }
      }
    }
  } else if (warp < 12) {
    // ===================== weight unpack: thread = (row, half of the super-block) =====================
    // Warps 4-7 turn the first half (elements 0..127) of every super-block of their 128 rows into stage 2u, warps
    // 8-11 the second half into stage 2u + 1 and the min-term bytes.  (84-byte super-blocks are 4-byte aligned at
    // best and row strides are not multiples of 16 for K % 1024 != 0 -- no TMA box: the warps stage them themselves.)
    const int h = (warp - 4) >> 2;
    const int r = (threadIdx.x - 128) & 127;
    const uint32_t r7 = (uint32_t)(r & 7);
    const uint32_t row_off = (uint32_t)(r >> 3) * 1024u + r7 * 128u;
    // Raw weights: a warp copies 16 rows per super-block, lanes 0..20 one 32-bit word each (cp.async, 84 contiguous
    // bytes = 1-2 cache lines per instruction), into a staging slot whose 21-word row pitch is conflict-free for
    // the per-row reads that follow.  (A first version had every thread __ldg its own row: 32 cache lines per
    // warp-load, 13 loads per unit -- the L1 data pipe was 60 % busy with them and the kernel ran at 365-627 TOPS.)
    // Rows beyond M re-read row M - 1 (their TMEM lanes are never stored).  Two slots: unit u + 2 is requested as
    // soon as every thread has its words of unit u in registers.
    const int uw = warp - 4;                                      // 0..7: rows 16 uw .. 16 uw + 15
    const uint32_t raw_a = smem_u32(smem + kXRawOff);
    // rows of this warp that exist; the copy pointer steps by the row stride only while the next row exists
    const int rows_valid = max(1, min(16, p.M - (m0 + 16 * uw)));
    const uint8_t *ubase = p.w + (int64_t)min(m0 + 16 * uw, p.M - 1) * p.w_stride + lane * 4;    // unit 0, row 0 of the warp
    const uint32_t dst0 = raw_a + (uint32_t)(16 * uw * 84 + lane * 4);
    const int64_t wstride = p.w_stride;
    auto request = [&](int u) {
      if (u < n_units && lane < 21) {
        const uint8_t *src = ubase + (int64_t)u * 84;
        const uint32_t dst = dst0 + (uint32_t)((u & 1) * kXRawUnit);
#pragma unroll
        for (int i = 0; i < 16; i++) {
          asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst + (uint32_t)(i * 84)), "l"(src) : "memory");
          src += (i + 1 < rows_valid) ? wstride : 0;
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");        // every thread commits (possibly empty) groups
    };
    request(0);
    request(1);
    for (int u = 0; u < n_units; u++) {
      asm volatile("cp.async.wait_group 1;" ::: "memory");        // this thread's copies of unit u have landed
      asm volatile("bar.sync 1, 256;" ::: "memory");              // ... and everybody else's
      // words of a super-block: [0, 4) scales, [4, 20) qs, 20 = {d, dmin}.  This thread needs the half's 8 qs words,
      // all 4 scale words (h == 1 builds the min-term bytes from them) and, for h == 0, the {d, dmin} word.
      uint32_t qs[8], sc[4], dd;
      const uint32_t rb = raw_a + (uint32_t)((u & 1) * kXRawUnit + r * 84);
#pragma unroll
      for (int i = 0; i < 4; i++) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(sc[i]) : "r"(rb + 4u * i));
#pragma unroll
      for (int i = 0; i < 8; i++) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(qs[i]) : "r"(rb + 16u + 32u * h + 4u * i));
      asm volatile("ld.shared.u32 %0, [%1];" : "=r"(dd) : "r"(rb + 80u));
      asm volatile("bar.sync 2, 256;" ::: "memory");              // the slot has been read by all: refill it
      request(u + 2);
      const int st = 2 * u + h, s = st % S;
      const int g = u >> 2, mt = g & 1, q = u & 3;
      if (h == 1 && q == 0) mbar_wait(&m_empty[mt], ((uint32_t)(g >> 1) & 1u) ^ 1u);
      mbar_wait(&empty[s], ((uint32_t)(st / S) & 1u) ^ 1u);
      const uint32_t sch[2] = {h ? sc[2] : sc[0], h ? sc[3] : sc[1]};     // (a run-time index would put sc[] in local memory)
      q2k_half_to_i8(qs, sch, smem_u32(smem + kXAOff + s * kXTile) + row_off, r7);
      if (h == 0) {
        const uint32_t slot = (uint32_t)((u % R) * GM + r) * 4u;
        asm volatile("st.shared.f32 [%0], %1;" ::"r"(smem_u32(dws) + slot), "f"(half_bits_to_float(dd & 0xffffu)) : "memory");
        asm volatile("st.shared.f32 [%0], %1;" ::"r"(smem_u32(mws) + slot), "f"(half_bits_to_float(dd >> 16)) : "memory");
      } else {
        // A'': chunk 2 q = the 16 mins, chunk 2 q + 1 = 16 x the mins (both fit u8: <= 15 / <= 240)
        const uint32_t a2 = smem_u32(smem + kXA2Off + mt * kXTile) + row_off;
        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(a2 + (((uint32_t)(2 * q) ^ r7) << 4)),
                     "r"((sc[0] >> 4) & 0x0F0F0F0Fu), "r"((sc[1] >> 4) & 0x0F0F0F0Fu), "r"((sc[2] >> 4) & 0x0F0F0F0Fu),
                     "r"((sc[3] >> 4) & 0x0F0F0F0Fu)
                     : "memory");
        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(a2 + (((uint32_t)(2 * q + 1) ^ r7) << 4)),
                     "r"(sc[0] & 0xF0F0F0F0u), "r"(sc[1] & 0xF0F0F0F0u), "r"(sc[2] & 0xF0F0F0F0u), "r"(sc[3] & 0xF0F0F0F0u)
                     : "memory");
      }
      fence_proxy_async();               // generic-proxy stores -> visible to the tensor core
      __syncwarp();
      if (lane == 0) mbar_arrive(&full_a[s]);
    }
  } else {
    // ===================== drain (thread = weight row x 32 activation rows) =====================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 80;");
    const int q4 = warp & 3, cs = (warp - 12) >> 2;
    const int row = q4 * 32 + lane;
    float2 acc[16];
#pragma unroll
    for (int i = 0; i < 16; i++) acc[i] = make_float2(0.f, 0.f);
    const uint32_t dxs_a = smem_u32(dxs) + (uint32_t)(cs * 32) * 4u;
    const uint32_t dws_a = smem_u32(dws) + (uint32_t)row * 4u, mws_a = smem_u32(mws) + (uint32_t)row * 4u;
    const uint32_t tlane = tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(cs * 32);
    const uint32_t tfull_a = smem_u32(t_full), tempty_a = smem_u32(t_empty);
#pragma unroll 1
    for (int u = 0; u < n_units; u++) {
      const uint32_t tb = (uint32_t)u & 1u, slot = (uint32_t)u & (uint32_t)(R - 1);
      mbar_wait_addr(tfull_a + 8u * tb, ((uint32_t)u >> 1) & 1u);
      tc_fence_after();
      const float wd = lds_f32(dws_a + slot * (uint32_t)(GM * 4)), wm = -lds_f32(mws_a + slot * (uint32_t)(GM * 4));
      const uint32_t dx_a = dxs_a + slot * (uint32_t)(GN * 4);
      const uint32_t tb0 = tlane + tb * (uint32_t)(2 * GN);
#pragma unroll
      for (int hh = 0; hh < 2; hh++) {
        uint32_t i1[16], i2[16];
        tmem_ld16(tb0 + (uint32_t)(16 * hh), i1);
        tmem_ld16(tb0 + (uint32_t)(GN + 16 * hh), i2);
        tmem_ld_wait();
        if (hh == 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_addr(tempty_a + 8u * tb);
        }
#pragma unroll
        for (int c = 0; c < 4; c++) {
          const float4 xd = lds_f4(dx_a + 16u * (4 * hh + c));
          // C += x.d * (d * isum - dmin * summs): src/lamm_kernel_q2_k.hpp:68-74 regrouped; the integers are exact.
          // (s32 -> f32 by I2FP on the XU pipe.  A magic-number conversion -- as_float(v + 0x4B400000) - 1.5 * 2^23 on the
          // ALU / FMA pipes -- for both sums or for summs only measured within +-2 %: profiles/r2l/.)
          float2 t0 = fmul2s(wd, make_float2((float)(int)i1[4 * c], (float)(int)i1[4 * c + 1]));
          float2 t1 = fmul2s(wd, make_float2((float)(int)i1[4 * c + 2], (float)(int)i1[4 * c + 3]));
          t0 = ffma2s(wm, make_float2((float)(int)i2[4 * c], (float)(int)i2[4 * c + 1]), t0);
          t1 = ffma2s(wm, make_float2((float)(int)i2[4 * c + 2], (float)(int)i2[4 * c + 3]), t1);
          acc[8 * hh + 2 * c] = ffma2(make_float2(xd.x, xd.y), t0, acc[8 * hh + 2 * c]);
          acc[8 * hh + 2 * c + 1] = ffma2(make_float2(xd.z, xd.w), t1, acc[8 * hh + 2 * c + 1]);
        }
      }
    }
    // epilogue: C[n * ldc + m]; lanes = consecutive m
    const int m = m0 + row;
    if (m < p.M) {
#pragma unroll
      for (int c = 0; c < 16; c++) {
        const int n = n0 + cs * 32 + 2 * c;
        if (n < p.N) x_store<PEERS>(p, (int64_t)n * p.ldc + m, acc[c].x);
        if (n + 1 < p.N) x_store<PEERS>(p, (int64_t)(n + 1) * p.ldc + m, acc[c].y);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 512);
}


// ---- activation operands ---------------------------------------------------------------------------------
// Xq (ggml block_q8_K rows, the bit-exact output of lamm_quantize.cu) ->
//   X8  : int8 [Npad][K]          the quants, K contiguous (TMA-able: the 292-byte blocks are not)
//   XB2 : int8 [Npad][K / 8]      per super-block 32 bytes: lo_0..15 = bsums & 15, hi_0..15 = bsums >> 4
//   xdT : f32  [K / 256][Npad]    x.d
// rows n >= N are zeros.  One warp per (row, super-block).
__global__ void __launch_bounds__(256) prep_q8k_i8_kernel(const uint8_t *__restrict__ xq, int64_t xq_stride,
                                                         uint8_t *__restrict__ x8, uint8_t *__restrict__ xb2,
                                                         float *__restrict__ xdT, int N, int Npad, int K) {
  const int lane = threadIdx.x & 31;
  const int nsb = K / 256;
  const int64_t total = (int64_t)Npad * nsb;
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t i = warp0; i < total; i += n_warps) {
    const int n = (int)(i / nsb), sb = (int)(i - (int64_t)n * nsb);
    uint32_t q0 = 0u, q1 = 0u;
    int bs = 0;
    float d = 0.f;
    if (n < N) {
      const uint8_t *blk = xq + (int64_t)n * xq_stride + (int64_t)sb * 292;
      const uint32_t *qs = reinterpret_cast<const uint32_t *>(blk + Q8_K_QS + lane * 8);
      q0 = qs[0];
      q1 = qs[1];
      if (lane < 16) bs = *reinterpret_cast<const int16_t *>(blk + Q8_K_BSUMS + 2 * lane);
      if (lane == 0) d = *reinterpret_cast<const float *>(blk);
    }
    *reinterpret_cast<uint2 *>(x8 + (int64_t)n * K + (int64_t)sb * 256 + lane * 8) = make_uint2(q0, q1);
    // lanes 0..15 hold bsums[lane]: byte `lane` = lo, byte 16 + lane = hi (arithmetic shift: bsums = 16 hi + lo)
    uint8_t *b2 = xb2 + (int64_t)n * (K / 8) + (int64_t)sb * 32;
    if (lane < 16) {
      b2[lane] = (uint8_t)(bs & 15);
      b2[16 + lane] = (uint8_t)(signed char)(bs >> 4);
    }
    if (lane == 0) xdT[(int64_t)sb * Npad + n] = d;
  }
}

// ---- host side ----------------------------------------------------------------------------------------
bool gemm_x_supported(int w_type, const void *w, int64_t w_stride, int64_t M, int64_t N, int64_t K) {
  if (w_type != LAMMCU_TYPE_Q2_K) return false;
  if (K <= 0 || K % 256 || M <= 0 || N <= 0 || M > 0x3fffffff || N > 0x3fffffff) return false;
  return (((uintptr_t)w | (uintptr_t)w_stride) & 3) == 0;
}

size_t gemm_x_ws_bytes(int64_t N, int64_t K) {
  const int64_t npad = (N + GN - 1) / GN * GN;
  return (size_t)(npad * K + npad * (K / 8) + (K / 256) * npad * 4 + 1024);
}

// xq: quantized activations (ggml block_q8_K rows, row stride xq_stride); ws: gemm_x_ws_bytes() of scratch
int launch_gemm_x(int w_type, const void *w, int64_t w_stride, const void *xq, int64_t xq_stride, float *c, int64_t ldc,
                  int64_t M, int64_t N, int64_t K, void *ws, cudaStream_t st) {
  if (w_type != LAMMCU_TYPE_Q2_K) { set_error("gemm_x: weight type %d not supported", w_type); return LAMMCU_ERR_TYPE; }
  const int64_t npad = (N + GN - 1) / GN * GN;
  uint8_t *base = (uint8_t *)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
  uint8_t *x8 = base, *xb2 = base + npad * K;
  float *xdT = (float *)(xb2 + npad * (K / 8));
  {
    const int64_t warps = npad * (K / 256);
    int blocks = (int)((warps + 7) / 8);
    const int cap = sm_count_cached() * 8;
    if (blocks > cap) blocks = cap;
    prep_q8k_i8_kernel<<<blocks, 256, 0, st>>>((const uint8_t *)xq, xq_stride, x8, xb2, xdT, (int)N, (int)npad, (int)K);
    g_launch_count++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("prep_q8k_i8 launch: %s", cudaGetErrorString(e)); return LAMMCU_ERR_CUDA; }
  }
  XParams p = {};
  p.dxT = xdT;  p.c = c;  p.ldc = ldc;  p.w = (const uint8_t *)w;  p.w_stride = w_stride;
  p.M = (int)M;  p.N = (int)N;  p.Npad = (int)npad;  p.K = (int)K;
  CUtensorMap tm_x, tm_b2;
  int rc;
  if ((rc = make_map_2d(&tm_x, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, x8, (uint64_t)K, (uint64_t)npad, (uint64_t)K, 128, GN,
                        CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  if ((rc = make_map_2d(&tm_b2, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, xb2, (uint64_t)(K / 8), (uint64_t)npad, (uint64_t)(K / 8), 128, GN,
                        CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  constexpr int smem = kXTotal + 1024;
  static_assert(smem <= 227 * 1024, "shared memory budget");
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_q2k_i8_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_q2k_i8_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) {
      set_error("gemm_q2k_i8: cudaFuncSetAttribute(%d): %s", smem, cudaGetErrorString(e));
      return LAMMCU_ERR_CUDA;
    }
    attr_set = true;
  }
  if (const GemmPeers *peers = gemm_peers_take()) {
    p.n_peer = peers->n;
    for (int g = 0; g < peers->n; g++) p.cp[g] = peers->c[g];
  }
  const int tiles = (int)(((M + GM - 1) / GM) * (npad / GN));
  if (p.n_peer) gemm_q2k_i8_kernel<true><<<tiles, kXThreads, smem, st>>>(tm_x, tm_b2, p);
  else gemm_q2k_i8_kernel<false><<<tiles, kXThreads, smem, st>>>(tm_x, tm_b2, p);
  g_launch_count++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("gemm_q2k_i8 launch: %s", cudaGetErrorString(e));
    return LAMMCU_ERR_CUDA;
  }
  return LAMMCU_OK;
}

}  // namespace lamm
