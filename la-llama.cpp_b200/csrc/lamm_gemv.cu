// lamm_gemv.cu -- decode path (N <= 8): C[j, i] = dot(W row i, Xq row j).
//
// Replaces LAMMImpl<T>::matmul + lamm_kernel_*.hpp (src/lamm_impl.hpp:20-147,
// src/lamm_kernel_{f32,q2_k,q4_0,q4_1,q5_0,q5_1,q8_0}.hpp) for small N.
// Memory-bound: every weight byte is read from HBM exactly once, in ggml's own
// packed block layout (18/20/22/24/34/84-byte blocks, no repacking).
//
// One persistent CTA = 1 producer warp + 8 consumer warps.
//   producer : claims row tiles from a global atomic counter and streams each
//              tile's rows HBM -> shared memory with cp.async.bulk (TMA 1-D,
//              SASS UBLKCP) into a ring of stages guarded by full/empty
//              mbarriers.  Row starts are only 2/4-byte aligned, so each copy
//              covers the enclosing 16-byte-aligned span and the consumer adds
//              the (< 16 B) offset back.
//   consumers: the quantized activations live in shared memory for the whole
//              kernel (re-laid-out so 128-bit reads are conflict free); each
//              warp owns one (row, K-segment) of the tile, lanes stride over
//              blocks, unpack nibbles / 5th bits / K-quant sub-scales in
//              registers, dp4a against the int8 activations, f32 FMA of the
//              block scales, then a warp-shuffle reduction per row.
#include "lamm_common.cuh"

namespace lamm {

constexpr int kConsumerWarps = 8;
constexpr int kThreads = (kConsumerWarps + 1) * 32;
constexpr int kMaxStages = 8;
constexpr int kCtrlBytes = 1024;          // barriers, tile ids, partial sums

struct GemvParams {
  const uint8_t *w;      int64_t w_stride;     // bytes
  const uint8_t *xq;     int64_t xq_stride;    // bytes (F32: the f32 X itself)
  float *c;              int64_t ldc;          // floats
  int M, n_valid;        // rows of W, live columns (<= NC)
  int units;             // work units per row (blocks; Q2_K: 8 per super-block; F32: float4s)
  int kb;                // blocks per row
  int row_bytes;         // payload bytes per row
  int R, KS;             // rows per tile, K-segments per row (R*KS == 8)
  int num_tiles, stages;
  int slot_pitch;        // bytes between row slots inside a stage
  int x_bytes;           // shared-memory bytes of the activation area
  unsigned int *sched;   // [0] next tile, [1] finished CTAs (self-resetting)
};

// ---- integer dot products ----------------------------------------------------
__device__ __forceinline__ int dp4a_us(uint32_t a_u8x4, uint32_t b_s8x4, int c) {
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a_u8x4), "r"(b_s8x4), "r"(c));
  return d;
}
__device__ __forceinline__ int dp4a_ss(uint32_t a_s8x4, uint32_t b_s8x4, int c) {
  int d;
  asm("dp4a.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a_s8x4), "r"(b_s8x4), "r"(c));
  return d;
}

// N consecutive 32-bit words starting at a 2-byte-aligned shared address
template <int N>
__device__ __forceinline__ void lds_words_a2(const uint8_t *p, uint32_t (&out)[N]) {
  const uint32_t addr = smem_u32(p);
  const uint32_t base = addr & ~3u;
  const uint32_t sh = (addr & 2u) * 8u;
  uint32_t w[N + 1];
#pragma unroll
  for (int i = 0; i <= N; i++)
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w[i]) : "r"(base + 4u * i));
#pragma unroll
  for (int i = 0; i < N; i++) out[i] = __funnelshift_r(w[i], w[i + 1], sh);
}
template <int N>
__device__ __forceinline__ void lds_words_a4(const uint8_t *p, uint32_t (&out)[N]) {
  const uint32_t addr = smem_u32(p);
#pragma unroll
  for (int i = 0; i < N; i++)
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(out[i]) : "r"(addr + 4u * i));
}
__device__ __forceinline__ uint32_t lds_u16(const uint8_t *p) {
  uint16_t v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(smem_u32(p)));
  return v;
}

// 4 bits -> bit 4 of 4 consecutive bytes (the Q5 "5th bit", element j <- bit j)
__device__ __forceinline__ uint32_t spread4_to_bit4(uint32_t nib) {
  return (nib * 0x02040810u) & 0x10101010u;
}

// ---- one decoded 32-element weight block -----------------------------------------
// lo[i] holds elements 4i..4i+3, hi[i] elements 16+4i..16+4i+3, as bytes.
struct WBlock32 {
  uint32_t lo[4], hi[4];
  float d, m;
};

template <int WT>
__device__ __forceinline__ void read_wblock(const uint8_t *p, WBlock32 &b) {
  if constexpr (WT == LAMMCU_TYPE_Q4_0) {                 // {f16 d; u8 qs[16]}
    uint32_t q[4];
    b.d = half_bits_to_float(lds_u16(p));
    b.m = 0.f;
    lds_words_a2<4>(p + 2, q);
#pragma unroll
    for (int i = 0; i < 4; i++) { b.lo[i] = q[i] & 0x0F0F0F0Fu; b.hi[i] = (q[i] >> 4) & 0x0F0F0F0Fu; }
  } else if constexpr (WT == LAMMCU_TYPE_Q4_1) {          // {f16 d, m; u8 qs[16]}
    uint32_t q[5];
    lds_words_a2<5>(p, q);
    b.d = half_bits_to_float(q[0] & 0xffffu);
    b.m = half_bits_to_float(q[0] >> 16);
#pragma unroll
    for (int i = 0; i < 4; i++) { b.lo[i] = q[i + 1] & 0x0F0F0F0Fu; b.hi[i] = (q[i + 1] >> 4) & 0x0F0F0F0Fu; }
  } else if constexpr (WT == LAMMCU_TYPE_Q5_0) {          // {f16 d; u8 qh[4]; u8 qs[16]}
    uint32_t q[5];
    b.d = half_bits_to_float(lds_u16(p));
    b.m = 0.f;
    lds_words_a2<5>(p + 2, q);
    const uint32_t qh = q[0];
#pragma unroll
    for (int i = 0; i < 4; i++) {
      b.lo[i] = (q[i + 1] & 0x0F0F0F0Fu) | spread4_to_bit4((qh >> (4 * i)) & 0xFu);
      b.hi[i] = ((q[i + 1] >> 4) & 0x0F0F0F0Fu) | spread4_to_bit4((qh >> (16 + 4 * i)) & 0xFu);
    }
  } else if constexpr (WT == LAMMCU_TYPE_Q5_1) {          // {f16 d, m; u8 qh[4]; u8 qs[16]}
    uint32_t q[6];
    lds_words_a2<6>(p, q);
    b.d = half_bits_to_float(q[0] & 0xffffu);
    b.m = half_bits_to_float(q[0] >> 16);
    const uint32_t qh = q[1];
#pragma unroll
    for (int i = 0; i < 4; i++) {
      b.lo[i] = (q[i + 2] & 0x0F0F0F0Fu) | spread4_to_bit4((qh >> (4 * i)) & 0xFu);
      b.hi[i] = ((q[i + 2] >> 4) & 0x0F0F0F0Fu) | spread4_to_bit4((qh >> (16 + 4 * i)) & 0xFu);
    }
  } else {                                                // Q8_0 {f16 d; i8 qs[32]}
    uint32_t q[8];
    b.d = half_bits_to_float(lds_u16(p));
    b.m = 0.f;
    lds_words_a2<8>(p + 2, q);
#pragma unroll
    for (int i = 0; i < 4; i++) { b.lo[i] = q[i]; b.hi[i] = q[i + 4]; }
  }
}

// ---- activations in shared memory -------------------------------------------------
// 32-element formats: per column n and block kb
//   xlo[n*kb_n + kb] : uint4, elements 0..15     (16-B stride: conflict-free LDS.128)
//   xhi[n*kb_n + kb] : uint4, elements 16..31
//   xds[n*kb_n + kb] : float2 (d, s) for Q8_1; (d, bit-cast int sum of qs) for Q8_0
// The sum of the activation quants turns sum((q-8)*x) into sum(q*x) - 8*sum(x).
struct XS32 {
  const uint4 *lo, *hi;
  const float2 *ds;
};

template <int VT, int NC>
__device__ __forceinline__ void stage_x32(const GemvParams &p, uint8_t *xs, int tid /*0..255*/) {
  constexpr int XB = (VT == LAMMCU_TYPE_Q8_0) ? 34 : 36;
  uint32_t *lo = reinterpret_cast<uint32_t *>(xs);
  uint32_t *hi = lo + (size_t)NC * p.kb * 4;
  float2 *ds = reinterpret_cast<float2 *>(hi + (size_t)NC * p.kb * 4);
  const int total = NC * p.kb;                 // blocks incl. dead columns
  const int live = p.n_valid * p.kb;
  const int i = tid & 7;
  for (int base = 0; base < total; base += 32) {
    const int g = base + (tid >> 3);
    uint32_t word = 0;
    float d = 0.f, s = 0.f;
    if (g < live) {
      const int n = g / p.kb, kb = g - n * p.kb;
      const uint8_t *src = p.xq + (int64_t)n * p.xq_stride + (int64_t)kb * XB;
      if (VT == LAMMCU_TYPE_Q8_1) {
        word = *reinterpret_cast<const uint32_t *>(src + 4 + 4 * i);
        if (i == 0) {
          const uint32_t dsb = *reinterpret_cast<const uint32_t *>(src);
          d = half_bits_to_float(dsb & 0xffffu);
          s = half_bits_to_float(dsb >> 16);
        }
      } else {
        const uint16_t *q = reinterpret_cast<const uint16_t *>(src + 2 + 4 * i);
        word = (uint32_t)q[0] | ((uint32_t)q[1] << 16);
        if (i == 0) d = half_bits_to_float(*reinterpret_cast<const uint16_t *>(src));
      }
    }
    int xsum = dp4a_ss(word, 0x01010101u, 0);
    xsum += __shfl_xor_sync(0xffffffffu, xsum, 4);
    xsum += __shfl_xor_sync(0xffffffffu, xsum, 2);
    xsum += __shfl_xor_sync(0xffffffffu, xsum, 1);
    if (g < total) {
      ((i < 4) ? lo : hi)[(size_t)g * 4 + (i & 3)] = word;
      if (i == 0) ds[g] = make_float2(d, (VT == LAMMCU_TYPE_Q8_1) ? s : __int_as_float(xsum));
    }
  }
}

template <int WT, int NC>
__device__ __forceinline__ void dot_unit32(const uint8_t *rowp, int kb, const GemvParams &p,
                                           const uint8_t *xs, float (&acc)[NC]) {
  constexpr int BS = type_info(WT).size;
  const uint4 *xlo = reinterpret_cast<const uint4 *>(xs);
  const uint4 *xhi = xlo + (size_t)NC * p.kb;
  const float2 *xds = reinterpret_cast<const float2 *>(xhi + (size_t)NC * p.kb);
  WBlock32 b;
  read_wblock<WT>(rowp + kb * BS, b);
#pragma unroll
  for (int n = 0; n < NC; n++) {
    const uint4 l = xlo[n * p.kb + kb], h = xhi[n * p.kb + kb];
    const float2 ds = xds[n * p.kb + kb];
    int s = 0;
    if constexpr (WT == LAMMCU_TYPE_Q8_0) {
      s = dp4a_ss(b.lo[0], l.x, s); s = dp4a_ss(b.lo[1], l.y, s);
      s = dp4a_ss(b.lo[2], l.z, s); s = dp4a_ss(b.lo[3], l.w, s);
      s = dp4a_ss(b.hi[0], h.x, s); s = dp4a_ss(b.hi[1], h.y, s);
      s = dp4a_ss(b.hi[2], h.z, s); s = dp4a_ss(b.hi[3], h.w, s);
    } else {
      s = dp4a_us(b.lo[0], l.x, s); s = dp4a_us(b.lo[1], l.y, s);
      s = dp4a_us(b.lo[2], l.z, s); s = dp4a_us(b.lo[3], l.w, s);
      s = dp4a_us(b.hi[0], h.x, s); s = dp4a_us(b.hi[1], h.y, s);
      s = dp4a_us(b.hi[2], h.z, s); s = dp4a_us(b.hi[3], h.w, s);
    }
    if constexpr (WT == LAMMCU_TYPE_Q4_0) s -= 8 * __float_as_int(ds.y);     // (q-8)
    if constexpr (WT == LAMMCU_TYPE_Q5_0) s -= 16 * __float_as_int(ds.y);    // (q-16)
    acc[n] = fmaf(b.d * ds.x, (float)s, acc[n]);
    if constexpr (WT == LAMMCU_TYPE_Q4_1 || WT == LAMMCU_TYPE_Q5_1) acc[n] = fmaf(b.m, ds.y, acc[n]);
  }
}

// ---- Q2_K x Q8_K --------------------------------------------------------------------
// Work unit u = sb*8 + half*4 + c: the 8 bytes qs[32*half + 8c .. +8) of super-block
// sb, i.e. 32 elements spread over the 4 bit-planes j (element 128*half + 32*j + 8c + t,
// sub-block g = 8*half + 2*j + (c>=2); src/lamm_kernel_q2_k.hpp:49-71).  The unit
// also owns the min-term of sub-blocks 2*(u&7) and 2*(u&7)+1.
// Shared layout per column n (U = 8 * super-blocks):
//   xq[n][j][U]  : uint2  = the 8 int8 activations matching plane j of unit u
//   xbs[n][U]    : uint32 = bsums[2*(u&7)], bsums[2*(u&7)+1] (2 x int16)
//   xd[n][U/8]   : float  = block_q8_K.d
template <int NC>
__device__ __forceinline__ void stage_xq2k(const GemvParams &p, uint8_t *xs, int tid) {
  const int U = p.units, SB = p.kb;
  uint2 *xq = reinterpret_cast<uint2 *>(xs);
  uint32_t *xbs = reinterpret_cast<uint32_t *>(xq + (size_t)NC * 4 * U);
  float *xd = reinterpret_cast<float *>(xbs + (size_t)NC * U);
  for (int idx = tid; idx < NC * 4 * U; idx += kConsumerWarps * 32) {
    const int n = idx / (4 * U), r = idx - n * 4 * U, j = r / U, u = r - j * U;
    uint2 v = make_uint2(0u, 0u);
    if (n < p.n_valid) {
      const int sb = u >> 3, half = (u >> 2) & 1, c = u & 3;
      const uint8_t *src = p.xq + (int64_t)n * p.xq_stride + (int64_t)sb * 292 + Q8_K_QS +
                           128 * half + 32 * j + 8 * c;
      const uint32_t *s32 = reinterpret_cast<const uint32_t *>(src);   // 4-B aligned only
      v = make_uint2(s32[0], s32[1]);
    }
    xq[idx] = v;
  }
  for (int idx = tid; idx < NC * U; idx += kConsumerWarps * 32) {
    const int n = idx / U, u = idx - n * U;
    uint32_t v = 0;
    if (n < p.n_valid) {
      const uint8_t *src = p.xq + (int64_t)n * p.xq_stride + (int64_t)(u >> 3) * 292 + Q8_K_BSUMS +
                           4 * (u & 7);
      v = *reinterpret_cast<const uint32_t *>(src);
    }
    xbs[idx] = v;
  }
  for (int idx = tid; idx < NC * SB; idx += kConsumerWarps * 32) {
    const int n = idx / SB, sb = idx - n * SB;
    xd[idx] = (n < p.n_valid)
                  ? *reinterpret_cast<const float *>(p.xq + (int64_t)n * p.xq_stride + (int64_t)sb * 292)
                  : 0.f;
  }
}

template <int NC>
__device__ __forceinline__ void dot_unit_q2k(const uint8_t *rowp, int u, const GemvParams &p,
                                             const uint8_t *xs, float (&acc)[NC]) {
  const int U = p.units;
  const uint2 *xq = reinterpret_cast<const uint2 *>(xs);
  const uint32_t *xbs = reinterpret_cast<const uint32_t *>(xq + (size_t)NC * 4 * U);
  const float *xd = reinterpret_cast<const float *>(xbs + (size_t)NC * U);
  const int sb = u >> 3, half = (u >> 2) & 1, c = u & 3, u8 = u & 7;
  const uint8_t *blk = rowp + sb * 84;                    // {u8 scales[16]; u8 qs[64]; f16 d, dmin}
  uint32_t q[2], scw[2], dm[1], sc2;
  lds_words_a4<2>(blk + 16 + 32 * half + 8 * c, q);
  lds_words_a4<2>(blk + 8 * half, scw);                   // scales[8*half .. +8)
  lds_words_a4<1>(blk + 80, dm);
  sc2 = lds_u16(blk + 2 * u8);                            // scales[2*u8], scales[2*u8+1]
  const float wd = half_bits_to_float(dm[0] & 0xffffu), wdmin = half_bits_to_float(dm[0] >> 16);
  // scale byte of plane j: scales[8*half + 2*j + (c>=2)]
  const uint32_t sel = (c >> 1);
  int sc[4];
  sc[0] = (scw[0] >> (8 * sel)) & 0xF;
  sc[1] = (scw[0] >> (16 + 8 * sel)) & 0xF;
  sc[2] = (scw[1] >> (8 * sel)) & 0xF;
  sc[3] = (scw[1] >> (16 + 8 * sel)) & 0xF;
  const int m0 = (sc2 >> 4) & 0xF, m1 = (sc2 >> 12) & 0xF;
#pragma unroll
  for (int n = 0; n < NC; n++) {
    int isum = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const uint2 x = xq[(size_t)(n * 4 + j) * U + u];
      const uint32_t mask = 0x03030303u << (2 * j);
      // (q & mask) = value << 2j as u8; the dot is exactly 4^j times too large
      int t = dp4a_us(q[0] & mask, x.x, 0);
      t = dp4a_us(q[1] & mask, x.y, t);
      isum += sc[j] * (t >> (2 * j));
    }
    const uint32_t bs = xbs[n * U + u];
    const int summs = (int)(short)(bs & 0xffffu) * m0 + (int)(short)(bs >> 16) * m1;
    const float d8 = xd[n * p.kb + sb];
    acc[n] = fmaf(d8 * wd, (float)isum, acc[n]);
    acc[n] = fmaf(-(d8 * wdmin), (float)summs, acc[n]);
  }
}

// ---- Q3_K / Q4_K / Q5_K / Q6_K x Q8_K ---------------------------------------
// Not in the reference's table: the behaviour matched is ggml's own ggml_vec_dot_q{3,4,5,6}_K_q8_K
// (ggml-quants.c:6127, 7000, 7609, 8305; scalar forms :6583-6646, 7301-7358, 7968-8029, 8695-8737).
// Same work unit and the same shared activation layout as Q2_K above: unit u = sb*8 + half*4 + c
// covers elements 128*half + 32*j + 8c + t (plane j < 4, t < 8) of super-block sb, and owns the
// bsums terms of the 16-element sub-blocks 2*(u&7), 2*(u&7)+1 (= 32-element group u&7).
__device__ __forceinline__ int lds_s8(const uint8_t *p) {
  int v;
  asm volatile("ld.shared.s8 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)));
  return v;
}
__device__ __forceinline__ int byte_of(uint32_t w, int i) { return (int)((w >> (8 * i)) & 0xFFu); }

template <int WT, int NC>
__device__ __forceinline__ void dot_unit_kq(const uint8_t *rowp, int u, const GemvParams &p,
                                            const uint8_t *xs, float (&acc)[NC]) {
  const int U = p.units;
  const uint2 *xq = reinterpret_cast<const uint2 *>(xs);
  const uint32_t *xbs = reinterpret_cast<const uint32_t *>(xq + (size_t)NC * 4 * U);
  const float *xd = reinterpret_cast<const float *>(xbs + (size_t)NC * U);
  const int sb = u >> 3, half = (u >> 2) & 1, c = u & 3;
  const uint8_t *blk = rowp + sb * type_info(WT).size;
  uint32_t pl[4][2];          // the 4 planes of the unit as unsigned bytes
  int sc[4];                  // integer scale of each plane's sub-block
  int ms0, ms1;               // multipliers of bsums[2*(u&7)], bsums[2*(u&7)+1]
  float wd, wdmin = 0.f;
  if constexpr (WT == LAMMCU_TYPE_Q4_K || WT == LAMMCU_TYPE_Q5_K) {
    constexpr int QS = (WT == LAMMCU_TYPE_Q4_K) ? 16 : 48;
    uint32_t hdr[4], a[2], b[2];
    lds_words_a4<4>(blk, hdr);                                  // d | dmin, scales[0..11]
    lds_words_a4<2>(blk + QS + 64 * half + 8 * c, a);           // pair 2*half:   planes 0 (lo), 1 (hi)
    lds_words_a4<2>(blk + QS + 64 * half + 32 + 8 * c, b);      // pair 2*half+1: planes 2 (lo), 3 (hi)
    wd = half_bits_to_float(hdr[0] & 0xffffu);
    wdmin = half_bits_to_float(hdr[0] >> 16);
    // get_scale_min_k4 for groups 4*half .. 4*half+3, four at a time
    const uint32_t scw = half ? ((hdr[3] & 0x0F0F0F0Fu) | (((hdr[1] >> 6) & 0x03030303u) << 4)) : (hdr[1] & 0x3F3F3F3Fu);
    const uint32_t mnw = half ? (((hdr[3] >> 4) & 0x0F0F0F0Fu) | (((hdr[2] >> 6) & 0x03030303u) << 4))
                              : (hdr[2] & 0x3F3F3F3Fu);
#pragma unroll
    for (int j = 0; j < 4; j++) sc[j] = byte_of(scw, j);
    ms0 = ms1 = byte_of(mnw, c);
#pragma unroll
    for (int k = 0; k < 2; k++) {
      pl[0][k] = a[k] & 0x0F0F0F0Fu;  pl[1][k] = (a[k] >> 4) & 0x0F0F0F0Fu;
      pl[2][k] = b[k] & 0x0F0F0F0Fu;  pl[3][k] = (b[k] >> 4) & 0x0F0F0F0Fu;
    }
    if constexpr (WT == LAMMCU_TYPE_Q5_K) {
      uint32_t h[2];
      lds_words_a4<2>(blk + 16 + 8 * c, h);                     // bit g of qh[l]: 5th bit of element 32g + l
#pragma unroll
      for (int j = 0; j < 4; j++)
#pragma unroll
        for (int k = 0; k < 2; k++) pl[j][k] |= ((h[k] >> (4 * half + j)) & 0x01010101u) << 4;
    }
  } else if constexpr (WT == LAMMCU_TYPE_Q6_K) {
    uint32_t a[2], b[2], h[2];
    lds_words_a2<2>(blk + 64 * half + 8 * c, a);                // ql: planes 0 (lo), 2 (hi)
    lds_words_a2<2>(blk + 64 * half + 32 + 8 * c, b);           // ql: planes 1 (lo), 3 (hi)
    lds_words_a2<2>(blk + 128 + 32 * half + 8 * c, h);          // qh: 2 bits per plane
    wd = half_bits_to_float(lds_u16(blk + 208));
#pragma unroll
    for (int k = 0; k < 2; k++) {
      pl[0][k] = (a[k] & 0x0F0F0F0Fu) | ((h[k] & 0x03030303u) << 4);
      pl[1][k] = (b[k] & 0x0F0F0F0Fu) | (((h[k] >> 2) & 0x03030303u) << 4);
      pl[2][k] = ((a[k] >> 4) & 0x0F0F0F0Fu) | (((h[k] >> 4) & 0x03030303u) << 4);
      pl[3][k] = ((b[k] >> 4) & 0x0F0F0F0Fu) | (((h[k] >> 6) & 0x03030303u) << 4);
    }
    const uint8_t *scp = blk + 192 + 8 * half;                  // int8 scales of this half's 8 sub-blocks
#pragma unroll
    for (int j = 0; j < 4; j++) sc[j] = lds_s8(scp + 2 * j + (c >> 1));
    ms0 = 32 * lds_s8(scp + 2 * c);                             // sum (q - 32) x = sum q x - 32 bsum
    ms1 = 32 * lds_s8(scp + 2 * c + 1);
  } else {   // Q3_K
    uint32_t q[2], h[2], s3[3];
    lds_words_a2<2>(blk + 32 + 32 * half + 8 * c, q);           // low 2 bits, 4 planes per byte
    lds_words_a2<2>(blk + 8 * c, h);                            // hmask: bit 4*half + j belongs to plane j
    lds_words_a2<3>(blk + 96, s3);
    wd = half_bits_to_float(lds_u16(blk + 108));
    // the 16 6-bit scales as bytes of aux[0..3] (ggml-quants.c:2570-2575)
    uint32_t aux[4];
    aux[0] = (s3[0] & 0x0F0F0F0Fu) | ((s3[2] & 0x03030303u) << 4);
    aux[1] = (s3[1] & 0x0F0F0F0Fu) | (((s3[2] >> 2) & 0x03030303u) << 4);
    aux[2] = ((s3[0] >> 4) & 0x0F0F0F0Fu) | (((s3[2] >> 4) & 0x03030303u) << 4);
    aux[3] = ((s3[1] >> 4) & 0x0F0F0F0Fu) | (((s3[2] >> 6) & 0x03030303u) << 4);
    const uint32_t lo = half ? aux[2] : aux[0], hi = half ? aux[3] : aux[1];     // sub-blocks 8*half .. +7
#pragma unroll
    for (int j = 0; j < 4; j++) {
      sc[j] = byte_of(j < 2 ? lo : hi, 2 * (j & 1) + (c >> 1)) - 32;
#pragma unroll
      for (int k = 0; k < 2; k++)
        pl[j][k] = ((q[k] >> (2 * j)) & 0x03030303u) | (((h[k] >> (4 * half + j)) & 0x01010101u) << 2);
    }
    const uint32_t own = (c >> 1) ? hi : lo;                    // sub-blocks 8*half + 2c, + 2c + 1
    ms0 = 4 * (byte_of(own, 2 * (c & 1)) - 32);                 // value = (q | hbit << 2) - 4
    ms1 = 4 * (byte_of(own, 2 * (c & 1) + 1) - 32);
  }
#pragma unroll
  for (int n = 0; n < NC; n++) {
    int isum = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const uint2 x = xq[(size_t)(n * 4 + j) * U + u];
      int t = dp4a_us(pl[j][0], x.x, 0);
      t = dp4a_us(pl[j][1], x.y, t);
      isum += sc[j] * t;
    }
    const uint32_t bs = xbs[n * U + u];
    const int summs = (int)(short)(bs & 0xffffu) * ms0 + (int)(short)(bs >> 16) * ms1;
    const float d8 = xd[n * p.kb + sb];
    if constexpr (WT == LAMMCU_TYPE_Q4_K || WT == LAMMCU_TYPE_Q5_K) {
      acc[n] = fmaf(d8 * wd, (float)isum, acc[n]);
      acc[n] = fmaf(-(d8 * wdmin), (float)summs, acc[n]);
    } else {
      acc[n] = fmaf(d8 * wd, (float)(isum - summs), acc[n]);
    }
  }
}

// ---- F32 x F32 ------------------------------------------------------------------------
// unit = one float4 of the row; X rows are copied to shared memory as they are.
template <int NC>
__device__ __forceinline__ void stage_xf32(const GemvParams &p, uint8_t *xs, int tid) {
  float *x = reinterpret_cast<float *>(xs);
  const int K = p.units * 4;
  for (int idx = tid; idx < NC * K; idx += kConsumerWarps * 32) {
    const int n = idx / K, k = idx - n * K;
    x[idx] = (n < p.n_valid)
                 ? *reinterpret_cast<const float *>(p.xq + (int64_t)n * p.xq_stride + 4 * (int64_t)k)
                 : 0.f;
  }
}
template <int NC>
__device__ __forceinline__ void dot_unit_f32(const uint8_t *rowp, int u, const GemvParams &p,
                                             const uint8_t *xs, float (&acc)[NC]) {
  uint32_t w[4];
  lds_words_a4<4>(rowp + 16 * u, w);
  const float4 *x = reinterpret_cast<const float4 *>(xs);
#pragma unroll
  for (int n = 0; n < NC; n++) {
    const float4 v = x[n * p.units + u];
    acc[n] = fmaf(__uint_as_float(w[0]), v.x, acc[n]);
    acc[n] = fmaf(__uint_as_float(w[1]), v.y, acc[n]);
    acc[n] = fmaf(__uint_as_float(w[2]), v.z, acc[n]);
    acc[n] = fmaf(__uint_as_float(w[3]), v.w, acc[n]);
  }
}

// ---- F16 weights: ggml_vec_dot_f16 (ggml.c:1414-1457) on src1 rounded to fp16 -------------
// unit = 16 bytes = 8 fp16 weights.  Xq holds the fp16 activations (ggml_fp32_to_fp16_row); they stay fp16 in shared
// memory (one byte of activation read per byte of weight: as f32 the dot phase read 4 B of x per 2 B of w and the
// kernel was shared-memory-bound at 30-48 % of the HBM peak) and both operands are widened in registers, so every
// product w * x is exact in f32 (11 + 11 significant bits) and only the f32 accumulation order differs from ggml's
// SIMD lanes.
template <int NC>
__device__ __forceinline__ void stage_xf16(const GemvParams &p, uint8_t *xs, int tid) {
  uint4 *x = reinterpret_cast<uint4 *>(xs);
  const int U = p.units;                                         // 16-byte chunks per row
  for (int idx = tid; idx < NC * U; idx += kConsumerWarps * 32) {
    const int n = idx / U, u = idx - n * U;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (n < p.n_valid) {
      const uint8_t *src = p.xq + (int64_t)n * p.xq_stride + 16 * (int64_t)u;
      if ((((uintptr_t)p.xq | (uintptr_t)p.xq_stride) & 15) == 0) {
        v = *reinterpret_cast<const uint4 *>(src);
      } else {                                                   // 2-byte aligned rows
        const uint16_t *h = reinterpret_cast<const uint16_t *>(src);
        v = make_uint4(h[0] | ((uint32_t)h[1] << 16), h[2] | ((uint32_t)h[3] << 16), h[4] | ((uint32_t)h[5] << 16),
                       h[6] | ((uint32_t)h[7] << 16));
      }
    }
    x[idx] = v;
  }
}
template <int NC>
__device__ __forceinline__ void dot_unit_f16(const uint8_t *rowp, int u, const GemvParams &p,
                                             const uint8_t *xs, float (&acc)[NC]) {
  uint32_t w[4];
  lds_words_a4<4>(rowp + 16 * u, w);
  float2 wf[4];
#pragma unroll
  for (int i = 0; i < 4; i++) wf[i] = __half22float2(*reinterpret_cast<const __half2 *>(&w[i]));
  const uint4 *x = reinterpret_cast<const uint4 *>(xs);
#pragma unroll
  for (int n = 0; n < NC; n++) {
    const uint4 xv = x[n * p.units + u];
    const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const float2 xf = __half22float2(*reinterpret_cast<const __half2 *>(&xw[i]));
      acc[n] = fmaf(wf[i].x, xf.x, acc[n]);
      acc[n] = fmaf(wf[i].y, xf.y, acc[n]);
    }
  }
}

// ---- the kernel ---------------------------------------------------------------------
template <int WT, int NC>
__device__ __forceinline__ void stage_x(const GemvParams &p, uint8_t *xs, int tid) {
  if constexpr (WT == LAMMCU_TYPE_F32) stage_xf32<NC>(p, xs, tid);
  else if constexpr (WT == LAMMCU_TYPE_F16) stage_xf16<NC>(p, xs, tid);
  else if constexpr (WT == LAMMCU_TYPE_Q2_K || is_kquant_ext(WT)) stage_xq2k<NC>(p, xs, tid);
  else stage_x32<type_info(WT).vec_dot, NC>(p, xs, tid);
}
template <int WT, int NC>
__device__ __forceinline__ void dot_unit(const uint8_t *rowp, int u, const GemvParams &p,
                                         const uint8_t *xs, float (&acc)[NC]) {
  if constexpr (WT == LAMMCU_TYPE_F32) dot_unit_f32<NC>(rowp, u, p, xs, acc);
  else if constexpr (WT == LAMMCU_TYPE_F16) dot_unit_f16<NC>(rowp, u, p, xs, acc);
  else if constexpr (WT == LAMMCU_TYPE_Q2_K) dot_unit_q2k<NC>(rowp, u, p, xs, acc);
  else if constexpr (is_kquant_ext(WT)) dot_unit_kq<WT, NC>(rowp, u, p, xs, acc);
  else dot_unit32<WT, NC>(rowp, u, p, xs, acc);
}

__device__ __forceinline__ void consumer_bar() {   // the 8 consumer warps only
  asm volatile("bar.sync 1, %0;" ::"n"(kConsumerWarps * 32) : "memory");
}

template <int WT, int NC>
__global__ void __launch_bounds__(kThreads) gemv_stream_kernel(const GemvParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t *full = reinterpret_cast<uint64_t *>(smem);
  uint64_t *empty = full + kMaxStages;
  int *tile_of = reinterpret_cast<int *>(smem + 128);
  float *part = reinterpret_cast<float *>(smem + 256);     // [2][8 warps][8 cols]
  uint8_t *xs = smem + kCtrlBytes;
  uint8_t *stage0 = xs + p.x_bytes;
  const int stage_bytes = p.R * p.slot_pitch;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; s++) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], kConsumerWarps);
    }
    fence_barrier_init();
  }
  __syncthreads();

  if (warp == kConsumerWarps) {
    // ===================== producer =====================
    if (lane == 0) {
      const uint64_t pol = policy_evict_first();
      int s = 0;
      uint32_t ph = 0;
      for (;;) {
        const int tile = (int)atomicAdd(&p.sched[0], 1u);
        mbar_wait(&empty[s], ph ^ 1u);
        if (tile >= p.num_tiles) {
          tile_of[s] = -1;
          mbar_arrive(&full[s]);
          break;
        }
        tile_of[s] = tile;
        const int r0 = tile * p.R;
        const int rows = min(p.R, p.M - r0);
        uint32_t total = 0;
        for (int r = 0; r < rows; r++) {
          const uintptr_t a = (uintptr_t)(p.w + (int64_t)(r0 + r) * p.w_stride);
          total += (uint32_t)(((a + p.row_bytes + 15) & ~(uintptr_t)15) - (a & ~(uintptr_t)15));
        }
        mbar_arrive_expect_tx(&full[s], total);
        for (int r = 0; r < rows; r++) {
          const uintptr_t a = (uintptr_t)(p.w + (int64_t)(r0 + r) * p.w_stride);
          const uintptr_t a0 = a & ~(uintptr_t)15;
          const uint32_t bytes = (uint32_t)(((a + p.row_bytes + 15) & ~(uintptr_t)15) - a0);
          bulk_g2s_stream(stage0 + (size_t)s * stage_bytes + (size_t)r * p.slot_pitch,
                          reinterpret_cast<const void *>(a0), bytes, &full[s], pol);
        }
        if (++s == p.stages) { s = 0; ph ^= 1u; }
      }
      // self-resetting scheduler: the last CTA to finish zeroes both counters
      __threadfence();
      const unsigned int done = atomicAdd(&p.sched[1], 1u);
      if (done == gridDim.x - 1) {
        atomicExch(&p.sched[0], 0u);
        atomicExch(&p.sched[1], 0u);
      }
    }
    return;
  }

  // ===================== consumers =====================
  const int tid = threadIdx.x;                  // 0..255
  stage_x<WT, NC>(p, xs, tid);
  consumer_bar();

  const int rslot = warp % p.R, seg = warp / p.R;
  const int u_per_seg = (p.units + p.KS - 1) / p.KS;
  const int u0 = seg * u_per_seg, u1 = min(p.units, u0 + u_per_seg);
  int s = 0;
  uint32_t ph = 0, it = 0;
  for (;; it++) {
    mbar_wait(&full[s], ph);
    const int tile = tile_of[s];
    if (tile < 0) break;
    const int row = tile * p.R + rslot;
    float acc[NC];
#pragma unroll
    for (int n = 0; n < NC; n++) acc[n] = 0.f;
    if (row < p.M) {
      const uintptr_t a = (uintptr_t)(p.w + (int64_t)row * p.w_stride);
      const uint8_t *rowp = stage0 + (size_t)s * stage_bytes + (size_t)rslot * p.slot_pitch + (a & 15);
#pragma unroll 2
      for (int u = u0 + lane; u < u1; u += 32) dot_unit<WT, NC>(rowp, u, p, xs, acc);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[s]);      // weights of this stage are consumed
#pragma unroll
    for (int n = 0; n < NC; n++) acc[n] = warp_sum(acc[n]);
    if (p.KS == 1) {
      if (row < p.M) {
#pragma unroll
        for (int n = 0; n < NC; n++)
          if (lane == n && n < p.n_valid) p.c[(int64_t)n * p.ldc + row] = acc[n];
      }
    } else {
      float *pb = part + (it & 1u) * 64;
#pragma unroll
      for (int n = 0; n < NC; n++)
        if (lane == n) pb[warp * 8 + n] = acc[n];
      consumer_bar();
      if (seg == 0 && row < p.M && lane < p.n_valid) {
        float t = 0.f;
        for (int k = 0; k < p.KS; k++) t += pb[(rslot + k * p.R) * 8 + lane];
        p.c[(int64_t)lane * p.ldc + row] = t;
      }
    }
    if (++s == p.stages) { s = 0; ph ^= 1u; }
  }
}

// ---- host side ------------------------------------------------------------------------
constexpr int kSmemMax = 227 * 1024;

template <int WT, int NC>
static int launch_one(const GemvParams &p, int grid, int smem, cudaStream_t st) {
  static bool attr_set = false;    // per instantiation
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemv_stream_kernel<WT, NC>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemMax);
    if (e != cudaSuccess) {
      set_error("gemv: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return LAMMCU_ERR_CUDA;
    }
    attr_set = true;
  }
  gemv_stream_kernel<WT, NC><<<grid, kThreads, smem, st>>>(p);
  g_launch_count++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("gemv launch (type %d, nc %d, smem %d): %s", WT, NC, smem, cudaGetErrorString(e));
    return LAMMCU_ERR_CUDA;
  }
  return LAMMCU_OK;
}

template <int WT>
static int launch_nc(int nc, const GemvParams &p, int grid, int smem, cudaStream_t st) {
  switch (nc) {
  case 1: return launch_one<WT, 1>(p, grid, smem, st);
  case 2: return launch_one<WT, 2>(p, grid, smem, st);
  case 4: return launch_one<WT, 4>(p, grid, smem, st);
  default: return launch_one<WT, 8>(p, grid, smem, st);
  }
}

static int x_bytes_for(int w_type, int nc, int kb, int64_t K) {
  int b;
  if (w_type == LAMMCU_TYPE_F32) b = nc * (int)K * 4;
  else if (w_type == LAMMCU_TYPE_F16) b = nc * (int)K * 2;
  else if (type_info(w_type).blck == 256) b = nc * kb * 292;
  else b = nc * kb * 40;
  return (b + 127) & ~127;
}

int launch_gemv(int w_type, const void *w, int64_t w_stride, const void *xq, int64_t xq_stride,
                float *c, int64_t ldc, int64_t M, int64_t N, int64_t K, cudaStream_t st) {
  const TypeInfo ti = type_info(w_type);
  if (!lammcu_type_supported(w_type)) {
    set_error("gemv: weight type %d not supported", w_type);
    return LAMMCU_ERR_TYPE;
  }
  const int kmul = (w_type == LAMMCU_TYPE_F32 || w_type == LAMMCU_TYPE_F16) ? 8 : ti.blck;    // src/lamm_impl.hpp:123-126
  if (M <= 0 || N <= 0 || K <= 0 || K % kmul || M > 0x3fffffff || K > 0x3fffffff) {
    set_error("gemv: bad shape M=%lld N=%lld K=%lld", (long long)M, (long long)N, (long long)K);
    return LAMMCU_ERR_SHAPE;
  }
  const int walign = (w_type == LAMMCU_TYPE_F32 || w_type == LAMMCU_TYPE_F16 || w_type == LAMMCU_TYPE_Q2_K || w_type == LAMMCU_TYPE_Q4_K ||
                      w_type == LAMMCU_TYPE_Q5_K) ? 4 : 2;
  const int xalign = (ti.vec_dot == LAMMCU_TYPE_Q8_0 || ti.vec_dot == LAMMCU_TYPE_F16) ? 2 : 4;
  if (((uintptr_t)w | (uintptr_t)w_stride) % walign || ((uintptr_t)xq | (uintptr_t)xq_stride) % xalign ||
      ((uintptr_t)c & 3)) {
    set_error("gemv: operand alignment (W %d-byte, Xq %d-byte, C 4-byte) not met", walign, xalign);
    return LAMMCU_ERR_ALIGN;
  }
  const int kb = (int)(K / ti.blck);
  const int64_t row_bytes64 = (w_type == LAMMCU_TYPE_F32) ? K * 4 : (w_type == LAMMCU_TYPE_F16) ? K * 2 : (int64_t)kb * ti.size;
  const int slot_pitch = (int)(((row_bytes64 + 15 + 15) & ~(int64_t)15) + 16);
  if (kCtrlBytes + x_bytes_for(w_type, 1, kb, K) + 2 * slot_pitch > kSmemMax) {
    set_error("gemv: K=%lld too large for the shared-memory pipeline", (long long)K);
    return LAMMCU_ERR_SHAPE;
  }

  GemvParams p;
  p.w = (const uint8_t *)w;  p.w_stride = w_stride;
  p.xq_stride = xq_stride;   p.ldc = ldc;
  p.M = (int)M;              p.kb = kb;
  p.row_bytes = (int)row_bytes64;
  p.slot_pitch = slot_pitch;
  p.units = (w_type == LAMMCU_TYPE_F32) ? (int)(K / 4) : (w_type == LAMMCU_TYPE_F16) ? (int)(K / 8) : (ti.blck == 256 ? kb * 8 : kb);

  // rows per tile: the largest of 8/4/2/1 whose stage stays <= 24 KB
  int R = 8;
  while (R > 1 && R * slot_pitch > 24 * 1024) R >>= 1;
  // keep every lane busy: a K-segment should hold >= 32 units
  while (R < 8 && p.units / (kConsumerWarps / R) < 32) R <<= 1;
  p.R = R;
  p.KS = kConsumerWarps / R;
  p.num_tiles = (int)((M + R - 1) / R);
  const int stage_bytes = R * slot_pitch;
  const int sms = sm_count_cached();

  for (int64_t n0 = 0; n0 < N;) {
    // widest column group (1/2/4/8) that covers the remaining columns and fits
    int nc = 1;
    while (nc < 8 && nc < N - n0) nc <<= 1;
    while (nc > 1 && kCtrlBytes + x_bytes_for(w_type, nc, kb, K) + 2 * stage_bytes > kSmemMax) nc >>= 1;
    const int n_valid = (int)((N - n0 < nc) ? (N - n0) : nc);
    p.x_bytes = x_bytes_for(w_type, nc, kb, K);
    const int fixed = kCtrlBytes + p.x_bytes;
    // aim for 2 CTAs per SM when the activations leave room, else 1
    int budget = (fixed + 3 * stage_bytes <= kSmemMax / 2 - 1024) ? (kSmemMax / 2 - 1024) : kSmemMax;
    int stages = (budget - fixed) / stage_bytes;
    if (stages > kMaxStages) stages = kMaxStages;
    if (stages < 2) stages = 2;
    p.stages = stages;
    const int smem = fixed + stages * stage_bytes;
    int per_sm = kSmemMax / (smem + 1024);
    if (per_sm < 1) per_sm = 1;
    if (per_sm > 4) per_sm = 4;
    int grid = sms * per_sm;
    if (grid > p.num_tiles) grid = p.num_tiles;

    p.xq = (const uint8_t *)xq + n0 * xq_stride;
    p.c = c + n0 * ldc;
    p.n_valid = n_valid;
    p.sched = next_sched_slot();
    if (!p.sched) return LAMMCU_ERR_CUDA;
    int rc;
    switch (w_type) {
    case LAMMCU_TYPE_F32:  rc = launch_nc<LAMMCU_TYPE_F32>(nc, p, grid, smem, st); break;
    case LAMMCU_TYPE_F16:  rc = launch_nc<LAMMCU_TYPE_F16>(nc, p, grid, smem, st); break;
    case LAMMCU_TYPE_Q4_0: rc = launch_nc<LAMMCU_TYPE_Q4_0>(nc, p, grid, smem, st); break;
    case LAMMCU_TYPE_Q4_1: rc = launch_nc<LAMMCU_TYPE_Q4_1>(nc, p, grid, smem, st); break;
    case LAMMCU_TYPE_Q5_0: rc = launch_nc<LAMMCU_TYPE_Q5_0>(nc, p, grid, smem, st); break;
    case LAMMCU_TYPE_Q5_1: rc = launch_nc<LAMMCU_TYPE_Q5_1>(nc, p, grid, smem, st); break;
    case LAMMCU_TYPE_Q8_0: rc = launch_nc<LAMMCU_TYPE_Q8_0>(nc, p, grid, smem, st); break;
    case LAMMCU_TYPE_Q3_K: rc = launch_nc<LAMMCU_TYPE_Q3_K>(nc, p, grid, smem, st); break;
    case LAMMCU_TYPE_Q4_K: rc = launch_nc<LAMMCU_TYPE_Q4_K>(nc, p, grid, smem, st); break;
    case LAMMCU_TYPE_Q5_K: rc = launch_nc<LAMMCU_TYPE_Q5_K>(nc, p, grid, smem, st); break;
    case LAMMCU_TYPE_Q6_K: rc = launch_nc<LAMMCU_TYPE_Q6_K>(nc, p, grid, smem, st); break;
    default:               rc = launch_nc<LAMMCU_TYPE_Q2_K>(nc, p, grid, smem, st); break;
    }
    if (rc != LAMMCU_OK) return rc;
    n0 += n_valid;
  }
  return LAMMCU_OK;
}

}  // namespace lamm
