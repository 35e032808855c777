// lamm_decode.cu -- the decode path proper (N == 1): one launch per mul_mat.
//
// Replaces, for one activation row, ggml's INIT-phase quantize (llama.cpp-b2430/ggml.c:
// 10865-10887 -> quantize_row_q8_{0,1,K}) AND LAMMImpl<T>::matmul + lamm_kernel_*.hpp
// (src/lamm_impl.hpp:20-147) with a single HBM-bound kernel:
//
//   * every warp streams its own rows: lane 0 issues cp.async.bulk (TMA 1-D, SASS UBLKCP)
//     copies of (row, K-segment) spans HBM -> a private ring of shared-memory slots, each
//     guarded by an mbarrier; no producer warp, no cross-warp hand-off;
//   * the activation row is quantized once per CTA (bit-exact q8_0 / q8_1 / q8_K, shared
//     arithmetic with lamm_quantize.cu), laid out conflict-free in shared memory, and each
//     lane then keeps the quants of ITS blocks in registers for the whole kernel, so the
//     inner loop reads only weights from shared memory;
//   * a lane owns 4 consecutive 32-blocks (72/80/88/96/136 contiguous bytes, read with
//     LDS.64/LDS.128, fields picked with compile-time offsets) or half a Q2_K super-block;
//     nibbles / 5th bits / K-quant sub-scales are unpacked in registers, the block dot is
//     dp4a (dp2a for the Q2_K min term), the row finishes with a warp-shuffle reduction;
//   * rows are dealt round-robin to SMs so every SM streams the same number of bytes (+-1
//     row), and programmatic dependent launch lets the NEXT mul_mat of a chain start
//     streaming its (immutable) weights while this one drains -- activations are only read
//     after griddepcontrol.wait.
#include "lamm_quant_dev.cuh"
#include "lamm_xop.cuh"

#include <algorithm>
#include <vector>
#include <cstdlib>

namespace lamm {

constexpr int kDecWarps = 8;            // default consumer warps per CTA
constexpr int kDecMaxWarps = 10;        // K-segment counts that do not divide 8 use 9 or 10
constexpr int kDecMaxThreads = kDecMaxWarps * 32;
constexpr int kDecMaxSlots = 4;
constexpr int kDecCtrl = 1024;       // mbarriers [10][4] + partial sums

struct DecodeParams {
  const uint8_t *w;   int64_t w_stride;   // bytes
  const float *x;                         // f32 activation row      (XF32)
  const uint8_t *xq;                      // quantized activation row (!XF32), ggml layout
  uint8_t *xq_out;                        // optional: CTA 0 stores the quantized row here
  float *c;
  int M, kb;            // rows; blocks per row (Q2_K: super-blocks)
  int units;            // lane units per row (4 blocks, or half a super-block)
  int KS, seg_units;    // K segments per row, units per segment (<= 32)
  int RG;               // row groups per CTA (= active warps / KS)
  int slots, slot_bytes;
  int x_bytes;          // shared-memory bytes of the staged activations
  unsigned long long *trace;   // optional: 8 globaltimer stamps per CTA (lammcu_set_trace)
  // 16-warp kernel: rows per warp prefetched into L2 beyond the shared-memory ring.  0 (off) everywhere: 1-4 rows ahead cost
  // 4-10 % (round 1).  The never-taken branch stays in the refill path on purpose: without it ptxas lays the same kernel
  // out 3.4 % slower (1057 vs 1092 tok/s on the 7B chain, measured three times on one box: profiles/r2ac, r2aq, r2ar).
  int l2_ahead;
  // sibling batch (16-warp kernel): up to 4 weight matrices of one type, K and row stride that share
  // the activation row (q/k/v, gate/up) form one virtual matrix of M total rows.  Global rows
  // [mc[j-1], mc[j]) belong to matrix j; wb[j] / cb[j] are VIRTUAL bases (pointer minus the rows
  // before it), so a row's address is wb[j] + row * w_stride.  nbatch <= 1: w and c are used as is.
  const uint8_t *wb[4];   float *cb[4];   int mc[3];   int nbatch;
  // row-sharded multi-GPU (n_peer > 0): the epilogue stores every result as a {value, tag} pair
  // (one 64-bit store, "LL" format) into the full C vector of EVERY rank over NVLink.  The tag is
  // this launch's sequence number; a consumer polls the pairs of its activation row until every
  // tag matches, so data and "ready" arrive in one NVLink transaction: no fence, no flag, no
  // collective.  (See lammcu.h "row-sharded multi-GPU decode" for the buffer layout.)
  unsigned long long *ll_peer[LAMMCU_MAX_RANKS];   // LL pairs of the full C on each rank (peer-mapped)
  const unsigned long long *x_ll;          // XM == 2: activation row as LL pairs (a previous output)
  unsigned int *arrive_ctr;                // local: CTA arrival tickets of all sharded launches
  unsigned int *status;                    // local: [0] set to 1 if a poll timed out
  int x_back;                              // XM == 2: x was produced x_back sharded launches ago (>= 1)
  int n_peer, my_rank, row_begin;
  // sharded sibling batch: matrix j of the batch has its own LL vector on every rank;
  // llb[j][g] is a VIRTUAL base (pointer + first row of this rank's shard - rows before matrix j)
  unsigned long long *llb[4][LAMMCU_MAX_RANKS];
  // SURVEY 8f-3 (XM == 1 only): the op in front of the matmul, applied while the row is staged (lamm_xop.cuh)
  int xop_kind;  const float *x_aux;  float x_eps;  const uint16_t *silu_tab;
};

// ---- integer dots -------------------------------------------------------------------
__device__ __forceinline__ int dp4a_u8s8(uint32_t a, uint32_t b, int c) {
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
__device__ __forceinline__ int dp4a_s8s8(uint32_t a, uint32_t b, int c) {
  int d;
  asm("dp4a.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
// a: 2 x s16, b: 4 x u8 (lo: bytes 0,1; hi: bytes 2,3)
__device__ __forceinline__ int dp2a_lo_s16u8(uint32_t a, uint32_t b, int c) {
  int d;
  asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
__device__ __forceinline__ int dp2a_hi_s16u8(uint32_t a, uint32_t b, int c) {
  int d;
  asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

// ---- per-format constants -------------------------------------------------------------
template <int WT> struct DecFmt {
  static constexpr int BS = type_info(WT).size;          // bytes per 32-block
  static constexpr int NW = BS;                          // words per lane chunk (4 blocks)
  static constexpr int VT = type_info(WT).vec_dot;
  static constexpr bool kHasMin = (WT == LAMMCU_TYPE_Q4_1 || WT == LAMMCU_TYPE_Q5_1);
};

// NW consecutive words from a shared address aligned to 16 (NW % 4 == 0) or 8 bytes
template <int NW>
__device__ __forceinline__ void lds_chunk(uint32_t (&w)[NW], uint32_t addr) {
  if constexpr (NW % 4 == 0) {
#pragma unroll
    for (int i = 0; i < NW / 4; i++)
      asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                   : "=r"(w[4 * i]), "=r"(w[4 * i + 1]), "=r"(w[4 * i + 2]), "=r"(w[4 * i + 3])
                   : "r"(addr + 16u * i));
  } else {
#pragma unroll
    for (int i = 0; i < NW / 2; i++)
      asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(w[2 * i]), "=r"(w[2 * i + 1]) : "r"(addr + 8u * i));
  }
}
// the 32-bit word at compile-time byte offset OFF (OFF % 2 == 0) of the chunk
template <int OFF, int NW>
__device__ __forceinline__ uint32_t word_at(const uint32_t (&w)[NW]) {
  if constexpr (OFF % 4 == 0) return w[OFF / 4];
  else return __byte_perm(w[OFF / 4], w[OFF / 4 + 1], 0x5432);
}
template <int OFF, int NW>
__device__ __forceinline__ float half_at(const uint32_t (&w)[NW]) {
  if constexpr (OFF % 4 == 0) return half_bits_to_float(w[OFF / 4] & 0xffffu);
  else return half_bits_to_float(w[OFF / 4] >> 16);
}
// 4 bits -> bit 4 of 4 consecutive bytes (Q5 "5th bit": element j <- bit j)
__device__ __forceinline__ uint32_t spread_bit4(uint32_t nib) {
  return (nib * 0x02040810u) & 0x10101010u;
}

// One 32-block of the lane's chunk against the lane's register-resident activations.
// xq[0..3]: elements 0..15, xq[4..7]: elements 16..31; xd = f32(d_x);
// xs = f32(s_x) for q8_1, or bit-cast integer sum of the quants for q8_0.
template <int WT, int J, int NW>
__device__ __forceinline__ float block_dot8(const uint32_t (&w)[NW], const uint32_t (&xq)[8], float xd,
                                            float xs, float acc);

template <int WT, int J, int NW>
__device__ __forceinline__ float block_dot(const uint32_t (&w)[NW], const uint32_t (&xall)[32], float xd,
                                           float xs, float acc) {
  const uint32_t xq[8] = {xall[8 * J],     xall[8 * J + 1], xall[8 * J + 2], xall[8 * J + 3],
                          xall[8 * J + 4], xall[8 * J + 5], xall[8 * J + 6], xall[8 * J + 7]};
  return block_dot8<WT, J, NW>(w, xq, xd, xs, acc);
}

template <int WT, int J, int NW>
__device__ __forceinline__ float block_dot8(const uint32_t (&w)[NW], const uint32_t (&xq)[8], float xd,
                                            float xs, float acc) {
  constexpr int O = J * DecFmt<WT>::BS;
  if constexpr (WT == LAMMCU_TYPE_Q4_0 || WT == LAMMCU_TYPE_Q4_1) {
    constexpr int QS = O + (WT == LAMMCU_TYPE_Q4_0 ? 2 : 4);
    int slo = 0, shi = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const uint32_t q = (i == 0)   ? word_at<QS, NW>(w)
                         : (i == 1) ? word_at<QS + 4, NW>(w)
                         : (i == 2) ? word_at<QS + 8, NW>(w)
                                    : word_at<QS + 12, NW>(w);
      slo = dp4a_u8s8(q & 0x0F0F0F0Fu, xq[i], slo);
      shi = dp4a_u8s8(q & 0xF0F0F0F0u, xq[4 + i], shi);      // 16 x the high-nibble dot, exactly
    }
    int sumi = slo + (shi >> 4);
    if constexpr (WT == LAMMCU_TYPE_Q4_0) {
      sumi -= 8 * __float_as_int(xs);                          // sum((q-8) x) = sum(q x) - 8 sum(x)
      return fmaf(half_at<O, NW>(w) * xd, (float)sumi, acc);
    } else {
      acc = fmaf(half_at<O, NW>(w) * xd, (float)sumi, acc);
      return fmaf(half_at<O + 2, NW>(w), xs, acc);             // + m_w * s_x
    }
  } else if constexpr (WT == LAMMCU_TYPE_Q5_0 || WT == LAMMCU_TYPE_Q5_1) {
    constexpr int QH = O + (WT == LAMMCU_TYPE_Q5_0 ? 2 : 4);
    constexpr int QS = QH + 4;
    const uint32_t qh = word_at<QH, NW>(w);
    int sumi = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const uint32_t q = (i == 0)   ? word_at<QS, NW>(w)
                         : (i == 1) ? word_at<QS + 4, NW>(w)
                         : (i == 2) ? word_at<QS + 8, NW>(w)
                                    : word_at<QS + 12, NW>(w);
      const uint32_t lo = (q & 0x0F0F0F0Fu) | spread_bit4((qh >> (4 * i)) & 0xFu);
      const uint32_t hi = ((q >> 4) & 0x0F0F0F0Fu) | spread_bit4((qh >> (16 + 4 * i)) & 0xFu);
      sumi = dp4a_u8s8(lo, xq[i], sumi);
      sumi = dp4a_u8s8(hi, xq[4 + i], sumi);
    }
    if constexpr (WT == LAMMCU_TYPE_Q5_0) {
      sumi -= 16 * __float_as_int(xs);
      return fmaf(half_at<O, NW>(w) * xd, (float)sumi, acc);
    } else {
      acc = fmaf(half_at<O, NW>(w) * xd, (float)sumi, acc);
      return fmaf(half_at<O + 2, NW>(w), xs, acc);
    }
  } else {   // Q8_0
    constexpr int QS = O + 2;
    int sumi = 0;
    sumi = dp4a_s8s8(word_at<QS, NW>(w), xq[0], sumi);
    sumi = dp4a_s8s8(word_at<QS + 4, NW>(w), xq[1], sumi);
    sumi = dp4a_s8s8(word_at<QS + 8, NW>(w), xq[2], sumi);
    sumi = dp4a_s8s8(word_at<QS + 12, NW>(w), xq[3], sumi);
    sumi = dp4a_s8s8(word_at<QS + 16, NW>(w), xq[4], sumi);
    sumi = dp4a_s8s8(word_at<QS + 20, NW>(w), xq[5], sumi);
    sumi = dp4a_s8s8(word_at<QS + 24, NW>(w), xq[6], sumi);
    sumi = dp4a_s8s8(word_at<QS + 28, NW>(w), xq[7], sumi);
    return fmaf(half_at<O, NW>(w) * xd, (float)sumi, acc);
  }
}

// ---- activation staging: f32 (or ggml-quantized) row -> shared memory --------------------
// 32-element formats, U = units per row (kb / 4):
//   xqs : uint4 [8][U]   slot j of unit u = quant words (j&1)*4.. of block 4u + j/2
//   xds : float2 [kb]    (d, s) or (d, bit-cast int sum)
// Q2_K / Q8_K, U = 2 * super-blocks (unit = 128 elements):
//   xqs : uint4 [8][U]   slot j of unit u = quant words 4j..4j+3 of the unit (32 words)
//   xbs : uint4 [U]      8 x int16 bsums of the unit
//   xd  : float [kb]
// 8 consecutive activations starting at element e.  XM == 1: plain f32.  XM == 2: LL pairs written
// by the peer-store epilogues of all ranks; spins until all 8 tags equal `tag` (bounded: ~2 s,
// then status[0] = 1 and whatever is there is used, so a lost peer cannot hang the GPU).
__device__ __forceinline__ void ld_ll2(const unsigned long long *p, unsigned long long &a, unsigned long long &b) {
  asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}
template <int XM>
__device__ __forceinline__ void load_x8(const DecodeParams &p, size_t e, unsigned int tag, float4 &a, float4 &b) {
  if constexpr (XM == 2) {
    const unsigned long long *lp = p.x_ll + e;
    unsigned long long v[8];
    unsigned int spins = 0;
    unsigned long long t0 = 0;
    while (true) {
#pragma unroll
      for (int i = 0; i < 4; i++) ld_ll2(lp + 2 * i, v[2 * i], v[2 * i + 1]);
      bool ok = true;
#pragma unroll
      for (int i = 0; i < 8; i++) ok &= ((unsigned int)(v[i] >> 32) == tag);
      if (ok) break;
      __nanosleep(p.n_peer > 0 ? 100 : 20);   // a spinning warp would steal issue slots from a producer sharing this SM
      if ((++spins & 1023u) == 0) {
        unsigned long long now;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (t0 == 0) t0 = now;
        else if (now - t0 > 2000000000ull) { *p.status = 1u; break; }
      }
    }
    a = make_float4(__uint_as_float((unsigned int)v[0]), __uint_as_float((unsigned int)v[1]),
                    __uint_as_float((unsigned int)v[2]), __uint_as_float((unsigned int)v[3]));
    b = make_float4(__uint_as_float((unsigned int)v[4]), __uint_as_float((unsigned int)v[5]),
                    __uint_as_float((unsigned int)v[6]), __uint_as_float((unsigned int)v[7]));
  } else if constexpr (XM == 3) {
    // persistent chain kernel: x was written by OTHER CTAs of the same launch a moment ago and this SM's L1 may still
    // hold the buffer's previous contents (the activation vectors are reused across stages): read through L2
    const float *xp = p.x + e;
    asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w) : "l"(xp) : "memory");
    asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(xp + 4) : "memory");
  } else {
    const float4 *xp = reinterpret_cast<const float4 *>(p.x + e);
    a = xp[0];
    b = xp[1];
    // SURVEY 8f-3 (XM == 4 instantiations only: the plain kernel carries none of this): for f32 rows `tag` is free and
    // carries the bits of the row's RMS scale
    if constexpr (XM != 4) return;
    if (p.xop_kind == 1) {
      const float sc = __uint_as_float(tag);
      a = xop_norm4(a, sc, p.x_aux ? p.x_aux + e : nullptr);
      b = xop_norm4(b, sc, p.x_aux ? p.x_aux + e + 4 : nullptr);
    } else if (p.xop_kind == 2) {
      a = xop_silu4(a, p.x_aux + e, p.silu_tab);
      b = xop_silu4(b, p.x_aux + e + 4, p.silu_tab);
    }
  }
}

template <int VT, int XM>
__device__ __forceinline__ void stage_x_row(const DecodeParams &p, uint8_t *xs, unsigned int tag = 0u) {
  constexpr bool XF32 = (XM != 0);
  const int tid = threadIdx.x, lane = tid & 31;
  const int U = p.units;
  uint32_t *xqs = reinterpret_cast<uint32_t *>(xs);
  const bool store = (p.xq_out != nullptr) && blockIdx.x == 0 && XF32;
  if constexpr (VT == LAMMCU_TYPE_Q8_K) {
    uint32_t *xbs = xqs + (size_t)32 * U;
    float *xd = reinterpret_cast<float *>(xbs + (size_t)4 * U);
    for (int sb = tid >> 5; sb < p.kb; sb += (int)(blockDim.x >> 5)) {
      uint32_t q_lo, q_hi;
      float d;
      int pair_sum;
      if constexpr (XF32) {
        float4 a, b;
        load_x8<XM>(p, (size_t)sb * 256 + lane * 8, tag, a, b);
        const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        quant256_lane8(v, lane, q_lo, q_hi, d, pair_sum);
        if (store) {
          uint8_t *blk = p.xq_out + (size_t)sb * 292;
          uint32_t *qs = reinterpret_cast<uint32_t *>(blk + Q8_K_QS + lane * 8);
          qs[0] = q_lo;
          qs[1] = q_hi;
          if ((lane & 1) == 0) *reinterpret_cast<int16_t *>(blk + Q8_K_BSUMS + lane) = (int16_t)pair_sum;
          if (lane == 0) *reinterpret_cast<float *>(blk) = d;
        }
      } else {
        const uint8_t *blk = p.xq + (size_t)sb * 292;
        const uint32_t *qs = reinterpret_cast<const uint32_t *>(blk + Q8_K_QS + lane * 8);
        q_lo = qs[0];
        q_hi = qs[1];
        pair_sum = *reinterpret_cast<const int16_t *>(blk + Q8_K_BSUMS + (lane & ~1));
        d = *reinterpret_cast<const float *>(blk);
      }
      const int unit = 2 * sb + (lane >> 4), wi = (2 * lane) & 31;
      uint32_t *dst = xqs + ((size_t)(wi >> 2) * U + unit) * 4 + (wi & 3);
      dst[0] = q_lo;
      dst[1] = q_hi;
      if ((lane & 1) == 0) {
        const int g = lane >> 1;
        reinterpret_cast<int16_t *>(xbs)[(size_t)(2 * sb + (g >> 3)) * 8 + (g & 7)] = (int16_t)pair_sum;
      }
      if (lane == 0) xd[sb] = d;
    }
  } else {
    constexpr int XB = (VT == LAMMCU_TYPE_Q8_0) ? 34 : 36;
    float2 *xds = reinterpret_cast<float2 *>(xqs + (size_t)32 * U);
    if constexpr (XF32) {
      // 4 lanes x 8 floats per block, 64 blocks per pass, two passes' loads in flight
      const int l4 = tid & 3;
      const int per_pass = (int)(blockDim.x >> 2);
      for (int b0 = 0; b0 < p.kb; b0 += 2 * per_pass) {
        float4 va[2], vb[2];
#pragma unroll
        for (int h = 0; h < 2; h++) {
          const int b = b0 + h * per_pass + (tid >> 2);
          va[h] = vb[h] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (b < p.kb) load_x8<XM>(p, (size_t)b * 32 + l4 * 8, tag, va[h], vb[h]);
        }
#pragma unroll
        for (int h = 0; h < 2; h++) {
          const int b = b0 + h * per_pass + (tid >> 2);
          if (b0 + h * per_pass + ((tid & ~31) >> 2) >= p.kb) break;     // warp-uniform: nothing left
          uint32_t q_lo, q_hi;
          float d;
          int sum;
          quant32_lane8(va[h], vb[h], q_lo, q_hi, d, sum);
          if (b < p.kb) {
            const __half hd = __float2half_rn(d);
            __half hs = hd;
            float sf;
            if constexpr (VT == LAMMCU_TYPE_Q8_1) {
              hs = __float2half_rn(__fmul_rn((float)sum, d));
              sf = __half2float(hs);
            } else {
              sf = __int_as_float(sum);
            }
            const int u = b >> 2, j = (b & 3) * 2 + (l4 >> 1);
            *reinterpret_cast<uint2 *>(xqs + ((size_t)j * U + u) * 4 + (l4 & 1) * 2) = make_uint2(q_lo, q_hi);
            if (l4 == 0) xds[b] = make_float2(__half2float(hd), sf);
            if (store) {
              uint8_t *blk = p.xq_out + (size_t)b * XB;
              if constexpr (VT == LAMMCU_TYPE_Q8_1) {
                uint32_t *o = reinterpret_cast<uint32_t *>(blk + 4 + l4 * 8);
                o[0] = q_lo;
                o[1] = q_hi;
                if (l4 == 0) *reinterpret_cast<__half2 *>(blk) = __halves2half2(hd, hs);
              } else {
                uint16_t *o = reinterpret_cast<uint16_t *>(blk + 2 + l4 * 8);
                o[0] = (uint16_t)(q_lo & 0xffff);
                o[1] = (uint16_t)(q_lo >> 16);
                o[2] = (uint16_t)(q_hi & 0xffff);
                o[3] = (uint16_t)(q_hi >> 16);
                if (l4 == 0) *reinterpret_cast<__half *>(blk) = hd;
              }
            }
          }
        }
      }
      return;
    }
    const int l8 = tid & 7;
    for (int b0 = 0; b0 < p.kb; b0 += (int)(blockDim.x >> 3)) {
      const int b = b0 + (tid >> 3);
      const bool live = b < p.kb;
      uint32_t q = 0;
      float df = 0.f, sf = 0.f;
      if constexpr (XF32) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (live) v = *reinterpret_cast<const float4 *>(p.x + (size_t)b * 32 + l8 * 4);
        float d;
        int sum;
        q = quant32_lane4(v, d, sum);
        const __half hd = __float2half_rn(d);
        df = __half2float(hd);
        __half hs = __float2half_rn(0.f);
        if constexpr (VT == LAMMCU_TYPE_Q8_1) {
          hs = __float2half_rn(__fmul_rn((float)sum, d));
          sf = __half2float(hs);
        } else {
          sf = __int_as_float(sum);
        }
        if (store && live) {
          uint8_t *blk = p.xq_out + (size_t)b * XB;
          if constexpr (VT == LAMMCU_TYPE_Q8_1) {
            *reinterpret_cast<uint32_t *>(blk + 4 + l8 * 4) = q;
            if (l8 == 0) *reinterpret_cast<__half2 *>(blk) = __halves2half2(hd, hs);
          } else {
            uint16_t *o = reinterpret_cast<uint16_t *>(blk + 2 + l8 * 4);
            o[0] = (uint16_t)(q & 0xffff);
            o[1] = (uint16_t)(q >> 16);
            if (l8 == 0) *reinterpret_cast<__half *>(blk) = hd;
          }
        }
      } else {
        int sum = 0;
        if (live) {
          const uint8_t *blk = p.xq + (size_t)b * XB;
          if constexpr (VT == LAMMCU_TYPE_Q8_1) {
            q = *reinterpret_cast<const uint32_t *>(blk + 4 + 4 * l8);
            const uint32_t ds = *reinterpret_cast<const uint32_t *>(blk);
            df = half_bits_to_float(ds & 0xffffu);
            sf = half_bits_to_float(ds >> 16);
          } else {
            const uint16_t *h = reinterpret_cast<const uint16_t *>(blk + 2 + 4 * l8);
            q = (uint32_t)h[0] | ((uint32_t)h[1] << 16);
            df = half_bits_to_float(*reinterpret_cast<const uint16_t *>(blk));
          }
        }
        if constexpr (VT == LAMMCU_TYPE_Q8_0) {
          sum = dp4a_s8s8(q, 0x01010101u, 0);
          sum += __shfl_xor_sync(0xffffffffu, sum, 4);
          sum += __shfl_xor_sync(0xffffffffu, sum, 2);
          sum += __shfl_xor_sync(0xffffffffu, sum, 1);
          sf = __int_as_float(sum);
        }
      }
      if (live) {
        const int u = b >> 2, j = (b & 3) * 2 + (l8 >> 2);
        xqs[((size_t)j * U + u) * 4 + (l8 & 3)] = q;
        if (l8 == 0) xds[b] = make_float2(df, sf);
      }
    }
  }
}

// mbar_wait on a precomputed 32-bit shared address (no generic -> shared conversion on the critical path)
__device__ __forceinline__ void mbar_wait_u32(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "LAMM_WAITA_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra LAMM_DONEA_%=;\n\t"
      "bra LAMM_WAITA_%=;\n\t"
      "LAMM_DONEA_%=:\n\t"
      "}" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// event 0: %globaltimer (ns); events 1..7: SM cycle counter (cheap; deltas within a CTA only)
__device__ __forceinline__ void trace_stamp(unsigned long long *trace, int e) {
  if (trace && threadIdx.x == 0) {
    unsigned long long t;
    if (e == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    else asm volatile("mov.u64 %0, %%clock64;" : "=l"(t));
    trace[(size_t)blockIdx.x * 8 + e] = t;
  }
}

// ---- the kernel ----------------------------------------------------------------------
// One weight row-segment (the lane's unit of it) against the lane's activations.
template <int WT>
struct LaneX {                       // register-resident activations of one lane
  uint32_t q[32];
  float d[4], s[4];                  // Q2_K: d[0] = x.d
  uint32_t bs[4];                    // Q2_K: 8 x int16 bsums
};

template <int WT, int NW>
__device__ __forceinline__ void load_unit(uint32_t (&w)[NW], uint32_t slot_addr, int lane_c) {
  if constexpr (WT == LAMMCU_TYPE_Q2_K) {
    // half h of super-block lane_c/2: scales[8h..8h+8), qs[32h..32h+32), (d, dmin)
    const uint32_t blk = slot_addr + (uint32_t)(lane_c >> 1) * 84u, h = lane_c & 1;
#pragma unroll
    for (int k = 0; k < 8; k++) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w[k]) : "r"(blk + 16u + 32u * h + 4u * k));
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w[8]) : "r"(blk + 8u * h));
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w[9]) : "r"(blk + 8u * h + 4u));
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w[10]) : "r"(blk + 80u));
  } else {
    lds_chunk<NW>(w, slot_addr + (uint32_t)lane_c * (4u * NW));
  }
}

template <int WT, int NW>
__device__ __forceinline__ float dot_unit(const uint32_t (&w)[NW], const LaneX<WT> &x) {
  float acc = 0.f;
  if constexpr (WT == LAMMCU_TYPE_Q2_K) {
    const uint32_t sc0 = w[8] & 0x0F0F0F0Fu, sc1 = w[9] & 0x0F0F0F0Fu;
    const uint32_t mn0 = (w[8] >> 4) & 0x0F0F0F0Fu, mn1 = (w[9] >> 4) & 0x0F0F0F0Fu;
    int summs = dp2a_lo_s16u8(x.bs[0], mn0, 0);
    summs = dp2a_hi_s16u8(x.bs[1], mn0, summs);
    summs = dp2a_lo_s16u8(x.bs[2], mn1, summs);
    summs = dp2a_hi_s16u8(x.bs[3], mn1, summs);
    int isum = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {          // bit plane j: elements 32j.. of the unit
      const uint32_t mask = 0x03030303u << (2 * j);
#pragma unroll
      for (int e = 0; e < 2; e++) {        // sub-block 2j + e
        int t = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) t = dp4a_u8s8(w[4 * e + k] & mask, x.q[8 * j + 4 * e + k], t);
        const int g = 2 * j + e;
        const uint32_t scale = ((g < 4 ? sc0 : sc1) >> (8 * (g & 3))) & 0xFFu;
        isum += (int)scale * (t >> (2 * j));             // t is exactly 4^j x the plane dot
      }
    }
    const float wd = half_bits_to_float(w[10] & 0xffffu), wdmin = half_bits_to_float(w[10] >> 16);
    acc = fmaf(x.d[0] * wd, (float)isum, acc);
    acc = fmaf(-(x.d[0] * wdmin), (float)summs, acc);
  } else {
    acc = block_dot<WT, 0, NW>(w, x.q, x.d[0], x.s[0], acc);
    acc = block_dot<WT, 1, NW>(w, x.q, x.d[1], x.s[1], acc);
    acc = block_dot<WT, 2, NW>(w, x.q, x.d[2], x.s[2], acc);
    acc = block_dot<WT, 3, NW>(w, x.q, x.d[3], x.s[3], acc);
  }
  return acc;
}

// Sum each of a[0..3] over the 32 lanes with 6 shuffles; lane l (l < 4) returns the total of a[l].
__device__ __forceinline__ float warp_sum4(float a0, float a1, float a2, float a3, int lane) {
  const bool hi16 = lane & 16, hi8 = lane & 8;
  // lanes with bit 4 clear keep (a0, a1), the others (a2, a3)
  const float s0 = hi16 ? a0 : a2, s1 = hi16 ? a1 : a3;
  float k0 = (hi16 ? a2 : a0) + __shfl_xor_sync(0xffffffffu, s0, 16);
  float k1 = (hi16 ? a3 : a1) + __shfl_xor_sync(0xffffffffu, s1, 16);
  // bit 3 clear keeps k0, set keeps k1
  const float s = hi8 ? k0 : k1;
  float v = (hi8 ? k1 : k0) + __shfl_xor_sync(0xffffffffu, s, 8);
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  // lanes 0..7 hold a0 total, 8..15 a1, 16..23 a2, 24..31 a3
  return __shfl_sync(0xffffffffu, v, (lane & 3) * 8);
}

template <int WT, bool XF32, int MAXT>
__global__ void __launch_bounds__(MAXT, 2) decode_kernel(const DecodeParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr bool kQ2K = (WT == LAMMCU_TYPE_Q2_K);
  constexpr int VT = type_info(WT).vec_dot;
  constexpr int NW = kQ2K ? 11 : DecFmt<WT>::NW;
  constexpr int S = kDecMaxSlots;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint64_t *full = reinterpret_cast<uint64_t *>(smem) + warp * S;
  float *part = reinterpret_cast<float *>(smem + 512);          // [RG][2][KS][4]
  uint8_t *xs = smem + kDecCtrl;
  uint8_t *ring = xs + p.x_bytes + (size_t)warp * S * p.slot_bytes;

  const bool active = warp < p.RG * p.KS;
  const int rg = warp / p.KS, seg = warp - rg * p.KS;
  const int u0 = seg * p.seg_units;                              // first unit of the segment
  const int seg_n = max(0, min(p.seg_units, p.units - u0));      // live lanes
  constexpr int unit_bytes = kQ2K ? 42 : 4 * type_info(WT).size;
  // Q2_K: a unit is half a super-block, segments start on super-block boundaries
  const uint32_t seg_off = kQ2K ? (uint32_t)(u0 >> 1) * 84u : (uint32_t)u0 * unit_bytes;
  const uint32_t seg_bytes = kQ2K ? (uint32_t)((seg_n + 1) >> 1) * 84u : (uint32_t)seg_n * unit_bytes;
  const int row0 = blockIdx.x + gridDim.x * rg, row_step = gridDim.x * p.RG;
  const int n_rows = (row0 < p.M) ? (p.M - 1 - row0) / row_step + 1 : 0;      // rows of this warp
  const uint8_t *wseg = p.w + seg_off;
  const uint32_t ring_addr = smem_u32(ring);
  const uint32_t full_addr = smem_u32(full);

  auto issue = [&](int i, int s, uint64_t pol) {                 // lane 0: stream this warp's i-th row
    const uintptr_t a = (uintptr_t)(wseg + (int64_t)(row0 + i * row_step) * p.w_stride);
    const uintptr_t a0 = a & ~(uintptr_t)15;
    const uint32_t bytes = (uint32_t)(((a + seg_bytes + 15) & ~(uintptr_t)15) - a0);
    uint64_t *bar = full + s;
    mbar_arrive_expect_tx(bar, bytes);
    bulk_g2s_stream(ring + (size_t)s * p.slot_bytes, reinterpret_cast<const void *>(a0), bytes, bar, pol);
  };

  trace_stamp(p.trace, 0);
  trace_stamp(p.trace, 1);
  uint64_t pol = 0;
  if (active && lane == 0) {
#pragma unroll
    for (int s = 0; s < S; s++) mbar_init(&full[s], 1);
    fence_barrier_init();
    pol = policy_evict_first();
    // weights are immutable: start streaming before the previous kernel has finished
#pragma unroll
    for (int s = 0; s < S; s++)
      if (s < n_rows) issue(s, s, pol);
  }
  pdl_launch_dependents();
  pdl_wait();                                                    // activations are valid from here on
  trace_stamp(p.trace, 2);
  stage_x_row<VT, XF32 ? 1 : 0>(p, xs);
  __syncthreads();
  trace_stamp(p.trace, 3);
  if (!active) return;

  // ---- the lane's activations, register-resident for the whole kernel ----
  const int U = p.units, ug = u0 + lane;
  const bool lane_live = lane < seg_n;
  const uint4 *xqs = reinterpret_cast<const uint4 *>(xs);
  LaneX<WT> X;
#pragma unroll
  for (int j = 0; j < 8; j++) {
    const uint4 v = lane_live ? xqs[(size_t)j * U + ug] : make_uint4(0u, 0u, 0u, 0u);
    X.q[4 * j] = v.x; X.q[4 * j + 1] = v.y; X.q[4 * j + 2] = v.z; X.q[4 * j + 3] = v.w;
  }
#pragma unroll
  for (int j = 0; j < 4; j++) { X.d[j] = 0.f; X.s[j] = 0.f; X.bs[j] = 0u; }
  if constexpr (kQ2K) {
    const uint4 *xbs = xqs + (size_t)8 * U;
    const float *xdp = reinterpret_cast<const float *>(xbs + U);
    if (lane_live) {
      const uint4 b = xbs[ug];
      X.bs[0] = b.x; X.bs[1] = b.y; X.bs[2] = b.z; X.bs[3] = b.w;
      X.d[0] = xdp[ug >> 1];
    }
  } else {
    const float2 *xds = reinterpret_cast<const float2 *>(xqs + (size_t)8 * U);
    if (lane_live) {
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const float2 v = xds[4 * ug + j];
        X.d[j] = v.x;
        X.s[j] = v.y;
      }
    }
  }

  trace_stamp(p.trace, 5);
  const int lane_c = lane_live ? lane : 0;     // dead lanes re-read unit 0 (their activations are 0)
  // every row of this warp has the same (address & 15): w_stride % 16 == 0 or Q2_K 4-byte rows
  for (int i0 = 0, ph = 0; i0 < n_rows; i0 += S, ph ^= 1) {
    const int nb = min(S, n_rows - i0);          // rows in this batch (warp-uniform)
    float acc[S];
#pragma unroll
    for (int s = 0; s < S; s++) acc[s] = 0.f;
    const uint32_t a15 = (uint32_t)((uintptr_t)(wseg + (int64_t)row0 * p.w_stride) & 15);
#pragma unroll
    for (int pr = 0; pr < S; pr += 2) {          // two rows at a time: 2x the independent dp4a chains
      if (pr < nb) {
        const bool two = pr + 1 < nb;
        uint32_t w0[NW], w1[NW];
        mbar_wait(full + pr, (uint32_t)ph);
        if (two) mbar_wait(full + pr + 1, (uint32_t)ph);
        load_unit<WT, NW>(w0, ring_addr + (uint32_t)pr * p.slot_bytes + a15, lane_c);
        load_unit<WT, NW>(w1, ring_addr + (uint32_t)(two ? pr + 1 : pr) * p.slot_bytes + a15, lane_c);
        const float d0 = dot_unit<WT, NW>(w0, X);
        const float d1 = dot_unit<WT, NW>(w1, X);
        acc[pr] = d0;
        acc[pr + 1] = two ? d1 : 0.f;
        __syncwarp();                            // all lanes have consumed both slots
        if (i0 == 0) trace_stamp(p.trace, 6 + pr / 2);
        if (lane == 0) {
          if (i0 + S + pr < n_rows) issue(i0 + S + pr, pr, pol);
          if (i0 + S + pr + 1 < n_rows) issue(i0 + S + pr + 1, pr + 1, pol);
        }
      }
    }
    const float tot = warp_sum4(acc[0], acc[1], acc[2], acc[3], lane);
    if (p.KS == 1) {
      if (lane < nb) p.c[row0 + (i0 + lane) * row_step] = tot;
    } else {
      float *pb = part + (size_t)((rg * 2 + ph) * p.KS) * 4;
      if (lane < 4) pb[seg * 4 + lane] = tot;
      asm volatile("bar.sync %0, %1;" ::"r"(1 + rg), "r"(p.KS * 32) : "memory");
      if (seg == 0 && lane < nb) {
        float t = 0.f;
        for (int k = 0; k < p.KS; k++) t += pb[k * 4 + lane];
        p.c[row0 + (i0 + lane) * row_step] = t;
      }
    }
  }
  (void)full_addr;
  trace_stamp(p.trace, 4);
}

// ---- the 16-warp variant ---------------------------------------------------------------
// Same streaming structure, tuned for the latency of SHORT kernels (a 4096x4096 Q4_1 matrix is
// 71 KB per SM: the whole mul_mat is one DRAM round trip plus a serial phase after the
// activations arrive).  16 warps at <= 64 registers so that two CTAs -- this mul_mat and the
// next one of the chain, already prefetching -- are resident per SM, 4 warps per scheduler
// during the serial phase.  The lane keeps only the activation SCALES in registers and re-reads
// the activation quants from shared memory (conflict-free uint4 slots) for every row.
constexpr int kD16Warps = 16;
constexpr int kD16Slots = 2;

// lanes 0..15 return the total of a0, lanes 16..31 of a1 ... then lane l < 2 gets a_l
__device__ __forceinline__ float warp_sum2(float a0, float a1, int lane) {
  const bool hi16 = lane & 16;
  float v = (hi16 ? a1 : a0) + __shfl_xor_sync(0xffffffffu, hi16 ? a0 : a1, 16);
  v += __shfl_xor_sync(0xffffffffu, v, 8);
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  return __shfl_sync(0xffffffffu, v, (lane & 1) * 16);
}

__device__ __forceinline__ void lds128(uint32_t (&v)[4], uint32_t addr) {
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(addr));
}

// xaddr: shared address of uint4 slot 0 of the lane's unit; jstride: bytes between slots j
template <int WT, int NW>
__device__ __forceinline__ float dot_unit_sx(const uint32_t (&w)[NW], uint32_t xaddr, uint32_t jstride,
                                             const float (&xd)[4], const float (&xs)[4], const uint32_t (&xbs)[4]) {
  float acc = 0.f;
  if constexpr (WT == LAMMCU_TYPE_Q2_K) {
    const uint32_t sc0 = w[8] & 0x0F0F0F0Fu, sc1 = w[9] & 0x0F0F0F0Fu;
    const uint32_t mn0 = (w[8] >> 4) & 0x0F0F0F0Fu, mn1 = (w[9] >> 4) & 0x0F0F0F0Fu;
    int summs = dp2a_lo_s16u8(xbs[0], mn0, 0);
    summs = dp2a_hi_s16u8(xbs[1], mn0, summs);
    summs = dp2a_lo_s16u8(xbs[2], mn1, summs);
    summs = dp2a_hi_s16u8(xbs[3], mn1, summs);
    int isum = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const uint32_t mask = 0x03030303u << (2 * j);
#pragma unroll
      for (int e = 0; e < 2; e++) {
        uint32_t x[4];
        lds128(x, xaddr + (uint32_t)(2 * j + e) * jstride);
        int t = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) t = dp4a_u8s8(w[4 * e + k] & mask, x[k], t);
        const int g = 2 * j + e;
        const uint32_t scale = ((g < 4 ? sc0 : sc1) >> (8 * (g & 3))) & 0xFFu;
        isum += (int)scale * (t >> (2 * j));
      }
    }
    const float wd = half_bits_to_float(w[10] & 0xffffu), wdmin = half_bits_to_float(w[10] >> 16);
    acc = fmaf(xd[0] * wd, (float)isum, acc);
    acc = fmaf(-(xd[0] * wdmin), (float)summs, acc);
  } else {
    uint32_t a[4], b[4];
#define LAMM_D16_BLOCK(J)                                                                   \
    lds128(a, xaddr + (uint32_t)(2 * J) * jstride);                                         \
    lds128(b, xaddr + (uint32_t)(2 * J + 1) * jstride);                                     \
    {                                                                                       \
      const uint32_t xq[8] = {a[0], a[1], a[2], a[3], b[0], b[1], b[2], b[3]};              \
      acc = block_dot8<WT, J, NW>(w, xq, xd[J], xs[J], acc);                                \
    }
    LAMM_D16_BLOCK(0)
    LAMM_D16_BLOCK(1)
    LAMM_D16_BLOCK(2)
    LAMM_D16_BLOCK(3)
#undef LAMM_D16_BLOCK
  }
  return acc;
}

// ---- K-quants beyond the reference's table: Q3_K / Q4_K / Q5_K / Q6_K x Q8_K ----------
// Behaviour matched: ggml_vec_dot_q{3,4,5,6}_K_q8_K (ggml-quants.c:6127, 7000, 7609, 8305).  Same lane unit
// as Q2_K -- half a super-block, 128 consecutive elements = activation slots 0..7 (16 int8 each) -- but
// the weight words are pulled from the ring as they are needed (a unit is up to 112 bytes; holding it all
// would not fit the 64-register budget).  Q3_K / Q6_K super-blocks (110 / 210 bytes) are only 2-byte
// aligned in the ring: their words are assembled from aligned pairs with a funnel shift.
template <int N, bool A2>
__device__ __forceinline__ void lds_words(uint32_t (&out)[N], uint32_t addr) {
  if constexpr (!A2) {
    if constexpr (N % 4 == 0) {
#pragma unroll
      for (int i = 0; i < N / 4; i++)
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(out[4 * i]), "=r"(out[4 * i + 1]), "=r"(out[4 * i + 2]), "=r"(out[4 * i + 3])
                     : "r"(addr + 16u * i));
    } else {
#pragma unroll
      for (int i = 0; i < N; i++) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(out[i]) : "r"(addr + 4u * i));
    }
  } else {
    const uint32_t base = addr & ~3u, sh = (addr & 2u) * 8u;
    uint32_t w[N + 1];
#pragma unroll
    for (int i = 0; i <= N; i++) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w[i]) : "r"(base + 4u * i));
#pragma unroll
    for (int i = 0; i < N; i++) out[i] = __funnelshift_r(w[i], w[i + 1], sh);
  }
}
__device__ __forceinline__ int dot16(const uint32_t *q, const uint32_t (&x)[4], int acc) {
  acc = dp4a_u8s8(q[0], x[0], acc);
  acc = dp4a_u8s8(q[1], x[1], acc);
  acc = dp4a_u8s8(q[2], x[2], acc);
  return dp4a_u8s8(q[3], x[3], acc);
}
__device__ __forceinline__ int s16_lo(uint32_t v) { return (int)(short)(v & 0xffffu); }
__device__ __forceinline__ int s16_hi(uint32_t v) { return (int)v >> 16; }

// blk: shared address of the lane's super-block in the ring slot; h: which half; xaddr / jstride: the lane's
// activation slots; xd: x.d of the super-block; xbs: the 8 int16 bsums of the half
template <int WT>
__device__ __forceinline__ float dot_unit_kx(uint32_t blk, int h, uint32_t xaddr, uint32_t jstride, float xd,
                                             const uint32_t (&xbs)[4]) {
  uint32_t x[4];
  if constexpr (WT == LAMMCU_TYPE_Q4_K || WT == LAMMCU_TYPE_Q5_K) {
    // {f16 d, dmin; u8 scales[12]; (Q5_K: u8 qh[32];) u8 qs[128]}: 16-byte aligned
    constexpr uint32_t QS = (WT == LAMMCU_TYPE_Q4_K) ? 16u : 48u;
    uint32_t hdr[4];
    lds_words<4, false>(hdr, blk);
    // get_scale_min_k4 (ggml-quants.c:2027-2034) for groups 4h .. 4h+3, four at a time
    const uint32_t scw = h ? ((hdr[3] & 0x0F0F0F0Fu) | (((hdr[1] >> 6) & 0x03030303u) << 4)) : (hdr[1] & 0x3F3F3F3Fu);
    const uint32_t mnw = h ? (((hdr[3] >> 4) & 0x0F0F0F0Fu) | (((hdr[2] >> 6) & 0x03030303u) << 4))
                           : (hdr[2] & 0x3F3F3F3Fu);
    int isum = 0;
#pragma unroll
    for (int pr = 0; pr < 2; pr++) {               // 32 bytes: low nibbles = group 4h+2pr, high = 4h+2pr+1
      uint32_t q[8], lo[8], hi[8];
      lds_words<8, false>(q, blk + QS + 64u * h + 32u * pr);
#pragma unroll
      for (int i = 0; i < 8; i++) { lo[i] = q[i] & 0x0F0F0F0Fu; hi[i] = (q[i] >> 4) & 0x0F0F0F0Fu; }
      if constexpr (WT == LAMMCU_TYPE_Q5_K) {
        uint32_t qh[8];
        lds_words<8, false>(qh, blk + 16u);        // bit g of qh[l]: 5th bit of element 32g + l
#pragma unroll
        for (int i = 0; i < 8; i++) {
          const uint32_t b = qh[i] >> (4 * h + 2 * pr);
          lo[i] |= (b & 0x01010101u) << 4;
          hi[i] |= ((b >> 1) & 0x01010101u) << 4;
        }
      }
      int tl = 0, th = 0;
      lds128(x, xaddr + (uint32_t)(4 * pr) * jstride);      tl = dot16(lo, x, tl);
      lds128(x, xaddr + (uint32_t)(4 * pr + 1) * jstride);  tl = dot16(lo + 4, x, tl);
      lds128(x, xaddr + (uint32_t)(4 * pr + 2) * jstride);  th = dot16(hi, x, th);
      lds128(x, xaddr + (uint32_t)(4 * pr + 3) * jstride);  th = dot16(hi + 4, x, th);
      isum += (int)((scw >> (16 * pr)) & 0xFFu) * tl + (int)((scw >> (16 * pr + 8)) & 0xFFu) * th;
    }
    // min term: m[g] * (bsums[2g] + bsums[2g+1]) for the half's 4 groups
    const uint32_t m01 = __byte_perm(mnw, 0u, 0x1100), m23 = __byte_perm(mnw, 0u, 0x3322);
    int summs = dp2a_lo_s16u8(xbs[0], m01, 0);
    summs = dp2a_hi_s16u8(xbs[1], m01, summs);
    summs = dp2a_lo_s16u8(xbs[2], m23, summs);
    summs = dp2a_hi_s16u8(xbs[3], m23, summs);
    const float wd = half_bits_to_float(hdr[0] & 0xffffu), wdmin = half_bits_to_float(hdr[0] >> 16);
    float acc = (xd * wd) * (float)isum;
    return fmaf(-(xd * wdmin), (float)summs, acc);
  } else if constexpr (WT == LAMMCU_TYPE_Q6_K) {
    // {u8 ql[128]; u8 qh[64]; i8 scales[16]; f16 d}: 2-byte aligned.  Plane j (elements 32j.. of the half):
    // j = 0: ql[l] & 15, 1: ql[32 + l] & 15, 2: ql[l] >> 4, 3: ql[32 + l] >> 4, each | ((qh[l] >> 2j) & 3) << 4
    uint32_t hq[8], sc2[2], dw[1];
    lds_words<8, true>(hq, blk + 128u + 32u * h);
    lds_words<2, true>(sc2, blk + 192u + 8u * h);
    lds_words<1, true>(dw, blk + 208u);
    int isum = 0;
#pragma unroll
    for (int ab = 0; ab < 2; ab++) {               // ql[0..31] (planes 0, 2) then ql[32..63] (planes 1, 3)
      uint32_t q[8], lo[8], hi[8];
      lds_words<8, true>(q, blk + 64u * h + 32u * ab);
#pragma unroll
      for (int i = 0; i < 8; i++) {
        lo[i] = (q[i] & 0x0F0F0F0Fu) | (((hq[i] >> (2 * ab)) & 0x03030303u) << 4);
        hi[i] = ((q[i] >> 4) & 0x0F0F0F0Fu) | (((hq[i] >> (2 * ab + 4)) & 0x03030303u) << 4);
      }
      // sub-blocks 2j, 2j+1 of the half; sum (q - 32) x = sum q x - 32 bsum
      const int jl = ab, jh = ab + 2;
      int t;
      lds128(x, xaddr + (uint32_t)(2 * jl) * jstride);      t = dot16(lo, x, 0);
      isum += (int)(signed char)((sc2[jl >> 1] >> (16 * (jl & 1))) & 0xFFu) * (t - 32 * s16_lo(xbs[jl]));
      lds128(x, xaddr + (uint32_t)(2 * jl + 1) * jstride);  t = dot16(lo + 4, x, 0);
      isum += (int)(signed char)((sc2[jl >> 1] >> (16 * (jl & 1) + 8)) & 0xFFu) * (t - 32 * s16_hi(xbs[jl]));
      lds128(x, xaddr + (uint32_t)(2 * jh) * jstride);      t = dot16(hi, x, 0);
      isum += (int)(signed char)((sc2[jh >> 1] >> (16 * (jh & 1))) & 0xFFu) * (t - 32 * s16_lo(xbs[jh]));
      lds128(x, xaddr + (uint32_t)(2 * jh + 1) * jstride);  t = dot16(hi + 4, x, 0);
      isum += (int)(signed char)((sc2[jh >> 1] >> (16 * (jh & 1) + 8)) & 0xFFu) * (t - 32 * s16_hi(xbs[jh]));
    }
    return (xd * half_bits_to_float(dw[0] & 0xffffu)) * (float)isum;
  } else {
    // Q3_K {u8 hmask[32]; u8 qs[64]; u8 scales[12]; f16 d}: 2-byte aligned.  Plane j of half h:
    // (qs[32h + l] >> 2j) & 3, minus 4 unless bit 4h + j of hmask[l] is set
    uint32_t hm[8], q[8], s3[3], dw[1];
    lds_words<8, true>(hm, blk);
    lds_words<8, true>(q, blk + 32u + 32u * h);
    lds_words<3, true>(s3, blk + 96u);
    lds_words<1, true>(dw, blk + 108u);
    // the half's eight 6-bit scales as bytes of (slo, shi) (ggml-quants.c:2570-2575)
    const uint32_t slo = h ? (((s3[0] >> 4) & 0x0F0F0F0Fu) | (((s3[2] >> 4) & 0x03030303u) << 4))
                           : ((s3[0] & 0x0F0F0F0Fu) | ((s3[2] & 0x03030303u) << 4));
    const uint32_t shi = h ? (((s3[1] >> 4) & 0x0F0F0F0Fu) | (((s3[2] >> 6) & 0x03030303u) << 4))
                           : ((s3[1] & 0x0F0F0F0Fu) | (((s3[2] >> 2) & 0x03030303u) << 4));
    int isum = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      uint32_t v[8];
#pragma unroll
      for (int i = 0; i < 8; i++)
        v[i] = ((q[i] >> (2 * j)) & 0x03030303u) | (((hm[i] >> (4 * h + j)) & 0x01010101u) << 2);
      const uint32_t sw = (j < 2 ? slo : shi) >> (16 * (j & 1));
      int t;
      lds128(x, xaddr + (uint32_t)(2 * j) * jstride);      t = dot16(v, x, 0);
      isum += ((int)(sw & 0xFFu) - 32) * (t - 4 * s16_lo(xbs[j]));
      lds128(x, xaddr + (uint32_t)(2 * j + 1) * jstride);  t = dot16(v + 4, x, 0);
      isum += ((int)((sw >> 8) & 0xFFu) - 32) * (t - 4 * s16_hi(xbs[j]));
    }
    return (xd * half_bits_to_float(dw[0] & 0xffffu)) * (float)isum;
  }
}

__device__ __forceinline__ void store_c(const DecodeParams &p, int row, float v, unsigned int tag) {
  if (p.n_peer == 0) {
    float *c = p.c;
    if (p.nbatch > 1) {
      c = p.cb[0];
      if (row >= p.mc[0]) c = p.cb[1];
      if (row >= p.mc[1]) c = p.cb[2];
      if (row >= p.mc[2]) c = p.cb[3];
    }
    c[row] = v;
  } else {
    const unsigned long long pair = ((unsigned long long)tag << 32) | (unsigned long long)__float_as_uint(v);
    if (p.nbatch > 1) {
      int j = 0;
      if (row >= p.mc[0]) j = 1;
      if (row >= p.mc[1]) j = 2;
      if (row >= p.mc[2]) j = 3;
      for (int g = 0; g < p.n_peer; g++)
        asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p.llb[j][g] + row), "l"(pair) : "memory");
      return;
    }
    for (int g = 0; g < p.n_peer; g++)
      asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p.ll_peer[g] + p.row_begin + row), "l"(pair) : "memory");
  }
}

__device__ __forceinline__ void srv_stamp(const ServerParams *sp, unsigned int job, int e) {
  if (sp->trace && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    sp->trace[(size_t)(job & 63u) * 8 + e] = t;
  }
}

// SERVER: the body runs as one job of the resident kernel (decode_server_kernel below): bid / nblk
// are the job's virtual grid, the activations arrive as {value, tag} pairs staged from pinned host
// memory by CTA 0, the results go to the `cs` staging row, and the ring barriers are invalidated for the next job.
// Each warp keeps RD = 2 rows in flight (its private ring).  Deeper rings (3 / 4 slots where two CTAs still fit on an
// SM) and a two-rows-per-activation-read dot were measured in round 2 and removed: no format gained from the deeper
// ring (Q2_K / Q3_K lost 1-8 %, profiles/r2a/sweep_ring*.jsonl) and the row pair spilled (1094 -> 922 tok/s).
template <int WT, int XM, bool SERVER>
__device__ __forceinline__ void decode16_body(const DecodeParams &p, uint8_t *smem, const int bid, const int nblk,
                                              const ServerParams *sp, const unsigned int job_tag, float *cs) {
  constexpr bool kKX = is_kquant_ext(WT);                        // Q3_K..Q6_K: Q2_K's geometry, own dot
  constexpr bool kEarlyRefill = !(WT == LAMMCU_TYPE_Q5_0 || WT == LAMMCU_TYPE_Q5_1);
  constexpr bool kQ2K = (WT == LAMMCU_TYPE_Q2_K) || kKX;         // unit = half a super-block
  constexpr uint32_t SBB = (uint32_t)type_info(WT).size;         // bytes per super-block (84 for Q2_K)
  constexpr int VT = type_info(WT).vec_dot;
  constexpr int NW = kKX ? 4 : (kQ2K ? 11 : DecFmt<WT>::NW);
  constexpr int S = kD16Slots;                                   // rows per loop iteration (one shuffle tree)
  constexpr int RD = kD16Slots;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint64_t *full = reinterpret_cast<uint64_t *>(smem) + warp * RD;
  float *part = reinterpret_cast<float *>(smem + 512);          // [RG][2][KS][2]
  uint8_t *xs = smem + kDecCtrl;
  uint8_t *ring = xs + p.x_bytes + (size_t)warp * RD * p.slot_bytes;

  const bool active = warp < p.RG * p.KS;
  const int rg = warp / p.KS, seg = warp - rg * p.KS;
  const int u0 = seg * p.seg_units;
  const int seg_n = max(0, min(p.seg_units, p.units - u0));
  constexpr int unit_bytes = kQ2K ? 42 : 4 * type_info(WT).size;
  const uint32_t seg_off = kQ2K ? (uint32_t)(u0 >> 1) * SBB : (uint32_t)u0 * unit_bytes;
  const uint32_t seg_bytes = kQ2K ? (uint32_t)((seg_n + 1) >> 1) * SBB : (uint32_t)seg_n * unit_bytes;
  // rows of this warp: interleaved over the grid (a launch), or -- SERVER, where nblk is the number of
  // rows per CTA -- a contiguous chunk per CTA so that its results leave as one coalesced PCIe write
  const int chunk_lo = SERVER ? bid * nblk : 0;
  const int chunk_hi = SERVER ? min(p.M, chunk_lo + nblk) : p.M;
  const int row0 = SERVER ? chunk_lo + rg : bid + nblk * rg;
  const int row_step = SERVER ? p.RG : nblk * p.RG;
  const int n_rows = (row0 < chunk_hi) ? (chunk_hi - 1 - row0) / row_step + 1 : 0;
  const uint8_t *wseg = p.w + seg_off;
  const uint32_t ring_addr = smem_u32(ring);

  auto row_ptr = [&](int r) -> uintptr_t {       // this warp's K-segment of (virtual) row r
    const uint8_t *base = wseg;
    if (p.nbatch > 1) {
      base = p.wb[0];
      if (r >= p.mc[0]) base = p.wb[1];
      if (r >= p.mc[1]) base = p.wb[2];
      if (r >= p.mc[2]) base = p.wb[3];
      base += seg_off;
    }
    return (uintptr_t)(base + (int64_t)r * p.w_stride);
  };
  auto issue = [&](int i, int s, uint64_t pol) {
    const uintptr_t a = row_ptr(row0 + i * row_step);
    const uintptr_t a0 = a & ~(uintptr_t)15;
    const uint32_t bytes = (uint32_t)(((a + seg_bytes + 15) & ~(uintptr_t)15) - a0);
    uint64_t *bar = full + s;
    mbar_arrive_expect_tx(bar, bytes);
    bulk_g2s_stream(ring + (size_t)s * p.slot_bytes, reinterpret_cast<const void *>(a0), bytes, bar, pol);
    // lookahead beyond the ring (LAMM_DECODE_L2_AHEAD, default 0 = never taken; see DecodeParams::l2_ahead)
    if (!SERVER && p.l2_ahead > 0 && i + p.l2_ahead < n_rows) {
      const uintptr_t b = row_ptr(row0 + (i + p.l2_ahead) * row_step);
      const uintptr_t b0 = b & ~(uintptr_t)15;
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(reinterpret_cast<const void *>(b0)),
                   "r"((uint32_t)(((b + seg_bytes + 15) & ~(uintptr_t)15) - b0)) : "memory");
    }
  };

  trace_stamp(p.trace, 0);
  trace_stamp(p.trace, 1);
  uint64_t pol = 0;
  if (active && lane == 0) {
#pragma unroll
    for (int s = 0; s < RD; s++) mbar_init(&full[s], 1);
    fence_barrier_init();
    pol = policy_evict_first();
#pragma unroll
    for (int s = 0; s < RD; s++)
      if (s < n_rows) issue(s, s, pol);
  }
  // ---- everything a lane needs that does NOT depend on the activations is computed (and pinned in
  // registers) before the dependency resolves: the serial phase after griddepcontrol.wait is the launch's
  // critical path, 16 warps in lockstep, and ~80 instructions of index math there cost ~0.5 us
  const int U = p.units, ug = u0 + lane;
  const bool lane_live = lane < seg_n;
  const int lane_c = lane_live ? lane : 0;
  // dead lanes read unit u0's quants against zero scales (Q2_K: zero d and bsums)
  uint32_t xaddr = smem_u32(xs) + (uint32_t)(lane_live ? ug : u0) * 16u;
  uint32_t jstride = (uint32_t)U * 16u;
  // scales of the lane's unit: float2 [kb] behind the quants (Q8_K: uint4 bsums [U], then float d [kb])
  uint32_t xs_addr = smem_u32(xs) + 8u * jstride + (kQ2K ? (uint32_t)ug * 16u : (uint32_t)ug * 32u);
  uint32_t xd_addr = smem_u32(xs) + 9u * jstride + (uint32_t)(ug >> 1) * 4u;          // Q8_K only
  // every row of this warp has the same (address & 15) -- also across the matrices of a batch (host check)
  uint32_t a15 = (uint32_t)((uintptr_t)((p.nbatch > 1 ? p.wb[0] + seg_off : wseg) + (int64_t)row0 * p.w_stride) & 15);
  uint32_t full_u32 = smem_u32(full);
  asm volatile("" : "+r"(xaddr), "+r"(jstride), "+r"(xs_addr), "+r"(xd_addr), "+r"(a15), "+r"(full_u32));

  unsigned int tag = 0u, x_tag = 0u;
  if constexpr (SERVER) {
    x_tag = job_tag;                               // the server kernel has staged X as {value, tag} pairs
  } else if (p.n_peer > 0) {
    // This launch's tag = its sequence number on the communicator, from an arrival ticket: every CTA of
    // every sharded launch (all have gridDim.x == p.grid_fixed) adds 1 to a rank-local counter BEFORE it
    // lets the next launch start, so when a CTA of this launch draws its ticket all CTAs of all earlier
    // launches have drawn theirs: ticket / grid = number of earlier launches.  No dependence on the
    // previous grid's COMPLETION, so a launch whose activations are an LL vector never executes
    // griddepcontrol.wait: its inputs are validated element by element through their tags.
    volatile unsigned int *tag_s = reinterpret_cast<volatile unsigned int *>(smem + kDecCtrl - 16);
    if (threadIdx.x == 0) *tag_s = atomicAdd(p.arrive_ctr, 1u) / (unsigned int)nblk + 1u;
    __syncthreads();
    tag = *tag_s;
    x_tag = tag - (unsigned int)p.x_back;
    pdl_launch_dependents();
    if constexpr (XM != 2) pdl_wait();
    if (bid >= p.M) return;              // padding CTA (the grid is fixed): owns no row
    if constexpr (XM == 2) {
      // Sleep until the activations are (nearly) there: lane g of warp 0 watches the LAST element of
      // rank g's slice of x with a back-off while every other warp is parked at the barrier (no issue
      // slots, no L2 traffic against the producer that shares this SM).  This is only a wake-up hint:
      // stage_x_row still validates every element's tag.
      if (warp == 0 && lane < p.n_peer) {
        const int K = p.kb * type_info(WT).blck;
        const int per = (K + p.n_peer - 1) / p.n_peer;
        const unsigned long long *sp = p.x_ll + (min(K, (lane + 1) * per) - 1);
        unsigned long long v, t0 = 0;
        unsigned int spins = 0;
        while (true) {
          asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(sp) : "memory");
          if ((unsigned int)(v >> 32) == x_tag) break;
          __nanosleep(40);
          if ((++spins & 1023u) == 0) {
            unsigned long long now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            if (t0 == 0) t0 = now;
            else if (now - t0 > 2000000000ull) break;      // stage_x_row reports the timeout
          }
        }
      }
      __syncthreads();
    }
  } else {
    pdl_launch_dependents();
    pdl_wait();
  }
  trace_stamp(p.trace, 2);
  if constexpr (XM == 4 && !SERVER) {
    // llm_build_norm in front of this matmul: every CTA reduces the row itself (K floats, L2-resident)
    if (p.xop_kind == 1) x_tag = __float_as_uint(cta_rms_scale(p.x, p.kb * (int)type_info(WT).blck, p.x_eps, smem + 256));   // [256, 516): free until the first partial sum
  }
  if constexpr (SERVER) {
    // p.x_back != 0 (job word 15): the host found this call's activation row identical to the previous job's
    // (q/k/v and gate/up share src1), and this instance quantized that row into xs then: nothing crossed
    // PCIe for it and nothing is staged again (the ring and the result row never overlap xs)
    if (p.x_back == 0) stage_x_row<VT, XM>(p, xs, x_tag);
  } else {
    stage_x_row<VT, XM>(p, xs, x_tag);
  }
  __syncthreads();
  trace_stamp(p.trace, 3);
  if constexpr (SERVER) { if (bid == 0) srv_stamp(sp, job_tag, 2); }   // activations quantized
  if (active) {

  float xd[4] = {0.f, 0.f, 0.f, 0.f}, xsc[4] = {0.f, 0.f, 0.f, 0.f};
  uint32_t xbs[4] = {0u, 0u, 0u, 0u};
  if (lane_live) {
    if constexpr (kQ2K) {
      lds128(xbs, xs_addr);
      asm volatile("ld.shared.f32 %0, [%1];" : "=f"(xd[0]) : "r"(xd_addr));
    } else {
      uint32_t v0[4], v1[4];
      lds128(v0, xs_addr);
      lds128(v1, xs_addr + 16u);
      xd[0] = __uint_as_float(v0[0]); xsc[0] = __uint_as_float(v0[1]); xd[1] = __uint_as_float(v0[2]); xsc[1] = __uint_as_float(v0[3]);
      xd[2] = __uint_as_float(v1[0]); xsc[2] = __uint_as_float(v1[1]); xd[3] = __uint_as_float(v1[2]); xsc[3] = __uint_as_float(v1[3]);
    }
  }
  // row i0 + s lives in slot s and every slot flips its phase once per iteration
  for (int i0 = 0, ph = 0; i0 < n_rows; i0 += S, ph ^= 1) {
    const int nb = min(S, n_rows - i0);
    float acc[S] = {0.f, 0.f};
#pragma unroll
    for (int s = 0; s < S; s++) {
      if (s < nb) {
        uint32_t w[NW];
        const uint32_t sl = (uint32_t)s;                         // this row's ring slot
        if (i0 == 0 && s == 0) trace_stamp(p.trace, 5);          // lane setup done, about to wait for the weights
        mbar_wait_u32(full_u32 + 8u * s, (uint32_t)ph);
        if (i0 == 0 && s == 0) trace_stamp(p.trace, 6);          // first weights landed
        // SERVER rows step by RG: Q2_K rows (84 B per super-block) then differ in their 16-byte phase; so do
        // Q3_K / Q6_K rows (110 / 210 B per super-block) whenever RG * kb is odd
        const uint32_t a15r = (!SERVER && !kKX) ? a15 : (uint32_t)(row_ptr(row0 + (i0 + s) * row_step) & 15);
        if constexpr (kKX) {
          (void)w;
          acc[s] = dot_unit_kx<WT>(ring_addr + sl * p.slot_bytes + a15r + (uint32_t)(lane_c >> 1) * SBB,
                                   lane_c & 1, xaddr, jstride, xd[0], xbs);
        } else {
          load_unit<WT, NW>(w, ring_addr + sl * p.slot_bytes + a15r, lane_c);
          // The slot is free as soon as every lane holds its words in registers: refill it BEFORE the dot
          // product, not after (a slot is then busy for one DRAM round trip instead of round trip + dot).
          // The vote consumes the LAST word loaded (shared-memory loads of a warp return in order), so all
          // lanes' reads have completed before lane 0 hands the slot back to the copy engine.
          // Not for Q5_0 / Q5_1: their 22/24 live weight words + the refill's address math spill at 64
          // registers and the kernel gets 10 % slower (measured), so they refill after the dot.
          if constexpr (kEarlyRefill) {
            unsigned int seen;
            asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.u32 p, %1, 0;\n\tvote.sync.any.pred q, p, 0xffffffff;\n\t"
                         "selp.u32 %0, 1, 0, q;\n\t}" : "=r"(seen) : "r"(w[NW - 1]) : "memory");
            (void)seen;
            if (lane == 0 && i0 + RD + s < n_rows) issue(i0 + RD + s, (int)sl, pol);
          }
          acc[s] = dot_unit_sx<WT, NW>(w, xaddr, jstride, xd, xsc, xbs);
        }
        if constexpr (kKX || !kEarlyRefill) {
          __syncwarp();
          if (lane == 0 && i0 + RD + s < n_rows) issue(i0 + RD + s, (int)sl, pol);
        }
        if (i0 == 0 && s == 0) trace_stamp(p.trace, 7);          // first row's dot done
      }
    }
    const float tot = warp_sum2(acc[0], acc[1], lane);
    if (p.KS == 1) {
      if (lane < nb) {
        if constexpr (SERVER) cs[row0 + (i0 + lane) * row_step - chunk_lo] = tot;
        else store_c(p, row0 + (i0 + lane) * row_step, tot, tag);
      }
    } else {
      float *pb = part + (size_t)((rg * 2 + ph) * p.KS) * 2;
      if (lane < 2) pb[seg * 2 + lane] = tot;
      asm volatile("bar.sync %0, %1;" ::"r"(1 + rg), "r"(p.KS * 32) : "memory");
      if (seg == 0 && lane < nb) {
        float t = 0.f;
        for (int k = 0; k < p.KS; k++) t += pb[k * 2 + lane];
        if constexpr (SERVER) cs[row0 + (i0 + lane) * row_step - chunk_lo] = t;
        else store_c(p, row0 + (i0 + lane) * row_step, t, tag);
      }
    }
  }
  if constexpr (SERVER) {                    // every copy has been consumed: free the barriers for the next job
    __syncwarp();
    if (lane == 0)
      for (int s = 0; s < RD; s++) asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(smem_u32(&full[s])) : "memory");
  }
  }   // active
  trace_stamp(p.trace, 4);
}

template <int WT, int XM>
__global__ void __launch_bounds__(kD16Warps * 32, 2) decode16_kernel(const DecodeParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  decode16_body<WT, XM, false>(p, smem, (int)blockIdx.x, (int)gridDim.x, nullptr, 0u, nullptr);
}

// ---- the resident decode server --------------------------------------------------------------
// A synchronous host call (ggml's lamm_mul_mat on host tensors) pays ~10 us of launch + completion
// latency per kernel on top of ~5 us of work.  The server is ONE long-lived launch, a CTA per SM,
// that takes jobs from a mailbox in pinned host memory instead:
//   host:  memcpy X into the pinned row, write 16 {word, tag} pairs (the job: weight pointer +
//          launch plan) with tag = job number; then poll the result pairs while copying them out.
//   CTA 0: warp 0 polls the host mailbox over PCIe and republishes the pairs in a device mailbox
//          that the other CTAs poll in L2 (self-validating pairs: no flag, no fence).  It alone
//          ACCEPTS jobs, so an idle exit is one decision for the whole grid.  It then pulls X over
//          PCIe (one round trip, all 512 threads) and republishes it as {value, tag} pairs that
//          every CTA's stage_x_row validates element by element.
//   all:   run the decode body on a contiguous chunk of rows (weights prefetched while X crosses
//          PCIe) and write the chunk to pinned memory as {value, tag} pairs with coalesced 16-byte
//          stores (scattered 4-byte PCIe writes cost ~4 ns EACH).  Completion is carried by the
//          data: no arrival counter, no system fence (measured at 2.5 - 5 us per job).
// It gives up after idle_ns without a job (done_host[1] = the tag it was waiting for) so that other
// work can have the SMs; the host relaunches it on demand.  Every poll is bounded.

template <int WT>
__global__ void __launch_bounds__(kD16Warps * 32, 1) decode_server_kernel(const ServerParams sp) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ DecodeParams jp;
  __shared__ unsigned int jw[kSrvWords];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int bid = blockIdx.x;
  unsigned long long t_launch;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_launch));
  // CTA 0: pre_tag[slot] = tag of a job it has ALREADY republished into that device slot while accepting an earlier one
  __shared__ unsigned int pre_tag[kSrvSlots];
  // Every CTA: the job AFTER the one it is accepting, when that was already in the device mailbox (read by the idle
  // half of the warp in the same load): the next iteration starts from this stash without another L2 round trip.
  __shared__ unsigned int jn[kSrvWords];
  __shared__ unsigned int jn_tag;
  if (tid == 0) jn_tag = 0u;
  if (tid < kSrvSlots) pre_tag[tid] = 0u;
  __syncthreads();
  for (unsigned int s = sp.seq0;; ++s) {
    // ---- 1. the job
    const int slot = (int)(s % (unsigned int)kSrvSlots);
    if (warp == 0) {
      const bool pre = bid == 0 && pre_tag[slot] == s;     // accepted with an earlier job: in the device mailbox already
      const bool stashed = jn_tag == s;                   // (tags are never 0: lammcu_api.cu restarts them before the wrap)
      unsigned long long v = 0, t0, now;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
      __syncwarp();
      if (stashed) {
        if (lane < kSrvWords) jw[lane] = jn[lane];
        if (lane == 0) jn_tag = 0u;
      } else if (bid == 0 && !pre) {
        // The host mailbox, all kSrvSlots x 16 pairs per poll (one PCIe round trip either way): lane l holds pair l of
        // slots 0/1 (v0) and of slots 2/3 (v1).  Bounded lifetime: a caller that never pauses must not starve kernels
        // queued on other streams -- but a job that was republished is never given up (the other CTAs run it).
        bool give_up = s != sp.seq0 && t0 - t_launch > (unsigned long long)sp.life_ns;
        unsigned long long v0 = 0, v1 = 0;
        auto slot_full = [&](unsigned int tag_, int sl) {
          const unsigned int m0 = __ballot_sync(0xffffffffu, (unsigned int)(v0 >> 32) == tag_);
          const unsigned int m1 = __ballot_sync(0xffffffffu, (unsigned int)(v1 >> 32) == tag_);
          return (((sl < 2 ? m0 : m1) >> ((sl & 1) * 16)) & 0xffffu) == 0xffffu;
        };
        while (!give_up) {
          asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v0) : "l"(sp.mbox_host + lane) : "memory");
          asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v1) : "l"(sp.mbox_host + 32 + lane) : "memory");
          if (slot_full(s, slot)) break;
          asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
          give_up = __any_sync(0xffffffffu, now - t0 > (unsigned long long)sp.idle_ns);   // one decision for the warp
        }
        const int sl0 = lane >> 4, sl1 = 2 + (lane >> 4);      // slots of this lane's two pairs
        if (give_up) {
          v = ((unsigned long long)s << 32) | ((lane & 15) == 0 ? 2u : 0u);
          if (sl0 == slot || sl1 == slot) {
            jw[lane & 15] = (unsigned int)v;
            asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(sp.mbox_dev + slot * kSrvWords + (lane & 15)), "l"(v) : "memory");
          }
          if (lane == 0) {
            sp.done_host[1] = s;
            __threadfence_system();
          }
        } else {
          // this job, and the next two if the same read found them complete (their slots were free: the host posts a
          // tag only after the job kSrvSlots before it has been answered)
          unsigned int pub = 1u << slot;
#pragma unroll
          for (int d = 1; d <= 2; d++) {
            const int sl = (slot + d) % kSrvSlots;
            if (slot_full(s + (unsigned int)d, sl)) {
              pub |= 1u << sl;
              if (lane == 0) pre_tag[sl] = s + (unsigned int)d;
            }
          }
          // the half that holds the NEXT slot goes first, so that a CTA that sees this job usually sees its follower too
          const bool v1_first = slot == 1;                  // slots 1 -> 2 live in v0 -> v1; 3 -> 0 in v1 -> v0 (v0 first)
          if (v1_first && ((pub >> sl1) & 1u)) asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(sp.mbox_dev + 32 + lane), "l"(v1) : "memory");
          if ((pub >> sl0) & 1u) asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(sp.mbox_dev + lane), "l"(v0) : "memory");
          if (!v1_first && ((pub >> sl1) & 1u)) asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(sp.mbox_dev + 32 + lane), "l"(v1) : "memory");
          if (sl0 == slot) jw[lane & 15] = (unsigned int)v0;
          if (sl1 == slot) jw[lane & 15] = (unsigned int)v1;
          const int sln = (slot + 1) % kSrvSlots;
          if ((pub >> sln) & 1u) {
            if (sl0 == sln) jn[lane & 15] = (unsigned int)v0;
            if (sl1 == sln) jn[lane & 15] = (unsigned int)v1;
            if (lane == 0) jn_tag = s + 1u;
          }
        }
      } else {
        // every other CTA (and CTA 0 for a job it republished earlier): the device slot, in L2
        // lanes 0..15: this job's slot; lanes 16..31: the next slot, in the same load
        const unsigned long long *src = sp.mbox_dev + (lane < kSrvWords ? slot : (slot + 1) % kSrvSlots) * kSrvWords + (lane & 15);
        bool give_up = false;
        while (true) {
          asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(src) : "memory");
          if (__all_sync(0xffffffffu, lane >= kSrvWords || (unsigned int)(v >> 32) == s)) break;
          asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
          __nanosleep(40);
          if (__any_sync(0xffffffffu, now - t0 > 4000000000ull)) {   // one decision for the warp; CTA 0 always publishes something
            give_up = true;
            *sp.status = 2u;
            break;
          }
        }
        const bool nx = !give_up && __all_sync(0xffffffffu, lane < kSrvWords || (unsigned int)(v >> 32) == s + 1u);
        if (give_up) v = ((unsigned long long)s << 32) | (lane == 0 ? 2u : 0u);
        if (lane < kSrvWords) jw[lane] = (unsigned int)v;
        else if (nx) jn[lane - kSrvWords] = (unsigned int)v;
        if (nx && lane == 0) jn_tag = s + 1u;
      }
    }
    __syncthreads();
    if (bid == 0) srv_stamp(&sp, s, 0);            // job seen
    if (jw[0] != 1u) return;                       // 2: stop request or idle time-out
    // ---- 2. CTA 0: X, pinned host memory -> {value, tag} pairs in HBM (one PCIe round trip)
    if (bid == 0 && jw[15] == 0u) {                 // jw[15] != 0: same row as the previous job, still staged
      const int K4 = (int)jw[5] * (type_info(WT).blck / 4);
      const unsigned long long th = (unsigned long long)s << 32;
      for (int base = 0; base < K4; base += 4 * (int)blockDim.x) {
        uint4 xr[4];
#pragma unroll
        for (int r = 0; r < 4; r++) {
          const int i = base + r * (int)blockDim.x + tid;
          if (i < K4)
            asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                         : "=r"(xr[r].x), "=r"(xr[r].y), "=r"(xr[r].z), "=r"(xr[r].w)
                         : "l"(sp.x_host + 4 * (size_t)i) : "memory");
        }
#pragma unroll
        for (int r = 0; r < 4; r++) {
          const int i = base + r * (int)blockDim.x + tid;
          if (i < K4) {
            unsigned long long *dst = sp.x_ll + 4 * (size_t)i;
            asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(dst), "l"(th | xr[r].x), "l"(th | xr[r].y) : "memory");
            asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(dst + 2), "l"(th | xr[r].z), "l"(th | xr[r].w) : "memory");
          }
        }
      }
    }
    if (bid == 0) srv_stamp(&sp, s, 1);            // X forwarded (or known unchanged)
    if (tid == 0) {
      DecodeParams q = {};
      q.w = reinterpret_cast<const uint8_t *>(((unsigned long long)jw[2] << 32) | jw[1]);
      q.w_stride = (int64_t)jw[3];
      q.M = (int)jw[4];  q.kb = (int)jw[5];  q.units = (int)jw[6];  q.KS = (int)jw[7];  q.seg_units = (int)jw[8];
      q.RG = (int)jw[9];  q.slots = kD16Slots;  q.slot_bytes = (int)jw[10];  q.x_bytes = (int)jw[11];
      q.x_ll = sp.x_ll;  q.status = sp.status;
      q.x_back = (int)jw[15];                      // server jobs: "activations unchanged" (see decode16_body)
      jp = q;
    }
    __syncthreads();
    // ---- 3. the rows of this CTA's chunk
    const int R = (int)jw[12];                     // rows per CTA (multiple of 4)
    float *cs = reinterpret_cast<float *>(smem + jw[13]);
    // jw[14] != 0: CTA 0 only accepts jobs and forwards X (it is the last to start on its rows, so its chunk
    // is the call's critical path); the row chunks are dealt to CTAs 1 .. grid-1 (virtual id vb)
    const int vb = jw[14] ? (bid == 0 ? (int)gridDim.x - 1 : bid - 1) : bid;
    if (!(jw[14] && bid == 0)) decode16_body<WT, 2, true>(jp, smem, vb, R, &sp, s, cs);
    __syncthreads();
    if (vb == 0) srv_stamp(&sp, s, 3);             // first chunk's rows done
    // ---- 4. results: the chunk as {value, tag} pairs, coalesced, straight to pinned memory
    {
      const int lo = vb * R, n = (jw[14] && bid == 0) ? 0 : max(0, min(jp.M, lo + R) - lo);
      const unsigned long long th = (unsigned long long)s << 32;
      unsigned long long *dst = sp.c_ll_host + (size_t)slot * sp.c_slot_pairs + lo;
      for (int i = 2 * tid; i < n; i += 2 * (int)blockDim.x) {
        const unsigned long long a = th | __float_as_uint(cs[i]);
        if (i + 1 < n) {
          const unsigned long long b = th | __float_as_uint(cs[i + 1]);
          asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(dst + i), "l"(a), "l"(b) : "memory");
        } else {
          asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(dst + i), "l"(a) : "memory");
        }
      }
    }
    if (vb == 0) srv_stamp(&sp, s, 4);             // first chunk written
    __syncthreads();
  }
}

// ---- the persistent decode chain --------------------------------------------------------------
// A token's matmuls are a CHAIN (q,k,v = W x; o = Wo v; g,u = W o; x' = Wd u; ...): each launch of decode16_kernel
// starts with a serial part -- wait for the previous grid, read + quantize the activation row, set the lanes up --
// during which its weight ring is full and HBM idles: ~2 us x 129 launches of a 0.9 ms token (DESIGN d1).  Here ONE
// launch runs the whole list.  A warp's TMA ring does not know about matmul boundaries: lane 0 walks the warp's rows
// of stage s, then of stage s + 1, ... and keeps `slots` copies in flight (weights are immutable, so a later stage's
// rows may arrive long before its activations exist); with one CTA per SM the ring is ~200 KB per SM, ~30 MB per
// GPU = 4.6 us of streaming, several times the serial part.  Stages are separated by a grid barrier on a counter
// (zeroed by a memset node in front of the launch): after its last row a CTA does bar.sync, fence, atomicAdd; before
// reading the next activation row it polls the counter (bounded: 2 s, then status = 1).  All CTAs are co-resident
// (grid = SM count, cooperative launch), so the barrier cannot deadlock.  Arithmetic per row is decode16_body's.
//
// MEASURED (round 2, LLaMA-7B Q4_1 chain, profiles/r2n .. r2q, tools/trace_chain.py): bit-identical to one launch per
// op, but 920 tok/s against 1057 for 129 PDL-chained decode16 launches.  The stamps: per stage ~0.8 us from "rows done"
// to the next stage's start (fence + atomic), ~1.0 us barrier wait, ~1.0 us activation quantize -- a 2.8 us serial
// part against 2.15 us for griddepcontrol.wait + quantize -- and the rows phase consumes ~0.9 us per row per warp =
// 16 rows/us/SM, which is just the DRAM rate: with 16 warps per SM the dot phase is latency-bound, the ring stays
// full and nothing is prefetched during the serial part (two resident CTAs give the launch-per-op path 32 warps).
// It stays as an entry point for callers that want one launch per token; the bench's default is the PDL chain.
struct ChainStage {
  const uint8_t *wb[4];   float *cb[4];   int mc[3];   int nbatch;   // as DecodeParams (virtual bases)
  const float *x;
  int64_t w_stride;
  int M, kb, units, KS, seg_units, RG, x_bytes, inv_KS;     // inv_KS: (warp * inv_KS) >> 16 == warp / KS for warp < 32
};
struct ChainParams {
  const ChainStage *stages;
  int n_stages;
  unsigned int *done_ctr, *status;
  int slots, slot_bytes, x_bytes_max;
  unsigned long long *trace;   // diagnostics (lammcu_set_trace): CTA 0 stamps [stage][4] = barrier wait begins, barrier
                               // passed, activations quantized, rows done (globaltimer ns)
};
constexpr int kChainCtrl = 2048;          // [warps][slots] mbarriers (1 KB), partial sums at +1024, slot phases at +1536
constexpr int kChainWarps = 16;           // measured: 16 warps (105-128 registers, no spills) 920 tok/s on the 7B Q4_1 chain;
                                          // 32 warps at 64 registers (20-200 bytes of spills) 681 (profiles/r2q/)
constexpr int kChainMaxSlots = 8;

template <int WT>
__global__ void __launch_bounds__(kChainWarps * 32, 1) decode_chain_kernel(const ChainParams cp) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ DecodeParams spb[2];                // the current stage's parameters; the next stage's are built ahead
  constexpr bool kKX = is_kquant_ext(WT);
  constexpr bool kEarlyRefill = !(WT == LAMMCU_TYPE_Q5_0 || WT == LAMMCU_TYPE_Q5_1);
  constexpr bool kQ2K = (WT == LAMMCU_TYPE_Q2_K) || kKX;
  constexpr uint32_t SBB = (uint32_t)type_info(WT).size;
  constexpr int VT = type_info(WT).vec_dot;
  constexpr int NW = kKX ? 4 : (kQ2K ? 11 : DecFmt<WT>::NW);
  constexpr int unit_bytes = kQ2K ? 42 : 4 * type_info(WT).size;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int bid = (int)blockIdx.x, nblk = (int)gridDim.x;
  const int D = cp.slots;
  uint64_t *full = reinterpret_cast<uint64_t *>(smem) + warp * kChainMaxSlots;
  float *part = reinterpret_cast<float *>(smem + 1024);
  // 16-byte phase of the row in each ring slot, written by lane 0 when it issues the copy (ordered by the mbarrier)
  const uint32_t a15_u32 = smem_u32(smem + 1536) + (uint32_t)warp * (kChainMaxSlots * 4);
  uint8_t *xs = smem + kChainCtrl;
  uint8_t *ring = xs + cp.x_bytes_max + (size_t)warp * D * cp.slot_bytes;
  const uint32_t ring_addr = smem_u32(ring), full_u32 = smem_u32(full);

  // ---- the warp's rows of one stage: which K segment, which rows
  struct Geo { int n_rows, row0, row_step; uint32_t seg_off, seg_bytes; int u0, seg_n, rg, seg; };
  auto geometry = [&](const ChainStage &S) -> Geo {
    Geo g;
    const bool active = warp < S.RG * S.KS;
    g.rg = (warp * S.inv_KS) >> 16;  g.seg = warp - g.rg * S.KS;
    g.u0 = g.seg * S.seg_units;
    g.seg_n = max(0, min(S.seg_units, S.units - g.u0));
    g.seg_off = kQ2K ? (uint32_t)(g.u0 >> 1) * SBB : (uint32_t)g.u0 * unit_bytes;
    g.seg_bytes = kQ2K ? (uint32_t)((g.seg_n + 1) >> 1) * SBB : (uint32_t)g.seg_n * unit_bytes;
    g.row0 = bid + nblk * g.rg;
    g.row_step = nblk * S.RG;
    g.n_rows = (active && g.row0 < S.M) ? (S.M - 1 - g.row0) / g.row_step + 1 : 0;
    return g;
  };
  // the matrix of a sibling batch that (virtual) row r belongs to: its base + the first row of the NEXT matrix
  auto select_base = [&](const ChainStage &S, uint32_t seg_off, int r, const uint8_t *&base, int &lim) {
    base = S.wb[0];
    lim = 0x7fffffff;
    if (S.nbatch > 1) {
      lim = S.mc[0];
      if (r >= S.mc[0]) { base = S.wb[1]; lim = S.mc[1]; }
      if (r >= S.mc[1]) { base = S.wb[2]; lim = S.mc[2]; }
      if (r >= S.mc[2]) { base = S.wb[3]; lim = 0x7fffffff; }
    }
    base += seg_off;
  };

  // ---- the copy cursor (lane 0): runs up to D rows ahead of the consumer, across stage boundaries.  Everything a
  // copy needs sits in registers (the stage table is read when the cursor ENTERS a stage or a new matrix of a batch):
  // the refill is on the consuming warp's critical path, once per row.
  int is = -1, ii = 0, in_rows = 0, i_row = 0, i_step = 0, i_lim = 0;
  uint32_t i_off = 0, i_bytes = 0, islot = 0;
  const uint8_t *i_base = nullptr;
  int64_t i_stride = 0;
  uint64_t pol = 0;
  auto next_stage = [&]() {                       // first stage after `is` in which this warp owns rows
    for (++is; is < cp.n_stages; ++is) {
      const ChainStage &S = cp.stages[is];
      const Geo g = geometry(S);
      if (g.n_rows > 0) {
        ii = 0; in_rows = g.n_rows; i_row = g.row0; i_step = g.row_step; i_off = g.seg_off; i_bytes = g.seg_bytes;
        i_stride = S.w_stride;
        select_base(S, i_off, i_row, i_base, i_lim);
        return;
      }
    }
  };
  auto issue_one = [&]() {                        // lane 0, is < n_stages
    if (i_row >= i_lim) select_base(cp.stages[is], i_off, i_row, i_base, i_lim);
    const uintptr_t a = (uintptr_t)(i_base + (int64_t)i_row * i_stride);
    const uintptr_t a0 = a & ~(uintptr_t)15;
    const uint32_t bytes = (uint32_t)(((a + i_bytes + 15) & ~(uintptr_t)15) - a0);
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(a15_u32 + 4u * islot), "r"((uint32_t)(a & 15)) : "memory");
    uint64_t *bar = full + islot;
    mbar_arrive_expect_tx(bar, bytes);
    bulk_g2s_stream(ring + (size_t)islot * cp.slot_bytes, reinterpret_cast<const void *>(a0), bytes, bar, pol);
    islot = (islot + 1u == (uint32_t)D) ? 0u : islot + 1u;
    i_row += i_step;
    if (++ii >= in_rows) next_stage();
  };
  if (lane == 0) {
    for (int s_ = 0; s_ < D; s_++) mbar_init(&full[s_], 1);
    fence_barrier_init();
    pol = policy_evict_first();
    next_stage();
    for (int s_ = 0; s_ < D && is < cp.n_stages; s_++) issue_one();
  }
  uint32_t cslot = 0, cph = 0;                    // consumer: slot of the next row, parity of the current lap

  auto build_params = [&](int st_) {               // one thread: stage st_'s parameters as the decode helpers want them
    const ChainStage &S = cp.stages[st_];
    DecodeParams q = {};
    q.x = S.x;  q.c = S.cb[0];  q.M = S.M;  q.kb = S.kb;  q.units = S.units;  q.KS = S.KS;  q.seg_units = S.seg_units;
    q.RG = S.RG;  q.x_bytes = S.x_bytes;  q.w_stride = S.w_stride;  q.nbatch = S.nbatch;
    for (int j = 0; j < 4; j++) { q.wb[j] = S.wb[j];  q.cb[j] = S.cb[j]; }
    for (int j = 0; j < 3; j++) q.mc[j] = S.mc[j];
    spb[st_ & 1] = q;
  };
  if (tid == 0) build_params(0);

  for (int st = 0; st < cp.n_stages; st++) {
    const ChainStage &S = cp.stages[st];
    const DecodeParams &sp = spb[st & 1];
    if (tid == 0) {
      if (cp.trace && bid == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(cp.trace[st * 4 + 0]));
      if (st > 0) {                               // grid barrier: every CTA has stored its rows of stage st - 1
        const unsigned int target = (unsigned int)nblk * (unsigned int)st;
        unsigned int v, spins = 0;
        unsigned long long t0 = 0;
        while (true) {
          asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(cp.done_ctr) : "memory");
          if (v >= target) break;
          if ((++spins & 255u) == 0) {
            unsigned long long now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            if (t0 == 0) t0 = now;
            else if (now - t0 > 2000000000ull) { *cp.status = 1u; break; }
          }
        }
        __threadfence();
      }
    }
    __syncthreads();
    if (cp.trace && bid == 0 && tid == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(cp.trace[st * 4 + 1]));
    stage_x_row<VT, 3>(sp, xs);
    __syncthreads();
    if (cp.trace && bid == 0 && tid == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(cp.trace[st * 4 + 2]));
    // the next stage's parameters (its buffer was last read in stage st - 1: two barriers ago); the last warp has the fewest rows
    if (tid == kChainWarps * 32 - 1 && st + 1 < cp.n_stages) build_params(st + 1);
    const Geo g = geometry(S);
    if (g.n_rows > 0) {
      const int U = S.units, ug = g.u0 + lane;
      const bool lane_live = lane < g.seg_n;
      const int lane_c = lane_live ? lane : 0;
      const uint32_t xaddr = smem_u32(xs) + (uint32_t)(lane_live ? ug : g.u0) * 16u;
      const uint32_t jstride = (uint32_t)U * 16u;
      const uint32_t xs_addr = smem_u32(xs) + 8u * jstride + (kQ2K ? (uint32_t)ug * 16u : (uint32_t)ug * 32u);
      const uint32_t xd_addr = smem_u32(xs) + 9u * jstride + (uint32_t)(ug >> 1) * 4u;          // Q8_K only
      float xd[4] = {0.f, 0.f, 0.f, 0.f}, xsc[4] = {0.f, 0.f, 0.f, 0.f};
      uint32_t xbs[4] = {0u, 0u, 0u, 0u};
      if (lane_live) {
        if constexpr (kQ2K) {
          lds128(xbs, xs_addr);
          asm volatile("ld.shared.f32 %0, [%1];" : "=f"(xd[0]) : "r"(xd_addr));
        } else {
          uint32_t v0[4], v1[4];
          lds128(v0, xs_addr);
          lds128(v1, xs_addr + 16u);
          xd[0] = __uint_as_float(v0[0]); xsc[0] = __uint_as_float(v0[1]); xd[1] = __uint_as_float(v0[2]); xsc[1] = __uint_as_float(v0[3]);
          xd[2] = __uint_as_float(v1[0]); xsc[2] = __uint_as_float(v1[1]); xd[3] = __uint_as_float(v1[2]); xsc[3] = __uint_as_float(v1[3]);
        }
      }
      for (int i0 = 0, ph = 0; i0 < g.n_rows; i0 += 2, ph ^= 1) {
        const int nb = min(2, g.n_rows - i0);
        float acc[2] = {0.f, 0.f};
#pragma unroll
        for (int s_ = 0; s_ < 2; s_++) {
          if (s_ < nb) {
            uint32_t w[NW];
            const uint32_t sl = cslot;
            mbar_wait_u32(full_u32 + 8u * sl, cph);
            cslot = (sl + 1u == (uint32_t)D) ? 0u : sl + 1u;
            if (cslot == 0u) cph ^= 1u;
            uint32_t a15r;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(a15r) : "r"(a15_u32 + 4u * sl) : "memory");
            if constexpr (kKX) {
              (void)w;
              acc[s_] = dot_unit_kx<WT>(ring_addr + sl * cp.slot_bytes + a15r + (uint32_t)(lane_c >> 1) * SBB, lane_c & 1, xaddr,
                                        jstride, xd[0], xbs);
            } else {
              load_unit<WT, NW>(w, ring_addr + sl * cp.slot_bytes + a15r, lane_c);
              if constexpr (kEarlyRefill) {       // the slot is free once every lane holds its words (see decode16_body)
                unsigned int seen;
                asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.u32 p, %1, 0;\n\tvote.sync.any.pred q, p, 0xffffffff;\n\t"
                             "selp.u32 %0, 1, 0, q;\n\t}" : "=r"(seen) : "r"(w[NW - 1]) : "memory");
                (void)seen;
                if (lane == 0 && is < cp.n_stages) issue_one();
              }
              acc[s_] = dot_unit_sx<WT, NW>(w, xaddr, jstride, xd, xsc, xbs);
            }
            if constexpr (kKX || !kEarlyRefill) {
              __syncwarp();
              if (lane == 0 && is < cp.n_stages) issue_one();
            }
          }
        }
        const float tot = warp_sum2(acc[0], acc[1], lane);
        if (S.KS == 1) {
          if (lane < nb) store_c(sp, g.row0 + (i0 + lane) * g.row_step, tot, 0u);
        } else {
          float *pb = part + (size_t)((g.rg * 2 + ph) * S.KS) * 2;
          if (lane < 2) pb[g.seg * 2 + lane] = tot;
          asm volatile("bar.sync %0, %1;" ::"r"(1 + g.rg), "r"(S.KS * 32) : "memory");
          if (g.seg == 0 && lane < nb) {
            float t = 0.f;
            for (int k = 0; k < S.KS; k++) t += pb[k * 2 + lane];
            store_c(sp, g.row0 + (i0 + lane) * g.row_step, t, 0u);
          }
        }
      }
    }
    __syncthreads();                                // all rows of the stage stored; xs and sp may be overwritten
    if (cp.trace && bid == 0 && tid == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(cp.trace[st * 4 + 3]));
    if (tid == 0 && st + 1 < cp.n_stages) {
      __threadfence();
      atomicAdd(cp.done_ctr, 1u);
    }
  }
}

// ---- host side --------------------------------------------------------------------------
constexpr int kDecSmemMax = 227 * 1024;

template <int WT, bool XF32, int MAXT>
static int launch_decode_tt(const DecodeParams &p, int grid, int threads, int smem, cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(decode_kernel<WT, XF32, MAXT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kDecSmemMax);
    if (e != cudaSuccess) {
      set_error("decode: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return LAMMCU_ERR_CUDA;
    }
    attr_set = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(threads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, decode_kernel<WT, XF32, MAXT>, p);
  g_launch_count++;
  if (e != cudaSuccess) {
    set_error("decode launch (type %d, smem %d): %s", WT, smem, cudaGetErrorString(e));
    return LAMMCU_ERR_CUDA;
  }
  return LAMMCU_OK;
}

template <int WT, bool XF32>
static int launch_decode_t(const DecodeParams &p, int grid, int threads, int smem, cudaStream_t st) {
  // 8 warps: 128 registers per thread; 9-10 warps (K-segment counts 3, 5): 102
  if (threads <= kDecWarps * 32) return launch_decode_tt<WT, XF32, kDecWarps * 32>(p, grid, threads, smem, st);
  return launch_decode_tt<WT, XF32, kDecMaxThreads>(p, grid, threads, smem, st);
}

template <int WT, int XF32>
static int launch_decode16_t(const DecodeParams &p, int grid, int threads, int smem, cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(decode16_kernel<WT, XF32>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kDecSmemMax);
    if (e != cudaSuccess) {
      set_error("decode16: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return LAMMCU_ERR_CUDA;
    }
    attr_set = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(threads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, decode16_kernel<WT, XF32>, p);
  g_launch_count++;
  if (e != cudaSuccess) {
    set_error("decode16 launch (type %d, smem %d): %s", WT, smem, cudaGetErrorString(e));
    return LAMMCU_ERR_CUDA;
  }
  return LAMMCU_OK;
}

template <int XF32>
static int launch_decode16_x(int w_type, const DecodeParams &p, int grid, int threads, int smem, cudaStream_t st) {
  switch (w_type) {
  case LAMMCU_TYPE_Q4_0: return launch_decode16_t<LAMMCU_TYPE_Q4_0, XF32>(p, grid, threads, smem, st);
  case LAMMCU_TYPE_Q4_1: return launch_decode16_t<LAMMCU_TYPE_Q4_1, XF32>(p, grid, threads, smem, st);
  case LAMMCU_TYPE_Q5_0: return launch_decode16_t<LAMMCU_TYPE_Q5_0, XF32>(p, grid, threads, smem, st);
  case LAMMCU_TYPE_Q5_1: return launch_decode16_t<LAMMCU_TYPE_Q5_1, XF32>(p, grid, threads, smem, st);
  case LAMMCU_TYPE_Q8_0: return launch_decode16_t<LAMMCU_TYPE_Q8_0, XF32>(p, grid, threads, smem, st);
  case LAMMCU_TYPE_Q3_K: return launch_decode16_t<LAMMCU_TYPE_Q3_K, XF32>(p, grid, threads, smem, st);
  case LAMMCU_TYPE_Q4_K: return launch_decode16_t<LAMMCU_TYPE_Q4_K, XF32>(p, grid, threads, smem, st);
  case LAMMCU_TYPE_Q5_K: return launch_decode16_t<LAMMCU_TYPE_Q5_K, XF32>(p, grid, threads, smem, st);
  case LAMMCU_TYPE_Q6_K: return launch_decode16_t<LAMMCU_TYPE_Q6_K, XF32>(p, grid, threads, smem, st);
  default:               return launch_decode16_t<LAMMCU_TYPE_Q2_K, XF32>(p, grid, threads, smem, st);
  }
}

template <bool XF32>
static int launch_decode_x(int w_type, const DecodeParams &p, int grid, int threads, int smem, cudaStream_t st) {
  switch (w_type) {
  case LAMMCU_TYPE_Q4_0: return launch_decode_t<LAMMCU_TYPE_Q4_0, XF32>(p, grid, threads, smem, st);
  case LAMMCU_TYPE_Q4_1: return launch_decode_t<LAMMCU_TYPE_Q4_1, XF32>(p, grid, threads, smem, st);
  case LAMMCU_TYPE_Q5_0: return launch_decode_t<LAMMCU_TYPE_Q5_0, XF32>(p, grid, threads, smem, st);
  case LAMMCU_TYPE_Q5_1: return launch_decode_t<LAMMCU_TYPE_Q5_1, XF32>(p, grid, threads, smem, st);
  case LAMMCU_TYPE_Q8_0: return launch_decode_t<LAMMCU_TYPE_Q8_0, XF32>(p, grid, threads, smem, st);
  default:               return launch_decode_t<LAMMCU_TYPE_Q2_K, XF32>(p, grid, threads, smem, st);
  }
}

// Whether the single-launch decode kernel can take this operand set (else the caller uses
// the generic GEMV of lamm_gemv.cu, which handles any 2/4-byte alignment and any K).
bool decode_supported(int w_type, const void *w, int64_t w_stride, const void *x_or_xq, int64_t M, int64_t K) {
  if (w_type == LAMMCU_TYPE_F32 || w_type == LAMMCU_TYPE_F16 || !lammcu_type_supported(w_type)) return false;
  if (M <= 0 || M > 0x3fffffff || K <= 0 || K > (1 << 22)) return false;
  if (w_type == LAMMCU_TYPE_Q3_K || w_type == LAMMCU_TYPE_Q6_K) {      // 110 / 210-byte super-blocks: 16-bit words
    if (K % 256) return false;
    return (((uintptr_t)w | (uintptr_t)w_stride) & 1) == 0 && ((uintptr_t)x_or_xq & 15) == 0;
  }
  if (w_type == LAMMCU_TYPE_Q2_K) {
    if (K % 256) return false;
    return (((uintptr_t)w | (uintptr_t)w_stride) & 3) == 0 && ((uintptr_t)x_or_xq & 15) == 0;
  }
  if (K % 256) return false;                       // 8 blocks: every lane chunk and segment is 16-B aligned
  return (((uintptr_t)w | (uintptr_t)w_stride) & 15) == 0 && ((uintptr_t)x_or_xq & 15) == 0;
}

static int plan_k_split(DecodeParams &p, const TypeInfo &ti, bool q2k, int64_t K) {
  p.kb = (int)(K / ti.blck);
  p.units = q2k ? 2 * p.kb : p.kb / 4;
  p.KS = (p.units + 31) / 32;
  if (p.KS > kDecMaxWarps) {
    set_error("decode: K=%lld needs more than %d K-segments", (long long)K, kDecMaxWarps);
    return LAMMCU_ERR_SHAPE;
  }
  const int gran = q2k ? 8 : 2;                    // keeps every segment start 16-byte aligned
  p.seg_units = (((p.units + p.KS - 1) / p.KS) + gran - 1) / gran * gran;
  if (p.seg_units > 32) p.seg_units = 32;
  p.KS = (p.units + p.seg_units - 1) / p.seg_units;
  return LAMMCU_OK;
}

// shared-memory shape of the 16-warp kernel for a planned K split; returns the dynamic bytes
static int plan_d16_smem(DecodeParams &p, const TypeInfo &ti, bool q2k, int *warps_out) {
  const int warps = (kD16Warps / p.KS) * p.KS;
  p.RG = warps / p.KS;
  const int seg_bytes = q2k ? (p.seg_units / 2) * ti.size : p.seg_units * 4 * ti.size;
  p.slot_bytes = ((seg_bytes + 15) & ~15) + 16;
  p.x_bytes = q2k ? (p.units * (128 + 16) + p.kb * 4) : (p.units * 128 + p.kb * 8);
  p.x_bytes = (p.x_bytes + 127) & ~127;
  p.slots = kD16Slots;
  *warps_out = warps;
  return kDecCtrl + p.x_bytes + warps * p.slots * p.slot_bytes;
}

constexpr int kSrvSmem = 227 * 1024 - 2048;        // dynamic bytes of the server (static: the job's DecodeParams)

// the job the host posts to the resident server for one N == 1 mul_mat (kSrvWords 32-bit words)
int decode_server_job(int w_type, const void *w, int64_t w_stride, int64_t M, int64_t K, unsigned int *words) {
  if (!decode_supported(w_type, w, w_stride, nullptr, M, K) || w_stride > 0xffffffffll) {
    set_error("decode server: operands not supported");
    return LAMMCU_ERR_SHAPE;
  }
  const TypeInfo ti = type_info(w_type);
  const bool q2k = (ti.blck == 256);               // super-block formats: unit = half a super-block
  DecodeParams p = {};
  int rc = plan_k_split(p, ti, q2k, K);
  if (rc) return rc;
  if (p.KS > kD16Warps) { set_error("decode server: K too large"); return LAMMCU_ERR_SHAPE; }
  int warps = 0;
  const int body_smem = (plan_d16_smem(p, ti, q2k, &warps) + 127) & ~127;
  // CTA 0 accepts jobs and forwards X but computes no rows (default; LAMM_SERVER_IO_CTA=0 gives it a chunk too)
  static const int io_cta = [] { const char *e = getenv("LAMM_SERVER_IO_CTA"); return e ? atoi(e) : 1; }();
  const int workers = sm_count_cached() - (io_cta && sm_count_cached() > 1 ? 1 : 0);
  const int64_t R = ((M + workers - 1) / workers + 3) & ~(int64_t)3;                         // rows per CTA
  if (K > 65536 || body_smem + R * 4 > kSrvSmem) { set_error("decode server: shape too large for shared memory"); return LAMMCU_ERR_SHAPE; }
  for (int i = 0; i < kSrvWords; i++) words[i] = 0u;
  words[0] = 1u;
  words[1] = (unsigned int)((uintptr_t)w & 0xffffffffu);
  words[2] = (unsigned int)((uintptr_t)w >> 32);
  words[3] = (unsigned int)w_stride;
  words[4] = (unsigned int)M;   words[5] = (unsigned int)p.kb;        words[6] = (unsigned int)p.units;
  words[7] = (unsigned int)p.KS; words[8] = (unsigned int)p.seg_units; words[9] = (unsigned int)p.RG;
  words[10] = (unsigned int)p.slot_bytes;  words[11] = (unsigned int)p.x_bytes;  words[12] = (unsigned int)R;
  words[13] = (unsigned int)body_smem;
  words[14] = (io_cta && sm_count_cached() > 1) ? 1u : 0u;
  return LAMMCU_OK;
}

template <int WT>
static int launch_server_t(const ServerParams &sp, cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(decode_server_kernel<WT>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSrvSmem);
    if (e != cudaSuccess) { set_error("decode server: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return LAMMCU_ERR_CUDA; }
    attr_set = true;
  }
  decode_server_kernel<WT><<<sm_count_cached(), kD16Warps * 32, kSrvSmem, st>>>(sp);
  g_launch_count++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("decode server launch: %s", cudaGetErrorString(e)); return LAMMCU_ERR_CUDA; }
  return LAMMCU_OK;
}

int launch_decode_server(int w_type, const ServerParams &sp, cudaStream_t st) {
  switch (w_type) {
  case LAMMCU_TYPE_Q4_0: return launch_server_t<LAMMCU_TYPE_Q4_0>(sp, st);
  case LAMMCU_TYPE_Q4_1: return launch_server_t<LAMMCU_TYPE_Q4_1>(sp, st);
  case LAMMCU_TYPE_Q5_0: return launch_server_t<LAMMCU_TYPE_Q5_0>(sp, st);
  case LAMMCU_TYPE_Q5_1: return launch_server_t<LAMMCU_TYPE_Q5_1>(sp, st);
  case LAMMCU_TYPE_Q8_0: return launch_server_t<LAMMCU_TYPE_Q8_0>(sp, st);
  case LAMMCU_TYPE_Q2_K: return launch_server_t<LAMMCU_TYPE_Q2_K>(sp, st);
  case LAMMCU_TYPE_Q3_K: return launch_server_t<LAMMCU_TYPE_Q3_K>(sp, st);
  case LAMMCU_TYPE_Q4_K: return launch_server_t<LAMMCU_TYPE_Q4_K>(sp, st);
  case LAMMCU_TYPE_Q5_K: return launch_server_t<LAMMCU_TYPE_Q5_K>(sp, st);
  case LAMMCU_TYPE_Q6_K: return launch_server_t<LAMMCU_TYPE_Q6_K>(sp, st);
  default: set_error("decode server: type %d", w_type); return LAMMCU_ERR_TYPE;
  }
}

int launch_decode(int w_type, const void *w, int64_t w_stride, const float *x, const void *xq, void *xq_out,
                  float *c, int64_t M, int64_t K, cudaStream_t st, const ShardInfo *sh, const BatchInfo *bt,
                  const XOp *xop) {
  const TypeInfo ti = type_info(w_type);
  const bool q2k = (ti.blck == 256);               // super-block formats: unit = half a super-block
  DecodeParams p = {};
  p.w = (const uint8_t *)w;  p.w_stride = w_stride;
  p.x = x;  p.xq = (const uint8_t *)xq;  p.xq_out = (uint8_t *)xq_out;
  p.c = c;  p.M = (int)M;
  int rc = plan_k_split(p, ti, q2k, K);
  if (rc) return rc;
  static const int variant_env = [] {
    const char *e = getenv("LAMM_DECODE_VARIANT");          // tuning knob: 8 or 16 (default) warps
    return (e && atoi(e) == 8) ? 8 : 16;
  }();
  // peer stores, sibling batches and the K-quants beyond Q2_K live in the 16-warp kernel
  static const int l2_ahead = [] {
    const char *e = getenv("LAMM_DECODE_L2_AHEAD");          // tuning knob: rows per warp prefetched into L2 (0 = off)
    const int v = e ? atoi(e) : 0;
    return v < 0 ? 0 : (v > 16 ? 16 : v);
  }();
  p.l2_ahead = l2_ahead;
  const int variant = (sh || bt || xop || is_kquant_ext(w_type)) ? 16 : variant_env;
  if (xop && xop->kind) {
    if (!x || (sh && sh->x_ll)) { set_error("decode: a fused op needs a plain f32 activation row"); return LAMMCU_ERR_ARG; }
    if (xop->kind == 1 && !(xop->eps > 0.f)) { set_error("decode: rms_norm needs eps > 0"); return LAMMCU_ERR_ARG; }
    if (((uintptr_t)xop->aux & 15) || (xop->kind == 2 && !xop->aux)) { set_error("decode: fused op operand NULL or not 16-byte aligned"); return LAMMCU_ERR_ALIGN; }
    p.xop_kind = xop->kind;  p.x_aux = xop->aux;  p.x_eps = xop->eps;
    if (xop->kind == 2 && !(p.silu_tab = silu_table_dev())) return LAMMCU_ERR_CUDA;
  }
  if (bt) {
    p.nbatch = bt->count;
    int64_t first[4] = {0, 0, 0, 0};
    for (int j = 1; j < bt->count; j++) first[j] = first[j - 1] + bt->M[j - 1];
    for (int j = 0; j < 4; j++) {
      const int jj = j < bt->count ? j : bt->count - 1;      // unused slots repeat the last matrix
      p.wb[j] = (const uint8_t *)bt->w[jj] - first[jj] * w_stride;
      p.cb[j] = sh ? nullptr : bt->c[jj] - first[jj];
      if (j > 0) p.mc[j - 1] = j < bt->count ? (int)first[j] : 0x7fffffff;
      if (sh)
        for (int g = 0; g < sh->n_peer; g++) p.llb[j][g] = sh->ll_batch[jj][g] + sh->row_begin_b[jj] - first[jj];
    }
  }
  int xm = x ? 1 : 0;
  if (sh) {
    for (int g = 0; g < sh->n_peer; g++) p.ll_peer[g] = sh->ll_peer[g];
    p.arrive_ctr = sh->arrive_ctr;  p.status = sh->status;
    p.n_peer = sh->n_peer;  p.my_rank = sh->my_rank;  p.row_begin = sh->row_begin;
    if (sh->x_ll) { p.x_ll = sh->x_ll;  p.x_back = sh->x_back;  xm = 2; }
  }
  int warps;
  if (variant == 16) {
    warps = (kD16Warps / p.KS) * p.KS;
  } else {
    warps = kDecWarps;                             // the most warps <= 10 that K-segments tile exactly
    if (kDecWarps % p.KS) warps = (kDecMaxWarps / p.KS) * p.KS;
  }
  p.RG = warps / p.KS;
  const int seg_bytes = q2k ? (p.seg_units / 2) * ti.size : p.seg_units * 4 * ti.size;
  p.slot_bytes = ((seg_bytes + 15) & ~15) + 16;
  p.x_bytes = q2k ? (p.units * (128 + 16) + p.kb * 4) : (p.units * 128 + p.kb * 8);
  p.x_bytes = (p.x_bytes + 127) & ~127;
  p.slots = (variant == 16) ? kD16Slots : kDecMaxSlots;
  int smem = kDecCtrl + p.x_bytes + warps * p.slots * p.slot_bytes;
  if (smem > kDecSmemMax) {
    set_error("decode: K=%lld too large for shared memory", (long long)K);
    return LAMMCU_ERR_SHAPE;
  }
  static const int per_sm = [] {
    const char *e = getenv("LAMM_DECODE_CTAS_PER_SM");       // tuning knob: 1 (default) or 2
    return (e && atoi(e) == 2) ? 2 : 1;
  }();
  int grid = sm_count_cached() * per_sm;
  const int64_t need = (M + p.RG - 1) / p.RG;
  if (grid > need) grid = (int)need;
  if (sh) grid = sm_count_cached();                // fixed grid: the arrival-ticket tag divides by it
  p.trace = next_trace_slot(grid);
  if (variant == 16) {
    if (xm == 2) return launch_decode16_x<2>(w_type, p, grid, warps * 32, smem, st);
    if (p.xop_kind) return launch_decode16_x<4>(w_type, p, grid, warps * 32, smem, st);    // the op-in-front instantiations
    return xm ? launch_decode16_x<1>(w_type, p, grid, warps * 32, smem, st)
              : launch_decode16_x<0>(w_type, p, grid, warps * 32, smem, st);
  }
  return x ? launch_decode_x<true>(w_type, p, grid, warps * 32, smem, st)
           : launch_decode_x<false>(w_type, p, grid, warps * 32, smem, st);
}

// ---- persistent decode chain: host side ------------------------------------------------------------
struct ChainPlan {
  int w_type, n_stages, grid, smem;
  ChainStage *stages_dev;
  unsigned int *ctrs;           // [0] stage-done counter (zeroed before every launch), [2] status
  ChainParams params;
};

template <int WT>
static int chain_launch_t(const ChainPlan *pl, cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(decode_chain_kernel<WT>, cudaFuncAttributeMaxDynamicSharedMemorySize, kDecSmemMax - 2048);
    if (e != cudaSuccess) { set_error("decode chain: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return LAMMCU_ERR_CUDA; }
    attr_set = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(pl->grid);
  cfg.blockDim = dim3(kChainWarps * 32);
  cfg.dynamicSmemBytes = pl->smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;      // all CTAs co-resident: the grid barrier cannot deadlock
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, decode_chain_kernel<WT>, pl->params);
  g_launch_count++;
  if (e != cudaSuccess) {
    set_error("decode chain launch (type %d, smem %d): %s", WT, pl->smem, cudaGetErrorString(e));
    return LAMMCU_ERR_CUDA;
  }
  return LAMMCU_OK;
}

int chain_run(ChainPlan *pl, cudaStream_t st) {
  pl->params.trace = (pl->n_stages * 4 <= pl->grid * 8) ? next_trace_slot(pl->grid) : nullptr;   // NULL unless armed
  cudaError_t e = cudaMemsetAsync(pl->ctrs, 0, 4, st);
  if (e != cudaSuccess) { set_error("decode chain: cudaMemsetAsync: %s", cudaGetErrorString(e)); return LAMMCU_ERR_CUDA; }
  switch (pl->w_type) {
  case LAMMCU_TYPE_Q4_0: return chain_launch_t<LAMMCU_TYPE_Q4_0>(pl, st);
  case LAMMCU_TYPE_Q4_1: return chain_launch_t<LAMMCU_TYPE_Q4_1>(pl, st);
  case LAMMCU_TYPE_Q5_0: return chain_launch_t<LAMMCU_TYPE_Q5_0>(pl, st);
  case LAMMCU_TYPE_Q5_1: return chain_launch_t<LAMMCU_TYPE_Q5_1>(pl, st);
  case LAMMCU_TYPE_Q8_0: return chain_launch_t<LAMMCU_TYPE_Q8_0>(pl, st);
  case LAMMCU_TYPE_Q2_K: return chain_launch_t<LAMMCU_TYPE_Q2_K>(pl, st);
  case LAMMCU_TYPE_Q3_K: return chain_launch_t<LAMMCU_TYPE_Q3_K>(pl, st);
  case LAMMCU_TYPE_Q4_K: return chain_launch_t<LAMMCU_TYPE_Q4_K>(pl, st);
  case LAMMCU_TYPE_Q5_K: return chain_launch_t<LAMMCU_TYPE_Q5_K>(pl, st);
  case LAMMCU_TYPE_Q6_K: return chain_launch_t<LAMMCU_TYPE_Q6_K>(pl, st);
  default: set_error("decode chain: type %d", pl->w_type); return LAMMCU_ERR_TYPE;
  }
}

void chain_destroy(ChainPlan *pl) {
  if (!pl) return;
  cudaFree(pl->stages_dev);
  cudaFree(pl->ctrs);
  delete pl;
}

int chain_create(int w_type, int n_ops, const ChainOp *ops, ChainPlan **out) {
  if (!ops || !out || n_ops < 1 || n_ops > 65536) { set_error("decode chain: bad argument"); return LAMMCU_ERR_ARG; }
  const TypeInfo ti = type_info(w_type);
  const bool q2k = (ti.blck == 256);
  std::vector<ChainStage> stages((size_t)n_ops);
  int slot_bytes = 0, x_bytes_max = 0;
  for (int i = 0; i < n_ops; i++) {
    const ChainOp &op = ops[i];
    ChainStage &S = stages[(size_t)i];
    S = ChainStage();
    static_assert(kChainWarps * kChainMaxSlots * 8 <= 1024 && kChainWarps * kChainMaxSlots * 4 <= 512, "control area");
    if (op.count < 1 || op.count > 4 || !op.x) { set_error("decode chain: op %d: 1..4 matrices and an activation row", i); return LAMMCU_ERR_ARG; }
    int64_t first[4] = {0, 0, 0, 0}, total = 0;
    for (int j = 0; j < op.count; j++) {
      if (!op.w[j] || !op.c[j] || op.M[j] <= 0) { set_error("decode chain: op %d: bad operand %d", i, j); return LAMMCU_ERR_ARG; }
      // same conditions as the sibling batch: every matrix the kernel can take, all rows at one 16-byte phase
      if (!decode_supported(w_type, op.w[j], op.w_stride, op.x, op.M[j], op.K) ||
          (((uintptr_t)total * (uintptr_t)op.w_stride) & 15) != 0 || (((uintptr_t)op.w[j] ^ (uintptr_t)op.w[0]) & 15) != 0) {
        set_error("decode chain: op %d matrix %d: needs a quantized type, K %% 256 == 0, 16-byte aligned operands and row "
                  "counts that keep every matrix of a batch at the same 16-byte phase", i, j);
        return LAMMCU_ERR_SHAPE;
      }
      first[j] = total;
      total += op.M[j];
    }
    if (total > 0x3fffffff) { set_error("decode chain: op %d: too many rows", i); return LAMMCU_ERR_SHAPE; }
    DecodeParams p = {};
    int rc = plan_k_split(p, ti, q2k, op.K);
    if (rc) return rc;
    if (p.KS > kD16Warps) { set_error("decode chain: op %d: K too large", i); return LAMMCU_ERR_SHAPE; }
    int warps = 0;
    plan_d16_smem(p, ti, q2k, &warps);
    // 32 warps: twice the row groups of the 16-warp plan; K-split row groups reduce through named barriers 1..15
    p.RG = kChainWarps / p.KS;
    if (p.KS > 1 && p.RG > 15) p.RG = 15;
    S.nbatch = op.count;
    for (int j = 0; j < 4; j++) {
      const int jj = j < op.count ? j : op.count - 1;
      S.wb[j] = (const uint8_t *)op.w[jj] - first[jj] * op.w_stride;
      S.cb[j] = op.c[jj] - first[jj];
      if (j > 0) S.mc[j - 1] = j < op.count ? (int)first[j] : 0x7fffffff;
    }
    S.x = op.x;  S.w_stride = op.w_stride;  S.M = (int)total;
    S.kb = p.kb;  S.units = p.units;  S.KS = p.KS;  S.seg_units = p.seg_units;  S.RG = p.RG;  S.x_bytes = p.x_bytes;
    S.inv_KS = (65536 + p.KS - 1) / p.KS;
    for (int wv = 0; wv < kChainWarps; wv++)
      if (((wv * S.inv_KS) >> 16) != wv / p.KS) { set_error("decode chain: reciprocal of KS=%d is not exact", p.KS); return LAMMCU_ERR_SHAPE; }
    slot_bytes = std::max(slot_bytes, p.slot_bytes);
    x_bytes_max = std::max(x_bytes_max, p.x_bytes);
  }
  const int budget = kDecSmemMax - 2048 /* static: the stage's DecodeParams */ - kChainCtrl - x_bytes_max;
  int slots = budget / (kChainWarps * slot_bytes);
  if (slots > kChainMaxSlots) slots = kChainMaxSlots;
  if (slots < 2) { set_error("decode chain: K too large for the shared-memory ring"); return LAMMCU_ERR_SHAPE; }
  ChainPlan *pl = new ChainPlan();
  pl->w_type = w_type;  pl->n_stages = n_ops;  pl->grid = sm_count_cached();
  pl->smem = kChainCtrl + x_bytes_max + kChainWarps * slots * slot_bytes;
  pl->stages_dev = nullptr;  pl->ctrs = nullptr;
  cudaError_t e = cudaMalloc(&pl->stages_dev, sizeof(ChainStage) * (size_t)n_ops);
  if (e == cudaSuccess) e = cudaMalloc(&pl->ctrs, 256);
  if (e == cudaSuccess) e = cudaMemcpy(pl->stages_dev, stages.data(), sizeof(ChainStage) * (size_t)n_ops, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemset(pl->ctrs, 0, 256);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    set_error("decode chain: %s", cudaGetErrorString(e));
    chain_destroy(pl);
    return LAMMCU_ERR_CUDA;
  }
  pl->params.stages = pl->stages_dev;  pl->params.n_stages = n_ops;
  pl->params.done_ctr = pl->ctrs;  pl->params.status = pl->ctrs + 2;
  pl->params.slots = slots;  pl->params.slot_bytes = slot_bytes;  pl->params.x_bytes_max = x_bytes_max;
  *out = pl;
  return LAMMCU_OK;
}

int chain_status(ChainPlan *pl, int *timed_out) {
  unsigned int v = 0;
  cudaError_t e = cudaMemcpy(&v, pl->ctrs + 2, 4, cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) { set_error("decode chain status: %s", cudaGetErrorString(e)); return LAMMCU_ERR_CUDA; }
  *timed_out = (int)v;
  return LAMMCU_OK;
}

}  // namespace lamm
