// lamm_quant_dev.cuh -- the activation-quantize arithmetic, shared by the stand-alone
// quantize kernels (lamm_quantize.cu) and the fused decode kernel (lamm_decode.cu).
//
// Bit-exact with the scalar reference quantizers (llama.cpp-b2430/ggml-quants.c:1182-1205,
// 1396-1429, 3981-4018; nearest_int :1767-1772).  Every f32 operation is an explicitly
// rounded intrinsic (__fdiv_rn, __fmul_rn, __fadd_rn): no FMA contraction, no approximate
// division; max/abs are exact and order-free.
#pragma once
#include "lamm_common.cuh"

namespace lamm {

// ---- the per-block arithmetic, shared with the fused GEMV prologue ---------

// 8 lanes x 4 elements = one 32-block.  Returns the 4 quants packed little-
// endian; d (f32, before fp16 rounding) and the integer sum of the block's
// quants are broadcast to the 8 lanes.
__device__ __forceinline__ uint32_t quant32_lane4(float4 v, float &d_out, int &sum_out) {
  float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
#pragma unroll
  for (int o = 4; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
  const float d = __fdiv_rn(amax, 127.0f);
  const float id = (d != 0.0f) ? __fdiv_rn(1.0f, d) : 0.0f;
  // roundf: half away from zero, as C's roundf
  const int q0 = (int)roundf(__fmul_rn(v.x, id));
  const int q1 = (int)roundf(__fmul_rn(v.y, id));
  const int q2 = (int)roundf(__fmul_rn(v.z, id));
  const int q3 = (int)roundf(__fmul_rn(v.w, id));
  int s = q0 + q1 + q2 + q3;
#pragma unroll
  for (int o = 4; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  d_out = d;
  sum_out = s;
  return (uint32_t)(q0 & 0xff) | ((uint32_t)(q1 & 0xff) << 8) | ((uint32_t)(q2 & 0xff) << 16) |
         ((uint32_t)(q3 & 0xff) << 24);
}

__device__ __forceinline__ int nearest_int_rne(float v) {
  const float t = __fadd_rn(v, 12582912.0f);           // 1.5 * 2^23
  return (__float_as_int(t) & 0x007fffff) - 0x00400000;
}

template <bool VEC>
__device__ __forceinline__ float4 load4(const float *p) {
  if (VEC) return *reinterpret_cast<const float4 *>(p);
  return make_float4(p[0], p[1], p[2], p[3]);
}


// One warp = one 256-super-block of Q8_K; lane owns elements [8*lane, 8*lane+8).
// Returns the lane's 8 quants packed in (q_lo, q_hi); d (f32) is broadcast; pair_sum is the
// sum of the 16-element group the lane belongs to (lanes 2g, 2g+1 -> bsums[g]).
__device__ __forceinline__ void quant256_lane8(const float (&v)[8], int lane, uint32_t &q_lo,
                                               uint32_t &q_hi, float &d_out, int &pair_sum) {
  // first index attaining the largest |x| (strict '>' scan in the reference)
  float amax = 0.0f;
  int idx = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) {
    const float ax = fabsf(v[j]);
    if (ax > amax) { amax = ax; idx = j; }
  }
  idx += lane * 8;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float oa = __shfl_xor_sync(0xffffffffu, amax, o);
    const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
    if (oa > amax || (oa == amax && oi < idx)) { amax = oa; idx = oi; }
  }
  float mx = 0.0f;
#pragma unroll
  for (int j = 0; j < 8; j++) mx = ((idx & 7) == j) ? v[j] : mx;
  mx = __shfl_sync(0xffffffffu, mx, idx >> 3);
  q_lo = 0; q_hi = 0;
  int sum = 0;
  float d = 0.0f;
  if (amax != 0.0f) {
    const float iscale = __fdiv_rn(-127.0f, mx);
    int q[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
      q[j] = min(127, nearest_int_rne(__fmul_rn(iscale, v[j])));
      sum += q[j];
    }
    q_lo = (uint32_t)(q[0] & 0xff) | ((uint32_t)(q[1] & 0xff) << 8) |
           ((uint32_t)(q[2] & 0xff) << 16) | ((uint32_t)(q[3] & 0xff) << 24);
    q_hi = (uint32_t)(q[4] & 0xff) | ((uint32_t)(q[5] & 0xff) << 8) |
           ((uint32_t)(q[6] & 0xff) << 16) | ((uint32_t)(q[7] & 0xff) << 24);
    d = __fdiv_rn(1.0f, iscale);
  }
  sum += __shfl_xor_sync(0xffffffffu, sum, 1);       // 16-element group = 2 lanes
  d_out = d;
  pair_sum = sum;
}

// pack 4 ints in [-128, 127] into little-endian int8 x 4 (2 instructions)
__device__ __forceinline__ uint32_t pack_s8x4(int q0, int q1, int q2, int q3) {
  uint32_t hi, r;
  asm("cvt.pack.sat.s8.s32.b32 %0, %1, %2, %3;" : "=r"(hi) : "r"(q3), "r"(q2), "r"(0));
  asm("cvt.pack.sat.s8.s32.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(q1), "r"(q0), "r"(hi));
  return r;
}

// roundf (half away from zero) for |t| <= 256, as an int: trunc(t + copysign(0.5 - 2^-25, t)).
// Exhaustively checked against C roundf for every float in [-256, 256] ("quantizers").
__device__ __forceinline__ int round_half_away_small(float t) {
  return __float2int_rz(__fadd_rn(t, copysignf(0.49999997f, t)));
}

// 4 lanes x 8 elements = one 32-block (lane l4 owns elements [8*l4, 8*l4 + 8)).  Same arithmetic
// as quant32_lane4; returns the lane's quants as two packed words, d and the block's quant sum.
__device__ __forceinline__ void quant32_lane8(const float4 a, const float4 b, uint32_t &q_lo, uint32_t &q_hi,
                                              float &d_out, int &sum_out) {
  float amax = fmaxf(fmaxf(fmaxf(fabsf(a.x), fabsf(a.y)), fmaxf(fabsf(a.z), fabsf(a.w))),
                     fmaxf(fmaxf(fabsf(b.x), fabsf(b.y)), fmaxf(fabsf(b.z), fabsf(b.w))));
  amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
  amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
  const float d = __fdiv_rn(amax, 127.0f);
  const float id = (d != 0.0f) ? __fdiv_rn(1.0f, d) : 0.0f;
  q_lo = pack_s8x4(round_half_away_small(__fmul_rn(a.x, id)), round_half_away_small(__fmul_rn(a.y, id)),
                   round_half_away_small(__fmul_rn(a.z, id)), round_half_away_small(__fmul_rn(a.w, id)));
  q_hi = pack_s8x4(round_half_away_small(__fmul_rn(b.x, id)), round_half_away_small(__fmul_rn(b.y, id)),
                   round_half_away_small(__fmul_rn(b.z, id)), round_half_away_small(__fmul_rn(b.w, id)));
  int s = __dp4a((int)q_lo, 0x01010101, 0);
  s = __dp4a((int)q_hi, 0x01010101, s);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  d_out = d;
  sum_out = s;
}

}  // namespace lamm
