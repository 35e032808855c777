// lamm_gemm_prep.cu -- operand preparation for the prefill GEMM (lamm_gemm.cu).
//
// The tensor-core GEMM multiplies EXACT small integers held in fp16: the int8 activation
// quants q_x in [-127, 127] and the unpacked weight quants are exactly representable, their
// products and the per-block sums (< 2^24) are exact in the tensor core's f32 accumulator,
// so every 32-element block sum equals the reference's integer sumi bit for bit
// (src/lamm_kernel_q4_1.hpp:12-24: sumi, then d_w * d_x * sumi + m_w * s_x).
//
// This kernel turns the quantized activation rows Xq (ggml block_q8_0 / block_q8_1, the
// bit-exact output of lamm_quantize.cu) into the GEMM's B operand:
//   X16 : fp16 [Npad][K]   the quants as fp16, K contiguous.  Inside every 32-block the
//                          elements are stored in the order the weight unpacker produces
//                          them cheaply (perm 1, "nibble order", for Q4/Q5: 16-byte chunk w
//                          = elements 4w, 4w+2, 16+4w, 16+4w+2, 4w+1, 4w+3, 16+4w+1,
//                          16+4w+3; perm 0 = natural order for Q8_0).  A dot product does
//                          not care about the order as long as both operands agree.
//   dxT : f32 [Kb][Npad]   f32(d_x) of block kb of row n (transposed: one K-block's scales of
//                          all rows are contiguous, broadcast-read by the drain warps)
//   sxT : f32 [Kb][Npad]   f32(s_x) (q8_1 only)
// Rows n >= N are written as zeros.
#include "lamm_common.cuh"
#include "lamm_quant_dev.cuh"


namespace lamm {

template <int VT, int PERM>
__global__ void __launch_bounds__(256) gemm_prep_kernel(const uint8_t *__restrict__ xq, int64_t xq_stride,
                                                       __half *__restrict__ x16, float *__restrict__ dxT,
                                                       float *__restrict__ sxT, int N, int Npad, int kb) {
  constexpr int XB = (VT == LAMMCU_TYPE_Q8_0) ? 34 : 36;
  const int64_t total = (int64_t)Npad * kb;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int n = (int)(i / kb), b = (int)(i - (int64_t)n * kb);
    uint32_t x[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
    float d = 0.f, s = 0.f;
    if (n < N) {
      const uint8_t *blk = xq + (int64_t)n * xq_stride + (int64_t)b * XB;
      if constexpr (VT == LAMMCU_TYPE_Q8_1) {
        const uint32_t ds = *reinterpret_cast<const uint32_t *>(blk);
        d = half_bits_to_float(ds & 0xffffu);
        s = half_bits_to_float(ds >> 16);
#pragma unroll
        for (int j = 0; j < 8; j++) x[j] = *reinterpret_cast<const uint32_t *>(blk + 4 + 4 * j);
      } else {
        d = half_bits_to_float(*reinterpret_cast<const uint16_t *>(blk));
#pragma unroll
        for (int j = 0; j < 8; j++) {
          const uint16_t *h = reinterpret_cast<const uint16_t *>(blk + 2 + 4 * j);
          x[j] = (uint32_t)h[0] | ((uint32_t)h[1] << 16);
        }
      }
    }
    auto el = [&](int e) -> __half {                 // element e of the block as fp16
      const int v = (int)(signed char)((x[e >> 2] >> (8 * (e & 3))) & 0xffu);
      return __int2half_rn(v);
    };
    uint4 *dst = reinterpret_cast<uint4 *>(x16 + ((int64_t)n * kb + b) * 32);
#pragma unroll
    for (int w = 0; w < 4; w++) {
      __half h[8];
      if constexpr (PERM == 1) {
        h[0] = el(4 * w);      h[1] = el(4 * w + 2);      h[2] = el(16 + 4 * w);     h[3] = el(16 + 4 * w + 2);
        h[4] = el(4 * w + 1);  h[5] = el(4 * w + 3);      h[6] = el(16 + 4 * w + 1); h[7] = el(16 + 4 * w + 3);
      } else {
#pragma unroll
        for (int j = 0; j < 8; j++) h[j] = el(8 * w + j);
      }
      uint4 v;
      v.x = (uint32_t)__half_as_ushort(h[0]) | ((uint32_t)__half_as_ushort(h[1]) << 16);
      v.y = (uint32_t)__half_as_ushort(h[2]) | ((uint32_t)__half_as_ushort(h[3]) << 16);
      v.z = (uint32_t)__half_as_ushort(h[4]) | ((uint32_t)__half_as_ushort(h[5]) << 16);
      v.w = (uint32_t)__half_as_ushort(h[6]) | ((uint32_t)__half_as_ushort(h[7]) << 16);
      dst[w] = v;
    }
    dxT[(int64_t)b * Npad + n] = d;
    if (VT == LAMMCU_TYPE_Q8_1) sxT[(int64_t)b * Npad + n] = s;
  }
}

int launch_gemm_prep(int vt, int perm, const void *xq, int64_t xq_stride, void *x16, float *dxT, float *sxT,
                     int64_t N, int64_t Npad, int64_t K, cudaStream_t st) {
  const int kb = (int)(K / 32);
  const int64_t total = Npad * kb;
  int blocks = (int)((total + 255) / 256);
  const int cap = sm_count_cached() * 8;
  if (blocks > cap) blocks = cap;
  const uint8_t *p = (const uint8_t *)xq;
  __half *o = (__half *)x16;
  if (vt == LAMMCU_TYPE_Q8_1) {
    if (perm) gemm_prep_kernel<LAMMCU_TYPE_Q8_1, 1><<<blocks, 256, 0, st>>>(p, xq_stride, o, dxT, sxT, (int)N, (int)Npad, kb);
    else gemm_prep_kernel<LAMMCU_TYPE_Q8_1, 0><<<blocks, 256, 0, st>>>(p, xq_stride, o, dxT, sxT, (int)N, (int)Npad, kb);
  } else if (vt == LAMMCU_TYPE_Q8_0) {
    if (perm) gemm_prep_kernel<LAMMCU_TYPE_Q8_0, 1><<<blocks, 256, 0, st>>>(p, xq_stride, o, dxT, sxT, (int)N, (int)Npad, kb);
    else gemm_prep_kernel<LAMMCU_TYPE_Q8_0, 0><<<blocks, 256, 0, st>>>(p, xq_stride, o, dxT, sxT, (int)N, (int)Npad, kb);
  } else {
    set_error("gemm_prep: activation type %d not supported", vt);
    return LAMMCU_ERR_TYPE;
  }
  g_launch_count++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("gemm_prep launch: %s", cudaGetErrorString(e));
    return LAMMCU_ERR_CUDA;
  }
  return LAMMCU_OK;
}

// ---- scaled-operand variant (lamm_gemm_so.cu) ------------------------------------------------
// The B operand carries the activation scale: v = f32(d_x) * q_x is exact in f32 (11 + 8 significant
// bits).  Each activation row n is normalised by an exact power of two 2^-E_n (E_n from the row's
// largest block scale, so that |v| 2^-E_n < 2^11: any activation range fits fp16) and split into two
// fp16 terms, v 2^-E_n ~= hi + lo, |residue| <= 2^-22 |v| (or 2^-25 absolute for elements far below
// the row maximum).  The epilogue multiplies column n by colscale[n] = 2^E_n.
//   XH, XL   : fp16 [Npad][K]      same element order as X16 above (perm 1 / perm 0)
//   SH       : fp16 [Npad][Kbpad]  s_x 2^-E_n (q8_1 only; s_x is an fp16 field, so one term is exact)
//                                  for the m_w * s_x extra stages; Kbpad = Kb rounded up to 64, zero filled
//   colscale : f32 [Npad]
// One warp per activation row.
template <int VT, int PERM>
__global__ void __launch_bounds__(256) gemm_prep_so_kernel(const uint8_t *__restrict__ xq, int64_t xq_stride,
                                                          __half *__restrict__ xh, __half *__restrict__ xl,
                                                          __half *__restrict__ sh, float *__restrict__ colscale,
                                                          int N, int Npad, int kb, int kbpad) {
  constexpr int XB = (VT == LAMMCU_TYPE_Q8_0) ? 34 : 36;
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int n = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; n < Npad; n += warps) {
    const uint8_t *row = xq + (int64_t)n * xq_stride;
    float dmax = 0.f;
    if (n < N)
      for (int b = lane; b < kb; b += 32)
        dmax = fmaxf(dmax, fabsf(half_bits_to_float(*reinterpret_cast<const uint16_t *>(row + (int64_t)b * XB))));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dmax = fmaxf(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
    int E = 0;
    if (dmax > 0.f && dmax < 3.0e38f) E = ilogbf(dmax) - 3;         // |d q| < 2^(ilogb + 8)  ->  < 2^11 after 2^-E
    const float down = ldexpf(1.0f, -E);                            // exact power of two
    if (lane == 0) colscale[n] = ldexpf(1.0f, E);
    for (int b = lane; b < kbpad; b += 32) {
      if (b >= kb) {                                                // zero tail of the s_x tile row
        if (VT == LAMMCU_TYPE_Q8_1) sh[(int64_t)n * kbpad + b] = __float2half_rn(0.f);
        continue;
      }
      uint32_t x[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
      float d = 0.f, s = 0.f;
      if (n < N) {
        const uint8_t *blk = row + (int64_t)b * XB;
        if constexpr (VT == LAMMCU_TYPE_Q8_1) {
          const uint32_t ds = *reinterpret_cast<const uint32_t *>(blk);
          d = half_bits_to_float(ds & 0xffffu);
          s = half_bits_to_float(ds >> 16);
#pragma unroll
          for (int j = 0; j < 8; j++) x[j] = *reinterpret_cast<const uint32_t *>(blk + 4 + 4 * j);
        } else {
          d = half_bits_to_float(*reinterpret_cast<const uint16_t *>(blk));
#pragma unroll
          for (int j = 0; j < 8; j++) {
            const uint16_t *h = reinterpret_cast<const uint16_t *>(blk + 2 + 4 * j);
            x[j] = (uint32_t)h[0] | ((uint32_t)h[1] << 16);
          }
        }
      }
      const float ds_ = d * down;                       // exact
      auto el = [&](int e) -> float {                   // d_x * q_x * 2^-E of element e, exact in f32
        const int v = (int)(signed char)((x[e >> 2] >> (8 * (e & 3))) & 0xffu);
        return __fmul_rn(ds_, (float)v);
      };
      uint4 *dh = reinterpret_cast<uint4 *>(xh + ((int64_t)n * kb + b) * 32);
      uint4 *dl = reinterpret_cast<uint4 *>(xl + ((int64_t)n * kb + b) * 32);
#pragma unroll
      for (int w = 0; w < 4; w++) {
        float f[8];
        if constexpr (PERM == 1) {
          f[0] = el(4 * w);      f[1] = el(4 * w + 2);      f[2] = el(16 + 4 * w);     f[3] = el(16 + 4 * w + 2);
          f[4] = el(4 * w + 1);  f[5] = el(4 * w + 3);      f[6] = el(16 + 4 * w + 1); f[7] = el(16 + 4 * w + 3);
        } else {
#pragma unroll
          for (int j = 0; j < 8; j++) f[j] = el(8 * w + j);
        }
        uint32_t hw[4], lw[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const __half h0 = __float2half_rn(f[2 * j]), h1 = __float2half_rn(f[2 * j + 1]);
          const __half l0 = __float2half_rn(f[2 * j] - __half2float(h0));
          const __half l1 = __float2half_rn(f[2 * j + 1] - __half2float(h1));
          hw[j] = (uint32_t)__half_as_ushort(h0) | ((uint32_t)__half_as_ushort(h1) << 16);
          lw[j] = (uint32_t)__half_as_ushort(l0) | ((uint32_t)__half_as_ushort(l1) << 16);
        }
        dh[w] = make_uint4(hw[0], hw[1], hw[2], hw[3]);
        dl[w] = make_uint4(lw[0], lw[1], lw[2], lw[3]);
      }
      if (VT == LAMMCU_TYPE_Q8_1) sh[(int64_t)n * kbpad + b] = __float2half_rn(s * down);
    }
  }
}

int launch_gemm_prep_so(int vt, int perm, const void *xq, int64_t xq_stride, void *xh, void *xl, void *sh_,
                        float *colscale, int64_t N, int64_t Npad, int64_t K, cudaStream_t st) {
  const int kb = (int)(K / 32), kbpad = (kb + 63) / 64 * 64;
  __half *sh = (__half *)sh_;
  int blocks = (int)((Npad + 7) / 8);                 // 8 warps per CTA, one row per warp
  const int cap = sm_count_cached() * 8;
  if (blocks > cap) blocks = cap;
  const uint8_t *p = (const uint8_t *)xq;
  __half *h = (__half *)xh, *l = (__half *)xl;
  if (vt == LAMMCU_TYPE_Q8_1) {
    if (perm) gemm_prep_so_kernel<LAMMCU_TYPE_Q8_1, 1><<<blocks, 256, 0, st>>>(p, xq_stride, h, l, sh, colscale, (int)N, (int)Npad, kb, kbpad);
    else gemm_prep_so_kernel<LAMMCU_TYPE_Q8_1, 0><<<blocks, 256, 0, st>>>(p, xq_stride, h, l, sh, colscale, (int)N, (int)Npad, kb, kbpad);
  } else if (vt == LAMMCU_TYPE_Q8_0) {
    if (perm) gemm_prep_so_kernel<LAMMCU_TYPE_Q8_0, 1><<<blocks, 256, 0, st>>>(p, xq_stride, h, l, sh, colscale, (int)N, (int)Npad, kb, kbpad);
    else gemm_prep_so_kernel<LAMMCU_TYPE_Q8_0, 0><<<blocks, 256, 0, st>>>(p, xq_stride, h, l, sh, colscale, (int)N, (int)Npad, kb, kbpad);
  } else {
    set_error("gemm_prep_so: activation type %d not supported", vt);
    return LAMMCU_ERR_TYPE;
  }
  g_launch_count++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("gemm_prep_so launch: %s", cudaGetErrorString(e));
    return LAMMCU_ERR_CUDA;
  }
  return LAMMCU_OK;
}

// Fused activation quantize + operand prep: f32 X -> (bit-exact Q8_0 / Q8_1 quantization, in registers)
// -> XH, XL, SH, colscale as gemm_prep_so_kernel writes them; the ggml-layout Xq is only written when
// the caller wants it (xq_out != NULL).  Same arithmetic as quantize_row_q8_{0,1}_reference
// (ggml-quants.c:1182-1205, 1396-1429): d = amax / 127, id = 1 / d, q = roundf(x * id), y.d = fp16(d),
// y.s = fp16(sum * d).  One warp per row, one thread per 32-block.
template <int VT, int PERM>
__global__ void __launch_bounds__(256) quant_prep_so_kernel(const float *__restrict__ x, int64_t ldx,
                                                           uint8_t *__restrict__ xq_out, int64_t xq_stride,
                                                           __half *__restrict__ xh, __half *__restrict__ xl,
                                                           __half *__restrict__ sh, float *__restrict__ colscale,
                                                           int N, int Npad, int kb, int kbpad) {
  constexpr int XB = (VT == LAMMCU_TYPE_Q8_0) ? 34 : 36;
  __shared__ float red[8];
  const int lane = threadIdx.x & 31, tid = threadIdx.x;
  for (int n = blockIdx.x; n < Npad; n += gridDim.x) {          // one CTA per activation row
    const float *row = x + (int64_t)n * ldx;
    float dmax = 0.f;
    if (n < N) {
      for (int b = tid; b < kb; b += 256) {
        const float4 *src = reinterpret_cast<const float4 *>(row + (int64_t)b * 32);
        float amax = 0.f;
#pragma unroll
        for (int j = 0; j < 8; j++) {
          const float4 f = src[j];
          amax = fmaxf(amax, fmaxf(fmaxf(fabsf(f.x), fabsf(f.y)), fmaxf(fabsf(f.z), fabsf(f.w))));
        }
        dmax = fmaxf(dmax, __half2float(__float2half_rn(__fdiv_rn(amax, 127.0f))));
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dmax = fmaxf(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
    __syncthreads();                                              // red[] of the previous row is consumed
    if (lane == 0) red[tid >> 5] = dmax;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 8; i++) dmax = fmaxf(dmax, red[i]);
    int E = 0;
    if (dmax > 0.f && dmax < 3.0e38f) E = ilogbf(dmax) - 3;
    const float down = ldexpf(1.0f, -E);
    if (tid == 0) colscale[n] = ldexpf(1.0f, E);
    for (int b = tid; b < kbpad; b += 256) {
      if (b >= kb) {
        if (VT == LAMMCU_TYPE_Q8_1) sh[(int64_t)n * kbpad + b] = __float2half_rn(0.f);
        continue;
      }
      int q[32];
      float dsc = 0.f;
      __half hd = __float2half_rn(0.f), hs = __float2half_rn(0.f);
#pragma unroll
      for (int e = 0; e < 32; e++) q[e] = 0;
      if (n < N) {
        const float4 *src = reinterpret_cast<const float4 *>(row + (int64_t)b * 32);
        float v[32];
        float amax = 0.f;
#pragma unroll
        for (int j = 0; j < 8; j++) {
          const float4 f = src[j];
          v[4 * j] = f.x;  v[4 * j + 1] = f.y;  v[4 * j + 2] = f.z;  v[4 * j + 3] = f.w;
          amax = fmaxf(amax, fmaxf(fmaxf(fabsf(f.x), fabsf(f.y)), fmaxf(fabsf(f.z), fabsf(f.w))));
        }
        const float d = __fdiv_rn(amax, 127.0f);
        const float id = (d != 0.0f) ? __fdiv_rn(1.0f, d) : 0.0f;
        int sum = 0;
#pragma unroll
        for (int e = 0; e < 32; e++) {
          q[e] = round_half_away_small(__fmul_rn(v[e], id));
          sum += q[e];
        }
        hd = __float2half_rn(d);
        if (VT == LAMMCU_TYPE_Q8_1) hs = __float2half_rn(__fmul_rn((float)sum, d));
        dsc = __half2float(hd) * down;
        if (xq_out) {                                     // ggml block, for the bit-exact check
          uint8_t *blk = xq_out + (int64_t)n * xq_stride + (int64_t)b * XB;
          uint16_t *o16 = reinterpret_cast<uint16_t *>(blk);
          o16[0] = __half_as_ushort(hd);
          int off = 1;
          if (VT == LAMMCU_TYPE_Q8_1) { o16[1] = __half_as_ushort(hs); off = 2; }
#pragma unroll
          for (int e = 0; e < 16; e++)
            o16[off + e] = (uint16_t)((uint32_t)(q[2 * e] & 0xff) | ((uint32_t)(q[2 * e + 1] & 0xff) << 8));
        }
      }
      auto el = [&](int e) -> float { return __fmul_rn(dsc, (float)q[e]); };
      uint4 *dh = reinterpret_cast<uint4 *>(xh + ((int64_t)n * kb + b) * 32);
      uint4 *dl = reinterpret_cast<uint4 *>(xl + ((int64_t)n * kb + b) * 32);
#pragma unroll
      for (int w = 0; w < 4; w++) {
        float f[8];
        if constexpr (PERM == 1) {
          f[0] = el(4 * w);      f[1] = el(4 * w + 2);      f[2] = el(16 + 4 * w);     f[3] = el(16 + 4 * w + 2);
          f[4] = el(4 * w + 1);  f[5] = el(4 * w + 3);      f[6] = el(16 + 4 * w + 1); f[7] = el(16 + 4 * w + 3);
        } else {
#pragma unroll
          for (int j = 0; j < 8; j++) f[j] = el(8 * w + j);
        }
        uint32_t hw[4], lw[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const __half h0 = __float2half_rn(f[2 * j]), h1 = __float2half_rn(f[2 * j + 1]);
          const __half l0 = __float2half_rn(f[2 * j] - __half2float(h0));
          const __half l1 = __float2half_rn(f[2 * j + 1] - __half2float(h1));
          hw[j] = (uint32_t)__half_as_ushort(h0) | ((uint32_t)__half_as_ushort(h1) << 16);
          lw[j] = (uint32_t)__half_as_ushort(l0) | ((uint32_t)__half_as_ushort(l1) << 16);
        }
        dh[w] = make_uint4(hw[0], hw[1], hw[2], hw[3]);
        dl[w] = make_uint4(lw[0], lw[1], lw[2], lw[3]);
      }
      if (VT == LAMMCU_TYPE_Q8_1) sh[(int64_t)n * kbpad + b] = __float2half_rn(__half2float(hs) * down);
    }
  }
}

int launch_quant_prep_so(int vt, int perm, const float *x, int64_t ldx, void *xq_out, int64_t xq_stride, void *xh,
                         void *xl, void *sh_, float *colscale, int64_t N, int64_t Npad, int64_t K, cudaStream_t st) {
  const int kb = (int)(K / 32), kbpad = (kb + 63) / 64 * 64;
  __half *sh = (__half *)sh_;
  int blocks = (int)Npad;
  const int cap = sm_count_cached() * 16;
  if (blocks > cap) blocks = cap;
  __half *h = (__half *)xh, *l = (__half *)xl;
  uint8_t *xo = (uint8_t *)xq_out;
#define LAMM_QP(VT_, P_) quant_prep_so_kernel<VT_, P_><<<blocks, 256, 0, st>>>(x, ldx, xo, xq_stride, h, l, sh, colscale, (int)N, (int)Npad, kb, kbpad)
  if (vt == LAMMCU_TYPE_Q8_1) { if (perm) LAMM_QP(LAMMCU_TYPE_Q8_1, 1); else LAMM_QP(LAMMCU_TYPE_Q8_1, 0); }
  else if (vt == LAMMCU_TYPE_Q8_0) { if (perm) LAMM_QP(LAMMCU_TYPE_Q8_0, 1); else LAMM_QP(LAMMCU_TYPE_Q8_0, 0); }
  else {
    set_error("quant_prep_so: activation type %d not supported", vt);
    return LAMMCU_ERR_TYPE;
  }
#undef LAMM_QP
  g_launch_count++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("quant_prep_so launch: %s", cudaGetErrorString(e));
    return LAMMCU_ERR_CUDA;
  }
  return LAMMCU_OK;
}

// ---- direct variant (gemm_direct_kernel): Q2_K / F32 weights ------------------------------------
__device__ __forceinline__ void split_store32(const float (&v)[32], __half *hi, __half *lo) {
  uint4 *dh = reinterpret_cast<uint4 *>(hi), *dl = reinterpret_cast<uint4 *>(lo);
#pragma unroll
  for (int w = 0; w < 4; w++) {
    uint32_t hw[4], lw[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const float a = v[8 * w + 2 * j], b = v[8 * w + 2 * j + 1];
      const __half h0 = __float2half_rn(a), h1 = __float2half_rn(b);
      const __half l0 = __float2half_rn(a - __half2float(h0)), l1 = __float2half_rn(b - __half2float(h1));
      hw[j] = (uint32_t)__half_as_ushort(h0) | ((uint32_t)__half_as_ushort(h1) << 16);
      lw[j] = (uint32_t)__half_as_ushort(l0) | ((uint32_t)__half_as_ushort(l1) << 16);
    }
    dh[w] = make_uint4(hw[0], hw[1], hw[2], hw[3]);
    dl[w] = make_uint4(lw[0], lw[1], lw[2], lw[3]);
  }
}

// ---- K-quants beyond the reference's table: one 32-element group g of a super-block ----
// Layouts ggml-common.h:228-313; element order and scale packing as dequantize_row_q{3,4,5,6}_K
// (ggml-quants.c:2550-2598, 2851-2874, 3134-3160, 3383-3412).  Blocks are only 2-byte aligned for
// Q3_K / Q6_K, so everything is read as bytes or 16-bit words.
__device__ __forceinline__ uint32_t ldg16(const uint8_t *p) { return *reinterpret_cast<const uint16_t *>(p); }
// get_scale_min_k4 (ggml-quants.c:2027-2034): 6-bit scale and min of group j from the 12 packed bytes
__device__ __forceinline__ void scale_min_k4(int j, const uint8_t *q, int &sc, int &mn) {
  if (j < 4) {
    sc = q[j] & 63;
    mn = q[j + 4] & 63;
  } else {
    sc = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4);
    mn = (q[j + 4] >> 4) | ((q[j] >> 6) << 4);
  }
}
// 6-bit Q3_K scale j (0..15), before the -32 (ggml-quants.c:2570-2575)
__device__ __forceinline__ int scale_q3k(int j, const uint8_t *s) {
  const int lo = (j < 8) ? (s[j] & 0xF) : (s[j - 8] >> 4);
  const int hi = (s[8 + (j & 3)] >> (2 * (j >> 2))) & 3;
  return lo | (hi << 4);
}
template <int WT>
__device__ __forceinline__ void kquant_group32(const uint8_t *blk, int g, float S, float (&v)[32]) {
  if constexpr (WT == LAMMCU_TYPE_Q4_K || WT == LAMMCU_TYPE_Q5_K) {
    // {f16 d, dmin; u8 scales[12]; (Q5_K: u8 qh[32];) u8 qs[128]}
    const float d = half_bits_to_float(ldg16(blk)) * S, dmin = half_bits_to_float(ldg16(blk + 2)) * S;
    int sc, mn;
    scale_min_k4(g, blk + 4, sc, mn);
    const float dl = d * (float)sc, ml = dmin * (float)mn;
    const uint8_t *qs = blk + (WT == LAMMCU_TYPE_Q4_K ? 16 : 48) + 32 * (g >> 1);
    const int sh = 4 * (g & 1);
#pragma unroll
    for (int l = 0; l < 32; l += 2) {
      const uint32_t q2 = ldg16(qs + l) >> sh;
      uint32_t q0 = q2 & 15u, q1 = (q2 >> 8) & 15u;
      if constexpr (WT == LAMMCU_TYPE_Q5_K) {
        const uint32_t h2 = ldg16(blk + 16 + l) >> g;
        q0 |= (h2 & 1u) << 4;
        q1 |= ((h2 >> 8) & 1u) << 4;
      }
      v[l] = fmaf(dl, (float)q0, -ml);
      v[l + 1] = fmaf(dl, (float)q1, -ml);
    }
  } else if constexpr (WT == LAMMCU_TYPE_Q6_K) {
    // {u8 ql[128]; u8 qh[64]; i8 scales[16]; f16 d}
    const int n = g >> 2, j = g & 3;
    const float d = half_bits_to_float(ldg16(blk + 208)) * S;
    const signed char *scp = reinterpret_cast<const signed char *>(blk + 192 + 8 * n + 2 * j);
    const float d0 = d * (float)scp[0], d1 = d * (float)scp[1];
    const uint8_t *ql = blk + 64 * n + 32 * (j & 1), *qh = blk + 128 + 32 * n;
    const int shl = 4 * (j >> 1), shh = 2 * j;
#pragma unroll
    for (int l = 0; l < 32; l += 2) {
      const uint32_t a = ldg16(ql + l) >> shl, h = ldg16(qh + l) >> shh;
      const int q0 = (int)((a & 15u) | ((h & 3u) << 4)) - 32;
      const int q1 = (int)(((a >> 8) & 15u) | (((h >> 8) & 3u) << 4)) - 32;
      v[l] = (l < 16 ? d0 : d1) * (float)q0;
      v[l + 1] = (l < 16 ? d0 : d1) * (float)q1;
    }
  } else {   // Q3_K {u8 hmask[32]; u8 qs[64]; u8 scales[12]; f16 d}
    const int n = g >> 2, j = g & 3;
    const float d = half_bits_to_float(ldg16(blk + 108)) * S;
    const float d0 = d * (float)(scale_q3k(8 * n + 2 * j, blk + 96) - 32);
    const float d1 = d * (float)(scale_q3k(8 * n + 2 * j + 1, blk + 96) - 32);
    const uint8_t *qs = blk + 32 + 32 * n;
    const int hb = 4 * n + j;
#pragma unroll
    for (int l = 0; l < 32; l += 2) {
      const uint32_t a = ldg16(qs + l) >> (2 * j), h = ldg16(blk + l) >> hb;
      const int q0 = (int)(a & 3u) - ((h & 1u) ? 0 : 4);
      const int q1 = (int)((a >> 8) & 3u) - (((h >> 8) & 1u) ? 0 : 4);
      v[l] = (l < 16 ? d0 : d1) * (float)q0;
      v[l + 1] = (l < 16 ? d0 : d1) * (float)q1;
    }
  }
}

// W (Q2_K .. Q6_K super-blocks or f32 rows) -> AH, AL fp16 [Mpad][K], value * 2^F_m, natural K order, and
// rowscale[m] = 2^F_m: an exact power of two per weight row that lifts the row's largest |w| to [2^14, 2^15) so that
// neither large weights (|w| >= 1024 used to overflow a fixed 2^6 pre-scale) nor tiny ones (lo term lost to fp16
// subnormals) leave the 22-bit hi + lo representation; the GEMM epilogue multiplies row m by 2^-F_m.
// Q2_K element (ggml-common.h:214-225, src/lamm_kernel_q2_k.hpp:49-71): for half n of the super-block and
// shift s in {0,2,4,6}, elements 128n + 16s .. + 31 are (qs[32n + l] >> s) & 3 with scale bytes
// scales[8n + s] (l < 16) and scales[8n + s + 1]; w = d * (sc & 15) * q - dmin * (sc >> 4).
template <int WT>
__device__ __forceinline__ void dequant_group32(const uint8_t *row, int b, float S, float (&v)[32]) {
  if constexpr (WT == LAMMCU_TYPE_Q2_K) {
    const uint8_t *blk = row + (int64_t)(b >> 3) * 84;
    const int ib = b & 7, n = ib >> 2, shift = 2 * (ib & 3);
    const uint32_t dd = *reinterpret_cast<const uint32_t *>(blk + 80);
    const float d = half_bits_to_float(dd & 0xffffu) * S, dmin = half_bits_to_float(dd >> 16) * S;
    const uint32_t sc2 = *reinterpret_cast<const uint16_t *>(blk + 2 * ib);
    const float d0 = d * (float)(sc2 & 15u), m0 = dmin * (float)((sc2 >> 4) & 15u);
    const float d1 = d * (float)((sc2 >> 8) & 15u), m1 = dmin * (float)(sc2 >> 12);
    const uint32_t *qs = reinterpret_cast<const uint32_t *>(blk + 16 + 32 * n);
#pragma unroll
    for (int j = 0; j < 8; j++) {
      const uint32_t qw = qs[j] >> shift;
#pragma unroll
      for (int e = 0; e < 4; e++) {
        const float qv = (float)((qw >> (8 * e)) & 3u);
        v[4 * j + e] = (j < 4) ? fmaf(d0, qv, -m0) : fmaf(d1, qv, -m1);
      }
    }
  } else if constexpr (is_kquant_ext(WT)) {
    kquant_group32<WT>(row + (int64_t)(b >> 3) * type_info(WT).size, b & 7, S, v);
  } else {
    const float4 *src = reinterpret_cast<const float4 *>(row + (int64_t)b * 128);
#pragma unroll
    for (int j = 0; j < 8; j++) {
      const float4 f = src[j];
      v[4 * j] = f.x * S;  v[4 * j + 1] = f.y * S;  v[4 * j + 2] = f.z * S;  v[4 * j + 3] = f.w * S;
    }
  }
}

// 2^F with 2^14 <= vmax 2^F < 2^15 (1 for an all-zero or non-finite row); 2^F and 2^-F both stay normal f32
__device__ __forceinline__ float lift_scale(float vmax) {
  if (!(vmax > 0.f) || !(vmax < 3.0e38f)) return 1.0f;
  const int F = 14 - ilogbf(vmax);
  return ldexpf(1.0f, F > 120 ? 120 : F);
}

template <int WT>
__global__ void __launch_bounds__(256) dequant_hilo_kernel(const uint8_t *__restrict__ w, int64_t w_stride,
                                                          __half *__restrict__ ah, __half *__restrict__ al,
                                                          float *__restrict__ rowscale, int M, int Mpad, int K) {
  const int kb32 = K / 32;
  __shared__ float red[8];
  const int tid = threadIdx.x, lane = tid & 31;
  for (int m = blockIdx.x; m < Mpad; m += gridDim.x) {              // one CTA per weight row
    const uint8_t *row = w + (int64_t)m * w_stride;
    float vmax = 0.f;
    if (m < M) {
      for (int b = tid; b < kb32; b += 256) {
        float v[32];
        dequant_group32<WT>(row, b, 1.0f, v);
#pragma unroll
        for (int l = 0; l < 32; l++) vmax = fmaxf(vmax, fabsf(v[l]));
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
    __syncthreads();                                                // red[] of the previous row is consumed
    if (lane == 0) red[tid >> 5] = vmax;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 8; i++) vmax = fmaxf(vmax, red[i]);
    const float S = lift_scale(vmax);
    if (tid == 0) rowscale[m] = S;
    for (int b = tid; b < kb32; b += 256) {
      float v[32];
#pragma unroll
      for (int l = 0; l < 32; l++) v[l] = 0.f;
      if (m < M) dequant_group32<WT>(row, b, S, v);
      split_store32(v, ah + ((int64_t)m * kb32 + b) * 32, al + ((int64_t)m * kb32 + b) * 32);
    }
  }
}

// The scaled-operand kernel unpacks W on the fly, so its per-row lift comes from a pass over the block headers:
// rowscale[m] = 2^F_m from max_kb max(|d_w| * QMAX, |m_w|).  One warp per row.
template <int WT>
__global__ void __launch_bounds__(256) wrow_scale_kernel(const uint8_t *__restrict__ w, int64_t w_stride,
                                                        float *__restrict__ rowscale, int M, int Mpad, int kb) {
  constexpr int BS = type_info(WT).size;
  constexpr float QMAX = (WT == LAMMCU_TYPE_Q4_0) ? 8.f : (WT == LAMMCU_TYPE_Q4_1) ? 15.f : (WT == LAMMCU_TYPE_Q5_0) ? 16.f
                         : (WT == LAMMCU_TYPE_Q5_1) ? 31.f : 128.f;
  constexpr bool kMin = (WT == LAMMCU_TYPE_Q4_1 || WT == LAMMCU_TYPE_Q5_1);
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int m = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; m < Mpad; m += warps) {
    float vmax = 0.f;
    if (m < M) {
      const uint8_t *row = w + (int64_t)m * w_stride;
      for (int b = lane; b < kb; b += 32) {
        const uint8_t *blk = row + (int64_t)b * BS;
        vmax = fmaxf(vmax, fabsf(half_bits_to_float(ldg16(blk))) * QMAX);
        if (kMin) vmax = fmaxf(vmax, fabsf(half_bits_to_float(ldg16(blk + 2))));
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
    if (lane == 0) rowscale[m] = lift_scale(vmax);
  }
}

int launch_wrow_scale(int w_type, const void *w, int64_t w_stride, float *rowscale, int64_t M, int64_t Mpad, int64_t K,
                      cudaStream_t st) {
  const int kb = (int)(K / 32);
  int blocks = (int)((Mpad + 7) / 8);
  const int cap = sm_count_cached() * 8;
  if (blocks > cap) blocks = cap;
  const uint8_t *p = (const uint8_t *)w;
  switch (w_type) {
#define LAMM_WR(T) case T: wrow_scale_kernel<T><<<blocks, 256, 0, st>>>(p, w_stride, rowscale, (int)M, (int)Mpad, kb); break;
  LAMM_WR(LAMMCU_TYPE_Q4_0) LAMM_WR(LAMMCU_TYPE_Q4_1) LAMM_WR(LAMMCU_TYPE_Q5_0) LAMM_WR(LAMMCU_TYPE_Q5_1) LAMM_WR(LAMMCU_TYPE_Q8_0)
#undef LAMM_WR
  default: set_error("wrow_scale: weight type %d not supported", w_type); return LAMMCU_ERR_TYPE;
  }
  g_launch_count++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("wrow_scale launch: %s", cudaGetErrorString(e));
    return LAMMCU_ERR_CUDA;
  }
  return LAMMCU_OK;
}
int launch_dequant_hilo(int w_type, const void *w, int64_t w_stride, void *ah, void *al, float *rowscale, int64_t M,
                        int64_t Mpad, int64_t K, cudaStream_t st) {
  int blocks = (int)Mpad;
  const int cap = sm_count_cached() * 16;
  if (blocks > cap) blocks = cap;
  if (w_type == LAMMCU_TYPE_Q2_K)
    dequant_hilo_kernel<LAMMCU_TYPE_Q2_K><<<blocks, 256, 0, st>>>((const uint8_t *)w, w_stride, (__half *)ah, (__half *)al,
                                                               rowscale, (int)M, (int)Mpad, (int)K);
  else if (w_type == LAMMCU_TYPE_F32)
    dequant_hilo_kernel<LAMMCU_TYPE_F32><<<blocks, 256, 0, st>>>((const uint8_t *)w, w_stride, (__half *)ah, (__half *)al,
                                                              rowscale, (int)M, (int)Mpad, (int)K);
#define LAMM_DQ_CASE(T)                                                                                             \
  else if (w_type == T)                                                                                             \
    dequant_hilo_kernel<T><<<blocks, 256, 0, st>>>((const uint8_t *)w, w_stride, (__half *)ah, (__half *)al, rowscale, \
                                                   (int)M, (int)Mpad, (int)K);
  LAMM_DQ_CASE(LAMMCU_TYPE_Q3_K)
  LAMM_DQ_CASE(LAMMCU_TYPE_Q4_K)
  LAMM_DQ_CASE(LAMMCU_TYPE_Q5_K)
  LAMM_DQ_CASE(LAMMCU_TYPE_Q6_K)
#undef LAMM_DQ_CASE
  else {
    set_error("dequant_hilo: weight type %d not supported", w_type);
    return LAMMCU_ERR_TYPE;
  }
  g_launch_count++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("dequant_hilo launch: %s", cudaGetErrorString(e));
    return LAMMCU_ERR_CUDA;
  }
  return LAMMCU_OK;
}

// Activations for the direct variant: Q8_K blocks (x.d is f32, ggml-common.h:316-321) or plain f32 rows
// -> XH, XL fp16 [Npad][K] (natural order), row-normalised by 2^-E_n, and colscale[n] = 2^E_n.
template <int VT>
__global__ void __launch_bounds__(256) prep_direct_kernel(const uint8_t *__restrict__ xq, int64_t xq_stride,
                                                         __half *__restrict__ xh, __half *__restrict__ xl,
                                                         float *__restrict__ colscale, int N, int Npad, int K) {
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const int kb32 = K / 32;
  for (int n = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; n < Npad; n += warps) {
    const uint8_t *row = xq + (int64_t)n * xq_stride;
    float vmax = 0.f;                                   // upper bound of |value| in the row
    if (n < N) {
      if constexpr (VT == LAMMCU_TYPE_Q8_K) {
        for (int sb = lane; sb < K / 256; sb += 32)
          vmax = fmaxf(vmax, fabsf(*reinterpret_cast<const float *>(row + (int64_t)sb * 292)) * 128.f);
      } else {
        const float *x = reinterpret_cast<const float *>(row);
        for (int k = lane; k < K; k += 32) vmax = fmaxf(vmax, fabsf(x[k]));
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
    int E = 0;
    if (vmax > 0.f && vmax < 3.0e38f) E = ilogbf(vmax) - 10;       // |v| 2^-E < 2^11
    const float down = ldexpf(1.0f, -E);
    if (lane == 0) colscale[n] = ldexpf(1.0f, E);
    for (int b = lane; b < kb32; b += 32) {
      float v[32];
#pragma unroll
      for (int l = 0; l < 32; l++) v[l] = 0.f;
      if (n < N) {
        if constexpr (VT == LAMMCU_TYPE_Q8_K) {
          const uint8_t *blk = row + (int64_t)(b >> 3) * 292;
          const float d = *reinterpret_cast<const float *>(blk) * down;
          const uint32_t *qs = reinterpret_cast<const uint32_t *>(blk + 4 + 32 * (b & 7));
#pragma unroll
          for (int j = 0; j < 8; j++) {
            const uint32_t qw = qs[j];
#pragma unroll
            for (int e = 0; e < 4; e++) v[4 * j + e] = d * (float)(int)(signed char)((qw >> (8 * e)) & 0xffu);
          }
        } else {
          const float4 *src = reinterpret_cast<const float4 *>(row + (int64_t)b * 128);
#pragma unroll
          for (int j = 0; j < 8; j++) {
            const float4 f = src[j];
            v[4 * j] = f.x * down;  v[4 * j + 1] = f.y * down;  v[4 * j + 2] = f.z * down;  v[4 * j + 3] = f.w * down;
          }
        }
      }
      split_store32(v, xh + ((int64_t)n * kb32 + b) * 32, xl + ((int64_t)n * kb32 + b) * 32);
    }
  }
}

int launch_prep_direct(int vt, const void *xq, int64_t xq_stride, void *xh, void *xl, float *colscale, int64_t N,
                       int64_t Npad, int64_t K, cudaStream_t st) {
  int blocks = (int)((Npad + 7) / 8);
  const int cap = sm_count_cached() * 8;
  if (blocks > cap) blocks = cap;
  if (vt == LAMMCU_TYPE_Q8_K)
    prep_direct_kernel<LAMMCU_TYPE_Q8_K><<<blocks, 256, 0, st>>>((const uint8_t *)xq, xq_stride, (__half *)xh, (__half *)xl,
                                                              colscale, (int)N, (int)Npad, (int)K);
  else if (vt == LAMMCU_TYPE_F32)
    prep_direct_kernel<LAMMCU_TYPE_F32><<<blocks, 256, 0, st>>>((const uint8_t *)xq, xq_stride, (__half *)xh, (__half *)xl,
                                                             colscale, (int)N, (int)Npad, (int)K);
  else {
    set_error("prep_direct: activation type %d not supported", vt);
    return LAMMCU_ERR_TYPE;
  }
  g_launch_count++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("prep_direct launch: %s", cudaGetErrorString(e));
    return LAMMCU_ERR_CUDA;
  }
  return LAMMCU_OK;
}

}  // namespace lamm
