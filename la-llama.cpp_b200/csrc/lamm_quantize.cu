// lamm_quantize.cu -- activation re-quantization f32 -> Q8_0 / Q8_1 / Q8_K.
//
// Bit-exact with the scalar reference quantizers the reference platform runs
// in ggml's INIT phase (llama.cpp-b2430/ggml.c:10865-10887):
//   quantize_row_q8_0_reference  ggml-quants.c:1182-1205
//   quantize_row_q8_1_reference  ggml-quants.c:1396-1429
//   quantize_row_q8_K_reference  ggml-quants.c:3981-4018 (+ nearest_int :1767)
// Every f32 operation below is an explicitly rounded intrinsic (__fdiv_rn,
// __fmul_rn, __fadd_rn) so nvcc can neither contract a*b+c into an FMA nor
// substitute an approximate division; max/abs are exact and order-free.
//
// HBM-bound: 4 B read + ~1.06 B written per element.  One lane owns 4 (Q8_0/1)
// or 8 (Q8_K) consecutive elements, read as float4; 8 / 32 lanes form one block.
#include "lamm_quant_dev.cuh"

namespace lamm {

// ---- Q8_0 / Q8_1 ------------------------------------------------------------
// grid-stride over groups of 4 blocks (one warp = 4 blocks of one row).
template <int VT, bool VEC>
__global__ void __launch_bounds__(256) quantize_q8_01_kernel(const float *__restrict__ x,
                                                            int64_t ldx, uint8_t *__restrict__ xq,
                                                            int64_t n_rows, int64_t kb_per_row) {
  constexpr int BS = (VT == LAMMCU_TYPE_Q8_0) ? 34 : 36;
  constexpr int QS = (VT == LAMMCU_TYPE_Q8_0) ? Q8_0_QS : Q8_1_QS;
  const int lane = threadIdx.x & 31;
  const int sub = lane >> 3, l8 = lane & 7;
  const int64_t groups_per_row = (kb_per_row + 3) >> 2;
  const int64_t n_groups = n_rows * groups_per_row;
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t g = warp0; g < n_groups; g += n_warps) {
    const int64_t row = g / groups_per_row;
    const int64_t kb = (g - row * groups_per_row) * 4 + sub;
    const bool live = kb < kb_per_row;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (live) v = load4<VEC>(x + row * ldx + kb * 32 + l8 * 4);
    float d;
    int sum;
    const uint32_t q = quant32_lane4(v, d, sum);
    if (!live) continue;
    uint8_t *blk = xq + (row * kb_per_row + kb) * BS;
    if (VT == LAMMCU_TYPE_Q8_1) {
      *reinterpret_cast<uint32_t *>(blk + QS + l8 * 4) = q;            // 36-B blocks: 4-B aligned
      if (l8 == 0) {
        const float s = __fmul_rn((float)sum, d);
        __half2 ds = __halves2half2(__float2half_rn(d), __float2half_rn(s));
        *reinterpret_cast<__half2 *>(blk) = ds;
      }
    } else {
      uint16_t *p = reinterpret_cast<uint16_t *>(blk + QS + l8 * 4);   // 34-B blocks: 2-B aligned
      p[0] = (uint16_t)(q & 0xffff);
      p[1] = (uint16_t)(q >> 16);
      if (l8 == 0) *reinterpret_cast<__half *>(blk) = __float2half_rn(d);
    }
  }
}

// ---- Q8_K -------------------------------------------------------------------
// one warp = one 256-super-block; lane owns elements [8*lane, 8*lane+8).
template <bool VEC>
__global__ void __launch_bounds__(256) quantize_q8_K_kernel(const float *__restrict__ x,
                                                           int64_t ldx, uint8_t *__restrict__ xq,
                                                           int64_t n_rows, int64_t kb_per_row) {
  const int lane = threadIdx.x & 31;
  const int64_t n_blocks = n_rows * kb_per_row;
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t b = warp0; b < n_blocks; b += n_warps) {
    const int64_t row = b / kb_per_row, kb = b - row * kb_per_row;
    const float *xp = x + row * ldx + kb * 256 + lane * 8;
    const float4 a = load4<VEC>(xp), c = load4<VEC>(xp + 4);
    const float v[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
    uint32_t q_lo, q_hi;
    float d;
    int sum;
    quant256_lane8(v, lane, q_lo, q_hi, d, sum);
    uint8_t *blk = xq + b * 292;
    uint32_t *qs = reinterpret_cast<uint32_t *>(blk + Q8_K_QS + lane * 8);   // 292 % 4 == 0
    qs[0] = q_lo;
    qs[1] = q_hi;
    // all-zero block: the reference leaves bsums untouched; we write zeros.
    if ((lane & 1) == 0)
      *reinterpret_cast<int16_t *>(blk + Q8_K_BSUMS + lane) = (int16_t)sum;  // bsums[lane/2]
    if (lane == 0) *reinterpret_cast<float *>(blk) = d;
  }
}

// ---- F16 ---------------------------------------------------------
// ggml converts src1 to the F16 vec-dot type with ggml_fp32_to_fp16_row (ggml.c:10865-10887 -> :378-: IEEE round to
// nearest even, F16C on x86): __float2half_rn is bit-identical.  4 elements per thread.
__global__ void __launch_bounds__(256) convert_f16_kernel(const float *__restrict__ x, int64_t ldx, __half *__restrict__ y,
                                                         int64_t n_rows, int64_t k) {
  const int64_t k4 = k >> 2, total = n_rows * k4;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = i / k4, c = (i - row * k4) * 4;
    const float *p = x + row * ldx + c;
    const __half2 a = __floats2half2_rn(p[0], p[1]), b = __floats2half2_rn(p[2], p[3]);
    uint2 o;
    o.x = *reinterpret_cast<const uint32_t *>(&a);
    o.y = *reinterpret_cast<const uint32_t *>(&b);
    *reinterpret_cast<uint2 *>(y + row * k + c) = o;
  }
}

int launch_quantize_act(int vt, const float *x, int64_t ldx, void *xq, int64_t n_rows, int64_t k,
                        cudaStream_t st) {
  if (vt == LAMMCU_TYPE_F16) {
    if (n_rows <= 0 || k <= 0 || k % 8 || ((uintptr_t)x & 3) || ((uintptr_t)xq & 7)) {
      set_error("quantize_act: F16 needs K %% 8 == 0 and an 8-byte aligned output");
      return LAMMCU_ERR_SHAPE;
    }
    int64_t blocks = (n_rows * (k / 4) + 255) / 256;
    const int64_t cap = (int64_t)sm_count_cached() * 16;
    if (blocks > cap) blocks = cap;
    convert_f16_kernel<<<(int)blocks, 256, 0, st>>>(x, ldx, (__half *)xq, n_rows, k);
    g_launch_count++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("convert_f16 launch: %s", cudaGetErrorString(e)); return LAMMCU_ERR_CUDA; }
    return LAMMCU_OK;
  }
  const TypeInfo ti = type_info(vt);
  if (vt != LAMMCU_TYPE_Q8_0 && vt != LAMMCU_TYPE_Q8_1 && vt != LAMMCU_TYPE_Q8_K) {
    set_error("quantize_act: type %d is not an activation type", vt);
    return LAMMCU_ERR_TYPE;
  }
  if (n_rows <= 0 || k <= 0 || k % ti.blck) {
    set_error("quantize_act: bad shape rows=%lld k=%lld", (long long)n_rows, (long long)k);
    return LAMMCU_ERR_SHAPE;
  }
  if (((uintptr_t)x & 3) || ((uintptr_t)xq & 3)) {
    set_error("quantize_act: x and xq must be 4-byte aligned");
    return LAMMCU_ERR_ALIGN;
  }
  const bool vec = (((uintptr_t)x & 15) == 0) && (ldx % 4 == 0);
  const int64_t kb = k / ti.blck;
  const int64_t warps = (vt == LAMMCU_TYPE_Q8_K) ? n_rows * kb : n_rows * ((kb + 3) / 4);
  int64_t blocks = (warps + 7) / 8;
  const int64_t cap = (int64_t)sm_count_cached() * 8;
  if (blocks > cap) blocks = cap;
  uint8_t *out = (uint8_t *)xq;
  if (vt == LAMMCU_TYPE_Q8_0) {
    if (vec) quantize_q8_01_kernel<LAMMCU_TYPE_Q8_0, true><<<blocks, 256, 0, st>>>(x, ldx, out, n_rows, kb);
    else quantize_q8_01_kernel<LAMMCU_TYPE_Q8_0, false><<<blocks, 256, 0, st>>>(x, ldx, out, n_rows, kb);
  } else if (vt == LAMMCU_TYPE_Q8_1) {
    if (vec) quantize_q8_01_kernel<LAMMCU_TYPE_Q8_1, true><<<blocks, 256, 0, st>>>(x, ldx, out, n_rows, kb);
    else quantize_q8_01_kernel<LAMMCU_TYPE_Q8_1, false><<<blocks, 256, 0, st>>>(x, ldx, out, n_rows, kb);
  } else {
    if (vec) quantize_q8_K_kernel<true><<<blocks, 256, 0, st>>>(x, ldx, out, n_rows, kb);
    else quantize_q8_K_kernel<false><<<blocks, 256, 0, st>>>(x, ldx, out, n_rows, kb);
  }
  g_launch_count++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("quantize_act launch: %s", cudaGetErrorString(e));
    return LAMMCU_ERR_CUDA;
  }
  return LAMMCU_OK;
}

}  // namespace lamm
