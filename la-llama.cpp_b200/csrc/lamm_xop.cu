// lamm_xop.cu -- SURVEY 8f-3: RMS-norm (x norm weight) and SiLU(gate) x up fused into the activation quantizer.
// See lamm_xop.cuh for what is fused and why the result is ggml's bit for bit.  N == 1 decode does not come here: the
// decode kernel applies the same device functions while it stages the row (lamm_decode.cu, load_x8).
//
// One CTA per activation row (grid-stride): HBM-bound, 4 (+4 for up) bytes read and ~1.06 written per element, the
// norm's second pass over the row is served by L1/L2.  Rows are 16-byte aligned (float4 loads).
#include "lamm_quant_dev.cuh"
#include "lamm_xop.cuh"

#include <cmath>
#include <mutex>
#include <vector>

namespace lamm {

template <int VT>
__global__ void __launch_bounds__(256) xop_quantize_kernel(XOp op, const float *__restrict__ x, int64_t ldx,
                                                          uint8_t *__restrict__ xq, int64_t n_rows, int k) {
  __shared__ __align__(16) uint8_t scratch[512];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NW = 8;
  for (int64_t row = blockIdx.x; row < n_rows; row += gridDim.x) {
    const float *xr = x + row * ldx;
    const float *ar = (op.kind == 2) ? op.aux + row * op.ld_aux : op.aux;
    float scale = 1.f;
    if (op.kind == 1) scale = cta_rms_scale(xr, k, op.eps, scratch);
    auto ld4 = [&](int e) -> float4 {
      float4 v = *reinterpret_cast<const float4 *>(xr + e);
      if (op.kind == 1) v = xop_norm4(v, scale, ar ? ar + e : nullptr);
      else if (op.kind == 2) v = xop_silu4(v, ar + e, op.silu);
      return v;
    };
    if constexpr (VT == LAMMCU_TYPE_F32) {
      float *out = reinterpret_cast<float *>(xq) + row * (int64_t)k;
      for (int e = tid * 4; e < k; e += 256 * 4) *reinterpret_cast<float4 *>(out + e) = ld4(e);
    } else if constexpr (VT == LAMMCU_TYPE_F16) {
      __half *out = reinterpret_cast<__half *>(xq) + row * (int64_t)k;
      for (int e = tid * 4; e < k; e += 256 * 4) {
        const float4 v = ld4(e);
        const __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
        uint2 o;
        o.x = *reinterpret_cast<const uint32_t *>(&a);
        o.y = *reinterpret_cast<const uint32_t *>(&b);
        *reinterpret_cast<uint2 *>(out + e) = o;
      }
    } else if constexpr (VT == LAMMCU_TYPE_Q8_K) {
      const int kb = k >> 8;
      uint8_t *out = xq + row * (int64_t)kb * 292;
      for (int sb = warp; sb < kb; sb += NW) {
        const float4 a = ld4(sb * 256 + lane * 8), c = ld4(sb * 256 + lane * 8 + 4);
        const float v[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
        uint32_t q_lo, q_hi;
        float d;
        int sum;
        quant256_lane8(v, lane, q_lo, q_hi, d, sum);
        uint8_t *blk = out + (int64_t)sb * 292;
        uint32_t *qs = reinterpret_cast<uint32_t *>(blk + Q8_K_QS + lane * 8);
        qs[0] = q_lo;
        qs[1] = q_hi;
        if ((lane & 1) == 0) *reinterpret_cast<int16_t *>(blk + Q8_K_BSUMS + lane) = (int16_t)sum;
        if (lane == 0) *reinterpret_cast<float *>(blk) = d;
      }
    } else {
      constexpr int BS = (VT == LAMMCU_TYPE_Q8_0) ? 34 : 36;
      constexpr int QS = (VT == LAMMCU_TYPE_Q8_0) ? Q8_0_QS : Q8_1_QS;
      const int kb = k >> 5, sub = lane >> 3, l8 = lane & 7;
      uint8_t *out = xq + row * (int64_t)kb * BS;
      for (int g = warp; g * 4 < kb; g += NW) {
        const int b = g * 4 + sub;
        const bool live = b < kb;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (live) v = ld4(b * 32 + l8 * 4);
        float d;
        int sum;
        const uint32_t q = quant32_lane4(v, d, sum);
        if (!live) continue;
        uint8_t *blk = out + (int64_t)b * BS;
        if constexpr (VT == LAMMCU_TYPE_Q8_1) {
          *reinterpret_cast<uint32_t *>(blk + QS + l8 * 4) = q;
          if (l8 == 0) {
            const float s = __fmul_rn((float)sum, d);
            *reinterpret_cast<__half2 *>(blk) = __halves2half2(__float2half_rn(d), __float2half_rn(s));
          }
        } else {
          uint16_t *p = reinterpret_cast<uint16_t *>(blk + QS + l8 * 4);
          p[0] = (uint16_t)(q & 0xffff);
          p[1] = (uint16_t)(q >> 16);
          if (l8 == 0) *reinterpret_cast<__half *>(blk) = __float2half_rn(d);
        }
      }
    }
  }
}

// ggml_init fills ggml_table_silu_f16 on the host with the host's libm (ggml.c:2705-2720); so does this, once per
// device, and the 128 KB table lives in device memory from then on (served from L2 / L1 by the lookups).
const uint16_t *silu_table_dev() {
  static std::mutex mu;
  static uint16_t *tab = nullptr;
  static int tab_dev = -1;
  std::lock_guard<std::mutex> lk(mu);
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) { set_error("silu table: cudaGetDevice failed"); return nullptr; }
  if (tab && tab_dev == dev) return tab;
  std::vector<uint16_t> host(1 << 16);
  for (int i = 0; i < (1 << 16); i++) {
    const float f = __half2float(__ushort_as_half((unsigned short)i));
    host[i] = __half_as_ushort(__float2half_rn(f / (1.0f + expf(-f))));       // ggml_silu_f32, ggml.c:1883-1885
  }
  uint16_t *p = nullptr;
  cudaError_t e = cudaMalloc(&p, host.size() * 2);
  if (e == cudaSuccess) e = cudaMemcpy(p, host.data(), host.size() * 2, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) { set_error("silu table: %s", cudaGetErrorString(e)); if (p) cudaFree(p); return nullptr; }
  tab = p;                                    // an older device's table is left to that device's context
  tab_dev = dev;
  return tab;
}

int launch_xop_quantize(const XOp &op_in, int vt, const float *x, int64_t ldx, void *xq, int64_t n_rows, int64_t k,
                        cudaStream_t st) {
  XOp op = op_in;
  if (op.kind != 1 && op.kind != 2) { set_error("xop: kind %d (1 = rms_norm x weight, 2 = silu(gate) x up)", op.kind); return LAMMCU_ERR_ARG; }
  if (op.kind == 1 && !(op.eps > 0.f)) { set_error("xop: rms_norm needs eps > 0 (ggml asserts it, ggml.c:10377)"); return LAMMCU_ERR_ARG; }
  if (op.kind == 2 && !op.aux) { set_error("xop: silu(gate) x up needs the up rows"); return LAMMCU_ERR_ARG; }
  const int blck = (vt == LAMMCU_TYPE_F32 || vt == LAMMCU_TYPE_F16) ? 8 : type_info(vt).blck;
  if (vt != LAMMCU_TYPE_F32 && vt != LAMMCU_TYPE_F16 && vt != LAMMCU_TYPE_Q8_0 && vt != LAMMCU_TYPE_Q8_1 &&
      vt != LAMMCU_TYPE_Q8_K) {
    set_error("xop: type %d is not an activation type", vt);
    return LAMMCU_ERR_TYPE;
  }
  if (n_rows <= 0 || k <= 0 || k % blck || k > (1 << 24)) {
    set_error("xop: bad shape rows=%lld k=%lld", (long long)n_rows, (long long)k);
    return LAMMCU_ERR_SHAPE;
  }
  if ((((uintptr_t)x | (uintptr_t)op.aux) & 15) || ldx % 4 || (op.kind == 2 && op.ld_aux % 4) || ((uintptr_t)xq & 7)) {
    set_error("xop: rows of x / aux must be 16-byte aligned, the output 8-byte aligned");
    return LAMMCU_ERR_ALIGN;
  }
  if (op.kind == 2) {
    op.silu = silu_table_dev();
    if (!op.silu) return LAMMCU_ERR_CUDA;
  }
  int64_t blocks = n_rows;
  const int64_t cap = (int64_t)sm_count_cached() * 8;
  if (blocks > cap) blocks = cap;
  uint8_t *out = (uint8_t *)xq;
  switch (vt) {
  case LAMMCU_TYPE_F32:  xop_quantize_kernel<LAMMCU_TYPE_F32><<<(int)blocks, 256, 0, st>>>(op, x, ldx, out, n_rows, (int)k); break;
  case LAMMCU_TYPE_F16:  xop_quantize_kernel<LAMMCU_TYPE_F16><<<(int)blocks, 256, 0, st>>>(op, x, ldx, out, n_rows, (int)k); break;
  case LAMMCU_TYPE_Q8_0: xop_quantize_kernel<LAMMCU_TYPE_Q8_0><<<(int)blocks, 256, 0, st>>>(op, x, ldx, out, n_rows, (int)k); break;
  case LAMMCU_TYPE_Q8_1: xop_quantize_kernel<LAMMCU_TYPE_Q8_1><<<(int)blocks, 256, 0, st>>>(op, x, ldx, out, n_rows, (int)k); break;
  default:               xop_quantize_kernel<LAMMCU_TYPE_Q8_K><<<(int)blocks, 256, 0, st>>>(op, x, ldx, out, n_rows, (int)k); break;
  }
  g_launch_count++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("xop quantize launch: %s", cudaGetErrorString(e)); return LAMMCU_ERR_CUDA; }
  return LAMMCU_OK;
}

}  // namespace lamm
