// lamm_xop.cuh -- SURVEY 8f-3: the ops llama.cpp's graph puts in front of a matmul, fused into the activation
// quantizer so that their f32 result never makes a round trip through HBM:
//   kind 1  llm_build_norm (llama.cpp:5171-5201): ggml_rms_norm(x, eps) (ggml.c:10355-10405), then ggml_mul by the
//           norm weight (aux[K], may be NULL)
//   kind 2  llm_build_ffn, SiLU-gated (llama.cpp:5203-5290): ggml_silu(gate) (ggml.c:1894-1902: fp16 table) times up
//           (x = gate, aux = up)
// Every f32 operation is the same IEEE operation ggml performs, explicitly rounded (no FMA contraction, no
// approximate division or square root), so the activation the quantizer sees is ggml's bit for bit -- with one
// caveat: the sum of squares is a double like ggml's, but added in a tree instead of element order, so the f32 mean
// can differ in its last bit when the double sum lands within ~1e-16 (relative) of an f32 rounding boundary.
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>

namespace lamm {

struct XOp {
  int kind;               // 0: none
  const float *aux;       // kind 1: norm weight [K] or NULL; kind 2: up [N][ld_aux]
  int64_t ld_aux;         // kind 2: floats between rows of aux
  float eps;              // kind 1
  const uint16_t *silu;   // kind 2: device copy of ggml_table_silu_f16 (filled by silu_table_dev())
};

// ggml_vec_silu_f32 with GGML_SILU_FP16: round to fp16, look the pattern up, widen
__device__ __forceinline__ float silu_lookup(const uint16_t *tab, float g) {
  const unsigned short h = __half_as_ushort(__float2half_rn(g));
  return __half2float(__ushort_as_half(__ldg(tab + h)));
}

__device__ __forceinline__ float4 xop_norm4(float4 v, float scale, const float *w) {
  v.x = __fmul_rn(v.x, scale);  v.y = __fmul_rn(v.y, scale);
  v.z = __fmul_rn(v.z, scale);  v.w = __fmul_rn(v.w, scale);          // ggml_vec_scale_f32
  if (w) {                                                             // ggml_mul
    const float4 m = *reinterpret_cast<const float4 *>(w);
    v.x = __fmul_rn(v.x, m.x);  v.y = __fmul_rn(v.y, m.y);
    v.z = __fmul_rn(v.z, m.z);  v.w = __fmul_rn(v.w, m.w);
  }
  return v;
}
__device__ __forceinline__ float4 xop_silu4(float4 g, const float *up, const uint16_t *tab) {
  const float4 u = *reinterpret_cast<const float4 *>(up);
  g.x = __fmul_rn(silu_lookup(tab, g.x), u.x);  g.y = __fmul_rn(silu_lookup(tab, g.y), u.y);
  g.z = __fmul_rn(silu_lookup(tab, g.z), u.z);  g.w = __fmul_rn(silu_lookup(tab, g.w), u.w);
  return g;
}

// 1 / sqrtf(mean(x^2) + eps) of one row, computed by the whole CTA (every thread returns it).  x 16-byte aligned,
// K % 4 == 0.  scratch: 32 doubles + 1 float of shared memory.  Ends with a __syncthreads().
__device__ __forceinline__ float cta_rms_scale(const float *x, int K, float eps, uint8_t *scratch) {
  double s = 0.0;
  for (int i = (int)threadIdx.x * 4; i < K; i += (int)blockDim.x * 4) {
    const float4 v = *reinterpret_cast<const float4 *>(x + i);
    s += (double)__fmul_rn(v.x, v.x);  s += (double)__fmul_rn(v.y, v.y);
    s += (double)__fmul_rn(v.z, v.z);  s += (double)__fmul_rn(v.w, v.w);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  double *wp = reinterpret_cast<double *>(scratch);
  float *out = reinterpret_cast<float *>(scratch + 256);
  const int warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  if ((threadIdx.x & 31) == 0) wp[warp] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < nw; w++) t += wp[w];
    const float mean = (float)(t / (double)K);
    *out = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
  }
  __syncthreads();
  return *out;
}

const uint16_t *silu_table_dev();          // NULL (and an error message) if it cannot be built
int launch_xop_quantize(const XOp &op, int vt, const float *x, int64_t ldx, void *xq, int64_t n_rows, int64_t k,
                        cudaStream_t st);

}  // namespace lamm
