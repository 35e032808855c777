// lamm_gemm.cu -- prefill path (N > gemv_max_n): the one dense contraction of the hot path,
// on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM).
//
// Replaces LAMMImpl<T>::matmul_simd_block + lamm_simd_block_kernel<4,4>
// (src/lamm_impl.hpp:90-147, src/lamm_kernel_*.hpp) for many activation rows.
//
// C[n, m] = sum_kb d_w[m,kb] * d_x[n,kb] * sumi[m,n,kb]  (+ m_w[m,kb] * s_x[n,kb])
// The scale changes every 32-element block on BOTH operands, so the integer block sum has to
// leave the tensor core once per block.  Design:
//   * kind::f16 MMA on exact small integers: the unpacked weight quants and the int8
//     activation quants are exactly representable in fp16, the block sums are exact in the
//     f32 accumulator (|sumi| < 2^24) -- bit-identical to the reference's int32 sumi -- and
//     arrive in registers as floats: the drain is one packed FMUL2 + one packed FFMA2 per TWO
//     accumulator elements instead of I2F + FMUL + FFMA per element for an s32 accumulator.
//   * one CTA = 128 weight rows x 128 activation rows, 16 warps:
//       warp 0      TMA producer: raw packed weight blocks (2-D box over bytes), the fp16
//                   activation tile (SWIZZLE_128B), the per-block activation scales
//       warp 1      MMA issuer (one thread): per 32-block 2 x tcgen05.mma M128 N128 K16; two
//                   blocks (256 TMEM columns) per hand-off, ring of 2, tcgen05.commit -> mbarrier
//       warp 2      TMEM allocation
//       warps 4-7   weight unpack (thread = row): nibbles / 5th bits -> fp16 with the
//                   0x6400-magic trick, written in the UMMA K-major SWIZZLE_128B layout
//       warps 8-15  drain (thread = TMEM lane = weight row, 64 columns each): tcgen05.ld,
//                   acc += (d_w * d_x[n]) * S  with f32x2 math, accumulators in registers
//   * the epilogue stores C[n*ldc + m] straight from registers (consecutive lanes =
//     consecutive m: coalesced).
#include "lamm_tc.cuh"

#include <cstdlib>

namespace lamm {

constexpr int kGStages = 4;                    // A16 / B16 ring
constexpr int kSStages = 2;                    // TMEM accumulator ring (2 x 128 columns each)
constexpr int kRawStages = 2;                  // raw packed-weight ring (256 K elements each)
constexpr int kScaleRing = 16;                 // K-blocks of d_w / d_x kept in shared memory
constexpr int kGemmThreads = 512;

struct GemmParams {
  const float *dxT, *sxT;     // [Kb][Npad]
  float *c;
  int64_t ldc;
  int M, N, Npad, K;
  int debug;                  // diagnostics (LAMM_GEMM_DEBUG): 4 = no unpack
};

// One pipeline stage (64 K = blocks 2Q, 2Q+1 of the raw chunk) of the thread's row r.
template <int WT, int Q, int NWR>
__device__ __forceinline__ void unpack_stage(const uint32_t (&raw)[NWR], uint8_t *a_stage, int r, float *dws,
                                             float *mws, int kbi0) {
  const uint32_t row_base = smem_u32(a_stage) + (uint32_t)(r >> 3) * 1024u + (uint32_t)(r & 7) * 128u;
#pragma unroll
  for (int j = 0; j < 2; j++) {
    uint4 out[4];
    float d, m;
    if (j == 0) unpack_block<WT, 2 * Q, NWR>(raw, out, d, m);
    else unpack_block<WT, 2 * Q + 1, NWR>(raw, out, d, m);
#pragma unroll
    for (int w = 0; w < 4; w++) {
      const uint32_t c = (uint32_t)(4 * j + w) ^ (uint32_t)(r & 7);
      asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(row_base + c * 16u), "r"(out[w].x), "r"(out[w].y),
                   "r"(out[w].z), "r"(out[w].w)
                   : "memory");
    }
    const uint32_t slot = (uint32_t)(((kbi0 + j) % kScaleRing) * GM + r) * 4u;
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(smem_u32(dws) + slot), "f"(d) : "memory");
    if constexpr (WT == LAMMCU_TYPE_Q4_1 || WT == LAMMCU_TYPE_Q5_1)
      asm volatile("st.shared.f32 [%0], %1;" ::"r"(smem_u32(mws) + slot), "f"(m) : "memory");
  }
}

// ---- the kernel ----------------------------------------------------------------------------
// shared memory map (offsets from the 1024-aligned base)
template <int WT> struct GemmSmem {
  static constexpr int BS = type_info(WT).size;
  static constexpr int kRawRowBytes = 8 * BS;                       // 8 blocks = 256 K elements
  static constexpr int kRawBytes = GM * kRawRowBytes;
  static constexpr int kAOff = 0;
  static constexpr int kBOff = kAOff + kGStages * GM * GK * 2;
  static constexpr int kRawOff = kBOff + kGStages * GN * GK * 2;
  static constexpr int kDxOff = kRawOff + kRawStages * ((kRawBytes + 127) / 128 * 128);
  static constexpr bool kMin = (WT == LAMMCU_TYPE_Q4_1 || WT == LAMMCU_TYPE_Q5_1);   // s_x / m_w rings only then
  static constexpr int kSxOff = kDxOff + kScaleRing * GN * 4;
  static constexpr int kDwOff = kSxOff + (kMin ? kScaleRing * GN * 4 : 0);
  static constexpr int kMwOff = kDwOff + kScaleRing * GM * 4;
  static constexpr int kBarOff = kMwOff + (kMin ? kScaleRing * GM * 4 : 0);
  static constexpr int kTotal = kBarOff + 256;
};

template <int WT>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_x, const GemmParams p) {
  using L = GemmSmem<WT>;
  constexpr bool kMin = (WT == LAMMCU_TYPE_Q4_1 || WT == LAMMCU_TYPE_Q5_1);
  constexpr int NWR = 2 * L::BS;                                    // raw words per row chunk
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + L::kBarOff);
  uint64_t *full_b = bars, *full_a = bars + 4, *empty = bars + 8, *raw_full = bars + 12, *raw_empty = bars + 14,
           *s_full = bars + 16, *s_empty = bars + 19;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 24);
  float *dxs = reinterpret_cast<float *>(smem + L::kDxOff), *sxs = reinterpret_cast<float *>(smem + L::kSxOff);
  float *dws = reinterpret_cast<float *>(smem + L::kDwOff), *mws = reinterpret_cast<float *>(smem + L::kMwOff);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tiles = (p.N + GN - 1) / GN;
  const int m0 = (blockIdx.x / n_tiles) * GM, n0 = (blockIdx.x % n_tiles) * GN;
  const int n_stages = p.K / GK, n_kb = p.K / 32;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kGStages; i++) {
      mbar_init(&full_b[i], 1);
      mbar_init(&full_a[i], 4);          // one arrival per unpack warp
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < kRawStages; i++) {
      mbar_init(&raw_full[i], 1);
      mbar_init(&raw_empty[i], 4);
    }
    for (int i = 0; i < kSStages; i++) {
      mbar_init(&s_full[i], 1);
      mbar_init(&s_empty[i], 8);         // one arrival per drain warp
    }
    fence_barrier_init();
    prefetch_tmap(&tm_w);
    prefetch_tmap(&tm_x);
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      for (int st = 0; st < n_stages; st++) {
        if ((st & 3) == 0) {
          const int ms = st >> 2, r = ms % kRawStages;
          mbar_wait(&raw_empty[r], ((uint32_t)(ms / kRawStages) & 1u) ^ 1u);
          mbar_arrive_expect_tx(&raw_full[r], (uint32_t)L::kRawBytes);
          tma_load_2d(smem + L::kRawOff + r * ((L::kRawBytes + 127) / 128 * 128), &tm_w, ms * (L::kRawRowBytes / 2), m0,
                      &raw_full[r]);
        }
        const int s = st % kGStages;
        mbar_wait(&empty[s], ((uint32_t)(st / kGStages) & 1u) ^ 1u);
        mbar_arrive_expect_tx(&full_b[s], (uint32_t)(GN * GK * 2 + 2 * GN * 4 * (kMin ? 2 : 1)));
        tma_load_2d(smem + L::kBOff + s * (GN * GK * 2), &tm_x, st * GK, n0, &full_b[s]);
#pragma unroll
        for (int j = 0; j < 2; j++) {
          const int kbi = 2 * st + j, slot = (kbi % kScaleRing) * GN;
          bulk_g2s(dxs + slot, p.dxT + (int64_t)kbi * p.Npad + n0, GN * 4, &full_b[s]);
          if (kMin) bulk_g2s(sxs + slot, p.sxT + (int64_t)kbi * p.Npad + n0, GN * 4, &full_b[s]);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      for (int st = 0; st < n_stages; st++) {
        const int s = st % kGStages;
        const uint32_t ph = (uint32_t)(st / kGStages) & 1u;
        mbar_wait(&full_a[s], ph);
        mbar_wait(&full_b[s], ph);
        tc_fence_after();
        const uint64_t adesc = make_sw128_desc(smem_u32(smem + L::kAOff + s * (GM * GK * 2)));
        const uint64_t bdesc = make_sw128_desc(smem_u32(smem + L::kBOff + s * (GN * GK * 2)));
        // one TMEM stage = the two 32-blocks of this pipeline stage (2 x 128 columns)
        const int sr = st % kSStages;
        mbar_wait(&s_empty[sr], ((uint32_t)(st / kSStages) & 1u) ^ 1u);
        tc_fence_after();
#pragma unroll
        for (int j = 0; j < 2; j++) {
          // the K offset inside the 128-byte swizzle atom advances the start address (16-byte units)
          const uint32_t d = tmem_base + (uint32_t)(sr * 2 * GN + j * GN);
          umma_f16(d, adesc + (uint64_t)(j * 4), bdesc + (uint64_t)(j * 4), kIdescF16, 0u);
          umma_f16(d, adesc + (uint64_t)(j * 4 + 2), bdesc + (uint64_t)(j * 4 + 2), kIdescF16, 1u);
        }
        umma_commit(&s_full[sr]);
        umma_commit(&empty[s]);
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // ===================== weight unpack (thread = row) =====================
    const int r = threadIdx.x - 128;
    for (int ms = 0; ms < n_stages / 4; ms++) {
      const int rs = ms % kRawStages;
      mbar_wait(&raw_full[rs], (uint32_t)(ms / kRawStages) & 1u);
      uint32_t raw[NWR];
      const uint32_t src = smem_u32(smem + L::kRawOff + rs * ((L::kRawBytes + 127) / 128 * 128)) + (uint32_t)r * L::kRawRowBytes;
#pragma unroll
      for (int i = 0; i < NWR / 4; i++)
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(raw[4 * i]), "=r"(raw[4 * i + 1]), "=r"(raw[4 * i + 2]), "=r"(raw[4 * i + 3])
                     : "r"(src + 16u * i));
      __syncwarp();
      if (lane == 0) mbar_arrive_after_loads(&raw_empty[rs], raw, (uint32_t)p.K & 1u);   // (lamm_common.cuh: not before the loads)
#pragma unroll
      for (int q = 0; q < 4; q++) {
        const int st = ms * 4 + q, s = st % kGStages;
        mbar_wait(&empty[s], ((uint32_t)(st / kGStages) & 1u) ^ 1u);
        uint8_t *a_stage = smem + L::kAOff + s * (GM * GK * 2);
        if (p.debug & 4) {
        } else if (q == 0) unpack_stage<WT, 0, NWR>(raw, a_stage, r, dws, mws, 2 * st);
        else if (q == 1) unpack_stage<WT, 1, NWR>(raw, a_stage, r, dws, mws, 2 * st);
        else if (q == 2) unpack_stage<WT, 2, NWR>(raw, a_stage, r, dws, mws, 2 * st);
        else unpack_stage<WT, 3, NWR>(raw, a_stage, r, dws, mws, 2 * st);
        fence_proxy_async();             // generic-proxy stores -> visible to the tensor core
        __syncwarp();
        if (lane == 0) mbar_arrive(&full_a[s]);
      }
    }
  } else if (warp >= 8) {
    // ===================== drain (thread = weight row, 64 activation rows) =====================
    const int q4 = warp & 3, h = (warp - 8) >> 2;
    const int row = q4 * 32 + lane;
    float2 acc[32];
#pragma unroll
    for (int i = 0; i < 32; i++) acc[i] = make_float2(0.f, 0.f);
    const uint32_t dxs_a = smem_u32(dxs), sxs_a = smem_u32(sxs), dws_a = smem_u32(dws), mws_a = smem_u32(mws);
    // 16-column TMEM loads, double-buffered: the load of chunk i+1 is in flight during the math of chunk i
    auto ld16 = [&](uint32_t taddr, uint32_t (&r)[16]) {
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
          "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
          : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
            "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
          : "r"(taddr)
          : "memory");
    };
    for (int st = 0; st < n_stages; st++) {
      const int sr = st % kSStages;
      mbar_wait(&s_full[sr], (uint32_t)(st / kSStages) & 1u);
      tc_fence_after();
      const uint32_t tbase = tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(sr * 2 * GN + h * 64);
      uint32_t sv[2][16];
      ld16(tbase, sv[0]);
      tmem_ld_wait();
#pragma unroll
      for (int ch = 0; ch < 8; ch++) {           // chunk = (K-block j = ch / 4, columns 16 * (ch % 4) ..)
        const int j = ch >> 2, c16 = ch & 3;
        if (ch + 1 < 8) ld16(tbase + (uint32_t)(((ch + 1) >> 2) * GN + ((ch + 1) & 3) * 16), sv[(ch + 1) & 1]);
        const int slot = (2 * st + j) % kScaleRing;
        const float dw = lds_f32(dws_a + (uint32_t)(slot * GM + row) * 4u);
        const float2 dw2 = make_float2(dw, dw);
        const uint32_t dx_a = dxs_a + (uint32_t)(slot * GN + h * 64 + c16 * 16) * 4u;
        const uint32_t sx_a = sxs_a + (uint32_t)(slot * GN + h * 64 + c16 * 16) * 4u;
        float2 mw2 = make_float2(0.f, 0.f);
        if constexpr (kMin) {
          const float mw = lds_f32(mws_a + (uint32_t)(slot * GM + row) * 4u);
          mw2 = make_float2(mw, mw);
        }
        const uint32_t(&v)[16] = sv[ch & 1];
#pragma unroll
        for (int c = 0; c < 4; c++) {
          const float4 dx = lds_f4(dx_a + 16u * c);
          float2 &a0 = acc[8 * c16 + 2 * c], &a1 = acc[8 * c16 + 2 * c + 1];
          a0 = ffma2(fmul2(dw2, make_float2(dx.x, dx.y)),
                     make_float2(__uint_as_float(v[4 * c]), __uint_as_float(v[4 * c + 1])), a0);
          a1 = ffma2(fmul2(dw2, make_float2(dx.z, dx.w)),
                     make_float2(__uint_as_float(v[4 * c + 2]), __uint_as_float(v[4 * c + 3])), a1);
          if constexpr (kMin) {
            const float4 sx = lds_f4(sx_a + 16u * c);
            a0 = ffma2(mw2, make_float2(sx.x, sx.y), a0);
            a1 = ffma2(mw2, make_float2(sx.z, sx.w), a1);
          }
        }
        tmem_ld_wait();                          // chunk ch + 1 has landed
        if (ch == 6) {                           // the last load of this TMEM stage is in registers
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&s_empty[sr]);
        }
      }
    }
    // epilogue: C[n * ldc + m]; lanes = consecutive m
    const int m = m0 + row;
    if (m < p.M) {
#pragma unroll
      for (int c = 0; c < 32; c++) {
        const int n = n0 + h * 64 + 2 * c;
        if (n < p.N) p.c[(int64_t)n * p.ldc + m] = acc[c].x;
        if (n + 1 < p.N) p.c[(int64_t)(n + 1) * p.ldc + m] = acc[c].y;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 512);
}

// ---- host side ------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled encode_fn() {
  static PFN_encodeTiled fn = [] {
    void *f = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qr) != cudaSuccess ||
        qr != cudaDriverEntryPointSuccess)
      f = nullptr;
    return (PFN_encodeTiled)f;
  }();
  return fn;
}

int make_map_2d(CUtensorMap *map, CUtensorMapDataType dt, int elem_bytes, const void *base, uint64_t dim0,
                       uint64_t dim1, uint64_t stride1_bytes, uint32_t box0, uint32_t box1, CUtensorMapSwizzle sw) {
  PFN_encodeTiled fn = encode_fn();
  if (!fn) {
    set_error("gemm: cuTensorMapEncodeTiled is not available");
    return LAMMCU_ERR_CUDA;
  }
  cuuint64_t dims[2] = {dim0, dim1};
  cuuint64_t strides[1] = {stride1_bytes};
  cuuint32_t box[2] = {box0, box1};
  cuuint32_t estr[2] = {1, 1};
  (void)elem_bytes;
  CUresult r = fn(map, dt, 2, const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("gemm: cuTensorMapEncodeTiled failed (%d)", (int)r);
    return LAMMCU_ERR_CUDA;
  }
  return LAMMCU_OK;
}

template <int WT>
static int launch_gemm_tc(const CUtensorMap &tm_w, const CUtensorMap &tm_x, const GemmParams &p, cudaStream_t st) {
  using L = GemmSmem<WT>;
  constexpr int smem = L::kTotal + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<WT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) {
      set_error("gemm: cudaFuncSetAttribute(%d): %s", smem, cudaGetErrorString(e));
      return LAMMCU_ERR_CUDA;
    }
    attr_set = true;
  }
  const int tiles = ((p.M + GM - 1) / GM) * ((p.N + GN - 1) / GN);
  gemm_tc_kernel<WT><<<tiles, kGemmThreads, smem, st>>>(tm_w, tm_x, p);
  g_launch_count++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("gemm launch (type %d): %s", WT, cudaGetErrorString(e));
    return LAMMCU_ERR_CUDA;
  }
  return LAMMCU_OK;
}

bool gemm_tc_supported(int w_type, const void *w, int64_t w_stride, int64_t M, int64_t N, int64_t K) {
  if (w_type != LAMMCU_TYPE_Q4_0 && w_type != LAMMCU_TYPE_Q4_1 && w_type != LAMMCU_TYPE_Q5_0 &&
      w_type != LAMMCU_TYPE_Q5_1 && w_type != LAMMCU_TYPE_Q8_0)
    return false;
  if (K <= 0 || K % 256 || M <= 0 || N <= 0 || M > 0x3fffffff || N > 0x3fffffff) return false;
  return (((uintptr_t)w | (uintptr_t)w_stride) & 15) == 0;
}

size_t gemm_ws_bytes(int64_t N, int64_t K) {
  const int64_t npad = (N + GN - 1) / GN * GN, kb = K / 32;
  return (size_t)(npad * K * 2 + 2 * kb * npad * 4 + 1024);
}

// xq: quantized activations (ggml blocks, row stride xq_stride); ws: gemm_ws_bytes() of scratch
int launch_gemm_tc_any(int w_type, const void *w, int64_t w_stride, const void *xq, int64_t xq_stride, float *c,
                       int64_t ldc, int64_t M, int64_t N, int64_t K, void *ws, cudaStream_t st) {
  const TypeInfo ti = type_info(w_type);
  const int64_t npad = (N + GN - 1) / GN * GN, kb = K / 32;
  uint8_t *base = (uint8_t *)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
  void *x16 = base;
  float *dxT = (float *)(base + npad * K * 2);
  float *sxT = dxT + kb * npad;
  const int perm = (w_type == LAMMCU_TYPE_Q8_0) ? 0 : 1;
  int rc = launch_gemm_prep(ti.vec_dot, perm, xq, xq_stride, x16, dxT, sxT, N, npad, K, st);
  if (rc) return rc;
  CUtensorMap tm_w, tm_x;
  // raw weights as uint16 [M][row_bytes / 2], box = 8 blocks x 128 rows
  rc = make_map_2d(&tm_w, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, w, (uint64_t)(kb * ti.size / 2), (uint64_t)M,
                   (uint64_t)w_stride, (uint32_t)(4 * ti.size), GM, CU_TENSOR_MAP_SWIZZLE_NONE);
  if (rc) return rc;
  rc = make_map_2d(&tm_x, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, x16, (uint64_t)K, (uint64_t)npad, (uint64_t)K * 2, GK, GN,
                   CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  GemmParams p;
  p.dxT = dxT;  p.sxT = sxT;  p.c = c;  p.ldc = ldc;
  p.M = (int)M;  p.N = (int)N;  p.Npad = (int)npad;  p.K = (int)K;
  static const int dbg = [] { const char *e = getenv("LAMM_GEMM_DEBUG"); return e ? atoi(e) : 0; }();
  p.debug = dbg;
  switch (w_type) {
  case LAMMCU_TYPE_Q4_0: return launch_gemm_tc<LAMMCU_TYPE_Q4_0>(tm_w, tm_x, p, st);
  case LAMMCU_TYPE_Q4_1: return launch_gemm_tc<LAMMCU_TYPE_Q4_1>(tm_w, tm_x, p, st);
  case LAMMCU_TYPE_Q5_0: return launch_gemm_tc<LAMMCU_TYPE_Q5_0>(tm_w, tm_x, p, st);
  case LAMMCU_TYPE_Q5_1: return launch_gemm_tc<LAMMCU_TYPE_Q5_1>(tm_w, tm_x, p, st);
  default:               return launch_gemm_tc<LAMMCU_TYPE_Q8_0>(tm_w, tm_x, p, st);
  }
}

// Prefill entry: tensor-core GEMM where it applies, else the column-grouped GEMV (correct for
// every type / alignment; re-reads W once per 8 columns).  Two tensor-core kernels: the
// scaled-operand mainloop of lamm_gemm_so.cu (default) and the exact per-block drain above
// (lammcu_set_gemm_exact(1): integer block sums bit-identical to the reference's sumi).
static int g_gemm_exact = [] { const char *e = getenv("LAMM_GEMM_EXACT"); return e ? atoi(e) : 0; }();
void set_gemm_exact(int on) { g_gemm_exact = on ? 1 : 0; }
int get_gemm_exact() { return g_gemm_exact; }

// Q2_K / F32 weights: dequantize-once + TMA-fed GEMM (lamm_gemm_so.cu, direct variant)
static bool gemm_direct_supported(int w_type, const void *w, int64_t w_stride, const void *xq, int64_t xq_stride,
                                  int64_t M, int64_t N, int64_t K) {
  if (w_type != LAMMCU_TYPE_Q2_K && w_type != LAMMCU_TYPE_F32 && !is_kquant_ext(w_type)) return false;
  if (K <= 0 || K % 256 || M <= 0 || N <= 0 || M > 0x3fffffff || N > 0x3fffffff) return false;
  const uintptr_t al = (w_type == LAMMCU_TYPE_F32) ? 15 : 3;
  // Q3_K / Q6_K blocks (110 / 210 bytes) are only 2-byte aligned: their dequantizer reads 16-bit words
  const uintptr_t alw = (w_type == LAMMCU_TYPE_Q3_K || w_type == LAMMCU_TYPE_Q6_K) ? 1 : al;
  return (((uintptr_t)w | (uintptr_t)w_stride) & alw) == 0 && (((uintptr_t)xq | (uintptr_t)xq_stride) & al) == 0;
}

int launch_gemm(int w_type, const void *w, int64_t w_stride, const void *xq, int64_t xq_stride, float *c,
                int64_t ldc, int64_t M, int64_t N, int64_t K, cudaStream_t st) {
  // Q2_K: the int8 fold is exact AND the fastest kernel, whatever the mode (LAMM_GEMM_Q2K_I8=0: the round-1
  // dequantize-once path, for comparison)
  static const int q2k_i8 = [] { const char *e = getenv("LAMM_GEMM_Q2K_I8"); return e ? atoi(e) : 1; }();
  if (q2k_i8 && gemm_x_supported(w_type, w, w_stride, M, N, K) && (((uintptr_t)xq | (uintptr_t)xq_stride) & 3) == 0) {
    void *ws = gemm_workspace(gemm_x_ws_bytes(N, K));
    if (!ws) return LAMMCU_ERR_NOMEM;
    return launch_gemm_x(w_type, w, w_stride, xq, xq_stride, c, ldc, M, N, K, ws, st);
  }
  if (w_type == LAMMCU_TYPE_F16 && gemm_f16_supported(w, w_stride, xq, xq_stride, M, N, K)) {
    if (gemm_f16_pair_supported(M, N, K)) return launch_gemm_f16_pair(w, w_stride, xq, xq_stride, c, ldc, M, N, K, st);
    return launch_gemm_f16(w, w_stride, xq, xq_stride, c, ldc, M, N, K, st);
  }
  if (gemm_tc_supported(w_type, w, w_stride, M, N, K)) {
    if (!g_gemm_exact) {
      void *ws = gemm_workspace(gemm_so_ws_bytes(M, N, K));
      if (!ws) return LAMMCU_ERR_NOMEM;
      return launch_gemm_so(w_type, w, w_stride, xq, xq_stride, c, ldc, M, N, K, ws, st);
    }
    void *ws = gemm_workspace(gemm_ws_bytes(N, K));
    if (!ws) return LAMMCU_ERR_NOMEM;
    return launch_gemm_tc_any(w_type, w, w_stride, xq, xq_stride, c, ldc, M, N, K, ws, st);
  }
  if (!g_gemm_exact && gemm_direct_supported(w_type, w, w_stride, xq, xq_stride, M, N, K)) {
    void *ws = gemm_workspace(gemm_direct_ws_bytes(M, N, K));
    if (!ws) return LAMMCU_ERR_NOMEM;
    return launch_gemm_direct(w_type, w, w_stride, xq, xq_stride, c, ldc, M, N, K, ws, st);
  }
  return launch_gemv(w_type, w, w_stride, xq, xq_stride, c, ldc, M, N, K, st);
}

}  // namespace lamm
