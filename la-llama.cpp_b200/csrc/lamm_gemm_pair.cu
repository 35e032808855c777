// lamm_gemm_pair.cu -- prefill GEMM on CTA PAIRS (tcgen05 cta_group::2): two CTAs of one cluster (the two SMs of a
// TPC) compute a 256 x 256 tile with ONE MMA stream issued by the leader CTA.
//
// Why: a single-CTA M128 x N128 x K16 MMA reads 8 KB of shared memory per 64 cycles = the SM's whole 128 B/clk of
// shared-memory bandwidth, and a 128 x 128 tile pulls 32 KB from L2 per 256 tensor cycles (the F16 kernel of
// lamm_gemm_so.cu stops at ~1050 TFLOPS = 47 % of the measured f16 peak for that reason).  In a pair each SM keeps
// its own 128 rows of A and HALF of B (128 of the 256 activation rows): per 128 cycles of tensor time an SM reads
// 8 KB for twice the flops, and the pair loads 64 KB per 256 x 256 x 64 step instead of 128 KB.
//
// Layout per CTA (rank r of the pair): A tile = weight rows [m0 + 128 r, +128), B tile = activation rows
// [n0 + 128 r, +128); the accumulator rows of CTA r (its 128 weight rows x all 256 activation columns) live in its
// own TMEM, so each CTA's epilogue drains and stores its own half.  Barriers: the LEADER's `full` barrier counts the
// bytes of both CTAs' TMA loads (the peer's loads complete_tx on it through .cta_group::2 + mapa); `empty` and
// `acc_full` are armed in both CTAs by one multicast tcgen05.commit.
#include "lamm_tc.cuh"

#include <cstdlib>

namespace lamm {

struct PairParams {
  float *c;
  int64_t ldc;
  int M, N, K;
  float *cp[LAMMCU_MAX_RANKS];      // row-sharded multi-GPU prefill: full [N][ldc] result of every rank (see SoParams)
  int n_peer;
};

constexpr int kPairStages = 6;
constexpr int kPairStageBytes = 2 * GM * GK * 2;        // A 16 KB + B 16 KB per CTA
constexpr int kPairSmem = kPairStages * kPairStageBytes + 256 + 1024;

template <bool PEERS>
__device__ __forceinline__ void pair_store(const PairParams &p, int64_t off, float v) {
  if constexpr (!PEERS) {
    p.c[off] = v;
  } else {
#pragma unroll
    for (int g = 0; g < LAMMCU_MAX_RANKS; g++)
      if (g < p.n_peer) p.cp[g][off] = v;
  }
}

// F16 weights x fp16 activations: one MMA per K step, no rescaling.
// PERSISTENT (as gemm_so_kernel): a pair walks tiles cl, cl + ncl, ...; the TMA ring and its barrier phases run on across
// tiles and the accumulator is double-buffered in TMEM (2 x 256 columns), so a tile's drain and the next tile's first
// loads hide behind the mainloops.
template <bool PEERS>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(192, 1)
gemm_f16_pair_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_x, const PairParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + kPairStages * kPairStageBytes);
  uint64_t *full = bars, *empty = bars + kPairStages, *acc_full = bars + 2 * kPairStages, *acc_empty = acc_full + 2;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(acc_empty + 2);
  static_assert((2 * kPairStages + 4) * 8 + 4 <= 256, "barrier area");
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int cl = (int)(blockIdx.x >> 1), ncl = (int)(gridDim.x >> 1);
  const int tiles_n = (p.N + 255) / 256, n_tiles = ((p.M + 255) / 256) * tiles_n;
  const int n_stages = p.K / GK;
  // tile tl: this CTA's weight rows (A) = its accumulator rows; first accumulator column; its half of the activation rows (B)
  auto tile_m0 = [&](int tl) { return (tl / tiles_n) * 256 + (int)rank * 128; };
  auto tile_n0 = [&](int tl) { return (tl % tiles_n) * 256; };
  if (threadIdx.x == 0) {
    for (int i = 0; i < kPairStages; i++) {
      mbar_init(&full[i], 1);                          // the leader's own arrive.expect_tx (bytes of both CTAs)
      mbar_init(&empty[i], 1);                         // multicast MMA commit
    }
    for (int i = 0; i < 2; i++) {
      mbar_init(&acc_full[i], 1);                      // multicast MMA commit: accumulator i holds a finished tile
      mbar_init(&acc_empty[i], 8);                     // the four epilogue warps of both CTAs have drained it (leader's copy counts)
    }
    fence_barrier_init();
    prefetch_tmap(&tm_a);
    prefetch_tmap(&tm_x);
  }
  __syncthreads();
  cluster_sync_all();                                  // barriers of both CTAs exist before anything remote touches them
  if (warp == 1) tmem_alloc_pair(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (warp == 0) {
    if (lane == 0) {
      int gst = 0;                                     // ring stages so far, over all tiles
      for (int tl = cl; tl < n_tiles; tl += ncl) {
        const int m0 = tile_m0(tl), nb = tile_n0(tl) + (int)rank * 128;
        for (int st = 0; st < n_stages; st++, gst++) {
          const int s = gst % kPairStages;
          uint8_t *stage = smem + s * kPairStageBytes;
          mbar_wait(&empty[s], ((uint32_t)(gst / kPairStages) & 1u) ^ 1u);
          if (rank == 0) mbar_arrive_expect_tx(&full[s], (uint32_t)(2 * kPairStageBytes));
          const uint32_t bar = mapa_rank(smem_u32(&full[s]), 0u);
          tma_load_2d_pair(stage, &tm_a, st * GK, m0, bar);
          tma_load_2d_pair(stage + GM * GK * 2, &tm_x, st * GK, nb, bar);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && rank == 0) {
      int gst = 0, it = 0;
      for (int tl = cl; tl < n_tiles; tl += ncl, it++) {
        const int ab = it & 1;                         // accumulator buffer of this tile
        mbar_wait_cluster(&acc_empty[ab], ((uint32_t)(it >> 1) & 1u) ^ 1u);
        tc_fence_after();
        const uint32_t tacc = tmem_base + (uint32_t)(ab * 256);
        for (int st = 0; st < n_stages; st++, gst++) {
          const int s = gst % kPairStages;
          mbar_wait(&full[s], (uint32_t)(gst / kPairStages) & 1u);
          tc_fence_after();
          const uint32_t base = smem_u32(smem + s * kPairStageBytes);
          const uint64_t a = make_sw128_desc(base), b = make_sw128_desc(base + GM * GK * 2);
#pragma unroll
          for (int kk = 0; kk < 4; kk++)
            umma_f16_pair(tacc, a + (uint64_t)(kk * 2), b + (uint64_t)(kk * 2), kIdescF16Pair, (st | kk) ? 1u : 0u);
          umma_commit_pair(&empty[s]);
        }
        umma_commit_pair(&acc_full[ab]);
      }
    }
  } else {
    // epilogue warps 2..5: TMEM lane quarter = warp & 3; 256 accumulator columns = activation rows n0 .. n0 + 255
    const int q4 = warp & 3, row = q4 * 32 + lane;
    int it = 0;
    for (int tl = cl; tl < n_tiles; tl += ncl, it++) {
      const int ab = it & 1, m = tile_m0(tl) + row, n0 = tile_n0(tl);
      mbar_wait(&acc_full[ab], (uint32_t)(it >> 1) & 1u);
      tc_fence_after();
      const uint32_t tbase = tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(ab * 256);
#pragma unroll 2
      for (int c16 = 0; c16 < 16; c16++) {
        uint32_t v[16];
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
              "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
            : "r"(tbase + (uint32_t)(c16 * 16))
            : "memory");
        tmem_ld_wait();
        if (m < p.M) {
#pragma unroll
          for (int i = 0; i < 16; i++) {
            const int n = n0 + c16 * 16 + i;
            if (n < p.N) pair_store<PEERS>(p, (int64_t)n * p.ldc + m, __uint_as_float(v[i]));
          }
        }
      }
      // the buffer goes back to the MMA issuer: every lane's tcgen05.ld has completed (wait::ld above)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (rank == 0) mbar_arrive(&acc_empty[ab]);
        else mbar_arrive_cluster(mapa_rank(smem_u32(&acc_empty[ab]), 0u));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                                  // the peer's shared memory and TMEM outlive the leader's MMAs
  if (warp == 1) tmem_dealloc_pair(tmem_base, 512);
}

bool gemm_f16_pair_supported(int64_t M, int64_t N, int64_t K) {
  const char *e = getenv("LAMM_GEMM_PAIR");              // per call: tests compare both forms in one process
  const int on = e ? atoi(e) : 1;
  return on && N > 128 && M > 128 && K % GK == 0;
}

int launch_gemm_f16_pair(const void *w, int64_t w_stride, const void *xq, int64_t xq_stride, float *c, int64_t ldc,
                         int64_t M, int64_t N, int64_t K, cudaStream_t st) {
  CUtensorMap tm_a, tm_x;
  int rc;
  if ((rc = make_map_2d(&tm_a, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, w, (uint64_t)K, (uint64_t)M, (uint64_t)w_stride, GK, GM,
                        CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  if ((rc = make_map_2d(&tm_x, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, xq, (uint64_t)K, (uint64_t)N, (uint64_t)xq_stride, GK, GN,
                        CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  PairParams p = {};
  p.c = c;  p.ldc = ldc;  p.M = (int)M;  p.N = (int)N;  p.K = (int)K;
  if (const GemmPeers *peers = gemm_peers_take()) {
    p.n_peer = peers->n;
    for (int g = 0; g < peers->n; g++) p.cp[g] = peers->c[g];
  }
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_f16_pair_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kPairSmem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_f16_pair_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kPairSmem);
    if (e != cudaSuccess) { set_error("gemm_f16_pair: cudaFuncSetAttribute(%d): %s", kPairSmem, cudaGetErrorString(e)); return LAMMCU_ERR_CUDA; }
    attr_set = true;
  }
  const int pairs = (int)(((M + 255) / 256) * ((N + 255) / 256)), slots = sm_count_cached() / 2;
  const int grid = 2 * (pairs < slots ? pairs : slots);                // persistent: one pair per TPC walks the tiles
  if (p.n_peer) gemm_f16_pair_kernel<true><<<grid, 192, kPairSmem, st>>>(tm_a, tm_x, p);
  else gemm_f16_pair_kernel<false><<<grid, 192, kPairSmem, st>>>(tm_a, tm_x, p);
  g_launch_count++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("gemm_f16_pair launch: %s", cudaGetErrorString(e)); return LAMMCU_ERR_CUDA; }
  return LAMMCU_OK;
}

}  // namespace lamm
