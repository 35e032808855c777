// lamm_gemm_so.cu -- prefill GEMM, "scaled operand" variant: the per-block scales ride INSIDE the
// tensor-core operands, so the accumulator never leaves TMEM during the K loop.
//
// Replaces LAMMImpl<T>::matmul_simd_block + lamm_simd_block_kernel<4,4> (src/lamm_impl.hpp:90-147,
// src/lamm_kernel_*.hpp) for many activation rows, like lamm_gemm.cu, but trades the bit-exact
// integer block sums of that kernel for a plain GEMM mainloop:
//
//   C[n, m] = sum_k (d_w[m,b(k)] * q_w[m,k]) * (d_x[n,b(k)] * q_x[n,k])   (+ sum_b m_w[m,b] * s_x[n,b])
//
//   A' = d_w * q_w * 2^F_m   : (F_m: per-row lift to [2^14, 2^15), lamm_gemm_prep.cu) split EXACTLY into two fp16 terms hi + lo (hi = rn(d q), lo = fma(d, q, -hi))
//   B' = d_x * q_x * 2^-E_n  : exact in f32 (11 + 8 bits), row-normalised by a power of two and split
//                              into two fp16 terms by lamm_gemm_prep.cu (residue <= 2^-22)
//   A'B' ~= Ah*Bh + Ah*Bl + Al*Bh   -- three kind::f16 MMAs per K step, f32 accumulate in TMEM; the
//   dropped Al*Bl term is <= 2^-22 relative per product (mixed fp16 x bf16 operands are not accepted by
//   the hardware, hence fp16 on both sides and the exact power-of-two scalings), against the
//   path's tolerance of 1e-5 * sqrt(K) * max|C| (>= 1.6e-4): tests/test_gpu_gemm.py checks both modes.
//   The m_w * s_x term of Q4_1 / Q5_1 (src/lamm_kernel_q4_1.hpp:24, s_x is the fp16 field of
//   block_q8_1 and must be used as stored) is a rank-(K/32) update: the unpack threads collect the
//   m_w of 64 consecutive blocks in an extra A tile (fp16, exact), lamm_gemm_prep.cu lays the s_x out
//   the same way (fp16, exact), and the MMA issuer folds the pair in as one extra stage per 64 blocks.
//
// PERSISTENT (round 2): a CTA (or CTA pair, below) walks tiles cl, cl + ncl, ... -- the rings and barrier phases run on
// across tiles and the accumulator is double-buffered in TMEM, so tile i's epilogue and tile i + 1's prologue hide
// behind the mainloops (13B Q4_1 N = 2048: 490 -> 519 TOPS, profiles/r2an).
// One CTA = 128 weight rows x 128 activation rows per tile, 16 warps:
//   warp 0      TMA producer: raw packed weight blocks (2-D box over bytes, 8 blocks per row), the two
//               fp16 activation tiles (SWIZZLE_128B), the s_x tiles of the extra stages
//   warp 1      MMA issuer (one thread): per 64-K stage 4 x 3 tcgen05.mma M128 N128 K16 into ONE
//               128-column accumulator; tcgen05.commit frees the stage; a final commit ends the tile
//   warp 2      TMEM allocation
//   warps 4-11  weight unpack (thread = row x one 32-block of the stage): nibbles / 5th bits -> small
//               ints (0x6400 magic) * d_w -> fp16 hi / lo in the UMMA K-major SWIZZLE_128B layout
//   warps 12-15 epilogue: tcgen05.ld the accumulator (thread = TMEM lane = weight row), store C coalesced
#include "lamm_tc.cuh"

#include <cstdlib>

namespace lamm {

// ring depths: unpacked weight tiles (hi + lo) / activation tiles (hi + lo): 2 / 3 in both forms.  Measured on the pair
// form at 13B Q4_1 N = 2048 (profiles/r2y): 2 / 3 490 TOPS, 3 / 2 465, 3 / 3 (fits Q4_0 only) 496 vs 498 -- the third B
// stage (TMA latency) matters, a third A stage does not; two unpack groups of 8 warps on alternate stages (768
// threads) were slower, 450 (profiles/r2x).  Barrier slots below assume <= 3 stages of either kind.
constexpr double kSoPairGain = 1.4;            // per-tile speed of the pair form over the single form (519 vs 373 TOPS, profiles/r2an)
#ifndef LAMM_SO_PAIR_NA
#define LAMM_SO_PAIR_NA 2
#endif
#ifndef LAMM_SO_PAIR_NB
#define LAMM_SO_PAIR_NB 3
#endif
// ATM (pair form only): the unpacked weight tiles live in TENSOR MEMORY (3 stages x 64 columns behind ONE 256-column
// accumulator) instead of shared memory, which leaves room for a deeper activation ring
#ifndef LAMM_SO_ATM_NB
#define LAMM_SO_ATM_NB 4
#endif
constexpr int kSoAtmStages = 3;
template <bool PAIR, bool ATM = false> struct SoRing {
  static constexpr int A = ATM ? kSoAtmStages : PAIR ? LAMM_SO_PAIR_NA : 2, B = ATM ? LAMM_SO_ATM_NB : PAIR ? LAMM_SO_PAIR_NB : 3;
};
constexpr int kSoMaxA = 3, kSoMaxB = 6;       // barrier slots
constexpr int kSoThreads = 512;
constexpr int kTileBytes = GM * GK * 2;        // 16 KB: one 128 x 64 fp16 operand tile

struct SoParams {
  const float *colscale;      // [Npad]: 2^E_n, undoes the activation rows' power-of-two normalisation
  // [Mpad]: 2^F_m, an exact power of two per WEIGHT row that lifts the row's largest |d_w q_w| (and |m_w|) to
  // 2^14 <= . < 2^15 before the fp16 hi/lo split, so that no fp16 scale -- 6e-8 .. 65504 -- overflows the split or
  // loses its lo term to subnormals; the epilogue multiplies row m by 2^-F_m (wrow_scale_kernel, lamm_gemm_prep.cu)
  const float *rowscale;
  float *c;
  int64_t ldc;
  int M, N, Npad, K;
  // row-sharded multi-GPU prefill: the epilogue stores every C element into the full [N][ldc] result
  // of EVERY rank (peer-mapped, already offset to this rank's first column) instead of `c`
  float *cp[LAMMCU_MAX_RANKS];
  int n_peer;
};

static thread_local const GemmPeers *t_peers = nullptr;
static thread_local bool t_peers_used = false;
void gemm_set_peers(const GemmPeers *peers) { t_peers = peers;  t_peers_used = false; }
bool gemm_peers_used() { return t_peers_used; }
const GemmPeers *gemm_peers_take() {
  if (t_peers) t_peers_used = true;
  return t_peers;
}
static void fill_peers(SoParams &p) {
  p.n_peer = 0;
  for (int g = 0; g < LAMMCU_MAX_RANKS; g++) p.cp[g] = nullptr;
  if (t_peers) {
    p.n_peer = t_peers->n;
    for (int g = 0; g < t_peers->n; g++) p.cp[g] = t_peers->c[g];
    t_peers_used = true;
  }
}

// one accumulator element to its destination(s)
// PEERS is a template parameter of the kernels: the single-GPU epilogue must stay the plain coalesced store
// (a run-time branch + dynamically indexed pointer array in the 16-wide store loop cost 25 % of the GEMM)
template <bool PEERS>
__device__ __forceinline__ void store_out(const SoParams &p, int64_t idx, float v) {
  if constexpr (!PEERS) {
    p.c[idx] = v;
  } else {
#pragma unroll
    for (int g = 0; g < LAMMCU_MAX_RANKS; g++)
      if (g < p.n_peer) p.cp[g][idx] = v;
  }
}

template <int WT, int kSoAStagesAll, int kSoBStages, bool ATM = false> struct SoSmem {
  static constexpr int kSoAStages = ATM ? 0 : kSoAStagesAll;         // A tiles in shared memory
  static constexpr int BS = type_info(WT).size;
  static constexpr int kRawChunkBytes = 8 * BS;                     // 8 blocks = 256 K elements = 4 stages
  // Row pitch of the raw chunk in shared memory.  Thread = row reads its row with ld.shared.v4, so the pitch decides the
  // bank conflicts: 144 / 176 / 272 bytes (Q4_0, Q5_0, Q8_0) are conflict-free, but 160 (Q4_1) is 2-way and 192 (Q5_1)
  // 4-way (16 wavefronts per load; ncu counted 19 M conflicts per 13B launch).  Those two get a 16-byte pad: the TMA box is
  // simply 16 bytes wider (it reads into the next chunk, or gets zeros past the row's end), 176 / 208-byte pitch.
  static constexpr int kRawRowBytes = kRawChunkBytes + ((kRawChunkBytes / 16) % 2 == 0 ? 16 : 0);
  static constexpr int kRawBytes = GM * kRawRowBytes;
  static constexpr bool kMin = (WT == LAMMCU_TYPE_Q4_1 || WT == LAMMCU_TYPE_Q5_1);
  static constexpr int kAhOff = 0;
  static constexpr int kAlOff = kAhOff + kSoAStages * kTileBytes;
  static constexpr int kBhOff = kAlOff + kSoAStages * kTileBytes;
  static constexpr int kBlOff = kBhOff + kSoBStages * kTileBytes;
  static constexpr int kEhOff = kBlOff + kSoBStages * kTileBytes;   // m_w of 64 blocks (kMin only)
  static constexpr int kRawOff = kEhOff + (kMin ? kTileBytes : 0);
  static constexpr int kBarOff = kRawOff + (kRawBytes + 127) / 128 * 128;
  static constexpr int kTotal = kBarOff + 256;
  static_assert(kSoAStagesAll <= kSoMaxA && kSoBStages <= kSoMaxB, "barrier slots");
};

// 8 fp16 small integers (one 16-byte chunk of unpack_block) * d -> 8 fp16 hi + 8 fp16 lo, EXACT:
// hi = rn(q * d) in fp16, lo = fma(q, d, -hi) is the rounding residual (<= 8 significant bits, exactly
// representable as long as it stays above the fp16 subnormal step 2^-24).  d2 = (d, d) carries the
// per-row lift 2^F_m (SoParams::rowscale), which puts the row's largest |d q| in [2^14, 2^15): no fp16 scale
// overflows the split and only blocks > 2^24 below the row's largest lose bits; the epilogue takes it out.
__device__ __forceinline__ void scale_split8(const uint4 &q, __half2 d2, uint4 &hi, uint4 &lo) {
  const uint32_t in[4] = {q.x, q.y, q.z, q.w};
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const __half2 qq = *reinterpret_cast<const __half2 *>(&in[i]);
    const __half2 hh = __hmul2(qq, d2);
    const __half2 ll = __hfma2(qq, d2, __hneg2(hh));
    h[i] = *reinterpret_cast<const uint32_t *>(&hh);
    l[i] = *reinterpret_cast<const uint32_t *>(&ll);
  }
  hi = make_uint4(h[0], h[1], h[2], h[3]);
  lo = make_uint4(l[0], l[1], l[2], l[3]);
}

// Block JB of the raw chunk (j = JB & 1: first / second 32-block of its 64-K stage) of the thread's row r.
template <int WT, int JB, int NWR>
__device__ __forceinline__ void unpack_block_so(const uint32_t (&raw)[NWR], uint32_t ah_stage, uint32_t al_stage, int r,
                                                float rs) {
  constexpr int j = JB & 1;
  const uint32_t row_off = (uint32_t)(r >> 3) * 1024u + (uint32_t)(r & 7) * 128u;
  uint4 out[4];
  float d, m;
  unpack_block<WT, JB, NWR>(raw, out, d, m);
  const __half2 d2 = __float2half2_rn(d * rs);      // d is an fp16 value, rs a power of two: exact unless the block's
                                                    // scale is > 2^24 below the row's largest (then it is noise)
#pragma unroll
  for (int w = 0; w < 4; w++) {
    uint4 hi, lo;
    scale_split8(out[w], d2, hi, lo);
    const uint32_t c = ((uint32_t)(4 * j + w) ^ (uint32_t)(r & 7)) * 16u;
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(ah_stage + row_off + c), "r"(hi.x), "r"(hi.y),
                 "r"(hi.z), "r"(hi.w)
                 : "memory");
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(al_stage + row_off + c), "r"(lo.x), "r"(lo.y),
                 "r"(lo.z), "r"(lo.w)
                 : "memory");
  }
}

// ATM: the same block into the thread's TMEM lane: 16 words hi at column 16 j of the stage, 16 words lo 32 columns on
template <int WT, int JB, int NWR>
__device__ __forceinline__ void unpack_block_tm(const uint32_t (&raw)[NWR], uint32_t t_stage, float rs) {
  constexpr int j = JB & 1;
  uint4 out[4];
  float d, m;
  unpack_block<WT, JB, NWR>(raw, out, d, m);
  const __half2 d2 = __float2half2_rn(d * rs);
  uint32_t h[16], l[16];
#pragma unroll
  for (int w = 0; w < 4; w++) {
    uint4 hi, lo;
    scale_split8(out[w], d2, hi, lo);
    h[4 * w] = hi.x;  h[4 * w + 1] = hi.y;  h[4 * w + 2] = hi.z;  h[4 * w + 3] = hi.w;
    l[4 * w] = lo.x;  l[4 * w + 1] = lo.y;  l[4 * w + 2] = lo.z;  l[4 * w + 3] = lo.w;
  }
  tmem_st16(t_stage + (uint32_t)(16 * j), h);
  tmem_st16(t_stage + (uint32_t)(32 + 16 * j), l);
}

// the 8 m_w of the raw chunk (Q4_1 / Q5_1: fp16 at byte 2 of every block) times the row's lift S: one 16-byte
// chunk of the extra A tile (m_w is an fp16 field, the power-of-two scale keeps it exact)
template <int WT, int NWR>
__device__ __forceinline__ uint4 chunk_mins(const uint32_t (&raw)[NWR], float S) {
  constexpr int BS = type_info(WT).size;
  const __half2 a = __floats2half2_rn(raw_half<0 * BS + 2, NWR>(raw) * S, raw_half<1 * BS + 2, NWR>(raw) * S);
  const __half2 b = __floats2half2_rn(raw_half<2 * BS + 2, NWR>(raw) * S, raw_half<3 * BS + 2, NWR>(raw) * S);
  const __half2 c = __floats2half2_rn(raw_half<4 * BS + 2, NWR>(raw) * S, raw_half<5 * BS + 2, NWR>(raw) * S);
  const __half2 d = __floats2half2_rn(raw_half<6 * BS + 2, NWR>(raw) * S, raw_half<7 * BS + 2, NWR>(raw) * S);
  return make_uint4(*reinterpret_cast<const uint32_t *>(&a), *reinterpret_cast<const uint32_t *>(&b),
                    *reinterpret_cast<const uint32_t *>(&c), *reinterpret_cast<const uint32_t *>(&d));
}

// PAIR: the CTA-pair form (cta_group::2, launched as clusters of 2; see lamm_gemm_pair.cu for the protocol): each CTA
// still unpacks its own 128 weight rows and loads 128 activation rows, but the pair's leader issues ONE M256 x N256
// MMA stream over both, so per unit of work an SM reads half the shared memory and pulls half the activation bytes
// from L2 (the 128 x 128 form is bound by exactly those two: 125 B/clk of MMA operand reads + 42 B/clk of B tiles).
template <int WT, bool PEERS, bool PAIR, bool ATM = false>
__global__ void __launch_bounds__(kSoThreads, 1)
gemm_so_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_xh,
               const __grid_constant__ CUtensorMap tm_xl, const __grid_constant__ CUtensorMap tm_sh, const SoParams p) {
  static_assert(PAIR || !ATM, "A in tensor memory: pair form only");
  constexpr int kSoAStages = SoRing<PAIR, ATM>::A, kSoBStages = SoRing<PAIR, ATM>::B;
  using L = SoSmem<WT, kSoAStages, kSoBStages, ATM>;
  constexpr bool kMin = L::kMin;
  constexpr int NWR = 2 * L::BS;                                    // raw words per row chunk
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + L::kBarOff);
  uint64_t *full_a = bars, *empty_a = full_a + kSoMaxA, *full_b = empty_a + kSoMaxA, *empty_b = full_b + kSoMaxB,
           *raw_full = empty_b + kSoMaxB, *raw_empty = raw_full + 1, *acc_full = raw_empty + 1, *acc_empty = acc_full + 2,
           *ext_full = acc_empty + 2, *ext_empty = ext_full + 1;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(ext_empty + 1);
  static_assert((2 * kSoMaxA + 2 * kSoMaxB + 8 + 1) * 8 <= 256, "barrier area");

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int TN = PAIR ? 256 : GN;                                // accumulator columns per CTA
  constexpr int TM = PAIR ? 256 : GM;                                // weight rows per tile (pair: 128 per CTA)
  // accumulators in TMEM: two (tile i drains while tile i + 1 accumulates), or ONE followed by the A stages (ATM)
  constexpr int kAccBufs = ATM ? 1 : 2;
  constexpr uint32_t kTmemCols = ATM ? 512u : (uint32_t)(2 * TN), kTmemA = (uint32_t)TN;
  static_assert(!ATM || TN + kSoAStages * 64 <= 512, "tensor memory budget");
  const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
  const bool leader = (rank == 0u);
  const int n_stages = p.K / GK, n_chunks = n_stages / 4;
  // PERSISTENT: a CTA (pair) walks tiles cl, cl + ncl, ...; rings and barrier phases run on across tiles (global
  // counters), and the accumulator is double-buffered in TMEM (2 x TN columns), so the epilogue of tile i and the
  // prologue of tile i + 1 hide behind the mainloops instead of adding ~10 % per tile.
  const int tiles_n = (p.N + TN - 1) / TN, n_tiles = ((p.M + TM - 1) / TM) * tiles_n;
  const int cl = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x, ncl = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int n_ext = kMin ? (n_chunks + 7) / 8 : 0;                   // extra (m_w x s_x) stages per tile
  // tile t: m0 = this CTA's weight rows = accumulator rows; n0 = first accumulator column; nb = activation rows it loads
  auto tile_m0 = [&](int tl) { return (tl / tiles_n) * TM + (int)rank * GM; };
  auto tile_n0 = [&](int tl) { return (tl % tiles_n) * TN; };

  // Q4_1 / Q5_1: an extra stage (m_w tile x s_x tile) is folded in before main stage 4c for every raw
  // chunk c that completes a group of 64 blocks (c % 8 == 7) or is the last one.  The m_w tile is
  // complete as soon as the unpack warps START chunk c, i.e. four stages before the group's end.
  auto ext_before = [&](int st) -> bool {
    if (!kMin || (st & 3)) return false;
    const int c = st >> 2;
    return (c & 7) == 7 || c == n_chunks - 1;
  };

  if (threadIdx.x == 0) {
    for (int i = 0; i < kSoAStages; i++) {
      mbar_init(&full_a[i], PAIR ? 16 : 8);            // one arrival per unpack warp (PAIR: of both CTAs, on the leader's)
      mbar_init(&empty_a[i], 1);                       // MMA commit
    }
    for (int i = 0; i < kSoBStages; i++) {
      mbar_init(&full_b[i], 1);
      mbar_init(&empty_b[i], 1);
    }
    mbar_init(raw_full, 1);
    mbar_init(raw_empty, 8);
    for (int i = 0; i < kAccBufs; i++) {
      mbar_init(&acc_full[i], 1);                      // MMA commit: accumulator i holds a finished tile
      mbar_init(&acc_empty[i], PAIR ? 8 : 4);          // the epilogue warps (of both CTAs) have drained it
    }
    mbar_init(ext_full, PAIR ? 8 : 4);                 // the four unpack warps that own m_w (j == 0) (PAIR: of both CTAs)
    mbar_init(ext_empty, 1);
    fence_barrier_init();
    prefetch_tmap(&tm_w);
    prefetch_tmap(&tm_xh);
    prefetch_tmap(&tm_xl);
    if (kMin) prefetch_tmap(&tm_sh);
  }
  if constexpr (kMin) {
    // stale shared memory may hold NaN patterns: unwritten columns of the m_w tile meet zero s_x
    const uint32_t e0 = smem_u32(smem + L::kEhOff) + threadIdx.x * 32u;
#pragma unroll
    for (int i = 0; i < 2; i++)
      asm volatile("st.shared.v4.u32 [%0], {%1, %1, %1, %1};" ::"r"(e0 + 16u * i), "r"(0u) : "memory");
    fence_proxy_async();
  }
  if constexpr (PAIR) {
    __syncthreads();
    cluster_sync_all();                                // both CTAs' barriers exist before anything remote arrives
    if (warp == 2) tmem_alloc_pair(tmem_slot, kTmemCols);
  } else {
    if (warp == 2) tmem_alloc(tmem_slot, kTmemCols);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // PAIR: "full" barriers live in the leader; arrivals / TMA completions of the peer go there through mapa
  auto load_b = [&](void *dst, const CUtensorMap *map, int x, int y, uint64_t *bar) {
    if constexpr (PAIR) tma_load_2d_pair(dst, map, x, y, mapa_rank(smem_u32(bar), 0u));
    else tma_load_2d(dst, map, x, y, bar);
  };
  auto arrive_full = [&](uint64_t *bar) {
    if constexpr (PAIR) {
      if (leader) mbar_arrive(bar);
      else mbar_arrive_cluster(mapa_rank(smem_u32(bar), 0u));
    } else {
      mbar_arrive(bar);
    }
  };
  auto commit = [&](uint64_t *bar) {
    if constexpr (PAIR) umma_commit_pair(bar);
    else umma_commit(bar);
  };

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int bseq = 0;                                    // B-ring sequence number (main + extra stages), over all tiles
      int gchunk = 0;                                  // raw chunks loaded so far
      for (int tl = cl; tl < n_tiles; tl += ncl) {
        const int m0 = tile_m0(tl), nb = tile_n0(tl) + (int)rank * GN;
        for (int st = 0; st < n_stages; st++) {
          if ((st & 3) == 0) {
            const int ms = st >> 2;
            mbar_wait(raw_empty, ((uint32_t)gchunk & 1u) ^ 1u);
            mbar_arrive_expect_tx(raw_full, (uint32_t)L::kRawBytes);
            tma_load_2d(smem + L::kRawOff, &tm_w, ms * (L::kRawChunkBytes / 2), m0, raw_full);
            gchunk++;
          }
          if (ext_before(st)) {
            const int s = bseq % kSoBStages, e = st >> 5;
            mbar_wait(&empty_b[s], ((uint32_t)(bseq / kSoBStages) & 1u) ^ 1u);
            if (leader) mbar_arrive_expect_tx(&full_b[s], (uint32_t)((PAIR ? 2 : 1) * kTileBytes));
            load_b(smem + L::kBhOff + s * kTileBytes, &tm_sh, e * 64, nb, &full_b[s]);
            bseq++;
          }
          const int s = bseq % kSoBStages;
          mbar_wait(&empty_b[s], ((uint32_t)(bseq / kSoBStages) & 1u) ^ 1u);
          if (leader) mbar_arrive_expect_tx(&full_b[s], (uint32_t)((PAIR ? 4 : 2) * kTileBytes));
          load_b(smem + L::kBhOff + s * kTileBytes, &tm_xh, st * GK, nb, &full_b[s]);
          load_b(smem + L::kBlOff + s * kTileBytes, &tm_xl, st * GK, nb, &full_b[s]);
          bseq++;
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (PAIR: the leader's, for both CTAs) =====================
    if (lane == 0 && leader) {
      int bseq = 0, gst = 0, gext = 0, it = 0;         // B stages, A stages, extra stages, tiles so far
      for (int tl = cl; tl < n_tiles; tl += ncl, it++) {
        const int ab = ATM ? 0 : (it & 1);             // accumulator buffer of this tile
        const uint32_t aph = (uint32_t)(ATM ? it : (it >> 1)) & 1u;
        if constexpr (PAIR) mbar_wait_cluster(&acc_empty[ab], aph ^ 1u);
        else mbar_wait(&acc_empty[ab], aph ^ 1u);
        tc_fence_after();
        const uint32_t tacc = tmem_base + (uint32_t)(ab * TN);
        auto mma = [&](uint64_t a, uint64_t b, uint32_t acc) {
          if constexpr (PAIR) umma_f16_pair(tacc, a, b, kIdescF16Pair, acc);
          else umma_f16(tacc, a, b, kIdescF16, acc);
        };
        uint32_t acc_on = 0u;                          // 0 only for the very first MMA of the tile
        for (int st = 0; st < n_stages; st++) {
          if (ext_before(st)) {
            const int sb = bseq % kSoBStages;
            if constexpr (PAIR) mbar_wait_cluster(ext_full, (uint32_t)gext & 1u);
            else mbar_wait(ext_full, (uint32_t)gext & 1u);
            mbar_wait(&full_b[sb], (uint32_t)(bseq / kSoBStages) & 1u);
            tc_fence_after();
            const uint64_t ah = make_sw128_desc(smem_u32(smem + L::kEhOff));
            const uint64_t bh = make_sw128_desc(smem_u32(smem + L::kBhOff + sb * kTileBytes));
#pragma unroll
            for (int kk = 0; kk < 4; kk++) {
              mma(ah + (uint64_t)(kk * 2), bh + (uint64_t)(kk * 2), acc_on);
              acc_on = 1u;
            }
            commit(ext_empty);
            commit(&empty_b[sb]);
            bseq++;
            gext++;
          }
          const int sa = gst % kSoAStages, sb = bseq % kSoBStages;
          if constexpr (PAIR) mbar_wait_cluster(&full_a[sa], (uint32_t)(gst / kSoAStages) & 1u);
          else mbar_wait(&full_a[sa], (uint32_t)(gst / kSoAStages) & 1u);
          mbar_wait(&full_b[sb], (uint32_t)(bseq / kSoBStages) & 1u);
          tc_fence_after();
          const uint64_t bh = make_sw128_desc(smem_u32(smem + L::kBhOff + sb * kTileBytes));
          const uint64_t bl = make_sw128_desc(smem_u32(smem + L::kBlOff + sb * kTileBytes));
          if constexpr (ATM) {
            // A from tensor memory: hi = 32 columns (64 fp16 per lane), lo the next 32; a K16 step is 8 columns
            const uint32_t th = tmem_base + kTmemA + (uint32_t)(sa * 64), tl_ = th + 32u;
#pragma unroll
            for (int kk = 0; kk < 4; kk++) {
              const uint64_t o = (uint64_t)(kk * 2);
              const uint32_t oa = (uint32_t)(kk * 8);
              umma_f16_pair_ts(tacc, th + oa, bh + o, kIdescF16Pair, acc_on);
              acc_on = 1u;
              umma_f16_pair_ts(tacc, th + oa, bl + o, kIdescF16Pair, 1u);
              umma_f16_pair_ts(tacc, tl_ + oa, bh + o, kIdescF16Pair, 1u);
            }
          } else {
            const uint64_t ah = make_sw128_desc(smem_u32(smem + L::kAhOff + sa * kTileBytes));
            const uint64_t al = make_sw128_desc(smem_u32(smem + L::kAlOff + sa * kTileBytes));
#pragma unroll
            for (int kk = 0; kk < 4; kk++) {
              // the K offset inside the 128-byte swizzle atom advances the start address (16-byte units)
              const uint64_t o = (uint64_t)(kk * 2);
              mma(ah + o, bh + o, acc_on);
              acc_on = 1u;
              mma(ah + o, bl + o, 1u);
              mma(al + o, bh + o, 1u);
            }
          }
          commit(&empty_a[sa]);
          commit(&empty_b[sb]);
          bseq++;
          gst++;
        }
        commit(&acc_full[ab]);
      }
    }
  } else if (warp >= 4 && warp < 12) {
    // ===================== weight unpack (thread = row x block j of the stage) =====================
    const int r = (threadIdx.x - 128) & 127, j = (warp - 4) >> 2;
    const int mpad = (p.M + GM - 1) / GM * GM;                        // rowscale has [Mpad] entries: rows beyond M carry 1
    int gchunk = 0, gst0 = 0, it = 0;                                 // raw chunks, A stages before this tile, tiles so far
    for (int tl = cl; tl < n_tiles; tl += ncl, it++, gst0 += n_stages) {
      const int m0 = tile_m0(tl);
      const float rs = (m0 + r < mpad) ? __ldg(p.rowscale + m0 + r) : 1.f;   // (PAIR: the peer's tile may lie beyond Mpad)
      for (int ms = 0; ms < n_chunks; ms++, gchunk++) {
        mbar_wait(raw_full, (uint32_t)gchunk & 1u);
        uint32_t raw[NWR];
        const uint32_t src = smem_u32(smem + L::kRawOff) + (uint32_t)r * L::kRawRowBytes;
#pragma unroll
        for (int i = 0; i < NWR / 4; i++)
          asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                       : "=r"(raw[4 * i]), "=r"(raw[4 * i + 1]), "=r"(raw[4 * i + 2]), "=r"(raw[4 * i + 3])
                       : "r"(src + 16u * i));
__syncwarp();
        if (lane == 0) mbar_arrive_after_loads(raw_empty, raw, (uint32_t)p.K & 1u);   // K is a multiple of 64
        if constexpr (kMin) {
          if (j == 0) {
            if ((ms & 7) == 0) {                       // a new group of 64 blocks: the MMA must be done with the previous m_w tile
              const int gg = it * n_ext + (ms >> 3);
              if (gg > 0) mbar_wait(ext_empty, (uint32_t)(gg - 1) & 1u);
            }
            const uint4 mh = chunk_mins<WT, NWR>(raw, rs);
            const uint32_t off = (uint32_t)(r >> 3) * 1024u + (uint32_t)(r & 7) * 128u + (((uint32_t)(ms & 7) ^ (uint32_t)(r & 7)) * 16u);
            asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(smem_u32(smem + L::kEhOff) + off), "r"(mh.x),
                         "r"(mh.y), "r"(mh.z), "r"(mh.w)
                         : "memory");
            if ((ms & 7) == 7 || ms == n_chunks - 1) {
              fence_proxy_async();
              __syncwarp();
              if (lane == 0) arrive_full(ext_full);
            }
          }
        }
#pragma unroll
        for (int q = 0; q < 4; q++) {
          const int st = gst0 + ms * 4 + q, s = st % kSoAStages;
          mbar_wait(&empty_a[s], ((uint32_t)(st / kSoAStages) & 1u) ^ 1u);
          if constexpr (ATM) {
            // thread = row = TMEM lane (warp % 4 is the lane quarter); the MMA of the stage's previous use has completed
            tc_fence_after();
            const uint32_t ts = tmem_base + ((uint32_t)(r & ~31) << 16) + kTmemA + (uint32_t)(s * 64);
            if (j == 0) {
              if (q == 0) unpack_block_tm<WT, 0, NWR>(raw, ts, rs);
              else if (q == 1) unpack_block_tm<WT, 2, NWR>(raw, ts, rs);
              else if (q == 2) unpack_block_tm<WT, 4, NWR>(raw, ts, rs);
              else unpack_block_tm<WT, 6, NWR>(raw, ts, rs);
            } else {
              if (q == 0) unpack_block_tm<WT, 1, NWR>(raw, ts, rs);
              else if (q == 1) unpack_block_tm<WT, 3, NWR>(raw, ts, rs);
              else if (q == 2) unpack_block_tm<WT, 5, NWR>(raw, ts, rs);
              else unpack_block_tm<WT, 7, NWR>(raw, ts, rs);
            }
            tmem_st_wait();                  // the stores have landed in tensor memory
            tc_fence_before();
          } else {
            const uint32_t ah = smem_u32(smem + L::kAhOff + s * kTileBytes), al = smem_u32(smem + L::kAlOff + s * kTileBytes);
            if (j == 0) {
              if (q == 0) unpack_block_so<WT, 0, NWR>(raw, ah, al, r, rs);
              else if (q == 1) unpack_block_so<WT, 2, NWR>(raw, ah, al, r, rs);
              else if (q == 2) unpack_block_so<WT, 4, NWR>(raw, ah, al, r, rs);
              else unpack_block_so<WT, 6, NWR>(raw, ah, al, r, rs);
            } else {
              if (q == 0) unpack_block_so<WT, 1, NWR>(raw, ah, al, r, rs);
              else if (q == 1) unpack_block_so<WT, 3, NWR>(raw, ah, al, r, rs);
              else if (q == 2) unpack_block_so<WT, 5, NWR>(raw, ah, al, r, rs);
              else unpack_block_so<WT, 7, NWR>(raw, ah, al, r, rs);
            }
            fence_proxy_async();             // generic-proxy stores -> visible to the tensor core
          }
          __syncwarp();
          if (lane == 0) arrive_full(&full_a[s]);
        }
      }
    }
  } else if (warp >= 12) {
    // ===================== epilogue (thread = TMEM lane = weight row) =====================
    const int q4 = warp & 3, row = q4 * 32 + lane;
    int it = 0;
    for (int tl = cl; tl < n_tiles; tl += ncl, it++) {
      const int ab = ATM ? 0 : (it & 1), m = tile_m0(tl) + row, n0 = tile_n0(tl);
      const float rinv = (m < p.M) ? 1.0f / __ldg(p.rowscale + m) : 1.f;    // exact: a power of two
      mbar_wait(&acc_full[ab], (uint32_t)(ATM ? it : (it >> 1)) & 1u);
      tc_fence_after();
      const uint32_t tbase = tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(ab * TN);
#pragma unroll 2
      for (int c16 = 0; c16 < TN / 16; c16++) {
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
            if (n < p.N) store_out<PEERS>(p, (int64_t)n * p.ldc + m, __uint_as_float(v[i]) * (__ldg(p.colscale + n) * rinv));
          }
        }
      }
      // the accumulator buffer goes back to the MMA issuer: every lane's tcgen05.ld has completed (wait::ld above)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) arrive_full(&acc_empty[ab]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if constexpr (PAIR) {
    cluster_sync_all();                                // the peer's shared memory and TMEM outlive the leader's MMAs
    if (warp == 2) tmem_dealloc_pair(tmem_base, kTmemCols);
  } else {
    if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
  }
}

// ---- direct variant: both operands pre-split in global memory --------------------------------
// Q2_K and F32 weights: lamm_gemm_prep.cu dequantizes W once per call into fp16 hi / lo [M][K]
// (w * 2^F_m, the per-row lift; Q2_K: w = d * sc * q - dmin * mn evaluated in f32, which is the reference's
// d*isum - dmin*summs regrouped, src/lamm_kernel_q2_k.hpp:28-75), so the mainloop is a plain TMA-fed
// GEMM: no unpack warps, no extra stages.  Same three MMAs per K step, same epilogue.
constexpr int kDirStages = 3;

// F16W (SURVEY 8f-2, F16 weights x activations rounded to fp16 = ggml_vec_dot_f16's operands): both operands ARE fp16,
// so there is no lo tile, one MMA per K step and no rescaling in the epilogue; the A map sits on the caller's W.
template <bool PEERS, bool F16W = false>
__global__ void __launch_bounds__(192, 1)
gemm_direct_kernel(const __grid_constant__ CUtensorMap tm_ah, const __grid_constant__ CUtensorMap tm_al,
                   const __grid_constant__ CUtensorMap tm_xh, const __grid_constant__ CUtensorMap tm_xl, const SoParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  constexpr int kStageBytes = (F16W ? 2 : 4) * kTileBytes;        // Ah, Al, Bh, Bl (F16W: A, B)
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + kDirStages * kStageBytes);
  uint64_t *full = bars, *empty = bars + kDirStages, *acc_full = bars + 2 * kDirStages;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 8);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tiles = (p.N + GN - 1) / GN;
  const int m0 = (blockIdx.x / n_tiles) * GM, n0 = (blockIdx.x % n_tiles) * GN;
  const int n_stages = p.K / GK;
  if (threadIdx.x == 0) {
    for (int i = 0; i < kDirStages; i++) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_init(acc_full, 1);
    fence_barrier_init();
    prefetch_tmap(&tm_ah);
    prefetch_tmap(&tm_al);
    prefetch_tmap(&tm_xh);
    prefetch_tmap(&tm_xl);
  }
  if (warp == 1) tmem_alloc(tmem_slot, 128);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (warp == 0) {
    if (lane == 0) {
      for (int st = 0; st < n_stages; st++) {
        const int s = st % kDirStages;
        uint8_t *stage = smem + s * kStageBytes;
        mbar_wait(&empty[s], ((uint32_t)(st / kDirStages) & 1u) ^ 1u);
        mbar_arrive_expect_tx(&full[s], (uint32_t)kStageBytes);
        tma_load_2d(stage, &tm_ah, st * GK, m0, &full[s]);
        if constexpr (F16W) {
          tma_load_2d(stage + kTileBytes, &tm_xh, st * GK, n0, &full[s]);
        } else {
          tma_load_2d(stage + kTileBytes, &tm_al, st * GK, m0, &full[s]);
          tma_load_2d(stage + 2 * kTileBytes, &tm_xh, st * GK, n0, &full[s]);
          tma_load_2d(stage + 3 * kTileBytes, &tm_xl, st * GK, n0, &full[s]);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      for (int st = 0; st < n_stages; st++) {
        const int s = st % kDirStages;
        mbar_wait(&full[s], (uint32_t)(st / kDirStages) & 1u);
        tc_fence_after();
        const uint32_t base = smem_u32(smem + s * kStageBytes);
        const uint64_t ah = make_sw128_desc(base), al = make_sw128_desc(base + kTileBytes);
        const uint64_t bh = make_sw128_desc(base + (F16W ? 1 : 2) * kTileBytes), bl = make_sw128_desc(base + 3 * kTileBytes);
#pragma unroll
        for (int kk = 0; kk < 4; kk++) {
          const uint64_t o = (uint64_t)(kk * 2);
          umma_f16(tmem_base, ah + o, bh + o, kIdescF16, (st | kk) ? 1u : 0u);
          if constexpr (!F16W) {
            umma_f16(tmem_base, ah + o, bl + o, kIdescF16, 1u);
            umma_f16(tmem_base, al + o, bh + o, kIdescF16, 1u);
          }
        }
        umma_commit(&empty[s]);
      }
      umma_commit(acc_full);
    }
  } else {
    // epilogue warps 2..5: TMEM lane quarter = warp & 3
    const int q4 = warp & 3, row = q4 * 32 + lane, m = m0 + row;
    const float rinv = F16W ? 1.0f : 1.0f / __ldg(p.rowscale + m);                  // exact: a power of two ([Mpad] entries)
    mbar_wait(acc_full, 0u);
    tc_fence_after();
    const uint32_t tbase = tmem_base + ((uint32_t)(q4 * 32) << 16);
#pragma unroll 2
    for (int c16 = 0; c16 < 8; c16++) {
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
          if (n < p.N) store_out<PEERS>(p, (int64_t)n * p.ldc + m, F16W ? __uint_as_float(v[i]) : __uint_as_float(v[i]) * (__ldg(p.colscale + n) * rinv));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 128);
}

size_t gemm_direct_ws_bytes(int64_t M, int64_t N, int64_t K) {
  const int64_t npad = (N + GN - 1) / GN * GN, mpad = (M + GM - 1) / GM * GM;
  return (size_t)(2 * mpad * K * 2 + 2 * npad * K * 2 + npad * 4 + mpad * 4 + 2048);
}

// W: Q2_K blocks or f32; X: Q8_K blocks (Q2_K) or f32 (F32 weights)
int launch_gemm_direct(int w_type, const void *w, int64_t w_stride, const void *xq, int64_t xq_stride, float *c,
                       int64_t ldc, int64_t M, int64_t N, int64_t K, void *ws, cudaStream_t st) {
  const int64_t npad = (N + GN - 1) / GN * GN, mpad = (M + GM - 1) / GM * GM;
  uint8_t *base = (uint8_t *)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
  void *ah = base, *al = base + mpad * K * 2;
  void *xh = base + 2 * mpad * K * 2, *xl = (uint8_t *)xh + npad * K * 2;
  float *colscale = (float *)((uint8_t *)xl + npad * K * 2);
  float *rowscale = colscale + npad;
  int rc = launch_dequant_hilo(w_type, w, w_stride, ah, al, rowscale, M, mpad, K, st);
  if (rc) return rc;
  rc = launch_prep_direct(type_info(w_type).vec_dot, xq, xq_stride, xh, xl, colscale, N, npad, K, st);
  if (rc) return rc;
  CUtensorMap tm_ah, tm_al, tm_xh, tm_xl;
  if ((rc = make_map_2d(&tm_ah, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, ah, (uint64_t)K, (uint64_t)mpad, (uint64_t)K * 2, GK, GM,
                        CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  if ((rc = make_map_2d(&tm_al, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, al, (uint64_t)K, (uint64_t)mpad, (uint64_t)K * 2, GK, GM,
                        CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  if ((rc = make_map_2d(&tm_xh, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, xh, (uint64_t)K, (uint64_t)npad, (uint64_t)K * 2, GK, GN,
                        CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  if ((rc = make_map_2d(&tm_xl, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, xl, (uint64_t)K, (uint64_t)npad, (uint64_t)K * 2, GK, GN,
                        CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  SoParams p;
  p.colscale = colscale;  p.rowscale = rowscale;  p.c = c;  p.ldc = ldc;
  p.M = (int)M;  p.N = (int)N;  p.Npad = (int)npad;  p.K = (int)K;
  fill_peers(p);
  constexpr int smem = kDirStages * 4 * kTileBytes + 256 + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_direct_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(gemm_direct_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) {
      set_error("gemm_direct: cudaFuncSetAttribute(%d): %s", smem, cudaGetErrorString(e));
      return LAMMCU_ERR_CUDA;
    }
    attr_set = true;
  }
  const int tiles = (int)((mpad / GM) * (npad / GN));
  if (p.n_peer) gemm_direct_kernel<true><<<tiles, 192, smem, st>>>(tm_ah, tm_al, tm_xh, tm_xl, p);
  else gemm_direct_kernel<false><<<tiles, 192, smem, st>>>(tm_ah, tm_al, tm_xh, tm_xl, p);
  g_launch_count++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("gemm_direct launch: %s", cudaGetErrorString(e));
    return LAMMCU_ERR_CUDA;
  }
  return LAMMCU_OK;
}

// F16 weights x fp16 activations (both as the caller has them): W [M][K] fp16 rows (16-byte aligned, stride % 16 == 0),
// xq [N][K] fp16 rows.  Rows beyond M / N in the last tiles are zero-filled by the TMA unit.
bool gemm_f16_supported(const void *w, int64_t w_stride, const void *xq, int64_t xq_stride, int64_t M, int64_t N, int64_t K) {
  if (K <= 0 || K % GK || M <= 0 || N <= 0 || M > 0x3fffffff || N > 0x3fffffff) return false;
  return (((uintptr_t)w | (uintptr_t)w_stride | (uintptr_t)xq | (uintptr_t)xq_stride) & 15) == 0;
}
int launch_gemm_f16(const void *w, int64_t w_stride, const void *xq, int64_t xq_stride, float *c, int64_t ldc, int64_t M,
                    int64_t N, int64_t K, cudaStream_t st) {
  CUtensorMap tm_a, tm_x;
  int rc;
  if ((rc = make_map_2d(&tm_a, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, w, (uint64_t)K, (uint64_t)M, (uint64_t)w_stride, GK, GM,
                        CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  if ((rc = make_map_2d(&tm_x, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, xq, (uint64_t)K, (uint64_t)N, (uint64_t)xq_stride, GK, GN,
                        CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  SoParams p;
  p.colscale = nullptr;  p.rowscale = nullptr;  p.c = c;  p.ldc = ldc;
  p.M = (int)M;  p.N = (int)N;  p.Npad = (int)((N + GN - 1) / GN * GN);  p.K = (int)K;
  fill_peers(p);
  constexpr int smem = kDirStages * 2 * kTileBytes + 256 + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_direct_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_direct_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { set_error("gemm_f16: cudaFuncSetAttribute(%d): %s", smem, cudaGetErrorString(e)); return LAMMCU_ERR_CUDA; }
    attr_set = true;
  }
  const int tiles = (int)(((M + GM - 1) / GM) * ((N + GN - 1) / GN));
  if (p.n_peer) gemm_direct_kernel<true, true><<<tiles, 192, smem, st>>>(tm_a, tm_a, tm_x, tm_x, p);
  else gemm_direct_kernel<false, true><<<tiles, 192, smem, st>>>(tm_a, tm_a, tm_x, tm_x, p);
  g_launch_count++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("gemm_f16 launch: %s", cudaGetErrorString(e)); return LAMMCU_ERR_CUDA; }
  return LAMMCU_OK;
}

// ---- host side ------------------------------------------------------------------------------
template <int WT>
static int launch_so(const CUtensorMap &tm_w, const CUtensorMap &tm_xh, const CUtensorMap &tm_xl,
                     const CUtensorMap &tm_sh, const SoParams &p, cudaStream_t st) {
  constexpr int smem = SoSmem<WT, SoRing<false>::A, SoRing<false>::B>::kTotal + 1024;
  constexpr int smem_pair = SoSmem<WT, SoRing<true>::A, SoRing<true>::B>::kTotal + 1024;
  static_assert(smem <= 227 * 1024, "shared memory budget");
  constexpr int smem_atm = SoSmem<WT, SoRing<true, true>::A, SoRing<true, true>::B, true>::kTotal + 1024;
  constexpr bool pair_fits = smem_pair <= 227 * 1024;
  static_assert(smem_atm <= 227 * 1024, "shared memory budget (A in tensor memory)");
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_so_kernel<WT, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(gemm_so_kernel<WT, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e == cudaSuccess && pair_fits)
      e = cudaFuncSetAttribute(gemm_so_kernel<WT, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_pair);
    if (e == cudaSuccess && pair_fits)
      e = cudaFuncSetAttribute(gemm_so_kernel<WT, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_pair);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(gemm_so_kernel<WT, false, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_atm);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(gemm_so_kernel<WT, true, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_atm);
    if (e != cudaSuccess) {
      set_error("gemm_so: cudaFuncSetAttribute(%d): %s", smem, cudaGetErrorString(e));
      return LAMMCU_ERR_CUDA;
    }
    attr_set = true;
  }
  // CTA pairs (256 x 256 per cluster of 2) whenever both dimensions have more than one 128-tile; LAMM_GEMM_PAIR=0: never
  // (read per call, like LAMM_GEMM_ATM below: tests force each form on ragged shapes in one process)
  const char *pair_env = getenv("LAMM_GEMM_PAIR");
  const int pair_on = pair_env ? atoi(pair_env) : 1;
  // Unpacked weight tiles of the pair form in TENSOR MEMORY (ATM, opt-in) or in shared memory (default).  Measured at
  // 13B N = 2048 (profiles/r2bc, r2bd, r2bh): ATM halves the shared-memory traffic and draws less power (1582 vs 1485 MHz
  // under the cap) but needs more cycles per tile (ONE accumulator: the epilogue is exposed; tcgen05.st + wait::st sit in
  // the unpack chain) -- Q4_1 499 vs 527 TOPS, Q4_0 501 vs 522, Q8_0 491 vs 514, Q5_1 491 vs 512 (it led Q5_1 only while
  // that format's raw rows were 4-way bank-conflicted).  LAMM_GEMM_ATM=1 selects it (read per call: tests run both).
  const char *atm_env = getenv("LAMM_GEMM_ATM");
  const int atm_on = atm_env ? (atoi(atm_env) != 0) : !pair_fits;
  cudaError_t e;
  // no persistent scheduler yet, so the form that needs less time for its waves wins: a pair tile is 4x the work of a
  // single tile on 2 SMs = 2 single-tile times, divided by the measured per-tile gain of the pair form (LAMM_GEMM_PAIR=2:
  // always pairs)
  const int pairs = ((p.M + 255) / 256) * ((p.N + 255) / 256);
  const int singles = ((p.M + GM - 1) / GM) * ((p.N + GN - 1) / GN);
  const int sms = sm_count_cached(), slots = sms / 2;
  const double t_pair = (double)((pairs + slots - 1) / slots) * 2.0 / kSoPairGain;
  const double t_single = (double)((singles + sms - 1) / sms);
  if ((pair_fits || atm_on) && pair_on && p.M > GM && p.N > GN && (pair_on > 1 || t_pair < t_single)) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * (pairs < slots ? pairs : slots));         // persistent: one pair per TPC walks the tiles
    cfg.blockDim = dim3(kSoThreads);
    cfg.dynamicSmemBytes = atm_on ? smem_atm : smem_pair;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (atm_on) {
      if (p.n_peer) e = cudaLaunchKernelEx(&cfg, gemm_so_kernel<WT, true, true, true>, tm_w, tm_xh, tm_xl, tm_sh, p);
      else e = cudaLaunchKernelEx(&cfg, gemm_so_kernel<WT, false, true, true>, tm_w, tm_xh, tm_xl, tm_sh, p);
    } else if (p.n_peer) {
      e = cudaLaunchKernelEx(&cfg, gemm_so_kernel<WT, true, true>, tm_w, tm_xh, tm_xl, tm_sh, p);
    } else {
      e = cudaLaunchKernelEx(&cfg, gemm_so_kernel<WT, false, true>, tm_w, tm_xh, tm_xl, tm_sh, p);
    }
  } else {
    const int grid = singles < sms ? singles : sms;                    // persistent: one CTA per SM walks the tiles
    if (p.n_peer) gemm_so_kernel<WT, true, false><<<grid, kSoThreads, smem, st>>>(tm_w, tm_xh, tm_xl, tm_sh, p);
    else gemm_so_kernel<WT, false, false><<<grid, kSoThreads, smem, st>>>(tm_w, tm_xh, tm_xl, tm_sh, p);
    e = cudaGetLastError();
  }
  g_launch_count++;
  if (e != cudaSuccess) {
    set_error("gemm_so launch (type %d): %s", WT, cudaGetErrorString(e));
    return LAMMCU_ERR_CUDA;
  }
  return LAMMCU_OK;
}

// box width of the raw-weight tensor map = the shared-memory row pitch of the kernel's raw chunk (SoSmem::kRawRowBytes)
static int so_raw_row_bytes(int w_type) {
  switch (w_type) {
  case LAMMCU_TYPE_Q4_0: return SoSmem<LAMMCU_TYPE_Q4_0, 2, 3>::kRawRowBytes;
  case LAMMCU_TYPE_Q4_1: return SoSmem<LAMMCU_TYPE_Q4_1, 2, 3>::kRawRowBytes;
  case LAMMCU_TYPE_Q5_0: return SoSmem<LAMMCU_TYPE_Q5_0, 2, 3>::kRawRowBytes;
  case LAMMCU_TYPE_Q5_1: return SoSmem<LAMMCU_TYPE_Q5_1, 2, 3>::kRawRowBytes;
  default:               return SoSmem<LAMMCU_TYPE_Q8_0, 2, 3>::kRawRowBytes;
  }
}

size_t gemm_so_ws_bytes(int64_t M, int64_t N, int64_t K) {
  const int64_t npad = (N + GN - 1) / GN * GN, kbpad = (K / 32 + 63) / 64 * 64, mpad = (M + GM - 1) / GM * GM;
  return (size_t)(2 * npad * K * 2 + npad * kbpad * 2 + npad * 4 + mpad * 4 + 1024);
}

// xq: quantized activations (ggml blocks, row stride xq_stride); ws: gemm_so_ws_bytes() of scratch
// or x != NULL: f32 activations, quantized on the fly (the ggml-layout Xq is written to xq iff xq != NULL)
int launch_gemm_so(int w_type, const void *w, int64_t w_stride, const void *xq, int64_t xq_stride, float *c,
                   int64_t ldc, int64_t M, int64_t N, int64_t K, void *ws, cudaStream_t st, const float *x, int64_t ldx) {
  const TypeInfo ti = type_info(w_type);
  const int64_t npad = (N + GN - 1) / GN * GN, kb = K / 32, kbpad = (kb + 63) / 64 * 64;
  uint8_t *base = (uint8_t *)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
  void *xh = base, *xl = base + npad * K * 2;
  void *sh = base + 2 * npad * K * 2;
  float *colscale = (float *)((uint8_t *)sh + npad * kbpad * 2);
  float *rowscale = colscale + npad;
  const int perm = (w_type == LAMMCU_TYPE_Q8_0) ? 0 : 1;
  int rc = launch_wrow_scale(w_type, w, w_stride, rowscale, M, (M + GM - 1) / GM * GM, K, st);
  if (rc) return rc;
  rc = x ? launch_quant_prep_so(ti.vec_dot, perm, x, ldx, const_cast<void *>(xq), xq_stride, xh, xl, sh, colscale, N,
                                    npad, K, st)
             : launch_gemm_prep_so(ti.vec_dot, perm, xq, xq_stride, xh, xl, sh, colscale, N, npad, K, st);
  if (rc) return rc;
  CUtensorMap tm_w, tm_xh, tm_xl, tm_sh;
  // raw weights as uint16 [M][row_bytes / 2], box = 8 blocks x 128 rows
  rc = make_map_2d(&tm_w, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, w, (uint64_t)(kb * ti.size / 2), (uint64_t)M,
                   (uint64_t)w_stride, (uint32_t)(so_raw_row_bytes(w_type) / 2), GM, CU_TENSOR_MAP_SWIZZLE_NONE);
  if (rc) return rc;
  rc = make_map_2d(&tm_xh, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, xh, (uint64_t)K, (uint64_t)npad, (uint64_t)K * 2, GK, GN,
                   CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  rc = make_map_2d(&tm_xl, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, xl, (uint64_t)K, (uint64_t)npad, (uint64_t)K * 2, GK, GN,
                   CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  // s_x of the extra stages (Q4_1 / Q5_1; a harmless dummy otherwise)
  rc = make_map_2d(&tm_sh, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, sh, (uint64_t)kbpad, (uint64_t)npad, (uint64_t)kbpad * 2,
                   GK, GN, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  SoParams p;
  p.colscale = colscale;  p.rowscale = rowscale;  p.c = c;  p.ldc = ldc;
  p.M = (int)M;  p.N = (int)N;  p.Npad = (int)npad;  p.K = (int)K;
  fill_peers(p);
  switch (w_type) {
  case LAMMCU_TYPE_Q4_0: return launch_so<LAMMCU_TYPE_Q4_0>(tm_w, tm_xh, tm_xl, tm_sh, p, st);
  case LAMMCU_TYPE_Q4_1: return launch_so<LAMMCU_TYPE_Q4_1>(tm_w, tm_xh, tm_xl, tm_sh, p, st);
  case LAMMCU_TYPE_Q5_0: return launch_so<LAMMCU_TYPE_Q5_0>(tm_w, tm_xh, tm_xl, tm_sh, p, st);
  case LAMMCU_TYPE_Q5_1: return launch_so<LAMMCU_TYPE_Q5_1>(tm_w, tm_xh, tm_xl, tm_sh, p, st);
  default:               return launch_so<LAMMCU_TYPE_Q8_0>(tm_w, tm_xh, tm_xl, tm_sh, p, st);
  }
}

}  // namespace lamm
