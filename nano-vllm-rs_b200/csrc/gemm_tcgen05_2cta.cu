// gemm_tcgen05_2cta.cu — K5 for M > 128: the same contract as gemm_tcgen05.cu on CTA PAIRS (tcgen05 cta_group::2).
//
// GEMM throughput on B200 is bound by the L2 -> SM operand feed (a 128 x 256 single-CTA tile moves 1 byte per
// ~43 MACs; the measured 1.35 PFLOP/s of gemm_tcgen05.cu matches that bound).  A CTA pair computes a 256 x BN tile:
// each CTA stages its own 128 rows of x and only HALF of the w tile, the leader CTA issues one
// tcgen05.mma.cta_group::2 (M = 256) that reads both CTAs' shared memory and writes 128 accumulator rows into each
// CTA's TMEM -> 64 MACs per byte fetched, i.e. 1.5x less L2 traffic per FLOP.
//   both CTAs, warp 0 lane 0   TMA producer (cp.async.bulk.tensor ... cta_group::2: bytes complete on the LEADER's barrier)
//   leader,    warp 1 lane 0   MMA issuer; tcgen05.commit ... multicast releases the smem stage / publishes the tile in BOTH CTAs
//   both CTAs, warp 2          TMEM allocator (cta_group::2)
//   both CTAs, warps 4-7       epilogue on the CTA's own 128 TMEM lanes; drained-signal goes to the leader (remote mbarrier arrive)
#include "tc.cuh"

namespace nvllm {

template <int BN>
struct Gemm2Cfg {
  static constexpr int BM = 256, BM_CTA = 128, BK = 64;
  static constexpr int kABytes = BM_CTA * BK * 2;
  static constexpr int kBBytes = (BN / 2) * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;  // per CTA
  static constexpr int kStagesRaw = (200 * 1024) / kStageBytes;
  static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
  static constexpr int kTmemCols = (2 * BN <= 64) ? 64 : (2 * BN <= 128) ? 128 : (2 * BN <= 256) ? 256 : 512;
  static constexpr int kSmemBytes = kStages * kStageBytes + (2 * kStages + 4) * 8 + 16 + 1024;
};

template <typename OutT>
__device__ __forceinline__ void store_row32_2(OutT* dst, const uint32_t (&r)[32], int col0, int n) {
  if constexpr (sizeof(OutT) == 2) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (col0 + j * 8 < n) {
        uint4 v;
        v.x = pack_bf16x2(__uint_as_float(r[8 * j + 0]), __uint_as_float(r[8 * j + 1]));
        v.y = pack_bf16x2(__uint_as_float(r[8 * j + 2]), __uint_as_float(r[8 * j + 3]));
        v.z = pack_bf16x2(__uint_as_float(r[8 * j + 4]), __uint_as_float(r[8 * j + 5]));
        v.w = pack_bf16x2(__uint_as_float(r[8 * j + 6]), __uint_as_float(r[8 * j + 7]));
        *reinterpret_cast<uint4*>(dst + j * 8) = v;
      }
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (col0 + j * 4 < n)
        *reinterpret_cast<uint4*>(dst + j * 4) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
  }
}

// SILU == true: w is the packed [gate (N rows) | up (N rows)] matrix of MergedColumnParallelLinear (linear.rs:421-433);
// the leader CTA stages BN/2 gate rows, its peer the matching BN/2 up rows, so accumulator columns [0, BN/2) are gate and
// [BN/2, BN) are up for the same BN/2 output features, and the epilogue writes y[:, n_tile*BN/2 + c] = silu(gate)*up
// (SiluAndMul::forward, activation.rs:46-63) without the [M, 2N] intermediate ever reaching HBM.  N = output features.
template <int BN, typename OutT, bool SILU>
__global__ void __launch_bounds__(256, 1)
gemm_tcgen05_2cta_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                         OutT* __restrict__ y, int M, int N, int K, int64_t ldy, int num_m_tiles, int num_tiles) {
  using C = Gemm2Cfg<BN>;
  constexpr int STAGES = C::kStages;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sA = smem_base;
  const uint32_t sB = smem_base + STAGES * C::kABytes;
  const uint32_t bars = smem_base + STAGES * C::kStageBytes;
  auto full_bar = [&](int s) { return bars + s * 8; };
  auto empty_bar = [&](int s) { return bars + (STAGES + s) * 8; };
  auto tfull_bar = [&](int a) { return bars + (2 * STAGES + a) * 8; };
  auto tempty_bar = [&](int a) { return bars + (2 * STAGES + 2 + a) * 8; };
  const uint32_t tmem_slot = bars + (2 * STAGES + 4) * 8;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int num_kb = K / C::BK;
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
  pdl_launch_dependents();

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_w);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);   // leader: its producer's arrive.expect_tx (+ both CTAs' TMA bytes)
      mbar_init(empty_bar(s), 1);  // each CTA: multicast tcgen05.commit from the leader
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);     // each CTA: multicast commit
      mbar_init(tempty_bar(a), 256);  // leader: 128 epilogue threads of each CTA
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc_cg2<C::kTmemCols>(tmem_slot);
  tc_fence_before();
  cluster_sync_all();  // barriers of BOTH CTAs are initialised before any remote signal can arrive
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0 && lane == 0) {
    // ===== TMA producer (both CTAs) =====
    pdl_wait();
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      const int m_tile = tile % num_m_tiles, n_tile = tile / num_m_tiles;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(empty_bar(stage), phase ^ 1);
        if (leader) mbar_expect_tx(full_bar(stage), 2 * C::kStageBytes);
        tma_load_2d_cg2(sA + stage * C::kABytes, &tmap_x, full_bar(stage), kb * C::BK, m_tile * C::BM + rank * C::BM_CTA);
        tma_load_2d_cg2(sB + stage * C::kBBytes, &tmap_w, full_bar(stage), kb * C::BK,
                        SILU ? static_cast<int>(rank) * N + n_tile * (BN / 2) : n_tile * BN + static_cast<int>(rank) * (BN / 2));
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1 && lane == 0 && leader) {
    // ===== MMA issuer (leader CTA, single thread) =====
    constexpr uint32_t idesc = umma_idesc_bf16(C::BM, BN, 0, 0);
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    pdl_wait();  // keep the wait for the predecessor out of the bounded mbarrier waits
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      mbar_wait(tempty_bar(acc), acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * BN;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(full_bar(stage), phase);
        tc_fence_after();
        const uint32_t a0 = sA + stage * C::kABytes, b0 = sB + stage * C::kBBytes;
#pragma unroll
        for (int k = 0; k < C::BK / 16; ++k) {
          const uint64_t adesc = umma_desc_sw128(a0 + k * 32, 16, 1024);
          const uint64_t bdesc = umma_desc_sw128(b0 + k * 32, 16, 1024);
          umma_bf16_ss_cg2(tmem_d, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
        }
        umma_commit_cg2(empty_bar(stage), 0b11);
        if (kb == num_kb - 1) umma_commit_cg2(tfull_bar(acc), 0b11);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else if (warp >= 4) {
    // ===== epilogue (both CTAs): own 128 TMEM lanes -> global =====
    pdl_wait();
    const int q = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      const int m_tile = tile % num_m_tiles, n_tile = tile / num_m_tiles;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const int row = m_tile * C::BM + rank * C::BM_CTA + q * 32 + lane;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;
      if constexpr (SILU) {
#pragma unroll 1
        for (int c = 0; c < BN / 64; ++c) {
          uint32_t g[32], u[32];
          tmem_ld_32x32b_x32(taddr + c * 32, g);
          tmem_ld_32x32b_x32(taddr + BN / 2 + c * 32, u);
          tmem_ld_wait();
#pragma unroll
          for (int e = 0; e < 32; e += 2) {
            // same rounding points as the unfused path: gate_up is a bf16 tensor, silu(gate) is a bf16 tensor
            const uint32_t gb = pack_bf16x2(__uint_as_float(g[e]), __uint_as_float(g[e + 1]));
            const uint32_t ub = pack_bf16x2(__uint_as_float(u[e]), __uint_as_float(u[e + 1]));
            const float g0 = bf16_lo(gb), g1 = bf16_hi(gb);
            const uint32_t sb = pack_bf16x2(g0 / (1.0f + expf(-g0)), g1 / (1.0f + expf(-g1)));
            g[e] = __float_as_uint(bf16_lo(sb) * bf16_lo(ub));
            g[e + 1] = __float_as_uint(bf16_hi(sb) * bf16_hi(ub));
          }
          const int col0 = n_tile * (BN / 2) + c * 32;
          if (row < M && col0 < N) store_row32_2<OutT>(y + row * ldy + col0, g, col0, N);
        }
      } else {
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t r[32];
          tmem_ld_32x32b_x32(taddr + c * 32, r);
          tmem_ld_wait();
          const int col0 = n_tile * BN + c * 32;
          if (row < M && col0 < N) store_row32_2<OutT>(y + row * ldy + col0, r, col0, N);
        }
      }
      tc_fence_before();
      mbar_arrive_cluster(tempty_bar(acc), 0);  // tell the leader this CTA's half of the accumulator is drained
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  cluster_sync_all();  // neither CTA may exit (or free TMEM) while its peer can still touch its smem / barriers
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_cg2<C::kTmemCols>(tmem_base);
  }
}

template <int BN, typename OutT, bool SILU = false>
static int launch_gemm2(const CUtensorMap& tx, const CUtensorMap& tw, void* y, int M, int N, int K, int64_t ldy,
                        cudaStream_t st) {
  using C = Gemm2Cfg<BN>;
  auto kern = gemm_tcgen05_2cta_kernel<BN, OutT, SILU>;
  NVLLM_SMEM_ATTR_ONCE(kern, C::kSmemBytes);
  const int num_m_tiles = (M + C::BM - 1) / C::BM;
  const int num_n_tiles = SILU ? (N + BN / 2 - 1) / (BN / 2) : (N + BN - 1) / BN;
  const int num_tiles = num_m_tiles * num_n_tiles;
  const int max_clusters = sm_count() / 2;
  const int clusters = num_tiles < max_clusters ? num_tiles : max_clusters;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(2 * clusters);
  cfg.blockDim = dim3(256);
  cfg.dynamicSmemBytes = C::kSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  NVLLM_CUDA(cudaLaunchKernelEx(&cfg, kern, tx, tw, static_cast<OutT*>(y), M, N, K, ldy, num_m_tiles, num_tiles));
  return 0;
}

// Tile width for the CTA-pair kernel: small problems take the narrowest tile that still fits in one wave of 74
// clusters; otherwise the widest tile whose last wave is reasonably full.
static int pick_bn_2cta(int M, int N) {
  const int clusters = sm_count() / 2;
  const long mt = (M + 255) / 256;
  const int small[3] = {64, 128, 256};
  for (int i = 0; i < 3; ++i)
    if (mt * ((N + small[i] - 1) / small[i]) <= clusters) return small[i];
  const int wide[3] = {256, 128, 64};
  int best = 256;
  double best_score = -1.0;
  for (int i = 0; i < 3; ++i) {
    const long tiles = mt * ((N + wide[i] - 1) / wide[i]);
    const long waves = (tiles + clusters - 1) / clusters;
    const double eff = static_cast<double>(tiles) / static_cast<double>(waves * clusters);
    const double reuse = wide[i] == 256 ? 1.0 : (wide[i] == 128 ? 0.8 : 0.6);  // MACs per byte fetched
    if (eff * reuse > best_score) { best_score = eff * reuse; best = wide[i]; }
  }
  return best;
}

int gemm_2cta_dispatch(void* y, const void* x, const void* w, int m, int n, int k, int64_t ldx, int64_t ldw,
                       int64_t ldy, int out_dtype, cudaStream_t st) {
  const int forced = knobs().gemm_bn;
  const int bn = (forced == 64 || forced == 128 || forced == 256) ? forced : pick_bn_2cta(m, n);
  CUtensorMap tx, tw;
  if (int rc = make_tmap_bf16_2d(&tx, x, m, k, ldx, 128)) return rc;
  if (int rc = make_tmap_bf16_2d(&tw, w, n, k, ldw, bn / 2)) return rc;
#define NVLLM_G2(BNV)                                                                            \
  case BNV:                                                                                      \
    return out_dtype == NVLLM_BF16 ? launch_gemm2<BNV, bf16>(tx, tw, y, m, n, k, ldy, st)        \
                                   : launch_gemm2<BNV, float>(tx, tw, y, m, n, k, ldy, st)
  switch (bn) {
    NVLLM_G2(256);
    NVLLM_G2(128);
    default:
      NVLLM_G2(64);
  }
#undef NVLLM_G2
}

// y[M, inter] = silu(x . Wgate^T) * (x . Wup^T) with w = [Wgate | Wup] packed row-wise ([2*inter, K]).
int gemm_silu_2cta_dispatch(void* y, const void* x, const void* w, int m, int inter, int k, int64_t ldx, int64_t ldw,
                            int64_t ldy, cudaStream_t st) {
  // accumulator tile BN covers BN/2 output features; reuse the tile heuristic on the 2*inter accumulator columns
  const int bn = pick_bn_2cta(m, 2 * inter);
  NVLLM_CHECK_ARG(inter % (bn / 2) == 0 || true, NVLLM_ERR_ARG, "gemm_silu: inter");
  CUtensorMap tx, tw;
  if (int rc = make_tmap_bf16_2d(&tx, x, m, k, ldx, 128)) return rc;
  if (int rc = make_tmap_bf16_2d(&tw, w, 2 * static_cast<int64_t>(inter), k, ldw, bn / 2)) return rc;
  switch (bn) {
    case 256: return launch_gemm2<256, bf16, true>(tx, tw, y, m, inter, k, ldy, st);
    case 128: return launch_gemm2<128, bf16, true>(tx, tw, y, m, inter, k, ldy, st);
    default: return launch_gemm2<64, bf16, true>(tx, tw, y, m, inter, k, ldy, st);
  }
}

}  // namespace nvllm
