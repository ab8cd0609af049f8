// gemm_splitk.cu — K5'': split-K variant of the tcgen05 GEMM for the decode-sized projections (M <= 512 rows).
//
// Why: measured in-graph (profiles/r01_gemm_ingraph_v1.jsonl) the small-M GEMMs cost ~0.31 us per 64-wide k-block per
// CTA whatever the tile width — each SM can only ingest ~35-45 B/clk through TMA, and every CTA of a 128 x BN tile has
// to pull (128 + BN) x K x 2 bytes on its own (the x tile is re-read by every n-tile; multicasting it does not help,
// the bytes still have to enter each SM).  o_proj / down_proj of Qwen3-0.6B (N = 1024, K = 2048 / 3072) run on only
// 64 CTAs at 12-17 us while their weights would stream from HBM in ~1 us.
// What: a cluster of SK CTAs shares one 128 x BN output tile and splits K: per-CTA ingest drops to
// (128 + BN) x K / SK x 2 bytes, all SMs pull in parallel and wide tiles (good x reuse) become affordable.  Each CTA
// accumulates its partial tile in TMEM, parks it as f32 in its own shared memory (over the drained operand ring), and
// after one cluster barrier every CTA reduces BN / SK columns of the tile over distributed shared memory in FIXED rank
// order (deterministic, independent of timing) and writes its slice of y.  No workspace, no atomics, no second kernel.
// Same contract as nvllm_gemm_bf16 / nvllm_gemm_silu_mul_bf16 with bf16 output (LinearBase::forward, linear.rs:228-239;
// MergedColumnParallelLinear + SiluAndMul, linear.rs:437-439 + activation.rs:46-63).
#include "tc.cuh"

namespace nvllm {
#ifdef NVLLM_TIMELINE
// Debug build only (make timeline -> lib/libnvllm_timeline.so, tools/gemm_timeline.py): every CTA records %globaltimer at
// its phase boundaries into a host-provided buffer, to decompose the fixed per-launch cost of the decode-sized GEMMs.
__device__ unsigned long long* g_tl_buf = nullptr;   // [cap][12]: {blockIdx, t0..t9, smid}
__device__ unsigned int g_tl_cursor = 0;
__device__ unsigned int g_tl_cap = 0;
__device__ __forceinline__ unsigned long long tl_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t));
  return t;
}
#define TL_DECL __shared__ unsigned long long tl_s[12];
#define TL(i) do { tl_s[i] = tl_now(); } while (0)
#else
#define TL_DECL
#define TL(i) do { } while (0)
#endif
namespace {

template <int BN>
struct SkCfg {
  static constexpr int BM = 128, BK = 64;
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kMaxStages = (192 * 1024) / kStageBytes > 8 ? 8 : (192 * 1024) / kStageBytes;
  // f32 partial tile, column-major [BN][128]: a warp (32 consecutive rows of one column) touches 128 contiguous bytes, for
  // the local stores (conflict-free) and for the DSMEM loads (one coalesced request per warp — strided 16-byte DSMEM
  // loads measured 2.4 us slower per kernel, profiles/r01_gemm_ingraph_v3_splitk_v2.jsonl)
  static constexpr int kColPitch = BM * 4;
  static constexpr int kPartialBytes = BN * kColPitch;
  static constexpr int kTmemCols = BN < 32 ? 32 : BN;
  static constexpr int kMaxBars = 2 * kMaxStages + 1;
  // The operand ring uses `stages` stages (a launch parameter; the k-loop is latency-bound, so the default is the deepest
  // ring that fits: shallower rings that let two CTAs share an SM measured slower); the partial tile is parked over the
  // drained ring.
  static constexpr int ring_bytes(int stages) {
    return stages * kStageBytes > kPartialBytes ? stages * kStageBytes : ((kPartialBytes + 1023) / 1024) * 1024;
  }
  static constexpr int smem_bytes(int stages) { return ring_bytes(stages) + kMaxBars * 8 + 16 + 1024; }
};

__device__ __forceinline__ uint32_t dsmem_addr(uint32_t local, uint32_t cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(r) : "r"(local), "r"(cta));
  return r;
}
__device__ __forceinline__ float ld_dsmem_f32(uint32_t addr) {
  float v;
  asm volatile("ld.shared::cluster.f32 %0, [%1];\n" : "=f"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void st_smem_f32(uint32_t addr, float v) {
  asm volatile("st.shared.f32 [%0], %1;\n" ::"r"(addr), "f"(v) : "memory");
}

// One cluster of SK CTAs per (m_tile, n_tile); cluster rank r accumulates k-blocks [r * kb_per, (r + 1) * kb_per).
// SILU: w = [gate (N rows) | up (N rows)], the B tile is BN/2 gate rows + the matching BN/2 up rows, y[:, BN/2 per tile].
template <int BN, int SK, bool SILU>
__global__ void __launch_bounds__(256, 1)
gemm_splitk_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                   bf16* __restrict__ y, int M, int N, int K, int64_t ldy, int num_m_tiles, int STAGES, const RowPush ps) {
  using C = SkCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sA = smem_base;
  const uint32_t sB = smem_base + STAGES * C::kABytes;
  const uint32_t bars = smem_base + C::ring_bytes(STAGES);
  auto full_bar = [&](int s) { return bars + s * 8; };
  auto empty_bar = [&](int s) { return bars + (STAGES + s) * 8; };
  const uint32_t tfull_bar = bars + 2 * STAGES * 8;
  const uint32_t tmem_slot = bars + (2 * STAGES + 1) * 8;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = static_cast<int>(cluster_ctarank());
  const int tile = static_cast<int>(blockIdx.x) / SK;
  const int m_tile = tile % num_m_tiles, n_tile = tile / num_m_tiles;
  const int kb_per = (K / C::BK) / SK, kb0 = rank * kb_per;
  TL_DECL
  if (threadIdx.x == 0) TL(0);   // kernel entry
  pdl_launch_dependents();

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_w);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    mbar_init(tfull_bar, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<C::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  if (threadIdx.x == 0) TL(1);   // prologue (barrier init, TMEM alloc) done

  if (warp == 0 && lane == 0) {
    // ===== TMA producer: weights first (independent of the predecessor kernel), x after the dependency wait =====
    auto load_w = [&](int stage, int kb) {
      if constexpr (SILU) {
        tma_load_2d(sB + stage * C::kBBytes, &tmap_w, full_bar(stage), kb * C::BK, n_tile * (BN / 2));
        tma_load_2d(sB + stage * C::kBBytes + (BN / 2) * 128, &tmap_w, full_bar(stage), kb * C::BK, N + n_tile * (BN / 2));
      } else {
        tma_load_2d(sB + stage * C::kBBytes, &tmap_w, full_bar(stage), kb * C::BK, n_tile * BN);
      }
    };
    const int pre = kb_per < STAGES ? kb_per : STAGES;
    for (int i = 0; i < pre; ++i) {
      mbar_expect_tx(full_bar(i), C::kStageBytes);
      load_w(i, kb0 + i);
    }
    pdl_wait();
    TL(2);   // predecessor kernel complete (x may be loaded)
    int stage = 0;
    uint32_t phase = 0;
    for (int i = 0; i < kb_per; ++i) {
      if (i >= pre) {
        mbar_wait(empty_bar(stage), phase ^ 1);
        mbar_expect_tx(full_bar(stage), C::kStageBytes);
        load_w(stage, kb0 + i);
      }
      tma_load_2d(sA + stage * C::kABytes, &tmap_x, full_bar(stage), (kb0 + i) * C::BK, m_tile * C::BM);
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1 && lane == 0) {
    // ===== MMA issuer =====
    constexpr uint32_t idesc = umma_idesc_bf16(C::BM, BN, 0, 0);
    int stage = 0;
    uint32_t phase = 0;
    pdl_wait();  // keep the wait for the predecessor out of the bounded mbarrier waits
    for (int i = 0; i < kb_per; ++i) {
      mbar_wait(full_bar(stage), phase);
      if (i == 0) TL(3);             // first operand stage landed
      if (i == kb_per - 1) TL(4);    // last operand stage landed
      tc_fence_after();
      const uint32_t a0 = sA + stage * C::kABytes, b0 = sB + stage * C::kBBytes;
#pragma unroll
      for (int k = 0; k < C::BK / 16; ++k)
        umma_bf16_ss(tmem_base, umma_desc_sw128(a0 + k * 32, 16, 1024), umma_desc_sw128(b0 + k * 32, 16, 1024), idesc,
                     (i | k) != 0 ? 1u : 0u);
      umma_commit(empty_bar(stage));
      if (i == kb_per - 1) umma_commit(tfull_bar);
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
  } else if (warp >= 4) {
    // ===== park the partial tile: TMEM -> f32 shared memory, column-major (lanes = rows: conflict-free) =====
    const int q = warp & 3;
    pdl_wait();               // (also needed below: y may still be read by an earlier kernel)
    mbar_wait(tfull_bar, 0);  // every MMA has retired: the operand ring is dead and can be overwritten
    if (threadIdx.x == 128) TL(5);   // accumulator complete
    tc_fence_after();
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const uint32_t prow = smem_base + (q * 32 + lane) * 4;
#pragma unroll 1
    for (int c = 0; c < BN / 32; ++c) {
      uint32_t r[32];
      tmem_ld_32x32b_x32(taddr + c * 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int e = 0; e < 32; ++e) st_smem_f32(prow + (c * 32 + e) * C::kColPitch, __uint_as_float(r[e]));
    }
    tc_fence_before();
  }
  __syncwarp();
  if (threadIdx.x == 128) TL(6);     // partial tile parked in shared memory
  cluster_sync_all();  // all SK partial tiles are parked and visible cluster-wide
  if (threadIdx.x == 128) TL(7);     // first cluster barrier passed

  if (warp >= 4) {
    // ===== reduce this CTA's column slice over the cluster in fixed rank order, then the usual epilogue =====
    pdl_wait();  // y may still be read by an earlier kernel until the predecessor has completed
    const int q = warp & 3;
    const int row_in = q * 32 + lane;
    const int row = m_tile * C::BM + row_in;
    uint32_t peer[SK];
#pragma unroll
    for (int r = 0; r < SK; ++r) peer[r] = dsmem_addr(smem_base + row_in * 4, static_cast<uint32_t>(r));
    if constexpr (SILU) {
      constexpr int CW = BN / (2 * SK);  // output columns of this CTA
      static_assert(CW >= 8 && CW % 8 == 0, "slice must be whole 16-byte vectors");
      bf16* dst = y + row * ldy + n_tile * (BN / 2) + rank * CW;
#pragma unroll 1
      for (int j = 0; j < CW; j += 8) {
        const int cg = rank * CW + j, cu = BN / 2 + cg;
        float vg[SK][8], vu[SK][8];
#pragma unroll
        for (int r = 0; r < SK; ++r)  // all 16 * SK loads of the chunk in flight before the first add
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            vg[r][e] = ld_dsmem_f32(peer[r] + (cg + e) * C::kColPitch);
            vu[r][e] = ld_dsmem_f32(peer[r] + (cu + e) * C::kColPitch);
          }
        float g[8], u[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          float sg = 0.f, su = 0.f;
#pragma unroll
          for (int r = 0; r < SK; ++r) { sg += vg[r][e]; su += vu[r][e]; }  // fixed rank order
          g[e] = sg; u[e] = su;
        }
        uint32_t o[4];
#pragma unroll
        for (int e = 0; e < 8; e += 2) {  // rounding points of the unfused path: bf16 gate_up, bf16 silu(gate), bf16 product
          const uint32_t gb = pack_bf16x2(g[e], g[e + 1]);
          const uint32_t ub = pack_bf16x2(u[e], u[e + 1]);
          const float g0 = bf16_lo(gb), g1 = bf16_hi(gb);
          const uint32_t sb = pack_bf16x2(g0 / (1.0f + expf(-g0)), g1 / (1.0f + expf(-g1)));
          o[e / 2] = pack_bf16x2(bf16_lo(sb) * bf16_lo(ub), bf16_hi(sb) * bf16_hi(ub));
        }
        if (row < M) *reinterpret_cast<uint4*>(dst + j) = make_uint4(o[0], o[1], o[2], o[3]);
      }
    } else {
      constexpr int CW = BN / SK;
      static_assert(CW >= 8 && CW % 8 == 0, "slice must be whole 16-byte vectors");
      constexpr int CH = (CW % 16 == 0 && SK <= 4) ? 16 : 8;  // columns per iteration: CH * SK coalesced loads in flight
      bf16* dst = y + row * ldy + n_tile * BN + rank * CW;
      bool remote = false;  // TP: the row is reduced by another rank -> straight into its low-latency stage buffer
      if (ps.ctrl != nullptr && row < M) {
        if (bf16* pr = push_row(ps, row)) { dst = pr + n_tile * BN + rank * CW; remote = true; }
      }
#pragma unroll 1
      for (int j = 0; j < CW; j += CH) {
        const int c0 = rank * CW + j;
        float v[SK][CH];
#pragma unroll
        for (int r = 0; r < SK; ++r)
#pragma unroll
          for (int e = 0; e < CH; ++e) v[r][e] = ld_dsmem_f32(peer[r] + (c0 + e) * C::kColPitch);
        float a[CH];
#pragma unroll
        for (int e = 0; e < CH; ++e) {
          float sum = 0.f;
#pragma unroll
          for (int r = 0; r < SK; ++r) sum += v[r][e];  // fixed rank order
          a[e] = sum;
        }
        if (row < M) {
#pragma unroll
          for (int e = 0; e < CH; e += 8) {
            const uint4 o = make_uint4(pack_bf16x2(a[e], a[e + 1]), pack_bf16x2(a[e + 2], a[e + 3]),
                                       pack_bf16x2(a[e + 4], a[e + 5]), pack_bf16x2(a[e + 6], a[e + 7]));
            if (remote) st_sys_16(dst + j + e, ll_scrub(o));
            else *reinterpret_cast<uint4*>(dst + j + e) = o;
          }
        }
      }
    }
  }
  __syncwarp();
  if (threadIdx.x == 128) TL(8);     // DSMEM reduce + global stores done
  cluster_sync_all();  // nobody leaves while a peer may still read its partial tile
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<C::kTmemCols>(tmem_base);
  }
#ifdef NVLLM_TIMELINE
  if (threadIdx.x == 128) TL(9);
  __syncthreads();
  if (threadIdx.x == 0 && g_tl_buf != nullptr) {
    const unsigned int slot = atomicAdd(&g_tl_cursor, 1u);
    if (slot < g_tl_cap) {
      unsigned long long* e = g_tl_buf + static_cast<size_t>(slot) * 12;
      unsigned int smid;
      asm volatile("mov.u32 %0, %%smid;\n" : "=r"(smid));
      e[0] = blockIdx.x;
      for (int i = 0; i < 10; ++i) e[1 + i] = tl_s[i];
      e[11] = smid;
    }
  }
#endif
}

// Stages of the operand ring: NVLLM_GEMM_SK_STAGES (benchmark knob) or the deepest that fits, never more than the
// k-blocks of one CTA.
template <int BN>
int splitk_stages(int kb_per) {
  using C = SkCfg<BN>;
  int want = knobs().sk_stages > 0 ? knobs().sk_stages : C::kMaxStages;
  if (want < 2) want = 2;
  if (want > C::kMaxStages) want = C::kMaxStages;
  return want < kb_per ? want : (kb_per < 1 ? 1 : kb_per);
}

template <int BN, int SK, bool SILU>
int launch_splitk(const CUtensorMap& tx, const CUtensorMap& tw, void* y, int M, int N, int K, int64_t ldy, int tiles,
                  int num_m_tiles, cudaStream_t st, const RowPush& ps) {
  using C = SkCfg<BN>;
  auto kern = gemm_splitk_kernel<BN, SK, SILU>;
  NVLLM_SMEM_ATTR_ONCE(kern, C::smem_bytes(C::kMaxStages));
  const int stages = splitk_stages<BN>((K / C::BK) / SK);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(tiles * SK);
  cfg.blockDim = dim3(256);
  cfg.dynamicSmemBytes = C::smem_bytes(stages);
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = SK; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  NVLLM_CUDA(cudaLaunchKernelEx(&cfg, kern, tx, tw, static_cast<bf16*>(y), M, N, K, ldy, num_m_tiles, stages, ps));
  return 0;
}

// Cost of one CTA in "bytes through the SM's TMA port" (~40-60 B/clk): the operands, plus the reduce-scatter of the
// partial tile over DSMEM — (SK-1)/SK of a BN x 128 f32 tile per CTA at ~20 B/clk (B300_MICROARCH.md: 17-21 B/clk), i.e.
// weighted x4 — which is why narrow tiles with SK = 4 win for the 0.6B projections.  Clusters of 8 measured 3-6 us
// slower than clusters of 4 at equal bytes (fewer can be co-scheduled per GPC): flat penalty.
// Data: profiles/r01_gemm_ingraph_v2_splitk.jsonl, r01_gemm_ingraph_v3_splitk_v2.jsonl.
long splitk_cost(int bn, int sk, int k) {
  return 2l * (128 + bn) * (k / sk) + (sk > 1 ? 4l * 512 * bn * (sk - 1) / sk : 0) + (sk >= 8 ? 600l * 1024 : 0);
}

}  // namespace

// Chooses (BN, SK) for y[m, n] = x . w^T (accumulator width n_acc = n, or 2 * inter for the SiLU form); returns false
// when no split-K configuration beats the single-CTA tile `bn_default` by a clear margin.  NVLLM_GEMM_SPLITK:
// "0" = never, "BN,SK" = force that configuration where it is legal, unset / "1" = the cost model below.
bool splitk_choose(int m, int n_acc, int k, bool silu, int bn_default, int* out_bn, int* out_sk) {
  const Knobs& kn = knobs();
  if (kn.splitk_mode == 0) return false;
  if (m > 512 || k % 64 != 0) return false;  // up to four 128-row tile rows (the TP decode batches of 512 sequences)
  const int sms = sm_count();
  const int mt = (m + 127) / 128, num_kb = k / 64;
  auto legal = [&](int bn, int sk) {
    if (n_acc % bn != 0 || num_kb % sk != 0 || num_kb / sk < 1) return false;
    if ((silu ? bn / (2 * sk) : bn / sk) < 8 || (bn == 64 && sk == 8)) return false;
    return static_cast<long>(mt) * (n_acc / bn) * sk <= sms;
  };
  if (kn.splitk_mode == 2) {
    const int fbn = kn.splitk_bn, fsk = kn.splitk_sk;
    const bool known = (fbn == 64 || fbn == 128 || fbn == 256) && (fsk == 2 || fsk == 4 || fsk == 8);
    if (known && legal(fbn, fsk)) { *out_bn = fbn; *out_sk = fsk; return true; }
    return false;
  }
  const int bns[3] = {256, 128, 64}, sks[3] = {8, 4, 2};
  long best = -1;
  for (int bn : bns)
    for (int sk : sks) {
      if (!legal(bn, sk)) continue;
      const long c = splitk_cost(bn, sk, k);
      if (best < 0 || c < best) { best = c; *out_bn = bn; *out_sk = sk; }
    }
  if (best < 0) return false;
  const long tiles_d = static_cast<long>(mt) * ((n_acc + bn_default - 1) / bn_default);
  const long waves_d = (tiles_d + sms - 1) / sms;
  return best * 10 < splitk_cost(bn_default, 1, k) * waves_d * 9;  // at least 10 % fewer bytes per SM
}

#ifdef NVLLM_TIMELINE
}  // namespace nvllm
extern "C" int nvllmdbg_timeline(unsigned long long* d_buf, unsigned int cap) {
  unsigned int zero = 0;
  if (cudaMemcpyToSymbol(nvllm::g_tl_buf, &d_buf, sizeof(d_buf)) != cudaSuccess) return 1;
  if (cudaMemcpyToSymbol(nvllm::g_tl_cap, &cap, sizeof(cap)) != cudaSuccess) return 1;
  if (cudaMemcpyToSymbol(nvllm::g_tl_cursor, &zero, sizeof(zero)) != cudaSuccess) return 1;
  return 0;
}
namespace nvllm {
#endif

int gemm_splitk_dispatch(void* y, const void* x, const void* w, int m, int n, int k, int64_t ldx, int64_t ldw, int64_t ldy,
                         bool silu, int bn, int sk, cudaStream_t st, const RowPush& ps) {
  CUtensorMap tx, tw;
  const int n_acc = silu ? 2 * n : n;
  if (int rc = make_tmap_bf16_2d(&tx, x, m, k, ldx, 128)) return rc;
  if (int rc = make_tmap_bf16_2d(&tw, w, n_acc, k, ldw, silu ? bn / 2 : bn)) return rc;
  const int mt = (m + 127) / 128, tiles = mt * (n_acc / bn);
#define NVLLM_SK(BNV, SKV)                                                                                   \
  if (bn == BNV && sk == SKV)                                                                                \
    return silu ? launch_splitk<BNV, SKV, true>(tx, tw, y, m, n, k, ldy, tiles, mt, st, RowPush{})                      \
                : launch_splitk<BNV, SKV, false>(tx, tw, y, m, n, k, ldy, tiles, mt, st, ps)
  NVLLM_SK(256, 8); NVLLM_SK(256, 4); NVLLM_SK(256, 2);
  NVLLM_SK(128, 8); NVLLM_SK(128, 4); NVLLM_SK(128, 2);
  NVLLM_SK(64, 4); NVLLM_SK(64, 2);
#undef NVLLM_SK
  NVLLM_CHECK_ARG(false, NVLLM_ERR_ARG, "gemm_splitk: no kernel for bn=%d sk=%d", bn, sk);
}

}  // namespace nvllm
