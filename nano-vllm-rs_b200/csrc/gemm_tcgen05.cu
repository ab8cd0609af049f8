// gemm_tcgen05.cu — K5: y[M,N] = x[M,K] . w[N,K]^T, bf16 x bf16 -> fp32 on the 5th-gen tensor cores.
//
// Replaces candle_nn::Linear behind LinearBase::forward (reference src/layers/linear.rs:228-239,354-356,
// 437-439) and ParallelLMHead::compute_logits (src/layers/embed_head.rs:292-306).
//
// Persistent, warp-specialised, one CTA per SM:
//   warp 0 (one lane)  TMA producer: cp.async.bulk.tensor 2-D tiles of x (128 x 64) and w (BN x 64) into a
//                      STAGES-deep shared-memory ring, SWIZZLE_128B, completion on `full` mbarriers
//   warp 1 (one lane)  MMA issuer: tcgen05.mma.cta_group::1.kind::f16, M=128, N=BN, K=16 x4 per stage,
//                      accumulators in TMEM (2 x BN columns, double buffered); tcgen05.commit releases
//                      the smem stage (`empty`) and, after the last k-block, publishes the tile (`tmem_full`)
//   warp 2             TMEM allocator
//   warps 4-7          epilogue: tcgen05.ld (32 lanes x 32 columns per warp) -> bf16/fp32 -> 16-byte global
//                      stores, overlapped with the next tile's MMAs through the second TMEM buffer
// Both operands are K-major (row-major x and row-major w), the canonical TN case: no transposes anywhere.
// Tiles are enumerated m-fastest so CTAs running together share the same w tile in L2; w is read from
// HBM once.  M and N tails: TMA zero-fills out-of-bounds rows, the epilogue predicates its stores.
#include "tc.cuh"

namespace nvllm {

template <int BN>
struct GemmCfg {
  static constexpr int BM = 128, BK = 64;
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStagesRaw = (200 * 1024) / kStageBytes;
  static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
  static constexpr int kTmemCols = (2 * BN <= 32) ? 32 : (2 * BN <= 64) ? 64 : (2 * BN <= 128) ? 128 : (2 * BN <= 256) ? 256 : 512;
  static constexpr int kBarBytes = (2 * kStages + 4) * 8 + 16;
  static constexpr int kSmemBytes = kStages * kStageBytes + kBarBytes + 1024;  // + alignment slack
};

template <typename OutT>
__device__ __forceinline__ void store_row32(OutT* dst, const uint32_t (&r)[32], int col0, int n);

template <>
__device__ __forceinline__ void store_row32<bf16>(bf16* dst, const uint32_t (&r)[32], int col0, int n) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    if (col0 + j * 8 < n) {
      uint4 v;
      v.x = pack_bf16x2(__uint_as_float(r[8 * j + 0]), __uint_as_float(r[8 * j + 1]));
      v.y = pack_bf16x2(__uint_as_float(r[8 * j + 2]), __uint_as_float(r[8 * j + 3]));
      v.z = pack_bf16x2(__uint_as_float(r[8 * j + 4]), __uint_as_float(r[8 * j + 5]));
      v.w = pack_bf16x2(__uint_as_float(r[8 * j + 6]), __uint_as_float(r[8 * j + 7]));
      *reinterpret_cast<uint4*>(dst + j * 8) = v;
    }
  }
}
// The same 32 columns into a peer's low-latency stage buffer (RowPush): scrubbed, system-scope stores.
__device__ __forceinline__ void push_row32(bf16* dst, const uint32_t (&r)[32], int col0, int n) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    if (col0 + j * 8 < n) {
      uint4 v;
      v.x = pack_bf16x2(__uint_as_float(r[8 * j + 0]), __uint_as_float(r[8 * j + 1]));
      v.y = pack_bf16x2(__uint_as_float(r[8 * j + 2]), __uint_as_float(r[8 * j + 3]));
      v.z = pack_bf16x2(__uint_as_float(r[8 * j + 4]), __uint_as_float(r[8 * j + 5]));
      v.w = pack_bf16x2(__uint_as_float(r[8 * j + 6]), __uint_as_float(r[8 * j + 7]));
      st_sys_16(dst + j * 8, ll_scrub(v));
    }
  }
}
template <>
__device__ __forceinline__ void store_row32<float>(float* dst, const uint32_t (&r)[32], int col0, int n) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    if (col0 + j * 4 < n) {
      *reinterpret_cast<uint4*>(dst + j * 4) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
    }
  }
}

// SILU == true: w = [gate (N rows) | up (N rows)]; the B tile is staged as BN/2 gate rows + the matching BN/2 up rows and
// the epilogue writes y[:, n_tile*BN/2 + c] = silu(gate) * up (see gemm_tcgen05_2cta.cu for the CTA-pair variant).
// Optional qkv epilogue (ROPE mode, BN == 128 == head_dim): each output tile is one head of the packed [q | k | v] row,
// and each epilogue thread owns that head for one token.  It applies what Qwen3Attention::forward does between qkv_proj
// and Attention::forward (qwen3.rs:219-229) plus Attention::store_kv_cache (attention.rs:150-174): round to bf16 (the
// qkv tensor), per-head RMSNorm (q/k), bf16, half-split RoPE from the f32 table, bf16; q/k/v rows go to the qkv buffer
// and k/v additionally into the paged cache at slot_mapping[token] — same rounding points as the stand-alone kernel.
struct RopeEpi {
  const int64_t* positions;
  const float* cos_sin;     // [max_position, 128]
  const bf16* q_norm_w;     // may be NULL (no qk-norm)
  const bf16* k_norm_w;
  bf16* k_cache;
  bf16* v_cache;
  const int32_t* slot_mapping;
  int H, Hkv, max_position, block_size;
  int64_t blk_stride, tok_stride, head_stride;
  float eps;
};

// CN > 1: clusters of CN CTAs along N work on the same 128-row x tile.  Each CTA loads 128/CN rows of it and MULTICASTS
// them into all CN shared memories (L2 -> SM traffic for x drops by CN: the bound for small-N, long-K shards such as the
// TP=8 qkv / o / down projections); a stage is released only when all CN CTAs' MMAs have retired (multicast commit).
template <int BN, typename OutT, bool SILU, int CN, bool ROPE = false>
__global__ void __launch_bounds__(256, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                    OutT* __restrict__ y, int M, int N, int K, int64_t ldy, int num_m_tiles, int num_tiles,
                    const RopeEpi re = RopeEpi{}, int STAGES = GemmCfg<BN>::kStages, const RowPush ps = RowPush{}) {
  using C = GemmCfg<BN>;  // STAGES <= C::kStages: ring depth of this launch (fewer stages -> two CTAs per SM, see launch_gemm)
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B tiles: 1024-byte aligned
  const uint32_t sA = smem_base;
  const uint32_t sB = smem_base + STAGES * C::kABytes;
  const uint32_t bars = smem_base + STAGES * C::kStageBytes;
  auto full_bar = [&](int s) { return bars + s * 8; };
  auto empty_bar = [&](int s) { return bars + (STAGES + s) * 8; };
  auto tfull_bar = [&](int a) { return bars + (2 * STAGES + a) * 8; };
  auto tempty_bar = [&](int a) { return bars + (2 * STAGES + 2 + a) * 8; };
  const uint32_t tmem_slot = bars + (2 * STAGES + 4) * 8;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_kb = K / C::BK;
  // num_tiles counts CLUSTER tiles: (m_tile, group of CN consecutive n-tiles); this CTA owns n-tile group*CN + rank
  const int rank = CN > 1 ? static_cast<int>(cluster_ctarank()) : 0;
  const int cta0 = static_cast<int>(blockIdx.x) / CN, cta_stride = static_cast<int>(gridDim.x) / CN;
  constexpr uint16_t kMask = static_cast<uint16_t>((1u << CN) - 1u);
  pdl_launch_dependents();  // everything up to pdl_wait() below is independent of the previous kernel

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_w);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), CN);  // every CTA of the cluster must have drained the stage before it is refilled
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 128);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<C::kTmemCols>(tmem_slot);
  tc_fence_before();
  if constexpr (CN > 1) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0 && lane == 0) {
    // ===== TMA producer =====
    // The weights do not depend on the previous kernel: the first ring of w tiles is requested BEFORE the
    // programmatic-dependency wait, so their HBM latency hides behind the predecessor's tail; x follows after it.
    auto load_w = [&](int stage, int kb, int n_tile) {
      if constexpr (SILU) {
        tma_load_2d(sB + stage * C::kBBytes, &tmap_w, full_bar(stage), kb * C::BK, n_tile * (BN / 2));
        tma_load_2d(sB + stage * C::kBBytes + (BN / 2) * 128, &tmap_w, full_bar(stage), kb * C::BK, N + n_tile * (BN / 2));
      } else {
        tma_load_2d(sB + stage * C::kBBytes, &tmap_w, full_bar(stage), kb * C::BK, n_tile * BN);
      }
    };
    auto load_x = [&](int stage, int kb, int m_tile) {
      if constexpr (CN > 1) {
        constexpr int kRows = C::BM / CN;  // this CTA's share of the x tile, multicast to the whole cluster
        tma_load_2d_mcast(sA + stage * C::kABytes + rank * kRows * 128, &tmap_x, full_bar(stage), kb * C::BK,
                          m_tile * C::BM + rank * kRows, kMask);
      } else {
        tma_load_2d(sA + stage * C::kABytes, &tmap_x, full_bar(stage), kb * C::BK, m_tile * C::BM);
      }
    };
    const int pre = num_kb < STAGES ? num_kb : STAGES;
    {
      const int n_tile0 = (cta0 / num_m_tiles) * CN + rank;
      for (int kb = 0; kb < pre; ++kb) {
        mbar_expect_tx(full_bar(kb), C::kStageBytes);
        load_w(kb, kb, n_tile0);
      }
    }
    pdl_wait();
    int stage = 0;
    uint32_t phase = 0;
    bool first = true;
    for (int tile = cta0; tile < num_tiles; tile += cta_stride) {
      const int m_tile = tile % num_m_tiles, n_tile = (tile / num_m_tiles) * CN + rank;
      for (int kb = 0; kb < num_kb; ++kb) {
        if (first && kb < pre) {  // w tile already in flight, barrier already armed
          load_x(stage, kb, m_tile);
        } else {
          mbar_wait(empty_bar(stage), phase ^ 1);
          mbar_expect_tx(full_bar(stage), C::kStageBytes);
          load_x(stage, kb, m_tile);
          load_w(stage, kb, n_tile);
        }
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      first = false;
    }
  } else if (warp == 1 && lane == 0) {
    // ===== MMA issuer (single thread) =====
    constexpr uint32_t idesc = umma_idesc_bf16(C::BM, BN, 0, 0);
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    pdl_wait();  // no operand can land before the predecessor is done: keep that wait out of the bounded mbarrier waits
    for (int tile = cta0; tile < num_tiles; tile += cta_stride) {
      mbar_wait(tempty_bar(acc), acc_phase ^ 1);  // epilogue has drained this TMEM buffer
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
          umma_bf16_ss(tmem_d, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
        }
        if constexpr (CN > 1) umma_commit_mcast(empty_bar(stage), kMask);  // tell every producer of the cluster
        else umma_commit(empty_bar(stage));                 // smem stage reusable once these MMAs retire
        if (kb == num_kb - 1) umma_commit(tfull_bar(acc));  // accumulator complete
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else if (warp >= 4) {
    // ===== epilogue: TMEM -> registers -> global =====
    pdl_wait();              // y may still be read by an earlier kernel until the predecessor has completed
    const int q = warp & 3;  // TMEM lane quadrant this warp may access
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = cta0; tile < num_tiles; tile += cta_stride) {
      const int m_tile = tile % num_m_tiles, n_tile = (tile / num_m_tiles) * CN + rank;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const int row = m_tile * C::BM + q * 32 + lane;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;
      if constexpr (ROPE) {
        static_assert(!ROPE || (BN == 128 && sizeof(OutT) == 2), "ROPE epilogue: one 128-wide head per tile, bf16 out");
        float v[128];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t r[32];
          tmem_ld_32x32b_x32(taddr + c * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int e = 0; e < 32; ++e) v[c * 32 + e] = __bfloat162float(__float2bfloat16_rn(__uint_as_float(r[e])));  // bf16 qkv
        }
        if (row < M) {
          const int head = n_tile;  // packed row: H q heads, Hkv k heads, Hkv v heads
          const int kind = head < re.H ? 0 : (head < re.H + re.Hkv ? 1 : 2);
          if (kind < 2) {
            const bf16* w = kind == 0 ? re.q_norm_w : re.k_norm_w;
            if (w != nullptr) {
              float ss = 0.f;
#pragma unroll
              for (int e = 0; e < 128; ++e) ss += v[e] * v[e];
              const float inv = 1.0f / sqrtf(ss * (1.0f / 128.0f) + re.eps);
#pragma unroll
              for (int e = 0; e < 128; ++e) v[e] = __bfloat162float(__float2bfloat16_rn(v[e] * inv * __bfloat162float(w[e])));
            }
            int64_t pos = re.positions[row];
            pos = pos < 0 ? 0 : (pos >= re.max_position ? re.max_position - 1 : pos);
            const float4* cs = reinterpret_cast<const float4*>(re.cos_sin + pos * 128);
#pragma unroll
            for (int e4 = 0; e4 < 16; ++e4) {
              const float4 c4 = __ldg(cs + e4), s4 = __ldg(cs + 16 + e4);
              const float cc[4] = {c4.x, c4.y, c4.z, c4.w}, sn[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const int e = e4 * 4 + i;
                const float x1 = v[e], x2 = v[64 + e];
                v[e] = x1 * cc[i] - x2 * sn[i];
                v[64 + e] = x2 * cc[i] + x1 * sn[i];
              }
            }
          }
          uint4 o[16];
#pragma unroll
          for (int c = 0; c < 16; ++c) {
            o[c].x = pack_bf16x2(v[c * 8 + 0], v[c * 8 + 1]); o[c].y = pack_bf16x2(v[c * 8 + 2], v[c * 8 + 3]);
            o[c].z = pack_bf16x2(v[c * 8 + 4], v[c * 8 + 5]); o[c].w = pack_bf16x2(v[c * 8 + 6], v[c * 8 + 7]);
          }
          uint4* dst = reinterpret_cast<uint4*>(y + row * ldy + head * 128);
#pragma unroll
          for (int c = 0; c < 16; ++c) dst[c] = o[c];
          if (kind >= 1 && re.k_cache != nullptr) {
            const int slot = __ldg(re.slot_mapping + row);
            if (slot >= 0) {
              const int blk = slot / re.block_size, off = slot - blk * re.block_size;
              const int kvh = kind == 1 ? head - re.H : head - re.H - re.Hkv;
              uint4* cd = reinterpret_cast<uint4*>((kind == 1 ? re.k_cache : re.v_cache) + blk * re.blk_stride +
                                                   off * re.tok_stride + kvh * re.head_stride);
#pragma unroll
              for (int c = 0; c < 16; ++c) cd[c] = o[c];
            }
          }
        }
      } else if constexpr (SILU) {
#pragma unroll 1
        for (int c = 0; c < BN / 64; ++c) {
          uint32_t g[32], u[32];
          tmem_ld_32x32b_x32(taddr + c * 32, g);
          tmem_ld_32x32b_x32(taddr + BN / 2 + c * 32, u);
          tmem_ld_wait();
#pragma unroll
          for (int e = 0; e < 32; e += 2) {  // rounding points of the unfused path: bf16 gate_up, bf16 silu(gate)
            const uint32_t gb = pack_bf16x2(__uint_as_float(g[e]), __uint_as_float(g[e + 1]));
            const uint32_t ub = pack_bf16x2(__uint_as_float(u[e]), __uint_as_float(u[e + 1]));
            const float g0 = bf16_lo(gb), g1 = bf16_hi(gb);
            const uint32_t sb = pack_bf16x2(g0 / (1.0f + expf(-g0)), g1 / (1.0f + expf(-g1)));
            g[e] = __float_as_uint(bf16_lo(sb) * bf16_lo(ub));
            g[e + 1] = __float_as_uint(bf16_hi(sb) * bf16_hi(ub));
          }
          const int col0 = n_tile * (BN / 2) + c * 32;
          if (row < M && col0 < N) store_row32<OutT>(y + row * ldy + col0, g, col0, N);
        }
      } else {
        bf16* remote = nullptr;  // TP: this row is reduced by another rank -> straight into its stage buffer
        if constexpr (sizeof(OutT) == 2) {
          if (ps.ctrl != nullptr && row < M) remote = push_row(ps, row);
        }
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t r[32];
          tmem_ld_32x32b_x32(taddr + c * 32, r);
          tmem_ld_wait();
          const int col0 = n_tile * BN + c * 32;
          if (row < M && col0 < N) {
            if (remote != nullptr) push_row32(remote + col0, r, col0, N);
            else store_row32<OutT>(y + row * ldy + col0, r, col0, N);
          }
        }
      }
      tc_fence_before();
      mbar_arrive(tempty_bar(acc));
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  if constexpr (CN > 1) cluster_sync_all(); else __syncthreads();  // peers may still multicast into / signal this CTA
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<C::kTmemCols>(tmem_base);
  }
}

template <int BN, typename OutT, bool SILU = false, int CN = 1, bool ROPE = false>
static int launch_gemm(const CUtensorMap& tx, const CUtensorMap& tw, void* y, int M, int N, int K, int64_t ldy,
                       cudaStream_t st, const RopeEpi& re = RopeEpi{}, const RowPush& ps = RowPush{}) {
  using C = GemmCfg<BN>;
  auto kern = gemm_tcgen05_kernel<BN, OutT, SILU, CN, ROPE>;
  NVLLM_SMEM_ATTR_ONCE(kern, C::kSmemBytes);
  const int num_m_tiles = (M + C::BM - 1) / C::BM;
  const int num_n_tiles = SILU ? (N + BN / 2 - 1) / (BN / 2) : (N + BN - 1) / BN;
  const int num_tiles = num_m_tiles * (num_n_tiles / CN);  // cluster tiles (the caller guarantees num_n_tiles % CN == 0)
  const int max_clusters = sm_count() / CN;
  const int clusters = num_tiles < max_clusters ? num_tiles : max_clusters;
  // Ring depth: NVLLM_GEMM_STAGES (benchmark knob) caps it for single-wave (decode-sized) launches: with <= ~100 KB of
  // shared memory two CTAs fit on an SM, so the NEXT kernel's CTAs become resident (prologue + weight prefetch under PDL)
  // while this kernel's are still running.
  int stages = C::kStages;
  if (CN == 1 && num_tiles <= max_clusters) {
    const int want = knobs().gemm_stages;
    if (want >= 2 && want < stages) stages = want;
  }
  const int smem = C::kSmemBytes - (C::kStages - stages) * C::kStageBytes;
  if constexpr (CN == 1) {
    NVLLM_CUDA(launch_pdl(kern, dim3(clusters), dim3(256), smem, st, tx, tw, static_cast<OutT*>(y), M, N, K, ldy,
                          num_m_tiles, num_tiles, re, stages, ps));
  } else {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(clusters * CN);
    cfg.blockDim = dim3(256);
    cfg.dynamicSmemBytes = C::kSmemBytes;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CN; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    NVLLM_CUDA(cudaLaunchKernelEx(&cfg, kern, tx, tw, static_cast<OutT*>(y), M, N, K, ldy, num_m_tiles, num_tiles, re,
                                  static_cast<int>(C::kStages), ps));
  }
  return 0;
}

static int pick_bn(int M, int N) {
  const int forced = knobs().gemm_bn;  // benchmarking knob (nvllm_reload_env() lets one process sweep it)
  if (forced == 32 || forced == 64 || forced == 128 || forced == 192 || forced == 256) return forced;
  const int sms = sm_count();
  const long mt = (M + 127) / 128;
  const int cands[5] = {32, 64, 128, 192, 256};
  // Small problems (decode): the narrowest tile that still fits in ONE wave — maximum number of CTAs streaming the
  // weights concurrently, no second (serialised, latency-bound) pass over the persistent loop.  (192 exists for the
  // 512-row TP batches: 4 x N/256 tiles often fill only 2/3 of the SMs, or spill a third of a wave into a second one.)
  for (int i = 0; i < 5; ++i) {
    const long tiles = mt * ((N + cands[i] - 1) / cands[i]);
    if (tiles <= sms) return cands[i];
  }
  // A few waves (M = 512 decode batches on the wide projections): the makespan is waves x tile width.
  long best_cost = 0;
  int best = 256;
  for (int i = 4; i >= 2; --i) {
    const long tiles = mt * ((N + cands[i] - 1) / cands[i]);
    const long waves = (tiles + sms - 1) / sms;
    if (waves > 4) continue;
    const long cost = waves * cands[i];
    if (best_cost == 0 || cost < best_cost) { best_cost = cost; best = cands[i]; }  // ties keep the wider tile
  }
  if (best_cost != 0) return best;
  // Large problems (prefill, LM head): widest tile = highest MMA efficiency and operand reuse.
  return 256;
}

int gemm_2cta_dispatch(void* y, const void* x, const void* w, int m, int n, int k, int64_t ldx, int64_t ldw,
                       int64_t ldy, int out_dtype, cudaStream_t st);  // gemm_tcgen05_2cta.cu

int gemm_silu_2cta_dispatch(void* y, const void* x, const void* w, int m, int inter, int k, int64_t ldx, int64_t ldw,
                            int64_t ldy, cudaStream_t st);  // gemm_tcgen05_2cta.cu

// gemm_splitk.cu: cluster split-K for decode-sized projections (m <= 256)
bool splitk_choose(int m, int n_acc, int k, bool silu, int bn_default, int* out_bn, int* out_sk);
int gemm_splitk_dispatch(void* y, const void* x, const void* w, int m, int n, int k, int64_t ldx, int64_t ldw, int64_t ldy,
                         bool silu, int bn, int sk, cudaStream_t st, const RowPush& ps = RowPush{});

static bool use_2cta();
// CTA pairs pay off once there are at least two 256-row tiles or the problem is large enough to fill the machine with
// 256 x 256 tiles; between 129 and 511 rows the single-CTA kernel (more, narrower, single-wave tiles) measured faster
// inside the decode step.  Skinny problems with a very long K (8B down_proj at decode: M = 512, K = 12288 > 2N) stream x
// from L2 once per n-tile and also do better on single-CTA tiles until a split-K variant exists.
static bool want_2cta(int m, int n, int k) {
  if (m <= 128 || !use_2cta()) return false;
  if (knobs().gemm_2cta == 2) return true;  // benchmark knob: CTA pairs for every m > 128
  // In-graph measurements (profiles/r01_gemm_ingraph_v2_splitk.jsonl): at M = 512 (the TP decode batches) the single-CTA
  // kernel is 7-17 % faster on all four projections; CTA pairs pay off from two 256-row tile rows upwards (prefill).
  if (m <= 512) return false;
  (void)n; (void)k;
  return true;
}

static bool use_mcast() {
  // x-multicast clusters: OFF by default.  In-graph they measured 5-7 % SLOWER than plain tiles on every shape they
  // apply to (the bytes still have to enter every SM; the bound is the SM's TMA ingest, not L2 reads) and the long-K
  // narrow-N projections they were built for now go to the cluster split-K kernel.  NVLLM_GEMM_MCAST=1 enables them.
  return knobs().gemm_mcast != 0;
}

static bool use_2cta() {
  return knobs().gemm_2cta != 0;  // A/B knob: 0 keeps every shape on the single-CTA kernel
}

}  // namespace nvllm

extern "C" size_t nvllm_gemm_workspace_bytes(int32_t, int32_t, int32_t) { return 0; }

namespace nvllm {
// y = x . w^T.  ps.ctrl != nullptr (bf16 output, n == ldy == ps.hidden, m <= 512): rows another TP rank reduces are
// pushed into that rank's low-latency stage buffer by the epilogue instead of being written to y (see RowPush).
int gemm_bf16_impl(void* y, const void* x, const void* w, int32_t m, int32_t n, int32_t k, int64_t ldx, int64_t ldw,
                   int64_t ldy, int32_t out_dtype, void* stream, const RowPush& ps) {
  const bool push = ps.ctrl != nullptr;
  NVLLM_CHECK_ARG(m >= 0 && n > 0 && k > 0, NVLLM_ERR_ARG, "gemm: m=%d n=%d k=%d", m, n, k);
  NVLLM_CHECK_ARG(k % 64 == 0 && n % 8 == 0, NVLLM_ERR_ARG, "gemm: need k %% 64 == 0 and n %% 8 == 0 (k=%d n=%d)", k, n);
  NVLLM_CHECK_ARG(ldx >= k && ldw >= k && ldy >= n && ldx % 8 == 0 && ldw % 8 == 0 && ldy % 8 == 0, NVLLM_ERR_ALIGN,
                  "gemm: leading dimensions ldx=%lld ldw=%lld ldy=%lld", (long long)ldx, (long long)ldw, (long long)ldy);
  NVLLM_CHECK_ARG(out_dtype == NVLLM_BF16 || out_dtype == NVLLM_F32, NVLLM_ERR_DTYPE, "gemm: out_dtype %d", out_dtype);
  if (m == 0) return 0;
  NVLLM_CHECK_ARG(y && x && w && aligned16(y) && aligned16(x) && aligned16(w), NVLLM_ERR_ALIGN, "gemm: pointers");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (out_dtype == NVLLM_BF16) {  // decode-sized, long-K projections: split K over a cluster (gemm_splitk.cu)
    int sbn = 0, ssk = 0;
    if (splitk_choose(m, n, k, false, pick_bn(m, n), &sbn, &ssk))
      return gemm_splitk_dispatch(y, x, w, m, n, k, ldx, ldw, ldy, false, sbn, ssk, st, ps);
  }
  if (!push && want_2cta(m, n, k)) return gemm_2cta_dispatch(y, x, w, m, n, k, ldx, ldw, ldy, out_dtype, st);
  const int bn = pick_bn(m, n);
  CUtensorMap tx, tw;
  // x-multicast clusters of 4 along N: worth it when x (re-read once per n-tile) dominates the L2 -> SM traffic,
  // i.e. several n-tiles of a narrow width and a long K.
  if (!push && use_mcast() && bn <= 64 && n % (4 * bn) == 0 && k >= 2048 && out_dtype == NVLLM_BF16) {
    if (int rc = make_tmap_bf16_2d(&tx, x, m, k, ldx, 32)) return rc;
    if (int rc = make_tmap_bf16_2d(&tw, w, n, k, ldw, bn)) return rc;
    return bn == 64 ? launch_gemm<64, bf16, false, 4>(tx, tw, y, m, n, k, ldy, st)
                    : launch_gemm<32, bf16, false, 4>(tx, tw, y, m, n, k, ldy, st);
  }
  if (int rc = make_tmap_bf16_2d(&tx, x, m, k, ldx, 128)) return rc;
  if (int rc = make_tmap_bf16_2d(&tw, w, n, k, ldw, bn)) return rc;
#define NVLLM_GEMM_CASE(BNV)                                                                     \
  case BNV:                                                                                      \
    return out_dtype == NVLLM_BF16 ? launch_gemm<BNV, bf16>(tx, tw, y, m, n, k, ldy, st, RopeEpi{}, ps) \
                                   : launch_gemm<BNV, float>(tx, tw, y, m, n, k, ldy, st)
  switch (bn) {
    NVLLM_GEMM_CASE(256);
    NVLLM_GEMM_CASE(192);
    NVLLM_GEMM_CASE(128);
    NVLLM_GEMM_CASE(64);
    default:
      NVLLM_GEMM_CASE(32);
  }
#undef NVLLM_GEMM_CASE
}
}  // namespace nvllm

extern "C" int nvllm_gemm_bf16(void* y, const void* x, const void* w, int32_t m, int32_t n, int32_t k,
                               int64_t ldx, int64_t ldw, int64_t ldy, int32_t out_dtype, void* workspace,
                               size_t workspace_bytes, void* stream) {
  (void)workspace; (void)workspace_bytes;
  return nvllm::gemm_bf16_impl(y, x, w, m, n, k, ldx, ldw, ldy, out_dtype, stream, nvllm::RowPush{});
}

// MergedColumnParallelLinear::forward + SiluAndMul::forward in one call (linear.rs:437-439 + activation.rs:46-63).
extern "C" int nvllm_gemm_silu_mul_bf16(void* out, const void* x, const void* w_gate_up, int32_t m, int32_t inter,
                                        int32_t k, int64_t ldx, int64_t ldw, int64_t ldout, void* workspace,
                                        size_t workspace_bytes, void* stream) {
  using namespace nvllm;
  NVLLM_CHECK_ARG(m >= 0 && inter > 0 && k > 0 && k % 64 == 0 && inter % 8 == 0, NVLLM_ERR_ARG,
                  "gemm_silu_mul: m=%d inter=%d k=%d (k %% 64, inter %% 8)", m, inter, k);
  NVLLM_CHECK_ARG(ldx >= k && ldw >= k && ldout >= inter && ldx % 8 == 0 && ldw % 8 == 0 && ldout % 8 == 0,
                  NVLLM_ERR_ALIGN, "gemm_silu_mul: leading dimensions");
  if (m == 0) return 0;
  NVLLM_CHECK_ARG(out && x && w_gate_up && aligned16(out) && aligned16(x) && aligned16(w_gate_up), NVLLM_ERR_ALIGN,
                  "gemm_silu_mul: pointers");
  (void)workspace; (void)workspace_bytes;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (inter % 128 == 0) {  // fused epilogue: the [m, 2*inter] tensor never exists
    const int bn = pick_bn(m, 2 * inter) < 64 ? 64 : pick_bn(m, 2 * inter);
    {
      int sbn = 0, ssk = 0;
      if (splitk_choose(m, 2 * inter, k, true, bn, &sbn, &ssk))
        return gemm_splitk_dispatch(out, x, w_gate_up, m, inter, k, ldx, ldw, ldout, true, sbn, ssk, st);
    }
    if (want_2cta(m, 2 * inter, k)) return gemm_silu_2cta_dispatch(out, x, w_gate_up, m, inter, k, ldx, ldw, ldout, st);
    CUtensorMap tx, tw;
    if (int rc = make_tmap_bf16_2d(&tx, x, m, k, ldx, 128)) return rc;
    if (int rc = make_tmap_bf16_2d(&tw, w_gate_up, 2 * static_cast<int64_t>(inter), k, ldw, bn / 2)) return rc;
    switch (bn) {
      case 256: return launch_gemm<256, bf16, true>(tx, tw, out, m, inter, k, ldout, st);
      case 192: return launch_gemm<192, bf16, true>(tx, tw, out, m, inter, k, ldout, st);
      case 128: return launch_gemm<128, bf16, true>(tx, tw, out, m, inter, k, ldout, st);
      default: return launch_gemm<64, bf16, true>(tx, tw, out, m, inter, k, ldout, st);
    }
  }
  // odd intermediate sizes: plain GEMM into the workspace, then the SiLU*mul kernel
  const size_t need = static_cast<size_t>(m) * 2 * inter * 2;
  NVLLM_CHECK_ARG(workspace && workspace_bytes >= need && aligned16(workspace), NVLLM_ERR_WORKSPACE,
                  "gemm_silu_mul: workspace %zu < %zu bytes", workspace_bytes, need);
  NVLLM_CHECK_ARG(ldout == inter, NVLLM_ERR_ARG, "gemm_silu_mul: unfused path needs a contiguous output");
  if (int rc = nvllm_gemm_bf16(workspace, x, w_gate_up, m, 2 * inter, k, ldx, ldw, 2 * inter, NVLLM_BF16, nullptr, 0, stream)) return rc;
  return nvllm_silu_mul(out, workspace, m, inter, NVLLM_BF16, stream);
}

// QKVParallelLinear::forward (linear.rs:354-356) + q/k norm + RotaryEmbedding::forward + Attention::store_kv_cache
// (qwen3.rs:219-229, attention.rs:150-174) in one kernel: qkv = x . Wqkv^T with the RopeEpi epilogue above.
extern "C" int nvllm_gemm_qkv_rope_store(void* qkv, const void* x, const void* w_qkv, int32_t m, int32_t k, int64_t ldx,
                                         int64_t ldw, const int64_t* positions, const float* cos_sin, const void* q_norm_w,
                                         const void* k_norm_w, float eps, int32_t num_heads, int32_t num_kv_heads,
                                         int32_t head_dim, int32_t max_position, void* k_cache, void* v_cache,
                                         const int32_t* slot_mapping, const nvllm_kv_layout* layout, void* stream) {
  using namespace nvllm;
  NVLLM_CHECK_ARG(head_dim == 128, NVLLM_ERR_ARG, "gemm_qkv_rope_store: head_dim %d (128 only)", head_dim);
  NVLLM_CHECK_ARG(m >= 0 && k > 0 && k % 64 == 0 && num_heads > 0 && num_kv_heads > 0 && max_position > 0, NVLLM_ERR_ARG,
                  "gemm_qkv_rope_store: sizes");
  if (m == 0) return 0;
  NVLLM_CHECK_ARG(qkv && x && w_qkv && positions && cos_sin && aligned16(qkv) && aligned16(x) && aligned16(w_qkv),
                  NVLLM_ERR_ALIGN, "gemm_qkv_rope_store: pointers");
  NVLLM_CHECK_ARG((q_norm_w == nullptr) == (k_norm_w == nullptr), NVLLM_ERR_ARG, "gemm_qkv_rope_store: q/k norm weights");
  RopeEpi re{};
  re.positions = positions; re.cos_sin = cos_sin;
  re.q_norm_w = static_cast<const bf16*>(q_norm_w); re.k_norm_w = static_cast<const bf16*>(k_norm_w);
  re.H = num_heads; re.Hkv = num_kv_heads; re.max_position = max_position; re.eps = eps; re.block_size = 1;
  if (k_cache != nullptr) {
    NVLLM_CHECK_ARG(layout && v_cache && slot_mapping && layout->head_dim == 128 && layout->num_kv_heads == num_kv_heads &&
                        layout->block_size > 0 && layout->token_stride % 8 == 0 && layout->head_stride % 8 == 0 &&
                        layout->block_stride % 8 == 0,
                    NVLLM_ERR_ARG, "gemm_qkv_rope_store: cache layout");
    re.k_cache = static_cast<bf16*>(k_cache); re.v_cache = static_cast<bf16*>(v_cache); re.slot_mapping = slot_mapping;
    re.block_size = layout->block_size; re.blk_stride = layout->block_stride; re.tok_stride = layout->token_stride;
    re.head_stride = layout->head_stride;
  }
  const int n = (num_heads + 2 * num_kv_heads) * 128;
  CUtensorMap tx, tw;
  if (int rc = make_tmap_bf16_2d(&tx, x, m, k, ldx, 128)) return rc;
  if (int rc = make_tmap_bf16_2d(&tw, w_qkv, n, k, ldw, 128)) return rc;
  return launch_gemm<128, bf16, false, 1, true>(tx, tw, qkv, m, n, k, n, static_cast<cudaStream_t>(stream), re);
}
