// prefill_attn.cu — K3 / K4: varlen causal flash-attention forward on tcgen05 / TMEM (tensor-bound).
//
// Replaces Attention::flash_attention_varlen + create_causal_mask (reference src/layers/attention.rs:177-208,
// 321-339: per-sequence loop, O(L^2) host-built mask, dense matmul) and flash_attention_varlen_with_cache
// (:211-222, prefill of the uncached suffix against the paged prefix).
//
// One CTA = 256 query rows of one (sequence, head): two row groups (RG) of 128 threads, one query row per thread
// (== one TMEM lane), sharing the K/V tiles of the KV head; KV tiles of 128 keys.
//   warp 8 (one lane)   TMA producer: Q once (2 x 128 x 128), then K_0, V_0, K_1, V_1, ... through a ring of NSLOT
//                       32 KiB slots; dense K/V via a 2-D map, paged K/V via a 4-D map (d, token, head, block) + block_tables
//   warp 9 (one lane)   MMA issuer for both groups: S = Q . K^T (8 x tcgen05.mma M128 N128 K16, both operands K-major in
//                       shared memory) and O += P . V (8 x tcgen05.mma with A = P read from TENSOR MEMORY, V MN-major)
//   RG threads          tcgen05.ld S row -> row max -> exp2 (3/4 on MUFU, 1/4 by an XU-free polynomial on the FMA pipe) ->
//                       P as packed bf16 written BACK INTO TMEM over the S columns (tcgen05.st; each thread only touches its
//                       own lane, so no cross-thread hazard) -> mbarrier arrive.  No shared-memory round trip for P.
//                       O accumulates in TMEM across tiles; the reference maximum is refreshed lazily (only when a tile
//                       exceeds it by 2^8), and only then is O read, scaled and written back.
// While one group runs its softmax the tensor core executes the other group's MMAs.
// TMEM: 512 columns = 2 RG x (S/P 128 + O 128).  Causal mask is bottom-right aligned (row r of Lq sees keys <= Lk - Lq + r)
// so the same kernel serves Lq == Lk and suffix-over-prefix.  CTAs are ordered longest-first.
#include "tc.cuh"

namespace nvllm {

int check_layout(const nvllm_kv_layout* L);  // store_kvcache.cu

namespace pf {
constexpr int D = 128, BM = 128, BN = 128, NSLOT = 4;
constexpr int kQBytes = BM * D * 2;          // one RG's Q tile: 2 halves of [128 x 64]
constexpr int kTileBytes = BN * D * 2;       // one K (or V) tile: 2 halves of [128 x 64]
constexpr int kSmemQ = 0;
constexpr int kSmemKV = kSmemQ + 2 * kQBytes;
constexpr int kSmemBar = kSmemKV + NSLOT * kTileBytes;
constexpr int kNumBars = 1 + 2 * NSLOT + 6;
constexpr int kSmemBytes = kSmemBar + kNumBars * 8 + 16 + 1024;
constexpr int kThreads = 384;  // 2 softmax warpgroups + 1 warpgroup holding the TMA and MMA warps (2 idle warps)
}  // namespace pf

struct PrefillParams {
  bf16* out;
  const int32_t* cu_q;
  const int32_t* cu_k;
  const int32_t* tables;  // NULL: dense K/V
  int H, Hkv, max_blocks, block_size, box_tok;
  int64_t out_stride;
  float scale_log2;
};

// POLY: how many of every 4 exponentials go to a polynomial instead of MUFU.EX2 (NVLLM_PREFILL_POLY): 1 / 2 / 3 with the
// floor-based polynomial (the first version; its FRND / F2I also run on the XU pipe, so 2 and 3 gain nothing over 1),
// 4 / 5 / 6 = 1/4, 1/2, 3/4 with the XU-free polynomial.  Default 4 (measured best: profiles/r01_prefill_microbench_v5_polysplit.jsonl).
template <int POLY>
__global__ void __launch_bounds__(pf::kThreads, 1)
prefill_attn_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                    const __grid_constant__ CUtensorMap tmap_v, const PrefillParams p) {
  using namespace pf;
  pdl_launch_dependents();
  pdl_wait();  // q / k / v (and the cache) come from the previous kernel
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t bars = smem_base + kSmemBar;
  const uint32_t q_full = bars;
  auto slot_full = [&](int s) { return bars + (1 + s) * 8; };
  auto slot_empty = [&](int s) { return bars + (1 + NSLOT + s) * 8; };
  auto s_full = [&](int g) { return bars + (1 + 2 * NSLOT + g) * 8; };       // S_j of group g is in TMEM
  auto p_ready = [&](int g) { return bars + (1 + 2 * NSLOT + 2 + g) * 8; };  // P_j of group g is in TMEM (128 arrivals)
  auto o_done = [&](int g) { return bars + (1 + 2 * NSLOT + 4 + g) * 8; };   // PV_j of group g has retired
  const uint32_t tmem_slot = bars + kNumBars * 8;

  const int seq = blockIdx.z, head = blockIdx.y;
  const int qblk = gridDim.x - 1 - blockIdx.x;  // longest (latest rows) first: shorter tail
  const int kvh = head / (p.H / p.Hkv);
  const int q0 = __ldg(p.cu_q + seq), Lq = __ldg(p.cu_q + seq + 1) - q0;
  const int k0 = __ldg(p.cu_k + seq), Lk = __ldg(p.cu_k + seq + 1) - k0;
  if (qblk * 2 * BM >= Lq) return;  // uniform per CTA: nothing allocated yet
  const int shift = Lk - Lq;        // bottom-right alignment
  auto tiles_for = [&](int g) {     // number of KV tiles a row group needs (0 when it has no valid row)
    const int first = qblk * 2 * BM + g * BM;
    if (first >= Lq) return 0;
    const int last = min(first + BM, Lq) - 1;
    const int visible = max(1, min(Lk, shift + last + 1));
    return (visible + BN - 1) / BN;
  };
  const int n_rg[2] = {tiles_for(0), tiles_for(1)};
  const int n_cta = max(n_rg[0], n_rg[1]);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&tmap_q); tma_prefetch_desc(&tmap_k); tma_prefetch_desc(&tmap_v);
    mbar_init(q_full, 1);
    for (int s = 0; s < NSLOT; ++s) { mbar_init(slot_full(s), 1); mbar_init(slot_empty(s), 1); }
    for (int g = 0; g < 2; ++g) { mbar_init(s_full(g), 1); mbar_init(p_ready(g), BM); mbar_init(o_done(g), 1); }
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<uint32_t*>(smem_gen + kSmemBar + kNumBars * 8);

  // Register reallocation (setmaxnreg, per warpgroup): the producer / MMA warpgroup keeps 88 registers per thread and
  // donates the rest; each softmax thread holds a full 128-key S row (128 registers) and gets 208.
  if (warp >= 8) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 88;\n");  // must dominate the role code below (and only it)
  }
  if (warp >= 10) {
    // idle warps of the third warpgroup
  } else if (warp == 8) {
    if (lane == 0) {
      // ===== TMA producer =====
      mbar_expect_tx(q_full, 2 * kQBytes);
      for (int g = 0; g < 2; ++g)
        for (int hf = 0; hf < 2; ++hf)
          tma_load_2d(smem_base + kSmemQ + g * kQBytes + hf * (BM * 128), &tmap_q, q_full, head * D + hf * 64,
                      q0 + qblk * 2 * BM + g * BM);
      const int32_t* table = p.tables ? p.tables + static_cast<int64_t>(seq) * p.max_blocks : nullptr;
      for (int t = 0; t < 2 * n_cta; ++t) {  // t even: K_{t/2}, t odd: V_{t/2}
        const int j = t >> 1, slot = t % NSLOT;
        const CUtensorMap* map = (t & 1) ? &tmap_v : &tmap_k;
        mbar_wait(slot_empty(slot), ((t / NSLOT) & 1) ^ 1);
        mbar_expect_tx(slot_full(slot), kTileBytes);
        const uint32_t dst = smem_base + kSmemKV + slot * kTileBytes;
        if (table == nullptr) {
          for (int hf = 0; hf < 2; ++hf)
            tma_load_2d(dst + hf * (BN * 128), map, slot_full(slot), kvh * D + hf * 64, k0 + j * BN);
        } else {
          for (int c = 0; c < BN / p.box_tok; ++c) {  // box_tok = min(128, block_size) tokens per TMA box
            const int tok = j * BN + c * p.box_tok;
            const int bi = tok / p.block_size;
            const int blk = bi < p.max_blocks ? __ldg(table + bi) : -1;
            const int blk_c = blk < 0 ? 0x3fffff : blk;  // past the sequence end: far out of bounds -> TMA zero-fills
            for (int hf = 0; hf < 2; ++hf)
              tma_load_4d(dst + hf * (BN * 128) + c * p.box_tok * 128, map, slot_full(slot), hf * 64,
                          tok - bi * p.block_size, kvh, blk_c);
          }
        }
      }
    }
  } else if (warp == 9) {
    if (lane == 0) {
      // ===== MMA issuer (both row groups) =====
      constexpr uint32_t idesc_s = umma_idesc_bf16(BM, BN, 0, 0);  // Q K-major, K K-major
      constexpr uint32_t idesc_o = umma_idesc_bf16(BM, D, 0, 1);   // P K-major (TMEM), V MN-major
      auto tile_full = [&](int t) { mbar_wait(slot_full(t % NSLOT), (t / NSLOT) & 1); tc_fence_after(); };
      auto issue_s = [&](int g, int j) {  // S_j^g = Q^g . K_j^T
        const uint32_t sQ = smem_base + kSmemQ + g * kQBytes;
        const uint32_t sK = smem_base + kSmemKV + ((2 * j) % NSLOT) * kTileBytes;
#pragma unroll
        for (int ks = 0; ks < D / 16; ++ks) {
          const uint64_t ad = umma_desc_sw128(sQ + (ks >> 2) * (BM * 128) + (ks & 3) * 32, 16, 1024);
          const uint64_t bd = umma_desc_sw128(sK + (ks >> 2) * (BN * 128) + (ks & 3) * 32, 16, 1024);
          umma_bf16_ss(tmem_base + g * 256, ad, bd, idesc_s, ks != 0);
        }
        umma_commit(s_full(g));
      };
      mbar_wait(q_full, 0);
      if (n_cta > 0) {
        tile_full(0);
        for (int g = 0; g < 2; ++g)
          if (n_rg[g] > 0) issue_s(g, 0);
        umma_commit(slot_empty(0));  // K_0 consumed once these MMAs retire
      }
      for (int j = 0; j < n_cta; ++j) {
        bool v_waited = false, k_waited = false;
        for (int g = 0; g < 2; ++g) {
          if (j >= n_rg[g]) continue;
          mbar_wait(p_ready(g), j & 1);  // softmax of group g wrote P_j (and rescaled O if it had to)
          tc_fence_after();
          if (!v_waited) { tile_full(2 * j + 1); v_waited = true; }
          const uint32_t sV = smem_base + kSmemKV + ((2 * j + 1) % NSLOT) * kTileBytes;
#pragma unroll
          for (int ks = 0; ks < BN / 16; ++ks) {
            // P: 16 keys = 8 packed columns per step.  V tile: rows = keys, 64 contiguous d per row; the two d halves
            // (MN groups) are 16 KiB apart, 8 keys per 1 KiB.
            const uint64_t bd = umma_desc_sw128(sV + ks * 2048, BN * 128, 1024);
            umma_bf16_ts(tmem_base + g * 256 + 128, tmem_base + g * 256 + ks * 8, bd, idesc_o, (j | ks) != 0);
          }
          umma_commit(o_done(g));
          if (j + 1 < n_rg[g]) {
            if (!k_waited) { tile_full(2 * j + 2); k_waited = true; }
            issue_s(g, j + 1);  // executes after PV_j^g in the pipe: P_j (aliased with S) is no longer needed by then
          }
        }
        // both groups' uses of V_j and K_{j+1} are issued: release the slots once they retire
        umma_commit(slot_empty((2 * j + 1) % NSLOT));
        if (j + 1 < n_cta) umma_commit(slot_empty((2 * j + 2) % NSLOT));
      }
    }
  } else {
    // ===== row groups: softmax + output =====
    asm volatile("setmaxnreg.inc.sync.aligned.u32 208;\n");
    const int g = warp >> 2;               // row group 0 / 1
    const int t = threadIdx.x & 127;       // row within the group's tile == TMEM lane
    const int n_my = n_rg[g];
    const int row = qblk * 2 * BM + g * BM + t;  // row within the sequence's queries
    const int limit = shift + row;               // last visible key index
    const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t tS = tmem_base + g * 256 + lane_off;  // S (fp32, 128 columns) / P (packed bf16, first 64 columns)
    const uint32_t tO = tS + 128;
    constexpr float kRescaleThreshold = 8.0f;  // log2 units: P stays <= 2^8 between refreshes
    // 2^x for x <= 0 on the FMA pipe: x = n + f, f in [0, 1); 2^f by a degree-3 polynomial (rel. err 2.3e-4, far below
    // bf16's 2^-9), 2^n by adding n to the exponent field.
    auto exp2_poly = [](float x) {
      x = fmaxf(x, -126.f);
      const float n = floorf(x);
      const float f = x - n;
      float pfr = fmaf(f, 0.0790209f, 0.2241543f);
      pfr = fmaf(pfr, f, 0.6965186f);
      pfr = fmaf(pfr, f, 0.9998430f);
      return __int_as_float(__float_as_int(pfr) + (static_cast<int>(n) << 23));
    };

    // The same without any XU-pipe instruction (the ncu source page of round 1 shows FRND.FLOOR and F2I.FLOOR of the
    // version above on the XU pipe next to MUFU.EX2 — it relieves nothing): round-to-nearest by the 1.5 * 2^23 magic add,
    // f in [-0.5, 0.5], degree-3 minimax polynomial (rel. err 7.5e-5), n taken from the low mantissa bits of the sum.
    auto exp2_poly2 = [](float x) {
      x = fmaxf(x, -126.f);
      const float t = x + 12582912.f;
      const float f = x - (t - 12582912.f);
      float pfr = fmaf(f, 0.05517210811376572f, 0.24261118471622467f);
      pfr = fmaf(pfr, f, 0.693260908126831f);
      pfr = fmaf(pfr, f, 0.9999280571937561f);
      return __int_as_float(__float_as_int(pfr) + (__float_as_int(t) << 23));
    };

    float m_ref = -INFINITY, l_run = 0.f;  // m_ref: the (stale-by-design) maximum the exponentials are taken against
    for (int j = 0; j < n_my; ++j) {
      mbar_wait(s_full(g), j & 1);
      // S_j was issued after PV_{j-1} and MMAs of one thread retire in order: this wait is free.  It consumes the o_done
      // phases strictly in order (a parity wait may never lag two phases behind), and O is quiescent from here on.
      if (j > 0) mbar_wait(o_done(g), (j - 1) & 1);
      tc_fence_after();
      uint32_t sr[4][32];
#pragma unroll
      for (int c = 0; c < 4; ++c) tmem_ld_32x32b_x32(tS + c * 32, sr[c]);
      tmem_ld_wait();
      const int kbase = j * BN;
      if ((kbase + BN - 1 > limit) || (kbase + BN > Lk)) {
#pragma unroll
        for (int c = 0; c < BN; ++c)
          if (kbase + c > limit || kbase + c >= Lk) sr[c >> 5][c & 31] = 0xff800000u;  // -inf
      }
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};  // 4 independent chains
#pragma unroll
      for (int c = 0; c < BN; ++c) mx4[c & 3] = fmaxf(mx4[c & 3], __uint_as_float(sr[c >> 5][c & 31]));
      const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
      // lazy rescale: refresh the reference maximum only when this tile exceeds it by more than the threshold
      const bool refresh = (mx - m_ref) * p.scale_log2 > kRescaleThreshold;  // also true for the first finite tile
      float corr = 1.f;
      if (refresh) {
        corr = (m_ref == -INFINITY) ? 0.f : fast_exp2((m_ref - mx) * p.scale_log2);
        m_ref = mx;
        l_run *= corr;
      }
      if (j > 0 && __any_sync(0xffffffffu, refresh)) {  // tcgen05.ld/st are warp-collective: the whole warp takes part
#pragma unroll
        for (int c = 0; c < D / 32; ++c) {
          uint32_t orr[32];
          tmem_ld_32x32b_x32(tO + c * 32, orr);
          tmem_ld_wait();
#pragma unroll
          for (int e = 0; e < 32; ++e) orr[e] = __float_as_uint(__uint_as_float(orr[e]) * corr);
          tmem_st_32x32b_x32(tO + c * 32, orr);
        }
      }
      const float moff = (m_ref == -INFINITY) ? 0.f : m_ref * p.scale_log2;
      float ls4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int h2 = 0; h2 < 2; ++h2) {  // 2 x 64 keys -> 2 x 32 packed columns
        uint32_t pk[32];
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          const int c = h2 * 64 + e * 2;
          const float a0 = fmaf(__uint_as_float(sr[c >> 5][c & 31]), p.scale_log2, -moff);
          const float a1 = fmaf(__uint_as_float(sr[(c + 1) >> 5][(c + 1) & 31]), p.scale_log2, -moff);
          float p0, p1;
          if constexpr (POLY <= 3) {
            p0 = (POLY >= 3 && (e & 1)) ? exp2_poly(a0) : fast_exp2(a0);               // MUFU (POLY 3: half of them FMA)
            p1 = (POLY >= 2 || (POLY == 1 && (e & 1))) ? exp2_poly(a1) : fast_exp2(a1);  // FMA-pipe share per POLY
          } else {  // POLY 4 / 5 / 6: 1/4, 1/2, 3/4 of the exponentials on the XU-free polynomial
            p0 = (POLY >= 6 && (e & 1)) ? exp2_poly2(a0) : fast_exp2(a0);
            p1 = (POLY >= 5 || (e & 1)) ? exp2_poly2(a1) : fast_exp2(a1);
          }
          ls4[e & 3] += p0 + p1;
          pk[e] = pack_bf16x2(p0, p1);      // low half = even key: consecutive K elements share a 32-bit column
        }
        tmem_st_32x32b_x32(tS + h2 * 32, pk);  // P over this thread's own S lane
      }
      l_run += (ls4[0] + ls4[1]) + (ls4[2] + ls4[3]);
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(p_ready(g));
    }
    float acc[D];
    if (n_my > 0) {
      mbar_wait(o_done(g), (n_my - 1) & 1);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < D / 32; ++c) {
        uint32_t orr[32];
        tmem_ld_32x32b_x32(tO + c * 32, orr);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; ++e) acc[c * 32 + e] = __uint_as_float(orr[e]);
      }
      tc_fence_before();
    }
    // ---- output -------------------------------------------------------------------------------------
    if (n_my > 0 && row < Lq) {
      const float inv = l_run > 0.f ? 1.0f / l_run : 0.f;
      bf16* orow = p.out + static_cast<int64_t>(q0 + row) * p.out_stride + head * D;
#pragma unroll
      for (int c = 0; c < D / 8; ++c) {
        uint4 v;
        v.x = pack_bf16x2(acc[c * 8 + 0] * inv, acc[c * 8 + 1] * inv);
        v.y = pack_bf16x2(acc[c * 8 + 2] * inv, acc[c * 8 + 3] * inv);
        v.z = pack_bf16x2(acc[c * 8 + 4] * inv, acc[c * 8 + 5] * inv);
        v.w = pack_bf16x2(acc[c * 8 + 6] * inv, acc[c * 8 + 7] * inv);
        *reinterpret_cast<uint4*>(orow + c * 8) = v;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

}  // namespace nvllm

extern "C" int nvllm_varlen_prefill_attention(void* out, const void* q, const void* k, const void* v,
                                              const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k,
                                              int32_t batch, int32_t total_q, int32_t total_k,
                                              int32_t max_seqlen_q, int32_t max_seqlen_k, int32_t num_heads,
                                              int32_t num_kv_heads, int32_t head_dim, int64_t q_token_stride,
                                              int64_t kv_token_stride, int64_t out_token_stride,
                                              const void* k_cache, const void* v_cache,
                                              const int32_t* block_tables, int32_t max_blocks_per_seq,
                                              const nvllm_kv_layout* layout, float scale, int32_t dtype,
                                              void* stream) {
  using namespace nvllm;
  NVLLM_CHECK_ARG(dtype == NVLLM_BF16, NVLLM_ERR_DTYPE, "prefill_attention: dtype %d (bf16 only)", dtype);
  NVLLM_CHECK_ARG(head_dim == pf::D, NVLLM_ERR_ARG, "prefill_attention: head_dim %d (128 only)", head_dim);
  NVLLM_CHECK_ARG(batch >= 0 && batch <= 65535 && total_q >= 0 && max_seqlen_q >= 0, NVLLM_ERR_ARG,
                  "prefill_attention: batch %d total_q %d", batch, total_q);
  if (batch == 0 || total_q == 0 || max_seqlen_q == 0) return 0;
  NVLLM_CHECK_ARG(num_heads > 0 && num_kv_heads > 0 && num_heads % num_kv_heads == 0 && num_heads <= 65535,
                  NVLLM_ERR_ARG, "prefill_attention: heads %d / kv heads %d", num_heads, num_kv_heads);
  NVLLM_CHECK_ARG(out && q && cu_seqlens_q && cu_seqlens_k, NVLLM_ERR_ARG, "prefill_attention: NULL pointer");
  NVLLM_CHECK_ARG(aligned16(out) && out_token_stride % 8 == 0 && out_token_stride >= (int64_t)num_heads * head_dim,
                  NVLLM_ERR_ALIGN, "prefill_attention: out alignment / stride");
  const bool paged = block_tables != nullptr;
  CUtensorMap tq, tk, tv;
  if (int rc = make_tmap_bf16_2d(&tq, q, total_q, (int64_t)num_heads * head_dim, q_token_stride, pf::BM)) return rc;
  PrefillParams p;
  p.box_tok = pf::BN;
  if (!paged) {
    NVLLM_CHECK_ARG(k && v && total_k > 0, NVLLM_ERR_ARG, "prefill_attention: dense k/v missing");
    if (int rc = make_tmap_bf16_2d(&tk, k, total_k, (int64_t)num_kv_heads * head_dim, kv_token_stride, pf::BN)) return rc;
    if (int rc = make_tmap_bf16_2d(&tv, v, total_k, (int64_t)num_kv_heads * head_dim, kv_token_stride, pf::BN)) return rc;
    p.block_size = 1;
    p.max_blocks = 0;
  } else {
    if (int rc = check_layout(layout)) return rc;
    NVLLM_CHECK_ARG(k_cache && v_cache && max_blocks_per_seq > 0, NVLLM_ERR_ARG, "prefill_attention: paged cache missing");
    NVLLM_CHECK_ARG(layout->num_kv_heads == num_kv_heads && layout->head_dim == head_dim, NVLLM_ERR_ARG,
                    "prefill_attention: layout heads/dim mismatch");
    const int bs = layout->block_size;
    NVLLM_CHECK_ARG(bs % 16 == 0 && (bs >= pf::BN ? bs % pf::BN == 0 : pf::BN % bs == 0), NVLLM_ERR_ARG,
                    "prefill_attention: block_size %d must divide or be a multiple of %d", bs, pf::BN);
    p.box_tok = bs < pf::BN ? bs : pf::BN;
    const uint64_t dims[4] = {(uint64_t)head_dim, (uint64_t)bs, (uint64_t)num_kv_heads, 0x3fffffull};
    const uint64_t strides[3] = {(uint64_t)layout->token_stride * 2, (uint64_t)layout->head_stride * 2,
                                 (uint64_t)layout->block_stride * 2};
    const uint32_t box[4] = {64, (uint32_t)p.box_tok, 1, 1};
    if (int rc = make_tmap_bf16_nd(&tk, k_cache, 4, dims, strides, box)) return rc;
    if (int rc = make_tmap_bf16_nd(&tv, v_cache, 4, dims, strides, box)) return rc;
    p.block_size = bs;
    p.max_blocks = max_blocks_per_seq;
  }
  p.out = static_cast<bf16*>(out);
  p.cu_q = cu_seqlens_q;
  p.cu_k = cu_seqlens_k;
  p.tables = block_tables;
  p.H = num_heads;
  p.Hkv = num_kv_heads;
  p.out_stride = out_token_stride;
  p.scale_log2 = scale * 1.4426950408889634f;
  {
    static std::once_flag once;
    static cudaError_t once_err = cudaSuccess;
    std::call_once(once, [] {
      const auto attr = cudaFuncAttributeMaxDynamicSharedMemorySize;
      cudaError_t e = cudaFuncSetAttribute(prefill_attn_kernel<1>, attr, pf::kSmemBytes);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(prefill_attn_kernel<2>, attr, pf::kSmemBytes);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(prefill_attn_kernel<3>, attr, pf::kSmemBytes);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(prefill_attn_kernel<4>, attr, pf::kSmemBytes);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(prefill_attn_kernel<5>, attr, pf::kSmemBytes);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(prefill_attn_kernel<6>, attr, pf::kSmemBytes);
      once_err = e;
    });
    NVLLM_CUDA(once_err);
  }
  (void)max_seqlen_k;
  dim3 grid((max_seqlen_q + 2 * pf::BM - 1) / (2 * pf::BM), num_heads, batch);
  const int poly = knobs().prefill_poly;  // XU-free polynomial on 1/4 of the exponentials: +26 % at 4K, 1060 TFLOP/s at 32K
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (poly == 2) NVLLM_CUDA(launch_pdl(prefill_attn_kernel<2>, grid, dim3(pf::kThreads), pf::kSmemBytes, st, tq, tk, tv, p));
  else if (poly == 3) NVLLM_CUDA(launch_pdl(prefill_attn_kernel<3>, grid, dim3(pf::kThreads), pf::kSmemBytes, st, tq, tk, tv, p));
  else if (poly == 4) NVLLM_CUDA(launch_pdl(prefill_attn_kernel<4>, grid, dim3(pf::kThreads), pf::kSmemBytes, st, tq, tk, tv, p));
  else if (poly == 5) NVLLM_CUDA(launch_pdl(prefill_attn_kernel<5>, grid, dim3(pf::kThreads), pf::kSmemBytes, st, tq, tk, tv, p));
  else if (poly == 6) NVLLM_CUDA(launch_pdl(prefill_attn_kernel<6>, grid, dim3(pf::kThreads), pf::kSmemBytes, st, tq, tk, tv, p));
  else NVLLM_CUDA(launch_pdl(prefill_attn_kernel<1>, grid, dim3(pf::kThreads), pf::kSmemBytes, st, tq, tk, tv, p));
  return 0;
}
