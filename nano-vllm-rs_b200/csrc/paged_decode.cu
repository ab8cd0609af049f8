// paged_decode.cu — K2: split-K paged decode attention for GQA heads (HBM-bound).
//
// Replaces Attention::flash_attention_decode -> compute_attention_with_cache -> gather_cached_kv ->
// compute_attention (reference src/layers/attention.rs:225-318,238-261): the reference walks
// block_tables on the host, concatenates whole KV blocks (a full copy of the context) and runs a
// dense matmul + softmax per sequence.  Here every KV byte is read from HBM exactly once.
//
// Work decomposition
//   grid = (num_splits, Hkv * ceil(g/8), batch); one CTA owns one (sequence, kv head, q-head group of
//   <= 8, split of the KV length).  A CTA iterates over tiles of 16*NWARPS consecutive tokens; warp w
//   owns rows [16w, 16w+16) of every tile and runs its own cp.async pipeline (STAGES deep, 8 KiB per
//   stage for D=128), so the main loop contains no block-level barrier at all.
// Staging
//   cp.async (LDGSTS) 16-byte copies, global -> shared, XOR-swizzled on the 16-byte chunk index so the
//   ldmatrix reads are bank-conflict free.  Rows past context_lens are zero-filled.
// Math (per warp, per 16 tokens), transposed so that the GQA group sits on the MMA N dimension:
//   S^T[16 tok x 8 heads] = K[16 x D] . Q^T[D x 8]          (mma.sync m16n8k16, A = K via ldmatrix)
//   online softmax over tokens: quad/warp shuffles (xor 4,8,16), exp2 with scale*log2(e) folded in
//   P^T -> B operand through movmatrix.trans (8x8 b16 transpose inside the warp)
//   O^T[D x 8 heads] += V^T[D x 16 tok] . P^T[16 x 8]        (A = V via ldmatrix.trans)
//   No padding waste for g = 8, 2x for g = 4, 4x for g = 2 — the tensor pipe is idle otherwise.
// Epilogue: warps merge (m, l, O) through shared memory; with one split the CTA writes bf16 output,
//   otherwise fp32 partials + (m, l) go to the workspace and paged_decode_reduce_kernel combines.
//
// Algorithmic bytes per launch: sum_b 2*L_b*Hkv*D*2 + 2*B*H*D*2 + 4*sum_b ceil(L_b/bs) + 4*B.
#include <math.h>

#include <atomic>

#include "common.cuh"

namespace nvllm {

int check_layout(const nvllm_kv_layout* L);  // store_kvcache.cu

struct DecodeParams {
  bf16* out;
  const bf16* q;
  const bf16* kc;
  const bf16* vc;
  const int32_t* tables;
  const int32_t* ctx;
  float* ws_o;   // [B, H, S, D]
  float* ws_ml;  // [B, H, S, 2]
  int batch, H, Hkv, g, ngrp, max_blocks, block_size, num_splits;
  int64_t q_stride, out_stride, blk_stride, tok_stride, head_stride;
  float scale_log2;
  // FUSE only (nvllm_paged_decode_attention_fused): q points at the packed [q | k | v] rows of this step; the kernel applies
  // q/k-norm + RoPE itself, writes the step's K/V row into the cache and attends to it from shared memory.
  bf16* kc_w;
  bf16* vc_w;
  const bf16* qnw;       // may be NULL (no qk-norm)
  const bf16* knw;
  const float* cos_sin;  // [max_position, D]: D/2 cosines then D/2 sines per position
  const int64_t* pos;
  const int32_t* slots;
  float eps;
  int max_position;
};

template <int D, int STAGES, int NWARPS>
struct DecodeCfg {
  static constexpr int kChunks = D / 8;                // 16-byte chunks per row
  static constexpr int kRowBytes = D * 2;
  static constexpr int kSubTileBytes = 16 * kRowBytes;  // 16 tokens of K (or V)
  static constexpr int kStageBytes = 2 * kSubTileBytes; // K + V
  static constexpr int kWarpBytes = STAGES * kStageBytes;
  static constexpr int kPipeBytes = NWARPS * kWarpBytes;
  static constexpr int kOStride = D + 4;                // fp32 words per head row in the merge buffer
  static constexpr int kMergeBytes = NWARPS * (8 * kOStride + 16) * 4;
  static constexpr int kSmemBytes = kPipeBytes > kMergeBytes ? kPipeBytes : kMergeBytes;
  static constexpr int kRowsPerLane = 16 * kChunks / 32;  // cp.async per lane per K sub-tile
  static constexpr int kRowStep = 32 / kChunks;
};

// FUSE == true: Qwen3Attention::forward's per-head q/k RMSNorm + RotaryEmbedding::forward (qwen3.rs:219-229,
// rotary_embedding.rs:23-48) and Attention::store_kv_cache (attention.rs:150-174) run inside this kernel: every warp norms
// and rotates its Q fragments in registers; the context seen through the cache is context_lens - 1 tokens; warp 0 of the
// CTA owning the last non-empty split builds the step's own K (normed, rotated) / V row, stores it at slot_mapping[b] and
// processes it as one extra 16-token tile (rows 1..15 masked) straight from shared memory.  Same rounding points as the
// stand-alone qknorm_rope_store kernel (bf16 after the norm, bf16 after the rotation).
template <int D, int STAGES, int NWARPS, bool FUSE>
__global__ void __launch_bounds__(NWARPS * 32, (DecodeCfg<D, STAGES, NWARPS>::kSmemBytes <= 72 * 1024) ? 3 : (DecodeCfg<D, STAGES, NWARPS>::kSmemBytes <= 110 * 1024) ? 2 : 1)
paged_decode_kernel(const DecodeParams p) {
  using C = DecodeCfg<D, STAGES, NWARPS>;
  constexpr int KS = D / 16;  // k-steps of QK^T == m-tiles of PV
  extern __shared__ __align__(128) uint8_t smem[];

  pdl_launch_dependents();
  const int split = blockIdx.x;
  const int kvh = blockIdx.y / p.ngrp;
  const int grp = blockIdx.y - kvh * p.ngrp;
  const int b = blockIdx.z;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  pdl_wait();  // q and the newest KV row were written by the previous kernel
  const int ctx_all = __ldg(p.ctx + b);        // tokens attended to, including (FUSE) this step's own
  const int ctx = FUSE ? max(ctx_all - 1, 0) : ctx_all;  // tokens read from the cache
  const int gvalid = min(8, p.g - grp * 8);   // q heads handled by this CTA
  const int qh0 = kvh * p.g + grp * 8;         // first q head

  constexpr int TILE = 16 * NWARPS;
  const int tiles_total = (ctx + TILE - 1) / TILE;
  const int tps = (tiles_total + p.num_splits - 1) / p.num_splits;
  const int t0 = split * tps;
  const int t1 = min(t0 + tps, tiles_total);
  const int n = max(t1 - t0, 0);

  const uint32_t warp_smem = smem_u32(smem) + warp * C::kWarpBytes;
  const int32_t* table = p.tables + static_cast<int64_t>(b) * p.max_blocks;
  const bf16* kbase = p.kc + kvh * p.head_stride;
  const bf16* vbase = p.vc + kvh * p.head_stride;

  auto issue = [&](int i) {
    const int s = i % STAGES;
    const int tok_base = ((t0 + i) * NWARPS + warp) * 16;
    int blk = -1, off = 0;
    if (tok_base < ctx) {
      const int bi = tok_base / p.block_size;
      off = tok_base - bi * p.block_size;
      if (bi < p.max_blocks) blk = __ldg(table + bi);
    }
    const uint32_t sK = warp_smem + s * C::kStageBytes;
    const uint32_t sV = sK + C::kSubTileBytes;
    const int c = lane % C::kChunks;
    const int r0 = lane / C::kChunks;
    const int64_t gofs0 = static_cast<int64_t>(blk < 0 ? 0 : blk) * p.blk_stride + c * 8;
#pragma unroll
    for (int j = 0; j < C::kRowsPerLane; ++j) {
      const int r = r0 + j * C::kRowStep;
      const bool valid = (blk >= 0) && (tok_base + r < ctx);
      const int64_t gofs = valid ? gofs0 + static_cast<int64_t>(off + r) * p.tok_stride : 0;
      const uint32_t dst = r * C::kRowBytes + ((c ^ (r & 7)) << 4);
      cp_async_16(sK + dst, kbase + gofs, valid ? 16 : 0);
      cp_async_16(sV + dst, vbase + gofs, valid ? 16 : 0);
    }
  };

  // The first STAGES-1 tiles are on their way before anything that depends on q is touched.
#pragma unroll
  for (int s = 0; s < STAGES - 1; ++s) {
    if (s < n) issue(s);
    cp_async_commit();
  }

  // ---- Q^T B-fragments: head n = lane/4, dims 16ks + 2(lane%4) + {0,1} (+8) --------------------
  uint32_t qb[KS][2];
  {
    const int hn = lane >> 2;
    const bool hv = hn < gvalid;
    const bf16* qrow = p.q + b * p.q_stride + static_cast<int64_t>(qh0 + (hv ? hn : 0)) * D;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      const int d0 = ks * 16 + (lane & 3) * 2;
      qb[ks][0] = hv ? *reinterpret_cast<const uint32_t*>(qrow + d0) : 0u;
      qb[ks][1] = hv ? *reinterpret_cast<const uint32_t*>(qrow + d0 + 8) : 0u;
    }
    if constexpr (FUSE) {
      static_assert(!FUSE || D == 128, "fused q/k-norm + RoPE: head_dim 128");
      // this lane holds dims 16ks + 2(lane%4) + {0,1} (+8) of head lane/4: the 4 lanes of a quad hold the whole head,
      // and the RoPE partner d + 64 of every dim sits in the same lane at ks + 4
      float x[KS][4];
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        x[ks][0] = bf16_lo(qb[ks][0]); x[ks][1] = bf16_hi(qb[ks][0]);
        x[ks][2] = bf16_lo(qb[ks][1]); x[ks][3] = bf16_hi(qb[ks][1]);
      }
      if (p.qnw != nullptr) {
        float ss = 0.f;
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) ss += x[ks][0] * x[ks][0] + x[ks][1] * x[ks][1] + x[ks][2] * x[ks][2] + x[ks][3] * x[ks][3];
        ss += __shfl_xor_sync(0xffffffffu, ss, 1);
        ss += __shfl_xor_sync(0xffffffffu, ss, 2);
        const float inv = 1.0f / sqrtf(ss / static_cast<float>(D) + p.eps);
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
          const int d0 = ks * 16 + (lane & 3) * 2;
          const uint32_t w0 = __ldg(reinterpret_cast<const uint32_t*>(p.qnw + d0));
          const uint32_t w1 = __ldg(reinterpret_cast<const uint32_t*>(p.qnw + d0 + 8));
          x[ks][0] = __bfloat162float(__float2bfloat16_rn(x[ks][0] * inv * bf16_lo(w0)));
          x[ks][1] = __bfloat162float(__float2bfloat16_rn(x[ks][1] * inv * bf16_hi(w0)));
          x[ks][2] = __bfloat162float(__float2bfloat16_rn(x[ks][2] * inv * bf16_lo(w1)));
          x[ks][3] = __bfloat162float(__float2bfloat16_rn(x[ks][3] * inv * bf16_hi(w1)));
        }
      }
      int64_t pos = __ldg(p.pos + b);
      pos = pos < 0 ? 0 : (pos >= p.max_position ? p.max_position - 1 : pos);
      const float* cs = p.cos_sin + pos * D;
#pragma unroll
      for (int ks = 0; ks < KS / 2; ++ks) {
        const int d0 = ks * 16 + (lane & 3) * 2;   // < 64
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int d = d0 + (e & 1) + (e >> 1) * 8;
          const float c = __ldg(cs + d), sn = __ldg(cs + D / 2 + d);
          const float x1 = x[ks][e], x2 = x[ks + KS / 2][e];
          x[ks][e] = x1 * c - x2 * sn;
          x[ks + KS / 2][e] = x2 * c + x1 * sn;
        }
      }
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        qb[ks][0] = hv ? pack_bf16x2(x[ks][0], x[ks][1]) : 0u;
        qb[ks][1] = hv ? pack_bf16x2(x[ks][2], x[ks][3]) : 0u;
      }
    }
  }

  // ---- FUSE: this step's own token is built while the first tiles fly: stored into the cache, and kept as row 0 of a
  // private 16-token sub-tile (after the pipeline / merge buffers) that warp 0 of the last split scores after the loop ----
  const int last_split = tiles_total == 0 ? 0 : (tiles_total - 1) / tps;  // the last split that owns any cached tile
  const bool own_new = FUSE && split == last_split && warp == 0 && ctx_all > 0;
  const uint32_t new_off = C::kSmemBytes;  // K sub-tile, then V sub-tile
  if constexpr (FUSE) {
    if (own_new) {
      const uint4 z = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
      for (int j = 0; j < 2 * C::kSubTileBytes / (32 * 16); ++j)   // zero K and V sub-tiles (rows 1..15 must be finite)
        *reinterpret_cast<uint4*>(smem + new_off + (j * 32 + lane) * 16) = z;
      __syncwarp();
      // lane l owns elements (2l, 2l+1) of each half of the row: exactly the RoPE pairs (j, j + D/2)
      const bf16* krow = p.q + b * p.q_stride + static_cast<int64_t>(p.H + kvh) * D;
      const bf16* vrow = p.q + b * p.q_stride + static_cast<int64_t>(p.H + p.Hkv + kvh) * D;
      const uint32_t ka = *reinterpret_cast<const uint32_t*>(krow + 2 * lane);
      const uint32_t kb2 = *reinterpret_cast<const uint32_t*>(krow + D / 2 + 2 * lane);
      const uint32_t va = *reinterpret_cast<const uint32_t*>(vrow + 2 * lane);
      const uint32_t vb = *reinterpret_cast<const uint32_t*>(vrow + D / 2 + 2 * lane);
      float x1[2] = {bf16_lo(ka), bf16_hi(ka)}, x2[2] = {bf16_lo(kb2), bf16_hi(kb2)};
      if (p.knw != nullptr) {
        float ss = x1[0] * x1[0] + x2[0] * x2[0] + x1[1] * x1[1] + x2[1] * x2[1];
        ss = warp_sum(ss);
        const float inv = 1.0f / sqrtf(ss / static_cast<float>(D) + p.eps);
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          x1[e] = __bfloat162float(__float2bfloat16_rn(x1[e] * inv * __bfloat162float(p.knw[2 * lane + e])));
          x2[e] = __bfloat162float(__float2bfloat16_rn(x2[e] * inv * __bfloat162float(p.knw[D / 2 + 2 * lane + e])));
        }
      }
      int64_t pos = __ldg(p.pos + b);
      pos = pos < 0 ? 0 : (pos >= p.max_position ? p.max_position - 1 : pos);
      const float* cs = p.cos_sin + pos * D;
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const float c = __ldg(cs + 2 * lane + e), sn = __ldg(cs + D / 2 + 2 * lane + e);
        const float o1 = x1[e] * c - x2[e] * sn, o2 = x2[e] * c + x1[e] * sn;
        x1[e] = o1; x2[e] = o2;
      }
      const uint32_t k1 = pack_bf16x2(x1[0], x1[1]), k2 = pack_bf16x2(x2[0], x2[1]);
      // row 0 of the sub-tile (no swizzle for row 0: chunk c ^ 0)
      uint32_t* k0s = reinterpret_cast<uint32_t*>(smem + new_off);
      uint32_t* v0s = reinterpret_cast<uint32_t*>(smem + new_off + C::kSubTileBytes);
      k0s[lane] = k1; k0s[D / 4 + lane] = k2;
      v0s[lane] = va; v0s[D / 4 + lane] = vb;
      if (grp == 0) {
        const int slot = __ldg(p.slots + b);
        if (slot >= 0) {
          const int blk = slot / p.block_size, off = slot - blk * p.block_size;
          const int64_t o = blk * p.blk_stride + off * p.tok_stride + kvh * p.head_stride;
          *reinterpret_cast<uint32_t*>(p.kc_w + o + 2 * lane) = k1;
          *reinterpret_cast<uint32_t*>(p.kc_w + o + D / 2 + 2 * lane) = k2;
          *reinterpret_cast<uint32_t*>(p.vc_w + o + 2 * lane) = va;
          *reinterpret_cast<uint32_t*>(p.vc_w + o + D / 2 + 2 * lane) = vb;
        }
      }
      __syncwarp();
    }
  }

  float m_run[2] = {-INFINITY, -INFINITY};
  float l_run[2] = {0.f, 0.f};
  float acc[KS][4];
#pragma unroll
  for (int mt = 0; mt < KS; ++mt) acc[mt][0] = acc[mt][1] = acc[mt][2] = acc[mt][3] = 0.f;

  // One 16-token sub-tile of this warp: S^T = K.Q^T, online softmax, O^T += V^T.P^T.  Tokens >= lim are masked.
  auto process = [&](uint32_t sK, uint32_t sV, int tok_base, int lim) {
    const int mat = lane >> 3;
    const int sw = lane & 7;

    // ---- S^T = K . Q^T --------------------------------------------------------------------------
    float sacc[4] = {0.f, 0.f, 0.f, 0.f};
    {
      const int r = sw + (mat & 1) * 8;
      const uint32_t rowaddr = sK + r * C::kRowBytes;
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        uint32_t a[4];
        const int c = 2 * ks + (mat >> 1);
        ldmatrix_x4(a, rowaddr + ((c ^ sw) << 4));
        mma_bf16_16816(sacc, a, qb[ks][0], qb[ks][1]);
      }
    }

    // ---- online softmax over tokens (rows of S^T): lanes sharing lane%4 hold the same heads ------
    const int tokA = tok_base + (lane >> 2);
    float s0 = sacc[0] * p.scale_log2, s1 = sacc[1] * p.scale_log2;
    float s2 = sacc[2] * p.scale_log2, s3 = sacc[3] * p.scale_log2;
    if (tokA >= lim) s0 = s1 = -INFINITY;
    if (tokA + 8 >= lim) s2 = s3 = -INFINITY;
    float mx0 = fmaxf(s0, s2), mx1 = fmaxf(s1, s3);
#pragma unroll
    for (int o = 4; o < 32; o <<= 1) {
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, o));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, o));
    }
    const float mn0 = fmaxf(m_run[0], mx0), mn1 = fmaxf(m_run[1], mx1);
    const float ms0 = (mn0 == -INFINITY) ? 0.f : mn0, ms1 = (mn1 == -INFINITY) ? 0.f : mn1;
    const float corr0 = fast_exp2(m_run[0] - ms0), corr1 = fast_exp2(m_run[1] - ms1);
    const float p0 = fast_exp2(s0 - ms0), p1 = fast_exp2(s1 - ms1);
    const float p2 = fast_exp2(s2 - ms0), p3 = fast_exp2(s3 - ms1);
    l_run[0] = l_run[0] * corr0 + p0 + p2;
    l_run[1] = l_run[1] * corr1 + p1 + p3;
    m_run[0] = mn0;
    m_run[1] = mn1;
    const uint32_t pb0 = movmatrix_trans(pack_bf16x2(p0, p1));  // tokens 0-7  -> B rows k
    const uint32_t pb1 = movmatrix_trans(pack_bf16x2(p2, p3));  // tokens 8-15

    // ---- O^T += V^T . P^T -------------------------------------------------------------------------
    {
      const int r = sw + (mat >> 1) * 8;
      const uint32_t rowaddr = sV + r * C::kRowBytes;
      const bool rescale = __any_sync(0xffffffffu, (corr0 != 1.f) || (corr1 != 1.f));
#pragma unroll
      for (int mt = 0; mt < KS; ++mt) {
        if (rescale) {
          acc[mt][0] *= corr0;
          acc[mt][1] *= corr1;
          acc[mt][2] *= corr0;
          acc[mt][3] *= corr1;
        }
        uint32_t a[4];
        const int c = 2 * mt + (mat & 1);
        ldmatrix_x4_trans(a, rowaddr + ((c ^ sw) << 4));
        mma_bf16_16816(acc[mt], a, pb0, pb1);
      }
    }
  };

  for (int i = 0; i < n; ++i) {
    cp_async_wait<STAGES - 2>();
    __syncwarp();
    if (i + STAGES - 1 < n) issue(i + STAGES - 1);
    cp_async_commit();
    const uint32_t sK = warp_smem + (i % STAGES) * C::kStageBytes;
    process(sK, sK + C::kSubTileBytes, ((t0 + i) * NWARPS + warp) * 16, ctx);
  }

  if constexpr (FUSE) {
    if (own_new) process(smem_u32(smem) + new_off, smem_u32(smem) + new_off + C::kSubTileBytes, ctx, ctx_all);
  }

  // ---- merge the NWARPS partial states through shared memory ------------------------------------
  cp_async_wait<0>();
#pragma unroll
  for (int o = 4; o < 32; o <<= 1) {
    l_run[0] += __shfl_xor_sync(0xffffffffu, l_run[0], o);
    l_run[1] += __shfl_xor_sync(0xffffffffu, l_run[1], o);
  }
  __syncthreads();  // every warp is done with its pipeline buffers
  float* so = reinterpret_cast<float*>(smem);                   // [NWARPS][8][kOStride]
  float* sml = so + NWARPS * 8 * C::kOStride;                   // [NWARPS][8][2]
  {
    const int h0 = (lane & 3) * 2;
    float* sw_o = so + warp * 8 * C::kOStride;
#pragma unroll
    for (int mt = 0; mt < KS; ++mt) {
      const int d = mt * 16 + (lane >> 2);
      sw_o[h0 * C::kOStride + d] = acc[mt][0];
      sw_o[(h0 + 1) * C::kOStride + d] = acc[mt][1];
      sw_o[h0 * C::kOStride + d + 8] = acc[mt][2];
      sw_o[(h0 + 1) * C::kOStride + d + 8] = acc[mt][3];
    }
    if (lane < 4) {
      float* w_ml = sml + warp * 16;
      w_ml[h0 * 2 + 0] = m_run[0];
      w_ml[h0 * 2 + 1] = l_run[0];
      w_ml[(h0 + 1) * 2 + 0] = m_run[1];
      w_ml[(h0 + 1) * 2 + 1] = l_run[1];
    }
  }
  __syncthreads();

  for (int idx = threadIdx.x; idx < gvalid * D; idx += NWARPS * 32) {
    const int hn = idx / D;
    const int d = idx - hn * D;
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < NWARPS; ++w) M = fmaxf(M, sml[w * 16 + hn * 2]);
    float o = 0.f, L = 0.f;
    if (M != -INFINITY) {
#pragma unroll
      for (int w = 0; w < NWARPS; ++w) {
        const float mw = sml[w * 16 + hn * 2];
        const float f = (mw == -INFINITY) ? 0.f : fast_exp2(mw - M);
        o += so[(w * 8 + hn) * C::kOStride + d] * f;
        L += sml[w * 16 + hn * 2 + 1] * f;
      }
    }
    const int qh = qh0 + hn;
    if (p.num_splits == 1) {
      p.out[b * p.out_stride + static_cast<int64_t>(qh) * D + d] =
          __float2bfloat16_rn(L > 0.f ? o / L : 0.f);
    } else {
      const int64_t slot = (static_cast<int64_t>(b) * p.H + qh) * p.num_splits + split;
      p.ws_o[slot * D + d] = o;
      if (d == 0) {
        p.ws_ml[slot * 2 + 0] = M;
        p.ws_ml[slot * 2 + 1] = L;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------------------------
// Persistent variant (one split): gridDim.x = min(items, resident CTAs) CTAs walk the (sequence, kv head, q group) items
// round-robin, and every warp's cp.async pipeline runs ACROSS item boundaries: while the four warps merge their partial
// states of item j and write its output, the first STAGES-1 sub-tiles of item j+1 are already in flight (its Q fragments
// and context length were requested an item earlier).  The per-CTA ramp (launch, first table lookup, pipeline fill) and
// drain of the one-CTA-per-item kernel — ~10 % of a 393 KB item at the cfg2 shape (769-token contexts) — are paid once
// per CTA instead of once per item.  The merge buffer is separate from the pipeline buffers for that reason.
template <int D, int STAGES, int NWARPS>
__global__ void __launch_bounds__(NWARPS * 32, 2)
paged_decode_persist_kernel(const DecodeParams p) {
  using C = DecodeCfg<D, STAGES, NWARPS>;
  constexpr int KS = D / 16;
  constexpr int TILE = 16 * NWARPS;
  extern __shared__ __align__(128) uint8_t smem[];
  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int per_b = p.Hkv * p.ngrp;
  const int items = p.batch * per_b;
  const uint32_t warp_smem = smem_u32(smem) + warp * C::kWarpBytes;
  float* so = reinterpret_cast<float*>(smem + C::kPipeBytes);  // [NWARPS][8][kOStride]
  float* sml = so + NWARPS * 8 * C::kOStride;                  // [NWARPS][8][2]
  pdl_wait();  // q and the newest KV row were written by the previous kernel

  // ---- producer cursor (per warp): item, tile within it, and that item's metadata ---------------------------------
  int p_it = blockIdx.x;          // item the next load belongs to
  int p_t = 0, p_n = 0, p_ctx = 0;
  const int32_t* p_table = nullptr;
  const bf16* p_k = nullptr;
  const bf16* p_v = nullptr;
  int nxt_ctx = 0;                // context length of the item after p_it (requested one item ahead)
  auto ctx_of = [&](int it) { return it < items ? __ldg(p.ctx + it / per_b) : 0; };
  auto enter = [&](int ctx) {     // producer enters item p_it whose context length is ctx
    if (p_it < items) {
      const int b = p_it / per_b;
      const int kvh = (p_it - b * per_b) / p.ngrp;
      p_ctx = ctx;
      p_n = (ctx + TILE - 1) / TILE;
      p_table = p.tables + static_cast<int64_t>(b) * p.max_blocks;
      p_k = p.kc + kvh * p.head_stride;
      p_v = p.vc + kvh * p.head_stride;
    } else {
      p_n = 0;
    }
    p_t = 0;
    nxt_ctx = ctx_of(p_it + gridDim.x);
  };
  enter(ctx_of(p_it));
  int issued = 0;                 // groups committed so far (stage = issued % STAGES)
  auto issue = [&]() {            // one commit group: the next sub-tile of this warp, or nothing when the work is done
    while (p_it < items && p_t >= p_n) {  // skip to the next non-empty item
      p_it += gridDim.x;
      enter(nxt_ctx);
    }
    if (p_it < items) {
      const int s = issued % STAGES;
      const int tok_base = (p_t * NWARPS + warp) * 16;
      int blk = -1, off = 0;
      if (tok_base < p_ctx) {
        const int bi = tok_base / p.block_size;
        off = tok_base - bi * p.block_size;
        if (bi < p.max_blocks) blk = __ldg(p_table + bi);
      }
      const uint32_t sK = warp_smem + s * C::kStageBytes;
      const uint32_t sV = sK + C::kSubTileBytes;
      const int c = lane % C::kChunks;
      const int r0 = lane / C::kChunks;
      const int64_t gofs0 = static_cast<int64_t>(blk < 0 ? 0 : blk) * p.blk_stride + c * 8;
#pragma unroll
      for (int j = 0; j < C::kRowsPerLane; ++j) {
        const int r = r0 + j * C::kRowStep;
        const bool valid = (blk >= 0) && (tok_base + r < p_ctx);
        const int64_t gofs = valid ? gofs0 + static_cast<int64_t>(off + r) * p.tok_stride : 0;
        const uint32_t dst = r * C::kRowBytes + ((c ^ (r & 7)) << 4);
        cp_async_16(sK + dst, p_k + gofs, valid ? 16 : 0);
        cp_async_16(sV + dst, p_v + gofs, valid ? 16 : 0);
      }
      ++p_t;
    }
    cp_async_commit();
    ++issued;
  };
#pragma unroll
  for (int s = 0; s < STAGES - 1; ++s) issue();

  // ---- Q^T B-fragments of an item (head n = lane/4, dims 16ks + 2(lane%4) + {0,1} (+8)) ---------------------------
  auto load_q = [&](int it, uint32_t (&qf)[KS][2]) {
    const int itc = it < items ? it : 0;
    const int b = itc / per_b;
    const int rem = itc - b * per_b;
    const int kvh = rem / p.ngrp, grp = rem - kvh * p.ngrp;
    const int gvalid = min(8, p.g - grp * 8);
    const int hn = lane >> 2;
    const bool hv = hn < gvalid && it < items;
    const bf16* qrow = p.q + b * p.q_stride + static_cast<int64_t>(kvh * p.g + grp * 8 + (hv ? hn : 0)) * D;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      const int d0 = ks * 16 + (lane & 3) * 2;
      qf[ks][0] = hv ? *reinterpret_cast<const uint32_t*>(qrow + d0) : 0u;
      qf[ks][1] = hv ? *reinterpret_cast<const uint32_t*>(qrow + d0 + 8) : 0u;
    }
  };
  uint32_t qb[KS][2], qn[KS][2];
  load_q(blockIdx.x, qb);
  int c_ctx = ctx_of(blockIdx.x);
  int consumed = 0;  // sub-tiles consumed so far (stage = consumed % STAGES)

  for (int it = blockIdx.x; it < items; it += gridDim.x) {
    const int b = it / per_b;
    const int rem = it - b * per_b;
    const int kvh = rem / p.ngrp, grp = rem - kvh * p.ngrp;
    const int gvalid = min(8, p.g - grp * 8);
    const int qh0 = kvh * p.g + grp * 8;
    const int ctx = c_ctx;
    const int n = (ctx + TILE - 1) / TILE;
    load_q(it + gridDim.x, qn);            // next item's Q and context length: needed an item from now
    c_ctx = ctx_of(it + gridDim.x);

    float m_run[2] = {-INFINITY, -INFINITY};
    float l_run[2] = {0.f, 0.f};
    float acc[KS][4];
#pragma unroll
    for (int mt = 0; mt < KS; ++mt) acc[mt][0] = acc[mt][1] = acc[mt][2] = acc[mt][3] = 0.f;

    for (int i = 0; i < n; ++i) {
      cp_async_wait<STAGES - 2>();
      __syncwarp();
      issue();
      const uint32_t sK = warp_smem + (consumed % STAGES) * C::kStageBytes;
      const uint32_t sV = sK + C::kSubTileBytes;
      ++consumed;
      const int tok_base = (i * NWARPS + warp) * 16;
      const int mat = lane >> 3;
      const int sw = lane & 7;
      // ---- S^T = K . Q^T ------------------------------------------------------------------------
      float sacc[4] = {0.f, 0.f, 0.f, 0.f};
      {
        const int r = sw + (mat & 1) * 8;
        const uint32_t rowaddr = sK + r * C::kRowBytes;
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
          uint32_t a[4];
          const int c = 2 * ks + (mat >> 1);
          ldmatrix_x4(a, rowaddr + ((c ^ sw) << 4));
          mma_bf16_16816(sacc, a, qb[ks][0], qb[ks][1]);
        }
      }
      // ---- online softmax over tokens --------------------------------------------------------------
      const int tokA = tok_base + (lane >> 2);
      float s0 = sacc[0] * p.scale_log2, s1 = sacc[1] * p.scale_log2;
      float s2 = sacc[2] * p.scale_log2, s3 = sacc[3] * p.scale_log2;
      if (tokA >= ctx) s0 = s1 = -INFINITY;
      if (tokA + 8 >= ctx) s2 = s3 = -INFINITY;
      float mx0 = fmaxf(s0, s2), mx1 = fmaxf(s1, s3);
#pragma unroll
      for (int o = 4; o < 32; o <<= 1) {
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, o));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, o));
      }
      const float mn0 = fmaxf(m_run[0], mx0), mn1 = fmaxf(m_run[1], mx1);
      const float ms0 = (mn0 == -INFINITY) ? 0.f : mn0, ms1 = (mn1 == -INFINITY) ? 0.f : mn1;
      const float corr0 = fast_exp2(m_run[0] - ms0), corr1 = fast_exp2(m_run[1] - ms1);
      const float p0 = fast_exp2(s0 - ms0), p1 = fast_exp2(s1 - ms1);
      const float p2 = fast_exp2(s2 - ms0), p3 = fast_exp2(s3 - ms1);
      l_run[0] = l_run[0] * corr0 + p0 + p2;
      l_run[1] = l_run[1] * corr1 + p1 + p3;
      m_run[0] = mn0;
      m_run[1] = mn1;
      const uint32_t pb0 = movmatrix_trans(pack_bf16x2(p0, p1));
      const uint32_t pb1 = movmatrix_trans(pack_bf16x2(p2, p3));
      // ---- O^T += V^T . P^T -------------------------------------------------------------------------
      {
        const int r = sw + (mat >> 1) * 8;
        const uint32_t rowaddr = sV + r * C::kRowBytes;
        const bool rescale = __any_sync(0xffffffffu, (corr0 != 1.f) || (corr1 != 1.f));
#pragma unroll
        for (int mt = 0; mt < KS; ++mt) {
          if (rescale) {
            acc[mt][0] *= corr0;
            acc[mt][1] *= corr1;
            acc[mt][2] *= corr0;
            acc[mt][3] *= corr1;
          }
          uint32_t a[4];
          const int c = 2 * mt + (mat & 1);
          ldmatrix_x4_trans(a, rowaddr + ((c ^ sw) << 4));
          mma_bf16_16816(acc[mt], a, pb0, pb1);
        }
      }
    }

    // ---- merge the NWARPS partial states (the next item's sub-tiles keep arriving meanwhile) ----------------------
#pragma unroll
    for (int o = 4; o < 32; o <<= 1) {
      l_run[0] += __shfl_xor_sync(0xffffffffu, l_run[0], o);
      l_run[1] += __shfl_xor_sync(0xffffffffu, l_run[1], o);
    }
    {
      const int h0 = (lane & 3) * 2;
      float* sw_o = so + warp * 8 * C::kOStride;
#pragma unroll
      for (int mt = 0; mt < KS; ++mt) {
        const int d = mt * 16 + (lane >> 2);
        sw_o[h0 * C::kOStride + d] = acc[mt][0];
        sw_o[(h0 + 1) * C::kOStride + d] = acc[mt][1];
        sw_o[h0 * C::kOStride + d + 8] = acc[mt][2];
        sw_o[(h0 + 1) * C::kOStride + d + 8] = acc[mt][3];
      }
      if (lane < 4) {
        float* w_ml = sml + warp * 16;
        w_ml[h0 * 2 + 0] = m_run[0];
        w_ml[h0 * 2 + 1] = l_run[0];
        w_ml[(h0 + 1) * 2 + 0] = m_run[1];
        w_ml[(h0 + 1) * 2 + 1] = l_run[1];
      }
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < gvalid * D; idx += NWARPS * 32) {
      const int hn = idx / D;
      const int d = idx - hn * D;
      float M = -INFINITY;
#pragma unroll
      for (int w = 0; w < NWARPS; ++w) M = fmaxf(M, sml[w * 16 + hn * 2]);
      float o = 0.f, L = 0.f;
      if (M != -INFINITY) {
#pragma unroll
        for (int w = 0; w < NWARPS; ++w) {
          const float mw = sml[w * 16 + hn * 2];
          const float f = (mw == -INFINITY) ? 0.f : fast_exp2(mw - M);
          o += so[(w * 8 + hn) * C::kOStride + d] * f;
          L += sml[w * 16 + hn * 2 + 1] * f;
        }
      }
      p.out[b * p.out_stride + static_cast<int64_t>(qh0 + hn) * D + d] = __float2bfloat16_rn(L > 0.f ? o / L : 0.f);
    }
    __syncthreads();  // the merge buffer is free again
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) { qb[ks][0] = qn[ks][0]; qb[ks][1] = qn[ks][1]; }
  }
  cp_async_wait<0>();
}

// Combines the split-K partials: grid (H, B), D threads.
template <int D>
__global__ void __launch_bounds__(D) paged_decode_reduce_kernel(const DecodeParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const int qh = blockIdx.x, b = blockIdx.y, d = threadIdx.x;
  const int64_t base = (static_cast<int64_t>(b) * p.H + qh) * p.num_splits;
  float M = -INFINITY;
  for (int s = 0; s < p.num_splits; ++s) M = fmaxf(M, __ldg(p.ws_ml + (base + s) * 2));
  float o = 0.f, L = 0.f;
  if (M != -INFINITY) {
    for (int s = 0; s < p.num_splits; ++s) {
      const float ms = __ldg(p.ws_ml + (base + s) * 2);
      if (ms == -INFINITY) continue;
      const float f = fast_exp2(ms - M);
      o += p.ws_o[(base + s) * D + d] * f;
      L += __ldg(p.ws_ml + (base + s) * 2 + 1) * f;
    }
  }
  p.out[b * p.out_stride + static_cast<int64_t>(qh) * D + d] = __float2bfloat16_rn(L > 0.f ? o / L : 0.f);
}

}  // namespace nvllm

// ------------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------------
namespace nvllm {

template <int D, int STAGES, int NWARPS, bool FUSE = false>
static int launch_decode(const DecodeParams& p, cudaStream_t stream) {
  using C = DecodeCfg<D, STAGES, NWARPS>;
  auto kern = paged_decode_kernel<D, STAGES, NWARPS, FUSE>;
  constexpr int kBytes = C::kSmemBytes + (FUSE ? 2 * C::kSubTileBytes : 0);  // + the step's own K/V sub-tile
  NVLLM_SMEM_ATTR_ONCE(kern, kBytes);
  dim3 grid(p.num_splits, p.Hkv * p.ngrp, p.batch);
  NVLLM_CUDA(launch_pdl(kern, grid, dim3(NWARPS * 32), kBytes, stream, p));
  if (p.num_splits > 1)
    NVLLM_CUDA(launch_pdl(paged_decode_reduce_kernel<D>, dim3(p.H, p.batch), dim3(D), 0, stream, p));
  return 0;
}

template <int D, int STAGES, int NWARPS>
static int launch_decode_persist(const DecodeParams& p, cudaStream_t stream) {
  using C = DecodeCfg<D, STAGES, NWARPS>;
  auto kern = paged_decode_persist_kernel<D, STAGES, NWARPS>;
  constexpr int kBytes = C::kPipeBytes + C::kMergeBytes;
  NVLLM_SMEM_ATTR_ONCE(kern, kBytes);
  static std::atomic<int> per_sm_cached{0};
  int per_sm = per_sm_cached.load(std::memory_order_relaxed);
  if (per_sm == 0) {
    NVLLM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, NWARPS * 32, kBytes));
    if (per_sm < 1) per_sm = 1;
    per_sm_cached.store(per_sm, std::memory_order_relaxed);
  }
  const long items = static_cast<long>(p.batch) * p.Hkv * p.ngrp;
  const long slots = static_cast<long>(per_sm) * sm_count();
  NVLLM_CUDA(launch_pdl(kern, dim3(static_cast<unsigned>(items < slots ? items : slots)), dim3(NWARPS * 32), kBytes, stream, p));
  return 0;
}

// Tuning knob for benchmarking only (bench.py --decode-cfg): 0 = default.
static int decode_cfg_id() { return knobs().decode_cfg; }

}  // namespace nvllm

extern "C" size_t nvllm_paged_decode_workspace_bytes(int32_t batch, int32_t num_heads,
                                                     int32_t head_dim, int32_t num_splits) {
  if (batch <= 0 || num_heads <= 0 || head_dim <= 0) return 0;
  if (num_splits <= 0) num_splits = 64;  // upper bound of nvllm_paged_decode_choose_splits
  if (num_splits == 1) return 0;
  const size_t slots = static_cast<size_t>(batch) * num_heads * num_splits;
  return slots * head_dim * sizeof(float) + slots * 2 * sizeof(float);
}

extern "C" int nvllm_paged_decode_choose_splits(int32_t batch, int32_t num_heads,
                                                int32_t num_kv_heads, int32_t max_context_len) {
  if (batch <= 0 || num_heads <= 0 || num_kv_heads <= 0) return 1;
  const int g = num_heads / num_kv_heads;
  const int ngrp = (g + 7) / 8;
  const long base = static_cast<long>(batch) * num_kv_heads * ngrp;
  const long slots = 2L * nvllm::sm_count();  // 2 resident CTAs per SM
  long max_s = max_context_len / 256;          // keep >= 256 tokens per split
  if (max_s < 1) max_s = 1;
  if (max_s > 64) max_s = 64;
  // Measured (profiles/r02_decode_split_sweep.jsonl, 512 seqs x 1536 tokens, Hkv = 1 / 2 / 4 / 8 per rank): once the grid
  // holds a full wave, ONE split wins at every shard shape — 5.94 vs 5.16 TB/s at Hkv = 1 (512 CTAs = 1.73 waves), where
  // round 1's wave-filling model picked 4.  A split costs its f32 partials, a second launch and shorter CTAs (6-24 % at
  // 2-8 splits); dynamic CTA scheduling already hides most of a ragged last wave.  Split only to FILL the first wave
  // (few sequences x long contexts: 64 x 32.5K at Hkv = 1 -> 4 splits = 256 CTAs, the measured best).
  if (base >= slots) return 1;
  long sp = slots / base;
  if (sp > max_s) sp = max_s;
  return sp < 1 ? 1 : static_cast<int>(sp);
}

namespace nvllm {
struct FuseArgs {  // non-NULL positions == fused entry point
  const int64_t* positions = nullptr;
  const float* cos_sin = nullptr;
  const void* q_norm_w = nullptr;
  const void* k_norm_w = nullptr;
  const int32_t* slot_mapping = nullptr;
  float eps = 0.f;
  int max_position = 0;
};

static int decode_dispatch(void* out, const void* q, void* k_cache, void* v_cache, const int32_t* block_tables,
                           const int32_t* context_lens, int32_t batch, int32_t num_heads, int32_t max_blocks_per_seq,
                           int64_t q_token_stride, int64_t out_token_stride, const nvllm_kv_layout* layout, float scale,
                           void* workspace, size_t workspace_bytes, int32_t num_splits, int32_t dtype, void* stream,
                           const FuseArgs& fa) {
  const bool fuse = fa.positions != nullptr;
  NVLLM_CHECK_ARG(dtype == NVLLM_BF16, NVLLM_ERR_DTYPE, "paged_decode: dtype %d (bf16 only)", dtype);
  if (int rc = check_layout(layout)) return rc;
  NVLLM_CHECK_ARG(batch >= 0 && batch <= 65535, NVLLM_ERR_ARG, "paged_decode: batch %d", batch);
  if (batch == 0) return 0;
  const int Hkv = layout->num_kv_heads, D = layout->head_dim;
  NVLLM_CHECK_ARG(num_heads > 0 && num_heads % Hkv == 0, NVLLM_ERR_ARG,
                  "paged_decode: num_heads %d not a multiple of num_kv_heads %d", num_heads, Hkv);
  NVLLM_CHECK_ARG(layout->block_size % 16 == 0, NVLLM_ERR_ARG,
                  "paged_decode: block_size %d must be a multiple of 16", layout->block_size);
  NVLLM_CHECK_ARG(max_blocks_per_seq > 0, NVLLM_ERR_ARG, "paged_decode: max_blocks_per_seq %d",
                  max_blocks_per_seq);
  NVLLM_CHECK_ARG(out && q && k_cache && v_cache && block_tables && context_lens, NVLLM_ERR_ARG,
                  "paged_decode: NULL pointer");
  NVLLM_CHECK_ARG(aligned16(out) && aligned16(q) && aligned16(k_cache) && aligned16(v_cache),
                  NVLLM_ERR_ALIGN, "paged_decode: pointers must be 16-byte aligned");
  const int64_t q_row = fuse ? static_cast<int64_t>(num_heads + 2 * Hkv) * D : static_cast<int64_t>(num_heads) * D;
  NVLLM_CHECK_ARG(q_token_stride % 8 == 0 && out_token_stride % 8 == 0 && q_token_stride >= q_row &&
                      out_token_stride >= static_cast<int64_t>(num_heads) * D,
                  NVLLM_ERR_ALIGN, "paged_decode: q/out token stride");
  if (fuse) {
    NVLLM_CHECK_ARG(D == 128, NVLLM_ERR_ARG, "paged_decode_fused: head_dim %d (128 only)", D);
    NVLLM_CHECK_ARG(fa.cos_sin && fa.slot_mapping && fa.max_position > 0 && (fa.q_norm_w == nullptr) == (fa.k_norm_w == nullptr),
                    NVLLM_ERR_ARG, "paged_decode_fused: cos_sin / slot_mapping / norm weights");
  }
  const int g = num_heads / Hkv;
  if (num_splits <= 0)
    num_splits = nvllm_paged_decode_choose_splits(batch, num_heads, Hkv,
                                                  max_blocks_per_seq * layout->block_size);
  NVLLM_CHECK_ARG(num_splits <= 1024, NVLLM_ERR_ARG, "paged_decode: num_splits %d", num_splits);
  DecodeParams p;
  p.out = static_cast<bf16*>(out);
  p.q = static_cast<const bf16*>(q);
  p.kc = static_cast<const bf16*>(k_cache);
  p.vc = static_cast<const bf16*>(v_cache);
  p.kc_w = static_cast<bf16*>(k_cache);
  p.vc_w = static_cast<bf16*>(v_cache);
  p.qnw = static_cast<const bf16*>(fa.q_norm_w);
  p.knw = static_cast<const bf16*>(fa.k_norm_w);
  p.cos_sin = fa.cos_sin;
  p.pos = fa.positions;
  p.slots = fa.slot_mapping;
  p.eps = fa.eps;
  p.max_position = fa.max_position;
  p.tables = block_tables;
  p.ctx = context_lens;
  p.ws_o = nullptr;
  p.ws_ml = nullptr;
  if (num_splits > 1) {
    const size_t need = nvllm_paged_decode_workspace_bytes(batch, num_heads, D, num_splits);
    NVLLM_CHECK_ARG(workspace != nullptr && workspace_bytes >= need, NVLLM_ERR_WORKSPACE,
                    "paged_decode: workspace %zu < %zu bytes", workspace_bytes, need);
    NVLLM_CHECK_ARG(aligned16(workspace), NVLLM_ERR_ALIGN, "paged_decode: workspace alignment");
    const size_t slots = static_cast<size_t>(batch) * num_heads * num_splits;
    p.ws_o = static_cast<float*>(workspace);
    p.ws_ml = p.ws_o + slots * D;
  }
  p.batch = batch;
  p.H = num_heads;
  p.Hkv = Hkv;
  p.g = g;
  p.ngrp = (g + 7) / 8;
  p.max_blocks = max_blocks_per_seq;
  p.block_size = layout->block_size;
  p.num_splits = num_splits;
  p.q_stride = q_token_stride;
  p.out_stride = out_token_stride;
  p.blk_stride = layout->block_stride;
  p.tok_stride = layout->token_stride;
  p.head_stride = layout->head_stride;
  p.scale_log2 = scale * 1.4426950408889634f;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (fuse) return launch_decode<128, 3, 4, true>(p, st);
  // One split and more items than resident CTAs: the persistent kernel (NVLLM_DECODE_PERSIST=0: one CTA per item).
  if (D == 128 && num_splits == 1 && knobs().decode_persist && decode_cfg_id() == 0 &&
      static_cast<long>(batch) * Hkv * p.ngrp > 2L * sm_count())
    return launch_decode_persist<128, 3, 4>(p, st);
  if (D == 128) {
    switch (decode_cfg_id()) {
      case 1: return launch_decode<128, 4, 4>(p, st);
      case 2: return launch_decode<128, 3, 8>(p, st);
      case 3: return launch_decode<128, 2, 4>(p, st);   // 64 KiB smem: 3 CTAs / SM
      case 4: return launch_decode<128, 4, 2>(p, st);   // 2 warps x 4 stages, 64 KiB: 3 CTAs / SM
      case 5: return launch_decode<128, 3, 2>(p, st);   // 2 warps x 3 stages, 48 KiB: 4 CTAs / SM
      case 6: return launch_decode<128, 6, 2>(p, st);   // 2 warps x 6 stages, 96 KiB: 2 CTAs / SM, deeper per-warp pipeline
      default: return launch_decode<128, 3, 4>(p, st);
    }
  }
  return launch_decode<64, 4, 4>(p, st);
}
}  // namespace nvllm

extern "C" int nvllm_paged_decode_attention(void* out, const void* q, const void* k_cache,
                                            const void* v_cache, const int32_t* block_tables,
                                            const int32_t* context_lens, int32_t batch,
                                            int32_t num_heads, int32_t max_blocks_per_seq,
                                            int64_t q_token_stride, int64_t out_token_stride,
                                            const nvllm_kv_layout* layout, float scale,
                                            void* workspace, size_t workspace_bytes,
                                            int32_t num_splits, int32_t dtype, void* stream) {
  return nvllm::decode_dispatch(out, q, const_cast<void*>(k_cache), const_cast<void*>(v_cache), block_tables, context_lens,
                                batch, num_heads, max_blocks_per_seq, q_token_stride, out_token_stride, layout, scale, workspace,
                                workspace_bytes, num_splits, dtype, stream, nvllm::FuseArgs{});
}

extern "C" int nvllm_paged_decode_attention_fused(void* out, const void* qkv, const int64_t* positions, const float* cos_sin,
                                                  const void* q_norm_w, const void* k_norm_w, float eps, int32_t max_position,
                                                  void* k_cache, void* v_cache, const int32_t* slot_mapping,
                                                  const int32_t* block_tables, const int32_t* context_lens, int32_t batch,
                                                  int32_t num_heads, int32_t max_blocks_per_seq, int64_t qkv_token_stride,
                                                  int64_t out_token_stride, const nvllm_kv_layout* layout, float scale,
                                                  void* workspace, size_t workspace_bytes, int32_t num_splits, int32_t dtype,
                                                  void* stream) {
  NVLLM_CHECK_ARG(positions != nullptr, NVLLM_ERR_ARG, "paged_decode_fused: positions is NULL");
  nvllm::FuseArgs fa;
  fa.positions = positions; fa.cos_sin = cos_sin; fa.q_norm_w = q_norm_w; fa.k_norm_w = k_norm_w;
  fa.slot_mapping = slot_mapping; fa.eps = eps; fa.max_position = max_position;
  return nvllm::decode_dispatch(out, qkv, k_cache, v_cache, block_tables, context_lens, batch, num_heads, max_blocks_per_seq,
                                qkv_token_stride, out_token_stride, layout, scale, workspace, workspace_bytes, num_splits, dtype,
                                stream, fa);
}

// ------------------------------------------------------------------------------------------------
// Split-K decode attention for callers that cannot hand us a workspace.
//
// Who this is for: code that reaches the library from a place where device memory cannot be obtained —
// a host callback or plugin hook that only receives the tensors of one operator, a runtime whose memory pool
// is frozen once serving starts, a C or Rust embedder without a CUDA allocator of its own.  Such a caller has its
// inputs and its output and nothing else, yet the split-K path needs somewhere to park the f32 partial sums between
// its two launches.  The array below is that place: it lives in the library's own device image, so it exists on
// every GPU that loads libnvllm.so from the moment the module is loaded until the process ends, and nobody has to
// allocate or free it.
//
// The price, stated once: kStaticScratchBytes (8 MiB) of HBM per GPU for the life of the process; one array per
// GPU shared by every call, so two calls that may run at the same time on one GPU (different streams, or
// different host threads on different streams) overwrite each other's partials; and a hard size limit.
// ------------------------------------------------------------------------------------------------
namespace nvllm {
constexpr size_t kStaticScratchBytes = size_t(8) << 20;
constexpr int kStaticScratchMaxDevices = 64;
static __device__ __align__(256) float g_decode_static_scratch[kStaticScratchBytes / sizeof(float)];

// In host code `g_decode_static_scratch` is only a handle the runtime uses to look the array up; its value is not a
// device address.  The real address differs per GPU, so it is asked for once per device, with that device current.
static int decode_static_scratch_ptr(float** out) {
  static std::atomic<float*> per_device[kStaticScratchMaxDevices];
  int dev = -1;
  NVLLM_CUDA(cudaGetDevice(&dev));
  NVLLM_CHECK_ARG(dev >= 0 && dev < kStaticScratchMaxDevices, NVLLM_ERR_ARG,
                  "paged_decode_static: device ordinal %d outside [0, %d)", dev, kStaticScratchMaxDevices);
  float* ptr = per_device[dev].load(std::memory_order_acquire);
  if (ptr == nullptr) {
    void* sym = nullptr;
    NVLLM_CUDA(cudaGetSymbolAddress(&sym, g_decode_static_scratch));
    NVLLM_CHECK_ARG(sym != nullptr, NVLLM_ERR_STATE, "paged_decode_static: cudaGetSymbolAddress returned NULL on device %d", dev);
    ptr = static_cast<float*>(sym);
    per_device[dev].store(ptr, std::memory_order_release);  // two threads racing here store the same address
  }
  *out = ptr;
  return 0;
}
}  // namespace nvllm

extern "C" size_t nvllm_paged_decode_static_scratch_bytes(void) { return nvllm::kStaticScratchBytes; }

extern "C" int nvllm_paged_decode_attention_static(void* out, const void* q, const void* k_cache, const void* v_cache,
                                                   const int32_t* block_tables, const int32_t* context_lens,
                                                   int32_t batch, int32_t num_heads, int32_t max_blocks_per_seq,
                                                   int64_t q_token_stride, int64_t out_token_stride,
                                                   const nvllm_kv_layout* layout, float scale, int32_t num_splits,
                                                   int32_t dtype, void* stream) {
  NVLLM_CHECK_ARG(layout != nullptr, NVLLM_ERR_ARG, "paged_decode_static: layout is NULL");
  if (num_splits <= 0)
    num_splits = nvllm_paged_decode_choose_splits(batch, num_heads, layout->num_kv_heads,
                                                  max_blocks_per_seq * layout->block_size);
  const size_t need = nvllm_paged_decode_workspace_bytes(batch, num_heads, layout->head_dim, num_splits);
  NVLLM_CHECK_ARG(need <= nvllm::kStaticScratchBytes, NVLLM_ERR_WORKSPACE,
                  "paged_decode_static: batch %d x heads %d x splits %d x head_dim %d needs %zu bytes of partials, the "
                  "built-in scratch holds %zu; use fewer splits or nvllm_paged_decode_attention with your own workspace",
                  batch, num_heads, num_splits, layout->head_dim, need, nvllm::kStaticScratchBytes);
  float* scratch = nullptr;
  if (int rc = nvllm::decode_static_scratch_ptr(&scratch)) return rc;
  // Nothing has to be cleared: the first launch writes every (sequence, head, split) partial the second one reads,
  // empty splits included (m = -inf, l = 0), so what an earlier call left behind is never looked at.
  return nvllm::decode_dispatch(out, q, const_cast<void*>(k_cache), const_cast<void*>(v_cache), block_tables, context_lens,
                                batch, num_heads, max_blocks_per_seq, q_token_stride, out_token_stride, layout, scale,
                                scratch, nvllm::kStaticScratchBytes, num_splits, dtype, stream, nvllm::FuseArgs{});
}
