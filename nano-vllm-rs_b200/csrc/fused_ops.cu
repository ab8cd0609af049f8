// fused_ops.cu — the HBM-bound element-wise / reduction kernels of the layer:
//   K6  add + RMSNorm            (layernorm.rs:58-97, qwen3.rs:382,389)
//   K7  q/k RMSNorm + RoPE (+K1 KV-cache store)  (rotary_embedding.rs:23-48,136-158; attention.rs:78-81)
//   K8  SiLU * up                (activation.rs:46-63)
//   K9  embedding gather, K10 last-token gather  (embed_head.rs:77-97,272-289)
//   K11 greedy argmax            (sampler.rs:109-112)
// Every kernel touches each byte once with 16-byte (or full-warp 128-byte) accesses; rounding points
// follow the reference's tensor boundaries so the CPU oracle (oracle/layers.py) can match them.
#include <math.h>

#include "common.cuh"

namespace nvllm {

int check_layout(const nvllm_kv_layout* L);  // store_kvcache.cu

// ------------------------------------------------------------------------------------------------
// K6: one CTA per token row; each thread keeps <= 2 16-byte vectors of the row in registers.
// ------------------------------------------------------------------------------------------------
template <bool HAS_RES>
__global__ void __launch_bounds__(1024)
add_rmsnorm_kernel(bf16* __restrict__ out, const bf16* __restrict__ x, bf16* __restrict__ residual,
                   const bf16* __restrict__ weight, int hidden, float eps) {
  pdl_launch_dependents();
  pdl_wait();
  const int row = blockIdx.x;
  const int nvec = hidden >> 3;
  const uint4* xr = reinterpret_cast<const uint4*>(x + static_cast<int64_t>(row) * hidden);
  uint4* rr = HAS_RES ? reinterpret_cast<uint4*>(residual + static_cast<int64_t>(row) * hidden) : nullptr;
  uint4* orow = reinterpret_cast<uint4*>(out + static_cast<int64_t>(row) * hidden);
  const uint4* wv = reinterpret_cast<const uint4*>(weight);

  float v[2][8];
  float ss = 0.f;
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int i = threadIdx.x + k * blockDim.x;
    if (i < nvec) {
      uint4 a = xr[i];
      const uint32_t aw[4] = {a.x, a.y, a.z, a.w};
      if (HAS_RES) {
        uint4 r = rr[i];
        const uint32_t rw[4] = {r.x, r.y, r.z, r.w};
        uint32_t sw[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          // bf16 tensor add (qwen3.rs:382,389): round the sum to bf16, then upcast for the norm
          sw[j] = pack_bf16x2(bf16_lo(aw[j]) + bf16_lo(rw[j]), bf16_hi(aw[j]) + bf16_hi(rw[j]));
          v[k][2 * j] = bf16_lo(sw[j]);
          v[k][2 * j + 1] = bf16_hi(sw[j]);
        }
        rr[i] = make_uint4(sw[0], sw[1], sw[2], sw[3]);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          v[k][2 * j] = bf16_lo(aw[j]);
          v[k][2 * j + 1] = bf16_hi(aw[j]);
        }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) ss += v[k][j] * v[k][j];
    }
  }
  __shared__ float red[32];
  __shared__ float s_inv;
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : 0.f;
    t = warp_sum(t);
    if (threadIdx.x == 0) s_inv = 1.0f / sqrtf(t / static_cast<float>(hidden) + eps);
  }
  __syncthreads();
  const float inv = s_inv;
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int i = threadIdx.x + k * blockDim.x;
    if (i < nvec) {
      uint4 w = __ldg(wv + i);
      const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        o[j] = pack_bf16x2(v[k][2 * j] * inv * bf16_lo(ww[j]), v[k][2 * j + 1] * inv * bf16_hi(ww[j]));
      orow[i] = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// K7 (+K1): one warp per (token, head) of the packed qkv row.  Lane l owns elements
// [EPL*l, EPL*l+EPL) of each half (EPL = D/64), i.e. exactly the RoPE pairs (j, j + D/2).
// ------------------------------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(256)
qknorm_rope_store_kernel(bf16* __restrict__ qkv, const int64_t* __restrict__ positions,
                         const float* __restrict__ cos_sin, const bf16* __restrict__ qw,
                         const bf16* __restrict__ kw, float eps, int num_tokens, int H, int Hkv,
                         int max_position, bf16* __restrict__ k_cache, bf16* __restrict__ v_cache,
                         const int32_t* __restrict__ slot_mapping, int block_size,
                         int64_t blk_stride, int64_t tok_stride, int64_t head_stride) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int EPL = D / 64;
  constexpr int HALF = D / 2;
  const int heads = H + 2 * Hkv;
  const int64_t gw = (blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x) >> 5;
  if (gw >= static_cast<int64_t>(num_tokens) * heads) return;
  const int lane = threadIdx.x & 31;
  const int tok = static_cast<int>(gw / heads);
  const int h = static_cast<int>(gw - static_cast<int64_t>(tok) * heads);
  bf16* row = qkv + (static_cast<int64_t>(tok) * heads + h) * D;
  const int kind = h < H ? 0 : (h < H + Hkv ? 1 : 2);  // q, k, v

  float x1[EPL], x2[EPL];
  if constexpr (EPL == 2) {
    const uint32_t a = *reinterpret_cast<const uint32_t*>(row + 2 * lane);
    const uint32_t b = *reinterpret_cast<const uint32_t*>(row + HALF + 2 * lane);
    x1[0] = bf16_lo(a); x1[EPL - 1] = bf16_hi(a);
    x2[0] = bf16_lo(b); x2[EPL - 1] = bf16_hi(b);
  } else {
    x1[0] = __bfloat162float(row[lane]);
    x2[0] = __bfloat162float(row[HALF + lane]);
  }

  if (kind < 2) {
    const bf16* w = kind == 0 ? qw : kw;
    if (w != nullptr) {  // per-head RMSNorm over D (Qwen3 q_norm / k_norm), output rounded to bf16
      float ss = 0.f;
#pragma unroll
      for (int e = 0; e < EPL; ++e) ss += x1[e] * x1[e] + x2[e] * x2[e];
      ss = warp_sum(ss);
      const float inv = 1.0f / sqrtf(ss / static_cast<float>(D) + eps);
#pragma unroll
      for (int e = 0; e < EPL; ++e) {
        const float w1 = __bfloat162float(w[EPL * lane + e]);
        const float w2 = __bfloat162float(w[HALF + EPL * lane + e]);
        x1[e] = __bfloat162float(__float2bfloat16_rn(x1[e] * inv * w1));
        x2[e] = __bfloat162float(__float2bfloat16_rn(x2[e] * inv * w2));
      }
    }
    int64_t pos = positions[tok];
    if (pos < 0) pos = 0;
    if (pos >= max_position) pos = max_position - 1;
    const float* cs = cos_sin + pos * D;
#pragma unroll
    for (int e = 0; e < EPL; ++e) {
      const float c = __ldg(cs + EPL * lane + e);
      const float s = __ldg(cs + HALF + EPL * lane + e);
      const float o1 = x1[e] * c - x2[e] * s;
      const float o2 = x2[e] * c + x1[e] * s;
      x1[e] = o1;
      x2[e] = o2;
    }
  }

  uint32_t o1p = 0, o2p = 0;
  bf16 o1s, o2s;
  if constexpr (EPL == 2) {
    o1p = pack_bf16x2(x1[0], x1[EPL - 1]);
    o2p = pack_bf16x2(x2[0], x2[EPL - 1]);
  } else {
    o1s = __float2bfloat16_rn(x1[0]);
    o2s = __float2bfloat16_rn(x2[0]);
  }
  if (kind < 2) {
    if constexpr (EPL == 2) {
      *reinterpret_cast<uint32_t*>(row + 2 * lane) = o1p;
      *reinterpret_cast<uint32_t*>(row + HALF + 2 * lane) = o2p;
    } else {
      row[lane] = o1s;
      row[HALF + lane] = o2s;
    }
  }
  if (kind >= 1 && k_cache != nullptr) {
    const int slot = __ldg(slot_mapping + tok);
    if (slot >= 0) {
      const int blk = slot / block_size;
      const int off = slot - blk * block_size;
      const int kvh = kind == 1 ? h - H : h - H - Hkv;
      bf16* dst = (kind == 1 ? k_cache : v_cache) + blk * blk_stride + off * tok_stride + kvh * head_stride;
      if constexpr (EPL == 2) {
        *reinterpret_cast<uint32_t*>(dst + 2 * lane) = o1p;
        *reinterpret_cast<uint32_t*>(dst + HALF + 2 * lane) = o2p;
      } else {
        dst[lane] = o1s;
        dst[HALF + lane] = o2s;
      }
    }
  }
}

__global__ void rope_table_kernel(float* __restrict__ cos_sin, int max_position, int D, double theta) {
  const int half = D / 2;
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= static_cast<int64_t>(max_position) * half) return;
  const int pos = static_cast<int>(i / half);
  const int j = static_cast<int>(i - static_cast<int64_t>(pos) * half);
  const double inv_freq = 1.0 / pow(theta, static_cast<double>(2 * j) / static_cast<double>(D));
  const double ang = static_cast<double>(pos) * inv_freq;
  cos_sin[static_cast<int64_t>(pos) * D + j] = static_cast<float>(cos(ang));
  cos_sin[static_cast<int64_t>(pos) * D + half + j] = static_cast<float>(sin(ang));
}

}  // namespace nvllm

namespace nvllm {

// K8: out[t, j] = bf16( bf16(silu(g)) * u ) — two roundings, as the reference's two tensor ops.
__global__ void __launch_bounds__(256)
silu_mul_kernel(uint4* __restrict__ out, const uint4* __restrict__ x, int64_t total_vec, int inter_vec) {
  pdl_launch_dependents();
  pdl_wait();
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total_vec;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t t = i / inter_vec;
    const int j = static_cast<int>(i - t * inter_vec);
    const uint4 g = ld_nc_16(x + t * 2 * inter_vec + j);
    const uint4 u = ld_nc_16(x + t * 2 * inter_vec + inter_vec + j);
    const uint32_t gw[4] = {g.x, g.y, g.z, g.w};
    const uint32_t uw[4] = {u.x, u.y, u.z, u.w};
    uint32_t o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float g0 = bf16_lo(gw[k]), g1 = bf16_hi(gw[k]);
      const uint32_t s = pack_bf16x2(g0 / (1.0f + expf(-g0)), g1 / (1.0f + expf(-g1)));
      o[k] = pack_bf16x2(bf16_lo(s) * bf16_lo(uw[k]), bf16_hi(s) * bf16_hi(uw[k]));
    }
    out[i] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// K9 / K10: row gather, one warp per output row, 16-byte vectors.
template <typename IndexFn>
__device__ __forceinline__ void gather_rows(uint4* __restrict__ out, const uint4* __restrict__ src,
                                            int rows, int row_vec, IndexFn index_of) {
  const int64_t gw = (blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x) >> 5;
  if (gw >= rows) return;
  const int lane = threadIdx.x & 31;
  const int64_t r = index_of(static_cast<int>(gw));
  const uint4* s = src + r * row_vec;
  uint4* d = out + gw * row_vec;
  for (int i = lane; i < row_vec; i += 32) d[i] = __ldg(s + i);
}
__global__ void __launch_bounds__(256)
embedding_gather_kernel(uint4* out, const uint4* table, const int64_t* __restrict__ ids, int rows,
                        int row_vec, int vocab) {
  pdl_launch_dependents();
  pdl_wait();
  gather_rows(out, table, rows, row_vec, [&](int i) {
    int64_t id = ids[i];
    return id < 0 ? 0 : (id >= vocab ? static_cast<int64_t>(vocab - 1) : id);
  });
}
__global__ void __launch_bounds__(256)
gather_last_kernel(uint4* out, const uint4* x, const int32_t* __restrict__ cu, int rows, int row_vec) {
  pdl_launch_dependents();
  pdl_wait();
  gather_rows(out, x, rows, row_vec, [&](int i) { return static_cast<int64_t>(cu[i + 1] - 1); });
}

// K11: one CTA per row; first maximum wins (== candle argmax / oracle np.argmax).
template <typename T>
__device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<bf16>(bf16 v) { return __bfloat162float(v); }

template <typename T>
__global__ void __launch_bounds__(1024)
argmax_kernel(int64_t* __restrict__ out_ids, float* __restrict__ out_max, const T* __restrict__ logits,
              int vocab, int64_t ld, int64_t vocab_offset) {
  pdl_launch_dependents();
  pdl_wait();
  const T* row = logits + blockIdx.x * ld;
  float best = -INFINITY;
  int bidx = 0x7fffffff;
  constexpr int VEC = 16 / sizeof(T);  // elements per 16-byte load
  int head = 0;                        // scalar prologue up to the first 16-byte aligned element
  {
    const uintptr_t a = reinterpret_cast<uintptr_t>(row) & 15u;
    if (a) head = min(vocab, static_cast<int>((16 - a) / sizeof(T)));
  }
  for (int i = threadIdx.x; i < head; i += blockDim.x) {
    const float v = to_f32<T>(row[i]);
    if (v > best || (v == best && i < bidx)) { best = v; bidx = i; }
  }
  const int nvec = (vocab - head) / VEC;
  const uint4* rv = reinterpret_cast<const uint4*>(row + head);
  for (int j = threadIdx.x; j < nvec; j += blockDim.x) {
    const uint4 q = ld_nc_16(rv + j);
    T e[VEC];
    *reinterpret_cast<uint4*>(e) = q;
#pragma unroll
    for (int k = 0; k < VEC; ++k) {  // ascending index inside the vector keeps the first-max rule
      const float v = to_f32<T>(e[k]);
      const int i = head + j * VEC + k;
      if (v > best || (v == best && i < bidx)) { best = v; bidx = i; }
    }
  }
  for (int i = head + nvec * VEC + threadIdx.x; i < vocab; i += blockDim.x) {
    const float v = to_f32<T>(row[i]);
    if (v > best || (v == best && i < bidx)) { best = v; bidx = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bidx, o);
    if (ov > best || (ov == best && oi < bidx)) { best = ov; bidx = oi; }
  }
  __shared__ float sv[32];
  __shared__ int si[32];
  if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = bidx; }
  __syncthreads();
  if (threadIdx.x < 32) {
    const int nw = blockDim.x >> 5;
    best = threadIdx.x < nw ? sv[threadIdx.x] : -INFINITY;
    bidx = threadIdx.x < nw ? si[threadIdx.x] : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bidx, o);
      if (ov > best || (ov == best && oi < bidx)) { best = ov; bidx = oi; }
    }
    if (threadIdx.x == 0) {
      if (bidx == 0x7fffffff) bidx = 0;  // all -inf / NaN row
      out_ids[blockIdx.x] = static_cast<int64_t>(bidx) + vocab_offset;
      if (out_max) out_max[blockIdx.x] = best;
    }
  }
}

__global__ void argmax_combine_kernel(int64_t* __restrict__ out_ids, const float* __restrict__ all_max,
                                      const int64_t* __restrict__ all_ids, int tp, int batch) {
  pdl_launch_dependents();
  pdl_wait();
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  float best = all_max[b];
  int64_t id = all_ids[b];
  for (int r = 1; r < tp; ++r) {
    const float v = all_max[r * batch + b];
    if (v > best) { best = v; id = all_ids[r * batch + b]; }
  }
  out_ids[b] = id;
}

// Seeded init shared bit-for-bit with oracle/weights.py.
__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
  x ^= x >> 27; x *= 0x94D049BB133111EBull;
  x ^= x >> 31;
  return x;
}
template <typename T>
__global__ void __launch_bounds__(256)
init_weight_kernel(T* __restrict__ w, int64_t count, uint64_t key, float scale) {
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < count;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const uint64_t r = splitmix64(key + static_cast<uint64_t>(i + 1) * 0x9E3779B97F4A7C15ull);
    const int s = static_cast<int>((r & 0xffff) + ((r >> 16) & 0xffff) + ((r >> 32) & 0xffff) + (r >> 48)) - 131070;
    const float v = __fmul_rn(static_cast<float>(s), scale);
    if constexpr (sizeof(T) == 2) w[i] = __float2bfloat16_rn(v); else w[i] = v;
  }
}

static inline unsigned grid_for(int64_t work_items, int threads, int per_sm) {
  int64_t blocks = (work_items + threads - 1) / threads;
  const int64_t cap = static_cast<int64_t>(sm_count()) * per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<unsigned>(blocks);
}

}  // namespace nvllm

using namespace nvllm;

extern "C" int nvllm_add_rmsnorm(void* out, const void* x, void* residual, const void* weight,
                                 int32_t num_tokens, int32_t hidden, float eps, int32_t dtype,
                                 void* stream) {
  NVLLM_CHECK_ARG(dtype == NVLLM_BF16, NVLLM_ERR_DTYPE, "add_rmsnorm: dtype %d (bf16 only)", dtype);
  NVLLM_CHECK_ARG(num_tokens >= 0 && hidden > 0 && hidden % 8 == 0 && hidden <= 16384, NVLLM_ERR_ARG,
                  "add_rmsnorm: num_tokens=%d hidden=%d (hidden %% 8 == 0, <= 16384)", num_tokens, hidden);
  if (num_tokens == 0) return 0;
  NVLLM_CHECK_ARG(out && x && weight, NVLLM_ERR_ARG, "add_rmsnorm: NULL pointer");
  NVLLM_CHECK_ARG(aligned16(out) && aligned16(x) && aligned16(weight) && aligned16(residual),
                  NVLLM_ERR_ALIGN, "add_rmsnorm: pointers must be 16-byte aligned");
  const int nvec = hidden / 8;
  int threads = ((nvec + 1) / 2 + 31) / 32 * 32;  // 2 vectors per thread when the row is long
  if (nvec <= 256) threads = (nvec + 31) / 32 * 32;
  if (threads > 1024) threads = 1024;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (residual)
    NVLLM_CUDA(launch_pdl(add_rmsnorm_kernel<true>, dim3(num_tokens), dim3(threads), 0, st, static_cast<bf16*>(out),
                          static_cast<const bf16*>(x), static_cast<bf16*>(residual), static_cast<const bf16*>(weight),
                          hidden, eps));
  else
    NVLLM_CUDA(launch_pdl(add_rmsnorm_kernel<false>, dim3(num_tokens), dim3(threads), 0, st, static_cast<bf16*>(out),
                          static_cast<const bf16*>(x), static_cast<bf16*>(nullptr), static_cast<const bf16*>(weight),
                          hidden, eps));
  NVLLM_LAUNCH_CHECK("add_rmsnorm_kernel");
  return 0;
}

extern "C" int nvllm_qknorm_rope_store(void* qkv, const int64_t* positions, const float* cos_sin,
                                       const void* q_norm_w, const void* k_norm_w, float eps,
                                       int32_t num_tokens, int32_t num_heads, int32_t num_kv_heads,
                                       int32_t head_dim, int32_t max_position, void* k_cache,
                                       void* v_cache, const int32_t* slot_mapping,
                                       const nvllm_kv_layout* layout, int32_t dtype, void* stream) {
  NVLLM_CHECK_ARG(dtype == NVLLM_BF16, NVLLM_ERR_DTYPE, "qknorm_rope_store: dtype %d", dtype);
  NVLLM_CHECK_ARG(head_dim == 64 || head_dim == 128, NVLLM_ERR_ARG, "qknorm_rope_store: head_dim %d", head_dim);
  NVLLM_CHECK_ARG(num_tokens >= 0 && num_heads > 0 && num_kv_heads > 0 && max_position > 0,
                  NVLLM_ERR_ARG, "qknorm_rope_store: bad sizes");
  if (num_tokens == 0) return 0;
  NVLLM_CHECK_ARG(qkv && positions && cos_sin, NVLLM_ERR_ARG, "qknorm_rope_store: NULL pointer");
  NVLLM_CHECK_ARG((q_norm_w == nullptr) == (k_norm_w == nullptr), NVLLM_ERR_ARG,
                  "qknorm_rope_store: q_norm_w and k_norm_w must both be set or both NULL");
  int64_t bs = 1, s_blk = 0, s_tok = 0, s_head = 0;
  if (k_cache != nullptr) {
    if (int rc = check_layout(layout)) return rc;
    NVLLM_CHECK_ARG(v_cache && slot_mapping, NVLLM_ERR_ARG, "qknorm_rope_store: v_cache/slot_mapping NULL");
    NVLLM_CHECK_ARG(layout->num_kv_heads == num_kv_heads && layout->head_dim == head_dim, NVLLM_ERR_ARG,
                    "qknorm_rope_store: layout heads/dim mismatch");
    bs = layout->block_size; s_blk = layout->block_stride; s_tok = layout->token_stride;
    s_head = layout->head_stride;
  }
  const int64_t warps = static_cast<int64_t>(num_tokens) * (num_heads + 2 * num_kv_heads);
  const unsigned blocks = static_cast<unsigned>((warps + 7) / 8);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define NVLLM_ROPE_LAUNCH(DD)                                                                         \
  NVLLM_CUDA(launch_pdl(qknorm_rope_store_kernel<DD>, dim3(blocks), dim3(256), 0, st,                 \
      static_cast<bf16*>(qkv), positions, cos_sin, static_cast<const bf16*>(q_norm_w),                \
      static_cast<const bf16*>(k_norm_w), eps, num_tokens, num_heads, num_kv_heads, max_position,     \
      static_cast<bf16*>(k_cache), static_cast<bf16*>(v_cache), slot_mapping, static_cast<int>(bs),   \
      s_blk, s_tok, s_head))
  if (head_dim == 128) NVLLM_ROPE_LAUNCH(128); else NVLLM_ROPE_LAUNCH(64);
#undef NVLLM_ROPE_LAUNCH
  NVLLM_LAUNCH_CHECK("qknorm_rope_store_kernel");
  return 0;
}

extern "C" int nvllm_rope_table(float* cos_sin, int32_t max_position, int32_t head_dim, double theta,
                                void* stream) {
  NVLLM_CHECK_ARG(cos_sin && max_position > 0 && head_dim > 0 && head_dim % 2 == 0 && theta > 0,
                  NVLLM_ERR_ARG, "rope_table: bad arguments");
  const int64_t n = static_cast<int64_t>(max_position) * (head_dim / 2);
  rope_table_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      cos_sin, max_position, head_dim, theta);
  NVLLM_LAUNCH_CHECK("rope_table_kernel");
  return 0;
}

extern "C" int nvllm_silu_mul(void* out, const void* x, int32_t num_tokens, int32_t inter,
                              int32_t dtype, void* stream) {
  NVLLM_CHECK_ARG(dtype == NVLLM_BF16, NVLLM_ERR_DTYPE, "silu_mul: dtype %d", dtype);
  NVLLM_CHECK_ARG(num_tokens >= 0 && inter > 0 && inter % 8 == 0, NVLLM_ERR_ARG,
                  "silu_mul: num_tokens=%d inter=%d (inter %% 8 == 0)", num_tokens, inter);
  if (num_tokens == 0) return 0;
  NVLLM_CHECK_ARG(out && x && aligned16(out) && aligned16(x), NVLLM_ERR_ALIGN, "silu_mul: pointers");
  const int64_t total = static_cast<int64_t>(num_tokens) * (inter / 8);
  NVLLM_CUDA(launch_pdl(silu_mul_kernel, dim3(grid_for(total, 256, 8)), dim3(256), 0, static_cast<cudaStream_t>(stream),
                        static_cast<uint4*>(out), static_cast<const uint4*>(x), total, inter / 8));
  NVLLM_LAUNCH_CHECK("silu_mul_kernel");
  return 0;
}

extern "C" int nvllm_embedding_gather(void* out, const void* table, const int64_t* ids,
                                      int32_t num_tokens, int32_t hidden, int32_t vocab,
                                      int32_t dtype, void* stream) {
  NVLLM_CHECK_ARG(dtype == NVLLM_BF16, NVLLM_ERR_DTYPE, "embedding_gather: dtype %d", dtype);
  NVLLM_CHECK_ARG(num_tokens >= 0 && hidden > 0 && hidden % 8 == 0 && vocab > 0, NVLLM_ERR_ARG,
                  "embedding_gather: bad sizes");
  if (num_tokens == 0) return 0;
  NVLLM_CHECK_ARG(out && table && ids && aligned16(out) && aligned16(table), NVLLM_ERR_ALIGN,
                  "embedding_gather: pointers");
  NVLLM_CUDA(launch_pdl(embedding_gather_kernel, dim3((num_tokens + 7) / 8), dim3(256), 0, static_cast<cudaStream_t>(stream),
                        static_cast<uint4*>(out), static_cast<const uint4*>(table), ids, num_tokens, hidden / 8, vocab));
  NVLLM_LAUNCH_CHECK("embedding_gather_kernel");
  return 0;
}

extern "C" int nvllm_gather_last_tokens(void* out, const void* x, const int32_t* cu_seqlens_q,
                                        int32_t batch, int32_t hidden, int32_t dtype, void* stream) {
  NVLLM_CHECK_ARG(dtype == NVLLM_BF16, NVLLM_ERR_DTYPE, "gather_last_tokens: dtype %d", dtype);
  NVLLM_CHECK_ARG(batch >= 0 && hidden > 0 && hidden % 8 == 0, NVLLM_ERR_ARG, "gather_last_tokens: sizes");
  if (batch == 0) return 0;
  NVLLM_CHECK_ARG(out && x && cu_seqlens_q && aligned16(out) && aligned16(x), NVLLM_ERR_ALIGN,
                  "gather_last_tokens: pointers");
  NVLLM_CUDA(launch_pdl(gather_last_kernel, dim3((batch + 7) / 8), dim3(256), 0, static_cast<cudaStream_t>(stream),
                        static_cast<uint4*>(out), static_cast<const uint4*>(x), cu_seqlens_q, batch, hidden / 8));
  NVLLM_LAUNCH_CHECK("gather_last_kernel");
  return 0;
}

extern "C" int nvllm_argmax(int64_t* out_ids, float* out_max, const void* logits, int32_t batch,
                            int32_t vocab, int64_t ld, int64_t vocab_offset, int32_t logits_dtype,
                            void* stream) {
  NVLLM_CHECK_ARG(batch >= 0 && vocab > 0 && ld >= vocab, NVLLM_ERR_ARG, "argmax: sizes");
  if (batch == 0) return 0;
  NVLLM_CHECK_ARG(out_ids && logits, NVLLM_ERR_ARG, "argmax: NULL pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (logits_dtype == NVLLM_F32)
    NVLLM_CUDA(launch_pdl(argmax_kernel<float>, dim3(batch), dim3(1024), 0, st, out_ids, out_max, static_cast<const float*>(logits), vocab, ld, vocab_offset));
  else if (logits_dtype == NVLLM_BF16)
    NVLLM_CUDA(launch_pdl(argmax_kernel<bf16>, dim3(batch), dim3(1024), 0, st, out_ids, out_max, static_cast<const bf16*>(logits), vocab, ld, vocab_offset));
  else
    NVLLM_CHECK_ARG(false, NVLLM_ERR_DTYPE, "argmax: dtype %d", logits_dtype);
  NVLLM_LAUNCH_CHECK("argmax_kernel");
  return 0;
}

extern "C" int nvllm_argmax_combine(int64_t* out_ids, const float* all_max, const int64_t* all_ids,
                                    int32_t tp, int32_t batch, void* stream) {
  NVLLM_CHECK_ARG(tp > 0 && batch >= 0, NVLLM_ERR_ARG, "argmax_combine: sizes");
  if (batch == 0) return 0;
  NVLLM_CHECK_ARG(out_ids && all_max && all_ids, NVLLM_ERR_ARG, "argmax_combine: NULL pointer");
  NVLLM_CUDA(launch_pdl(argmax_combine_kernel, dim3((batch + 127) / 128), dim3(128), 0, static_cast<cudaStream_t>(stream),
                        out_ids, all_max, all_ids, tp, batch));
  NVLLM_LAUNCH_CHECK("argmax_combine_kernel");
  return 0;
}

// out[b, r * vshard + j] = in[r, b, j]: the all-gathered vocab shards of a tensor-parallel LM head put back into full rows
// (the gather the reference leaves as a TODO, embed_head.rs:321-336).  16-byte vectors when vshard % 4 == 0.
namespace nvllm {
__global__ void assemble_vocab_shards_kernel(float* __restrict__ out, const float* __restrict__ in, int tp, int rows,
                                             int vshard, int64_t total_vec, int vec) {
  pdl_launch_dependents();
  pdl_wait();
  const int per_row = vshard / vec;  // vectors per (rank, row)
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total_vec;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t rb = i / per_row;            // r * rows + b
    const int j = static_cast<int>(i - rb * per_row);
    const int r = static_cast<int>(rb / rows), b = static_cast<int>(rb - static_cast<int64_t>(r) * rows);
    const int64_t dst = (static_cast<int64_t>(b) * tp + r) * vshard + static_cast<int64_t>(j) * vec;
    if (vec == 4) *reinterpret_cast<uint4*>(out + dst) = ld_nc_16(in + i * 4);
    else out[dst] = in[i];
  }
}
}  // namespace nvllm

extern "C" int nvllm_assemble_vocab_shards(float* out, const float* gathered, int32_t tp, int32_t rows, int32_t vshard,
                                           void* stream) {
  using namespace nvllm;
  NVLLM_CHECK_ARG(tp > 0 && rows >= 0 && vshard > 0, NVLLM_ERR_ARG, "assemble_vocab_shards: sizes");
  if (rows == 0) return 0;
  NVLLM_CHECK_ARG(out && gathered && out != gathered, NVLLM_ERR_ARG, "assemble_vocab_shards: pointers");
  const int vec = (vshard % 4 == 0 && aligned16(out) && aligned16(gathered)) ? 4 : 1;
  const int64_t total_vec = static_cast<int64_t>(tp) * rows * (vshard / vec);
  NVLLM_CUDA(launch_pdl(assemble_vocab_shards_kernel, dim3(grid_for(total_vec, 256, 8)), dim3(256), 0,
                        static_cast<cudaStream_t>(stream), out, gathered, static_cast<int>(tp), static_cast<int>(rows),
                        static_cast<int>(vshard), total_vec, vec));
  return 0;
}

extern "C" int nvllm_init_weight(void* w, int64_t count, uint64_t seed, uint64_t tensor_id, float std,
                                 int32_t dtype, void* stream) {
  NVLLM_CHECK_ARG(count >= 0, NVLLM_ERR_ARG, "init_weight: count");
  if (count == 0) return 0;
  NVLLM_CHECK_ARG(w != nullptr, NVLLM_ERR_ARG, "init_weight: NULL pointer");
  const uint64_t key = seed ^ (tensor_id << 40);
  const float scale = std * 2.6428998e-05f;  // 1 / std of the 4-field Irwin-Hall sum (37837.23)
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == NVLLM_BF16)
    init_weight_kernel<bf16><<<grid_for(count, 256, 8), 256, 0, st>>>(static_cast<bf16*>(w), count, key, scale);
  else if (dtype == NVLLM_F32)
    init_weight_kernel<float><<<grid_for(count, 256, 8), 256, 0, st>>>(static_cast<float*>(w), count, key, scale);
  else
    NVLLM_CHECK_ARG(false, NVLLM_ERR_DTYPE, "init_weight: dtype %d", dtype);
  NVLLM_LAUNCH_CHECK("init_weight_kernel");
  return 0;
}
