// sampler.cu — K12: Sampler::forward / sample_single (reference src/layers/sampler.rs:26-219) on the device.
//
// The reference samples one row at a time on the host: a blocking scalar read per parameter, two full sorts of the
// vocabulary (top-k, top-p), a softmax and an unseeded Tensor::rand.  Here one CTA owns one row of the f32 logits and
// never sorts: both filters are *selections*, done MSB-first over the order-preserving integer image of the scaled
// logits with 256-bin shared-memory histograms (4 passes each over a row that stays in L2):
//   top-k  (sampler.rs:115-149)  k-th largest key by COUNT; ties at the threshold keep the lowest indices (the
//                                reference's stable sort), resolved by an ordered ballot scan only when they straddle k
//   top-p  (sampler.rs:152-191)  smallest prefix of the descending order whose probability mass reaches p, found by
//                                MASS: exp(s - max) in 32.32 fixed point, so every sum is an integer sum (order-free,
//                                deterministic); target = ceil(p * total)
//   Gumbel-max (sampler.rs:194-219) argmax(s + g), g = -log(-log(u)), first maximum wins
// Rows with temperature == 0 take the same final pass without filters or noise (greedy, sampler.rs:79-81,109-112).
// Seeded contract for u (the reference's draws are unseeded and cannot be reproduced): see oracle/layers.py.
// HBM/L2-bound integer + compare work; no tensor cores.
#include <math.h>

#include "common.cuh"

namespace nvllm {
namespace {

constexpr int kThreads = 512;
constexpr int kWarps = kThreads / 32;
constexpr uint64_t kGolden = 0x9E3779B97F4A7C15ull;
constexpr unsigned kFull = 0xffffffffu;

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
  x ^= x >> 27; x *= 0x94D049BB133111EBull;
  x ^= x >> 31;
  return x;
}
// order-preserving float -> uint32 (larger float <=> larger key); -0 is folded onto +0 first
__device__ __forceinline__ uint32_t okey(float f) {
  const uint32_t u = __float_as_uint(f + 0.0f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float okey_inv(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}
__device__ __forceinline__ float scaled(float x, float T) { return T == 1.0f ? x : __fdiv_rn(x, T); }  // sampler.rs:84-88
// exp(s - m) in 32.32 fixed point (s <= m, so the value is in [0, 2^32])
__device__ __forceinline__ unsigned long long fx_exp(float s, float m) {
  return __float2ull_rz(expf(s - m) * 4294967296.0f);
}

// keep(key, i) <=> key > sel.key || (key == sel.key && i <= sel.cut)
struct Sel {
  uint32_t key, cut;
};
__device__ __forceinline__ bool keeps(const Sel& s, uint32_t key, uint32_t i) {
  return key > s.key || (key == s.key && i <= s.cut);
}

struct Shared {
  unsigned long long hist[256];
  unsigned long long need, eq;
  float redf[32];
  int redi[32];
  uint32_t cnt[kWarps];
  int sel_bin;
  uint32_t cut;
};

// MSB-first radix selection over the keys of one row, largest first.  weight(key, i) is the element's contribution
// (0 = not a candidate).  Finds the key K with  sum{w : key > K} < target <= sum{w : key >= K}  and returns how much of
// the target the group key == K has to supply (`need`, >= 1) and that group's total weight (`eq`).
// MASS == true: the target is ceil(p * total weight), fixed after the first pass.
template <bool MASS, typename WF>
__device__ void radix_select(Shared& sh, const float* __restrict__ row, int V, float T, WF weight, unsigned long long target,
                             double p, uint32_t& out_key, unsigned long long& out_need, unsigned long long& out_eq) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint32_t prefix = 0, mask = 0;
  unsigned long long rem = target;
  for (int pass = 0; pass < 4; ++pass) {
    const int shift = 24 - 8 * pass;
    for (int b = tid; b < 256; b += kThreads) sh.hist[b] = 0ull;
    __syncthreads();
    for (int i = tid; i < V; i += kThreads) {
      const uint32_t key = okey(scaled(row[i], T));
      if ((key & mask) == prefix) {
        const unsigned long long w = weight(key, static_cast<uint32_t>(i));
        if (w) atomicAdd(&sh.hist[(key >> shift) & 255u], w);
      }
    }
    __syncthreads();
    if (warp == 0) {  // lane l owns bins 255-8l .. 248-8l, scanned from the top
      unsigned long long loc[8], s = 0ull;
#pragma unroll
      for (int j = 0; j < 8; ++j) { loc[j] = sh.hist[255 - 8 * lane - j]; s += loc[j]; }
      unsigned long long incl = s;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long t = __shfl_up_sync(kFull, incl, o);
        if (lane >= o) incl += t;
      }
      const unsigned long long total = __shfl_sync(kFull, incl, 31);
      if (MASS && pass == 0) {
        const double t = ceil(p * static_cast<double>(total));
        rem = t < 1.0 ? 1ull : static_cast<unsigned long long>(t);
      }
      if (rem > total) rem = total;  // cannot happen for valid inputs; keeps the search well defined
      if (rem < 1ull) rem = 1ull;
      if (total == 0ull && lane == 0) { sh.sel_bin = 0; sh.need = 1ull; sh.eq = 0ull; }  // no candidate at all
      const unsigned long long excl = incl - s;
      if (excl < rem && rem <= incl) {  // exactly one lane
        unsigned long long c = excl;
        int bsel = -1;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if (bsel < 0) {
            if (c + loc[j] >= rem) { bsel = 255 - 8 * lane - j; sh.need = rem - c; sh.eq = loc[j]; }
            else c += loc[j];
          }
        }
        sh.sel_bin = bsel;
      }
    }
    __syncthreads();
    prefix |= static_cast<uint32_t>(sh.sel_bin) << shift;
    mask |= 255u << shift;
    rem = sh.need;
    out_eq = sh.eq;
    __syncthreads();
  }
  out_key = prefix;
  out_need = rem;
}

// Smallest index I with #{i <= I : pred(i)} == need (pred true for at least `need` elements).  Each warp counts a
// contiguous chunk with ballots, the warp holding the need-th match walks its chunk again.
template <typename PF>
__device__ uint32_t ordered_cut(Shared& sh, int V, PF pred, uint32_t need) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int chunk = ((V + kWarps - 1) / kWarps + 31) & ~31;
  const int start = warp * chunk, end = min(V, start + chunk);
  uint32_t cnt = 0;
  for (int base = start; base < end; base += 32) {
    const int i = base + lane;
    cnt += __popc(__ballot_sync(kFull, i < end && pred(static_cast<uint32_t>(i))));
  }
  if (lane == 0) sh.cnt[warp] = cnt;
  if (tid == 0) sh.cut = 0xffffffffu;
  __syncthreads();
  uint32_t run = 0;
  int wsel = -1;
  for (int w = 0; w < kWarps; ++w) {
    if (wsel < 0) {
      if (run + sh.cnt[w] >= need) wsel = w; else run += sh.cnt[w];
    }
  }
  if (warp == wsel) {
    uint32_t c = run;
    for (int base = start; base < end; base += 32) {
      const int i = base + lane;
      const bool p = i < end && pred(static_cast<uint32_t>(i));
      const unsigned bal = __ballot_sync(kFull, p);
      const uint32_t n = __popc(bal);
      if (c + n >= need) {
        const uint32_t r = need - c;  // 1-based rank inside this ballot
        if (p && static_cast<uint32_t>(__popc(bal & (kFull >> (31 - lane)))) == r) sh.cut = static_cast<uint32_t>(i);
        break;
      }
      c += n;
    }
  }
  __syncthreads();
  const uint32_t cut = sh.cut;
  __syncthreads();
  return cut;
}

__device__ float block_max(Shared& sh, float v) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_max(v);
  if (lane == 0) sh.redf[warp] = v;
  __syncthreads();
  float r = sh.redf[0];
  for (int w = 1; w < kWarps; ++w) r = fmaxf(r, sh.redf[w]);
  __syncthreads();
  return r;
}

__global__ void __launch_bounds__(kThreads)
sample_kernel(int64_t* __restrict__ out_ids, const float* __restrict__ logits, int V, int64_t ld,
              const float* __restrict__ temperatures, const int32_t* __restrict__ top_k, const float* __restrict__ top_p,
              const uint64_t* __restrict__ row_seeds, int32_t* __restrict__ kept_count, uint8_t* __restrict__ kept_mask) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ Shared sh;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.x;
  const float* __restrict__ row = logits + b * ld;
  const float T = temperatures[b];
  const bool greedy = T == 0.0f;
  Sel sk{0u, 0xffffffffu}, sp{0u, 0xffffffffu};  // key >= 0 always holds: no filtering

  if (!greedy) {
    // ---- top-k: the k-th largest key by count (sampler.rs:115-149) ----
    const int k = top_k != nullptr ? top_k[b] : 0;
    if (k > 0 && k < V) {
      uint32_t key;
      unsigned long long need, eq;
      radix_select<false>(sh, row, V, T, [](uint32_t, uint32_t) { return 1ull; }, static_cast<unsigned long long>(k), 0.0, key,
                          need, eq);
      sk.key = key;
      if (need < eq)  // ties straddle k: the stable sort keeps the lowest indices
        sk.cut = ordered_cut(sh, V, [&](uint32_t i) { return okey(scaled(row[i], T)) == key; }, static_cast<uint32_t>(need));
    }
    // ---- top-p: smallest descending prefix whose softmax mass reaches p (sampler.rs:152-191) ----
    const float p = top_p != nullptr ? top_p[b] : 0.0f;
    if (p > 0.0f && p < 1.0f) {
      float m = -INFINITY;  // the maximum always survives top-k
      for (int i = tid; i < V; i += kThreads) m = fmaxf(m, scaled(row[i], T));
      m = block_max(sh, m);
      const Sel skc = sk;
      auto w_of = [&](uint32_t key, uint32_t i) { return keeps(skc, key, i) ? fx_exp(okey_inv(key), m) : 0ull; };
      uint32_t key;
      unsigned long long need, eq;
      radix_select<true>(sh, row, V, T, w_of, 0ull, static_cast<double>(p), key, need, eq);
      sp.key = key;
      const unsigned long long w1 = fx_exp(okey_inv(key), m);  // every element of the tie group weighs the same
      if (w1 > 0ull) {
        const unsigned long long need_cnt = (need + w1 - 1ull) / w1, eq_cnt = eq / w1;
        if (need_cnt < eq_cnt)
          sp.cut = ordered_cut(sh, V, [&](uint32_t i) { const uint32_t kk = okey(scaled(row[i], T)); return kk == key && keeps(skc, kk, i); },
                               static_cast<uint32_t>(need_cnt));
      }
    }
  }

  // ---- final pass: Gumbel-max over the survivors (sampler.rs:194-219), or plain argmax for greedy rows ----
  const uint64_t seed = greedy ? 0ull : row_seeds[b];
  float best = -INFINITY;
  int bidx = 0x7fffffff, kept = 0;
  for (int i = tid; i < V; i += kThreads) {
    const float x = row[i];
    float score;
    bool keep = true;
    if (greedy) {
      score = x;
    } else {
      const float s = scaled(x, T);
      const uint32_t key = okey(s);
      keep = keeps(sk, key, static_cast<uint32_t>(i)) && keeps(sp, key, static_cast<uint32_t>(i));
      const uint64_t h = mix64(seed + static_cast<uint64_t>(i + 1) * kGolden);
      const float u = (static_cast<float>(h >> 41) + 0.5f) * 1.1920928955078125e-07f;  // 2^-23; in (0, 1)
      const float uc = fminf(fmaxf(u, 1e-8f), 1.0f - 1e-8f);                             // sampler.rs:211-212
      score = keep ? s + (-logf(-logf(uc))) : -INFINITY;
    }
    if (kept_mask != nullptr) kept_mask[static_cast<int64_t>(b) * V + i] = keep ? 1 : 0;
    kept += keep ? 1 : 0;
    if (score > best || (score == best && i < bidx)) { best = score; bidx = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(kFull, best, o);
    const int oi = __shfl_xor_sync(kFull, bidx, o);
    kept += __shfl_xor_sync(kFull, kept, o);
    if (ov > best || (ov == best && oi < bidx)) { best = ov; bidx = oi; }
  }
  if (lane == 0) { sh.redf[warp] = best; sh.redi[warp] = bidx; sh.cnt[warp] = static_cast<uint32_t>(kept); }
  __syncthreads();
  if (tid == 0) {
    int total = 0;
    for (int w = 0; w < kWarps; ++w) {
      total += static_cast<int>(sh.cnt[w]);
      if (sh.redf[w] > best || (sh.redf[w] == best && sh.redi[w] < bidx)) { best = sh.redf[w]; bidx = sh.redi[w]; }
    }
    out_ids[b] = bidx == 0x7fffffff ? 0 : bidx;
    if (kept_count != nullptr) kept_count[b] = total;
  }
}

}  // namespace
}  // namespace nvllm

// Sampler::forward (sampler.rs:26-67) for a whole batch in one launch; all pointers are device pointers.
extern "C" int nvllm_sample(int64_t* out_ids, const float* logits, int32_t batch, int32_t vocab, int64_t ld,
                            const float* temperatures, const int32_t* top_k, const float* top_p, const uint64_t* row_seeds,
                            int32_t* kept_count, uint8_t* kept_mask, void* stream) {
  using namespace nvllm;
  NVLLM_CHECK_ARG(batch >= 0 && vocab > 0 && ld >= vocab, NVLLM_ERR_ARG, "sample: batch=%d vocab=%d ld=%lld", batch, vocab,
                  (long long)ld);
  if (batch == 0) return 0;
  NVLLM_CHECK_ARG(out_ids && logits && temperatures && row_seeds, NVLLM_ERR_ARG, "sample: NULL pointer");
  NVLLM_CUDA(launch_pdl(sample_kernel, dim3(batch), dim3(kThreads), 0, static_cast<cudaStream_t>(stream), out_ids, logits,
                        static_cast<int>(vocab), ld, temperatures, top_k, top_p, row_seeds, kept_count, kept_mask));
  return 0;
}
