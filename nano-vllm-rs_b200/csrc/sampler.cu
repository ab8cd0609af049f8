// sampler.cu — K12: Sampler::forward / sample_single (reference src/layers/sampler.rs:26-219) on the device.
//
// The reference samples one row at a time on the host: a blocking scalar read per parameter, two full sorts of the
// vocabulary (top-k, top-p), a softmax and an unseeded Tensor::rand.  Here one CTA owns one row of the f32 logits and
// never sorts: both filters are *selections*, done MSB-first over the order-preserving integer image of the scaled
// logits with shared-memory histograms (3 passes each over 8 + 12 + 12 key bits, 16-byte loads, warp-aggregated updates):
//   top-k  (sampler.rs:115-149)  k-th largest key by COUNT; ties at the threshold keep the lowest indices (the
//                                reference's stable sort), resolved by an ordered ballot scan only when they straddle k
//   top-p  (sampler.rs:152-191)  smallest prefix of the descending order whose probability mass reaches p, found by
//                                MASS: exp(s - max) in 0.32 fixed point, so every sum is an integer sum (order-free,
//                                deterministic); target = ceil(p * total)
//   Gumbel-max (sampler.rs:194-219) argmax(s + g), g = -log(-log(u)), first maximum wins
// Rows with temperature == 0 take the same final pass without filters or noise (greedy, sampler.rs:79-81,109-112).
// Seeded contract for u (the reference's draws are unseeded and cannot be reproduced): see oracle/layers.py.
// HBM/L2-bound integer + compare work; no tensor cores.
#include <math.h>

#include "common.cuh"

namespace nvllm {
namespace {

constexpr int kThreads = 1024;
constexpr int kWarps = kThreads / 32;
constexpr int kBins = 4096;  // three selection passes over 8 + 12 + 12 key bits (see radix_select)
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
// exp(s - m) in 0.32 fixed point (s <= m): [0, 2^32 - 1] — the maximum itself is capped one unit below 2^32 so that a
// weight fits 32 bits (two 16-bit warp reductions per histogram update)
__device__ __forceinline__ uint32_t fx_exp(float s, float m) {
  const unsigned long long w = __float2ull_rz(expf(s - m) * 4294967296.0f);
  return w > 0xffffffffull ? 0xffffffffu : static_cast<uint32_t>(w);
}

// keep(key, i) <=> key > sel.key || (key == sel.key && i <= sel.cut)
struct Sel {
  uint32_t key, cut;
};
__device__ __forceinline__ bool keeps(const Sel& s, uint32_t key, uint32_t i) {
  return key > s.key || (key == s.key && i <= s.cut);
}

struct Shared {
  unsigned long long hist[kBins];
  unsigned long long need, eq;
  float redf[32];
  int redi[32];
  uint32_t cnt[kWarps];
  int sel_bin;
  uint32_t cut;
};

// Visits the row with 16-byte loads (two in flight per thread).  f(v, i0, n): v[0..n) are elements i0 .. i0+n-1; f is
// called the SAME number of times by every thread of the CTA (n == 0 for idle slots), so it may use full-warp intrinsics.
template <typename F>
__device__ __forceinline__ void for_each4(const float* __restrict__ row, int V, F f) {
  const int tid = threadIdx.x;
  if ((reinterpret_cast<uintptr_t>(row) & 15u) == 0) {
    const int nvec = V >> 2;
    const float4* r4 = reinterpret_cast<const float4*>(row);
    for (int base = 0; base < nvec; base += 2 * kThreads) {
      const int j0 = base + tid, j1 = j0 + kThreads;
      float4 q0 = make_float4(0.f, 0.f, 0.f, 0.f), q1 = q0;
      if (j0 < nvec) q0 = __ldg(r4 + j0);
      if (j1 < nvec) q1 = __ldg(r4 + j1);
      const float v0[4] = {q0.x, q0.y, q0.z, q0.w}, v1[4] = {q1.x, q1.y, q1.z, q1.w};
      f(v0, 4 * j0, j0 < nvec ? 4 : 0);
      if (base + kThreads < nvec) f(v1, 4 * j1, j1 < nvec ? 4 : 0);  // block-uniform condition
    }
    const int t0 = nvec << 2;
    if (t0 < V) {  // V % 4 trailing elements: thread 0 takes them, everybody makes the call
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      int n = 0;
      if (tid == 0) { n = V - t0; for (int e = 0; e < n; ++e) v[e] = row[t0 + e]; }
      f(v, t0, n);
    }
  } else {
    for (int base = 0; base < V; base += kThreads) {
      const int i = base + tid;
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      if (i < V) v[0] = row[i];
      f(v, i, i < V ? 1 : 0);
    }
  }
}

// MSB-first radix selection over the keys of one row, largest first, 8 + 12 + 12 bits.  weight(key, i) is the element's
// contribution (0 = not a candidate).  Finds the key K with  sum{w : key > K} < target <= sum{w : key >= K}  and returns
// how much of the target the group key == K has to supply (`need`, >= 1) and that group's total weight (`eq`).
// MASS == true: weights are 32-bit masses and the target is ceil(p * total), fixed after the first pass; MASS == false:
// weights are 0/1.  Histogram updates are aggregated per warp (match_any on the bin: logits cluster in a handful of
// exponent bins, un-aggregated shared-memory atomics serialise on them).  row_max (optional) receives max(scaled row).
template <bool MASS, typename WF>
__device__ void radix_select(Shared& sh, const float* __restrict__ row, int V, float T, WF weight, unsigned long long target,
                             double p, uint32_t& out_key, unsigned long long& out_need, unsigned long long& out_eq,
                             float* row_max) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint32_t prefix = 0, mask = 0;
  unsigned long long rem = target;
  float mx = -INFINITY;
#pragma unroll 1
  for (int pass = 0; pass < 3; ++pass) {
    // First digit = sign + 7 exponent bits only: every element takes part in the first pass, and the warp-aggregated
    // update costs ~ the number of DISTINCT bins in a warp (a handful at this width); later passes see few candidates.
    const int shift = pass == 0 ? 24 : (pass == 1 ? 12 : 0);
    const int nbins = pass == 0 ? 256 : 4096;
    for (int b = tid; b < nbins; b += kThreads) sh.hist[b] = 0ull;
    __syncthreads();
    for_each4(row, V, [&](const float (&v)[4], int i0, int n) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        bool part = e < n;
        const float s = scaled(v[e], T);
        const uint32_t key = okey(s);
        if (pass == 0 && row_max != nullptr && part) mx = fmaxf(mx, s);
        part = part && (key & mask) == prefix;
        const uint32_t w = part ? static_cast<uint32_t>(weight(key, static_cast<uint32_t>(i0 + e))) : 0u;
        part = part && w != 0u;
        if (__any_sync(kFull, part)) {
          const uint32_t bin = part ? ((key >> shift) & static_cast<uint32_t>(nbins - 1)) : 0xffffffffu;
          const unsigned peers = __match_any_sync(kFull, bin);
          unsigned long long sum;
          if (MASS) {
            const uint32_t lo = __reduce_add_sync(peers, w & 0xffffu), hi = __reduce_add_sync(peers, w >> 16);
            sum = (static_cast<unsigned long long>(hi) << 16) + lo;
          } else {
            sum = static_cast<unsigned long long>(__popc(peers));
          }
          if (part && lane == __ffs(peers) - 1) atomicAdd(&sh.hist[bin], sum);
        }
      }
    });
    __syncthreads();
    if (warp == 0) {  // lane l owns bins nbins-1-per*l .. nbins-per*(l+1), scanned from the top
      const int per = nbins / 32, top = nbins - 1 - per * lane;
      unsigned long long s = 0ull;
      for (int j = 0; j < per; ++j) s += sh.hist[top - j];
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
        for (int j = 0; j < per; ++j) {
          const unsigned long long h = sh.hist[top - j];
          if (c + h >= rem) { sh.sel_bin = top - j; sh.need = rem - c; sh.eq = h; break; }
          c += h;
        }
      }
    }
    __syncthreads();
    prefix |= static_cast<uint32_t>(sh.sel_bin) << shift;
    mask |= static_cast<uint32_t>(nbins - 1) << shift;
    rem = sh.need;
    out_eq = sh.eq;
    __syncthreads();
  }
  out_key = prefix;
  out_need = rem;
  if (row_max != nullptr) *row_max = mx;  // per-thread partial; the caller reduces it
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
    const int k = top_k != nullptr ? top_k[b] : 0;
    const float p = top_p != nullptr ? top_p[b] : 0.0f;
    const bool do_k = k > 0 && k < V, do_p = p > 0.0f && p < 1.0f;
    float m = -INFINITY;  // max of the scaled row (it always survives top-k), needed by top-p only
    bool have_m = false;
    // ---- top-k: the k-th largest key by count (sampler.rs:115-149) ----
    if (do_k) {
      uint32_t key;
      unsigned long long need, eq;
      radix_select<false>(sh, row, V, T, [](uint32_t, uint32_t) { return 1u; }, static_cast<unsigned long long>(k), 0.0, key,
                          need, eq, do_p ? &m : nullptr);
      have_m = do_p;
      sk.key = key;
      if (need < eq)  // ties straddle k: the stable sort keeps the lowest indices
        sk.cut = ordered_cut(sh, V, [&](uint32_t i) { return okey(scaled(row[i], T)) == key; }, static_cast<uint32_t>(need));
    }
    // ---- top-p: smallest descending prefix whose softmax mass reaches p (sampler.rs:152-191) ----
    if (do_p) {
      if (!have_m)
        for_each4(row, V, [&](const float (&v)[4], int, int n) {
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (e < n) m = fmaxf(m, scaled(v[e], T));
        });
      m = block_max(sh, m);
      const Sel skc = sk;
      auto w_of = [&](uint32_t key, uint32_t i) { return keeps(skc, key, i) ? fx_exp(okey_inv(key), m) : 0u; };
      uint32_t key;
      unsigned long long need, eq;
      radix_select<true>(sh, row, V, T, w_of, 0ull, static_cast<double>(p), key, need, eq, nullptr);
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
  for_each4(row, V, [&](const float (&v)[4], int i0, int n) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      if (e < n) {
        const int i = i0 + e;
        float score;
        bool keep = true;
        if (greedy) {
          score = v[e];
        } else {
          const float s = scaled(v[e], T);
          const uint32_t key = okey(s);
          keep = keeps(sk, key, static_cast<uint32_t>(i)) && keeps(sp, key, static_cast<uint32_t>(i));
          score = -INFINITY;
          if (keep) {  // the noise is only needed for survivors
            const uint64_t h = mix64(seed + static_cast<uint64_t>(i + 1) * kGolden);
            const float u = (static_cast<float>(h >> 41) + 0.5f) * 1.1920928955078125e-07f;  // 2^-23; in (0, 1)
            const float uc = fminf(fmaxf(u, 1e-8f), 1.0f - 1e-8f);                             // sampler.rs:211-212
            score = s + (-logf(-logf(uc)));
          }
        }
        if (kept_mask != nullptr) kept_mask[static_cast<int64_t>(b) * V + i] = keep ? 1 : 0;
        kept += keep ? 1 : 0;
        if (score > best || (score == best && i < bidx)) { best = score; bidx = i; }
      }
    }
  });
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
