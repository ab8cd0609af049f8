// model_runner.cu — ModelRunner (reference src/engine/model_runner.rs:19-448) + Qwen3Model
// (src/models/qwen3.rs:127-591) as a C++ host runtime over the kernels of nvllm.h.
//
// What the reference does on the host per step — build tensors, set a global Context, walk the layer objects, one
// blocking scalar read per token / block / row — becomes: one packed H2D copy of the step metadata, ~10 kernel
// launches per layer on one stream (captured once per batch-size bucket as a CUDA graph for decode, the stub of
// model_runner.rs:303-361 made real), one 8-byte-per-sequence D2H copy of the greedy tokens.
// Weights are this rank's tensor-parallel shard (qwen3.rs:157-183,273-292 arithmetic): q/k/v rows packed as
// [q | k | v] (linear.rs:331-340), gate/up rows packed as [gate | up] (linear.rs:421-433), o_proj / down_proj
// column slices followed by an all-reduce (the linear.rs:236 TODO), embedding replicated, LM head vocab-sharded
// with a (max, id) all-gather (embed_head.rs:321-336 TODO).
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "../common.cuh"
#include "scheduler.hpp"

namespace nvllm {
int comm_world(const nvllm_comm* c);
int comm_rank(const nvllm_comm* c);

namespace {

#define RUN(expr)                   \
  do {                              \
    int _rc = (expr);               \
    if (_rc != 0) return _rc;       \
  } while (0)

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
  x ^= x >> 27; x *= 0x94D049BB133111EBull;
  x ^= x >> 31;
  return x;
}
// Shard [rows, cols] (row pitch ld) of the logical tensor whose element (r, c) has flat index r*full_cols + c:
// same generator as nvllm_init_weight / oracle/weights.py, so every rank holds slices of identical tensors.
__global__ void init_shard_kernel(bf16* __restrict__ w, int64_t rows, int64_t cols, int64_t ld, int64_t full_cols,
                                  int64_t row0, int64_t col0, uint64_t key, float scale) {
  const int64_t total = rows * cols;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t r = i / cols, c = i - r * cols;
    const uint64_t idx = static_cast<uint64_t>((row0 + r) * full_cols + col0 + c);
    const uint64_t h = mix64(key + (idx + 1) * 0x9E3779B97F4A7C15ull);
    const int s = static_cast<int>((h & 0xffff) + ((h >> 16) & 0xffff) + ((h >> 32) & 0xffff) + (h >> 48)) - 131070;
    w[r * ld + c] = __float2bfloat16_rn(__fmul_rn(static_cast<float>(s), scale));
  }
}
__global__ void fill_bf16_kernel(bf16* __restrict__ w, int64_t n, float v) {
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x)
    w[i] = __float2bfloat16_rn(v);
}

struct LayerWeights {
  bf16 *input_ln = nullptr, *qkv = nullptr, *q_norm = nullptr, *k_norm = nullptr, *o = nullptr,
       *post_ln = nullptr, *gate_up = nullptr, *down = nullptr;
};

struct Graph {
  cudaGraphExec_t exec = nullptr;
  int uses = 0;
  int64_t nodes = 0;
};

}  // namespace
}  // namespace nvllm

using namespace nvllm;

struct nvllm_runner {
  nvllm_config cfg{};
  nvllm_comm* comm = nullptr;
  cudaStream_t stream = nullptr;
  // per-rank dimensions
  int tp = 1, rank = 0, H = 0, Hkv = 0, D = 0, hidden = 0, inter = 0, vocab = 0, vshard = 0, layers = 0, qkv_dim = 0;
  int max_tokens = 0, max_seqs = 0, table_cap = 0, num_blocks = 0;
  nvllm_kv_layout layout{};
  // weights
  bf16 *embed = nullptr, *final_norm = nullptr, *lm_head = nullptr;
  std::vector<LayerWeights> lw;
  float* rope = nullptr;
  std::vector<bf16*> kc, vc;
  // activations
  bf16 *h = nullptr, *normed = nullptr, *qkv = nullptr, *attn = nullptr, *proj = nullptr, *gate_up = nullptr,
       *act = nullptr, *last_hidden = nullptr;
  float* logits = nullptr;
  float *gathered_logits = nullptr, *full_logits = nullptr;  // tp > 1: [tp, seqs, V/tp] all-gathered, then [seqs, V] (sampling)
  float *local_max = nullptr, *all_max = nullptr;
  int64_t *local_ids = nullptr, *all_ids = nullptr, *tokens = nullptr;
  void* decode_ws = nullptr;
  size_t decode_ws_bytes = 0;
  // step metadata: one device block + one pinned staging block
  uint8_t *d_meta = nullptr, *h_meta = nullptr;
  size_t meta_bytes = 0;
  int64_t* h_tokens = nullptr;  // pinned
  // sampler parameters of the current step (only used when some sequence has temperature != 0):
  // [temperature f32 | top_k i32 | top_p f32 | row_seed u64] x max_seqs, pinned staging + device copy
  uint8_t *d_samp = nullptr, *h_samp = nullptr;
  // pointers into d_meta for the current step
  int64_t *d_ids = nullptr, *d_pos = nullptr;
  int32_t *d_slots = nullptr, *d_cu_q = nullptr, *d_cu_k = nullptr, *d_ctx = nullptr, *d_tables = nullptr;
  std::map<int, Graph> graphs;
  std::vector<void*> allocs;
  int64_t launches = 0, last_launches = 0;
  // last decode batch (for nvllm_runner_time_decode_attention)
  int last_decode_b = 0, last_decode_splits = 1;
  std::vector<int32_t> last_ctx;
  bool weights_ready = false;
  // load_weight bookkeeping: every tensor the architecture needs must have arrived before the first execute
  // (a partially loaded checkpoint would otherwise run on uninitialised layer weights)
  std::vector<uint32_t> loaded;  // per layer: bit mask of the 11 per-layer tensors; last entry: embed | lm_head | norm
  // per-launch profile (nvllm_runner_profile): eager steps record a CUDA event after every launch; an event between two
  // launches also breaks their programmatic overlap, so the differences are serialised per-kernel times — the
  // multi-rank-safe counterpart of an ncu launch list (kernels that wait for peers cannot run under ncu's serialisation)
  bool profiling = false;
  std::vector<std::pair<const char*, cudaEvent_t>> prof_marks;
  std::vector<cudaEvent_t> prof_pool;
  std::map<std::string, std::pair<double, int64_t>> prof_acc;  // name -> (total us, launches)
  bool fuse_rope = false; // NVLLM_FUSE_ROPE=1: qkv GEMM epilogue does q/k norm + RoPE + KV store for decode-sized batches
  bool fuse_decode = false; // NVLLM_FUSE_DECODE=1 (opt-in, measured slower): q/k-norm + RoPE + KV store inside the paged attention kernel
  bool fused_tp = false;  // h / normed / proj live in NVLink-mapped symmetric memory; all-reduce fused with add + RMSNorm

  ~nvllm_runner() {
    for (auto& g : graphs)
      if (g.second.exec) cudaGraphExecDestroy(g.second.exec);
    for (void* p : allocs) cudaFree(p);
    for (auto& m : prof_marks) cudaEventDestroy(m.second);
    for (cudaEvent_t e : prof_pool) cudaEventDestroy(e);
    if (h_meta) cudaFreeHost(h_meta);
    if (h_tokens) cudaFreeHost(h_tokens);
    if (h_samp) cudaFreeHost(h_samp);
    if (stream) cudaStreamDestroy(stream);
  }

  template <typename T>
  int dalloc(T** p, size_t count) {
    void* q = nullptr;
    const size_t bytes = (count * sizeof(T) + 255) & ~static_cast<size_t>(255);
    cudaError_t e = cudaMalloc(&q, bytes ? bytes : 256);
    if (e != cudaSuccess) {
      set_error("cudaMalloc of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
      return static_cast<int>(e);
    }
    allocs.push_back(q);
    *p = static_cast<T*>(q);
    return 0;
  }

  int create(const nvllm_config& c, nvllm_comm* cm);
  int init_weights(uint64_t seed, float std);
  int load_weight(const char* name, const void* h_bf16, int64_t rows, int64_t cols);
  int upload(const nvllm_batch& b, int padded);
  int forward(int n_tokens, int n_seqs, bool is_prefill, bool paged_prefill, const nvllm_batch& b);
  int execute_async(const nvllm_batch& b);
  int fetch(int64_t* out, int n);
  int sample(const nvllm_batch& b);
  void mark(const char* name) {  // after a launch: no-op unless profiling an eager (non-captured) step
    if (!profiling) return;
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(stream, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) return;
    cudaEvent_t e = nullptr;
    if (!prof_pool.empty()) { e = prof_pool.back(); prof_pool.pop_back(); }
    else if (cudaEventCreate(&e) != cudaSuccess) return;
    cudaEventRecord(e, stream);
    prof_marks.emplace_back(name, e);
  }
  void prof_collect() {  // stream must be idle
    for (size_t i = 1; i < prof_marks.size(); ++i) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, prof_marks[i - 1].second, prof_marks[i].second) == cudaSuccess) {
        auto& a = prof_acc[prof_marks[i].first];
        a.first += ms * 1e3; a.second += 1;
      }
    }
    for (auto& m : prof_marks) prof_pool.push_back(m.second);
    prof_marks.clear();
  }
};

// ------------------------------------------------------------------------------------------------ create
static int bucket_for(int b, int max_seqs) {
  int p = 1;
  while (p < b) p <<= 1;
  return p > max_seqs ? max_seqs : p;
}

int nvllm_runner::create(const nvllm_config& c, nvllm_comm* cm) {
  RUN(nvllm_config_validate(&c));
  NVLLM_CHECK_ARG(nvllm_device_ok() == 1, NVLLM_ERR_STATE, "runner_create: no sm_100 CUDA device (there is no CPU path)");
  cfg = c;
  comm = cm;
  tp = c.tensor_parallel_size;
  rank = c.tp_rank;
  NVLLM_CHECK_ARG((tp == 1) == (cm == nullptr) || tp == 1, NVLLM_ERR_ARG, "runner_create: tp=%d needs a communicator", tp);
  NVLLM_CHECK_ARG(cm == nullptr || (comm_world(cm) == tp && comm_rank(cm) == rank), NVLLM_ERR_ARG,
                  "runner_create: communicator rank/world do not match the config");
  H = c.num_attention_heads / tp;            // qwen3.rs:158
  Hkv = c.num_key_value_heads / tp;          // qwen3.rs:159, model_runner.rs:372
  D = c.head_dim;
  hidden = c.hidden_size;
  inter = c.intermediate_size / tp;          // linear.rs:401
  vocab = c.vocab_size;
  vshard = vocab / tp;
  layers = c.num_hidden_layers;
  qkv_dim = (H + 2 * Hkv) * D;               // linear.rs:303-307
  max_seqs = c.max_num_seqs;
  max_tokens = c.max_num_batched_tokens > max_seqs ? c.max_num_batched_tokens : max_seqs;
  num_blocks = c.num_kvcache_blocks > 0 ? c.num_kvcache_blocks : 1000;  // model_runner.rs:84
  const int bs = c.kvcache_block_size;
  table_cap = (c.max_model_len + bs - 1) / bs;
  layout.block_size = bs; layout.num_kv_heads = Hkv; layout.head_dim = D; layout.reserved = 0;
  layout.block_stride = static_cast<int64_t>(bs) * Hkv * D;
  if (c.kv_layout == 1) { layout.token_stride = static_cast<int64_t>(Hkv) * D; layout.head_stride = D; }   // reference
  else { layout.token_stride = D; layout.head_stride = static_cast<int64_t>(bs) * D; }                     // head-major

  fuse_rope = knobs().fuse_rope != 0;
  fuse_decode = knobs().fuse_decode != 0 && c.head_dim == 128;
  NVLLM_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
  // weights
  RUN(dalloc(&embed, static_cast<size_t>(vocab) * hidden));
  RUN(dalloc(&final_norm, hidden));
  if (c.tie_word_embeddings) lm_head = embed + static_cast<size_t>(rank) * vshard * hidden;
  else RUN(dalloc(&lm_head, static_cast<size_t>(vshard) * hidden));
  lw.resize(layers);
  for (auto& w : lw) {
    RUN(dalloc(&w.input_ln, hidden)); RUN(dalloc(&w.post_ln, hidden));
    RUN(dalloc(&w.q_norm, D)); RUN(dalloc(&w.k_norm, D));
    RUN(dalloc(&w.qkv, static_cast<size_t>(qkv_dim) * hidden));
    RUN(dalloc(&w.o, static_cast<size_t>(hidden) * H * D));
    RUN(dalloc(&w.gate_up, static_cast<size_t>(2) * inter * hidden));
    RUN(dalloc(&w.down, static_cast<size_t>(hidden) * inter));
  }
  RUN(dalloc(&rope, static_cast<size_t>(c.max_position_embeddings) * D));
  RUN(nvllm_rope_table(rope, c.max_position_embeddings, D, c.rope_theta, stream));
  // KV cache: per layer K and V, zero-initialised like the reference (model_runner.rs:379-389)
  kc.resize(layers); vc.resize(layers);
  const size_t cache_elems = static_cast<size_t>(num_blocks) * layout.block_stride;
  for (int l = 0; l < layers; ++l) {
    RUN(dalloc(&kc[l], cache_elems)); RUN(dalloc(&vc[l], cache_elems));
    NVLLM_CUDA(cudaMemsetAsync(kc[l], 0, cache_elems * 2, stream));
    NVLLM_CUDA(cudaMemsetAsync(vc[l], 0, cache_elems * 2, stream));
  }
  // activations
  const size_t T = max_tokens;
  if (tp > 1) {
    if (knobs().fused_tp != 0) {  // A/B knob NVLLM_FUSED_TP=0: ncclAllReduce + separate add_rmsnorm
      const size_t one = (T * hidden * 2 + 255) & ~static_cast<size_t>(255);
      void* base = nullptr;
      if (nvllm_comm_symm_alloc(comm, 3 * one, &base) == 0) {
        h = static_cast<bf16*>(base);
        normed = reinterpret_cast<bf16*>(static_cast<uint8_t*>(base) + one);
        proj = reinterpret_cast<bf16*>(static_cast<uint8_t*>(base) + 2 * one);
        fused_tp = true;
      }
    }
  }
  if (!fused_tp) { RUN(dalloc(&h, T * hidden)); RUN(dalloc(&normed, T * hidden)); RUN(dalloc(&proj, T * hidden)); }
  RUN(dalloc(&qkv, T * qkv_dim)); RUN(dalloc(&attn, T * H * D));
  RUN(dalloc(&gate_up, T * 2 * inter)); RUN(dalloc(&act, T * inter));
  RUN(dalloc(&last_hidden, static_cast<size_t>(max_seqs) * hidden));
  RUN(dalloc(&logits, static_cast<size_t>(max_seqs) * vshard));
  // gathered_logits / full_logits (tp > 1, 2 x max_seqs x vocab f32 = 2 x 311 MB at 512 x 151936) are taken on the first
  // SAMPLED batch (sample() below): greedy-only serving, the benchmarked TP configuration included, never pays for them.
  RUN(dalloc(&tokens, max_seqs)); RUN(dalloc(&local_ids, max_seqs)); RUN(dalloc(&local_max, max_seqs));
  RUN(dalloc(&all_ids, static_cast<size_t>(max_seqs) * tp)); RUN(dalloc(&all_max, static_cast<size_t>(max_seqs) * tp));
  // split-K workspace: the largest any decode bucket can ask for
  for (int b = 1;; b <<= 1) {
    const int bb = b > max_seqs ? max_seqs : b;
    const int s = nvllm_paged_decode_choose_splits(bb, H, Hkv, table_cap * bs);
    const size_t need = nvllm_paged_decode_workspace_bytes(bb, H, D, s);
    if (need > decode_ws_bytes) decode_ws_bytes = need;
    if (b >= max_seqs) break;
  }
  if (decode_ws_bytes) { uint8_t* p = nullptr; RUN(dalloc(&p, decode_ws_bytes)); decode_ws = p; }
  // metadata
  const size_t S = max_seqs;
  const size_t prefill_meta = 20 * T + 8 * (S + 1) + 4 * S * table_cap;
  const size_t decode_meta = S * (24 + 4 * static_cast<size_t>(table_cap));
  meta_bytes = (prefill_meta > decode_meta ? prefill_meta : decode_meta) + 64;
  RUN(dalloc(&d_meta, meta_bytes));
  NVLLM_CUDA(cudaMallocHost(reinterpret_cast<void**>(&h_meta), meta_bytes));
  NVLLM_CUDA(cudaMallocHost(reinterpret_cast<void**>(&h_tokens), sizeof(int64_t) * S));
  RUN(dalloc(&d_samp, 24 * S));
  NVLLM_CUDA(cudaMallocHost(reinterpret_cast<void**>(&h_samp), 24 * S));
  NVLLM_CUDA(cudaStreamSynchronize(stream));
  return 0;
}

// ------------------------------------------------------------------------------------------------ weights
// Logical tensor ids of the seeded init (mirrored by oracle/qwen3.py): 1 embed, 2 lm_head, 16*(l+1) + {0 q, 1 k,
// 2 v, 3 o, 4 gate, 5 up, 6 down}.  Norm weights are 1.
int nvllm_runner::init_weights(uint64_t seed, float std) {
  const float scale = std * 2.6428998e-05f;
  auto shard = [&](bf16* dst, int64_t rows, int64_t cols, int64_t ld, int64_t full_cols, int64_t row0, int64_t col0,
                   uint64_t id) {
    const int64_t total = rows * cols;
    int64_t blocks = (total + 255) / 256;
    const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
    if (blocks > cap) blocks = cap;
    init_shard_kernel<<<static_cast<unsigned>(blocks), 256, 0, stream>>>(dst, rows, cols, ld, full_cols, row0, col0,
                                                                        seed ^ (id << 40), scale);
  };
  auto ones = [&](bf16* dst, int64_t n) { fill_bf16_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, stream>>>(dst, n, 1.0f); };
  const int Hf = cfg.num_attention_heads, Hkvf = cfg.num_key_value_heads, If = cfg.intermediate_size;
  shard(embed, vocab, hidden, hidden, hidden, 0, 0, 1);
  if (!cfg.tie_word_embeddings) shard(lm_head, vshard, hidden, hidden, hidden, static_cast<int64_t>(rank) * vshard, 0, 2);
  ones(final_norm, hidden);
  for (int l = 0; l < layers; ++l) {
    LayerWeights& w = lw[l];
    const uint64_t base = 16ull * (l + 1);
    ones(w.input_ln, hidden); ones(w.post_ln, hidden); ones(w.q_norm, D); ones(w.k_norm, D);
    // qkv_proj rows: [q heads of this rank | k heads | v heads] (linear.rs:303-307,331-340)
    shard(w.qkv, static_cast<int64_t>(H) * D, hidden, hidden, hidden, static_cast<int64_t>(rank) * H * D, 0, base + 0);
    shard(w.qkv + static_cast<int64_t>(H) * D * hidden, static_cast<int64_t>(Hkv) * D, hidden, hidden, hidden,
          static_cast<int64_t>(rank) * Hkv * D, 0, base + 1);
    shard(w.qkv + static_cast<int64_t>(H + Hkv) * D * hidden, static_cast<int64_t>(Hkv) * D, hidden, hidden, hidden,
          static_cast<int64_t>(rank) * Hkv * D, 0, base + 2);
    // o_proj [hidden, Hf*D]: column slice (linear.rs:202,249-267)
    shard(w.o, hidden, static_cast<int64_t>(H) * D, static_cast<int64_t>(H) * D, static_cast<int64_t>(Hf) * D, 0,
          static_cast<int64_t>(rank) * H * D, base + 3);
    // gate_up rows: [gate shard | up shard] (linear.rs:401,421-433)
    shard(w.gate_up, inter, hidden, hidden, hidden, static_cast<int64_t>(rank) * inter, 0, base + 4);
    shard(w.gate_up + static_cast<int64_t>(inter) * hidden, inter, hidden, hidden, hidden,
          static_cast<int64_t>(rank) * inter, 0, base + 5);
    // down_proj [hidden, If]: column slice
    shard(w.down, hidden, inter, inter, If, 0, static_cast<int64_t>(rank) * inter, base + 6);
  }
  (void)Hkvf;
  NVLLM_LAUNCH_CHECK("init_weights");
  NVLLM_CUDA(cudaStreamSynchronize(stream));
  weights_ready = true;
  return 0;
}

// HF tensor names -> this rank's slice (utils/loader.rs:175-198 packed-module mapping; linear.rs:154-172,249-267).
int nvllm_runner::load_weight(const char* name_c, const void* src, int64_t rows, int64_t cols) {
  NVLLM_CHECK_ARG(name_c && src, NVLLM_ERR_ARG, "load_weight: NULL");
  const std::string name(name_c);
  const int Hf = cfg.num_attention_heads, Hkvf = cfg.num_key_value_heads, If = cfg.intermediate_size;
  // copies rows [r0, r0+nr) x cols [c0, c0+nc) of the host tensor (pitch `cols`) into dst (pitch nc)
  auto slice = [&](bf16* dst, int64_t exp_rows, int64_t exp_cols, int64_t r0, int64_t nr, int64_t c0, int64_t nc) -> int {
    NVLLM_CHECK_ARG(rows == exp_rows && cols == exp_cols, NVLLM_ERR_ARG, "load_weight(%s): expected [%lld, %lld], got [%lld, %lld]",
                    name_c, (long long)exp_rows, (long long)exp_cols, (long long)rows, (long long)cols);
    const uint8_t* s = static_cast<const uint8_t*>(src) + (r0 * cols + c0) * 2;
    NVLLM_CUDA(cudaMemcpy2DAsync(dst, nc * 2, s, cols * 2, nc * 2, nr, cudaMemcpyHostToDevice, stream));
    NVLLM_CUDA(cudaStreamSynchronize(stream));
    return 0;
  };
  if (loaded.empty()) loaded.assign(layers + 1, 0u);
  // marks tensor `bit` of group `grp` as present and re-evaluates completeness
  auto done = [&](int rc, int grp, int bit) -> int {
    if (rc != 0) return rc;
    loaded[grp] |= 1u << bit;
    const uint32_t top = cfg.tie_word_embeddings ? 0x5u : 0x7u;  // embed | [lm_head] | norm
    bool all = (loaded[layers] & top) == top;
    const uint32_t need = cfg.qk_norm ? 0x7ffu : 0x7f3u;
    for (int l = 0; l < layers && all; ++l) all = (loaded[l] & need) == need;
    weights_ready = all;
    return 0;
  };
  if (name == "model.embed_tokens.weight") return done(slice(embed, vocab, hidden, 0, vocab, 0, hidden), layers, 0);
  if (name == "lm_head.weight") {
    // tied checkpoints often ship lm_head.weight as a copy of the embedding: the engine reads the head out of the
    // embedding table in that case (create()), so the tensor is accepted and ignored instead of aborting the load
    if (cfg.tie_word_embeddings) return 0;
    return done(slice(lm_head, vocab, hidden, static_cast<int64_t>(rank) * vshard, vshard, 0, hidden), layers, 1);
  }
  if (name == "model.norm.weight") return done(slice(final_norm, 1, hidden, 0, 1, 0, hidden), layers, 2);
  const std::string pre = "model.layers.";
  NVLLM_CHECK_ARG(name.compare(0, pre.size(), pre) == 0, NVLLM_ERR_ARG, "load_weight: unknown tensor '%s'", name_c);
  const size_t dot = name.find('.', pre.size());
  NVLLM_CHECK_ARG(dot != std::string::npos, NVLLM_ERR_ARG, "load_weight: bad name '%s'", name_c);
  const int l = atoi(name.substr(pre.size(), dot - pre.size()).c_str());
  NVLLM_CHECK_ARG(l >= 0 && l < layers, NVLLM_ERR_ARG, "load_weight: layer %d out of range", l);
  const std::string rest = name.substr(dot + 1);
  LayerWeights& w = lw[l];
  const int64_t HD = static_cast<int64_t>(H) * D, KD = static_cast<int64_t>(Hkv) * D;
  if (rest == "input_layernorm.weight") return done(slice(w.input_ln, 1, hidden, 0, 1, 0, hidden), l, 0);
  if (rest == "post_attention_layernorm.weight") return done(slice(w.post_ln, 1, hidden, 0, 1, 0, hidden), l, 1);
  if (rest == "self_attn.q_norm.weight") return done(slice(w.q_norm, 1, D, 0, 1, 0, D), l, 2);
  if (rest == "self_attn.k_norm.weight") return done(slice(w.k_norm, 1, D, 0, 1, 0, D), l, 3);
  if (rest == "self_attn.q_proj.weight") return done(slice(w.qkv, static_cast<int64_t>(Hf) * D, hidden, rank * HD, HD, 0, hidden), l, 4);
  if (rest == "self_attn.k_proj.weight") return done(slice(w.qkv + HD * hidden, static_cast<int64_t>(Hkvf) * D, hidden, rank * KD, KD, 0, hidden), l, 5);
  if (rest == "self_attn.v_proj.weight") return done(slice(w.qkv + (HD + KD) * hidden, static_cast<int64_t>(Hkvf) * D, hidden, rank * KD, KD, 0, hidden), l, 6);
  if (rest == "self_attn.o_proj.weight") return done(slice(w.o, hidden, static_cast<int64_t>(Hf) * D, 0, hidden, rank * HD, HD), l, 7);
  if (rest == "mlp.gate_proj.weight") return done(slice(w.gate_up, If, hidden, static_cast<int64_t>(rank) * inter, inter, 0, hidden), l, 8);
  if (rest == "mlp.up_proj.weight") return done(slice(w.gate_up + static_cast<int64_t>(inter) * hidden, If, hidden, static_cast<int64_t>(rank) * inter, inter, 0, hidden), l, 9);
  if (rest == "mlp.down_proj.weight") return done(slice(w.down, hidden, If, 0, hidden, static_cast<int64_t>(rank) * inter, inter), l, 10);
  NVLLM_CHECK_ARG(false, NVLLM_ERR_ARG, "load_weight: unknown tensor '%s'", name_c);
}

// ------------------------------------------------------------------------------------------------ step metadata
// Packs the batch into the pinned block and issues ONE H2D copy.  Decode rows are padded to the graph bucket with
// (id 0, pos 0, slot -1, ctx 0, table -1): slot -1 skips the KV store, ctx 0 makes attention write zeros.
int nvllm_runner::upload(const nvllm_batch& b, int padded) {
  const size_t cap = table_cap;
  size_t used = 0;
  if (b.is_prefill) {
    const size_t N = b.num_tokens, S = b.num_seqs;
    NVLLM_CHECK_ARG(N <= static_cast<size_t>(max_tokens) && S <= static_cast<size_t>(max_seqs), NVLLM_ERR_ARG,
                    "execute: batch of %zu tokens / %zu seqs exceeds the runner's capacity", N, S);
    size_t o = 0;
    auto put = [&](const void* src, size_t bytes) { memcpy(h_meta + o, src, bytes); size_t at = o; o += bytes; return at; };
    d_ids = reinterpret_cast<int64_t*>(d_meta + put(b.input_ids, 8 * N));
    d_pos = reinterpret_cast<int64_t*>(d_meta + put(b.positions, 8 * N));
    d_slots = reinterpret_cast<int32_t*>(d_meta + put(b.slot_mapping, 4 * N));
    d_cu_q = reinterpret_cast<int32_t*>(d_meta + put(b.cu_seqlens_q, 4 * (S + 1)));
    d_cu_k = reinterpret_cast<int32_t*>(d_meta + put(b.cu_seqlens_k, 4 * (S + 1)));
    d_ctx = nullptr;
    d_tables = nullptr;
    if (b.has_block_tables) {
      NVLLM_CHECK_ARG(static_cast<size_t>(b.max_blocks) <= cap, NVLLM_ERR_ARG, "execute: block table wider than max_model_len allows");
      d_tables = reinterpret_cast<int32_t*>(d_meta + o);
      int32_t* t = reinterpret_cast<int32_t*>(h_meta + o);
      for (size_t i = 0; i < S; ++i)
        for (size_t j = 0; j < cap; ++j) t[i * cap + j] = j < static_cast<size_t>(b.max_blocks) ? b.block_tables[i * b.max_blocks + j] : -1;
      o += 4 * S * cap;
    }
    used = o;
  } else {
    const size_t B = b.num_seqs, P = padded;
    NVLLM_CHECK_ARG(P <= static_cast<size_t>(max_seqs) && static_cast<size_t>(b.max_blocks) <= cap, NVLLM_ERR_ARG,
                    "execute: decode batch %zu (padded %zu) / table width %d exceed capacity", B, P, b.max_blocks);
    int64_t* ids = reinterpret_cast<int64_t*>(h_meta);
    int64_t* pos = reinterpret_cast<int64_t*>(h_meta + 8 * P);
    int32_t* slots = reinterpret_cast<int32_t*>(h_meta + 16 * P);
    int32_t* ctx = reinterpret_cast<int32_t*>(h_meta + 20 * P);
    int32_t* tab = reinterpret_cast<int32_t*>(h_meta + 24 * P);
    for (size_t i = 0; i < P; ++i) {
      const bool real = i < B;
      ids[i] = real ? b.input_ids[i] : 0;
      pos[i] = real ? b.positions[i] : 0;
      slots[i] = real ? b.slot_mapping[i] : -1;
      ctx[i] = real ? b.context_lens[i] : 0;
      for (size_t j = 0; j < cap; ++j)
        tab[i * cap + j] = (real && j < static_cast<size_t>(b.max_blocks)) ? b.block_tables[i * b.max_blocks + j] : -1;
    }
    d_ids = reinterpret_cast<int64_t*>(d_meta);
    d_pos = reinterpret_cast<int64_t*>(d_meta + 8 * P);
    d_slots = reinterpret_cast<int32_t*>(d_meta + 16 * P);
    d_ctx = reinterpret_cast<int32_t*>(d_meta + 20 * P);
    d_tables = reinterpret_cast<int32_t*>(d_meta + 24 * P);
    d_cu_q = d_cu_k = nullptr;
    used = P * (24 + 4 * cap);
    last_decode_b = static_cast<int>(P);
    last_ctx.assign(ctx, ctx + P);
  }
  NVLLM_CUDA(cudaMemcpyAsync(d_meta, h_meta, used, cudaMemcpyHostToDevice, stream));
  return 0;
}

// ------------------------------------------------------------------------------------------------ forward
// Qwen3Model::forward (qwen3.rs:487-505) -> Qwen3Layer::forward (:372-392) -> Qwen3Attention::forward (:208-240) /
// Qwen3MLP::forward (:305-314) -> ParallelLMHead::forward (embed_head.rs:250-269) -> greedy sample (sampler.rs:109-112).
int nvllm_runner::forward(int N, int S, bool is_prefill, bool paged_prefill, const nvllm_batch& b) {
  void* st = stream;
  const float eps = cfg.rms_norm_eps;
  const float scale = 1.0f / sqrtf(static_cast<float>(D));  // attention.rs:45
  const int HD = H * D;
  launches = 0;
  mark("(step start)");
  auto gemm = [&](void* y, const void* x, const void* w, int n, int k, int64_t ldy, int out_dt, const char* what) {
    ++launches;
    const int rc = nvllm_gemm_bf16(y, x, w, N, n, k, k, k, ldy, out_dt, nullptr, 0, st);
    mark(what);
    return rc;
  };
  // h += all-reduce(proj); normed = rmsnorm(h) * w — the reference's RowParallelLinear all-reduce TODO (linear.rs:236)
  // followed by the residual add + norm: one fused kernel over NVLink peer memory, or NCCL + add_rmsnorm.
  // Decode-sized batches under TP: o_proj / down_proj push the rows another rank reduces straight from their epilogue
  // into that rank's stage buffer (nvllm_gemm_bf16_push), and the fused all-reduce skips its own push phase.
  const bool push = tp > 1 && fused_tp && knobs().gemm_push && N <= 512 && nvllm_comm_ll_applies(comm, N, hidden);
  auto row_gemm = [&](const void* x, const void* w, int k, const char* what) -> int {
    if (!push) return gemm(proj, x, w, hidden, k, hidden, NVLLM_BF16, what);
    ++launches;
    const int rc = nvllm_gemm_bf16_push(comm, proj, x, w, N, hidden, k, k, k, hidden, st);
    mark(what);
    return rc;
  };
  auto add_norm = [&](const bf16* w) -> int {
    if (tp > 1 && fused_tp) {
      ++launches;
      const int rc = push ? nvllm_allreduce_add_rmsnorm_pushed(comm, normed, proj, h, w, N, hidden, eps, NVLLM_BF16, st)
                          : nvllm_allreduce_add_rmsnorm(comm, normed, proj, h, w, N, hidden, eps, NVLLM_BF16, st);
      mark("allreduce_add_rmsnorm (fused TP epilogue)");
      return rc;
    }
    if (tp > 1) { RUN(nvllm_allreduce(comm, proj, static_cast<size_t>(N) * hidden, NVLLM_BF16, st)); ++launches; mark("ncclAllReduce"); }
    ++launches;
    const int rc = nvllm_add_rmsnorm(normed, proj, h, w, N, hidden, eps, NVLLM_BF16, st);
    mark("add_rmsnorm");
    return rc;
  };
  RUN(nvllm_embedding_gather(h, embed, d_ids, N, hidden, vocab, NVLLM_BF16, st)); ++launches; mark("embedding_gather");
  const int splits = is_prefill ? 1 : nvllm_paged_decode_choose_splits(N, H, Hkv, table_cap * layout.block_size);
  if (!is_prefill) last_decode_splits = splits;
  for (int l = 0; l < layers; ++l) {
    const LayerWeights& w = lw[l];
    // input_layernorm; for l > 0 fused with the residual add of the previous MLP output (qwen3.rs:389 -> :378)
    // (and, under TP, with the all-reduce of that output: see add_norm below)
    if (l == 0) { RUN(nvllm_add_rmsnorm(normed, h, nullptr, w.input_ln, N, hidden, eps, NVLLM_BF16, st)); ++launches; mark("add_rmsnorm"); }
    else RUN(add_norm(w.input_ln));
    const bf16* qnw = cfg.qk_norm ? w.q_norm : nullptr;
    const bf16* knw = cfg.qk_norm ? w.k_norm : nullptr;
    bool roped = false;  // q/k-norm + RoPE + KV store already applied to the qkv buffer / cache
    if (fuse_rope && D == 128 && N <= 256) {
      // decode-sized batches: qkv GEMM with q/k norm + RoPE + KV-cache store in its epilogue (one head per 128-wide tile)
      ++launches;
      RUN(nvllm_gemm_qkv_rope_store(qkv, normed, w.qkv, N, hidden, hidden, hidden, d_pos, rope, qnw, knw, eps, H, Hkv, D,
                                    cfg.max_position_embeddings, kc[l], vc[l], d_slots, &layout, st));
      mark("qkv_proj GEMM + rope/store epilogue");
      roped = true;
    } else {
      RUN(gemm(qkv, normed, w.qkv, qkv_dim, hidden, qkv_dim, NVLLM_BF16, "qkv_proj GEMM"));
      if (is_prefill || !fuse_decode) {
        // q/k norm + RoPE in place, K/V of this step into the cache (store before attention: note N2)
        RUN(nvllm_qknorm_rope_store(qkv, d_pos, rope, qnw, knw, eps, N, H, Hkv, D, cfg.max_position_embeddings, kc[l], vc[l],
                                    d_slots, &layout, NVLLM_BF16, st)); ++launches;
        mark("qknorm_rope_store");
        roped = true;
      }
    }
    if (!is_prefill) {
      if (!roped) {  // one launch: q/k-norm + RoPE + KV store + split-K paged attention (qwen3.rs:219-233)
        RUN(nvllm_paged_decode_attention_fused(attn, qkv, d_pos, rope, qnw, knw, eps, cfg.max_position_embeddings, kc[l], vc[l],
                                               d_slots, d_tables, d_ctx, N, H, table_cap, qkv_dim, HD, &layout, scale,
                                               decode_ws, decode_ws_bytes, splits, NVLLM_BF16, st));
      } else {
        RUN(nvllm_paged_decode_attention(attn, qkv, kc[l], vc[l], d_tables, d_ctx, N, H, table_cap, qkv_dim, HD, &layout,
                                         scale, decode_ws, decode_ws_bytes, splits, NVLLM_BF16, st));
      }
      launches += splits > 1 ? 2 : 1;
      mark(roped ? "paged_decode_attention" : "paged_decode_attention (+ q/k-norm, RoPE, KV store)");
    } else {
      RUN(nvllm_varlen_prefill_attention(attn, qkv, qkv + HD, qkv + HD + Hkv * D, d_cu_q, d_cu_k, S, N, N, b.max_seqlen_q,
                                         b.max_seqlen_k, H, Hkv, D, qkv_dim, qkv_dim, HD, paged_prefill ? kc[l] : nullptr,
                                         paged_prefill ? vc[l] : nullptr, paged_prefill ? d_tables : nullptr, table_cap,
                                         &layout, scale, NVLLM_BF16, st)); ++launches;
      mark(paged_prefill ? "prefill_attention (paged prefix)" : "prefill_attention");
    }
    RUN(row_gemm(attn, w.o, HD, "o_proj GEMM"));
    RUN(add_norm(w.post_ln));                                                                                          // qwen3.rs:382-385
    // gate_up GEMM with the SiLU*mul in its epilogue (MergedColumnParallelLinear + SiluAndMul, qwen3.rs:307-310)
    RUN(nvllm_gemm_silu_mul_bf16(act, normed, w.gate_up, N, inter, hidden, hidden, hidden, inter, gate_up,
                                 static_cast<size_t>(max_tokens) * 2 * inter * 2, st));
    launches += (inter % 128 == 0) ? 1 : 2;
    mark("gate_up_proj GEMM + SiLU*mul");
    RUN(row_gemm(act, w.down, inter, "down_proj GEMM"));
  }
  RUN(add_norm(final_norm));                                                                                           // qwen3.rs:389,501
  const bf16* head_in = normed;
  int rows = N;
  if (is_prefill) {  // only the last token of each sequence feeds the head (embed_head.rs:250-289)
    RUN(nvllm_gather_last_tokens(last_hidden, normed, d_cu_q, S, hidden, NVLLM_BF16, st)); ++launches; mark("gather_last_tokens");
    head_in = last_hidden;
    rows = S;
  }
  ++launches;
  RUN(nvllm_gemm_bf16(logits, head_in, lm_head, rows, vshard, hidden, hidden, hidden, vshard, NVLLM_F32, nullptr, 0, st));
  mark("lm_head GEMM");
  if (tp == 1) {
    RUN(nvllm_argmax(tokens, nullptr, logits, rows, vshard, vshard, 0, NVLLM_F32, st)); ++launches; mark("argmax");
  } else {  // sharded argmax: local (max, id) -> all-gather -> first-max combine
    RUN(nvllm_argmax(local_ids, local_max, logits, rows, vshard, vshard, static_cast<int64_t>(rank) * vshard, NVLLM_F32, st));
    RUN(nvllm_allgather(comm, local_max, all_max, sizeof(float) * rows, st));
    RUN(nvllm_allgather(comm, local_ids, all_ids, sizeof(int64_t) * rows, st));
    RUN(nvllm_argmax_combine(tokens, all_max, all_ids, tp, rows, st));
    launches += 4;
    mark("sharded argmax (argmax + 2 all-gathers + combine)");
  }
  return 0;
}

int nvllm_runner::execute_async(const nvllm_batch& b) {
  NVLLM_CHECK_ARG(weights_ready, NVLLM_ERR_STATE,
                  "execute: weights not initialised (init_weights, or load_weight for EVERY tensor of the architecture)");
  NVLLM_CHECK_ARG(b.num_seqs > 0 && b.num_tokens > 0, NVLLM_ERR_ARG, "execute: empty batch");
  bool sampled = false;
  for (int i = 0; i < b.num_seqs && b.temperatures != nullptr; ++i) sampled |= b.temperatures[i] != 0.0f;
  if (b.is_prefill) {
    RUN(upload(b, 0));
    RUN(forward(b.num_tokens, b.num_seqs, true, b.has_block_tables != 0, b));
    last_launches = launches;
  } else {
    const int padded = bucket_for(b.num_seqs, max_seqs);
    RUN(upload(b, padded));
    Graph& g = graphs[padded];
    ++g.uses;
    if (cfg.enforce_eager || g.uses == 1) {
      // first use of a bucket runs eagerly (lazy per-kernel attribute setup happens outside capture)
      RUN(forward(padded, padded, false, false, b));
      last_launches = launches;
    } else {
      if (g.exec == nullptr) {
        cudaGraph_t graph = nullptr;
        NVLLM_CUDA(cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal));
        const int rc = forward(padded, padded, false, false, b);
        cudaError_t e = cudaStreamEndCapture(stream, &graph);
        if (rc != 0) { if (graph) cudaGraphDestroy(graph); return rc; }
        NVLLM_CUDA(e);
        size_t nn = 0;
        cudaGraphGetNodes(graph, nullptr, &nn);
        g.nodes = static_cast<int64_t>(nn);
        NVLLM_CUDA(cudaGraphInstantiate(&g.exec, graph, 0));
        cudaGraphDestroy(graph);
      }
      NVLLM_CUDA(cudaGraphLaunch(g.exec, stream));
      last_launches = g.nodes;
    }
  }
  if (sampled) RUN(sample(b));  // overwrites the greedy tokens of the forward (outside the decode graph)
  NVLLM_CUDA(cudaMemcpyAsync(h_tokens, tokens, sizeof(int64_t) * b.num_seqs, cudaMemcpyDeviceToHost, stream));
  return 0;
}

// ModelRunner::sample_tokens -> Sampler::batch_sample (model_runner.rs:131-156, sampler.rs:221-254): the per-sequence
// parameters travel as four small arrays in one H2D copy; one kernel samples the whole batch from the f32 logits.
int nvllm_runner::sample(const nvllm_batch& b) {
  const size_t S = static_cast<size_t>(max_seqs), n = static_cast<size_t>(b.num_seqs);
  float* t = reinterpret_cast<float*>(h_samp);
  int32_t* k = reinterpret_cast<int32_t*>(h_samp + 4 * S);
  float* p = reinterpret_cast<float*>(h_samp + 8 * S);
  uint64_t* sd = reinterpret_cast<uint64_t*>(h_samp + 16 * S);
  auto mix = [](uint64_t x) {
    x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
    x ^= x >> 27; x *= 0x94D049BB133111EBull;
    x ^= x >> 31;
    return x;
  };
  const uint64_t es = static_cast<uint64_t>(static_cast<int64_t>(cfg.sampling_seed));
  for (size_t i = 0; i < n; ++i) {
    t[i] = b.temperatures[i];
    k[i] = b.top_k ? b.top_k[i] : 0;
    p[i] = b.top_p ? b.top_p[i] : 0.0f;
    const int64_t len = b.is_prefill ? b.cu_seqlens_k[i + 1] - b.cu_seqlens_k[i] : b.context_lens[i];
    const uint64_t sid = b.seq_ids ? static_cast<uint64_t>(b.seq_ids[i]) : i;
    sd[i] = mix(es ^ mix(sid * 0x9E3779B97F4A7C15ull + static_cast<uint64_t>(len)));
  }
  NVLLM_CUDA(cudaMemcpyAsync(d_samp, h_samp, 24 * S, cudaMemcpyHostToDevice, stream));
  ++last_launches;
  if (tp > 1) {
    if (gathered_logits == nullptr) {  // first sampled batch: every rank gets here in the same step (same batches everywhere);
      // outside any graph capture (sampling runs after the decode graph), one-time cudaMalloc device synchronisation
      RUN(dalloc(&gathered_logits, static_cast<size_t>(max_seqs) * vocab));
      RUN(dalloc(&full_logits, static_cast<size_t>(max_seqs) * vocab));
    }
    // Vocab-sharded head (embed_head.rs:321-336 TODO): all-gather the f32 shards, put the rows back together, and run the
    // SAME deterministic sampler on the full rows on every rank — identical tokens everywhere, no token broadcast.
    RUN(nvllm_allgather(comm, logits, gathered_logits, sizeof(float) * n * vshard, stream));
    RUN(nvllm_assemble_vocab_shards(full_logits, gathered_logits, tp, b.num_seqs, vshard, stream));
    last_launches += 2;
    return nvllm_sample(tokens, full_logits, b.num_seqs, vocab, vocab, reinterpret_cast<const float*>(d_samp),
                        reinterpret_cast<const int32_t*>(d_samp + 4 * S), reinterpret_cast<const float*>(d_samp + 8 * S),
                        reinterpret_cast<const uint64_t*>(d_samp + 16 * S), nullptr, nullptr, stream);
  }
  return nvllm_sample(tokens, logits, b.num_seqs, vshard, vshard, reinterpret_cast<const float*>(d_samp),
                      reinterpret_cast<const int32_t*>(d_samp + 4 * S), reinterpret_cast<const float*>(d_samp + 8 * S),
                      reinterpret_cast<const uint64_t*>(d_samp + 16 * S), nullptr, nullptr, stream);
}

int nvllm_runner::fetch(int64_t* out, int n) {
  NVLLM_CUDA(cudaStreamSynchronize(stream));
  if (!prof_marks.empty()) prof_collect();
  if (comm != nullptr) RUN(nvllm_comm_error(comm));  // a fused TP kernel gave up waiting for a peer: tokens are garbage
  if (out) memcpy(out, h_tokens, sizeof(int64_t) * n);
  return 0;
}

// ------------------------------------------------------------------------------------------------ C ABI
#define NVLLM_TRY_NEW(ptr, expr)                 \
  try {                                          \
    ptr = (expr);                                \
  } catch (const std::exception& e) {            \
    set_error("%s", e.what());                   \
    return NVLLM_ERR_STATE;                      \
  }

struct nvllm_scheduler;  // host_capi.cu
extern "C" {

int nvllm_runner_create(nvllm_runner** r, const nvllm_config* cfg, nvllm_comm* comm) {
  NVLLM_CHECK_ARG(r && cfg, NVLLM_ERR_ARG, "runner_create: NULL");
  nvllm_runner* p = nullptr;
  NVLLM_TRY_NEW(p, new nvllm_runner());
  const int rc = p->create(*cfg, comm);
  if (rc != 0) { delete p; return rc; }
  *r = p;
  return 0;
}
int nvllm_runner_destroy(nvllm_runner* r) {
  if (r) { cudaStreamSynchronize(r->stream); delete r; }
  return 0;
}
int nvllm_runner_init_weights(nvllm_runner* r, uint64_t seed, float std) {
  NVLLM_CHECK_ARG(r, NVLLM_ERR_ARG, "runner_init_weights: NULL");
  return r->init_weights(seed, std);
}
int nvllm_runner_load_weight(nvllm_runner* r, const char* name, const void* h_bf16, int64_t rows, int64_t cols) {
  NVLLM_CHECK_ARG(r, NVLLM_ERR_ARG, "runner_load_weight: NULL");
  return r->load_weight(name, h_bf16, rows, cols);
}
int nvllm_runner_execute_async(nvllm_runner* r, const nvllm_batch* b) {
  NVLLM_CHECK_ARG(r && b, NVLLM_ERR_ARG, "runner_execute: NULL");
  return r->execute_async(*b);
}
int nvllm_runner_fetch_tokens(nvllm_runner* r, int64_t* h_tokens, int32_t n) {
  NVLLM_CHECK_ARG(r && n >= 0 && n <= r->max_seqs, NVLLM_ERR_ARG, "runner_fetch_tokens: arguments");
  return r->fetch(h_tokens, n);
}
int nvllm_runner_execute(nvllm_runner* r, const nvllm_batch* b, int64_t* h_tokens, float* h_logits) {
  NVLLM_CHECK_ARG(r && b, NVLLM_ERR_ARG, "runner_execute: NULL");
  RUN(r->execute_async(*b));
  if (h_logits)
    NVLLM_CUDA(cudaMemcpyAsync(h_logits, r->logits, sizeof(float) * b->num_seqs * r->vshard, cudaMemcpyDeviceToHost, r->stream));
  return r->fetch(h_tokens, b->num_seqs);
}
void* nvllm_runner_stream(nvllm_runner* r) { return r ? r->stream : nullptr; }
int nvllm_runner_profile(nvllm_runner* r, int32_t enable) {
  NVLLM_CHECK_ARG(r, NVLLM_ERR_ARG, "runner_profile: NULL");
  NVLLM_CUDA(cudaStreamSynchronize(r->stream));
  if (!r->prof_marks.empty()) r->prof_collect();
  if (enable && !r->profiling) r->prof_acc.clear();
  r->profiling = enable != 0;
  return 0;
}
int64_t nvllm_runner_profile_report(nvllm_runner* r, char* h_buf, size_t cap) {
  if (!r) return 0;
  std::string out;
  char line[256];
  for (const auto& kv : r->prof_acc) {
    snprintf(line, sizeof(line), "%s\t%lld\t%.3f\n", kv.first.c_str(), (long long)kv.second.second, kv.second.first);
    out += line;
  }
  if (h_buf && cap > 0) {
    const size_t n = out.size() < cap - 1 ? out.size() : cap - 1;
    memcpy(h_buf, out.data(), n);
    h_buf[n] = 0;
  }
  return static_cast<int64_t>(out.size());
}
int64_t nvllm_runner_last_launches(const nvllm_runner* r) { return r ? r->last_launches : 0; }
int nvllm_runner_kv_cache(nvllm_runner* r, int32_t layer, void** k_cache, void** v_cache, nvllm_kv_layout* layout) {
  NVLLM_CHECK_ARG(r && layer >= 0 && layer < r->layers, NVLLM_ERR_ARG, "runner_kv_cache: layer %d", layer);
  if (k_cache) *k_cache = r->kc[layer];
  if (v_cache) *v_cache = r->vc[layer];
  if (layout) *layout = r->layout;
  return 0;
}

// Replays the last decode step `iters` times with its metadata already resident in HBM (no host work, no copies):
// the device-only cost of one step.  KV writes land in the same slots every time (idempotent).
int nvllm_runner_replay_decode(nvllm_runner* r, int32_t iters, float* ms_total) {
  NVLLM_CHECK_ARG(r && iters > 0 && ms_total, NVLLM_ERR_ARG, "replay_decode: arguments");
  NVLLM_CHECK_ARG(r->last_decode_b > 0 && r->d_ctx != nullptr, NVLLM_ERR_STATE, "replay_decode: no decode step has run yet");
  const int P = r->last_decode_b;
  auto it = r->graphs.find(P);
  const bool use_graph = !r->cfg.enforce_eager && it != r->graphs.end() && it->second.exec != nullptr;
  nvllm_batch dummy;
  memset(&dummy, 0, sizeof(dummy));
  cudaEvent_t e0, e1;
  NVLLM_CUDA(cudaEventCreate(&e0));
  NVLLM_CUDA(cudaEventCreate(&e1));
  NVLLM_CUDA(cudaEventRecord(e0, r->stream));
  for (int i = 0; i < iters; ++i) {
    if (use_graph) NVLLM_CUDA(cudaGraphLaunch(it->second.exec, r->stream));
    else RUN(r->forward(P, P, false, false, dummy));
  }
  NVLLM_CUDA(cudaEventRecord(e1, r->stream));
  NVLLM_CUDA(cudaEventSynchronize(e1));
  NVLLM_CUDA(cudaEventElapsedTime(ms_total, e0, e1));
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return 0;
}

// Re-launches the paged decode attention on the last decode batch's metadata (still resident), cycling through the
// layers' caches so that consecutive launches never re-read the same KV from L2; CUDA events on the runner's stream.
int nvllm_runner_time_decode_attention(nvllm_runner* r, int32_t iters, float* ms_per_launch, int64_t* alg_bytes) {
  NVLLM_CHECK_ARG(r && iters > 0 && ms_per_launch, NVLLM_ERR_ARG, "time_decode_attention: arguments");
  NVLLM_CHECK_ARG(r->last_decode_b > 0 && r->d_ctx != nullptr, NVLLM_ERR_STATE, "time_decode_attention: no decode step has run yet");
  const int B = r->last_decode_b, HD = r->H * r->D;
  const float scale = 1.0f / sqrtf(static_cast<float>(r->D));
  const bool fused = r->fuse_decode && !(r->fuse_rope && r->D == 128 && B <= 256);  // the kernel the decode step launches
  auto launch = [&](int l) {
    if (fused) {  // re-stores the step's K/V rows into the same slots: idempotent
      const auto& w = r->lw[l];
      return nvllm_paged_decode_attention_fused(r->attn, r->qkv, r->d_pos, r->rope, r->cfg.qk_norm ? w.q_norm : nullptr,
                                                r->cfg.qk_norm ? w.k_norm : nullptr, r->cfg.rms_norm_eps,
                                                r->cfg.max_position_embeddings, r->kc[l], r->vc[l], r->d_slots, r->d_tables,
                                                r->d_ctx, B, r->H, r->table_cap, r->qkv_dim, HD, &r->layout, scale,
                                                r->decode_ws, r->decode_ws_bytes, r->last_decode_splits, NVLLM_BF16, r->stream);
    }
    return nvllm_paged_decode_attention(r->attn, r->qkv, r->kc[l], r->vc[l], r->d_tables, r->d_ctx, B, r->H, r->table_cap,
                                        r->qkv_dim, HD, &r->layout, scale, r->decode_ws, r->decode_ws_bytes,
                                        r->last_decode_splits, NVLLM_BF16, r->stream);
  };
  for (int i = 0; i < 3; ++i) RUN(launch(i % r->layers));
  cudaEvent_t e0, e1;
  NVLLM_CUDA(cudaEventCreate(&e0));
  NVLLM_CUDA(cudaEventCreate(&e1));
  NVLLM_CUDA(cudaEventRecord(e0, r->stream));
  for (int i = 0; i < iters; ++i) RUN(launch(i % r->layers));
  NVLLM_CUDA(cudaEventRecord(e1, r->stream));
  NVLLM_CUDA(cudaEventSynchronize(e1));
  float ms = 0.f;
  NVLLM_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  *ms_per_launch = ms / iters;
  if (alg_bytes) {  // sum_b 2*L_b*Hkv*D*2 + 2*B*H*D*2 + 4*sum_b ceil(L_b/bs) + 4*B
    int64_t bytes = 0;
    for (int32_t L : r->last_ctx)
      bytes += 2ll * L * r->Hkv * r->D * 2 + 4ll * ((L + r->layout.block_size - 1) / r->layout.block_size);
    bytes += 2ll * B * r->H * r->D * 2 + 4ll * B;
    *alg_bytes = bytes;
  }
  return 0;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------ LLMEngine
// LLMEngine::step (llm_engine.rs:155-197): schedule -> execute_model -> sample_tokens -> postprocess.
extern "C" {
int nvllm_scheduler_create(nvllm_scheduler** s, const nvllm_config* cfg);
int nvllm_scheduler_destroy(nvllm_scheduler* s);
}

struct nvllm_engine {
  nvllm_scheduler* sched = nullptr;
  nvllm_runner* runner = nullptr;
  std::vector<int64_t> toks;
};

extern "C" {

int nvllm_engine_create(nvllm_engine** e, const nvllm_config* cfg, nvllm_comm* comm) {
  NVLLM_CHECK_ARG(e && cfg, NVLLM_ERR_ARG, "engine_create: NULL");
  nvllm_engine* p = nullptr;
  NVLLM_TRY_NEW(p, new nvllm_engine());
  int rc = nvllm_scheduler_create(&p->sched, cfg);
  if (rc == 0) rc = nvllm_runner_create(&p->runner, cfg, comm);
  if (rc != 0) {
    if (p->sched) nvllm_scheduler_destroy(p->sched);
    delete p;
    return rc;
  }
  p->toks.resize(cfg->max_num_seqs);
  *e = p;
  return 0;
}
int nvllm_engine_destroy(nvllm_engine* e) {
  if (!e) return 0;
  nvllm_runner_destroy(e->runner);
  nvllm_scheduler_destroy(e->sched);
  delete e;
  return 0;
}
nvllm_scheduler* nvllm_engine_scheduler(nvllm_engine* e) { return e ? e->sched : nullptr; }
nvllm_runner* nvllm_engine_runner(nvllm_engine* e) { return e ? e->runner : nullptr; }
int64_t nvllm_engine_add_request(nvllm_engine* e, const int64_t* h_prompt, int32_t n, const nvllm_sampling_params* sp) {
  NVLLM_CHECK_ARG(e, NVLLM_ERR_ARG, "engine_add_request: NULL");
  return nvllm_scheduler_add_sequence(e->sched, h_prompt, n, sp);
}
int nvllm_engine_is_finished(const nvllm_engine* e) { return e ? nvllm_scheduler_is_finished(e->sched) : 1; }
int nvllm_engine_step(nvllm_engine* e, int32_t* out_is_prefill, int32_t* out_num_seqs, int32_t* out_num_tokens) {
  NVLLM_CHECK_ARG(e, NVLLM_ERR_ARG, "engine_step: NULL");
  nvllm_batch b;
  RUN(nvllm_scheduler_schedule(e->sched, &b));
  RUN(nvllm_runner_execute(e->runner, &b, e->toks.data(), nullptr));
  RUN(nvllm_scheduler_postprocess(e->sched, e->toks.data(), b.num_seqs));
  if (out_is_prefill) *out_is_prefill = b.is_prefill;
  if (out_num_seqs) *out_num_seqs = b.num_seqs;
  if (out_num_tokens) *out_num_tokens = b.num_tokens;
  return 0;
}

}  // extern "C"
