// host_capi.cu — extern "C" surface of the pure-host layer (include/nvllm_engine.h): Config, SamplingParams,
// compute_hash and the Scheduler.  No CUDA calls: these entry points work on a machine without a GPU.
#include <string.h>

#include "../common.cuh"
#include "scheduler.hpp"

using nvllm::host::Scheduler;

struct nvllm_scheduler {
  Scheduler impl;
  explicit nvllm_scheduler(const nvllm_config& c) : impl(c) {}
};

#define NVLLM_TRY(body)                         \
  try {                                         \
    body;                                       \
  } catch (const std::invalid_argument& e) {    \
    nvllm::set_error("%s", e.what());           \
    return NVLLM_ERR_ARG;                       \
  } catch (const std::exception& e) {           \
    nvllm::set_error("%s", e.what());           \
    return NVLLM_ERR_STATE;                     \
  }

extern "C" {

int nvllm_config_init(nvllm_config* c, const char* arch) {
  NVLLM_CHECK_ARG(c != nullptr && arch != nullptr, NVLLM_ERR_ARG, "config_init: NULL");
  memset(c, 0, sizeof(*c));
  // Config::default (config.rs:54-71)
  c->max_num_batched_tokens = 32768; c->max_num_seqs = 512; c->max_model_len = 4096;
  c->tensor_parallel_size = 1; c->enforce_eager = 0; c->kvcache_block_size = 256; c->num_kvcache_blocks = 0;
  c->tp_rank = 0; c->eos_token_id = -1;
  // common Qwen3 values: vocab 151936, eps 1e-6, head_dim 128, qk-norm, theta 1e6
  c->vocab_size = 151936; c->head_dim = 128; c->max_position_embeddings = 40960; c->qk_norm = 1;
  c->rms_norm_eps = 1e-6f; c->rope_theta = 1e6; c->kv_layout = 0;
  if (!strcmp(arch, "qwen3-0.6b")) {
    c->hidden_size = 1024; c->intermediate_size = 3072; c->num_hidden_layers = 28;
    c->num_attention_heads = 16; c->num_key_value_heads = 8; c->tie_word_embeddings = 1;
  } else if (!strcmp(arch, "qwen3-8b")) {
    c->hidden_size = 4096; c->intermediate_size = 12288; c->num_hidden_layers = 36;
    c->num_attention_heads = 32; c->num_key_value_heads = 8;
  } else if (!strcmp(arch, "qwen3-32b")) {
    c->hidden_size = 5120; c->intermediate_size = 25600; c->num_hidden_layers = 64;
    c->num_attention_heads = 64; c->num_key_value_heads = 8;
  } else if (!strcmp(arch, "reference-default")) {  // Qwen3Config::default (qwen3.rs:70-89)
    c->hidden_size = 4096; c->intermediate_size = 11008; c->num_hidden_layers = 32;
    c->num_attention_heads = 32; c->num_key_value_heads = 32; c->head_dim = 128; c->qk_norm = 0;
    c->rope_theta = 1e4; c->max_position_embeddings = 32768;
  } else if (!strcmp(arch, "tiny")) {
    c->vocab_size = 1024; c->hidden_size = 256; c->intermediate_size = 512; c->num_hidden_layers = 2;
    c->num_attention_heads = 4; c->num_key_value_heads = 2; c->max_position_embeddings = 4096;
    c->tie_word_embeddings = 0;
  } else if (!strcmp(arch, "qwen3-0.6b-2l")) {  // two layers at the real Qwen3-0.6B widths (HF golden fixture), small vocab
    c->vocab_size = 2048; c->hidden_size = 1024; c->intermediate_size = 3072; c->num_hidden_layers = 2;
    c->num_attention_heads = 16; c->num_key_value_heads = 8; c->max_position_embeddings = 4096;
    c->tie_word_embeddings = 0;
  } else if (!strcmp(arch, "tiny-tp8")) {  // 2-layer test model whose head counts divide by 8
    c->vocab_size = 2048; c->hidden_size = 512; c->intermediate_size = 1024; c->num_hidden_layers = 2;
    c->num_attention_heads = 16; c->num_key_value_heads = 8; c->max_position_embeddings = 4096;
    c->tie_word_embeddings = 0;
  } else {
    NVLLM_CHECK_ARG(false, NVLLM_ERR_ARG, "config_init: unknown architecture '%s'", arch);
  }
  return 0;
}

int nvllm_config_validate(const nvllm_config* c) {
  NVLLM_CHECK_ARG(c != nullptr, NVLLM_ERR_ARG, "config_validate: NULL");
  // config.rs:94 demands block_size %% 256 == 0, yet the reference's own scheduler / block-manager tests run with
  // 16 and 4 (scheduler.rs:383, block_manager.rs:427); the kernels need a multiple of 16.
  NVLLM_CHECK_ARG(c->kvcache_block_size > 0 && c->kvcache_block_size % 16 == 0, NVLLM_ERR_ARG,
                  "KV cache block size must be a positive multiple of 16, got %d", c->kvcache_block_size);
  // the paged-prefix prefill kernel walks the cache in 128-key tiles (prefill_attn.cu): a block must divide a tile or be
  // a whole number of tiles — refused here, not at the first prefix-cached prefill with scheduler state already mutated
  NVLLM_CHECK_ARG(c->kvcache_block_size >= 128 ? c->kvcache_block_size % 128 == 0 : 128 % c->kvcache_block_size == 0,
                  NVLLM_ERR_ARG, "KV cache block size %d must divide 128 or be a multiple of 128", c->kvcache_block_size);
  NVLLM_CHECK_ARG(c->tensor_parallel_size >= 1 && c->tensor_parallel_size <= 8, NVLLM_ERR_ARG,
                  "Tensor parallel size must be between 1 and 8, got %d", c->tensor_parallel_size);  // config.rs:99-101
  NVLLM_CHECK_ARG(c->tp_rank >= 0 && c->tp_rank < c->tensor_parallel_size, NVLLM_ERR_ARG, "tp_rank %d", c->tp_rank);
  NVLLM_CHECK_ARG(c->max_num_seqs > 0 && c->max_num_batched_tokens > 0 && c->max_model_len > 0, NVLLM_ERR_ARG,
                  "max_num_seqs / max_num_batched_tokens / max_model_len must be positive");
  NVLLM_CHECK_ARG(c->num_kvcache_blocks >= 0, NVLLM_ERR_ARG, "num_kvcache_blocks %d", c->num_kvcache_blocks);
  const int tp = c->tensor_parallel_size;
  // Qwen3Config::validate (qwen3.rs:106-123)
  NVLLM_CHECK_ARG(c->num_attention_heads > 0 && c->num_attention_heads % tp == 0, NVLLM_ERR_ARG,
                  "Number of attention heads must be divisible by tensor parallel size");
  NVLLM_CHECK_ARG(c->num_key_value_heads > 0 && c->num_key_value_heads % tp == 0, NVLLM_ERR_ARG,
                  "Number of key-value heads must be divisible by tensor parallel size");
  NVLLM_CHECK_ARG(c->intermediate_size > 0 && c->intermediate_size % tp == 0, NVLLM_ERR_ARG,
                  "Intermediate size must be divisible by tensor parallel size");
  NVLLM_CHECK_ARG(c->num_attention_heads % c->num_key_value_heads == 0, NVLLM_ERR_ARG,
                  "num_attention_heads must be a multiple of num_key_value_heads");
  // every Qwen3 named architecture has head_dim 128 and the tcgen05 prefill kernel is built for it (prefill_attn.cu)
  NVLLM_CHECK_ARG(c->head_dim == 128, NVLLM_ERR_ARG, "head_dim %d is not supported by the engine (128 only)", c->head_dim);
  NVLLM_CHECK_ARG(c->hidden_size > 0 && c->hidden_size % 64 == 0 && (c->intermediate_size / tp) % 64 == 0 &&
                      c->vocab_size > 0 && c->vocab_size % (8 * tp) == 0 && c->num_hidden_layers > 0,
                  NVLLM_ERR_ARG, "hidden %% 64, (intermediate/tp) %% 64, vocab %% (8 tp) must hold");
  NVLLM_CHECK_ARG(c->max_position_embeddings >= c->max_model_len, NVLLM_ERR_ARG,
                  "max_position_embeddings %d < max_model_len %d", c->max_position_embeddings, c->max_model_len);
  return 0;
}

int nvllm_sampling_params_init(nvllm_sampling_params* sp) {
  NVLLM_CHECK_ARG(sp != nullptr, NVLLM_ERR_ARG, "sampling_params_init: NULL");
  memset(sp, 0, sizeof(*sp));
  sp->temperature = 1.0f; sp->max_tokens = 64; sp->ignore_eos = 0;  // sampling_params.rs defaults
  return 0;
}

uint64_t nvllm_compute_hash(const int64_t* h_tokens, size_t n, int32_t has_prefix, uint64_t prefix) {
  return nvllm::host::compute_hash(h_tokens, n, has_prefix != 0, prefix);
}

int nvllm_scheduler_create(nvllm_scheduler** s, const nvllm_config* cfg) {
  NVLLM_CHECK_ARG(s && cfg, NVLLM_ERR_ARG, "scheduler_create: NULL");
  NVLLM_CHECK_ARG(cfg->kvcache_block_size > 0 && cfg->max_num_seqs > 0 && cfg->max_num_batched_tokens > 0,
                  NVLLM_ERR_ARG, "scheduler_create: block size / limits must be positive");
  NVLLM_TRY(*s = new nvllm_scheduler(*cfg));
  return 0;
}
int nvllm_scheduler_destroy(nvllm_scheduler* s) { delete s; return 0; }

int64_t nvllm_scheduler_add_sequence(nvllm_scheduler* s, const int64_t* h_prompt, int32_t n,
                                     const nvllm_sampling_params* sp) {
  NVLLM_CHECK_ARG(s && sp && (n == 0 || h_prompt), NVLLM_ERR_ARG, "add_sequence: NULL");
  NVLLM_CHECK_ARG(n > 0, NVLLM_ERR_ARG, "add_sequence: empty prompt");
  int64_t id = -1;
  NVLLM_TRY(id = s->impl.add_sequence(h_prompt, n, *sp));
  return id;
}
int nvllm_scheduler_is_finished(const nvllm_scheduler* s) { return s && s->impl.is_finished() ? 1 : 0; }

int nvllm_scheduler_schedule(nvllm_scheduler* s, nvllm_batch* out) {
  NVLLM_CHECK_ARG(s && out, NVLLM_ERR_ARG, "schedule: NULL");
  NVLLM_CHECK_ARG(!s->impl.is_finished(), NVLLM_ERR_STATE, "schedule: nothing to schedule");
  NVLLM_TRY(s->impl.schedule(out));
  return 0;
}
int nvllm_scheduler_postprocess(nvllm_scheduler* s, const int64_t* h_tokens, int32_t n) {
  NVLLM_CHECK_ARG(s && (n == 0 || h_tokens), NVLLM_ERR_ARG, "postprocess: NULL");
  NVLLM_TRY(s->impl.postprocess(h_tokens, n));
  return 0;
}
int nvllm_scheduler_stats(const nvllm_scheduler* s, nvllm_sched_stats* out) {
  NVLLM_CHECK_ARG(s && out, NVLLM_ERR_ARG, "stats: NULL");
  s->impl.stats(out);
  return 0;
}
int nvllm_scheduler_block_info(const nvllm_scheduler* s, int32_t block_id, int32_t* ref_count, int32_t* has_hash,
                               uint64_t* hash) {
  NVLLM_CHECK_ARG(s, NVLLM_ERR_ARG, "block_info: NULL");
  const auto& bm = s->impl.block_manager();
  NVLLM_CHECK_ARG(block_id >= 0 && block_id < bm.num_blocks(), NVLLM_ERR_ARG, "block_info: id %d", block_id);
  const auto& b = bm.block(block_id);
  if (ref_count) *ref_count = b.ref_count;
  if (has_hash) *has_hash = b.has_hash ? 1 : 0;
  if (hash) *hash = b.hash;
  return 0;
}
int32_t nvllm_scheduler_free_list(const nvllm_scheduler* s, int32_t* h_out, int32_t cap) {
  if (!s) return 0;
  const auto& bm = s->impl.block_manager();
  int32_t n = 0;
  for (int32_t id = bm.free_front(); id >= 0; id = bm.free_next(id), ++n)
    if (h_out && n < cap) h_out[n] = id;
  return n;
}
int32_t nvllm_scheduler_sequence(const nvllm_scheduler* s, int64_t seq_id, int64_t* h_tokens, int32_t cap,
                                 int32_t* num_prompt_tokens, int32_t* status) {
  const nvllm::host::Sequence* qp = s ? s->impl.find(seq_id) : nullptr;
  if (qp == nullptr) return -1;
  const auto& q = *qp;
  if (num_prompt_tokens) *num_prompt_tokens = q.num_prompt_tokens;
  if (status) *status = static_cast<int32_t>(q.status);
  if (h_tokens)
    for (int32_t i = 0; i < q.len() && i < cap; ++i) h_tokens[i] = q.token_ids[i];
  return q.len();
}

int nvllm_scheduler_release_sequence(nvllm_scheduler* s, int64_t seq_id) {
  NVLLM_CHECK_ARG(s, NVLLM_ERR_ARG, "release_sequence: NULL");
  NVLLM_CHECK_ARG(s->impl.release(seq_id), NVLLM_ERR_STATE, "release_sequence: sequence %lld is unknown or not finished",
                  (long long)seq_id);
  return 0;
}

}  // extern "C"
