/*
 * nvllm_engine.h — C ABI of the C++ host runtime that mirrors the reference's Rust engine types for the
 * paged-attention decode hot path: Config / Qwen3Config, SamplingParams, Sequence, BlockManager, Scheduler,
 * ModelRunner, LLMEngine.  The reference toolchain (Rust) is absent from this image, so the host side above the
 * kernel ABI (nvllm.h) is C++; a Rust crate would bind either level (rust/ffi.rs).
 * Citations are file:line under the reference's src/.
 *
 * Error convention: as nvllm.h (0 OK, <0 argument/state error, >0 cudaError_t, nvllm_last_error()).
 * Threading: one caller thread per object; no global state (the reference's process-global Context,
 * utils/context.rs:155, is replaced by the explicit nvllm_batch).
 */
#ifndef NVLLM_ENGINE_H_
#define NVLLM_ENGINE_H_

#include "nvllm.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Config (config.rs:16-71) + Qwen3Config (models/qwen3.rs:26-68) with the named-architecture fields the
 * reference lacks (head_dim, qk_norm — SURVEY.md §8 note N1). */
typedef struct nvllm_config {
  /* Config */
  int32_t max_num_batched_tokens; /* 32768 */
  int32_t max_num_seqs;           /* 512   */
  int32_t max_model_len;          /* 4096  */
  int32_t tensor_parallel_size;   /* 1     */
  int32_t enforce_eager;          /* 0: decode steps run as captured CUDA graphs */
  int32_t kvcache_block_size;     /* 256   */
  int32_t num_kvcache_blocks;     /* 0 = None -> 1000 (scheduler.rs:72, model_runner.rs:84) */
  int32_t tp_rank;                /* this process's rank (the reference always builds rank 0, model_runner.rs:75) */
  int64_t eos_token_id;           /* -1 = None */
  /* Qwen3Config */
  int32_t vocab_size, hidden_size, intermediate_size, num_hidden_layers;
  int32_t num_attention_heads, num_key_value_heads, head_dim, max_position_embeddings;
  int32_t tie_word_embeddings, qk_norm;
  int32_t kv_layout;              /* 0 = head-major [blk,Hkv,bs,D] (engine default), 1 = reference [blk,bs,Hkv,D] */
  int32_t sampling_seed;          /* engine seed of the sampler's counter-based uniforms (nvllm.h, nvllm_sample) */
  float rms_norm_eps;
  float reserved_f;
  double rope_theta;
} nvllm_config;

/* Fills the reference defaults (config.rs:54-71) and a named Qwen3 architecture:
 * "qwen3-0.6b" | "qwen3-8b" | "qwen3-32b" | "reference-default" (qwen3.rs:70-89: hidden 4096, 32 layers, 32/32 heads,
 * theta 1e4, no qk-norm) | "tiny" / "tiny-tp8" (2-layer test models). */
int nvllm_config_init(nvllm_config* cfg, const char* arch);
/* Config::validate (config.rs:83-119) + Qwen3Config::validate (qwen3.rs:106-123) on the fields above. */
int nvllm_config_validate(const nvllm_config* cfg);

/* SamplingParams (sampling_params.rs:10-88).  temperature == 0 is greedy (sampler.rs:79-81); otherwise temperature /
 * top-k / top-p / Gumbel-max sampling runs on the device (nvllm_sample).  Under tensor parallelism the vocab-sharded
 * logits are all-gathered and re-assembled (nvllm_assemble_vocab_shards) so that every rank samples the same token from
 * the same full rows.  repetition_penalty is carried but unused, as in the reference (sampler.rs:26-67 never reads it). */
typedef struct nvllm_sampling_params {
  float temperature; /* reference default 1.0 */
  int32_t max_tokens; /* 64 */
  int32_t ignore_eos; /* 0 */
  int32_t top_k;      /* 0 = None */
  float top_p;        /* 0 = None */
  float repetition_penalty; /* 0 = None */
} nvllm_sampling_params;
int nvllm_sampling_params_init(nvllm_sampling_params* sp);

/* BlockManager::compute_hash (block_manager.rs:109-123): XXH64(seed 0) over [prefix u64 LE]? || tokens i64 LE. */
uint64_t nvllm_compute_hash(const int64_t* h_tokens, size_t n, int32_t has_prefix, uint64_t prefix);

/* One scheduled step: what ModelRunner::prepare_inputs + create_context build (model_runner.rs:159-300), i.e. the
 * fields of utils/context.rs:11-35 plus ids/positions, as HOST arrays owned by the scheduler (valid until its next
 * schedule call).  i64: seq_ids, input_ids, positions; i32: everything else. */
typedef struct nvllm_batch {
  int32_t is_prefill, num_seqs, num_tokens, max_seqlen_q, max_seqlen_k, max_blocks, has_block_tables, reserved;
  const int64_t* seq_ids;      /* [num_seqs] */
  const int64_t* input_ids;    /* [num_tokens] */
  const int64_t* positions;    /* [num_tokens] */
  const int32_t* slot_mapping; /* [num_tokens] */
  const int32_t* cu_seqlens_q; /* [num_seqs+1] (prefill) */
  const int32_t* cu_seqlens_k; /* [num_seqs+1] (prefill) */
  const int32_t* context_lens; /* [num_seqs]   (decode)  */
  const int32_t* block_tables; /* [num_seqs, max_blocks] padded with -1 (decode; prefill iff has_block_tables) */
  const float* temperatures;   /* [num_seqs] */
  const int32_t* top_k;        /* [num_seqs], 0 = None */
  const float* top_p;          /* [num_seqs], 0 = None */
} nvllm_batch;

/* Scheduler (scheduler.rs:14-334) owning a BlockManager (block_manager.rs:69-331) and the Sequences
 * (sequence.rs:50-237).  Pure host code: works without a GPU. */
typedef struct nvllm_scheduler nvllm_scheduler;
int nvllm_scheduler_create(nvllm_scheduler** s, const nvllm_config* cfg);
int nvllm_scheduler_destroy(nvllm_scheduler* s);
/* Scheduler::add_sequence(Sequence::new(prompt, params)); returns the sequence id (>= 0) or a negative error.
 * Admission bounds the reference never enforces (config.rs:27,60 carry max_model_len, nothing reads it): a prompt longer
 * than max_model_len or than max_num_batched_tokens is refused with NVLLM_ERR_ARG before any state changes, and
 * max_tokens is clamped to max_model_len - n + 1 so no sequence can outgrow the runner's block tables. */
int64_t nvllm_scheduler_add_sequence(nvllm_scheduler* s, const int64_t* h_prompt, int32_t n,
                                     const nvllm_sampling_params* sp);
int nvllm_scheduler_is_finished(const nvllm_scheduler* s);
/* Scheduler::schedule (prefill-first, else decode with tail preemption) + the runner's metadata builders. */
int nvllm_scheduler_schedule(nvllm_scheduler* s, nvllm_batch* out);
/* Scheduler::postprocess: append one token per scheduled sequence, stop / deallocate finished ones. */
int nvllm_scheduler_postprocess(nvllm_scheduler* s, const int64_t* h_tokens, int32_t n);
typedef struct nvllm_sched_stats {
  int64_t total_sequences, waiting, running, finished, preemptions, prefill_batches, decode_batches;
  int64_t total_blocks, free_blocks, used_blocks, cached_blocks;
} nvllm_sched_stats;
int nvllm_scheduler_stats(const nvllm_scheduler* s, nvllm_sched_stats* out);
/* Block introspection for the bit-exact tests: ref_count, hash (has_hash = 0 when None). */
int nvllm_scheduler_block_info(const nvllm_scheduler* s, int32_t block_id, int32_t* ref_count,
                               int32_t* has_hash, uint64_t* hash);
/* Copies the free list (front first) into h_out[cap]; returns its length. */
int32_t nvllm_scheduler_free_list(const nvllm_scheduler* s, int32_t* h_out, int32_t cap);
/* Sequence read-back: number of tokens (prompt + completion), copies up to cap of them; status: 0 Waiting, 1 Running,
 * 2 Finished, 3 Preempted. */
int32_t nvllm_scheduler_sequence(const nvllm_scheduler* s, int64_t seq_id, int64_t* h_tokens, int32_t cap,
                                 int32_t* num_prompt_tokens, int32_t* status);

/* Frees the storage of a FINISHED sequence and recycles its slot (the reference returns finished sequences to the caller
 * and drops them, llm_engine.rs:186-193).  Its id is never reused; nvllm_scheduler_sequence returns -1 for it afterwards. */
int nvllm_scheduler_release_sequence(nvllm_scheduler* s, int64_t seq_id);

/* ModelRunner (model_runner.rs:19-448) + Qwen3Model (models/qwen3.rs): device weights (this rank's TP shard),
 * per-layer paged KV cache, activation buffers, decode CUDA graphs per batch-size bucket. */
typedef struct nvllm_runner nvllm_runner;
int nvllm_runner_create(nvllm_runner** r, const nvllm_config* cfg, nvllm_comm* comm /* NULL when tp == 1 */);
int nvllm_runner_destroy(nvllm_runner* r);
/* Seeded synthetic weights (nvllm_init_weight semantics per logical tensor, sliced per rank; norm weights = 1). */
int nvllm_runner_init_weights(nvllm_runner* r, uint64_t seed, float std);
/* Loads one full (unsharded) HF-named tensor from host bf16 memory; the runner keeps its TP slice and packs
 * q/k/v -> qkv_proj, gate/up -> gate_up_proj (utils/loader.rs:175-198; linear.rs:154-172,249-267). */
int nvllm_runner_load_weight(nvllm_runner* r, const char* name, const void* h_bf16, int64_t rows, int64_t cols);
/* ModelRunner::execute_model + sample_tokens (model_runner.rs:105-156): runs the forward for `batch`, writes the
 * sampled token of every sequence (greedy argmax, or nvllm_sample when any temperature != 0; the row seed of
 * sequence s at length n is mix64(sampling_seed ^ mix64(s * 0x9E3779B97F4A7C15 + n))) to h_tokens[num_seqs].  h_logits (nullable): this rank's f32 logits
 * [num_seqs, vocab/tp] copied back for parity tests. Synchronises the runner's stream before returning. */
int nvllm_runner_execute(nvllm_runner* r, const nvllm_batch* batch, int64_t* h_tokens, float* h_logits);
/* Same step split in two, for device-side timing: enqueue (no sync) / wait + fetch tokens. */
int nvllm_runner_execute_async(nvllm_runner* r, const nvllm_batch* batch);
int nvllm_runner_fetch_tokens(nvllm_runner* r, int64_t* h_tokens, int32_t n);
void* nvllm_runner_stream(nvllm_runner* r);
/* Per-launch profile of EAGER steps (enforce_eager, prefill, or the first use of a decode bucket): while enabled the
 * runner records a CUDA event after every launch; consecutive differences are serialised per-kernel times (an event
 * between two launches also removes their programmatic overlap).  This is the launch list for multi-rank runs, where
 * kernels that wait for peers cannot be serialised by ncu.  enable = 1 clears the accumulators and starts, 0 stops.
 * nvllm_runner_profile_report writes "name\tlaunches\ttotal_us\n" lines into h_buf (NUL-terminated, truncated to cap) and
 * returns the full length. */
int nvllm_runner_profile(nvllm_runner* r, int32_t enable);
int64_t nvllm_runner_profile_report(nvllm_runner* r, char* h_buf, size_t cap);
/* Number of kernels launched by the last execute (decode graph node count, or eager launches). */
int64_t nvllm_runner_last_launches(const nvllm_runner* r);
/* Device pointers of layer `layer`'s K / V cache and the layout in use (tests). */
int nvllm_runner_kv_cache(nvllm_runner* r, int32_t layer, void** k_cache, void** v_cache, nvllm_kv_layout* layout);
/* Re-runs only the paged decode attention of layer 0 on the last decode batch `iters` times between two CUDA
 * events (for the roofline line of bench.py); returns the average milliseconds per launch. */
int nvllm_runner_time_decode_attention(nvllm_runner* r, int32_t iters, float* ms_per_launch, int64_t* alg_bytes);

/* Replays the last decode step `iters` times with its inputs already resident in HBM (graph launches only; no host
 * scheduling, no copies) between two CUDA events on the runner's stream; writes the total milliseconds. */
int nvllm_runner_replay_decode(nvllm_runner* r, int32_t iters, float* ms_total);

/* LLMEngine (llm_engine.rs:23-362): scheduler + runner; step() = schedule -> execute_model -> sample_tokens ->
 * postprocess (llm_engine.rs:155-197).  The async/streaming/tokeniser facade is out of scope. */
typedef struct nvllm_engine nvllm_engine;
int nvllm_engine_create(nvllm_engine** e, const nvllm_config* cfg, nvllm_comm* comm);
int nvllm_engine_destroy(nvllm_engine* e);
nvllm_scheduler* nvllm_engine_scheduler(nvllm_engine* e);
nvllm_runner* nvllm_engine_runner(nvllm_engine* e);
int64_t nvllm_engine_add_request(nvllm_engine* e, const int64_t* h_prompt, int32_t n, const nvllm_sampling_params* sp);
/* One engine step.  out_is_prefill / out_num_seqs / out_num_tokens describe the batch that ran (nullable). */
int nvllm_engine_step(nvllm_engine* e, int32_t* out_is_prefill, int32_t* out_num_seqs, int32_t* out_num_tokens);
int nvllm_engine_is_finished(const nvllm_engine* e);

#ifdef __cplusplus
}
#endif
#endif /* NVLLM_ENGINE_H_ */
