/*
 * nvllm.h — C ABI of the B200-native (sm_100a) paged-attention decode hot path.
 *
 * This is the drop-in boundary for ssvgopal/nano-vllm-rs.  The reference has no FFI of its own
 * (pure Rust on candle); each entry point below replaces one Rust method on the hot path and cites
 * it as  <file>:<lines>  relative to the reference tree.  The Rust-side binding a maintainer
 * would add is shown in INTEGRATION.md and rust/ffi.rs.
 *
 * Conventions (all entry points):
 *   - return 0 on success, <0 for an argument error (shape / alignment / dtype), >0 = cudaError_t;
 *     nvllm_last_error() returns a thread-local message for the last non-zero return.
 *   - the caller owns every buffer (inputs, outputs, KV cache, workspace) and the stream; kernels
 *     never allocate, free or synchronise and keep no global state, so every call is legal inside
 *     CUDA-graph capture.  Pointers are device pointers unless the name starts with `h_`.
 *     (Stated exceptions: nvllm_paged_decode_attention_static owns a per-GPU scratch array and is outside this
 *     paragraph's promises — read its own comment; nvllm_reload_env is not re-entrant; nvllm_comm_* are collectives.)
 *   - activations / weights / KV cache are bf16 (NVLLM_BF16); accumulation is fp32.
 *   - strides are in ELEMENTS.  The innermost (head_dim) dimension is always contiguous and every
 *     row start must be 16-byte aligned.
 *   - `stream` is a cudaStream_t passed as void*.
 */
#ifndef NVLLM_H_
#define NVLLM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NVLLM_VERSION 100 /* 0.1.0 */

enum nvllm_dtype { NVLLM_BF16 = 0, NVLLM_F32 = 1 };

enum nvllm_status {
  NVLLM_OK = 0,
  NVLLM_ERR_ARG = -1,       /* bad shape / null pointer / unsupported size */
  NVLLM_ERR_ALIGN = -2,     /* pointer or stride not 16-byte aligned        */
  NVLLM_ERR_DTYPE = -3,     /* dtype other than the documented one          */
  NVLLM_ERR_WORKSPACE = -4, /* workspace too small                          */
  NVLLM_ERR_STATE = -5,     /* engine-level call in the wrong state         */
  NVLLM_ERR_COMM = -6       /* NCCL / peer-memory failure                   */
};

int nvllm_version(void);
const char* nvllm_last_error(void);
/* 1 when a CUDA device of compute capability 10.x is visible, else 0 (never throws). */
int nvllm_device_ok(void);
/* The NVLLM_* environment knobs (A/B measurement switches) are parsed once per process, never on a
 * dispatch path; benchmark tools that sweep a knob inside one process call this after changing the environment.
 * NOT re-entrant, unlike every other entry point: it rewrites the knob set in place, so call it only while no other thread
 * is inside the library (a dispatch running concurrently could see half of the old set and half of the new). */
int nvllm_reload_env(void);

/* ------------------------------------------------------------------------------------------------
 * KV-cache layout descriptor.  The reference allocates each per-layer cache as
 * [num_blocks, block_size, num_kv_heads, head_dim]  (src/engine/model_runner.rs:379-389); that is
 *   block_stride = block_size*Hkv*D, token_stride = Hkv*D, head_stride = D.
 * The engine in this repo uses the head-major variant [num_blocks, Hkv, block_size, D]
 *   block_stride = Hkv*block_size*D, token_stride = D, head_stride = block_size*D
 * (64 KiB contiguous per (block, head)).  Kernels accept either through these strides.
 * slot s  ->  block s / block_size, offset s % block_size   (slot_mapping semantics of
 * src/layers/attention.rs:150-174 with a flat [num_blocks*block_size, ...] view).
 * ---------------------------------------------------------------------------------------------- */
typedef struct nvllm_kv_layout {
  int32_t block_size;    /* tokens per block (reference: 256, src/config.rs:94)           */
  int32_t num_kv_heads;  /* KV heads held by this rank (Hkv / tp)                           */
  int32_t head_dim;      /* 64 or 128                                                       */
  int32_t reserved;
  int64_t block_stride;  /* elements between consecutive blocks                            */
  int64_t token_stride;  /* elements between consecutive tokens inside a block             */
  int64_t head_stride;   /* elements between consecutive KV heads of one token             */
} nvllm_kv_layout;

/* K1 — replaces Attention::store_kv_cache (src/layers/attention.rs:150-174).
 * cache[slot_mapping[i]] = k[i], v[i] for i in [0, num_tokens); slot < 0 is skipped (padding rows
 * of a captured decode graph).  k, v: [num_tokens, Hkv, D] with `kv_token_stride` elements between
 * tokens (so they may be views into a packed qkv buffer).  Bit-exact copy. */
int nvllm_store_kvcache(const void* k, const void* v, void* k_cache, void* v_cache,
                        const int32_t* slot_mapping, int32_t num_tokens, int64_t kv_token_stride,
                        const nvllm_kv_layout* layout, int32_t dtype, void* stream);

/* K2 — replaces Attention::flash_attention_decode -> compute_attention_with_cache ->
 * gather_cached_kv -> compute_attention (src/layers/attention.rs:225-318,238-261).
 * out[b,h,:] = softmax(scale * q[b,h,:] . K[b,h/g]^T) . V[b,h/g] over the first context_lens[b]
 * slots of the blocks listed in block_tables[b,:] (entries < 0 are skipped, attention.rs:306).
 * q/out: [batch, H, D] with `q_token_stride` / `out_token_stride` elements between sequences.
 * block_tables: i32 [batch, max_blocks_per_seq] padded with -1 (model_runner.rs:280-293);
 * context_lens: i32 [batch] (model_runner.rs:275).  Split-K: `num_splits` >= 1 partitions of the
 * KV length per (sequence, kv head); 0 lets the library choose.  Workspace holds the fp32 partials
 * (only touched when the effective split count is > 1). */
size_t nvllm_paged_decode_workspace_bytes(int32_t batch, int32_t num_heads, int32_t head_dim,
                                          int32_t num_splits);
int nvllm_paged_decode_choose_splits(int32_t batch, int32_t num_heads, int32_t num_kv_heads,
                                     int32_t max_context_len);
int nvllm_paged_decode_attention(void* out, const void* q, const void* k_cache, const void* v_cache,
                                 const int32_t* block_tables, const int32_t* context_lens,
                                 int32_t batch, int32_t num_heads, int32_t max_blocks_per_seq,
                                 int64_t q_token_stride, int64_t out_token_stride,
                                 const nvllm_kv_layout* layout, float scale, void* workspace,
                                 size_t workspace_bytes, int32_t num_splits, int32_t dtype,
                                 void* stream);

/* nvllm_paged_decode_attention for callers that cannot provide a workspace.  NO COUNTERPART IN THE REFERENCE (its
 * decode attention is candle tensor ops that allocate freely); this exists for integrations, not for parity.
 *
 * Who needs it: a caller that sits where device memory cannot be obtained — an operator hook or host callback that is
 * handed only the tensors of one op, an embedder whose memory pool is fixed once serving starts, a C / Rust host with
 * no CUDA allocator of its own.  It still passes q, the caches, the tables and `out`; the f32 split-K partials go to an
 * array built into libnvllm.so's device image.  The call performs no cudaMalloc / cudaFree and calls no allocating
 * library routine.  Same arguments, kernels and results as nvllm_paged_decode_attention (bit for bit at an equal
 * num_splits), minus `workspace` / `workspace_bytes`.
 *
 * Limits and rules:
 *   - the array is nvllm_paged_decode_static_scratch_bytes() = 8 MiB, present on EVERY GPU that loads the library until
 *     the process exits.  A call that needs more — nvllm_paged_decode_workspace_bytes(batch, num_heads, head_dim,
 *     num_splits) > 8 MiB; num_splits = 0 resolves through nvllm_paged_decode_choose_splits first — returns
 *     NVLLM_ERR_WORKSPACE with the sizes in nvllm_last_error() and launches nothing.  One split needs no scratch at all.
 *   - the first call on each device looks up that device's address of the array (cudaGetSymbolAddress, with the calling
 *     thread's current device, which must be the device of `stream` and of every buffer) and remembers it.  Make that
 *     first call during set-up, before any CUDA stream capture begins.
 *   - all calls on a GPU share the one array.  Calls ordered on one stream are safe; calls that can overlap on the GPU
 *     (two streams, or two threads with a stream each) race on the partials and return garbage — serialise them or use
 *     the workspace variant.  The array keeps whatever the last call left; nothing needs clearing, because every
 *     partial a call reads was written by that same call. */
size_t nvllm_paged_decode_static_scratch_bytes(void);
int nvllm_paged_decode_attention_static(void* out, const void* q, const void* k_cache, const void* v_cache,
                                        const int32_t* block_tables, const int32_t* context_lens, int32_t batch,
                                        int32_t num_heads, int32_t max_blocks_per_seq, int64_t q_token_stride,
                                        int64_t out_token_stride, const nvllm_kv_layout* layout, float scale,
                                        int32_t num_splits, int32_t dtype, void* stream);

/* K7 + K1 + K2 in ONE launch for decode steps — replaces, per layer, what Qwen3Attention::forward does between qkv_proj
 * and o_proj (src/models/qwen3.rs:219-233): per-head q/k RMSNorm + RotaryEmbedding::forward
 * (src/layers/rotary_embedding.rs:23-48,145-158), Attention::store_kv_cache (src/layers/attention.rs:150-174) and
 * Attention::flash_attention_decode (:225-318).  qkv: the packed [q | k | v] rows of this step ([batch, (H+2Hkv)*D], read
 * only — nothing is written back into it); the step's K (normed, rotated) and V rows are stored at slot_mapping[b] (< 0:
 * skipped) and attended to from shared memory, so context_lens[b] COUNTS this step's token (context_lens - 1 tokens are
 * read through block_tables).  Same rounding points as nvllm_qknorm_rope_store followed by nvllm_paged_decode_attention
 * (bf16 after the norm, bf16 after the rotation); head_dim 128 only.  q_norm_w / k_norm_w NULL = no qk-norm. */
int nvllm_paged_decode_attention_fused(void* out, const void* qkv, const int64_t* positions, const float* cos_sin,
                                       const void* q_norm_w, const void* k_norm_w, float eps, int32_t max_position,
                                       void* k_cache, void* v_cache, const int32_t* slot_mapping,
                                       const int32_t* block_tables, const int32_t* context_lens, int32_t batch,
                                       int32_t num_heads, int32_t max_blocks_per_seq, int64_t qkv_token_stride,
                                       int64_t out_token_stride, const nvllm_kv_layout* layout, float scale,
                                       void* workspace, size_t workspace_bytes, int32_t num_splits, int32_t dtype,
                                       void* stream);

/* K3 / K4 — replaces Attention::flash_attention_varlen (+ create_causal_mask) and
 * flash_attention_varlen_with_cache (src/layers/attention.rs:177-222,321-339).
 * Causal attention per sequence i over q rows [cu_seqlens_q[i], cu_seqlens_q[i+1]) and keys
 * [cu_seqlens_k[i], cu_seqlens_k[i+1]); the mask is bottom-right aligned (query row j of Lq sees
 * keys 0 .. Lk-Lq+j), which is the plain causal mask when Lq == Lk.
 * If block_tables == NULL keys/values come from the dense k, v: [total_k, Hkv, D];
 * otherwise (prefix-cache prefill) they are read from the paged cache through block_tables[i,:]
 * (the new K/V of this step must already have been stored — note N2 in SURVEY.md §8).
 * total_q / total_k: rows of q / of the dense k, v (bounds of the TMA descriptors; total_k is ignored when paged).
 * Requires Lk >= Lq per sequence and head_dim == 128; runs on tcgen05 tensor cores. */
int nvllm_varlen_prefill_attention(void* out, const void* q, const void* k, const void* v,
                                   const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k,
                                   int32_t batch, int32_t total_q, int32_t total_k,
                                   int32_t max_seqlen_q, int32_t max_seqlen_k, int32_t num_heads, int32_t num_kv_heads, int32_t head_dim,
                                   int64_t q_token_stride, int64_t kv_token_stride,
                                   int64_t out_token_stride, const void* k_cache,
                                   const void* v_cache, const int32_t* block_tables,
                                   int32_t max_blocks_per_seq, const nvllm_kv_layout* layout,
                                   float scale, int32_t dtype, void* stream);

/* K5 — replaces LinearBase::forward of ReplicatedLinear / ColumnParallelLinear / RowParallelLinear /
 * QKVParallelLinear / MergedColumnParallelLinear and ParallelLMHead::compute_logits
 * (src/layers/linear.rs:228-239,354-356,437-439; src/layers/embed_head.rs:292-306).
 * y[M,N] = x[M,K] . w[N,K]^T   (x, w bf16 row-major with leading dimensions ldx, ldw; fp32
 * accumulate on tcgen05 tensor cores; y is bf16 or fp32 per `out_dtype`, leading dimension ldy).
 * K % 64 == 0, N % 8 == 0, ldx/ldw % 8 == 0.  workspace: nvllm_gemm_workspace_bytes() bytes
 * (TMA descriptors are passed by value, the workspace is only used for split-K partials; may be
 * NULL when that function returns 0). */
size_t nvllm_gemm_workspace_bytes(int32_t m, int32_t n, int32_t k);
int nvllm_gemm_bf16(void* y, const void* x, const void* w, int32_t m, int32_t n, int32_t k,
                    int64_t ldx, int64_t ldw, int64_t ldy, int32_t out_dtype, void* workspace,
                    size_t workspace_bytes, void* stream);
/* MergedColumnParallelLinear::forward followed by SiluAndMul::forward (linear.rs:437-439, activation.rs:46-63) in one
 * call: out[M, inter] = silu(x . Wgate^T) * (x . Wup^T), w_gate_up = [Wgate | Wup] packed row-wise ([2*inter, K]).
 * When inter %% 128 == 0 the activation runs in the GEMM epilogue (the B tile holds BN/2 gate rows + the matching BN/2
 * up rows; with CTA pairs one CTA stages the gate rows, its peer the up rows) and the [M, 2*inter] intermediate never
 * reaches HBM; other sizes use `workspace` (>= M*2*inter*2 bytes) for it.
 * Same rounding points as nvllm_gemm_bf16 + nvllm_silu_mul (bit-identical results). */
int nvllm_gemm_silu_mul_bf16(void* out, const void* x, const void* w_gate_up, int32_t m, int32_t inter, int32_t k,
                             int64_t ldx, int64_t ldw, int64_t ldout, void* workspace, size_t workspace_bytes,
                             void* stream);
/* QKVParallelLinear::forward (linear.rs:354-356) + the q/k norm, RotaryEmbedding::forward and Attention::store_kv_cache
 * that follow it (qwen3.rs:219-229, attention.rs:150-174) in ONE kernel: qkv[M, (H+2Hkv)*128] = x . Wqkv^T, each 128-wide
 * output tile being one head whose epilogue thread applies the same math (and rounding points) as
 * nvllm_qknorm_rope_store before writing the qkv row and, for k / v heads, the paged cache.  head_dim 128 only.
 * k_cache may be NULL (no store).  Equivalent to nvllm_gemm_bf16 + nvllm_qknorm_rope_store. */
int nvllm_gemm_qkv_rope_store(void* qkv, const void* x, const void* w_qkv, int32_t m, int32_t k, int64_t ldx, int64_t ldw,
                              const int64_t* positions, const float* cos_sin, const void* q_norm_w,
                              const void* k_norm_w, float eps, int32_t num_heads, int32_t num_kv_heads,
                              int32_t head_dim, int32_t max_position, void* k_cache, void* v_cache,
                              const int32_t* slot_mapping, const nvllm_kv_layout* layout, void* stream);
/* Plain CUDA-core version of the same contract; test infrastructure for the kernel above (used to
 * check large shapes on the device where the CPU oracle is too slow).  Never on the product path. */
int nvllm_gemm_bf16_simt(void* y, const void* x, const void* w, int32_t m, int32_t n, int32_t k,
                         int64_t ldx, int64_t ldw, int64_t ldy, int32_t out_dtype, void* stream);

/* K6 — replaces RMSNorm::forward_simple / forward_with_residual (src/layers/layernorm.rs:58-97)
 * and the residual adds of Qwen3Layer::forward (src/models/qwen3.rs:382,389).
 *   residual == NULL : out = rmsnorm(x) * w
 *   residual != NULL : residual <- bf16(residual + x) (in place);  out = rmsnorm(residual) * w
 * f32 math: v * w / sqrt(mean(v^2) + eps).  x/out: [num_tokens, hidden], contiguous rows. */
int nvllm_add_rmsnorm(void* out, const void* x, void* residual, const void* weight,
                      int32_t num_tokens, int32_t hidden, float eps, int32_t dtype, void* stream);

/* K7 (+K1 fused) — replaces RotaryEmbedding::forward / apply_rotary_emb_single
 * (src/layers/rotary_embedding.rs:23-48,136-158), the per-head q_norm/k_norm of the named Qwen3
 * architecture (absent in the reference, SURVEY.md §8 note N1) and, when k_cache != NULL, the
 * store_kv_cache call that follows (attention.rs:78-81).
 * In place on the packed qkv buffer [num_tokens, (H + 2*Hkv) * D] (q | k | v, linear.rs:331-340):
 *   q_h <- rope(rmsnorm_D(q_h) * q_norm_w),  k_h <- rope(rmsnorm_D(k_h) * k_norm_w)
 *   rope: half-split (NeoX) rotation [x1*c - x2*s | x2*c + x1*s], c,s = cos,sin(pos * theta^(-2j/D))
 *   taken from the fp32 table cos_sin: [max_position, D] = (cos[0..D/2) | sin[0..D/2)).
 * q_norm_w / k_norm_w may be NULL (no qk-norm: the reference's behaviour).  positions: i64. */
int nvllm_qknorm_rope_store(void* qkv, const int64_t* positions, const float* cos_sin,
                            const void* q_norm_w, const void* k_norm_w, float eps,
                            int32_t num_tokens, int32_t num_heads, int32_t num_kv_heads,
                            int32_t head_dim, int32_t max_position, void* k_cache, void* v_cache,
                            const int32_t* slot_mapping, const nvllm_kv_layout* layout,
                            int32_t dtype, void* stream);
/* Fills cos_sin [max_position, D] fp32 exactly as the oracle does (f64 inv_freq and angle, rounded
 * to f32) — rotary_embedding.rs:82-110 with the table kept in f32 (note N1). */
int nvllm_rope_table(float* cos_sin, int32_t max_position, int32_t head_dim, double theta,
                     void* stream);

/* K8 — replaces SiluAndMul::forward (src/layers/activation.rs:46-63).
 * out[t, j] = silu(x[t, j]) * x[t, inter + j],  x: [num_tokens, 2*inter], out: [num_tokens, inter]. */
int nvllm_silu_mul(void* out, const void* x, int32_t num_tokens, int32_t inter, int32_t dtype,
                   void* stream);

/* K9 / K10 — replaces VocabParallelEmbedding::forward (src/layers/embed_head.rs:77-97) and
 * ParallelLMHead::extract_last_tokens (embed_head.rs:272-289).
 * gather rows: out[i,:] = table[ids[i],:]  (ids i64).  With `row_index` semantic = last token of each
 * sequence use nvllm_gather_last_tokens: out[i,:] = x[cu_seqlens_q[i+1]-1, :]. */
int nvllm_embedding_gather(void* out, const void* table, const int64_t* ids, int32_t num_tokens,
                           int32_t hidden, int32_t vocab, int32_t dtype, void* stream);
int nvllm_gather_last_tokens(void* out, const void* x, const int32_t* cu_seqlens_q, int32_t batch,
                             int32_t hidden, int32_t dtype, void* stream);

/* K11 — replaces Sampler::greedy_sample (src/layers/sampler.rs:109-112): argmax over the vocab,
 * first maximum wins.  logits: [batch, vocab] (bf16 or f32, leading dimension ld).  Writes the token
 * id (+ vocab_offset, for vocab-sharded heads) to out_ids[b] (i64) and, if out_max != NULL, the
 * maximum as f32 to out_max[b] (used for the cross-rank combine, embed_head.rs:321-336). */
int nvllm_argmax(int64_t* out_ids, float* out_max, const void* logits, int32_t batch, int32_t vocab,
                 int64_t ld, int64_t vocab_offset, int32_t logits_dtype, void* stream);
/* Cross-rank combine of per-shard (max, id) pairs gathered as [tp, batch]: picks the largest max,
 * lowest shard on ties (== first maximum over the concatenated vocab). */
int nvllm_argmax_combine(int64_t* out_ids, const float* all_max, const int64_t* all_ids,
                         int32_t tp, int32_t batch, void* stream);
/* Puts the all-gathered vocab shards of a tensor-parallel LM head back into full rows — the gather the reference leaves as
 * a TODO (src/layers/embed_head.rs:321-336): out[b, r*vshard + j] = gathered[r, b, j] (f32).  With it every rank holds the
 * full [rows, vocab] logits and runs the same deterministic nvllm_sample: temperature / top-k / top-p under TP. */
int nvllm_assemble_vocab_shards(float* out, const float* gathered, int32_t tp, int32_t rows, int32_t vshard,
                                void* stream);

/* K12 — replaces Sampler::forward / batch_sample / sample_single (src/layers/sampler.rs:26-106,221-254) with its
 * helpers apply_top_k (:115-149), apply_top_p (:152-191), multinomial_sample (:194-203) and sample_gumbel (:205-219)
 * for a whole batch in one launch (one CTA per row; selections instead of the reference's two full sorts).
 * logits: f32 [batch, vocab] (leading dimension ld).  Per row b (all arrays are DEVICE pointers):
 *   temperatures[b] == 0  -> greedy argmax, first maximum wins (sampler.rs:79-81)
 *   otherwise s = logits / temperature; top_k[b] in (0, vocab) keeps the first k entries of the stable descending
 *   order (0 or NULL array = None); top_p[b] in (0, 1) keeps the shortest descending prefix whose softmax mass
 *   reaches p (other values or NULL array = None); token = argmax(s + gumbel), gumbel = -log(-log(u)).
 * The reference draws u from an unseeded RNG; here u[j] = ((mix64(row_seeds[b] + (j+1)*0x9E3779B97F4A7C15) >> 41)
 * + 0.5) * 2^-23 with mix64 = the splitmix64 finaliser — the same contract as oracle/layers.py, so a (seed, logits)
 * pair always yields the same token.  kept_count (i32 [batch]) / kept_mask (u8 [batch, vocab]) are optional
 * outputs describing the surviving set (parity tests); pass NULL in production. */
int nvllm_sample(int64_t* out_ids, const float* logits, int32_t batch, int32_t vocab, int64_t ld,
                 const float* temperatures, const int32_t* top_k, const float* top_p, const uint64_t* row_seeds,
                 int32_t* kept_count, uint8_t* kept_mask, void* stream);

/* C1 — the all-reduce the reference leaves as a TODO after o_proj / down_proj
 * (src/layers/linear.rs:236).  One communicator per process (one process per GPU).
 * nvllm_comm_unique_id writes NVLLM_UNIQUE_ID_BYTES bytes on rank 0; the caller broadcasts them
 * (torch.distributed / any transport) and every rank calls nvllm_comm_init. */
#define NVLLM_UNIQUE_ID_BYTES 128
typedef struct nvllm_comm nvllm_comm;
int nvllm_comm_unique_id(void* h_id_out);
int nvllm_comm_init(nvllm_comm** comm, const void* h_id, int32_t rank, int32_t world);
/* Collective tear-down: closes the peer mappings, meets the other ranks once more (an all-reduce of a scratch word, with a
 * host-side deadline of 6 x NVLLM_COMM_TIMEOUT_MS) so nobody frees a region a peer still has mapped, then frees.  When the
 * communicator has already reported NVLLM_ERR_COMM, or the meeting times out, it does NOT wait: the region is leaked, the
 * NCCL communicator aborted, and NVLLM_ERR_COMM returned — a dead peer can never keep this call from returning. */
int nvllm_comm_destroy(nvllm_comm* comm);
/* In-place sum over ranks of buf[count] (bf16 or f32), enqueued on `stream` (graph-capturable). */
int nvllm_allreduce(nvllm_comm* comm, void* buf, size_t count, int32_t dtype, void* stream);
int nvllm_allgather(nvllm_comm* comm, const void* send, void* recv, size_t bytes_per_rank,
                    void* stream);
/* Symmetric (peer-mapped) memory: a collective call — every rank allocates `bytes`, the CUDA IPC handles are exchanged
 * through the communicator and each peer's region is mapped over NVLink.  *local_base = this rank's region. */
int nvllm_comm_symm_alloc(nvllm_comm* comm, size_t bytes, void** local_base);
/* C1 fused with what always follows it (RowParallelLinear all-reduce TODO linear.rs:236 + the residual add and RMSNorm
 * of Qwen3Layer::forward, qwen3.rs:382-385,389): one two-shot kernel over NVLink peer memory —
 *   residual <- bf16(residual + bf16(sum over ranks of x));  out <- rmsnorm(residual) * weight.
 * `out` is complete on EVERY rank afterwards; `residual` is updated only for the rows this rank owns
 * (rows [rank*R, (rank+1)*R), R = ceil(num_tokens / world)) — it is a row-sharded stream whose only readers are
 * subsequent calls of this function with the same num_tokens.
 * out, x, residual: [num_tokens, hidden] bf16 buffers at identical offsets inside every rank's symmetric region
 * (required by the flag protocol; the push protocol only needs 16-byte alignment).
 * Bit-identical results on all ranks (each row is reduced once, by its owner, in fixed rank order). */
int nvllm_allreduce_add_rmsnorm(nvllm_comm* comm, void* out, void* x, void* residual, const void* weight,
                                int32_t num_tokens, int32_t hidden, float eps, int32_t dtype, void* stream);
/* Two protocols sit behind the call above.  Messages up to NVLLM_AR_LL_MAX_BYTES (8 MiB, decode batches) take the
 * low-latency PUSH protocol: no flags, no fences — every 32-bit word of a not-yet-written buffer holds 0x80008000 (two bf16
 * -0.0; senders map that one pattern to +0.0), partial rows are stored straight into the owner's stage buffer, the owner
 * polls the data itself, reduces, and stores the normed rows into every peer's gather buffer, which the peer polls and
 * copies to `out`; three buffer pairs rotate per call (NVLLM_AR_LL=0 disables it).  Larger messages (prefill) take the
 * flag protocol with peer loads.  nvllm_comm_ll_applies: 1 when a [num_tokens, hidden] call takes the push protocol. */
int nvllm_comm_ll_applies(const nvllm_comm* comm, int32_t num_tokens, int32_t hidden);
/* RowParallelLinear::forward (linear.rs:228-239) whose all-reduce is the call below: y = x . w^T (bf16, m <= 512,
 * n == ldy == hidden, nvllm_comm_ll_applies(comm, m, n)); the GEMM's epilogue stores the rows another rank reduces straight
 * into that rank's stage buffer (they never touch y), so the transfer runs tile by tile under the GEMM's tail. */
int nvllm_gemm_bf16_push(nvllm_comm* comm, void* y, const void* x, const void* w, int32_t m, int32_t n, int32_t k,
                         int64_t ldx, int64_t ldw, int64_t ldy, void* stream);
/* nvllm_allreduce_add_rmsnorm for an x produced by nvllm_gemm_bf16_push (same num_tokens / hidden): skips the push. */
int nvllm_allreduce_add_rmsnorm_pushed(nvllm_comm* comm, void* out, void* x, void* residual, const void* weight,
                                       int32_t num_tokens, int32_t hidden, float eps, int32_t dtype, void* stream);

/* 0, or NVLLM_ERR_COMM once a fused kernel of this communicator gave up waiting for a peer (a rank died, or stalled for
 * longer than NVLLM_COMM_TIMEOUT_MS, default 10 min).  The kernel neither hangs nor traps: it raises this host-visible
 * flag, stops waiting in later calls and runs to completion; results from then on are undefined.  Sticky. */
int nvllm_comm_error(const nvllm_comm* comm);

/* Deterministic seeded weight init shared bit-for-bit with the CPU oracle (oracle/weights.py):
 * element i of tensor `tensor_id` = bf16( (sum of the four 16-bit fields of
 * splitmix64(seed ^ tensor_id<<40, i)  - 131070) * std_scale ).  Integer arithmetic + one exact
 * int->f32 multiply, so host and device agree bit-exactly. */
int nvllm_init_weight(void* w, int64_t count, uint64_t seed, uint64_t tensor_id, float std,
                      int32_t dtype, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Engine-level C ABI: the C++ host runtime that mirrors the reference's Rust types one-to-one
 * (Config, SamplingParams, Sequence, BlockManager, Scheduler, ModelRunner, Qwen3Model, LLMEngine).
 * Declared in nvllm_engine.h; kept separate so a Rust caller that only wants the kernels binds
 * this header alone.
 * ---------------------------------------------------------------------------------------------- */

#ifdef __cplusplus
} /* extern "C" */
#endif

#endif /* NVLLM_H_ */
