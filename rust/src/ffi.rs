//! Safe(r) wrappers with the reference's method surface over the generated `extern "C"` declarations (ffi_gen.rs): one
//! type per layer of nano-vllm-rs whose `forward` the B200 library replaces, plus RAII handles for the
//! host-runtime level (§2).  Tensors stay owned by Rust (candle); the wrappers borrow raw device pointers for the duration
//! of the enqueued work, so every call is `unsafe` and takes the CUDA stream explicitly.
//! COMPILE-UNVERIFIED (no Rust toolchain in the build image); the same ABI is exercised through ctypes by
//! nano-vllm-rs_b200/_ffi.py and every test under tests/.
use std::ffi::{c_void, CStr, CString};

pub use crate::ffi_gen::*; // every extern "C" declaration + #[repr(C)] struct of include/nvllm.h and nvllm_engine.h

pub const NVLLM_BF16: i32 = 0;
pub const NVLLM_F32: i32 = 1;

/// Non-zero return -> anyhow::Error carrying nvllm_last_error() (the reference's `map_err(anyhow!)` convention,
/// model_runner.rs:123-124).
pub fn check(rc: i32) -> anyhow::Result<()> {
    if rc == 0 {
        return Ok(());
    }
    let msg = unsafe { CStr::from_ptr(nvllm_last_error()) }.to_string_lossy().into_owned();
    anyhow::bail!("nvllm error {rc}: {msg}")
}

/// The `Context` fields Attention::forward reads (utils/context.rs:11-35), as device pointers.
#[derive(Clone, Copy)]
pub struct AttnContext {
    pub is_prefill: bool,
    pub slot_mapping: *const i32,
    pub cu_seqlens_q: *const i32,
    pub cu_seqlens_k: *const i32,
    pub max_seqlen_q: i32,
    pub max_seqlen_k: i32,
    pub context_lens: *const i32,
    pub block_tables: *const i32, // null: no prefix cache (prefill) — attention.rs:84-90
    pub max_blocks: i32,
    pub num_seqs: i32,
    pub num_tokens: i32,
}

/// `Attention` (src/layers/attention.rs:27-99): store_kv_cache + the three flash-attention branches of `forward`.
pub struct Attention {
    pub num_heads: i32,
    pub num_kv_heads: i32,
    pub head_dim: i32,
    pub scale: f32, // 1 / sqrt(head_dim), attention.rs:45
    pub layout: nvllm_kv_layout,
}

impl Attention {
    /// The reference's own cache allocation `[num_blocks, block_size, Hkv, D]` (model_runner.rs:364-396).
    pub fn reference_layout(block_size: i32, num_kv_heads: i32, head_dim: i32) -> nvllm_kv_layout {
        nvllm_kv_layout {
            block_size, num_kv_heads, head_dim, reserved: 0,
            block_stride: block_size as i64 * num_kv_heads as i64 * head_dim as i64,
            token_stride: num_kv_heads as i64 * head_dim as i64,
            head_stride: head_dim as i64,
        }
    }

    /// attention.rs:65-99.  q / k / v: the packed qkv rows of this step (token strides in elements); `out`: [tokens, H*D].
    #[allow(clippy::too_many_arguments)]
    pub unsafe fn forward(&self, out: *mut c_void, q: *const c_void, k: *const c_void, v: *const c_void, qkv_stride: i64,
                          k_cache: *mut c_void, v_cache: *mut c_void, ctx: &AttnContext, workspace: *mut c_void,
                          workspace_bytes: usize, stream: *mut c_void) -> anyhow::Result<()> {
        let hd = self.num_heads as i64 * self.head_dim as i64;
        // attention.rs:78-81  store_kv_cache (slot < 0 = padding, skipped)
        check(nvllm_store_kvcache(k, v, k_cache, v_cache, ctx.slot_mapping, ctx.num_tokens, qkv_stride, &self.layout,
                                  NVLLM_BF16, stream))?;
        if ctx.is_prefill {
            // attention.rs:84-90  flash_attention_varlen / flash_attention_varlen_with_cache (bottom-right causal mask)
            let paged = !ctx.block_tables.is_null();
            check(nvllm_varlen_prefill_attention(
                out, q, k, v, ctx.cu_seqlens_q, ctx.cu_seqlens_k, ctx.num_seqs, ctx.num_tokens, ctx.num_tokens,
                ctx.max_seqlen_q, ctx.max_seqlen_k, self.num_heads, self.num_kv_heads, self.head_dim, qkv_stride,
                qkv_stride, hd, if paged { k_cache as *const c_void } else { std::ptr::null() },
                if paged { v_cache as *const c_void } else { std::ptr::null() }, ctx.block_tables, ctx.max_blocks,
                &self.layout, self.scale, NVLLM_BF16, stream))
        } else {
            // attention.rs:92-95  flash_attention_decode (num_splits 0 = the library's choice)
            check(nvllm_paged_decode_attention(out, q, k_cache, v_cache, ctx.block_tables, ctx.context_lens, ctx.num_seqs,
                                               self.num_heads, ctx.max_blocks, qkv_stride, hd, &self.layout, self.scale,
                                               workspace, workspace_bytes, 0, NVLLM_BF16, stream))
        }
    }

    /// The decode branch of `forward` for a host that has no device allocator to give it a workspace (a plugin hook that
    /// only sees one op's tensors, a pool sealed after start-up): the split-K partials live in an 8 MiB array inside
    /// libnvllm.so.  No counterpart in the reference.  Errors with NVLLM_ERR_WORKSPACE when the batch needs more than
    /// `nvllm_paged_decode_static_scratch_bytes()`; call it once during set-up (before any stream capture), and never from
    /// two streams at once — all calls on a GPU share the array.  The KV store of this step must already have happened.
    #[allow(clippy::too_many_arguments)]
    pub unsafe fn forward_decode_no_alloc(&self, out: *mut c_void, q: *const c_void, qkv_stride: i64, k_cache: *const c_void,
                                          v_cache: *const c_void, ctx: &AttnContext, stream: *mut c_void) -> anyhow::Result<()> {
        let hd = self.num_heads as i64 * self.head_dim as i64;
        check(nvllm_paged_decode_attention_static(out, q, k_cache, v_cache, ctx.block_tables, ctx.context_lens, ctx.num_seqs,
                                                  self.num_heads, ctx.max_blocks, qkv_stride, hd, &self.layout, self.scale,
                                                  0, NVLLM_BF16, stream))
    }

    /// Bytes `forward` may need in `workspace` for a decode batch (split-K partials).
    pub fn decode_workspace_bytes(&self, max_batch: i32) -> usize {
        unsafe { nvllm_paged_decode_workspace_bytes(max_batch, self.num_heads, self.head_dim, 0) }
    }
}

/// `LinearBase::forward` for every linear of the model (src/layers/linear.rs:228-239,354-356,437-439): y = x . W^T, weight
/// `[out, in]` row-major bf16 = this rank's shard.
pub struct Linear {
    pub weight: *const c_void,
    pub in_features: i32,
    pub out_features: i32,
}

impl Linear {
    pub unsafe fn forward(&self, y: *mut c_void, x: *const c_void, num_tokens: i32, stream: *mut c_void) -> anyhow::Result<()> {
        let (n, k) = (self.out_features, self.in_features);
        check(nvllm_gemm_bf16(y, x, self.weight, num_tokens, n, k, k as i64, k as i64, n as i64, NVLLM_BF16,
                              std::ptr::null_mut(), 0, stream))
    }

    /// MergedColumnParallelLinear::forward + SiluAndMul::forward (linear.rs:437-439 + activation.rs:46-63, i.e.
    /// Qwen3MLP::forward qwen3.rs:307-310) in one launch; `self.out_features` = 2 * intermediate (gate rows, then up rows).
    pub unsafe fn forward_silu_mul(&self, act: *mut c_void, x: *const c_void, num_tokens: i32, workspace: *mut c_void,
                                   workspace_bytes: usize, stream: *mut c_void) -> anyhow::Result<()> {
        let (inter, k) = (self.out_features / 2, self.in_features);
        check(nvllm_gemm_silu_mul_bf16(act, x, self.weight, num_tokens, inter, k, k as i64, k as i64, inter as i64,
                                       workspace, workspace_bytes, stream))
    }
}

/// `RMSNorm::forward` (src/layers/layernorm.rs:50-97) fused with the residual add that precedes it in Qwen3Layer::forward
/// (qwen3.rs:382,389): residual <- residual + x, out <- rmsnorm(residual) * weight (null `residual`: plain norm of x).
pub struct RmsNorm {
    pub weight: *const c_void,
    pub hidden: i32,
    pub eps: f32,
}

impl RmsNorm {
    pub unsafe fn forward(&self, out: *mut c_void, x: *const c_void, residual: *mut c_void, num_tokens: i32,
                          stream: *mut c_void) -> anyhow::Result<()> {
        check(nvllm_add_rmsnorm(out, x, residual, self.weight, num_tokens, self.hidden, self.eps, NVLLM_BF16, stream))
    }

    /// Under tensor parallelism: the all-reduce RowParallelLinear::forward leaves as a TODO (linear.rs:236) + the add + the
    /// norm in one kernel over NVLink peer memory.  out / x / residual live in the communicator's symmetric region.
    pub unsafe fn forward_allreduce(&self, comm: *mut nvllm_comm, out: *mut c_void, x: *mut c_void, residual: *mut c_void,
                                    num_tokens: i32, stream: *mut c_void) -> anyhow::Result<()> {
        check(nvllm_allreduce_add_rmsnorm(comm, out, x, residual, self.weight, num_tokens, self.hidden, self.eps, NVLLM_BF16,
                                          stream))?;
        check(nvllm_comm_error(comm)) // sticky, host-visible: a peer that never arrived in an EARLIER call
    }
}

/// `RotaryEmbedding::forward` (src/layers/rotary_embedding.rs:136-158) applied in place to the packed qkv rows, with the
/// per-head q/k RMSNorm of Qwen3 before it and `Attention::store_kv_cache` after it (qwen3.rs:219-229) in the same launch.
pub struct RotaryEmbedding {
    pub cos_sin: *const f32, // [max_position, head_dim] f32: head_dim/2 cosines then head_dim/2 sines per position
    pub max_position: i32,
    pub head_dim: i32,
}

impl RotaryEmbedding {
    /// Fills `cos_sin` (rotary_embedding.rs:75-118; kept in f32 — the reference's cast to the model dtype breaks bf16).
    pub unsafe fn build_table(cos_sin: *mut f32, max_position: i32, head_dim: i32, theta: f64, stream: *mut c_void) -> anyhow::Result<()> {
        check(nvllm_rope_table(cos_sin, max_position, head_dim, theta, stream))
    }

    #[allow(clippy::too_many_arguments)]
    pub unsafe fn forward_qkv(&self, qkv: *mut c_void, positions: *const i64, q_norm_w: *const c_void, k_norm_w: *const c_void,
                              eps: f32, num_tokens: i32, num_heads: i32, num_kv_heads: i32, k_cache: *mut c_void,
                              v_cache: *mut c_void, slot_mapping: *const i32, layout: &nvllm_kv_layout, stream: *mut c_void)
                              -> anyhow::Result<()> {
        check(nvllm_qknorm_rope_store(qkv, positions, self.cos_sin, q_norm_w, k_norm_w, eps, num_tokens, num_heads, num_kv_heads,
                                      self.head_dim, self.max_position, k_cache, v_cache, slot_mapping, layout, NVLLM_BF16, stream))
    }
}

/// `VocabParallelEmbedding::forward` (src/layers/embed_head.rs:77-97; the table is replicated, so no all-reduce) and
/// `ParallelLMHead::forward` (:250-306): last-token gather for prefill batches, then logits = hidden . W^T in f32.
pub struct EmbedHead {
    pub embed: *const c_void,
    pub lm_head: *const c_void, // this rank's vocabulary shard [vocab_shard, hidden]
    pub vocab: i32,
    pub vocab_shard: i32,
    pub hidden: i32,
}

impl EmbedHead {
    pub unsafe fn embed(&self, out: *mut c_void, ids: *const i64, num_tokens: i32, stream: *mut c_void) -> anyhow::Result<()> {
        check(nvllm_embedding_gather(out, self.embed, ids, num_tokens, self.hidden, self.vocab, NVLLM_BF16, stream))
    }

    /// `last_hidden`: scratch [num_seqs, hidden], used when `cu_seqlens_q` is non-null (prefill: embed_head.rs:272-289).
    pub unsafe fn logits(&self, logits: *mut f32, hidden_states: *const c_void, cu_seqlens_q: *const i32, num_seqs: i32,
                         last_hidden: *mut c_void, stream: *mut c_void) -> anyhow::Result<()> {
        let mut x = hidden_states;
        if !cu_seqlens_q.is_null() {
            check(nvllm_gather_last_tokens(last_hidden, hidden_states, cu_seqlens_q, num_seqs, self.hidden, NVLLM_BF16, stream))?;
            x = last_hidden as *const c_void;
        }
        let (n, k) = (self.vocab_shard, self.hidden);
        check(nvllm_gemm_bf16(logits as *mut c_void, x, self.lm_head, num_seqs, n, k, k as i64, k as i64, n as i64, NVLLM_F32,
                              std::ptr::null_mut(), 0, stream))
    }
}

/// What `ModelRunner::sample_tokens` -> `Sampler::batch_sample` (src/engine/model_runner.rs:131-156,
/// src/layers/sampler.rs:221-254) becomes: the per-sequence parameters are device arrays, one launch samples the batch.
pub struct DeviceSampler;

impl DeviceSampler {
    /// `Sampler::greedy_sample` (sampler.rs:109-112): first maximum wins; `vocab_offset` = rank * vocab_shard under TP.
    pub unsafe fn greedy(out_ids: *mut i64, out_max: *mut f32, logits: *const f32, batch: i32, vocab: i32, vocab_offset: i64,
                         stream: *mut c_void) -> anyhow::Result<()> {
        check(nvllm_argmax(out_ids, out_max, logits as *const c_void, batch, vocab, vocab as i64, vocab_offset, NVLLM_F32, stream))
    }

    /// logits: f32 [batch, vocab] on the device; temperatures / top_k / top_p / row_seeds: device arrays of `batch`
    /// entries (top_k / top_p may be null = None for every row); out_ids: i64 [batch] on the device.
    #[allow(clippy::too_many_arguments)]
    pub unsafe fn forward(out_ids: *mut i64, logits: *const f32, batch: i32, vocab: i32, temperatures: *const f32,
                          top_k: *const i32, top_p: *const f32, row_seeds: *const u64, stream: *mut c_void) -> anyhow::Result<()> {
        check(nvllm_sample(out_ids, logits, batch, vocab, vocab as i64, temperatures, top_k, top_p, row_seeds,
                           std::ptr::null_mut(), std::ptr::null_mut(), stream))
    }
}

// ------------------------------------------------------------------------------------------------ host-runtime level
/// `Config` (src/config.rs) for a named architecture ("qwen3-0.6b", "qwen3-8b", "qwen3-32b"); fields are public.
pub fn config_for(arch: &str) -> anyhow::Result<nvllm_config> {
    let name = CString::new(arch)?;
    let mut cfg: nvllm_config = unsafe { std::mem::zeroed() };
    check(unsafe { nvllm_config_init(&mut cfg, name.as_ptr()) })?;
    Ok(cfg)
}

/// `LLMEngine` (src/engine/llm_engine.rs:71-357, synchronous): owns a Scheduler (block manager, prefix cache) and a
/// ModelRunner (weights, KV cache, decode graphs).  `step` is `LLMEngine::step` (:155-197).
pub struct Engine {
    raw: *mut nvllm_engine,
}

impl Engine {
    pub fn new(cfg: &nvllm_config, comm: *mut nvllm_comm) -> anyhow::Result<Self> {
        check(unsafe { nvllm_config_validate(cfg) })?;
        let mut raw = std::ptr::null_mut();
        check(unsafe { nvllm_engine_create(&mut raw, cfg, comm) })?;
        Ok(Engine { raw })
    }

    /// Seeded weights (tests, benchmarks); real checkpoints go through `load_weight` with the HF tensor names.
    pub fn init_weights(&mut self, seed: u64, std: f32) -> anyhow::Result<()> {
        check(unsafe { nvllm_runner_init_weights(nvllm_engine_runner(self.raw), seed, std) })
    }

    pub fn load_weight(&mut self, hf_name: &str, bf16_rows: &[u16], rows: i64, cols: i64) -> anyhow::Result<()> {
        anyhow::ensure!(bf16_rows.len() as i64 == rows * cols, "load_weight: {} elements for {rows} x {cols}", bf16_rows.len());
        let name = CString::new(hf_name)?;
        check(unsafe { nvllm_runner_load_weight(nvllm_engine_runner(self.raw), name.as_ptr(), bf16_rows.as_ptr() as *const c_void, rows, cols) })
    }

    /// `LLMEngine::add_request` (llm_engine.rs:131-152): returns the sequence id; fails for prompts the configuration
    /// can never serve (longer than max_model_len / max_num_batched_tokens).
    pub fn add_request(&mut self, prompt: &[i64], sp: &nvllm_sampling_params) -> anyhow::Result<i64> {
        let id = unsafe { nvllm_engine_add_request(self.raw, prompt.as_ptr(), prompt.len() as i32, sp) };
        if id < 0 {
            check(id as i32)?;
        }
        Ok(id)
    }

    /// One schedule -> execute -> postprocess round; returns (is_prefill, num_seqs, num_tokens) of the batch it ran.
    pub fn step(&mut self) -> anyhow::Result<(bool, i32, i32)> {
        let (mut p, mut s, mut t) = (0i32, 0i32, 0i32);
        check(unsafe { nvllm_engine_step(self.raw, &mut p, &mut s, &mut t) })?;
        Ok((p != 0, s, t))
    }

    pub fn is_finished(&self) -> bool {
        unsafe { nvllm_engine_is_finished(self.raw) != 0 }
    }

    /// Tokens of a sequence (prompt + completion) and whether it has finished; releases it from the scheduler when done.
    pub fn take_if_finished(&mut self, seq_id: i64, max_len: usize) -> anyhow::Result<Option<Vec<i64>>> {
        let sched = unsafe { nvllm_engine_scheduler(self.raw) };
        let mut toks = vec![0i64; max_len];
        let (mut n_prompt, mut status) = (0i32, 0i32);
        let n = unsafe { nvllm_scheduler_sequence(sched, seq_id, toks.as_mut_ptr(), max_len as i32, &mut n_prompt, &mut status) };
        if n < 0 {
            check(n)?;
        }
        if status != 2 {
            return Ok(None); // 0 Waiting, 1 Running, 2 Finished, 3 Preempted (SequenceStatus, sequence.rs:10-16)
        }
        toks.truncate(n as usize);
        check(unsafe { nvllm_scheduler_release_sequence(sched, seq_id) })?;
        Ok(Some(toks))
    }
}

impl Drop for Engine {
    fn drop(&mut self) {
        unsafe { nvllm_engine_destroy(self.raw) };
    }
}
