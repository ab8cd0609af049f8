//! `extern "C"` declarations for include/nvllm.h + the safe wrappers with the reference's method signatures.
//! COMPILE-UNVERIFIED (no Rust toolchain in the build image); the same ABI is exercised through ctypes by
//! nano-vllm-rs_b200/_ffi.py and every test under tests/.
#![allow(non_camel_case_types)]
use std::ffi::{c_char, c_void, CStr};

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct nvllm_kv_layout {
    pub block_size: i32,
    pub num_kv_heads: i32,
    pub head_dim: i32,
    pub reserved: i32,
    pub block_stride: i64,
    pub token_stride: i64,
    pub head_stride: i64,
}

pub const NVLLM_BF16: i32 = 0;
pub const NVLLM_F32: i32 = 1;

extern "C" {
    pub fn nvllm_version() -> i32;
    pub fn nvllm_last_error() -> *const c_char;
    pub fn nvllm_store_kvcache(k: *const c_void, v: *const c_void, k_cache: *mut c_void, v_cache: *mut c_void,
                               slot_mapping: *const i32, num_tokens: i32, kv_token_stride: i64,
                               layout: *const nvllm_kv_layout, dtype: i32, stream: *mut c_void) -> i32;
    pub fn nvllm_paged_decode_workspace_bytes(batch: i32, num_heads: i32, head_dim: i32, num_splits: i32) -> usize;
    pub fn nvllm_paged_decode_attention(out: *mut c_void, q: *const c_void, k_cache: *const c_void, v_cache: *const c_void,
                                        block_tables: *const i32, context_lens: *const i32, batch: i32, num_heads: i32,
                                        max_blocks_per_seq: i32, q_token_stride: i64, out_token_stride: i64,
                                        layout: *const nvllm_kv_layout, scale: f32, workspace: *mut c_void,
                                        workspace_bytes: usize, num_splits: i32, dtype: i32, stream: *mut c_void) -> i32;
    pub fn nvllm_varlen_prefill_attention(out: *mut c_void, q: *const c_void, k: *const c_void, v: *const c_void,
                                          cu_seqlens_q: *const i32, cu_seqlens_k: *const i32, batch: i32, total_q: i32,
                                          total_k: i32, max_seqlen_q: i32, max_seqlen_k: i32, num_heads: i32,
                                          num_kv_heads: i32, head_dim: i32, q_token_stride: i64, kv_token_stride: i64,
                                          out_token_stride: i64, k_cache: *const c_void, v_cache: *const c_void,
                                          block_tables: *const i32, max_blocks_per_seq: i32,
                                          layout: *const nvllm_kv_layout, scale: f32, dtype: i32, stream: *mut c_void) -> i32;
    pub fn nvllm_gemm_bf16(y: *mut c_void, x: *const c_void, w: *const c_void, m: i32, n: i32, k: i32, ldx: i64, ldw: i64,
                           ldy: i64, out_dtype: i32, workspace: *mut c_void, workspace_bytes: usize, stream: *mut c_void) -> i32;
    pub fn nvllm_add_rmsnorm(out: *mut c_void, x: *const c_void, residual: *mut c_void, weight: *const c_void,
                             num_tokens: i32, hidden: i32, eps: f32, dtype: i32, stream: *mut c_void) -> i32;
    pub fn nvllm_qknorm_rope_store(qkv: *mut c_void, positions: *const i64, cos_sin: *const f32, q_norm_w: *const c_void,
                                   k_norm_w: *const c_void, eps: f32, num_tokens: i32, num_heads: i32, num_kv_heads: i32,
                                   head_dim: i32, max_position: i32, k_cache: *mut c_void, v_cache: *mut c_void,
                                   slot_mapping: *const i32, layout: *const nvllm_kv_layout, dtype: i32,
                                   stream: *mut c_void) -> i32;
    pub fn nvllm_silu_mul(out: *mut c_void, x: *const c_void, num_tokens: i32, inter: i32, dtype: i32, stream: *mut c_void) -> i32;
    pub fn nvllm_argmax(out_ids: *mut i64, out_max: *mut f32, logits: *const c_void, batch: i32, vocab: i32, ld: i64,
                        vocab_offset: i64, logits_dtype: i32, stream: *mut c_void) -> i32;
    // ... the remaining entry points of nvllm.h / nvllm_engine.h follow the same pattern.
}

/// Non-zero return -> anyhow::Error carrying nvllm_last_error() (the reference's `map_err(anyhow!)` convention,
/// model_runner.rs:123-124).
pub fn check(rc: i32) -> anyhow::Result<()> {
    if rc == 0 {
        return Ok(());
    }
    let msg = unsafe { CStr::from_ptr(nvllm_last_error()) }.to_string_lossy().into_owned();
    anyhow::bail!("nvllm error {rc}: {msg}")
}

/// What `Attention::forward` (src/layers/attention.rs:65-99) becomes in the decode branch: the candle tensors stay
/// owned by Rust, the kernels borrow their device pointers for the duration of the enqueued work.
pub struct DecodeAttention {
    pub num_heads: i32,
    pub layout: nvllm_kv_layout,
    pub scale: f32, // 1 / sqrt(head_dim), attention.rs:45
}

impl DecodeAttention {
    /// q: [batch, H, D] bf16; caches as allocated by ModelRunner::create_kv_cache (model_runner.rs:364-396);
    /// block_tables / context_lens / slot_mapping are the Context fields (utils/context.rs:11-35).
    #[allow(clippy::too_many_arguments)]
    pub unsafe fn forward(&self, out: *mut c_void, q: *const c_void, k_new: *const c_void, v_new: *const c_void,
                          k_cache: *mut c_void, v_cache: *mut c_void, slot_mapping: *const i32, block_tables: *const i32,
                          context_lens: *const i32, batch: i32, max_blocks: i32, workspace: *mut c_void,
                          workspace_bytes: usize, stream: *mut c_void) -> anyhow::Result<()> {
        let d = self.layout.head_dim as i64;
        // attention.rs:78-81  store_kv_cache
        check(nvllm_store_kvcache(k_new, v_new, k_cache, v_cache, slot_mapping, batch,
                                  self.layout.num_kv_heads as i64 * d, &self.layout, NVLLM_BF16, stream))?;
        // attention.rs:92-95  flash_attention_decode
        check(nvllm_paged_decode_attention(out, q, k_cache, v_cache, block_tables, context_lens, batch, self.num_heads,
                                           max_blocks, self.num_heads as i64 * d, self.num_heads as i64 * d, &self.layout,
                                           self.scale, workspace, workspace_bytes, 0, NVLLM_BF16, stream))
    }
}
