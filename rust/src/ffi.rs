//! Safe wrappers with the reference's method signatures over the generated `extern "C"` surface (ffi_gen.rs).
//! COMPILE-UNVERIFIED (no Rust toolchain in the build image); the same ABI is exercised through ctypes by
//! nano-vllm-rs_b200/_ffi.py and every test under tests/.
use std::ffi::{c_void, CStr};

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

/// What `ModelRunner::sample_tokens` -> `Sampler::batch_sample` (src/engine/model_runner.rs:131-156,
/// src/layers/sampler.rs:221-254) becomes: the per-sequence parameters are device arrays, one launch samples the batch.
pub struct DeviceSampler;

impl DeviceSampler {
    /// logits: f32 [batch, vocab] on the device; temperatures / top_k / top_p / row_seeds: device arrays of `batch`
    /// entries (top_k / top_p may be null = None for every row); out_ids: i64 [batch] on the device.
    #[allow(clippy::too_many_arguments)]
    pub unsafe fn forward(out_ids: *mut i64, logits: *const f32, batch: i32, vocab: i32, temperatures: *const f32,
                          top_k: *const i32, top_p: *const f32, row_seeds: *const u64, stream: *mut c_void) -> anyhow::Result<()> {
        check(nvllm_sample(out_ids, logits, batch, vocab, vocab as i64, temperatures, top_k, top_p, row_seeds,
                           std::ptr::null_mut(), std::ptr::null_mut(), stream))
    }
}
