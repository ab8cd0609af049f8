"""One Python function per kernel entry point of include/nvllm.h.

Arguments are torch CUDA tensors (PyTorch is only the device-memory / stream plumbing); every call is
enqueued on torch's current stream and goes through the C ABI — there is no eager / CPU fallback.
Names mirror the reference layer methods they replace (file:line in include/nvllm.h)."""
import math

import torch

from . import _ffi
from ._ffi import KvLayout, check

BF16, F32 = 0, 1


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _p(t):
    return None if t is None else t.data_ptr()


def _need(t, dtype, name):
    if not t.is_cuda:
        raise ValueError(f"{name}: expected a CUDA tensor (no CPU path exists)")
    if t.dtype != dtype:
        raise ValueError(f"{name}: expected {dtype}, got {t.dtype}")


def store_kvcache(k, v, k_cache, v_cache, slot_mapping, layout: KvLayout):
    """Attention::store_kv_cache (attention.rs:150-174). k, v: [N, Hkv, D] (last two dims contiguous)."""
    _need(k, torch.bfloat16, "k"); _need(v, torch.bfloat16, "v"); _need(slot_mapping, torch.int32, "slot_mapping")
    assert k.stride(-1) == 1 and k.stride(1) == k.shape[2] and k.stride() == v.stride()
    check(_ffi.lib().nvllm_store_kvcache(_p(k), _p(v), _p(k_cache), _p(v_cache), _p(slot_mapping), k.shape[0],
                                         k.stride(0), layout, BF16, _stream()))


def paged_decode_workspace_bytes(batch, num_heads, head_dim, num_splits=0):
    return _ffi.lib().nvllm_paged_decode_workspace_bytes(batch, num_heads, head_dim, num_splits)


def paged_decode_attention(q, k_cache, v_cache, block_tables, context_lens, layout: KvLayout, scale=None,
                           num_splits=0, out=None, workspace=None):
    """Attention::flash_attention_decode (attention.rs:225-318). q: [B, H, D] -> out [B, H, D]."""
    _need(q, torch.bfloat16, "q"); _need(block_tables, torch.int32, "block_tables")
    _need(context_lens, torch.int32, "context_lens")
    B, H, D = q.shape
    assert q.stride(2) == 1 and q.stride(1) == D and block_tables.is_contiguous()
    if out is None:
        out = torch.empty((B, H, D), dtype=torch.bfloat16, device=q.device)
    if scale is None:
        scale = 1.0 / math.sqrt(D)
    if num_splits == 0:
        num_splits = _ffi.lib().nvllm_paged_decode_choose_splits(B, H, layout.num_kv_heads,
                                                                 block_tables.shape[1] * layout.block_size)
    need = paged_decode_workspace_bytes(B, H, D, num_splits)
    if need and (workspace is None or workspace.numel() * workspace.element_size() < need):
        workspace = torch.empty(need, dtype=torch.uint8, device=q.device)
    check(_ffi.lib().nvllm_paged_decode_attention(
        _p(out), _p(q), _p(k_cache), _p(v_cache), _p(block_tables), _p(context_lens), B, H, block_tables.shape[1],
        q.stride(0), out.stride(0), layout, scale, _p(workspace),
        0 if workspace is None else workspace.numel() * workspace.element_size(), num_splits, BF16, _stream()))
    return out


def paged_decode_attention_fused(qkv, positions, cos_sin, q_norm_w, k_norm_w, eps, k_cache, v_cache, slot_mapping, block_tables,
                                 context_lens, num_heads, layout: KvLayout, scale=None, num_splits=0, out=None, workspace=None):
    """q/k-norm + RoPE + store_kv_cache + flash_attention_decode in one launch (qwen3.rs:219-233).  qkv: [B, (H+2Hkv)*D]
    packed rows of this step (read only); context_lens counts this step's token.  -> out [B, H, D]."""
    _need(qkv, torch.bfloat16, "qkv"); _need(positions, torch.int64, "positions"); _need(cos_sin, torch.float32, "cos_sin")
    _need(block_tables, torch.int32, "block_tables"); _need(context_lens, torch.int32, "context_lens")
    _need(slot_mapping, torch.int32, "slot_mapping")
    B, H, D = qkv.shape[0], num_heads, layout.head_dim
    assert qkv.is_contiguous() and qkv.shape[1] == (H + 2 * layout.num_kv_heads) * D and block_tables.is_contiguous()
    if out is None:
        out = torch.empty((B, H, D), dtype=torch.bfloat16, device=qkv.device)
    if scale is None:
        scale = 1.0 / math.sqrt(D)
    if num_splits == 0:
        num_splits = _ffi.lib().nvllm_paged_decode_choose_splits(B, H, layout.num_kv_heads,
                                                                 block_tables.shape[1] * layout.block_size)
    need = paged_decode_workspace_bytes(B, H, D, num_splits)
    if need and (workspace is None or workspace.numel() * workspace.element_size() < need):
        workspace = torch.empty(need, dtype=torch.uint8, device=qkv.device)
    check(_ffi.lib().nvllm_paged_decode_attention_fused(
        _p(out), _p(qkv), _p(positions), _p(cos_sin), _p(q_norm_w), _p(k_norm_w), eps, cos_sin.shape[0], _p(k_cache),
        _p(v_cache), _p(slot_mapping), _p(block_tables), _p(context_lens), B, H, block_tables.shape[1], qkv.stride(0),
        out.stride(0), layout, scale, _p(workspace),
        0 if workspace is None else workspace.numel() * workspace.element_size(), num_splits, BF16, _stream()))
    return out


def varlen_prefill_attention(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, scale=None,
                             k_cache=None, v_cache=None, block_tables=None, layout: KvLayout = None, out=None):
    """Attention::flash_attention_varlen / _with_cache (attention.rs:177-222). q: [N, H, D]; k, v: [Nk, Hkv, D]."""
    _need(q, torch.bfloat16, "q"); _need(cu_seqlens_q, torch.int32, "cu_seqlens_q")
    N, H, D = q.shape
    if scale is None:
        scale = 1.0 / math.sqrt(D)
    if out is None:
        out = torch.empty((N, H, D), dtype=torch.bfloat16, device=q.device)
    paged = block_tables is not None
    Hkv = layout.num_kv_heads if paged else k.shape[1]
    check(_ffi.lib().nvllm_varlen_prefill_attention(
        _p(out), _p(q), _p(k), _p(v), _p(cu_seqlens_q), _p(cu_seqlens_k), cu_seqlens_q.numel() - 1,
        N, 0 if k is None else k.shape[0], max_seqlen_q, max_seqlen_k, H, Hkv, D, q.stride(0), 0 if k is None else k.stride(0), out.stride(0),
        _p(k_cache), _p(v_cache), _p(block_tables), block_tables.shape[1] if paged else 0,
        layout if paged else None, scale, BF16, _stream()))
    return out


def gemm_bf16(x, w, out=None, out_dtype=torch.bfloat16, simt=False):
    """LinearBase::forward (linear.rs): y[M,N] = x[M,K] @ w[N,K]^T on tcgen05 (simt=True: test-only CUDA-core path)."""
    _need(x, torch.bfloat16, "x"); _need(w, torch.bfloat16, "w")
    M, K = x.shape
    N = w.shape[0]
    assert w.shape[1] == K and x.stride(1) == 1 and w.stride(1) == 1
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=x.device)
    od = BF16 if out.dtype == torch.bfloat16 else F32
    L = _ffi.lib()
    if simt:
        check(L.nvllm_gemm_bf16_simt(_p(out), _p(x), _p(w), M, N, K, x.stride(0), w.stride(0), out.stride(0), od, _stream()))
        return out
    need = L.nvllm_gemm_workspace_bytes(M, N, K)
    ws = torch.empty(need, dtype=torch.uint8, device=x.device) if need else None
    check(L.nvllm_gemm_bf16(_p(out), _p(x), _p(w), M, N, K, x.stride(0), w.stride(0), out.stride(0), od, _p(ws), need, _stream()))
    return out


def gemm_silu_mul(x, w_gate_up, out=None):
    """MergedColumnParallelLinear::forward + SiluAndMul::forward fused (linear.rs:437-439, activation.rs:46-63)."""
    _need(x, torch.bfloat16, "x"); _need(w_gate_up, torch.bfloat16, "w_gate_up")
    M, K = x.shape
    inter = w_gate_up.shape[0] // 2
    if out is None:
        out = torch.empty((M, inter), dtype=torch.bfloat16, device=x.device)
    ws = torch.empty((M, 2 * inter), dtype=torch.bfloat16, device=x.device)
    check(_ffi.lib().nvllm_gemm_silu_mul_bf16(_p(out), _p(x), _p(w_gate_up), M, inter, K, x.stride(0), w_gate_up.stride(0),
                                              out.stride(0), _p(ws), ws.numel() * 2, _stream()))
    return out


def gemm_qkv_rope_store(x, w_qkv, positions, cos_sin, q_norm_w, k_norm_w, eps, num_heads, num_kv_heads, head_dim,
                        k_cache=None, v_cache=None, slot_mapping=None, layout: KvLayout = None, out=None):
    """qkv_proj + q/k norm + RoPE + store_kv_cache in one kernel (linear.rs:354-356, qwen3.rs:219-229, attention.rs:150-174)."""
    _need(x, torch.bfloat16, "x"); _need(w_qkv, torch.bfloat16, "w_qkv"); _need(positions, torch.int64, "positions")
    M, K = x.shape
    n = (num_heads + 2 * num_kv_heads) * head_dim
    if out is None:
        out = torch.empty((M, n), dtype=torch.bfloat16, device=x.device)
    check(_ffi.lib().nvllm_gemm_qkv_rope_store(_p(out), _p(x), _p(w_qkv), M, K, x.stride(0), w_qkv.stride(0), _p(positions),
                                               _p(cos_sin), _p(q_norm_w), _p(k_norm_w), eps, num_heads, num_kv_heads, head_dim,
                                               cos_sin.shape[0], _p(k_cache), _p(v_cache), _p(slot_mapping), layout, _stream()))
    return out


def add_rmsnorm(x, weight, eps, residual=None, out=None):
    """RMSNorm::forward (layernorm.rs:50-97). With `residual`, it is updated in place to bf16(residual + x)."""
    _need(x, torch.bfloat16, "x")
    assert x.is_contiguous() and (residual is None or residual.is_contiguous())
    if out is None:
        out = torch.empty_like(x)
    check(_ffi.lib().nvllm_add_rmsnorm(_p(out), _p(x), _p(residual), _p(weight), x.shape[0], x.shape[1], eps, BF16, _stream()))
    return out


def rope_table(max_position, head_dim, theta, device="cuda"):
    t = torch.empty((max_position, head_dim), dtype=torch.float32, device=device)
    check(_ffi.lib().nvllm_rope_table(_p(t), max_position, head_dim, float(theta), _stream()))
    return t


def qknorm_rope_store(qkv, positions, cos_sin, q_norm_w, k_norm_w, eps, num_heads, num_kv_heads, head_dim,
                      k_cache=None, v_cache=None, slot_mapping=None, layout: KvLayout = None):
    """RotaryEmbedding::forward (+ q/k norm, + store_kv_cache), in place on qkv [N, (H+2Hkv)*D]."""
    _need(qkv, torch.bfloat16, "qkv"); _need(positions, torch.int64, "positions"); _need(cos_sin, torch.float32, "cos_sin")
    assert qkv.is_contiguous() and qkv.shape[1] == (num_heads + 2 * num_kv_heads) * head_dim
    check(_ffi.lib().nvllm_qknorm_rope_store(_p(qkv), _p(positions), _p(cos_sin), _p(q_norm_w), _p(k_norm_w), eps,
                                             qkv.shape[0], num_heads, num_kv_heads, head_dim, cos_sin.shape[0],
                                             _p(k_cache), _p(v_cache), _p(slot_mapping), layout, BF16, _stream()))
    return qkv


def silu_mul(x, out=None):
    """SiluAndMul::forward (activation.rs:46-63). x: [N, 2I] -> [N, I]."""
    _need(x, torch.bfloat16, "x")
    assert x.is_contiguous()
    inter = x.shape[1] // 2
    if out is None:
        out = torch.empty((x.shape[0], inter), dtype=torch.bfloat16, device=x.device)
    check(_ffi.lib().nvllm_silu_mul(_p(out), _p(x), x.shape[0], inter, BF16, _stream()))
    return out


def embedding_gather(table, ids, out=None):
    """VocabParallelEmbedding::forward (embed_head.rs:77-97)."""
    _need(table, torch.bfloat16, "table"); _need(ids, torch.int64, "ids")
    if out is None:
        out = torch.empty((ids.numel(), table.shape[1]), dtype=torch.bfloat16, device=table.device)
    check(_ffi.lib().nvllm_embedding_gather(_p(out), _p(table), _p(ids), ids.numel(), table.shape[1], table.shape[0], BF16, _stream()))
    return out


def gather_last_tokens(x, cu_seqlens_q, out=None):
    """ParallelLMHead::extract_last_tokens (embed_head.rs:272-289)."""
    _need(x, torch.bfloat16, "x"); _need(cu_seqlens_q, torch.int32, "cu_seqlens_q")
    B = cu_seqlens_q.numel() - 1
    if out is None:
        out = torch.empty((B, x.shape[1]), dtype=torch.bfloat16, device=x.device)
    check(_ffi.lib().nvllm_gather_last_tokens(_p(out), _p(x), _p(cu_seqlens_q), B, x.shape[1], BF16, _stream()))
    return out


def argmax(logits, vocab_offset=0, return_max=False):
    """Sampler::greedy_sample (sampler.rs:109-112)."""
    assert logits.is_cuda and logits.stride(1) == 1
    B, V = logits.shape
    ids = torch.empty(B, dtype=torch.int64, device=logits.device)
    mx = torch.empty(B, dtype=torch.float32, device=logits.device) if return_max else None
    dt = F32 if logits.dtype == torch.float32 else BF16
    check(_ffi.lib().nvllm_argmax(_p(ids), _p(mx), _p(logits), B, V, logits.stride(0), vocab_offset, dt, _stream()))
    return (ids, mx) if return_max else ids


def argmax_combine(all_max, all_ids):
    tp, B = all_max.shape
    out = torch.empty(B, dtype=torch.int64, device=all_max.device)
    check(_ffi.lib().nvllm_argmax_combine(_p(out), _p(all_max), _p(all_ids), tp, B, _stream()))
    return out


def sample(logits, temperatures, row_seeds, top_k=None, top_p=None, debug=False):
    """Sampler::forward (sampler.rs:26-106): temperature / top-k / top-p / Gumbel-max per row, greedy where T == 0.
    logits f32 [B, V]; temperatures f32 [B]; row_seeds: uint64 values (any int64/uint64 tensor of bit patterns) [B];
    top_k i32 [B] / top_p f32 [B] optional.  debug=True also returns (kept_count i32 [B], kept_mask u8 [B, V])."""
    _need(logits, torch.float32, "logits"); _need(temperatures, torch.float32, "temperatures")
    assert logits.stride(1) == 1 and row_seeds.element_size() == 8 and row_seeds.is_cuda
    if top_k is not None:
        _need(top_k, torch.int32, "top_k")
    if top_p is not None:
        _need(top_p, torch.float32, "top_p")
    B, V = logits.shape
    ids = torch.empty(B, dtype=torch.int64, device=logits.device)
    cnt = torch.empty(B, dtype=torch.int32, device=logits.device) if debug else None
    mask = torch.empty((B, V), dtype=torch.uint8, device=logits.device) if debug else None
    check(_ffi.lib().nvllm_sample(_p(ids), _p(logits), B, V, logits.stride(0), _p(temperatures), _p(top_k), _p(top_p),
                                  _p(row_seeds), _p(cnt), _p(mask), _stream()))
    return (ids, cnt, mask) if debug else ids


def init_weight(shape, seed, tensor_id, std, dtype=torch.bfloat16, device="cuda"):
    t = torch.empty(shape, dtype=dtype, device=device)
    check(_ffi.lib().nvllm_init_weight(_p(t), t.numel(), seed, tensor_id, std, BF16 if dtype == torch.bfloat16 else F32, _stream()))
    return t
