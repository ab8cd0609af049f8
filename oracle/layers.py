"""numpy restatement of the reference's layer math (TEST INFRASTRUCTURE — see oracle/__init__.py).

All tensors are float32 numpy arrays holding bf16-representable values ("bf16 tensors"); `round_bf16`
is applied wherever the reference materialises a tensor in the model dtype.  Citations are
file:line in /root/reference/src."""
import numpy as np

from .bf16 import round_bf16

# ---------------------------------------------------------------------------- KV cache (K1)


def store_kv_cache(k, v, k_cache, v_cache, slot_mapping):
    """Attention::store_kv_cache, layers/attention.rs:150-174 — per-token loop `cache[slot] = x[i]`.
    k, v: [N, Hkv, D]; caches: [num_blocks, block_size, Hkv, D] viewed flat over slots
    (model_runner.rs:379-389).  slot < 0 is skipped (padding rows of a captured graph; the reference
    never produces them).  In place (the reference discards slice_assign's result — intent is in place)."""
    kc = k_cache.reshape(-1, *k_cache.shape[2:])
    vc = v_cache.reshape(-1, *v_cache.shape[2:])
    for i in range(k.shape[0]):  # attention.rs:162
        slot = int(slot_mapping[i])  # attention.rs:163
        if slot < 0:
            continue
        kc[slot] = k[i]  # attention.rs:169
        vc[slot] = v[i]  # attention.rs:170


# ---------------------------------------------------------------------------- attention core


def create_causal_mask(lq, lk=None):
    """Attention::create_causal_mask, attention.rs:329-339: mask[i][j] = -inf for j > i.
    Generalised to Lk >= Lq with bottom-right alignment (row i sees keys 0 .. Lk-Lq+i), which is what
    prefill over a cached prefix needs (attention.rs:211-222 intent)."""
    lk = lq if lk is None else lk
    i = np.arange(lq)[:, None]
    j = np.arange(lk)[None, :]
    return np.where(j > i + (lk - lq), -np.inf, 0.0).astype(np.float32)


def softmax_last_dim(x):
    """candle_nn::ops::softmax_last_dim (attention.rs:257): exp(x - max) / sum."""
    m = np.max(x, axis=-1, keepdims=True)
    m = np.where(np.isfinite(m), m, 0.0)
    e = np.exp(x - m)
    s = np.sum(e, axis=-1, keepdims=True)
    return e / np.where(s == 0, 1.0, s)


def compute_attention(q, k, v, scale, causal):
    """Attention::compute_attention, attention.rs:238-261, per head (the reference's transpose(1,2) on
    [L,H,D] batches over tokens — SURVEY.md §8 a4; the intent, per-head attention, is restated).
    q: [Lq, H, D], k, v: [Lk, Hkv, D]; GQA map h -> h // (H/Hkv) (expand_kv_heads, attention.rs:419-435).
    f32 math on bf16 inputs, output rounded to bf16."""
    lq, H, D = q.shape
    lk, Hkv, _ = k.shape
    g = H // Hkv
    qh = q.transpose(1, 0, 2).reshape(Hkv, g * lq, D)  # [Hkv, g*Lq, D] (head-major inside the group)
    kh = k.transpose(1, 0, 2)  # [Hkv, Lk, D]
    vh = v.transpose(1, 0, 2)
    scores = np.matmul(qh, kh.transpose(0, 2, 1)) * np.float32(scale)  # attention.rs:246-247
    scores = scores.reshape(Hkv, g, lq, lk)
    if causal:
        scores = scores + create_causal_mask(lq, lk)[None, None]  # attention.rs:250-251,321-326
    w = softmax_last_dim(scores)  # attention.rs:257
    out = np.matmul(w.reshape(Hkv, g * lq, lk), vh)  # attention.rs:260
    out = out.reshape(Hkv * g, lq, D).transpose(1, 0, 2)
    return round_bf16(out.astype(np.float32))


def gather_cached_kv(k_cache, v_cache, block_table, context_len=None):
    """Attention::gather_cached_kv, attention.rs:293-318: concatenate the blocks listed in the table,
    skipping ids < 0 (:306).  The reference then attends to every slot of the last block including
    unwritten zeros; the intent (note N1) truncates to context_len."""
    ks, vs = [], []
    for bid in block_table:
        bid = int(bid)
        if bid >= 0:
            ks.append(k_cache[bid])
            vs.append(v_cache[bid])
    if not ks:
        shp = (0,) + k_cache.shape[2:]
        return np.zeros(shp, np.float32), np.zeros(shp, np.float32)
    k = np.concatenate(ks, 0)
    v = np.concatenate(vs, 0)
    if context_len is not None:
        k, v = k[:context_len], v[:context_len]
    return k, v


def flash_attention_decode(q, k_cache, v_cache, block_tables, context_lens, scale):
    """Attention::flash_attention_decode -> compute_attention_with_cache, attention.rs:225-235,264-290:
    per-sequence loop, gather, non-causal attention of the single query token.
    q: [B, H, D]; caches [num_blocks, block_size, Hkv, D]; block_tables [B, max_blocks] (pad -1)."""
    outs = []
    for i in range(q.shape[0]):  # attention.rs:278
        L = int(context_lens[i])
        k, v = gather_cached_kv(k_cache, v_cache, block_tables[i], L)
        if k.shape[0] == 0:
            outs.append(np.zeros((1,) + q.shape[1:], np.float32))
            continue
        outs.append(compute_attention(q[i : i + 1], k, v, scale, causal=False))
    return np.concatenate(outs, 0)


def flash_attention_varlen(q, k, v, cu_seqlens_q, cu_seqlens_k, scale):
    """Attention::flash_attention_varlen, attention.rs:177-208: per-sequence causal attention."""
    outs = []
    for i in range(len(cu_seqlens_q) - 1):  # attention.rs:193
        qs, qe = int(cu_seqlens_q[i]), int(cu_seqlens_q[i + 1])
        ks, ke = int(cu_seqlens_k[i]), int(cu_seqlens_k[i + 1])
        outs.append(compute_attention(q[qs:qe], k[ks:ke], v[ks:ke], scale, causal=True))  # :203
    return np.concatenate(outs, 0) if outs else np.zeros_like(q)


def flash_attention_varlen_with_cache(q, k_cache, v_cache, block_tables, cu_seqlens_q, cu_seqlens_k, scale):
    """Attention::flash_attention_varlen_with_cache, attention.rs:211-222 (intent): the
    uncached suffix of each sequence attends, bottom-right causally, to prefix + itself read back from the
    paged cache through block_tables (this step's K/V already stored — note N2)."""
    outs = []
    for i in range(len(cu_seqlens_q) - 1):
        qs, qe = int(cu_seqlens_q[i]), int(cu_seqlens_q[i + 1])
        lk = int(cu_seqlens_k[i + 1]) - int(cu_seqlens_k[i])
        k, v = gather_cached_kv(k_cache, v_cache, block_tables[i], lk)
        outs.append(compute_attention(q[qs:qe], k, v, scale, causal=True))
    return np.concatenate(outs, 0) if outs else np.zeros_like(q)


# ---------------------------------------------------------------------------- norms / rope / act


def rms_norm(x, weight, eps):
    """RMSNorm::forward_simple, layernorm.rs:58-75: f32, x / sqrt(mean(x^2) + eps) * w, cast back."""
    x = x.astype(np.float32)
    ms = np.mean(x * x, axis=-1, keepdims=True, dtype=np.float32)
    rms = np.sqrt(ms + np.float32(eps))
    return round_bf16((x / rms) * weight.astype(np.float32))


def add_rms_norm(x, residual, weight, eps):
    """RMSNorm::forward_with_residual, layernorm.rs:78-97 + Qwen3Layer residual adds, qwen3.rs:382,389:
    s = bf16(x + residual) (a bf16 tensor add), then the norm of s.  Returns (normed, s)."""
    s = round_bf16(x.astype(np.float32) + residual.astype(np.float32))
    return rms_norm(s, weight, eps), s


def rope_cos_sin_table(max_position, head_dim, theta):
    """RotaryEmbedding::new / compute_cos_sin_cache, rotary_embedding.rs:82-110: inv_freq_j = 1/theta^(2j/D)
    in f64; angle = pos * inv_freq; table [max_pos, D] = cos | sin.  Kept in f32 (the reference casts
    inv_freq and positions to the model dtype, which breaks for bf16 — note N1)."""
    j = np.arange(0, head_dim, 2, dtype=np.float64)
    inv_freq = 1.0 / np.power(np.float64(theta), j / np.float64(head_dim))
    ang = np.arange(max_position, dtype=np.float64)[:, None] * inv_freq[None, :]
    return np.concatenate([np.cos(ang), np.sin(ang)], axis=1).astype(np.float32)


def apply_rotary_emb_single(x, cos, sin):
    """apply_rotary_emb_single, rotary_embedding.rs:23-48: halves x1 = x[..., :D/2], x2 = x[..., D/2:];
    out = [x1*cos - x2*sin | x2*cos + x1*sin].  x: [N, heads, D]; cos, sin: [N, D/2] (broadcast over heads).
    f32 math, one rounding to bf16 at the output (note N1)."""
    half = x.shape[-1] // 2
    x1, x2 = x[..., :half].astype(np.float32), x[..., half:].astype(np.float32)
    c, s = cos[:, None, :], sin[:, None, :]
    return round_bf16(np.concatenate([x1 * c - x2 * s, x2 * c + x1 * s], axis=-1))


def qknorm_rope(q, k, positions, table, q_norm_w, k_norm_w, eps):
    """Qwen3Attention::forward steps between split_qkv and Attention::forward (qwen3.rs:208-240) with the
    named architecture's per-head q_norm / k_norm (absent in the reference — note N1).
    q: [N, H, D], k: [N, Hkv, D], positions i64 [N]."""
    if q_norm_w is not None:
        q = rms_norm(q, q_norm_w, eps)
        k = rms_norm(k, k_norm_w, eps)
    half = q.shape[-1] // 2
    cs = table[np.asarray(positions, dtype=np.int64)]  # index_select, rotary_embedding.rs:136-143
    cos, sin = cs[:, :half], cs[:, half:]
    return apply_rotary_emb_single(q, cos, sin), apply_rotary_emb_single(k, cos, sin)


def silu(x):
    """silu, activation.rs:12-15: x * sigmoid(x)."""
    x = x.astype(np.float32)
    return x / (np.float32(1.0) + np.exp(-x))


def silu_and_mul(x):
    """SiluAndMul::forward, activation.rs:46-63: split the last dim in halves, silu(gate) * up.
    Two tensor ops in the model dtype -> two bf16 roundings."""
    inter = x.shape[-1] // 2
    gate, up = x[..., :inter], x[..., inter:]
    return round_bf16(round_bf16(silu(gate)) * up.astype(np.float32))


# ---------------------------------------------------------------------------- linear / embed / head


def linear(x, w, out_bf16=True):
    """LinearBase::forward, linear.rs:228-239,354-356,437-439; ParallelLMHead::compute_logits,
    embed_head.rs:292-306: y = x . W^T (bf16 inputs, f32 accumulate)."""
    y = np.matmul(np.asarray(x, dtype=np.float32), np.asarray(w, dtype=np.float32).T)  # no copy of f32 weights
    return round_bf16(y) if out_bf16 else np.asarray(y, dtype=np.float32)


def embedding(table, ids):
    """VocabParallelEmbedding::forward (tp == 1), embed_head.rs:77-97: plain row gather."""
    return table[np.asarray(ids, dtype=np.int64)]


def extract_last_tokens(x, cu_seqlens_q):
    """ParallelLMHead::extract_last_tokens, embed_head.rs:272-289: row cu_seqlens_q[i+1]-1 per sequence."""
    idx = np.asarray(cu_seqlens_q[1:], dtype=np.int64) - 1
    return x[idx]


def greedy_sample(logits):
    """Sampler::greedy_sample, sampler.rs:109-112: argmax over the vocab (first maximum)."""
    return np.argmax(logits, axis=-1).astype(np.int64)


# ---------------------------------------------------------------------------- sampler
#
# Sampler::forward / sample_single / apply_top_k / apply_top_p / multinomial_sample / sample_gumbel, sampler.rs:26-219.
# The reference draws its uniforms from an UNSEEDED Tensor::rand (sampler.rs:208), so its random draws cannot be
# reproduced by anyone; the build fixes a seeded, counter-based contract instead (identical here and on the GPU):
#   row_seed  = mix64(engine_seed ^ mix64(seq_id * GOLDEN + num_tokens_of_the_sequence))
#   u[j]      = ((mix64(row_seed + (j + 1) * GOLDEN) >> 41) + 0.5) * 2^-23          in (0, 1), exact in f32
#   gumbel[j] = -log(-log(clamp(u[j], 1e-8, 1 - 1e-8)))                              sampler.rs:211-216, f32
# Other deliberate readings of the reference: top_k == 0 means "no top-k" (batch_sample maps None to 0, which the
# reference would then index at k - 1 = usize::MAX, sampler.rs:243,131); top_p >= 1 or <= 0 means "no top-p".

GOLDEN = np.uint64(0x9E3779B97F4A7C15)


def mix64(x):
    """splitmix64 finaliser on uint64 (arrays or scalars), wrap-around arithmetic."""
    with np.errstate(over="ignore"):
        x = np.asarray(x, dtype=np.uint64)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return x ^ (x >> np.uint64(31))


def sampler_row_seed(engine_seed, seq_id, num_tokens):
    with np.errstate(over="ignore"):
        inner = mix64(np.uint64(seq_id) * GOLDEN + np.uint64(num_tokens))
        return int(mix64(np.uint64(int(engine_seed) & 0xFFFFFFFFFFFFFFFF) ^ inner))  # i32 seed sign-extended to u64


def sampler_uniform(row_seed, n):
    with np.errstate(over="ignore"):
        j = np.arange(1, n + 1, dtype=np.uint64)
        h = mix64(np.uint64(row_seed) + j * GOLDEN)
    return ((h >> np.uint64(41)).astype(np.float32) + np.float32(0.5)) * np.float32(2.0 ** -23)


def sample_gumbel(u):
    """sample_gumbel, sampler.rs:205-219 (f32)."""
    eps = np.float32(1e-8)
    uc = np.clip(u.astype(np.float32), eps, np.float32(1.0) - eps)
    return -np.log(-np.log(uc))


def apply_top_k(logits, k):
    """apply_top_k, sampler.rs:115-149: the first k entries of the stable descending sort survive, the rest -> -inf."""
    v = logits.shape[0]
    k = min(int(k), v)
    order = np.argsort(-logits, kind="stable")  # sort_by(b.partial_cmp(a)) is stable: ties keep index order
    out = np.full(v, -np.inf, dtype=np.float32)
    out[order[:k]] = logits[order[:k]]
    return out


def apply_top_p(logits, p):
    """apply_top_p, sampler.rs:152-191: softmax, stable descending sort, sequential f32 cumulative sum, keep entries up to
    and including the first one at which the running sum reaches p (all of them if it never does)."""
    v = logits.shape[0]
    m = np.max(logits)
    e = np.exp(logits - m).astype(np.float32)
    probs = e / np.sum(e, dtype=np.float32)
    order = np.argsort(-probs, kind="stable")
    cum = np.cumsum(probs[order], dtype=np.float32)  # numpy's cumsum is the sequential recurrence
    hit = np.nonzero(cum >= np.float32(p))[0]
    cutoff = int(hit[0]) + 1 if hit.size else v
    out = np.full(v, -np.inf, dtype=np.float32)
    out[order[:cutoff]] = logits[order[:cutoff]]
    return out


def sample_single(logits, temperature, top_p, top_k, row_seed):
    """sample_single, sampler.rs:70-106.  Returns (token, filtered logits, scores = filtered + gumbel or None)."""
    logits = np.asarray(logits, dtype=np.float32)
    if temperature == 0.0:
        return int(np.argmax(logits)), logits, None
    s = logits / np.float32(temperature) if temperature != 1.0 else logits.copy()
    if top_k:
        s = apply_top_k(s, top_k)
    if top_p is not None and 0.0 < top_p < 1.0:
        s = apply_top_p(s, top_p)
    scores = s + sample_gumbel(sampler_uniform(row_seed, logits.shape[0]))
    return int(np.argmax(scores)), s, scores


def sampler_forward(logits, temperatures, top_p, top_k, row_seeds):
    """Sampler::forward / batch_sample, sampler.rs:26-67,221-254: one sample_single per row."""
    out = np.zeros(len(temperatures), dtype=np.int64)
    for i in range(len(temperatures)):
        out[i], _, _ = sample_single(logits[i], float(temperatures[i]), None if top_p is None else float(top_p[i]),
                                     None if top_k is None else int(top_k[i]), int(row_seeds[i]))
    return out
