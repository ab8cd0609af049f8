"""CPU tests pinning the oracle's floating-point layer functions: (i) every formula-level known answer the reference's
in-file unit tests state, (ii) an independent torch-CPU fp32 implementation of the same math, and
(iii) the committed golden vectors under tests/golden/ (generator: tests/golden/make_layers_golden.py).
The reference holds no numeric golden outputs for attention / logits, so (ii)+(iii) are cross-checks, not reference pins."""
import json
import os

import numpy as np
import torch

from oracle import layers as O
from oracle.bf16 import round_bf16

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_scale_attention_rs_456():
    assert 1.0 / np.sqrt(np.float32(64)) == np.float32(0.125)


def test_causal_mask_pattern_attention_rs_470_485():
    m = O.create_causal_mask(3)
    ninf = -np.inf
    assert m.tolist() == [[0.0, ninf, ninf], [0.0, 0.0, ninf], [0.0, 0.0, 0.0]]
    # bottom-right alignment for Lk > Lq (prefix-cache prefill): row i sees keys 0 .. Lk-Lq+i
    assert O.create_causal_mask(2, 4).tolist() == [[0.0, 0.0, 0.0, ninf], [0.0, 0.0, 0.0, 0.0]]


def test_gqa_head_mapping_attention_rs_419_435():
    """expand_kv_heads is a repeat-interleave: q head h uses kv head h // (H / Hkv)."""
    rng = np.random.default_rng(0)
    q = round_bf16(rng.standard_normal((1, 8, 64)).astype(np.float32))
    k = round_bf16(rng.standard_normal((5, 2, 64)).astype(np.float32)); v = round_bf16(rng.standard_normal((5, 2, 64)).astype(np.float32))
    out = O.compute_attention(q, k, v, 0.125, causal=False)
    k_exp, v_exp = np.repeat(k, 4, axis=1), np.repeat(v, 4, axis=1)   # [5, 8, 64]: the reference's expansion
    ref = O.compute_attention(q, k_exp, v_exp, 0.125, causal=False)
    assert np.array_equal(out, ref)


def test_rope_identities_rotary_embedding_rs_322_386():
    table = O.rope_cos_sin_table(16, 8, 10000.0)
    assert np.all(table[0, :4] == 1.0) and np.all(table[0, 4:] == 0.0)          # cos(0) = 1, sin(0) = 0
    x = round_bf16(np.random.default_rng(1).standard_normal((3, 2, 8)).astype(np.float32))
    out = O.apply_rotary_emb_single(x, np.ones((3, 4), np.float32), np.zeros((3, 4), np.float32))
    assert np.array_equal(out, x)                                                # identity for c = 1, s = 0
    # a 90-degree rotation swaps the halves with a sign: [x1*0 - x2*1 | x2*0 + x1*1]
    out = O.apply_rotary_emb_single(x, np.zeros((3, 4), np.float32), np.ones((3, 4), np.float32))
    assert np.array_equal(out, np.concatenate([-x[..., 4:], x[..., :4]], -1))
    # inv_freq_j = theta^(-2j/D)
    assert np.allclose(table[1, :4], np.cos(10000.0 ** (-np.arange(0, 8, 2) / 8)), atol=1e-7)


def test_rmsnorm_unit_rms_layernorm_rs_195_221():
    x = np.array([[1, 2, 3, 4], [5, 6, 7, 8]], np.float32)
    out = O.rms_norm(x, np.ones(4, np.float32), 1e-6)
    rms = np.sqrt((out.astype(np.float64) ** 2).mean(-1))
    assert np.all(np.abs(rms - 1.0) < 0.1)


def test_silu_and_mul_activation_rs_264_290():
    x = np.array([[1.0, 2.0, 3.0, 0.5, 1.5, 2.5]], np.float32)
    out = O.silu_and_mul(x)
    exp = O.silu(x[:, :3]) * x[:, 3:]
    assert out.shape == (1, 3)
    assert np.all(np.abs(out - exp) <= 2 ** -7 * np.abs(exp))     # the oracle rounds to bf16; the reference test runs in f32
    assert np.allclose(O.silu(np.array([1.0, 2.0, 3.0], np.float32)) * np.array([0.5, 1.5, 2.5], np.float32), exp[0], atol=1e-6)


def test_greedy_sampler_rs_333_342():
    assert O.greedy_sample(np.array([[1.0, 2.0, 5.0, 1.5]], np.float32)).tolist() == [2]
    assert O.greedy_sample(np.array([[3.0, 7.0, 7.0]], np.float32)).tolist() == [1]      # first maximum


# ---------------------------------------------------------------------------- independent torch-CPU cross-checks

def test_attention_vs_torch_sdpa():
    rng = np.random.default_rng(2)
    H, Hkv, D = 8, 2, 64
    for lq, lk, causal in [(1, 37, False), (16, 16, True), (5, 21, True)]:
        q = round_bf16(rng.standard_normal((lq, H, D)).astype(np.float32))
        k = round_bf16(rng.standard_normal((lk, Hkv, D)).astype(np.float32)); v = round_bf16(rng.standard_normal((lk, Hkv, D)).astype(np.float32))
        out = O.compute_attention(q, k, v, 1 / np.sqrt(D), causal)
        tq = torch.from_numpy(q).permute(1, 0, 2); tk = torch.from_numpy(k).permute(1, 0, 2).repeat_interleave(H // Hkv, 0)
        tv = torch.from_numpy(v).permute(1, 0, 2).repeat_interleave(H // Hkv, 0)
        mask = None
        if causal:
            mask = torch.ones(lq, lk, dtype=torch.bool).tril(diagonal=lk - lq)
        ref = torch.nn.functional.scaled_dot_product_attention(tq, tk, tv, attn_mask=mask).permute(1, 0, 2).numpy()
        assert np.abs(out - ref).max() <= 2 ** -7 * np.abs(ref).max() + 1e-6, (lq, lk, causal)


def test_paged_decode_equals_dense_attention():
    """Walking block_tables (gather_cached_kv, attention.rs:293-318) == attention over the dense per-sequence K/V."""
    from tests.util import make_paged_kv
    rng = np.random.default_rng(3)
    ctx = [5, 40, 16, 33]
    kc, vc, tables, ks, vs = make_paged_kv(rng, ctx, 16, 2, 64)
    q = round_bf16(rng.standard_normal((4, 4, 64)).astype(np.float32))
    out = O.flash_attention_decode(q, kc, vc, tables, np.asarray(ctx, np.int32), 0.125)
    for b in range(4):
        ref = O.compute_attention(q[b:b + 1], ks[b], vs[b], 0.125, causal=False)
        assert np.array_equal(out[b:b + 1], ref)


def test_rmsnorm_rope_vs_torch():
    rng = np.random.default_rng(4)
    x = round_bf16(rng.standard_normal((7, 256)).astype(np.float32)); w = round_bf16(1 + 0.1 * rng.standard_normal(256).astype(np.float32))
    t = torch.from_numpy(x)
    ref = (t * torch.rsqrt(t.pow(2).mean(-1, keepdim=True) + 1e-6) * torch.from_numpy(w)).numpy()
    assert np.abs(O.rms_norm(x, w, 1e-6) - ref).max() <= 2 ** -7 * np.abs(ref).max()
    # HF-style rotate_half RoPE (transformers Qwen3: q*cos + rotate_half(q)*sin with cos/sin duplicated over both halves)
    D = 64
    q = round_bf16(rng.standard_normal((7, 3, D)).astype(np.float32))
    pos = rng.integers(0, 500, 7)
    table = O.rope_cos_sin_table(512, D, 1e6)
    cos, sin = table[pos, : D // 2], table[pos, D // 2 :]
    out = O.apply_rotary_emb_single(q, cos, sin)
    tq = torch.from_numpy(q)
    c2 = torch.from_numpy(np.concatenate([cos, cos], -1))[:, None, :]; s2 = torch.from_numpy(np.concatenate([sin, sin], -1))[:, None, :]
    rot = torch.cat([-tq[..., D // 2 :], tq[..., : D // 2]], -1)
    ref = (tq * c2 + rot * s2).numpy()
    assert np.abs(out - ref).max() <= 2 ** -7 * np.abs(ref).max()


def test_layer_golden_vectors():
    g = json.load(open(os.path.join(ROOT, "tests", "golden", "layers.json")))
    a = lambda k: np.asarray(g[k], np.float32)
    assert np.array_equal(O.rms_norm(a("x"), a("w"), 1e-6), a("rms_norm"))
    assert np.array_equal(O.silu_and_mul(a("gate_up")), a("silu_and_mul"))
    q, k, v = a("q"), a("k"), a("v")
    assert np.array_equal(O.compute_attention(q, k, v, 0.125, True), a("attention_causal"))
    qr, kr = O.qknorm_rope(q, k, np.asarray(g["positions"], np.int64), O.rope_cos_sin_table(64, 64, 1e6), None, None, 1e-6)
    assert np.array_equal(qr, a("q_rope")) and np.array_equal(kr, a("k_rope"))
