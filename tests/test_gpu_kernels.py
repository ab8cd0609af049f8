"""-m gpu parity tests: every HBM-bound kernel behind include/nvllm.h against the CPU oracle, through the
C ABI (ctypes).  Bit-exact for data movement / integer work; bf16 tolerance stated per test."""
import numpy as np
import pytest
import torch

from oracle import layers as O
from oracle.bf16 import round_bf16
from oracle.weights import init_weight_bf16_bits, init_weight_f32
from tests.util import (assert_close, cache_from_layout, cache_to_layout, make_paged_kv, randn_bf16, to_dev_bf16,
                        to_np_f32)

pytestmark = pytest.mark.gpu

# attention tolerance: |d| <= 1e-2 + 1.6e-2 |ref|  (~2 bf16 ulps of the output + P rounding)
ATT_ATOL, ATT_RTOL = 1e-2, 1.6e-2


def _layout(kind, bs, hkv, d):
    from nano_vllm_rs_b200._ffi import KvLayout
    return KvLayout.reference(bs, hkv, d) if kind == "reference" else KvLayout.head_major(bs, hkv, d)


@pytest.mark.parametrize("kind", ["reference", "head_major"])
@pytest.mark.parametrize("n,hkv,d,bs", [(1, 8, 128, 256), (37, 8, 128, 256), (512, 2, 128, 16), (130, 4, 64, 32), (0, 8, 128, 256)])
def test_store_kvcache_bit_exact(cuda, kind, n, hkv, d, bs):
    from nano_vllm_rs_b200 import ops
    rng = np.random.default_rng(n * 7 + hkv)
    nblk = max(2, (n + bs - 1) // bs + 2)
    k = randn_bf16(rng, (n, hkv, d)); v = randn_bf16(rng, (n, hkv, d))
    kc = randn_bf16(rng, (nblk, bs, hkv, d)); vc = randn_bf16(rng, (nblk, bs, hkv, d))
    slots = rng.permutation(nblk * bs)[:n].astype(np.int32)
    if n > 4:
        slots[3] = -1  # padding row: must be skipped
    kd, vd = cache_to_layout(kc, kind, cuda), cache_to_layout(vc, kind, cuda)
    # k, v as strided views into a packed qkv-like buffer (token stride > Hkv*D)
    packed = torch.zeros((max(n, 1), 3 * hkv * d), dtype=torch.bfloat16, device=cuda)
    kview = packed[:n, hkv * d : 2 * hkv * d].view(n, hkv, d) if n else packed[:0, : hkv * d].view(0, hkv, d)
    vview = packed[:n, 2 * hkv * d :].view(n, hkv, d) if n else packed[:0, : hkv * d].view(0, hkv, d)
    if n:
        kview.copy_(to_dev_bf16(k, cuda)); vview.copy_(to_dev_bf16(v, cuda))
    ops.store_kvcache(kview, vview, kd, vd, torch.from_numpy(slots).to(cuda), _layout(kind, bs, hkv, d))
    torch.cuda.synchronize()
    O.store_kv_cache(k, v, kc, vc, slots)
    assert np.array_equal(cache_from_layout(kd, kind), kc)
    assert np.array_equal(cache_from_layout(vd, kind), vc)


DECODE_CASES = [
    # (context_lens, H, Hkv, D, block_size, num_splits)
    ([1], 16, 8, 128, 256, 1),
    ([5, 300, 256, 257, 1024, 17], 16, 8, 128, 256, 1),   # Qwen3-0.6B heads, ragged, block edges
    ([5, 300, 256, 257, 1024, 17], 16, 8, 128, 256, 3),   # same through split-K + reduce
    ([700, 64, 1], 32, 8, 128, 256, 0),                   # Qwen3-8B group 4, library-chosen splits
    ([2000, 33], 8, 1, 128, 256, 4),                      # TP=8 shard of Qwen3-32B: 1 kv head, group 8
    ([129, 47, 16, 15], 4, 2, 128, 16, 2),                # reference scheduler-test block size 16
    ([90, 200], 12, 1, 128, 32, 1),                       # group 12 > 8: two q-head groups per kv head
    ([100, 31], 8, 4, 64, 32, 1),                         # head_dim 64
    ([0, 40], 16, 8, 128, 256, 1),                        # empty context -> zeros
]


@pytest.mark.parametrize("kind", ["reference", "head_major"])
@pytest.mark.parametrize("case", DECODE_CASES, ids=[str(i) for i in range(len(DECODE_CASES))])
def test_paged_decode_attention_vs_oracle(cuda, kind, case):
    from nano_vllm_rs_b200 import ops
    ctx, H, Hkv, D, bs, splits = case
    rng = np.random.default_rng(len(ctx) * 1000 + H + bs)
    kc, vc, tables, _, _ = make_paged_kv(rng, ctx, bs, Hkv, D)
    q = randn_bf16(rng, (len(ctx), H, D))
    scale = 1.0 / np.sqrt(D)
    ref = O.flash_attention_decode(q, kc, vc, tables, np.asarray(ctx, np.int32), scale)
    out = ops.paged_decode_attention(to_dev_bf16(q, cuda), cache_to_layout(kc, kind, cuda), cache_to_layout(vc, kind, cuda),
                                     torch.from_numpy(tables).to(cuda), torch.tensor(ctx, dtype=torch.int32, device=cuda),
                                     _layout(kind, bs, Hkv, D), scale=scale, num_splits=splits)
    torch.cuda.synchronize()
    assert_close(to_np_f32(out), ref, ATT_ATOL, ATT_RTOL, f"paged decode {case}")


PERSIST_CASES = [
    # (batch, max_ctx, H, Hkv, block_size): more (sequence, kv head, group) items than resident CTAs -> the persistent kernel
    (80, 700, 16, 8, 256),      # 640 items, Qwen3-0.6B heads, ragged contexts incl. empty rows
    (330, 300, 4, 1, 256),      # TP=8 shard shape (one kv head per rank, group 4): 330 items on 296 CTAs
    (50, 260, 24, 2, 16),       # group 12: two q-head groups per kv head (200 items... x ngrp 2 = 200), block size 16
    (700, 90, 8, 8, 32),        # many short items: item boundaries dominate
]


@pytest.mark.parametrize("kind", ["reference", "head_major"])
@pytest.mark.parametrize("case", PERSIST_CASES, ids=[str(i) for i in range(len(PERSIST_CASES))])
def test_paged_decode_persistent_kernel(cuda, kind, case, monkeypatch):
    """The persistent variant of K2 (pipelines run across item boundaries) vs the oracle, and bit-identical to the
    one-CTA-per-item kernel (same per-warp token order, same merge).  Ragged contexts with empty rows (padded graph rows)."""
    from nano_vllm_rs_b200 import ops
    B, L, H, Hkv, bs = case
    D = 128
    rng = np.random.default_rng(B * 7 + L)
    ctx = rng.integers(1, L + 1, B).tolist()
    for i in range(0, B, 17):
        ctx[i] = 0                                     # empty rows anywhere in the item sequence
    ctx[-1] = L
    kc, vc, tables, _, _ = make_paged_kv(rng, ctx, bs, Hkv, D)
    q = randn_bf16(rng, (B, H, D))
    scale = 1.0 / np.sqrt(D)
    args = (to_dev_bf16(q, cuda), cache_to_layout(kc, kind, cuda), cache_to_layout(vc, kind, cuda),
            torch.from_numpy(tables).to(cuda), torch.tensor(ctx, dtype=torch.int32, device=cuda), _layout(kind, bs, Hkv, D))
    monkeypatch.setenv("NVLLM_DECODE_PERSIST", "1")   # opt-in: it measured within +-1 % of the default kernel
    out = ops.paged_decode_attention(*args, scale=scale, num_splits=1)
    monkeypatch.setenv("NVLLM_DECODE_PERSIST", "0")
    plain = ops.paged_decode_attention(*args, scale=scale, num_splits=1)
    torch.cuda.synchronize()
    assert torch.equal(out, plain)
    ref = O.flash_attention_decode(q, kc, vc, tables, np.asarray(ctx, np.int32), scale)
    assert_close(to_np_f32(out), ref, ATT_ATOL, ATT_RTOL, f"persistent paged decode {case}")


FUSED_CASES = [
    # (context_lens INCLUDING this step's token, H, Hkv, block_size, num_splits, qk_norm)
    ([1], 16, 8, 256, 1, True),                             # nothing cached: the step's own token only
    ([6, 301, 257, 258, 1025, 18, 513], 16, 8, 256, 1, True),  # Qwen3-0.6B heads; new token opens a block (257, 513, 1025)
    ([6, 301, 257, 258, 1025, 18, 513], 16, 8, 256, 3, True),  # through split-K: the last non-empty split owns the new token
    ([701, 65, 2, 17], 32, 8, 256, 0, True),                # Qwen3-8B group 4, library-chosen splits
    ([2001, 34], 8, 1, 256, 4, True),                       # TP=8 shard of Qwen3-32B: 1 kv head, group 8
    ([130, 48, 17, 16], 4, 2, 16, 2, True),                 # block size 16: new token first / last in its block
    ([91, 201], 12, 1, 32, 1, False),                       # group 12 > 8 (two q-head groups, one stores), no qk-norm
    ([0, 41], 16, 8, 256, 1, True),                         # padded graph row: ctx 0, slot -1 -> zeros, nothing stored
]


@pytest.mark.parametrize("kind", ["reference", "head_major"])
@pytest.mark.parametrize("case", FUSED_CASES, ids=[str(i) for i in range(len(FUSED_CASES))])
def test_paged_decode_attention_fused_vs_oracle(cuda, kind, case):
    """nvllm_paged_decode_attention_fused = q/k-norm + RoPE + store_kv_cache + paged decode attention in one launch, against
    the oracle's three functions in sequence.  Attention output within the attention tolerance; the stored K row within the
    RoPE tolerance (2 bf16 ulps), the stored V row and every other cache byte bit-exact."""
    from nano_vllm_rs_b200 import ops
    ctx, H, Hkv, bs, splits, qknorm = case
    D, B, maxpos = 128, len(ctx), 4096
    rng = np.random.default_rng(B * 977 + H + bs)
    cached = [max(c - 1, 0) for c in ctx]
    kc, vc, tables, _, _ = make_paged_kv(rng, [c + 1 for c in cached], bs, Hkv, D)   # room for the new token's block
    # the slot of the new token must hold garbage before the call (make_paged_kv wrote c+1 tokens): scrub it to a marker
    qkv = randn_bf16(rng, (B, (H + 2 * Hkv) * D))
    pos = np.asarray([max(c - 1, 0) for c in ctx], np.int64)
    slots = np.asarray([(tables[b, cached[b] // bs] * bs + cached[b] % bs) if ctx[b] > 0 else -1 for b in range(B)], np.int32)
    qw = round_bf16(1.0 + 0.1 * rng.standard_normal(D).astype(np.float32)) if qknorm else None
    kw = round_bf16(1.0 + 0.1 * rng.standard_normal(D).astype(np.float32)) if qknorm else None
    table = O.rope_cos_sin_table(maxpos, D, 1e6)
    scale = 1.0 / np.sqrt(D)
    # ---- oracle: qknorm_rope -> store_kv_cache -> flash_attention_decode
    q = qkv[:, : H * D].reshape(B, H, D); k = qkv[:, H * D : (H + Hkv) * D].reshape(B, Hkv, D)
    v = qkv[:, (H + Hkv) * D :].reshape(B, Hkv, D)
    q_ref, k_ref = O.qknorm_rope(q, k, pos, table, qw, kw, 1e-6)
    kc_ref, vc_ref = kc.copy(), vc.copy()
    O.store_kv_cache(k_ref, v, kc_ref, vc_ref, slots)
    ref = O.flash_attention_decode(q_ref, kc_ref, vc_ref, tables, np.asarray(ctx, np.int32), scale)
    # ---- device
    kd, vd = cache_to_layout(kc, kind, cuda), cache_to_layout(vc, kind, cuda)
    qkv_d = to_dev_bf16(qkv, cuda)
    out = ops.paged_decode_attention_fused(
        qkv_d, torch.from_numpy(pos).to(cuda), torch.from_numpy(table).to(cuda), None if qw is None else to_dev_bf16(qw, cuda),
        None if kw is None else to_dev_bf16(kw, cuda), 1e-6, kd, vd, torch.from_numpy(slots).to(cuda),
        torch.from_numpy(tables).to(cuda), torch.tensor(ctx, dtype=torch.int32, device=cuda), H, _layout(kind, bs, Hkv, D),
        scale=scale, num_splits=splits)
    torch.cuda.synchronize()
    assert np.array_equal(to_np_f32(qkv_d), qkv), "the packed qkv rows are read only"
    assert_close(to_np_f32(out), ref, ATT_ATOL, ATT_RTOL, f"fused paged decode {case}")
    kc_got, vc_got = cache_from_layout(kd, kind), cache_from_layout(vd, kind)
    assert np.array_equal(vc_got, vc_ref), "V rows are copied bit-exactly, nothing else in the cache may change"
    mask = np.zeros(kc.shape[:2], bool)
    for b in range(B):
        if slots[b] >= 0:
            mask[slots[b] // bs, slots[b] % bs] = True
    assert np.array_equal(kc_got[~mask], kc[~mask]), "only the new tokens' slots may change"
    assert_close(kc_got[mask], kc_ref[mask], 2e-3, 2 ** -6, "stored K rows (norm + RoPE)")


def test_paged_decode_full_size_properties(cuda):
    """BASELINE config 2 shape (256 seqs x 1024 ctx, 16/8 heads): size-independent checks —
    (i) V == const c per head  =>  out == c exactly representable in bf16;  (ii) split-K invariance;
    (iii) layout invariance (reference vs head-major physical layouts give identical bits)."""
    from nano_vllm_rs_b200 import ops
    from nano_vllm_rs_b200._ffi import KvLayout
    B, L, H, Hkv, D, bs = 256, 1024, 16, 8, 128, 256
    g = torch.Generator(device="cuda").manual_seed(11)
    nblk = B * L // bs
    kc = torch.randn((nblk, bs, Hkv, D), generator=g, device=cuda, dtype=torch.float32).to(torch.bfloat16)
    vc = torch.randn((nblk, bs, Hkv, D), generator=g, device=cuda, dtype=torch.float32).to(torch.bfloat16)
    q = torch.randn((B, H, D), generator=g, device=cuda, dtype=torch.float32).to(torch.bfloat16)
    perm = torch.randperm(nblk, generator=torch.Generator().manual_seed(7)).to(torch.int32).view(B, L // bs).to(cuda)
    ctx = torch.full((B,), L, dtype=torch.int32, device=cuda)
    lay = KvLayout.reference(bs, Hkv, D)
    o1 = ops.paged_decode_attention(q, kc, vc, perm, ctx, lay, num_splits=1)
    o4 = ops.paged_decode_attention(q, kc, vc, perm, ctx, lay, num_splits=4)
    kh, vh = kc.permute(0, 2, 1, 3).contiguous(), vc.permute(0, 2, 1, 3).contiguous()
    oh = ops.paged_decode_attention(q, kh, vh, perm, ctx, KvLayout.head_major(bs, Hkv, D), num_splits=1)
    torch.cuda.synchronize()
    assert torch.equal(o1, oh), "layout invariance"
    assert_close(to_np_f32(o4), to_np_f32(o1), 4e-3, 1.6e-2, "split-K invariance")
    # cross-check a slice against torch on the device (fp32 math on the same bf16 data)
    for b in (0, 101, 255):
        k = kc[perm[b].long()].reshape(L, Hkv, D).float(); v = vc[perm[b].long()].reshape(L, Hkv, D).float()
        qq = q[b].float().view(Hkv, H // Hkv, D)
        s = torch.einsum("kgd,lkd->kgl", qq, k) / np.sqrt(D)
        ref = torch.einsum("kgl,lkd->kgd", torch.softmax(s, -1), v).reshape(H, D)
        assert_close(to_np_f32(o1[b]), to_np_f32(ref), ATT_ATOL, ATT_RTOL, f"decode full-size row {b}")
    vconst = torch.full_like(vc, 0.75)
    oc = ops.paged_decode_attention(q, kc, vconst, perm, ctx, lay, num_splits=2)
    torch.cuda.synchronize()
    assert_close(to_np_f32(oc), np.full((B, H, D), 0.75), 4e-3, 0, "constant-V property")


@pytest.mark.parametrize("n,hidden", [(1, 1024), (256, 1024), (77, 4096), (5, 5120), (3, 64)])
@pytest.mark.parametrize("with_res", [False, True])
def test_add_rmsnorm_vs_oracle(cuda, n, hidden, with_res):
    from nano_vllm_rs_b200 import ops
    rng = np.random.default_rng(n + hidden)
    x = randn_bf16(rng, (n, hidden)); res = randn_bf16(rng, (n, hidden), 3.0)
    w = round_bf16(1.0 + 0.1 * rng.standard_normal(hidden).astype(np.float32))
    eps = 1e-6
    xd, wd = to_dev_bf16(x, cuda), to_dev_bf16(w, cuda)
    if with_res:
        rd = to_dev_bf16(res, cuda)
        out = ops.add_rmsnorm(xd, wd, eps, residual=rd)
        ref, s = O.add_rms_norm(x, res, w, eps)
        torch.cuda.synchronize()
        assert np.array_equal(to_np_f32(rd), s), "residual stream must be bit-exact (one bf16 add)"
    else:
        out = ops.add_rmsnorm(xd, wd, eps)
        ref = O.rms_norm(x, w, eps)
    torch.cuda.synchronize()
    # one bf16 ulp: f32 reduction order differs from numpy's pairwise sum
    assert_close(to_np_f32(out), ref, 1e-6, 2 ** -7, "rmsnorm")


def test_rope_table_vs_oracle(cuda):
    from nano_vllm_rs_b200 import ops
    t = ops.rope_table(4096, 128, 1e6)
    ref = O.rope_cos_sin_table(4096, 128, 1e6)
    torch.cuda.synchronize()
    assert_close(to_np_f32(t), ref, 2e-7, 0, "rope table (f64 math, f32 storage: <= 1 ulp of 1.0)")


@pytest.mark.parametrize("kind", ["reference", "head_major"])
@pytest.mark.parametrize("n,H,Hkv,D,qknorm", [(64, 16, 8, 128, True), (7, 32, 8, 128, False), (33, 4, 2, 64, True), (1, 8, 1, 128, True)])
def test_qknorm_rope_store_vs_oracle(cuda, kind, n, H, Hkv, D, qknorm):
    from nano_vllm_rs_b200 import ops
    rng = np.random.default_rng(n + H)
    bs, nblk, maxpos = 16, (n + 15) // 16 + 2, 2048
    qkv = randn_bf16(rng, (n, (H + 2 * Hkv) * D))
    pos = rng.integers(0, maxpos, n).astype(np.int64)
    qw = round_bf16(1.0 + 0.1 * rng.standard_normal(D).astype(np.float32)) if qknorm else None
    kw = round_bf16(1.0 + 0.1 * rng.standard_normal(D).astype(np.float32)) if qknorm else None
    table = O.rope_cos_sin_table(maxpos, D, 1e6)
    slots = rng.permutation(nblk * bs)[:n].astype(np.int32)
    kc = np.zeros((nblk, bs, Hkv, D), np.float32); vc = np.zeros_like(kc)
    q = qkv[:, : H * D].reshape(n, H, D); k = qkv[:, H * D : (H + Hkv) * D].reshape(n, Hkv, D)
    v = qkv[:, (H + Hkv) * D :].reshape(n, Hkv, D)
    q_ref, k_ref = O.qknorm_rope(q, k, pos, table, qw, kw, 1e-6)
    kc_ref, vc_ref = kc.copy(), vc.copy()
    O.store_kv_cache(k_ref, v, kc_ref, vc_ref, slots)

    qkv_d = to_dev_bf16(qkv, cuda)
    kd, vd = cache_to_layout(kc, kind, cuda), cache_to_layout(vc, kind, cuda)
    ops.qknorm_rope_store(qkv_d, torch.from_numpy(pos).to(cuda), torch.from_numpy(table).to(cuda),
                          None if qw is None else to_dev_bf16(qw, cuda), None if kw is None else to_dev_bf16(kw, cuda),
                          1e-6, H, Hkv, D, kd, vd, torch.from_numpy(slots).to(cuda), _layout(kind, bs, Hkv, D))
    torch.cuda.synchronize()
    got = to_np_f32(qkv_d)
    # bf16 tolerance: 1 ulp from the norm's reduction order feeding the rotation (<= 2 ulps at the output)
    assert_close(got[:, : H * D].reshape(n, H, D), q_ref, 2e-3, 2 ** -6, "q rope")
    assert_close(got[:, H * D : (H + Hkv) * D].reshape(n, Hkv, D), k_ref, 2e-3, 2 ** -6, "k rope")
    assert np.array_equal(got[:, (H + Hkv) * D :], qkv[:, (H + Hkv) * D :]), "v must be untouched"
    # the cache holds exactly what the kernel wrote back in place (bit-exact data movement)
    kc_got, vc_got = cache_from_layout(kd, kind), cache_from_layout(vd, kind)
    kc_exp, vc_exp = np.zeros_like(kc), np.zeros_like(vc)
    O.store_kv_cache(got[:, H * D : (H + Hkv) * D].reshape(n, Hkv, D), v, kc_exp, vc_exp, slots)
    assert np.array_equal(kc_got, kc_exp) and np.array_equal(vc_got, vc_exp)
    assert np.array_equal(vc_got, vc_ref)


@pytest.mark.parametrize("n,inter", [(1, 3072), (256, 3072), (19, 12288), (4, 8)])
def test_silu_mul_vs_oracle(cuda, n, inter):
    from nano_vllm_rs_b200 import ops
    rng = np.random.default_rng(n + inter)
    x = randn_bf16(rng, (n, 2 * inter), 2.0)
    out = ops.silu_mul(to_dev_bf16(x, cuda))
    torch.cuda.synchronize()
    # expf vs numpy exp differ by <= 1 f32 ulp, which can flip a bf16 rounding: 1 bf16 ulp
    assert_close(to_np_f32(out), O.silu_and_mul(x), 1e-6, 2 ** -7, "silu_mul")


def test_embedding_and_last_token_gather_bit_exact(cuda):
    from nano_vllm_rs_b200 import ops
    rng = np.random.default_rng(5)
    table = randn_bf16(rng, (1000, 1024))
    ids = rng.integers(0, 1000, 300).astype(np.int64)
    out = ops.embedding_gather(to_dev_bf16(table, cuda), torch.from_numpy(ids).to(cuda))
    x = randn_bf16(rng, (300, 1024))
    cu = np.array([0, 1, 50, 51, 300], np.int32)
    last = ops.gather_last_tokens(to_dev_bf16(x, cuda), torch.from_numpy(cu).to(cuda))
    torch.cuda.synchronize()
    assert np.array_equal(to_np_f32(out), O.embedding(table, ids))
    assert np.array_equal(to_np_f32(last), O.extract_last_tokens(x, cu))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_argmax_first_max_bit_exact(cuda, dtype):
    from nano_vllm_rs_b200 import ops
    rng = np.random.default_rng(9)
    logits = randn_bf16(rng, (37, 151936))
    logits[3, 777] = logits[3, 150000] = 50.0     # tie: first index wins (np.argmax / candle semantics)
    logits[5, :] = -1.0
    logits[0] = [1.0, 2.0, 5.0, 1.5] + [0.0] * (151936 - 4)   # sampler.rs:333-342 known answer -> 2
    t = torch.from_numpy(logits).to(cuda).to(dtype)
    ids, mx = ops.argmax(t, vocab_offset=0, return_max=True)
    ids_off = ops.argmax(t, vocab_offset=1000)
    torch.cuda.synchronize()
    ref = O.greedy_sample(logits)
    assert ref[0] == 2 and ref[3] == 777 and ref[5] == 0
    assert np.array_equal(ids.cpu().numpy(), ref)
    assert np.array_equal(ids_off.cpu().numpy(), ref + 1000)
    assert np.array_equal(mx.cpu().numpy(), logits.max(-1))
    # cross-shard combine (embed_head.rs:321-336 TODO): 4 vocab shards == argmax over the full vocab
    shards = t.view(37, 4, -1)
    am, ai = zip(*[[x for x in reversed(ops.argmax(shards[:, r].contiguous(), vocab_offset=r * shards.shape[2], return_max=True))] for r in range(4)])
    comb = ops.argmax_combine(torch.stack(am), torch.stack(ai))
    torch.cuda.synchronize()
    assert np.array_equal(comb.cpu().numpy(), ref)


def test_init_weight_matches_oracle_bit_exact(cuda):
    from nano_vllm_rs_b200 import ops
    w = ops.init_weight((1000, 333), seed=1234, tensor_id=17, std=0.02)
    wf = ops.init_weight((4097,), seed=99, tensor_id=2, std=1.0, dtype=torch.float32)
    torch.cuda.synchronize()
    assert np.array_equal(w.view(torch.int16).cpu().numpy().view(np.uint16), init_weight_bf16_bits((1000, 333), 1234, 17, 0.02))
    assert np.array_equal(wf.cpu().numpy(), init_weight_f32(4097, 99, 2, 1.0))
