"""Shared helpers for the parity tests: bf16 <-> numpy bridges and synthetic paged-KV builders."""
import numpy as np
import torch

from oracle.bf16 import round_bf16


def to_dev_bf16(a, dev):
    """float32 numpy array of bf16-representable values -> torch bf16 CUDA tensor (exact)."""
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev).to(torch.bfloat16)


def to_np_f32(t):
    return t.detach().float().cpu().numpy()


def randn_bf16(rng, shape, scale=1.0):
    return round_bf16((rng.standard_normal(shape) * scale).astype(np.float32))


def assert_close(got, ref, atol, rtol, what=""):
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    err = np.abs(got - ref)
    tol = atol + rtol * np.abs(ref)
    bad = err > tol
    if bad.any():
        i = np.unravel_index(np.argmax(err - tol), err.shape)
        raise AssertionError(f"{what}: {bad.sum()}/{bad.size} outside |d| <= {atol} + {rtol}|ref|; worst at {i}: "
                             f"got {got[i]} ref {ref[i]} (max abs err {err.max():.4g})")


def make_paged_kv(rng, context_lens, block_size, num_kv_heads, head_dim, extra_blocks=3, shuffle=True):
    """Random K/V for every sequence scattered into a reference-layout cache
    [num_blocks, block_size, Hkv, D] through a (shuffled) block table. Returns
    (k_cache, v_cache, block_tables[B, max_blocks] padded with -1, per-seq dense K, V lists)."""
    B = len(context_lens)
    nblk_seq = [(L + block_size - 1) // block_size for L in context_lens]
    total = sum(nblk_seq) + extra_blocks
    ids = np.arange(total)
    if shuffle:
        rng.shuffle(ids)
    k_cache = randn_bf16(rng, (total, block_size, num_kv_heads, head_dim))  # garbage where unwritten
    v_cache = randn_bf16(rng, (total, block_size, num_kv_heads, head_dim))
    max_blocks = max(1, max(nblk_seq))
    tables = np.full((B, max_blocks), -1, dtype=np.int32)
    ks, vs = [], []
    pos = 0
    for b, L in enumerate(context_lens):
        k = randn_bf16(rng, (L, num_kv_heads, head_dim))
        v = randn_bf16(rng, (L, num_kv_heads, head_dim))
        for j in range(nblk_seq[b]):
            bid = int(ids[pos]); pos += 1
            tables[b, j] = bid
            n = min(block_size, L - j * block_size)
            k_cache[bid, :n] = k[j * block_size : j * block_size + n]
            v_cache[bid, :n] = v[j * block_size : j * block_size + n]
        ks.append(k); vs.append(v)
    return k_cache, v_cache, tables, ks, vs


def cache_to_layout(cache_ref_np, layout_kind, dev):
    """Reference-layout numpy cache -> device tensor in the requested physical layout."""
    t = to_dev_bf16(cache_ref_np, dev)  # [nblk, bs, Hkv, D]
    if layout_kind == "head_major":
        t = t.permute(0, 2, 1, 3).contiguous()  # [nblk, Hkv, bs, D]
    return t


def cache_from_layout(t, layout_kind):
    if layout_kind == "head_major":
        t = t.permute(0, 2, 1, 3)
    return to_np_f32(t.contiguous())
