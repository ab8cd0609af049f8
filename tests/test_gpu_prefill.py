"""-m gpu parity tests for the tcgen05 varlen prefill attention (K3 dense, K4 paged prefix) vs the CPU oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import layers as O
from tests.util import assert_close, cache_to_layout, make_paged_kv, randn_bf16, to_dev_bf16, to_np_f32

pytestmark = pytest.mark.gpu
ATT_ATOL, ATT_RTOL = 1e-2, 1.6e-2   # SURVEY.md §8c attention tolerance


def _cu(lens):
    return np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)


DENSE = [
    ([1], 4, 2), ([5], 16, 8), ([64], 4, 4), ([65, 1, 128], 4, 2), ([129, 256, 257], 8, 2), ([300, 513, 7], 16, 8),
    ([1024], 32, 8), ([128] * 4, 16, 8),   # the last one is BASELINE config 1's prefill shape
]


@pytest.mark.parametrize("lens,H,Hkv", DENSE, ids=[str(i) for i in range(len(DENSE))])
def test_varlen_prefill_dense_vs_oracle(cuda, lens, H, Hkv):
    from nano_vllm_rs_b200 import ops
    D = 128
    rng = np.random.default_rng(sum(lens) + H)
    n = sum(lens)
    # q/k/v as column slices of one packed qkv buffer (how the model calls it: strided rows)
    qkv = randn_bf16(rng, (n, (H + 2 * Hkv) * D))
    q = qkv[:, : H * D].reshape(n, H, D); k = qkv[:, H * D : (H + Hkv) * D].reshape(n, Hkv, D)
    v = qkv[:, (H + Hkv) * D :].reshape(n, Hkv, D)
    cu = _cu(lens)
    scale = 1 / np.sqrt(D)
    ref = O.flash_attention_varlen(q, k, v, cu, cu, scale)
    t = to_dev_bf16(qkv, cuda)
    qd = t[:, : H * D].view(n, H, D); kd = t[:, H * D : (H + Hkv) * D].view(n, Hkv, D); vd = t[:, (H + Hkv) * D :].view(n, Hkv, D)
    cud = torch.from_numpy(cu).to(cuda)
    out = ops.varlen_prefill_attention(qd, kd, vd, cud, cud, max(lens), max(lens), scale=scale)
    torch.cuda.synchronize()
    assert_close(to_np_f32(out), ref, ATT_ATOL, ATT_RTOL, f"prefill dense {lens}")


PAGED = [
    # (prefix_lens, suffix_lens, H, Hkv, block_size)
    ([256], [44], 4, 2, 256), ([0, 512, 256], [300, 1, 129], 16, 8, 256), ([64, 128], [16, 70], 4, 2, 64),
    ([32, 0, 48], [17, 33, 100], 4, 1, 16), ([4096], [256], 8, 8, 256),
]


@pytest.mark.parametrize("kind", ["reference", "head_major"])
@pytest.mark.parametrize("case", PAGED, ids=[str(i) for i in range(len(PAGED))])
def test_varlen_prefill_paged_prefix_vs_oracle(cuda, kind, case):
    """flash_attention_varlen_with_cache (attention.rs:211-222 intent): the suffix queries attend to prefix + suffix
    read back from the paged cache (all K/V of the step already stored — note N2)."""
    from nano_vllm_rs_b200 import ops
    from nano_vllm_rs_b200._ffi import KvLayout
    pre, suf, H, Hkv, bs = case
    D = 128
    rng = np.random.default_rng(sum(pre) + sum(suf) + bs)
    tot = [a + b for a, b in zip(pre, suf)]
    kc, vc, tables, ks, vs = make_paged_kv(rng, tot, bs, Hkv, D)
    nq = sum(suf)
    q = randn_bf16(rng, (nq, H, D))
    cu_q, cu_k = _cu(suf), _cu(tot)
    scale = 1 / np.sqrt(D)
    ref = O.flash_attention_varlen_with_cache(q, kc, vc, tables, cu_q, cu_k, scale)
    lay = KvLayout.reference(bs, Hkv, D) if kind == "reference" else KvLayout.head_major(bs, Hkv, D)
    out = ops.varlen_prefill_attention(to_dev_bf16(q, cuda), None, None, torch.from_numpy(cu_q).to(cuda),
                                       torch.from_numpy(cu_k).to(cuda), max(suf), max(tot), scale=scale,
                                       k_cache=cache_to_layout(kc, kind, cuda), v_cache=cache_to_layout(vc, kind, cuda),
                                       block_tables=torch.from_numpy(tables).to(cuda), layout=lay)
    torch.cuda.synchronize()
    assert_close(to_np_f32(out), ref, ATT_ATOL, ATT_RTOL, f"prefill paged {case}")


def test_varlen_prefill_full_size_vs_torch(cuda):
    """BASELINE config 2 prefill step: 64 seqs x 512 tokens, 16/8 heads — against torch SDPA (fp32 math on the
    device) on sampled sequences, plus the paged == dense consistency property."""
    from nano_vllm_rs_b200 import ops
    B, L, H, Hkv, D = 64, 512, 16, 8, 128
    g = torch.Generator(device="cuda").manual_seed(3)
    q = torch.randn((B * L, H, D), generator=g, device=cuda).to(torch.bfloat16)
    k = torch.randn((B * L, Hkv, D), generator=g, device=cuda).to(torch.bfloat16)
    v = torch.randn((B * L, Hkv, D), generator=g, device=cuda).to(torch.bfloat16)
    cu = torch.arange(0, B * L + 1, L, dtype=torch.int32, device=cuda)
    out = ops.varlen_prefill_attention(q, k, v, cu, cu, L, L)
    torch.cuda.synchronize()
    for b in (0, 17, 63):
        sl = slice(b * L, (b + 1) * L)
        qq = q[sl].float().permute(1, 0, 2); kk = k[sl].float().permute(1, 0, 2).repeat_interleave(H // Hkv, 0)
        vv = v[sl].float().permute(1, 0, 2).repeat_interleave(H // Hkv, 0)
        ref = torch.nn.functional.scaled_dot_product_attention(qq, kk, vv, is_causal=True).permute(1, 0, 2)
        assert_close(to_np_f32(out[sl]), to_np_f32(ref), ATT_ATOL, ATT_RTOL, f"prefill full-size seq {b}")
