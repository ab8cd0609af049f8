"""Generates tests/golden/layers.json: small fixed inputs and the outputs of an INDEPENDENT torch-CPU fp32 implementation
of the layer math (rounded to bf16 where the reference materialises a bf16 tensor).  Run from the repo root."""
import json
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from oracle.bf16 import round_bf16  # only the rounding helper; the math below is torch


def bf(t):
    return torch.from_numpy(round_bf16(t.numpy().astype(np.float32)))


def main():
    g = torch.Generator().manual_seed(0)
    x = bf(torch.randn(3, 32, generator=g)); w = bf(1 + 0.1 * torch.randn(32, generator=g))
    rms = bf(x / torch.sqrt(x.pow(2).mean(-1, keepdim=True) + 1e-6) * w)
    gu = bf(torch.randn(2, 16, generator=g) * 2)
    sm = bf(bf(gu[:, :8] / (1 + torch.exp(-gu[:, :8]))) * gu[:, 8:])
    q = bf(torch.randn(4, 4, 64, generator=g)); k = bf(torch.randn(4, 2, 64, generator=g)); v = bf(torch.randn(4, 2, 64, generator=g))
    kk, vv = k.repeat_interleave(2, 1), v.repeat_interleave(2, 1)
    s = torch.einsum("qhd,khd->hqk", q, kk) * 0.125 + torch.triu(torch.full((4, 4), float("-inf")), 1)
    att = bf(torch.einsum("hqk,khd->qhd", torch.softmax(s, -1), vv))
    pos = [0, 3, 17, 63]
    inv = 1.0 / (1e6 ** (torch.arange(0, 64, 2, dtype=torch.float64) / 64))
    ang = torch.tensor(pos, dtype=torch.float64)[:, None] * inv[None]
    cos, sin = torch.cos(ang).float()[:, None, :], torch.sin(ang).float()[:, None, :]
    rope = lambda t: bf(torch.cat([t[..., :32] * cos - t[..., 32:] * sin, t[..., 32:] * cos + t[..., :32] * sin], -1))
    out = dict(x=x, w=w, rms_norm=rms, gate_up=gu, silu_and_mul=sm, q=q, k=k, v=v, attention_causal=att, q_rope=rope(q), k_rope=rope(k))
    js = {n: t.numpy().astype(np.float64).tolist() for n, t in out.items()}
    js["positions"] = pos
    json.dump(js, open("tests/golden/layers.json", "w"))


if __name__ == "__main__":
    main()
