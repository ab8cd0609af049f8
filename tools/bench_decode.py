"""Micro-benchmark of the paged decode attention kernel (K2): achieved algorithmic HBM GB/s.

  python tools/bench_decode.py [--batch 256 --ctx 1024 --heads 16 --kv-heads 8 --layout head_major --splits 0]

Algorithmic bytes per launch:
  sum_b 2*L_b*Hkv*D*2 + 2*B*H*D*2 + 4*sum_b ceil(L_b/bs) + 4*B      (split-K partials excluded)
KV (>= 1 GiB at the default shape) is far larger than the 126 MB L2, so every iteration streams from HBM."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nano_vllm_rs_b200 import ops  # noqa: E402
from nano_vllm_rs_b200._ffi import KvLayout  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--ctx", type=int, default=1024)
    ap.add_argument("--ragged", action="store_true", help="context lengths U[ctx/2, ctx]")
    ap.add_argument("--heads", type=int, default=16)
    ap.add_argument("--kv-heads", type=int, default=8)
    ap.add_argument("--head-dim", type=int, default=128)
    ap.add_argument("--block-size", type=int, default=256)
    ap.add_argument("--layout", default="head_major", choices=["head_major", "reference"])
    ap.add_argument("--splits", type=int, default=0)
    ap.add_argument("--sweep-splits", default="", help="comma list: one JSON line per split count (0 = the chooser)")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    B, L, H, Hkv, D, bs = a.batch, a.ctx, a.heads, a.kv_heads, a.head_dim, a.block_size
    g = torch.Generator(device="cuda").manual_seed(7)
    if a.ragged:
        ctx = torch.randint(L // 2, L + 1, (B,), generator=torch.Generator().manual_seed(7), dtype=torch.int32)
    else:
        ctx = torch.full((B,), L, dtype=torch.int32)
    nb = ((ctx + bs - 1) // bs)
    maxb = int(nb.max())
    nblk = int(nb.sum())
    shape = (nblk, Hkv, bs, D) if a.layout == "head_major" else (nblk, bs, Hkv, D)
    kc = torch.randn(shape, generator=g, device=dev, dtype=torch.bfloat16)
    vc = torch.randn(shape, generator=g, device=dev, dtype=torch.bfloat16)
    q = torch.randn((B, H, D), generator=g, device=dev, dtype=torch.bfloat16)
    perm = torch.randperm(nblk, generator=torch.Generator().manual_seed(7)).to(torch.int32)  # non-contiguous pages
    tables = torch.full((B, maxb), -1, dtype=torch.int32)
    o = 0
    for b in range(B):
        n = int(nb[b]); tables[b, :n] = perm[o:o + n]; o += n
    tables, ctx_d = tables.to(dev), ctx.to(dev)
    lay = KvLayout.head_major(bs, Hkv, D) if a.layout == "head_major" else KvLayout.reference(bs, Hkv, D)
    out = torch.empty((B, H, D), dtype=torch.bfloat16, device=dev)
    ws = torch.empty(max(16, ops.paged_decode_workspace_bytes(B, H, D, 64)), dtype=torch.uint8, device=dev)
    alg = int((2 * ctx.long() * Hkv * D * 2).sum()) + 2 * B * H * D * 2 + 4 * int(nb.sum()) + 4 * B
    from nano_vllm_rs_b200 import _ffi
    for splits in ([int(x) for x in a.sweep_splits.split(",")] if a.sweep_splits else [a.splits]):
        run = lambda: ops.paged_decode_attention(q, kc, vc, tables, ctx_d, lay, num_splits=splits, out=out, workspace=ws)
        for _ in range(a.warmup):
            run()
        torch.cuda.synchronize()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.iters)]
        for s, e in evs:
            s.record(); run(); e.record()
        torch.cuda.synchronize()
        ms = sorted(s.elapsed_time(e) for s, e in evs)
        med = ms[len(ms) // 2]
        chosen = splits or _ffi.lib().nvllm_paged_decode_choose_splits(B, H, Hkv, maxb * bs)
        print(json.dumps({"kernel": "paged_decode_attention", "cfg": os.environ.get("NVLLM_DECODE_CFG", "0"),
                          "layout": a.layout, "batch": B, "ctx": L, "ragged": a.ragged, "heads": H, "kv_heads": Hkv,
                          "splits": splits, "splits_used": chosen, "ms_median": round(med, 4), "ms_min": round(ms[0], 4),
                          "alg_bytes": alg, "GBps_median": round(alg / med / 1e6, 1), "GBps_best": round(alg / ms[0] / 1e6, 1)}),
              flush=True)


if __name__ == "__main__":
    main()
