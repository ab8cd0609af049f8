"""Micro-benchmark of the tcgen05 varlen prefill attention (K3): TFLOP/s = 4*H*D*sum_b sum_i (Lk-Lq+i+1) / time."""
import argparse, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nano_vllm_rs_b200 import ops  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--only", default="", help="substring of the shape name")
    a = ap.parse_args()
    for name, B, L, H, Hkv in [("0.6B cfg2 prefill step", 64, 512, 16, 8), ("8B cfg3 prefill step", 32, 1024, 32, 8),
                               ("8B 4K prompt", 8, 4096, 32, 8), ("32B/TP8 32K", 1, 32768, 8, 1)]:
        if a.only and a.only not in name:
            continue
        D = 128
        q = torch.randn((B * L, H, D), device="cuda").to(torch.bfloat16)
        k = torch.randn((B * L, Hkv, D), device="cuda").to(torch.bfloat16)
        v = torch.randn((B * L, Hkv, D), device="cuda").to(torch.bfloat16)
        cu = torch.arange(0, B * L + 1, L, dtype=torch.int32, device="cuda")
        out = torch.empty_like(q)
        for _ in range(3):
            ops.varlen_prefill_attention(q, k, v, cu, cu, L, L, out=out)
        torch.cuda.synchronize()
        evs = []
        for _ in range(a.iters):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); ops.varlen_prefill_attention(q, k, v, cu, cu, L, L, out=out); e.record(); evs.append((s, e))
        torch.cuda.synchronize()
        ms = sorted(s.elapsed_time(e) for s, e in evs)[len(evs) // 2]
        flops = 4.0 * H * D * B * (L * (L + 1) / 2)
        print(json.dumps({"shape": name, "B": B, "L": L, "H": H, "Hkv": Hkv, "ms": round(ms, 4), "TFLOPs": round(flops / ms / 1e9, 1)}))


if __name__ == "__main__":
    main()
