"""Micro-benchmark of the device sampler (K12, nvllm_sample): us per launch and effective GB/s (logit bytes actually
streamed per pass x passes) for the workload's batch (256 rows x 151936 f32 logits = 155.6 MB, larger than L2) under
greedy / temperature / top-k / top-p / top-k + top-p, next to the greedy argmax kernel (K11)."""
import argparse, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nano_vllm_rs_b200 import ops  # noqa: E402


def timeit(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


def main():
    ap = argparse.ArgumentParser(); ap.add_argument("--batch", type=int, default=256); ap.add_argument("--vocab", type=int, default=151936)
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    B, V = a.batch, a.vocab
    g = torch.Generator(device="cuda").manual_seed(0)
    logits = torch.randn((B, V), generator=g, device="cuda") * 3.0
    seeds = torch.randint(0, 2 ** 62, (B,), generator=g, device="cuda", dtype=torch.int64)
    f = lambda v: torch.full((B,), v, dtype=torch.float32, device="cuda")
    i = lambda v: torch.full((B,), v, dtype=torch.int32, device="cuda")
    row_bytes = B * V * 4
    cases = {  # name: (temperature, top_k, top_p, passes over the row)
        "greedy (T=0)": (f(0.0), None, None, 1), "temperature only": (f(0.8), None, None, 1),
        "top_k=50": (f(0.8), i(50), None, 5), "top_p=0.9": (f(0.8), None, f(0.9), 6), "top_k=50 + top_p=0.9": (f(0.8), i(50), f(0.9), 10),
    }
    us = timeit(lambda: ops.argmax(logits), a.iters)
    print(json.dumps({"kernel": "argmax (K11)", "batch": B, "vocab": V, "us": round(us, 1), "GBps": round(row_bytes / us / 1e3, 1)}))
    for name, (t, k, p, passes) in cases.items():
        us = timeit(lambda: ops.sample(logits, t, seeds, top_k=k, top_p=p), a.iters)
        print(json.dumps({"kernel": "sample (K12)", "case": name, "batch": B, "vocab": V, "us": round(us, 1), "passes": passes,
                          "GBps_effective": round(passes * row_bytes / us / 1e3, 1)}), flush=True)


if __name__ == "__main__":
    main()
