"""Micro-benchmark of the tcgen05 GEMM (K5): TFLOP/s and effective GB/s per Qwen3 shape, next to torch.matmul
(cuBLAS) as the library yard-stick.  L2 is flushed between iterations by rotating through > 126 MB of weights
when the weight itself is smaller than L2."""
import argparse, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nano_vllm_rs_b200 import ops  # noqa: E402

SHAPES = {  # name: (M, N, K)
    "0.6B qkv decode": (256, 4096, 1024), "0.6B o decode": (256, 1024, 2048), "0.6B gate_up decode": (256, 6144, 1024),
    "0.6B down decode": (256, 1024, 3072), "0.6B lm_head decode": (256, 151936, 1024),
    "0.6B qkv prefill": (32768, 4096, 1024), "0.6B gate_up prefill": (32768, 6144, 1024),
    "8B qkv decode": (512, 6144, 4096), "8B gate_up decode": (512, 24576, 4096), "8B down decode": (512, 4096, 12288),
    "8B gate_up prefill": (8192, 24576, 4096), "8B down prefill": (8192, 4096, 12288), "square 8192": (8192, 8192, 8192),
}


def timeit(fn, iters, warmup=3):
    for _ in range(warmup):
        fn(0)
    torch.cuda.synchronize()
    evs = []
    for i in range(iters):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(i); e.record(); evs.append((s, e))
    torch.cuda.synchronize()
    ms = sorted(s.elapsed_time(e) for s, e in evs)
    return ms[len(ms) // 2]


def main():
    ap = argparse.ArgumentParser(); ap.add_argument("--iters", type=int, default=10); ap.add_argument("--only", default="")
    a = ap.parse_args()
    for name, (M, N, K) in SHAPES.items():
        if a.only and a.only not in name:
            continue
        wbytes = N * K * 2
        copies = max(1, min(8, (256 << 20) // wbytes + 1))  # rotate weights so they do not sit in L2
        x = torch.randn((M, K), device="cuda").to(torch.bfloat16)
        ws = [(torch.randn((N, K), device="cuda") * 0.02).to(torch.bfloat16) for _ in range(copies)]
        y = torch.empty((M, N), dtype=torch.bfloat16, device="cuda")
        ours = timeit(lambda i: ops.gemm_bf16(x, ws[i % copies], out=y), a.iters)
        lib = timeit(lambda i: torch.matmul(x, ws[i % copies].t(), out=y), a.iters)
        fl = 2.0 * M * N * K; by = (M * K + N * K + M * N) * 2
        print(json.dumps({"shape": name, "M": M, "N": N, "K": K, "bn": os.environ.get("NVLLM_GEMM_BN", "auto"),
                          "ms": round(ours, 4), "TFLOPs": round(fl / ours / 1e9, 1), "GBps": round(by / ours / 1e6, 1),
                          "cublas_ms": round(lib, 4), "cublas_TFLOPs": round(fl / lib / 1e9, 1)}))
        del ws


if __name__ == "__main__":
    main()
