"""In-graph cost of the decode-step GEMMs (K5): each shape is captured `layers` times back to back into one CUDA graph
(a different weight tensor per launch, > L2 in total, PDL edges kept — exactly how the decode step issues them) and the
graph is replayed; us/launch = replay time / launches.  Per-launch event timing from Python cannot resolve these
(~25 us floor); this can.  Sweeps the dispatch knobs (NVLLM_GEMM_BN / MCAST / 2CTA / SPLITK) in one process."""
import argparse, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nano_vllm_rs_b200 import _ffi, ops  # noqa: E402

SHAPES = {  # name: (M, N, K, kind)
    "0.6B qkv": (256, 4096, 1024, "gemm"), "0.6B o": (256, 1024, 2048, "gemm"), "0.6B gate_up+silu": (256, 3072, 1024, "silu"),
    "0.6B down": (256, 1024, 3072, "gemm"),
    "8B m256 qkv": (256, 6144, 4096, "gemm"), "8B m256 o": (256, 4096, 4096, "gemm"), "8B m256 gate_up+silu": (256, 12288, 4096, "silu"),
    "8B m256 down": (256, 4096, 12288, "gemm"),
    "32B/tp8 m64 qkv": (64, 1280, 5120, "gemm"), "32B/tp8 m64 o": (64, 5120, 1024, "gemm"),
    "32B/tp8 m64 gate_up+silu": (64, 3200, 5120, "silu"), "32B/tp8 m64 down": (64, 5120, 3200, "gemm"),
    "8B/tp2 qkv": (512, 3072, 4096, "gemm"), "8B/tp2 o": (512, 4096, 2048, "gemm"), "8B/tp2 gate_up+silu": (512, 6144, 4096, "silu"),
    "8B/tp2 down": (512, 4096, 6144, "gemm"),
    "8B m512 qkv": (512, 6144, 4096, "gemm"), "8B m512 o": (512, 4096, 4096, "gemm"), "8B m512 gate_up+silu": (512, 12288, 4096, "silu"),
    "8B m512 down": (512, 4096, 12288, "gemm"),
    "8B/tp4 qkv": (512, 1536, 4096, "gemm"), "8B/tp4 o": (512, 4096, 1024, "gemm"), "8B/tp4 gate_up+silu": (512, 3072, 4096, "silu"),
    "8B/tp4 down": (512, 4096, 3072, "gemm"),
    "8B/tp8 qkv": (512, 768, 4096, "gemm"), "8B/tp8 o": (512, 4096, 512, "gemm"), "8B/tp8 gate_up+silu": (512, 1536, 4096, "silu"),
    "8B/tp8 down": (512, 4096, 1536, "gemm"),
}
S, T, G = "NVLLM_GEMM_SPLITK", "NVLLM_GEMM_SK_STAGES", "NVLLM_GEMM_STAGES"
KNOBS = [{}, {S: "0"}, {S: "64,4"}, {S: "64,2"}, {S: "128,4"}, {S: "128,2"}, {S: "256,4"}, {S: "256,2"},
         {S: "0", "NVLLM_GEMM_MCAST": "1"}, {S: "0", "NVLLM_GEMM_2CTA": "2"},
         {S: "0", "NVLLM_GEMM_BN": "128"}, {S: "0", "NVLLM_GEMM_BN": "192"}, {S: "0", "NVLLM_GEMM_BN": "256"}, {S: "128,8"}, {S: "256,8"}]
ALL = ("NVLLM_GEMM_BN", "NVLLM_GEMM_MCAST", "NVLLM_GEMM_2CTA", S, T, G)


def run(name, M, N, K, kind, layers, reps):
    x = torch.randn((M, K), device="cuda").to(torch.bfloat16)
    rows = 2 * N if kind == "silu" else N
    ws = [(torch.randn((rows, K), device="cuda") * 0.02).to(torch.bfloat16) for _ in range(layers)]
    y = torch.empty((M, N), dtype=torch.bfloat16, device="cuda")
    fn = (lambda w: ops.gemm_silu_mul(x, w, out=y)) if kind == "silu" else (lambda w: ops.gemm_bf16(x, w, out=y))
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn(ws[0])                      # lazy attribute setup outside capture
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for w in ws:
                fn(w)
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(reps):
            g.replay()
        e1.record(s)
        torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (reps * layers)
    wbytes = rows * K * 2
    return {"shape": name, "M": M, "N": N, "K": K, "kind": kind, "us_per_launch": round(us, 2),
            "weight_GBps": round(wbytes / us / 1e3, 1), "TFLOPs": round(2.0 * M * rows * K / us / 1e6, 1)}


def main():
    ap = argparse.ArgumentParser(); ap.add_argument("--layers", type=int, default=28); ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--only", default=""); ap.add_argument("--knobs", default="", help="comma list of indices into KNOBS")
    a = ap.parse_args()
    for name, (M, N, K, kind) in SHAPES.items():
        if a.only and not any(o in name for o in a.only.split(",")):
            continue
        for ki, kn in enumerate(KNOBS):
            if a.knobs and str(ki) not in a.knobs.split(","):
                continue
            for k in ALL:
                os.environ.pop(k, None)
            os.environ.update(kn)
            _ffi.lib().nvllm_reload_env()   # the library parses its knobs once; sweeps re-parse explicitly
            try:
                r = run(name, M, N, K, kind, a.layers, a.reps)
            except Exception as ex:  # a knob that does not apply to a shape
                r = {"shape": name, "error": str(ex)[:200]}
            r["knobs"] = kn
            print(json.dumps(r), flush=True)
    for k in ALL:
        os.environ.pop(k, None)
    _ffi.lib().nvllm_reload_env()


if __name__ == "__main__":
    main()
