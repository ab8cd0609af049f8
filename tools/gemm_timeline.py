"""Where the fixed ~5 us of a decode-sized GEMM launch goes: per-CTA %globaltimer stamps at the phase boundaries of the
cluster split-K kernel (debug build: make -C nano-vllm-rs_b200/csrc timeline -> lib/libnvllm_timeline.so), collected over a
CUDA graph of `layers` back-to-back launches on distinct weights (PDL edges kept), exactly how the decode step issues them.

  NVLLM_LIB_PATH=nano-vllm-rs_b200/lib/libnvllm_timeline.so python tools/gemm_timeline.py [--shape o|down|qkv]

Phases (ns, medians over CTAs and launches): gap = first CTA start - last CTA end of the previous launch; prologue = barrier
init + TMEM alloc + block barrier; pdl = until the predecessor is complete; first = until the first operand stage has landed;
stream = first -> last stage landed; mma = last stage -> accumulator complete; park = TMEM -> shared; sync1; reduce = DSMEM
reduce + global stores; sync2 = second cluster barrier + exit."""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nano_vllm_rs_b200 import _ffi, ops  # noqa: E402

SHAPES = {"o": (256, 1024, 2048), "down": (256, 1024, 3072), "qkv": (256, 4096, 1024), "tp8_o": (512, 4096, 512),
          "tp8_qkv": (512, 768, 4096)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="o")
    ap.add_argument("--layers", type=int, default=28)
    a = ap.parse_args()
    lib = _ffi.lib()
    assert hasattr(lib, "nvllmdbg_timeline"), "load the timeline build: NVLLM_LIB_PATH=.../libnvllm_timeline.so"
    lib.nvllmdbg_timeline.argtypes = [C.c_void_p, C.c_uint]
    M, N, K = SHAPES[a.shape]
    x = torch.randn((M, K), device="cuda").to(torch.bfloat16)
    ws = [(torch.randn((N, K), device="cuda") * 0.02).to(torch.bfloat16) for _ in range(a.layers)]
    y = torch.empty((M, N), dtype=torch.bfloat16, device="cuda")
    cap = 64 * 1024
    buf = torch.zeros((cap, 12), dtype=torch.int64, device="cuda")
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        ops.gemm_bf16(x, ws[0], out=y)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for w in ws:
                ops.gemm_bf16(x, w, out=y)
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        assert lib.nvllmdbg_timeline(buf.data_ptr(), cap) == 0
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s); g.replay(); e1.record(s)
        torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / a.layers
    t = buf.cpu().numpy()
    t = t[t[:, 1] != 0]
    if len(t) == 0:
        print(json.dumps({"error": "no timeline entries: this shape did not take the split-K kernel"}))
        return
    ctas = int(t[:, 0].max()) + 1
    launches = len(t) // ctas
    L = np.zeros((launches, ctas, 10), dtype=np.int64)
    smid = np.zeros((launches, ctas), dtype=np.int64)
    for b in range(ctas):
        rows = t[t[:, 0] == b]
        rows = rows[np.argsort(rows[:, 1])]
        for i in range(min(launches, len(rows))):
            L[i, b] = rows[i, 1:11]
            smid[i, b] = rows[i, 11]
    names = ["prologue", "pdl", "first", "stream", "mma", "park", "sync1", "reduce", "sync2"]
    segs = {"prologue": (0, 1), "pdl": (0, 2), "first": (0, 3), "stream": (3, 4), "mma": (4, 5), "park": (5, 6), "sync1": (6, 7),
            "reduce": (7, 8), "sync2": (8, 9)}
    out = {"shape": a.shape, "M": M, "N": N, "K": K, "ctas": ctas, "launches": launches, "us_per_launch_in_graph": round(us, 2),
           "distinct_sms_per_launch": int(np.median([len(set(smid[i])) for i in range(launches)]))}
    sel = slice(2, launches)   # skip the first launches of the graph
    for n in names:
        lo, hi = segs[n]
        d = (L[sel, :, hi] - L[sel, :, lo]).astype(np.float64)
        out[n + "_ns"] = {"median": float(np.median(d)), "p10": float(np.percentile(d, 10)), "p90": float(np.percentile(d, 90))}
    life = (L[sel, :, 9] - L[sel, :, 0]).astype(np.float64)
    out["cta_lifetime_ns"] = {"median": float(np.median(life)), "p90": float(np.percentile(life, 90))}
    start, end = L[:, :, 0].min(axis=1), L[:, :, 9].max(axis=1)
    out["gap_prev_end_to_first_start_ns"] = {"median": float(np.median(start[1:] - end[:-1])),
                                             "min": float((start[1:] - end[:-1]).min()), "max": float((start[1:] - end[:-1]).max())}
    out["start_skew_ns"] = float(np.median(L[sel, :, 0].max(axis=1) - L[sel, :, 0].min(axis=1)))
    out["launch_period_ns"] = float(np.median(np.diff(start)))
    out["kernel_span_ns"] = float(np.median(end - start))
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
