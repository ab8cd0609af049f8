"""Where the time of one low-latency fused all-reduce call goes: thread 0 of every CTA stamps %globaltimer at the phase
boundaries (debug build: make -C nano-vllm-rs_b200/csrc timeline), over a CUDA graph of back-to-back calls.

  NVLLM_LIB_PATH=nano-vllm-rs_b200/lib/libnvllm_timeline.so python -m torch.distributed.run --nnodes=1 --nproc-per-node N \
      --master-addr 127.0.0.1 --master-port P tools/ar_timeline.py

Per rank (clocks of different GPUs are not comparable): wait = entry -> predecessor complete, push = partial rows stored,
clear, arrive = clear issued -> all staged vectors of the first owned row present, reduce = until the owned rows are
broadcast, gather = until the foreign rows are present and copied, plus the call period and the span of one call."""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nano_vllm_rs_b200 import _ffi  # noqa: E402
from nano_vllm_rs_b200 import engine as E  # noqa: E402

SHAPES = [("8B decode, 512 seqs", 512, 4096), ("32B decode, 64 seqs", 64, 5120)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--calls", type=int, default=24)
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = E.lib()
    assert hasattr(lib, "nvllmdbg_ar_timeline"), "load the timeline build: NVLLM_LIB_PATH=.../libnvllm_timeline.so"
    lib.nvllmdbg_ar_timeline.argtypes = [C.c_void_p, C.c_uint]
    uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        buf = (C.c_uint8 * 128)()
        _ffi.check(lib.nvllm_comm_unique_id(buf))
        uid.copy_(torch.tensor(list(buf), dtype=torch.uint8))
    dist.broadcast(uid, 0)
    comm = E.vp()
    _ffi.check(lib.nvllm_comm_init(C.byref(comm), bytes(uid.cpu().tolist()), rank, world))
    one = (max(t * h for _, t, h in SHAPES) * 2 + 255) // 256 * 256
    base = E.vp()
    _ffi.check(lib.nvllm_comm_symm_alloc(comm, 3 * one, C.byref(base)))
    p_h, p_out, p_x = base.value, base.value + one, base.value + 2 * one
    s = torch.cuda.Stream()
    st = C.c_void_p(s.cuda_stream)
    cap = 64 * 1024
    tl = torch.zeros((cap, 12), dtype=torch.int64, device="cuda")
    for name, T, hid in SHAPES:
        w = torch.ones(hid, dtype=torch.bfloat16, device="cuda")
        call = lambda: _ffi.check(lib.nvllm_allreduce_add_rmsnorm(comm, p_out, p_x, p_h, w.data_ptr(), T, hid, 1e-6, 0, st))
        with torch.cuda.stream(s):
            call(); torch.cuda.synchronize(); dist.barrier()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=s):
                for _ in range(a.calls):
                    call()
            for _ in range(3):
                g.replay()
            torch.cuda.synchronize(); dist.barrier()
            tl.zero_()
            assert lib.nvllmdbg_ar_timeline(tl.data_ptr(), cap) == 0
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s); g.replay(); e1.record(s)
            torch.cuda.synchronize(); dist.barrier()
            assert lib.nvllmdbg_ar_timeline(None, 0) == 0
        us = e0.elapsed_time(e1) * 1e3 / a.calls
        t = tl.cpu().numpy()
        t = t[t[:, 1] != 0]
        ctas = int(t[:, 0].max()) + 1
        calls = len(t) // ctas
        L = np.zeros((calls, ctas, 8), dtype=np.int64)
        own = np.zeros((calls, ctas), dtype=np.int64)
        for b in range(ctas):
            rows = t[t[:, 0] == b]
            rows = rows[np.argsort(rows[:, 1])]
            for i in range(min(calls, len(rows))):
                L[i, b] = rows[i, 1:9]
                own[i, b] = rows[i, 10]
        sel = slice(2, calls)
        o = own[sel] > 0
        med = lambda x: float(np.median(x)) if x.size else None
        d = lambda hi, lo, m=None: med(((L[sel, :, hi] - L[sel, :, lo]).astype(np.float64))[m] if m is not None else (L[sel, :, hi] - L[sel, :, lo]).astype(np.float64))
        start, end = L[:, :, 1].min(axis=1), L[:, :, 7].max(axis=1)
        out = {"shape": name, "world": world, "rank": rank, "ctas": ctas, "calls": calls, "us_per_call_in_graph": round(us, 2),
               "wait_ns": d(1, 0), "push_ns": d(2, 1), "clear_ns": d(3, 2),
               "owners": {"n": int(o[0].sum()), "arrive_ns": d(4, 3, o), "reduce_bcast_ns": d(5, 4, o), "gather_ns": d(6, 5, o)},
               "others": {"gather_ns": d(6, 3, ~o)},
               "cta_lifetime_after_wait_ns": d(7, 1), "call_span_after_wait_ns": med((end - start)[2:].astype(np.float64)),
               "call_period_ns": med(np.diff(start).astype(np.float64)),
               "gap_prev_end_to_next_ready_ns": med((start[1:] - end[:-1]).astype(np.float64))}
        outs = [None] * world
        dist.all_gather_object(outs, out)
        if rank == 0:
            for x in outs:
                print(json.dumps(x), flush=True)
    dist.barrier()
    _ffi.check(lib.nvllm_comm_destroy(comm))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
