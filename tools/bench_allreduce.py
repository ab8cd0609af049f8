"""In-graph cost of the fused TP epilogue (C1: all-reduce + residual add + RMSNorm over NVLink peer memory) against
ncclAllReduce + the stand-alone add_rmsnorm kernel, at the message sizes of the decode step.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/bench_allreduce.py

`calls` launches are captured back to back into one CUDA graph (PDL edges kept — how the decode step issues them, minus the
GEMMs in between) and the graph is replayed; us/call = replay time / calls, max over ranks.  Also checks the fused result
against a plain f32 torch computation of the same sum (every rank, every row)."""
import argparse
import ctypes as C
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nano_vllm_rs_b200 import _ffi  # noqa: E402
from nano_vllm_rs_b200 import engine as E  # noqa: E402

SHAPES = [("8B decode, 512 seqs", 512, 4096), ("8B decode, 256 seqs", 256, 4096), ("32B decode, 64 seqs", 64, 5120),
          ("8B prefill chunk, 8192 tokens", 8192, 4096)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--calls", type=int, default=72)
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = E.lib()
    uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        buf = (C.c_uint8 * 128)()
        _ffi.check(lib.nvllm_comm_unique_id(buf))
        uid.copy_(torch.tensor(list(buf), dtype=torch.uint8))
    dist.broadcast(uid, 0)
    comm = E.vp()
    _ffi.check(lib.nvllm_comm_init(C.byref(comm), bytes(uid.cpu().tolist()), rank, world))
    max_elems = max(t * h for _, t, h in SHAPES)
    one = (max_elems * 2 + 255) // 256 * 256
    base = E.vp()
    _ffi.check(lib.nvllm_comm_symm_alloc(comm, 3 * one, C.byref(base)))
    p_h, p_out, p_x = base.value, base.value + one, base.value + 2 * one

    s = torch.cuda.Stream()
    st = C.c_void_p(s.cuda_stream)
    rt = C.CDLL("libcudart.so.12")   # the runtime libnvllm.so links against (already mapped)
    rt.cudaMemcpy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]

    def memcpy(dst, src, nbytes):   # device -> device
        assert rt.cudaMemcpy(dst, src, nbytes, 3) == 0
    for name, T, hid in SHAPES:
        n = T * hid
        g = torch.Generator(device="cuda"); g.manual_seed(100 + rank)
        x = (torch.randn(n, generator=g, device="cuda") * 0.5).to(torch.bfloat16)
        g2 = torch.Generator(device="cuda"); g2.manual_seed(7)
        res = torch.randn(n, generator=g2, device="cuda").to(torch.bfloat16)      # same residual on every rank
        w = (1 + 0.1 * torch.randn(hid, generator=g2, device="cuda")).to(torch.bfloat16)
        # ---- correctness: one fused call vs f32 torch (before and after the graph replays, both protocols) -------------
        xs = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(xs, x)
        tot = torch.zeros(n, dtype=torch.float32, device="cuda")
        for t in xs:
            tot += t.float()
        hnew = (tot.to(torch.bfloat16).float() + res.float()).to(torch.bfloat16).float().view(T, hid)
        ref = (hnew * torch.rsqrt((hnew * hnew).mean(-1, keepdim=True) + 1e-6) * w.float()).to(torch.bfloat16)

        def check():
            memcpy(p_x, x.data_ptr(), n * 2)
            memcpy(p_h, res.data_ptr(), n * 2)
            torch.cuda.synchronize(); dist.barrier()
            with torch.cuda.stream(s):
                _ffi.check(lib.nvllm_allreduce_add_rmsnorm(comm, p_out, p_x, p_h, w.data_ptr(), T, hid, 1e-6, 0, st))
            torch.cuda.synchronize(); dist.barrier()
            out = torch.empty(n, dtype=torch.bfloat16, device="cuda")
            memcpy(out.data_ptr(), p_out, n * 2)
            e = torch.tensor([(out.view(T, hid).float() - ref.float()).abs().max().item()], device="cuda")
            dist.all_reduce(e, op=dist.ReduceOp.MAX)
            return float(e.item())
        # ---- timing --------------------------------------------------------------------------------------------------
        def timed(fn):
            with torch.cuda.stream(s):
                fn(); torch.cuda.synchronize(); dist.barrier()
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr, stream=s):
                    for _ in range(a.calls):
                        fn()
                gr.replay(); torch.cuda.synchronize(); dist.barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(s)
                for _ in range(a.reps):
                    gr.replay()
                e1.record(s)
                torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) * 1e3 / (a.reps * a.calls)], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        call = lambda: _ffi.check(lib.nvllm_allreduce_add_rmsnorm(comm, p_out, p_x, p_h, w.data_ptr(), T, hid, 1e-6, 0, st))
        fused, errs = {}, {}
        for proto, key in (("1", "low_latency"), ("0", "flags")):
            os.environ["NVLLM_AR_LL"] = proto
            lib.nvllm_reload_env()
            e0 = check()
            fused[key] = timed(call)
            errs[key] = max(e0, check())   # 1 bf16 ulp at |v| < 8 (f32 summation order differs from torch's)
        os.environ["NVLLM_AR_LL"] = "1"
        lib.nvllm_reload_env()
        err = max(errs.values())
        ok = err <= 0.0625 and lib.nvllm_comm_error(comm) == 0
        y = torch.empty(n, dtype=torch.bfloat16, device="cuda")

        def nccl_path():
            _ffi.check(lib.nvllm_allreduce(comm, x.data_ptr(), n, 0, st))
            _ffi.check(lib.nvllm_add_rmsnorm(y.data_ptr(), x.data_ptr(), res.data_ptr(), w.data_ptr(), T, hid, 1e-6, 0, st))
        nccl = timed(nccl_path)
        if rank == 0:
            msg = T * hid * 2
            print(json.dumps({"shape": name, "tokens": T, "hidden": hid, "world": world, "message_bytes": msg,
                              "fused_us_per_call": round(min(fused.values()), 2),
                              "fused_low_latency_us_per_call": round(fused["low_latency"], 2),
                              "fused_flags_us_per_call": round(fused["flags"], 2), "nccl_plus_norm_us_per_call": round(nccl, 2),
                              "nvlink_bytes_per_rank_per_call": 2 * (world - 1) * msg // world,
                              "fused_nvlink_GBps_per_dir": round(2 * (world - 1) * msg / world / min(fused.values()) / 1e3, 1),
                              "low_latency_applies": world * ((T + world - 1) // world) * hid * 2 <= (8 << 20),
                              "max_abs_err_vs_f32_torch": err, "ok": ok, "calls_in_graph": a.calls}), flush=True)
        assert ok, (name, err)
    dist.barrier()
    _ffi.check(lib.nvllm_comm_destroy(comm))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
