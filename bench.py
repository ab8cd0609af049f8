#!/usr/bin/env python
"""bench.py — the driver's measurement contract for the paged-attention decode hot path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--model ...]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...      (N > 1, one rank per GPU)

A "step" is one decode step of the whole model over one batch (embedding -> 28/36 layers of norm / qkv GEMM / qk-norm
+ RoPE + KV store / split-K paged decode attention / o_proj / MLP -> LM head -> greedy argmax).
  N == 1 : BASELINE.json configs[1] — Qwen3-0.6B bf16, 256 concurrent seqs, paged KV block 256, 512-token prompts +
           512 decode steps.  The K timed steps are centred on the workload's mean context (768 tokens): the synthetic
           prompts are 768 - W - K/2 tokens long (the cost of a step depends on the context only through its length);
           with K >= 480 the prompts are exactly 512 tokens and the run IS the workload.
  N  > 1 : BASELINE.json configs[2] — Qwen3-8B bf16 tensor-parallel TP=N, 512 seqs x (1K prompt + 1K decode), timed
           window centred on context 1536; "scaling": "strong" (same job on more GPUs).
value  : decode tokens/s with the step's inputs already resident in HBM (CUDA events around K replays of the step).
e2e    : the same metric through the public API (LLMEngine.step: schedule -> H2D metadata -> graph -> D2H tokens ->
         postprocess) — host buffers in, host tokens out, wall clock between device synchronisations, max over ranks.
roofline: the dominant kernel (paged decode attention), algorithmic bytes / CUDA-event time of the kernel re-launched
         on the live batch, cycling through the layers' caches (inputs >> L2), vs MEASURED_PEAKS.json.
cpu_baseline / --impl reference: the CPU oracle (numpy port of the reference's algorithm; the reference itself cannot
         be built — DESIGN.md) timed on the host cores on a bounded sample of the same workload."""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

MODELS = {"qwen3-0.6b": dict(seqs=256, prompt=512, decode=512, blocks_per_seq=4),
          "qwen3-8b": dict(seqs=512, prompt=1024, decode=1024, blocks_per_seq=8),
          # BASELINE.json configs[4]: Qwen3-32B TP=8, 64 seqs at 32K context (long-context split-K paged decode stress)
          "qwen3-32b": dict(seqs=64, prompt=32256, decode=512, blocks_per_seq=129)}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default=None, help="override the workload model (qwen3-0.6b | qwen3-8b | qwen3-32b)")
    ap.add_argument("--seqs", type=int, default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--eager", action="store_true", help="enforce_eager (no CUDA graphs)")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = sorted(int(r[1]) for r in self.rows if len(r) >= 9 and r[1].isdigit())
        mx = max([int(r[2]) for r in self.rows if len(r) >= 9 and r[2].isdigit()] or [0])
        reasons = set()
        for r in self.rows:
            if len(r) < 9:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ CPU oracle arm
def cpu_decode_sample(model, ctx, sample_seqs, steps, warmup, budget_s=None):
    """Times the CPU oracle (oracle/qwen3.py: numpy restatement of the reference algorithm — per-sequence block gather +
    dense softmax attention, f32) on `sample_seqs` sequences of the workload at context `ctx`; weights and KV are random
    (timing only; drawn from one 16 Mi-element pool repeated cyclically, so that setting up 8B parameters takes seconds).
    `budget_s` bounds the timed steps (at least one runs).  Returns (tokens/s, cores, description, ms/step)."""
    import numpy as np
    from oracle import host as H
    from oracle import qwen3 as Q
    cfg = Q.arch_config(model)
    rng = np.random.default_rng(0)
    bs = 256
    nblk_seq = (ctx + steps + warmup + bs) // bs + 1
    w = {}
    shapes = Q.seeded_weights.__doc__  # noqa: F841 (ids documented there)
    H_, Hkv, D, hid, I, V = (cfg[k] for k in ("num_attention_heads", "num_key_value_heads", "head_dim", "hidden_size",
                                              "intermediate_size", "vocab_size"))
    pool = rng.standard_normal(1 << 24, dtype=np.float32) * np.float32(0.02)
    rnd = lambda *s: np.resize(pool, s)
    w["model.embed_tokens.weight"] = rnd(V, hid)
    w["model.norm.weight"] = np.ones(hid, np.float32)
    if not cfg["tie_word_embeddings"]:
        w["lm_head.weight"] = rnd(V, hid)
    for l in range(cfg["num_hidden_layers"]):
        p = f"model.layers.{l}."
        for n, shp in (("input_layernorm", (hid,)), ("post_attention_layernorm", (hid,)), ("self_attn.q_norm", (D,)),
                       ("self_attn.k_norm", (D,))):
            w[p + n + ".weight"] = np.ones(shp, np.float32)
        for n, shp in (("self_attn.q_proj", (H_ * D, hid)), ("self_attn.k_proj", (Hkv * D, hid)), ("self_attn.v_proj", (Hkv * D, hid)),
                       ("self_attn.o_proj", (hid, H_ * D)), ("mlp.gate_proj", (I, hid)), ("mlp.up_proj", (I, hid)),
                       ("mlp.down_proj", (hid, I))):
            w[p + n + ".weight"] = rnd(*shp)
    m = Q.Qwen3Oracle(cfg, w, sample_seqs * nblk_seq, bs)
    for l in range(cfg["num_hidden_layers"]):
        m.k_cache[l][:] = np.resize(pool, m.k_cache[l].shape) * np.float32(50.0)
        m.v_cache[l][:] = np.resize(pool[::-1], m.v_cache[l].shape) * np.float32(50.0)
    seqs = []
    for i in range(sample_seqs):
        s = H.Sequence(rng.integers(0, V, ctx).tolist(), H.SamplingParams(temperature=0.0, max_tokens=1 << 30), bs)
        s.block_table = list(range(i * nblk_seq, (i + 1) * nblk_seq))[: s.num_blocks()]
        seqs.append(s)
    times = []
    for it in range(warmup + steps):
        for i, s in enumerate(seqs):  # keep the block tables long enough as the context grows
            while len(s.block_table) < s.num_blocks():
                s.block_table.append(i * nblk_seq + len(s.block_table))
        t0 = time.perf_counter()
        logits = m.forward(H.prepare_decode(seqs, bs))
        toks = np.argmax(logits, -1)
        dt = time.perf_counter() - t0
        if it >= warmup:
            times.append(dt)
        for s, t in zip(seqs, toks):
            s.append_token(int(t))
        if budget_s is not None and times and sum(times) + times[-1] > budget_s:
            break
    cores = os.cpu_count() or 1
    tps = sample_seqs * len(times) / sum(times)
    desc = (f"{len(times)} decode step(s) of {sample_seqs} of the workload's sequences at context {ctx} "
            f"({model}, f32 numpy oracle, BLAS threads = {cores})")
    return tps, cores, desc, sum(times) / len(times) * 1e3


def tensor_kernels(peak_tflops):
    """Secondary rooflines (north_star (c)): the tensor-core kernels of the workload's PREFILL step, timed live with CUDA
    events on torch's stream at the shapes of configs[1] (64 sequences x 512 tokens per prefill step of Qwen3-0.6B).
    Returns None on any failure — the headline line never depends on it."""
    try:
        import torch
        from nano_vllm_rs_b200 import ops

        def ms_of(fn, iters=10):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                fn()
            e1.record(); torch.cuda.synchronize()
            return e0.elapsed_time(e1) / iters
        out = {}
        B, L, H, Hkv, D = 64, 512, 16, 8, 128
        q = torch.randn((B * L, H, D), device="cuda").to(torch.bfloat16)
        k = torch.randn((B * L, Hkv, D), device="cuda").to(torch.bfloat16)
        v = torch.randn((B * L, Hkv, D), device="cuda").to(torch.bfloat16)
        cu = torch.arange(0, B * L + 1, L, dtype=torch.int32, device="cuda")
        o = torch.empty_like(q)
        ms = ms_of(lambda: ops.varlen_prefill_attention(q, k, v, cu, cu, L, L, out=o))
        tf = 4.0 * H * D * B * (L * (L + 1) / 2) / ms / 1e9
        out["prefill_attention"] = {"shape": "64 x 512 tokens, 16/8 heads, causal", "ms": round(ms, 4), "achieved": round(tf, 1),
                                    "peak": peak_tflops, "unit": "TFLOP/s", "frac": round(tf / peak_tflops, 4)}
        M, N, K = B * L, 6144, 1024                                  # gate_up projection of the prefill step
        x = torch.randn((M, K), device="cuda").to(torch.bfloat16)
        w = (torch.randn((N, K), device="cuda") * 0.02).to(torch.bfloat16)
        y = torch.empty((M, N), dtype=torch.bfloat16, device="cuda")
        ms = ms_of(lambda: ops.gemm_bf16(x, w, out=y))
        tf = 2.0 * M * N * K / ms / 1e9
        out["prefill_gemm"] = {"shape": f"{M} x {N} x {K}", "ms": round(ms, 4), "achieved": round(tf, 1), "peak": peak_tflops,
                               "unit": "TFLOP/s", "frac": round(tf / peak_tflops, 4)}
        return out
    except Exception as ex:  # noqa: BLE001
        return {"error": str(ex)[:200]}


def workload(args):
    model = args.model or ("qwen3-0.6b" if args.gpus == 1 else "qwen3-8b")
    wl = dict(MODELS[model])
    if args.seqs:
        wl["seqs"] = args.seqs
    K, W = args.steps, args.warmup
    mid = wl["prompt"] + wl["decode"] // 2
    if K >= wl["decode"] - 32:   # the run is the workload itself
        prompt = wl["prompt"]
        K = min(K, wl["decode"] - W - 1)
    else:
        prompt = max(wl["prompt"], mid - W - K // 2)
    return model, wl, prompt, K, W


def workload_name(model, wl):
    return (f"{model} bf16 decode: {wl['seqs']} concurrent seqs, paged KV block 256, "
            f"{wl['prompt']}-token prompts + {wl['decode']} decode steps")


def reference_arm(args):
    """bench.py --impl reference: the reference's own CPU implementation of the path.  The Rust crate cannot be built
    here (no toolchain, invalid manifest, type errors on the path — DESIGN.md), so this times the oracle port."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    model, wl, prompt, K, W = workload(args)
    ctx = prompt + W + K // 2
    # bounded sample: the whole arm ends within a few minutes whatever --steps says (about 60 s of timed steps)
    big = model != "qwen3-0.6b"
    sample = 8 if big else (32 if K <= 8 else 16)
    K_cpu, W_cpu = max(1, min(K, 64)), 1
    tps, cores, desc, ms = cpu_decode_sample(model, ctx, sample, K_cpu, W_cpu, budget_s=60.0)
    K_cpu = int(desc.split(" decode step")[0])   # steps actually timed within the budget
    line = {"impl": "reference", "metric": "decode_tokens_per_s", "value": round(tps, 2), "unit": "tokens/s", "n_gpus": args.gpus,
            "steps": K_cpu, "warmup": W_cpu, "ms_per_step": round(ms, 2), "higher_is_better": True,
            "scaling": "strong" if args.gpus > 1 else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(model, wl) + " — CPU oracle port of the reference algorithm (f32) on a "
                                   "bounded sample of the workload's sequences", "sample": desc,
                       "parallelism": "host cores (the reference has no GPU or multi-GPU path that runs)"},
            "cpu_baseline": {"value": round(tps, 2), "unit": "tokens/s", "cores": cores, "kind": "port", "sample": desc},
            "e2e": {"value": round(tps, 2), "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    if args.impl == "reference":
        return reference_arm(args)
    import numpy as np
    import torch

    from nano_vllm_rs_b200 import _ffi
    from nano_vllm_rs_b200 import engine as E

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == args.gpus or world == 1 and args.gpus == 1, f"--gpus {args.gpus} but WORLD_SIZE={world}"
    torch.cuda.set_device(local)
    lib = E.lib()
    assert lib.nvllm_device_ok() == 1, "no sm_100 device: the product path has no fallback"
    comm = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            buf = (E.C.c_uint8 * 128)()
            _ffi.check(lib.nvllm_comm_unique_id(buf))
            uid.copy_(torch.tensor(list(buf), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        raw = bytes(uid.cpu().tolist())
        h = E.vp()
        _ffi.check(lib.nvllm_comm_init(E.C.byref(h), raw, rank, world))
        comm = h

    model, wl, prompt, K, W = workload(args)
    B = wl["seqs"]
    max_len = prompt + W + K + 8
    nblk = B * ((max_len + 255) // 256) + 8
    cfg = E.Config(model, tensor_parallel_size=world, tp_rank=rank, max_num_seqs=max(B, 8), num_kvcache_blocks=nblk,
                   max_model_len=max(4096, ((max_len + 255) // 256) * 256), enforce_eager=int(args.eager),
                   max_position_embeddings=max(40960, ((max_len + 255) // 256) * 256))
    eng = E.LLMEngine(cfg, comm=comm, seed=1234, std=0.02)
    rng = np.random.default_rng(1234)                      # token stream seed 1234, ids mod vocab
    sp = E.SamplingParams(temperature=0.0, max_tokens=W + 2 * K + 64, ignore_eos=True)
    for _ in range(B):
        eng.add_request(rng.integers(0, cfg.vocab_size, prompt), sp)
    t0 = time.perf_counter()
    n_prefill = 0
    while eng.scheduler.get_stats()["waiting"] > 0:         # prefill-first scheduling: all prompts, untimed
        eng.step(); n_prefill += 1
    prefill_s = time.perf_counter() - t0
    for _ in range(max(W, 3)):                             # warm-up decode steps (step 1 eager, step 2 captures the graph)
        eng.step()
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        return float(t.item())

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    # ---- e2e: K steps through the public API (host buffers in, host tokens out) -------------------------------
    ctx_first = eng.scheduler.sequence(0)["token_ids"].__len__()
    barrier()
    torch.cuda.cudart().cudaProfilerStart()   # ncu --profile-from-start off: capture only the timed steps
    t0 = time.perf_counter()
    for _ in range(K):
        is_prefill, nseq, _ = eng.step()
        assert not is_prefill and nseq == B
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    torch.cuda.cudart().cudaProfilerStop()
    barrier()
    ctx_last = len(eng.scheduler.sequence(0)["token_ids"]) - 1
    launches = eng.model_runner.last_launches()
    # ---- value: K replays of the (last) step with inputs resident in HBM, CUDA events on the runner's stream -----
    barrier()
    dev_ms = max_over_ranks(eng.model_runner.replay_decode(K))
    barrier()
    # ---- roofline of the dominant kernel, on the live batch ------------------------------------------------------
    att_ms, att_bytes = eng.model_runner.time_decode_attention(iters=4 * cfg.num_hidden_layers)
    clocks = sampler.stop() if rank == 0 else None
    if rank != 0:
        return
    peak, peak_src = peaks()
    achieved = att_bytes / (att_ms * 1e-3) / 1e9
    layers = cfg.num_hidden_layers
    traffic, traffic_capture = None, None
    prof = os.path.join(ROOT, "profiles", "decode_attention_ncu.json")  # dram bytes/launch from the committed ncu capture
    if os.path.exists(prof):
        try:
            pj = json.load(open(prof))
            if pj.get("model") == model and pj.get("batch") == B:
                traffic = pj.get("dram_bytes_per_launch")
                # the capture ran at its own context length: its algorithmic bytes are the ones to compare traffic with
                traffic_capture = {k: pj.get(k) for k in ("context", "alg_bytes_per_launch", "ratio")}
        except Exception:
            pass
    cap = cfg.max_model_len // 256
    line = {
        "metric": "decode_tokens_per_s", "value": round(B * K / (dev_ms * 1e-3), 1), "unit": "tokens/s", "n_gpus": world,
        "steps": K, "warmup": max(W, 3), "ms_per_step": round(dev_ms / K, 4), "higher_is_better": True,
        "scaling": "strong" if world > 1 else "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload_name(model, wl) + f"; timed steps at context "
                               f"{ctx_first}..{ctx_last} (workload mean {wl['prompt'] + wl['decode'] // 2})",
                   "seqs": B, "prompt_len_used": prompt, "parallelism": f"tp{world}",
                   "cuda_graphs": not args.eager, "l2_note": "inputs larger than L2: one step reads "
                   f"{round(att_bytes * layers / 1e9, 1)} GB of KV (L2 = 126 MB), no flush needed",
                   "prefill_steps_untimed": n_prefill, "prefill_s": round(prefill_s, 2)},
        "e2e": {"value": round(B * K / e2e_s, 1), "unit": "tokens/s", "ms_per_step": round(e2e_s / K * 1e3, 4),
                "h2d_bytes_per_step": B * (24 + 4 * cap), "d2h_bytes_per_step": 8 * B,
                "api": "LLMEngine.step (schedule -> H2D metadata -> decode graph -> D2H tokens -> postprocess)"},
        "gpu_launches": int(launches) * K,
        "launches_per_step": int(launches),
        "roofline": {"kernel": "paged_decode_kernel", "bound": "hbm", "achieved": round(achieved, 1), "peak": peak,
                     "unit": "GB/s", "frac": round(achieved / peak, 4), "frac_of_8TBps_nominal": round(achieved / 8000.0, 4),
                     "peak_source": peak_src, "traffic": traffic, "traffic_capture": traffic_capture,
                     "alg_bytes_per_launch": int(att_bytes),
                     "ms_per_launch": round(att_ms, 5),
                     "share_of_step": round(att_ms * layers / (dev_ms / K), 4),
                     "how": "CUDA events on the runner's stream around re-launches of the kernel on the live batch, "
                            "cycling through the layers' KV caches"},
        "clocks": clocks,
    }
    if world == 1:
        try:
            pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("bf16_tflops", 1657.6)
        except Exception:
            pk = 1657.6
        line["tensor_kernels"] = tensor_kernels(pk)
    if world == 1 and not args.no_cpu_baseline:
        ctx_mid = (ctx_first + ctx_last) // 2
        tps, cores, desc, _ = cpu_decode_sample(model, ctx_mid, 16, 2, 1)
        line["cpu_baseline"] = {"value": round(tps, 2), "unit": "tokens/s", "cores": cores, "kind": "port", "sample": desc}
    print(json.dumps(line), flush=True)
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
