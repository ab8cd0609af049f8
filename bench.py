#!/usr/bin/env python
"""bench.py — the driver's measurement contract for the paged-attention decode hot path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--model ...] [--workload decode|prefix]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...      (N > 1, one rank per GPU)

A "step" is one decode step of the whole model over one batch (embedding -> 28/36 layers of norm / qkv GEMM / qk-norm
+ RoPE + KV store / split-K paged decode attention / o_proj / MLP -> LM head -> greedy argmax).
  N == 1 : BASELINE.json configs[1] — Qwen3-0.6B bf16, 256 concurrent seqs, paged KV block 256, 512-token prompts +
           512 decode steps.  The K timed steps are centred on the workload's mean context (768 tokens): the synthetic
           prompts are 768 - W - K/2 tokens long (the cost of a step depends on the context only through its length);
           with K >= 480 the prompts are exactly 512 tokens and the run IS the workload.
  N  > 1 : BASELINE.json configs[2] — Qwen3-8B bf16 tensor-parallel TP=N, 512 seqs x (1K prompt + 1K decode), timed
           window centred on context 1536; "scaling": "strong" (same job on more GPUs; the 1-GPU point of the same job is
           `--gpus 1 --model qwen3-8b`, kept in profiles/ and quoted in every N > 1 line).
  --workload prefix : BASELINE.json configs[3] — Qwen3-8B, 256 seqs sharing a 4096-token prefix (+ 256 unique tokens),
           N = 1: prefix-cached varlen prefill (block dedup, paged-prefix attention) then decode at context 4353+.
value  : decode tokens/s with the step's inputs already resident in HBM (CUDA events around K replays of the step).
e2e    : the same metric through the public API (LLMEngine.step: schedule -> H2D metadata -> graph -> D2H tokens ->
         postprocess) — host buffers in, host tokens out, wall clock between device synchronisations, max over ranks.
roofline: the dominant kernel (paged decode attention), algorithmic bytes / CUDA-event time of the kernel re-launched
         on the live batch, cycling through the layers' caches (inputs >> L2), vs MEASURED_PEAKS.json.
parity : checked INSIDE the run, before the timed region, exit code 1 on failure —
         N == 1: the logits of sequence 0 of the first prefill batch of the benchmarked workload vs the CPU oracle;
         N  > 1: `tp_parity` — a 2-layer Qwen3 through ModelRunner(comm) with the fused all-reduce epilogue vs the TP=1
         oracle (vocab-sharded logits gathered on rank 0), and the sampled tokens of every warm-up step of the real
         workload compared across ranks.
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
# BASELINE.json configs[3]: 256 seqs = one shared 4096-token prefix (seed 4321) + 256 unique tokens each
PREFIX = dict(model="qwen3-8b", seqs=256, shared=4096, unique=256)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default=None, help="override the workload model (qwen3-0.6b | qwen3-8b | qwen3-32b)")
    ap.add_argument("--workload", default="decode", choices=["decode", "prefix"])
    ap.add_argument("--seqs", type=int, default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the in-run parity checks (profiling runs)")
    ap.add_argument("--eager", action="store_true", help="enforce_eager (no CUDA graphs)")
    ap.add_argument("--profile", action="store_true",
                    help="per-launch CUDA-event profile of eager decode steps (implies --eager); adds `launch_profile`")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def tensor_peak():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("bf16_tflops", 1590.0), "measured"
    except Exception:
        return 1590.0, "fallback"


_T0 = time.time()


_LAST = {"t": time.time(), "msg": "start"}


def progress(msg):
    """One timestamped line per phase on stderr (every rank): when a run stalls, the last line says where.  Every call
    also feeds the stall watchdog (arm_watchdog)."""
    _LAST["t"], _LAST["msg"] = time.time(), msg
    print(f"[bench rank {os.environ.get('RANK', '0')} +{time.time() - _T0:6.1f}s] {msg}", file=sys.stderr, flush=True)


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
def cpu_decode_sample(model, ctx, seqs, steps, warmup, budget_s=None, distinct=None):
    """Times the CPU oracle (oracle/qwen3.py: numpy restatement of the reference algorithm — per-sequence block gather +
    dense softmax attention, f32) on a decode batch of `seqs` sequences at context `ctx`; weights and KV are random
    (timing only; drawn from one 16 Mi-element pool repeated cyclically, so that setting up 8B parameters takes seconds).
    `distinct` (default: all) bounds the f32 KV cache held in host RAM: sequence i uses the paged blocks of sequence
    i % distinct — every sequence still gathers and attends over its full context (the work of the step is unchanged;
    the CPU merely sees a warmer cache, which favours the baseline).  `budget_s` bounds the timed steps (at least one
    runs).  Returns (tokens/s, cores, description, ms/step)."""
    import numpy as np
    from oracle import host as H
    from oracle import qwen3 as Q
    cfg = Q.arch_config(model)
    rng = np.random.default_rng(0)
    bs = 256
    distinct = min(seqs, distinct or seqs)
    nblk_seq = (ctx + steps + warmup + bs) // bs + 1
    w = {}
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
    m = Q.Qwen3Oracle(cfg, w, distinct * nblk_seq, bs)
    for l in range(cfg["num_hidden_layers"]):
        m.k_cache[l][:] = np.resize(pool, m.k_cache[l].shape) * np.float32(50.0)
        m.v_cache[l][:] = np.resize(pool[::-1], m.v_cache[l].shape) * np.float32(50.0)
    sq = []
    for i in range(seqs):
        s = H.Sequence(rng.integers(0, V, ctx).tolist(), H.SamplingParams(temperature=0.0, max_tokens=1 << 30), bs)
        s.block_table = list(range((i % distinct) * nblk_seq, (i % distinct + 1) * nblk_seq))[: s.num_blocks()]
        sq.append(s)
    times = []
    for it in range(warmup + steps):
        for i, s in enumerate(sq):  # keep the block tables long enough as the context grows
            while len(s.block_table) < s.num_blocks():
                s.block_table.append((i % distinct) * nblk_seq + len(s.block_table))
        t0 = time.perf_counter()
        logits = m.forward(H.prepare_decode(sq, bs))
        toks = np.argmax(logits, -1)
        dt = time.perf_counter() - t0
        if it >= warmup:
            times.append(dt)
        for s, t in zip(sq, toks):
            s.append_token(int(t))
        if budget_s is not None and times and sum(times) + times[-1] > budget_s:
            break
    cores = os.cpu_count() or 1
    tps = seqs * len(times) / sum(times)
    share = "" if distinct == seqs else f" (f32 KV of {distinct} distinct sequences held in host RAM, reused cyclically)"
    desc = (f"{len(times)} decode step(s) of {seqs} sequences at context {ctx}{share} "
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
    here (no toolchain, invalid manifest, type errors on the path — DESIGN.md), so this times the oracle port, on the SAME
    workload string as our arm: the full decode batch of configs[1] (256 sequences) for 2 timed steps after 1 warm-up
    (~10 s each); for the 8B / 32B workloads a bounded number of the batch's sequences (stated in `sample`)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    model, wl, prompt, K, W = workload(args)
    ctx = prompt + max(W, 3) + K // 2
    if model == "qwen3-0.6b":
        seqs, distinct, K_cpu = wl["seqs"], 16, 2           # the full batch; f32 KV of 16 sequences (3.7 GB) reused cyclically
    else:
        seqs, distinct, K_cpu = 8, 8, 2                      # 8B / 32B: f32 weights alone are 32+ GB of host RAM
    W_cpu = 1
    tps, cores, desc, ms = cpu_decode_sample(model, ctx, seqs, K_cpu, W_cpu, budget_s=150.0, distinct=distinct)
    K_cpu = int(desc.split(" decode step")[0])   # steps actually timed within the budget
    line = {"impl": "reference", "metric": "decode_tokens_per_s", "value": round(tps, 2), "unit": "tokens/s", "n_gpus": args.gpus,
            "steps": K_cpu, "warmup": W_cpu, "ms_per_step": round(ms, 2), "higher_is_better": True,
            "scaling": "strong" if args.gpus > 1 else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(model, wl), "seqs": seqs, "sample": desc,
                       "what": "CPU oracle port of the reference algorithm (f32 numpy; the Rust crate cannot be built)",
                       "parallelism": "host cores (the reference has no GPU or multi-GPU path that runs)"},
            "cpu_baseline": {"value": round(tps, 2), "unit": "tokens/s", "cores": cores, "kind": "port", "sample": desc},
            "e2e": {"value": round(tps, 2), "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ in-run parity checks
# Logit tolerance of tests/test_gpu_engine.py (activations are bf16 at every tensor boundary on both sides; accumulation
# order differs): |d| <= 0.035 * sqrt(layers) * sigma(row) + 1.6e-2 * |ref|.  Under tensor parallelism every rank's
# partial sum of o_proj / down_proj is rounded to bf16 before the all-reduce (a bf16 tensor per rank, linear.rs:228-239):
# + 0.05 sigma (tests/test_tp_gloo.py measures 0.08 sigma for the same effect on the CPU).
LOGIT_SIGMA_PER_SQRT_LAYER, LOGIT_RTOL, TP_EXTRA_SIGMA = 0.035, 1.6e-2, 0.05


def _to_inputs(b):
    from oracle import host as H
    return H.StepInputs(b.is_prefill, b.input_ids, b.positions, b.slot_mapping, b.cu_seqlens_q, b.cu_seqlens_k,
                        b.max_seqlen_q, b.max_seqlen_k, b.context_lens, b.block_tables)


def make_comm(E, _ffi, lib, rank, world, torch):
    """One nvllm communicator per call: rank 0 draws the NCCL unique id, torch.distributed carries it (plumbing only)."""
    import torch.distributed as dist
    uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        buf = (E.C.c_uint8 * 128)()
        _ffi.check(lib.nvllm_comm_unique_id(buf))
        uid.copy_(torch.tensor(list(buf), dtype=torch.uint8))
    dist.broadcast(uid, 0)
    h = E.vp()
    _ffi.check(lib.nvllm_comm_init(E.C.byref(h), bytes(uid.cpu().tolist()), rank, world))
    return h


def tp_parity_check(E, _ffi, lib, rank, world, torch):
    """The tensor-parallel path against the ORACLE, inside the benchmark run: a 2-layer Qwen3 whose head counts divide by 8
    ("tiny-tp8": hidden 512, 16/8 heads of 128, inter 1024, vocab 2048), seeded weights, prefill + 9 decode steps through
    Scheduler + ModelRunner(comm) — sharded weights, the fused all-reduce + add + RMSNorm epilogue over NVLink peer memory
    (eager, then captured, then replayed as a CUDA graph), vocab-sharded LM head, (max, id) combine — vs the TP=1 numpy
    oracle on the gathered logits.  Returns the `tp_parity` dict on every rank (ok is agreed through a broadcast)."""
    import numpy as np
    import torch.distributed as dist
    comm = make_comm(E, _ffi, lib, rank, world, torch)
    arch, bs, nblk, std, seed = "tiny-tp8", 16, 64, 0.05, 1234
    cfg = E.Config(arch, kvcache_block_size=bs, num_kvcache_blocks=nblk, tensor_parallel_size=world, tp_rank=rank)
    sched, runner = E.Scheduler(cfg), E.ModelRunner(cfg, comm=comm)
    runner.init_weights(seed, std)
    rng = np.random.default_rng(5)
    prompts = [rng.integers(0, 1024, n).tolist() for n in (7, 40, 16, 33)]
    for p in prompts:
        sched.add_sequence(p, E.SamplingParams(temperature=0.0, max_tokens=10, ignore_eos=True))
    oracle = None
    if rank == 0:
        from oracle import qwen3 as Q
        ocfg = Q.arch_config(arch, max_position_embeddings=cfg.max_position_embeddings)
        oracle = Q.Qwen3Oracle(ocfg, Q.seeded_weights(ocfg, seed, std), nblk, bs)
        layers = ocfg["num_hidden_layers"]
    res = dict(vs="oracle (TP=1 numpy restatement, same seeded weights)", arch=arch, world=world, steps=0, decode_steps=0,
               rows=0, strict_tokens=0, near_ties=0, max_err_sigma=0.0, ranks_agree=True, ok=True, why="")
    progress("tp_parity: communicator + runner up, stepping")
    while not sched.is_finished():
        b = sched.schedule()
        toks, logits = runner.execute_model(b, return_logits=True)
        progress(f"tp_parity: step {res['steps']} executed; gathering logits")
        lg = torch.from_numpy(np.ascontiguousarray(logits)).cuda()
        tk = torch.from_numpy(np.ascontiguousarray(toks)).cuda()
        all_lg = [torch.empty_like(lg) for _ in range(world)]
        all_tk = [torch.empty_like(tk) for _ in range(world)]
        dist.all_gather(all_lg, lg)
        dist.all_gather(all_tk, tk)
        if rank == 0:
            if not all(torch.equal(t, all_tk[0]) for t in all_tk):
                res["ranks_agree"], res["ok"], res["why"] = False, False, f"step {res['steps']}: ranks sampled different tokens"
            full = torch.cat(all_lg, dim=1).cpu().numpy()
            ref = oracle.forward(_to_inputs(b))
            sig = ref.std(axis=-1, keepdims=True)
            tol = (LOGIT_SIGMA_PER_SQRT_LAYER * np.sqrt(layers) + TP_EXTRA_SIGMA) * sig + LOGIT_RTOL * np.abs(ref)
            err = np.abs(full - ref)
            res["max_err_sigma"] = max(res["max_err_sigma"], float((err / sig).max()))
            if not (err <= tol).all():
                res["ok"], res["why"] = False, f"step {res['steps']}: logits off by {(err / sig).max():.4f} sigma"
            ref_tok = np.argmax(ref, -1)
            for i in range(len(toks)):
                res["rows"] += 1
                if toks[i] == ref_tok[i]:
                    res["strict_tokens"] += 1
                elif ref[i, ref_tok[i]] - ref[i, toks[i]] <= 2 * tol[i].max():
                    res["near_ties"] += 1       # the oracle itself is within tolerance of a tie between the two candidates
                else:
                    res["ok"], res["why"] = False, f"step {res['steps']} row {i}: token {toks[i]} vs oracle {ref_tok[i]}"
        sched.postprocess(toks)                 # teacher-forced with the GPU's tokens (identical on every rank by construction)
        res["steps"] += 1
        res["decode_steps"] += 0 if b.is_prefill else 1
    res["max_err_sigma"] = round(res["max_err_sigma"], 4)
    res["tolerance"] = (f"|d| <= ({LOGIT_SIGMA_PER_SQRT_LAYER} * sqrt(layers) + {TP_EXTRA_SIGMA}) * sigma + {LOGIT_RTOL} * |ref|; "
                        "tokens equal unless the oracle's own margin is below twice that")
    res["fused_epilogue"] = os.environ.get("NVLLM_FUSED_TP", "1") != "0"
    flag = torch.tensor([1 if res["ok"] else 0], dtype=torch.int32, device="cuda")
    dist.broadcast(flag, 0)
    res["ok"] = bool(flag.item())
    dist.barrier()
    progress("tp_parity: destroying its runner and communicator")
    del runner, sched
    _ffi.check(lib.nvllm_comm_destroy(comm))
    return res


def n1_prefill_parity(eng, model, std, seed):
    """End-to-end parity at the BENCHMARKED shape (N = 1): runs the first prefill batch of the real workload through
    Scheduler + ModelRunner, and compares the f32 logits of its first sequence with the CPU oracle model prefilling that
    prompt alone (same seeded weights).  Returns the `parity` dict; the step's tokens are post-processed as usual."""
    import numpy as np
    from oracle import host as H
    from oracle import qwen3 as Q
    b = eng.scheduler.schedule()
    assert b.is_prefill
    toks, logits = eng.model_runner.execute_model(b, return_logits=True)
    eng.scheduler.postprocess(toks)
    n0 = int(b.cu_seqlens_q[1])
    prompt = b.input_ids[:n0].tolist()
    ocfg = Q.arch_config(model, max_position_embeddings=eng.config.max_position_embeddings)
    t0 = time.perf_counter()
    oracle = Q.Qwen3Oracle(ocfg, Q.seeded_weights(ocfg, seed, std), (n0 + 255) // 256 + 1, 256)
    s = H.Sequence(prompt, H.SamplingParams(temperature=0.0, max_tokens=4), 256)
    s.block_table = list(range(s.num_blocks()))
    ref = oracle.forward(H.prepare_prefill([s], 256))[0]
    layers = ocfg["num_hidden_layers"]
    sig = float(ref.std())
    tol = LOGIT_SIGMA_PER_SQRT_LAYER * np.sqrt(layers) * sig + LOGIT_RTOL * np.abs(ref)
    err = np.abs(logits[0] - ref)
    top = int(np.argmax(ref))
    tok_ok = int(toks[0]) == top or float(ref[top] - ref[int(toks[0])]) <= 2 * float(tol.max())
    return {"vs": "oracle (numpy restatement, same seeded weights)", "what": f"logits of sequence 0 ({n0} tokens) of the first "
            f"prefill batch ({len(b.seq_ids)} seqs, {len(b.input_ids)} tokens) of the benchmarked workload",
            "max_err_sigma": round(float((err / sig).max()), 4), "tolerance_sigma": round(LOGIT_SIGMA_PER_SQRT_LAYER * float(np.sqrt(layers)), 4),
            "token": int(toks[0]), "oracle_token": top, "ok": bool((err <= tol).all() and tok_ok),
            "oracle_s": round(time.perf_counter() - t0, 1)}


# ------------------------------------------------------------------------------------------------ the decode benchmark
def run_decode(args, model, wl, prompt, K, W, rank, world, local, want_parity, want_cpu):
    """Builds the engine for `model` at TP = world, prefills the workload's prompts (untimed), and measures decode steps.
    Returns the JSON line (rank 0) or None (other ranks)."""
    import numpy as np
    import torch

    from nano_vllm_rs_b200 import _ffi
    from nano_vllm_rs_b200 import engine as E
    lib = E.lib()
    comm = make_comm(E, _ffi, lib, rank, world, torch) if world > 1 else None
    B = wl["seqs"]
    W = max(W, 3)
    max_len = prompt + W + K + 8
    nblk = B * ((max_len + 255) // 256) + 8
    eager = int(args.eager or args.profile)
    cfg = E.Config(model, tensor_parallel_size=world, tp_rank=rank, max_num_seqs=max(B, 8), num_kvcache_blocks=nblk,
                   max_model_len=max(4096, ((max_len + 255) // 256) * 256), enforce_eager=eager,
                   max_position_embeddings=max(40960, ((max_len + 255) // 256) * 256))
    seed, std = 1234, 0.02
    progress(f"building the engine: {model}, tp={world}, {B} sequences, {nblk} KV blocks")
    eng = E.LLMEngine(cfg, comm=comm, seed=seed, std=std)
    progress("engine ready (weights initialised, KV cache allocated); prefilling")
    rng = np.random.default_rng(1234)                      # token stream seed 1234, ids mod vocab
    sp = E.SamplingParams(temperature=0.0, max_tokens=W + 2 * K + 64, ignore_eos=True)
    sids = [eng.add_request(rng.integers(0, cfg.vocab_size, prompt), sp) for _ in range(B)]

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

    def ranks_agree_on_last_tokens():
        """Every rank runs its own scheduler on the same request stream: the sampled tokens must be identical everywhere."""
        last = torch.tensor([eng.scheduler.sequence(i)["token_ids"][-1] for i in sids], dtype=torch.int64, device="cuda")
        allt = [torch.empty_like(last) for _ in range(world)]
        torch.distributed.all_gather(allt, last)
        return all(torch.equal(t, allt[0]) for t in allt)

    parity = None
    t0 = time.perf_counter()
    n_prefill = 0
    if want_parity and world == 1 and model == "qwen3-0.6b":
        parity = n1_prefill_parity(eng, model, std, seed)   # runs (and checks) the first prefill batch
        n_prefill += 1
        t0 += parity["oracle_s"]                            # the oracle's CPU time is not prefill time
    while eng.scheduler.get_stats()["waiting"] > 0:         # prefill-first scheduling: all prompts, untimed
        eng.step(); n_prefill += 1
    torch.cuda.synchronize()
    prefill_s = time.perf_counter() - t0
    progress(f"prefill done: {n_prefill} steps in {prefill_s:.1f} s; warm-up decode steps")
    agree_steps, agree = 0, True
    if world > 1 and want_parity:
        agree &= ranks_agree_on_last_tokens(); agree_steps += 1            # the prefill's sampled tokens
    for _ in range(W):                                     # warm-up decode steps (step 1 eager, step 2 captures the graph)
        eng.step()
        if world > 1 and want_parity:
            agree &= ranks_agree_on_last_tokens(); agree_steps += 1
    torch.cuda.synchronize()

    progress("warm-up done (decode graph captured); timing")
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    # ---- e2e: K steps through the public API (host buffers in, host tokens out) -------------------------------
    ctx_first = len(eng.scheduler.sequence(sids[0])["token_ids"])
    if args.profile:
        eng.model_runner.profile(True)
    barrier()
    torch.cuda.cudart().cudaProfilerStart()   # ncu --profile-from-start off: capture only the timed steps
    t0 = time.perf_counter()
    for _ in range(K):
        is_prefill, nseq, _ = eng.step()
        assert not is_prefill and nseq == B
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    torch.cuda.cudart().cudaProfilerStop()
    progress(f"e2e region done ({K} steps through LLEngine.step); device-timed replays")
    barrier()
    launch_profile = None
    if args.profile:
        eng.model_runner.profile(False)
        rep = eng.model_runner.profile_report()
        tot = sum(us for _, us in rep.values()) or 1.0
        launch_profile = {"how": "CUDA events after every launch of eager steps on this rank (serialised, no PDL overlap): "
                                 "compare shares; kernels that wait for peers include the wait",
                          "steps": K, "us_per_step": round(tot / K, 1),
                          "kernels": {n: {"launches_per_step": round(c / K, 2), "us_per_launch": round(us / c, 2),
                                          "us_per_step": round(us / K, 1), "share": round(us / tot, 4)}
                                      for n, (c, us) in sorted(rep.items(), key=lambda kv: -kv[1][1])}}
    ctx_last = len(eng.scheduler.sequence(sids[0])["token_ids"]) - 1
    launches = eng.model_runner.last_launches()
    # ---- value: K replays of the (last) step with inputs resident in HBM, CUDA events on the runner's stream -----
    barrier()
    dev_ms = max_over_ranks(eng.model_runner.replay_decode(K))
    barrier()
    # ---- roofline of the dominant kernel, on the live batch ------------------------------------------------------
    att_ms, att_bytes = eng.model_runner.time_decode_attention(iters=4 * cfg.num_hidden_layers)
    clocks = sampler.stop() if rank == 0 else None
    layers, cap = cfg.num_hidden_layers, cfg.max_model_len // 256
    barrier()
    progress("measurements done; tearing the engine down")
    del eng
    if comm is not None:
        _ffi.check(lib.nvllm_comm_destroy(comm))
    progress("engine and communicator destroyed")
    if rank != 0:
        return None
    peak, peak_src = peaks()
    achieved = att_bytes / (att_ms * 1e-3) / 1e9
    traffic, traffic_capture = None, None
    prof = os.path.join(ROOT, "profiles", "decode_attention_ncu.json")  # dram bytes/launch from the committed ncu capture
    if os.path.exists(prof):
        try:
            pj = json.load(open(prof))
            for cap_ in (pj if isinstance(pj, list) else [pj]):
                if cap_.get("model") == model and cap_.get("batch") == B and cap_.get("tp", 1) == world:
                    traffic = cap_.get("dram_bytes_per_launch")
                    # the capture ran at its own context length: its algorithmic bytes are the ones to compare traffic with
                    traffic_capture = {k: cap_.get(k) for k in ("context", "alg_bytes_per_launch", "ratio", "source")}
        except Exception:
            pass
    line = {
        "metric": "decode_tokens_per_s", "value": round(B * K / (dev_ms * 1e-3), 1), "unit": "tokens/s", "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": round(dev_ms / K, 4), "higher_is_better": True,
        "scaling": "strong" if world > 1 else "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload_name(model, wl), "timed_context": f"{ctx_first}..{ctx_last} (workload mean "
                   f"{wl['prompt'] + wl['decode'] // 2})", "seqs": B, "prompt_len_used": prompt, "parallelism": f"tp{world}",
                   "cuda_graphs": not eager, "l2_note": "inputs larger than L2: one step reads "
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
    if parity is not None:
        line["parity"] = parity
    if world > 1 and want_parity:
        line["ranks_agree_on_tokens"] = {"ok": bool(agree), "steps_checked": agree_steps,
                                         "what": "last sampled token of all sequences, all-gathered after the prefill and every warm-up step"}
    if launch_profile is not None:
        line["launch_profile"] = launch_profile
    if world == 1 and want_cpu:
        ctx_mid = (ctx_first + ctx_last) // 2
        progress("cpu_baseline: the numpy oracle on a bounded sample")
        tps, cores, desc, _ = cpu_decode_sample(model, ctx_mid, 16, 2, 1)
        line["cpu_baseline"] = {"value": round(tps, 2), "unit": "tokens/s", "cores": cores, "kind": "port", "sample": desc}
    return line


# ------------------------------------------------------------------------------------------------ configs[3]: prefix caching
def run_prefix(args, rank, world, local):
    """BASELINE.json configs[3] on ONE GPU: Qwen3-8B, 256 sequences = one shared 4096-token prefix (seed 4321) + 256 unique
    tokens.  Prefill: sequence 0 computes all 4352 tokens, every other sequence hits the 16 shared blocks (ref_count 256)
    and computes only its 256-token suffix against the paged prefix (attention.rs:211-222, block_manager.rs:157-219);
    then decode at context 4353+.  The block accounting of SURVEY.md Appendix B is asserted (272 blocks after prefill,
    + 256 at the first decode step)."""
    import numpy as np
    import torch

    from nano_vllm_rs_b200 import engine as E
    assert world == 1, "--workload prefix is the single-GPU configs[3] run"
    model, B, shared, unique = PREFIX["model"], args.seqs or PREFIX["seqs"], PREFIX["shared"], PREFIX["unique"]
    K, W = args.steps, max(args.warmup, 3)
    L = shared + unique
    per_seq = (L + W + K + 8 + 255) // 256
    nblk = shared // 256 + B * (per_seq - shared // 256) + 8
    cfg = E.Config(model, max_num_seqs=max(B, 8), num_kvcache_blocks=nblk, max_model_len=8192,
                   max_num_batched_tokens=32768, enforce_eager=int(args.eager))
    eng = E.LLMEngine(cfg, seed=1234, std=0.02)
    rng = np.random.default_rng(4321)
    prefix = rng.integers(0, cfg.vocab_size, shared)
    sp = E.SamplingParams(temperature=0.0, max_tokens=W + 2 * K + 64, ignore_eos=True)
    sids = [eng.add_request(np.concatenate([prefix, rng.integers(0, cfg.vocab_size, unique)]), sp) for _ in range(B)]
    H, D, layers = cfg.num_attention_heads, cfg.head_dim, cfg.num_hidden_layers
    # ---- prefill (timed: wall clock between synchronisations, per-launch events on) ------------------------------
    eng.model_runner.profile(True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n_prefill, computed = 0, 0
    while eng.scheduler.get_stats()["waiting"] > 0:
        is_prefill, _, ntok = eng.step()
        assert is_prefill
        n_prefill += 1; computed += ntok
    torch.cuda.synchronize()
    prefill_s = time.perf_counter() - t0
    eng.model_runner.profile(False)
    rep = eng.model_runner.profile_report()
    st = eng.scheduler.get_stats()
    blocks_after_prefill = st["used_blocks"]
    tot_us = sum(us for _, us in rep.values()) or 1.0
    att_us = sum(us for n, (c, us) in rep.items() if n.startswith("prefill_attention"))
    # 4 * H * D * sum_b sum_i (Lk - Lq + i + 1) per layer: sequence 0 causal over 4352, the others 256 rows over 4096 + i + 1
    pairs = L * (L + 1) // 2 + (B - 1) * (unique * shared + unique * (unique + 1) // 2)
    att_flop = 4.0 * H * D * pairs * layers
    tpk, tpk_src = tensor_peak()
    att_tf = att_flop / (att_us * 1e-6) / 1e12 if att_us else 0.0
    prefill = {"steps": n_prefill, "seconds": round(prefill_s, 3), "computed_tokens": computed, "prompt_tokens": B * L,
               "computed_tokens_per_s": round(computed / prefill_s, 1), "prompt_tokens_per_s": round(B * L / prefill_s, 1),
               "blocks_after_prefill": blocks_after_prefill, "cached_blocks": st["cached_blocks"],
               "attention": {"kernel": "prefill_attn_kernel (paged-prefix mode: 4-D TMA over block tables)",
                             "share_of_prefill": round(att_us / tot_us, 4), "ms_total": round(att_us / 1e3, 2),
                             "achieved": round(att_tf, 1), "peak": tpk, "unit": "TFLOP/s", "frac": round(att_tf / tpk, 4),
                             "peak_source": f"{tpk_src} (MEASURED_PEAKS.json bf16_tflops, burst)",
                             "flops": "4*H*D*sum_b sum_i (Lk - Lq + i + 1) per layer, causal pairs counted exactly"},
               "kernel_shares": {n: round(us / tot_us, 4) for n, (c, us) in sorted(rep.items(), key=lambda kv: -kv[1][1])[:8]},
               "how": "wall clock between device synchronisations around all prefill steps through LLMEngine.step, a CUDA event "
                      "after every launch (serialises the programmatic overlap; shares are per-kernel device times)"}
    # ---- decode -------------------------------------------------------------------------------------------------
    eng.step()                                             # first decode step: every sequence takes one new block
    blocks_after_first_decode = eng.scheduler.get_stats()["used_blocks"]
    for _ in range(W - 1):
        eng.step()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    sampler.start()
    ctx_first = len(eng.scheduler.sequence(sids[0])["token_ids"])
    t0 = time.perf_counter()
    for _ in range(K):
        is_prefill, nseq, _ = eng.step()
        assert not is_prefill and nseq == B
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    ctx_last = len(eng.scheduler.sequence(sids[0])["token_ids"]) - 1
    launches = eng.model_runner.last_launches()
    dev_ms = eng.model_runner.replay_decode(K)
    att_ms, att_bytes = eng.model_runner.time_decode_attention(iters=4 * layers)
    clocks = sampler.stop()
    peak, peak_src = peaks()
    achieved = att_bytes / (att_ms * 1e-3) / 1e9
    cap = cfg.max_model_len // 256
    expect_prefill_blocks = shared // 256 + B * ((L + 255) // 256 - shared // 256)
    ok_blocks = (blocks_after_prefill == expect_prefill_blocks and blocks_after_first_decode == expect_prefill_blocks + B)
    line = {
        "metric": "decode_tokens_per_s", "value": round(B * K / (dev_ms * 1e-3), 1), "unit": "tokens/s", "n_gpus": 1,
        "steps": K, "warmup": W, "ms_per_step": round(dev_ms / K, 4), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": f"{model} bf16 prefix caching: {B} seqs sharing a {shared}-token prefix + {unique} unique tokens "
                               f"(BASELINE.json configs[3]), paged KV block 256; decode timed at context {ctx_first}..{ctx_last}",
                   "seqs": B, "parallelism": "tp1", "cuda_graphs": not args.eager,
                   "l2_note": f"inputs larger than L2: one decode step reads {round(att_bytes * layers / 1e9, 1)} GB of KV"},
        "prefill": prefill,
        "blocks": {"after_prefill": blocks_after_prefill, "after_first_decode_step": blocks_after_first_decode,
                   "expected": [expect_prefill_blocks, expect_prefill_blocks + B], "ok": ok_blocks,
                   "what": "SURVEY.md Appendix B cfg4: 16 shared blocks at ref_count 256 + one private block per sequence, "
                           "then one more per sequence at len 4353"},
        "e2e": {"value": round(B * K / e2e_s, 1), "unit": "tokens/s", "ms_per_step": round(e2e_s / K * 1e3, 4),
                "h2d_bytes_per_step": B * (24 + 4 * cap), "d2h_bytes_per_step": 8 * B, "api": "LLMEngine.step"},
        "gpu_launches": int(launches) * K, "launches_per_step": int(launches),
        "roofline": {"kernel": "paged_decode_kernel", "bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                     "frac": round(achieved / peak, 4), "frac_of_8TBps_nominal": round(achieved / 8000.0, 4),
                     "peak_source": peak_src, "traffic": None, "alg_bytes_per_launch": int(att_bytes),
                     "ms_per_launch": round(att_ms, 5), "share_of_step": round(att_ms * layers / (dev_ms / K), 4),
                     "note": "algorithmic bytes count every sequence's full context; the 16 shared blocks (4096 of 4353+ tokens) "
                             "are the same physical memory for all sequences, so most of these reads hit L2 — a fraction far "
                             "above 1.0 here is block dedup at work, not a measurement error"},
        "clocks": clocks,
    }
    del eng
    return line, ok_blocks


def arm_watchdog():
    """A bench run that stops making progress (a rank waiting for a peer that died, a kernel that never returns) must not
    sit there until someone else's limit kills it without a trace.  A daemon thread watches the time since the last
    progress() line: once one PHASE has lasted BENCH_WATCHDOG_S seconds (default 240; a whole normal run takes 40-60 s) it
    dumps every thread's Python stack to stderr, prints {"error": "watchdog", "phase": ...} as this process's JSON line on
    stdout — so the record names the phase even when stderr's tail is lost — and exits with status 3 (torchrun then stops
    the other ranks).  ctypes calls and torch's CUDA waits release the GIL, so the thread runs while the main thread is
    stuck in either; faulthandler's C-level timer 30 s later is the backstop for a stall that holds the GIL."""
    import faulthandler
    import threading
    limit = float(os.environ.get("BENCH_WATCHDOG_S", "240"))
    if limit <= 0:
        return

    def watch():
        while True:
            time.sleep(1.0)
            stalled = time.time() - _LAST["t"]
            if stalled > limit:
                faulthandler.dump_traceback(file=sys.stderr, all_threads=True)
                print(json.dumps({"error": "watchdog", "phase": _LAST["msg"], "stalled_s": round(stalled, 1),
                                  "rank": int(os.environ.get("RANK", "0")), "n_gpus": int(os.environ.get("WORLD_SIZE", "1")),
                                  "since_start_s": round(time.time() - _T0, 1)}), flush=True)
                sys.stderr.flush()
                os._exit(3)
            faulthandler.cancel_dump_traceback_later()
            faulthandler.dump_traceback_later(max(1.0, limit + 30.0 - stalled), exit=True)

    threading.Thread(target=watch, name="bench-watchdog", daemon=True).start()


def main():
    args = parse()
    arm_watchdog()
    if args.impl == "reference":
        return reference_arm(args)
    import torch

    from nano_vllm_rs_b200 import _ffi
    from nano_vllm_rs_b200 import engine as E

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == args.gpus or world == 1 and args.gpus == 1, f"--gpus {args.gpus} but WORLD_SIZE={world}"
    torch.cuda.set_device(local)
    lib = E.lib()
    assert lib.nvllm_device_ok() == 1, "no sm_100 device: the product path has no fallback"
    if args.workload == "prefix":
        line, ok = run_prefix(args, rank, world, local)
        print(json.dumps(line), flush=True)
        sys.exit(0 if ok else 1)
    tp_parity = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        if not args.no_parity:
            progress("process group up; tp_parity check (seeded tiny model vs the TP=1 oracle)")
            tp_parity = tp_parity_check(E, _ffi, lib, rank, world, torch)
            progress(f"tp_parity done: ok={tp_parity['ok']}")
    model, wl, prompt, K, W = workload(args)
    line = run_decode(args, model, wl, prompt, K, W, rank, world, local, want_parity=not args.no_parity,
                      want_cpu=not args.no_cpu_baseline)
    ok = True
    if rank == 0:
        if world == 1:
            pk, _ = tensor_peak()
            progress("tensor_kernels: prefill GEMM / attention micro-measurements")
            line["tensor_kernels"] = tensor_kernels(pk)
            ok = line.get("parity", {}).get("ok", True)
        else:
            line["tp_parity"] = tp_parity
            agree = line.get("ranks_agree_on_tokens", {}).get("ok", True)
            if tp_parity is not None:
                tp_parity["ranks_agree"] = bool(tp_parity["ranks_agree"] and agree)
                ok = tp_parity["ok"] and agree
            # the same job on ONE GPU (python bench.py --gpus 1 --model qwen3-8b; measured line kept under profiles/)
            ref1 = os.path.join(ROOT, "profiles", "r02_bench_8b_n1.json")
            if model == "qwen3-8b" and os.path.exists(ref1):
                try:
                    v1 = json.load(open(ref1))["value"]
                    line["config"]["tp1_same_workload_tokens_per_s"] = v1
                    line["config"]["strong_scaling_efficiency"] = round(line["value"] / (world * v1), 4)
                    line["config"]["tp1_source"] = "profiles/r02_bench_8b_n1.json (this repo, same bounded window, N = 1)"
                except Exception:
                    pass
        print(json.dumps(line), flush=True)
    if world > 1:
        flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device="cuda")
        torch.distributed.broadcast(flag, 0)
        ok = bool(flag.item())
        torch.distributed.destroy_process_group()
    if not ok:
        sys.exit(1)


if __name__ == "__main__":
    main()
