"""-m gpu, needs >= 2 GPUs (skipped otherwise): tensor-parallel engine (one process per GPU, fused all-reduce + add + RMSNorm
over NVLink peer memory after o_proj / down_proj, vocab-sharded argmax) against the CPU ORACLE (TP=1 numpy model, same seeded
weights) and against the single-GPU engine on the same requests.  On a 1-GPU box the same path is checked by bench.py's
`tp_parity` at N > 1 (the driver's scaling run)."""
import os
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _tp_worker(rank, world, port, arch, prompts, max_tokens, q, sps=None, std=0.05, sampling_seed=0):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import time
    t_start = time.time()
    say = lambda msg: print(f"[tp rank {rank}/{world} +{time.time() - t_start:5.1f}s] {msg}", flush=True)
    try:
        import torch.distributed as dist
        from nano_vllm_rs_b200 import _ffi
        from nano_vllm_rs_b200 import engine as E
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
        lib = E.lib()
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            buf = (E.C.c_uint8 * 128)()
            _ffi.check(lib.nvllm_comm_unique_id(buf))
            uid.copy_(torch.tensor(list(buf), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        h = E.vp()
        say("process group up")
        _ffi.check(lib.nvllm_comm_init(E.C.byref(h), bytes(uid.cpu().tolist()), rank, world))
        say("nvllm communicator up")
        cfg = E.Config(arch, kvcache_block_size=16, num_kvcache_blocks=64, tensor_parallel_size=world, tp_rank=rank,
                       sampling_seed=sampling_seed)
        sched, runner = E.Scheduler(cfg), E.ModelRunner(cfg, comm=h)
        runner.init_weights(1234, std)
        say("runner ready")
        for i, p in enumerate(prompts):
            sp = E.SamplingParams(**sps[i]) if sps else E.SamplingParams(temperature=0.0, max_tokens=max_tokens, ignore_eos=True)
            sched.add_sequence(p, sp)
        trace = []
        while not sched.is_finished():
            b = sched.schedule()
            toks, logits = runner.execute_model(b, return_logits=True)
            trace.append((toks.copy(), logits.copy()))
            sched.postprocess(toks)
        say(f"{len(trace)} steps done")
        q.put((rank, "ok", trace))
        q.close(); q.join_thread()        # results are with the parent before any teardown starts
        dist.barrier()
        del runner, sched
        _ffi.check(lib.nvllm_comm_destroy(h))
        say("communicator destroyed")
        os._exit(0)                       # skip interpreter / NCCL static teardown: nothing left to check
    except Exception:  # noqa: BLE001
        import traceback
        q.put((rank, "fail: " + traceback.format_exc(), None))
        q.close(); q.join_thread()
        os._exit(1)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_tp_engine_matches_single_gpu(cuda, world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    from nano_vllm_rs_b200 import engine as E
    import torch.multiprocessing as mp
    rng = np.random.default_rng(5)
    prompts = [rng.integers(0, 1024, n).tolist() for n in (7, 40, 16, 33)]
    max_tokens = 12
    arch = "tiny" if world == 2 else "tiny-tp8"
    # TP = 1 reference run (same seeded logical weights)
    cfg = E.Config(arch, kvcache_block_size=16, num_kvcache_blocks=64)
    sched, runner = E.Scheduler(cfg), E.ModelRunner(cfg)
    runner.init_weights(1234, 0.05)
    for p in prompts:
        sched.add_sequence(p, E.SamplingParams(temperature=0.0, max_tokens=max_tokens, ignore_eos=True))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 2000
    procs = [ctx.Process(target=_tp_worker, args=(r, world, port, arch, prompts, max_tokens, q), daemon=True)
             for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=300) for _ in procs], key=lambda r: r[0])
    for p in procs:           # teardown of a rank must never hold the test (or pytest's exit) hostage
        p.join(timeout=20)
        if p.is_alive():
            print(f"[tp test] rank process {p.pid} still alive 20 s after reporting: killing it", flush=True)
            p.kill()
            p.join(timeout=5)
    for rank, status, _ in res:
        assert status == "ok", f"rank {rank}: {status}"
    traces = [r[2] for r in res]
    from oracle import qwen3 as Q
    from tests.test_gpu_engine import LOGIT_RTOL, LOGIT_SIGMA_PER_SQRT_LAYER, to_inputs
    ocfg = Q.arch_config(arch, max_position_embeddings=cfg.max_position_embeddings)
    oracle = Q.Qwen3Oracle(ocfg, Q.seeded_weights(ocfg, 1234, 0.05), 64, 16)
    step, mismatched = 0, 0
    while not sched.is_finished():
        b = sched.schedule()
        toks1, logits1 = runner.execute_model(b, return_logits=True)
        # ---- vs the oracle: the usual bf16 bound + 0.05 sigma for the per-rank partial sums rounded to bf16 before the sum
        ref = oracle.forward(to_inputs(b))
        tp_full = np.concatenate([traces[r][step][1] for r in range(world)], axis=1)
        sig_r = ref.std(axis=-1, keepdims=True)
        tol = (LOGIT_SIGMA_PER_SQRT_LAYER * np.sqrt(ocfg["num_hidden_layers"]) + 0.05) * sig_r + LOGIT_RTOL * np.abs(ref)
        assert (np.abs(tp_full - ref) <= tol).all(), (step, float((np.abs(tp_full - ref) / sig_r).max()))
        ref_tok = np.argmax(ref, -1)
        for i, t in enumerate(traces[0][step][0]):
            assert t == ref_tok[i] or ref[i, ref_tok[i]] - ref[i, t] <= 2 * tol[i].max(), (step, i, t, ref_tok[i])
        tp_toks = traces[0][step][0]
        for r in range(1, world):
            assert np.array_equal(traces[r][step][0], tp_toks), "ranks disagree on the sampled tokens"
        tp_logits = np.concatenate([traces[r][step][1] for r in range(world)], axis=1)
        sigma = logits1.std()
        # the all-reduce sums bf16-rounded per-rank partials: a few bf16 ulps per layer vs TP=1
        assert np.abs(tp_logits - logits1).max() <= 0.1 * sigma, (step, np.abs(tp_logits - logits1).max() / sigma)
        mismatched += int((tp_toks != toks1).sum())
        sched.postprocess(tp_toks)   # teacher-force with the TP tokens so the two runs see the same inputs
        step += 1
    assert step == len(traces[0])
    assert mismatched <= 0.1 * step * len(prompts)


def _run_ranks(world, arch, prompts, max_tokens, **kw):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + (os.getpid() * 7 + world) % 2000
    procs = [ctx.Process(target=_tp_worker, args=(r, world, port, arch, prompts, max_tokens, q), kwargs=kw, daemon=True)
             for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=300) for _ in procs], key=lambda r: r[0])
    for p in procs:
        p.join(timeout=20)
        if p.is_alive():
            p.kill(); p.join(timeout=5)
    for rank, status, _ in res:
        assert status == "ok", f"rank {rank}: {status}"
    return [r[2] for r in res]


@pytest.mark.parametrize("world", [2, 8])
def test_tp_sampling_matches_oracle_sampler_on_gathered_logits(cuda, world):
    """Sampler::batch_sample (sampler.rs:221-254) over VOCAB-SHARDED logits (the gather TODO of embed_head.rs:321-336):
    temperature / top-k / top-p / mixed greedy rows under TP.  Every rank must draw the same token, and that token must be
    what the oracle sampler draws from the concatenated f32 logits with the row seed
    mix64(sampling_seed ^ mix64(seq_id * GOLDEN + len))."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    from nano_vllm_rs_b200 import engine as E
    from oracle import layers as O
    arch = "tiny" if world == 2 else "tiny-tp8"
    rng = np.random.default_rng(11)
    lens = (9, 20, 33, 5, 16)
    prompts = [rng.integers(0, 1024, n).tolist() for n in lens]
    sps = [dict(temperature=0.8, top_k=50, top_p=0.9, max_tokens=10, ignore_eos=True),
           dict(temperature=0.0, max_tokens=10, ignore_eos=True),
           dict(temperature=1.0, max_tokens=7, ignore_eos=True),
           dict(temperature=1.3, top_p=0.5, max_tokens=10, ignore_eos=True),
           dict(temperature=0.6, top_k=5, max_tokens=10, ignore_eos=True)]
    traces = _run_ranks(world, arch, prompts, 10, sps=sps, std=0.3, sampling_seed=77)
    # replay the schedule on the host (pure scheduler, teacher-forced with the TP tokens) to recover seq ids / lengths / params
    sched = E.Scheduler(E.Config(arch, kvcache_block_size=16, num_kvcache_blocks=64, sampling_seed=77))
    for p, sp in zip(prompts, sps):
        sched.add_sequence(p, E.SamplingParams(**sp))
    step = rows = strict = not_argmax = 0
    while not sched.is_finished():
        b = sched.schedule()
        toks = traces[0][step][0]
        for r in range(1, world):
            assert np.array_equal(traces[r][step][0], toks), f"step {step}: ranks drew different tokens"
        logits = np.concatenate([traces[r][step][1] for r in range(world)], axis=1)
        seq_lens = b.seq_lens()
        for i in range(len(toks)):
            T, k, p = float(b.temperatures[i]), int(b.top_k[i]), float(b.top_p[i])
            seed = O.sampler_row_seed(77, int(b.seq_ids[i]), int(seq_lens[i]))
            want, filt, scores = O.sample_single(logits[i], T, p if p > 0 else None, k if k > 0 else None, seed)
            rows += 1
            if toks[i] == want:
                strict += 1
            else:   # logf-ulp near-tie of the Gumbel scores, or the element exactly at the top-p cut
                assert T != 0.0, "greedy rows must match exactly"
                sc = logits[i] / np.float32(T)
                assert sc[toks[i]] >= sc[np.isfinite(filt)].min() - 1e-3, (step, i, toks[i], want)
                g = O.sample_gumbel(O.sampler_uniform(seed, logits.shape[1]))
                assert sc[toks[i]] + g[toks[i]] >= scores.max() - 1e-4, (step, i, toks[i], want)
            if T != 0.0 and toks[i] != int(np.argmax(logits[i])):
                not_argmax += 1
        sched.postprocess(toks)
        step += 1
    assert step == len(traces[0])
    print(dict(rows=rows, strict=strict, sampled_not_argmax=not_argmax))
    assert strict >= 0.97 * rows and not_argmax > 3


# ---------------------------------------------------------------------------------------------------------------------
# The fused all-reduce + add + RMSNorm kernels by themselves (C1), both protocols.
AR_SHAPES = [(512, 4096), (3, 1024), (64, 5120), (512, 4096), (1, 1024), (37, 2048), (256, 1024), (7, 4096), (512, 4096),
             (300, 1024), (16, 8192), (5, 1024)]


def _ar_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    try:
        import ctypes as C
        import torch.distributed as dist
        from nano_vllm_rs_b200 import _ffi
        from nano_vllm_rs_b200 import engine as E
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
        lib = E.lib()
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            buf = (C.c_uint8 * 128)()
            _ffi.check(lib.nvllm_comm_unique_id(buf))
            uid.copy_(torch.tensor(list(buf), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        comm = E.vp()
        _ffi.check(lib.nvllm_comm_init(C.byref(comm), bytes(uid.cpu().tolist()), rank, world))
        one = (max(t * h for t, h in AR_SHAPES) * 2 + 255) // 256 * 256
        base = E.vp()
        _ffi.check(lib.nvllm_comm_symm_alloc(comm, 3 * one, C.byref(base)))
        p_h, p_out, p_x = base.value, base.value + one, base.value + 2 * one
        rt = C.CDLL("libcudart.so.12")
        rt.cudaMemcpy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
        s = torch.cuda.Stream()
        st = C.c_void_p(s.cuda_stream)
        report = []
        for it, (T, hid) in enumerate(AR_SHAPES):
            n = T * hid
            g = torch.Generator(device="cuda"); g.manual_seed(1000 * it + rank)
            x = (torch.randn(n, generator=g, device="cuda") * 0.5).to(torch.bfloat16)
            g2 = torch.Generator(device="cuda"); g2.manual_seed(7 + it)
            res = torch.randn(n, generator=g2, device="cuda").to(torch.bfloat16)
            w = (1 + 0.1 * torch.randn(hid, generator=g2, device="cuda")).to(torch.bfloat16)
            # the bit pattern the low-latency protocol reserves: pairs of -0.0 in the partial sums, rows that sum to zero,
            # and a residual that cancels the sum (normed output = +-0)
            x.view(T, hid)[0, :16] = -0.0
            if T > 1:
                x.view(T, hid)[T - 1] = 0.0 if rank else -0.0
                res.view(T, hid)[T - 1, : hid // 2] = -0.0
            outs = {}
            for proto in ("1", "0"):      # low-latency push, then the flag protocol
                os.environ["NVLLM_AR_LL"] = proto
                lib.nvllm_reload_env()
                for _ in range(2 if proto == "1" else 1):   # twice: the pair rotation must not depend on the call count
                    assert rt.cudaMemcpy(p_x, x.data_ptr(), n * 2, 3) == 0
                    assert rt.cudaMemcpy(p_h, res.data_ptr(), n * 2, 3) == 0
                    assert rt.cudaMemcpy(p_out, torch.full((n,), 7.0, dtype=torch.bfloat16, device="cuda").data_ptr(), n * 2, 3) == 0
                    torch.cuda.synchronize(); dist.barrier()
                    with torch.cuda.stream(s):
                        _ffi.check(lib.nvllm_allreduce_add_rmsnorm(comm, p_out, p_x, p_h, w.data_ptr(), T, hid, 1e-6, 0, st))
                    torch.cuda.synchronize(); dist.barrier()
                assert lib.nvllm_comm_error(comm) == 0
                out = torch.empty(n, dtype=torch.bfloat16, device="cuda")
                hres = torch.empty(n, dtype=torch.bfloat16, device="cuda")
                assert rt.cudaMemcpy(out.data_ptr(), p_out, n * 2, 3) == 0
                assert rt.cudaMemcpy(hres.data_ptr(), p_h, n * 2, 3) == 0
                outs[proto] = (out, hres)
            xs = [torch.empty_like(x) for _ in range(world)]
            dist.all_gather(xs, x)
            tot = torch.zeros(n, dtype=torch.float32, device="cuda")
            for t in xs:
                tot += t.float()
            hnew = (tot.to(torch.bfloat16).float() + res.float()).to(torch.bfloat16).float().view(T, hid)
            ref = (hnew * torch.rsqrt((hnew * hnew).mean(-1, keepdim=True) + 1e-6) * w.float()).to(torch.bfloat16)
            R = (T + world - 1) // world
            mine = slice(min(T, rank * R), min(T, rank * R + R))
            o1, h1 = outs["1"]; o0, h0 = outs["0"]
            report.append(dict(T=T, hid=hid,
                               ll_err=float((o1.view(T, hid).float() - ref.float()).abs().max()),
                               flag_err=float((o0.view(T, hid).float() - ref.float()).abs().max()),
                               ll_eq_flag=bool(torch.equal(o1.float(), o0.float())),                 # -0 == +0
                               res_eq=bool(torch.equal(h1.view(T, hid)[mine].float(), h0.view(T, hid)[mine].float())),
                               res_ref=float((h1.view(T, hid)[mine].float() - hnew[mine]).abs().max()) if mine.stop > mine.start else 0.0))
        # every rank must hold the same normed tensor bit for bit (receivers get the owner's bytes)
        last = outs["1"][0].view(torch.uint8)
        alls = [torch.empty_like(last) for _ in range(world)]
        dist.all_gather(alls, last)
        same = all(torch.equal(a, alls[0]) for a in alls)
        q.put((rank, "ok", dict(report=report, ranks_identical=same)))
        q.close(); q.join_thread()
        dist.barrier()
        _ffi.check(lib.nvllm_comm_destroy(comm))
        os._exit(0)
    except Exception:  # noqa: BLE001
        import traceback
        q.put((rank, "fail: " + traceback.format_exc(), None))
        q.close(); q.join_thread()
        os._exit(1)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_fused_allreduce_low_latency_and_flag_protocols(cuda, world):
    """nvllm_allreduce_add_rmsnorm (the all-reduce linear.rs:236 leaves as a TODO, fused with qwen3.rs:382-389's add + norm):
    the low-latency push protocol and the flag protocol against an f32 torch computation of the same sum and against each
    other (equal as numbers everywhere — the push protocol maps the word 0x80008000, two -0.0, to +0.0), over a sequence of
    calls whose sizes change (ragged rows, fewer rows than ranks, buffer-pair rotation and clears of differing sizes)."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + (os.getpid() * 13 + world) % 2000
    procs = [ctx.Process(target=_ar_worker, args=(r, world, port, q), daemon=True) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=300) for _ in procs], key=lambda r: r[0])
    for p in procs:
        p.join(timeout=20)
        if p.is_alive():
            p.kill(); p.join(timeout=5)
    for rank, status, _ in res:
        assert status == "ok", f"rank {rank}: {status}"
    for rank, _, d in res:
        assert d["ranks_identical"], rank
        for e in d["report"]:
            assert e["ll_err"] <= 0.0625 and e["flag_err"] <= 0.0625, (rank, e)   # 1 bf16 ulp at |v| < 8
            assert e["ll_eq_flag"] and e["res_eq"], (rank, e)
            assert e["res_ref"] <= 0.0625, (rank, e)
