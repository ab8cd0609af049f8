"""CPU tests of bench.py: the reference arm (`--impl reference`, the CPU oracle port timed on the host cores) prints one JSON
line with the keys the driver's contract names, on the same metric / unit / workload wording as the GPU arm; the workload
window arithmetic follows BASELINE.json's configs.  (The GPU arm itself is exercised on the GPU box by the driver.)"""
import json
import os
import subprocess
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_the_contract_line():
    p = subprocess.run([sys.executable, "bench.py", "--impl", "reference", "--seqs", "4", "--steps", "8", "--warmup", "3"],
                       cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "decode_tokens_per_s" and d["unit"] == "tokens/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["value"] > 0 and d["vs_baseline"] is None
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "4 sequences" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["workload"].startswith("qwen3-0.6b bf16 decode: 4 concurrent seqs")


def test_reference_arm_other_ranks_do_no_work():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    p = subprocess.run([sys.executable, "bench.py", "--impl", "reference", "--gpus", "2"], cwd=ROOT, env=env,
                       capture_output=True, text=True, timeout=120)
    assert p.returncode == 0 and p.stdout.strip() == ""


def test_workload_windows_follow_the_baseline_configs():
    sys.path.insert(0, ROOT)
    import bench
    base = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    args = types.SimpleNamespace(model=None, gpus=1, seqs=None, steps=64, warmup=4)
    model, wl, prompt, K, W = bench.workload(args)
    assert model == "qwen3-0.6b" and wl["seqs"] == 256 and (K, W) == (64, 4)
    assert prompt + W + K // 2 == wl["prompt"] + wl["decode"] // 2          # the timed window is centred on the mean context
    name = bench.workload_name(model, wl)
    assert "256 concurrent seqs" in name and "512-token prompts + 512 decode steps" in name
    args.gpus = 8
    model, wl, prompt, K, W = bench.workload(args)
    assert model == "qwen3-8b" and wl["seqs"] == 512
    args.model = "qwen3-32b"
    model, wl, prompt, K, W = bench.workload(args)
    assert wl["seqs"] == 64 and wl["blocks_per_seq"] == 129
    assert "decode" in base["metric"] or "token" in base["metric"]              # the metric bench.py prints is BASELINE.json's


def test_watchdog_names_the_stalled_phase_on_stdout():
    """A phase that outlasts BENCH_WATCHDOG_S ends the process with status 3 and ONE JSON line on stdout that names the last
    progress message — the round-2 8-GPU run hung for 870 s and left nothing in the driver's record; this is what it would
    have left.  progress() re-arms the timer: a run made of many short phases is never cut."""
    code = ("import sys, time; sys.path.insert(0, %r); import bench; bench.arm_watchdog()\n"
            "for i in range(4):\n    time.sleep(0.7); bench.progress('short phase %%d' %% i)\n"
            "bench.progress('the phase that hangs'); time.sleep(30)\n") % ROOT
    env = dict(os.environ, BENCH_WATCHDOG_S="1.5", RANK="3", WORLD_SIZE="8")
    p = subprocess.run([sys.executable, "-c", code], cwd=ROOT, env=env, capture_output=True, text=True, timeout=60)
    assert p.returncode == 3, (p.returncode, p.stderr[-1500:])
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["error"] == "watchdog" and d["phase"] == "the phase that hangs" and d["rank"] == 3 and d["n_gpus"] == 8
    assert d["stalled_s"] >= 1.5 and "short phase 3" in p.stderr and "Thread" in p.stderr  # stacks went to stderr
