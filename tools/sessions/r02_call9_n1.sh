#!/bin/bash
# Round-2 GPU session 9 (1 GPU): GEMM tests (forced variants now really forced), 192-wide tile sweep on the TP shapes,
# the N=1 bench lines (0.6B headline, 8B on one GPU).
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_prefill.py tests/test_gpu_kernels.py -m gpu -q > gpurun_out/c9_pytest.log 2>&1; echo "pytest rc=$?"
tail -c 2500 gpurun_out/c9_pytest.log
timeout 900 python tools/bench_gemm_graph.py --only "8B m512,8B/tp2,8B/tp4,8B/tp8,8B m256" --knobs 0,1,10,11,12,13,14 > gpurun_out/c9_gemm_sweep.jsonl 2> gpurun_out/c9_gemm_sweep.err; echo "sweep rc=$?"
timeout 600 python bench.py --steps 64 --warmup 4 > gpurun_out/c9_bench_n1.json 2> gpurun_out/c9_bench_n1.err; echo "bench n1 rc=$?"
timeout 600 python bench.py --gpus 1 --model qwen3-8b --no-cpu-baseline > gpurun_out/c9_bench_8b_n1.json 2> gpurun_out/c9_bench_8b_n1.err; echo "bench 8b rc=$?"
python - <<'PY'
import json
best = {}
for l in open('gpurun_out/c9_gemm_sweep.jsonl'):
    r = json.loads(l)
    if 'error' in r: continue
    best.setdefault(r['shape'], []).append((r['us_per_launch'], json.dumps(r['knobs'])))
for k, v in best.items():
    d = [x for x in v if x[1] == '{}'][0][0]
    print(k, 'default', d, '| best', sorted(v)[0], '|', ' '.join(f"{a}:{b}" for a, b in sorted(v)[:4]))
PY
for f in gpurun_out/c9_bench_n1.json gpurun_out/c9_bench_8b_n1.json; do grep '^{' $f | head -c 250; echo; done
