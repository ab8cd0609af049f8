#!/bin/bash
# Round-2 GPU session 1 (1 GPU): full -m gpu suite, the N=1 bench line, the 8B job on one GPU, configs[3], per-launch profile.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/c1_gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c1_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/c1_pytest.log
timeout 600 python bench.py --steps 64 --warmup 4 > gpurun_out/c1_bench_n1.json 2> gpurun_out/c1_bench_n1.err; echo "bench n1 rc=$?"
timeout 600 python bench.py --gpus 1 --model qwen3-8b --no-cpu-baseline > gpurun_out/c1_bench_8b_n1.json 2> gpurun_out/c1_bench_8b_n1.err; echo "bench 8b rc=$?"
timeout 600 python bench.py --workload prefix --steps 32 > gpurun_out/c1_bench_cfg4.json 2> gpurun_out/c1_bench_cfg4.err; echo "bench cfg4 rc=$?"
timeout 300 python bench.py --profile --steps 8 --no-cpu-baseline --no-parity > gpurun_out/c1_bench_n1_profile.json 2> gpurun_out/c1_profile.err; echo "profile rc=$?"
tail -c 600 gpurun_out/c1_pytest.log
head -c 1500 gpurun_out/c1_bench_n1.json
