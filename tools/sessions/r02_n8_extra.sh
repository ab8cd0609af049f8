#!/bin/bash
# 8 GPUs: phase timeline of the low-latency all-reduce at 8 ranks (debug build), then the cfg5 line (Qwen3-32B TP=8).
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
NVLLM_LIB_PATH=nano-vllm-rs_b200/lib/libnvllm_timeline.so timeout 300 $TR --master-port 29728 tools/ar_timeline.py > gpurun_out/c17_ar_timeline_w8.jsonl 2> gpurun_out/c17_ar_timeline_w8.err; echo "timeline rc=$?"
grep '^{' gpurun_out/c17_ar_timeline_w8.jsonl | cut -c1-700
timeout 900 $TR --master-port 29729 bench.py --gpus 8 --model qwen3-32b --steps 32 --warmup 4 > gpurun_out/c17_bench_32b_n8.json 2> gpurun_out/c17_bench_32b_n8.err; echo "bench 32b rc=$?"
tail -c 400 gpurun_out/c17_bench_32b_n8.err
grep '^{' gpurun_out/c17_bench_32b_n8.json | head -c 1500; echo
