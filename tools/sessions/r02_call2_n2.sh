#!/bin/bash
# Round-2 GPU session 2 (2 GPUs): TP tests vs the oracle, the N=2 bench line with tp_parity, the fused all-reduce micro-bench,
# a per-launch profile of the TP decode step.
mkdir -p gpurun_out
# the fused decode kernel (q/k-norm + RoPE + store inside the attention) is new in this snapshot: validate it first, fall back if red
if timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "fused or paged_decode" > gpurun_out/c2_pytest_fused.log 2>&1; then
  echo "fused decode kernel: green"
else
  echo "fused decode kernel: RED -> NVLLM_FUSE_DECODE=0 for the rest of this session"; export NVLLM_FUSE_DECODE=0
fi
tail -c 1500 gpurun_out/c2_pytest_fused.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gpu_tp.py -m gpu -x -q > gpurun_out/c2_pytest_tp.log 2>&1; echo "pytest tp rc=$?"
timeout 600 $TR --master-port 29711 bench.py --gpus 2 --steps 64 --warmup 4 > gpurun_out/c2_bench_n2.json 2> gpurun_out/c2_bench_n2.err; echo "bench n2 rc=$?"
timeout 600 $TR --master-port 29712 tools/bench_allreduce.py > gpurun_out/c2_allreduce_n2.jsonl 2> gpurun_out/c2_allreduce_n2.err; echo "allreduce rc=$?"
timeout 600 $TR --master-port 29713 bench.py --gpus 2 --steps 8 --warmup 4 --profile --no-parity > gpurun_out/c2_profile_n2.json 2> gpurun_out/c2_profile_n2.err; echo "profile rc=$?"
NVLLM_FUSED_TP=0 timeout 600 $TR --master-port 29714 bench.py --gpus 2 --steps 32 --warmup 4 --no-parity > gpurun_out/c2_bench_n2_nccl.json 2> gpurun_out/c2_bench_n2_nccl.err; echo "bench n2 nccl rc=$?"
tail -c 400 gpurun_out/c2_pytest_tp.log
cat gpurun_out/c2_allreduce_n2.jsonl
tail -c 300 gpurun_out/c2_bench_n2.err
