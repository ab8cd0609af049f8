#!/bin/bash
# Round-2 GPU session 4 (N GPUs, default 2): TP tests vs the oracle (incl. both fused all-reduce protocols), the fused
# all-reduce micro-bench, the bench line with tp_parity, the flag-protocol and NCCL A/B lines, a per-launch profile.
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_gpu_tp.py -m gpu -q > gpurun_out/c4_pytest_tp_n$N.log 2>&1; echo "pytest tp rc=$?"
tail -c 1500 gpurun_out/c4_pytest_tp_n$N.log
timeout 600 $TR --master-port 29712 tools/bench_allreduce.py > gpurun_out/c4_allreduce_n$N.jsonl 2> gpurun_out/c4_allreduce_n$N.err; echo "allreduce rc=$?"
cat gpurun_out/c4_allreduce_n$N.jsonl; tail -c 600 gpurun_out/c4_allreduce_n$N.err
timeout 600 $TR --master-port 29711 bench.py --gpus $N --steps 64 --warmup 4 > gpurun_out/c4_bench_n$N.json 2> gpurun_out/c4_bench_n$N.err; echo "bench rc=$?"
tail -c 400 gpurun_out/c4_bench_n$N.err
NVLLM_AR_LL=0 timeout 600 $TR --master-port 29715 bench.py --gpus $N --steps 32 --warmup 4 --no-parity > gpurun_out/c4_bench_n${N}_flags.json 2> gpurun_out/c4_bench_n${N}_flags.err; echo "bench flags rc=$?"
NVLLM_FUSED_TP=0 timeout 600 $TR --master-port 29714 bench.py --gpus $N --steps 32 --warmup 4 --no-parity > gpurun_out/c4_bench_n${N}_nccl.json 2> gpurun_out/c4_bench_n${N}_nccl.err; echo "bench nccl rc=$?"
timeout 600 $TR --master-port 29713 bench.py --gpus $N --steps 8 --warmup 4 --profile --no-parity > gpurun_out/c4_profile_n$N.json 2> gpurun_out/c4_profile_n$N.err; echo "profile rc=$?"
for f in gpurun_out/c4_bench_n$N.json gpurun_out/c4_bench_n${N}_flags.json gpurun_out/c4_bench_n${N}_nccl.json; do grep '^{' $f | head -c 700; echo; done
NVLLM_FUSE_DECODE=1 timeout 600 $TR --master-port 29716 bench.py --gpus $N --steps 32 --warmup 4 --no-parity > gpurun_out/c4_bench_n${N}_fusedec.json 2> gpurun_out/c4_bench_n${N}_fusedec.err; echo "bench fused-decode rc=$?"
grep '^{' gpurun_out/c4_bench_n${N}_fusedec.json | head -c 400; echo
