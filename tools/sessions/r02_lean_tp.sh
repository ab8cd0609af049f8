#!/bin/bash
# Lean TP session (N GPUs): fused all-reduce micro-bench + the bench line (tp_parity inside).  Tag = $2.
N=${1:-8}; TAG=${2:-c6}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29712 tools/bench_allreduce.py > gpurun_out/${TAG}_allreduce_n$N.jsonl 2> gpurun_out/${TAG}_allreduce_n$N.err; echo "allreduce rc=$?"
grep '^{' gpurun_out/${TAG}_allreduce_n$N.jsonl | cut -c1-420; tail -c 300 gpurun_out/${TAG}_allreduce_n$N.err
timeout 600 $TR --master-port 29711 bench.py --gpus $N --steps 64 --warmup 4 > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err; echo "bench rc=$?"
tail -c 300 gpurun_out/${TAG}_bench_n$N.err
grep '^{' gpurun_out/${TAG}_bench_n$N.json | head -c 500; echo
