#!/bin/bash
# GEMM-epilogue push validation (N GPUs): TP tests vs oracle, bench line with tp_parity, A/B against NVLLM_GEMM_PUSH=0.
N=${1:-2}; TAG=${2:-c8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_gpu_tp.py -m gpu -q -x > gpurun_out/${TAG}_pytest_tp_n$N.log 2>&1; echo "pytest tp rc=$?"
tail -c 1200 gpurun_out/${TAG}_pytest_tp_n$N.log
timeout 600 $TR --master-port 29711 bench.py --gpus $N --steps 64 --warmup 4 > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err; echo "bench rc=$?"
tail -c 300 gpurun_out/${TAG}_bench_n$N.err
NVLLM_GEMM_PUSH=0 timeout 600 $TR --master-port 29715 bench.py --gpus $N --steps 64 --warmup 4 --no-parity > gpurun_out/${TAG}_bench_n${N}_nopush.json 2> gpurun_out/${TAG}_bench_n${N}_nopush.err; echo "bench nopush rc=$?"
for f in gpurun_out/${TAG}_bench_n$N.json gpurun_out/${TAG}_bench_n${N}_nopush.json; do grep '^{' $f | head -c 300; echo; done
