#!/bin/bash
# LL all-reduce phase timeline at 2 and N ranks (debug build).
N=${1:-4}
mkdir -p gpurun_out
export NVLLM_LIB_PATH=nano-vllm-rs_b200/lib/libnvllm_timeline.so
for W in 2 $N; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port $((29720 + W)) tools/ar_timeline.py > gpurun_out/c7_ar_timeline_w$W.jsonl 2> gpurun_out/c7_ar_timeline_w$W.err; echo "timeline w=$W rc=$?"
  grep '^{' gpurun_out/c7_ar_timeline_w$W.jsonl | head -4 | cut -c1-900; tail -c 300 gpurun_out/c7_ar_timeline_w$W.err
done
