#!/bin/bash
# Paged decode attention: (stages, warps) instantiations at the TP-shard shapes with one split (1 GPU).
mkdir -p gpurun_out
O=gpurun_out/c11_decode_cfg_sweep.jsonl; : > $O
for cfg in 0 1 2 3 4 5 6; do
  for hk in "4 1" "8 2" "16 4"; do set -- $hk
    NVLLM_DECODE_CFG=$cfg timeout 100 python tools/bench_decode.py --batch 512 --ctx 1536 --heads $1 --kv-heads $2 --splits 1 >> $O 2>> gpurun_out/c11.err
  done
  NVLLM_DECODE_CFG=$cfg timeout 100 python tools/bench_decode.py --batch 256 --ctx 768 --heads 16 --kv-heads 8 --splits 1 >> $O 2>> gpurun_out/c11.err
done
python - <<'PY'
import json
for l in open('gpurun_out/c11_decode_cfg_sweep.jsonl'):
    r = json.loads(l)
    print('cfg', r['cfg'], r['batch'], r['ctx'], r['heads'], r['kv_heads'], 'ms', r['ms_median'], 'GB/s', r['GBps_median'])
PY
tail -c 300 gpurun_out/c11.err
