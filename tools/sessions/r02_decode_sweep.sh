#!/bin/bash
# Paged decode attention: split count sweep at the TP-shard shapes of the 8B / 32B workloads (1 GPU).
mkdir -p gpurun_out
O=gpurun_out/c10_decode_sweep.jsonl; : > $O
for hk in "4 1" "8 2" "16 4" "32 8"; do set -- $hk
  timeout 200 python tools/bench_decode.py --batch 512 --ctx 1536 --heads $1 --kv-heads $2 --sweep-splits 0,1,2,3,4,5,6,8 >> $O 2>> gpurun_out/c10_decode_sweep.err
done
timeout 200 python tools/bench_decode.py --batch 64 --ctx 32512 --heads 8 --kv-heads 1 --sweep-splits 0,4,5,8,9,10,14,16,19,24,28,32 >> $O 2>> gpurun_out/c10_decode_sweep.err
timeout 200 python tools/bench_decode.py --batch 256 --ctx 768 --heads 16 --kv-heads 8 --sweep-splits 0,1,2,3 >> $O 2>> gpurun_out/c10_decode_sweep.err
timeout 200 python tools/bench_decode.py --batch 256 --ctx 4400 --heads 32 --kv-heads 8 --sweep-splits 0,1,2,3,4 >> $O 2>> gpurun_out/c10_decode_sweep.err
python - <<'PY'
import json
for l in open('gpurun_out/c10_decode_sweep.jsonl'):
    r = json.loads(l)
    print(r['batch'], r['ctx'], r['heads'], r['kv_heads'], 'splits', r['splits'], 'used', r['splits_used'], 'ms', r['ms_median'], 'GB/s', r['GBps_median'])
PY
tail -c 300 gpurun_out/c10_decode_sweep.err
