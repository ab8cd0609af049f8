#!/bin/bash
# Persistent paged decode kernel (1 GPU): parity tests, micro-bench A/B at the bench shape and the TP shard shapes, bench line.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "paged_decode" > gpurun_out/c16_pytest.log 2>&1; echo "pytest rc=$?"
tail -c 1500 gpurun_out/c16_pytest.log
O=gpurun_out/c16_decode_persist.jsonl; : > $O
for pe in 1 0; do
  for sh in "256 769 16 8" "512 1536 4 1" "512 1536 8 2" "512 1536 16 4" "512 1536 32 8" "256 4400 32 8"; do set -- $sh
    NVLLM_DECODE_PERSIST=$pe timeout 100 python tools/bench_decode.py --batch $1 --ctx $2 --heads $3 --kv-heads $4 --splits 1 | sed "s/^{/{\"persist\": $pe, /" >> $O 2>> gpurun_out/c16.err
  done
  NVLLM_DECODE_PERSIST=$pe timeout 100 python tools/bench_decode.py --batch 256 --ctx 1024 --ragged --heads 16 --kv-heads 8 --splits 1 | sed "s/^{/{\"persist\": $pe, /" >> $O 2>> gpurun_out/c16.err
done
python - <<'PY'
import json
for l in open('gpurun_out/c16_decode_persist.jsonl'):
    r = json.loads(l)
    print('persist', r['persist'], r['batch'], r['ctx'], r['ragged'], r['heads'], r['kv_heads'], 'ms', r['ms_median'], 'GB/s', r['GBps_median'])
PY
timeout 600 python bench.py --steps 64 --warmup 4 --no-cpu-baseline > gpurun_out/c16_bench_n1.json 2> gpurun_out/c16_bench_n1.err; echo "bench rc=$?"
NVLLM_DECODE_PERSIST=0 timeout 600 python bench.py --steps 64 --warmup 4 --no-cpu-baseline --no-parity > gpurun_out/c16_bench_n1_plain.json 2> gpurun_out/c16_bench_n1_plain.err; echo "bench plain rc=$?"
for f in gpurun_out/c16_bench_n1.json gpurun_out/c16_bench_n1_plain.json; do grep '^{' $f | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['achieved'], d['roofline']['share_of_step'], d.get('parity',{}).get('ok'))"; done
tail -c 300 gpurun_out/c16.err gpurun_out/c16_bench_n1.err
