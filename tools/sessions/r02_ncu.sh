#!/bin/bash
# Round-2 ncu session (1 GPU): launch list of the bench command, --set full of the paged decode attention at the bench shape
# and at the TP shard shapes, of the prefill attention at 64 x 512 and 8 x 4096.  Every ncu run follows a plain run of the
# same command that exited 0.
mkdir -p gpurun_out
NCU="ncu --clock-control none"
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-parity"
$B > gpurun_out/n_plain_bench.log 2>&1 && $NCU --metrics gpu__time_duration.sum -c 3000 --csv --log-file gpurun_out/r02_launches.csv $B > gpurun_out/n_ncu_bench.log 2>&1; echo "launch list rc=$?"
i=0
for sh in "256 769 16 8" "512 1536 4 1" "512 1536 8 2" "512 1536 16 4"; do set -- $sh; i=$((i+1))
  D="python tools/bench_decode.py --batch $1 --ctx $2 --heads $3 --kv-heads $4 --splits 1 --iters 3 --warmup 2"
  $D > gpurun_out/n_plain_dec$i.log 2>&1 && $NCU --set full --import-source on -k regex:paged_decode_kernel -s 2 -c 1 -o gpurun_out/r02_decode_$1x$2_hkv$4 $D > gpurun_out/n_ncu_dec$i.log 2>&1; echo "decode $sh rc=$?"
  cat gpurun_out/n_plain_dec$i.log | tail -1 | cut -c1-300
done
for o in "cfg2" "4K"; do
  P="python tools/bench_prefill.py --iters 2 --only $o"
  $P > gpurun_out/n_plain_pre_$o.log 2>&1 && $NCU --set full --import-source on -k regex:prefill_attn -s 2 -c 1 -o gpurun_out/r02_prefill_$o $P > gpurun_out/n_ncu_pre_$o.log 2>&1; echo "prefill $o rc=$?"
  tail -1 gpurun_out/n_plain_pre_$o.log | cut -c1-200
done
ls -la gpurun_out/*.ncu-rep
