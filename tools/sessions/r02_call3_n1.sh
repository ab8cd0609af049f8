#!/bin/bash
# Round-2 GPU session 3 (1 GPU): the whole -m gpu suite (every file, no -x), smoke(), the fused decode kernel A/B on the N=1
# bench line, the split-K GEMM phase timeline, per-launch profile with the fused kernel.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/c3_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/c3_pytest.log
tail -c 1800 gpurun_out/c3_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c3_smoke.log 2>&1; echo "smoke rc=$?"; tail -c 300 gpurun_out/c3_smoke.log
if grep -q "fused" gpurun_out/c3_pytest.log && grep -E "FAILED.*fused" gpurun_out/c3_pytest.log; then export NVLLM_FUSE_DECODE=0; echo "fused decode RED: disabled"; fi
timeout 600 python bench.py --steps 64 --warmup 4 > gpurun_out/c3_bench_n1.json 2> gpurun_out/c3_bench_n1.err; echo "bench n1 rc=$?"
NVLLM_FUSE_DECODE=0 timeout 600 python bench.py --steps 64 --warmup 4 --no-cpu-baseline > gpurun_out/c3_bench_n1_unfused.json 2> gpurun_out/c3_bench_n1_unfused.err; echo "bench n1 unfused rc=$?"
timeout 300 python bench.py --profile --steps 8 --no-cpu-baseline --no-parity > gpurun_out/c3_bench_n1_profile.json 2> gpurun_out/c3_profile.err; echo "profile rc=$?"
for sh in o down qkv tp8_o tp8_qkv; do
  NVLLM_LIB_PATH=nano-vllm-rs_b200/lib/libnvllm_timeline.so timeout 120 python tools/gemm_timeline.py --shape $sh >> gpurun_out/c3_timeline.jsonl 2>> gpurun_out/c3_timeline.err; echo "timeline $sh rc=$?"
done
head -c 1200 gpurun_out/c3_bench_n1.json; echo
head -c 600 gpurun_out/c3_bench_n1_unfused.json; echo
cat gpurun_out/c3_timeline.jsonl
