#!/bin/bash
# One GPU-box visit: per-file parity tests (own process + timeout each), smoke, short bench.
# Everything is logged under gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt 2>&1
for f in tests/test_gpu_kernels.py tests/test_gpu_model.py; do
  n=$(basename $f .py)
  timeout 900 python -m pytest $f -q -m gpu -x --timeout=300 > gpurun_out/$n.log 2>&1
  echo "$n exit $?" | tee -a gpurun_out/summary.txt
  tail -5 gpurun_out/$n.log
done
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/summary.txt; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?" | tee -a gpurun_out/summary.txt
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
timeout 900 python bench.py --steps 10 --warmup 3 --precision bf16x3 > gpurun_out/bench_bf16x3.json 2> gpurun_out/bench_bf16x3.err; echo "bench bf16x3 exit $?" | tee -a gpurun_out/summary.txt
cat gpurun_out/bench_bf16x3.json; tail -5 gpurun_out/bench_bf16x3.err
