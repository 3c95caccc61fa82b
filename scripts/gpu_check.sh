#!/bin/bash
# One GPU-box visit: per-file parity tests (own process + timeout each), smoke, short bench,
# then the ncu launch list of the same bench command.  Everything is logged under gpurun_out/.
PREC=${PREC:-bf16x3}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt 2>&1
: > gpurun_out/summary.txt
if [ -z "$SKIP_TESTS" ]; then
for f in tests/test_gpu_kernels.py tests/test_gpu_model.py; do
  n=$(basename $f .py)
  timeout 900 python -m pytest $f -q -m gpu -x --timeout=300 > gpurun_out/$n.log 2>&1
  echo "$n exit $?" | tee -a gpurun_out/summary.txt
  tail -5 gpurun_out/$n.log
done
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/summary.txt; tail -3 gpurun_out/smoke.log
fi
for p in fp32 $PREC; do
timeout 900 python bench.py --steps 10 --warmup 3 --precision $p > gpurun_out/bench_$p.json 2> gpurun_out/bench_$p.err; echo "bench $p exit $?" | tee -a gpurun_out/summary.txt
cat gpurun_out/bench_$p.json; tail -5 gpurun_out/bench_$p.err
done
if [ -z "$SKIP_NCU" ]; then
timeout 600 python bench.py --steps 2 --warmup 3 --precision $PREC --skip-cpu > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
    --log-file gpurun_out/launches_$PREC.csv python bench.py --steps 2 --warmup 3 --precision $PREC --skip-cpu > gpurun_out/ncu.log 2>&1
echo "ncu exit $?" | tee -a gpurun_out/summary.txt
fi
