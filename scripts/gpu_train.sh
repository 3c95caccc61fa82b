#!/bin/bash
# One GPU-box visit for the training step: bench (bf16x3 + bf16), then the ncu launch list of a short run.
mkdir -p gpurun_out
for p in bf16x3 bf16; do
timeout 600 python bench_train.py --steps 10 --warmup 3 --precision $p ${SKIP_CPU:+--skip-cpu} > gpurun_out/bench_train_$p.json 2> gpurun_out/bench_train_$p.err; echo "train bench $p exit $?"
cat gpurun_out/bench_train_$p.json; tail -5 gpurun_out/bench_train_$p.err
done
if [ -z "$SKIP_NCU" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv \
    --log-file gpurun_out/launches_train.csv python bench_train.py --steps 1 --warmup 3 --precision bf16x3 --skip-cpu > gpurun_out/ncu_train.log 2>&1
echo "ncu exit $?"
fi
