#!/bin/bash
# ncu --set full of the tensor-core LSTM kernels and the attention backward (after a plain run exits 0).
mkdir -p gpurun_out
timeout 300 python bench_train.py --steps 1 --warmup 3 --skip-cpu > gpurun_out/plain_train.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_train.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lstm_mma_cluster_kernel -s 3 -c 1 \
    -o gpurun_out/prof_lstm_fwd_train -f python bench_train.py --steps 1 --warmup 3 --skip-cpu > gpurun_out/ncu_lstm_fwd.log 2>&1
echo "ncu lstm fwd (train) exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lstm_bwd_cluster_kernel -s 3 -c 1 \
    -o gpurun_out/prof_lstm_bwd -f python bench_train.py --steps 1 --warmup 3 --skip-cpu > gpurun_out/ncu_lstm_bwd.log 2>&1
echo "ncu lstm_bwd exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mha_bwd_tc_kernel -s 12 -c 1 \
    -o gpurun_out/prof_mha_bwd -f python bench_train.py --steps 1 --warmup 3 --skip-cpu > gpurun_out/ncu_mha_bwd.log 2>&1
echo "ncu mha_bwd exit $?"
timeout 300 python bench.py --steps 2 --warmup 3 --skip-cpu > gpurun_out/plain.log 2>&1 || { echo "plain infer run failed"; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lstm_mma_cluster_kernel -s 3 -c 1 \
    -o gpurun_out/prof_lstm_fwd_infer -f python bench.py --steps 2 --warmup 3 --skip-cpu > gpurun_out/ncu_lstm_infer.log 2>&1
echo "ncu lstm fwd (infer) exit $?"
ls -la gpurun_out/*.ncu-rep
