set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$TR --nproc-per-node 8 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/scale_c2_n8.json 2> gpurun_out/scale_c2_n8.err
$TR --nproc-per-node 8 --master-port 29512 bench.py --workload c3 --gpus 8 --steps 5 --warmup 3 > gpurun_out/scale_c3_n8.json 2> gpurun_out/scale_c3_n8.err
$TR --nproc-per-node 4 --master-port 29513 bench.py --workload c3 --gpus 4 --steps 5 --warmup 3 > gpurun_out/scale_c3_n4.json 2> gpurun_out/scale_c3_n4.err
$TR --nproc-per-node 2 --master-port 29514 bench.py --workload c3 --gpus 2 --steps 4 --warmup 2 > gpurun_out/scale_c3_n2.json 2> gpurun_out/scale_c3_n2.err
for n in 2 4 8; do $TR --nproc-per-node $n --master-port 2952$n bench.py --workload c4 --gpus $n --steps 5 --warmup 3 > gpurun_out/scale_c4_n$n.json 2> gpurun_out/scale_c4_n$n.err; done
$TR --nproc-per-node 8 --master-port 29531 bench.py --workload c4 --recordings 4 --gpus 8 --steps 5 --warmup 3 > gpurun_out/scale_c4r4_n8.json 2> gpurun_out/scale_c4r4_n8.err
python bench.py --workload c3 --steps 4 --warmup 2 > gpurun_out/scale_c3_n1.json 2> gpurun_out/scale_c3_n1.err
python bench.py --workload c4 --steps 4 --warmup 2 > gpurun_out/scale_c4_n1.json 2> gpurun_out/scale_c4_n1.err
python bench.py --workload c4 --recordings 4 --steps 4 --warmup 2 --skip-cpu > gpurun_out/scale_c4r4_n1.json 2> gpurun_out/scale_c4r4_n1.err
echo finished
