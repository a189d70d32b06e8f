#!/bin/bash
# after the per-sample kernel restructuring: strong-scaling forward and fitting records at N = $1 GPUs
set -u
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
$TR --master-port 29511 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/r02b_bench_strong_n$N.json 2> gpurun_out/err_n$N.log; echo "strong rc=$?"; cut -c1-300 gpurun_out/r02b_bench_strong_n$N.json
$TR --master-port 29512 bench.py --gpus $N --mode fit --steps 400 --warmup 16 --graph --allreduce-every 8 > gpurun_out/r02b_bench_fit_graph_k8_n$N.json 2>> gpurun_out/err_n$N.log; echo "fit k8 rc=$?"; cut -c1-300 gpurun_out/r02b_bench_fit_graph_k8_n$N.json
$TR --master-port 29513 bench.py --gpus $N --mode fit --steps 400 --warmup 16 --graph --allreduce-every 32 > gpurun_out/r02b_bench_fit_graph_k32_n$N.json 2>> gpurun_out/err_n$N.log; echo "fit k32 rc=$?"; cut -c1-300 gpurun_out/r02b_bench_fit_graph_k32_n$N.json
$TR --master-port 29514 bench.py --gpus $N --mode fit --steps 400 --warmup 16 --graph --allreduce-every 1 > gpurun_out/r02b_bench_fit_graph_k1_n$N.json 2>> gpurun_out/err_n$N.log; echo "fit k1 rc=$?"; cut -c1-300 gpurun_out/r02b_bench_fit_graph_k1_n$N.json
tail -3 gpurun_out/err_n$N.log
