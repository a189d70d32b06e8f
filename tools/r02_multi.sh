#!/bin/bash
# multi-GPU records: $1 = number of GPUs on the box; runs N = 2 .. $1
set -u
mkdir -p gpurun_out
NMAX=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
P=29500
for N in 2 4 8; do
  [ $N -le $NMAX ] || continue
  P=$((P+1))
  $TR --nproc-per-node $N --master-port $P bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/r02_bench_strong_n$N.json 2> gpurun_out/err_n$N.log; echo "strong n=$N rc=$?"; cut -c1-330 gpurun_out/r02_bench_strong_n$N.json
  P=$((P+1))
  $TR --nproc-per-node $N --master-port $P bench.py --gpus $N --steps 20 --warmup 5 --scaling weak --no-e2e > gpurun_out/r02_bench_weak_n$N.json 2>> gpurun_out/err_n$N.log; echo "weak n=$N rc=$?"; cut -c1-250 gpurun_out/r02_bench_weak_n$N.json
  P=$((P+1))
  $TR --nproc-per-node $N --master-port $P bench.py --gpus $N --mode sweep --steps 10 --no-cpu-baseline > gpurun_out/r02_sweep_n$N.json 2>> gpurun_out/err_n$N.log; echo "sweep n=$N rc=$?"
  P=$((P+1))
  $TR --nproc-per-node $N --master-port $P bench.py --gpus $N --mode fit --steps 400 --warmup 8 --graph --allreduce-every 8 > gpurun_out/r02_bench_fit_graph_k8_n$N.json 2>> gpurun_out/err_n$N.log; echo "fit k8 n=$N rc=$?"; cut -c1-330 gpurun_out/r02_bench_fit_graph_k8_n$N.json
  P=$((P+1))
  $TR --nproc-per-node $N --master-port $P bench.py --gpus $N --mode fit --steps 400 --warmup 8 --graph --allreduce-every 1 > gpurun_out/r02_bench_fit_graph_k1_n$N.json 2>> gpurun_out/err_n$N.log; echo "fit k1 n=$N rc=$?"; cut -c1-330 gpurun_out/r02_bench_fit_graph_k1_n$N.json
  tail -3 gpurun_out/err_n$N.log
done
nvidia-smi topo -m > gpurun_out/r02_topo_n$NMAX.txt 2>&1
