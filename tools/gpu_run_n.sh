#!/bin/bash
# multi-GPU records: forward bench + graph fit bench under torchrun on N GPUs of one box
set -u
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
timeout 300 $TR --master-port 29512 bench.py --gpus $N --mode fit --graph --steps 100 --warmup 5 > gpurun_out/bench_fit_graph_n$N.json 2> gpurun_out/bench_fit_graph_n$N.err
cut -c1-2400 gpurun_out/bench_n$N.json; tail -2 gpurun_out/bench_n$N.err; cut -c1-300 gpurun_out/bench_fit_graph_n$N.json; tail -2 gpurun_out/bench_fit_graph_n$N.err
