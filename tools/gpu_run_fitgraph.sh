#!/bin/bash
# multi-GPU fit step from a CUDA graph (kernels + NCCL all-reduce): tools/gpu_run_fitgraph.sh N
set -u
N=${1:-2}
mkdir -p gpurun_out
for g in "" "--graph"; do
  timeout 90 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $N --mode fit --steps 200 --warmup 5 $g > gpurun_out/bench_fit${g#--}_n$N.json 2> gpurun_out/bench_fit${g#--}_n$N.err
  cut -c1-260 gpurun_out/bench_fit${g#--}_n$N.json; tail -n 2 gpurun_out/bench_fit${g#--}_n$N.err | cut -c1-300
done
