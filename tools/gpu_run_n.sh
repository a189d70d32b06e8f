#!/bin/bash
# multi-GPU bench: tools/gpu_run_n.sh N
set -u
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/smi_n$N.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --mode fit --steps 100 --warmup 5 > gpurun_out/bench_fit_n$N.json 2> gpurun_out/bench_fit_n$N.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err
cut -c1-600 gpurun_out/bench_n$N.json; cut -c1-400 gpurun_out/bench_fit_n$N.json; cut -c1-200 gpurun_out/bench_ref_n$N.json; tail -n 3 gpurun_out/bench_n$N.err; tail -n 3 gpurun_out/bench_fit_n$N.err
