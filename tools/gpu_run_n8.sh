#!/bin/bash
# N-GPU forward bench only (short): tools/gpu_run_n8.sh N
set -u
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/smi_n$N.txt
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 20 --warmup 3 --e2e-steps 2 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $N --mode fit --steps 100 --warmup 5 > gpurun_out/bench_fit_n$N.json 2> gpurun_out/bench_fit_n$N.err
cut -c1-330 gpurun_out/bench_n$N.json; grep -o '"kernel_ms.*"roofline' gpurun_out/bench_n$N.json; grep -o '"e2e.*' gpurun_out/bench_n$N.json | cut -c1-200; cut -c1-300 gpurun_out/bench_fit_n$N.json; tail -n 2 gpurun_out/bench_n$N.err | cut -c1-300
