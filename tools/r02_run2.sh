#!/bin/bash
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r02_pytest_gpu.log
python bench.py --steps 10 --warmup 3 --cpu-seconds 3 > gpurun_out/r02_bench_n1.json 2> gpurun_out/bench_err.log; echo "bench rc=$?"; tail -3 gpurun_out/bench_err.log; cut -c1-1500 gpurun_out/r02_bench_n1.json
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --graph --batch 8192 > gpurun_out/r02_bench_n1_b8192_graph.json 2>> gpurun_out/bench_err.log; echo "bench graph rc=$?"; cut -c1-400 gpurun_out/r02_bench_n1_b8192_graph.json
python bench.py --mode fit --steps 200 --warmup 8 --graph --allreduce-every 8 > gpurun_out/r02_bench_fit_graph_n1.json 2>> gpurun_out/bench_err.log; echo "fit rc=$?"; cat gpurun_out/r02_bench_fit_graph_n1.json
python bench.py --mode fit --loss mesh --steps 20 --warmup 3 > gpurun_out/r02_bench_fit_mesh_n1.json 2>> gpurun_out/bench_err.log; echo "fitmesh rc=$?"; cat gpurun_out/r02_bench_fit_mesh_n1.json
tail -5 gpurun_out/bench_err.log
