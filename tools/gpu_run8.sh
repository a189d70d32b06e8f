#!/bin/bash
# full GPU suite + short forward bench + fit bench (after a pose_prep / GEMM change)
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/tests_all.log
tail -3 gpurun_out/tests_all.log
timeout 300 python bench.py --steps 30 --warmup 5 --no-e2e --no-cpu-baseline > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err
cut -c1-200 gpurun_out/bench_short.json; grep -o '"kernel_ms.*' gpurun_out/bench_short.json | cut -c1-700; tail -3 gpurun_out/bench_short.err
timeout 300 python bench.py --mode fit --graph --steps 100 --warmup 10 > gpurun_out/bench_fit_graph.json 2> gpurun_out/bench_fit.err
cut -c1-300 gpurun_out/bench_fit_graph.json; tail -3 gpurun_out/bench_fit.err
