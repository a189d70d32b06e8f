#!/bin/bash
# forward parity tests + short bench + sweep (after a blend kernel change)
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_forward_gpu.py -m gpu -x -q 2>&1 | tail -15 > gpurun_out/tests_fwd.log
tail -5 gpurun_out/tests_fwd.log
timeout 300 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err
cut -c1-1800 gpurun_out/bench_short.json; tail -3 gpurun_out/bench_short.err
if [ "${1:-}" = "sweep" ]; then
  timeout 600 python bench.py --mode sweep --steps 20 > gpurun_out/sweep_n1.json 2> gpurun_out/sweep_n1.err
  cut -c1-600 gpurun_out/sweep_n1.json; tail -3 gpurun_out/sweep_n1.err
fi
