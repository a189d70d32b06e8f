#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 120 python tools/diag_blend.py 300 > gpurun_out/diag.log 2>&1; echo "rc=$?" >> gpurun_out/diag.log
timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/tests2.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/bench_tc.json 2> gpurun_out/bench_tc.err
cat gpurun_out/diag.log; tail -5 gpurun_out/tests2.log; cat gpurun_out/bench_tc.json | cut -c1-1500
