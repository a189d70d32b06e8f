#!/bin/bash
# full GPU suite + default bench + fit bench
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -15 > gpurun_out/tests_all.log
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
timeout 300 python bench.py --mode fit --steps 100 --warmup 5 > gpurun_out/bench_fit.json 2> gpurun_out/bench_fit.err
timeout 300 python bench.py --mode fit --steps 100 --warmup 5 --graph > gpurun_out/bench_fit_graph.json 2>> gpurun_out/bench_fit.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
tail -5 gpurun_out/tests_all.log; cut -c1-2500 gpurun_out/bench_default.json; cat gpurun_out/bench_fit.json gpurun_out/bench_fit_graph.json gpurun_out/bench_ref.json | cut -c1-900; tail -3 gpurun_out/bench_fit.err
