#!/bin/bash
# blend GEMM operand-format change: diagnostic vs the CUDA-core path, forward parity tests, short bench
set -u
mkdir -p gpurun_out
timeout 300 python tools/diag_blend.py 600 > gpurun_out/diag_blend.log 2>&1; cat gpurun_out/diag_blend.log | cut -c1-300
timeout 900 python -m pytest tests/test_forward_gpu.py tests/test_reference_pin_gpu.py tests/test_fit_gpu.py -m gpu -x -q 2>&1 | tail -15 > gpurun_out/tests_fwd.log
tail -5 gpurun_out/tests_fwd.log
timeout 300 python bench.py --steps 30 --warmup 5 --no-e2e --no-cpu-baseline > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err
cut -c1-2000 gpurun_out/bench_short.json; tail -3 gpurun_out/bench_short.err
