#!/bin/bash
# reference-pin GPU tests (+ optional extra pytest args)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_reference_pin_gpu.py "$@" -m gpu -q 2>&1 | tail -60 > gpurun_out/tests_pin.log
cat gpurun_out/tests_pin.log
