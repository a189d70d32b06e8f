#!/bin/bash
# rasteriser parity tests (+ optional args: extra pytest files)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_raster_gpu.py tests/test_render_gpu.py "$@" -m gpu -x -q 2>&1 | tail -25 > gpurun_out/tests_raster.log
cat gpurun_out/tests_raster.log
