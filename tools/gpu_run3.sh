#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest ${1:-tests} -m gpu -q -x 2>&1 | tail -40 > gpurun_out/tests3.log
tail -40 gpurun_out/tests3.log
