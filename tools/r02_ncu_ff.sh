#!/bin/bash
set -u
mkdir -p gpurun_out
FF_TIME=65536 python tools/ff_check.py 8 > gpurun_out/plain_ff.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:fused_fwd -s 5 -c 1 -o gpurun_out/r02_ff_prof env FF_TIME=65536 python tools/ff_check.py 8 > gpurun_out/ncu_ff.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_ff.log; cat gpurun_out/plain_ff.log
