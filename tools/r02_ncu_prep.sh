#!/bin/bash
set -u
mkdir -p gpurun_out
BCMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$BCMD > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"pose_prep_kernel" -s 3 -c 1 -o gpurun_out/r02_prep $BCMD > gpurun_out/ncu_p.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_p.log
