#!/bin/bash
set -u
mkdir -p gpurun_out
BCMD="python bench.py --mode fit --no-graph --steps 4 --warmup 3"
$BCMD > gpurun_out/plain_fit.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"joint_bwd_kernel|loss_finish_kernel" -s 8 -c 2 -o gpurun_out/r02_fit $BCMD > gpurun_out/ncu_fit.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_fit.log
