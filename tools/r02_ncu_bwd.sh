#!/bin/bash
# ncu --set full of the backward kernels (second iteration of tools/bwd_probe.py); $1 = kernel regex
set -u
mkdir -p gpurun_out
PAT=${1:-skin_tc_kernel}
python tools/bwd_probe.py 65536 2 > gpurun_out/plain_bwd.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"$PAT" -s ${2:-2} -c ${3:-2} -o gpurun_out/r02_bwd_prof python tools/bwd_probe.py 65536 2 > gpurun_out/ncu_bwd.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_bwd.log
