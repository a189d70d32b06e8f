#!/bin/bash
# ncu: launch list + full capture of the kernels named in $1 (regex), on a short bench run.
set -u
mkdir -p gpurun_out
PAT=${1:-skin_tc_kernel}
TAG=${2:-prof}
BCMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$BCMD > gpurun_out/plain_$TAG.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_$TAG.csv $BCMD > gpurun_out/ncu_l_$TAG.log 2>&1
$BCMD > gpurun_out/plain2_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$PAT -s ${3:-3} -c ${4:-2} -o gpurun_out/$TAG $BCMD > gpurun_out/ncu_f_$TAG.log 2>&1
tail -3 gpurun_out/ncu_f_$TAG.log
