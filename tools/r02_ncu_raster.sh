#!/bin/bash
set -u
mkdir -p gpurun_out
BCMD="python bench.py --mode raster --steps 2 --warmup 3 --no-cpu-baseline"
$BCMD > gpurun_out/plain_r.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"raster_tri_kernel|raster_resolve4_kernel" -s 4 -c 2 -o gpurun_out/r02_raster $BCMD > gpurun_out/ncu_r.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_r.log
