#!/bin/bash
# round 2, GPU call 1: tcgen05 feature probe, TF32 peak, ncu --set full of the final forward kernels
set -u
mkdir -p gpurun_out
timeout 120 ./tools/mma_probe > gpurun_out/r02_mma_probe.txt 2>&1; echo "probe rc=$?"
timeout 200 python tools/measure_tf32_peak.py > gpurun_out/r02_tf32_peak.json 2> gpurun_out/tf32_err.log; echo "tf32 rc=$?"
BCMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$BCMD > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"blend_tc_kernel|skin_tc_kernel" -s 6 -c 2 -o gpurun_out/r02_fwd_final $BCMD > gpurun_out/ncu_f.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_f.log
cat gpurun_out/r02_mma_probe.txt
cat gpurun_out/r02_tf32_peak.json
