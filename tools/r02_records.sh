#!/bin/bash
# final records of the forward step: launch list + ncu --set full of blend_tc_kernel and skin_tc_kernel<0> (one ncu use per call)
set -u
mkdir -p gpurun_out
BCMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$BCMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 12 -c 8 --csv --log-file gpurun_out/r02_launches_fwd.csv $BCMD > gpurun_out/ncu_l.log 2>&1
echo "launch list rc=$?"
$BCMD > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"blend_tc_kernel|skin_tc_kernel" -s 6 -c 2 -o gpurun_out/r02_fwd_final $BCMD > gpurun_out/ncu_f.log 2>&1
echo "full rc=$?"; tail -2 gpurun_out/ncu_f.log
python tools/bwd_probe.py 65536 2 > gpurun_out/plain3.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed --clock-control none -s 8 -c 8 --csv --log-file gpurun_out/r02_launches_bwd.csv python tools/bwd_probe.py 65536 2 > gpurun_out/ncu_b.log 2>&1
echo "bwd list rc=$?"
