#!/bin/bash
# First GPU pass: parity tests, bench with the SIMT blend, launch list + one full ncu capture of skin_kernel.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt
python -m pytest tests -m gpu -q -x --deselect tests/test_forward_gpu.py::test_dropin_smpl_class -k "not 0-" 2>&1 | tail -40 > gpurun_out/tests1.log
python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/tests1_all.log
python bench.py --blend simt --steps 10 --warmup 3 > gpurun_out/bench_simt.json 2> gpurun_out/bench_simt.err
BCMD="python bench.py --blend simt --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$BCMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_simt.csv $BCMD > gpurun_out/ncu1.log 2>&1
$BCMD > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:skin_kernel -s 3 -c 1 -o gpurun_out/skin_v1 $BCMD > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out
