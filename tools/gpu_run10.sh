#!/bin/bash
# skinning unit-count scan (units = 54 vertex tiles x n_chunks, n_chunks = waves * 148 / 54)
set -u
mkdir -p gpurun_out
: > gpurun_out/skin_waves.txt
for w in 3 7 15 3 7 15; do
  SMPLB_SKIN_WAVES=$w timeout 300 python bench.py --steps 30 --warmup 5 --no-e2e --no-cpu-baseline > gpurun_out/sw.json 2> gpurun_out/sw.err
  echo "waves=$w $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/sw.json) $(grep -o '"skinning": [0-9.]*' gpurun_out/sw.json) $(grep -o '"blend_gemm": [0-9.]*' gpurun_out/sw.json)" | tee -a gpurun_out/skin_waves.txt
done
