#!/bin/bash
# everything: full GPU suite, smoke, every bench mode (records under gpurun_out/)
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -15 > gpurun_out/tests_all.log; tail -4 gpurun_out/tests_all.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 300 python bench.py --mode fit --steps 100 --warmup 5 > gpurun_out/bench_fit.json 2> gpurun_out/bench_fit.err
timeout 300 python bench.py --mode fit --steps 100 --warmup 5 --graph > gpurun_out/bench_fit_graph.json 2>> gpurun_out/bench_fit.err
timeout 300 python bench.py --mode fit --loss mesh --steps 50 --warmup 5 > gpurun_out/bench_fit_mesh.json 2>> gpurun_out/bench_fit.err
timeout 600 python bench.py --mode sweep --steps 20 > gpurun_out/sweep_n1.json 2> gpurun_out/sweep_n1.err
timeout 400 python bench.py --mode raster --steps 20 --warmup 3 --cpu-seconds 8 > gpurun_out/bench_raster.json 2> gpurun_out/bench_raster.err
timeout 400 python bench.py --mode normals --steps 10 --warmup 3 --cpu-seconds 6 > gpurun_out/bench_normals.json 2> gpurun_out/bench_normals.err
for f in bench_default bench_ref bench_fit bench_fit_graph bench_fit_mesh bench_raster bench_normals; do echo "== $f"; cut -c1-420 gpurun_out/$f.json; done
tail -2 gpurun_out/*.err | grep -v "^$" | tail -12
