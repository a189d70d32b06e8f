#!/bin/bash
set -u
mkdir -p gpurun_out
python bench.py > gpurun_out/r02b_bench_n1.json 2> gpurun_out/err_b.log; echo "bench rc=$?"; cut -c1-260 gpurun_out/r02b_bench_n1.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02b_bench_reference_n1.json 2>> gpurun_out/err_b.log; echo "ref rc=$?"; cut -c1-260 gpurun_out/r02b_bench_reference_n1.json
python bench.py --mode fit --graph --steps 400 --warmup 16 > gpurun_out/r02b_bench_fit_graph_n1.json 2>> gpurun_out/err_b.log; echo "fit rc=$?"; cut -c1-230 gpurun_out/r02b_bench_fit_graph_n1.json
python bench.py --mode fit --graph --allreduce-every 8 --steps 400 --warmup 16 > gpurun_out/r02b_bench_fit_graph_k8_n1.json 2>> gpurun_out/err_b.log; cut -c1-230 gpurun_out/r02b_bench_fit_graph_k8_n1.json
python bench.py --mode fit --no-graph --steps 400 --warmup 16 > gpurun_out/r02b_bench_fit_eager_n1.json 2>> gpurun_out/err_b.log; cut -c1-230 gpurun_out/r02b_bench_fit_eager_n1.json
python bench.py --mode fit --loss mesh --graph --steps 50 --warmup 5 > gpurun_out/r02b_bench_fit_mesh_n1.json 2>> gpurun_out/err_b.log; cut -c1-230 gpurun_out/r02b_bench_fit_mesh_n1.json
python bench.py --mode sweep --steps 10 > gpurun_out/r02b_sweep_n1.json 2>> gpurun_out/err_b.log; echo "sweep rc=$?"
tail -3 gpurun_out/err_b.log
