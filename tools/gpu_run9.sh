#!/bin/bash
# end-of-session records: full GPU suite, default bench (both arms), fit variants, sweep, launch list
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -15 > gpurun_out/tests_all.log
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
timeout 300 python bench.py --mode fit --steps 100 --warmup 5 > gpurun_out/bench_fit.json 2> gpurun_out/bench_fit.err
timeout 300 python bench.py --mode fit --steps 100 --warmup 5 --graph > gpurun_out/bench_fit_graph.json 2>> gpurun_out/bench_fit.err
timeout 300 python bench.py --mode fit --loss mesh --steps 50 --warmup 5 > gpurun_out/bench_fit_mesh.json 2>> gpurun_out/bench_fit.err
timeout 600 python bench.py --mode sweep --steps 20 > gpurun_out/sweep_n1.json 2> gpurun_out/sweep_n1.err
BCMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_final_fwd.csv $BCMD > gpurun_out/ncu_l_final.log 2>&1
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
tail -3 gpurun_out/tests_all.log; cut -c1-2600 gpurun_out/bench_default.json; cat gpurun_out/bench_fit.json gpurun_out/bench_fit_graph.json gpurun_out/bench_fit_mesh.json gpurun_out/bench_ref.json | cut -c1-400; tail -n 3 gpurun_out/bench_fit.err; tail -n 3 gpurun_out/sweep_n1.err
