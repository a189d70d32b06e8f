#!/bin/bash
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
python bench.py --mode sweep --sweep-min 8192 --steps 10 --no-cpu-baseline > gpurun_out/sweep_bwd.json 2> gpurun_out/sweep_err.log; echo "sweep rc=$?"; tail -3 gpurun_out/sweep_err.log
python - <<'PY'
import json
d=json.load(open('gpurun_out/sweep_bwd.json'))
for r in d['rows']:
    print(r['B_global'], 'fwd %.3f fwd+bwd %.3f bwd %.3f' % (r['fwd_ms'], r['fwd_bwd_ms'], r['bwd_ms']), {k: round(v,3) for k,v in r.get('kernel_ms_bwd_rank0',{}).items()}, {k: round(v,3) for k,v in r.get('kernel_ms_rank0',{}).items()})
PY
