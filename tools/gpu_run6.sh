#!/bin/bash
# full GPU suite on the default path, forward parity with the L2-chunked schedule, then a chunk-size scan of the bench.
# RECORD ONLY: the env-switched chunked schedule (SMPLB_FWD_CHUNK / SMPLB_FWD_PINGPONG in smplb_forward) was measured
# with this script, lost (profiles/r01_blend_v4_ncu_summary.md) and was removed again; the variables are no-ops now.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/tests_all.log
tail -3 gpurun_out/tests_all.log
SMPLB_FWD_CHUNK=256 timeout 600 python -m pytest tests/test_forward_gpu.py -m gpu -x -q 2>&1 | tail -8 > gpurun_out/tests_fwd_chunk.log
tail -3 gpurun_out/tests_fwd_chunk.log
: > gpurun_out/chunk_scan.jsonl
for cfg in "0 1" "512 1" "768 1" "1024 1" "1024 0" "1536 1" "2048 1" "4096 1"; do
  set -- $cfg
  echo "chunk=$1 pingpong=$2" >> gpurun_out/chunk_scan.jsonl
  SMPLB_FWD_CHUNK=$1 SMPLB_FWD_PINGPONG=$2 timeout 300 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline >> gpurun_out/chunk_scan.jsonl 2> gpurun_out/chunk_scan.err
done
python - <<'PY'
import json
for l in open('gpurun_out/chunk_scan.jsonl'):
    l = l.strip()
    if l.startswith('{'):
        d = json.loads(l); print('  ms/step', round(d['ms_per_step'], 3), d.get('kernel_ms'), d['clocks']['sm_mhz'])
    else:
        print(l)
PY
