#!/bin/bash
# compute-sanitizer over the hot-path kernels at small batches; ONE tool per gpurun call (B200_PROFILING.md):
#   tools/sanitize.sh memcheck | racecheck      -> gpurun_out/r02_sanitize_<tool>.log
set -u
TOOL=${1:-memcheck}
mkdir -p gpurun_out
python tools/sanitize_probe.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 900 compute-sanitizer --tool $TOOL --print-limit 30 python tools/sanitize_probe.py > gpurun_out/r02_sanitize_$TOOL.log 2>&1
echo "compute-sanitizer $TOOL rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|ok,|done|Error|hazard" gpurun_out/r02_sanitize_$TOOL.log | head -30
