#!/bin/bash
# gpurun with retries while the pod has no free slot (exit code 3 / "transient"); usage: tools/gpu.sh <timeout> <log> '<command>'
T=$1; LOG=$2; shift 2
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --timeout $T -- "$@" > $LOG 2>&1
  rc=$?
  if grep -q "status=transient\|status=busy" $LOG || [ $rc -eq 3 ]; then sleep 45; continue; fi
  break
done
exit $rc
