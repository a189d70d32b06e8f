#!/bin/bash
for d in 0 1 2 3; do
SMPLB_DBG=$d python bench.py --mode sweep --sweep-min 65536 --steps 10 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
r=d['rows'][-1]; print('DBG=$d', {k: round(v,3) for k,v in r.get('kernel_ms_bwd_rank0',{}).items()})"
done
