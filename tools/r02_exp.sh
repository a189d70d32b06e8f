#!/bin/bash
for d in 63 127 191 255; do echo "DBG=$d"; SMPLB_DBG=$d FF_TIME=65536 python tools/ff_check.py 8 2>&1 | grep "fused forward"; done
