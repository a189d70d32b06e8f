#!/bin/bash
# A/B of the programmatic launch of the per-sample fitting kernel at N GPUs: all-reduce every step (k=1) and every 8 steps
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
P=29610
for pm in 0 1; do for k in 1 8; do
  P=$((P+1))
  if [ $pm = 1 ]; then export SMPLB_PDL_MAIN=1; else unset SMPLB_PDL_MAIN; fi
  $TR --master-port $P bench.py --gpus $N --mode fit --steps 400 --warmup 16 --graph --allreduce-every $k 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('pdl_main', $pm, 'k', $k, d['value'], d['ms_per_step'])"
done; done
