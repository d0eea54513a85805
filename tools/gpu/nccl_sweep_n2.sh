#!/bin/bash
# N = 2 training step under different NCCL CTA budgets (does the collective take SMs from the persistent GEMMs?)
N=${1:-2}
for v in "X=1" "NCCL_MAX_CTAS=4" "NCCL_MAX_CTAS=8" "NCCL_MAX_CTAS=16" "SVPC_WGRAD_STREAM=0"; do
  env $v python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 20 --warmup 3 --task train --no-e2e --no-optimizer --no-kernel-timing --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', d['value'], d['ms_per_step'])"
done
