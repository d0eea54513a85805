#!/bin/bash
# N-rank training step under different NCCL CTA budgets / persistent GEMM grid caps (SVPC_TC_GRID)
N=${1:-2}
shift
for v in "$@"; do
  env $v python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 20 --warmup 3 --task train --no-e2e --no-optimizer --no-kernel-timing --no-cpu-baseline --no-parity 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', d['value'], d['ms_per_step'])"
done
