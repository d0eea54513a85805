#!/bin/bash
# N = 2: NCCL equivalence check, then the training step with this build / a base build of the library (SVPC_LIB_PATH), and the
# same at N = 1
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --verify-ddp > gpurun_out/ab_verify_n2.json 2> gpurun_out/ab_verify.err; echo "verify rc=$?"; python -c "
import json; v=json.loads(open('gpurun_out/ab_verify_n2.json').read().strip().splitlines()[-1])['verify_ddp']; print({k:(v[k].get('ok'), v[k].get('worst_grad_rel_l2')) for k in ('fp32','bf16')})"
Q="--task train --no-e2e --no-optimizer --no-kernel-timing --no-cpu-baseline --no-parity --steps 20 --warmup 3"
BASE=$PWD/svpc_b200/libsvpc_b200_base.so
for v in "X=1" "SVPC_LIB_PATH=$BASE" "SVPC_TC_DYNAMIC=0" "X=2" "SVPC_LIB_PATH=$BASE Y=2"; do
  env $v python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 $Q 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('N=2 ${v%%/*}', d['value'], d['ms_per_step'])"
done
for v in "X=1" "SVPC_LIB_PATH=$BASE" "SVPC_TC_DYNAMIC=0" "X=2" "SVPC_LIB_PATH=$BASE Y=2"; do
  env $v python bench.py $Q 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('N=1 ${v%%/*}', d['value'], d['ms_per_step'])"
done
