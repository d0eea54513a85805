#!/bin/bash
# N = 2: NCCL equivalence check + the training bench line under torchrun
T=${1:-r2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --verify-ddp > gpurun_out/${T}_verify_ddp_n2.json 2> gpurun_out/verify_ddp.err; echo "verify rc=$?"; cat gpurun_out/${T}_verify_ddp_n2.json | cut -c1-600
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 3 --task train > gpurun_out/${T}_bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/${T}_bench_n2.json
