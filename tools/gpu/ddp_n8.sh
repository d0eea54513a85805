#!/bin/bash
# N = 8: the driver's launch of the benchmark (reference arm first, then ours) + the NCCL equivalence check
T=${1:-r2}
N=${2:-8}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/${T}_bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err; echo "ref rc=$?"
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $N --steps 10 --warmup 3 ) > gpurun_out/${T}_bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/${T}_bench_n$N.json; tail -4 gpurun_out/bench_n$N.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus $N --verify-ddp > gpurun_out/${T}_verify_ddp_n$N.json 2> gpurun_out/verify_ddp_n$N.err; echo "verify rc=$?"; cut -c1-300 gpurun_out/${T}_verify_ddp_n$N.json
