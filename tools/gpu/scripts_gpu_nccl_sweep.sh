#!/bin/bash
# dp8 step time against the number of NCCL channels (CTAs the all-reduce kernels take away from the persistent GEMMs)
mkdir -p gpurun_out
N=${1:-8}
for ch in 8 16; do
NCCL_MAX_NCHANNELS=$ch timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$((ch % 10)) bench.py --gpus $N --steps 10 --warmup 3 --no-e2e --no-optimizer --no-kernel-timing > gpurun_out/bench_n${N}_ch$ch.json 2> gpurun_out/bench_n${N}_ch$ch.err; echo "ch=$ch rc=$?"
cut -c1-140 gpurun_out/bench_n${N}_ch$ch.json
done
