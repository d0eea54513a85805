#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
for mb in 8 16 64; do
SVPC_DDP_BUCKET_MB=$mb timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 10 --warmup 3 --no-e2e --no-kernel-timing --no-optimizer > gpurun_out/bench_n${N}_b$mb.json 2> gpurun_out/bench_n${N}_b$mb.err; echo "bucket $mb MB rc=$?"
cut -c1-170 gpurun_out/bench_n${N}_b$mb.json
done
