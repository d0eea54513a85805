#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench N=$N rc=$?"
tail -3 gpurun_out/bench_n$N.err; cut -c1-1500 gpurun_out/bench_n$N.json
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --task decode --steps 5 --warmup 3 > gpurun_out/bench_decode_n$N.json 2> gpurun_out/bench_decode_n$N.err; echo "decode N=$N rc=$?"
cut -c1-600 gpurun_out/bench_decode_n$N.json
