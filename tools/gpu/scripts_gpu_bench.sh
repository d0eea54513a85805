#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r1.json 2> gpurun_out/bench_r1.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_r1.err; cat gpurun_out/bench_r1.json
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-kernel-timing --no-optimizer"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu.log; wc -l gpurun_out/launches.csv
