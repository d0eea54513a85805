#!/bin/bash
# launch list with caches NOT flushed between kernels (closer to the durations inside a graph replay)
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-kernel-timing --no-optimizer --no-graph"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 6000 --csv --log-file gpurun_out/launches_warm.csv $CMD > gpurun_out/ncu_warm.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/launches_warm.csv
