#!/bin/bash
# ncu per-launch list of one bench process (4 steps), standard flags.
mkdir -p gpurun_out
T=${1:-r1_e}
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-kernel-timing --no-optimizer --no-graph"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 || { tail -5 gpurun_out/plain.log; exit 1; }
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/${T}_launches.csv $CMD > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
