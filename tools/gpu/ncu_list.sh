#!/bin/bash
# Per-launch list of the training step (eager launches, 4 steps in the process): gpurun_out/<tag>_launches.csv
# read with: python profiles/agg_launches.py gpurun_out/<tag>_launches.csv 4
T=${1:-r2}
shift
CMD="python bench.py --task train --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-kernel-timing --no-optimizer --no-graph --no-parity $@"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 || { tail -5 gpurun_out/plain.log; exit 1; }
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/${T}_launches.csv $CMD > gpurun_out/ncu.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/${T}_launches.csv
