#!/bin/bash
# ncu --set full with source correlation of ONE tcgen05 GEMM launch of tools/gemm_bench.py shape index $1 (default 5:
# 19200x768x768 bias+res); exports raw + source pages as CSV into gpurun_out/
IDX=${1:-5}
TAG=${2:-gemm$IDX}
timeout 600 python tools/gemm_bench.py 2 $IDX > gpurun_out/${TAG}_plain.txt 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 3 -c 1 -f -o /tmp/prof_${TAG} python tools/gemm_bench.py 2 $IDX > gpurun_out/${TAG}_ncu.log 2>&1
ncu -i /tmp/prof_${TAG}.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
ncu -i /tmp/prof_${TAG}.ncu-rep --page source --csv > gpurun_out/${TAG}_source.csv 2>/dev/null
ncu -i /tmp/prof_${TAG}.ncu-rep --page details > gpurun_out/${TAG}_details.txt 2>/dev/null
wc -l gpurun_out/${TAG}_source.csv
