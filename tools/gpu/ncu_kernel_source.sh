#!/bin/bash
# ncu --set full + source page of ONE launch of kernel regex $1 from command "$3..." ; tag $2
RX=$1; TAG=$2; shift 2
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$RX -s 2 -c 1 -f -o /tmp/prof_${TAG} "$@" > gpurun_out/${TAG}_ncu.log 2>&1
ncu -i /tmp/prof_${TAG}.ncu-rep --page source --csv > gpurun_out/${TAG}_source.csv 2>/dev/null
ncu -i /tmp/prof_${TAG}.ncu-rep --page details > gpurun_out/${TAG}_details.txt 2>/dev/null
wc -l gpurun_out/${TAG}_source.csv
