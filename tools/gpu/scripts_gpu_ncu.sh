#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-kernel-timing --no-graph"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 || exit 1
timeout 1500 ncu --set full --clock-control none -k regex:"lstm_step|reasoner|pointer|gumbel|asl" -s 108 -c 36 -f -o /tmp/prof_tail $CMD > gpurun_out/ncu1.log 2>&1
echo "ncu1 rc=$?"
ncu -i /tmp/prof_tail.ncu-rep --page raw --csv > gpurun_out/prof_tail_raw.csv 2>/dev/null
timeout 1500 ncu --set full --clock-control none -k regex:"gemm_tc|attn_tc" -s 420 -c 60 -f -o /tmp/prof_gemm $CMD > gpurun_out/ncu2.log 2>&1
echo "ncu2 rc=$?"
ncu -i /tmp/prof_gemm.ncu-rep --page raw --csv > gpurun_out/prof_gemm_raw.csv 2>/dev/null
ls -la gpurun_out/ /tmp/*.ncu-rep
