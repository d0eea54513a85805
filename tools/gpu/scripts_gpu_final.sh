#!/bin/bash
# Round-end verification + evidence: GPU tests, smoke, benches (train / reference arm / decode), ncu launch list (standard
# flags), ncu --set full of the dominant GEMM, the optimiser pass and the sequential / attention kernels.
mkdir -p gpurun_out
T=${1:-r1_d}
timeout 1500 python -m pytest tests -q -m gpu --tb=short --maxfail=20 > gpurun_out/tests.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/tests.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/${T}_bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cut -c1-400 gpurun_out/${T}_bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_reference.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cut -c1-300 gpurun_out/${T}_bench_reference.json
timeout 600 python bench.py --task decode --steps 5 --warmup 3 > gpurun_out/${T}_bench_decode.json 2> gpurun_out/bench_decode.err; echo "decode rc=$?"
cut -c1-300 gpurun_out/${T}_bench_decode.json
timeout 300 python tools/optim_bench.py > gpurun_out/${T}_optim_bench.json 2> gpurun_out/optim_bench.err; echo "optim rc=$?"; cat gpurun_out/${T}_optim_bench.json
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-kernel-timing --no-optimizer --no-graph"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 || exit 1
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/${T}_launches.csv $CMD > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 429 -c 12 -f -o /tmp/prof_gemm $CMD > gpurun_out/ncu_gemm.log 2>&1; echo "ncu gemm rc=$?"
ncu -i /tmp/prof_gemm.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_gemm_raw.csv 2>/dev/null
timeout 1200 ncu --set full --clock-control none -k regex:"lstm_seq|attn_tc|attn_k4|attn_q1|reasoner" -s 54 -c 18 -f -o /tmp/prof_tail $CMD > gpurun_out/ncu_tail.log 2>&1; echo "ncu tail rc=$?"
ncu -i /tmp/prof_tail.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_tail_raw.csv 2>/dev/null
timeout 600 ncu --set full --clock-control none -k regex:"optim_" -s 9 -c 3 -f -o /tmp/prof_optim python tools/optim_bench.py > gpurun_out/ncu_optim.log 2>&1; echo "ncu optim rc=$?"
ncu -i /tmp/prof_optim.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_optim_raw.csv 2>/dev/null
ls -la gpurun_out | tail -20
