#!/bin/bash
# Full GPU verification: all -m gpu tests, smoke, bench (train + decode), ncu launch list.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --tb=short --maxfail=20 > gpurun_out/tests.log 2>&1; echo "tests rc=$?"
tail -8 gpurun_out/tests.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cut -c1-3000 gpurun_out/bench.json
timeout 600 python bench.py --task decode --steps 5 --warmup 3 > gpurun_out/bench_decode.json 2> gpurun_out/bench_decode.err; echo "decode rc=$?"
cut -c1-1500 gpurun_out/bench_decode.json
./tools/gpu/scripts_gpu_ncu_list.sh
