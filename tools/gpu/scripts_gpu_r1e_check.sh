#!/bin/bash
# Quick check of the 8f-4 row on the GPU + the bench line of this build.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_features.py tests/test_checkpoint.py tests/test_optim_gpu.py -q -m gpu --tb=short > gpurun_out/tests_new.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/tests_new.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 200 python tools/feature_reader_bench.py > gpurun_out/r1_e_feature_reader_bench.json 2> gpurun_out/reader.err; echo "reader rc=$?"; cat gpurun_out/r1_e_feature_reader_bench.json
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r1_e_bench_pre.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cut -c1-1500 gpurun_out/r1_e_bench_pre.json
nproc; lscpu | grep -i "model name\|numa" | head -5
