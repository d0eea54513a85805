#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_dropout_gpu.py tests/test_model_gpu.py -q -m gpu --tb=short --maxfail=20 > gpurun_out/tests.log 2>&1; echo "tests rc=$?"
tail -15 gpurun_out/tests.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cut -c1-1200 gpurun_out/bench.json
./tools/gpu/scripts_gpu_ncu_list.sh
