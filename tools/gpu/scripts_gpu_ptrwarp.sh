#!/bin/bash
# final check of the build (default path), then the warp-per-row pointer forward behind its switch: tests + timing
mkdir -p gpurun_out
timeout 600 python -m pytest tests -q -m gpu --tb=short > gpurun_out/tests.log 2>&1; echo "tests(default) rc=$?"; tail -1 gpurun_out/tests.log
SVPC_PTR_WARP=2 timeout 300 python -m pytest tests/test_ops_gpu.py tests/test_model_gpu.py -q -m gpu --tb=short -k "pointer or golden or oracle or bf16" > gpurun_out/tests_ptrwarp.log 2>&1; echo "tests(ptr warp) rc=$?"; tail -3 gpurun_out/tests_ptrwarp.log
SVPC_PTR_WARP=1 timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-kernel-timing --no-optimizer --no-e2e > gpurun_out/q_ptrwarp.json 2> gpurun_out/q_ptrwarp.err; echo "bench(ptr warp) rc=$?"; cut -c1-150 gpurun_out/q_ptrwarp.json
timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-kernel-timing --no-optimizer --no-e2e > gpurun_out/q_default.json 2> gpurun_out/q_default.err; echo "bench(default) rc=$?"; cut -c1-150 gpurun_out/q_default.json
