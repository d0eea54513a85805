#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_metrics.py tests/test_optim_gpu.py -q -m gpu --tb=short > gpurun_out/metrics_tests.log 2>&1; echo "tests rc=$?"
tail -12 gpurun_out/metrics_tests.log
timeout 300 python tools/metrics_bench.py > gpurun_out/metrics_bench.json 2> gpurun_out/metrics_bench.err; echo "bench rc=$?"
tail -3 gpurun_out/metrics_bench.err; cat gpurun_out/metrics_bench.json
