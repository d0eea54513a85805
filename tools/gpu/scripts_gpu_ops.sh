#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu --tb=short --maxfail=10 -x > gpurun_out/tests.log 2>&1; echo "tests rc=$?"
tail -4 gpurun_out/tests.log
timeout 300 python tools/step_kernels.py > gpurun_out/step_kernels.txt 2>&1; head -70 gpurun_out/step_kernels.txt
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-kernel-timing 2>/dev/null | cut -c1-300
