#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ops_gpu.py -q -m gpu --tb=short -k "pointer or gumbel" > gpurun_out/q_tests.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/q_tests.log
./tools/gpu/scripts_gpu_klist.sh "pointer|gumbel"
