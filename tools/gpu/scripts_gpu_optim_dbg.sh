#!/bin/bash
mkdir -p gpurun_out
timeout 600 compute-sanitizer --tool memcheck python -m pytest tests/test_optim_gpu.py -q -m gpu --tb=short -x -k fused > gpurun_out/optim_san.log 2>&1; echo "rc=$?"
grep -n 'Invalid\|at \|by thread\|Address\|optim_' gpurun_out/optim_san.log | head -30
