#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ops_gpu.py -q -m gpu -k "gemm" --tb=short --maxfail=10 > gpurun_out/gemm_tests.log 2>&1; echo "gemm tests rc=$?"
tail -5 gpurun_out/gemm_tests.log
timeout 300 python tools/gemm_bench.py 20 > gpurun_out/gemm_bench.txt 2>&1; echo "gemm bench rc=$?"
cat gpurun_out/gemm_bench.txt
timeout 300 python tools/gemm_diag.py > gpurun_out/gemm_diag.txt 2>&1; cat gpurun_out/gemm_diag.txt
