#!/bin/bash
# first GPU pass: operator parity, whole-model parity, then the tcgen05 route (separately, so a hang there cannot hide the rest)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 600 python -m pytest tests/test_ops_gpu.py -q -m gpu -k "not gemm_tc" --tb=short --maxfail=30 > gpurun_out/ops.log 2>&1; echo "ops rc=$?"
tail -25 gpurun_out/ops.log
timeout 600 python -m pytest tests/test_model_gpu.py -q -m gpu -k "fp32 or no_cpu" --tb=short --maxfail=30 > gpurun_out/model_fp32.log 2>&1; echo "model fp32 rc=$?"
tail -25 gpurun_out/model_fp32.log
timeout 300 python -m pytest tests/test_ops_gpu.py -q -m gpu -k "gemm_tc" --tb=line --maxfail=100 > gpurun_out/tc.log 2>&1; echo "tc rc=$?"
tail -40 gpurun_out/tc.log
timeout 600 python -m pytest tests/test_model_gpu.py -q -m gpu -k "bf16" --tb=short > gpurun_out/model_bf16.log 2>&1; echo "model bf16 rc=$?"
tail -15 gpurun_out/model_bf16.log
