#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_model_gpu.py -q -m gpu -k "$1" --tb=short 2>&1 | tail -15
