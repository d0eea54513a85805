#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_optim_gpu.py -q -m gpu --tb=short > gpurun_out/optim_tests.log 2>&1; echo "optim tests rc=$?"
tail -25 gpurun_out/optim_tests.log
timeout 300 python tools/optim_bench.py > gpurun_out/optim_bench.json 2> gpurun_out/optim_bench.err; echo "optim bench rc=$?"
tail -3 gpurun_out/optim_bench.err; cat gpurun_out/optim_bench.json
