#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_optim_gpu.py tests/test_model_gpu.py -q -m gpu --tb=short > gpurun_out/mix_tests.log 2>&1; echo "tests rc=$?"
tail -6 gpurun_out/mix_tests.log
timeout 300 python tools/optim_bench.py > gpurun_out/optim_bench.json 2> gpurun_out/optim_bench.err; echo "optim bench rc=$?"
tail -3 gpurun_out/optim_bench.err; cat gpurun_out/optim_bench.json
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_mix.json 2> gpurun_out/bench_mix.err; echo "bench rc=$?"
cut -c1-1500 gpurun_out/bench_mix.json
