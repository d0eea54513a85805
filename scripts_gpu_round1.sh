#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --tb=short --maxfail=20 > gpurun_out/tests.log 2>&1; echo "tests rc=$?"
tail -40 gpurun_out/tests.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -3 gpurun_out/bench.err | cut -c1-300; cat gpurun_out/bench.json
timeout 600 python bench.py --steps 10 --warmup 3 --eval-mode --no-cpu-baseline > gpurun_out/bench_eval.json 2> gpurun_out/bench_eval.err; echo "bench-eval rc=$?"
cut -c1-400 gpurun_out/bench_eval.json
