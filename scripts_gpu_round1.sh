#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_decode_gpu.py -q -m gpu --tb=short --maxfail=20 > gpurun_out/tests.log 2>&1; echo "tests rc=$?"
tail -30 gpurun_out/tests.log
timeout 600 python bench.py --task decode --steps 5 --warmup 3 > gpurun_out/bench_decode.json 2> gpurun_out/bench_decode.err; echo "decode rc=$?"
tail -3 gpurun_out/bench_decode.err | cut -c1-300; cat gpurun_out/bench_decode.json
