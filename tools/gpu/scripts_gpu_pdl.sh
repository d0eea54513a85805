#!/bin/bash
# PDL on/off comparison: GPU tests with PDL on, then the training bench with SVPC_PDL=1 and =0.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --tb=short --maxfail=20 > gpurun_out/tests.log 2>&1; echo "tests rc=$?"
tail -8 gpurun_out/tests.log
for p in 1 0; do
SVPC_PDL=$p timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_pdl$p.json 2> gpurun_out/bench_pdl$p.err; echo "bench pdl=$p rc=$?"
tail -2 gpurun_out/bench_pdl$p.err; cut -c1-900 gpurun_out/bench_pdl$p.json
done
SVPC_PDL=1 timeout 600 python bench.py --task decode --steps 5 --warmup 3 > gpurun_out/bench_decode.json 2> gpurun_out/bench_decode.err; echo "decode rc=$?"
cut -c1-800 gpurun_out/bench_decode.json
