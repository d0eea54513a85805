#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -q -m gpu --tb=short > gpurun_out/tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/tests.log
timeout 200 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 300 python bench.py --steps 100 --warmup 3 --no-cpu-baseline --no-kernel-timing > gpurun_out/r1_e_bench_steps100.json 2> gpurun_out/b100.err; echo "bench100 rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r1_e_bench_steps100.json').read().strip().splitlines()[-1]); print('value', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], 'iter', d['iteration_with_optimizer']['value'])"
