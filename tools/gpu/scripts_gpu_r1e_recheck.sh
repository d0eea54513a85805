#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu --tb=short -x > gpurun_out/tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/tests.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r1_e_bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cut -c1-200 gpurun_out/r1_e_bench.json; python -c "
import json; d=json.loads(open('gpurun_out/r1_e_bench.json').read().strip().splitlines()[-1]); print('e2e', d['e2e']['value'], d['e2e']['ms_per_step'], 'iter', d['iteration_with_optimizer']['value'], 'roof', d['roofline']['frac'], 'cpu', d['cpu_baseline'])"
timeout 600 python bench.py --task decode --steps 5 --warmup 3 > gpurun_out/r1_e_bench_decode.json 2> gpurun_out/bench_decode.err; echo "decode rc=$?"
cut -c1-200 gpurun_out/r1_e_bench_decode.json
./tools/gpu/scripts_gpu_klist.sh "video_ln_pack|pointer_fwd|pointer_bwd_kernel"
