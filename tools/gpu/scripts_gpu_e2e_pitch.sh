#!/bin/bash
# e2e against the way the feature blocks go up: per-row 1-D copies (default), one pitched 2-D copy, whole blocks 1-D
mkdir -p gpurun_out
run() {
timeout 200 python bench.py --steps 30 --warmup 3 --no-cpu-baseline --no-kernel-timing --no-optimizer > gpurun_out/e2e_$1.json 2> gpurun_out/e2e_$1.err; echo "$1 rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/e2e_$1.json').read().strip().splitlines()[-1]); print('value', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e']['h2d_bytes_per_step'])"
}
SVPC_H2D_ROWS=1 run rows
run pitched2d
timeout 300 python -m pytest tests/test_model_gpu.py -q -m gpu -k "prefetch or arena" --tb=short 2>&1 | tail -2
