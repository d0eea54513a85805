#!/bin/bash
# quick check of a kernel change: selected tests ($1 = -k expression for tests/test_ops_gpu.py), headline parity, step time
K=${1:-pointer}
python -m pytest tests/test_ops_gpu.py -m gpu -q -x --tb=short -k "$K" 2>&1 | tail -8
python -m pytest tests/test_headline_gpu.py tests/test_model_gpu.py -m gpu -q -x --tb=short 2>&1 | tail -5
Q="--task train --no-cpu-baseline --no-e2e --no-kernel-timing --no-parity --steps 30 --warmup 5"
python bench.py $Q 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('step', d['ms_per_step'], d['value'])"
shift
for v in "$@"; do echo "== $v"; env $v python bench.py $Q 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('step', d['ms_per_step'], d['value'])"; done
