#!/bin/bash
# op tests selected by $1 (pytest -k), then a short bench line and the per-call CUDA-event table of the named kernels
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ops_gpu.py -q -m gpu --tb=short -k "$1" > gpurun_out/q_tests.log 2>&1; echo "tests rc=$?"
tail -6 gpurun_out/q_tests.log
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/q_bench.json").read().strip().splitlines()[-1])
print("value", d["value"], "ms", d["ms_per_step"], "e2e", d.get("e2e", {}).get("value"), "iter", d.get("iteration_with_optimizer", {}).get("value"))
kt = d.get("kernel_timing") or d.get("kernels") or {}
print(json.dumps(kt)[:1500])
PY
