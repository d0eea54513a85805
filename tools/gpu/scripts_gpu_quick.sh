#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu --tb=short -x > gpurun_out/tests.log 2>&1; echo "all tests rc=$?"
tail -3 gpurun_out/tests.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
tail -2 gpurun_out/bench_quick.err | cut -c1-300; cut -c1-250 gpurun_out/bench_quick.json; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_quick.json'))
print(d["ms_per_step"], d["e2e"], d.get("iteration_with_optimizer"))
for k in d["top_kernels_ms_per_step"]: print(k)
PY
