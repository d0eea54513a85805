#!/bin/bash
# pointer / gumbel kernels: tests, then ncu per-launch durations of the micro-benchmark (cold caches: read as relative)
python -m pytest tests/test_ops_gpu.py -m gpu -q -x --tb=short -k "pointer or gumbel" 2>&1 | tail -5
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"pointer_|gumbel_" -c 400 --csv --log-file gpurun_out/ptr_launches.csv python tools/pointer_bench.py > /dev/null 2>&1
python - <<'PY'
import csv, collections
rows = list(csv.reader(open('gpurun_out/ptr_launches.csv')))
i0 = next(i for i, r in enumerate(rows) if r and r[0] == 'ID')
hdr = rows[i0]; k = hdr.index('Kernel Name'); v = hdr.index('Metric Value')
d = collections.defaultdict(list)
for r in rows[i0 + 1:]:
    if len(r) > v: d[r[k][:60]].append(float(r[v].replace(',', '')) / 1e3)
for n, t in d.items(): print(f"{sum(t)/len(t):8.1f} us  n={len(t):3d}  {n}")
PY
