#!/bin/bash
# ncu per-launch durations (cold caches: compare before/after, not absolute) of the kernels matching regex $1 in one eager training step
RX=$1
CMD="python bench.py --task train --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-kernel-timing --no-optimizer --no-graph --no-parity"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"$RX" -c 400 --csv --log-file gpurun_out/k_launches.csv $CMD > gpurun_out/k_ncu.log 2>&1 || tail -5 gpurun_out/k_ncu.log
python - <<'PY'
import csv, collections
rows = list(csv.reader(open('gpurun_out/k_launches.csv')))
i0 = next(i for i, r in enumerate(rows) if r and r[0] == 'ID')
hdr = rows[i0]; k = hdr.index('Kernel Name'); v = hdr.index('Metric Value'); g = hdr.index('Grid Size')
d = collections.defaultdict(list)
for r in rows[i0 + 1:]:
    if len(r) > v: d[(r[k][:70], r[g])].append(float(r[v].replace(',', '')) / 1e3)
for n, t in d.items(): print(f"{sum(t)/len(t):8.1f} us  n={len(t):3d}  {n}")
PY
