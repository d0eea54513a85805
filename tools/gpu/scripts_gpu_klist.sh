#!/bin/bash
# ncu durations of the kernels matching $1 (regex) in one bench process
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-kernel-timing --no-optimizer --no-graph"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"$1" -c 400 --csv --log-file gpurun_out/klist.csv $CMD > gpurun_out/klist.log 2>&1; echo "rc=$?"
python - <<'PY'
import csv, collections
rows = list(csv.reader(l for l in open("gpurun_out/klist.csv") if l.startswith('"')))
h = rows[0]; ki = h.index("Kernel Name"); vi = h.index("Metric Value")
d = collections.defaultdict(list)
for r in rows[1:]:
    try: d[r[ki][:40]].append(float(r[vi].replace(",", "")) / 1e3)
    except ValueError: pass
for k, v in d.items(): print(f"{k:42s} n={len(v):3d} avg={sum(v)/len(v):8.1f} us  min={min(v):8.1f}")
PY
