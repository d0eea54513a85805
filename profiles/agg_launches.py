#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list: per (kernel, grid) and per kernel name.
usage: agg_launches.py launches.csv [n_steps_in_process] [top]"""
import collections
import csv
import sys

path = sys.argv[1]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 4
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
lines = [l for l in open(path) if not l.startswith("==")]
agg = collections.defaultdict(lambda: [0, 0.0])
for row in csv.DictReader(lines):
    v = float(row["Metric Value"].replace(",", ""))
    u = row["Metric Unit"]
    v = v / 1000 if u == "ns" else (v * 1000 if u == "ms" else v)
    k = (row["Kernel Name"][:64], row["Grid Size"], row["Block Size"])
    agg[k][0] += 1
    agg[k][1] += v
tot = sum(v[1] for v in agg.values())
print(f"total {tot / n:.1f} us/step over {n} steps")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{v[1] / n:9.1f} us/step n={v[0] / n:6.1f} avg={v[1] / v[0]:8.1f}  {k}")
by = collections.defaultdict(float)
for k, v in agg.items():
    by[k[0].split("(")[0][:44]] += v[1] / n
print("--- by kernel")
for k, v in sorted(by.items(), key=lambda kv: -kv[1])[:30]:
    print(f"{v:9.1f} {v / tot * n:6.3f} {k}")
