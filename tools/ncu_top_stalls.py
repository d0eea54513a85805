#!/usr/bin/env python
"""Top stall locations of an `ncu --page source --csv` export: usage ncu_top_stalls.py file.csv [n]"""
import csv
import sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr, body = rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
tot = sum(int(r[ix['# Samples']]) for r in body)
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
print("total samples", tot)
for r in sorted(body, key=lambda r: -int(r[ix['# Samples']]))[:n]:
    s = {h: int(r[ix[h]]) for h in stalls if int(r[ix[h]]) > 0}
    main = sorted(s.items(), key=lambda kv: -kv[1])[:2]
    print(f"{int(r[ix['# Samples']]):6d} {100 * int(r[ix['# Samples']]) / max(tot, 1):5.1f}%  {r[ix['Source']].strip()[:72]:72s} {main}")
agg = {}
for r in body:
    for h in stalls:
        agg[h] = agg.get(h, 0) + int(r[ix[h]])
print(sorted(agg.items(), key=lambda kv: -kv[1])[:8])
