#!/usr/bin/env python
"""Summarise an `ncu --page raw --csv` export: one line per profiled launch with the metrics that matter for a roofline
reading.  usage: ncu_raw_summary.py raw.csv [name filter]"""
import csv
import sys

path = sys.argv[1]
flt = sys.argv[2] if len(sys.argv) > 2 else ""
rows = list(csv.reader(open(path)))
hdr = rows[0]
idx = {h: i for i, h in enumerate(hdr)}
want = [("gpu__time_duration.sum", "dur_us"), ("dram__bytes_read.sum", "rd_MB"), ("dram__bytes_write.sum", "wr_MB"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"), ("launch__registers_per_thread", "regs"),
        ("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "ld_sect"), ("lts__t_sector_hit_rate.pct", "l2hit%"),
        ("smsp__cycles_active.avg", "cyc")]
units = rows[1]
print("kernel grid block " + " ".join(n for _, n in want))
for r in rows[2:]:
    name = r[idx["Kernel Name"]]
    if flt and flt not in name:
        continue
    out = []
    for m, n in want:
        if m not in idx:
            out.append("-")
            continue
        v = r[idx[m]].replace(",", "")
        try:
            v = float(v)
        except ValueError:
            out.append(v)
            continue
        u = units[idx[m]]
        if n == "dur_us":
            v = v / 1000 if u in ("ns", "nsecond") else (v * 1000 if u in ("ms", "msecond") else v)
        if n in ("rd_MB", "wr_MB"):
            f = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1, "Gbyte": 1e3}.get(u, 1e-6)
            v *= f
        out.append(f"{v:.1f}")
    print(name[:34], r[idx["Grid Size"]].replace(" ", ""), r[idx["Block Size"]].replace(" ", ""), " ".join(out))
