#!/usr/bin/env python
"""SASS opcode evidence for the built library: per kernel family, how many tcgen05 / TMEM / TMA / mma.sync instructions
the sm_100a code really contains (cuobjdump -sass of svpc_b200/libsvpc_b200.so).

    python tools/sass_histogram.py > profiles/r2_sass_opcodes.txt

UTCHMMA = tcgen05.mma kind::f16, UTCQMMA/UTCOMMA = other tcgen05 kinds, LDTM / STTM = tcgen05.ld / st (TMEM),
UTMALDG / UTMASTG = cp.async.bulk.tensor load / store (TMA), UTCBAR = tcgen05.commit, SYNCS = mbarrier, HMMA = mma.sync.
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "svpc_b200", "libsvpc_b200.so")
OPS = ("UTCHMMA", "UTCQMMA", "UTCOMMA", "UTCIMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "UTCBAR", "UTCATOMSWS",
       "SYNCS", "HMMA", "FFMA", "LDGSTS", "REDG", "RED", "ATOMG", "ACQBULK", "UCGABAR_ARV", "CCTL")


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    per = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
            cur = re.sub(r"\(.*", "", name)
            cur = re.sub(r"^void ", "", cur)
            per.setdefault(cur, collections.Counter())
            per[cur]["_variants"] += 1
            continue
        if cur is None:
            continue
        m = re.search(r"^\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m:
            op = m.group(1)
            per[cur]["_insts"] += 1
            for o in OPS:
                if op == o or op.startswith(o + "."):
                    per[cur][o] += 1
    fam = collections.OrderedDict()
    for k, c in per.items():
        f = re.sub(r"<.*", "", k)
        fam.setdefault(f, collections.Counter()).update(c)
    tot = collections.Counter()
    for c in fam.values():
        tot.update(c)
    print(f"# cuobjdump -sass {os.path.relpath(LIB, ROOT)}: {len(per)} kernels in {len(fam)} families, {tot['_insts']} instructions")
    print("# whole library: " + ", ".join(f"{o} x{tot[o]}" for o in OPS if tot[o]))
    print(f"{'kernel family':44s} {'variants':>8s} {'insts':>8s}  opcodes of interest")
    for f, c in sorted(fam.items(), key=lambda kv: -kv[1]["_insts"]):
        ops = ", ".join(f"{o} {c[o]}" for o in OPS if c[o] and o != "FFMA")
        print(f"{f[:44]:44s} {c['_variants']:8d} {c['_insts']:8d}  {ops}")


if __name__ == "__main__":
    sys.exit(main())
