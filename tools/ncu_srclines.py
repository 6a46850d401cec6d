#!/usr/bin/env python
"""Per-source-line totals of an ncu report (usage: ncu_srclines.py report.ncu-rep [N]): warp samples and executed
instructions summed over the SASS of every CUDA-C line (needs -lineinfo and --import-source on)."""
import collections
import csv
import subprocess
import sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
tot = collections.defaultdict(lambda: [0, 0, ""])
f, cur = "", None
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        f = r[1].split("/")[-1]
        continue
    if r[0] in ("Function Name", "Line No"):
        continue
    if r[0]:
        cur = (f, int(r[0]))
        tot[cur][2] = r[1].strip()
    elif cur and len(r) > 7:
        num = lambda v: int(v) if v.strip().isdigit() else 0
        tot[cur][0] += num(r[6])
        tot[cur][1] += num(r[7])
n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
S = sum(v[0] for v in tot.values()); I = sum(v[1] for v in tot.values())
print("samples", S, "instructions", I)
for k, v in sorted(tot.items(), key=lambda kv: -kv[1][1 if len(sys.argv) > 3 else 0])[:n]:
    print("%-14s %5d  smp %6d  inst %10d (%4.1f%%)  %s" % (k[0], k[1], v[0], v[1], 100.0 * v[1] / max(I, 1), v[2][:90]))
