#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time share per kernel."""
import csv
import re
import sys
from collections import defaultdict


def main(path, out=None):
    rows = []
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    rd = csv.DictReader(lines)
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1e-3)
        name = re.sub(r"\(.*", "", r["Kernel Name"])
        rows.append((name, v * scale, r.get("Grid Size", ""), r.get("Block Size", "")))
    agg = defaultdict(lambda: [0, 0.0])
    for n, us, *_ in rows:
        agg[n][0] += 1
        agg[n][1] += us
    tot = sum(v[1] for v in agg.values())
    lines = ["| kernel | launches | total us | share | avg us |", "|---|---:|---:|---:|---:|"]
    for n, (c, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        lines.append("| `%s` | %d | %.1f | %.1f%% | %.1f |" % (n[:90], c, us, 100 * us / tot, us / c))
    lines.append("| total | %d | %.1f | 100%% | |" % (len(rows), tot))
    text = "\n".join(lines)
    if out:
        open(out, "a").write(text + "\n")
    print(text)


if __name__ == "__main__":
    main(*sys.argv[1:])
