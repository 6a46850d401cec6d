#!/usr/bin/env python
"""Warp-sampling buckets of a tcgen05 kernel's SASS (usage: ncu_buckets.py source.csv): samples, instruction count and
top stall reasons between the landmark instructions (setmaxnreg, LDTM/STTM, UTMALDG, first UTCHMMA)."""
import csv
import sys
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]
data = rows[hi + 1:]
ia = hdr.index("Source")
isamp = hdr.index("# Samples")
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
seg, segname, ninst, st, out = 0, "start", 0, {}, []
for r in data:
    s = r[ia]
    n = int(r[isamp] or 0)
    key = None
    if "USETMAXREG" in s:
        key = s.split()[0]
    if "LDTM" in s or "STTM" in s:
        key = s[:30].strip()
    if "UTMALDG" in s:
        key = "UTMALDG"
    if "UTCHMMA" in s and out and "UTCHMMA" not in segname:
        key = "UTCHMMA"
    if key:
        out.append((segname, seg, ninst, st))
        seg, ninst, st, segname = 0, 0, {}, r[0][-5:] + " " + key
    seg += n
    ninst += 1
    for i in stall:
        v = int(r[i] or 0)
        if v:
            st[hdr[i][6:]] = st.get(hdr[i][6:], 0) + v
out.append((segname, seg, ninst, st))
tot = sum(b for a, b, c, d in out)
for a, b, c, d in out:
    if b > tot * 0.004:
        print(str(b).rjust(6), str(c).rjust(5), a.ljust(40), dict(sorted(d.items(), key=lambda kv: -kv[1])[:4]))
print(tot)
