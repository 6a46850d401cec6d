#!/usr/bin/env python
"""Top stalled SASS instructions of an ncu report (usage: ncu_topstall.py report.ncu-rep [N])."""
import csv,sys,collections,subprocess
rep=sys.argv[1]
out=subprocess.run(["ncu","-i",rep,"--page","source","--csv","--print-source","sass"],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
hi=[i for i,r in enumerate(rows) if r and r[0]=="Address"][0]
hdr=rows[hi]; data=rows[hi+1:]
ia=hdr.index("Source"); isamp=hdr.index("# Samples"); iex=hdr.index("Instructions Executed")
stall=[i for i,h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot=collections.Counter()
for r in data:
    if 'EXIT' in r[ia]: continue
    for i in stall: tot[hdr[i][6:]]+=int(r[i] or 0)
print("stalls:",tot.most_common(8))
print("instr executed:",sum(int(r[iex] or 0) for r in data))
top=sorted(data,key=lambda r:-int(r[isamp] or 0))[:int(sys.argv[2]) if len(sys.argv)>2 else 14]
for r in top:
    st={hdr[i][6:]:int(r[i] or 0) for i in stall if int(r[i] or 0)>0}
    st=dict(sorted(st.items(),key=lambda kv:-kv[1])[:2])
    print(r[0][-5:], r[isamp].rjust(6), r[ia][:64].ljust(64), st)
