#!/usr/bin/env python
"""Summarise `ncu --page source --csv` output: stall-reason totals and the hottest SASS instructions.
usage: ncu -i X.ncu-rep --page source --csv > src.csv ; python tools/ncu_stalls.py src.csv [topN] [kernel_index]"""
import csv, sys
path = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25; which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
rows = list(csv.reader(open(path)))
# split into kernels: a "Kernel Name" row followed by a header row
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
s = starts[which]; e = starts[which + 1] if which + 1 < len(starts) else len(rows)
print(rows[s][1][:120])
hdr = rows[s + 1]
data = [r for r in rows[s + 2:e] if len(r) == len(hdr)]
ia, isrc, isamp, iex = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
num = lambda v: int(v) if v.strip().isdigit() else 0
tot = sum(num(r[isamp]) for r in data)
print("total samples", tot, "instructions", len(data), "executed", sum(num(r[iex]) for r in data))
agg = {}
for r in data:
    for i in stall_cols:
        agg[hdr[i]] = agg.get(hdr[i], 0) + num(r[i])
print([(k, v, f"{100*v/max(tot,1):.1f}%") for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]])
for r in sorted(data, key=lambda r: -num(r[isamp]))[:topn]:
    st = {hdr[i][6:]: num(r[i]) for i in stall_cols if num(r[i]) > 0}
    print(r[ia][-5:], f"{num(r[isamp]):6d}", f"{num(r[iex]):7d}", r[isrc][:72].ljust(72), dict(sorted(st.items(), key=lambda kv: -kv[1])[:3]))
