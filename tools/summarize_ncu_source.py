#!/usr/bin/env python
"""Top stalled SASS lines of one kernel from `ncu -i X.ncu-rep --page source --csv` (tools/gpu_ncu_src.sh).
Usage: python tools/summarize_ncu_source.py X_source.csv [N lines]"""
import csv, sys, re
path = sys.argv[1]
rows = list(csv.reader(open(path)))
hdr = rows[1]; body = rows[2:]
ci = {h: i for i, h in enumerate(hdr)}
tot = sum(int(r[ci["# Samples"]]) for r in body)
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
print(rows[0][1][:90], "total samples", tot)
print({h[6:]: sum(int(r[ci[h]]) for r in body) for h in stall_cols})
top = sorted(body, key=lambda r: -int(r[ci["# Samples"]]))[:int(sys.argv[2]) if len(sys.argv) > 2 else 25]
for r in top:
    st = sorted(((int(r[ci[h]]), h[6:]) for h in stall_cols), reverse=True)[:2]
    idx = body.index(r)
    print(f"{idx:5d} {int(r[ci['# Samples']]):6d} {100*int(r[ci['# Samples']])/tot:5.1f}%  {r[ci['Source']].strip()[:80]:80s} {st}")
