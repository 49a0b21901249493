#!/usr/bin/env python
"""Top stalled SASS instructions of an `ncu --page source --csv` dump (usage: ncu_top.py file.csv [N])."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hi = next(i for i, r in enumerate(rows) if "# Samples" in r)
hdr = rows[hi]
iS, isrc, ie = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
stalls = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data = []
for k, r in enumerate(rows[hi + 1:]):
    if len(r) <= iS or not r[iS].isdigit():
        continue
    top = sorted(((int(r[i] or 0), h) for i, h in stalls), reverse=True)[:2]
    data.append((int(r[iS]), k, r[isrc][:90], r[ie], top))
tot = sum(d[0] for d in data)
print("total samples", tot, "instructions", len(data))
for d in sorted(data, key=lambda x: -x[0])[:n]:
    print(f"{d[0]:6d} {100*d[0]/tot:5.1f}% line{d[1]:5d} exec={d[3]:>8} {d[2]:90s} {d[4]}")
