#!/usr/bin/env python
"""Top stall sites of an `ncu -i X.ncu-rep --page source --csv` export: python scripts/ncu_top_stalls.py src.csv [N]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
H = rows[hdr]
col = {h: i for i, h in enumerate(H)}
stalls = [h for h in H if h.startswith("stall_") and "Not Issued" not in h]
data = []
for r in rows[hdr + 1:]:
    if len(r) < len(H):
        continue
    try:
        n = int(r[col["# Samples"]] or 0)
    except ValueError:
        continue
    data.append((n, r))
total = sum(n for n, _ in data)
print("total samples", total)
for n, r in sorted(data, key=lambda x: -x[0])[:top]:
    why = sorted(((int(r[col[s]] or 0), s[6:]) for s in stalls), reverse=True)[:3]
    print(f"{n:7d} {100 * n / max(total, 1):5.1f}%  {r[col['Address']][-5:]}  {r[col['Source']][:70]:70s} {[(w, c) for c, w in why if c]}")
