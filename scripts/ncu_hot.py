#!/usr/bin/env python
"""Top stall sites from `ncu -i rep --page source --csv` (SASS view): address, samples, dominant
stall reasons, instruction.  Usage: ncu_hot.py source.csv [top_n]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ix = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
data = []
tot = 0
for r in rows[hi + 1:]:
    if len(r) < len(hdr):
        continue
    try:
        n = int(r[ix["# Samples"]] or 0)
    except ValueError:
        continue
    tot += n
    st = sorted(((int(r[ix[c]] or 0), c[6:]) for c in stall_cols), reverse=True)[:3]
    data.append((n, r[ix["Address"]], r[ix["Source"]], st, len(data)))
print("total samples", tot)
agg = {}
for r in rows[hi + 1:]:
    if len(r) < len(hdr):
        continue
    for c in stall_cols:
        try:
            agg[c] = agg.get(c, 0) + int(r[ix[c]] or 0)
        except ValueError:
            pass
print("stall totals:", sorted(((v, k[6:]) for k, v in agg.items() if v), reverse=True)[:8])
for n, a, s, st, pos in sorted(data, reverse=True)[:top]:
    print(f"{n:7d} {100.0 * n / max(tot, 1):5.1f}%  #{pos:5d} {s[:70]:70s} {[(c, k) for c, k in st if c]}")
