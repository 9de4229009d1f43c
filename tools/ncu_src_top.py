#!/usr/bin/env python
"""Top stalled SASS instructions of one kernel from `ncu --page source --csv` output.
usage: ncu_src_top.py file.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = rows[1]
col = {h: i for i, h in enumerate(hdr)}
body = rows[2:]
for _k in range(len(body)):
    if body[_k] and body[_k][0] == "Kernel Name":
        body = body[:_k]
        break
body = [r for r in body if len(r) == len(hdr)]
tot = sum(int(r[col["# Samples"]] or 0) for r in body)
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
print("total samples", tot, "instructions", len(body))
agg = {h: sum(int(r[col[h]] or 0) for r in body) for h in stall_cols}
print("by reason:", {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
order = sorted(range(len(body)), key=lambda i: -int(body[i][col["# Samples"]] or 0))[:n]
for i in sorted(order):
    r = body[i]
    s = int(r[col["# Samples"]] or 0)
    top = sorted(((int(r[col[h]] or 0), h) for h in stall_cols), reverse=True)[:2]
    print(f"{i:5d} {s:7d} {100.0*s/max(tot,1):5.1f}%  {r[col['Source']].strip():70s} exec={r[col['Instructions Executed']]} {top}")
