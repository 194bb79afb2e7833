#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` dump: samples / instructions by SASS region and stall reason.

usage: ncu -i prof.ncu-rep --page source --csv > src.csv ; python tools/ncu_source_summary.py src.csv [nregions]
Regions are split at backward-branch targets (loops): prints the hottest contiguous address windows.
"""
import csv
import sys
from collections import defaultdict

path = sys.argv[1]
win = int(sys.argv[2]) if len(sys.argv) > 2 else 64
rows = list(csv.reader(open(path)))
hdr = rows[1]
col = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot_s = sum(int(r[col["# Samples"]] or 0) for r in data)
tot_i = sum(int(r[col["Instructions Executed"]] or 0) for r in data)
print(f"total samples {tot_s}, warp instructions {tot_i}")
agg = defaultdict(int)
for r in data:
    for s in stalls:
        agg[s] += int(r[col[s]] or 0)
print("stall reasons:", ", ".join(f"{k[6:]} {v / max(tot_s, 1):.1%}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:9]))
# windows of `win` instructions
print(f"hottest windows of {win} SASS instructions:")
wins = []
for b in range(0, len(data), win):
    blk = data[b:b + win]
    s = sum(int(r[col["# Samples"]] or 0) for r in blk)
    i = sum(int(r[col["Instructions Executed"]] or 0) for r in blk)
    wins.append((s, i, b, blk))
for s, i, b, blk in sorted(wins, key=lambda w: -w[0])[:14]:
    ops = defaultdict(int)
    for r in blk:
        ops[r[col["Source"]].split()[0].lstrip("@!P0123456789T").strip() if r[col["Source"]] else "?"] += int(r[col["Instructions Executed"]] or 0)
    top = ", ".join(f"{k}:{v / max(i, 1):.0%}" for k, v in sorted(ops.items(), key=lambda kv: -kv[1])[:5])
    st = defaultdict(int)
    for r in blk:
        for q in stalls:
            st[q] += int(r[col[q]] or 0)
    tst = ", ".join(f"{k[6:]}:{v / max(s, 1):.0%}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:4])
    print(f"  @{blk[0][col['Address']]} samples {s / max(tot_s, 1):6.1%} insts {i / max(tot_i, 1):6.1%}  [{top}]  stalls[{tst}]")
