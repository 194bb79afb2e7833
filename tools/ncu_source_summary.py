#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` dump: opcode mix, stall reasons and hottest SASS windows of one kernel.

usage: ncu -i prof.ncu-rep --page source --csv > src.csv
       python tools/ncu_source_summary.py src.csv [--kernel REGEX] [--window N] [--top K]
A dump holds one or two sections per kernel (SASS view, then source view); the first section of the first matching
kernel is used.
"""
import argparse
import csv
import re
from collections import Counter, defaultdict

ap = argparse.ArgumentParser()
ap.add_argument("csv")
ap.add_argument("--kernel", default=".")
ap.add_argument("--window", type=int, default=96)
ap.add_argument("--top", type=int, default=10)
a = ap.parse_args()

rows = list(csv.reader(open(a.csv)))
sections, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "data": []}
        sections.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = r
    elif cur is not None and len(r) == len(cur["hdr"]):
        cur["data"].append(r)
if not sections:  # single-kernel dump without the name line
    sections = [{"name": "?", "hdr": rows[1], "data": rows[2:]}]
sec = next(s for s in sections if re.search(a.kernel, s["name"]))
hdr, data = sec["hdr"], sec["data"]
col = {h: i for i, h in enumerate(hdr)}
num = lambda r, k: int(r[col[k]] or 0)  # noqa: E731
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot_s = sum(num(r, "# Samples") for r in data) or 1
tot_i = sum(num(r, "Instructions Executed") for r in data) or 1
tot_t = sum(num(r, "Thread Instructions Executed") for r in data)
print(f"kernel: {sec['name'][:90]}")
print(f"samples {tot_s}, warp instructions {tot_i}, avg active threads {tot_t / tot_i:.1f}")


def opcode(r):
    src = r[col["Source"]].split()
    if not src:
        return "?"
    op = src[1] if src[0].startswith("@") and len(src) > 1 else src[0]
    return op.split(".")[0]


mix, mix_s = Counter(), Counter()
for r in data:
    mix[opcode(r)] += num(r, "Instructions Executed")
    mix_s[opcode(r)] += num(r, "# Samples")
print("opcode mix (share of warp instructions / of samples):")
print("  " + ", ".join(f"{k} {v / tot_i:.1%}/{mix_s[k] / tot_s:.1%}" for k, v in mix.most_common(18)))
agg = defaultdict(int)
for r in data:
    for s in stalls:
        agg[s] += num(r, s)
print("stall reasons: " + ", ".join(f"{k[6:]} {v / tot_s:.1%}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
wins = []
for b in range(0, len(data), a.window):
    blk = data[b:b + a.window]
    wins.append((sum(num(r, "# Samples") for r in blk), sum(num(r, "Instructions Executed") for r in blk), blk))
print(f"hottest windows of {a.window} SASS instructions:")
for s, i, blk in sorted(wins, key=lambda w: -w[0])[:a.top]:
    ops = Counter()
    for r in blk:
        ops[opcode(r)] += num(r, "Instructions Executed")
    top = ", ".join(f"{k}:{v / max(i, 1):.0%}" for k, v in ops.most_common(5))
    st = defaultdict(int)
    for r in blk:
        for q in stalls:
            st[q] += num(r, q)
    tst = ", ".join(f"{k[6:]}:{v / max(s, 1):.0%}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
    print(f"  @{blk[0][col['Address']][-6:]} samples {s / tot_s:6.1%} insts {i / tot_i:6.1%}  [{top}]  stalls[{tst}]")
