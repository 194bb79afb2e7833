#!/usr/bin/env python
"""Samples and instructions per SOURCE LINE from `ncu -i x.ncu-rep --page source --print-source cuda,sass --csv`.

usage: python tools/ncu_lines.py dump.csv [--top 40] [--file rtc_pool.cuh] [--bucket 1]
"""
import argparse
import csv
import sys
from collections import defaultdict

csv.field_size_limit(1 << 30)
ap = argparse.ArgumentParser()
ap.add_argument("csv")
ap.add_argument("--top", type=int, default=40)
ap.add_argument("--file", default=None)
a = ap.parse_args()

cur, hdr = None, None
samples = defaultdict(int)
insts = defaultdict(int)
text = {}
for r in csv.reader(open(a.csv)):
    if not r:
        continue
    if r[0] in ("File Name", "File Path"):
        cur, hdr = r[1].split("/")[-1], None
        continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Line No":
        hdr = {h: i for i, h in enumerate(r)}
        # "Source" appears twice (cuda line text, sass text): first index is the cuda text
        continue
    if hdr is None or cur is None:
        continue
    try:
        ln = int(r[0])
    except ValueError:
        continue  # (the SASS rows under a line: the line's own row already carries their sums)
    if r[1]:
        text[(cur, ln)] = r[1]
    try:
        s = int(r[hdr["# Samples"]] or 0)
        n = int(r[hdr["Instructions Executed"]] or 0)
    except (ValueError, IndexError, KeyError):
        continue
    samples[(cur, ln)] += s
    insts[(cur, ln)] += n
tot_s, tot_i = sum(samples.values()) or 1, sum(insts.values()) or 1
byfile = defaultdict(lambda: [0, 0])
for k in samples:
    byfile[k[0]][0] += samples[k]
    byfile[k[0]][1] += insts[k]
print("per file: samples %, instructions %")
for f, (s, n) in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
    print(f"  {f:32s} {100 * s / tot_s:5.1f}  {100 * n / tot_i:5.1f}")
keys = [k for k in samples if a.file is None or k[0] == a.file]
print(f"top lines ({a.file or 'all files'}):")
for k in sorted(keys, key=lambda k: -samples[k])[:a.top]:
    print(f"  {k[0]:18s}:{k[1]:4d}  samples {100 * samples[k] / tot_s:5.2f}%  insts {100 * insts[k] / tot_i:5.2f}%  | {text.get(k, '').strip()[:90]}")
