#!/usr/bin/env python
"""The bulk-copy (TMA) and mbarrier instructions of the streamed kernels, from the built library's SASS.

usage: python tools/sass_tma.py > profiles/rNN_sass_stream_tma.txt
For every kernel instantiation that streams scene tiles through shared memory (cp.async.bulk + mbarrier, rtc_tma.cuh)
prints the UBLKCP (bulk copy global -> shared), SYNCS (mbarrier arrive / expect_tx / try_wait) and fence lines with a
few instructions of context -- the evidence DESIGN.md section 4 cites for "tiles are moved by the TMA unit"."""
import os
import re
import subprocess

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(root, "raytracingc_b200", "librtc_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
want = {"render_paths_kernelILb1ELi3ELb0E": "render_paths_kernel<STREAM=true, MODE=3, GROUPS=false>  (serial form, streamed)",
        "render_pool_kernelILb1ELi3ELi16E": "render_pool_kernel<STREAM=true, MODE=3, WARPS=16>  (pool form, streamed)",
        "render_skyblock_kernel": "render_skyblock_kernel  (primary rays, block bounds; streams the edge-function table when it exceeds a tile)",
        "render_blocktask_kernel": "render_blocktask_kernel"}
cur, lines = None, {}
for l in txt.splitlines():
    if "Function :" in l:
        cur = next((k for k in want if k in l), None)
        continue
    if cur and re.match(r"^\s+/\*[0-9a-f]{4,5}\*/", l):
        lines.setdefault(cur, []).append(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l.rstrip()))
print("sm_100a SASS (cuobjdump -sass raytracingc_b200/librtc_b200.so): bulk asynchronous copies and mbarrier traffic of the")
print("streamed kernels.  UBLKCP = cp.async.bulk.shared.global (the TMA unit moves a tile, no registers staged);")
print("SYNCS.ARRIVE.TRANS64 / SYNCS.PHASECHK.TRANS64.TRYWAIT = mbarrier.arrive.expect_tx / mbarrier.try_wait.parity.\n")
for k, title in want.items():
    ls = lines.get(k, [])
    hits = [i for i, l in enumerate(ls) if re.search(r"\b(UBLKCP|SYNCS|FENCE\.VIEW\.ASYNC|UTMA)", l)]
    ops = {}
    for i in hits:
        op = re.sub(r"/\*[0-9a-f]+\*/", "", ls[i]).split()
        op = op[1] if op[0].startswith("@") else op[0]
        ops[op] = ops.get(op, 0) + 1
    print(f"== {title}: {len(ls)} instructions; " + ", ".join(f"{o} x{n}" for o, n in sorted(ops.items())))
    shown = set()
    for i in hits:
        for j in range(max(0, i - 1), min(len(ls), i + 2)):
            if j not in shown:
                shown.add(j)
                print(ls[j])
        print("    ...")
    print()
