#!/usr/bin/env python
"""Extract the branch-free first phase of the path kernel's sweep (the hot loop) from the built library's SASS.

usage: python tools/sass_sweep.py [MODE] [serial|pool] > profiles/rNN_sass_sweep.txt
       (MODE 3 = packed quick records, 1 = scalar quick records, 0 = reference arithmetic; pool = render_pool_kernel, two rays per lane)
The loop is found as the call-free, convergence-barrier-free backward branch whose body has the most FP32 arithmetic."""
import os
import re
import subprocess
import sys
from collections import Counter

mode = sys.argv[1] if len(sys.argv) > 1 else "3"
form = sys.argv[2] if len(sys.argv) > 2 else "serial"
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(root, "raytracingc_b200", "librtc_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
name = f"_ZN3rtc19render_paths_kernelILb0ELi{mode}E" if form == "serial" else f"_ZN3rtc18render_pool_kernelILb0ELi{mode}ELi12E"
lines, on = [], False
for l in txt.splitlines():
    if "Function :" in l:
        on = name in l
        continue
    if on and re.match(r"^\s+/\*[0-9a-f]{4,5}\*/", l):
        lines.append(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l.rstrip()))
addr = lambda l: int(re.search(r"/\*([0-9a-f]+)\*/", l).group(1), 16)  # noqa: E731
best = None
for i, l in enumerate(lines):
    m = re.search(r"BRA\S*\s+(?:!?U?P\d,\s*)?(0x[0-9a-f]+)", l)
    if m and int(m.group(1), 16) < addr(l):
        n = (addr(l) - int(m.group(1), 16)) // 16 + 1
        body = lines[i - n + 1:i + 1]
        fp = sum(bool(re.search(r"\b(FMUL2?|FADD2?|FFMA2?)\b", b)) for b in body)
        if n < 400 and not any(re.search(r"\b(CALL|BSSY|LDG)\b", b) for b in body) and (best is None or fp > best[0]):
            best = (fp, body)
body = best[1]
ops = Counter()
for l in body:
    t = re.sub(r"/\*[0-9a-f]+\*/", "", l).split()
    ops[(t[1] if t[0].startswith("@") else t[0]).split(".")[0]] += 1
per = max(1, ops["LDS"] // (3 if mode == "0" else 4))  # 4 LDS.128 of quick record per triangle (3 of geometry in mode 0)
if form == "pool":
    print(f"render_pool_kernel<false, {mode}, 12> (sm_100a, cuobjdump -sass of raytracingc_b200/librtc_b200.so): the branch-free")
    print(f"first phase of sweep_two (rtc_pool.cuh), {per} triangles x TWO rays per iteration (every LDS.128 feeds both rays): "
          f"{len(body)} instructions = {len(body) / (2 * per):.2f} per (ray, triangle) pair.")
else:
    print(f"render_paths_kernel<false, {mode}> (sm_100a, cuobjdump -sass of raytracingc_b200/librtc_b200.so): the branch-free")
    print(f"first phase of sweep_chunked (rtc_paths.cuh), {per} triangles per iteration: {len(body)} instructions = "
          f"{len(body) / per:.2f} per (ray, triangle) pair.")
if mode in ("3", "4"):
    print("Two triangles per packed instruction (FFMA2 / FMUL2 / FADD2 on pair-interleaved records, the ray's scalars as the")
    print("broadcast operand Rx.F32); per-pair data by warp-uniform LDS.128; survivors set one bit of the lane's mask through")
    print("a chained predicate (FSETP ... .AND/.OR P) and a predicated add; phase 2 (tri_test_exact) follows the loop.")
else:
    print("Per-triangle data by warp-uniform LDS.128; survivors set one bit of the lane's mask through a chained predicate")
    print("(FSETP ... .AND/.OR P) and a predicated add; phase 2 (tri_test_exact on the survivors) follows the loop.")
print("opcode counts: " + ", ".join(f"{k} {v}" for k, v in ops.most_common()))
print()
print("\n".join(body))
