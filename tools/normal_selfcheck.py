#!/usr/bin/env python
"""Exhaustive check of the normal deviate's fast path: all 2^32 generator states (every input the reference can reach),
device fast path against the device's exact sequence, bit for bit.  usage: python tools/normal_selfcheck.py > profiles/rNN_normal_selfcheck.txt"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from raytracingc_b200 import api  # noqa: E402

tot = {"different": 0, "declined": 0, "states": 0}
t = time.perf_counter()
for part in range(16):  # 16 launches of 2^28 states
    r = api.probe_normal_selfcheck(part << 28, 1 << 28)
    for k in tot:
        tot[k] += r[k]
    print(f"states [{part << 28:#010x}, +2^28): different {r['different']}, declined {r['declined']}", flush=True)
print(f"ALL {tot['states']} states (2^32 = {1 << 32}): fast path differs from the exact sequence on {tot['different']} states; "
      f"declined (fell back to the exact sequence) on {tot['declined']} = {tot['declined'] / tot['states']:.2e} of them; "
      f"{time.perf_counter() - t:.1f} s")
sys.exit(0 if tot["different"] == 0 and tot["states"] == 1 << 32 else 1)
