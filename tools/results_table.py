#!/usr/bin/env python
"""Markdown table of the bench lines committed under profiles/ (usage: results_table.py r1d)."""
import json
import os
import sys

tag = sys.argv[1] if len(sys.argv) > 1 else "r1d"
root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles")
names = [("c4", "C4 ultracomplex 3840×2160, 64 spp, 4 bounces"), ("c3", "C3 complex 1920×1080, 64 spp (triangles only)"),
         ("c3s", "C3 with the sphere"), ("c2", "C2 suze 1280×720, 16 spp, 3 bounces"),
         ("default", "default (14 triangles + sphere) 640×480, 16 spp"), ("c1", "C1 simplest 320×240, 1 spp, 1 bounce")]
print("| workload | frame ms (`value`) | Mrays/s | e2e ms | e2e Mrays/s | path-kernel ms | pairs/s (path kernel) | "
      "roofline frac (staged flops) | brute-force frac (51 flop/pair) |")
print("|---|---|---|---|---|---|---|---|---|")
for key, label in names:
    f = os.path.join(root, f"{tag}_bench_{key}.json")
    if not os.path.exists(f):
        continue
    d = json.loads(open(f).read())
    r = d["roofline"]
    print(f"| {label} | {d['ms_per_step']:.2f} | {d['value']:.0f} | {d['e2e']['ms_per_step']:.2f} | {d['e2e']['value']:.0f} | "
          f"{r['kernel_ms']:.2f} | {r['pairs_per_s']:.2e} | {r['frac']:.3f} | {r['brute_force_frac']:.3f} |")
f = os.path.join(root, f"{tag}_bench_reference_c4.json")
if os.path.exists(f):
    d = json.loads(open(f).read())
    print(f"| C4, reference CPU arm ({d['cpu_baseline']['cores']} host cores, {d['cpu_baseline']['sample'].split('(')[1].split(',')[0]}) "
          f"| — | {d['value']:.1f} | — | {d['value']:.1f} | — | — | — | — |")
f = os.path.join(root, f"{tag}_bench_c4.json")
d = json.loads(open(f).read())
if d.get("modes"):
    print("\nC4 modes: " + ", ".join(f"{k} {v['frame_ms']:.1f} ms" for k, v in d["modes"].items()) +
          f"; FP32 FMA peak measured in the run {d['roofline']['peak']:.1f} TFLOP/s; clocks {d['clocks']}")
