#!/usr/bin/env python
"""Markdown table of the bench lines committed under profiles/ (usage: results_table.py r1d)."""
import json
import os
import sys

tag = sys.argv[1] if len(sys.argv) > 1 else "r2"
root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles")
names = [("c4", "C4 ultracomplex 3840×2160, 64 spp, 4 bounces"), ("c3", "C3 complex 1920×1080, 64 spp (triangles only)"),
         ("c3s", "C3 with the sphere"), ("c2", "C2 suze 1280×720, 16 spp, 3 bounces"),
         ("default", "default (14 triangles + sphere) 640×480, 16 spp"), ("c1", "C1 simplest 320×240, 1 spp, 1 bounce"),
         ("c4close", "C4 scene, close-up camera (-p 0 -2 -6 -t 0 -1 0), 3840×2160, 64 spp"),
         ("c5l", "C5 reduced: 1 M synthetic triangles, 960×540 lattice of the 4K frame, 16 spp")]
print("| workload | frame ms (`value`) | Mrays/s | e2e ms | e2e Mrays/s | path-kernel ms | bounce Mrays/s (path kernel) | pairs/s (path kernel) | "
      "roofline frac (staged flops) | brute-force frac (51 flop/pair) | frame = golden |")
print("|---|---|---|---|---|---|---|---|---|---|---|")
for key, label in names:
    f = os.path.join(root, f"{tag}_bench_{key}.json")
    if not os.path.exists(f):
        continue
    d = json.loads(open(f).read())
    r = d["roofline"]
    pk = d.get("path_kernel", {})
    print(f"| {label} | {d['ms_per_step']:.2f} | {d['value']:.1f} | {d['e2e']['ms_per_step']:.2f} | {d['e2e']['value']:.1f} | "
          f"{r['kernel_ms']:.2f} | {pk.get('bounce_Mrays_per_s', 0):.1f} | {r['pairs_per_s']:.2e} | {r['frac']:.3f} | {r['brute_force_frac']:.3f} | "
          f"{d.get('parity_fnv_ok')} |")
f = os.path.join(root, f"{tag}_bench_reference_c4.json")
if os.path.exists(f):
    d = json.loads(open(f).read())
    print(f"| C4, reference CPU arm ({d['cpu_baseline']['cores']} host cores, {d['cpu_baseline']['sample'].split('(')[1].split(',')[0]}) "
          f"| — | {d['value']:.1f} | — | {d['value']:.1f} | — | — | — | — | — | (is the golden) |")
f = os.path.join(root, f"{tag}_bench_c4.json")
d = json.loads(open(f).read())
if d.get("modes"):
    print("\nC4 modes: " + ", ".join(f"{k} {v['frame_ms']:.1f} ms" for k, v in d["modes"].items()) +
          f"; paths_only (every pair of the frame through the path kernel's sweep): {d['modes']['paths_only']['pairs_per_s']:.3e} pairs/s, "
          f"brute-force frac {d['modes']['paths_only']['brute_force_frac']:.3f}"
          f"; FP32 FMA peak measured in the run {d['roofline']['peak']:.1f} TFLOP/s; clocks {d['clocks']}")
    cb = d.get("cpu_baseline")
    if cb:
        print(f"CPU baseline in the same run: {cb['value']:.1f} Mrays/s ({cb['kind']}, {cb['cores']} cores; {cb['sample']}); "
              f"e2e / cpu = {d['vs_cpu_baseline']['e2e']:.0f}x, value / cpu = {d['vs_cpu_baseline']['value']:.0f}x")
print()
print("| GPUs | frame ms (`value`) | Mrays/s | speed-up | efficiency | path kernel ms (rank 0) | primary ms | e2e ms `render_cuda(ngpus=N)` | e2e speed-up | frame = golden |")
print("|---|---|---|---|---|---|---|---|---|---|")
base = None
for n in (1, 2, 4, 8):
    f = os.path.join(root, f"{tag}_scale_n{n}.json")
    if not os.path.exists(f):
        continue
    d = json.loads(open(f).read())
    base = base or d
    print(f"| {n} | {d['ms_per_step']:.2f} | {d['value']:.0f} | {base['ms_per_step'] / d['ms_per_step']:.2f} | "
          f"{base['ms_per_step'] / d['ms_per_step'] / n:.2f} | {d['path_kernel']['ms']:.2f} | {d['primary_kernels']['ms']:.2f} | "
          f"{d['e2e']['ms_per_step']:.2f} | {base['e2e']['ms_per_step'] / d['e2e']['ms_per_step']:.2f} | {d['parity_fnv_ok']} |")
for name, label in (("scale_n8_nccl", "8, NCCL all_gather + un-interleave instead of peer writes"), ("c5l_n8_samples", "8, C5 lattice, sample split")):
    f = os.path.join(root, f"{tag}_{name}.json")
    if os.path.exists(f):
        d = json.loads(open(f).read())
        print(f"| {label} | {d['ms_per_step']:.2f} | {d['value']:.1f} | | | {d['path_kernel']['ms']:.2f} | {d['primary_kernels']['ms']:.2f} | "
              f"{d['e2e']['ms_per_step']:.2f} | | {d['parity_fnv_ok']} |")
