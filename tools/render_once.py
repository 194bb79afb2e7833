#!/usr/bin/env python
"""Render N frames of a bench workload through the C ABI (for ncu / quick timing); prints kernel ms."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import raytracingc_b200 as rtc  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="c4")
ap.add_argument("--frames", type=int, default=2)
ap.add_argument("--flags", type=int, default=0)
a = ap.parse_args()
wl = bench.WORKLOADS[a.workload]
W, H, spp, bounce, tonly = wl["W"], wl["H"], wl["spp"], wl["bounce"], wl["tonly"]
tris = bench.load_tris(wl["scene"], wl.get("synth", 0))
cam = rtc.camera_look_at(*wl["camera"]) if wl.get("camera") else rtc.camera_look_at()
for i in range(a.frames):
    rgb, _, st = rtc.render(tris, W, H, spp, bounce, triangles_only=tonly, flags=a.flags, camera=cam)
    print(f"{a.workload} frame {i}: kernel {st['kernel_ms']:.3f} ms, rays {st['rays']}, upload {st['upload_ms']:.3f} ms")
