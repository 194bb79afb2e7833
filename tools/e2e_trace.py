#!/usr/bin/env python
"""Host timeline of render_cuda() with pinned buffers (what bench.py's e2e leg times): wall per call, the library's own
RTC_TRACE lines (stderr), and its device-side numbers.  usage: RTC_TRACE=1 python tools/e2e_trace.py [--workload c4] [--frames 8]"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
import raytracingc_b200 as rtc  # noqa: E402
from raytracingc_b200 import api  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="c4")
ap.add_argument("--frames", type=int, default=8)
a = ap.parse_args()
wl = bench.WORKLOADS[a.workload]
W, H, spp, bounce, tonly = wl["W"], wl["H"], wl["spp"], wl["bounce"], wl["tonly"]
tris = bench.load_tris(wl["scene"], wl.get("synth", 0))
cam = rtc.camera_look_at(*wl["camera"]) if wl.get("camera") else rtc.camera_look_at()
sky = rtc.make_scene()
host_img = torch.empty((H, W, 3), dtype=torch.uint8).pin_memory()
host_tris = torch.from_numpy(tris.view(np.uint8).copy()).pin_memory().numpy().view(api.TRI_DTYPE)
walls = []
for i in range(a.frames):
    t = time.perf_counter()
    _, _, st = rtc.render(host_tris, W, H, spp, bounce, triangles_only=tonly, camera=cam, scene=sky, out=host_img.numpy())
    walls.append((time.perf_counter() - t) * 1e3)
    print(f"frame {i}: wall {walls[-1]:.3f} ms, kernels {st['kernel_ms']:.3f} ms, upload {st['upload_ms']:.3f} ms, launches {st['kernels_launched']}")
print(f"median wall {np.median(walls[2:]):.3f} ms over {len(walls) - 2} frames; image fnv {api.fnv1a64(host_img.numpy())}")
