#!/usr/bin/env python
"""One rank's share of a band-split frame on ONE GPU (what rank r of G renders in bench.py --gpus G): kernel times.

usage: python tools/rank_emul.py --gpus 8 [--rank 0] [--workload c4] [--frames 5]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
import raytracingc_b200 as rtc  # noqa: E402
from raytracingc_b200 import api, split  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="c4")
ap.add_argument("--gpus", type=int, default=8)
ap.add_argument("--rank", type=int, default=0)
ap.add_argument("--frames", type=int, default=5)
a = ap.parse_args()
wl = bench.WORKLOADS[a.workload]
W, H, spp, bounce, tonly = wl["W"], wl["H"], wl["spp"], wl["bounce"], wl["tonly"]
tris = bench.load_tris(wl["scene"], wl.get("synth", 0))
sc = rtc.DeviceScene(tris, device=0)
job = rtc.DeviceJob(W, H, spp, bounce, triangles_only=tonly, subset=split.band_subset(a.rank, a.gpus),
                    flags=api.RTC_FLAG_TIME_KERNELS)
job.to_c()
dev = torch.device("cuda", 0)
rgb = torch.zeros((job.out_h, job.out_w, 3), dtype=torch.uint8, device=dev)
stats = torch.zeros(24, dtype=torch.int64, device=dev)
cam = rtc.camera_look_at(*wl["camera"]) if wl.get("camera") else rtc.camera_look_at()
sky = rtc.make_scene()
st = torch.cuda.current_stream()
for i in range(a.frames):
    stats.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    sc.render(job, sky, cam, d_rgb=rgb.data_ptr(), d_stats=stats.data_ptr(), stream=st.cuda_stream)
    e1.record(st)
    torch.cuda.synchronize()
    pm, qm = sc.kernel_times()
    s = stats.cpu().numpy()
    print(f"rank {a.rank}/{a.gpus} frame {i}: {e0.elapsed_time(e1):.3f} ms (primary {pm:.3f}, paths {qm:.3f}), rays {int(s[0])}, "
          f"hit pixels {int(s[9])}, check failures {int(s[11])}; pool: samples evaluated {int(s[12])}, shade passes {int(s[13])} "
          f"({int(s[14]) / max(int(s[13]), 1):.1f} lanes), sweep passes {int(s[15])} ({int(s[16]) / max(int(s[15]), 1):.1f} rays), "
          f"iterations {int(s[17])}")
