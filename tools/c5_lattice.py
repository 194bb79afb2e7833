#!/usr/bin/env python
"""BASELINE.json config 5 on a sparse pixel lattice of the FULL 3840x2160 frame (every pixel owns its seed, so the
lattice is exact): synthetic 1 M-triangle scene, 256 spp, 4 bounces, optional sample split into sub-streams.
Prints device time, rays and ray-triangle pairs per second of the streamed (TMA-tiled) kernels."""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import raytracingc_b200 as rtc  # noqa: E402
from raytracingc_b200 import api  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--step", type=int, default=60)
ap.add_argument("--spp", type=int, default=256)
ap.add_argument("--ntris", type=int, default=1 << 20)
ap.add_argument("--nsplit", type=int, default=1)
ap.add_argument("--flags", type=int, default=0)
a = ap.parse_args()
tris = api.synth_scene(a.ntris)
sc = rtc.DeviceScene(tris, device=0)
job = rtc.DeviceJob(3840, 2160, a.spp, 4, triangles_only=1, nsplit=a.nsplit,
                    subset=(a.step // 2, a.step, a.step // 2, a.step, 1), flags=a.flags)
job.to_c()
dev = torch.device("cuda", 0)
rgb = torch.zeros((job.out_h, job.out_w, 3), dtype=torch.uint8, device=dev)
acc = torch.zeros((max(a.nsplit, 1), job.out_h, job.out_w, 3), dtype=torch.float32, device=dev)
stats = torch.zeros(24, dtype=torch.int64, device=dev)
cam, sky = rtc.camera_look_at(), rtc.make_scene()
st = torch.cuda.current_stream()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
t0 = time.perf_counter()
e0.record(st)
sc.render(job, sky, cam, d_rgb=rgb.data_ptr(), d_accum=acc.data_ptr() if a.nsplit > 1 else 0, d_stats=stats.data_ptr(),
          stream=st.cuda_stream)
if a.nsplit > 1:
    api.sum_quantize(0, acc.data_ptr(), a.nsplit, job.out_h * job.out_w * 3, d_rgb=rgb.data_ptr(), stream=st.cuda_stream)
e1.record(st)
torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
s = stats.cpu().numpy()
print(f"C5 lattice step {a.step}: {job.out_w}x{job.out_h} px of 3840x2160, {a.ntris} triangles, {a.spp} spp, nsplit "
      f"{a.nsplit}: {ms:.1f} ms, rays {int(s[0])}, pairs {int(s[1]):.3e}, {s[1] / ms / 1e6:.1f} G pairs/s, "
      f"{s[0] / ms / 1e3:.3f} Mrays/s, mismatches {int(s[11])}, mean image {rgb.float().mean().item():.2f}")
if a.flags & api.RTC_FLAG_TIME_KERNELS:
    pm, qm = sc.kernel_times()
    nprim = job.out_w * job.out_h * a.spp
    print(f"  primary kernels {pm:.1f} ms ({nprim * a.ntris / pm / 1e6:.1f} G pairs/s over {nprim} primary rays), "
          f"path kernel {qm:.1f} ms ({(int(s[0]) - nprim) * a.ntris / qm / 1e6:.1f} G pairs/s over {int(s[0]) - nprim} bounce rays)")
