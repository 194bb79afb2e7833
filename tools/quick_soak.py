#!/usr/bin/env python
"""Soak of the conservative quick test (packed form by default, RTC_QUICK_FORM=1 for the scalar one): many random
scenes of the kinds tests/test_gpu_quick.py uses, each rendered with RTC_FLAG_VERIFY_QUICK (every pair also runs the
exact test; accepted pairs whose quick bit was clear are counted and must be 0) and compared with the exact-only kernel.

usage: python tools/quick_soak.py [--seeds 100]"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracingc_b200 as rtc  # noqa: E402
from raytracingc_b200 import api  # noqa: E402
from tests.test_gpu_quick import SCENES, make_scene  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--seeds", type=int, default=100)
a = ap.parse_args()
pairs = fails = diff = 0
for seed in range(a.seeds):
    rng = np.random.default_rng(90000 + seed)
    name, n, extent, size, normals, offset = SCENES[seed % len(SCENES)]
    n = int(rng.integers(1, 2 * n + 1))
    size = float(size * rng.uniform(0.3, 3.0))
    tris = make_scene(rng, n, extent, size, normals, offset)
    look = np.array(offset, np.float64)
    eye = look + rng.uniform(-1, 1, 3) * extent + np.array([0.0, 0.0, -extent - 3.0])
    cam = rtc.camera_look_at(tuple(eye), tuple(look), float(rng.uniform(0.6, 1.6)))
    kw = dict(width=128, height=72, spp=3, max_bounce=5, triangles_only=1, camera=cam, want_ids=True)
    rgb_e, ids_e, st_e = rtc.render(tris, flags=api.RTC_FLAG_EXACT_ONLY, **kw)
    for base in (0, api.RTC_FLAG_PATHS_ONLY):
        rgb_v, ids_v, st_v = rtc.render(tris, flags=base | api.RTC_FLAG_VERIFY_QUICK, **kw)
        rgb_f, ids_f, st_f = rtc.render(tris, flags=base, **kw)
        fails += st_v["check_failures"]
        pairs += st_v["rays"] * n
        diff += int((rgb_v != rgb_e).any()) + int((ids_v != ids_e).any()) + int((rgb_f != rgb_e).any()) + int(st_f["rays"] != st_e["rays"])
print(f"quick soak, form {os.environ.get('RTC_QUICK_FORM', '3')}: {a.seeds} scenes x 2 modes, {pairs:.3e} (ray, triangle) pairs verified, "
      f"{fails} accepted pairs dropped by the quick test, {diff} image/id/ray-count differences against the exact-only kernel")
sys.exit(1 if fails or diff else 0)
