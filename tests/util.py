"""Shared helpers of the test-suite: golden fixtures, parity metrics."""
from __future__ import annotations

import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")

from oracle.oracle_api import TRI_DTYPE, Job  # noqa: E402


def kat() -> dict:
    return json.load(open(os.path.join(GOLDEN, "kat.json")))


def scene_tris(name: str) -> np.ndarray:
    """Triangle[] the reference's loader produced for a bundled scene (tests/golden/make_golden.py)."""
    return np.fromfile(os.path.join(GOLDEN, "scenes", name + ".tris"), dtype=TRI_DTYPE)


def golden_image(name: str):
    z = np.load(os.path.join(GOLDEN, "images", name + ".npz"))
    return z["rgb"], z["ids"]


def golden_lattice(name: str):
    z = np.load(os.path.join(GOLDEN, "lattice", name + ".npz"))
    return int(z["step"]), z["rgb"], z["ids"]


def job_of(entry: dict) -> Job:
    kw = dict(entry["job"])
    for k in ("origin", "look_at"):
        if k in kw:
            kw[k] = tuple(kw[k])
    return Job(**kw)


def psnr(a: np.ndarray, b: np.ndarray) -> float:
    mse = np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)
    return float("inf") if mse == 0 else 10.0 * np.log10(255.0 ** 2 / mse)


def parity_report(got: np.ndarray, want: np.ndarray) -> dict:
    """north_star tolerance: per pixel max channel difference."""
    d = np.abs(got.astype(np.int16) - want.astype(np.int16)).max(axis=-1)
    return {"exact_frac": float((d == 0).mean()), "within1_frac": float((d <= 1).mean()), "max_diff": int(d.max()),
            "psnr": psnr(got, want), "n_gt1": int((d > 1).sum())}


def assert_image_parity(got: np.ndarray, want: np.ndarray, what: str = ""):
    """BASELINE.json north_star: 8-bit pixels within +-1 LSB on >= 99.9 % of pixels, PSNR >= 50 dB."""
    r = parity_report(got, want)
    assert r["within1_frac"] >= 0.999, f"{what}: only {r['within1_frac']:.5f} of pixels within 1 LSB ({r})"
    assert r["psnr"] >= 50.0, f"{what}: PSNR {r['psnr']:.2f} dB < 50 ({r})"
    return r


def gpu_kwargs(job: Job) -> dict:
    """Job (oracle vocabulary) -> keyword arguments of raytracingc_b200.render()."""
    import raytracingc_b200 as rtc
    return dict(width=job.width, height=job.height, spp=job.spp, max_bounce=job.max_bounce,
                triangles_only=job.triangles_only, nsplit=job.nsplit,
                camera=rtc.camera_look_at(job.origin, job.look_at, job.fov),
                scene=rtc.make_scene(job.sun, job.sun_focus, job.sun_intensity, job.horizon, job.zenith, job.ground))
