"""-m gpu: BASELINE.json's full-size configs through the C ABI.

A full CPU render of these takes minutes (C4: ~9 core-minutes), so the checks
are: (1) the primary-hit id map of EVERY pixel hashes to the golden value
generated from the reference (bit-exact ids), (2) the ray count equals the
reference run's (identical hit / Russian-roulette decisions everywhere), (3)
the image hash equals the golden hash when the libm-dependent sky matches to the
bit, else (4) the pixels of the committed 8x8 golden lattice are within the
north_star tolerance.
"""
import os

import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu

FULL = ["c2_suze_1280x720_16spp_3b", "c3_complex_1920x1080_64spp_4b", "c3s_complex_sphere_1920x1080_64spp_4b",
        "c4_ultracomplex_3840x2160_64spp_4b", "default_640x480_16spp_4b"]


@pytest.mark.parametrize("name", FULL)
def test_full_config(name):
    import raytracingc_b200 as rtc
    from oracle.oracle_api import fnv1a64_bytes, fnv1a64_words
    e = util.kat()[name]
    job = util.job_of(e)
    tris = util.scene_tris(e["scene"])
    rgb, ids, st = rtc.render(tris, want_ids=True, **util.gpu_kwargs(job))
    assert fnv1a64_words(ids) == e["id_fnv"], "primary-hit ids are not bit-exact"
    assert st["hit_pixels"] == e["hit_pixels"]
    assert st["rays"] == e["rays"], f"ray count {st['rays']} != reference run {e['rays']}"
    step, want_rgb, want_ids = util.golden_lattice(name)
    assert (ids[::step, ::step] == want_ids).all()
    r = util.assert_image_parity(rgb[::step, ::step], want_rgb, name)
    assert st["check_failures"] == 0
    # The whole frame, bit for bit: the hash of all W*H*3 bytes equals the reference run's.  The one term that is not
    # reproducible to the last bit offline is glibc's powf in the sky (2.5e-4 of its results differ in the last bit,
    # tests/test_gpu_math.py), and a last-bit difference of a sky value moves an 8-bit pixel with probability ~1e-5:
    # should that ever flip a pixel, set RTC_ALLOW_SKY_LSB=1 to fall back to the +-1 LSB rule on the lattice above --
    # the fallback is reported, never silent.
    exact = fnv1a64_bytes(rgb) == e["image_fnv"]
    if not exact and os.environ.get("RTC_ALLOW_SKY_LSB") == "1":
        print(name, "IMAGE HASH DIFFERS from the reference run; +-1 LSB lattice rule applied instead:", r)
    else:
        assert exact, f"{name}: image hash {fnv1a64_bytes(rgb)} != reference run {e['image_fnv']} (lattice parity {r})"
    print(name, "image hash identical to the reference run:", exact, r, f"kernel {st['kernel_ms']:.2f} ms")
