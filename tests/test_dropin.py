"""The drop-in claim, against the reference's own main.c.

oracle/dropin.py applies the INTEGRATION.md patch to a scratch copy of /root/reference (accessor for raytracing.c's
private triangle array, main.c:246 -> heap image, main.c:285-302 -> one render_cuda() call) and links it against
librtc_b200.so.  CPU: the patch applies, the program builds, and without a GPU it fails loudly (no CPU fallback).
GPU: the patched reference binary renders the one configuration the STOCK binary reproduces too (RNG-free:
simplest.obj 64x48, the stock 4000 spp, 1 bounce) and its BMP equals the golden image bit for bit.
"""
import os
import struct
import subprocess

import numpy as np
import pytest

from tests import util

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
DROPIN_DIR = os.path.join(ROOT, "oracle", "_ref", "dropin")
BINARY = os.path.join(DROPIN_DIR, "rayt_dropin")
ARGS = ["-i", "3Dmodels/simplest.obj", "-s", "64", "48", "-b", "1"]


def read_bmp24(path: str) -> np.ndarray:
    """24-bit uncompressed BMP (what stbi_write_bmp writes, reference main.c:305) -> uint8 [H, W, 3] RGB, row 0 on top"""
    raw = open(path, "rb").read()
    assert raw[:2] == b"BM"
    off = struct.unpack_from("<I", raw, 10)[0]
    w, h = struct.unpack_from("<ii", raw, 18)
    bpp = struct.unpack_from("<H", raw, 28)[0]
    assert bpp == 24
    stride = (3 * w + 3) & ~3
    rows = np.frombuffer(raw, np.uint8, count=stride * abs(h), offset=off).reshape(abs(h), stride)[:, :3 * w]
    img = rows.reshape(abs(h), w, 3)[:, :, ::-1]  # BGR -> RGB
    return np.ascontiguousarray(img[::-1] if h > 0 else img)  # positive height = bottom-up


def test_integration_patch_builds_against_the_reference(tmp_path):
    if not os.path.exists("/root/reference/main.c"):
        pytest.skip("/root/reference only exists in the build container")
    from oracle import dropin
    binary = dropin.build()
    assert os.access(binary, os.X_OK)
    import raytracingc_b200 as rtc
    if rtc.device_count() == 0:
        # no GPU here: the patched reference must fail loudly -- there is no CPU rendering path behind render_cuda()
        r = subprocess.run([binary, *ARGS, "-o", str(tmp_path / "x.bmp")], cwd=DROPIN_DIR, capture_output=True, text=True)
        assert r.returncode == 1 and "no CUDA device" in r.stderr
        assert not (tmp_path / "x.bmp").exists()


@pytest.mark.gpu
def test_patched_reference_binary_renders_the_golden_image(tmp_path):
    if not os.path.exists(BINARY):
        pytest.skip("oracle/_ref/dropin/rayt_dropin not built (python -m oracle.dropin, needs /root/reference)")
    from oracle.oracle_api import fnv1a64_bytes
    out = tmp_path / "dropin.bmp"
    r = subprocess.run([BINARY, *ARGS, "-o", str(out)], cwd=DROPIN_DIR, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    rgb = read_bmp24(str(out))
    e = util.kat()["simplest_64x48_4000spp_1b"]
    want_rgb, _ = util.golden_image("simplest_64x48_4000spp_1b")
    assert rgb.shape == want_rgb.shape
    assert (rgb == want_rgb).all(), f"{(rgb != want_rgb).any(axis=-1).sum()} pixels differ from the reference's image"
    assert fnv1a64_bytes(rgb) == e["image_fnv"] == "5134a7dbc5bd16bb"  # the hash the STOCK rayt.out produces too
