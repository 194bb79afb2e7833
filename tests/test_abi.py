"""-m "not gpu": the C-ABI library loads, exports every symbol include/rtc_b200.h declares, keeps the reference's
struct layouts, and FAILS LOUDLY without a GPU (no CPU fallback).  No GPU compute is called here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from tests import util

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def rtc():
    import raytracingc_b200 as rtc
    from raytracingc_b200 import build
    build.build_lib()
    rtc.load_library()
    return rtc


def declared_functions():
    """every function prototype of include/rtc_b200.h"""
    src = open(os.path.join(ROOT, "include", "rtc_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"typedef struct \w+ \{.*?\} \w+;", "", src, flags=re.S)
    names = re.findall(r"\b(\w+)\s*\([^;{}]*\)\s*;", src)
    return sorted(set(n for n in names if n.startswith(("rtc_", "render_cuda"))))


def test_every_declared_symbol_is_exported(rtc):
    from raytracingc_b200 import api
    lib = rtc.load_library()
    names = declared_functions()
    assert len(names) >= 17 and "render_cuda" in names and "rtc_render_device" in names
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/rtc_b200.h but not exported by librtc_b200.so"
    assert sorted(api.EXPORTS) == names, "api.EXPORTS is out of date with the header"


def test_struct_layouts_match_the_reference(rtc):
    from raytracingc_b200 import api
    # reference raytracing.h: Scene 56 B, Triangle 68 B, Sphere 36 B; Camera = 4 vec3 + fov
    assert C.sizeof(api.Scene) == 56 and C.sizeof(api.Camera) == 52
    assert api.TRI_DTYPE.itemsize == 68 and api.SPHERE_DTYPE.itemsize == 36
    assert api.TRI_DTYPE.fields["normal"][1] == 36 and api.TRI_DTYPE.fields["color"][1] == 48
    assert C.sizeof(api.RtcSubset) == 20 and C.sizeof(api.RtcDeviceJob) == 68


def test_version_and_error_strings(rtc):
    lib = rtc.load_library()
    assert b"sm_100a" in lib.rtc_version()
    assert isinstance(lib.rtc_last_error(), bytes)


def test_camera_basis_equals_the_oracle(rtc, port_oracle):
    """rtc_camera_look_at is host code (main.c:252-255): bit-identical to the checker's"""
    for origin, look, fov in [((-4.75, -1.5, -4.75), (0.9, -1.2, 1.0), 1.0), ((0, -2, -6), (0, -1, 0), 2.0),
                              ((3, 1, 2), (-1, 0.5, 7), 0.5)]:
        assert bytes(rtc.camera_look_at(origin, look, fov)) == bytes(port_oracle.camera(origin, look, fov))
    sun = (-30.0, -85.0, 100.0)
    assert np.array(rtc.api.normalize(sun), np.float32).tobytes() == \
        np.array(port_oracle.normalize(sun), np.float32).tobytes()


def test_subset_size(rtc):
    from raytracingc_b200 import api, split
    assert api.subset_size(3840, 2160, api.RtcSubset(0, 1, 0, 1, 1)) == (3840, 2160)
    assert api.subset_size(101, 67, api.RtcSubset(0, 8, 0, 8, 1)) == (13, 9)
    for world in (2, 3, 4, 8):
        for H in (2160, 67, 9, 8):
            total = 0
            for g in range(world):
                w, h = api.subset_size(64, H, api.RtcSubset(*split.band_subset(g, world)))
                assert w == 64 and h == len(split.rows_of(g, world, H))
                total += h
            assert total == H


def test_no_gpu_means_error_not_fallback(rtc):
    if rtc.device_count() > 0:
        pytest.skip("a GPU is visible")
    tris = util.scene_tris("simplest")
    with pytest.raises(rtc.RtcError) as e:
        rtc.render(tris, 32, 24, 1, 1)
    assert e.value.code == 2  # RTC_ERR_NO_GPU
    with pytest.raises(rtc.RtcError):
        rtc.DeviceScene(tris)


def test_argument_checks(rtc):
    tris = util.scene_tris("simplest")
    for kw in (dict(width=0, height=24), dict(width=32, height=1), dict(width=32, height=24, spp=0),
               dict(width=32, height=24, max_bounce=-1)):
        args = dict(width=32, height=24, spp=1, max_bounce=1)
        args.update(kw)
        with pytest.raises(rtc.RtcError) as e:
            rtc.render(tris, **args)
        assert e.value.code == 1, kw  # RTC_ERR_ARG comes before any device work


def test_sample_split_needs_an_explicit_nsplit(rtc):
    """The sample split's sub-stream seeds are this library's own definition: render_cuda refuses to pick a count
    from the number of GPUs (the image would silently depend on it).  Checked before any device work."""
    from raytracingc_b200 import api
    tris = np.zeros(1, api.TRI_DTYPE)
    for nsplit in (0, 1, 3):
        with pytest.raises(rtc.RtcError) as e:
            rtc.render(tris, 32, 24, 4, 1, ngpus=2, split_mode=api.RTC_SPLIT_SAMPLES, nsplit=nsplit)
        assert e.value.code == 1 and "nsplit" in str(e.value)


def test_fnv_helper_equals_the_oracle_hash(rtc):
    from oracle.oracle_api import fnv1a64_bytes
    from raytracingc_b200 import api
    a = np.random.default_rng(7).integers(0, 256, 10007, dtype=np.uint8)
    assert api.fnv1a64(a) == fnv1a64_bytes(a)


def test_download_bands_argument_checks(rtc):
    """rtc_download_bands rejects bad arguments before any device work (no GPU needed)."""
    from raytracingc_b200 import api
    buf = np.zeros(64, np.uint8)
    p = buf.ctypes.data
    for args in ((0, 0, p, 24, 8, 2, 8, 0), (0, p, 0, 24, 8, 2, 8, 0), (0, p, p, 0, 8, 2, 8, 0), (0, p, p, 24, 0, 2, 8, 0),
                 (0, p, p, 24, 8, 0, 8, 0), (0, p, p, 24, 8, 2, 0, 0), (0, p, p, 24, 8, 2, 8, 2), (0, p, p, 24, 8, 2, 8, -1)):
        with pytest.raises(rtc.RtcError) as e:
            api.download_bands(*args)
        assert e.value.code == 1, args


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under raytracingc_b200/ or include/ may reference it"""
    for d in ("raytracingc_b200", "include"):
        for base, _, files in os.walk(os.path.join(ROOT, d)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".c")):
                    txt = open(os.path.join(base, f)).read()
                    assert "oracle_api" not in txt and "liboracle" not in txt and "libref_oracle" not in txt, f
