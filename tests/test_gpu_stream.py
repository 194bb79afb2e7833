"""-m gpu: the streamed (TMA-tiled) variants of the sky and path kernels, and the paths-only mode.

Small tiles are forced through the RTC_*_TILE / RTC_FORCE_STREAM knobs so the multi-tile double-buffered
pipeline (cp.async.bulk + mbarrier) runs on the golden scenes; results must stay bit-identical."""
import os

import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu

NAMES = ["suze_160x90_16spp_3b", "default_64x64_16spp_4b", "complex_sphere_160x90_16spp_4b",
         "suze_nsplit4_96x64_16spp_3b", "suze_odd_101x67_5spp_2b", "rsuzanne_128x96_32spp_5b"]


def _render_and_check(name, flags=0):
    import raytracingc_b200 as rtc
    e = util.kat()[name]
    job = util.job_of(e)
    tris = util.scene_tris(e["scene"])
    rgb, ids, st = rtc.render(tris, want_ids=True, flags=flags, **util.gpu_kwargs(job))
    want_rgb, want_ids = util.golden_image(name)
    assert (ids == want_ids).all(), f"{name}: {(ids != want_ids).sum()} ids differ"
    assert (rgb == want_rgb).all(), f"{name}: {(rgb != want_rgb).any(axis=-1).sum()} pixels differ"
    assert st["rays"] == e["rays"] and st["hit_pixels"] == e["hit_pixels"]


@pytest.fixture
def small_tiles():
    old = {k: os.environ.get(k) for k in ("RTC_FORCE_STREAM", "RTC_PATH_TILE", "RTC_SKY_TILE")}
    os.environ.update(RTC_FORCE_STREAM="1", RTC_PATH_TILE="24", RTC_SKY_TILE="20")
    yield
    for k, v in old.items():
        if v is None:
            os.environ.pop(k, None)
        else:
            os.environ[k] = v


@pytest.mark.parametrize("name", NAMES)
def test_streamed_tiles_bit_identical(small_tiles, name):
    _render_and_check(name)


@pytest.mark.parametrize("name", NAMES[:3])
@pytest.mark.parametrize("form,tile", [("1", "24"), ("1", "23"), ("3", "23"), ("3", "2")])
def test_streamed_tiles_record_layouts(small_tiles, monkeypatch, name, form, tile):
    """Both quick-record layouts (per triangle / per pair for the packed FP32 evaluation) through the TMA tiles; an
    odd tile request is rounded up to a whole pair in the packed layout."""
    monkeypatch.setenv("RTC_QUICK_FORM", form)
    monkeypatch.setenv("RTC_PATH_TILE", tile)
    _render_and_check(name)


@pytest.mark.parametrize("name", NAMES[:4])
def test_paths_only_mode(name):
    from raytracingc_b200 import api
    _render_and_check(name, flags=api.RTC_FLAG_PATHS_ONLY)


@pytest.mark.parametrize("name", NAMES[:3])
def test_paths_only_streamed(small_tiles, name):
    from raytracingc_b200 import api
    _render_and_check(name, flags=api.RTC_FLAG_PATHS_ONLY)


def test_midsize_scene_streams_natively():
    """suzannes.obj, 3868 triangles: table (242 KB) and geometry (181 KB) exceed the resident limits"""
    _render_and_check("suzannes_96x64_4spp_3b")
    from raytracingc_b200 import api
    _render_and_check("suzannes_96x64_4spp_3b", flags=api.RTC_FLAG_PATHS_ONLY)
