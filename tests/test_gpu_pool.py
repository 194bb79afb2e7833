"""-m gpu: the two forms of the path kernel against the golden vectors and against each other.

The pool form (rtc_pool.cuh) evaluates the samples of a pixel OUT OF ORDER (speculatively, from the RNG state 7*j draws
after the pixel's seed) and consumes them in order; the serial form (rtc_paths.cuh) runs them one after the other like
the reference.  Both must reproduce the reference's images bit for bit and its ray count exactly -- whatever the pool
geometry, the share of pixels a warp holds, with and without speculation, resident and streamed.
"""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu

KATS = ["suze_160x90_16spp_3b", "ultracomplex_160x90_16spp_4b", "complex_sphere_160x90_16spp_4b", "default_64x64_16spp_4b",
        "default_96x64_8spp_10b", "rsuzanne_128x96_32spp_5b", "suze_closeup_128x96_16spp_6b", "suze_odd_101x67_5spp_2b",
        "simplest_64x48_4000spp_1b", "suze_nsplit4_96x64_16spp_3b"]


@pytest.fixture(scope="module")
def rtc():
    import raytracingc_b200 as rtc
    assert rtc.device_count() > 0, "no CUDA device: the product has no CPU fallback"
    return rtc


def render_kat(rtc, name, flags=0):
    e = util.kat()[name]
    job = util.job_of(e)
    tris = util.scene_tris(e["scene"])
    rgb, ids, st = rtc.render(tris, want_ids=True, flags=flags, **util.gpu_kwargs(job))
    return e, rgb, ids, st


@pytest.mark.parametrize("form", ["pool", "serial"])
@pytest.mark.parametrize("name", KATS)
def test_each_form_reproduces_the_reference(rtc, name, form, monkeypatch):
    monkeypatch.setenv("RTC_PATHS", form)
    e, rgb, ids, st = render_kat(rtc, name)
    want_rgb, want_ids = util.golden_image(name)
    assert (ids == want_ids).all()
    assert (rgb == want_rgb).all(), f"{name} [{form}]: {(rgb != want_rgb).any(axis=-1).sum()} pixels differ from the reference's"
    assert st["rays"] == e["rays"], f"{name} [{form}]: rays {st['rays']} != reference {e['rays']}"
    assert st["check_failures"] == 0
    if form == "pool":
        # the pool form ran (its counters moved) and evaluated at least every consumed sample
        hit_samples = e["hit_pixels"] * e["job"]["spp"]
        assert st["pool_samples_evaluated"] >= hit_samples > 0 or e["hit_pixels"] == 0
    else:
        assert st["pool_samples_evaluated"] == 0


@pytest.mark.parametrize("env", [{"RTC_NO_SPEC": "1"}, {"RTC_POOL_SHARE_DIV": "1"}, {"RTC_POOL_SHARE_DIV": "64"},
                                 {"RTC_POOL_CFG": "1"}, {"RTC_FORCE_STREAM": "1", "RTC_PATH_TILE": "32"},
                                 {"RTC_REF_QUICK": "1"}, {"RTC_SKY_SEGS": "4"}, {"RTC_SKY_SEGS": "16"}],
                         ids=lambda d: ",".join(f"{k}={v}" for k, v in d.items()))
@pytest.mark.parametrize("name", ["ultracomplex_160x90_16spp_4b", "default_96x64_8spp_10b", "suze_closeup_128x96_16spp_6b"])
def test_pool_variants_are_bit_identical(rtc, name, env, monkeypatch):
    """scheduler knobs, kernel configurations (16-warp CTAs, streamed tiles), the reference-arithmetic quick test and
    the primary kernel's sample segments change nothing that reaches the image or the ray count"""
    monkeypatch.setenv("RTC_PATHS", "pool")
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    e, rgb, ids, st = render_kat(rtc, name)
    want_rgb, want_ids = util.golden_image(name)
    assert (ids == want_ids).all() and (rgb == want_rgb).all()
    assert st["rays"] == e["rays"] and st["check_failures"] == 0


@pytest.mark.parametrize("name", ["ultracomplex_160x90_16spp_4b", "complex_sphere_160x90_16spp_4b", "default_64x64_16spp_4b"])
def test_pool_quick_test_never_drops_an_accepted_pair(rtc, name, monkeypatch):
    """RTC_FLAG_VERIFY_QUICK in the pool form: every (ray, triangle) pair also runs the exact test; an accepted pair
    whose conservative quick bit was clear would be counted in check_failures"""
    from raytracingc_b200 import api
    monkeypatch.setenv("RTC_PATHS", "pool")
    e, rgb, ids, st = render_kat(rtc, name, flags=api.RTC_FLAG_VERIFY_QUICK)
    want_rgb, _ = util.golden_image(name)
    assert st["check_failures"] == 0
    assert (rgb == want_rgb).all() and st["rays"] == e["rays"]


def test_ref_quick_flag_is_pinned_to_the_reference(rtc, monkeypatch):
    """RTC_FLAG_REF_QUICK (first phase of the sweep in the reference's own arithmetic) against the golden image itself,
    in both forms -- not only against the other modes"""
    from raytracingc_b200 import api
    for form in ("pool", "serial"):
        monkeypatch.setenv("RTC_PATHS", form)
        for name in ("suze_160x90_16spp_3b", "default_64x64_16spp_4b"):
            e, rgb, ids, st = render_kat(rtc, name, flags=api.RTC_FLAG_REF_QUICK)
            want_rgb, want_ids = util.golden_image(name)
            assert (rgb == want_rgb).all() and (ids == want_ids).all() and st["rays"] == e["rays"], (form, name)


def test_auto_selection_takes_the_pool_for_short_lists_and_the_serial_form_for_long_ones(rtc, monkeypatch):
    monkeypatch.delenv("RTC_PATHS", raising=False)
    e, rgb, ids, st = render_kat(rtc, "ultracomplex_160x90_16spp_4b")      # 530 hit pixels: far fewer than lanes
    assert st["pool_samples_evaluated"] > 0
    monkeypatch.setenv("RTC_POOL_MAX_FACTOR", "0")                          # every list counts as long
    e, rgb2, ids2, st2 = render_kat(rtc, "ultracomplex_160x90_16spp_4b")
    assert st2["pool_samples_evaluated"] == 0 and (rgb2 == rgb).all() and st2["rays"] == st["rays"] == e["rays"]
