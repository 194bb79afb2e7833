"""-m gpu: the CUDA path (through the C ABI, render_cuda) against the golden vectors and the CPU oracle."""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu

SMALL = [k for k, v in util.kat().items() if not k.startswith("_") and
         v["job"]["width"] * v["job"]["height"] <= 320 * 240]


@pytest.fixture(scope="module")
def rtc():
    import raytracingc_b200 as rtc
    assert rtc.device_count() > 0, "no CUDA device: the product has no CPU fallback"
    return rtc


@pytest.mark.parametrize("name", SMALL)
def test_small_kat(rtc, name):
    from oracle.oracle_api import fnv1a64_words
    e = util.kat()[name]
    job = util.job_of(e)
    tris = util.scene_tris(e["scene"])
    rgb, ids, st = rtc.render(tris, want_ids=True, **util.gpu_kwargs(job))
    want_rgb, want_ids = util.golden_image(name)
    # primary-hit ids: bit exact
    assert (ids == want_ids).all(), f"{name}: {(ids != want_ids).sum()} primary ids differ"
    assert fnv1a64_words(ids) == e["id_fnv"]
    assert st["hit_pixels"] == e["hit_pixels"]
    # pixels: within tolerance (and normally identical)
    r = util.assert_image_parity(rgb, want_rgb, name)
    # ray count is a strong check of the RNG stream / path decisions: identical decisions => identical count
    assert st["rays"] == e["rays"], f"{name}: rays {st['rays']} != oracle {e['rays']} ({r})"
    print(name, r)
