"""-m gpu: render_cuda(ngpus = G) -- one process driving G devices, NCCL all-gather over NVLink.

Needs >= 2 visible devices (skipped on a 1-GPU box).  The output must be bit-identical for every G and both split
modes' definitions must match the oracle's (tile split: same image; sample split: the oracle's nsplit image)."""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def rtc():
    import raytracingc_b200 as rtc
    if rtc.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    return rtc


def gpu_counts(rtc):
    return [g for g in (2, 4, 8) if g <= rtc.device_count()]


@pytest.mark.parametrize("gather", ["0", "1"])
@pytest.mark.parametrize("name", ["suze_odd_101x67_5spp_2b", "default_64x64_16spp_4b", "ultracomplex_160x90_16spp_4b"])
def test_tile_split_is_bit_identical(rtc, name, gather, monkeypatch):
    """Both ways of assembling the frame: every device copies its own bands to the host image (default), or NCCL
    all-gather + un-interleave on device 0 and one copy (RTC_GATHER=1)."""
    monkeypatch.setenv("RTC_GATHER", gather)
    e = util.kat()[name]
    job = util.job_of(e)
    tris = util.scene_tris(e["scene"])
    want_rgb, want_ids = util.golden_image(name)
    for G in gpu_counts(rtc):
        rgb, ids, st = rtc.render(tris, want_ids=True, ngpus=G, **util.gpu_kwargs(job))
        assert st["ngpus_used"] == G
        assert (ids == want_ids).all(), f"G={G}"
        assert (rgb == want_rgb).all(), f"G={G}: {(rgb != want_rgb).any(axis=-1).sum()} pixels differ"
        assert st["rays"] == e["rays"] and st["hit_pixels"] == e["hit_pixels"]


@pytest.mark.parametrize("name", ["suze_nsplit4_96x64_16spp_3b", "default_nsplit2_64x48_8spp_4b"])
def test_sample_split_matches_the_oracle_definition(rtc, name):
    from raytracingc_b200 import api
    e = util.kat()[name]
    job = util.job_of(e)
    tris = util.scene_tris(e["scene"])
    want_rgb, want_ids = util.golden_image(name)
    one, _, st1 = rtc.render(tris, want_ids=True, **util.gpu_kwargs(job))  # all sub-streams on one GPU
    assert (one == want_rgb).all()
    for G in gpu_counts(rtc):
        if job.nsplit % G:
            continue
        rgb, ids, st = rtc.render(tris, want_ids=True, ngpus=G, split_mode=api.RTC_SPLIT_SAMPLES,
                                  **util.gpu_kwargs(job))
        assert (ids == want_ids).all()
        assert (rgb == want_rgb).all(), f"G={G}: ordered sum of sub-stream planes must be bit-exact"
        assert st["rays"] == e["rays"]


def test_full_size_c2_over_all_gpus(rtc):
    from oracle.oracle_api import fnv1a64_bytes, fnv1a64_words
    e = util.kat()["c2_suze_1280x720_16spp_3b"]
    job = util.job_of(e)
    tris = util.scene_tris(e["scene"])
    for G in gpu_counts(rtc):
        rgb, ids, st = rtc.render(tris, want_ids=True, ngpus=G, **util.gpu_kwargs(job))
        assert fnv1a64_words(ids) == e["id_fnv"] and fnv1a64_bytes(rgb) == e["image_fnv"]
        assert st["rays"] == e["rays"]
        print(f"C2 on {G} GPUs: kernel {st['kernel_ms']:.2f} ms, gather {st['gather_ms']:.2f} ms")


@pytest.mark.parametrize("name", ["ultracomplex_160x90_16spp_4b", "suze_odd_101x67_5spp_2b"])
def test_ranks_write_their_bands_in_place_over_peer_memory(rtc, name):
    """RtcDeviceJob.rgb_full: every device stores its pixels at their final address of ONE frame in device 0's memory
    (peer access over NVLink) -- the frame is assembled by the kernels' own stores, no gather, no un-interleave."""
    import torch
    from raytracingc_b200 import api, split
    e = util.kat()[name]
    job = util.job_of(e)
    tris = util.scene_tris(e["scene"])
    want_rgb, _ = util.golden_image(name)
    kw = util.gpu_kwargs(job)
    for G in gpu_counts(rtc):
        image = torch.zeros((job.height, job.width, 3), dtype=torch.uint8, device="cuda:0")
        scenes = []
        for g in range(G):
            api.enable_peer(g, 0)
            sc = rtc.DeviceScene(tris, device=g)
            scenes.append(sc)
            dj = rtc.DeviceJob(job.width, job.height, job.spp, job.max_bounce, triangles_only=job.triangles_only,
                               subset=split.band_subset(g, G), rgb_full=1)
            sc.render(dj, kw["scene"], kw["camera"], d_rgb=image.data_ptr())
        for g in range(G):
            torch.cuda.synchronize(g)
        got = image.cpu().numpy()
        assert (got == want_rgb).all(), f"G={G}: {(got != want_rgb).any(axis=-1).sum()} pixels differ"
        for sc in scenes:
            sc.close()
