"""-m gpu, collected LAST (pytest orders files by name): rtc_render_paths_noalloc, the entry point that allocates nothing.

No counterpart in the reference; added for callers that cannot call cudaMalloc
(CUDA graph capture, tight serving loops).  The path kernel's scratch -- its work cursor and the per-scene quick
records -- lives in ONE static device array of the library (include/rtc_b200.h: RTC_NOALLOC_MAX_TRIS,
RTC_NOALLOC_SCRATCH_BYTES) that every call shares and that keeps the last call's contents.  So the test calls it
through the C ABI (ctypes) several times IN A ROW, on alternating scenes, and compares every image with the existing
entry point's (rtc_render_device), bit for bit -- the tolerance all image tests of this suite use is equality -- and
with the golden image from the reference.  It must pass alone in a fresh process and after all other tests.
"""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def rtc():
    import raytracingc_b200 as rtc
    assert rtc.device_count() > 0, "no CUDA device: the product has no CPU fallback"
    return rtc


def _render_both(rtc, torch, sc, job, kw, want_stats=True):
    """(existing entry point, no-alloc entry point) -> images, id maps, ray counts; device buffers are the caller's"""
    from raytracingc_b200 import api
    dev = torch.device("cuda", 0)
    j = rtc.DeviceJob(job.width, job.height, job.spp, job.max_bounce, triangles_only=job.triangles_only)
    j.to_c()
    out = []
    for which in ("device", "noalloc"):
        rgb = torch.full((j.out_h, j.out_w, 3), 0x5A, dtype=torch.uint8, device=dev)  # poisoned: every pixel must be written
        ids = torch.full((j.out_h, j.out_w), -77, dtype=torch.int32, device=dev)
        stats = torch.zeros(api.STATS_WORDS, dtype=torch.int64, device=dev)
        st = torch.cuda.current_stream()
        if which == "device":
            sc.render(j, kw["scene"], kw["camera"], d_rgb=rgb.data_ptr(), d_ids=ids.data_ptr(), d_stats=stats.data_ptr(),
                      stream=st.cuda_stream)
        else:
            sc.render_noalloc(j, kw["scene"], kw["camera"], d_rgb=rgb.data_ptr(), d_ids=ids.data_ptr(),
                              d_stats=stats.data_ptr(), stream=st.cuda_stream)
        torch.cuda.synchronize()
        s = stats.cpu().numpy()
        out.append((rgb.cpu().numpy(), ids.cpu().numpy(), int(s[0]), int(s[11])))
    return out


def test_no_alloc_entry_point_equals_the_existing_one_on_calls_in_a_row(rtc):
    import torch
    names = ["suze_160x90_16spp_3b", "ultracomplex_160x90_16spp_4b", "suze_160x90_16spp_3b", "default_64x64_16spp_4b",
             "suze_closeup_128x96_16spp_6b"]
    scenes = {}
    for round_ in range(2):  # the whole sequence twice: the shared array always holds another scene's records
        for name in names:
            e = util.kat()[name]
            job = util.job_of(e)
            if e["scene"] not in scenes:
                scenes[e["scene"]] = rtc.DeviceScene(util.scene_tris(e["scene"]), device=0)
            kw = util.gpu_kwargs(job)
            (rgb0, ids0, rays0, bad0), (rgb1, ids1, rays1, bad1) = _render_both(rtc, torch, scenes[e["scene"]], job, kw)
            assert (rgb1 == rgb0).all(), f"{name} (pass {round_}): {(rgb1 != rgb0).any(axis=-1).sum()} pixels differ from rtc_render_device"
            assert (ids1 == ids0).all() and rays1 == rays0 and bad0 == 0 and bad1 == 0
            want_rgb, want_ids = util.golden_image(name)  # and both equal the reference's image
            assert (rgb1 == want_rgb).all() and (ids1 == want_ids).all() and rays1 == e["rays"]
    for sc in scenes.values():
        sc.close()


def test_no_alloc_streamed_scene_and_back_to_back_on_one_stream(rtc):
    """more triangles than fit resident (tiles through shared memory), still within the limit; then two no-alloc calls
    queued back to back on one stream without a synchronisation in between: ordered on the GPU, so no race"""
    import torch
    from oracle.oracle_api import Job
    from raytracingc_b200 import api
    dev = torch.device("cuda", 0)
    tris = api.synth_scene(1500)
    # the generator's triangles are 0.08 across (made for a million of them): stretch every triangle about its vertex A
    # (same plane, same stored normal) so that 1500 of them are visible at 96 x 64
    tris["posB"] = tris["posA"] + np.float32(8) * (tris["posB"] - tris["posA"])
    tris["posC"] = tris["posA"] + np.float32(8) * (tris["posC"] - tris["posA"])
    job = Job(width=96, height=64, spp=4, max_bounce=3)
    kw = util.gpu_kwargs(job)
    big = rtc.DeviceScene(tris, device=0)
    (rgb0, ids0, rays0, _), (rgb1, ids1, rays1, bad1) = _render_both(rtc, torch, big, job, kw)
    assert (rgb1 == rgb0).all() and (ids1 == ids0).all() and rays1 == rays0 and bad1 == 0
    assert (ids1 != -1).any(), "the streamed test scene should be visible"
    e = util.kat()["suze_160x90_16spp_3b"]
    sjob = util.job_of(e)
    skw = util.gpu_kwargs(sjob)
    small = rtc.DeviceScene(util.scene_tris(e["scene"]), device=0)
    ja = rtc.DeviceJob(job.width, job.height, job.spp, job.max_bounce)
    jb = rtc.DeviceJob(sjob.width, sjob.height, sjob.spp, sjob.max_bounce)
    ja.to_c(), jb.to_c()
    ra = torch.zeros((ja.out_h, ja.out_w, 3), dtype=torch.uint8, device=dev)
    rb = torch.zeros((jb.out_h, jb.out_w, 3), dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream()
    big.render_noalloc(ja, kw["scene"], kw["camera"], d_rgb=ra.data_ptr(), stream=st.cuda_stream)
    small.render_noalloc(jb, skw["scene"], skw["camera"], d_rgb=rb.data_ptr(), stream=st.cuda_stream)
    torch.cuda.synchronize()
    want_rgb, _ = util.golden_image("suze_160x90_16spp_3b")
    assert (ra.cpu().numpy() == rgb0).all() and (rb.cpu().numpy() == want_rgb).all()
    big.close()
    small.close()


def test_no_alloc_rejects_what_its_scratch_cannot_hold(rtc):
    import torch
    from oracle.oracle_api import Job
    from raytracingc_b200 import api
    dev = torch.device("cuda", 0)
    job = Job(width=32, height=24, spp=2, max_bounce=2)
    kw = util.gpu_kwargs(job)
    rgb = torch.zeros((24, 32, 3), dtype=torch.uint8, device=dev)
    j = rtc.DeviceJob(32, 24, 2, 2)
    too_big = rtc.DeviceScene(api.synth_scene(api.RTC_NOALLOC_MAX_TRIS + 1), device=0)
    with pytest.raises(rtc.RtcError) as e:
        too_big.render_noalloc(j, kw["scene"], kw["camera"], d_rgb=rgb.data_ptr())
    assert e.value.code == 1 and "RTC_NOALLOC_MAX_TRIS" in str(e.value)
    at_limit = rtc.DeviceScene(api.synth_scene(api.RTC_NOALLOC_MAX_TRIS), device=0)
    at_limit.render_noalloc(j, kw["scene"], kw["camera"], d_rgb=rgb.data_ptr())  # exactly the maximum is accepted
    torch.cuda.synchronize()
    for bad in (rtc.DeviceJob(32, 24, 2, 2, nsplit=2), rtc.DeviceJob(32, 24, 2, 0)):
        with pytest.raises(rtc.RtcError) as e:
            at_limit.render_noalloc(bad, kw["scene"], kw["camera"], d_rgb=rgb.data_ptr())
        assert e.value.code == 1
    with pytest.raises(rtc.RtcError):
        at_limit.render_noalloc(j, kw["scene"], kw["camera"], d_rgb=0)
    too_big.close()
    at_limit.close()
