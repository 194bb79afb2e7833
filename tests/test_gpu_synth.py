"""-m gpu: BASELINE.json config 5 -- the synthetic random-triangle scene (rtc_synth_scene, seed 0xB200).

The full config (1 M triangles, 3840x2160, 256 spp) is ~2e15 ray-triangle pairs: hours on a GPU, weeks on the
CPU.  Every pixel owns its seed, so a sparse pixel lattice at the FULL image size is exact: the GPU
(streamed TMA tiles, all kernels) is compared there with the CPU oracle, bit for bit, and the filtered kernels are
compared with the exact-only kernel on denser subsets."""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu


def _oracle_lattice(tris, job):
    from oracle.oracle_api import Oracle
    o = Oracle("port")
    o.set_triangles(tris)
    rgb, ids, st = o.render(job)
    return rgb[job.row_start::job.row_step, job.col_start::job.col_step], \
        ids[job.row_start::job.row_step, job.col_start::job.col_step], st


def _gpu_lattice(tris, job, flags=0, nsplit=1, ngpus=1, split_mode=0):
    """render only the lattice pixels through the device API (compact output)"""
    import torch
    import raytracingc_b200 as rtc
    sc = rtc.DeviceScene(tris, device=0)
    dj = rtc.DeviceJob(job.width, job.height, job.spp, job.max_bounce, triangles_only=job.triangles_only,
                       nsplit=nsplit, subset=(job.row_start, job.row_step, job.col_start, job.col_step, 1), flags=flags)
    dj.to_c()
    dev = torch.device("cuda", 0)
    rgb = torch.zeros((dj.out_h, dj.out_w, 3), dtype=torch.uint8, device=dev)
    ids = torch.full((dj.out_h, dj.out_w), -7, dtype=torch.int32, device=dev)
    stats = torch.zeros(24, dtype=torch.int64, device=dev)
    acc = torch.zeros((max(nsplit, 1), dj.out_h, dj.out_w, 3), dtype=torch.float32, device=dev)
    cam, sky = rtc.camera_look_at(job.origin, job.look_at, job.fov), rtc.make_scene()
    st = torch.cuda.current_stream().cuda_stream
    sc.render(dj, sky, cam, d_rgb=rgb.data_ptr(), d_accum=acc.data_ptr() if nsplit > 1 else 0, d_ids=ids.data_ptr(),
              d_stats=stats.data_ptr(), stream=st)
    if nsplit > 1:
        rtc.api.sum_quantize(0, acc.data_ptr(), nsplit, dj.out_h * dj.out_w * 3, d_rgb=rgb.data_ptr(), stream=st)
    torch.cuda.synchronize()
    sc.close()
    return rgb.cpu().numpy(), ids.cpu().numpy(), stats.cpu().numpy()


def test_generator_is_deterministic_and_in_range():
    from raytracingc_b200 import api
    a, b = api.synth_scene(4096), api.synth_scene(4096)
    assert a.tobytes() == b.tobytes()
    assert a.tobytes() != api.synth_scene(4096, seed=1).tobytes()
    assert (a["posA"][:, 0] >= -8).all() and (a["posA"][:, 0] <= 8).all() and (a["posA"][:, 2] >= -2).all()
    assert np.allclose(np.linalg.norm(a["normal"], axis=1), 1, atol=1e-6)
    assert set(np.unique(a["emissionStrength"])) <= {0.0, 4.0}
    assert api.synth_scene(1 << 20)[:4096].tobytes() == a.tobytes()  # a prefix of the full scene


@pytest.mark.parametrize("ntris,step,spp,bounce", [(20000, 30, 4, 3), (1 << 20, 480, 2, 4)])
def test_lattice_matches_cpu_oracle(ntris, step, spp, bounce):
    from oracle.oracle_api import Job
    from raytracingc_b200 import api
    tris = api.synth_scene(ntris)
    W, H = (480, 270) if ntris < 100000 else (3840, 2160)
    job = Job(W, H, spp, bounce, row_start=step // 2, row_step=step, col_start=step // 2, col_step=step)
    want_rgb, want_ids, ost = _oracle_lattice(tris, job)
    for flags in (0, api.RTC_FLAG_PER_PAIR_FILTER, api.RTC_FLAG_PATHS_ONLY):
        rgb, ids, st = _gpu_lattice(tris, job, flags)
        assert ids.shape == want_ids.shape
        assert (ids == want_ids).all(), f"flags {flags}: {(ids != want_ids).sum()} ids differ"
        assert (rgb == want_rgb).all(), f"flags {flags}: {(rgb != want_rgb).any(axis=-1).sum()} pixels differ"
        assert int(st[0]) == ost["rays"] and int(st[9]) == ost["hit_pixels"]
        assert int(st[11]) == 0  # no sample disagreed with the classification
    assert (want_ids >= 0).sum() > 0  # the scene is in front of the camera


def test_dense_region_filtered_equals_exact():
    """a dense window of the 1 M-triangle frame (every 8th pixel of 3840x2160 is too slow; use 16th, few spp)"""
    from oracle.oracle_api import Job
    from raytracingc_b200 import api
    tris = api.synth_scene(1 << 17)
    job = Job(960, 540, 2, 3, row_start=3, row_step=24, col_start=5, col_step=24)
    ref = _gpu_lattice(tris, job, api.RTC_FLAG_EXACT_ONLY)
    for flags in (0, api.RTC_FLAG_PER_PAIR_FILTER, api.RTC_FLAG_PATHS_ONLY):
        got = _gpu_lattice(tris, job, flags)
        assert (got[1] == ref[1]).all() and (got[0] == ref[0]).all() and got[2][0] == ref[2][0]


def test_sample_split_planes_on_synthetic():
    """config 5's sample split: G sub-streams per pixel, ordered sum -- identical to the oracle's nsplit image"""
    from oracle.oracle_api import Job
    from raytracingc_b200 import api
    tris = api.synth_scene(20000)
    job = Job(480, 270, 8, 3, row_start=15, row_step=30, col_start=15, col_step=30, nsplit=4)
    want_rgb, want_ids, ost = _oracle_lattice(tris, job)
    rgb, ids, st = _gpu_lattice(tris, job, 0, nsplit=4)
    assert (ids == want_ids).all() and (rgb == want_rgb).all() and int(st[0]) == ost["rays"]
