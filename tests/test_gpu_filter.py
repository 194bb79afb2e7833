"""-m gpu: the production kernel (conservative pre-filter + exact survivors) against the exact-only kernel.

The pre-filter may only reject pairs the reference's own arithmetic rejects, so both kernels must give
identical ids, pixels and ray counts on ANY scene -- checked here on random and adversarial geometry."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def random_scene(rng, n, extent, size, emit_frac=0.1):
    import raytracingc_b200 as rtc
    t = np.zeros(n, dtype=rtc.TRI_DTYPE)
    c = rng.uniform(-extent, extent, (n, 3)).astype(np.float32)
    c[:, 2] = rng.uniform(-2, 14, n)
    e1 = rng.uniform(-size, size, (n, 3)).astype(np.float32)
    e2 = rng.uniform(-size, size, (n, 3)).astype(np.float32)
    t["posA"], t["posB"], t["posC"] = c, c + e1, c + e2
    nrm = np.cross(t["posB"] - t["posA"], t["posC"] - t["posA"]).astype(np.float32)
    ln = np.linalg.norm(nrm, axis=1, keepdims=True).astype(np.float32)
    ln[ln == 0] = 1
    t["normal"] = nrm / ln
    # some normals deliberately inconsistent with the winding (OBJ vertex normals are)
    flip = rng.random(n) < 0.2
    t["normal"][flip] *= -1
    t["color"] = rng.uniform(0.2, 1.0, (n, 3))
    t["emissionStrength"] = np.where(rng.random(n) < emit_frac, 4.0, 0.0)
    t["smoothness"] = rng.uniform(0, 1, n)
    return t


CASES = [
    # n, extent, size, W, H, spp, bounce, origin, look_at, fov
    (200, 6.0, 2.0, 256, 128, 4, 3, (-4.75, -1.5, -4.75), (0.9, -1.2, 1.0), 1.0),
    (1000, 8.0, 0.3, 320, 200, 2, 4, (0.0, 0.0, -6.0), (0.0, 0.0, 1.0), 1.5),
    (50, 3.0, 30.0, 200, 120, 4, 5, (1.0, 2.0, -3.0), (0.0, 0.0, 4.0), 0.7),   # huge triangles, camera inside
    (300, 5.0, 1.0, 129, 67, 3, 2, (0.0, -1.0, 5.0), (0.0, 0.0, 0.0), 2.5),    # odd sizes, camera among triangles
    (64, 4.0, 1e-3, 160, 90, 2, 3, (-4.75, -1.5, -4.75), (0.9, -1.2, 1.0), 1.0),  # sub-pixel slivers (|det| < EPS)
]


@pytest.mark.parametrize("case", range(len(CASES)))
def test_fast_equals_exact_random(case):
    import raytracingc_b200 as rtc
    from raytracingc_b200 import api
    n, extent, size, W, H, spp, bounce, origin, look_at, fov = CASES[case]
    rng = np.random.default_rng(1234 + case)
    tris = random_scene(rng, n, extent, size)
    cam = rtc.camera_look_at(origin, look_at, fov)
    for tonly in (1, 0):
        kw = dict(width=W, height=H, spp=spp, max_bounce=bounce, triangles_only=tonly, camera=cam, want_ids=True)
        rgb_e, ids_e, st_e = rtc.render(tris, flags=api.RTC_FLAG_EXACT_ONLY, **kw)
        for fl in (0, api.RTC_FLAG_PER_PAIR_FILTER, api.RTC_FLAG_PATHS_ONLY):  # block filter, per-pair filter, none
            rgb_f, ids_f, st_f = rtc.render(tris, flags=fl, **kw)
            assert (ids_f == ids_e).all(), f"case {case} flags {fl}: {(ids_f != ids_e).sum()} ids differ"
            assert st_f["rays"] == st_e["rays"]
            assert (rgb_f == rgb_e).all(), f"case {case} flags {fl}: {(rgb_f != rgb_e).any(axis=-1).sum()} pixels differ"


def test_pixel_centre_edges():
    """Triangle edges that pass (to rounding) through pixel centres: vertices placed ON primary rays."""
    import raytracingc_b200 as rtc
    from raytracingc_b200 import api
    W, H = 96, 64
    cam = rtc.camera_look_at()
    o = np.array([cam.origin.x, cam.origin.y, cam.origin.z], np.float32)
    ex = np.array([cam.ex.x, cam.ex.y, cam.ex.z], np.float32)
    ey = np.array([cam.ey.x, cam.ey.y, cam.ey.z], np.float32)
    ez = np.array([cam.ez.x, cam.ez.y, cam.ez.z], np.float32)

    def on_ray(x, y, t):
        dx = np.float32((x - W // 2)) / np.float32(H // 2)
        dy = np.float32((y - H // 2)) / np.float32(H // 2)
        d = ex * dx + ey * dy + ez
        return (o + np.float32(t) * d / np.linalg.norm(d)).astype(np.float32)

    rng = np.random.default_rng(7)
    n = 120
    tris = np.zeros(n, dtype=rtc.TRI_DTYPE)
    for i in range(n):
        px = rng.integers(0, W, 3)
        py = rng.integers(0, H, 3)
        tt = rng.uniform(3, 9, 3)
        A, B, C = (on_ray(px[k], py[k], tt[k]) for k in range(3))
        nrm = np.cross(B - A, C - A)
        ln = np.linalg.norm(nrm)
        if ln == 0:
            nrm, ln = np.array([0, 0, 1.0]), 1.0
        nrm = nrm / ln
        if np.dot(nrm, A - o) > 0:
            nrm = -nrm
        tris[i] = (A, B, C, nrm.astype(np.float32), (0.8, 0.7, 0.6), 0.0, 0.3)
    kw = dict(width=W, height=H, spp=3, max_bounce=3, camera=cam, want_ids=True)
    rgb_e, ids_e, st_e = rtc.render(tris, flags=api.RTC_FLAG_EXACT_ONLY, **kw)
    for fl in (0, api.RTC_FLAG_PER_PAIR_FILTER):
        rgb_f, ids_f, st_f = rtc.render(tris, flags=fl, **kw)
        assert (ids_f == ids_e).all() and (rgb_f == rgb_e).all() and st_f["rays"] == st_e["rays"]
    assert (ids_e >= 0).sum() > 100


def test_sphere_block_bound_adversarial():
    """spheres decided by the block bound of the discriminant: many spheres, camera inside one, one behind the
    camera, sub-pixel and grazing ones -- block kernel == per-pair kernel == exact-only kernel, bit for bit"""
    import raytracingc_b200 as rtc
    from raytracingc_b200 import api
    rng = np.random.default_rng(99)
    tris = random_scene(rng, 60, 5.0, 1.5)
    sph = np.zeros(9, dtype=rtc.SPHERE_DTYPE)
    sph["pos"] = [(0, 1, 0), (-4.75, -1.5, -4.75), (-12, -3, -12), (2, -1, 3), (1.5, -1.2, 2.0), (0.9, -1.2, 1.0),
                  (3, 0, 8), (-2, -2, 1), (40, 0, 40)]
    sph["r"] = [2.5, 0.5, 3.0, 0.004, 0.8, 0.05, 4.0, 1e-4, 30.0]
    sph["color"] = rng.uniform(0.3, 1.0, (9, 3))
    sph["emissionStrength"] = [0, 0, 0, 5, 0, 2, 0, 0, 0]
    sph["smoothness"] = rng.uniform(0, 1, 9)
    for (W, H, origin, look, fov) in [(192, 108, (-4.75, -1.5, -4.75), (0.9, -1.2, 1.0), 1.0),   # inside sphere 1
                                      (160, 120, (-7.0, -2.0, -7.0), (0.9, -1.2, 1.0), 1.3),
                                      (129, 67, (0.0, -6.0, 0.5), (0.0, 1.0, 0.0), 0.6)]:
        cam = rtc.camera_look_at(origin, look, fov)
        kw = dict(width=W, height=H, spp=3, max_bounce=3, triangles_only=0, camera=cam, want_ids=True, spheres=sph)
        rgb_e, ids_e, st_e = rtc.render(tris, flags=api.RTC_FLAG_EXACT_ONLY, **kw)
        for fl in (0, api.RTC_FLAG_PER_PAIR_FILTER, api.RTC_FLAG_PATHS_ONLY):
            rgb_f, ids_f, st_f = rtc.render(tris, flags=fl, **kw)
            assert (ids_f == ids_e).all(), f"flags {fl}: {(ids_f != ids_e).sum()} ids differ"
            assert st_f["rays"] == st_e["rays"] and (rgb_f == rgb_e).all()
        assert (ids_e < -1).sum() > 0
