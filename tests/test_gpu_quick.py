"""-m gpu: the path kernel's conservative quick test (quick records, rtc_paths.cuh) may only drop (ray, triangle)
pairs the reference's arithmetic rejects.

RTC_FLAG_VERIFY_QUICK makes the kernel run the reference's exact test on EVERY pair and count accepted pairs whose
quick-test bit was clear (RenderStats.check_failures, must be 0); the images must also equal those of the
exact-only kernel and of the reference-arithmetic quick test (RTC_FLAG_REF_QUICK) bit for bit."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["packed", "scalar"])
def quick_layout(request, monkeypatch):
    """Both evaluations of the quick records: two triangles at a time with packed FP32 (default), one at a time."""
    monkeypatch.setenv("RTC_QUICK_FORM", {"packed": "3", "scalar": "1"}[request.param])
    return request.param


def make_scene(rng, n, extent, size, normals="exact", offset=(0.0, 0.0, 6.0)):
    import raytracingc_b200 as rtc
    t = np.zeros(n, dtype=rtc.TRI_DTYPE)
    c = (rng.uniform(-extent, extent, (n, 3)) + np.array(offset)).astype(np.float32)
    e1 = rng.uniform(-size, size, (n, 3)).astype(np.float32)
    e2 = rng.uniform(-size, size, (n, 3)).astype(np.float32)
    t["posA"], t["posB"], t["posC"] = c, c + e1, c + e2
    nrm = np.cross((t["posB"] - t["posA"]).astype(np.float64), (t["posC"] - t["posA"]).astype(np.float64))
    ln = np.linalg.norm(nrm, axis=1, keepdims=True)
    ln[ln == 0] = 1
    nrm = nrm / ln
    if normals == "rounded":      # OBJ text: 4 decimals
        nrm = np.round(nrm, 4)
    elif normals == "smooth":     # vertex normals: well off the face normal
        nrm = nrm + rng.normal(0, 0.3, nrm.shape)
        nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    elif normals == "flipped":    # against the winding
        nrm = -nrm
    elif normals == "mixed":
        k = rng.integers(0, 4, n)
        nrm[k == 1] *= -1
        nrm[k == 2] += rng.normal(0, 0.2, (int((k == 2).sum()), 3))
        nrm[k == 3] *= rng.uniform(0.1, 3.0, (int((k == 3).sum()), 1))  # not unit length
    t["normal"] = nrm.astype(np.float32)
    t["color"] = rng.uniform(0.3, 1.0, (n, 3))
    t["emissionStrength"] = np.where(rng.random(n) < 0.1, 4.0, 0.0)
    t["smoothness"] = rng.uniform(0, 1, n)
    return t


SCENES = [
    # name, n, extent, size, normals, offset
    ("plain", 300, 5.0, 1.5, "exact", (0, 0, 6)),
    ("rounded", 300, 5.0, 1.5, "rounded", (0, 0, 6)),
    ("smooth", 200, 4.0, 2.0, "smooth", (0, 0, 6)),
    ("flipped", 200, 4.0, 2.0, "flipped", (0, 0, 6)),
    ("mixed", 400, 6.0, 1.0, "mixed", (0, 0, 6)),
    ("huge", 60, 3.0, 60.0, "exact", (0, 0, 6)),            # edges ~100: wide rounding margins
    ("far", 300, 5.0, 1.0, "rounded", (900.0, -700.0, 800.0)),  # far from the origin: |s| ~ 1e3
    ("slivers", 400, 4.0, 2e-2, "exact", (0, 0, 6)),        # dets around EPS
    ("tiny", 400, 3.0, 1e-3, "rounded", (0, 0, 6)),         # everything below EPS
    ("dense", 1500, 2.0, 0.5, "exact", (0, 0, 5)),          # many hits per ray, long bounce chains
    ("odd", 333, 5.0, 1.5, "mixed", (0, 0, 6)),             # odd count: the last pair record is completed by a pad
    ("one", 1, 0.5, 3.0, "exact", (0, 0, 6)),
]


@pytest.mark.parametrize("scene", range(len(SCENES)))
def test_quick_never_drops_a_hit(scene):
    import raytracingc_b200 as rtc
    from raytracingc_b200 import api
    name, n, extent, size, normals, offset = SCENES[scene]
    rng = np.random.default_rng(500 + scene)
    tris = make_scene(rng, n, extent, size, normals, offset)
    look = np.array(offset, np.float64)
    cam = rtc.camera_look_at(tuple(look + np.array([0.5, -0.3, -extent - 4.0])), tuple(look), 1.2)
    kw = dict(width=160, height=96, spp=3, max_bounce=5, triangles_only=1, camera=cam, want_ids=True)
    rgb_e, ids_e, st_e = rtc.render(tris, flags=api.RTC_FLAG_EXACT_ONLY, **kw)
    for base in (0, api.RTC_FLAG_PATHS_ONLY):  # bounce rays only / primary rays (camera origin) too
        rgb_v, ids_v, st_v = rtc.render(tris, flags=base | api.RTC_FLAG_VERIFY_QUICK, **kw)
        assert st_v["check_failures"] == 0, f"{name}: {st_v['check_failures']} accepted pairs were dropped"
        assert st_v["rays"] == st_e["rays"]
        assert (ids_v == ids_e).all() and (rgb_v == rgb_e).all(), name
        for fl in (0, api.RTC_FLAG_REF_QUICK):
            rgb_f, ids_f, st_f = rtc.render(tris, flags=base | fl, **kw)
            assert st_f["rays"] == st_e["rays"] and st_f["check_failures"] == 0
            assert (ids_f == ids_e).all() and (rgb_f == rgb_e).all(), f"{name} flags {base | fl}"


def test_quick_degenerate_records():
    """Zero-area triangles, zero / NaN / huge normals, NaN and infinite vertices: records fall back to 'always keep'."""
    import raytracingc_b200 as rtc
    from raytracingc_b200 import api
    rng = np.random.default_rng(77)
    tris = make_scene(rng, 64, 3.0, 1.5, "exact")
    tris["posC"][0] = tris["posB"][0]                       # zero area
    tris["posB"][1] = tris["posA"][1]
    tris["normal"][2] = 0.0
    tris["normal"][3] = np.nan
    tris["normal"][4] *= 1e20
    tris["posA"][5, 0] = np.nan
    tris["posB"][6, 1] = np.inf
    tris["posC"][7] = tris["posA"][7] + np.float32(1e-30)   # denormal edges
    tris["posA"][8] *= np.float32(1e18)
    cam = rtc.camera_look_at((0.5, -0.3, -4.0), (0, 0, 6), 1.2)
    kw = dict(width=128, height=80, spp=2, max_bounce=4, triangles_only=1, camera=cam, want_ids=True)
    rgb_e, ids_e, st_e = rtc.render(tris, flags=api.RTC_FLAG_EXACT_ONLY, **kw)
    for base in (0, api.RTC_FLAG_PATHS_ONLY):
        rgb_v, ids_v, st_v = rtc.render(tris, flags=base | api.RTC_FLAG_VERIFY_QUICK, **kw)
        assert st_v["check_failures"] == 0
        assert (ids_v == ids_e).all() and (rgb_v.view(np.uint8) == rgb_e).all()


def test_quick_with_spheres_and_bundled_scene():
    """The bundled C2 scene (normals from the OBJ text) with the default sphere: verify + equality."""
    import os
    import raytracingc_b200 as rtc
    from raytracingc_b200 import api
    path = os.path.join(os.path.dirname(__file__), "golden", "scenes", "suze.tris")
    tris = np.fromfile(path, dtype=rtc.TRI_DTYPE)
    kw = dict(width=320, height=180, spp=4, max_bounce=3, triangles_only=0, want_ids=True)
    rgb_e, ids_e, st_e = rtc.render(tris, flags=api.RTC_FLAG_EXACT_ONLY, **kw)
    rgb_v, ids_v, st_v = rtc.render(tris, flags=api.RTC_FLAG_VERIFY_QUICK, **kw)
    assert st_v["check_failures"] == 0 and st_v["rays"] == st_e["rays"]
    assert (ids_v == ids_e).all() and (rgb_v == rgb_e).all()
    rgb_f, ids_f, st_f = rtc.render(tris, **kw)
    assert (rgb_f == rgb_e).all()


@pytest.mark.parametrize("height", [128, 135, 200, 333])
def test_pipelined_copy_matches_plain_copy(height, monkeypatch):
    """render_cuda copies the image to the host in row chunks behind the kernels (per-chunk completion flags);
    the result must equal the plain copy at the end of the frame, for heights that do not divide into the chunks,
    with pageable and with page-locked host buffers."""
    import os
    import torch
    import raytracingc_b200 as rtc
    path = os.path.join(os.path.dirname(__file__), "golden", "scenes", "ultracomplex.tris")
    tris = np.fromfile(path, dtype=rtc.TRI_DTYPE)
    kw = dict(width=257, height=height, spp=5, max_bounce=3, triangles_only=1)
    monkeypatch.setenv("RTC_PIPELINE_MIN_BYTES", "0")  # by default only frames of 4 MB and more are pipelined
    monkeypatch.setenv("RTC_NO_PIPELINE", "1")
    ref, _, st0 = rtc.render(tris, **kw)
    monkeypatch.delenv("RTC_NO_PIPELINE")
    out, _, st1 = rtc.render(tris, **kw)
    assert st0["rays"] == st1["rays"]
    assert (out == ref).all()
    pinned = torch.empty((height, 257, 3), dtype=torch.uint8).pin_memory()
    pinned.fill_(7)
    out2, _, _ = rtc.render(tris, out=pinned.numpy(), **kw)
    assert (pinned.numpy() == ref).all()
