"""-m "not gpu": the CPU checkers against the golden vectors, and the restatement against the reference.

The reference ships no tests for this path, so the pins are:
  * tests/golden/kat.json + images/: generated from the reference's OWN functions (oracle/_ref, compiled from
    /root/reference by oracle/Makefile) by tests/golden/make_golden.py;
  * SURVEY.md section 8c's hashes, obtained independently during the survey with a different driver.
Here the plain-C restatement (oracle/liboracle.so, "port") must reproduce all of them bit for bit, and -- when
oracle/_ref is present -- agree with the reference function by function.
"""
import numpy as np
import pytest

from oracle.oracle_api import Job, fnv1a64_bytes, fnv1a64_words
from tests import util

KAT = util.kat()
SMALL = sorted(k for k, v in KAT.items() if not k.startswith("_") and
               v["job"]["width"] * v["job"]["height"] <= 320 * 240 and v["job"]["spp"] <= 64)

# SURVEY.md section 8c table (image FNV, primary-id FNV, primary-hit pixels) -- independent of make_golden.py
SURVEY_8C = {
    "c1_simplest_320x240_1spp_1b": ("2d41b9713063a3a7", "4ab508f9f4fcabd9", 360),
    "suze_160x90_16spp_3b": ("9991ec05322ca540", "d8a5f5cce7f20281", 3156),
    "ultracomplex_160x90_16spp_4b": ("3af618e27c5b1ccd", "23f20496561bbfbe", 530),
    "default_64x64_16spp_4b": ("004d33e25eb3d070", "d7006780d317a01f", 4096),
    "c2_suze_1280x720_16spp_3b": ("a42d830a998d055a", "e7bc6dbbb41ab427", 203093),
    "c3_complex_1920x1080_64spp_4b": ("fdcf9eaffae2ae17", "e3bf15bd50704566", 75432),
    "c4_ultracomplex_3840x2160_64spp_4b": ("580c773a42e7955b", "630af82958b339e4", 307491),
    "default_640x480_16spp_4b": ("1b34559a244e00c5", "8d240c97b161e072", 307200),
    "simplest_64x48_4000spp_1b": ("5134a7dbc5bd16bb", "3b0274ad77881e0f", 20),
}


def test_kat_matches_survey_table():
    """the committed vectors are the survey's (two independent drivers of the reference's functions agree)"""
    for name, (img, ids, hits) in SURVEY_8C.items():
        e = KAT[name]
        assert (e["image_fnv"], e["id_fnv"], e["hit_pixels"]) == (img, ids, hits), name
        assert e["port_equals_reference"] is True


def test_scene_fixtures_have_the_reference_counts():
    for scene, n in KAT["_scenes"].items():
        t = util.scene_tris(scene)
        assert len(t) == n and t.dtype.itemsize == 68
    # Appendix B: simplest.obj is one triangle whose normal (0,1,0) is flipped to (0,-1,0) by raytracing.c:123-135
    t = util.scene_tris("simplest")
    assert tuple(t["normal"][0]) == (0.0, -1.0, 0.0)
    # triangles.txt: two emissive ceiling triangles (emission 10)
    d = util.scene_tris("default")
    assert (d["emissionStrength"] == 10).sum() == 2


def _check_against_golden(oracle, name):
    e = KAT[name]
    oracle.set_triangles(util.scene_tris(e["scene"]))
    rgb, ids, st = oracle.render(util.job_of(e))
    want_rgb, want_ids = util.golden_image(name)
    assert (ids == want_ids).all()
    assert (rgb == want_rgb).all(), f"{name}: {(rgb != want_rgb).any(axis=-1).sum()} pixels differ"
    assert fnv1a64_bytes(rgb) == e["image_fnv"]
    assert fnv1a64_words(ids) == e["id_fnv"]
    return st


@pytest.mark.parametrize("name", SMALL)
def test_port_reproduces_golden(port_oracle, name):
    e = KAT[name]
    st = _check_against_golden(port_oracle, name)
    assert st["rays"] == e["rays"] and st["hit_pixels"] == e["hit_pixels"]
    assert st["tri_exit"] == e["tri_exit"] and st["rng_draws"] == e["rng_draws"]
    assert st["sky_evals"] == e["sky_evals"] and st["sphere_tests"] == e["sphere_tests"]
    assert sum(st["tri_exit"]) == st["rays"] * KAT["_scenes"][e["scene"]]  # brute force: every ray tests every triangle


@pytest.mark.parametrize("name", ["suze_160x90_16spp_3b", "default_64x64_16spp_4b", "suze_nsplit4_96x64_16spp_3b",
                                  "suze_odd_101x67_5spp_2b"])
def test_reference_reproduces_golden(ref_oracle, name):
    _check_against_golden(ref_oracle, name)


def test_port_lattice_of_full_configs(port_oracle):
    """BASELINE.json's full-size configs on the committed 8x8 golden lattice (every pixel owns its seed, so a
    sparse subset is exact); sized to stay within seconds on the CPU."""
    for name in ["c2_suze_1280x720_16spp_3b", "default_640x480_16spp_4b"]:
        e = KAT[name]
        step, want_rgb, want_ids = util.golden_lattice(name)
        job = util.job_of(e)
        sub = 4 * step  # every 4th lattice pixel
        job.row_step = job.col_step = sub
        port_oracle.set_triangles(util.scene_tris(e["scene"]))
        rgb, ids, _ = port_oracle.render(job)
        assert (ids[::sub, ::sub] == want_ids[::4, ::4]).all()
        assert (rgb[::sub, ::sub] == want_rgb[::4, ::4]).all()


def test_subset_and_full_render_agree(port_oracle):
    """pixel independence (main.c:95 seeds per pixel): rows k, k+P, ... rendered separately equal the full frame"""
    e = KAT["suze_160x90_16spp_3b"]
    port_oracle.set_triangles(util.scene_tris(e["scene"]))
    want_rgb, _ = util.golden_image("suze_160x90_16spp_3b")
    job = util.job_of(e)
    out = np.zeros_like(want_rgb)
    rays = 0
    for k in range(3):
        job.row_start, job.row_step = k, 3
        rgb, _, st = port_oracle.render(job, want_ids=False)
        out[k::3] = rgb[k::3]
        rays += st["rays"]
    assert (out == want_rgb).all() and rays == e["rays"]


# ---- function-by-function: restatement vs the reference's own functions ---------------------------------------

def test_rng_stream(port_oracle, ref_oracle):
    rs = np.random.default_rng(5)
    states = [0, 1, 2, 0xFFFFFFFF, 0x80000000, 12345] + [int(x) for x in rs.integers(0, 2 ** 32, 2000)]
    for s in states:
        a, sa = port_oracle.random_value(s)
        b, sb = ref_oracle.random_value(s)
        assert sa == sb and np.float32(a).tobytes() == np.float32(b).tobytes()
        assert 0.0 <= a <= 1.0
    for s in states[:500]:
        (da, sa), (db, sb) = port_oracle.random_direction(s), ref_oracle.random_direction(s)
        assert sa == sb
        assert np.array(da, np.float32).tobytes() == np.array(db, np.float32).tobytes()


def test_rng_known_values(port_oracle):
    """moremath.c:90-95 worked by hand for state 0: s = 2891336453; hash as PCG-RXS-M-XS; / 4294967295.0"""
    s = 2891336453
    r = (((s >> ((s >> 28) + 4)) ^ s) * 277803737) & 0xFFFFFFFF
    r = (r >> 22) ^ r
    v, s2 = port_oracle.random_value(0)
    assert s2 == s and np.float32(v) == np.float32(r / 4294967295.0)


def test_environment_light(port_oracle, ref_oracle):
    sc = port_oracle.job_struct(Job(8, 8, 1, 1)).scene
    rs = np.random.default_rng(9)
    dirs = rs.normal(size=(400, 3)).astype(np.float32)
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    dirs = np.vstack([dirs, [[0, -1, 0], [0, 1, 0], [1, 0, 0], [0, -1e-4, 1], [0, 1e-4, 1]]]).astype(np.float32)
    for d in dirs:
        a = port_oracle.environment_light(tuple(map(float, d)), sc)
        b = ref_oracle.environment_light(tuple(map(float, d)), sc)
        assert np.array(a, np.float32).tobytes() == np.array(b, np.float32).tobytes(), d


def test_ray_triangle_and_sphere(port_oracle, ref_oracle):
    tris = util.scene_tris("ultracomplex")
    port_oracle.set_triangles(tris)
    ref_oracle.set_triangles(tris)
    rs = np.random.default_rng(11)
    o = (-4.75, -1.5, -4.75)
    nhit = 0
    for _ in range(300):
        tgt = rs.uniform(-2, 2, 3)
        d = np.array(tgt, np.float32) - np.array(o, np.float32)
        d = tuple(map(float, (d / np.linalg.norm(d)).astype(np.float32)))
        for i in range(0, len(tris), 7):
            (ha, da), (hb, db) = port_oracle.ray_triangle(o, d, i), ref_oracle.ray_triangle(o, d, i)
            assert ha == hb and (not ha or np.float32(da).tobytes() == np.float32(db).tobytes())
            nhit += ha
        sa, sb = port_oracle.ray_sphere(o, d, 0), ref_oracle.ray_sphere(o, d, 0)
        assert sa[0] == sb[0]
        if sa[0]:
            assert np.float32(sa[1]).tobytes() == np.float32(sb[1]).tobytes()
            assert np.array(sa[2], np.float32).tobytes() == np.array(sb[2], np.float32).tobytes()
    assert nhit > 0


def test_float_to_uint_edge_cases(port_oracle, ref_oracle):
    """moremath.c:25-30 and SURVEY H7: NaN -> 0, +inf -> 255, -inf -> 0, 0.999999 -> 254, 0.5 -> 127 (truncation)"""
    cases = {float("nan"): 0, float("inf"): 255, float("-inf"): 0, 0.999999: 254, 0.5: 127, 1.0: 255, 0.0: 0,
             -0.0: 0, 1e-9: 0, 0.00392157: 1, -3.0: 0, 7.5: 255}
    for f, want in cases.items():
        assert port_oracle.float_to_uint(f) == want, f
        assert ref_oracle.float_to_uint(f) == want, f


def test_camera_basis(port_oracle, ref_oracle):
    for origin, look, fov in [((-4.75, -1.5, -4.75), (0.9, -1.2, 1.0), 1.0), ((0, -2, -6), (0, -1, 0), 2.0),
                              ((3, 1, 2), (-1, 0.5, 7), 0.5)]:
        a, b = port_oracle.camera(origin, look, fov), ref_oracle.camera(origin, look, fov)
        assert bytes(a) == bytes(b)
