"""-m gpu: the FP64 islands of the device code against the reference's own functions, at scale.

RandomDiretion (PCG + Box-Muller in double: log, cos, sqrt, narrowed where the reference's float locals narrow) and
getEnvironmentLight (smoothstep, two powf, with the device's exact shortcuts around them) are evaluated on millions
of inputs by the device probes of the C ABI and compared bit for bit with oracle/_ref (the reference's functions on
glibc) when it is there, else with the CPU restatement."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _oracle():
    from oracle import oracle_api as oa
    return oa.Oracle("reference" if oa.have_reference_lib() else "port")


def test_random_direction_bits_at_scale():
    from raytracingc_b200 import api
    rng = np.random.default_rng(2024)
    states = np.concatenate([np.arange(0, 1 << 18, dtype=np.uint32),                       # the seeds pixels start from
                             rng.integers(0, 1 << 32, 1 << 21, dtype=np.uint64).astype(np.uint32),
                             np.array([0xFFFFFFFF, 0x80000000, 1, 12345], dtype=np.uint32)])
    want, want_st = _oracle().random_direction_batch(states)
    got, got_st = api.probe_random_direction(states)
    assert (got_st == want_st).all()
    bad = (got.view(np.uint32) != want.view(np.uint32)).any(axis=1)
    assert not bad.any(), f"{bad.sum()} of {len(states)} directions differ, first at state {states[bad][0]:#x}"


def test_environment_light_bits_at_scale():
    from raytracingc_b200 import api
    rng = np.random.default_rng(7)
    d = rng.normal(size=(1 << 20, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = d.astype(np.float32)
    special = np.array([[0, 0, 1], [1, 0, 0], [0, 1, 0], [0, -1, 0], [0, -0.0, 1], [0, 1e-9, 1], [0, -1e-9, 1],
                        [0, -0.74, 0.67], [0, -0.7400001, 0.67], [0, -0.0099, 1], [0, -0.01, 1], [0.3, -0.5, 2.0],
                        [0, -1e-3, 0], [-0.28, -0.8, 0.94], [0.28, 0.8, -0.94]], dtype=np.float32)
    # not all unit length: bounce directions are lerps of unit vectors (raytracing.c:276)
    short = (d[:4096] * rng.uniform(0.2, 1.0, (4096, 1))).astype(np.float32)
    dirs = np.concatenate([d, special, short])
    sky = api.make_scene()
    want = _oracle().environment_light_batch(dirs, sky)
    got = api.probe_environment_light(dirs, sky)
    # glibc's powf is accurate to < 1 ULP but not correctly rounded; the device evaluates pow in double and narrows
    # (correctly rounded): the two differ in the last bit of a float for ~2.5e-4 of the directions that reach a powf
    # (measured: 255 of 1 052 687; one of them by 2 ulp: both powf of that direction differ and the two terms are
    # added), and never where no powf is involved (dir.y >= 0)
    ulp = np.abs(got.view(np.int32).astype(np.int64) - want.view(np.int32).astype(np.int64))
    assert ulp.max() <= 3, f"sky value off by {ulp.max()} ulp for direction {dirs[ulp.max(axis=1).argmax()]}"
    bad = (ulp != 0).any(axis=1)
    assert bad.mean() < 1e-3, f"{bad.sum()} of {len(dirs)} sky values differ in the last bit"
    assert not bad[dirs[:, 1] >= 0].any(), "a direction below the horizon line differs: no powf is involved there"
