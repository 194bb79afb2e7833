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
    # glibc's powf is accurate to < 1 ULP but not correctly rounded.  Round 1 evaluated pow in double and narrowed
    # (correctly rounded): 255 of these 1 052 687 sky values differed from the reference's in the last bit.  The device
    # now restates glibc's own algorithm (csrc/rtc_powf.cuh, tools/powf_host.cpp): the tolerance is equality.
    bad = (got.view(np.uint32) != want.view(np.uint32)).any(axis=1)
    assert not bad.any(), f"{bad.sum()} of {len(dirs)} sky values differ, first for direction {dirs[bad][0]}"


@pytest.mark.parametrize("focus,intensity", [(0.5, 0.75), (3.0, 2.0), (300.0, 0.75), (5000.0, 10.0), (0.0, 1.0)])
def test_environment_light_other_sun_exponents(focus, intensity):
    """powf(d, sunFocus) over its whole range -- small and huge exponents (underflow to subnormals and to zero for d < 1,
    overflow for the d slightly above 1 that un-normalised bounce directions give), the zero exponent -- against the
    reference's function on glibc, bit for bit (NaNs: both NaN)."""
    from raytracingc_b200 import api
    rng = np.random.default_rng(int(focus * 10) + 1)
    sky = api.make_scene(sun_focus=focus, sun_intensity=intensity)
    sun = np.array([sky.normalizedSunDirection.x, sky.normalizedSunDirection.y, sky.normalizedSunDirection.z], dtype=np.float64)
    sun /= np.linalg.norm(sun)
    d = rng.normal(size=(1 << 17, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    near = sun + rng.normal(size=(1 << 17, 3)) * rng.choice([1e-1, 1e-2, 1e-3, 1e-4], size=(1 << 17, 1))
    near /= np.linalg.norm(near, axis=1, keepdims=True)
    long_ = near[:8192] * rng.uniform(1.0, 1.2, (8192, 1))  # d > 1: lerps of unit vectors can be longer than 1 too
    dirs = np.concatenate([d, near, -near[:8192], long_]).astype(np.float32)
    want = _oracle().environment_light_batch(dirs, sky)
    got = api.probe_environment_light(dirs, sky)
    same = (got.view(np.uint32) == want.view(np.uint32)) | (np.isnan(got) & np.isnan(want))
    bad = ~same.all(axis=1)
    assert not bad.any(), f"{bad.sum()} of {len(dirs)} sky values differ, first for direction {dirs[bad][0]}: {got[bad][0]} != {want[bad][0]}"


def test_normal_deviate_fast_path_equals_the_exact_sequence():
    """The Ziv-checked fast evaluation of RandomValueNormalDistrubtion (rtc_normal_fast.cuh) against the reference's own
    sequence on 2^27 generator states spread over the state space (tools/normal_selfcheck.py runs all 2^32:
    profiles/r2_normal_selfcheck.txt): never different, and it falls back rarely enough to pay."""
    from raytracingc_b200 import api
    tot = {"different": 0, "declined": 0, "states": 0}
    for first in (0, 0x12345678, 0x80000000, 0xfffff000):
        r = api.probe_normal_selfcheck(first, 1 << 25)
        for k in tot:
            tot[k] += r[k]
    assert tot["states"] == 4 << 25
    assert tot["different"] == 0
    assert 0 < tot["declined"] < tot["states"] * 1e-3, tot
