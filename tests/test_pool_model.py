"""CPU: the executable model of the path kernel's pool scheduler (tools/pool_model.py, csrc/rtc_pool.cuh).

The pool form evaluates the samples of a pixel out of order: a sample's light is a pure function of the RNG state it
starts from (7 draws per hit-bounce, none for a miss: raytracing.c:272-286, moremath.c:89-108), every reachable start
state is the seed advanced by 7 j steps, and the pixel is the pointer walk j <- j + h_j summed in sample order.  The
model runs the kernel's consume / refill / stage steps on synthetic scripts and must reproduce the serial walk's
float sums, ray counts and final states for every configuration -- with and without speculation, one context or many,
more contexts than slots' worth of work or fewer -- and never stall."""
import importlib.util
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _model():
    spec = importlib.util.spec_from_file_location("pool_model", os.path.join(ROOT, "tools", "pool_model.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_pool_scheduler_model_equals_the_serial_walk(capsys):
    m = _model()
    m.selftest(trials=120, seed=7)
    assert "identical to the serial walk" in capsys.readouterr().out


def test_lcg_skip_ahead_is_one_multiply_add():
    """the kernel jumps 7 LCG steps at once (lcg_mul(7), lcg_add(7) in rtc_pool.cuh, built by the same recurrence):
    the composition of seven steps of moremath.c:91 is s * M + C for every state tried"""
    m = _model()
    mul, add = 1, 0
    for _ in range(7):
        mul, add = (mul * 747796405) & 0xFFFFFFFF, (add * 747796405 + 2891336453) & 0xFFFFFFFF
    for s in (0, 1, 12345, 0xFFFFFFFF, 0x80000000, 0x12345678):
        assert m.adv7(s) == (s * mul + add) & 0xFFFFFFFF
