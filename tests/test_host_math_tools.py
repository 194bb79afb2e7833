"""CPU: the two device headers that restate library arithmetic, compiled FOR THE HOST and checked against glibc.

csrc/rtc_powf.cuh (glibc 2.39's powf algorithm) and csrc/rtc_normal_fast.cuh (the Ziv-tested fast path of the normal
deviate) are written so that the same source compiles with g++ (tools/powf_host.cpp, tools/normal_fast_host.cpp give
the handful of intrinsics).  Here they run against this machine's powf / log / cos / sqrt -- the library the
reference links (raytracing.c:153,156; moremath.c:98-103) -- on 10^8 arguments in a few seconds.  The device side of
the same claims is tests/test_gpu_math.py."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _has_fma():
    try:
        with open("/proc/cpuinfo") as f:
            return " fma " in f.read().replace("\n", " ")
    except OSError:
        return False


pytestmark = [pytest.mark.skipif(shutil.which("g++") is None, reason="no g++"),
              pytest.mark.skipif(not _has_fma(), reason="glibc selects its FMA variants only on a CPU with FMA; "
                                                        "the tools are built with -mfma")]


def _build(tmp_path, name):
    exe = str(tmp_path / name)
    subprocess.run(["g++", "-O2", "-mfma", "-fopenmp", "-ffp-contract=off", "-o", exe,
                    os.path.join(ROOT, "tools", name + ".cpp")], check=True)
    return exe


def test_restated_powf_equals_glibc(tmp_path):
    r = subprocess.run([_build(tmp_path, "powf_host"), "200000000"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert " 0 differ" in r.stdout and "EXP2_TAB" not in r.stdout, r.stdout


def test_fast_normal_deviate_equals_the_reference_sequence_on_glibc(tmp_path):
    r = subprocess.run([_build(tmp_path, "normal_fast_host"), "27"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "(glibc) on 0;" in r.stdout, r.stdout
    declined = int(r.stdout.split("declined")[1].split()[0])
    assert 0 < declined < (1 << 27) * 1e-3  # it has to pay: about one deviate in 10^4 takes the exact sequence
