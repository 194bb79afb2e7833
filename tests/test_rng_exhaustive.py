"""-m "not gpu": the GPU's division-free RandomValue conversion equals the reference's double division
(moremath.c:94) for ALL 2^32 hash outputs -- tools/verify_rng_div.c, run in 8 parallel slices (~5 s)."""
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_rng_division_identity_all_2_pow_32(tmp_path):
    exe = str(tmp_path / "verify_rng_div")
    subprocess.run(["gcc", "-O2", "-ffp-contract=off", os.path.join(ROOT, "tools", "verify_rng_div.c"), "-o", exe,
                    "-lm"], check=True)
    n = 8
    chunk = (1 << 32) // n

    def run(i):
        return subprocess.run([exe, str(i * chunk), str(chunk)], capture_output=True, text=True)

    with ThreadPoolExecutor(n) as ex:
        for r in ex.map(run, range(n)):
            assert r.returncode == 0 and "mismatches 0" in r.stdout, r.stdout
