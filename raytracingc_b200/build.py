"""Build librtc_b200.so (CUDA, sm_100a) and the host C program in-tree.

nvcc cross-compiles for sm_100a without a GPU, so this runs in the build
container; the built .so travels to the GPU box with the repository snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "librtc_b200.so")
HOST_DIR = os.path.join(ROOT, "host")
HOST_BIN = os.path.join(HOST_DIR, "rayt_b200")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    # host code must not contract a*b+c either (camera basis, rtc_normalize)
    "-Xcompiler", "-fPIC,-ffp-contract=off",
    "-shared", "-I" + os.path.join(ROOT, "include"), "-I" + CSRC,
]


def _nvcc() -> str:
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: cannot build librtc_b200.so")
    return nvcc


def _stale(target: str, sources: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources if os.path.exists(s))


def lib_sources() -> list[str]:
    src = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cu", ".cuh"))]
    src += [os.path.join(ROOT, "include", f) for f in ("rtc_b200.h", "rtc_types.h")]
    return src


def build_lib(force: bool = False, verbose: bool = False) -> str:
    if force or _stale(LIB, lib_sources()):
        cmd = [_nvcc(), *NVCC_FLAGS, "-o", LIB, os.path.join(CSRC, "rtc_api.cu"), "-ldl"]
        if verbose:
            cmd += ["-Xptxas", "-v"]
            print(" ".join(cmd), file=sys.stderr)
        subprocess.run(cmd, check=True)
    return LIB


def build_variant(name: str, defines: list[str], verbose: bool = False) -> str:
    """An alternative build of the library with extra -D flags (A/B measurements: run with RTC_B200_LIB=<path>)."""
    out = os.path.join(PKG, f"librtc_b200_{name}.so")
    cmd = [_nvcc(), *NVCC_FLAGS, *[f"-D{d}" for d in defines], "-o", out, os.path.join(CSRC, "rtc_api.cu"), "-ldl"]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    subprocess.run(cmd, check=True)
    return out


def build_host(force: bool = False) -> str | None:
    """The host C program (CLI + OBJ loader + image writer) that links librtc_b200.so."""
    mk = os.path.join(HOST_DIR, "Makefile")
    if not os.path.exists(mk):
        return None
    subprocess.run(["make", "-s", "-C", HOST_DIR] + (["-B"] if force else []), check=True)
    return HOST_BIN


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_lib(force, verbose)
    build_host(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
