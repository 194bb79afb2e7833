"""TEST INFRASTRUCTURE -- ctypes access to the two CPU checkers in oracle/.

Only tests/, ``__graft_entry__.smoke()`` and bench.py's CPU-baseline legs may
import this module; the product package (raytracingc_b200/) never does.

``Oracle("port")``       oracle/liboracle.so           plain-C restatement
``Oracle("reference")``  oracle/_ref/libref_oracle.so  the reference's own functions
                         (compiled here from /root/reference; prebuilt on the GPU box)

Both expose the same calls, so a test can run one against the other.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_LIB = os.path.join(HERE, "liboracle.so")
REF_LIB = os.path.join(HERE, "_ref", "libref_oracle.so")
REFERENCE_ROOT = "/root/reference"

TRI_DTYPE = np.dtype(
    [("posA", "<f4", 3), ("posB", "<f4", 3), ("posC", "<f4", 3), ("normal", "<f4", 3),
     ("color", "<f4", 3), ("emissionStrength", "<f4"), ("smoothness", "<f4")])
SPHERE_DTYPE = np.dtype(
    [("pos", "<f4", 3), ("r", "<f4"), ("color", "<f4", 3), ("emissionStrength", "<f4"), ("smoothness", "<f4")])
assert TRI_DTYPE.itemsize == 68 and SPHERE_DTYPE.itemsize == 36


class Vec3(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float)]

    def tup(self):
        return (self.x, self.y, self.z)


def _v(t) -> Vec3:
    return Vec3(*[float(np.float32(c)) for c in t])


class SceneC(C.Structure):
    _fields_ = [("sun", Vec3), ("horizon", Vec3), ("zenith", Vec3), ("ground", Vec3),
                ("sunFocus", C.c_float), ("sunIntensity", C.c_float)]


class CameraC(C.Structure):
    _fields_ = [("origin", Vec3), ("ex", Vec3), ("ey", Vec3), ("ez", Vec3), ("fov", C.c_float)]


class JobC(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("spp", C.c_int32), ("maxBounce", C.c_int32),
                ("trianglesOnly", C.c_int32), ("row_start", C.c_int32), ("row_step", C.c_int32),
                ("col_start", C.c_int32), ("col_step", C.c_int32), ("nsplit", C.c_int32),
                ("scene", SceneC), ("cam", CameraC)]


class StatsC(C.Structure):
    _fields_ = [("pixels", C.c_uint64), ("hit_pixels", C.c_uint64), ("rays", C.c_uint64),
                ("sphere_tests", C.c_uint64), ("tri_exit", C.c_uint64 * 6), ("rng_draws", C.c_uint64),
                ("sky_evals", C.c_uint64)]

    def as_dict(self):
        return {"pixels": self.pixels, "hit_pixels": self.hit_pixels, "rays": self.rays,
                "sphere_tests": self.sphere_tests, "tri_exit": list(self.tri_exit),
                "rng_draws": self.rng_draws, "sky_evals": self.sky_evals}


# defaults of the reference program: main.c:14,21-28 (scene) and main.c:114-116 (camera)
DEFAULT_SUN = (-30.0, -85.0, 100.0)
DEFAULT_ORIGIN = (-4.75, -1.5, -4.75)
DEFAULT_LOOK_AT = (0.9, -1.2, 1.0)
# the one sphere the reference hard-codes, scene.h:17-19
DEFAULT_SPHERES = np.array([((0, 1, 0), 2.5, (1, 1, 1), 0, 0)], dtype=SPHERE_DTYPE)


@dataclass
class Job:
    width: int
    height: int
    spp: int
    max_bounce: int
    triangles_only: int = 1
    row_start: int = 0
    row_step: int = 1
    col_start: int = 0
    col_step: int = 1
    nsplit: int = 1
    origin: tuple = DEFAULT_ORIGIN
    look_at: tuple = DEFAULT_LOOK_AT
    fov: float = 1.0
    sun: tuple = DEFAULT_SUN
    sun_focus: float = 22.0
    sun_intensity: float = 0.75
    horizon: tuple = (1.0, 1.0, 1.0)
    zenith: tuple = (0.263, 0.969, 0.871)
    ground: tuple = (0.66, 0.66, 0.66)


def build(ref: bool = True) -> None:
    """(Re)build the checkers; building the checker is not using it."""
    subprocess.run(["make", "-s", "-C", HERE, "all" if ref else os.path.join(HERE, "liboracle.so")], check=True)


def have_reference_lib() -> bool:
    return os.path.exists(REF_LIB)


class Oracle:
    def __init__(self, kind: str = "port"):
        assert kind in ("port", "reference")
        self.kind = kind
        path = PORT_LIB if kind == "port" else REF_LIB
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} missing: run `make -C oracle`")
        self.lib = C.CDLL(path)
        self.p = "oracle_" if kind == "port" else "ref_"
        self._ntris = 0
        f = self._fn
        f("render").argtypes = [C.POINTER(JobC), C.c_void_p, C.c_void_p, C.POINTER(StatsC)]
        f("render").restype = None
        f("camera_look_at").argtypes = [C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_float, C.POINTER(CameraC)]
        f("camera_look_at").restype = None
        f("normalize").argtypes = [C.POINTER(C.c_float), C.POINTER(C.c_float)]
        f("random_value").argtypes = [C.POINTER(C.c_uint32)]
        f("random_value").restype = C.c_float
        f("random_direction").argtypes = [C.POINTER(C.c_uint32), C.POINTER(C.c_float)]
        f("environment_light").argtypes = [C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(SceneC),
                                           C.POINTER(C.c_float)]
        f("ray_triangle").argtypes = [C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_int, C.POINTER(C.c_float)]
        f("ray_triangle").restype = C.c_int
        f("ray_sphere").argtypes = [C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_int, C.POINTER(C.c_float),
                                    C.POINTER(C.c_float)]
        f("ray_sphere").restype = C.c_int
        f("float_to_uint").argtypes = [C.c_float]
        f("float_to_uint").restype = C.c_uint8
        f("set_triangles").argtypes = [C.c_void_p, C.c_int]
        if kind == "port":
            f("set_spheres").argtypes = [C.c_void_p, C.c_int]
            self.set_spheres(DEFAULT_SPHERES)
        else:
            f("load_obj").argtypes = [C.c_char_p]
            f("load_obj").restype = C.c_int
            f("load_triangles_txt").argtypes = [C.c_char_p]
            f("load_triangles_txt").restype = C.c_int
            f("get_triangles").argtypes = [C.c_void_p, C.c_int]
            f("get_triangles").restype = C.c_int
            f("get_spheres").argtypes = [C.c_void_p, C.c_int]
            f("get_spheres").restype = C.c_int
            f("triangle_count").restype = C.c_int
            f("sphere_count").restype = C.c_int

    def _fn(self, name):
        return getattr(self.lib, self.p + name)

    # ---- scene -----------------------------------------------------------
    def set_triangles(self, tris: np.ndarray) -> None:
        tris = np.ascontiguousarray(tris, dtype=TRI_DTYPE)
        self._fn("set_triangles")(tris.ctypes.data, len(tris))
        self._ntris = len(tris)

    def set_spheres(self, sph: np.ndarray) -> None:
        if self.kind != "port":
            raise RuntimeError("the reference hard-codes its sphere list (scene.h:17-19)")
        sph = np.ascontiguousarray(sph, dtype=SPHERE_DTYPE)
        self._fn("set_spheres")(sph.ctypes.data, len(sph))

    def load_obj(self, path: str) -> np.ndarray:
        """reference only: loadOBJTriangles (raytracing.c:100) -> Triangle[]"""
        n = self._fn("load_obj")(path.encode())
        if n < 0:
            raise FileNotFoundError(path)
        return self.get_triangles()

    def load_triangles_txt(self, path: str) -> np.ndarray:
        """reference only: parseTriangleFile (raytracing.c:76); writes <path>.parsed"""
        n = self._fn("load_triangles_txt")(path.encode())
        if n <= 0:
            raise FileNotFoundError(path)
        return self.get_triangles()

    def get_triangles(self) -> np.ndarray:
        n = self._fn("triangle_count")()
        out = np.zeros(n, dtype=TRI_DTYPE)
        self._fn("get_triangles")(out.ctypes.data, n)
        self._ntris = n
        return out

    def get_spheres(self) -> np.ndarray:
        n = self._fn("sphere_count")()
        out = np.zeros(n, dtype=SPHERE_DTYPE)
        self._fn("get_spheres")(out.ctypes.data, n)
        return out

    # ---- camera / scene params ------------------------------------------
    def camera(self, origin, look_at, fov) -> CameraC:
        cam = CameraC()
        o = (C.c_float * 3)(*origin)
        t = (C.c_float * 3)(*look_at)
        self._fn("camera_look_at")(o, t, C.c_float(fov), C.byref(cam))
        return cam

    def normalize(self, v):
        i = (C.c_float * 3)(*v)
        o = (C.c_float * 3)()
        self._fn("normalize")(i, o)
        return tuple(o)

    def job_struct(self, job: Job) -> JobC:
        j = JobC()
        j.width, j.height, j.spp, j.maxBounce = job.width, job.height, job.spp, job.max_bounce
        j.trianglesOnly = job.triangles_only
        j.row_start, j.row_step, j.col_start, j.col_step = job.row_start, job.row_step, job.col_start, job.col_step
        j.nsplit = job.nsplit
        j.scene.sun = _v(self.normalize(job.sun))
        j.scene.horizon, j.scene.zenith, j.scene.ground = _v(job.horizon), _v(job.zenith), _v(job.ground)
        j.scene.sunFocus, j.scene.sunIntensity = job.sun_focus, job.sun_intensity
        j.cam = self.camera(job.origin, job.look_at, job.fov)
        return j

    # ---- render ----------------------------------------------------------
    def render(self, job: Job, want_ids: bool = True, out=None):
        """-> (rgb uint8 [H,W,3], ids int32 [H,W] or None, stats dict)."""
        j = self.job_struct(job)
        if out is None:
            rgb = np.zeros((job.height, job.width, 3), dtype=np.uint8)
            ids = np.full((job.height, job.width), -9, dtype=np.int32) if want_ids else None
        else:
            rgb, ids = out
        st = StatsC()
        self._fn("render")(C.byref(j), rgb.ctypes.data, ids.ctypes.data if ids is not None else None, C.byref(st))
        return rgb, ids, st.as_dict()

    # ---- probes ----------------------------------------------------------
    def random_value(self, state: int):
        s = C.c_uint32(state)
        v = self._fn("random_value")(C.byref(s))
        return v, s.value

    def random_direction(self, state: int):
        s = C.c_uint32(state)
        o = (C.c_float * 3)()
        self._fn("random_direction")(C.byref(s), o)
        return tuple(o), s.value

    def random_direction_batch(self, states: np.ndarray):
        """RandomDiretion on n independent states -> (float32 [n, 3], uint32 [n] states after)."""
        st = np.ascontiguousarray(states, dtype=np.uint32)
        out = np.empty((len(st), 3), dtype=np.float32)
        after = np.empty(len(st), dtype=np.uint32)
        f = self._fn("random_direction_batch")
        f.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        f.restype = None
        f(st.ctypes.data, len(st), out.ctypes.data, after.ctypes.data)
        return out, after

    def environment_light_batch(self, dirs: np.ndarray, scene) -> np.ndarray:
        """getEnvironmentLight on n directions; `scene` is any ctypes struct with the reference Scene's layout."""
        d = np.ascontiguousarray(dirs, dtype=np.float32).reshape(-1, 3)
        out = np.empty_like(d)
        f = self._fn("environment_light_batch")
        f.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        f.restype = None
        f(d.ctypes.data, len(d), C.addressof(scene), out.ctypes.data)
        return out

    def environment_light(self, direction, scene: SceneC):
        p = (C.c_float * 3)(0, 0, 0)
        d = (C.c_float * 3)(*direction)
        o = (C.c_float * 3)()
        self._fn("environment_light")(p, d, C.byref(scene), o)
        return tuple(o)

    def ray_triangle(self, pos, direction, i):
        p = (C.c_float * 3)(*pos)
        d = (C.c_float * 3)(*direction)
        dst = C.c_float()
        hit = self._fn("ray_triangle")(p, d, i, C.byref(dst))
        return hit, dst.value

    def ray_sphere(self, pos, direction, k):
        p = (C.c_float * 3)(*pos)
        d = (C.c_float * 3)(*direction)
        dst = C.c_float()
        n = (C.c_float * 3)()
        hit = self._fn("ray_sphere")(p, d, k, C.byref(dst), n)
        return hit, dst.value, tuple(n)

    def float_to_uint(self, f):
        return self._fn("float_to_uint")(C.c_float(f))


# ---- hashes used by the known-answer vectors ---------

def _hashlib():
    lib = C.CDLL(PORT_LIB)
    lib.oracle_fnv1a64_bytes.argtypes = [C.c_void_p, C.c_uint64]
    lib.oracle_fnv1a64_bytes.restype = C.c_uint64
    lib.oracle_fnv1a64_words.argtypes = [C.c_void_p, C.c_uint64]
    lib.oracle_fnv1a64_words.restype = C.c_uint64
    return lib


def fnv1a64_bytes(data) -> str:
    """FNV-1a-64 over the bytes of an array / bytes object."""
    a = np.ascontiguousarray(np.frombuffer(data, dtype=np.uint8) if isinstance(data, (bytes, bytearray)) else data)
    a = a.view(np.uint8).ravel()
    return f"{_hashlib().oracle_fnv1a64_bytes(a.ctypes.data, a.size):016x}"


def fnv1a64_words(words: np.ndarray) -> str:
    """The survey's id hash: xor each WHOLE uint32 id, then multiply."""
    a = np.ascontiguousarray(words).astype(np.int64).astype(np.uint32).ravel()
    return f"{_hashlib().oracle_fnv1a64_words(a.ctypes.data, a.size):016x}"
