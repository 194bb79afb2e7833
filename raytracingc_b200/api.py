"""ctypes binding of librtc_b200.so (include/rtc_b200.h).

This is plumbing for tests and bench.py; the product is the C ABI itself.
There is no Python or CPU rendering path here: if the CUDA library is missing
or no GPU is visible, calls raise.

Reference citations are into /root/reference.
"""
from __future__ import annotations

import ctypes as C
import os
import sys
from dataclasses import dataclass, field

import numpy as np

PKG = os.path.dirname(os.path.abspath(__file__))
# RTC_B200_LIB: an alternative build of the same library (A/B measurements of compile-time variants)
LIB_PATH = os.environ.get("RTC_B200_LIB") or os.path.join(PKG, "librtc_b200.so")

# reference raytracing.h:41-45 (Triangle, 68 B) and :34-39 (Sphere, 36 B)
TRI_DTYPE = np.dtype(
    [("posA", "<f4", 3), ("posB", "<f4", 3), ("posC", "<f4", 3), ("normal", "<f4", 3),
     ("color", "<f4", 3), ("emissionStrength", "<f4"), ("smoothness", "<f4")])
SPHERE_DTYPE = np.dtype(
    [("pos", "<f4", 3), ("r", "<f4"), ("color", "<f4", 3), ("emissionStrength", "<f4"), ("smoothness", "<f4")])

RTC_SPLIT_TILES, RTC_SPLIT_SAMPLES = 0, 1
RTC_FLAG_COUNT_EXITS, RTC_FLAG_EXACT_ONLY, RTC_FLAG_PATHS_ONLY, RTC_FLAG_PER_PAIR_FILTER, RTC_FLAG_TIME_KERNELS = 1, 2, 4, 8, 16
RTC_FLAG_VERIFY_QUICK, RTC_FLAG_REF_QUICK, RTC_FLAG_SERIAL_PATHS = 32, 64, 128
STATS_WORDS = 24
RTC_NOALLOC_MAX_TRIS = 2048  # include/rtc_b200.h
ERRORS = {1: "RTC_ERR_ARG", 2: "RTC_ERR_NO_GPU", 3: "RTC_ERR_CUDA", 4: "RTC_ERR_NCCL", 5: "RTC_ERR_NOMEM"}

# the sphere the reference hard-codes (scene.h:17-19); used when trianglesOnly == 0
DEFAULT_SPHERES = np.array([((0, 1, 0), 2.5, (1, 1, 1), 0, 0)], dtype=SPHERE_DTYPE)


class Vec3(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float)]


class Scene(C.Structure):  # reference raytracing.h:7-11
    _fields_ = [("normalizedSunDirection", Vec3), ("skyColorHorizon", Vec3), ("skyColorZenith", Vec3),
                ("groundColor", Vec3), ("sunFocus", C.c_float), ("sunIntensity", C.c_float)]


class Camera(C.Structure):  # reference threadArgs, main.c:30-38
    _fields_ = [("origin", Vec3), ("ex", Vec3), ("ey", Vec3), ("ez", Vec3), ("fov", C.c_float)]


class RenderStats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("tri_pairs", C.c_uint64), ("sphere_pairs", C.c_uint64),
                ("tri_exit", C.c_uint64 * 6), ("hit_pixels", C.c_uint64), ("upload_ms", C.c_float),
                ("kernel_ms", C.c_float), ("gather_ms", C.c_float), ("download_ms", C.c_float),
                ("kernels_launched", C.c_int32), ("ngpus_used", C.c_int32), ("check_failures", C.c_uint64),
                ("pool_samples_evaluated", C.c_uint64), ("pool_shade_passes", C.c_uint64), ("pool_shade_lanes", C.c_uint64),
                ("pool_sweep_passes", C.c_uint64), ("pool_sweep_rays", C.c_uint64), ("pool_iterations", C.c_uint64)]

    def as_dict(self):
        return {k: (list(getattr(self, k)) if k == "tri_exit" else getattr(self, k)) for k, _ in self._fields_}


class RenderParams(C.Structure):
    _fields_ = [("tris", C.c_void_p), ("ntris", C.c_int32), ("spheres", C.c_void_p), ("nspheres", C.c_int32),
                ("width", C.c_int32), ("height", C.c_int32), ("spp", C.c_int32), ("maxBounce", C.c_int32),
                ("trianglesOnly", C.c_int32), ("ngpus", C.c_int32), ("split_mode", C.c_int32),
                ("nsplit", C.c_int32), ("out_primary_id", C.c_void_p), ("stats", C.POINTER(RenderStats)),
                ("flags", C.c_uint32)]


class RtcSubset(C.Structure):
    _fields_ = [("row_start", C.c_int32), ("row_step", C.c_int32), ("col_start", C.c_int32),
                ("col_step", C.c_int32), ("band_h", C.c_int32)]


class RtcDeviceJob(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("spp", C.c_int32), ("maxBounce", C.c_int32),
                ("trianglesOnly", C.c_int32), ("nsplit", C.c_int32), ("split_first", C.c_int32),
                ("split_count", C.c_int32), ("subset", RtcSubset), ("out_w", C.c_int32), ("out_h", C.c_int32),
                ("flags", C.c_uint32), ("rgb_full", C.c_int32)]


class RtcError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"{ERRORS.get(code, code)}: {msg}")
        self.code = code


_lib = None

# every symbol include/rtc_b200.h declares
EXPORTS = ["render_cuda", "render_cuda_shutdown", "rtc_camera_look_at", "rtc_normalize", "rtc_device_count",
           "rtc_last_error", "rtc_version", "rtc_scene_create", "rtc_scene_destroy", "rtc_subset_size",
           "rtc_render_device", "rtc_sum_quantize_device", "rtc_quantize_device", "rtc_render_launch_count",
           "rtc_uninterleave_device", "rtc_download_bands", "rtc_fp32_peak_probe", "rtc_scene_update", "rtc_scene_kernel_times", "rtc_synth_scene",
           "rtc_probe_random_direction", "rtc_probe_environment_light", "rtc_launches_total", "rtc_ipc_export",
           "rtc_ipc_open", "rtc_ipc_close", "rtc_enable_peer", "rtc_fnv1a64", "rtc_render_paths_noalloc", "rtc_probe_normal_selfcheck"]


def load_library() -> C.CDLL:
    """Load librtc_b200.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(f"{LIB_PATH} not built: run `python -m raytracingc_b200.build` "
                                "(there is no CPU fallback)")
    if not os.environ.get("RTC_NCCL_LIB"):
        # render_cuda(ngpus > 1) dlopens NCCL: make it the copy torch is linked against, so that whichever of the
        # two is loaded first the process holds ONE libnccl.so.2
        for d in sys.path:
            cand = os.path.join(d, "nvidia", "nccl", "lib", "libnccl.so.2")
            if os.path.exists(cand):
                os.environ["RTC_NCCL_LIB"] = cand
                break
    lib = C.CDLL(LIB_PATH)
    P = C.POINTER
    lib.render_cuda.argtypes = [P(Scene), P(Camera), C.c_void_p, P(RenderParams)]
    lib.render_cuda.restype = C.c_int
    lib.render_cuda_shutdown.restype = None
    lib.rtc_camera_look_at.argtypes = [P(C.c_float), P(C.c_float), C.c_float, P(Camera)]
    lib.rtc_camera_look_at.restype = None
    lib.rtc_normalize.argtypes = [P(C.c_float), P(C.c_float)]
    lib.rtc_normalize.restype = None
    lib.rtc_device_count.restype = C.c_int
    lib.rtc_last_error.restype = C.c_char_p
    lib.rtc_version.restype = C.c_char_p
    lib.rtc_scene_create.argtypes = [C.c_int, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, P(C.c_void_p)]
    lib.rtc_scene_create.restype = C.c_int
    lib.rtc_scene_destroy.argtypes = [C.c_void_p]
    lib.rtc_scene_update.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p]
    lib.rtc_scene_update.restype = C.c_int
    lib.rtc_scene_destroy.restype = None
    lib.rtc_subset_size.argtypes = [C.c_int32, C.c_int32, P(RtcSubset), P(C.c_int32), P(C.c_int32)]
    lib.rtc_subset_size.restype = None
    lib.rtc_render_device.argtypes = [C.c_void_p, P(Scene), P(Camera), P(RtcDeviceJob), C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_void_p]
    lib.rtc_render_device.restype = C.c_int
    lib.rtc_sum_quantize_device.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p,
                                            C.c_void_p]
    lib.rtc_sum_quantize_device.restype = C.c_int
    lib.rtc_quantize_device.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    lib.rtc_quantize_device.restype = C.c_int
    lib.rtc_render_launch_count.argtypes = [C.c_void_p, P(RtcDeviceJob), C.c_int]
    lib.rtc_render_launch_count.restype = C.c_int
    lib.rtc_uninterleave_device.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                            C.c_int, C.c_size_t, C.c_void_p]
    lib.rtc_uninterleave_device.restype = C.c_int
    lib.rtc_download_bands.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int,
                                       C.c_void_p]
    lib.rtc_download_bands.restype = C.c_int
    lib.rtc_scene_kernel_times.argtypes = [C.c_void_p, P(C.c_float), P(C.c_float)]
    lib.rtc_scene_kernel_times.restype = C.c_int
    lib.rtc_synth_scene.argtypes = [C.c_uint64, C.c_int32, C.c_void_p]
    lib.rtc_synth_scene.restype = None
    lib.rtc_fp32_peak_probe.argtypes = [C.c_int, C.c_int, P(C.c_double), P(C.c_double)]
    lib.rtc_fp32_peak_probe.restype = C.c_int
    lib.rtc_probe_random_direction.argtypes = [C.c_int, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]
    lib.rtc_probe_random_direction.restype = C.c_int
    lib.rtc_probe_environment_light.argtypes = [C.c_int, P(Scene), C.c_void_p, C.c_int32, C.c_void_p]
    lib.rtc_probe_environment_light.restype = C.c_int
    lib.rtc_launches_total.restype = C.c_uint64
    lib.rtc_ipc_export.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
    lib.rtc_ipc_export.restype = C.c_int
    lib.rtc_ipc_open.argtypes = [C.c_int, C.c_void_p, P(C.c_void_p)]
    lib.rtc_ipc_open.restype = C.c_int
    lib.rtc_ipc_close.argtypes = [C.c_int, C.c_void_p]
    lib.rtc_ipc_close.restype = C.c_int
    lib.rtc_enable_peer.argtypes = [C.c_int, C.c_int]
    lib.rtc_enable_peer.restype = C.c_int
    lib.rtc_fnv1a64.argtypes = [C.c_void_p, C.c_size_t]
    lib.rtc_fnv1a64.restype = C.c_uint64
    lib.rtc_render_paths_noalloc.argtypes = [C.c_void_p, P(Scene), P(Camera), P(RtcDeviceJob), C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_void_p]
    lib.rtc_render_paths_noalloc.restype = C.c_int
    lib.rtc_probe_normal_selfcheck.argtypes = [C.c_int, C.c_uint32, C.c_uint64, P(C.c_uint64)]
    lib.rtc_probe_normal_selfcheck.restype = C.c_int
    _lib = lib
    return lib


def _check(rc: int) -> None:
    if rc != 0:
        raise RtcError(rc, load_library().rtc_last_error().decode())


def _f3(v):
    return (C.c_float * 3)(*[float(c) for c in v])


def _vec(v) -> Vec3:
    return Vec3(*[float(np.float32(c)) for c in v])


def device_count() -> int:
    return load_library().rtc_device_count()


def normalize(v):
    """normalized() of the reference (moremath.c:12-18), computed by the library's host code."""
    out = (C.c_float * 3)()
    load_library().rtc_normalize(_f3(v), out)
    return tuple(out)


def camera_look_at(origin=(-4.75, -1.5, -4.75), look_at=(0.9, -1.2, 1.0), fov=1.0) -> Camera:
    """Camera basis as reference main.c:252-255; defaults are main.c:114-116."""
    cam = Camera()
    load_library().rtc_camera_look_at(_f3(origin), _f3(look_at), C.c_float(fov), C.byref(cam))
    return cam


def make_scene(sun=(-30.0, -85.0, 100.0), sun_focus=22.0, sun_intensity=0.75, horizon=(1.0, 1.0, 1.0),
               zenith=(0.263, 0.969, 0.871), ground=(0.66, 0.66, 0.66)) -> Scene:
    """Scene with the reference's defaults (main.c:14,21-28); the sun is normalised as main.c:247."""
    s = Scene()
    s.normalizedSunDirection = _vec(normalize(sun))
    s.skyColorHorizon, s.skyColorZenith, s.groundColor = _vec(horizon), _vec(zenith), _vec(ground)
    s.sunFocus, s.sunIntensity = sun_focus, sun_intensity
    return s


def render(tris: np.ndarray, width: int, height: int, spp: int, max_bounce: int, *, spheres: np.ndarray | None = None,
           triangles_only: int = 1, camera: Camera | None = None, scene: Scene | None = None, ngpus: int = 1,
           split_mode: int = RTC_SPLIT_TILES, nsplit: int = 1, want_ids: bool = False, flags: int = 0,
           out: np.ndarray | None = None):
    """One frame through render_cuda() with HOST buffers (the drop-in call for reference main.c:285-302).

    Returns (rgb uint8 [H, W, 3], ids int32 [H, W] or None, stats dict).
    """
    lib = load_library()
    tris = np.ascontiguousarray(tris, dtype=TRI_DTYPE)
    sph = DEFAULT_SPHERES if spheres is None else np.ascontiguousarray(spheres, dtype=SPHERE_DTYPE)
    cam = camera if camera is not None else camera_look_at()
    sc = scene if scene is not None else make_scene()
    rgb = out if out is not None else np.empty((height, width, 3), dtype=np.uint8)
    ids = np.empty((height, width), dtype=np.int32) if want_ids else None
    st = RenderStats()
    p = RenderParams()
    p.tris, p.ntris = tris.ctypes.data, len(tris)
    p.spheres, p.nspheres = sph.ctypes.data, len(sph)
    p.width, p.height, p.spp, p.maxBounce, p.trianglesOnly = width, height, spp, max_bounce, triangles_only
    p.ngpus, p.split_mode, p.nsplit = ngpus, split_mode, nsplit
    p.out_primary_id = ids.ctypes.data if ids is not None else None
    p.stats = C.pointer(st)
    p.flags = flags
    _check(lib.render_cuda(C.byref(sc), C.byref(cam), rgb.ctypes.data, C.byref(p)))
    return rgb, ids, st.as_dict()


def shutdown() -> None:
    load_library().render_cuda_shutdown()


def subset_size(width: int, height: int, subset: RtcSubset):
    w, h = C.c_int32(), C.c_int32()
    load_library().rtc_subset_size(width, height, C.byref(subset), C.byref(w), C.byref(h))
    return w.value, h.value


@dataclass
class DeviceJob:
    width: int
    height: int
    spp: int
    max_bounce: int
    triangles_only: int = 1
    nsplit: int = 1
    split_first: int = 0
    split_count: int = 0
    subset: tuple = (0, 1, 0, 1, 1)  # row_start, row_step, col_start, col_step, band_h
    flags: int = 0
    rgb_full: int = 0  # 1: d_rgb is the full image; pixels go to their final row and column
    out_w: int = field(default=0, init=False)
    out_h: int = field(default=0, init=False)

    def to_c(self) -> RtcDeviceJob:
        j = RtcDeviceJob()
        j.width, j.height, j.spp, j.maxBounce = self.width, self.height, self.spp, self.max_bounce
        j.trianglesOnly, j.nsplit = self.triangles_only, self.nsplit
        j.split_first, j.split_count = self.split_first, self.split_count
        j.subset = RtcSubset(*self.subset)
        j.out_w, j.out_h = subset_size(self.width, self.height, j.subset)
        self.out_w, self.out_h = j.out_w, j.out_h
        j.flags = self.flags
        j.rgb_full = self.rgb_full
        return j

    @property
    def planes(self) -> int:
        g = max(self.nsplit, 1)
        return self.split_count if self.split_count > 0 else g - self.split_first


class DeviceScene:
    """A scene flattened into HBM on one device (rtc_scene_create); render() takes DEVICE pointers."""

    def __init__(self, tris: np.ndarray, spheres: np.ndarray | None = None, device: int = 0):
        lib = load_library()
        tris = np.ascontiguousarray(tris, dtype=TRI_DTYPE)
        sph = DEFAULT_SPHERES if spheres is None else np.ascontiguousarray(spheres, dtype=SPHERE_DTYPE)
        h = C.c_void_p()
        _check(lib.rtc_scene_create(device, tris.ctypes.data, len(tris), sph.ctypes.data, len(sph), C.byref(h)))
        self.handle, self.device, self.ntris, self.nspheres = h, device, len(tris), len(sph)

    def render(self, job: DeviceJob, scene: Scene, camera: Camera, d_rgb: int = 0, d_accum: int = 0, d_ids: int = 0,
               d_stats: int = 0, stream: int = 0) -> RtcDeviceJob:
        """Queue the render; arguments are raw device addresses (e.g. torch.Tensor.data_ptr()) and a cudaStream_t."""
        j = job.to_c()
        _check(load_library().rtc_render_device(self.handle, C.byref(scene), C.byref(camera), C.byref(j),
                                                d_rgb or None, d_accum or None, d_ids or None, d_stats or None,
                                                stream or None))
        return j

    def render_noalloc(self, job: DeviceJob, scene: Scene, camera: Camera, d_rgb: int, d_ids: int = 0, d_stats: int = 0,
                       stream: int = 0) -> RtcDeviceJob:
        """rtc_render_paths_noalloc: the path kernel with its scratch in the library's static device array -- no
        allocation inside the call (no counterpart in the reference).  <= RTC_NOALLOC_MAX_TRIS
        triangles; calls that overlap on one GPU race on the shared array."""
        j = job.to_c()
        _check(load_library().rtc_render_paths_noalloc(self.handle, C.byref(scene), C.byref(camera), C.byref(j), d_rgb or None,
                                                       d_ids or None, d_stats or None, stream or None))
        return j

    def update(self, tris: np.ndarray, spheres: np.ndarray | None = None, stream: int = 0):
        """Host -> device re-upload + flatten, queued on `stream` (arrays must stay alive until it has run)."""
        tris = np.ascontiguousarray(tris, dtype=TRI_DTYPE)
        sph = DEFAULT_SPHERES if spheres is None else np.ascontiguousarray(spheres, dtype=SPHERE_DTYPE)
        self._keep = (tris, sph)
        _check(load_library().rtc_scene_update(self.handle, tris.ctypes.data, len(tris), sph.ctypes.data, len(sph),
                                               stream or None))
        self.ntris, self.nspheres = len(tris), len(sph)

    def kernel_times(self):
        """(primary_ms, paths_ms) of the last render launched with RTC_FLAG_TIME_KERNELS (waits for it)."""
        p, q = C.c_float(), C.c_float()
        _check(load_library().rtc_scene_kernel_times(self.handle, C.byref(p), C.byref(q)))
        return p.value, q.value

    def launch_count(self, job: DeviceJob, want_ids: bool = False) -> int:
        return load_library().rtc_render_launch_count(self.handle, C.byref(job.to_c()), int(want_ids))

    def close(self):
        if self.handle:
            load_library().rtc_scene_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def sum_quantize(device: int, d_planes: int, nplanes: int, n: int, d_rgb: int = 0, d_sum: int = 0, stream: int = 0):
    _check(load_library().rtc_sum_quantize_device(device, d_planes, nplanes, n, d_rgb or None, d_sum or None,
                                                  stream or None))


def uninterleave(device: int, d_parts: int, d_img: int, elem_size: int, row_elems: int, height: int, nparts: int,
                 band_h: int, part_elems: int, stream: int = 0):
    _check(load_library().rtc_uninterleave_device(device, d_parts, d_img, elem_size, row_elems, height, nparts,
                                                  band_h, part_elems, stream or None))


def download_bands(device: int, d_part: int, h_img: int, row_bytes: int, height: int, nparts: int, band_h: int,
                   part: int, stream: int = 0):
    """This part's rows of a band split straight into a host image (asynchronous on `stream`)."""
    _check(load_library().rtc_download_bands(device, d_part, h_img, row_bytes, height, nparts, band_h, part,
                                             stream or None))


def fnv1a64(a: np.ndarray) -> str:
    """FNV-1a-64 of an array's bytes, as 16 hex digits (the image hash of tests/golden/kat.json)."""
    a = np.ascontiguousarray(a)
    return f"{load_library().rtc_fnv1a64(a.ctypes.data, a.nbytes):016x}"


def launches_total() -> int:
    """Kernels the library has launched since it was loaded (difference two readings)."""
    return int(load_library().rtc_launches_total())


def ipc_export(device: int, d_ptr: int) -> bytes:
    """64-byte CUDA IPC handle of a device buffer of this process."""
    h = (C.c_ubyte * 64)()
    _check(load_library().rtc_ipc_export(device, d_ptr, h))
    return bytes(h)


def ipc_open(device: int, handle: bytes) -> int:
    """Another process's buffer as a device pointer of this one."""
    h = (C.c_ubyte * 64).from_buffer_copy(handle)
    p = C.c_void_p()
    _check(load_library().rtc_ipc_open(device, h, C.byref(p)))
    return int(p.value)


def ipc_close(device: int, d_ptr: int) -> None:
    _check(load_library().rtc_ipc_close(device, d_ptr))


def enable_peer(device: int, peer: int) -> None:
    """Let `device` read and write `peer`'s memory (one process driving several GPUs)."""
    _check(load_library().rtc_enable_peer(device, peer))


def synth_scene(ntris: int, seed: int = 0xB200) -> np.ndarray:
    """Triangle[] of BASELINE.json's synthetic config 5 (host code in the library)."""
    out = np.zeros(ntris, dtype=TRI_DTYPE)
    load_library().rtc_synth_scene(seed, ntris, out.ctypes.data)
    return out


def fp32_peak_tflops(device: int = 0, reps: int = 5) -> float:
    """Measured dependent-free FFMA throughput of the device (2 flop per FFMA)."""
    t, ms = C.c_double(), C.c_double()
    _check(load_library().rtc_fp32_peak_probe(device, reps, C.byref(t), C.byref(ms)))
    return t.value


def probe_random_direction(states: np.ndarray, device: int = 0):
    """Device RandomDiretion on n independent RNG states -> (directions float32 [n, 3], states after uint32 [n])."""
    st = np.ascontiguousarray(states, dtype=np.uint32)
    out = np.empty((len(st), 3), dtype=np.float32)
    after = np.empty(len(st), dtype=np.uint32)
    _check(load_library().rtc_probe_random_direction(device, st.ctypes.data, len(st), out.ctypes.data, after.ctypes.data))
    return out, after


def probe_normal_selfcheck(first_state: int = 0, count: int = 1 << 32, device: int = 0) -> dict:
    """Fast path of the normal deviate against the exact sequence on `count` generator states (default: all 2^32)."""
    out = (C.c_uint64 * 3)()
    _check(load_library().rtc_probe_normal_selfcheck(device, first_state & 0xFFFFFFFF, count, out))
    return {"different": int(out[0]), "declined": int(out[1]), "states": int(out[2])}


def probe_environment_light(dirs: np.ndarray, scene: Scene | None = None, device: int = 0) -> np.ndarray:
    """Device getEnvironmentLight on n directions -> float32 [n, 3]."""
    d = np.ascontiguousarray(dirs, dtype=np.float32).reshape(-1, 3)
    sc = scene if scene is not None else make_scene()
    out = np.empty_like(d)
    _check(load_library().rtc_probe_environment_light(device, C.byref(sc), d.ctypes.data, len(d), out.ctypes.data))
    return out
