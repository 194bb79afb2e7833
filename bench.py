#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200-native render loop.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c4|c2|...]

A "step" is one frame of the workload (default: BASELINE.json's target config,
ultracomplex.obj 3840x2160, 64 spp, 4 bounces, default camera).  Prints ONE JSON
line (contract in the task statement; DESIGN.md section "Measurement").

  value     Mrays/s, whole job, scene resident in HBM, timed with CUDA events
            on the launching stream; ray = one closest-hit sweep (a call of
            calculateRayCollision, reference raytracing.c:270), counted by the
            kernel itself.
  e2e       same metric through the C ABI with HOST buffers: scene upload and
            image download inside the timed region.
  roofline  FP32 pipe: algorithmic flops of the reference's own early exits
             / kernel time, against the FFMA peak measured here.
  cpu_baseline  the reference's own C functions (oracle/_ref) on the host cores.

Under torchrun (N > 1) every rank renders its interleaved 8-row bands and the
parts are exchanged with NCCL all_gather (strong scaling of one frame).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
BAND_H = 8

# name -> (scene fixture, width, height, spp, bounces, trianglesOnly); BASELINE.json configs
WORKLOADS = {
    "c1": ("simplest", 320, 240, 1, 1, 1),
    "c2": ("suze", 1280, 720, 16, 3, 1),
    "c3": ("complex", 1920, 1080, 64, 4, 1),
    "c3s": ("complex", 1920, 1080, 64, 4, 0),
    "c4": ("ultracomplex", 3840, 2160, 64, 4, 1),
    "default": ("default", 640, 480, 16, 4, 0),
}
WORKLOAD_LABEL = {
    "c1": "3Dmodels/simplest.obj 320x240, 1 spp, 1 bounce",
    "c2": "3Dmodels/suze.obj 1280x720, 16 spp, 3 bounces",
    "c3": "3Dmodels/complex.obj 1920x1080, 64 spp, 4 bounces, trianglesOnly=1",
    "c3s": "3Dmodels/complex.obj 1920x1080, 64 spp, 4 bounces, spheres + triangles",
    "c4": "3Dmodels/ultracomplex.obj 3840x2160, 64 spp, 4 bounces",
    "default": "triangles.txt + sphere 640x480, 16 spp, 4 bounces",
}
# reference exits (raytracing.c:189-208, AB/AC hoisted): cumulative FP32 flops at each exit, SURVEY 8d
STAGE_FLOPS = [5, 19, 29, 45, 51, 51]
SPHERE_FLOPS = 17  # to the delta < 0 exit; 39 on a hit (not separated by the counters: lower bound used)


def load_tris(scene: str) -> np.ndarray:
    from raytracingc_b200 import TRI_DTYPE
    return np.fromfile(os.path.join(GOLDEN, "scenes", scene + ".tris"), dtype=TRI_DTYPE)


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons with NVML while the timed region runs."""

    def __init__(self, index: int, period: float = 0.02):
        super().__init__(daemon=True)
        self.period, self.index = period, index
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._stop_ev = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # noqa: BLE001
            self.err = repr(e)

    NAMES = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
             0x80: "hw_power_brake", 0x2: "applications_clocks_setting", 0x100: "display_clock_setting"}

    def run(self):
        if not self.ok:
            return
        while not self._stop_ev.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                r = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(
                    self.nv, "nvmlDeviceGetCurrentClocksEventReasons") else \
                    self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self.NAMES.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(self.period)

    def stop(self) -> dict:
        self._stop_ev.set()
        if self.is_alive():
            self.join(timeout=2)
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "NVML unavailable"}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ---------------------------------------------------------------------------------------
# CPU baseline: the reference's own functions (oracle/_ref), one process per host core,
# rows interleaved as the reference's threads are (main.c:84)
# ---------------------------------------------------------------------------------------

def _cpu_rows_worker(args):
    kind, scene, W, H, spp, bounce, tonly, rows = args
    from oracle.oracle_api import Job, Oracle
    o = Oracle(kind)
    o.set_triangles(load_tris(scene))
    rgb = np.zeros((H, W, 3), np.uint8)
    t = time.perf_counter()
    for (start, step) in rows:
        o.render(Job(W, H, spp, bounce, triangles_only=tonly, row_start=start, row_step=step), want_ids=False,
                 out=(rgb, None))
    return time.perf_counter() - t


def _count_rays_port(args):
    scene, W, H, spp, bounce, tonly, start, step = args
    from oracle.oracle_api import Job, Oracle
    o = Oracle("port")
    o.set_triangles(load_tris(scene))
    _, _, st = o.render(Job(W, H, spp, bounce, triangles_only=tonly, row_start=start, row_step=step), want_ids=False)
    return st["rays"]


class CpuBaseline:
    """Times a bounded, row-interleaved sample of the workload on all host cores."""

    def __init__(self, workload: str, target_core_seconds: float = 60.0):
        from oracle.oracle_api import have_reference_lib
        self.kind = "reference" if have_reference_lib() else "port"
        self.workload = workload
        self.scene, self.W, self.H, self.spp, self.bounce, self.tonly = WORKLOADS[workload]
        self.cores = os.cpu_count() or 1
        # survey: ~100 M ray-triangle pairs / s / core; choose the row stride so the sample is ~target core-seconds
        ntris = len(load_tris(self.scene))
        est_core_s = self.W * self.H * self.spp * 1.1 * max(ntris, 1) / 1.0e8
        stride = 1
        while est_core_s / stride > target_core_seconds and stride < self.H // self.cores:
            stride *= 2
        self.stride = stride
        self.rows = list(range(0, self.H, stride))
        self._rays = None
        self._pool = None

    def pool(self):
        import multiprocessing as mp
        if self._pool is None:
            self._pool = mp.get_context("spawn").Pool(self.cores)  # spawn: the parent may hold a CUDA context
        return self._pool

    def sample_desc(self) -> str:
        return (f"every {self.stride}th row of {WORKLOAD_LABEL[self.workload]} ({len(self.rows)} of {self.H} rows, "
                f"full spp), {self.cores} processes, rows interleaved")

    def rays(self) -> int:
        """Rays of the sample, counted by the plain-C restatement (the pristine reference cannot count)."""
        if self._rays is None:
            P = self.cores
            jobs = [(self.scene, self.W, self.H, self.spp, self.bounce, self.tonly, k * self.stride, P * self.stride)
                    for k in range(P)]
            self._rays = int(sum(self.pool().map(_count_rays_port, jobs)))
        return self._rays

    def step(self) -> float:
        """One pass over the sample; wall seconds."""
        P = self.cores
        jobs = [(self.kind, self.scene, self.W, self.H, self.spp, self.bounce, self.tonly,
                 [(k * self.stride, P * self.stride)]) for k in range(P)]
        t = time.perf_counter()
        self.pool().map(_cpu_rows_worker, jobs)
        return time.perf_counter() - t

    def close(self):
        if self._pool is not None:
            self._pool.close()
            self._pool.join()
            self._pool = None


# ---------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------

def algorithmic_flops(stats: np.ndarray) -> tuple[float, float]:
    """(stage-weighted flops, brute-force-ceiling flops) of one frame from the kernel's exit counters."""
    exits = [float(stats[3 + e]) for e in range(6)]
    staged = sum(n * w for n, w in zip(exits, STAGE_FLOPS)) + float(stats[2]) * SPHERE_FLOPS
    brute = float(stats[1]) * 51.0 + float(stats[2]) * SPHERE_FLOPS
    return staged, brute


def run_ours(a) -> dict:
    import torch
    import torch.distributed as dist

    import raytracingc_b200 as rtc
    from raytracingc_b200 import api

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != a.gpus and world > 1:
        raise SystemExit(f"--gpus {a.gpus} but WORLD_SIZE={world}")
    if a.gpus > 1 and world == 1:
        raise SystemExit("N > 1 is launched as one process per GPU: python -m torch.distributed.run "
                         "--nproc-per-node N bench.py --gpus N ...")
    if not torch.cuda.is_available() or rtc.device_count() <= local:
        raise SystemExit("bench.py: no CUDA device (the product has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    wl = a.workload
    scene_name, W, H, spp, bounce, tonly = WORKLOADS[wl]
    tris = load_tris(scene_name)
    cam, sky = rtc.camera_look_at(), rtc.make_scene()
    G = world
    subset = (rank * BAND_H, G * BAND_H, 0, 1, BAND_H) if G > 1 else (0, 1, 0, 1, 1)
    job = rtc.DeviceJob(W, H, spp, bounce, triangles_only=tonly, subset=subset)
    job.to_c()
    job0 = rtc.DeviceJob(W, H, spp, bounce, triangles_only=tonly, subset=(0, G * BAND_H, 0, 1, BAND_H) if G > 1 else subset)
    job0.to_c()
    part_h = job0.out_h  # rank 0 owns the most rows; parts are padded to it for the all_gather
    stream = torch.cuda.current_stream()
    sh = stream.cuda_stream

    dscene = rtc.DeviceScene(tris, device=local)
    part = torch.zeros((part_h, W, 3), dtype=torch.uint8, device=dev)
    gathered = torch.zeros((G, part_h, W, 3), dtype=torch.uint8, device=dev) if G > 1 else None
    image = torch.zeros((H, W, 3), dtype=torch.uint8, device=dev) if G > 1 else part
    stats = torch.zeros(api.STATS_WORDS, dtype=torch.int64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    host_img = torch.empty((H, W, 3), dtype=torch.uint8).pin_memory()
    host_tris = torch.from_numpy(tris.view(np.uint8).copy()).pin_memory()
    host_tris_np = host_tris.numpy().view(api.TRI_DTYPE)

    # N > 1, end-to-end: ONE host image shared by the ranks (POSIX shared memory, page-locked in every process); each
    # rank copies its own bands into it over its own PCIe link (rtc_download_bands) -- no collective on the data path
    shared = None
    if G > 1:
        shm_path = f"/dev/shm/rtc_b200_bench_{os.environ.get('MASTER_PORT', '0')}_{W}x{H}"
        nbytes = W * H * 3 + 8 * G
        try:
            if rank == 0:
                with open(shm_path, "wb") as f:
                    f.truncate(nbytes)
            dist.barrier()
            mm = np.memmap(shm_path, dtype=np.uint8, mode="r+", shape=(nbytes,))
            rc = torch.cuda.cudart().cudaHostRegister(mm.ctypes.data, nbytes, 0)
            ok = torch.tensor([1 if int(rc) == 0 else 0], device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if int(ok[0]) == 1:
                shared = {"mm": mm, "img": mm[:W * H * 3].reshape(H, W, 3), "flags": mm[W * H * 3:].view(np.int64),
                          "frame": 0}
                shared["flags"][rank] = 0
            dist.barrier()
        except (OSError, RuntimeError) as exc:  # no shared memory here: the gather path below still measures e2e
            print(f"bench.py: shared host image unavailable ({exc}); e2e goes through the all-gather", file=sys.stderr)
            shared = None

    launches = [0]
    nk = dscene.launch_count(job)  # kernels per rtc_render_device call (table preparation, primary kernel, path kernel)

    def frame(count_stats: bool = False, flags: int = 0):
        """device-resident step: render this rank's bands (+ gather and un-interleave when N > 1)"""
        job.flags = flags
        dscene.render(job, sky, cam, d_rgb=part.data_ptr(), d_stats=stats.data_ptr() if count_stats else 0, stream=sh)
        launches[0] += nk
        if G > 1:
            dist.all_gather_into_tensor(gathered, part)
            if rank == 0:
                api.uninterleave(local, gathered.data_ptr(), image.data_ptr(), 1, W * 3, H, G, BAND_H,
                                 part_h * W * 3, sh)
                launches[0] += 1

    def frame_e2e():
        """public-API step with HOST buffers: upload + flatten the scene, render, gather, download"""
        if G == 1:
            rtc.render(host_tris_np, W, H, spp, bounce, triangles_only=tonly, camera=cam, scene=sky,
                       out=host_img.numpy())
            launches[0] += 2 + nk  # flatten (triangles, spheres) + render kernels
        elif shared is not None:
            dscene.update(host_tris_np, stream=sh)
            job.flags = 0
            dscene.render(job, sky, cam, d_rgb=part.data_ptr(), stream=sh)
            launches[0] += 2 + nk
            api.download_bands(local, part.data_ptr(), shared["mm"].ctypes.data, W * 3, H, G, BAND_H, rank, sh)
            stream.synchronize()
            shared["frame"] += 1
            shared["flags"][rank] = shared["frame"]  # this rank's rows of frame n are in the host image
            if rank == 0:                            # the caller's view: the frame is complete when every rank says so
                t_wait = time.perf_counter()
                while int(shared["flags"].min()) < shared["frame"]:
                    if time.perf_counter() - t_wait > 30.0:
                        raise SystemExit("bench.py: a rank did not deliver its rows of the frame within 30 s")
        else:
            dscene.update(host_tris_np, stream=sh)
            launches[0] += 2
            frame()
            if rank == 0:
                host_img.copy_(image, non_blocking=True)
            stream.synchronize()

    def sync_all():
        torch.cuda.synchronize()
        if G > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # counted frames (exact kernel with exit counters; decisions are identical under parity, so the histogram
    # is the production kernels' too): the whole frame, and the primary rays alone (maxBounce = 1)
    stats.zero_()
    frame(count_stats=True, flags=api.RTC_FLAG_COUNT_EXITS)
    torch.cuda.synchronize()
    st_local = stats.clone()
    if G > 1:
        dist.all_reduce(stats)
    st = stats.cpu().numpy().astype(np.float64)
    rays_frame = float(st[0])
    staged_flops, brute_flops = algorithmic_flops(st)
    st_l = st_local.cpu().numpy().astype(np.float64)
    staged_local, brute_local = algorithmic_flops(st_l)
    job1 = rtc.DeviceJob(W, H, spp, 1, triangles_only=tonly, subset=subset, flags=api.RTC_FLAG_COUNT_EXITS)
    stats.zero_()
    dscene.render(job1, sky, cam, d_rgb=part.data_ptr(), d_stats=stats.data_ptr(), stream=sh)
    torch.cuda.synchronize()
    st_p = stats.cpu().numpy().astype(np.float64)          # primary rays of this rank's share
    staged_primary, brute_primary = algorithmic_flops(st_p)
    st_b = st_l - st_p                                      # bounce rays of this rank's share
    staged_bounce, brute_bounce = algorithmic_flops(st_b)

    peak_tflops = api.fp32_peak_tflops(local)

    for _ in range(max(a.warmup, 3)):
        frame()
    sync_all()

    sampler = ClockSampler(local)
    sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True),
           torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    launches[0] = 0
    for i in range(a.steps):
        flush.fill_(i & 0xFF)  # L2 flush between timed steps (outside the event pair)
        if G > 1:
            dist.barrier()
        ev[i][0].record(stream)
        job.flags = 0
        dscene.render(job, sky, cam, d_rgb=part.data_ptr(), stream=sh)
        launches[0] += nk
        ev[i][1].record(stream)
        if G > 1:
            dist.all_gather_into_tensor(gathered, part)
            if rank == 0:
                api.uninterleave(local, gathered.data_ptr(), image.data_ptr(), 1, W * 3, H, G, BAND_H,
                                 part_h * W * 3, sh)
                launches[0] += 1
        ev[i][2].record(stream)
    sync_all()
    clocks = sampler.stop()
    step_ms = [e[0].elapsed_time(e[2]) for e in ev]
    kern_ms = [e[0].elapsed_time(e[1]) for e in ev]
    total_ms = float(sum(step_ms))
    kernel_ms_avg = float(np.mean(kern_ms))
    n_launch = launches[0]

    # end-to-end through the public API with host buffers
    for _ in range(2):
        frame_e2e()
    sync_all()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        frame_e2e()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if shared is not None:
        # outside the timed region: the host image the ranks assembled equals the gathered device frame
        sync_all()
        frame()
        sync_all()
        if rank == 0 and not bool((torch.from_numpy(np.ascontiguousarray(shared["img"])) == image.cpu()).all()):
            raise SystemExit("bench.py: the shared host image differs from the gathered device frame")
        dist.barrier()
        torch.cuda.cudart().cudaHostUnregister(shared["mm"].ctypes.data)
        if rank == 0:
            os.unlink(shm_path)

    # device time of the two render kernels (events recorded by the library around each launch)
    prim_ms, path_ms = [], []
    for _ in range(3):
        job.flags = api.RTC_FLAG_TIME_KERNELS
        dscene.render(job, sky, cam, d_rgb=part.data_ptr(), stream=sh)
        try:
            a_ms, b_ms = dscene.kernel_times()
            prim_ms.append(a_ms)
            path_ms.append(b_ms)
        except rtc.RtcError:
            break
    job.flags = 0
    prim_ms = float(np.mean(prim_ms)) if prim_ms else None
    path_ms = float(np.mean(path_ms)) if path_ms else None

    # the same frame in the library's other modes (reported, not the headline)
    modes = {}
    if G == 1 and not a.no_modes:
        for name, fl in (("per_pair_filter", api.RTC_FLAG_PER_PAIR_FILTER), ("paths_only", api.RTC_FLAG_PATHS_ONLY),
                         ("exact_only", api.RTC_FLAG_EXACT_ONLY)):
            job.flags = fl
            dscene.render(job, sky, cam, d_rgb=part.data_ptr(), stream=sh)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            dscene.render(job, sky, cam, d_rgb=part.data_ptr(), stream=sh)
            e1.record(stream)
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            modes[name] = {"frame_ms": ms, "Mrays_per_s": rays_frame / (ms * 1e-3) / 1e6,
                           "algorithmic_frac_of_fp32_peak": staged_local / (ms * 1e-3) / 1e12 / peak_tflops}
        job.flags = 0

    if G > 1:
        t = torch.tensor([total_ms, e2e_s, kernel_ms_avg], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, e2e_s = float(t[0]), float(t[1])

    ms_per_step = total_ms / a.steps
    value = rays_frame / (ms_per_step * 1e-3) / 1e6
    e2e_value = rays_frame * a.steps / e2e_s / 1e6
    line = {
        "metric": "Mrays/s (ray = one closest-hit sweep over all triangles, reference raytracing.c:270)",
        "value": value, "unit": "Mrays/s", "n_gpus": G, "steps": a.steps, "warmup": max(a.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "bundled scene fixture (tests/golden/scenes, produced by the reference's loader)",
        "config": {"workload": WORKLOAD_LABEL[wl], "triangles": int(len(tris)), "rays_per_frame": int(rays_frame),
                   "camera": "reference default (main.c:114-116)", "split": "8-row bands interleaved over ranks, "
                   "NCCL all_gather + un-interleave" if G > 1 else "single GPU",
                   "l2": "flushed between timed steps (256 MiB fill, outside the event pair); scene is 8 KB",
                   "timing": "CUDA events per step on the launching stream, max over ranks"},
        "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": int(tris.nbytes + 36),
                "d2h_bytes_per_step": int(W * H * 3), "ms_per_step": e2e_s / a.steps * 1e3,
                "api": "render_cuda() with pinned host buffers" if G == 1 else
                       ("rtc_scene_update + rtc_render_device + rtc_download_bands: every rank copies its bands into one "
                        "shared page-locked host image over its own PCIe link, no collective" if shared is not None else
                        "rtc_scene_update + rtc_render_device + all_gather + rtc_uninterleave_device + D2H")},
        "gpu_launches": n_launch,
        "clocks": clocks,
        "frame_ms": ms_per_step,
    }
    # roofline of the dominant kernel.  The path kernel is the brute-force intersection kernel proper: every
    # (bounce ray, triangle) pair goes through its conservative quick test (FMUL/FFMA chains) and the survivors
    # through the reference's arithmetic (non-contracted FMUL/FADD).  The primary kernel decides most pairs with
    # conservative block bounds, so its algorithmic rate is not bounded by the FP32 peak.
    # traffic: DRAM bytes per launch of that kernel from the committed ncu --set full capture of this workload
    # (profiles/ncu_traffic.json: dram__bytes_read.sum + dram__bytes_write.sum), null when there is none.
    traffic = {}
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            traffic = json.load(f).get(wl, {})
    except OSError:
        pass

    def roof(kernel, ms, staged, brute, pairs, note):
        tr = next((v for k, v in traffic.items() if k in kernel), None)
        return {"bound": "fp32", "unit": "TFLOP/s", "kernel": kernel, "kernel_ms": ms,
                "achieved": staged / (ms * 1e-3) / 1e12, "peak": peak_tflops,
                "frac": staged / (ms * 1e-3) / 1e12 / peak_tflops, "traffic": tr,
                "brute_force_frac": brute / (ms * 1e-3) / 1e12 / peak_tflops,
                "pairs_per_s": pairs / (ms * 1e-3), "what": note}
    what = ("algorithmic FP32 flops of this kernel's (ray, triangle) pairs by the reference's own exits "
            "(5/19/29/45/51 per pair, SURVEY 8d) / its CUDA-event time; peak = dependent-free FFMA probe measured in "
            "this run (2 flop x FFMA/s; MEASURED_PEAKS.json has no FP32 entry); brute_force_frac = 51 flop x pairs")
    if path_ms and prim_ms:
        rp = roof("rtc::render_paths_kernel (bounce rays: brute-force sweep, conservative quick test + exact survivors, + shading)", path_ms, staged_bounce,
                  brute_bounce, float(st_b[1]), what)
        rq = roof("primary kernel (rtc::render_skyblock_kernel / render_sky_kernel: conservative bounds + exact "
                  "survivors, every primary ray every sample)", prim_ms, staged_primary, brute_primary, float(st_p[1]),
                  what + "; > 1 is expected here: pairs are decided by block bounds, not by the reference's flops")
        line["roofline"] = rp if path_ms >= prim_ms else rq
        line["roofline_other_kernel"] = rq if path_ms >= prim_ms else rp
    line["roofline_frame"] = roof("whole frame (all kernels)", kernel_ms_avg, staged_local, brute_local, float(st_l[1]),
                                  what)
    line["roofline_frame"]["nominal_peak"] = 2 * 128 * 148 * 1.965e9 / 1e12
    if "roofline" not in line:
        line["roofline"] = line["roofline_frame"]
    if modes:
        line["modes"] = modes
    if rank == 0 and G == 1 and not a.no_cpu:
        cb = CpuBaseline(wl)
        rays = cb.rays()  # also spawns and warms the worker pool
        secs = cb.step()
        line["cpu_baseline"] = {"value": rays / secs / 1e6, "unit": "Mrays/s", "cores": cb.cores, "kind": cb.kind,
                                "sample": cb.sample_desc(), "seconds": secs}
        cb.close()
    if world > 1:
        dist.destroy_process_group()
    return line if rank == 0 else None


# ---------------------------------------------------------------------------------------
# reference arm: the reference's own CPU implementation of the path on the host cores
# ---------------------------------------------------------------------------------------

def run_reference(a) -> dict | None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None  # rank 0 alone runs the CPU arm
    wl = a.workload
    cb = CpuBaseline(wl)
    rays = cb.rays()
    for _ in range(min(a.warmup, 1)):
        cb.step()
    secs = [cb.step() for _ in range(a.steps)]
    cb.close()
    ms = float(np.mean(secs)) * 1e3
    value = rays / (ms * 1e-3) / 1e6
    scene_name = WORKLOADS[wl][0]
    return {
        "impl": "reference",
        "metric": "Mrays/s (ray = one closest-hit sweep over all triangles, reference raytracing.c:270)",
        "value": value, "unit": "Mrays/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": min(a.warmup, 1),
        "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "bundled scene fixture (tests/golden/scenes, produced by the reference's loader)",
        "config": {"workload": WORKLOAD_LABEL[wl], "triangles": int(len(load_tris(scene_name))),
                   "rays_per_step": int(rays), "camera": "reference default (main.c:114-116)",
                   "note": "the reference's rayTriangle/calcColor/... compiled from /root/reference as one "
                           "translation unit (oracle/_ref); each step is a bounded row sample of the frame"},
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cb.cores, "kind": cb.kind,
                         "sample": cb.sample_desc()},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c4", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (N=1)")
    ap.add_argument("--no-modes", action="store_true", help="skip timing the other library modes")
    a = ap.parse_args()
    line = run_reference(a) if a.impl == "reference" else run_ours(a)
    if line is not None:
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
