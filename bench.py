#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200-native render loop.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c4|c4close|c5l|c2|...]
                    [--split tiles|samples] [--gather peer|nccl]

A "step" is one frame of the workload (default: BASELINE.json's target config, ultracomplex.obj 3840x2160, 64 spp,
4 bounces, default camera).  Prints ONE JSON line (contract in the task statement; DESIGN.md section "Measurement").

  value     Mrays/s, whole job, scene resident in HBM, timed with CUDA events on the launching stream; ray = one
            closest-hit sweep (a call of calculateRayCollision, reference raytracing.c:270), counted by the kernels.
            97 % of C4's rays are primary re-sweeps decided by block bounds: `path_kernel` beside it is the figure for
            the brute-force intersection kernel alone (bounce rays and ray-triangle pairs per second).
  e2e       same metric through the C ABI, render_cuda(), with HOST buffers: scene upload and image download inside
            the timed region -- at N > 1 too (render_cuda(ngpus = N), one host thread per device inside the library).
  roofline  FP32 pipe: algorithmic flops of the reference's own early exits / kernel time, against the
            FFMA peak measured here.
  cpu_baseline  the reference's own C functions (oracle/_ref) on the host cores.
  parity_fnv_ok  every run checks the frame it produced (device-resident and end-to-end) against the golden vector
            of tests/golden/kat.json, outside the timed region; a mismatch exits non-zero.

Under torchrun (N > 1) every rank renders its interleaved 8-row bands and writes its pixels straight into rank 0's
frame over NVLink (CUDA IPC peer pointer, RtcDeviceJob.rgb_full): no collective, no un-interleave kernel
(--gather nccl keeps the all-gather + un-interleave as the checked alternative).  --split samples (the synthetic
1 M-triangle workload c5l): every rank renders all pixels for its RNG sub-streams, float planes are all-gathered and
summed in sub-stream order (bit-exact for any N).
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
CLOSE_CAM = ((0.0, -2.0, -6.0), (0.0, -1.0, 0.0))  # -p 0 -2 -6 -t 0 -1 0

# BASELINE.json configs (c1..c4, default) + SURVEY 8d's extra datapoint (c4close) + config 5 reduced to a stated lattice (c5l)
WORKLOADS = {
    "c1": dict(scene="simplest", W=320, H=240, spp=1, bounce=1, tonly=1, kat="c1_simplest_320x240_1spp_1b",
               label="3Dmodels/simplest.obj 320x240, 1 spp, 1 bounce"),
    "c2": dict(scene="suze", W=1280, H=720, spp=16, bounce=3, tonly=1, kat="c2_suze_1280x720_16spp_3b",
               label="3Dmodels/suze.obj 1280x720, 16 spp, 3 bounces"),
    "c3": dict(scene="complex", W=1920, H=1080, spp=64, bounce=4, tonly=1, kat="c3_complex_1920x1080_64spp_4b",
               label="3Dmodels/complex.obj 1920x1080, 64 spp, 4 bounces, trianglesOnly=1"),
    "c3s": dict(scene="complex", W=1920, H=1080, spp=64, bounce=4, tonly=0, kat="c3s_complex_sphere_1920x1080_64spp_4b",
                label="3Dmodels/complex.obj 1920x1080, 64 spp, 4 bounces, spheres + triangles"),
    "c4": dict(scene="ultracomplex", W=3840, H=2160, spp=64, bounce=4, tonly=1, kat="c4_ultracomplex_3840x2160_64spp_4b",
               label="3Dmodels/ultracomplex.obj 3840x2160, 64 spp, 4 bounces"),
    "c4close": dict(scene="ultracomplex", W=3840, H=2160, spp=64, bounce=4, tonly=1, camera=CLOSE_CAM,
                    kat="c4close_ultracomplex_3840x2160_64spp_4b",
                    label="3Dmodels/ultracomplex.obj 3840x2160, 64 spp, 4 bounces, close-up camera -p 0 -2 -6 -t 0 -1 0"),
    "default": dict(scene="default", W=640, H=480, spp=16, bounce=4, tonly=0, kat="default_640x480_16spp_4b",
                    label="triangles.txt + sphere 640x480, 16 spp, 4 bounces"),
    # config 5 (1 M synthetic triangles, 3840x2160, 256 spp) is hours of brute force per frame: the bench line is the
    # same scene on every 4th pixel of the 4K frame (960 x 540 pixels) at 16 spp, 4 bounces
    "c5l": dict(scene="synth", synth=1 << 20, W=3840, H=2160, spp=16, bounce=4, tonly=1, lattice=4,
                kat="c5l_synth1m_lattice240_16spp_4b_nsplit{nsplit}", kat_step=240,
                label="synthetic random-triangle scene (1M tris, rtc_synth_scene seed 0xB200) on the 960x540 lattice of "
                      "every 4th pixel of 3840x2160, 16 spp, 4 bounces"),
}
# reference exits (raytracing.c:189-208, AB/AC hoisted): cumulative FP32 flops at each exit, SURVEY 8d
STAGE_FLOPS = [5, 19, 29, 45, 51, 51]
SPHERE_FLOPS = 17  # to the delta < 0 exit; 39 on a hit (not separated by the counters: lower bound used)
METRIC = "Mrays/s (ray = one closest-hit sweep over all triangles, reference raytracing.c:270)"


def load_tris(scene: str, synth: int = 0) -> np.ndarray:
    from raytracingc_b200 import TRI_DTYPE, api
    if scene == "synth":
        return api.synth_scene(synth)
    return np.fromfile(os.path.join(GOLDEN, "scenes", scene + ".tris"), dtype=TRI_DTYPE)


def kat_entry(wl: dict, nsplit: int = 1):
    if not wl.get("kat"):
        return None
    with open(os.path.join(GOLDEN, "kat.json")) as f:
        return json.load(f).get(wl["kat"].format(nsplit=nsplit))


def make_config(name: str, wl: dict, G: int, split: str, gather: str, ntris: int, rays_per_frame) -> dict:
    """`config` of the JSON line; identical for both arms of a run (the reference arm times the same workload)."""
    if G == 1:
        how = "single GPU"
    elif split == "samples":
        how = f"sample split: {G} RNG sub-streams per pixel, one per rank; float planes all-gathered, summed in sub-stream order"
    elif gather == "nccl":
        how = f"8-row bands interleaved over {G} ranks, NCCL all_gather + un-interleave kernel"
    else:
        how = f"8-row bands interleaved over {G} ranks, every rank writes its pixels into rank 0's frame over NVLink (peer pointer)"
    return {"workload": wl["label"], "name": name, "triangles": int(ntris),
            "rays_per_frame": None if rays_per_frame is None else int(rays_per_frame),
            "camera": "close-up (SURVEY 8d: -p 0 -2 -6 -t 0 -1 0)" if wl.get("camera") else "reference default (main.c:114-116)",
            "split": how,
            "l2": "flushed between timed steps (256 MiB fill, outside the event pair)",
            "timing": "CUDA events per step on the launching stream, max over ranks"}


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

def _oracle_job(wl: dict, start: int, step: int):
    from oracle.oracle_api import Job
    kw = dict(triangles_only=wl["tonly"], row_start=start, row_step=step)
    if wl.get("camera"):
        kw.update(origin=wl["camera"][0], look_at=wl["camera"][1])
    if wl.get("cpu_col_step"):
        kw.update(col_step=wl["cpu_col_step"])
    return Job(wl["W"], wl["H"], wl["spp"], wl["bounce"], **kw)


def _cpu_rows_worker(args):
    kind, wl, rows = args
    from oracle.oracle_api import Oracle
    o = Oracle(kind)
    o.set_triangles(load_tris(wl["scene"], wl.get("synth", 0)))
    rgb = np.zeros((wl["H"], wl["W"], 3), np.uint8)
    t = time.perf_counter()
    for (start, step) in rows:
        o.render(_oracle_job(wl, start, step), want_ids=False, out=(rgb, None))
    return time.perf_counter() - t


def _count_rays_port(args):
    wl, start, step = args
    from oracle.oracle_api import Oracle
    o = Oracle("port")
    o.set_triangles(load_tris(wl["scene"], wl.get("synth", 0)))
    _, _, st = o.render(_oracle_job(wl, start, step), want_ids=False)
    return st["rays"]


class CpuBaseline:
    """Times a bounded, row-interleaved sample of the workload on all host cores."""

    def __init__(self, name: str, target_core_seconds: float = 60.0):
        from oracle.oracle_api import have_reference_lib
        self.kind = "reference" if have_reference_lib() else "port"
        self.wl = dict(WORKLOADS[name])
        self.cores = os.cpu_count() or 1
        wl = self.wl
        H = wl["H"]
        if wl.get("kat_step"):
            # the lattice workload: the CPU sample is the golden sub-lattice (every kat_step-th pixel in x and y, full spp)
            self.stride = wl["kat_step"]
            self.wl["cpu_col_step"] = wl["kat_step"]
        else:
            # survey: ~100 M ray-triangle pairs / s / core; choose the row stride so the sample is ~target core-seconds
            ntris = len(load_tris(wl["scene"], wl.get("synth", 0)))
            est_core_s = wl["W"] * H * wl["spp"] * (3.0 if wl.get("camera") else 1.1) * max(ntris, 1) / 1.0e8
            stride = 1
            while est_core_s / stride > target_core_seconds and stride < H // self.cores:
                stride *= 2
            self.stride = stride
        self.rows = list(range(0, H, self.stride))
        self._rays = None
        self._pool = None

    def pool(self):
        import multiprocessing as mp
        if self._pool is None:
            self._pool = mp.get_context("spawn").Pool(self.cores)  # spawn: the parent may hold a CUDA context
        return self._pool

    def sample_desc(self) -> str:
        what = f"every {self.stride}th pixel in x and y" if self.wl.get("kat_step") else f"every {self.stride}th row"
        return (f"{what} of {self.wl['label']} ({len(self.rows)} of {self.wl['H']} rows, full spp), {self.cores} processes, "
                f"rows interleaved")

    def _jobs(self):
        P = self.cores
        return [(k * self.stride, P * self.stride) for k in range(P) if k * self.stride < self.wl["H"]]

    def rays(self) -> int:
        """Rays of the sample, counted by the plain-C restatement (the pristine reference cannot count)."""
        if self._rays is None:
            self._rays = int(sum(self.pool().map(_count_rays_port, [(self.wl, s, p) for s, p in self._jobs()])))
        return self._rays

    def step(self) -> float:
        """One pass over the sample; wall seconds."""
        jobs = [(self.kind, self.wl, [(s, p)]) for s, p in self._jobs()]
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
    cpu_group = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        # host-side barrier for the end-to-end leg: while rank 0 drives all N devices through render_cuda(), the other
        # ranks must WAIT ON THE CPU -- an NCCL barrier spins in a kernel on their GPU, and two processes on one GPU are
        # time-sliced (measured: 7.1 ms instead of 2.6 ms per frame at N = 2)
        cpu_group = dist.new_group(backend="gloo")

    name = a.workload
    wl = WORKLOADS[name]
    W, H, spp, bounce, tonly = wl["W"], wl["H"], wl["spp"], wl["bounce"], wl["tonly"]
    tris = load_tris(wl["scene"], wl.get("synth", 0))
    cam = rtc.camera_look_at(*wl["camera"]) if wl.get("camera") else rtc.camera_look_at()
    sky = rtc.make_scene()
    G = world
    lat = wl.get("lattice", 0)
    split = a.split or ("samples" if (lat and G > 1) else "tiles")
    if lat and G > 1 and split != "samples":
        raise SystemExit("the lattice workload is split by samples (a band split of a lattice is not a frame a caller can ask for)")
    samples = split == "samples" and G > 1
    nsplit = G if samples else 1
    peer = G > 1 and not samples and a.gather == "peer"
    stream = torch.cuda.current_stream()
    sh = stream.cuda_stream

    # ---- this rank's job and buffers ----
    if lat:
        subset = (0, lat, 0, lat, 1)
    elif G > 1 and not samples:
        subset = (rank * BAND_H, G * BAND_H, 0, 1, BAND_H)
    else:
        subset = (0, 1, 0, 1, 1)
    job = rtc.DeviceJob(W, H, spp, bounce, triangles_only=tonly, subset=subset, nsplit=nsplit,
                        split_first=rank if samples else 0, split_count=1 if samples else 0, rgb_full=1 if peer else 0)
    job.to_c()
    out_w, out_h = job.out_w, job.out_h
    full_w, full_h = (out_w, out_h) if (lat or samples or G == 1) else (W, H)  # the frame the caller gets
    dscene = rtc.DeviceScene(tris, device=local)
    stats = torch.zeros(api.STATS_WORDS, dtype=torch.int64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    image = torch.zeros((full_h, full_w, 3), dtype=torch.uint8, device=dev) if (rank == 0 or not peer) else None
    part = gathered = planes = all_planes = None
    d_rgb = d_accum = 0
    peer_ptr = 0
    if samples:
        planes = torch.zeros((1, out_h, out_w, 3), dtype=torch.float32, device=dev)
        all_planes = torch.zeros((G, out_h, out_w, 3), dtype=torch.float32, device=dev)
        d_accum = planes.data_ptr()
    elif peer:
        # rank 0 owns the frame; the others open it over CUDA IPC and write their pixels at their final addresses
        handle = [api.ipc_export(local, image.data_ptr()) if rank == 0 else None]
        dist.broadcast_object_list(handle, src=0)
        ok = 1
        try:
            if rank != 0:
                if os.environ.get("RTC_BENCH_FORCE_IPC_FAIL"):
                    raise rtc.RtcError(3, "forced by RTC_BENCH_FORCE_IPC_FAIL")
                peer_ptr = api.ipc_open(local, handle[0])
        except rtc.RtcError as exc:  # no peer access between these devices: every rank falls back to the gather
            ok = 0
            print(f"bench.py: rank {rank} cannot map rank 0's frame ({exc}); falling back to --gather nccl", file=sys.stderr)
        t_ok = torch.tensor([ok], device=dev)
        dist.all_reduce(t_ok, op=dist.ReduceOp.MIN)
        if int(t_ok[0]) == 1:
            d_rgb = image.data_ptr() if rank == 0 else peer_ptr
        else:
            if peer_ptr:
                api.ipc_close(local, peer_ptr)
                peer_ptr = 0
            peer = False
            a.gather = "nccl"
            job.rgb_full = 0
            job.to_c()
            if image is None:
                image = torch.zeros((full_h, full_w, 3), dtype=torch.uint8, device=dev)
    if G > 1 and not samples and not peer:  # --gather nccl: compact parts, all_gather, un-interleave on rank 0
        job0 = rtc.DeviceJob(W, H, spp, bounce, triangles_only=tonly, subset=(0, G * BAND_H, 0, 1, BAND_H))
        job0.to_c()
        part_h = job0.out_h  # rank 0 owns the most rows; parts are padded to it for the all_gather
        part = torch.zeros((part_h, W, 3), dtype=torch.uint8, device=dev)
        gathered = torch.zeros((G, part_h, W, 3), dtype=torch.uint8, device=dev)
        d_rgb = part.data_ptr()
    if G == 1:
        d_rgb = image.data_ptr()
    host_img = torch.empty((full_h, full_w, 3), dtype=torch.uint8).pin_memory()
    host_tris = torch.from_numpy(tris.view(np.uint8).copy()).pin_memory()
    host_tris_np = host_tris.numpy().view(api.TRI_DTYPE)

    def render_part(flags: int = 0, want_stats: bool = False, j=None):
        j = j or job
        j.flags = flags
        dscene.render(j, sky, cam, d_rgb=d_rgb, d_accum=d_accum, d_stats=stats.data_ptr() if want_stats else 0, stream=sh)

    def assemble():
        """what follows the render when the frame is not assembled by the stores themselves"""
        if samples:
            dist.all_gather_into_tensor(all_planes, planes)
            if rank == 0:
                api.sum_quantize(local, all_planes.data_ptr(), G, out_h * out_w * 3, d_rgb=image.data_ptr(), stream=sh)
        elif G > 1 and not peer:
            dist.all_gather_into_tensor(gathered, part)
            if rank == 0:
                api.uninterleave(local, gathered.data_ptr(), image.data_ptr(), 1, W * 3, H, G, BAND_H, part.numel(), sh)

    def sync_all():
        torch.cuda.synchronize()
        if G > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- counted frames (exact kernel with exit counters; decisions are identical under parity, so the histogram
    #      is the production kernels' too): the whole frame, and the primary rays alone (maxBounce = 1) ----
    stats.zero_()
    render_part(api.RTC_FLAG_COUNT_EXITS, True)
    torch.cuda.synchronize()
    st_local = stats.clone()
    if G > 1:
        dist.all_reduce(stats)
    st = stats.cpu().numpy().astype(np.float64)
    rays_frame = float(st[0])
    st_l = st_local.cpu().numpy().astype(np.float64)
    staged_local, brute_local = algorithmic_flops(st_l)
    job1 = rtc.DeviceJob(W, H, spp, 1, triangles_only=tonly, subset=subset, nsplit=nsplit, split_first=job.split_first,
                         split_count=job.split_count, rgb_full=job.rgb_full)
    stats.zero_()
    render_part(api.RTC_FLAG_COUNT_EXITS, True, job1)
    torch.cuda.synchronize()
    st_p = stats.cpu().numpy().astype(np.float64)          # primary rays of this rank's share
    staged_primary, brute_primary = algorithmic_flops(st_p)
    st_b = st_l - st_p                                      # bounce rays of this rank's share
    staged_bounce, brute_bounce = algorithmic_flops(st_b)
    peak_tflops = api.fp32_peak_tflops(local)

    for _ in range(max(a.warmup, 3)):
        render_part()
        assemble()
    sync_all()

    # ---- timed region: device-resident frame ----
    sampler = ClockSampler(local)
    sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True),
           torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    l0 = api.launches_total()
    for i in range(a.steps):
        flush.fill_(i & 0xFF)  # L2 flush between timed steps (outside the event pair)
        if G > 1:
            dist.barrier()
        ev[i][0].record(stream)
        render_part()
        ev[i][1].record(stream)
        assemble()
        ev[i][2].record(stream)
    sync_all()
    n_launch = api.launches_total() - l0
    clocks = sampler.stop()
    step_ms = [e[0].elapsed_time(e[2]) for e in ev]
    kern_ms = [e[0].elapsed_time(e[1]) for e in ev]
    total_ms = float(sum(step_ms))
    kernel_ms_avg = float(np.mean(kern_ms))

    # ---- parity of the device-resident frame (outside the timed region) ----
    kat = kat_entry(wl, nsplit)
    parity = {}

    def check_frame(img: np.ndarray, what: str):
        if kat is None:
            parity[what] = None
            return
        if wl.get("kat_step"):
            k = wl["kat_step"] // lat
            z = np.load(os.path.join(GOLDEN, "lattice", wl["kat"].format(nsplit=nsplit) + ".npz"))
            parity[what] = bool((img[::k, ::k] == z["rgb"]).all())
        else:
            parity[what] = api.fnv1a64(img) == kat["image_fnv"]

    sync_all()
    if rank == 0:
        check_frame(image.cpu().numpy(), "device_frame")

    # ---- end-to-end through the public API with HOST buffers: render_cuda(), all N devices driven by rank 0 ----
    def frame_e2e():
        rtc.render(host_tris_np, W, H, spp, bounce, triangles_only=tonly, camera=cam, scene=sky, ngpus=G,
                   split_mode=api.RTC_SPLIT_SAMPLES if samples else api.RTC_SPLIT_TILES, nsplit=nsplit,
                   out=host_img.numpy())

    e2e_s = None
    e2e_launches = 0
    if G > 1:
        torch.cuda.synchronize()
        dist.barrier(group=cpu_group)
    if lat:
        # the lattice is a subset only the device-resident API can ask for (render_cuda renders whole frames): the
        # end-to-end leg is upload + render + gather + download through that API
        def frame_e2e():  # noqa: F811
            dscene.update(host_tris_np, stream=sh)
            render_part()
            assemble()
            if rank == 0:
                host_img.copy_(image, non_blocking=True)
            stream.synchronize()
        for _ in range(2):
            frame_e2e()
        sync_all()
        t0 = time.perf_counter()
        for _ in range(a.steps):
            if G > 1:
                dist.barrier()
            frame_e2e()
        sync_all()
        e2e_s = time.perf_counter() - t0
    else:
        if rank == 0:
            for _ in range(2):
                frame_e2e()
            l1 = api.launches_total()
            t0 = time.perf_counter()
            for _ in range(a.steps):
                frame_e2e()
            e2e_s = time.perf_counter() - t0
            e2e_launches = api.launches_total() - l1
        if G > 1:
            dist.barrier(group=cpu_group)
    if rank == 0:
        check_frame(host_img.numpy(), "e2e_frame")

    # device time of the two render kernels (events recorded by the library around each launch)
    prim_ms, path_ms = [], []
    for _ in range(3):
        render_part(api.RTC_FLAG_TIME_KERNELS)
        try:
            a_ms, b_ms = dscene.kernel_times()
            prim_ms.append(a_ms)
            path_ms.append(b_ms)
        except rtc.RtcError:
            break
    prim_ms = float(np.mean(prim_ms)) if prim_ms else None
    path_ms = float(np.mean(path_ms)) if path_ms else None
    # scheduler counters of the path kernel for one frame
    stats.zero_()
    render_part(0, True)
    torch.cuda.synchronize()
    pool = stats.cpu().numpy()[12:18].astype(np.int64)

    # the same frame in the library's other modes (reported, not the headline)
    modes = {}
    if G == 1 and not a.no_modes and not lat:
        for mname, fl in (("per_pair_filter", api.RTC_FLAG_PER_PAIR_FILTER), ("paths_only", api.RTC_FLAG_PATHS_ONLY),
                          ("exact_only", api.RTC_FLAG_EXACT_ONLY), ("serial_paths", api.RTC_FLAG_SERIAL_PATHS)):
            render_part(fl)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            render_part(fl)
            e1.record(stream)
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            modes[mname] = {"frame_ms": ms, "Mrays_per_s": rays_frame / (ms * 1e-3) / 1e6}
            if mname == "exact_only":
                # the one mode that executes the reference's own operation sequence for every pair: its staged flops are real
                modes[mname]["staged_frac_of_fp32_peak"] = staged_local / (ms * 1e-3) / 1e12 / peak_tflops
            if mname == "paths_only":
                # every pair of the frame goes through the path kernel's sweep: the brute-force figure north_star asks for
                modes[mname]["pairs_per_s"] = float(st_l[1]) / (ms * 1e-3)
                modes[mname]["brute_force_frac"] = brute_local / (ms * 1e-3) / 1e12 / peak_tflops
        render_part(0)
        torch.cuda.synchronize()

    if G > 1:
        t = torch.tensor([total_ms, e2e_s if e2e_s is not None else 0.0, kernel_ms_avg, float(n_launch)],
                         dtype=torch.float64, device=dev)
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        total_ms, e2e_all = float(tmax[0]), float(tmax[1])
        e2e_s = e2e_all if e2e_s is None or lat else e2e_s
        n_launch = int(t[3])
        if peer_ptr:
            api.ipc_close(local, peer_ptr)

    ms_per_step = total_ms / a.steps
    value = rays_frame / (ms_per_step * 1e-3) / 1e6
    line = None
    if rank == 0:
        e2e_value = rays_frame * a.steps / e2e_s / 1e6
        h2d = int(tris.nbytes + 36)
        line = {
            "metric": METRIC,
            "value": value, "unit": "Mrays/s", "n_gpus": G, "steps": a.steps, "warmup": max(a.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32",
            "data": ("synthetic scene (rtc_synth_scene, splitmix64 seed 0xB200; SURVEY 8d defines it)" if wl["scene"] == "synth" else
                     "bundled scene fixture (tests/golden/scenes, produced by the reference's loader)"),
            "config": make_config(name, wl, G, split, a.gather, len(tris), rays_frame),
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": h2d * (G if not lat else 1),
                    "d2h_bytes_per_step": int(full_w * full_h * 3), "ms_per_step": e2e_s / a.steps * 1e3,
                    "api": ("rtc_scene_update + rtc_render_device + gather + D2H (the lattice is a subset only the "
                            "device-resident API renders)" if lat else
                            f"render_cuda(ngpus={G}) with pinned host buffers" +
                            (": one host thread per device, every device copies its bands into the caller's image" if G > 1 and not samples else "")),
                    "gpu_launches": int(e2e_launches)},
            "parity_fnv_ok": (None if not parity or any(v is None for v in parity.values()) else all(parity.values())),
            "parity": parity,
            "gpu_launches": int(n_launch),
            "gpu_launches_how": "counted by the library (rtc_launches_total: every launch site increments), timed region, all ranks",
            "clocks": clocks,
            "frame_ms": ms_per_step,
        }
        # roofline of the dominant kernel.  The path kernel is the brute-force intersection kernel proper: every
        # (bounce ray, triangle) pair goes through its conservative quick test (FMUL/FFMA chains) and the survivors
        # through the reference's arithmetic (non-contracted FMUL/FADD).  The primary kernel decides most pairs with
        # conservative block bounds; its figures are reported without a roofline fraction (they are not arithmetic).
        traffic = {}
        try:
            with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
                traffic = json.load(f).get(name, {})
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
                "this run (2 flop x FFMA/s; MEASURED_PEAKS.json has no FP32 entry); brute_force_frac = 51 flop x pairs; "
                "rank 0's share at N > 1")
        if path_ms:
            line["roofline"] = roof("path kernel (rtc::render_pool_kernel / render_paths_kernel: every bounce ray against every "
                                    "triangle, conservative quick test + exact survivors, + all shading)", path_ms,
                                    staged_bounce, brute_bounce, float(st_b[1]), what)
            # the figure for the intersection kernel alone, beside the headline: every one of these pairs really runs
            line["path_kernel"] = {"ms": path_ms, "bounce_rays": float(st_b[0]), "bounce_Mrays_per_s": float(st_b[0]) / (path_ms * 1e-3) / 1e6,
                                   "pairs": float(st_b[1]), "pairs_per_s": float(st_b[1]) / (path_ms * 1e-3),
                                   "share_of_frame": path_ms / kernel_ms_avg if kernel_ms_avg else None,
                                   "pool": {"samples_evaluated": int(pool[0]), "shade_passes": int(pool[1]),
                                            "lanes_per_shade_pass": float(pool[2]) / max(int(pool[1]), 1),
                                            "sweep_passes": int(pool[3]), "rays_per_sweep_pass": float(pool[4]) / max(int(pool[3]), 1),
                                            "scheduler_iterations": int(pool[5])}}
        if prim_ms:
            line["primary_kernels"] = {"ms": prim_ms, "rays": float(st_p[0]), "pairs_covered": float(st_p[1]),
                                       "note": "conservative block bounds decide most (primary ray, triangle) pairs: a rate, not a "
                                               "roofline fraction (rays * triangles / time would exceed the FP32 peak by design)"}
        if "roofline" not in line:
            line["roofline"] = roof("whole frame (all kernels)", kernel_ms_avg, staged_local, brute_local, float(st_l[1]), what)
        line["roofline"]["nominal_peak"] = 2 * 128 * 148 * 1.965e9 / 1e12
        if modes:
            line["modes"] = modes
        if G == 1 and not a.no_cpu:
            cb = CpuBaseline(name)
            rays = cb.rays()  # also spawns and warms the worker pool
            secs = cb.step()
            line["cpu_baseline"] = {"value": rays / secs / 1e6, "unit": "Mrays/s", "cores": cb.cores, "kind": cb.kind,
                                    "sample": cb.sample_desc(), "seconds": secs}
            line["vs_cpu_baseline"] = {"e2e": e2e_value / (rays / secs / 1e6), "value": value / (rays / secs / 1e6),
                                       "note": "same ray definition on both sides (closest-hit sweeps counted by the restated oracle / by the kernels)"}
            cb.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None and line["parity_fnv_ok"] is False:
        print(json.dumps(line), flush=True)
        raise SystemExit(f"bench.py: the frame differs from the golden vector: {parity}")
    return line


# ---------------------------------------------------------------------------------------
# reference arm: the reference's own CPU implementation of the path on the host cores
# ---------------------------------------------------------------------------------------

def run_reference(a) -> dict | None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None  # rank 0 alone runs the CPU arm
    name = a.workload
    wl = WORKLOADS[name]
    cb = CpuBaseline(name)
    rays = cb.rays()
    warm = max(a.warmup, 3)
    for _ in range(warm):
        cb.step()
    secs = [cb.step() for _ in range(a.steps)]
    cb.close()
    ms = float(np.mean(secs)) * 1e3
    value = rays / (ms * 1e-3) / 1e6
    G = a.gpus
    split = a.split or ("samples" if (wl.get("lattice") and G > 1) else "tiles")
    kat = kat_entry(wl, G if split == "samples" and G > 1 else 1)
    ntris = wl.get("synth") or len(load_tris(wl["scene"]))
    rays_frame = kat.get("rays") if (kat and not wl.get("kat_step")) else None
    return {
        "impl": "reference",
        "metric": METRIC,
        "value": value, "unit": "Mrays/s", "n_gpus": G, "steps": a.steps, "warmup": warm,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": ("synthetic scene (rtc_synth_scene, splitmix64 seed 0xB200; SURVEY 8d defines it)" if wl["scene"] == "synth" else
                 "bundled scene fixture (tests/golden/scenes, produced by the reference's loader)"),
        "config": make_config(name, wl, G, split, a.gather, ntris, rays_frame),
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cb.cores, "kind": cb.kind, "sample": cb.sample_desc(),
                         "rays_per_step": int(rays),
                         "note": "the reference's rayTriangle/calcColor/... compiled from /root/reference as one translation "
                                 "unit (oracle/_ref); each step is a bounded row sample of the frame, all host cores"},
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
    ap.add_argument("--split", default=None, choices=["tiles", "samples"], help="N > 1: how the frame is split (default: tiles; samples for c5l)")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"],
                    help="N > 1, tile split: ranks write into rank 0's frame over NVLink (peer) or NCCL all_gather + un-interleave")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (N=1)")
    ap.add_argument("--no-modes", action="store_true", help="skip timing the other library modes")
    a = ap.parse_args()
    line = run_reference(a) if a.impl == "reference" else run_ours(a)
    if line is not None:
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
