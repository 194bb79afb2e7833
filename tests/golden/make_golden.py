#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ from the REFERENCE itself.

Run in the build container, where /root/reference exists:

    make -C oracle && python tests/golden/make_golden.py [--full]

It drives oracle/_ref/libref_oracle.so -- the reference's own C functions,
compiled from /root/reference (see oracle/ref_harness.c) -- and writes

  scenes/<name>.tris   the Triangle[] (68 B records) the reference's loaders
                       produce for its bundled scenes (loadOBJTriangles,
                       raytracing.c:100; parseTriangleFile, raytracing.c:76).
                       The GPU box has no /root/reference, so these are the
                       scene inputs of the -m gpu tests and of bench.py.
  kat.json             known-answer vectors: FNV-1a-64 of the RGB image and of
                       the primary-hit id map, hit-pixel count, ray count, for
                       every config below (the survey's table, SURVEY.md 8c,
                       regenerated -- not copied).
  images/<kat>.npz     the small golden images/id maps themselves.
  lattice/<kat>.npz    for the full-size BASELINE.json configs (--full): the
                       pixels of an 8x8 lattice, enough for a tolerance check
                       without a ten-minute CPU render on the GPU box.

Nothing here is product code; the generating reference build is never shipped.
"""
import argparse
import json
import multiprocessing as mp
import os
import shutil
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.oracle_api import Job, Oracle, fnv1a64_bytes, fnv1a64_words  # noqa: E402

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))

OBJ_SCENES = ["simplest", "suze", "complex", "ultracomplex", "suzannes", "rsuzanne", "cube", "4geoms"]

# name -> (scene, Job kwargs). triangles_only=1 is what the stock CLI does for an
# OBJ (main.c:241); "default" is triangles.txt + the hard-coded sphere.
SMALL = {
    "c1_simplest_320x240_1spp_1b": ("simplest", dict(width=320, height=240, spp=1, max_bounce=1)),
    "simplest_64x48_4000spp_1b": ("simplest", dict(width=64, height=48, spp=4000, max_bounce=1)),
    "suze_160x90_16spp_3b": ("suze", dict(width=160, height=90, spp=16, max_bounce=3)),
    "ultracomplex_160x90_16spp_4b": ("ultracomplex", dict(width=160, height=90, spp=16, max_bounce=4)),
    "complex_160x90_16spp_4b": ("complex", dict(width=160, height=90, spp=16, max_bounce=4)),
    "complex_sphere_160x90_16spp_4b": ("complex", dict(width=160, height=90, spp=16, max_bounce=4, triangles_only=0)),
    "default_64x64_16spp_4b": ("default", dict(width=64, height=64, spp=16, max_bounce=4, triangles_only=0)),
    "default_96x64_8spp_10b": ("default", dict(width=96, height=64, spp=8, max_bounce=10, triangles_only=0)),
    "rsuzanne_128x96_32spp_5b": ("rsuzanne", dict(width=128, height=96, spp=32, max_bounce=5)),
    "suzannes_96x64_4spp_3b": ("suzannes", dict(width=96, height=64, spp=4, max_bounce=3)),
    "suze_closeup_128x96_16spp_6b": ("suze", dict(width=128, height=96, spp=16, max_bounce=6,
                                                  origin=(0.0, -2.0, -6.0), look_at=(0.0, -1.0, 0.0))),
    "suze_odd_101x67_5spp_2b": ("suze", dict(width=101, height=67, spp=5, max_bounce=2)),
    "suze_nsplit4_96x64_16spp_3b": ("suze", dict(width=96, height=64, spp=16, max_bounce=3, nsplit=4)),
    "default_nsplit2_64x48_8spp_4b": ("default", dict(width=64, height=48, spp=8, max_bounce=4, triangles_only=0,
                                                      nsplit=2)),
}

FULL = {
    "c2_suze_1280x720_16spp_3b": ("suze", dict(width=1280, height=720, spp=16, max_bounce=3)),
    "c3_complex_1920x1080_64spp_4b": ("complex", dict(width=1920, height=1080, spp=64, max_bounce=4)),
    "c3s_complex_sphere_1920x1080_64spp_4b": ("complex", dict(width=1920, height=1080, spp=64, max_bounce=4,
                                                              triangles_only=0)),
    "c4_ultracomplex_3840x2160_64spp_4b": ("ultracomplex", dict(width=3840, height=2160, spp=64, max_bounce=4)),
    "default_640x480_16spp_4b": ("default", dict(width=640, height=480, spp=16, max_bounce=4, triangles_only=0)),
}
LATTICE = 8

# SURVEY 8d's extra datapoint: the close-up camera (-p 0 -2 -6 -t 0 -1 0), where most pixels hit geometry and the path
# kernel, not the block bounds, decides the frame (--extra; ~1 CPU-hour)
EXTRA = {
    "c4close_ultracomplex_3840x2160_64spp_4b": ("ultracomplex", dict(width=3840, height=2160, spp=64, max_bounce=4,
                                                                     origin=(0.0, -2.0, -6.0), look_at=(0.0, -1.0, 0.0))),
}

# BASELINE.json config 5 reduced to what a CPU can check: the synthetic 1 M-triangle scene (rtc_synth_scene, seed 0xB200;
# host code of the library, SURVEY 8d defines it) at 16 spp, 4 bounces, on the pixels x % 240 == 0, y % 240 == 0 of the 4K
# frame (16 x 9 pixels: every 60th pixel of the 960 x 540 lattice bench.py --workload c5l renders), for the reference's one
# stream per pixel and for 2, 4, 8 sub-streams (the sample split's own definition) (--c5; ~1 CPU-minute each)
C5_TRIS, C5_SEED, C5_STEP = 1 << 20, 0xB200, 240
C5L = {f"c5l_synth1m_lattice240_16spp_4b_nsplit{g}": dict(width=3840, height=2160, spp=16, max_bounce=4, nsplit=g,
                                                           row_step=C5_STEP, col_step=C5_STEP) for g in (1, 2, 4, 8)}


def load_scene(ref: Oracle, name: str) -> np.ndarray:
    if name == "default":
        tmp = tempfile.mkdtemp()
        try:
            shutil.copy(os.path.join(REF, "triangles.txt"), tmp)  # the loader writes <file>.parsed beside it
            return ref.load_triangles_txt(os.path.join(tmp, "triangles.txt"))
        finally:
            shutil.rmtree(tmp)
    return ref.load_obj(os.path.join(REF, "3Dmodels", name + ".obj"))


def _rows_worker(args):
    scene, kw, k, P = args
    ref = Oracle("reference")
    port = Oracle("port")
    tris = np.fromfile(os.path.join(OUT, "scenes", scene + ".tris"), dtype=np.uint8)
    tris = tris.view(ref_tri_dtype())
    ref.set_triangles(tris)
    port.set_triangles(tris)
    job = Job(row_start=k, row_step=P, **kw)
    rgb, ids, _ = ref.render(job)
    rgb2, ids2, st = port.render(job)
    return k, rgb[k::P], ids[k::P], rgb2[k::P], ids2[k::P], st


def ref_tri_dtype():
    from oracle.oracle_api import TRI_DTYPE
    return TRI_DTYPE


def render_parallel(scene: str, kw: dict, P: int):
    """Rows interleaved over P processes like main.c:84; reference AND port, which must agree."""
    H, W = kw["height"], kw["width"]
    rgb = np.zeros((H, W, 3), np.uint8)
    ids = np.zeros((H, W), np.int32)
    rgb2 = np.zeros_like(rgb)
    ids2 = np.zeros_like(ids)
    stats = None
    with mp.Pool(P) as pool:
        for k, a, b, c, d, st in pool.imap_unordered(_rows_worker, [(scene, kw, k, P) for k in range(P)]):
            rgb[k::P], ids[k::P], rgb2[k::P], ids2[k::P] = a, b, c, d
            if stats is None:
                stats = st
            else:
                for key, v in st.items():
                    stats[key] = [x + y for x, y in zip(stats[key], v)] if isinstance(v, list) else stats[key] + v
    return rgb, ids, rgb2, ids2, stats


def _c5_worker(args):
    name, kw = args
    from raytracingc_b200 import api  # host code only: the scene generator (no GPU needed)
    tris = api.synth_scene(C5_TRIS, C5_SEED)
    ref, port = Oracle("reference"), Oracle("port")
    ref.set_triangles(tris)
    port.set_triangles(tris)
    job = Job(**kw)
    rgb, ids, _ = ref.render(job)
    rgb2, ids2, st = port.render(job)
    return name, rgb[::C5_STEP, ::C5_STEP].copy(), ids[::C5_STEP, ::C5_STEP].copy(), \
        bool((rgb == rgb2).all() and (ids == ids2).all()), st


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--full", action="store_true", help="also the full-size BASELINE.json configs (CPU-minutes)")
    ap.add_argument("--extra", action="store_true", help="only the close-up 4K frame (about a CPU-hour)")
    ap.add_argument("--c5", action="store_true", help="only the lattice vectors of the synthetic 1 M-triangle scene")
    ap.add_argument("--procs", type=int, default=os.cpu_count())
    a = ap.parse_args()

    os.makedirs(os.path.join(OUT, "scenes"), exist_ok=True)
    os.makedirs(os.path.join(OUT, "images"), exist_ok=True)
    os.makedirs(os.path.join(OUT, "lattice"), exist_ok=True)
    ref = Oracle("reference")
    counts = {}
    for name in OBJ_SCENES + ["default"]:
        tris = load_scene(ref, name)
        tris.tofile(os.path.join(OUT, "scenes", name + ".tris"))
        counts[name] = len(tris)
    print("scenes:", counts)

    kat_path = os.path.join(OUT, "kat.json")
    kat = json.load(open(kat_path)) if os.path.exists(kat_path) else {}
    kat["_scenes"] = counts
    kat["_how"] = "tests/golden/make_golden.py from oracle/_ref (reference functions, gcc -O3 -ffp-contract=off)"

    if a.c5:
        with mp.Pool(min(a.procs, len(C5L))) as pool:
            for name, rgb, ids, same, st in pool.imap_unordered(_c5_worker, list(C5L.items())):
                kat[name] = dict(scene=f"synth:{C5_TRIS}:{C5_SEED:#x}", job=C5L[name], image_fnv=fnv1a64_bytes(rgb),
                                 id_fnv=fnv1a64_words(ids), hit_pixels=int((ids != -1).sum()), rays=int(st["rays"]),
                                 port_equals_reference=same, lattice_step=C5_STEP)
                np.savez_compressed(os.path.join(OUT, "lattice", name + ".npz"), step=C5_STEP, rgb=rgb, ids=ids)
                print(f"{name}: image {kat[name]['image_fnv']} hits {kat[name]['hit_pixels']} rays {kat[name]['rays']} "
                      f"port==ref {same}", flush=True)
                json.dump(kat, open(kat_path, "w"), indent=1, sort_keys=True)
        return
    todo = dict(SMALL)
    if a.full:
        todo.update(FULL)
    if a.extra:
        todo = dict(EXTRA)
    for name, (scene, kw) in todo.items():
        t = time.time()
        rgb, ids, rgb2, ids2, st = render_parallel(scene, kw, a.procs)
        same = bool((rgb == rgb2).all() and (ids == ids2).all())
        entry = dict(scene=scene, job=kw, image_fnv=fnv1a64_bytes(rgb), id_fnv=fnv1a64_words(ids),
                     hit_pixels=int((ids != -1).sum()), rays=int(st["rays"]), tri_exit=[int(v) for v in st["tri_exit"]],
                     sphere_tests=int(st["sphere_tests"]), rng_draws=int(st["rng_draws"]),
                     sky_evals=int(st["sky_evals"]), port_equals_reference=same)
        kat[name] = entry
        if name in SMALL:
            np.savez_compressed(os.path.join(OUT, "images", name + ".npz"), rgb=rgb, ids=ids)
        else:
            np.savez_compressed(os.path.join(OUT, "lattice", name + ".npz"), step=LATTICE,
                                rgb=rgb[::LATTICE, ::LATTICE].copy(), ids=ids[::LATTICE, ::LATTICE].copy())
        print(f"{name}: {time.time() - t:.1f}s image {entry['image_fnv']} ids {entry['id_fnv']} "
              f"hits {entry['hit_pixels']} rays {entry['rays']} port==ref {same}", flush=True)
        json.dump(kat, open(kat_path, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
