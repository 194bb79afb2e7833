"""-m "not gpu": the N > 1 path on CPU, world_size 2 and 3 over gloo.

bench.py under torchrun and render_cuda(ngpus > 1) share one rule (raytracingc_b200/split.py): BAND_H-row bands
interleaved over ranks, parts padded to rank 0's size, all-gather, un-interleave.  Here each rank's "render" is
the CPU checker restricted to the rank's rows (the GPU kernel is covered by -m gpu); what is under test is the
host bookkeeping and the collective plumbing: the gathered image must equal the golden single-rank image."""
import os
import socket

import numpy as np
import pytest

from tests import util


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, name, q):
    import torch
    import torch.distributed as dist
    from oracle.oracle_api import Oracle
    from raytracingc_b200 import split
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    e = util.kat()[name]
    job = util.job_of(e)
    W, H = job.width, job.height
    o = Oracle("port")
    o.set_triangles(util.scene_tris(e["scene"]))
    rows = split.rows_of(rank, world, H)
    ph = split.part_rows(world, H)
    part = np.zeros((ph, W, 3), np.uint8)
    full = np.zeros((H, W, 3), np.uint8)
    # bands of this rank: rows y with (y // BAND_H) % world == rank, rendered band by band
    for b0 in range(rank * split.BAND_H, H, world * split.BAND_H):
        for y in range(b0, min(b0 + split.BAND_H, H)):
            job.row_start, job.row_step = y, H  # exactly one row
            o.render(job, want_ids=False, out=(full, None))
    part[:len(rows)] = full[rows]
    gathered = [torch.zeros((ph, W, 3), dtype=torch.uint8) for _ in range(world)]
    dist.all_gather(gathered, torch.from_numpy(part))
    img = split.uninterleave_host(np.stack([g.numpy() for g in gathered]), H)
    want, _ = util.golden_image(name)
    ok = bool((img == want).all())
    flag = torch.tensor([1 if ok else 0])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        q.put((int(flag[0]), int(len(rows))))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,name", [(2, "suze_odd_101x67_5spp_2b"), (3, "default_64x64_16spp_4b")])
def test_band_split_all_gather(world, name, port_oracle):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, name, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    ok, nrows = q.get(timeout=5)
    assert ok == 1
    assert nrows > 0


def _worker_in_place(rank, world, port, name, shm_path, q):
    """every rank writes its rows straight into ONE frame (here: a file in /dev/shm mapped by all ranks) at their final
    addresses -- the host mirror of the peer-write assembly (RtcDeviceJob.rgb_full): no gather, no un-interleave"""
    import torch
    import torch.distributed as dist
    from oracle.oracle_api import Oracle
    from raytracingc_b200 import split
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    e = util.kat()[name]
    job = util.job_of(e)
    W, H = job.width, job.height
    o = Oracle("port")
    o.set_triangles(util.scene_tris(e["scene"]))
    frame = np.memmap(shm_path, dtype=np.uint8, mode="r+", shape=(H, W, 3))
    rows = split.rows_of(rank, world, H)
    scratch = np.zeros((H, W, 3), np.uint8)
    for j in range(len(rows)):
        y = split.image_row(rank, world, j)
        assert y == rows[j]
        job.row_start, job.row_step = y, H
        o.render(job, want_ids=False, out=(scratch, None))
        frame[y] = scratch[y]  # the rank's own store, at the row's final address
    frame.flush()
    dist.barrier()
    if rank == 0:
        want, _ = util.golden_image(name)
        q.put(bool((np.asarray(frame) == want).all()))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,name", [(2, "suze_odd_101x67_5spp_2b"), (4, "default_64x64_16spp_4b")])
def test_band_split_written_in_place(world, name, port_oracle, tmp_path):
    import torch.multiprocessing as mp
    e = util.kat()[name]
    job = util.job_of(e)
    shm = f"/dev/shm/rtc_b200_test_{os.getpid()}_{world}"
    with open(shm, "wb") as f:
        f.truncate(job.width * job.height * 3)
    try:
        ctx = mp.get_context("spawn")
        q = ctx.Queue()
        port = _free_port()
        procs = [ctx.Process(target=_worker_in_place, args=(r, world, port, name, shm, q)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(180)
            assert p.exitcode == 0
        assert q.get(timeout=5) is True
    finally:
        os.unlink(shm)


def test_split_bookkeeping():
    from raytracingc_b200 import split
    for world in (1, 2, 3, 4, 8):
        for H in (8, 9, 67, 240, 2160):
            seen = np.concatenate([split.rows_of(g, world, H) for g in range(world)])
            assert sorted(seen.tolist()) == list(range(H))
            assert split.part_rows(world, H) == max(len(split.rows_of(g, world, H)) for g in range(world))
    # un-interleave is the inverse of the band split
    H, W, world = 67, 5, 3
    img = np.arange(H * W, dtype=np.int32).reshape(H, W)
    ph = split.part_rows(world, H)
    parts = np.zeros((world, ph, W), np.int32)
    for g in range(world):
        rows = split.rows_of(g, world, H)
        parts[g, :len(rows)] = img[rows]
    assert (split.uninterleave_host(parts, H) == img).all()
    # sample split: sub-stream ranges tile [0, spp) and ranks tile the sub-streams
    for spp, G in ((64, 8), (5, 4), (256, 8), (7, 7)):
        cover = []
        for g in range(G):
            s0, s1 = split.sample_range(g, G, spp)
            cover += list(range(s0, s1))
        assert cover == list(range(spp))
    assert [split.substreams_of(r, 4, 8) for r in range(4)] == [(0, 2), (2, 2), (4, 2), (6, 2)]
    with pytest.raises(ValueError):
        split.substreams_of(0, 3, 8)
