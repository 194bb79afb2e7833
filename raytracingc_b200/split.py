"""Host-side split of one frame over ranks (one process per GPU) -- SURVEY.md section 8e.

Pure bookkeeping: which pixels a rank renders and how the gathered parts become the image.  The reference
interleaves ROWS over its 12 threads for balance (main.c:84); here BAND_H-row bands are interleaved over ranks
(band b belongs to rank b % world), which keeps the warp tiles of the render kernel whole.  Nothing in this
module renders; it is shared by bench.py and the multi-rank tests (gloo on CPU, NCCL on the box).
"""
from __future__ import annotations

import numpy as np

BAND_H = 8  # rows per band: two 4-row warp tiles (rtc_api.cu uses the same constant for ngpus > 1)


def band_subset(rank: int, world: int, band_h: int = BAND_H) -> tuple:
    """RtcSubset (row_start, row_step, col_start, col_step, band_h) of `rank`'s share of a tile split."""
    if world <= 1:
        return (0, 1, 0, 1, 1)
    return (rank * band_h, world * band_h, 0, 1, band_h)


def rows_of(rank: int, world: int, height: int, band_h: int = BAND_H) -> np.ndarray:
    """image rows rank `rank` owns, in the order they appear in its compact part"""
    if world <= 1:
        return np.arange(height)
    y = np.arange(height)
    return y[(y // band_h) % world == rank]


def part_rows(world: int, height: int, band_h: int = BAND_H) -> int:
    """rows of the LARGEST part (rank 0's); every rank pads its part to this for the all-gather"""
    return len(rows_of(0, world, height, band_h))


def uninterleave_host(parts: np.ndarray, height: int, band_h: int = BAND_H) -> np.ndarray:
    """parts [world][part_rows][...] -> image [height][...]; numpy mirror of rtc_uninterleave_device"""
    world = parts.shape[0]
    img = np.empty((height,) + parts.shape[2:], dtype=parts.dtype)
    for g in range(world):
        rows = rows_of(g, world, height, band_h)
        img[rows] = parts[g, :len(rows)]
    return img


def sample_range(g: int, nsplit: int, spp: int) -> tuple:
    """samples [s0, s1) of sub-stream g of a sample split (sub-stream g is seeded x + y*W + g*W*H)"""
    return (g * spp) // nsplit, ((g + 1) * spp) // nsplit


def substreams_of(rank: int, world: int, nsplit: int) -> tuple:
    """(first, count) of the RNG sub-streams rank `rank` renders in a sample split; nsplit % world == 0"""
    if nsplit % world:
        raise ValueError(f"nsplit ({nsplit}) must be a multiple of the rank count ({world})")
    per = nsplit // world
    return rank * per, per


def image_row(rank: int, world: int, j: int, band_h: int = BAND_H) -> int:
    """Image row of compact row `j` of `rank`'s part: where a rank that writes IN PLACE (RtcDeviceJob.rgb_full, the
    frame assembled by the kernels' own stores into one buffer) puts it -- the host mirror of pixel_offset() in
    csrc/rtc_scene.cuh: y = row_start + (j // band_h) * row_step + j % band_h."""
    if world <= 1:
        return j
    return rank * band_h + (j // band_h) * world * band_h + j % band_h
