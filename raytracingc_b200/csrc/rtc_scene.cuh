// rtc_scene.cuh -- HBM layout of a flattened scene and the kernel argument block.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "rtc_math.cuh"

namespace rtc {

// Flattened scene (one per device).  The reference keeps Triangle[] as a 68-byte
// array of structs and copies each one by value per test (raytracing.c:231);
// here geometry is three 16-byte-aligned float4 per triangle, contiguous (48 B), so one
// warp-uniform 128-bit load brings a third of a triangle and a tile of triangles is ONE
// contiguous range for a bulk (TMA) copy into shared memory:
//     geo[3i+0] = { A.x,  A.y,  A.z,  n.x }
//     geo[3i+1] = { AB.x, AB.y, AB.z, n.y }      AB = B - A   (raytracing.c:191, hoisted:
//     geo[3i+2] = { AC.x, AC.y, AC.z, n.z }      AC = C - A    the same float subtraction)
// 48 B of geometry per triangle are read per (CTA sweep, triangle); the
// material (20 B in the reference) is read once per closest hit:
//     m0[i] = { color.r, color.g, color.b, emissionStrength }    m1[i] = smoothness
// Spheres: s0[k] = { centre.xyz, r }, s1[k] = { color, emission }, s2[k] = smoothness.
struct DeviceScene {
    const float4 *geo;
    const float4 *m0;
    const float *m1;
    int ntris;
    const float4 *s0, *s1;
    const float *s2;
    int nsph;
};

struct RenderArgs {
    DeviceScene sc;
    SkyParams sky;
    f3 origin, ex, ey, ez;
    float fov;
    int W, H;           // full image
    int spp, maxBounce, trianglesOnly;
    int nsplit;         // >= 1
    int split_first;    // blockIdx.z + split_first = sub-stream index
    // pixel subset -> compact output
    int row_start, row_step, col_start, col_step, band_h;
    int out_w, out_h;
    uint8_t *rgb;       // [out_h][out_w][3] or null
    float *accum;       // [planes][out_h][out_w][3] or null
    int32_t *ids;       // [out_h][out_w] or null
    unsigned long long *stats; // RTC_DEVICE_STATS_WORDS or null
    // Row-chunk completion for the pipelined device->host copy (render_cuda): prog[0..K) = primary-hit pixels per
    // chunk of prog_rows output rows (counted by the classification), [K..2K) = finished ones (path kernel),
    // [2K..3K) = flags set to 1 when a chunk's pixels are all final.  null: not tracked.
    unsigned *prog;
    int prog_rows, prog_chunks;
    // 1: rgb is the FULL image (W x H, possibly a peer device's memory): a pixel is stored at its final row and column
    // instead of its compact position -- the ranks of a tile split write straight into the destination frame.
    int rgb_full;
};

// byte offset of compact pixel `pix` (cj * out_w + ci) in a.rgb
__device__ __forceinline__ size_t pixel_offset(const RenderArgs &a, int pix)
{
    if (!a.rgb_full)
        return (size_t)pix * 3;
    const int ci = pix % a.out_w, cj = pix / a.out_w;
    const int x = a.col_start + ci * a.col_step;
    const int y = a.row_start + (cj / a.band_h) * a.row_step + (cj % a.band_h);
    return ((size_t)y * a.W + x) * 3;
}

// stats words
enum { ST_RAYS = 0, ST_TRI_PAIRS = 1, ST_SPH_PAIRS = 2, ST_EXIT0 = 3, /* ..8 */ ST_HIT_PIXELS = 9, ST_SLOW_PAIRS = 10, ST_MISMATCH = 11,
       // scheduler counters of the pool kernel (rtc_pool.cuh): samples evaluated (consumed ones + dropped speculative
       // ones), shade passes and the lanes they filled, sweep passes and the rays they carried, scheduler iterations
       ST_POOL_EVAL = 12, ST_POOL_SHADE_PASSES = 13, ST_POOL_SHADE_LANES = 14, ST_POOL_SWEEP_PASSES = 15,
       ST_POOL_SWEEP_RAYS = 16, ST_POOL_ITERS = 17 };

} // namespace rtc
