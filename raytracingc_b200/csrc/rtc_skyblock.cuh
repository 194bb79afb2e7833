// rtc_skyblock.cuh -- primary-ray kernel, block form (the default).
//
// Owns EVERY primary ray of the frame: the first sample classifies the pixels (primary-hit pixels go to the path
// kernel's work list together with their hit record {distance, id}; sky pixels are finished here), and every
// sample re-sweeps every pixel's primary ray, as the reference does (main.c:98-99), cross-checking the result
// against the classification.  The path kernel (rtc_paths.cuh) then only sweeps bounce rays.  Compared with
// render_sky_kernel (rtc_sky.cuh) the difference is the granularity of the conservative test.
// The four edge functions of a (camera, triangle) pair are affine in the pixel's (dx, dy), so over a block of pixels
//         max E_k  <=  E_k(block centre) + |a_k| * half_width + |b_k| * half_height ,
// and the right-hand side (with the rounding margins of rtc_fast.cuh folded into its constant) is ONE pair of
// FMAs.  A lane owns an 8 x 4 pixel block: if some E_k is negative at that bound, none of the block's 32 primary
// rays can pass the reference's tests for this triangle (two-level conservative rasterisation: nothing is
// approximated, every decision that reaches the image is taken by the reference's exact sequence).
// One level above, lane j first bounds triangle j of each 32-triangle chunk over the warp's whole 64 x 16 region
// (same inequality, the spread of the block centres added), so only triangles near the warp are block-tested.
// Blocks that survive are expanded by the whole warp, one lane per pixel: per-pixel edge functions, then
// tri_test_exact on the survivors, in triangle order (first index wins ties, raytracing.c:232).
//
// Warp = 8 x 4 lane-blocks = 64 x 16 pixels; CTA = 4 warps stacked in y = 64 x 64 pixels.
#pragma once

#include "rtc_sky.cuh"

namespace rtc {

constexpr int BLK_W = 8, BLK_H = 4;                 // pixels per lane-block
#ifndef RTC_SKYB_WARPS // warps per CTA of the block kernel (8 KB of hit records each): A/B knob
#define RTC_SKYB_WARPS 4
#endif
constexpr int SKYB_WARPS = RTC_SKYB_WARPS, SKYB_THREADS = 32 * SKYB_WARPS;
constexpr int SKYB_CHUNK = 32;                      // triangles per candidate mask

__host__ __device__ inline size_t skyblock_smem_bytes(int ntris, bool stream)
{
    return sky_smem_bytes(ntris, stream) + (size_t)SKYB_WARPS * 32 * 32 * sizeof(float2);
}

__global__ void __launch_bounds__(SKYB_THREADS) render_skyblock_kernel(const RenderArgs a,
                                                                       const float4 *__restrict__ tab,
                                                                       const SkyWork w)
{
    extern __shared__ float4 smem_dyn[];
    __shared__ uint64_t bars[2];
    const int ntris = a.sc.ntris;
    const int nsph = a.trianglesOnly ? 0 : a.sc.nsph; // <= 32 (host-checked)
    const bool stream = w.stream != 0;
    const int tile = stream ? w.tile : (ntris > 0 ? ntris : 1);
    const int ntiles = (ntris + tile - 1) / tile;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float4 *sF = smem_dyn;
    float2 *hitbuf = reinterpret_cast<float2 *>(sF + (stream ? 2 * SKY_TILE * TAB_F4 : TAB_F4 * (ntris > 0 ? ntris : 1))) +
                     (size_t)warp * 32 * 32; // [lane-block][pixel] {best, id}

    if (!stream) {
        for (int i = threadIdx.x; i < TAB_F4 * ntris; i += SKYB_THREADS)
            sF[i] = tab[i];
    } else if (threadIdx.x == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        fence_barrier_init();
    }
    __syncthreads();

    // this lane's block
    const int ci0 = (blockIdx.x * 8 + (lane & 7)) * BLK_W;
    const int cj0 = ((blockIdx.y * SKYB_WARPS + warp) * 4 + (lane >> 3)) * BLK_H;
    // blockIdx.z = plane * segs + segment: segment 0 owns the classification sample and everything that is written;
    // the other segments re-sweep their share of the remaining samples (counted and cross-checked, never written)
    const int segs = w.segs > 0 ? w.segs : 1;
    const int plane = (int)blockIdx.z / segs, seg = (int)blockIdx.z % segs;
    const int g = a.split_first + plane;
    unsigned valid = 0;
#pragma unroll
    for (int p = 0; p < 32; ++p)
        if (ci0 + (p & 7) < a.out_w && cj0 + (p >> 3) < a.out_h)
            valid |= 1u << p;

    const float fh = (float)(a.H / 2);
    auto x_of = [&](int ci) { return a.col_start + ci * a.col_step; };
    auto y_of = [&](int cj) { return a.row_start + (cj / a.band_h) * a.row_step + (cj % a.band_h); };
    auto dx_at = [&](int ci) { return xdiv((float)(x_of(ci) - a.W / 2), fh); };
    auto dy_at = [&](int cj) { return xdiv((float)(y_of(cj) - a.H / 2), fh); };
    // block centre in (dx, dy): midpoint of the first and last pixel centres
    const float cx = 0.5f * (dx_at(ci0) + dx_at(ci0 + BLK_W - 1));
    const float cy = 0.5f * (dy_at(cj0) + dy_at(cj0 + BLK_H - 1));
    // the warp's region: block centres span [wcx - ehw, wcx + ehw] x [wcy - ehh, wcy + ehh] (inflated for rounding)
    const float cx_lo = __shfl_sync(0xffffffffu, cx, 0), cx_hi = __shfl_sync(0xffffffffu, cx, 7);
    const float cy_lo = __shfl_sync(0xffffffffu, cy, 0), cy_hi = __shfl_sync(0xffffffffu, cy, 24);
    const float wcx = 0.5f * (cx_lo + cx_hi), wcy = 0.5f * (cy_lo + cy_hi);
    const float ehw = 0.5f * (cx_hi - cx_lo) * 1.001f + 1e-6f * (1.0f + fabsf(wcx));
    const float ehh = 0.5f * (cy_hi - cy_lo) * 1.001f + 1e-6f * (1.0f + fabsf(wcy));

    const int s0 = (int)((long long)g * a.spp / a.nsplit), s1 = (int)((long long)(g + 1) * a.spp / a.nsplit);
    // this segment's samples [sa, sb): segment 0 takes the first share (classification included)
    const int sa = s0 + (int)((long long)(s1 - s0) * seg / segs), sb = s0 + (int)((long long)(s1 - s0) * (seg + 1) / segs);
    const float wgt = (float)(1.0 / (double)a.spp);
    unsigned hits0 = 0;    // classification: pixels of this block whose primary ray hits
    bool touched0 = false; // classification: some triangle passed this block's bound
    unsigned badblocks = 0;
    unsigned long long nrays = 0, nslow = 0, nhit = 0;
    uint32_t k0 = 0;

    auto issue = [&](uint32_t k, int t) {
        const int n = min(tile, ntris - t * tile);
        const uint32_t bytes = (uint32_t)n * (uint32_t)TAB_F4 * (uint32_t)sizeof(float4);
        fence_proxy_async();
        mbar_arrive_expect_tx(&bars[k & 1], bytes);
        tma_load_1d(sF + (size_t)(k & 1) * SKY_TILE * TAB_F4, tab + (size_t)t * tile * TAB_F4, bytes, &bars[k & 1]);
    };

    for (int s = sa; s < sb; ++s) {
        bool touched = false; // this block's hitbuf entries are initialised for this sample
        unsigned hits = 0;    // pixels of this block whose primary ray hit (this sample)
        if (stream && threadIdx.x == 0) {
            issue(k0, 0);
            if (ntiles > 1)
                issue(k0 + 1, 1);
        }
        const bool warp_busy = __any_sync(0xffffffffu, valid != 0);
        // ---- spheres come first in the reference's closest-hit loop (raytracing.c:219-228): bound the
        //      discriminant over the block (prepare_spheres_kernel), expand the blocks a sphere may touch ----
        if (nsph > 0 && warp_busy) {
            unsigned smask = 0;
            for (int k = 0; k < nsph; ++k) {
                const float *q = w.sphq + 8 * k;
                const float A = __ldg(q), B = __ldg(q + 1), C = __ldg(q + 2), E = __ldg(q + 3), F_ = __ldg(q + 4),
                            G = __ldg(q + 5);
                const float q0 = fmaf(cx, fmaf(A, cx, fmaf(C, cy, E)), fmaf(cy, fmaf(B, cy, F_), G));
                const float qx = fmaf(2.0f * A, cx, fmaf(C, cy, E)), qy = fmaf(2.0f * B, cy, fmaf(C, cx, F_));
                const float bound = fmaf(fabsf(qx), w.hw, fmaf(fabsf(qy), w.hh, q0));
                smask |= !(bound < 0.0f) ? (1u << k) : 0u;
            }
            if (valid == 0 || touched0)
                smask = 0;
            if (s != sa) {
                badblocks += smask != 0;
                smask = 0;
            } else if (seg != 0 && smask != 0) {
                touched = true; // a block with candidates: every sample after the classification is the block-task kernel's
                smask = 0;
            }
            unsigned pending = __ballot_sync(0xffffffffu, smask != 0);
            while (pending) {
                const int src = __ffs(pending) - 1;
                pending &= pending - 1;
                unsigned m = __shfl_sync(0xffffffffu, smask, src);
                const int bci0 = __shfl_sync(0xffffffffu, ci0, src), bcj0 = __shfl_sync(0xffffffffu, cj0, src);
                const unsigned bvalid = __shfl_sync(0xffffffffu, valid, src);
                const bool mine = (bvalid >> lane) & 1u;
                const f3 d = primary_dir(a, dx_at(bci0 + (lane & 7)), dy_at(bcj0 + (lane >> 3)));
                float best = 999999.0f; // the sphere stage is the first to touch a block in a sample
                int id = -1;
                while (m) {
                    const int k = __ffs(m) - 1;
                    m &= m - 1;
                    if (mine)
                        sphere_test_exact(a.origin, d, __ldg(&a.sc.s0[k]), k, best, id);
                }
                hitbuf[src * 32 + lane] = make_float2(best, __int_as_float(id));
                const unsigned hm = __ballot_sync(0xffffffffu, id != -1 && mine);
                if (lane == src) {
                    touched = true;
                    hits = hm;
                }
                __syncwarp();
            }
        }
        for (int t = 0; t < ntiles; ++t) {
            const float4 *F = sF;
            int n = ntris, tbase = 0;
            if (stream) {
                const uint32_t k = k0 + (uint32_t)t;
                mbar_wait(&bars[k & 1], (k >> 1) & 1u);
                F = sF + (size_t)(k & 1) * SKY_TILE * TAB_F4;
                n = min(tile, ntris - t * tile);
                tbase = t * tile;
            }
            if (warp_busy) {
                for (int c0 = 0; c0 < n; c0 += SKYB_CHUNK) {
                    const int cn = min(SKYB_CHUNK, n - c0);
                    // ---- level 1: lane j bounds triangle c0 + j over the WARP's whole 64 x 16 pixel region ----
                    bool wpass = false;
                    if (lane < cn) {
                        const float4 *f = F + TAB_F4 * (c0 + lane);
                        const float4 f0 = f[0], f1 = f[1], f2 = f[2], f3_ = f[3];
                        const float e0 = fmaf(fabsf(f0.x), ehw, fmaf(fabsf(f0.y), ehh, fmaf(f0.x, wcx, fmaf(f0.y, wcy, f0.z))));
                        const float e1 = fmaf(fabsf(f1.x), ehw, fmaf(fabsf(f1.y), ehh, fmaf(f1.x, wcx, fmaf(f1.y, wcy, f1.z))));
                        const float e2 = fmaf(fabsf(f2.x), ehw, fmaf(fabsf(f2.y), ehh, fmaf(f2.x, wcx, fmaf(f2.y, wcy, f2.z))));
                        const float e3 =
                            fmaf(fabsf(f3_.x), ehw, fmaf(fabsf(f3_.y), ehh, fmaf(f3_.x, wcx, fmaf(f3_.y, wcy, f3_.z))));
                        wpass = (__float_as_int(e0) | __float_as_int(e1) | __float_as_int(e2) | __float_as_int(e3)) >= 0;
                    }
                    unsigned wm = __ballot_sync(0xffffffffu, wpass);
                    // ---- level 2: every lane bounds the surviving triangles over its own 8 x 4 pixel block ----
                    unsigned mask = 0;
                    while (wm) {
                        const int j = __ffs(wm) - 1;
                        wm &= wm - 1;
                        const float4 *f = F + TAB_F4 * (c0 + j);
                        const float4 f0 = f[0], f1 = f[1], f2 = f[2], f3_ = f[3];
                        const float e0 = fmaf(f0.x, cx, fmaf(f0.y, cy, f0.z));
                        const float e1 = fmaf(f1.x, cx, fmaf(f1.y, cy, f1.z));
                        const float e2 = fmaf(f2.x, cx, fmaf(f2.y, cy, f2.z));
                        const float e3 = fmaf(f3_.x, cx, fmaf(f3_.y, cy, f3_.z));
                        const int bits = __float_as_int(e0) | __float_as_int(e1) | __float_as_int(e2) | __float_as_int(e3);
                        mask |= (bits >= 0) ? (1u << j) : 0u;
                    }
                    if (valid == 0 || touched0)
                        mask = 0; // a block with candidates at the first sample is re-swept by the block-task kernel
                    if (s != sa) {
                        // same rays, same arithmetic: a block without candidates at the first sample has none now
                        badblocks += mask != 0;
                        mask = 0;
                    } else if (seg != 0 && mask != 0) {
                        touched = true; // (see the sphere stage)
                        mask = 0;
                    }
                    // ---- first sample: expand surviving blocks, the warp takes one block at a time, one lane
                    //      per pixel ----
                    unsigned pending = __ballot_sync(0xffffffffu, mask != 0);
                    while (pending) {
                        const int src = __ffs(pending) - 1;
                        pending &= pending - 1;
                        unsigned m = __shfl_sync(0xffffffffu, mask, src);
                        const int bci0 = __shfl_sync(0xffffffffu, ci0, src);
                        const int bcj0 = __shfl_sync(0xffffffffu, cj0, src);
                        const unsigned bsky = __shfl_sync(0xffffffffu, valid, src);
                        const bool btouched = __shfl_sync(0xffffffffu, (int)touched, src) != 0;
                        const bool mine = (bsky >> lane) & 1u; // this lane's pixel of the block is swept
                        const float dxr = dx_at(bci0 + (lane & 7)), dyr = dy_at(bcj0 + (lane >> 3));
                        float2 *hb = hitbuf + src * 32 + lane;
                        float best = 999999.0f;
                        int id = -1;
                        if (btouched) {
                            const float2 h = *hb;
                            best = h.x;
                            id = __float_as_int(h.y);
                        }
                        f3 d = primary_dir(a, dxr, dyr);
                        while (m) {
                            const int j = __ffs(m) - 1;
                            m &= m - 1;
                            const int ti = tbase + c0 + j;
                            const float4 *f = F + TAB_F4 * (c0 + j);
                            const float4 f0 = f[0], f1 = f[1], f2 = f[2], f3_ = f[3];
                            const float t0 = fmaf(f0.x, dxr, fmaf(f0.y, dyr, f0.w));
                            const float t1 = fmaf(f1.x, dxr, fmaf(f1.y, dyr, f1.w));
                            const float t2 = fmaf(f2.x, dxr, fmaf(f2.y, dyr, f2.w));
                            const float t3 = fmaf(f3_.x, dxr, fmaf(f3_.y, dyr, f3_.w));
                            const int bits =
                                __float_as_int(t0) | __float_as_int(t1) | __float_as_int(t2) | __float_as_int(t3);
                            if (bits >= 0 && mine) {
                                unsigned dummy[1];
                                tri_test_exact<false>(a.origin, d, __ldg(&a.sc.geo[3 * ti]), __ldg(&a.sc.geo[3 * ti + 1]),
                                                      __ldg(&a.sc.geo[3 * ti + 2]), ti, best, id, dummy);
                                nslow++;
                            }
                        }
                        *hb = make_float2(best, __int_as_float(id));
                        const unsigned hm = __ballot_sync(0xffffffffu, id != -1 && mine);
                        if (lane == src) {
                            touched = true;
                            hits = hm; // best only improves: a later chunk's ballot includes earlier hits
                            mask = 0;
                        }
                        __syncwarp();
                    }
                }
            }
            if (stream) {
                __syncthreads();
                if (threadIdx.x == 0 && t + 2 < ntiles)
                    issue(k0 + (uint32_t)t + 2, t + 2);
            }
        }
        k0 += (uint32_t)ntiles;
        if ((s == s0 && seg == 0) || !(s == sa ? touched : touched0))
            nrays += __popc(valid); // (the other samples of a block with candidates are the block-task kernel's)

        if (s == sa) {
            // ---- classification: ids are final, hit pixels go to the path kernel's list ----
            if (g == a.split_first && seg == 0) {
                if (a.ids != nullptr) {
                    for (int b = 0; b < 32; ++b) { // lane-block b of this warp, one lane per pixel
                        const unsigned bvalid = __shfl_sync(0xffffffffu, valid, b);
                        if (bvalid == 0)
                            continue;
                        const bool bt = __shfl_sync(0xffffffffu, (int)touched, b) != 0;
                        const int bci0 = __shfl_sync(0xffffffffu, ci0, b), bcj0 = __shfl_sync(0xffffffffu, cj0, b);
                        const int id = bt ? __float_as_int(hitbuf[b * 32 + lane].y) : -1;
                        if ((bvalid >> lane) & 1u)
                            a.ids[(size_t)(bcj0 + (lane >> 3)) * a.out_w + bci0 + (lane & 7)] = id;
                    }
                }
                nhit = __popc(hits);
                if (w.list != nullptr) {
                    unsigned cnt = __popc(hits), incl = cnt;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        unsigned v = __shfl_up_sync(0xffffffffu, incl, o);
                        if (lane >= o)
                            incl += v;
                    }
                    const unsigned tot = __shfl_sync(0xffffffffu, incl, 31);
                    if (a.prog != nullptr && cnt) // prog_rows is a multiple of the block height: one chunk per block
                        atomicAdd(&a.prog[cj0 / a.prog_rows], cnt);
                    if (tot) {
                        unsigned base = 0;
                        if (lane == 31)
                            base = atomicAdd(w.list_count, tot);
                        base = __shfl_sync(0xffffffffu, base, 31);
                        unsigned off = base + incl - cnt, hm = hits;
                        while (hm) {
                            const int p = __ffs(hm) - 1;
                            hm &= hm - 1;
                            w.list[off] = (cj0 + (p >> 3)) * a.out_w + ci0 + (p & 7);
                            w.rec[off] = hitbuf[lane * 32 + p]; // {distance, id} of the primary hit
                            ++off;
                        }
                    }
                }
            }
            hits0 = hits;
            touched0 = touched;
            // blocks with candidate triangles: their remaining samples are re-swept pixel by pixel by the
            // block-task kernel (dynamic work queue: these are the expensive blocks, a tile full of them would
            // otherwise be one warp's serial work)
            if (seg == 0 && touched && s + 1 < s1 && w.btask != nullptr) {
                const unsigned pos = atomicAdd(w.btask_count, 1u);
                w.btask[pos] = make_int4(ci0, cj0, (int)hits, plane);
            }
        }
    }

    // ---- sky pixels: spp identical contributions (same ray, no RNG draw), the reference's serial float adds;
    //      one lane per pixel so the serial chains of 32 pixels run side by side ----
    const unsigned sky = seg == 0 ? valid & ~hits0 : 0u;
    for (int b = 0; b < 32; ++b) {
        const unsigned bsky = __shfl_sync(0xffffffffu, sky, b);
        if (bsky == 0)
            continue;
        const int bci0 = __shfl_sync(0xffffffffu, ci0, b), bcj0 = __shfl_sync(0xffffffffu, cj0, b);
        if (!((bsky >> lane) & 1u))
            continue;
        const int ci = bci0 + (lane & 7), cj = bcj0 + (lane >> 3);
        f3 d = primary_dir(a, dx_at(ci), dy_at(cj));
        f3 env = environment_light(d, a.sky);
        f3 light = xadd3(mk3(0.f, 0.f, 0.f), xmul3(env, mk3(1.f, 1.f, 1.f))); // raytracing.c:291
        f3 c = xscale3(light, wgt);
        f3 acc = mk3(0.f, 0.f, 0.f);
        for (int s = s0; s < s1; ++s)
            acc = xadd3(acc, c);
        const size_t pix = (size_t)cj * a.out_w + ci;
        if (a.accum != nullptr) {
            float *p = a.accum + ((size_t)plane * a.out_w * a.out_h + pix) * 3;
            p[0] = acc.x;
            p[1] = acc.y;
            p[2] = acc.z;
        }
        if (a.rgb != nullptr && a.nsplit == 1) {
            uint8_t *p = a.rgb + pixel_offset(a, (int)pix);
            p[0] = (uint8_t)float_to_uint8(acc.x);
            p[1] = (uint8_t)float_to_uint8(acc.y);
            p[2] = (uint8_t)float_to_uint8(acc.z);
        }
    }

    if (a.stats != nullptr) {
        unsigned long long r_ = nrays, hp = nhit, sl = nslow, bad = badblocks;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            r_ += __shfl_xor_sync(0xffffffffu, r_, o);
            hp += __shfl_xor_sync(0xffffffffu, hp, o);
            sl += __shfl_xor_sync(0xffffffffu, sl, o);
            bad += __shfl_xor_sync(0xffffffffu, bad, o);
        }
        if (lane == 0) {
            atomicAdd(&a.stats[ST_RAYS], r_);
            atomicAdd(&a.stats[ST_TRI_PAIRS], r_ * (unsigned long long)ntris);
            atomicAdd(&a.stats[ST_SPH_PAIRS], r_ * (unsigned long long)nsph);
            atomicAdd(&a.stats[ST_SLOW_PAIRS], sl);
            if (a.ids != nullptr)
                atomicAdd(&a.stats[ST_HIT_PIXELS], hp);
            if (bad)
                atomicAdd(&a.stats[ST_MISMATCH], bad);
        }
    }
}

// ---- block-task kernel: the remaining samples of the blocks that have candidate triangles -------------------
// Work queue filled by render_skyblock_kernel's classification; persistent warps fetch one (block, plane) task
// at a time.  The warp owns the block for all its samples: lane p = pixel p (direction computed once), and per
// sample every triangle is bounded over the block (lane j bounds triangle j of each 32-triangle chunk, the same
// inequality as above), candidates run the per-pixel edge functions and tri_test_exact in triangle order.  The
// per-sample result is compared with the classification (it is the same ray); a difference is counted.
struct BlockTaskWork {
    const int4 *btask;
    const unsigned *btask_count;
    unsigned *cursor;
    int tile, stream;
};

constexpr int BTASK_THREADS = 128;

__global__ void __launch_bounds__(BTASK_THREADS) render_blocktask_kernel(const RenderArgs a,
                                                                         const float4 *__restrict__ tab,
                                                                         const BlockTaskWork w)
{
    extern __shared__ float4 sF[];
    __shared__ uint64_t bars[2];
    const int ntris = a.sc.ntris;
    const int nsph = a.trianglesOnly ? 0 : a.sc.nsph;
    const bool stream = w.stream != 0;
    const int tile = stream ? w.tile : (ntris > 0 ? ntris : 1);
    const int ntiles = (ntris + tile - 1) / tile;
    const int lane = threadIdx.x & 31;
    if (!stream) {
        for (int i = threadIdx.x; i < TAB_F4 * ntris; i += BTASK_THREADS)
            sF[i] = tab[i];
    } else if (threadIdx.x == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        fence_barrier_init();
    }
    __syncthreads();
    const unsigned ntasks = *w.btask_count;
    const float fh = (float)(a.H / 2);
    auto x_of = [&](int ci) { return a.col_start + ci * a.col_step; };
    auto y_of = [&](int cj) { return a.row_start + (cj / a.band_h) * a.row_step + (cj % a.band_h); };
    auto dx_at = [&](int ci) { return xdiv((float)(x_of(ci) - a.W / 2), fh); };
    auto dy_at = [&](int cj) { return xdiv((float)(y_of(cj) - a.H / 2), fh); };
    auto issue = [&](uint32_t k, int t) {
        const int n = min(tile, ntris - t * tile);
        const uint32_t bytes = (uint32_t)n * (uint32_t)TAB_F4 * (uint32_t)sizeof(float4);
        fence_proxy_async();
        mbar_arrive_expect_tx(&bars[k & 1], bytes);
        tma_load_1d(sF + (size_t)(k & 1) * SKY_TILE * TAB_F4, tab + (size_t)t * tile * TAB_F4, bytes, &bars[k & 1]);
    };

    // task state (warp-uniform except the per-pixel values)
    bool have = false, exhausted = false;
    int s = 0, s_end = 0;
    unsigned hits0 = 0;
    bool pvalid = false;
    float cx = 0.f, cy = 0.f, dxr = 0.f, dyr = 0.f;
    f3 d = mk3(0.f, 0.f, 0.f);
    unsigned long long nrays = 0, nslow = 0, bad = 0;
    uint32_t k0 = 0;

    while (true) {
        if (!have && !exhausted) {
            unsigned t = 0;
            if (lane == 0)
                t = atomicAdd(w.cursor, 1u);
            t = __shfl_sync(0xffffffffu, t, 0);
            if (t < ntasks) {
                const int4 bt = w.btask[t];
                const int ci0 = bt.x, cj0 = bt.y, g = a.split_first + bt.w;
                hits0 = (unsigned)bt.z;
                s = (int)((long long)g * a.spp / a.nsplit) + 1; // the first sample was the classification
                s_end = (int)((long long)(g + 1) * a.spp / a.nsplit);
                cx = 0.5f * (dx_at(ci0) + dx_at(ci0 + BLK_W - 1));
                cy = 0.5f * (dy_at(cj0) + dy_at(cj0 + BLK_H - 1));
                const int ci = ci0 + (lane & 7), cj = cj0 + (lane >> 3);
                pvalid = ci < a.out_w && cj < a.out_h;
                dxr = dx_at(ci);
                dyr = dy_at(cj);
                d = primary_dir(a, dxr, dyr);
                have = s < s_end;
            } else {
                exhausted = true;
            }
        }
        const bool any = stream ? (__syncthreads_or(have) != 0) : have;
        if (!any)
            break;
        // ---- one sample of this warp's block ----
        float best = 999999.0f;
        int id = -1;
        if (have && pvalid)
            for (int k = 0; k < nsph; ++k) // spheres first (raytracing.c:219-228)
                sphere_test_exact(a.origin, d, __ldg(&a.sc.s0[k]), k, best, id);
        if (stream && threadIdx.x == 0) {
            issue(k0, 0);
            if (ntiles > 1)
                issue(k0 + 1, 1);
        }
        for (int t = 0; t < ntiles; ++t) {
            const float4 *F = sF;
            int n = ntris, tbase = 0;
            if (stream) {
                const uint32_t k = k0 + (uint32_t)t;
                mbar_wait(&bars[k & 1], (k >> 1) & 1u);
                F = sF + (size_t)(k & 1) * SKY_TILE * TAB_F4;
                n = min(tile, ntris - t * tile);
                tbase = t * tile;
            }
            if (have) {
                for (int c0 = 0; c0 < n; c0 += 32) {
                    bool bpass = false;
                    if (c0 + lane < n) {
                        const float4 *f = F + TAB_F4 * (c0 + lane);
                        const float4 f0 = f[0], f1 = f[1], f2 = f[2], f3_ = f[3];
                        const float e0 = fmaf(f0.x, cx, fmaf(f0.y, cy, f0.z));
                        const float e1 = fmaf(f1.x, cx, fmaf(f1.y, cy, f1.z));
                        const float e2 = fmaf(f2.x, cx, fmaf(f2.y, cy, f2.z));
                        const float e3 = fmaf(f3_.x, cx, fmaf(f3_.y, cy, f3_.z));
                        bpass = (__float_as_int(e0) | __float_as_int(e1) | __float_as_int(e2) | __float_as_int(e3)) >= 0;
                    }
                    unsigned m = __ballot_sync(0xffffffffu, bpass);
                    while (m) {
                        const int j = __ffs(m) - 1;
                        m &= m - 1;
                        const int ti = tbase + c0 + j;
                        const float4 *f = F + TAB_F4 * (c0 + j);
                        const float4 f0 = f[0], f1 = f[1], f2 = f[2], f3_ = f[3];
                        const float t0 = fmaf(f0.x, dxr, fmaf(f0.y, dyr, f0.w));
                        const float t1 = fmaf(f1.x, dxr, fmaf(f1.y, dyr, f1.w));
                        const float t2 = fmaf(f2.x, dxr, fmaf(f2.y, dyr, f2.w));
                        const float t3 = fmaf(f3_.x, dxr, fmaf(f3_.y, dyr, f3_.w));
                        const int bits = __float_as_int(t0) | __float_as_int(t1) | __float_as_int(t2) | __float_as_int(t3);
                        if (bits >= 0 && pvalid) {
                            unsigned dummy[1];
                            tri_test_exact<false>(a.origin, d, __ldg(&a.sc.geo[3 * ti]), __ldg(&a.sc.geo[3 * ti + 1]),
                                                  __ldg(&a.sc.geo[3 * ti + 2]), ti, best, id, dummy);
                            nslow++;
                        }
                    }
                }
            }
            if (stream) {
                __syncthreads();
                if (threadIdx.x == 0 && t + 2 < ntiles)
                    issue(k0 + (uint32_t)t + 2, t + 2);
            }
        }
        k0 += (uint32_t)ntiles;
        if (have) {
            const unsigned hm = __ballot_sync(0xffffffffu, id != -1 && pvalid);
            if (lane == 0)
                bad += hm != hits0;
            nrays += pvalid ? 1 : 0;
            if (++s >= s_end)
                have = false;
        }
    }

    if (a.stats != nullptr) {
        unsigned long long r_ = nrays, sl = nslow, b_ = bad;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            r_ += __shfl_xor_sync(0xffffffffu, r_, o);
            sl += __shfl_xor_sync(0xffffffffu, sl, o);
            b_ += __shfl_xor_sync(0xffffffffu, b_, o);
        }
        if (lane == 0 && (r_ | b_)) {
            atomicAdd(&a.stats[ST_RAYS], r_);
            atomicAdd(&a.stats[ST_TRI_PAIRS], r_ * (unsigned long long)ntris);
            atomicAdd(&a.stats[ST_SPH_PAIRS], r_ * (unsigned long long)nsph);
            atomicAdd(&a.stats[ST_SLOW_PAIRS], sl);
            if (b_)
                atomicAdd(&a.stats[ST_MISMATCH], b_);
        }
    }
}

} // namespace rtc
