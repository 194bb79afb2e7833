// rtc_paths.cuh -- the path kernel: one lane = one pixel's RNG stream, exact arithmetic for every ray.
//
// A lane is a state machine over (sample, bounce) of "its" pixel (the stream is serial per pixel: seeded once,
// main.c:95, continuing across samples, main.c:98-99).  The warp iterates
//     sweep the lane's current ray over ALL triangles (calculateRayCollision, raytracing.c:216-240)  ->  shade
// so lanes with different path lengths never wait for each other, and a lane whose pixel is finished fetches
// the next pixel from a global work list (persistent lanes: no tail beyond one pixel's own serial chain).
//
// Work list: either the compact list of primary-hit pixels produced by the sky kernel (rtc_sky.cuh), or, with
// list == nullptr, every pixel of the subset.  A task is (pixel, plane); plane = sub-stream of a sample split.
//
// Geometry (48 B per triangle, rtc_scene.cuh) lives in shared memory:
//   STREAM = false  the whole scene is resident (loaded once per CTA); warps run independently.
//   STREAM = true   scenes of any size: PATH_TILE-triangle tiles are streamed by the TMA unit
//                   (cp.async.bulk + mbarrier, double-buffered); the CTA's warps sweep in lock-step.
#pragma once

#include "rtc_kernels.cuh"
#include "rtc_tma.cuh"

namespace rtc {

constexpr int PATH_THREADS = 128;
constexpr int PATH_TILE = 256; // triangles per streamed tile: 12 KB per stage, 2 stages (occupancy stays register-bound)

struct PathWork {
    const int *list;            // compact pixel indices (cj * out_w + ci) or nullptr = all pixels
    const unsigned *list_count; // device counter written by the sky kernel (used when list != nullptr)
    unsigned *cursor;           // device work cursor, zeroed before the launch
    int planes;                 // sub-streams rendered by this launch
    int write_ids;              // 1: this kernel owns the primary-id map (no sky kernel ran)
    int tile;                   // triangles per streamed tile (<= PATH_TILE), STREAM only
    int coop;                   // 1: cooperative sweep of the last live lanes (0 only for A/B measurements)
    const float2 *rec;          // with a list: {distance, id bits} of each listed pixel's primary hit, found by the
                                // primary kernel, which also re-sweeps that ray every sample; nullptr: the path
                                // kernel sweeps primary rays itself
};

__host__ __device__ inline size_t path_smem_bytes(int ntris, bool stream)
{
    return stream ? (size_t)2 * PATH_TILE * 3 * sizeof(float4) : (size_t)(ntris > 0 ? ntris : 1) * 3 * sizeof(float4);
}

// The reference's ray-triangle test arranged for SIMT, in two phases per chunk of 32 triangles.
//
// Phase 1 (tri_quick, branch-free, warp-uniform shared-memory reads): the reference's own values c = d.n,
// h = d x AC, det = AB.h, s = o - A, s.h -- same operations, same order -- are computed for every pair and
// the pair is dropped when the reference provably drops it:
//   * c >= 0                                  (cull, raytracing.c:189)
//   * |det| < EPS                             (raytracing.c:195; -EPS < det < EPS  <=>  |det| < EPS)
//   * t = s.h with det's sign;  t < -2^-100 |det|   =>  u = fl(s.h * fl(1/det)) < 0, |u| >= 2^-101: cannot round to -0
//   *                           t > |det| * 1.00001  =>  u > 1   (margin >> the 3 roundings of u)
// NaNs never drop a pair (every comparison is written so that an unordered result keeps it).
// Phase 2: each lane walks the set bits of its survivor mask in index order and runs the reference's full
// sequence (tri_test_exact) on them, so distances, ties (first index wins) and ids are the reference's.
__device__ __forceinline__ bool tri_quick(const f3 &o, const f3 &d, const float4 &p0, const float4 &p1, const float4 &p2)
{
    const f3 n = mk3(p0.w, p1.w, p2.w);
    const f3 AB = mk3(p1.x, p1.y, p1.z), AC = mk3(p2.x, p2.y, p2.z);
    const float c = xdot(d, n);
    const f3 h = xcross(d, AC);
    const float det = xdot(AB, h);
    const f3 s = xsub3(o, mk3(p0.x, p0.y, p0.z));
    const float sh = xdot(s, h);
    const float adet = fabsf(det);
    const float t = __int_as_float(__float_as_int(sh) ^ (__float_as_int(det) & 0x80000000)); // s.h * sign(det)
    const float hi = adet * 1.00001f;               // t > hi            =>  u > 1
    const float lo = adet * -7.8886090522101181e-31f; // t < -2^-100 |det| =>  u < 0 and |u| >= 2^-101: never rounds to -0
    bool keep = !(t > hi);
    keep = keep && !(t < lo);
    keep = keep && !(adet < 0.001f);
    keep = keep && !(c >= 0.0f);
    return keep;
}

// Same decision, folded into the survivor mask: the four compares chain through ONE predicate (FSETP ... .AND P)
// and a predicated OR sets the bit, instead of four materialised booleans (FSETP + SEL each) and a shifted select.
template <unsigned BIT>
__device__ __forceinline__ void tri_quick_bit(const f3 &o, const f3 &d, const float4 &p0, const float4 &p1, const float4 &p2,
                                              unsigned &m)
{
    const f3 n = mk3(p0.w, p1.w, p2.w);
    const f3 AB = mk3(p1.x, p1.y, p1.z), AC = mk3(p2.x, p2.y, p2.z);
    const float c = xdot(d, n);
    const f3 h = xcross(d, AC);
    const float det = xdot(AB, h);
    const f3 s = xsub3(o, mk3(p0.x, p0.y, p0.z));
    const float sh = xdot(s, h);
    const float adet = fabsf(det);
    const float t = __int_as_float(__float_as_int(sh) ^ (__float_as_int(det) & 0x80000000));
    const float hi = adet * 1.00001f;
    const float lo = adet * -7.8886090522101181e-31f;
    asm("{\n\t.reg .pred p;\n\t"
        "setp.leu.f32 p, %1, %2;\n\t"                // !(t > hi)
        "setp.geu.and.f32 p, %1, %3, p;\n\t"         // !(t < lo)
        "setp.geu.and.f32 p, %4, 0f3A83126F, p;\n\t" // !(adet < 0.001f)
        "setp.ltu.and.f32 p, %5, 0f00000000, p;\n\t" // !(c >= 0)
        "@p or.b32 %0, %0, %6;\n\t}"
        : "+r"(m)
        : "f"(t), "f"(hi), "f"(lo), "f"(adet), "f"(c), "n"(BIT));
}

template <typename GeoPtr>
__device__ __forceinline__ void sweep_chunked(const GeoPtr S, int n, int base, bool active, const f3 &o, const f3 &d,
                                              float &best, int &bestId)
{
    for (int c0 = 0; c0 < n; c0 += 32) {
        const int cn = min(32, n - c0);
        unsigned m = 0;
        if (active) {
            int j = 0;
            for (; j + 4 <= cn; j += 4) {
                const int t0 = 3 * (c0 + j);
                unsigned m4 = 0;
                tri_quick_bit<1u>(o, d, S[t0], S[t0 + 1], S[t0 + 2], m4);
                tri_quick_bit<2u>(o, d, S[t0 + 3], S[t0 + 4], S[t0 + 5], m4);
                tri_quick_bit<4u>(o, d, S[t0 + 6], S[t0 + 7], S[t0 + 8], m4);
                tri_quick_bit<8u>(o, d, S[t0 + 9], S[t0 + 10], S[t0 + 11], m4);
                m |= m4 << j;
            }
            for (; j < cn; ++j) {
                unsigned m1 = 0;
                tri_quick_bit<1u>(o, d, S[3 * (c0 + j)], S[3 * (c0 + j) + 1], S[3 * (c0 + j) + 2], m1);
                m |= m1 << j;
            }
        }
        while (m) {
            const int j = __ffs(m) - 1;
            m &= m - 1;
            unsigned dummy[1];
            tri_test_exact<false>(o, d, S[3 * (c0 + j)], S[3 * (c0 + j) + 1], S[3 * (c0 + j) + 2], base + c0 + j, best,
                                  bestId, dummy);
        }
    }
}

// Cooperative sweep for the tail of the work list: when a warp has few live lanes left (and no work to fetch),
// the 32 lanes form na groups of S lanes; group q sweeps the ray of the q-th live lane, member r testing the
// triangles j = r, r + S, ... of the tile (exact sequence), and the group's (distance, index) minimum -- lowest
// index among equal distances, which is what the reference's in-order strict-< scan keeps -- goes back to the
// owner.  The serial chain of the last pixels (what bounds a frame split over many GPUs) shrinks by ~S.
struct CoopCtx {
    int S, r, leader; // group size, member index, lane holding this lane's group result after the reduction
    bool member;      // this lane's group has a ray
    f3 o, d;
    float best;
    int id;
};

__device__ __forceinline__ void coop_tile(CoopCtx &c, const float4 *S_, int n, int base)
{
    if (!c.member)
        return;
    for (int j = c.r; j < n; j += c.S) {
        if (tri_quick(c.o, c.d, S_[3 * j], S_[3 * j + 1], S_[3 * j + 2])) {
            unsigned dummy[1];
            tri_test_exact<false>(c.o, c.d, S_[3 * j], S_[3 * j + 1], S_[3 * j + 2], base + j, c.best, c.id, dummy);
        }
    }
}

template <bool STREAM>
__global__ void __launch_bounds__(PATH_THREADS) render_paths_kernel(const RenderArgs a, const PathWork w)
{
    extern __shared__ float4 sGeo[];
    __shared__ uint64_t bars[2];
    const int ntris = a.sc.ntris;
    const int nsph = a.trianglesOnly ? 0 : a.sc.nsph;
    const int lane = threadIdx.x & 31;
    const int tile = w.tile;
    const int ntiles = STREAM ? (ntris + tile - 1) / tile : 1;

    if (!STREAM) {
        for (int i = threadIdx.x; i < 3 * ntris; i += PATH_THREADS)
            sGeo[i] = a.sc.geo[i];
    } else if (threadIdx.x == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        fence_barrier_init();
    }
    __syncthreads();

    const unsigned npix = w.list ? *w.list_count : (unsigned)(a.out_w * a.out_h);
    const unsigned total = npix * (unsigned)w.planes;
    // Group mode: with fewer tasks than lanes (a frame split over many GPUs, a small image) S lanes share one
    // pixel.  They hold identical state (same task, same deterministic shading), each sweeps the triangles
    // j = r, r + S, ... and the group takes the (distance, index) minimum: the pixel's serial chain, which bounds
    // the kernel when every lane has at most one task, gets ~S times shorter in its sweep part.
    int S = 1;
    if (w.coop)
        while (S < 32 && (unsigned long long)total * (unsigned)(2 * S) <= (unsigned long long)gridDim.x * PATH_THREADS)
            S *= 2;
    const int gr = lane & (S - 1), gbase = lane & ~(S - 1);
    const float fh = (float)(a.H / 2);
    const float wgt = (float)(1.0 / (double)a.spp); // main.c:99

    Lane L;
    f3 pdir = mk3(0.f, 0.f, 0.f);
    int pix = 0, plane = 0, s_end = 0;
    bool active = false, exhausted = false, first = false, fresh = false;
    const bool has_rec = w.rec != nullptr;
    float best0 = 999999.0f, best = 999999.0f;
    int id0 = -1, bestId = -1;
    bool pend = false; // (best, bestId) of the last sweep is a hit that is not shaded yet
    unsigned long long nrays = 0, nhitpix = 0;
    uint32_t k0 = 0; // tiles consumed so far (STREAM): stage = k & 1, parity = (k >> 1) & 1

    auto issue = [&](uint32_t k, int t) {
        const int n = min(tile, ntris - t * tile);
        const uint32_t bytes = (uint32_t)n * 3u * (uint32_t)sizeof(float4);
        fence_proxy_async();
        mbar_arrive_expect_tx(&bars[k & 1], bytes);
        tma_load_1d(sGeo + (size_t)(k & 1) * PATH_TILE * 3, a.sc.geo + (size_t)t * tile * 3, bytes, &bars[k & 1]);
    };

    // a path has ended: accumulate it (main.c:99) and start the next sample, or finish the pixel
    auto finish_sample = [&]() {
        L.acc = xadd3(L.acc, xscale3(L.light, wgt));
        L.sample++;
        if (L.sample >= s_end) {
            active = false;
            if (a.accum != nullptr && gr == 0) {
                float *p = a.accum + ((size_t)plane * a.out_w * a.out_h + pix) * 3;
                p[0] = L.acc.x;
                p[1] = L.acc.y;
                p[2] = L.acc.z;
            }
            if (a.rgb != nullptr && a.nsplit == 1 && gr == 0) { // vec3ToColor, raytracing.c:11-15
                uint8_t *p = a.rgb + (size_t)pix * 3;
                p[0] = (uint8_t)float_to_uint8(L.acc.x);
                p[1] = (uint8_t)float_to_uint8(L.acc.y);
                p[2] = (uint8_t)float_to_uint8(L.acc.z);
            }
        } else {
            L.pos = a.origin;
            L.dir = pdir;
            L.tint = mk3(1.f, 1.f, 1.f);
            L.light = mk3(0.f, 0.f, 0.f);
            L.bounce = 0;
            fresh = true;
        }
    };

    while (true) {
        // ---- idle lanes fetch the next (pixel, plane) task; repeated while a lane finishes its pixel in the
        //      shading pass below, so a warp never leaves (or sweeps with idle lanes) while work is left ----
        do {
        {
            const bool need = !active && !exhausted; // identical on the S lanes of a group
            const unsigned m = __ballot_sync(0xffffffffu, need && gr == 0);
            if (m) {
                unsigned base = 0;
                const int leader = __ffs(m) - 1;
                if (lane == leader)
                    base = atomicAdd(w.cursor, (unsigned)__popc(m));
                base = __shfl_sync(0xffffffffu, base, leader);
                if (need) {
                    const unsigned t = base + __popc(m & ((1u << gbase) - 1u)); // one task per group
                    if (t < total) {
                        plane = (int)(t / npix);
                        pix = w.list ? w.list[t % npix] : (int)(t % npix);
                        const int ci = pix % a.out_w, cj = pix / a.out_w;
                        const int x = a.col_start + ci * a.col_step;
                        const int y = a.row_start + (cj / a.band_h) * a.row_step + (cj % a.band_h);
                        const int g = a.split_first + plane;
                        const float dx = xdiv((float)(x - a.W / 2), fh), dy = xdiv((float)(y - a.H / 2), fh);
                        pdir = xnormalize(xadd3(xadd3(xscale3(a.ex, dx), xscale3(a.ey, dy)), xscale3(a.ez, a.fov)));
                        L.pos = a.origin;
                        L.dir = pdir;
                        L.tint = mk3(1.f, 1.f, 1.f);
                        L.light = mk3(0.f, 0.f, 0.f);
                        L.acc = mk3(0.f, 0.f, 0.f);
                        L.rng = (uint32_t)(x + y * a.W) + (uint32_t)g * (uint32_t)(a.W * a.H); // main.c:95
                        L.sample = (int)((long long)g * a.spp / a.nsplit);
                        s_end = (int)((long long)(g + 1) * a.spp / a.nsplit);
                        L.bounce = 0;
                        first = true;
                        fresh = true;
                        if (has_rec) {
                            const float2 r = w.rec[t % npix];
                            best0 = r.x;
                            id0 = __float_as_int(r.y);
                        }
                        active = L.sample < s_end;
                        if (!active) { // empty sample range: the pixel's partial sum is zero
                            if (a.accum != nullptr && gr == 0) {
                                float *p = a.accum + ((size_t)plane * a.out_w * a.out_h + pix) * 3;
                                p[0] = p[1] = p[2] = 0.f;
                            }
                        }
                    } else {
                        exhausted = true;
                    }
                }
            }
        }
        // ---- hit shading, one pass for both kinds of pending hits: the hit the last sweep found, and the known
        //      primary hit of a lane that starts a sample (the primary kernel sweeps that ray every sample).  A lane
        //      whose path ends here starts its next sample in the same loop. ----
        while (active && (pend || (has_rec && fresh))) {
            if (!pend) {
                best = best0;
                bestId = id0;
            }
            pend = false;
            fresh = false;
            if (shade_hit(L, a, best, bestId))
                finish_sample();
        }
        } while (__any_sync(0xffffffffu, !active && !exhausted));
        const bool any = STREAM ? (__syncthreads_or(active) != 0) : __any_sync(0xffffffffu, active);
        if (!any)
            break;

        // ---- closest hit of the lane's current ray: spheres first, then triangles in order ----
        best = 999999.0f;
        bestId = -1;
        if (active)
            for (int k = 0; k < nsph; ++k)
                sphere_test_exact(L.pos, L.dir, __ldg(&a.sc.s0[k]), k, best, bestId);
        // few live lanes and nothing left to fetch: sweep cooperatively (see CoopCtx)
        const unsigned act = __ballot_sync(0xffffffffu, active);
        const int na = __popc(act);
        const bool coop = S == 1 && na > 0 && na <= 16 && w.coop;
        CoopCtx cc;
        if (coop) {
            cc.S = na <= 1 ? 32 : (na <= 2 ? 16 : (na <= 4 ? 8 : (na <= 8 ? 4 : 2)));
            const int q = lane / cc.S;
            cc.r = lane % cc.S;
            cc.member = q < na;
            const int src = cc.member ? (int)__fns(act, 0, q + 1) : lane;
            cc.o = mk3(__shfl_sync(0xffffffffu, L.pos.x, src), __shfl_sync(0xffffffffu, L.pos.y, src),
                       __shfl_sync(0xffffffffu, L.pos.z, src));
            cc.d = mk3(__shfl_sync(0xffffffffu, L.dir.x, src), __shfl_sync(0xffffffffu, L.dir.y, src),
                       __shfl_sync(0xffffffffu, L.dir.z, src));
            cc.best = 999999.0f;
            cc.id = -1;
            cc.leader = (int)__popc(act & ((1u << lane) - 1u)) * cc.S; // meaningful on live lanes
        }
        auto coop_finish = [&]() {
            // (distance, index) minimum over each group; all lanes take part in the shuffles
            for (int o_ = cc.S >> 1; o_ > 0; o_ >>= 1) {
                const float ob = __shfl_xor_sync(0xffffffffu, cc.best, o_);
                const int oi = __shfl_xor_sync(0xffffffffu, cc.id, o_);
                if (oi != -1 && (cc.id == -1 || ob < cc.best || (ob == cc.best && oi < cc.id))) {
                    cc.best = ob;
                    cc.id = oi;
                }
            }
            const float gb = __shfl_sync(0xffffffffu, cc.best, cc.leader & 31);
            const int gi = __shfl_sync(0xffffffffu, cc.id, cc.leader & 31);
            if (active && gi != -1 && gb < best) { // spheres were tested first by the owner: they win ties
                best = gb;
                bestId = gi;
            }
        };
        const bool grouped = S > 1;
        if (grouped) { // the group's lanes hold the same ray; member gr sweeps every S-th triangle
            cc.S = S;
            cc.r = gr;
            cc.member = active;
            cc.o = L.pos;
            cc.d = L.dir;
            cc.best = 999999.0f;
            cc.id = -1;
            cc.leader = lane;
        }
        if (!STREAM) {
            if (coop || grouped) {
                coop_tile(cc, sGeo, ntris, 0);
                coop_finish();
            } else {
                sweep_chunked(sGeo, ntris, 0, active, L.pos, L.dir, best, bestId);
            }
        } else {
            if (threadIdx.x == 0) {
                issue(k0, 0);
                if (ntiles > 1)
                    issue(k0 + 1, 1);
            }
            for (int t = 0; t < ntiles; ++t) {
                const uint32_t k = k0 + (uint32_t)t;
                mbar_wait(&bars[k & 1], (k >> 1) & 1u);
                if (coop || grouped)
                    coop_tile(cc, sGeo + (size_t)(k & 1) * PATH_TILE * 3, min(tile, ntris - t * tile), t * tile);
                else
                    sweep_chunked(sGeo + (size_t)(k & 1) * PATH_TILE * 3, min(tile, ntris - t * tile), t * tile,
                                  active, L.pos, L.dir, best, bestId);
                __syncthreads(); // every warp is done with this stage
                if (threadIdx.x == 0 && t + 2 < ntiles)
                    issue(k + 2, t + 2);
            }
            k0 += (uint32_t)ntiles;
            if (coop || grouped)
                coop_finish();
        }

        // ---- shade (calcColor's loop body), advance the state machine ----
        if (active) {
            nrays += gr == 0;
            if (first) {
                first = false;
                if (w.write_ids && plane == 0 && gr == 0) {
                    if (a.ids != nullptr)
                        a.ids[pix] = bestId;
                    nhitpix += bestId != -1;
                }
            }
            fresh = false;
            if (bestId == -1) {
                shade_miss(L, a);
                finish_sample();
            } else {
                pend = true; // shaded at the top of the next iteration, together with the lanes starting a sample
            }
        }
    }

    if (a.stats != nullptr) {
        unsigned long long r = nrays, hp = nhitpix;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            r += __shfl_xor_sync(0xffffffffu, r, o);
            hp += __shfl_xor_sync(0xffffffffu, hp, o);
        }
        if (lane == 0 && r) {
            atomicAdd(&a.stats[ST_RAYS], r);
            atomicAdd(&a.stats[ST_TRI_PAIRS], r * (unsigned long long)ntris);
            atomicAdd(&a.stats[ST_SPH_PAIRS], r * (unsigned long long)nsph);
            atomicAdd(&a.stats[ST_SLOW_PAIRS], r * (unsigned long long)ntris);
            if (w.write_ids && a.ids != nullptr)
                atomicAdd(&a.stats[ST_HIT_PIXELS], hp);
        }
    }
}

} // namespace rtc
