// rtc_pool.cuh -- the path kernel, second form: warp-private task pools, samples of a pixel evaluated out of order.
//
// WHY samples may be evaluated out of order although the reference's RNG stream is serial per pixel (seeded once,
// main.c:95, continuing across samples, main.c:98-99): every sample of a pixel re-traces the SAME primary ray
// (main.c:94-99), a hit-bounce draws exactly 7 random numbers (6 for RandomDiretion, moremath.c:98-109, 1 for the
// roulette, raytracing.c:285) and a miss draws none (raytracing.c:289-293).  So a sample's light is a pure function of
// the RNG state it starts from, and every reachable start state is the pixel's seed advanced by 7*j steps of the LCG
// (one multiply-add per step of 7).  With T[j] = (light, h_j) for "the sample that starts at offset j", h_j = its
// number of hit-bounces, the pixel is the pointer walk  j <- j + h_j  summed in sample order (the reference's float
// sum).  The T[j] are independent: they are evaluated speculatively, in parallel; a sample with h != 1 merely SKIPS
// offsets (their evaluations are dropped), nothing else is invalidated.  tools/pool_model.py is an executable model of
// the scheduler below, checked against the serial walk.
//
// HOW: a warp owns a pool in shared memory: POOL_NS evaluation slots (one path each: ray, throughput, light, RNG
// state, offset) and POOL_NC pixel contexts (accumulator, walk pointer j, next unassigned offset, primary ray + hit).
// The warp alternates between
//     consume   finished slots are consumed by their pixel in offset order (acc += light, j += h); slots whose offset
//               was skipped are freed; a pixel whose last sample is consumed is written
//     refill    new pixels from the global work list while the warp holds fewer than its share; every context with
//               nothing outstanding gets a slot (that alone is the serial stream of the reference); slots left over
//               become speculative children at the next unassigned offsets of the contexts that can still use them
//     shade     up to 32 slots whose ray hit something: calcColor's loop body (raytracing.c:272-287), one per lane
//     sweep     up to 64 slots with a ray to trace: closest hit over ALL triangles (calculateRayCollision,
//               raytracing.c:216-240), TWO rays per lane sharing every warp-uniform record load; misses are shaded
//               (sky) on the spot
// so both passes run with full warps whatever the paths' lengths are, the FP64 shading of one warp overlaps the FP32
// sweeps of the others, and a frame with few pixels per GPU (a frame split over 8 GPUs) spreads each pixel's samples
// over the idle slots instead of running them one after the other.  Ray counts are added at consumption, so
// RenderStats.rays is the reference's count whatever was speculated.
#pragma once

#include <type_traits>

#include "rtc_paths.cuh"

namespace rtc {

#ifndef RTC_POOL_NS
#define RTC_POOL_NS 96
#endif
#ifndef RTC_POOL_NC
#define RTC_POOL_NC 64
#endif
// Warps per CTA: 8 (two CTAs per SM) while the resident scene leaves room for two sets of pools, else 16 (one CTA per SM;
// always when the scene is streamed: the tiles' double buffer is shared by more warps).  16 warps x 96 slots per SM either way.
constexpr int POOL_WARPS_SMALL = 8, POOL_WARPS_LARGE = 16;
constexpr int POOL_NS = RTC_POOL_NS; // evaluation slots per warp (multiple of 32, <= 96 + 32 * k is fine)
constexpr int POOL_NC = RTC_POOL_NC; // pixel contexts per warp (multiple of 32, <= 256: the context index has 8 bits)
constexpr int POOL_SPL = POOL_NS / 32, POOL_CPL = POOL_NC / 32; // slots / contexts owned per lane
static_assert(POOL_NS % 32 == 0 && POOL_NC % 32 == 0 && POOL_NC <= 256 && POOL_NS <= 255, "pool geometry");
constexpr int POOL_MAX_BOUNCE = 255; // bounce and hit counters are 8-bit fields of the slot's meta word

// slot fields (one array of POOL_NS words each)
enum { E_PX, E_PY, E_PZ, E_DX, E_DY, E_DZ, E_TX, E_TY, E_TZ, E_LX, E_LY, E_LZ, E_RNG, E_BEST, E_ID, E_OFF, E_META, E_FIELDS };
// context fields (one array of POOL_NC words each)
enum { C_PIX, C_PLANE, C_REM, C_TOTAL, C_J, C_TOP, C_TSTATE, C_AX, C_AY, C_AZ, C_DX, C_DY, C_DZ, C_BEST0, C_ID0, C_FIELDS };
constexpr int POOL_WORDS = E_FIELDS * POOL_NS + C_FIELDS * POOL_NC + (2 * POOL_NS + POOL_NC + 3) / 4 + 1;

// meta word of a slot: [1:0] state, [2] first sweep of the pixel (ids), [10:3] context, [18:11] bounce, [26:19] hits
enum { SL_FREE = 0, SL_SHADE = 1, SL_SWEEP = 2, SL_DONE = 3 };
constexpr unsigned SL_FIRST = 4u;
__device__ __forceinline__ unsigned meta_make(unsigned st, unsigned ctx, unsigned bounce, unsigned hits, unsigned first = 0)
{
    return st | first | (ctx << 3) | (bounce << 11) | (hits << 19);
}

// 7 steps of the LCG of moremath.c:91 as one multiply-add
__host__ __device__ constexpr uint32_t lcg_mul(int steps)
{
    uint32_t m = 1u;
    for (int i = 0; i < steps; ++i)
        m *= 747796405u;
    return m;
}
__host__ __device__ constexpr uint32_t lcg_add(int steps)
{
    uint32_t c = 0u;
    for (int i = 0; i < steps; ++i)
        c = c * 747796405u + 2891336453u;
    return c;
}
constexpr uint32_t LCG7_M = lcg_mul(7), LCG7_C = lcg_add(7);

struct PoolWork {
    const int *list;            // compact pixel indices of the primary-hit pixels, or nullptr = every pixel
    const unsigned *list_count; // device counter written by the primary kernel (used when list != nullptr)
    const float2 *rec;          // {distance, id bits} of each listed pixel's primary hit; nullptr: this kernel sweeps primary rays
    unsigned *cursor;           // device work cursor, zeroed before the launch
    unsigned *error;            // set to 1 when a warp's scheduler trips its watchdog (a bug: never in a correct build)
    int planes;                 // sub-streams rendered by this launch
    int write_ids;              // 1: this kernel owns the primary-id map (no primary kernel ran)
    int tile;                   // triangles per streamed tile (<= PATH_TILE), STREAM only
    int spec;                   // 1: speculative children allowed (needs rec)
    int ctx_share_div;          // a warp holds at most ceil(tasks / (warps * div)) pixels at a time (>= 1, <= POOL_NC)
    const float4 *qrec;         // MODE >= 1: quick records
    float obound;
    unsigned long long watchdog; // scheduler iterations per warp after which the kernel gives up
};

__host__ __device__ inline size_t pool_smem_bytes(int ntris, bool stream, int mode, int warps)
{
    return path_smem_bytes(ntris, stream, mode) + (size_t)warps * POOL_WORDS * sizeof(uint32_t);
}

// ---- the sweep of NR rays per lane over one resident scene / one streamed tile -------------------------------
// Same two phases as sweep_chunked (rtc_paths.cuh): branch-free conservative quick test of every pair, then the
// reference's exact sequence on the survivors in index order.  With NR = 2 every warp-uniform LDS.128 of a pair
// record feeds both rays (a uniform LDS.128 costs two shared-memory wavefronts whatever it carries: the pipe that
// bounds the one-ray form).
template <int MODE, int NR, int UNROLL>
__device__ __forceinline__ void sweep_multi(const float4 *S, const float4 *geo, float obound, int n, int base,
                                            const bool (&act)[NR], const f3 (&o)[NR], const f3 (&d)[NR], float (&best)[NR],
                                            int (&bestId)[NR], unsigned &viol)
{
    bool rayok[NR];
    QuickRay2 q2[NR];
#pragma unroll
    for (int r = 0; r < NR; ++r) {
        rayok[r] = MODE == 0 || quick_ray_ok(o[r], d[r], obound);
        q2[r] = quick_ray2(quick_ray(o[r], d[r]));
    }
    for (int c0 = 0; c0 < n; c0 += 32) {
        const int cn = min(32, n - c0);
        const unsigned full = cn == 32 ? 0xffffffffu : (1u << cn) - 1u;
        unsigned m[NR];
#pragma unroll
        for (int r = 0; r < NR; ++r)
            m[r] = 0;
        if (MODE == 0) {
            int j = 0;
#pragma unroll 1
            for (; j + 2 <= cn; j += 2) {
                const int t0 = 3 * (c0 + j);
#pragma unroll
                for (int r = 0; r < NR; ++r) {
                    unsigned m2 = 0;
                    tri_quick_bit<1u>(o[r], d[r], S[t0], S[t0 + 1], S[t0 + 2], m2);
                    tri_quick_bit<2u>(o[r], d[r], S[t0 + 3], S[t0 + 4], S[t0 + 5], m2);
                    m[r] |= m2 << j;
                }
            }
            for (; j < cn; ++j) {
#pragma unroll
                for (int r = 0; r < NR; ++r) {
                    unsigned m1 = 0;
                    tri_quick_bit<1u>(o[r], d[r], S[3 * (c0 + j)], S[3 * (c0 + j) + 1], S[3 * (c0 + j) + 2], m1);
                    m[r] |= m1 << j;
                }
            }
        } else {
            static_assert(MODE == 0 || mode_packed(MODE), "the pool kernel evaluates quick records in packed form");
            constexpr int UP = UNROLL >= 2 ? UNROLL / 2 : 1; // pair records per iteration
            const float4 *P = S + QPAIR_F4 * (c0 >> 1);
            const int np = (cn + 1) >> 1;
            int j = 0;
#pragma unroll 1
            for (; j + UP <= np; j += UP) {
#pragma unroll
                for (int r = 0; r < NR; ++r) {
                    unsigned mg = 0;
                    quick2_pairs<0, UP>(q2[r], P + QPAIR_F4 * j, mg);
                    m[r] |= mg << (2 * j);
                }
            }
#pragma unroll 1
            for (; j < np; ++j) {
#pragma unroll
                for (int r = 0; r < NR; ++r) {
                    unsigned m1 = 0;
                    quick2_pairs<0, 1>(q2[r], P + QPAIR_F4 * j, m1);
                    m[r] |= m1 << (2 * j);
                }
            }
        }
#pragma unroll
        for (int r = 0; r < NR; ++r) {
            m[r] &= full; // an odd count's last pair is completed by a record that never survives
            if (!rayok[r])
                m[r] = full;
            if (!act[r])
                m[r] = 0;
        }
        const float4 *G = MODE == 0 ? S : geo;
        if (mode_verify(MODE)) {
#pragma unroll
            for (int r = 0; r < NR; ++r)
                if (act[r])
                    for (int j = 0; j < cn; ++j) {
                        const int i = c0 + j;
                        unsigned ex[6] = {0, 0, 0, 0, 0, 0};
                        float b_ = 999999.0f;
                        int id_ = -1;
                        tri_test_exact<true>(o[r], d[r], G[3 * i], G[3 * i + 1], G[3 * i + 2], base + i, b_, id_, ex);
                        if (ex[5] && !((m[r] >> j) & 1u))
                            viol++;
                    }
        }
#pragma unroll
        for (int r = 0; r < NR; ++r) {
            unsigned mm = m[r];
            while (mm) {
                const int i = c0 + __ffs(mm) - 1;
                mm &= mm - 1;
                unsigned dummy[1];
                tri_test_exact<false>(o[r], d[r], G[3 * i], G[3 * i + 1], G[3 * i + 2], base + i, best[r], bestId[r], dummy);
            }
        }
    }
}

// ---- the kernel ---------------------------------------------------------------------------------------------
template <bool STREAM, int MODE, int WARPS>
__global__ void __launch_bounds__(32 * WARPS, WARPS == POOL_WARPS_SMALL ? 2 : 1) render_pool_kernel(const RenderArgs a, const PoolWork w)
{
    constexpr int POOL_THREADS = 32 * WARPS;
    extern __shared__ float4 sGeo[];
    __shared__ uint64_t bars[2];
    const int ntris = a.sc.ntris;
    const int nsph = a.trianglesOnly ? 0 : a.sc.nsph;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    const int cap = STREAM ? PATH_TILE : ntris;
    const int qoff = (int)path_q_f4(MODE, cap);
    constexpr int STAGE_F4 = (int)path_q_f4(MODE, PATH_TILE) + 3 * PATH_TILE;
    const int tile = w.tile;
    const int ntiles = STREAM ? (ntris + tile - 1) / tile : 1;
    // pools live behind the scene stage(s)
    uint32_t *pool = reinterpret_cast<uint32_t *>(sGeo + path_smem_bytes(ntris, STREAM, MODE) / sizeof(float4)) +
                     (size_t)warp * POOL_WORDS;
    float *Ef = reinterpret_cast<float *>(pool);
    uint32_t *Eu = pool;
    float *Cf = reinterpret_cast<float *>(pool + E_FIELDS * POOL_NS);
    uint32_t *Cu = pool + E_FIELDS * POOL_NS;
    unsigned char *freelist = reinterpret_cast<unsigned char *>(pool + E_FIELDS * POOL_NS + C_FIELDS * POOL_NC);
    unsigned char *stagelist = freelist + POOL_NS;
    unsigned char *ctxfree = stagelist + POOL_NS;
#define SLF(f, s) Ef[(f) * POOL_NS + (s)]
#define SLU(f, s) Eu[(f) * POOL_NS + (s)]
#define CXF(f, c) Cf[(f) * POOL_NC + (c)]
#define CXU(f, c) Cu[(f) * POOL_NC + (c)]

    if (!STREAM) {
        for (int i = threadIdx.x; i < qoff; i += POOL_THREADS)
            sGeo[i] = w.qrec[i];
        for (int i = threadIdx.x; i < 3 * ntris; i += POOL_THREADS)
            sGeo[qoff + i] = a.sc.geo[i];
    } else if (threadIdx.x == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        fence_barrier_init();
    }
#pragma unroll
    for (int k = 0; k < POOL_SPL; ++k)
        SLU(E_META, lane + 32 * k) = SL_FREE;
#pragma unroll
    for (int q = 0; q < POOL_CPL; ++q) {
        CXU(C_REM, lane + 32 * q) = 0u;
        CXU(C_J, lane + 32 * q) = 0xffffffffu;
    }
    __syncthreads();

    const unsigned npix = w.list ? *w.list_count : (unsigned)(a.out_w * a.out_h);
    const unsigned total = npix * (unsigned)w.planes;
    if (a.prog != nullptr && blockIdx.x == 0 && (int)threadIdx.x < a.prog_chunks && a.prog[threadIdx.x] == 0)
        *(volatile unsigned *)&a.prog[2 * a.prog_chunks + threadIdx.x] = 1u; // chunks without hit pixels are final already
    const bool has_rec = w.rec != nullptr;
    const bool spec = has_rec && w.spec != 0;
    const unsigned nwarps = gridDim.x * WARPS;
    // pixels this warp holds at a time: its share of the list, so that a short list is spread over all warps (the
    // free slots then run the pixels' samples in parallel) and a long one is taken a few pixels at a time
    int maxctx = (int)min((unsigned long long)POOL_NC,
                          ((unsigned long long)total + (unsigned long long)nwarps * w.ctx_share_div - 1) /
                              ((unsigned long long)nwarps * w.ctx_share_div));
    if (maxctx < 1)
        maxctx = 1;
    if (!spec)
        maxctx = POOL_NC;
    const float fh = (float)(a.H / 2);
    const float wgt = (float)(1.0 / (double)a.spp); // main.c:99

    bool exhausted = total == 0;
    unsigned seen = 0; // the list position after this warp's last fetch
    unsigned long long nrays = 0, nhitpix = 0, iters = 0;
    unsigned long long st_eval = 0, st_shade_pass = 0, st_shade_lanes = 0, st_sweep_pass = 0, st_sweep_rays = 0;
    unsigned viol = 0;
    uint32_t k0 = 0; // tiles consumed so far (STREAM)
    bool tripped = false;

    auto issue = [&](uint32_t k, int t) {
        const int n = min(tile, ntris - t * tile);
        const uint32_t bq = (uint32_t)path_q_f4(MODE, n) * (uint32_t)sizeof(float4), bg = (uint32_t)n * 3u * (uint32_t)sizeof(float4);
        float4 *stage = sGeo + (size_t)(k & 1) * STAGE_F4;
        fence_proxy_async();
        mbar_arrive_expect_tx(&bars[k & 1], bq + bg);
        if (MODE != 0)
            tma_load_1d(stage, w.qrec + path_q_f4(MODE, t * tile), bq, &bars[k & 1]);
        tma_load_1d(stage + qoff, a.sc.geo + (size_t)t * tile * 3, bg, &bars[k & 1]);
    };

    // the pixel's last sample is consumed: write it (vec3ToColor, raytracing.c:11-15)
    auto finish_pixel = [&](int c) {
        const int pix = (int)CXU(C_PIX, c), plane = (int)CXU(C_PLANE, c);
        const f3 acc = mk3(CXF(C_AX, c), CXF(C_AY, c), CXF(C_AZ, c));
        if (a.accum != nullptr) {
            float *p = a.accum + ((size_t)plane * a.out_w * a.out_h + pix) * 3;
            p[0] = acc.x;
            p[1] = acc.y;
            p[2] = acc.z;
        }
        if (a.rgb != nullptr && a.nsplit == 1) {
            uint8_t *p = a.rgb + pixel_offset(a, pix);
            p[0] = (uint8_t)float_to_uint8(acc.x);
            p[1] = (uint8_t)float_to_uint8(acc.y);
            p[2] = (uint8_t)float_to_uint8(acc.z);
            if (a.prog != nullptr) { // the chunk's last finished pixel releases its rows to the copy engine
                const int ch = (pix / a.out_w) / a.prog_rows, K = a.prog_chunks;
                __threadfence();
                if (atomicAdd(&a.prog[K + ch], 1u) + 1u == a.prog[ch]) {
                    __threadfence_system();
                    *(volatile unsigned *)&a.prog[2 * K + ch] = 1u;
                }
            }
        }
    };

    // ---- consume: finished slots in offset order, skipped offsets dropped -----------------------------------
    auto consume = [&]() {
        while (true) {
            int act[POOL_SPL]; // 0 nothing, 1 cancel, 2 consume
            bool any = false;
#pragma unroll
            for (int k = 0; k < POOL_SPL; ++k) {
                const int s = lane + 32 * k;
                const unsigned m = SLU(E_META, s);
                act[k] = 0;
                if ((m & 3u) != SL_FREE) {
                    const unsigned j = CXU(C_J, (m >> 3) & 255u), off = SLU(E_OFF, s);
                    if (off < j)
                        act[k] = 1; // a skipped offset, or the pixel is finished (j = 0xffffffff)
                    else if ((m & 3u) == SL_DONE && off == j)
                        act[k] = 2;
                    any = any || act[k] != 0;
                }
            }
            if (!__any_sync(0xffffffffu, any))
                break;
            __syncwarp();
#pragma unroll
            for (int k = 0; k < POOL_SPL; ++k) {
                const int s = lane + 32 * k;
                if (act[k] == 2) { // one slot per context at most: offsets of a context are distinct
                    const unsigned m = SLU(E_META, s);
                    const int c = (int)((m >> 3) & 255u);
                    const f3 acc = xadd3(mk3(CXF(C_AX, c), CXF(C_AY, c), CXF(C_AZ, c)), mk3(SLF(E_LX, s), SLF(E_LY, s), SLF(E_LZ, s))); // main.c:99
                    CXF(C_AX, c) = acc.x;
                    CXF(C_AY, c) = acc.y;
                    CXF(C_AZ, c) = acc.z;
                    nrays += SLU(E_BEST, s); // sweeps of this sample
                    const unsigned nj = CXU(C_J, c) + (spec ? ((m >> 19) & 255u) : 1u);
                    if (nj >= CXU(C_TOP, c)) {
                        CXU(C_TOP, c) = nj;
                        CXU(C_TSTATE, c) = SLU(E_RNG, s); // the state after this sample = the state at offset nj
                    }
                    const unsigned rem = CXU(C_REM, c) - 1u;
                    CXU(C_REM, c) = rem;
                    CXU(C_J, c) = nj;
                    if (rem == 0u) {
                        finish_pixel(c);
                        CXU(C_J, c) = 0xffffffffu; // the context's other slots drop out at the next pass
                    }
                }
                if (act[k] != 0)
                    SLU(E_META, s) = SL_FREE;
            }
            __syncwarp();
        }
    };

    // ---- refill -------------------------------------------------------------------------------------------
    auto refill = [&]() {
        unsigned fb[POOL_SPL];
        int nfree = 0;
#pragma unroll
        for (int k = 0; k < POOL_SPL; ++k) {
            fb[k] = __ballot_sync(0xffffffffu, (SLU(E_META, lane + 32 * k) & 3u) == SL_FREE);
            if ((fb[k] >> lane) & 1u)
                freelist[nfree + __popc(fb[k] & lt)] = (unsigned char)(lane + 32 * k);
            nfree += __popc(fb[k]);
        }
        if (nfree == 0)
            return;
        // contexts: live ones, and those with nothing outstanding
        unsigned lb[POOL_CPL];
        int nlive = 0, nd1 = 0, nfreectx = 0;
#pragma unroll
        for (int q = 0; q < POOL_CPL; ++q) {
            const int c = lane + 32 * q;
            const bool live = CXU(C_REM, c) != 0u;
            lb[q] = __ballot_sync(0xffffffffu, live);
            nd1 += __popc(__ballot_sync(0xffffffffu, live && CXU(C_TOP, c) == CXU(C_J, c)));
            if (!live)
                ctxfree[nfreectx + __popc(~lb[q] & lt)] = (unsigned char)c;
            nlive += __popc(lb[q]);
            nfreectx += __popc(~lb[q]);
        }
        __syncwarp(); // freelist / ctxfree are read by other lanes below
        // new pixels from the list
        int nnew = min(min(nfree - nd1, maxctx - nlive), nfreectx);
        if (!exhausted && nnew > 0) {
            // guided: towards the end of the list the warps take fewer pixels at a time
            const unsigned left = total - min(seen, total);
            const unsigned guide = max(1u, left / (2u * nwarps));
            nnew = (int)min((unsigned)nnew, guide);
            unsigned base = 0;
            if (lane == 0)
                base = atomicAdd(w.cursor, (unsigned)nnew);
            base = __shfl_sync(0xffffffffu, base, 0);
            seen = base + (unsigned)nnew;
            if (base >= total) {
                exhausted = true;
                nnew = 0;
            } else if (base + (unsigned)nnew > total) {
                nnew = (int)(total - base);
            }
            for (int i = lane; i < nnew; i += 32) {
                const int c = ctxfree[i];
                const unsigned t = base + (unsigned)i;
                const int plane = (int)(t / npix);
                const unsigned li = t % npix;
                const int pix = w.list ? w.list[li] : (int)li;
                const int ci = pix % a.out_w, cj = pix / a.out_w;
                const int x = a.col_start + ci * a.col_step;
                const int y = a.row_start + (cj / a.band_h) * a.row_step + (cj % a.band_h);
                const int g = a.split_first + plane;
                const float dx = xdiv((float)(x - a.W / 2), fh), dy = xdiv((float)(y - a.H / 2), fh); // main.c:88-94
                const f3 pdir = xnormalize(xadd3(xadd3(xscale3(a.ex, dx), xscale3(a.ey, dy)), xscale3(a.ez, a.fov)));
                const int s_begin = (int)((long long)g * a.spp / a.nsplit), s_end = (int)((long long)(g + 1) * a.spp / a.nsplit);
                const unsigned rem = (unsigned)(s_end - s_begin);
                CXU(C_PIX, c) = (unsigned)pix;
                CXU(C_PLANE, c) = (unsigned)plane;
                CXU(C_REM, c) = rem;
                CXU(C_TOTAL, c) = rem;
                CXU(C_J, c) = 0u;
                CXU(C_TOP, c) = 0u;
                CXU(C_TSTATE, c) = (uint32_t)(x + y * a.W) + (uint32_t)g * (uint32_t)(a.W * a.H); // main.c:95
                CXF(C_AX, c) = 0.f;
                CXF(C_AY, c) = 0.f;
                CXF(C_AZ, c) = 0.f;
                CXF(C_DX, c) = pdir.x;
                CXF(C_DY, c) = pdir.y;
                CXF(C_DZ, c) = pdir.z;
                if (has_rec) {
                    const float2 r = w.rec[li];
                    CXF(C_BEST0, c) = r.x;
                    CXU(C_ID0, c) = __float_as_uint(r.y);
                }
                if (rem == 0u) { // empty sample range: the partial sum is zero, nothing to trace
                    if (a.accum != nullptr) {
                        float *p = a.accum + ((size_t)plane * a.out_w * a.out_h + pix) * 3;
                        p[0] = p[1] = p[2] = 0.f;
                    }
                    CXU(C_J, c) = 0xffffffffu;
                }
            }
            __syncwarp();
        }
        // grants: one slot for every live context with nothing outstanding, then speculative children
        int g1[POOL_CPL], g2[POOL_CPL], want2[POOL_CPL];
        int n1 = 0, nwant2 = 0, mine1 = 0;
#pragma unroll
        for (int q = 0; q < POOL_CPL; ++q) {
            const int c = lane + 32 * q;
            const unsigned rem = CXU(C_REM, c), j = CXU(C_J, c), top = CXU(C_TOP, c);
            const bool live = rem != 0u;
            g1[q] = (live && top == j) ? 1 : 0;
            want2[q] = 0;
            if (spec && live) {
                // offsets below j + rem can still be needed; a pixel that keeps skipping (closed scenes) runs serially
                const unsigned consumed = CXU(C_TOTAL, c) - rem, skipped = j - consumed;
                if (skipped * 2u <= consumed + 4u)
                    want2[q] = max(0, (int)(j + rem - top) - g1[q]);
            }
            n1 += __popc(__ballot_sync(0xffffffffu, g1[q] != 0));
            nwant2 += __popc(__ballot_sync(0xffffffffu, want2[q] > 0));
            mine1 += g1[q];
        }
        const int left = nfree - n1;
        const int quota = (left > 0 && nwant2 > 0) ? (left + nwant2 - 1) / nwant2 : 0;
        int mine2 = 0;
#pragma unroll
        for (int q = 0; q < POOL_CPL; ++q) {
            g2[q] = min(want2[q], quota);
            mine2 += g2[q];
        }
        // positions in the free list: class 1 first, then class 2, each in (lane, q) order
        int p1 = mine1, p2 = mine2;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v1 = __shfl_up_sync(0xffffffffu, p1, o), v2 = __shfl_up_sync(0xffffffffu, p2, o);
            if (lane >= o) {
                p1 += v1;
                p2 += v2;
            }
        }
        const int tot2 = __shfl_sync(0xffffffffu, p2, 31);
        p1 -= mine1;              // exclusive
        p2 = n1 + (p2 - mine2);   // exclusive, behind class 1
        const int nassigned = min(nfree, n1 + tot2);
#pragma unroll
        for (int q = 0; q < POOL_CPL; ++q) {
            const int c = lane + 32 * q;
            const int ga = (g1[q] && p1 < nfree) ? 1 : 0;
            const int gb = max(0, min(g2[q], nfree - p2));
            if (ga + gb > 0) {
                unsigned top = CXU(C_TOP, c), ts = CXU(C_TSTATE, c);
                for (int i = 0; i < ga + gb; ++i) {
                    const int s = freelist[i < ga ? p1 : p2 + (i - ga)];
                    SLU(E_OFF, s) = top;
                    SLU(E_RNG, s) = ts;
                    SLU(E_META, s) = 0x80000000u | ((unsigned)c << 3); // to be initialised below
                    top += 1u;
                    if (spec)
                        ts = ts * LCG7_M + LCG7_C; // the state 7 draws later = the next offset's
                }
                CXU(C_TOP, c) = top;
                CXU(C_TSTATE, c) = ts;
            }
            p1 += g1[q];
            p2 += g2[q];
        }
        __syncwarp();
        // initialise the assigned slots, one per lane at a time
        for (int i = lane; i < nassigned; i += 32) {
            const int s = freelist[i];
            const unsigned m = SLU(E_META, s);
            if (!(m & 0x80000000u))
                continue; // (a class-1 position that was not granted: cannot happen, n1 <= nfree is not guaranteed)
            const int c = (int)((m >> 3) & 255u);
            SLF(E_PX, s) = a.origin.x;
            SLF(E_PY, s) = a.origin.y;
            SLF(E_PZ, s) = a.origin.z;
            SLF(E_DX, s) = CXF(C_DX, c);
            SLF(E_DY, s) = CXF(C_DY, c);
            SLF(E_DZ, s) = CXF(C_DZ, c);
            SLF(E_TX, s) = 1.f;
            SLF(E_TY, s) = 1.f;
            SLF(E_TZ, s) = 1.f;
            SLF(E_LX, s) = 0.f;
            SLF(E_LY, s) = 0.f;
            SLF(E_LZ, s) = 0.f;
            if (has_rec) { // the primary hit is known (the primary kernel sweeps that ray every sample)
                SLF(E_BEST, s) = CXF(C_BEST0, c);
                SLU(E_ID, s) = CXU(C_ID0, c);
                SLU(E_META, s) = meta_make(SL_SHADE, (unsigned)c, 0u, 0u);
            } else {
                const bool first = w.write_ids && SLU(E_OFF, s) == 0u && CXU(C_PLANE, c) == 0u;
                SLU(E_META, s) = meta_make(SL_SWEEP, (unsigned)c, 0u, 0u, first ? SL_FIRST : 0u);
            }
            st_eval++;
        }
        __syncwarp();
    };

    // ---- shade: calcColor's loop body for up to 32 slots whose ray hit ------------------------------------
    auto shade_stage = [&](const unsigned (&sb)[POOL_SPL]) {
        int n = 0;
#pragma unroll
        for (int k = 0; k < POOL_SPL; ++k) {
            if ((sb[k] >> lane) & 1u) {
                const int pos = n + __popc(sb[k] & lt);
                if (pos < 32)
                    stagelist[pos] = (unsigned char)(lane + 32 * k);
            }
            n += __popc(sb[k]);
        }
        n = min(n, 32);
        __syncwarp();
        st_shade_pass++;
        st_shade_lanes += lane < n;
        if (lane < n) {
            const int s = stagelist[lane];
            const unsigned m = SLU(E_META, s);
            Lane L;
            L.pos = mk3(SLF(E_PX, s), SLF(E_PY, s), SLF(E_PZ, s));
            L.dir = mk3(SLF(E_DX, s), SLF(E_DY, s), SLF(E_DZ, s));
            L.tint = mk3(SLF(E_TX, s), SLF(E_TY, s), SLF(E_TZ, s));
            L.light = mk3(SLF(E_LX, s), SLF(E_LY, s), SLF(E_LZ, s));
            L.rng = SLU(E_RNG, s);
            L.bounce = (int)((m >> 11) & 255u);
            const unsigned hits = ((m >> 19) & 255u) + 1u;
            const bool done = shade_hit(L, a, SLF(E_BEST, s), (int)SLU(E_ID, s));
            SLU(E_RNG, s) = L.rng;
            if (done) { // Russian roulette or the bounce limit: the sample's light, scaled as main.c:99 does
                const f3 c = xscale3(L.light, wgt);
                SLF(E_LX, s) = c.x;
                SLF(E_LY, s) = c.y;
                SLF(E_LZ, s) = c.z;
                SLU(E_BEST, s) = hits - (has_rec ? 1u : 0u); // sweeps of this sample by this kernel
                SLU(E_META, s) = meta_make(SL_DONE, (m >> 3) & 255u, 0u, hits);
            } else {
                SLF(E_PX, s) = L.pos.x;
                SLF(E_PY, s) = L.pos.y;
                SLF(E_PZ, s) = L.pos.z;
                SLF(E_DX, s) = L.dir.x;
                SLF(E_DY, s) = L.dir.y;
                SLF(E_DZ, s) = L.dir.z;
                SLF(E_TX, s) = L.tint.x;
                SLF(E_TY, s) = L.tint.y;
                SLF(E_TZ, s) = L.tint.z;
                SLF(E_LX, s) = L.light.x;
                SLF(E_LY, s) = L.light.y;
                SLF(E_LZ, s) = L.light.z;
                SLU(E_META, s) = meta_make(SL_SWEEP, (m >> 3) & 255u, (unsigned)L.bounce, hits);
            }
        }
        __syncwarp();
    };

    // ---- sweep: closest hit of up to 64 rays, NR per lane; misses take the sky at once -------------------------
    auto sweep_stage = [&](const unsigned (&wb)[POOL_SPL], auto nr_tag) {
        constexpr int NR = decltype(nr_tag)::value;
        int n = 0;
#pragma unroll
        for (int k = 0; k < POOL_SPL; ++k) {
            if ((wb[k] >> lane) & 1u) {
                const int pos = n + __popc(wb[k] & lt);
                if (pos < 32 * NR)
                    stagelist[pos] = (unsigned char)(lane + 32 * k);
            }
            n += __popc(wb[k]);
        }
        n = min(n, 32 * NR);
        __syncwarp();
        int s_[NR];
        bool act[NR];
        f3 o[NR], d[NR];
        float best[NR];
        int bestId[NR];
#pragma unroll
        for (int r = 0; r < NR; ++r) {
            act[r] = lane + 32 * r < n;
            s_[r] = act[r] ? (int)stagelist[lane + 32 * r] : 0;
            o[r] = mk3(SLF(E_PX, s_[r]), SLF(E_PY, s_[r]), SLF(E_PZ, s_[r]));
            d[r] = mk3(SLF(E_DX, s_[r]), SLF(E_DY, s_[r]), SLF(E_DZ, s_[r]));
            best[r] = 999999.0f;
            bestId[r] = -1;
            if (act[r]) // spheres come first in the reference's closest-hit loop (raytracing.c:219-228)
                for (int k = 0; k < nsph; ++k)
                    sphere_test_exact(o[r], d[r], __ldg(&a.sc.s0[k]), k, best[r], bestId[r]);
        }
        st_sweep_pass++;
#pragma unroll
        for (int r = 0; r < NR; ++r)
            st_sweep_rays += act[r];
        if (!STREAM) {
            sweep_multi<MODE, NR, NR == 2 ? 4 : 8>(sGeo, sGeo + qoff, w.obound, ntris, 0, act, o, d, best, bestId, viol);
        } else {
            if (threadIdx.x == 0) {
                issue(k0, 0);
                if (ntiles > 1)
                    issue(k0 + 1, 1);
            }
            for (int t = 0; t < ntiles; ++t) {
                const uint32_t k = k0 + (uint32_t)t;
                mbar_wait(&bars[k & 1], (k >> 1) & 1u);
                const float4 *stage = sGeo + (size_t)(k & 1) * STAGE_F4;
                if (n > 0)
                    sweep_multi<MODE, NR, 4>(stage, stage + qoff, w.obound, min(tile, ntris - t * tile), t * tile, act, o, d,
                                             best, bestId, viol);
                __syncthreads(); // every warp is done with this stage
                if (threadIdx.x == 0 && t + 2 < ntiles)
                    issue(k + 2, t + 2);
            }
            k0 += (uint32_t)ntiles;
        }
        // results go to the slots first, so that the (divergent, FP64) epilogue below is one copy of the code looping
        // over the lane's rays without indexing registers dynamically
#pragma unroll
        for (int r = 0; r < NR; ++r)
            if (act[r]) {
                SLF(E_BEST, s_[r]) = best[r];
                SLU(E_ID, s_[r]) = (unsigned)bestId[r];
            }
#pragma unroll 1
        for (int r = 0; r < NR; ++r) {
            if (lane + 32 * r >= n)
                continue;
            const int s = (int)stagelist[lane + 32 * r];
            const unsigned m = SLU(E_META, s);
            const int c = (int)((m >> 3) & 255u);
            const int id = (int)SLU(E_ID, s);
            if (m & SL_FIRST) { // this kernel owns the id map: the pixel's first primary sweep
                if (a.ids != nullptr)
                    a.ids[CXU(C_PIX, c)] = id;
                nhitpix += id != -1;
            }
            if (id == -1) { // miss: sky, the path ends (raytracing.c:289-293)
                Lane L;
                L.dir = mk3(SLF(E_DX, s), SLF(E_DY, s), SLF(E_DZ, s));
                L.tint = mk3(SLF(E_TX, s), SLF(E_TY, s), SLF(E_TZ, s));
                L.light = mk3(SLF(E_LX, s), SLF(E_LY, s), SLF(E_LZ, s));
                shade_miss(L, a);
                const f3 cc = xscale3(L.light, wgt);
                SLF(E_LX, s) = cc.x;
                SLF(E_LY, s) = cc.y;
                SLF(E_LZ, s) = cc.z;
                const unsigned hits = (m >> 19) & 255u;
                SLU(E_BEST, s) = hits + (has_rec ? 0u : 1u); // sweeps of this sample by this kernel
                SLU(E_META, s) = meta_make(SL_DONE, (unsigned)c, 0u, hits);
            } else {
                SLU(E_META, s) = (m & ~(3u | SL_FIRST)) | SL_SHADE;
            }
        }
        __syncwarp();
    };

    // ---- the scheduler ------------------------------------------------------------------------------------
    while (true) {
        bool sweep_due = false;
        unsigned sb[POOL_SPL], wb[POOL_SPL];
        int nsweep = 0;
        while (true) {
            if (++iters > w.watchdog) {
                tripped = true;
                break;
            }
            consume();
            refill();
            int nshade = 0;
            nsweep = 0;
#pragma unroll
            for (int k = 0; k < POOL_SPL; ++k) {
                const unsigned st = SLU(E_META, lane + 32 * k) & 3u;
                sb[k] = __ballot_sync(0xffffffffu, st == SL_SHADE);
                wb[k] = __ballot_sync(0xffffffffu, st == SL_SWEEP);
                nshade += __popc(sb[k]);
                nsweep += __popc(wb[k]);
            }
            // the pass that fills its lanes better; a streamed sweep costs a pass over the whole scene: fill it
            const int esw = STREAM ? (nsweep >= 64 || nshade == 0 ? 64 : 0) : min(64, nsweep > 32 ? nsweep : 2 * nsweep);
            const int esh = 2 * min(nshade, 32);
            if (nsweep > 0 && esw >= esh) {
                sweep_due = true;
                break;
            }
            if (nshade > 0) {
                shade_stage(sb);
                continue;
            }
            // nothing to run: finished when the list is exhausted and no pixel is live (else refill has work to hand out)
            bool live = false;
#pragma unroll
            for (int q = 0; q < POOL_CPL; ++q)
                live = live || CXU(C_REM, lane + 32 * q) != 0u;
            if (exhausted && !__any_sync(0xffffffffu, live))
                break;
        }
        if (STREAM) {
            if (!__syncthreads_or((sweep_due || tripped) ? 1 : 0))
                break;
            if (__syncthreads_or(tripped ? 1 : 0))
                break;
        } else if (!sweep_due) {
            break;
        }
        if (STREAM || nsweep > 32)
            sweep_stage(wb, std::integral_constant<int, 2>());
        else
            sweep_stage(wb, std::integral_constant<int, 1>());
    }
    if (tripped && lane == 0)
        atomicExch(w.error, 1u);

    if (a.stats != nullptr) {
        unsigned long long v[8] = {nrays, nhitpix, st_eval, st_shade_pass, st_shade_lanes, st_sweep_pass, st_sweep_rays, iters};
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
                v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
        if (mode_verify(MODE) && viol)
            atomicAdd(&a.stats[ST_MISMATCH], (unsigned long long)viol);
        if (tripped && lane == 0)
            atomicAdd(&a.stats[ST_MISMATCH], 1ull << 40);
        if (lane == 0) {
            if (v[0]) {
                atomicAdd(&a.stats[ST_RAYS], v[0]);
                atomicAdd(&a.stats[ST_TRI_PAIRS], v[0] * (unsigned long long)ntris);
                atomicAdd(&a.stats[ST_SPH_PAIRS], v[0] * (unsigned long long)nsph);
                atomicAdd(&a.stats[ST_SLOW_PAIRS], v[0] * (unsigned long long)ntris);
            }
            if (w.write_ids && a.ids != nullptr && v[1])
                atomicAdd(&a.stats[ST_HIT_PIXELS], v[1]);
            // scheduler counters (per-lane values were summed over the warp: passes and iterations count 32 per warp)
            atomicAdd(&a.stats[ST_POOL_EVAL], v[2]);
            atomicAdd(&a.stats[ST_POOL_SHADE_PASSES], v[3] / 32);
            atomicAdd(&a.stats[ST_POOL_SHADE_LANES], v[4]);
            atomicAdd(&a.stats[ST_POOL_SWEEP_PASSES], v[5] / 32);
            atomicAdd(&a.stats[ST_POOL_SWEEP_RAYS], v[6]);
            atomicAdd(&a.stats[ST_POOL_ITERS], v[7] / 32);
        }
    }
#undef SLF
#undef SLU
#undef CXF
#undef CXU
}

} // namespace rtc
