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
//     consume   lane c walks its pixel context: finished slots are consumed in offset order (acc += light, j += h), the
//               slots of skipped offsets are freed, a pixel whose last sample is consumed is written
//     refill    idle contexts take new pixels from the global work list (at most the warp's share at a time); every
//               context with nothing outstanding gets a slot (that alone is the serial stream of the reference); slots
//               left over become speculative children at the next unassigned offsets of the contexts that can use them
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
#define RTC_POOL_NS 64
#endif
// Warps per CTA and CTAs per SM of the resident form (A/B knobs); one CTA of POOL_WARPS_LARGE warps per SM when the
// scene is large or streamed (the tiles' double buffer is then shared by more warps).
#ifndef RTC_POOL_WARPS_SMALL
#define RTC_POOL_WARPS_SMALL 12
#endif
#ifndef RTC_POOL_MINB_SMALL
#define RTC_POOL_MINB_SMALL 2
#endif
#ifndef RTC_POOL_DONE_BATCH // finished samples a warp lets accumulate before it runs consume + refill while its passes are full
#define RTC_POOL_DONE_BATCH 16
#endif
#ifndef RTC_POOL_UP // pair records per iteration of the two-ray quick test
#define RTC_POOL_UP 1
#endif
constexpr int POOL_WARPS_SMALL = RTC_POOL_WARPS_SMALL, POOL_WARPS_LARGE = 16, POOL_MINB_SMALL = RTC_POOL_MINB_SMALL;
constexpr int POOL_NS = RTC_POOL_NS; // evaluation slots per warp (multiple of 32)
constexpr int POOL_NC = 32;          // pixel contexts per warp: context c belongs to lane c
constexpr int POOL_DONE_BATCH = RTC_POOL_DONE_BATCH;
constexpr int POOL_SPL = (POOL_NS + 31) / 32; // slots a lane looks after (lane, lane + 32, ...; the last row may be cut)
#ifndef RTC_POOL_RING
#define RTC_POOL_RING 32
#endif
constexpr int POOL_RING = RTC_POOL_RING; // outstanding offsets per context (a power of two)
static_assert(POOL_NS % 4 == 0 && POOL_NS >= 64 && POOL_NS <= 224 && (POOL_RING & (POOL_RING - 1)) == 0, "pool geometry");
constexpr int POOL_MAX_BOUNCE = 255; // bounce and hit counters are 8-bit fields of the slot's meta word

// slot fields (one array of POOL_NS words each)
enum { E_PX, E_PY, E_PZ, E_DX, E_DY, E_DZ, E_TX, E_TY, E_TZ, E_LX, E_LY, E_LZ, E_RNG, E_BEST, E_ID, E_META, E_FIELDS };
// context fields (one array of POOL_NC words each)
enum { C_PIX, C_PLANE, C_REM, C_TOTAL, C_J, C_TOP, C_TSTATE, C_AX, C_AY, C_AZ, C_DX, C_DY, C_DZ, C_BEST0, C_ID0, C_FIELDS };
constexpr int POOL_WORDS = E_FIELDS * POOL_NS + C_FIELDS * POOL_NC + (POOL_NC * POOL_RING + 2 * POOL_NS) / 4;

// meta word of a slot: [1:0] state, [18:11] bounce, [26:19] hits
enum { SL_FREE = 0, SL_SHADE = 1, SL_SWEEP = 2, SL_DONE = 3 };
__device__ __forceinline__ unsigned meta_make(unsigned st, unsigned bounce, unsigned hits) { return st | (bounce << 11) | (hits << 19); }

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
    const int *list;            // compact pixel indices of the primary-hit pixels
    const unsigned *list_count; // device counter written by the primary kernel
    const float2 *rec;          // {distance, id bits} of each listed pixel's primary hit
    unsigned *cursor;           // device work cursor, zeroed before the launch
    unsigned *error;            // set to 1 when a warp's scheduler trips its watchdog (a bug: never in a correct build)
    int planes;                 // sub-streams rendered by this launch
    int tile;                   // triangles per streamed tile (<= PATH_TILE), STREAM only
    int spec;                   // 1: speculative children allowed
    int ctx_share_div;          // a warp holds at most ceil(tasks / (warps * div)) pixels at a time (>= 1, <= POOL_NC)
    const float4 *qrec;         // MODE >= 1: quick records
    float obound;
    unsigned long long watchdog; // scheduler iterations per warp after which the kernel gives up
    unsigned max_tasks;          // the kernel returns at once when the list holds this many (pixel, sub-stream) tasks or
                                 // more: long lists are the serial form's (rtc_paths.cuh), the host launches both
};

__host__ __device__ inline size_t pool_smem_bytes(int ntris, bool stream, int mode, int warps)
{
    return path_smem_bytes(ntris, stream, mode) + (size_t)warps * POOL_WORDS * sizeof(uint32_t);
}

// ---- the sweep of TWO rays per lane over one resident scene / one streamed tile -------------------------------
// Same two phases as sweep_chunked (rtc_paths.cuh): branch-free conservative quick test of every pair, then the
// reference's exact sequence on the survivors in index order.  Every warp-uniform LDS.128 of a pair record feeds both
// rays (a uniform LDS.128 costs two shared-memory wavefronts whatever it carries: the pipe that bounds the one-ray
// form).  The loops are kept small on purpose: the SM's L0 instruction cache holds ~6 KB and the
// warps of an SM sub-partition are in different phases of this kernel.
template <int MODE>
__device__ __forceinline__ void sweep_two(const float4 *S, const float4 *geo, float obound, int n, int base, const bool (&act)[2],
                                          const f3 (&o)[2], const f3 (&d)[2], float (&best)[2], int (&bestId)[2], unsigned &viol)
{
    bool rayok[2];
    QuickRay2 q2[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        rayok[r] = MODE == 0 || quick_ray_ok(o[r], d[r], obound);
        q2[r] = quick_ray2(quick_ray(o[r], d[r]));
    }
#pragma unroll 1
    for (int c0 = 0; c0 < n; c0 += 32) {
        const int cn = min(32, n - c0);
        const unsigned full = cn == 32 ? 0xffffffffu : (1u << cn) - 1u;
        unsigned m[2] = {0u, 0u};
        if (MODE == 0) {
#pragma unroll 1
            for (int j = 0; j < cn; ++j) {
                const int t0 = 3 * (c0 + j);
#pragma unroll
                for (int r = 0; r < 2; ++r) {
                    unsigned m1 = 0;
                    tri_quick_bit<1u>(o[r], d[r], S[t0], S[t0 + 1], S[t0 + 2], m1);
                    m[r] |= m1 << j;
                }
            }
        } else {
            static_assert(MODE == 0 || mode_packed(MODE), "the pool kernel evaluates quick records in packed form");
            constexpr int UP = RTC_POOL_UP; // pair records per iteration
            const float4 *P = S + QPAIR_F4 * (c0 >> 1);
            const int np = (cn + 1) >> 1;
            int j = 0;
#pragma unroll 1
            for (; j + UP <= np; j += UP) {
#pragma unroll
                for (int r = 0; r < 2; ++r) {
                    unsigned mg = 0;
                    quick2_pairs<0, UP>(q2[r], P + QPAIR_F4 * j, mg);
                    m[r] |= mg << (2 * j);
                }
            }
            if (UP > 1) {
#pragma unroll 1
                for (; j < np; ++j) {
#pragma unroll
                    for (int r = 0; r < 2; ++r) {
                        unsigned m1 = 0;
                        quick2_pairs<0, 1>(q2[r], P + QPAIR_F4 * j, m1);
                        m[r] |= m1 << (2 * j);
                    }
                }
            }
        }
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            m[r] &= full; // an odd count's last pair is completed by a record that never survives
            if (!rayok[r])
                m[r] = full;
            if (!act[r])
                m[r] = 0;
        }
        const float4 *G = MODE == 0 ? S : geo;
        // phase 2, ONE copy of the exact sequence for both rays (selects, no dynamically indexed registers)
#pragma unroll 1
        for (int r = 0; r < 2; ++r) {
            unsigned mm = r ? m[1] : m[0];
            if (mode_verify(MODE))
                mm = act[r ? 1 : 0] ? full : 0u;
            if (mm == 0u)
                continue;
            const unsigned quick = r ? m[1] : m[0];
            const f3 oo = r ? o[1] : o[0], dd = r ? d[1] : d[0];
            float b = r ? best[1] : best[0];
            int bi = r ? bestId[1] : bestId[0];
            while (mm) {
                const int jj = __ffs(mm) - 1;
                const int i = c0 + jj;
                mm &= mm - 1;
                if (mode_verify(MODE)) { // every pair runs the exact test: an accepted pair the quick test dropped is counted
                    unsigned ex[6] = {0, 0, 0, 0, 0, 0};
                    tri_test_exact<true>(oo, dd, G[3 * i], G[3 * i + 1], G[3 * i + 2], base + i, b, bi, ex);
                    if (ex[5] && !((quick >> jj) & 1u))
                        viol++;
                } else {
                    unsigned dummy[1];
                    tri_test_exact<false>(oo, dd, G[3 * i], G[3 * i + 1], G[3 * i + 2], base + i, b, bi, dummy);
                }
            }
            if (r) {
                best[1] = b;
                bestId[1] = bi;
            } else {
                best[0] = b;
                bestId[0] = bi;
            }
        }
    }
}

// ---- the kernel ---------------------------------------------------------------------------------------------
template <bool STREAM, int MODE, int WARPS>
__global__ void __launch_bounds__(32 * WARPS, WARPS == POOL_WARPS_SMALL ? POOL_MINB_SMALL : 1) render_pool_kernel(const RenderArgs a, const PoolWork w)
{
    constexpr int POOL_THREADS = 32 * WARPS;
    if ((unsigned long long)*w.list_count * (unsigned)w.planes >= (unsigned long long)w.max_tasks)
        return;
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
    unsigned char *ring = reinterpret_cast<unsigned char *>(pool + E_FIELDS * POOL_NS + C_FIELDS * POOL_NC); // [ctx][offset & 15] -> slot
    unsigned char *freelist = ring + POOL_NC * POOL_RING;
    unsigned char *stagelist = freelist + POOL_NS;
#define SLF(f, s) Ef[(f) * POOL_NS + (s)]
#define SLU(f, s) Eu[(f) * POOL_NS + (s)]
#define CXF(f) Cf[(f) * POOL_NC + lane]
#define CXU(f) Cu[(f) * POOL_NC + lane]
#define RING(off) ring[lane * POOL_RING + ((off) & (POOL_RING - 1))]

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
        if (lane + 32 * k < POOL_NS)
            SLU(E_META, lane + 32 * k) = SL_FREE;
    CXU(C_REM) = 0u;
    __syncthreads();

    const unsigned npix = *w.list_count;
    const unsigned total = npix * (unsigned)w.planes;
    if (a.prog != nullptr && blockIdx.x == 0 && (int)threadIdx.x < a.prog_chunks && a.prog[threadIdx.x] == 0)
        *(volatile unsigned *)&a.prog[2 * a.prog_chunks + threadIdx.x] = 1u; // chunks without hit pixels are final already
    const bool spec = w.spec != 0;
    const unsigned nwarps = gridDim.x * WARPS;
    // pixels this warp holds at a time: its share of the list, so that a short list is spread over all warps (the
    // free slots then run the pixels' samples in parallel) and a long one is taken a few pixels at a time
    const int maxctx = (int)max(1ull, min((unsigned long long)POOL_NC,
                                          ((unsigned long long)total + (unsigned long long)nwarps * w.ctx_share_div - 1) /
                                              ((unsigned long long)nwarps * w.ctx_share_div)));
    const float fh = (float)(a.H / 2);
    const float wgt = (float)(1.0 / (double)a.spp); // main.c:99
    const float inv2w = 0.5f / (float)nwarps;

    bool exhausted = total == 0;
    unsigned seen = 0; // the list position after this warp's last fetch
    unsigned long long nrays = 0, iters = 0;
    unsigned st_eval = 0, st_shade_pass = 0, st_shade_lanes = 0, st_sweep_pass = 0, st_sweep_rays = 0;
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

    // ---- consume: lane c walks its context; finished slots in offset order, skipped offsets dropped --------------
    auto consume = [&]() {
        unsigned rem = CXU(C_REM);
        if (rem == 0u)
            return;
        unsigned j = CXU(C_J), top = CXU(C_TOP);
        if (j == top || (SLU(E_META, RING(j)) & 3u) != SL_DONE)
            return;
        f3 acc = mk3(CXF(C_AX), CXF(C_AY), CXF(C_AZ));
        unsigned ts = CXU(C_TSTATE);
        while (j < top) {
            const int s = RING(j);
            const unsigned m = SLU(E_META, s);
            if ((m & 3u) != SL_DONE)
                break;
            acc = xadd3(acc, mk3(SLF(E_LX, s), SLF(E_LY, s), SLF(E_LZ, s))); // main.c:99, in sample order
            nrays += SLU(E_BEST, s);                                         // sweeps of this sample
            SLU(E_META, s) = SL_FREE;
            const unsigned nj = j + (spec ? ((m >> 19) & 255u) : 1u);
            for (unsigned o = j + 1u; o < nj && o < top; ++o) // offsets this sample skipped
                SLU(E_META, RING(o)) = SL_FREE;
            if (nj >= top) {
                top = nj;
                ts = SLU(E_RNG, s); // the state after this sample = the state at offset nj
            }
            j = nj;
            if (--rem == 0u)
                break;
        }
        CXF(C_AX) = acc.x;
        CXF(C_AY) = acc.y;
        CXF(C_AZ) = acc.z;
        CXU(C_J) = j;
        CXU(C_TOP) = top;
        CXU(C_TSTATE) = ts;
        CXU(C_REM) = rem;
        if (rem == 0u) { // the pixel's last sample is consumed: drop what was speculated beyond it, write the pixel
            for (unsigned o = j; o < top; ++o)
                SLU(E_META, RING(o)) = SL_FREE;
            const int pix = (int)CXU(C_PIX), plane = (int)CXU(C_PLANE);
            if (a.accum != nullptr) {
                float *p = a.accum + ((size_t)plane * a.out_w * a.out_h + pix) * 3;
                p[0] = acc.x;
                p[1] = acc.y;
                p[2] = acc.z;
            }
            if (a.rgb != nullptr && a.nsplit == 1) { // vec3ToColor, raytracing.c:11-15
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
        }
    };

    // ---- refill: new pixels for idle contexts, slots for the offsets the contexts want ---------------------------
    auto refill = [&]() {
        int nfree = 0;
#pragma unroll
        for (int k = 0; k < POOL_SPL; ++k) {
            const unsigned fb = __ballot_sync(0xffffffffu, lane + 32 * k < POOL_NS && (SLU(E_META, lane + 32 * k) & 3u) == SL_FREE);
            if ((fb >> lane) & 1u)
                freelist[nfree + __popc(fb & lt)] = (unsigned char)(lane + 32 * k);
            nfree += __popc(fb);
        }
        unsigned rem = CXU(C_REM);
        const unsigned idle = __ballot_sync(0xffffffffu, rem == 0u);
        if (!exhausted && idle != 0u) {
            const int nlive = 32 - __popc(idle);
            // guided: towards the end of the list the warps take fewer pixels at a time
            const unsigned left = total - min(seen, total);
            const unsigned guide = max(1u, (unsigned)((float)left * inv2w)); // ~ left / (2 warps), no integer division
            const int nnew = min(max(0, min(__popc(idle), maxctx - nlive)), (int)min(guide, 32u));
            if (nnew > 0) {
                unsigned base = 0;
                if (lane == 0)
                    base = atomicAdd(w.cursor, (unsigned)nnew);
                base = __shfl_sync(0xffffffffu, base, 0);
                seen = base + (unsigned)nnew;
                if (base >= total)
                    exhausted = true;
                const unsigned t = base + (unsigned)__popc(idle & lt);
                if (rem == 0u && __popc(idle & lt) < nnew && t < total) {
                    const int plane = w.planes == 1 ? 0 : (int)(t / npix);
                    const unsigned li = w.planes == 1 ? t : t % npix;
                    const int pix = w.list[li];
                    const float2 hit = w.rec[li];
                    const int ci = pix % a.out_w, cj = pix / a.out_w;
                    const int x = a.col_start + ci * a.col_step;
                    const int y = a.row_start + (cj / a.band_h) * a.row_step + (cj % a.band_h);
                    const int g = a.split_first + plane;
                    const float dx = xdiv((float)(x - a.W / 2), fh), dy = xdiv((float)(y - a.H / 2), fh); // main.c:88-94
                    const f3 pdir = xnormalize(xadd3(xadd3(xscale3(a.ex, dx), xscale3(a.ey, dy)), xscale3(a.ez, a.fov)));
                    int s_begin = 0, s_end = a.spp;
                    if (a.nsplit != 1) { // (64-bit divisions: only for sub-streams)
                        s_begin = (int)((long long)g * a.spp / a.nsplit);
                        s_end = (int)((long long)(g + 1) * a.spp / a.nsplit);
                    }
                    rem = (unsigned)(s_end - s_begin);
                    CXU(C_PIX) = (unsigned)pix;
                    CXU(C_PLANE) = (unsigned)plane;
                    CXU(C_REM) = rem;
                    CXU(C_TOTAL) = rem;
                    CXU(C_J) = 0u;
                    CXU(C_TOP) = 0u;
                    CXU(C_TSTATE) = (uint32_t)(x + y * a.W) + (uint32_t)g * (uint32_t)(a.W * a.H); // main.c:95
                    CXF(C_AX) = 0.f;
                    CXF(C_AY) = 0.f;
                    CXF(C_AZ) = 0.f;
                    CXF(C_DX) = pdir.x;
                    CXF(C_DY) = pdir.y;
                    CXF(C_DZ) = pdir.z;
                    CXF(C_BEST0) = hit.x;
                    CXU(C_ID0) = __float_as_uint(hit.y);
                    if (rem == 0u && a.accum != nullptr) { // empty sample range: the partial sum is zero, nothing to trace
                        float *p = a.accum + ((size_t)plane * a.out_w * a.out_h + pix) * 3;
                        p[0] = p[1] = p[2] = 0.f;
                    }
                }
            }
        }
        if (nfree == 0)
            return;
        // demand of this lane's context: one slot when nothing is outstanding (the reference's serial stream), and
        // speculative children at the next unassigned offsets as far as they can still be needed.  A pixel that keeps
        // skipping offsets (a closed scene: most of its speculation is dropped) is served last, with the slots nobody
        // else wants -- idle slots cost nothing, and the slowest pixels are what a short work list waits for.
        unsigned j = 0, top = 0;
        int first = 0, extra = 0;
        bool good = true;
        if (rem != 0u) {
            j = CXU(C_J);
            top = CXU(C_TOP);
            first = top == j ? 1 : 0;
            if (spec) {
                const unsigned consumed = CXU(C_TOTAL) - rem, skipped = j - consumed;
                good = skipped * 2u <= consumed + 4u;
                extra = max(0, min((int)(j + rem - top), POOL_RING - (int)(top - j)) - first); // offsets below j + rem can be needed
            }
        }
        const unsigned fm = __ballot_sync(0xffffffffu, first != 0);
        const int n1 = __popc(fm);
        const int nwant_g = __popc(__ballot_sync(0xffffffffu, extra > 0 && good));
        const int nwant_b = __popc(__ballot_sync(0xffffffffu, extra > 0 && !good));
        const int left2 = max(0, nfree - n1);
        // (quotas by a float quotient rounded up: any value is safe, the positions below are clipped to the free list)
        if (good)
            extra = nwant_g > 0 ? min(extra, __float2int_ru(__fdividef((float)left2, (float)nwant_g))) : 0;
        const int tot_g = min(left2, (int)__reduce_add_sync(0xffffffffu, (unsigned)(good ? extra : 0)));
        const int left3 = left2 - tot_g;
        if (!good)
            extra = nwant_b > 0 ? min(extra, __float2int_ru(__fdividef((float)left3, (float)nwant_b))) : 0;
        // positions in the free list: the firsts, then the extras of the well-behaved pixels, then the others'
        unsigned sc = good ? (unsigned)extra : (unsigned)extra << 16; // both inclusive scans at once
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned v = __shfl_up_sync(0xffffffffu, sc, o);
            if (lane >= o)
                sc += v;
        }
        const int p2 = good ? n1 + (int)(sc & 0xffffu) - extra : n1 + tot_g + (int)(sc >> 16) - extra;
        const int p1 = __popc(fm & lt);
        const int ga = (first && p1 < nfree) ? 1 : 0;
        const int gb = max(0, min(extra, nfree - p2));
        __syncwarp(); // the free list is complete
        if (ga + gb > 0) {
            unsigned ts = CXU(C_TSTATE);
            const f3 pd = mk3(CXF(C_DX), CXF(C_DY), CXF(C_DZ));
            const float best0 = CXF(C_BEST0);
            const unsigned id0 = CXU(C_ID0);
            for (int i = 0; i < ga + gb; ++i) {
                const int s = freelist[i < ga ? p1 : p2 + (i - ga)];
                SLF(E_PX, s) = a.origin.x;
                SLF(E_PY, s) = a.origin.y;
                SLF(E_PZ, s) = a.origin.z;
                SLF(E_DX, s) = pd.x;
                SLF(E_DY, s) = pd.y;
                SLF(E_DZ, s) = pd.z;
                SLF(E_TX, s) = 1.f;
                SLF(E_TY, s) = 1.f;
                SLF(E_TZ, s) = 1.f;
                SLF(E_LX, s) = 0.f;
                SLF(E_LY, s) = 0.f;
                SLF(E_LZ, s) = 0.f;
                SLU(E_RNG, s) = ts;
                SLF(E_BEST, s) = best0; // the primary hit is known (the primary kernel sweeps that ray every sample)
                SLU(E_ID, s) = id0;
                SLU(E_META, s) = meta_make(SL_SHADE, 0u, 0u);
                RING(top) = (unsigned char)s;
                top += 1u;
                if (spec)
                    ts = ts * LCG7_M + LCG7_C; // the state 7 draws later = the next offset's
            }
            CXU(C_TOP) = top;
            CXU(C_TSTATE) = ts;
            st_eval += (unsigned)(ga + gb);
        }
    };

    // ---- shade: calcColor's loop body for up to 64 slots whose ray hit, two per lane one after the other ------------
    auto shade_stage = [&](const unsigned (&sb)[POOL_SPL]) {
        int n = 0;
#pragma unroll
        for (int k = 0; k < POOL_SPL; ++k) {
            if ((sb[k] >> lane) & 1u) {
                const int pos = n + __popc(sb[k] & lt);
                if (pos < 64)
                    stagelist[pos] = (unsigned char)(lane + 32 * k);
            }
            n += __popc(sb[k]);
        }
        n = n >= 64 ? 64 : min(n, 32); // full passes only
        __syncwarp();
#pragma unroll 1
        for (int r = 0; r < 2; ++r) {
            if (32 * r >= n)
                break;
            st_shade_pass++;
            st_shade_lanes += lane + 32 * r < n;
            if (lane + 32 * r >= n)
                continue;
            const int s = stagelist[lane + 32 * r];
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
                L.light = xscale3(L.light, wgt);
                SLU(E_BEST, s) = hits - 1u; // sweeps of this sample by this kernel
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
            }
            SLF(E_LX, s) = L.light.x;
            SLF(E_LY, s) = L.light.y;
            SLF(E_LZ, s) = L.light.z;
            SLU(E_META, s) = meta_make(done ? SL_DONE : SL_SWEEP, (unsigned)L.bounce, hits);
        }
    };

    // ---- sweep: closest hit of up to 64 rays, two per lane; misses take the sky at once -------------------------
    auto sweep_stage = [&](const unsigned (&wb)[POOL_SPL]) {
        int n = 0;
#pragma unroll
        for (int k = 0; k < POOL_SPL; ++k) {
            if ((wb[k] >> lane) & 1u) {
                const int pos = n + __popc(wb[k] & lt);
                if (pos < 64)
                    stagelist[pos] = (unsigned char)(lane + 32 * k);
            }
            n += __popc(wb[k]);
        }
        n = min(n, 64);
        __syncwarp();
        int s_[2];
        bool act[2];
        f3 o[2], d[2];
        float best[2];
        int bestId[2];
#pragma unroll
        for (int r = 0; r < 2; ++r) {
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
        st_sweep_rays += (unsigned)act[0] + (unsigned)act[1];
        if (!STREAM) {
            sweep_two<MODE>(sGeo, sGeo + qoff, w.obound, ntris, 0, act, o, d, best, bestId, viol);
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
                    sweep_two<MODE>(stage, stage + qoff, w.obound, min(tile, ntris - t * tile), t * tile, act, o, d, best, bestId, viol);
                __syncthreads(); // every warp is done with this stage
                if (threadIdx.x == 0 && t + 2 < ntiles)
                    issue(k + 2, t + 2);
            }
            k0 += (uint32_t)ntiles;
        }
        // results go to the slots first, so that the (divergent, FP64) epilogue below is one copy of the code looping
        // over the lane's rays without indexing registers dynamically
#pragma unroll
        for (int r = 0; r < 2; ++r)
            if (act[r]) {
                SLF(E_BEST, s_[r]) = best[r];
                SLU(E_ID, s_[r]) = (unsigned)bestId[r];
            }
#pragma unroll 1
        for (int r = 0; r < 2; ++r) {
            if (lane + 32 * r >= n)
                continue;
            const int s = (int)stagelist[lane + 32 * r];
            const unsigned m = SLU(E_META, s);
            if ((int)SLU(E_ID, s) == -1) { // miss: sky, the path ends (raytracing.c:289-293)
                Lane L;
                L.dir = mk3(SLF(E_DX, s), SLF(E_DY, s), SLF(E_DZ, s));
                L.tint = mk3(SLF(E_TX, s), SLF(E_TY, s), SLF(E_TZ, s));
                L.light = mk3(SLF(E_LX, s), SLF(E_LY, s), SLF(E_LZ, s));
                shade_miss(L, a);
                const f3 cc = xscale3(L.light, wgt);
                SLF(E_LX, s) = cc.x;
                SLF(E_LY, s) = cc.y;
                SLF(E_LZ, s) = cc.z;
                SLU(E_BEST, s) = (m >> 19) & 255u; // sweeps of this sample by this kernel
                SLU(E_META, s) = (m & ~3u) | SL_DONE;
            } else {
                SLU(E_META, s) = (m & ~3u) | SL_SHADE;
            }
        }
    };

    // ---- the scheduler ------------------------------------------------------------------------------------
    while (true) {
        bool sweep_due = false;
        unsigned sb[POOL_SPL], wb[POOL_SPL];
        while (true) {
            if (++iters > w.watchdog) {
                tripped = true;
                break;
            }
            // census of the slots; consume + refill (a few hundred instructions) only when enough samples have finished to
            // be worth it, or when the runnable work would not fill a pass -- they hand out work, they do not create it
            int nshade = 0, nsweep = 0, ndone = 0, nfree = 0;
            auto census = [&]() {
                nshade = nsweep = ndone = nfree = 0;
#pragma unroll
                for (int k = 0; k < POOL_SPL; ++k) {
                    const bool in = lane + 32 * k < POOL_NS;
                    const unsigned st = in ? SLU(E_META, lane + 32 * k) & 3u : (unsigned)SL_FREE;
                    sb[k] = __ballot_sync(0xffffffffu, st == SL_SHADE);
                    wb[k] = __ballot_sync(0xffffffffu, st == SL_SWEEP);
                    nshade += __popc(sb[k]);
                    nsweep += __popc(wb[k]);
                    ndone += __popc(__ballot_sync(0xffffffffu, st == SL_DONE));
                    nfree += __popc(__ballot_sync(0xffffffffu, in && st == SL_FREE));
                }
            };
            __syncwarp();
            census();
            if ((ndone >= POOL_DONE_BATCH || (nsweep < 64 && nshade < 32)) && (ndone > 0 || nfree > 0)) {
                consume();
                __syncwarp();
                refill();
                __syncwarp();
                census();
            }
            // the pass that fills its lanes better; a streamed sweep costs a pass over the whole scene: fill it
            const int esw = STREAM ? (nsweep >= 64 || nshade == 0 ? 64 : 0) : min(64, nsweep);
            const int esh = 2 * min(nshade, 32); // (a shade pass takes 64 slots when there are 64, else up to 32)
            if (nsweep > 0 && esw >= esh) {
                sweep_due = true;
                break;
            }
            if (nshade > 0) {
                shade_stage(sb);
                continue;
            }
            // nothing to run: finished when the list is exhausted and no pixel is live (else refill has work to hand out)
            if (exhausted && !__any_sync(0xffffffffu, CXU(C_REM) != 0u))
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
        sweep_stage(wb);
    }
    if (tripped && lane == 0)
        atomicExch(w.error, 1u);

    if (a.stats != nullptr) {
        unsigned long long v[7] = {nrays, st_eval, st_shade_pass, st_shade_lanes, st_sweep_pass, st_sweep_rays, iters};
#pragma unroll
        for (int i = 0; i < 7; ++i)
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
            // scheduler counters (per-lane values were summed over the warp: passes and iterations count 32 per warp)
            atomicAdd(&a.stats[ST_POOL_EVAL], v[1]);
            atomicAdd(&a.stats[ST_POOL_SHADE_PASSES], v[2] / 32);
            atomicAdd(&a.stats[ST_POOL_SHADE_LANES], v[3]);
            atomicAdd(&a.stats[ST_POOL_SWEEP_PASSES], v[4] / 32);
            atomicAdd(&a.stats[ST_POOL_SWEEP_RAYS], v[5]);
            atomicAdd(&a.stats[ST_POOL_ITERS], v[6] / 32);
        }
    }
#undef SLF
#undef SLU
#undef CXF
#undef CXU
#undef RING
}

} // namespace rtc
