// rtc_sky.cuh -- the primary-ray kernel: classification + every sample's primary sweep of the pixels whose
// primary ray escapes ("sky pixels").
//
// Primary rays share the camera origin, so a (primary ray, triangle) pair is decided by the four conservative
// edge functions prepared per frame (rtc_fast.cuh: prepare_primary_kernel): a pair is rejected only when the
// reference's own arithmetic provably rejects it; every other pair runs the reference's exact sequence.
//
//  * sample s0 classifies: pixels whose primary ray hits are appended to a compact work list for the path
//    kernel (rtc_paths.cuh), which owns them entirely (their RNG stream is serial); the primary-id map is
//    written here.
//  * a sky pixel draws no random number (raytracing.c:289-293: miss -> sky, 0 draws), so its spp samples are
//    spp identical sweeps of the same ray.  Each one is executed (the reference re-traces the primary ray per
//    sample, main.c:98-99; a counted ray is a sweep that ran) and cross-checked against the classification.
//
// Lane layout: 8 x 4 lanes per warp, R horizontally adjacent pixels per lane (warp tile 8R x 4), 4 warps per
// CTA stacked in y.  The edge-function table (64 B per triangle) is staged in shared memory: resident when it
// fits one tile, otherwise streamed per sample by the TMA unit (cp.async.bulk + mbarrier, double-buffered).
#pragma once

#include "rtc_fast.cuh"
#include "rtc_tma.cuh"

namespace rtc {

constexpr int SKY_TILE = 512; // triangles per table tile: 32 KB per stage

struct SkyWork {
    int *list;            // out: compact indices of primary-hit pixels (nullptr: do not build a list)
    unsigned *list_count; // out: number of entries (zeroed before the launch)
    float2 *rec;          // out (block kernel): {distance, id bits} of each listed pixel's primary hit
    int4 *btask;          // out (block kernel): blocks to re-sweep per pixel {ci0, cj0, hit mask, plane}
    const float *sphq;    // block kernel: 8 floats per sphere (prepare_spheres_kernel)
    float hw, hh;         // block kernel: half extents of a pixel block in (dx, dy), as used for the tables
    unsigned *btask_count;
    int tile;             // triangles per streamed tile (<= SKY_TILE)
    int stream;           // 1: stream the table per sample; 0: resident (needs ntris <= SKY_TILE)
    int segs;             // block kernel: the samples of a sub-stream are split over this many CTAs (blockIdx.z = plane *
                          // segs + segment) when the pixel grid alone does not fill the machine; >= 1
};

__host__ __device__ inline size_t sky_smem_bytes(int ntris, bool stream)
{
    const int n = stream ? 2 * SKY_TILE : (ntris > 0 ? ntris : 1);
    return (size_t)n * TAB_F4 * sizeof(float4);
}

template <int R>
__global__ void __launch_bounds__(FAST_THREADS) render_sky_kernel(const RenderArgs a, const float4 *__restrict__ tab,
                                                                  const SkyWork w)
{
    extern __shared__ float4 sF[];
    __shared__ uint64_t bars[2];
    const int ntris = a.sc.ntris;
    const int nsph = a.trianglesOnly ? 0 : a.sc.nsph;
    const bool stream = w.stream != 0;
    const int tile = stream ? w.tile : (ntris > 0 ? ntris : 1);
    const int ntiles = (ntris + tile - 1) / tile;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    if (!stream) {
        for (int i = threadIdx.x; i < TAB_F4 * ntris; i += FAST_THREADS)
            sF[i] = tab[i];
    } else if (threadIdx.x == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        fence_barrier_init();
    }
    __syncthreads();

    const int ci0 = (blockIdx.x * FAST_LX + (lane % FAST_LX)) * R;
    const int cj = (blockIdx.y * FAST_WARPS + warp) * FAST_LY + lane / FAST_LX;
    const int g = a.split_first + blockIdx.z;
    const int x0 = a.col_start + ci0 * a.col_step;
    const int y = a.row_start + (cj / a.band_h) * a.row_step + (cj % a.band_h);
    unsigned valid = 0;
#pragma unroll
    for (int r = 0; r < R; ++r)
        if (ci0 + r < a.out_w && cj < a.out_h)
            valid |= 1u << r;

    const float fh = (float)(a.H / 2);
    const float dy = xdiv((float)(y - a.H / 2), fh);
    const float dx0 = xdiv((float)(x0 - a.W / 2), fh);
    auto dx_of = [&](int r) { return xdiv((float)(x0 + r * a.col_step - a.W / 2), fh); };

    const int s0 = (int)((long long)g * a.spp / a.nsplit), s1 = (int)((long long)(g + 1) * a.spp / a.nsplit);
    const float wgt = (float)(1.0 / (double)a.spp);
    unsigned sky = valid;      // pixels this lane still sweeps
    unsigned mismatch = 0;     // a later sample disagreeing with the classification (impossible: same ray)
    unsigned long long nrays = 0, nslow = 0, nhit = 0;
    uint32_t k0 = 0;

    auto issue = [&](uint32_t k, int t) {
        const int n = min(tile, ntris - t * tile);
        const uint32_t bytes = (uint32_t)n * (uint32_t)TAB_F4 * (uint32_t)sizeof(float4);
        fence_proxy_async();
        mbar_arrive_expect_tx(&bars[k & 1], bytes);
        tma_load_1d(sF + (size_t)(k & 1) * SKY_TILE * TAB_F4, tab + (size_t)t * tile * TAB_F4, bytes, &bars[k & 1]);
    };

    for (int s = s0; s < s1; ++s) {
        float best[R];
        int id[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            best[r] = 999999.0f;
            id[r] = -1;
        }
        if (nsph > 0) {
#pragma unroll
            for (int r = 0; r < R; ++r)
                if ((sky >> r) & 1u) {
                    f3 d = primary_dir(a, dx_of(r), dy);
                    for (int k = 0; k < nsph; ++k)
                        sphere_test_exact(a.origin, d, __ldg(&a.sc.s0[k]), k, best[r], id[r]);
                }
        }
        if (stream && threadIdx.x == 0) {
            issue(k0, 0);
            if (ntiles > 1)
                issue(k0 + 1, 1);
        }
        const bool warp_busy = __any_sync(0xffffffffu, sky != 0);
        for (int t = 0; t < ntiles; ++t) {
            const float4 *F = sF;
            int n = ntris, base = 0;
            if (stream) {
                const uint32_t k = k0 + (uint32_t)t;
                mbar_wait(&bars[k & 1], (k >> 1) & 1u);
                F = sF + (size_t)(k & 1) * SKY_TILE * TAB_F4;
                n = min(tile, ntris - t * tile);
                base = t * tile;
            }
            if (warp_busy) {
#pragma unroll 2
                for (int i = 0; i < n; ++i) {
                    const float4 f0 = F[TAB_F4 * i], f1 = F[TAB_F4 * i + 1], f2 = F[TAB_F4 * i + 2], f3_ = F[TAB_F4 * i + 3];
                    float e0 = fmaf(f0.x, dx0, fmaf(f0.y, dy, f0.z));
                    float e1 = fmaf(f1.x, dx0, fmaf(f1.y, dy, f1.z));
                    float e2 = fmaf(f2.x, dx0, fmaf(f2.y, dy, f2.z));
                    float e3 = fmaf(f3_.x, dx0, fmaf(f3_.y, dy, f3_.z));
                    int acc = -1;
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        acc &= (__float_as_int(e0) | __float_as_int(e1) | __float_as_int(e2) | __float_as_int(e3));
                        if (r + 1 < R) {
                            e0 += f0.w;
                            e1 += f1.w;
                            e2 += f2.w;
                            e3 += f3_.w;
                        }
                    }
                    if (__any_sync(0xffffffffu, acc >= 0)) {
                        if (acc >= 0 && sky != 0) {
                            const int ti = base + i;
                            const float4 p0 = __ldg(&a.sc.geo[3 * ti]), p1 = __ldg(&a.sc.geo[3 * ti + 1]),
                                         p2 = __ldg(&a.sc.geo[3 * ti + 2]);
#pragma unroll
                            for (int r = 0; r < R; ++r) {
                                const float dxr = dx_of(r);
                                const float t0 = fmaf(f0.x, dxr, fmaf(f0.y, dy, f0.z));
                                const float t1 = fmaf(f1.x, dxr, fmaf(f1.y, dy, f1.z));
                                const float t2 = fmaf(f2.x, dxr, fmaf(f2.y, dy, f2.z));
                                const float t3 = fmaf(f3_.x, dxr, fmaf(f3_.y, dy, f3_.z));
                                const int bits =
                                    __float_as_int(t0) | __float_as_int(t1) | __float_as_int(t2) | __float_as_int(t3);
                                if (bits >= 0 && ((sky >> r) & 1u)) {
                                    unsigned dummy[1];
                                    f3 d = primary_dir(a, dxr, dy);
                                    tri_test_exact<false>(a.origin, d, p0, p1, p2, ti, best[r], id[r], dummy);
                                    nslow++;
                                }
                            }
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
        nrays += __popc(sky);

        unsigned hits = 0;
#pragma unroll
        for (int r = 0; r < R; ++r)
            if (id[r] != -1 && ((sky >> r) & 1u))
                hits |= 1u << r;
        if (s == s0) {
            // ---- classification: hit pixels go to the path kernel's list; ids are final ----
            if (g == a.split_first) {
                if (a.ids != nullptr) {
#pragma unroll
                    for (int r = 0; r < R; ++r)
                        if ((valid >> r) & 1u)
                            a.ids[(size_t)cj * a.out_w + ci0 + r] = id[r];
                }
                nhit = __popc(hits);
                if (w.list != nullptr) {
                    // warp-aggregated append (order is irrelevant: every pixel is independent)
                    unsigned cnt = __popc(hits), incl = cnt;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        unsigned v = __shfl_up_sync(0xffffffffu, incl, o);
                        if (lane >= o)
                            incl += v;
                    }
                    const unsigned tot = __shfl_sync(0xffffffffu, incl, 31);
                    unsigned base = 0;
                    if (tot) {
                        if (lane == 31)
                            base = atomicAdd(w.list_count, tot);
                        base = __shfl_sync(0xffffffffu, base, 31);
                        unsigned off = base + incl - cnt;
#pragma unroll
                        for (int r = 0; r < R; ++r)
                            if ((hits >> r) & 1u)
                                w.list[off++] = cj * a.out_w + ci0 + r;
                    }
                }
            }
            sky &= ~hits;
            nrays -= __popc(hits); // the classification sweep of a hit pixel is not a counted ray
            const bool more = stream ? (__syncthreads_or(sky != 0) != 0) : __any_sync(0xffffffffu, sky != 0);
            if (!more)
                break;
        } else {
            mismatch |= hits;
        }
    }

    // ---- sky pixels: spp identical contributions, added with the reference's serial float adds ----
    for (int r = 0; r < R; ++r) {
        if (!((sky >> r) & 1u))
            continue;
        f3 d = primary_dir(a, dx_of(r), dy);
        f3 env = environment_light(d, a.sky);
        f3 light = xadd3(mk3(0.f, 0.f, 0.f), xmul3(env, mk3(1.f, 1.f, 1.f))); // raytracing.c:291
        f3 c = xscale3(light, wgt);
        f3 acc = mk3(0.f, 0.f, 0.f);
        for (int s = s0; s < s1; ++s)
            acc = xadd3(acc, c);
        const size_t pix = (size_t)cj * a.out_w + ci0 + r;
        if (a.accum != nullptr) {
            float *p = a.accum + ((size_t)blockIdx.z * a.out_w * a.out_h + pix) * 3;
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
        unsigned long long r_ = nrays, hp = nhit, sl = nslow, bad = __popc(mismatch);
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

} // namespace rtc
