// rtc_kernels.cuh -- sm_100a kernels of the render loop.
//
//   flatten_kernel     Triangle[] AoS (68 B) -> SoA planes (+ AB/AC hoist)
//   render_kernel      the fused per-pixel loop: ray generation, closest-hit
//                      sweep, bounce shading with the reference RNG stream,
//                      radiance accumulation, quantise, store
//   sum_quantize_kernel  ordered sum of sample-split planes + quantise
//
// Mapping (DESIGN.md "Kernel"): one lane owns one pixel (one RNG stream: the
// stream is serial per pixel, reference main.c:95-99).  A lane is a small state
// machine over (sample, bounce); the warp iterates "sweep one ray per lane
// against every triangle, then shade", so lanes of different path lengths stay
// busy until their own pixel is finished (no per-sample lock-step).
#pragma once

#include "rtc_scene.cuh"

namespace rtc {

constexpr int TILE_W = 8;  // pixels per warp in x
constexpr int TILE_H = 4;  // pixels per warp in y
constexpr int CTA_WARPS_X = 2, CTA_WARPS_Y = 2;
constexpr int CTA_THREADS = 32 * CTA_WARPS_X * CTA_WARPS_Y;
constexpr int CTA_W = TILE_W * CTA_WARPS_X, CTA_H = TILE_H * CTA_WARPS_Y;

// ---- flatten ------------------------------------------------------------------
// One thread per triangle; raw is the caller's Triangle[] viewed as floats (17 per record).
__global__ void flatten_kernel(const float *__restrict__ raw, int n, float4 *geo, float4 *m0, float *m1)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    const float *t = raw + (size_t)i * 17;
    f3 A = mk3(t[0], t[1], t[2]), B = mk3(t[3], t[4], t[5]), C = mk3(t[6], t[7], t[8]);
    f3 AB = xsub3(B, A), AC = xsub3(C, A); // reference raytracing.c:191-192
    geo[3 * i] = make_float4(A.x, A.y, A.z, t[9]);
    geo[3 * i + 1] = make_float4(AB.x, AB.y, AB.z, t[10]);
    geo[3 * i + 2] = make_float4(AC.x, AC.y, AC.z, t[11]);
    m0[i] = make_float4(t[12], t[13], t[14], t[15]);
    m1[i] = t[16];
}

__global__ void flatten_spheres_kernel(const float *__restrict__ raw, int n, float4 *s0, float4 *s1, float *s2)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    const float *t = raw + (size_t)i * 9;
    s0[i] = make_float4(t[0], t[1], t[2], t[3]);
    s1[i] = make_float4(t[4], t[5], t[6], t[7]);
    s2[i] = t[8];
}

// ---- the reference's ray-triangle test, operation for operation ------------------
// raytracing.c:186-214 with AB/AC hoisted.  Comparisons are written exactly as
// the reference's so NaNs take the same branches; `x < 0.001` in double equals
// `x < 0.001f` for float x.
template <bool COUNT>
__device__ __forceinline__ void tri_test_exact(const f3 &o, const f3 &d, const float4 &p0, const float4 &p1,
                                               const float4 &p2, int idx, float &best, int &bestId,
                                               unsigned *exits)
{
    f3 n = mk3(p0.w, p1.w, p2.w);
    if (xdot(d, n) >= 0.0f) {
        if (COUNT) exits[0]++;
        return;
    }
    f3 AB = mk3(p1.x, p1.y, p1.z), AC = mk3(p2.x, p2.y, p2.z);
    f3 h = xcross(d, AC);
    float det = xdot(AB, h);
    if (-0.001f < det && det < 0.001f) {
        if (COUNT) exits[1]++;
        return;
    }
    float inv = xrcp(det);
    f3 s = xsub3(o, mk3(p0.x, p0.y, p0.z));
    float u = xmul(xdot(s, h), inv);
    if (u < 0.0f || u > 1.0f) {
        if (COUNT) exits[2]++;
        return;
    }
    f3 q = xcross(s, AB);
    float v = xmul(xdot(d, q), inv);
    if (v < 0.0f || xadd(u, v) > 1.0f) {
        if (COUNT) exits[3]++;
        return;
    }
    float dst = xmul(xdot(AC, q), inv);
    if (dst < 0.001f) {
        if (COUNT) exits[4]++;
        return;
    }
    if (COUNT) exits[5]++;
    if (dst < best) { // strict <: the first (lowest) index wins ties, raytracing.c:232
        best = dst;
        bestId = idx;
    }
}

// raytracing.c:162-184; the normal is recomputed by the shader for the winner only
__device__ __forceinline__ void sphere_test_exact(const f3 &o, const f3 &d, const float4 &s0, int k, float &best,
                                                  int &bestId)
{
    f3 off = xsub3(o, mk3(s0.x, s0.y, s0.z));
    float b = xdot(off, d);
    float c = xsub(xdot(off, off), xmul(s0.w, s0.w));
    float delta = xsub(xmul(b, b), c);
    if (delta < 0.0f)
        return;
    delta = xsqrt(delta);
    float dst = xsub(-b, delta);
    if (dst < 0.001f)
        dst = xadd(-b, delta);
    if (dst < 0.001f)
        return;
    if (dst < best) {
        best = dst;
        bestId = -2 - k;
    }
}

// ---- per-lane path state ---------------------------------------------------------
struct Lane {
    f3 pos, dir;     // current ray
    f3 tint, light;  // rayColor / incomingLight of the current path (raytracing.c:265-266)
    f3 acc;          // accumulatedColor of the pixel (main.c:96)
    uint32_t rng;    // rngState (scene.h:35), seeded per pixel (main.c:95)
    int sample, bounce;
};

// Shade one closest-hit result; returns true when the PATH (sample) is finished.
// Restates calcColor's loop body, raytracing.c:270-293.
// miss: sky, the path ends (raytracing.c:289-293)
__device__ __forceinline__ void shade_miss(Lane &L, const RenderArgs &a)
{
    f3 env = environment_light(L.dir, a.sky);
    L.light = xadd3(L.light, xmul3(env, L.tint));
}

// hit (bestId != -1): raytracing.c:272-287; returns true when the path is finished
__device__ __forceinline__ bool shade_hit(Lane &L, const RenderArgs &a, float best, int bestId)
{
    // closest.hitPoint = pos + dir*dst (raytracing.c:238)
    f3 hp = xadd3(L.pos, xscale3(L.dir, best));
    f3 n;
    float4 m0;
    float smooth;
    if (bestId >= 0) {
        n = mk3(__ldg(&a.sc.geo[3 * bestId]).w, __ldg(&a.sc.geo[3 * bestId + 1]).w, __ldg(&a.sc.geo[3 * bestId + 2]).w);
        m0 = __ldg(&a.sc.m0[bestId]);
        smooth = __ldg(&a.sc.m1[bestId]);
    } else {
        int k = -2 - bestId;
        float4 s0 = __ldg(&a.sc.s0[k]);
        n = xnormalize(xsub3(hp, mk3(s0.x, s0.y, s0.z))); // raytracing.c:181-182
        m0 = __ldg(&a.sc.s1[k]);
        smooth = __ldg(&a.sc.s2[k]);
    }
    f3 color = mk3(m0.x, m0.y, m0.z);
    const float4 rd = random_direction(L.rng);
    L.rng = __float_as_uint(rd.w);
    f3 diffuse = xnormalize(xadd3(n, mk3(rd.x, rd.y, rd.z))); // :274
    f3 specular = xreflect(L.dir, n);                           // :275
    L.dir = xlerp(diffuse, specular, smooth);                   // :276, not renormalised
    L.pos = hp;                                                 // :277
    f3 emitted = xscale3(color, m0.w);                          // :279
    L.light = xadd3(L.light, xmul3(emitted, L.tint));           // :280
    L.tint = xmul3(L.tint, color);                              // :281
    float p = fmaxf(fmaxf(L.tint.x, L.tint.y), L.tint.z);       // :284 fmax: NaN-ignoring, exact
    if (p < random_value(L.rng))                                // :285
        return true;
    L.tint = xscale3(L.tint, xrcp(p));                          // :287
    L.bounce++;
    return L.bounce >= a.maxBounce;                             // loop bound, :268
}

__device__ __forceinline__ bool shade(Lane &L, const RenderArgs &a, float best, int bestId)
{
    if (bestId == -1) {
        shade_miss(L, a);
        return true;
    }
    return shade_hit(L, a, best, bestId);
}

// ---- the fused render kernel (exact sweep) ---------------------------------------
template <bool COUNT>
__global__ void __launch_bounds__(CTA_THREADS) render_kernel(const RenderArgs a)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ci = blockIdx.x * CTA_W + (warp % CTA_WARPS_X) * TILE_W + (lane % TILE_W);
    const int cj = blockIdx.y * CTA_H + (warp / CTA_WARPS_X) * TILE_H + (lane / TILE_W);
    const int g = a.split_first + blockIdx.z; // sub-stream
    const bool valid = ci < a.out_w && cj < a.out_h;
    const int x = a.col_start + ci * a.col_step;
    const int y = a.row_start + (cj / a.band_h) * a.row_step + (cj % a.band_h);

    // main.c:88-94 ray generation
    float dx = xdiv((float)(x - a.W / 2), (float)(a.H / 2));
    float dy = xdiv((float)(y - a.H / 2), (float)(a.H / 2));
    f3 pdir = xadd3(xadd3(xscale3(a.ex, dx), xscale3(a.ey, dy)), xscale3(a.ez, a.fov));
    pdir = xnormalize(pdir);

    Lane L;
    L.pos = a.origin;
    L.dir = pdir;
    L.tint = mk3(1.f, 1.f, 1.f);
    L.light = mk3(0.f, 0.f, 0.f);
    L.acc = mk3(0.f, 0.f, 0.f);
    L.rng = (uint32_t)(x + y * a.W) + (uint32_t)g * (uint32_t)(a.W * a.H); // main.c:95 (+ sub-stream offset)
    const int s1 = (int)((long long)(g + 1) * a.spp / a.nsplit);
    L.sample = (int)((long long)g * a.spp / a.nsplit);
    L.bounce = 0;
    const float w = (float)(1.0 / (double)a.spp); // main.c:99, (float)(1./accumulationCount)

    bool active = valid && L.sample < s1 && a.maxBounce > 0;
    int firstId = -1;
    bool first = true;
    unsigned long long nrays = 0;
    unsigned exits[6] = {0, 0, 0, 0, 0, 0};
    const int ntris = a.sc.ntris;
    const int nsph = a.trianglesOnly ? 0 : a.sc.nsph;

    // a pixel whose loop never runs still accumulates spp zeros; nothing to do (acc stays 0)
    if (valid && a.ids != nullptr && g == a.split_first && !active) {
        // primary id is defined independently of spp/maxBounce: do one sweep for it
        float best = 999999.0f;
        int id = -1;
        for (int k = 0; k < nsph; ++k)
            sphere_test_exact(L.pos, L.dir, __ldg(&a.sc.s0[k]), k, best, id);
        for (int i = 0; i < ntris; ++i)
            tri_test_exact<false>(L.pos, L.dir, __ldg(&a.sc.geo[3 * i]), __ldg(&a.sc.geo[3 * i + 1]), __ldg(&a.sc.geo[3 * i + 2]), i, best,
                                  id, exits);
        firstId = id;
    }

    while (__any_sync(0xffffffffu, active)) {
        if (active) {
            // calculateRayCollision, raytracing.c:216-240: spheres first, then triangles in order
            float best = 999999.0f;
            int bestId = -1;
            for (int k = 0; k < nsph; ++k)
                sphere_test_exact(L.pos, L.dir, __ldg(&a.sc.s0[k]), k, best, bestId);
#pragma unroll 2
            for (int i = 0; i < ntris; ++i)
                tri_test_exact<COUNT>(L.pos, L.dir, __ldg(&a.sc.geo[3 * i]), __ldg(&a.sc.geo[3 * i + 1]), __ldg(&a.sc.geo[3 * i + 2]), i,
                                      best, bestId, exits);
            nrays++;
            if (first) {
                firstId = bestId;
                first = false;
            }
            if (shade(L, a, best, bestId)) {
                // main.c:99: acc = acc + light * (float)(1./spp)
                L.acc = xadd3(L.acc, xscale3(L.light, w));
                L.sample++;
                if (L.sample >= s1) {
                    active = false;
                } else {
                    L.pos = a.origin;
                    L.dir = pdir;
                    L.tint = mk3(1.f, 1.f, 1.f);
                    L.light = mk3(0.f, 0.f, 0.f);
                    L.bounce = 0;
                }
            }
        }
    }

    if (valid) {
        const size_t pix = (size_t)cj * a.out_w + ci;
        if (a.accum != nullptr) {
            float *p = a.accum + ((size_t)blockIdx.z * a.out_w * a.out_h + pix) * 3;
            p[0] = L.acc.x;
            p[1] = L.acc.y;
            p[2] = L.acc.z;
        }
        if (a.rgb != nullptr && a.nsplit == 1) { // vec3ToColor, raytracing.c:11-15
            uint8_t *p = a.rgb + pixel_offset(a, (int)pix);
            p[0] = (uint8_t)float_to_uint8(L.acc.x);
            p[1] = (uint8_t)float_to_uint8(L.acc.y);
            p[2] = (uint8_t)float_to_uint8(L.acc.z);
        }
        if (a.ids != nullptr && g == a.split_first)
            a.ids[pix] = firstId;
    }

    if (a.stats != nullptr) {
        unsigned long long hp = (valid && g == a.split_first && firstId != -1) ? 1ull : 0ull;
        unsigned long long r = nrays;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            r += __shfl_xor_sync(0xffffffffu, r, o);
            hp += __shfl_xor_sync(0xffffffffu, hp, o);
        }
        if (COUNT) {
#pragma unroll
            for (int e = 0; e < 6; ++e) {
                unsigned long long v = exits[e];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1)
                    v += __shfl_xor_sync(0xffffffffu, v, o);
                if (lane == 0 && v)
                    atomicAdd(&a.stats[ST_EXIT0 + e], v);
            }
        }
        if (lane == 0) {
            atomicAdd(&a.stats[ST_RAYS], r);
            atomicAdd(&a.stats[ST_TRI_PAIRS], r * (unsigned long long)ntris);
            atomicAdd(&a.stats[ST_SPH_PAIRS], r * (unsigned long long)nsph);
            if (a.ids != nullptr)
                atomicAdd(&a.stats[ST_HIT_PIXELS], hp);
        }
    }
}

// ---- ordered sum of sub-stream planes + quantise (sample split) ------------------
// total = acc_0 + acc_1 + ... in g order (oracle definition of nsplit), then
// floatToUint.  planes: [G][n] floats; rgb: [n] bytes; n = out_w*out_h*3.
__global__ void sum_quantize_kernel(const float *__restrict__ planes, int G, size_t n, uint8_t *rgb, float *sum_out)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    float t = planes[i];
    if (G > 1) {
        t = xadd(0.0f, t);
        for (int g = 1; g < G; ++g)
            t = xadd(t, planes[(size_t)g * n + i]);
    }
    if (rgb)
        rgb[i] = (uint8_t)float_to_uint8(t);
    if (sum_out)
        sum_out[i] = t;
}

} // namespace rtc
