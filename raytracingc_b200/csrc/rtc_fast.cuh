// rtc_fast.cuh -- the production render kernel for scenes whose tables fit in
// shared memory: screen-space conservative pre-filter for primary rays + the
// reference's exact ray-triangle sequence for every pair that survives it.
//
// Why a pre-filter is legal (DESIGN.md "Exactness"): all primary rays share the
// camera origin o, so for triangle (A, AB, AC, n) the reference's quantities are,
// in real arithmetic, linear in the ray direction d:
//     c   = d . n                    (cull:   c >= 0  -> miss)
//     det = d . (AC x AB)            u*det = d . (AC x s)      s = o - A
//     tq  = AC . (s x AB)  (const)   v*det = d . (s x AB)      t = tq / det
// and d is a positive multiple of D = ex*dx + ey*dy + ez*fov, which is affine in
// the pixel's (dx, dy).  So "the reference accepts this pair" implies four
// affine functions of (dx, dy) are >= -(rounding slack):
//     E0 = -(D.n)                  E1 = sg*(D.(AC x s))       sg = sign(tq)
//     E2 = sg*(D.(s x AB))         E3 = sg*(D.(AC x AB)) - E1 - E2
// prepare_primary_kernel computes their coefficients in double, adds a rigorous
// bound of (reference rounding error + our evaluation error) to the constant
// term, and the render kernel rejects a pair only if some E_k evaluates < 0.
// Rejected pairs are pairs the reference rejects too; every other pair runs
// tri_test_exact, so hit decisions, distances and ids are bit-identical.
#pragma once

#include "rtc_kernels.cuh"

namespace rtc {

struct PrimaryFrame {
    f3 origin, ex, ey, ez;
    float fov;
    double dxmax, dymax; // max |dx|, |dy| over the full image
    double Dmax;         // max |D| over the full image (a corner)
    double xstep;        // col_step / (H/2): dx distance between the R pixels of a lane
    int max_steps;       // R - 1
    int block;           // 1: table for the block kernel (rtc_skyblock.cuh): {a, b, c_block, c_pixel}
    double hw, hh;       // block mode: half extents of a lane's pixel block in (dx, dy), already inflated
};

// table: 4 x float4 per triangle, one per edge function: {a, b, c + margin, a * xstep}
// (block mode: {a, b, c + margin + |a| hw + |b| hh, c + margin}: the .z constant bounds the edge function over a
// whole pixel block from its value at the block centre, the .w constant is the per-pixel one)
__global__ void prepare_primary_kernel(const DeviceScene sc, const PrimaryFrame fr, float4 *__restrict__ tab)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= sc.ntris)
        return;
    const float4 p0 = sc.geo[3 * i], p1 = sc.geo[3 * i + 1], p2 = sc.geo[3 * i + 2];
    const f3 A = mk3(p0.x, p0.y, p0.z), AB = mk3(p1.x, p1.y, p1.z), AC = mk3(p2.x, p2.y, p2.z);
    const f3 n = mk3(p0.w, p1.w, p2.w);
    // exactly the reference's float values for a primary ray: s, q = s x AB, tq = AC . q
    const f3 s = xsub3(fr.origin, A);
    const f3 q = xcross(s, AB);
    const float tq = xdot(AC, q);

    const double u = 1.1920928955078125e-07; // 2^-23: twice the unit roundoff (safety factor 2)
    double X[4][3];
    double sg = tq > 0.0f ? 1.0 : -1.0;
    const double M[3] = {(double)AC.y * s.z - (double)AC.z * s.y, (double)AC.z * s.x - (double)AC.x * s.z,
                         (double)AC.x * s.y - (double)AC.y * s.x}; // AC x s
    const double Nab[3] = {(double)AC.y * AB.z - (double)AC.z * AB.y, (double)AC.z * AB.x - (double)AC.x * AB.z,
                           (double)AC.x * AB.y - (double)AC.y * AB.x}; // AC x AB
    const double Q[3] = {(double)q.x, (double)q.y, (double)q.z};
    const double N[3] = {(double)n.x, (double)n.y, (double)n.z};
    for (int k = 0; k < 3; ++k) {
        X[0][k] = -N[k];
        X[1][k] = sg * M[k];
        X[2][k] = sg * Q[k];
        X[3][k] = sg * (Nab[k] - M[k] - Q[k]);
    }
    auto l1 = [](double x, double y, double z) { return fabs(x) + fabs(y) + fabs(z); };
    const double n1 = l1(N[0], N[1], N[2]), s1 = l1(s.x, s.y, s.z), ab1 = l1(AB.x, AB.y, AB.z),
                 ac1 = l1(AC.x, AC.y, AC.z), m1 = l1(M[0], M[1], M[2]), q1 = l1(Q[0], Q[1], Q[2]),
                 nab1 = l1(Nab[0], Nab[1], Nab[2]);
    // slack of the reference's own float evaluation relative to real arithmetic, in units of |D|:
    //   direction error |d - D/|D||_inf <= 12u; a 3-term float dot errs <= 3u*sum|terms|; a float cross
    //   component errs <= 2u*sum|terms|
    const double B1 = 5.0 * s1 * ac1 + 12.0 * m1;
    const double B2 = 15.0 * q1;
    double delta[4];
    delta[0] = 15.0 * n1;
    delta[1] = B1;
    delta[2] = B2;
    delta[3] = 11.0 * ab1 * ac1 + 12.0 * nab1 + B1 + B2;
    const double tiny = 1e-38 * ab1 * ac1; // products that underflow to -0 pass the reference's `< 0` tests

    const double ex[3] = {fr.ex.x, fr.ex.y, fr.ex.z}, ey[3] = {fr.ey.x, fr.ey.y, fr.ey.z},
                 ez[3] = {fr.ez.x, fr.ez.y, fr.ez.z};
    bool usable = isfinite(tq) && tq != 0.0f;
    float4 out[4];
    for (int k = 0; k < 4; ++k) {
        double a = ex[0] * X[k][0] + ex[1] * X[k][1] + ex[2] * X[k][2];
        double b = ey[0] * X[k][0] + ey[1] * X[k][1] + ey[2] * X[k][2];
        double c = (double)fr.fov * (ez[0] * X[k][0] + ez[1] * X[k][1] + ez[2] * X[k][2]);
        double range = fabs(a) * fr.dxmax + fabs(b) * fr.dymax + fabs(c);
        // our own evaluation: coefficient rounding, two FMAs, up to max_steps incremental adds
        double eps = (8.0 + 2.0 * fr.max_steps) * range;
        double margin = 2.0 * u * (delta[k] * fr.Dmax + eps) + tiny * fr.Dmax + 1e-30;
        double cm = c + margin;
        if (fr.block) {
            // E(x, y) <= E(centre) + |a| hw + |b| hh over the block; the float coefficients (float)a, (float)b
            // are the ones the kernel multiplies with, their rounding is inside eps
            double ext = (fabs(a) * fr.hw + fabs(b) * fr.hh) * 1.0000005 + 2.0 * u * range;
            // the constants are rounded up so the float table entry is >= the real bound
            out[k] = make_float4((float)a, (float)b, __double2float_ru(cm + ext), __double2float_ru(cm));
        } else {
            out[k] = make_float4((float)a, (float)b, (float)cm, (float)(a * fr.xstep));
        }
        // too large / too small to reason about in float: let the exact path decide
        if (!isfinite(a) || !isfinite(b) || !isfinite(cm) || range > 1e30 || margin > 1e30)
            usable = false;
    }
    if (!usable) {
        if (isfinite(tq) && tq == 0.0f) {
            // t = tq * inv = +-0 -> `dst < EPSILON` holds: a primary ray can never hit this triangle
            for (int k = 0; k < 4; ++k)
                out[k] = make_float4(0.f, 0.f, -1.f, fr.block ? -1.f : 0.f);
        } else {
            for (int k = 0; k < 4; ++k)
                out[k] = make_float4(0.f, 0.f, 1.f, fr.block ? 1.f : 0.f); // always pass
        }
    }
    for (int k = 0; k < 4; ++k)
        tab[4 * i + k] = out[k];
}

// ---- spheres: conservative bound of the discriminant over a pixel block ---------------------------------------
// raySphere (raytracing.c:162-184) rejects a ray when delta = b*b - c < 0, b = off.d, c = off.off - r*r,
// off = o - centre.  For primary rays d = D/|D| with D affine in (dx, dy), so |D|^2 * delta is the quadratic
//     Q(dx, dy) = (D.off)^2 - |D|^2 * c = A dx^2 + B dy^2 + C dx dy + E dx + F dy + G ,
// and over a block centred at (cx, cy) with half extents (hw, hh), by the (exact) second-order Taylor expansion,
//     max Q <= Q(c) + |Qx(c)| hw + |Qy(c)| hh + |A| hw^2 + |B| hh^2 + |C| hw hh .
// The table entry holds A, B, C, E, F and G' = G + block term + margins, where the margins bound the reference's
// float error on delta (41 * 2^-23 * (|off|_1^2 + r^2), scaled by max |D|^2) and this kernel's own evaluation
// error.  max Q < 0  =>  the reference computes delta < 0 for every pixel of the block  =>  miss.
__global__ void prepare_spheres_kernel(const DeviceScene sc, const PrimaryFrame fr, float *__restrict__ sphq)
{
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= sc.nsph)
        return;
    const float4 s0 = sc.s0[k];
    const f3 offf = xsub3(fr.origin, mk3(s0.x, s0.y, s0.z)); // the reference's float subtraction
    const double off[3] = {offf.x, offf.y, offf.z};
    const double ex[3] = {fr.ex.x, fr.ex.y, fr.ex.z}, ey[3] = {fr.ey.x, fr.ey.y, fr.ey.z},
                 ez[3] = {fr.ez.x, fr.ez.y, fr.ez.z};
    auto dot = [](const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; };
    const double fov = fr.fov, r = s0.w;
    const double al = dot(ex, off), be = dot(ey, off), ga = fov * dot(ez, off);
    const double cc = dot(off, off) - r * r;
    const double A = al * al - dot(ex, ex) * cc, B = be * be - dot(ey, ey) * cc;
    const double C = 2 * al * be - 2 * dot(ex, ey) * cc;
    const double E = 2 * al * ga - 2 * fov * dot(ex, ez) * cc, F = 2 * be * ga - 2 * fov * dot(ey, ez) * cc;
    const double G = ga * ga - fov * fov * dot(ez, ez) * cc;
    const double X = fr.dxmax, Y = fr.dymax;
    const double l1 = fabs(off[0]) + fabs(off[1]) + fabs(off[2]);
    const double u = 1.1920928955078125e-07; // 2^-23
    const double S = fabs(A) * X * X + fabs(B) * Y * Y + fabs(C) * X * Y + fabs(E) * X + fabs(F) * Y + fabs(G) +
                     (2 * fabs(A) * X + fabs(C) * Y + fabs(E)) * fr.hw + (2 * fabs(B) * Y + fabs(C) * X + fabs(F)) * fr.hh;
    const double m_ref = 41.0 * u * (l1 * l1 + r * r) * fr.Dmax * fr.Dmax;
    const double m_eval = 64.0 * u * S;
    const double blk = (fabs(A) * fr.hw * fr.hw + fabs(B) * fr.hh * fr.hh + fabs(C) * fr.hw * fr.hh) * 1.000001;
    float *o = sphq + 8 * k;
    bool ok = isfinite(A) && isfinite(B) && isfinite(C) && isfinite(E) && isfinite(F) && isfinite(G) && S < 1e30;
    if (ok) {
        o[0] = (float)A; o[1] = (float)B; o[2] = (float)C; o[3] = (float)E; o[4] = (float)F;
        o[5] = __double2float_ru(G + blk + m_ref + m_eval + 1e-30);
        o[6] = o[7] = 0.f;
    } else { // cannot reason in float: every block is a candidate
        o[0] = o[1] = o[2] = o[3] = o[4] = 0.f;
        o[5] = 1.f;
        o[6] = o[7] = 0.f;
    }
}

// ---- fast render kernel -------------------------------------------------------------
// Lane layout: 8 x 4 lanes per warp, each lane owns R horizontally adjacent pixels
// (warp tile 8R x 4); a CTA is 4 warps stacked in y (CTA tile 8R x 16).
constexpr int FAST_LX = 8, FAST_LY = 4, FAST_WARPS = 4, FAST_THREADS = 32 * FAST_WARPS;

struct FastSmem {
    float4 *F;      // [4 * ntris] edge-function table
    float4 *G;      // [3 * ntris] geometry planes g0,g1,g2 interleaved
    float4 *state;  // [THREADS * R] {rng bits, acc.x, acc.y, acc.z}
    float2 *hit;    // [THREADS * R] {best, id bits}
};

__host__ __device__ inline size_t fast_smem_bytes(int ntris, int R)
{
    return (size_t)ntris * 7 * sizeof(float4) + (size_t)FAST_THREADS * R * (sizeof(float4) + sizeof(float2) + 1);
}

// primary ray direction of pixel (x, y): main.c:88-93, exact
__device__ __forceinline__ f3 primary_dir(const RenderArgs &a, float dx, float dy)
{
    f3 d = xadd3(xadd3(xscale3(a.ex, dx), xscale3(a.ey, dy)), xscale3(a.ez, a.fov));
    return xnormalize(d);
}

// The reference's ray-triangle test arranged for SIMT: the values the reference computes up to
// s.h (its c, det and s.h, same operations, same order) are evaluated straight-line for every pair,
// which gives the scheduler independent work across unrolled triangles; the exits are then taken on
// those exact values.  Two of them are decided without the division:
//   * sign(s.h) != sign(det), |s.h| >= 1e-30, |det| <= 2^20  =>  u = fl(s.h * fl(1/det)) < 0   (no underflow to -0)
//   * same sign and |s.h| > |det| * 1.00001                     =>  u > 1                       (margin >> 3 roundings)
// Every other pair continues with the reference's remaining operations, so decisions are identical.
__device__ __forceinline__ void tri_test_exact_sl(const f3 &o, const f3 &d, const float4 &p0, const float4 &p1,
                                                  const float4 &p2, int idx, float &best, int &bestId)
{
    const f3 n = mk3(p0.w, p1.w, p2.w);
    const f3 AB = mk3(p1.x, p1.y, p1.z), AC = mk3(p2.x, p2.y, p2.z);
    const float c = xdot(d, n);
    const f3 h = xcross(d, AC);
    const float det = xdot(AB, h);
    const f3 s = xsub3(o, mk3(p0.x, p0.y, p0.z));
    const float sh = xdot(s, h);
    const bool cont = !(c >= 0.0f) && !(-0.001f < det && det < 0.001f);
    const float adet = fabsf(det), ash = fabsf(sh);
    const bool opp = (__float_as_int(sh) ^ __float_as_int(det)) < 0;
    const bool rej = opp ? (ash >= 1e-30f && adet <= 1048576.0f) : (ash > xmul(adet, 1.00001f));
    if (cont && !rej) {
        const float inv = xrcp(det);
        const float u = xmul(sh, inv);
        if (u < 0.0f || u > 1.0f)
            return;
        const f3 q = xcross(s, AB);
        const float v = xmul(xdot(d, q), inv);
        if (v < 0.0f || xadd(u, v) > 1.0f)
            return;
        const float dst = xmul(xdot(AC, q), inv);
        if (dst < 0.001f)
            return;
        if (dst < best) {
            best = dst;
            bestId = idx;
        }
    }
}

// general exact closest hit against the shared-memory tables (spheres from global)
__device__ __noinline__ void closest_exact_smem(const RenderArgs &a, const float4 *__restrict__ G, int ntris, int nsph,
                                                f3 o, f3 d, float &best, int &bestId)
{
    best = 999999.0f;
    bestId = -1;
    for (int k = 0; k < nsph; ++k)
        sphere_test_exact(o, d, __ldg(&a.sc.s0[k]), k, best, bestId);
#pragma unroll 4
    for (int i = 0; i < ntris; ++i)
        tri_test_exact_sl(o, d, G[3 * i], G[3 * i + 1], G[3 * i + 2], i, best, bestId);
}

template <int R>
__global__ void __launch_bounds__(FAST_THREADS) render_fast_kernel(const RenderArgs a, const float4 *__restrict__ tab)
{
    extern __shared__ float4 smem_raw[];
    const int ntris = a.sc.ntris;
    const int nsph = a.trianglesOnly ? 0 : a.sc.nsph;
    float4 *sF = smem_raw;
    float4 *sG = sF + 4 * ntris;
    float4 *sState = sG + 3 * ntris;
    float2 *sHit = reinterpret_cast<float2 *>(sState + FAST_THREADS * R);
    unsigned char *sList = reinterpret_cast<unsigned char *>(sHit + FAST_THREADS * R); // [WARPS][32 * R]

    for (int i = threadIdx.x; i < 4 * ntris; i += FAST_THREADS)
        sF[i] = tab[i];
    for (int i = threadIdx.x; i < 3 * ntris; i += FAST_THREADS)
        sG[i] = a.sc.geo[i];
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
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

    float4 *myState = sState + threadIdx.x; // slot r at myState[r * FAST_THREADS]
    float2 *myHit = sHit + threadIdx.x;
    unsigned char *myList = sList + warp * 32 * R;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        uint32_t seed = (uint32_t)(x0 + r * a.col_step + y * a.W) + (uint32_t)g * (uint32_t)(a.W * a.H);
        myState[r * FAST_THREADS] = make_float4(__uint_as_float(seed), 0.f, 0.f, 0.f);
    }

    const int s0 = (int)((long long)g * a.spp / a.nsplit), s1 = (int)((long long)(g + 1) * a.spp / a.nsplit);
    const float w = (float)(1.0 / (double)a.spp);
    unsigned geom = 0;
    unsigned long long nrays = 0, nslow = 0;

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
            for (int r = 0; r < R; ++r) {
                f3 d = primary_dir(a, dx_of(r), dy);
                for (int k = 0; k < nsph; ++k)
                    sphere_test_exact(a.origin, d, __ldg(&a.sc.s0[k]), k, best[r], id[r]);
            }
        }
        // ---- primary sweep: conservative edge-function filter, exact test for survivors ----
#pragma unroll 2
        for (int i = 0; i < ntris; ++i) {
            const float4 f0 = sF[4 * i], f1 = sF[4 * i + 1], f2 = sF[4 * i + 2], f3_ = sF[4 * i + 3];
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
                if (acc >= 0) {
                    const float4 p0 = sG[3 * i], p1 = sG[3 * i + 1], p2 = sG[3 * i + 2];
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        const float dxr = dx_of(r);
                        const float t0 = fmaf(f0.x, dxr, fmaf(f0.y, dy, f0.z));
                        const float t1 = fmaf(f1.x, dxr, fmaf(f1.y, dy, f1.z));
                        const float t2 = fmaf(f2.x, dxr, fmaf(f2.y, dy, f2.z));
                        const float t3 = fmaf(f3_.x, dxr, fmaf(f3_.y, dy, f3_.z));
                        const int bits = __float_as_int(t0) | __float_as_int(t1) | __float_as_int(t2) | __float_as_int(t3);
                        if (bits >= 0 && ((valid >> r) & 1u)) {
                            unsigned dummy[1];
                            f3 d = primary_dir(a, dxr, dy);
                            tri_test_exact<false>(a.origin, d, p0, p1, p2, i, best[r], id[r], dummy);
                            nslow++;
                        }
                    }
                }
            }
        }
        nrays += __popc(valid);

        unsigned hits = 0;
#pragma unroll
        for (int r = 0; r < R; ++r)
            if (id[r] != -1 && ((valid >> r) & 1u))
                hits |= 1u << r;
        if (s == s0) {
            geom = hits;
            if (a.ids != nullptr && g == a.split_first) {
#pragma unroll
                for (int r = 0; r < R; ++r)
                    if ((valid >> r) & 1u)
                        a.ids[(size_t)cj * a.out_w + ci0 + r] = id[r];
            }
        }
        // ---- paths of the pixels whose primary ray hit (rare) ----
        // The warp pools its hit pixels and hands them out one per lane, so a silhouette tile with a few
        // hits costs one batch, not one serial path per owner lane.  A worker lane runs the whole path of
        // "its" pixel in registers; only {rng, acc} and the primary hit travel through shared memory.
        if (__any_sync(0xffffffffu, hits != 0)) {
            int n = 0;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const bool mine = (hits >> r) & 1u;
                const unsigned m = __ballot_sync(0xffffffffu, mine);
                if (mine) {
                    const int off = n + __popc(m & ((1u << lane) - 1u));
                    myList[off] = (unsigned char)(r * 32 + lane);
                    myHit[r * FAST_THREADS] = make_float2(best[r], __int_as_float(id[r]));
                }
                n += __popc(m);
            }
            __syncwarp();
            for (int b = 0; b < n; b += 32) {
                const int e = b + lane;
                if (e < n) {
                    const int slot = myList[e];
                    const int ol = slot & 31, r = slot >> 5; // owner lane, its pixel index
                    const int sidx = r * FAST_THREADS + warp * 32 + ol;
                    const int px = a.col_start + ((blockIdx.x * FAST_LX + (ol % FAST_LX)) * R + r) * a.col_step;
                    const int pcj = (blockIdx.y * FAST_WARPS + warp) * FAST_LY + ol / FAST_LX;
                    const int py = a.row_start + (pcj / a.band_h) * a.row_step + (pcj % a.band_h);
                    float4 st = sState[sidx];
                    const float2 h = sHit[sidx];
                    Lane L;
                    L.pos = a.origin;
                    L.dir = primary_dir(a, xdiv((float)(px - a.W / 2), fh), xdiv((float)(py - a.H / 2), fh));
                    L.tint = mk3(1.f, 1.f, 1.f);
                    L.light = mk3(0.f, 0.f, 0.f);
                    L.rng = __float_as_uint(st.x);
                    L.bounce = 0;
                    float bst = h.x;
                    int bid = __float_as_int(h.y);
                    while (!shade(L, a, bst, bid)) {
                        closest_exact_smem(a, sG, ntris, nsph, L.pos, L.dir, bst, bid);
                        nrays++;
                    }
                    st.x = __uint_as_float(L.rng);
                    st.y = xadd(st.y, xmul(L.light.x, w)); // main.c:99
                    st.z = xadd(st.z, xmul(L.light.y, w));
                    st.w = xadd(st.w, xmul(L.light.z, w));
                    sState[sidx] = st;
                }
                __syncwarp();
            }
        }
    }

    // ---- finish: sky pixels receive spp identical contributions (same ray, no RNG draw) ----
    for (int r = 0; r < R; ++r) {
        if (!((valid >> r) & 1u))
            continue;
        const float4 st = myState[r * FAST_THREADS];
        f3 acc = mk3(st.y, st.z, st.w);
        if (!((geom >> r) & 1u)) {
            f3 d = primary_dir(a, dx_of(r), dy);
            f3 env = environment_light(d, a.sky);
            f3 light = xadd3(mk3(0.f, 0.f, 0.f), xmul3(env, mk3(1.f, 1.f, 1.f))); // raytracing.c:291
            f3 c = xscale3(light, w);
            for (int s = s0; s < s1; ++s)
                acc = xadd3(acc, c);
        }
        const size_t pix = (size_t)cj * a.out_w + ci0 + r;
        if (a.accum != nullptr) {
            float *p = a.accum + ((size_t)blockIdx.z * a.out_w * a.out_h + pix) * 3;
            p[0] = acc.x;
            p[1] = acc.y;
            p[2] = acc.z;
        }
        if (a.rgb != nullptr && a.nsplit == 1) {
            uint8_t *p = a.rgb + pix * 3;
            p[0] = (uint8_t)float_to_uint8(acc.x);
            p[1] = (uint8_t)float_to_uint8(acc.y);
            p[2] = (uint8_t)float_to_uint8(acc.z);
        }
    }

    if (a.stats != nullptr) {
        unsigned long long hp = (g == a.split_first) ? __popc(geom) : 0, r_ = nrays, sl = nslow;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            r_ += __shfl_xor_sync(0xffffffffu, r_, o);
            hp += __shfl_xor_sync(0xffffffffu, hp, o);
            sl += __shfl_xor_sync(0xffffffffu, sl, o);
        }
        if (lane == 0) {
            atomicAdd(&a.stats[ST_RAYS], r_);
            atomicAdd(&a.stats[ST_TRI_PAIRS], r_ * (unsigned long long)ntris);
            atomicAdd(&a.stats[ST_SPH_PAIRS], r_ * (unsigned long long)nsph);
            atomicAdd(&a.stats[ST_SLOW_PAIRS], sl);
            if (a.ids != nullptr)
                atomicAdd(&a.stats[ST_HIT_PIXELS], hp);
        }
    }
}

} // namespace rtc
