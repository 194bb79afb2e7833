// rtc_fast.cuh -- per-frame tables of the conservative primary-ray pre-filter (triangles: edge functions; spheres:
// discriminant bound), used by the primary kernels in rtc_sky.cuh / rtc_skyblock.cuh; every pair that survives runs the
// reference exact sequence.
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

// float4 per triangle in the edge-function table: 4 edge functions + 1 pad.  The 80-byte stride makes the
// lane-per-triangle loads of the warp-level bound (lane j reads entry j) conflict-free in shared memory
// (a 64-byte stride puts every other lane of a quarter-warp on the same four banks).
constexpr int TAB_F4 = 5;

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
        tab[TAB_F4 * i + k] = out[k];
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

// ---- shared by the primary kernels (rtc_sky.cuh, rtc_skyblock.cuh) -------------------------------------------
// Lane layout of the per-pair kernel: 8 x 4 lanes per warp, R horizontally adjacent pixels per lane, 4 warps per CTA.
constexpr int FAST_LX = 8, FAST_LY = 4, FAST_WARPS = 4, FAST_THREADS = 32 * FAST_WARPS;

// primary ray direction of pixel (x, y): main.c:88-93, exact
__device__ __forceinline__ f3 primary_dir(const RenderArgs &a, float dx, float dy)
{
    f3 d = xadd3(xadd3(xscale3(a.ex, dx), xscale3(a.ey, dy)), xscale3(a.ez, a.fov));
    return xnormalize(d);
}


} // namespace rtc
