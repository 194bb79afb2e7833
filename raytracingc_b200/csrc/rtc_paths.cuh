// rtc_paths.cuh -- the path kernel: one lane = one pixel's RNG stream; every ray's result is the reference's.
//
// A lane is a state machine over (sample, bounce) of "its" pixel (the stream is serial per pixel: seeded once,
// main.c:95, continuing across samples, main.c:98-99).  A warp round is
//     fetch tasks -> shade pending hits (once) -> sweep the lane's current ray over ALL triangles
//     (calculateRayCollision, raytracing.c:216-240) -> shade misses
// so lanes with different path lengths never wait for each other, and a lane whose pixel is finished fetches
// the next pixel from a global work list (persistent lanes: no tail beyond one pixel's own serial chain).
//
// Work list: either the compact list of primary-hit pixels produced by the classification kernel
// (rtc_skyblock.cuh / rtc_sky.cuh) together with each pixel's primary-hit record, or, with list == nullptr, every
// pixel of the subset.  A task is (pixel, plane); plane = sub-stream of a sample split.
//
// The sweep has two phases per chunk of 32 triangles: a branch-free conservative quick test of every pair
// (MODE 0: in the reference's arithmetic; MODE 1: from per-triangle quick records, quick_records_kernel below;
// MODE 2: MODE 1 + verification) and the reference's exact test (tri_test_exact) on the survivors in index order.
// Shared memory holds the geometry (48 B per triangle, rtc_scene.cuh) and, in MODE 1/2, the 64 B quick records:
//   STREAM = false  the whole scene is resident (loaded once per CTA); warps run independently.
//   STREAM = true   scenes of any size: PATH_TILE-triangle tiles are streamed by the TMA unit
//                   (cp.async.bulk + mbarrier, double-buffered); the CTA's warps sweep in lock-step.
#pragma once

#include "rtc_kernels.cuh"
#include "rtc_tma.cuh"

namespace rtc {

constexpr int PATH_THREADS = 128;
constexpr int PATH_GROUP_CTAS = 5; // CTAs per SM of the group-mode instantiation
#ifndef RTC_QUICK_UNROLL
#define RTC_QUICK_UNROLL 8 // (ray, triangle) pairs per lane and iteration of the branch-free first phase; A/B knob (4: +1 %)
#endif
#ifdef RTC_PATH_MINBLOCKS // min resident CTAs per SM asked of the compiler (register cap); A/B knob
#define RTC_PATH_MINB RTC_PATH_MINBLOCKS
#else
#define RTC_PATH_MINB 0
#endif
#ifndef RTC_PATH_MINB3 // ... of the resident packed-quick-test instantiation: 6 CTAs/SM = 80 registers (84 uncapped)
#define RTC_PATH_MINB3 6
#endif
constexpr int PATH_TILE = 128; // triangles per streamed tile: 14 KB per stage (quick records + geometry), 2 stages

struct PathWork {
    const int *list;            // compact pixel indices (cj * out_w + ci) or nullptr = all pixels
    const unsigned *list_count; // device counter written by the sky kernel (used when list != nullptr)
    unsigned *cursor;           // device work cursor, zeroed before the launch
    int planes;                 // sub-streams rendered by this launch
    int write_ids;              // 1: this kernel owns the primary-id map (no sky kernel ran)
    int tile;                   // triangles per streamed tile (<= PATH_TILE), STREAM only
    int coop;                   // 1: cooperative sweep of the last live lanes (0 only for A/B measurements)
    unsigned group_lanes;       // lanes of the group-mode instantiation's grid (both instantiations decide with it)
    int spec_fails;             // consecutive failed batches after which a pixel goes back to one sample at a time
    int spec_retry;             // ... and the run of single-hit samples after which it tries batches again (0: never)
    int spec;                   // 1: group mode runs S consecutive samples of a pixel at once (see the kernel)
    const float4 *qrec;         // MODE >= 1: quick records (4 x float4 per triangle), built by quick_records_kernel
    float obound;               // MODE >= 1: the records are valid for ray origins with max |o_i| <= obound
    unsigned min_tasks;         // the kernel returns at once when the list holds fewer (pixel, sub-stream) tasks: short
                                // lists are the pool form's (rtc_pool.cuh), the host launches both
    const float2 *rec;          // with a list: {distance, id bits} of each listed pixel's primary hit, found by the
                                // primary kernel, which also re-sweeps that ray every sample; nullptr: the path
                                // kernel sweeps primary rays itself
};

// MODE of the sweep's first phase: 0 the reference's arithmetic, 1 quick records, 2 = 1 + verification,
// 3 quick records evaluated two triangles at a time with packed FP32 (FFMA2), 4 = 3 + verification.
// (5, 6: lane = triangle with the warp's rays two at a time from a per-warp scratch.  Measured and dropped from the
// kernel -- tools/sweep_probe.cu variant 6 keeps the loop: 31.7 clocks per (warp, 32 pairs) alone against 35.2 for
// mode 3, but 3.88 ms against 3.54 ms in the C4 path kernel; the layout helpers below still describe it.)
__host__ __device__ constexpr bool mode_packed(int mode) { return mode == 3 || mode == 4; }
__host__ __device__ constexpr bool mode_transposed(int mode) { return mode == 5 || mode == 6; }
__host__ __device__ constexpr bool mode_verify(int mode) { return mode == 2 || mode == 4 || mode == 6; }

// Quick records of n triangles, in float4: 4 per triangle (MODE 1, 2), or 9 per PAIR of consecutive triangles
// (MODE 3, 4): float4 j of a pair holds {slot 2j of the even triangle, slot 2j of the odd one, slot 2j+1 of the even
// one, slot 2j+1 of the odd one} -- one LDS.128 gives two packed operands -- and the 9th float4 is padding: lanes
// that read different pairs (cooperative sweep) are then 36 words apart and an LDS.128 of a quarter warp is free of
// bank conflicts.  An odd count is completed by a record that drops every ray (thr = +inf).
constexpr int QPAIR_F4 = 9;
// MODE 5, 6: 5 float4 per triangle (4 + padding): lane j reads triangle j's record, 20 words apart -- an LDS.128 of a
// quarter warp touches every bank once.
constexpr int QTRI_F4 = 5;
__host__ __device__ constexpr int mode_layout(int mode) { return mode_packed(mode) ? 2 : (mode_transposed(mode) ? 3 : (mode ? 1 : 0)); }
__host__ __device__ constexpr size_t path_q_f4(int mode, int n)
{
    return mode == 0 ? 0
                     : (mode_packed(mode) ? (size_t)QPAIR_F4 * (size_t)((n + 1) / 2)
                                          : (size_t)(mode_transposed(mode) ? QTRI_F4 : 4) * (size_t)n);
}
// float index of slot s (0..15) of triangle i's record in layout 1 (per triangle), 2 (per pair), 3 (per triangle, padded)
__host__ __device__ constexpr size_t path_q_slot(int layout, int i, int s)
{
    return layout == 2 ? (size_t)(i >> 1) * (4 * QPAIR_F4) + (size_t)((s >> 1) * 4 + (s & 1) * 2 + (i & 1))
                       : (size_t)i * (layout == 3 ? 4 * QTRI_F4 : 16) + s;
}
// per-warp scratch of the transposed sweep: the warp's 32 rays, two per slot: {dx dx' dy dy'}{dz dz' mx mx'}{my my' mz mz'}
// + the survivor masks of a chunk, 16 x {ray 2k, ray 2k + 1}
constexpr int PATH_SCRATCH_F4 = 16 * 3 + 8;

// shared memory (per stage when streaming): [quick records of the stage's triangles][geometry, 3 float4 per triangle]
__host__ __device__ inline size_t path_smem_bytes(int ntris, bool stream, int mode)
{
    const int n = stream ? PATH_TILE : (ntris > 0 ? ntris : 1);
    return ((stream ? 2 : 1) * (path_q_f4(mode, n) + (size_t)3 * n) +
            (mode_transposed(mode) ? (size_t)PATH_SCRATCH_F4 * (PATH_THREADS / 32) : 0)) * sizeof(float4);
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

// ---- MODE 1: conservative quick test from per-triangle quick records ----------------------------------------
//
// Phase 1 does not have to use the reference's arithmetic -- it only has to keep every pair the reference
// accepts; phase 2 (tri_test_exact, in index order) produces the result.  With s = o - A and the ray's moment
// m = o x d (computed once per sweep):
//     s.(d x AC) =  AC.m + (A x AC).d        d.(s x AB) = -AB.m - (A x AB).d        AB.(d x AC) = (AC x AB).d
// so the numerators of u and v and the determinant are three dot products against per-triangle constants: 18
// FMA-pipe operations per pair (the reference's u alone costs 29), and the test sees the whole triangle, not only
// the u strip.  The values differ from the reference's floats by rounding only; M (per triangle, built in double
// by quick_records_kernel) bounds 4 x ( |reference's float - exact| + |this code's float - exact| ) for every
// ray with max|d_i| <= QUICK_DBOUND and max|o_i| <= obound (other rays skip phase 1: all pairs go to phase 2):
//     |sh_r - U|, |vn_r - V| <= 2^-17 (obound + max|A_i|) Dmax Emax        (85 u (..), u = 2^-24)
//     |det_r - det|          <= 2^-18 Dmax Emax^2                           (55 u), Emax = max |AB_i|, |AC_i|
// The record holds its vectors multiplied by f = +-1 so that "det~ > 0" means "front facing by the stored
// normal" (n = lambda N + e with N = AC x AB, lambda and e computed in double; f = -sign(lambda)).
// A pair is dropped when
//   det~ < thr      det~ <= -X: c_r = fl(d.n) >= 0 (cull, raytracing.c:189), X = M + Dmax (2^-19 nmax + 3 emax)/|lambda|;
//                   |det~| < EPS - M: |det_r| < EPS (raytracing.c:195).  thr = EPS - M if X + M < EPS, else -X
//   det~ > M' (the sign of det_r is known), M' = EPS - thr >= M, and one of
//     U~ < -M'          u_r < 0  (M >= 1e-18 keeps fl(sh_r * inv_r) away from -0)
//     V~ < -M'          v_r < 0
//     U~ + V~ > det~ (1 + 2^-19) + M'      then u_r > 1, or fl(u_r + v_r) > 1 (u_r, v_r >= 0 carry <= 2.01 u each)
// Every comparison keeps the pair when unordered (NaN) and degenerate triangles get thr = -inf (M' = inf).
constexpr float QUICK_DBOUND = 1.0625f;
constexpr float QUICK_EPS = 0.0009999f; // just below the reference's 0.001f

__global__ void quick_records_kernel(const float4 *__restrict__ geo, int n, float obound, float4 *__restrict__ q, int layout)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    const float4 g0 = geo[3 * i], g1 = geo[3 * i + 1], g2 = geo[3 * i + 2];
    const double A[3] = {g0.x, g0.y, g0.z}, ab[3] = {g1.x, g1.y, g1.z}, ac[3] = {g2.x, g2.y, g2.z}, nr[3] = {g0.w, g1.w, g2.w};
    auto cross = [](const double *a, const double *b, double *c) {
        c[0] = a[1] * b[2] - a[2] * b[1];
        c[1] = a[2] * b[0] - a[0] * b[2];
        c[2] = a[0] * b[1] - a[1] * b[0];
    };
    double N[3], PU[3], PV[3];
    cross(ac, ab, N);  // AC x AB
    cross(A, ac, PU);  // A x AC
    cross(A, ab, PV);  // A x AB
    const double nn = nr[0] * N[0] + nr[1] * N[1] + nr[2] * N[2];
    const double NN = N[0] * N[0] + N[1] * N[1] + N[2] * N[2];
    const double Em = fmax(fmax(fmax(fabs(ab[0]), fabs(ab[1])), fabs(ab[2])), fmax(fmax(fabs(ac[0]), fabs(ac[1])), fabs(ac[2])));
    const double At = fmax(fmax(fabs(A[0]), fabs(A[1])), fabs(A[2]));
    const double nm = fmax(fmax(fabs(nr[0]), fabs(nr[1])), fabs(nr[2]));
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    const double Db = (double)QUICK_DBOUND;
    double thr = -inf, f = 1.0;
    // written so that NaN or inf anywhere leaves thr = -inf (the pair always goes to phase 2)
    if (NN > 0.0 && NN < 1e40 && Em < 1e10 && At < 1e10 && nm < 1e10 && (double)obound < 1e10) {
        const double lam = nn / NN;
        const double e0 = nr[0] - lam * N[0], e1 = nr[1] - lam * N[1], e2 = nr[2] - lam * N[2];
        const double em = fmax(fmax(fabs(e0), fabs(e1)), fabs(e2)) * (1.0 + 1e-9) + 1e-12 * nm;
        f = lam < 0.0 ? 1.0 : -1.0;
        const double mu = fmax(0x1p-17 * ((double)obound + At) * Db * Em, 0x1p-18 * Db * Em * Em) * 1.01 + 1e-18;
        const double M = 4.0 * mu;
        const double alam = fabs(lam);
        const double X = alam > 0.0 ? M + Db * (0x1p-19 * nm + 3.0 * em + 1e-30) / alam * 1.01 : inf;
        thr = (X + M < (double)QUICK_EPS) ? (double)QUICK_EPS - M : -X;
    }
    const float fs = (float)f;
    // slots 0-2 AC~ (x m), 3-5 PU~ (x d), 6-8 -AB~ (x m), 9-11 -PV~ (x d), 12-14 N~ (x d), 15 thr
    const float v[16] = {fs * g2.x,          fs * g2.y,          fs * g2.z,          fs * (float)PU[0],
                         fs * (float)PU[1],  fs * (float)PU[2],  -fs * g1.x,         -fs * g1.y,
                         -fs * g1.z,         -fs * (float)PV[0], -fs * (float)PV[1], -fs * (float)PV[2],
                         fs * (float)N[0],   fs * (float)N[1],   fs * (float)N[2],   __double2float_rd(thr)};
    float *qf = reinterpret_cast<float *>(q);
#pragma unroll
    for (int s = 0; s < 16; ++s)
        qf[path_q_slot(layout, i, s)] = v[s];
    if (layout == 2 && i == n - 1 && (n & 1)) { // the pair's missing triangle: a record that drops every ray
#pragma unroll
        for (int s = 0; s < 16; ++s)
            qf[path_q_slot(2, n, s)] = s == 15 ? __int_as_float(0x7f800000) : 0.f;
    }
}

// the ray's side of the quick test: direction and moment
struct QuickRay {
    f3 d, m;
};
__device__ __forceinline__ QuickRay quick_ray(const f3 &o, const f3 &d)
{
    QuickRay r;
    r.d = d;
    r.m = mk3(fmaf(o.y, d.z, -(o.z * d.y)), fmaf(o.z, d.x, -(o.x * d.z)), fmaf(o.x, d.y, -(o.y * d.x)));
    return r;
}

// the values of the quick test for one pair (see above); FMUL + FFMA chains, no attempt to match the reference's rounding
struct QuickVals {
    float U, V, det, sum, hi, M, thr;
};
__device__ __forceinline__ QuickVals quick_vals(const QuickRay &r, const float4 &r0, const float4 &r1, const float4 &r2,
                                                const float4 &r3)
{
    QuickVals v;
    v.U = fmaf(r1.y, r.d.z, fmaf(r1.x, r.d.y, fmaf(r0.w, r.d.x, fmaf(r0.z, r.m.z, fmaf(r0.y, r.m.y, r0.x * r.m.x)))));
    v.V = fmaf(r2.w, r.d.z, fmaf(r2.z, r.d.y, fmaf(r2.y, r.d.x, fmaf(r2.x, r.m.z, fmaf(r1.w, r.m.y, r1.z * r.m.x)))));
    v.det = fmaf(r3.z, r.d.z, fmaf(r3.y, r.d.y, r3.x * r.d.x));
    v.sum = v.U + v.V;
    v.thr = r3.w;
    v.M = QUICK_EPS - v.thr;
    v.hi = fmaf(v.det, 1.0000019073486328f, v.M);
    return v;
}

__device__ __forceinline__ bool tri_quick2(const QuickRay &r, const float4 &r0, const float4 &r1, const float4 &r2,
                                           const float4 &r3)
{
    const QuickVals v = quick_vals(r, r0, r1, r2, r3);
    bool p = !(v.U < -v.M);
    p = p && !(v.V < -v.M);
    p = p && !(v.sum > v.hi);
    p = p || !(v.det > v.M);
    p = p && !(v.det < v.thr);
    return p;
}

template <unsigned BIT>
__device__ __forceinline__ void tri_quick2_bit(const QuickRay &r, const float4 &r0, const float4 &r1, const float4 &r2,
                                               const float4 &r3, unsigned &m)
{
    const QuickVals v = quick_vals(r, r0, r1, r2, r3);
    const float nM = -v.M;
    asm("{\n\t.reg .pred p;\n\t"
        "setp.geu.f32 p, %1, %2;\n\t"        // !(U < -M)
        "setp.geu.and.f32 p, %3, %2, p;\n\t" // !(V < -M)
        "setp.leu.and.f32 p, %4, %5, p;\n\t" // !(U + V > hi)
        "setp.leu.or.f32 p, %6, %7, p;\n\t"  // or the sign of det is not known: !(det > M)
        "setp.geu.and.f32 p, %6, %8, p;\n\t" // !(det < thr)
        "@p or.b32 %0, %0, %9;\n\t}"
        : "+r"(m)
        : "f"(v.U), "f"(nM), "f"(v.V), "f"(v.sum), "f"(v.hi), "f"(v.det), "f"(v.M), "f"(v.thr), "n"(BIT));
}

// ---- MODE 3: the same quick test, two consecutive triangles at a time with packed FP32 (sm_100a FFMA2 / FMUL2 /
// FADD2: one issue slot for two FMAs).  Every half of a packed operation is the IEEE operation of the scalar code
// above, in the same order, so the values -- and with them the margins' derivation -- are those of MODE 1 bit for
// bit; only the issue count per pair changes (18 FMA-pipe slots -> 9).  The ray's scalars enter as the broadcast
// operand (SASS Rx.F32), the two triangles' constants come interleaved from the pair record (path_q_slot).
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi)
{
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpk2(f32x2 v, float &lo, float &hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c)
{
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}

struct QuickRay2 { // the ray's six scalars, each in both halves
    f32x2 dx, dy, dz, mx, my, mz;
};
__device__ __forceinline__ QuickRay2 quick_ray2(const QuickRay &r)
{
    QuickRay2 q;
    q.dx = pk2(r.d.x, r.d.x);
    q.dy = pk2(r.d.y, r.d.y);
    q.dz = pk2(r.d.z, r.d.z);
    q.mx = pk2(r.m.x, r.m.x);
    q.my = pk2(r.m.y, r.m.y);
    q.mz = pk2(r.m.z, r.m.z);
    return q;
}

// quick values of the two triangles of pair record P (8 float4); v0 = even triangle, v1 = odd triangle
__device__ __forceinline__ void quick_vals2(const QuickRay2 &r, const float4 *P, QuickVals &v0, QuickVals &v1)
{
    const ulonglong2 *Q = reinterpret_cast<const ulonglong2 *>(P);
    const ulonglong2 q0 = Q[0], q1 = Q[1], q2 = Q[2], q3 = Q[3], q4 = Q[4], q5 = Q[5], q6 = Q[6], q7 = Q[7];
    const f32x2 U = fma2(q2.y, r.dz, fma2(q2.x, r.dy, fma2(q1.y, r.dx, fma2(q1.x, r.mz, fma2(q0.y, r.my, mul2(q0.x, r.mx))))));
    const f32x2 V = fma2(q5.y, r.dz, fma2(q5.x, r.dy, fma2(q4.y, r.dx, fma2(q4.x, r.mz, fma2(q3.y, r.my, mul2(q3.x, r.mx))))));
    const f32x2 det = fma2(q7.x, r.dz, fma2(q6.y, r.dy, mul2(q6.x, r.dx)));
    const f32x2 sum = add2(U, V);
    const f32x2 M = sub2(pk2(QUICK_EPS, QUICK_EPS), q7.y);
    const f32x2 hi = fma2(det, pk2(1.0000019073486328f, 1.0000019073486328f), M);
    unpk2(U, v0.U, v1.U);
    unpk2(V, v0.V, v1.V);
    unpk2(det, v0.det, v1.det);
    unpk2(sum, v0.sum, v1.sum);
    unpk2(M, v0.M, v1.M);
    unpk2(hi, v0.hi, v1.hi);
    unpk2(q7.y, v0.thr, v1.thr);
}

// The same values for two pair records (four triangles), the FFMA2s ordered by RAY component: consecutive
// instructions then share their broadcast operand (operand-reuse cache), so each reads only the triangles' pair and
// the accumulator from the register file -- FFMA2 with three fresh 64-bit operands is bound by the register ports
// (tools/pipe_probe.cu: 1.26 FFMA2/clk/SM against 3.6 FFMA/clk/SM).
__device__ __forceinline__ void quick_vals2x2(const QuickRay2 &r, const float4 *P, QuickVals (&v)[4])
{
    const ulonglong2 *Qa = reinterpret_cast<const ulonglong2 *>(P), *Qb = reinterpret_cast<const ulonglong2 *>(P + QPAIR_F4);
    f32x2 Ua = mul2(Qa[0].x, r.mx), Ub = mul2(Qb[0].x, r.mx), Va = mul2(Qa[3].x, r.mx), Vb = mul2(Qb[3].x, r.mx);
    Ua = fma2(Qa[0].y, r.my, Ua), Ub = fma2(Qb[0].y, r.my, Ub), Va = fma2(Qa[3].y, r.my, Va), Vb = fma2(Qb[3].y, r.my, Vb);
    Ua = fma2(Qa[1].x, r.mz, Ua), Ub = fma2(Qb[1].x, r.mz, Ub), Va = fma2(Qa[4].x, r.mz, Va), Vb = fma2(Qb[4].x, r.mz, Vb);
    f32x2 Da = mul2(Qa[6].x, r.dx), Db = mul2(Qb[6].x, r.dx);
    Ua = fma2(Qa[1].y, r.dx, Ua), Ub = fma2(Qb[1].y, r.dx, Ub), Va = fma2(Qa[4].y, r.dx, Va), Vb = fma2(Qb[4].y, r.dx, Vb);
    Da = fma2(Qa[6].y, r.dy, Da), Db = fma2(Qb[6].y, r.dy, Db);
    Ua = fma2(Qa[2].x, r.dy, Ua), Ub = fma2(Qb[2].x, r.dy, Ub), Va = fma2(Qa[5].x, r.dy, Va), Vb = fma2(Qb[5].x, r.dy, Vb);
    Da = fma2(Qa[7].x, r.dz, Da), Db = fma2(Qb[7].x, r.dz, Db);
    Ua = fma2(Qa[2].y, r.dz, Ua), Ub = fma2(Qb[2].y, r.dz, Ub), Va = fma2(Qa[5].y, r.dz, Va), Vb = fma2(Qb[5].y, r.dz, Vb);
    const f32x2 eps = pk2(QUICK_EPS, QUICK_EPS), k = pk2(1.0000019073486328f, 1.0000019073486328f);
    const f32x2 ta = Qa[7].y, tb = Qb[7].y;
    const f32x2 Ma = sub2(eps, ta), Mb = sub2(eps, tb);
    const f32x2 Sa = add2(Ua, Va), Sb = add2(Ub, Vb);
    const f32x2 Ha = fma2(Da, k, Ma), Hb = fma2(Db, k, Mb);
    unpk2(Ua, v[0].U, v[1].U), unpk2(Ub, v[2].U, v[3].U);
    unpk2(Va, v[0].V, v[1].V), unpk2(Vb, v[2].V, v[3].V);
    unpk2(Da, v[0].det, v[1].det), unpk2(Db, v[2].det, v[3].det);
    unpk2(Sa, v[0].sum, v[1].sum), unpk2(Sb, v[2].sum, v[3].sum);
    unpk2(Ma, v[0].M, v[1].M), unpk2(Mb, v[2].M, v[3].M);
    unpk2(Ha, v[0].hi, v[1].hi), unpk2(Hb, v[2].hi, v[3].hi);
    unpk2(ta, v[0].thr, v[1].thr), unpk2(tb, v[2].thr, v[3].thr);
}

__device__ __forceinline__ bool quick_keep(const QuickVals &v)
{
    bool p = !(v.U < -v.M);
    p = p && !(v.V < -v.M);
    p = p && !(v.sum > v.hi);
    p = p || !(v.det > v.M);
    p = p && !(v.det < v.thr);
    return p;
}

template <unsigned BIT>
__device__ __forceinline__ void quick_keep_bit(const QuickVals &v, unsigned &m)
{
    const float nM = -v.M;
    asm("{\n\t.reg .pred p;\n\t"
        "setp.geu.f32 p, %1, %2;\n\t"        // !(U < -M)
        "setp.geu.and.f32 p, %3, %2, p;\n\t" // !(V < -M)
        "setp.leu.and.f32 p, %4, %5, p;\n\t" // !(U + V > hi)
        "setp.leu.or.f32 p, %6, %7, p;\n\t"  // or the sign of det is not known: !(det > M)
        "setp.geu.and.f32 p, %6, %8, p;\n\t" // !(det < thr)
        "@p or.b32 %0, %0, %9;\n\t}"
        : "+r"(m)
        : "f"(v.U), "f"(nM), "f"(v.V), "f"(v.sum), "f"(v.hi), "f"(v.det), "f"(v.M), "f"(v.thr), "n"(BIT));
}

// the same chain of compares, voted over the warp (lane = triangle): the ray's survivor mask
__device__ __forceinline__ unsigned quick_keep_ballot(const QuickVals &v)
{
    const float nM = -v.M;
    unsigned b;
    asm volatile("{\n\t.reg .pred p;\n\t"
                 "setp.geu.f32 p, %1, %2;\n\t"
                 "setp.geu.and.f32 p, %3, %2, p;\n\t"
                 "setp.leu.and.f32 p, %4, %5, p;\n\t"
                 "setp.leu.or.f32 p, %6, %7, p;\n\t"
                 "setp.geu.and.f32 p, %6, %8, p;\n\t"
                 "vote.sync.ballot.b32 %0, p, 0xffffffff;\n\t}"
                 : "=r"(b)
                 : "f"(v.U), "f"(nM), "f"(v.V), "f"(v.sum), "f"(v.hi), "f"(v.det), "f"(v.M), "f"(v.thr));
    return b;
}

// K-th .. (N-1)-th PAIR of a group of N consecutive pair records: bits 2K, 2K+1 of mg
template <int K, int N>
__device__ __forceinline__ void quick2_pairs(const QuickRay2 &r, const float4 *P, unsigned &mg)
{
    if constexpr (K < N) {
        QuickVals v0, v1;
        quick_vals2(r, P + QPAIR_F4 * K, v0, v1);
        quick_keep_bit<(1u << (2 * K))>(v0, mg);
        quick_keep_bit<(2u << (2 * K))>(v1, mg);
        quick2_pairs<K + 1, N>(r, P, mg);
    }
}

// K-th .. (N-1)-th pair of a group of N consecutive records: bit K of mg
template <int K, int N>
__device__ __forceinline__ void quick2_group(const QuickRay &r, const float4 *R, unsigned &mg)
{
    if constexpr (K < N) {
        tri_quick2_bit<(1u << K)>(r, R[4 * K], R[4 * K + 1], R[4 * K + 2], R[4 * K + 3], mg);
        quick2_group<K + 1, N>(r, R, mg);
    }
}

// a ray the quick records are valid for (see above); other rays send every pair to phase 2
__device__ __forceinline__ bool quick_ray_ok(const f3 &o, const f3 &d, float obound)
{
    const float dm = fmaxf(fmaxf(fabsf(d.x), fabsf(d.y)), fabsf(d.z));
    const float om = fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z));
    return dm <= QUICK_DBOUND && om <= obound; // false for NaN
}

// MODE 0: S = geo = geometry tile (3 float4 per triangle), quick test in the reference's arithmetic.
// MODE 1: S = quick-record tile (4 float4 per triangle), geo = geometry tile of the same triangles (phase 2).
// MODE 3: S = pair-record tile (QPAIR_F4 float4 per two triangles), packed quick test.
// MODE 2, 4: MODE 1, 3 + every pair also runs the exact test; an accepted pair the quick test dropped is counted (viol).
template <int MODE, int UNROLL = RTC_QUICK_UNROLL>
__device__ __forceinline__ void sweep_chunked(const float4 *S, const float4 *geo, float obound, int n, int base,
                                              bool active, const f3 &o, const f3 &d, float &best, int &bestId,
                                              unsigned &viol)
{
    const bool rayok = MODE == 0 || quick_ray_ok(o, d, obound);
    const QuickRay qr = quick_ray(o, d);
    const QuickRay2 qr2 = quick_ray2(qr);
    for (int c0 = 0; c0 < n; c0 += 32) {
        const int cn = min(32, n - c0);
        unsigned m = 0;
        if (active) {
            if (!rayok) {
                m = cn == 32 ? 0xffffffffu : (1u << cn) - 1u;
            } else if (MODE == 0) {
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
            } else if (mode_packed(MODE)) {
                constexpr int UP = UNROLL >= 2 ? UNROLL / 2 : 1; // pair records per iteration
                const float4 *P = S + QPAIR_F4 * (c0 >> 1);
                const int np = (cn + 1) >> 1;
                int j = 0;
#pragma unroll 1
                for (; j + UP <= np; j += UP) {
                    unsigned mg = 0;
                    quick2_pairs<0, UP>(qr2, P + QPAIR_F4 * j, mg);
                    m |= mg << (2 * j);
                }
#pragma unroll 1
                for (; j < np; ++j) {
                    unsigned m1 = 0;
                    quick2_pairs<0, 1>(qr2, P + QPAIR_F4 * j, m1);
                    m |= m1 << (2 * j);
                }
                if (cn < 32)
                    m &= (1u << cn) - 1u; // an odd count's last pair is completed by a record that never survives
            } else {
                constexpr int UN = UNROLL < 4 ? UNROLL : 4;
                int j = 0;
                for (; j + UN <= cn; j += UN) {
                    unsigned mg = 0;
                    quick2_group<0, UN>(qr, S + 4 * (c0 + j), mg);
                    m |= mg << j;
                }
                for (; j < cn; ++j) {
                    unsigned m1 = 0;
                    quick2_group<0, 1>(qr, S + 4 * (c0 + j), m1);
                    m |= m1 << j;
                }
            }
        }
        if (mode_verify(MODE) && active) {
            for (int j = 0; j < cn; ++j) {
                const int i = c0 + j;
                unsigned ex[6] = {0, 0, 0, 0, 0, 0};
                float b_ = 999999.0f;
                int id_ = -1;
                tri_test_exact<true>(o, d, geo[3 * i], geo[3 * i + 1], geo[3 * i + 2], base + i, b_, id_, ex);
                if (ex[5] && !((m >> j) & 1u))
                    viol++;
            }
        }
        while (m) {
            const int j = __ffs(m) - 1;
            m &= m - 1;
            unsigned dummy[1];
            if (MODE == 0) {
                tri_test_exact<false>(o, d, S[3 * (c0 + j)], S[3 * (c0 + j) + 1], S[3 * (c0 + j) + 2], base + c0 + j, best,
                                      bestId, dummy);
            } else {
                const int i = c0 + j;
                tri_test_exact<false>(o, d, geo[3 * i], geo[3 * i + 1], geo[3 * i + 2], base + i, best, bestId, dummy);
            }
        }
    }
}

// Cooperative sweep for the tail of the work list: when a warp has few live lanes left (and no work to fetch),
// the 32 lanes form na groups of S lanes; group q sweeps the ray of the q-th live lane, member r testing the
// triangles j = r, r + S, ... of the tile (quick test, then the exact sequence), and the group's (distance, index)
// minimum -- lowest index among equal distances, which is what the reference's in-order strict-< scan keeps --
// goes back to the owner.  The serial chain of the last pixels (what bounds a frame split over many GPUs)
// shrinks by ~S.
struct CoopCtx {
    int S, r, leader; // group size, member index, lane holding this lane's group result after the reduction
    bool member;      // this lane's group has a ray
    f3 o, d;
    float best;
    int id;
};

template <int MODE>
__device__ __forceinline__ void coop_tile(CoopCtx &c, const float4 *S_, const float4 *geo, float obound, int n, int base)
{
    if (!c.member)
        return;
    unsigned dummy[1];
    if (MODE == 0) {
        for (int j = c.r; j < n; j += c.S)
            if (tri_quick(c.o, c.d, S_[3 * j], S_[3 * j + 1], S_[3 * j + 2]))
                tri_test_exact<false>(c.o, c.d, S_[3 * j], S_[3 * j + 1], S_[3 * j + 2], base + j, c.best, c.id, dummy);
    } else if (mode_packed(MODE)) { // member r takes the pairs r, r + S, ...
        const bool rayok = quick_ray_ok(c.o, c.d, obound);
        const QuickRay2 qr2 = quick_ray2(quick_ray(c.o, c.d));
        for (int q = c.r; 2 * q < n; q += c.S) {
            bool k0 = true, k1 = true;
            if (rayok) {
                QuickVals v0, v1;
                quick_vals2(qr2, S_ + QPAIR_F4 * q, v0, v1);
                k0 = quick_keep(v0);
                k1 = quick_keep(v1);
            }
            const int j = 2 * q;
            if (k0)
                tri_test_exact<false>(c.o, c.d, geo[3 * j], geo[3 * j + 1], geo[3 * j + 2], base + j, c.best, c.id, dummy);
            if (k1 && j + 1 < n)
                tri_test_exact<false>(c.o, c.d, geo[3 * j + 3], geo[3 * j + 4], geo[3 * j + 5], base + j + 1, c.best, c.id, dummy);
        }
    } else {
        constexpr int QS = mode_transposed(MODE) ? QTRI_F4 : 4;
        const bool rayok = quick_ray_ok(c.o, c.d, obound);
        const QuickRay qr = quick_ray(c.o, c.d);
        for (int j = c.r; j < n; j += c.S)
            if (!rayok || tri_quick2(qr, S_[QS * j], S_[QS * j + 1], S_[QS * j + 2], S_[QS * j + 3]))
                tri_test_exact<false>(c.o, c.d, geo[3 * j], geo[3 * j + 1], geo[3 * j + 2], base + j, c.best, c.id, dummy);
    }
}

// GROUPS: the instantiation for frames with fewer tasks than half the lanes (group mode, S > 1).  The host launches
// both instantiations; each reads the task count on the device and the one it is not meant for returns at once, so
// the common case (S = 1) runs a kernel without the group-mode state in its registers.
// (the group instantiation is capped at the registers of PATH_GROUP_CTAS CTAs per SM and launched with exactly that
// grid: its S is chosen for a fully resident grid)
template <bool STREAM, int MODE, bool GROUPS>
__global__ void __launch_bounds__(PATH_THREADS, GROUPS ? PATH_GROUP_CTAS : (MODE == 3 ? (STREAM ? 5 : RTC_PATH_MINB3) : RTC_PATH_MINB)) render_paths_kernel(const RenderArgs a, const PathWork w)
{
    {
        const unsigned np = w.list ? *w.list_count : (unsigned)(a.out_w * a.out_h);
        if ((unsigned long long)np * (unsigned)w.planes < (unsigned long long)w.min_tasks)
            return;
        const bool groups = w.coop && (unsigned long long)np * (unsigned)w.planes * 2ull <= (unsigned long long)w.group_lanes;
        if (groups != GROUPS)
            return;
    }
    extern __shared__ float4 sGeo[];
    // shared memory (per stage when streaming): [quick records (MODE >= 1), path_q_f4][geometry, 3 float4 per triangle]
    const int cap = STREAM ? PATH_TILE : a.sc.ntris;       // triangles a stage holds
    const int qoff = (int)path_q_f4(MODE, cap);             // float4 offset of the stage's geometry
    constexpr int STAGE_F4 = (int)path_q_f4(MODE, PATH_TILE) + 3 * PATH_TILE;
    __shared__ uint64_t bars[2];
    const int ntris = a.sc.ntris;
    const int nsph = a.trianglesOnly ? 0 : a.sc.nsph;
    const int lane = threadIdx.x & 31;
    const int tile = w.tile;
    const int ntiles = STREAM ? (ntris + tile - 1) / tile : 1;

    if (!STREAM) {
        for (int i = threadIdx.x; i < qoff; i += PATH_THREADS)
            sGeo[i] = w.qrec[i];
        for (int i = threadIdx.x; i < 3 * ntris; i += PATH_THREADS)
            sGeo[qoff + i] = a.sc.geo[i];
    } else if (threadIdx.x == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        fence_barrier_init();
    }
    __syncthreads();

    const unsigned npix = w.list ? *w.list_count : (unsigned)(a.out_w * a.out_h);
    const unsigned total = npix * (unsigned)w.planes;
    if (a.prog != nullptr && blockIdx.x == 0 && (int)threadIdx.x < a.prog_chunks && a.prog[threadIdx.x] == 0)
        *(volatile unsigned *)&a.prog[2 * a.prog_chunks + threadIdx.x] = 1u; // chunks without hit pixels are final already
    // Group mode: with fewer tasks than lanes (a frame split over many GPUs, a small image) S lanes share one
    // pixel.  They hold identical state (same task, same deterministic shading), each sweeps the triangles
    // j = r, r + S, ... and the group takes the (distance, index) minimum: the pixel's serial chain, which bounds
    // the kernel when every lane has at most one task, gets ~S times shorter in its sweep part.
    int S = 1;
    if (GROUPS)
        while (S < 32 && (unsigned long long)total * (unsigned)(2 * S) <= (unsigned long long)w.group_lanes)
            S *= 2;
    const int gr = lane & (S - 1), gbase = lane & ~(S - 1);
    // Sample-parallel group mode (pixels whose primary hit is known): the S lanes run S CONSECUTIVE samples of the
    // pixel at once.  A sample whose path has exactly one hit-bounce draws exactly 7 random numbers (6 for the
    // direction, 1 for the roulette, moremath.c:98-109, raytracing.c:274-285), so lane r starts sample b + r from
    // the batch's base state advanced by 7 r steps of the LCG (a jump: state * M_r + C_r).  When all lanes have
    // finished, the batch is validated in order: samples up to and including the first one with more than one
    // hit-bounce started from the right state; their contributions are added in sample order (the reference's
    // float sum), the base state becomes that lane's final state, and everything after it is recomputed in the next
    // batch.  Nothing speculative is ever written.  Two failed batches in a row (a closed scene) switch the pixel
    // back to one sample at a time.
    const bool spec = GROUPS && S > 1 && w.rec != nullptr && w.spec != 0;
    uint32_t jm = 1u, jc = 0u; // 7 * gr steps of the LCG as one multiply-add
    if (spec)
        for (int k = 0; k < 7 * gr; ++k) {
            jc = jc * 747796405u + 2891336453u;
            jm *= 747796405u;
        }
    uint32_t bstate = 0;  // RNG state at the start of sample bsample (exact)
    int bsample = 0;      // first sample of the current batch
    bool parked = false;  // this lane's sample of the batch is finished (or it has none)
    bool has_sample = false, spec_on = false;
    int hcount = 0, fails = 0;
    unsigned sample_rays = 0;
    f3 contrib = mk3(0.f, 0.f, 0.f);
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
    unsigned viol = 0; // MODE 2, 4: accepted pairs the quick test dropped (must stay 0)
    uint32_t k0 = 0; // tiles consumed so far (STREAM): stage = k & 1, parity = (k >> 1) & 1

    auto issue = [&](uint32_t k, int t) {
        const int n = min(tile, ntris - t * tile);
        // (packed records: the host keeps the tile even, so a tile starts on a pair)
        const uint32_t bq = (uint32_t)path_q_f4(MODE, n) * (uint32_t)sizeof(float4), bg = (uint32_t)n * 3u * (uint32_t)sizeof(float4);
        float4 *stage = sGeo + (size_t)(k & 1) * STAGE_F4;
        fence_proxy_async();
        mbar_arrive_expect_tx(&bars[k & 1], bq + bg);
        if (MODE != 0)
            tma_load_1d(stage, w.qrec + path_q_f4(MODE, t * tile), bq, &bars[k & 1]);
        tma_load_1d(stage + qoff, a.sc.geo + (size_t)t * tile * 3, bg, &bars[k & 1]);
    };

    // the pixel's last sample is accumulated: write it
    auto finish_pixel = [&]() {
        {
            active = false;
            if (a.accum != nullptr && gr == 0) {
                float *p = a.accum + ((size_t)plane * a.out_w * a.out_h + pix) * 3;
                p[0] = L.acc.x;
                p[1] = L.acc.y;
                p[2] = L.acc.z;
            }
            if (a.rgb != nullptr && a.nsplit == 1 && gr == 0) { // vec3ToColor, raytracing.c:11-15
                uint8_t *p = a.rgb + pixel_offset(a, pix);
                p[0] = (uint8_t)float_to_uint8(L.acc.x);
                p[1] = (uint8_t)float_to_uint8(L.acc.y);
                p[2] = (uint8_t)float_to_uint8(L.acc.z);
                if (a.prog != nullptr) { // the chunk's last finished pixel releases its rows to the copy engine
                    const int c = (pix / a.out_w) / a.prog_rows, K = a.prog_chunks;
                    __threadfence();
                    if (atomicAdd(&a.prog[K + c], 1u) + 1u == a.prog[c]) {
                        __threadfence_system();
                        *(volatile unsigned *)&a.prog[2 * K + c] = 1u;
                    }
                }
            }
        }
    };
    auto reset_path = [&]() {
        L.pos = a.origin;
        L.dir = pdir;
        L.tint = mk3(1.f, 1.f, 1.f);
        L.light = mk3(0.f, 0.f, 0.f);
        L.bounce = 0;
        fresh = true;
        pend = false;
    };
    // sample-parallel group mode: this lane's sample of the batch that starts at (bsample, bstate)
    auto start_batch = [&]() {
        has_sample = (spec_on || gr == 0) && bsample + gr < s_end;
        L.sample = bsample + gr;
        L.rng = bstate * jm + jc;
        reset_path();
        hcount = 0;
        sample_rays = 0;
        contrib = mk3(0.f, 0.f, 0.f);
        parked = !has_sample;
    };
    // a path has ended: accumulate it (main.c:99) and start the next sample, or finish the pixel
    auto finish_sample = [&]() {
        if (spec) { // kept aside until the batch is validated
            contrib = xscale3(L.light, wgt);
            parked = true;
            return;
        }
        L.acc = xadd3(L.acc, xscale3(L.light, wgt));
        L.sample++;
        if (L.sample >= s_end)
            finish_pixel();
        else
            reset_path();
    };

    while (true) {
        // ---- idle lanes fetch the next (pixel, plane) task ----
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
                        if (spec && active) {
                            bsample = L.sample;
                            bstate = L.rng;
                            spec_on = true;
                            fails = 0;
                            start_batch();
                        }
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
        // ---- sample-parallel group mode: a group whose lanes have all finished their sample validates the batch
        //      (before the shading pass, so the next batch's first hits are shaded in this round) ----
        if (spec) {
            const unsigned gmask = (S == 32 ? 0xffffffffu : ((1u << S) - 1u)) << gbase;
            const unsigned pk = __ballot_sync(0xffffffffu, !active || parked);
            const bool all = active && (pk & gmask) == gmask;
            if (__any_sync(0xffffffffu, all)) {
                const unsigned dev = __ballot_sync(0xffffffffu, all && has_sample && hcount != 1) & gmask;
                const int nb = min(spec_on ? S : 1, s_end - bsample);
                const int nvalid = dev ? min(__ffs(dev) - gbase, nb) : nb; // up to and including the first deviation
                for (int j = 0; j < S; ++j) {
                    const f3 c = mk3(__shfl_sync(0xffffffffu, contrib.x, gbase + j), __shfl_sync(0xffffffffu, contrib.y, gbase + j),
                                     __shfl_sync(0xffffffffu, contrib.z, gbase + j));
                    const unsigned r = __shfl_sync(0xffffffffu, sample_rays, gbase + j);
                    if (all && j < nvalid) {
                        L.acc = xadd3(L.acc, c); // main.c:99, in sample order, identically on every lane of the group
                        nrays += gr == 0 ? r : 0u;
                    }
                }
                const uint32_t endstate = __shfl_sync(0xffffffffu, L.rng, gbase + max(nvalid - 1, 0));
                if (all) {
                    bstate = endstate;
                    bsample += nvalid;
                    if (nb > 1) {
                        fails = nvalid == 1 ? fails + 1 : 0;
                        if (fails >= w.spec_fails)
                            spec_on = false;
                    } else if (!spec_on && w.spec_retry > 0) {
                        // one sample at a time: samples that end at their first hit again re-arm the batches
                        fails = dev ? 0 : fails - 1;
                        if (fails <= -w.spec_retry) {
                            spec_on = true;
                            fails = w.spec_fails - 1;
                        }
                    }
                    if (bsample >= s_end)
                        finish_pixel();
                    else
                        start_batch();
                }
            }
        }
        // ---- hit shading, ONE pass per round for both kinds of pending hits: the hit the last sweep found, and the
        //      known primary hit of a lane that starts a sample (the primary kernel sweeps that ray every sample).
        //      A lane whose path ends here (Russian roulette, bounce limit) and whose next sample again starts from
        //      a known hit sits out this round's sweep: repeating the pass for those few lanes would cost a full
        //      shading pass at a few percent lane utilisation each time. ----
        if (active && !parked && (pend || (has_rec && fresh))) {
            if (!pend) {
                best = best0;
                bestId = id0;
            }
            pend = false;
            fresh = false;
            hcount++;
            if (shade_hit(L, a, best, bestId))
                finish_sample();
        }
        // lanes with a ray to sweep this round
        const bool sweeping = active && !parked && !(has_rec && fresh);
        // a lane that finished its pixel in the pass above fetches at the top of the next round: the warp leaves
        // only when no lane has work AND the list is exhausted
        const bool live = active || !exhausted;
        const bool any = STREAM ? (__syncthreads_or(live) != 0) : __any_sync(0xffffffffu, live);
        if (!any)
            break;

        // ---- closest hit of the lane's current ray: spheres first, then triangles in order ----
        best = 999999.0f;
        bestId = -1;
        if (sweeping)
            for (int k = 0; k < nsph; ++k)
                sphere_test_exact(L.pos, L.dir, __ldg(&a.sc.s0[k]), k, best, bestId);
        // few live lanes and nothing left to fetch: sweep cooperatively (see CoopCtx)
        const unsigned act = __ballot_sync(0xffffffffu, sweeping);
        const int na = __popc(act);
        const bool coop = (S == 1 || spec) && na > 0 && na <= 16 && w.coop;
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
            if (sweeping && gi != -1 && gb < best) { // spheres were tested first by the owner: they win ties
                best = gb;
                bestId = gi;
            }
        };
        const bool grouped = S > 1 && !spec;
        if (grouped) { // the group's lanes hold the same ray; member gr sweeps every S-th triangle
            cc.S = S;
            cc.r = gr;
            cc.member = sweeping;
            cc.o = L.pos;
            cc.d = L.dir;
            cc.best = 999999.0f;
            cc.id = -1;
            cc.leader = lane;
        }
        if (!STREAM) {
            if (coop || grouped) {
                coop_tile<MODE>(cc, sGeo, sGeo + qoff, w.obound, ntris, 0);
                coop_finish();
            } else {
                sweep_chunked<MODE>(sGeo, sGeo + qoff, w.obound, ntris, 0, sweeping, L.pos, L.dir, best, bestId, viol);
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
                const float4 *stage = sGeo + (size_t)(k & 1) * STAGE_F4;
                if (coop || grouped)
                    coop_tile<MODE>(cc, stage, stage + qoff, w.obound, min(tile, ntris - t * tile), t * tile);
                else
                    sweep_chunked<MODE, 4>(stage, stage + qoff, w.obound, min(tile, ntris - t * tile), t * tile,
                                        sweeping, L.pos, L.dir, best, bestId, viol);
                __syncthreads(); // every warp is done with this stage
                if (threadIdx.x == 0 && t + 2 < ntiles)
                    issue(k + 2, t + 2);
            }
            k0 += (uint32_t)ntiles;
            if (coop || grouped)
                coop_finish();
        }

        // ---- shade (calcColor's loop body), advance the state machine ----
        if (sweeping) {
            if (spec)
                sample_rays++; // counted when the batch validates the sample
            else
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
        if (mode_verify(MODE) && viol)
            atomicAdd(&a.stats[ST_MISMATCH], (unsigned long long)viol);
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
