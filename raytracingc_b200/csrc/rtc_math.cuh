// rtc_math.cuh -- device-side arithmetic of the render loop.
//
// Two families of operations live here and must not be mixed up:
//
//  * "exact" helpers (xmul, xadd, xsub, xdot, xcross, xnormalize, ...): every
//    call is ONE IEEE binary32 round-to-nearest operation that nvcc can never
//    contract into an FMA (__fmul_rn / __fadd_rn are contraction barriers).
//    They reproduce, operation for operation and in the same order, what gcc
//    emits for the reference's moremath.c on x86-64 (mulss/addss, no FMA,
//    FLT_EVAL_METHOD 0).  Everything that decides a hit, moves a ray or feeds
//    the RNG goes through these.  Where the reference widens to double
//    (`1./x`, sqrt(), `2.*x`, smoothstep's tail, Box-Muller) the helper either
//    widens too or uses the float instruction that is provably identical
//    (double rounding is innocuous for +,-,*,/,sqrt when 53 >= 2*24+2).
//
//  * plain float expressions (a*b+c, fmaf): used only by the conservative
//    pre-filter in rtc_kernels.cuh, whose result is a *superset* decision and
//    never reaches the image.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "rtc_normal_fast.cuh"
#include "rtc_powf.cuh"

namespace rtc {

struct f3 {
    float x, y, z;
};

__device__ __forceinline__ f3 mk3(float x, float y, float z) { return f3{x, y, z}; }

// ---- exact scalar ops -------------------------------------------------------
__device__ __forceinline__ float xmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float xadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float xsub(float a, float b) { return __fsub_rn(a, b); }
// (float)(1./(double)a): correctly rounded float reciprocal (reference moremath.c:14, raytracing.c:197,287)
__device__ __forceinline__ float xrcp(float a) { return __frcp_rn(a); }
// (float)sqrt((double)a) == correctly rounded float sqrt (reference moremath.c:9, raytracing.c:173)
__device__ __forceinline__ float xsqrt(float a) { return __fsqrt_rn(a); }
__device__ __forceinline__ float xdiv(float a, float b) { return __fdiv_rn(a, b); }

// ---- exact vec3 ops, reference moremath.c:33-87 -----------------------------
__device__ __forceinline__ f3 xadd3(f3 a, f3 b) { return mk3(xadd(a.x, b.x), xadd(a.y, b.y), xadd(a.z, b.z)); }
__device__ __forceinline__ f3 xsub3(f3 a, f3 b) { return mk3(xsub(a.x, b.x), xsub(a.y, b.y), xsub(a.z, b.z)); }
__device__ __forceinline__ f3 xscale3(f3 a, float s) { return mk3(xmul(a.x, s), xmul(a.y, s), xmul(a.z, s)); }
__device__ __forceinline__ f3 xmul3(f3 a, f3 b) { return mk3(xmul(a.x, b.x), xmul(a.y, b.y), xmul(a.z, b.z)); }
// dot: (ax*bx + ay*by) + az*bz, left to right (moremath.c:34)
__device__ __forceinline__ float xdot(f3 a, f3 b)
{
    return xadd(xadd(xmul(a.x, b.x), xmul(a.y, b.y)), xmul(a.z, b.z));
}
// cross (moremath.c:44-47)
__device__ __forceinline__ f3 xcross(f3 u, f3 v)
{
    return mk3(xsub(xmul(u.y, v.z), xmul(u.z, v.y)), xsub(xmul(u.z, v.x), xmul(u.x, v.z)),
               xsub(xmul(u.x, v.y), xmul(u.y, v.x)));
}
// normalized (moremath.c:7-18): len = sqrtf(sum), inv = 1/len, three multiplies
__device__ __forceinline__ f3 xnormalize(f3 v)
{
    float len = xsqrt(xadd(xadd(xmul(v.x, v.x), xmul(v.y, v.y)), xmul(v.z, v.z)));
    float inv = xrcp(len);
    return xscale3(v, inv);
}
// reflect (moremath.c:80-83): dir - normal * (float)(2. * dot); the doubling is exact
__device__ __forceinline__ f3 xreflect(f3 d, f3 n) { return xsub3(d, xscale3(n, xmul(2.0f, xdot(d, n)))); }
// lerp (moremath.c:85-88): a*(1-t) + b*t
__device__ __forceinline__ f3 xlerp(f3 a, f3 b, float t) { return xadd3(xscale3(a, xsub(1.0f, t)), xscale3(b, t)); }
__device__ __forceinline__ float xclamp01(float x) { return x < 0.0f ? 0.0f : (x > 1.0f ? 1.0f : x); }
// smoothstep (moremath.c:50-54): float quotient, float clamp, x*x in float,
// (3. - 2.*x) and the final product in double, narrowed on return.
__device__ __forceinline__ float xsmoothstep(float lo, float hi, float x)
{
    x = xclamp01(xdiv(xsub(x, lo), xsub(hi, lo)));
    double t = __dsub_rn(3.0, __dmul_rn(2.0, (double)x));
    return (float)__dmul_rn((double)xmul(x, x), t);
}

// ---- RNG, reference moremath.c:90-109 ---------------------------------------
// RandomValue(): state = state*747796405 + 2891336453, PCG-RXS-M-XS output
// hash, then (float)(r / 4294967295.0).  The double division is replaced by
//     (float)( fma((double)r, 2^-32, (double)r) * 2^-32 )
// which is bit-identical for all 2^32 values of r (both are r*2^-32*(1+eps),
// 0 < eps << 2^-24 relative, strictly between r and the next float rounding
// boundary; exhaustively verified by tests/test_rng_exhaustive.py via
// tools/verify_rng_div.c).
__device__ __forceinline__ uint32_t rng_next(uint32_t &state)
{
    state = state * 747796405u + 2891336453u;
    uint32_t r = ((state >> ((state >> 28) + 4u)) ^ state) * 277803737u;
    return (r >> 22) ^ r;
}
__device__ __forceinline__ float rng_to_unit(uint32_t r)
{
    // (double)r, exactly, without the I2F.F64 conversion (5 issue clocks on the XU pipe): 2^52 + r has r in its low word
    const double q = __dsub_rn(__hiloint2double(0x43300000, (int)r), 4503599627370496.0);
    return (float)__dmul_rn(__fma_rn(q, 2.3283064365386963e-10, q), 2.3283064365386963e-10);
}
__device__ __forceinline__ float random_value(uint32_t &state) { return rng_to_unit(rng_next(state)); }

// RandomValueNormalDistrubtion (moremath.c:98-103): all in double, narrowed
// three times exactly where the reference's `float` locals narrow.
// The reference's own sequence with the device's double libm: the definition of the result, and the slow path below.
__device__ __forceinline__ float normal_from_draws_exact(float u1, float u2)
{
    float theta = (float)__dmul_rn(6.2831853, (double)u1); // 2 * 3.14159265, folded in double
#if defined(RTC_WHATIF_FAST_SHADE) || defined(RTC_WHATIF_FAST_NORMAL) // TIMING EXPERIMENT ONLY (wrong images)
    return sqrtf(-2.0f * __logf(u2)) * __cosf(theta);
#endif
    float rho = (float)__dsqrt_rn(__dmul_rn(-2.0, log((double)u2)));
    return (float)__dmul_rn((double)rho, cos((double)theta));
}

// ---- the same float, cheaper: rtc_normal_fast.cuh (tailored double approximations + a Ziv test; falls back to the above) ----
__device__ __forceinline__ float normal_from_draws(float u1, float u2)
{
#ifndef RTC_NORMAL_EXACT_ONLY
    float v;
    if (nf::normal_fast(u1, u2, v))
        return v;
#endif
    return normal_from_draws_exact(u1, u2);
}

__device__ __forceinline__ float random_normal(uint32_t &state)
{
    float u1 = random_value(state); // theta's draw comes first
    float u2 = random_value(state);
    return normal_from_draws(u1, u2);
}
// RandomDiretion (moremath.c:105-109): x, then y, then z, then normalized.
// The call unit of the FP64 island: state in, {direction, state bits} out, all in registers.
__device__ __noinline__ float normal_exact_call(float u1, float u2) { return normal_from_draws_exact(u1, u2); }

__device__ __noinline__ float4 random_direction(uint32_t state)
{
#if !defined(RTC_NORMALS_LOOP) && !defined(RTC_NORMAL_EXACT_ONLY)
    // the six draws first (a serial but cheap chain), then the three fast paths as independent instruction streams --
    // a deviate is a chain of ~45 dependent FP64 instructions, three of them interleave: 960 clocks for a lone warp
    // against 1660 looped, 470 against 497 issue clocks (profiles/r2_shade_probe.txt), C4 frame -1 % -- and ONE
    // out-of-line copy of the exact sequence for whichever of them declined
    float u[6];
#pragma unroll
    for (int i = 0; i < 6; ++i)
        u[i] = random_value(state);
    float x, y, z;
    const bool kx = nf::normal_fast(u[0], u[1], x), ky = nf::normal_fast(u[2], u[3], y), kz = nf::normal_fast(u[4], u[5], z);
    if (!(kx && ky && kz)) {
        if (!kx)
            x = normal_exact_call(u[0], u[1]);
        if (!ky)
            y = normal_exact_call(u[2], u[3]);
        if (!kz)
            z = normal_exact_call(u[4], u[5]);
    }
#else
    // one copy of the FP64 code executed three times (with the libm sequence inline, round 1, the smaller instruction
    // footprint beat the overlap of three inlined chains: C4 path kernel 4.15 -> 3.95 ms)
    float x = 0.f, y = 0.f, z = 0.f;
#ifdef RTC_NORMALS_UNROLLED
#pragma unroll
#else
#pragma unroll 1
#endif
    for (int i = 0; i < 3; ++i) {
        x = y;
        y = z;
        z = random_normal(state);
    }
#endif
    f3 v = xnormalize(mk3(x, y, z));
    return make_float4(v.x, v.y, v.z, __uint_as_float(state));
}

// ---- sky, reference raytracing.c:151-160 ------------------------------------
struct SkyParams {
    f3 sun, horizon, zenith, ground;
    float sunFocus, sunIntensity;
    int sun_skip_ok; // host: sunFocus >= 0 and sunIntensity finite (see environment_light)
};

// powf as the reference's libm computes it: rtc_powf.cuh restates glibc 2.39's algorithm (two small tables, ~20 FP64
// instructions, one narrowing) and is bit-identical to it on every argument tried (2 * 10^9 on the host, 10^6 sky
// directions on the device).  Round 1 evaluated exp2(y * log2(x)) with CUDA's double libm instead: ~3.7 times the
// issue slots, and one sky value in 4000 off by one ulp.  Zero, negative, infinite or NaN bases and zero, infinite
// or NaN exponents take pow() in double (IEEE special values; never on the sky's path, see environment_light).
__device__ __noinline__ float xpowf(float x, float y)
{
#if defined(RTC_WHATIF_FAST_SHADE) || defined(RTC_WHATIF_FAST_POW) // TIMING EXPERIMENT ONLY (wrong images)
    return __powf(x, y);
#endif
#ifdef RTC_POW_ROUND1 // A/B only: round 1's evaluation (correctly rounded, so NOT always glibc's float)
    if (x > 0.0f && x < __int_as_float(0x7f800000) && y != 0.0f && fabsf(y) < 1e30f)
        return (float)exp2((double)y * log2((double)x));
#elif !defined(RTC_POW_LIBM)
    const uint32_t ix = __float_as_uint(x);
    if (ix - 1u < 0x7f7fffffu && y != 0.0f && fabsf(y) < __int_as_float(0x7f800000)) { // 0 < x < inf, y finite and non-zero
        // (a subnormal x: scaled by 2^23 and the exponent taken back, modulo 2^32, as glibc does)
        const uint32_t ib = ix >= 0x00800000u ? ix : __float_as_uint(__fmul_rn(x, 0x1p23f)) - (23u << 23);
        return pw::powf_glibc_bits(ib, (double)y);
    }
#endif
    return (float)pow((double)x, (double)y);
}

// Two exact shortcuts around the powf calls (the most expensive instructions of a miss):
//  * powf(+0, .35) = +0 and powf(1, .35) = 1 exactly (smoothstep clamps, so below the horizon line and above
//    -dir.y = .74 the argument IS 0 or 1);
//  * when sunMask = 0 the sun term is sun * 0: for d <= 1, sunFocus >= 0 and a finite intensity the power is a
//    finite number in [0, 1] whatever its value, so the product is a zero with the intensity's sign;
//    powf(+0, focus > 0) = +0 likewise.
__device__ __forceinline__ f3 environment_light(f3 dir, const SkyParams &s)
{
    const float x = xsmoothstep(0.0f, 0.74f, -dir.y);
    const float t = x == 0.0f ? 0.0f : (x == 1.0f ? 1.0f : xpowf(x, 0.35f));
    f3 grad = xlerp(s.horizon, s.zenith, t);
    float d = xdot(dir, s.sun);
    d = (d > 0.0f) ? d : 0.0f; // fmax(0, d): NaN -> 0
    const bool masked = !(dir.y < 0.0f);
    float sv;
    if (masked && d <= 1.0f && s.sun_skip_ok) {
        sv = copysignf(0.0f, s.sunIntensity);
    } else {
        const float pw = (d == 0.0f && s.sunFocus > 0.0f) ? 0.0f : xpowf(d, s.sunFocus);
        sv = xmul(xmul(pw, s.sunIntensity), masked ? 0.0f : 1.0f);
    }
    float g2s = xsmoothstep(-0.01f, 0.0f, -dir.y);
    f3 base = xlerp(s.ground, grad, g2s);
    return mk3(xadd(base.x, sv), xadd(base.y, sv), xadd(base.z, sv));
}

// ---- quantise, reference moremath.c:25-30 ------------------------------------
__device__ __forceinline__ uint32_t float_to_uint8(float f)
{
    if (f < 0.0f)
        return 0u;
    if (f >= 1.0f)
        return 255u;
    if (f != f)
        return 0u; // NaN: x86 cvttss2si -> 0x80000000 -> low byte 0
    return __float2uint_rz(xmul(f, 255.0f));
}

} // namespace rtc
