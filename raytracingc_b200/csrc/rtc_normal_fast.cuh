// rtc_normal_fast.cuh -- the fast path of the normal deviate (RandomValueNormalDistrubtion, moremath.c:98-103).
//
// The reference computes, with `float` locals and the double libm,
//     theta = (float)(6.2831853 * u1);  rho = (float)sqrt(-2 log(u2));  return (float)(rho * cos(theta));
// rtc_math.cuh's normal_from_draws_exact IS that sequence and defines the result.  It costs ~70 FP64 instructions and
// a dozen conversions per deviate, three deviates per hit.  This file computes the same two narrowings from tailored
// double approximations with a relative error below 2^-43 (term-by-term bounds in the comments below) and accepts
// them only when an error 30 times that cannot change the float (a Ziv test): rounding is monotonic, so when every double within the margin of the approximation rounds to
// the same float, so does the reference's own libm value (within 2^-52 of the truth, inside the margin too).  Else --
// about one deviate in 10^4, and for every input outside the ranges below -- the caller runs the exact sequence.
// The generator has only 2^32 states, so rtc_probe_normal_selfcheck compares the two for EVERY pair of draws the
// reference can ever make (tests/test_gpu_math.py: a slice; tools/normal_selfcheck.py: all of it,
// profiles/r2_normal_selfcheck.txt); tools/normal_fast_host.cpp checks the same source against glibc on the CPU.
//
// Second version.  The first one (short polynomials, but float<->double conversions wherever a float seed or a
// narrowing was needed: 19 F2F + FRND + F2I + I2F per deviate) was proven exact and measured NO faster than the libm
// sequence: FP64 arithmetic is cheap on this part (DFMA: 1.6 issue clocks per warp instruction) but conversions, FRND and
// MUFU share the XU pipe at 5 - 6.5 clocks each (profiles/r2_shade_probe.txt), which that version saturated.
// This one does every widening, narrowing, rounding test and integer conversion on the bit patterns (integer pipe):
// 2 MUFU + 1 F2F per deviate; its double constants sit in the constant bank (see K[] below).
//
// Host-compilable: tools/normal_fast_host.cpp defines RTC_NF_HOST and the few intrinsics used here.
#pragma once

#include <stdint.h>

#ifndef RTC_NF_HOST
#define RTC_NF_FN __device__ __forceinline__
#define RTC_NF_CONST static __constant__
namespace rtc {
namespace nf {
RTC_NF_FN double from_words(uint32_t hi, uint32_t lo) { return __hiloint2double((int)hi, (int)lo); }
RTC_NF_FN uint32_t hi_word(double d) { return (uint32_t)__double2hiint(d); }
RTC_NF_FN uint32_t lo_word(double d) { return (uint32_t)__double2loint(d); }
RTC_NF_FN uint32_t f_bits(float f) { return __float_as_uint(f); }
RTC_NF_FN float f_from_bits(uint32_t b) { return __uint_as_float(b); }
RTC_NF_FN double mul_rn(double a, double b) { return __dmul_rn(a, b); }
// one MUFU each, relative error <= 2^-22.4 including the rounding of the argument to float
RTC_NF_FN float rcp_seed(float a)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
    return r;
}
RTC_NF_FN float rsqrt_seed(float a)
{
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
    return r;
}
#else
namespace rtc {
namespace nf {
#endif

// The double constants live in the constant bank: an FP64 instruction takes a c[][] operand for free, while a literal
// costs two UMOVs each time it is used (a third of the first build's instructions).
RTC_NF_CONST double K[] = {
    /* 0 */ 1.0 / 17.0, 1.0 / 15.0, 1.0 / 13.0, 1.0 / 11.0, 1.0 / 9.0, 1.0 / 7.0, 1.0 / 5.0, 1.0 / 3.0,
    /* 8 */ 4503601774854144.0,      // 2^52 + 2^31
    /* 9 */ 1.9082149292705877e-10,  // ln 2, tail
    /* 10 */ 6.9314718036912382e-1,  // ln 2, head (32 bits)
    /* 11 */ 6755399441055744.0,     // 1.5 * 2^52
    /* 12 */ 1.5707963267948966, 0.31830988618379069,
    /* 14 */ 3.141592653589793,      // pi's double (0x400921FB54442D18)
    /* 15 */ 1.2246467991473532e-16, // pi - that double
    /* 16 */ -1.9572941063391263e-20, 8.2206352466243295e-18, -2.8114572543455206e-15, 7.6471637318198164e-13,
    /* 20 */ -1.6059043836821613e-10, 2.5052108385441720e-08, -2.7557319223985893e-06, 1.9841269841269841e-04,
    /* 24 */ -8.3333333333333332e-03, 1.6666666666666666e-01, // -1/21! ... 1/3!
    /* 26 */ 6.2831853,              // the reference's 2 * 3.14159265, folded in double
};

// distance (in ulps of the double) that an approximation must keep from a float rounding boundary to be accepted:
// 2^14 ulps is a relative 2^-39 .. 2^-38, more than 30 times the error bound of the values tested below.  (The
// result does not rest on that bound: every pair of draws the generator can make has been enumerated, see the top.)
constexpr int ZIV_ULPS = 1 << 14;

// (double)f for a positive normal float f, from its bits (exponent re-biased, mantissa shifted): exact
RTC_NF_FN double widen_bits(uint32_t b) { return from_words((b >> 3) + 0x38000000u, b << 29); }
// a float within 2^-23 of a positive double in the normal float range (truncated): for seeds only
RTC_NF_FN float narrow_trunc(double d) { return f_from_bits(((hi_word(d) - 0x38000000u) << 3) | (lo_word(d) >> 29)); }

// `out` = v rounded to float precision (still a double; |v| in the normal float range); true when every double within
// ZIV_ULPS of v rounds to that same float.  A float keeps the top 23 + 1 bits: the low 29 bits of the double decide,
// the boundary is at 0x10000000.  (An exact tie is never accepted, so rounding half up is all that is needed.)
RTC_NF_FN bool narrow_checked(double v, double &out)
{
    const uint32_t lo = lo_word(v), hi = hi_word(v);
    const int d = (int)(lo & 0x1FFFFFFFu) - 0x10000000;
    const uint32_t lo2 = lo + 0x10000000u;
    out = from_words(hi + (lo2 < lo ? 1u : 0u), lo2 & 0xE0000000u);
    return (d < 0 ? -d : d) > ZIV_ULPS;
}

// (double)(float)(6.2831853 * (double)u1), u1 = the normal float with bits b1, in [2^-32, 1]: the reference's theta, EXACTLY (round
// to nearest, ties to even, at bit 29 of the exactly rounded double product; theta >= 1.4e-9 is a normal float)
RTC_NF_FN double theta_as_double(uint32_t b1)
{
    const double t = mul_rn(K[26], widen_bits(b1));
    const uint32_t lo = lo_word(t), hi = hi_word(t);
    const uint32_t lo2 = lo + 0x0FFFFFFFu + ((lo >> 29) & 1u);
    return from_words(hi + (lo2 < lo ? 1u : 0u), lo2 & 0xE0000000u);
}

// sqrt(-2 ln x) for the float x with bits b2 in [2^-32, 1), relative error < 2^-44:
//   x = m 2^e, m in (sqrt(1/2), sqrt(2)];  ln m = 2 atanh(s), s = (m - 1) / (m + 1), |s| <= 0.1716: 8 terms, remainder 2^-50;
//   1 / (m + 1) from MUFU.RCP (2^-22.4) + one Newton step (2^-44.8, on a term that is at most half of the result);
//   e ln 2 in two pieces;  sqrt by MUFU.RSQ + two coupled Newton steps (2^-44, then 2^-52).
RTC_NF_FN double sqrt_m2log(uint32_t b2)
{
    int e = (int)(b2 >> 23) - 127;
    const uint32_t mant = b2 & 0x007fffffu;
    const uint32_t big = mant > 0x3504F3u ? 1u : 0u; // 1.mant > sqrt 2: halve it
    e += (int)big;
    const double m = from_words((0x3FF00000u - (big << 20)) | (mant >> 3), mant << 29);
    const double f = m - 1.0, a = m + 1.0; // both exact: m has 24 significant bits
    const float af = f_from_bits((0x3F800000u - (big << 23)) | mant) + 1.0f;
    double r = widen_bits(f_bits(rcp_seed(af)));
    r = fma(fma(-a, r, 1.0), r, r);
    const double s = f * r, z = s * s;
    double p = K[0]; // 1/17, 1/15 ... 1/3
    p = fma(p, z, K[1]);
    p = fma(p, z, K[2]);
    p = fma(p, z, K[3]);
    p = fma(p, z, K[4]);
    p = fma(p, z, K[5]);
    p = fma(p, z, K[6]);
    p = fma(p, z, K[7]);
    const double lm = fma(s * z, p, s); // atanh(s)
    // (double)e without a conversion: 2^52 + 2^31 + e has e in its low word
    const double ed = from_words(0x43300000u, 0x80000000u ^ (uint32_t)e) - K[8];
    // ln 2 = 6.93147180369123816490e-01 + 1.90821492927058770002e-10 (the classic 32-bit / tail split)
    double L = fma(ed, K[9], lm + lm);
    L = fma(ed, K[10], L);
    const double t = -2.0 * L; // in [2^-23.4, 44.4] for x in [2^-32, 1 - 2^-24]
    const double y = widen_bits(f_bits(rsqrt_seed(narrow_trunc(t))));
    double g = t * y, h = 0.5 * y;
    double c = fma(-g, h, 0.5);
    g = fma(g, c, g);
    h = fma(h, c, h);
    c = fma(-g, h, 0.5);
    return fma(g, c, g);
}

// cos(t) for a double t in [0, 6.3] that is a float value, relative error < 2^-48 (also next to the zeros of the
// cosine: t is a float, so |t - k pi/2| >= 2^-26):
//   cos t = sin(t + pi/2) = (-1)^n sin(r),  n = rint((t + pi/2) / pi) -- read from the low word of the sum with 1.5 * 2^52 --
//   r = t - (n - 1/2) pi in [-pi/2, pi/2]: the product inside an fma is exact, so the first fma rounds r once whatever
//   cancels, the second adds pi's tail;  Taylor to r^21: the next term is 2^-60 of the result.
RTC_NF_FN double cos_0_2pi(double t)
{
    const double MAGIC = K[11]; // 1.5 * 2^52: ulp 1
    const double v = fma(t + K[12], K[13], MAGIC);
    const double k = (v - MAGIC) - 0.5;
    double r = fma(-k, K[14], t); // pi's double
    r = fma(-k, K[15], r);        // pi - that double
    const double z = r * r;
    double q = K[16]; // -1/21!, 1/19!, ... 1/3!
    q = fma(q, z, K[17]);
    q = fma(q, z, K[18]);
    q = fma(q, z, K[19]);
    q = fma(q, z, K[20]);
    q = fma(q, z, K[21]);
    q = fma(q, z, K[22]);
    q = fma(q, z, K[23]);
    q = fma(q, z, K[24]);
    q = fma(q, z, K[25]);
    const double sn = fma(-(r * z), q, r);
    return from_words(hi_word(sn) ^ (lo_word(v) << 31), lo_word(sn)); // (-1)^n
}

// true when the fast path determined the reference's float (then `out` is it); u1, u2 are the two draws
RTC_NF_FN bool normal_fast(float u1, float u2, float &out)
{
    const uint32_t b1 = f_bits(u1), b2 = f_bits(u2);
    double rho, res;
    const bool ok1 = narrow_checked(sqrt_m2log(b2), rho);
    const bool ok2 = narrow_checked(rho * cos_0_2pi(theta_as_double(b1)), res);
    out = (float)res; // exact: res has float precision and |res| in [1e-12, 7]
    // u1 in [2^-32, 1], u2 in [2^-32, 1) -- every non-zero draw of the generator; anything else (0, 1 for u2, a
    // subnormal, which widen_bits does not handle, a NaN) is the exact sequence's business
    return ok1 && ok2 && (b1 - 0x2F800000u) <= 0x10000000u && (b2 - 0x2F800000u) < 0x10000000u;
}

} // namespace nf
} // namespace rtc
