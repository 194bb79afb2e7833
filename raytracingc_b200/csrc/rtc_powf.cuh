// rtc_powf.cuh -- powf(x, y) as the reference's C library computes it (getEnvironmentLight, raytracing.c:153,156).
//
// The reference calls libm's powf; the algorithm is not in /root/reference but in a third-party dependency:
//   GNU C Library 2.39 (Ubuntu GLIBC 2.39-0ubuntu8.5, the image's libm.so.6), sysdeps/ieee754/flt-32/e_powf.c --
//   the "optimized routines" powf (Szabolcs Nagy, 2017), selected on x86-64 CPUs with FMA as __powf_fma, i.e. compiled
//   with every a * b + c of the source contracted into one fused multiply-add.
// Published algorithm, restated here:
//   log2(x): x = 2^k z, z in [0x3f330000, 2 * that) as floats; a 16-entry table gives c near z with 1/c and log2(c);
//            r = z / c - 1 (|r| < 0.04), log2(x) = k + log2(c) + a degree-5 polynomial of r, all in double;
//   y log2(x) in double; outside (-150, 128) the result underflows to 0 / overflows to inf;
//   2^t: t = k'/32 + r' (the 2^52-shift rounding), 2^(i/32) from a 32-entry table, 2^r' a degree-3 polynomial;
//   the double product is narrowed to float ONCE.
// Its error before the narrowing is < 2^-58-ish, so it is correctly rounded almost always -- and WHEN it is not, the
// image contains glibc's float, not the correctly rounded one.  Round 1 evaluated exp2(y log2 x) with CUDA's double
// libm and matched glibc on all but 2.5e-4 of the sky directions; this restatement matches on all of them:
//   * tools/powf_host.cpp compiles THIS file for the host and compares it with the machine's powf: 2 * 10^9 random
//     cases (the sky's x in [0, 1] with y = .35 and large sun exponents, arguments near 1, random bit patterns incl.
//     overflow / underflow / subnormal results): 0 different; the 12 cases where the contracted and the un-contracted
//     form disagree all side with the contracted one (tests/test_host_math_tools.py runs 2 * 10^8 cases);
//   * tests/test_gpu_math.py compares the device's sky with oracle/_ref (the reference's own getEnvironmentLight
//     linked against glibc) on 10^6 directions: now bit for bit.
// The table values are libm's __powf_log2_data / __exp2f_data (numbers, read out of the image's libm.so.6; the 2^(i/32)
// column is the correctly rounded power minus i * 2^47, which tools/powf_host.cpp re-derives).  On a CPU without FMA
// glibc runs the un-contracted form, which differs from this one once in ~2 * 10^8 calls: the reference itself is
// machine-dependent there.
//
// Domain: x a positive finite float (subnormals through powf_glibc_bits), y finite and non-zero.  Everything else (zero,
// negative, infinite or NaN x; zero, infinite or NaN y) is the caller's slow path (pow() in double: IEEE special values).
#pragma once

#include <stdint.h>

#ifndef RTC_NF_HOST
#define RTC_PW_FN __device__ __forceinline__
#define RTC_PW_TABLE static __device__ const
namespace rtc {
namespace pw {
RTC_PW_FN double d_from_bits(uint64_t b) { return __longlong_as_double((long long)b); }
RTC_PW_FN uint64_t d_bits(double d) { return (uint64_t)__double_as_longlong(d); }
RTC_PW_FN uint32_t f_bits(float f) { return __float_as_uint(f); }
RTC_PW_FN double mul(double a, double b) { return __dmul_rn(a, b); } // never contracted with a neighbour
RTC_PW_FN double add(double a, double b) { return __dadd_rn(a, b); }
RTC_PW_FN double fmad(double a, double b, double c) { return __fma_rn(a, b, c); }
RTC_PW_FN double table(const double *t, uint32_t i) { return __ldg(t + i); } // a gather through L1, not the constant bank
RTC_PW_FN uint64_t table(const uint64_t *t, uint32_t i) { return (uint64_t)__ldg((const unsigned long long *)t + i); }
RTC_PW_FN float inf_f() { return __int_as_float(0x7f800000); }
#else
namespace rtc {
namespace pw {
#endif

// {1/c, log2(c)} for the 16 sub-intervals of [0x3f330000, 2 * 0x3f330000)
RTC_PW_TABLE double LOG2_TAB[32] = {
    0x1.661ec79f8f3bep+0, -0x1.efec65b963019p-2, 0x1.571ed4aaf883dp+0, -0x1.b0b6832d4fca4p-2,
    0x1.49539f0f010b0p+0, -0x1.7418b0a1fb77bp-2, 0x1.3c995b0b80385p+0, -0x1.39de91a6dcf7bp-2,
    0x1.30d190c8864a5p+0, -0x1.01d9bf3f2b631p-2, 0x1.25e227b0b8ea0p+0, -0x1.97c1d1b3b7af0p-3,
    0x1.1bb4a4a1a343fp+0, -0x1.2f9e393af3c9fp-3, 0x1.12358f08ae5bap+0, -0x1.960cbbf788d5cp-4,
    0x1.0953f419900a7p+0, -0x1.a6f9db6475fcep-5, 0x1.0000000000000p+0, 0x0.0p+0,
    0x1.e608cfd9a47acp-1, 0x1.338ca9f24f53dp-4,  0x1.ca4b31f026aa0p-1, 0x1.476a9543891bap-3,
    0x1.b2036576afce6p-1, 0x1.e840b4ac4e4d2p-3,  0x1.9c2d163a1aa2dp-1, 0x1.40645f0c6651cp-2,
    0x1.886e6037841edp-1, 0x1.88e9c2c1b9ff8p-2,  0x1.767dcf5534862p-1, 0x1.ce0a44eb17bccp-2,
};
// bit patterns of 2^(i/32) (correctly rounded doubles) minus i << 47, so that adding k' << 47 puts k' / 32 into the exponent
RTC_PW_TABLE uint64_t EXP2_TAB[32] = {
    0x3ff0000000000000ull, 0x3fefd9b0d3158574ull, 0x3fefb5586cf9890full, 0x3fef9301d0125b51ull,
    0x3fef72b83c7d517bull, 0x3fef54873168b9aaull, 0x3fef387a6e756238ull, 0x3fef1e9df51fdee1ull,
    0x3fef06fe0a31b715ull, 0x3feef1a7373aa9cbull, 0x3feedea64c123422ull, 0x3feece086061892dull,
    0x3feebfdad5362a27ull, 0x3feeb42b569d4f82ull, 0x3feeab07dd485429ull, 0x3feea47eb03a5585ull,
    0x3feea09e667f3bcdull, 0x3fee9f75e8ec5f74ull, 0x3feea11473eb0187ull, 0x3feea589994cce13ull,
    0x3feeace5422aa0dbull, 0x3feeb737b0cdc5e5ull, 0x3feec49182a3f090ull, 0x3feed503b23e255dull,
    0x3feee89f995ad3adull, 0x3feeff76f2fb5e47ull, 0x3fef199bdd85529cull, 0x3fef3720dcef9069ull,
    0x3fef5818dcfba487ull, 0x3fef7c97337b9b5full, 0x3fefa4afa2a490daull, 0x3fefd0765b6e4540ull,
};

// ix: the bits of a positive normal float x (for a subnormal x: the bits of x * 2^23, minus 23 << 23, modulo 2^32, as
// glibc does); y: finite, non-zero
RTC_PW_FN float powf_glibc_bits(uint32_t ix, double y)
{
    // x = 2^k z, z in [OFF, 2 OFF), OFF = 0x3f330000 (0.699...)
    const uint32_t tmp = ix - 0x3f330000u;
    const uint32_t i = (tmp >> 19) & 15u;
    const uint32_t top = tmp & 0xff800000u;
    const uint32_t iz = ix - top; // a normal float in [0.699, 1.398)
    const int k = (int)top >> 23;
    const double invc = table(LOG2_TAB, 2 * i), logc = table(LOG2_TAB, 2 * i + 1);
    // (double)z from its bits, (double)k by the 2^52 + 2^31 trick: both exact, no conversion instruction
    const double z = d_from_bits(((uint64_t)((iz >> 3) + 0x38000000u) << 32) | (uint64_t)(iz << 29));
    const double kd0 = add(d_from_bits(0x4330000000000000ull | (uint64_t)(0x80000000u ^ (uint32_t)k)), -4503601774854144.0);
    const double r = fmad(z, invc, -1.0);
    const double y0 = add(logc, kd0);
    const double r2 = mul(r, r);
    double p1 = fmad(0x1.27616c9496e0bp-2, r, -0x1.71969a075c67ap-2);
    const double p2 = fmad(0x1.ec70a6ca7baddp-2, r, -0x1.7154748bef6c8p-1);
    const double r4 = mul(r2, r2);
    double q = fmad(0x1.71547652ab82bp+0, r, y0);
    q = fmad(p2, r2, q);
    p1 = fmad(p1, r4, q); // log2(x)
    const double t = mul(y, p1);
    // |t| >= 126: overflow, underflow or still a float -- glibc's three cases
    if (((d_bits(t) >> 47) & 0xffffu) >= (0x405f800000000000ull >> 47)) {
        if (t > 0x1.fffffffd1d571p+6)
            return inf_f();
        if (t <= -150.0)
            return 0.0f;
    }
    const double SHIFT = 0x1.8p+47; // 1.5 * 2^52 / 32: the sum is rounded to a multiple of 1/32
    double kd = add(t, SHIFT);
    const uint64_t ki = d_bits(kd);
    kd = add(kd, -SHIFT);
    const double rr = add(t, -kd);
    const uint64_t sb = table(EXP2_TAB, (uint32_t)ki & 31u) + (ki << 47);
    const double s = d_from_bits(sb);
    const double zz = fmad(0x1.c6af84b912394p-5, rr, 0x1.ebfce50fac4f3p-3);
    const double rr2 = mul(rr, rr);
    double w = fmad(0x1.62e42ff0c52d6p-1, rr, 1.0);
    w = fmad(zz, rr2, w);
    return (float)mul(w, s); // the one narrowing (subnormal results included)
}

RTC_PW_FN float powf_glibc(float x, double y) { return powf_glibc_bits(f_bits(x), y); }

} // namespace pw
} // namespace rtc
