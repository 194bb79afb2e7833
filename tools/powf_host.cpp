// powf_host.cpp -- CPU check of csrc/rtc_powf.cuh (the SAME source, compiled for the host) against this machine's powf
// (glibc; the library the reference links, raytracing.c:153,156).
//
//   g++ -O2 -mfma -fopenmp -ffp-contract=off -o /tmp/powf_host tools/powf_host.cpp && /tmp/powf_host [cases]
//
// Four families of arguments, a quarter each: the sky's gradient (x in [0, 1), y = .35); sun exponents (x in [0, 1),
// y = .5 ... 1000); random bit patterns of x and y (overflow, underflow, subnormal results); x next to 1 with a large y.
// Prints the number of cases that differ (must be 0 on a CPU with FMA: see the header of rtc_powf.cuh) and checks the
// 2^(i/32) table against exp2() of this machine.  Exit status 1 when anything differs.
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <omp.h>

#define RTC_NF_HOST 1
#define RTC_PW_FN static inline
#define RTC_PW_TABLE static const
namespace rtc {
namespace pw {
static inline double d_from_bits(uint64_t b)
{
    double d;
    memcpy(&d, &b, 8);
    return d;
}
static inline uint64_t d_bits(double d)
{
    uint64_t b;
    memcpy(&b, &d, 8);
    return b;
}
static inline uint32_t f_bits(float f)
{
    uint32_t b;
    memcpy(&b, &f, 4);
    return b;
}
static inline double mul(double a, double b) { return a * b; } // -ffp-contract=off: never fused
static inline double add(double a, double b) { return a + b; }
static inline double fmad(double a, double b, double c) { return fma(a, b, c); }
static inline double table(const double *t, uint32_t i) { return t[i]; }
static inline uint64_t table(const uint64_t *t, uint32_t i) { return t[i]; }
static inline float inf_f() { return INFINITY; }
} // namespace pw
} // namespace rtc
#include "../raytracingc_b200/csrc/rtc_powf.cuh"

static inline float f_from_bits(uint32_t b)
{
    float f;
    memcpy(&f, &b, 4);
    return f;
}

int main(int argc, char **argv)
{
    const long n = argc > 1 ? atol(argv[1]) : 200000000;
    long bad = 0, tested = 0;
    for (int i = 0; i < 32; ++i) { // the table is the correctly rounded 2^(i/32), minus i << 47
        const uint64_t want = rtc::pw::d_bits(exp2(i / 32.0)) - ((uint64_t)i << 47);
        if (rtc::pw::EXP2_TAB[i] != want) {
            printf("EXP2_TAB[%d] = %#llx, exp2() gives %#llx\n", i, (unsigned long long)rtc::pw::EXP2_TAB[i], (unsigned long long)want);
            ++bad;
        }
    }
#pragma omp parallel reduction(+ : bad, tested)
    {
        uint64_t rs = 88172645463325252ull * (uint64_t)(omp_get_thread_num() * 2 + 1);
#pragma omp for schedule(static)
        for (long it = 0; it < n; ++it) {
            rs ^= rs << 13;
            rs ^= rs >> 7;
            rs ^= rs << 17;
            const uint64_t a = rs;
            float x, y;
            switch (it & 3) {
            case 0: x = (float)((a >> 8) & 0xffffff) * 0x1p-24f, y = 0.35f; break;
            case 1: x = (float)((a >> 8) & 0xffffff) * 0x1p-24f, y = (float)(1 + (a >> 40) % 2000) * 0.5f; break;
            case 2:
                x = f_from_bits(0x00800000u + (uint32_t)(a % (0x7f800000u - 0x00800000u)));
                y = f_from_bits((uint32_t)(a >> 32));
                if (!(fabsf(y) < 3e38f))
                    y = 3.7f;
                break;
            default: x = 1.0f - (float)((a >> 8) & 0xffff) * 0x1p-24f + (float)((a >> 50) & 7) * 0x1p-23f, y = (float)(1 + (a >> 40) % 3000); break;
            }
            if ((it & 1023) == 7)
                x = f_from_bits((uint32_t)(a >> 20) & 0x007fffffu); // a subnormal now and then
            if (!(x > 0.0f) || y == 0.0f) // the restatement's domain (the rest is pow()'s special values)
                continue;
            ++tested;
            const float want = powf(x, y);
            const float got = x >= 0x1p-126f ? rtc::pw::powf_glibc(x, (double)y)
                                             : rtc::pw::powf_glibc_bits(rtc::pw::f_bits(x * 0x1p23f) - (23u << 23), (double)y);
            if (memcmp(&want, &got, 4)) {
                if (bad < 10)
                    printf("x %a y %a: powf %a, restated %a\n", x, y, want, got);
                ++bad;
            }
        }
    }
    printf("%ld cases: %ld differ from this machine's powf\n", tested, bad);
    return bad != 0;
}
