// normal_fast_host.cpp -- CPU check of csrc/rtc_normal_fast.cuh (the SAME source, compiled for the host) against the
// reference's own sequence with glibc's double libm (moremath.c:89-103), over generator states.
//
//   g++ -O2 -mfma -fopenmp -ffp-contract=off -o /tmp/normal_fast_host tools/normal_fast_host.cpp && /tmp/normal_fast_host [log2 count] [first]
//
// For every state s: u1, u2 = the next two draws; prints how many states the fast path accepted with a float that
// differs from the reference's (must be 0) and how many it declined.  The two MUFU seeds are emulated by the correctly
// rounded float perturbed by up to +-2 ulp (pseudo-randomly), which covers the hardware's documented error.
// The whole 2^32 takes a few minutes on 4 cores; the GPU self-check (tools/normal_selfcheck.py) covers all states
// against the DEVICE's exact sequence in seconds -- this file is the pre-flight for it and pins glibc as well.
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define RTC_NF_HOST 1
#define RTC_NF_FN static inline
#define RTC_NF_CONST static const
static uint32_t g_perturb = 0;
#pragma omp threadprivate(g_perturb)
namespace rtc {
namespace nf {
static inline double from_words(uint32_t hi, uint32_t lo)
{
    uint64_t x = ((uint64_t)hi << 32) | lo;
    double d;
    memcpy(&d, &x, 8);
    return d;
}
static inline uint32_t hi_word(double d)
{
    uint64_t x;
    memcpy(&x, &d, 8);
    return (uint32_t)(x >> 32);
}
static inline uint32_t lo_word(double d)
{
    uint64_t x;
    memcpy(&x, &d, 8);
    return (uint32_t)x;
}
static inline uint32_t f_bits(float f)
{
    uint32_t b;
    memcpy(&b, &f, 4);
    return b;
}
static inline float f_from_bits(uint32_t b)
{
    float f;
    memcpy(&f, &b, 4);
    return f;
}
static inline double mul_rn(double a, double b) { return a * b; }
static inline float wobble(float r)
{
    g_perturb = g_perturb * 1664525u + 1013904223u;
    const int k = (int)((g_perturb >> 16) % 5u) - 2; // -2 .. +2 ulp
    return f_from_bits(f_bits(r) + (uint32_t)k);
}
static inline float rcp_seed(float a) { return wobble((float)(1.0 / (double)a)); }
static inline float rsqrt_seed(float a) { return wobble((float)(1.0 / sqrt((double)a))); }
} // namespace nf
} // namespace rtc
#include "../raytracingc_b200/csrc/rtc_normal_fast.cuh"

static inline uint32_t rng_next(uint32_t &state)
{
    state = state * 747796405u + 2891336453u;
    uint32_t r = ((state >> ((state >> 28) + 4u)) ^ state) * 277803737u;
    return (r >> 22) ^ r;
}

int main(int argc, char **argv)
{
    const int lg = argc > 1 ? atoi(argv[1]) : 26;
    const uint64_t first = argc > 2 ? strtoull(argv[2], 0, 0) : 0;
    const uint64_t count = 1ull << lg;
    const uint64_t stride = lg >= 32 ? 1 : 2654435761ull | 1; // a slice is spread over the whole state space
    uint64_t bad = 0, declined = 0;
#pragma omp parallel for reduction(+ : bad, declined) schedule(static)
    for (uint64_t i = 0; i < count; ++i) {
        uint32_t s = (uint32_t)(first + i * stride);
        const uint32_t s0 = s;
        const float u1 = (float)(rng_next(s) / 4294967295.0);
        const float u2 = (float)(rng_next(s) / 4294967295.0);
        // the reference is C: log(), sqrt(), cos() take and return double (the casts keep C++ from picking float overloads)
        float theta = 2 * 3.14159265 * u1;              // moremath.c:99
        float rho = sqrt(-2 * log((double)u2));         // moremath.c:100
        float want = (double)rho * cos((double)theta);  // moremath.c:101 (narrowed on return)
        float got;
        if (!rtc::nf::normal_fast(u1, u2, got)) {
            ++declined;
            continue;
        }
        if (memcmp(&got, &want, 4)) {
            if (bad < 10)
                printf("state %08x: u1 %a u2 %a  fast %a  reference %a\n", s0, u1, u2, got, want);
            ++bad;
        }
    }
    printf("%llu states: fast path differs from the reference sequence (glibc) on %llu; declined %llu = %.3g\n",
           (unsigned long long)count, (unsigned long long)bad, (unsigned long long)declined, (double)declined / (double)count);
    return bad != 0;
}
