/* Exhaustive check (all 2^32 inputs) that the division-free form used on the
 * GPU, (float)(fma((double)r, 2^-32, (double)r) * 2^-32), equals the reference's
 * (float)(r / 4294967295.0) (moremath.c:94).  Build: gcc -O2 -ffp-contract=off
 * tools/verify_rng_div.c -lm ; prints the number of mismatches (expect 0).
 * Optional args: first, count (to check a slice). */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
int main(int argc, char **argv)
{
    uint64_t first = argc > 1 ? strtoull(argv[1], 0, 0) : 0;
    uint64_t count = argc > 2 ? strtoull(argv[2], 0, 0) : (1ull << 32);
    uint64_t bad = 0;
    for (uint64_t i = first; i < first + count && i < (1ull << 32); ++i) {
        uint32_t r = (uint32_t)i;
        float a = (float)((double)r / 4294967295.0);
        double q = (double)r;
        float b = (float)(fma(q, 0x1p-32, q) * 0x1p-32);
        if (memcmp(&a, &b, 4)) {
            if (bad < 10)
                printf("mismatch r=%u ref=%a mine=%a\n", r, a, b);
            ++bad;
        }
    }
    printf("mismatches %llu\n", (unsigned long long)bad);
    return bad != 0;
}
