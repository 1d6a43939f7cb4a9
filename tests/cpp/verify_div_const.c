/* verify_div_const.c — host check of the division-by-window-size trick used by the cost filters
 * (stereo_matching_b200/csrc/cost_kernels.cu: div_win):
 *     q0 = RN(x * y), y = RN(1 / b);  r = fma(-b, q0, x);  q = copysign(fma(r, y, q0), x)
 * must equal RN(x / b) bit for bit.  usage: verify_div_const [stride]   (stride 1 = every finite float, ~50 s on 8 cores)
 * Prints the number of mismatches per divisor; exit status 0 iff there are none. */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

int main(int argc, char** argv) {
    const long long stride = argc > 1 ? atoll(argv[1]) : 1;
    const float divisors[5] = {1.0f, 2.0f, 3.0f, 4.0f, 5.0f};
    unsigned long long total_bad = 0;
    for (int t = 0; t < 5; ++t) {
        const float b = divisors[t], y = 1.0f / b;
        unsigned long long bad = 0;
#pragma omp parallel for reduction(+ : bad)
        for (long long i = 0; i < (1LL << 32); i += stride) {
            const uint32_t u = (uint32_t)i;
            float x;
            memcpy(&x, &u, 4);
            if (!isfinite(x)) continue;
            const float want = x / b;
            const float q0 = x * y;
            const float r = fmaf(-b, q0, x);
            const float q = copysignf(fmaf(r, y, q0), x);
            uint32_t a, c;
            memcpy(&a, &want, 4);
            memcpy(&c, &q, 4);
            bad += a != c;
        }
        printf("divisor %g: %llu mismatches (stride %lld)\n", b, bad, stride);
        total_bad += bad;
    }
    return total_bad ? 1 : 0;
}
