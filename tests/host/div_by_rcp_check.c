/* Host check of the division the CUDA kernels use where a reciprocal is at hand
 * (csrc/tf_chan_math.cuh: div_by_rcp): with y = RN(1/b), q0 = RN(a*y), r = a - b*q0
 * exactly (one fused multiply-add), RN(q0 + r*y) must BE the IEEE quotient a / b for
 * every pair, or vol and d would not stay bit-identical to the reference's.
 * usage: div_by_rcp_check N   -> prints "bad <count> of <pairs>" */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static uint64_t s = 88172645463325252ull;
static inline uint64_t rnd(void) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; }
static inline double urand(void) { return (rnd() >> 11) * (1.0 / 9007199254740992.0); }

static inline double div_by_rcp(double a, double b, double y) {
    const double q0 = a * y;
    const double r = fma(-q0, b, a);
    return fma(r, y, q0);
}

int main(int argc, char** argv) {
    const long N = argc > 1 ? atol(argv[1]) : 10000000L;
    long bad = 0, n = 0;
    for (int mode = 0; mode < 3; ++mode) {
        for (long i = 0; i < N; ++i) {
            double a, b;
            if (mode == 0) b = (double)(float)(1.0 + 1000.0 * urand());          /* ds: float32-valued */
            else if (mode == 1) b = 2.0 * tan(0.01 + 1.5 * urand());              /* 2 tan(angle) */
            else {                                                                /* any significand, half of them nearly all ones */
                uint64_t m = (rnd() & 0xfffffffffffffull) | 0x3ff0000000000000ull;
                if (i & 1) m |= 0xffffffffff000ull;
                memcpy(&b, &m, 8);
            }
            a = urand() * pow(2.0, (double)((int)(rnd() % 80) - 40));
            if (i % 7 == 0) { uint64_t m = (rnd() & 0xfffffffffffffull) | 0x3ff0000000000000ull; memcpy(&a, &m, 8); }
            if (div_by_rcp(a, b, 1.0 / b) != a / b) ++bad;
            ++n;
        }
    }
    printf("bad %ld of %ld\n", bad, n);
    return bad != 0;
}
