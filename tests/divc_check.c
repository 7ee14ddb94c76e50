/* divc_check.c - test infrastructure: is x / c reproduced bit for bit by the division-free sequence the kernels use?
 *
 *   q0 = x * y;  q1 = fma(fma(-q0, c, x), y, q0);  q = fma(fma(-q1, c, x), y, q1)      with y = RN(1 / c)
 *
 * (csrc/irs_common.cuh div_const / DIVC / div_small).  For every divisor on the command line the program feeds the
 * sequence adversarial numerators - values whose quotient lies within a few ulps of a rounding boundary (q = k + 1/2 ulp
 * neighbourhoods: x = RN(c * m) and its neighbours for random m, plus midpoints), random values across the exponent
 * range the kernels see, small integers and their neighbours - and compares with the hardware division.
 * Prints the number of mismatches per divisor; exit status 1 if any.
 *
 *   gcc -O2 -ffp-contract=off -o divc_check divc_check.c -lm && ./divc_check 2000000 3 5 40 100 120 135 1e6
 * DIVC_WEAK=1 replaces the sequence by the bare product x * y: the numerator set must then produce mismatches.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static uint64_t s_rng = 0x9E3779B97F4A7C15ull;
static uint64_t next_u64(void)
{
    uint64_t z = (s_rng += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static double from_bits(uint64_t b) { double d; memcpy(&d, &b, 8); return d; }
static uint64_t to_bits(double d) { uint64_t b; memcpy(&b, &d, 8); return b; }

static int s_weak = 0; /* self-test of the numerator set: the bare x * y must FAIL on it */
static double div_seq(double x, double c, double y)
{
    const double q0 = x * y;
    if (s_weak) return q0;
    const double q1 = fma(fma(-q0, c, x), y, q0);
    return fma(fma(-q1, c, x), y, q1);
}

static long check(double x, double c, double y)
{
    const double a = x / c, b = div_seq(x, c, y);
    if (to_bits(a) == to_bits(b)) return 0;
    if (a != a && b != b) return 0; /* both NaN */
    return 1;
}

int main(int argc, char** argv)
{
    if (argc < 3) { fprintf(stderr, "usage: %s n_per_class divisor...\n", argv[0]); return 2; }
    const long n = atol(argv[1]);
    s_weak = getenv("DIVC_WEAK") != NULL;
    long bad_total = 0;
    for (int a = 2; a < argc; ++a) {
        if (strncmp(argv[a], "rand:", 5) == 0) {
            /* run-time divisors (a sum of weights, a mean sent rate, an intent's threshold): y = 1 / c is one IEEE
             * division, every further numerator over the same c takes the sequence.  Random divisors over 70 binades,
             * among them the significands next to 1 and next to 2 (all ones), each with adversarial numerators. */
            const long nd = atol(argv[a] + 5);
            long bad = 0, tested = 0;
            for (long d = 0; d < nd; ++d) {
                uint64_t mant = next_u64() >> 12;
                const int kind = (int)(next_u64() % 8);
                if (kind == 0) mant = 0xFFFFFFFFFFFFFull - (next_u64() % 4);       /* 1.11...1x */
                else if (kind == 1) mant = next_u64() % 4;                           /* 1.00...0x */
                else if (kind == 2) mant &= ~((1ull << (int)(next_u64() % 52)) - 1); /* short significands */
                const int ex = (int)(next_u64() % 70) - 30;
                double c = ldexp(1.0 + (double)mant * 0x1p-52, ex);
                if (next_u64() & 1) c = (double)(int64_t)(1 + next_u64() % 100000); /* integers: counts, Mbit sizes */
                const double y = 1.0 / c;
                for (long i = 0; i < n; ++i) {
                    const int e2 = (int)(next_u64() % 60) - 30;
                    const double m = ldexp(1.0 + (double)(next_u64() >> 11) * 0x1p-53, e2);
                    uint64_t b = to_bits(m * c);
                    for (int k = -2; k <= 2; ++k) { bad += check(from_bits(b + (uint64_t)(int64_t)k), c, y); ++tested; }
                    const long double mid = (long double)m + (long double)ldexp(0x1p-53, e2);
                    b = to_bits((double)(mid * (long double)c));
                    for (int k = -2; k <= 2; ++k) { bad += check(from_bits(b + (uint64_t)(int64_t)k), c, y); ++tested; }
                    bad += check(ldexp(1.0 + (double)(next_u64() >> 11) * 0x1p-53, e2), c, y); ++tested;
                    bad += check((double)(int64_t)(next_u64() % 1000000), c, y); ++tested;
                    bad += check(-m, c, y); ++tested;
                }
                bad += check(0.0, c, y); ++tested;
            }
            printf("random divisors (%ld): %ld mismatches in %ld numerators\n", nd, bad, tested);
            bad_total += bad;
            continue;
        }
        const double c = atof(argv[a]), y = 1.0 / c;
        long bad = 0, tested = 0;
        for (long i = 0; i < n; ++i) {
            /* class 1: numerators next to exact multiples of c (quotients next to representable values) */
            const int ex = (int)(next_u64() % 80) - 40; /* magnitudes 2^-40 .. 2^40 */
            const double m = ldexp(1.0 + (double)(next_u64() >> 11) * 0x1p-53, ex);
            const double xm = m * c;
            uint64_t b = to_bits(xm);
            for (int d = -3; d <= 3; ++d) { bad += check(from_bits(b + (uint64_t)(int64_t)d), c, y); ++tested; }
            /* class 2: numerators whose quotient is next to a midpoint between two representable values */
            const long double mid = (long double)m + (long double)ldexp(0x1p-53, ex); /* m + half an ulp: exact in 64 bits */
            const double xh = (double)(mid * (long double)c);
            b = to_bits(xh);
            for (int d = -3; d <= 3; ++d) { bad += check(from_bits(b + (uint64_t)(int64_t)d), c, y); ++tested; }
            /* class 3: random bit patterns, positive and negative, exponents 2^-60 .. 2^60 */
            const uint64_t r = next_u64();
            const double xr = from_bits((r & 0x800FFFFFFFFFFFFFull) | ((uint64_t)(1023 - 60 + (int)((r >> 52) % 121)) << 52));
            bad += check(xr, c, y); ++tested;
            /* class 4: integers and counts (packets, RBs, UEs) times simple factors */
            const double xi = (double)(int64_t)(next_u64() % 100000000ull) * (double)(1 + (int)(next_u64() % 65536));
            bad += check(xi, c, y); ++tested;
            bad += check((double)(i % 100000), c, y); ++tested;
        }
        bad += check(0.0, c, y); ++tested;
        printf("divisor %.17g: %ld mismatches in %ld numerators\n", c, bad, tested);
        bad_total += bad;
    }
    return bad_total ? 1 : 0;
}
