// CPU check of openhyperflow2d_b200/csrc/hf2d_fmt.cuh (fmt_g6 = printf("%g") / ostream << double) against glibc printf: random bit patterns,
// moderate magnitudes, ties of the six-digit rounding, powers of ten and their neighbours, specials.  Built and run by tests/test_fmt_g6.py.
#include "hf2d_fmt.cuh"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#ifndef NRAND
#define NRAND 2000000
#endif
static long bad = 0, total = 0;
static void chk(double x) {
    char a[32], b[32];
    snprintf(a, 32, "%g", x);
    int n = fmt_g6(x, b); b[n] = 0;
    total++;
    if (strcmp(a, b)) { if (bad++ < 20) printf("MISMATCH %a (%.17g): printf '%s' ours '%s'\n", x, x, a, b); }
}
int main() {
    srand48(1);
    // random bit patterns (all exponents)
    for (long i = 0; i < NRAND; i++) { uint64_t u = ((uint64_t)mrand48() << 32) ^ (uint64_t)(mrand48() & 0xffffffff); double x; memcpy(&x, &u, 8); chk(x); }
    // moderate magnitudes
    for (long i = 0; i < NRAND; i++) { double x = ldexp(drand48() + 0.5, (int)(drand48() * 120) - 60); chk(i & 1 ? -x : x); }
    // decimal-looking values, ties of the six-digit rounding, powers of ten and their neighbours
    for (int e = -30; e <= 30; e++) for (long m = 100000; m < 100400; m++) {
        double s = pow(10., e);
        chk(m * s); chk((m + 0.5) * s); chk((m * 10 + 5) * s * 0.1); chk(nextafter((m + 0.5) * s, 0)); chk(nextafter((m + 0.5) * s, 1e300));
    }
    for (long m = 1000000; m < 1000000 + NRAND / 10; m++) { chk(m + 0.5); chk((double)m * 10 + 5); chk((2 * m + 1) / 1024.0); chk((2 * m + 1) / 64.0); }
    for (int e = -323; e <= 308; e++) { double s = pow(10., e); chk(s); chk(nextafter(s, 0)); chk(nextafter(s, 1e309)); chk(9.999995 * s); chk(9.9999949 * s); chk(1.5 * s); }
    chk(0.0); chk(-0.0); chk(1.0 / 0.0); chk(-1.0 / 0.0); chk(0.0 / 0.0); chk(-(0.0 / 0.0)); chk(5e-324); chk(1.7976931348623157e308); chk(2.2250738585072014e-308);
    chk(123456.5); chk(1234565.0); chk(0.0001234565); chk(999999.5); chk(9999995.0); chk(0.00001); chk(100000.); chk(1e-5); chk(999999.4999999999);
    printf("%ld mismatches of %ld\n", bad, total);
    return bad != 0;
}
