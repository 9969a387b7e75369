// hf2d_fmt.cuh -- "%g" with precision 6 (what std::ostream << double prints by default, i.e. what the reference's
// SaveData2D writes, deeps2d_core.cpp:2589-2673), correctly rounded, for device and host.
//
// x > 0 is scaled to y = x * 10^(5-e) in [1e5, 1e6) as an unevaluated sum p + pe:
//   |5-e| <= 22: 10^|5-e| is a double, so p + pe is EXACT (product split by FMA, or quotient + exact FMA remainder);
//                every tie of the six-digit rounding lives here (a value half-way between two six-digit decimals is
//                (2m+1) / (2^(n+1) 5^n) and binary only if 5^n divides 2m+1 < 2e6, i.e. n <= 9) and is resolved to even,
//                as glibc's exact printf does;
//   otherwise:   10^k = hi + lo from a table good to ~106 bits; no tie can occur, a wrong rounding would need the value
//                within ~1e-31 of a boundary.
#pragma once
#include <stdint.h>
#include <string.h>
#include <math.h>
#ifdef __CUDACC__            /* in the library: device code only; the same text compiled by g++ is the host check (tests/native) */
#define HF2D_FMT_HD __device__ __forceinline__
#define HF2D_FMT_CONST static __device__ const
#else
#define HF2D_FMT_HD static inline
#define HF2D_FMT_CONST static const
#endif

HF2D_FMT_CONST double P10_EXACT[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18,
                                       1e19, 1e20, 1e21, 1e22};
#include "hf2d_fmt_tab.inc"

// (p, pe) = (xh + xl) * 10^k, |k| <= 308
HF2D_FMT_HD void fmt_scale(double xh, double xl, int k, double& p, double& pe) {
    if (k >= 0 && k <= 22) {
        const double t = P10_EXACT[k];
        p = xh * t; pe = fma(xh, t, -p) + xl * t;
    } else if (k < 0 && k >= -22) {
        const double t = P10_EXACT[-k];
        p = xh / t; pe = (fma(-p, t, xh) + xl) / t;
    } else {
        const double h = k > 0 ? P10P_HI[k - 23] : P10N_HI[-k - 23], l = k > 0 ? P10P_LO[k - 23] : P10N_LO[-k - 23];
        p = xh * h; pe = fma(xh, h, -p) + (xh * l + xl * h);
    }
    const double s = p + pe;             // renormalise: |pe| <= ulp(p) / 2, which the rounding decision below relies on
    pe = pe - (s - p); p = s;
}

// writes the characters of printf("%g", x) to o (at most 13), returns their number; no terminator
HF2D_FMT_HD int fmt_g6(double x, char* o) {
    int n = 0;
    uint64_t b; memcpy(&b, &x, 8);
    const bool neg = (b >> 63) != 0;
    const uint64_t ab = b & 0x7fffffffffffffffull;
    if (neg) o[n++] = '-';
    if (ab > 0x7ff0000000000000ull) { o[n++] = 'n'; o[n++] = 'a'; o[n++] = 'n'; return n; }
    if (ab == 0x7ff0000000000000ull) { o[n++] = 'i'; o[n++] = 'n'; o[n++] = 'f'; return n; }
    if (ab == 0) { o[n++] = '0'; return n; }
    double ax; memcpy(&ax, &ab, 8);
    // decimal exponent: estimate from the binary one (never too high by more than one, never too low by more than one)
    int e2 = (int)(ab >> 52) - 1023;
    if ((ab >> 52) == 0) {                               // subnormal: position of the leading bit
        int lz = 0; uint64_t m_ = ab; while (!(m_ & (1ull << 51))) { m_ <<= 1; lz++; }
        e2 = -1023 - lz;
    }
    int e10 = (int)floor(e2 * 0.30102999566398120);
    double p = 0., pe = 0.;
    for (int it = 0; it < 4; it++) {
        const int k = 5 - e10;
        if (k > 308) { double q, qe; fmt_scale(ax, 0., 300, q, qe); fmt_scale(q, qe, k - 300, p, pe); }
        else fmt_scale(ax, 0., k, p, pe);
        if (p < 1e5 || (p == 1e5 && pe < 0.)) { e10--; continue; }
        if (p > 1e6 || (p == 1e6 && pe >= 0.)) { e10++; continue; }
        break;
    }
    // six significant digits, round half to even on the exact value t + f + pe
    const double t = floor(p), f = p - t;
    long m = (long)t;
    if (p >= 1e6) m = 999999 + 1;        // (p == 1e6 with pe < 0: just below 1e6 -> rounds up to it)
    else if (f > 0.5 || (f == 0.5 && (pe > 0. || (pe == 0. && (m & 1))))) m++;
    if (m >= 1000000) { m = 100000; e10++; }
    char d[6];
    for (int q = 5; q >= 0; q--) { d[q] = (char)('0' + m % 10); m /= 10; }
    int nd = 6; while (nd > 1 && d[nd - 1] == '0') nd--;         // significant digits left after stripping zeros
    const int X = e10;
    if (X < -4 || X >= 6) {
        o[n++] = d[0];
        if (nd > 1) { o[n++] = '.'; for (int q = 1; q < nd; q++) o[n++] = d[q]; }
        o[n++] = 'e'; o[n++] = X < 0 ? '-' : '+';
        int ax_ = X < 0 ? -X : X;
        if (ax_ >= 100) { o[n++] = (char)('0' + ax_ / 100); ax_ %= 100; }
        o[n++] = (char)('0' + ax_ / 10); o[n++] = (char)('0' + ax_ % 10);
    } else if (X >= 0) {
        for (int q = 0; q <= X; q++) o[n++] = d[q];              // (digits beyond nd are '0': d[] still holds them)
        if (nd > X + 1) { o[n++] = '.'; for (int q = X + 1; q < nd; q++) o[n++] = d[q]; }
    } else {
        o[n++] = '0'; o[n++] = '.';
        for (int q = 0; q < -X - 1; q++) o[n++] = '0';
        for (int q = 0; q < nd; q++) o[n++] = d[q];
    }
    return n;
}
