// glibc_log.cuh -- glibc 2.39's double-precision log(), FMA variant, operation for operation.
//
// The reference's partition scores are log2(x) := log(x) * (1.0 / log(2.0)) with libm's `log`
// (metacells/extensions.h:131-138; used by partitions.cpp:10-18,46,77,323-330,827-915).  glibc's log is
// not correctly rounded, so a bit-exact optimiser must evaluate the same polynomial and table in the same
// order: this file restates `__log_fma` (the ifunc variant every AVX2+FMA host selects) from its
// instruction sequence in the installed libm.so.6; each multiply-add below is one vfmadd there.  The
// constants come from libm's `__log_data` (scripts/extract_glibc_log_data.py -> glibc_log_data.inc).
// Compiles for the device (round-to-nearest intrinsics, so nvcc cannot re-associate or contract) and,
// for the CPU pin test, as plain C++ with -mfma -ffp-contract=off (tests/test_glibc_log.py).
#pragma once
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define GLIBC_LOG_HD __host__ __device__ __forceinline__
#define GLIBC_LOG_QUALIFIER static __device__ const
#else
#define GLIBC_LOG_HD static inline
#define GLIBC_LOG_QUALIFIER static const
#endif
#include "glibc_log_data.inc"

namespace glibc_log {

#if defined(__CUDA_ARCH__)
GLIBC_LOG_HD double mul(double a, double b) { return __dmul_rn(a, b); }
GLIBC_LOG_HD double add(double a, double b) { return __dadd_rn(a, b); }
GLIBC_LOG_HD double sub(double a, double b) { return __dsub_rn(a, b); }
GLIBC_LOG_HD double fma(double a, double b, double c) { return __fma_rn(a, b, c); }
GLIBC_LOG_HD double from_bits(unsigned long long bits) { return __longlong_as_double((long long)bits); }
GLIBC_LOG_HD unsigned long long to_bits(double value) { return (unsigned long long)__double_as_longlong(value); }
#else
GLIBC_LOG_HD double mul(double a, double b) { return a * b; }
GLIBC_LOG_HD double add(double a, double b) { return a + b; }
GLIBC_LOG_HD double sub(double a, double b) { return a - b; }
GLIBC_LOG_HD double fma(double a, double b, double c) { return __builtin_fma(a, b, c); }
GLIBC_LOG_HD double from_bits(unsigned long long bits) { double value; memcpy(&value, &bits, 8); return value; }
GLIBC_LOG_HD unsigned long long to_bits(double value) { unsigned long long bits; memcpy(&bits, &value, 8); return bits; }
#endif

// `tab` = glibc_log_tab (or a shared-memory copy of it): 128 x {invc, logc}; `poly` / `poly1` likewise.
GLIBC_LOG_HD double log(double x, const unsigned long long* tab, const unsigned long long* poly,
                        const unsigned long long* poly1) {
    unsigned long long ix = to_bits(x);
    // |x - 1| small: 1 - 0x1p-4 <= x < 1 + 0x1.09p-4  (ix - LO < HI - LO)
    if (ix - 0x3fee000000000000ULL < 0x3090000000000ULL) {
        if (ix == 0x3ff0000000000000ULL) return 0.0;
        const double r = sub(x, 1.0);
        const double r2 = mul(r, r);
        const double r3 = mul(r, r2);
        double p = fma(r, from_bits(poly1[2]), from_bits(poly1[1]));   // B1 + r B2
        double q = fma(r, from_bits(poly1[5]), from_bits(poly1[4]));   // B4 + r B5
        double s = fma(r, from_bits(poly1[8]), from_bits(poly1[7]));   // B7 + r B8
        p = fma(r2, from_bits(poly1[3]), p);                           // + r2 B3
        q = fma(r2, from_bits(poly1[6]), q);                           // + r2 B6
        s = fma(r2, from_bits(poly1[9]), s);                           // + r2 B9
        s = fma(r3, from_bits(poly1[10]), s);                          // + r3 B10
        s = fma(s, r3, q);
        s = fma(s, r3, p);                                             // the bracket; y = r3 * bracket
        const double two27 = 134217728.0;
        const double rw = fma(r, two27, r);                            // r + r * 0x1p27
        const double rhi = fma(-two27, r, rw);                         // (r + w) - w
        const double half = from_bits(poly1[0]);                       // B0 = -0.5
        const double rhi2 = mul(rhi, rhi);
        const double rlo = sub(r, rhi);
        const double hi = fma(rhi2, half, r);
        double lo = fma(rhi2, half, sub(r, hi));
        lo = fma(mul(half, rlo), add(r, rhi), lo);
        return add(hi, fma(s, r3, lo));
    }
    const unsigned top = (unsigned)(ix >> 48);
    if (top - 0x0010u >= 0x7ff0u - 0x0010u) {
        if ((ix << 1) == 0) return from_bits(0xfff0000000000000ULL);       // log(+-0) = -inf
        if (ix == 0x7ff0000000000000ULL) return x;                         // log(inf) = inf
        if ((top & 0x8000u) || (top & 0x7ff0u) == 0x7ff0u) return from_bits(0x7ff8000000000000ULL) ;  // x < 0 or nan
        ix = to_bits(mul(x, 4503599627370496.0));                          // subnormal: scale by 2^52
        ix -= 52ULL << 52;
    }
    const unsigned long long tmp = ix - 0x3fe6000000000000ULL;
    const int i = (int)((tmp >> 45) & 127);
    const int k = (int)((long long)tmp >> 52);
    const double z = from_bits(ix - (tmp & 0xfff0000000000000ULL));
    const double invc = from_bits(tab[2 * i]);
    const double logc = from_bits(tab[2 * i + 1]);
    const double kd = (double)k;
    const double w = fma(kd, from_bits(GLIBC_LOG_LN2HI), logc);
    const double r = fma(z, invc, -1.0);
    const double a12 = fma(r, from_bits(poly[2]), from_bits(poly[1]));   // A1 + r A2
    const double hi = add(r, w);
    const double r2 = mul(r, r);
    double lo = add(sub(w, hi), r);
    lo = fma(kd, from_bits(GLIBC_LOG_LN2LO), lo);
    const double r3 = mul(r, r2);
    const double a34 = fma(r, from_bits(poly[4]), from_bits(poly[3]));   // A3 + r A4
    lo = fma(r2, from_bits(poly[0]), lo);
    const double tail = fma(a34, r2, a12);
    return add(fma(r3, tail, lo), hi);
}

GLIBC_LOG_HD double log(double x) { return log(x, glibc_log_tab, glibc_log_poly, glibc_log_poly1); }

}  // namespace glibc_log
