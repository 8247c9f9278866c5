// fdlibm __ieee754_log restated for device code (and host, for the C++ tool layer).
//
// The reference computes entropy with Math.log (util/InfoContentStat.java:80,83;
// util/AttributeSplitStat.java:162,166,388,391,466,469).  StrictMath.log is specified as fdlibm's
// e_log.c algorithm; evaluating that algorithm with one IEEE rounding per operation (this file is
// compiled with -fmad=false, and uses no FMA intrinsics) gives bit-identical results on the GPU,
// on the host and in the CPU oracle.
//
// The algorithm and its constants are those of fdlibm's e_log.c (a third-party source that is not part of the reference
// tree), whose notice is preserved here:
//   Copyright (C) 1993 by Sun Microsystems, Inc. All rights reserved.
//   Developed at SunSoft, a Sun Microsystems, Inc. business.
//   Permission to use, copy, modify, and distribute this software is freely granted, provided that this notice is preserved.
#pragma once
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define AVT_HD __host__ __device__ __forceinline__
#else
#define AVT_HD inline
#endif

namespace avt {

AVT_HD double bits_to_double(uint32_t hi, uint32_t lo) {
    uint64_t u = ((uint64_t)hi << 32) | lo;
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)u);
#else
    double d;
    memcpy(&d, &u, 8);
    return d;
#endif
}

AVT_HD void double_to_bits(double d, int32_t& hi, uint32_t& lo) {
#if defined(__CUDA_ARCH__)
    uint64_t u = (uint64_t)__double_as_longlong(d);
#else
    uint64_t u;
    memcpy(&u, &d, 8);
#endif
    hi = (int32_t)(u >> 32);
    lo = (uint32_t)u;
}

AVT_HD double strict_log(double x) {
    const double ln2_hi = 6.93147180369123816490e-01, ln2_lo = 1.90821492927058770002e-10,
                 two54 = 1.80143985094819840000e+16, Lg1 = 6.666666666666735130e-01,
                 Lg2 = 3.999999999940941908e-01, Lg3 = 2.857142874366239149e-01,
                 Lg4 = 2.222219843214978396e-01, Lg5 = 1.818357216161805012e-01,
                 Lg6 = 1.531383769920937332e-01, Lg7 = 1.479819860511658591e-01;
    int32_t hx, k = 0, i, j;
    uint32_t lx;
    double_to_bits(x, hx, lx);
    if (hx < 0x00100000) {
        if (((hx & 0x7fffffff) | lx) == 0) return -two54 / 0.0;
        if (hx < 0) return (x - x) / 0.0;
        k -= 54;
        x *= two54;
        double_to_bits(x, hx, lx);
    }
    if (hx >= 0x7ff00000) return x + x;
    k += (hx >> 20) - 1023;
    hx &= 0x000fffff;
    i = (hx + 0x95f64) & 0x100000;
    x = bits_to_double((uint32_t)(hx | (i ^ 0x3ff00000)), lx);
    k += (i >> 20);
    double f = x - 1.0;
    if ((0x000fffff & (2 + hx)) < 3) {
        if (f == 0.0) {
            if (k == 0) return 0.0;
            double dk = (double)k;
            return dk * ln2_hi + dk * ln2_lo;
        }
        double R = f * f * (0.5 - 0.33333333333333333 * f);
        if (k == 0) return f - R;
        double dk = (double)k;
        return dk * ln2_hi - ((R - dk * ln2_lo) - f);
    }
    double s = f / (2.0 + f);
    double dk = (double)k;
    double z = s * s;
    i = hx - 0x6147a;
    double w = z * z;
    j = 0x6b851 - hx;
    double t1 = w * (Lg2 + w * (Lg4 + w * Lg6));
    double t2 = z * (Lg1 + w * (Lg3 + w * (Lg5 + w * Lg7)));
    i |= j;
    double R = t2 + t1;
    if (i > 0) {
        double hfsq = 0.5 * f * f;
        if (k == 0) return f - (hfsq - s * (hfsq + R));
        return dk * ln2_hi - ((hfsq - (s * (hfsq + R) + dk * ln2_lo)) - f);
    }
    if (k == 0) return f - s * (f - R);
    return dk * ln2_hi - ((s * (f - R) - dk * ln2_lo) - f);
}

}  // namespace avt
