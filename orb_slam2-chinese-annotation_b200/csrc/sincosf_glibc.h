// glibc >= 2.28 sincosf (sysdeps/ieee754/flt-32/s_sincosf.{c,h}, from ARM optimized-routines),
// restated for the range the ORB descriptor needs.  The reference calls cos(float)/sin(float)
// (src/ORBextractor.cc:124-125), which GCC folds into one sincosf call; its result depends on the
// host libm.  This restatement is bit-identical to glibc 2.39's sincosf for EVERY float in
// [0, 6.2832] (exhaustive sweep, tests/test_sincosf.py; both with and without FMA contraction of
// the double-precision polynomial, so the libm ifunc variant does not matter).
// Usable from host (tests) and device code.
#pragma once
#include <stdint.h>
#include <string.h>
#ifdef __CUDACC__
#define ORB_HD __host__ __device__ __forceinline__
#else
#define ORB_HD static inline
#include <math.h>
#endif

ORB_HD double orb_sc_mad(double a, double b, double c) {
#ifdef __CUDA_ARCH__
    return __fma_rn(a, b, c);
#else
    return fma(a, b, c);
#endif
}

// n & 1 swaps sin/cos; t selects the negated cosine polynomial (quadrants 2, 3).
ORB_HD void orb_sc_poly(double x, double x2, int t, int n, float* sinp, float* cosp) {
    const double sg = t ? -1.0 : 1.0;
    const double C0 = sg * 0x1p0, C1 = sg * -0x1.ffffffd0c621cp-2, C2 = sg * 0x1.55553e1068f19p-5,
                 C3 = sg * -0x1.6c087e89a359dp-10, C4 = sg * 0x1.99343027bf8c3p-16;
    const double S1 = -0x1.555545995a603p-3, S2 = 0x1.1107605230bc4p-7, S3 = -0x1.994eb3774cf24p-13;
    const double x4 = x2 * x2, x3 = x2 * x;
    const double c2 = orb_sc_mad(x2, C4, C3);
    const double s1 = orb_sc_mad(x2, S3, S2);
    const double c1 = orb_sc_mad(x2, C1, C0);
    const double x5 = x3 * x2, x6 = x4 * x2;
    const double s = orb_sc_mad(x3, S1, x);
    const double c = orb_sc_mad(x4, C2, c1);
    const float sv = (float)orb_sc_mad(x5, s1, s), cv = (float)orb_sc_mad(x6, c2, c);
    if (n & 1) { *cosp = sv; *sinp = cv; } else { *sinp = sv; *cosp = cv; }
}

// Valid for |y| < 120 (the descriptor's angles are in [0, 2*pi]).
ORB_HD void orb_sincosf(float y, float* sinp, float* cosp) {
    double x = (double)y;
    uint32_t u;
#ifdef __CUDA_ARCH__
    u = __float_as_uint(y);
#else
    memcpy(&u, &y, 4);
#endif
    const uint32_t top = (u >> 20) & 0x7ff;
    if (top < (0x3f490fdbu >> 20)) {  // |y| < pi/4
        if (top < (0x39800000u >> 20)) { *sinp = y; *cosp = 1.0f; return; }  // |y| < 2^-12
        orb_sc_poly(x, x * x, 0, 0, sinp, cosp);
    } else {
        const double r = x * 0x1.45F306DC9C883p+23;  // 2/pi * 2^24
        const int n = ((int32_t)r + 0x800000) >> 24;
        x = orb_sc_mad(-(double)n, 0x1.921FB54442D18p0, x);
        const double s = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
        orb_sc_poly(x * s, x * x, (n & 2) ? 1 : 0, n, sinp, cosp);
    }
}
