// pt_detmath.h — deterministic single-precision elementary functions.
//
// The reference integrator calls sin, cos, tan, atan, atan2, asin, powf and logf
// (device.cu:25-26, disney_specular.cuh:69-74,177, disney_clearcoat.cuh:19,26,
// disney_helper.cuh:6, sample_methods.hpp:38-41).  CUDA's libm and glibc do not
// agree bit for bit, so paths would diverge after the first bounce and the GPU could
// only be checked statistically.  These versions evaluate in binary64 with nothing
// but +,-,*,/,fma,sqrt,floor,frexp,ldexp (all exactly specified by IEEE-754 / exact by
// construction; the Horner steps of the polynomials are explicit fma() calls — one correctly rounded
// operation on both sides, half the binary64 instructions of a separate multiply and add, which matter on
// the GPU: they were a quarter of k_shade's instruction stream) and round ONCE to binary32.  The same source therefore returns the
// same bits from nvcc (-fmad=false) and from g++ (-ffp-contract=off), and the result
// is within 1 ulp of the true value (the binary64 intermediate is good to ~1e-14
// relative for the argument ranges the integrator produces: |x| < 1e4 for the
// trigonometric functions).
//
// tests/test_detmath.py checks every function against numpy float64 on dense grids.
#pragma once
#include "pt_defs.h"

namespace pt {
namespace dmi {  // internals, binary64

constexpr double kPio2Hi = 1.57079632679489655800e+00;
constexpr double kPio2Lo = 6.12323399573676603587e-17;
constexpr double kTwoOverPi = 6.36619772367581382433e-01;
constexpr double kPiD = 3.14159265358979311600e+00;
constexpr double kPio4 = 7.85398163397448278999e-01;
constexpr double kLn2Hi = 6.93147180369123816490e-01;
constexpr double kLn2Lo = 1.90821492927058770002e-10;
constexpr double kInvLn2 = 1.44269504088896338700e+00;

PT_HD double sin_poly(double r) {  // |r| <= pi/4, odd Taylor series to r^17
    double r2 = r * r;
    double p = 1.0 / 355687428096000.0;
    p = fma(p, r2, -(1.0 / 1307674368000.0));
    p = fma(p, r2, 1.0 / 6227020800.0);
    p = fma(p, r2, -(1.0 / 39916800.0));
    p = fma(p, r2, 1.0 / 362880.0);
    p = fma(p, r2, -(1.0 / 5040.0));
    p = fma(p, r2, 1.0 / 120.0);
    p = fma(p, r2, -(1.0 / 6.0));
    return r + r * (r2 * p);
}
PT_HD double cos_poly(double r) {  // |r| <= pi/4, even Taylor series to r^16
    double r2 = r * r;
    double p = 1.0 / 20922789888000.0;
    p = fma(p, r2, -(1.0 / 87178291200.0));
    p = fma(p, r2, 1.0 / 479001600.0);
    p = fma(p, r2, -(1.0 / 3628800.0));
    p = fma(p, r2, 1.0 / 40320.0);
    p = fma(p, r2, -(1.0 / 720.0));
    p = fma(p, r2, 1.0 / 24.0);
    p = fma(p, r2, -(0.5));
    return 1.0 + r2 * p;
}
// x = k*pi/2 + r.  Returns the quadrant k mod 4.
PT_HD int reduce_pio2(double x, double& r) {
    double k = floor(x * kTwoOverPi + 0.5);
    r = (x - k * kPio2Hi) - k * kPio2Lo;
    return (int)((long long)k & 3);
}
PT_HD void sincos_d(double x, double& s, double& c) {
    double r;
    int q = reduce_pio2(x, r);
    double sr = sin_poly(r), cr = cos_poly(r);
    switch (q) {
        case 0: s = sr; c = cr; break;
        case 1: s = cr; c = -sr; break;
        case 2: s = -sr; c = -cr; break;
        default: s = -cr; c = sr; break;
    }
}
PT_HD double atan_core(double a) {  // a >= 0 (may be +inf); result in [0, pi/2]
    bool inv = a > 1.0;
    double t = inv ? 1.0 / a : a;
    bool shift = t > 0.41421356237309503;  // tan(pi/8)
    double u = shift ? (t - 1.0) / (t + 1.0) : t;
    double u2 = u * u;
    double p = 1.0 / 37.0;
    p = fma(-u2, p, 1.0 / 35.0);
    p = fma(-u2, p, 1.0 / 33.0);
    p = fma(-u2, p, 1.0 / 31.0);
    p = fma(-u2, p, 1.0 / 29.0);
    p = fma(-u2, p, 1.0 / 27.0);
    p = fma(-u2, p, 1.0 / 25.0);
    p = fma(-u2, p, 1.0 / 23.0);
    p = fma(-u2, p, 1.0 / 21.0);
    p = fma(-u2, p, 1.0 / 19.0);
    p = fma(-u2, p, 1.0 / 17.0);
    p = fma(-u2, p, 1.0 / 15.0);
    p = fma(-u2, p, 1.0 / 13.0);
    p = fma(-u2, p, 1.0 / 11.0);
    p = fma(-u2, p, 1.0 / 9.0);
    p = fma(-u2, p, 1.0 / 7.0);
    p = fma(-u2, p, 1.0 / 5.0);
    p = fma(-u2, p, 1.0 / 3.0);
    double r = u - u * (u2 * p);
    if (shift) r = kPio4 + r;
    if (inv) r = (kPio2Hi - r) + kPio2Lo;
    return r;
}
PT_HD double log_core(double x) {  // x > 0, finite, normal
    int e;
    double m = frexp(x, &e);  // [0.5, 1)
    if (m < 0.70710678118654752) {
        m = m * 2.0;
        e = e - 1;
    }
    double s = (m - 1.0) / (m + 1.0);
    double s2 = s * s;
    double p = 1.0 / 23.0;
    p = fma(p, s2, 1.0 / 21.0);
    p = fma(p, s2, 1.0 / 19.0);
    p = fma(p, s2, 1.0 / 17.0);
    p = fma(p, s2, 1.0 / 15.0);
    p = fma(p, s2, 1.0 / 13.0);
    p = fma(p, s2, 1.0 / 11.0);
    p = fma(p, s2, 1.0 / 9.0);
    p = fma(p, s2, 1.0 / 7.0);
    p = fma(p, s2, 1.0 / 5.0);
    p = fma(p, s2, 1.0 / 3.0);
    double lm = 2.0 * (s + s * (s2 * p));
    double ed = (double)e;
    return ed * kLn2Hi + (ed * kLn2Lo + lm);
}
PT_HD double exp_core(double z) {
    if (z > 200.0) return 1e90;  // rounds to +inf as binary32
    if (z < -200.0) return 0.0;
    double k = floor(z * kInvLn2 + 0.5);
    double r = (z - k * kLn2Hi) - k * kLn2Lo;
    double p = 1.0 / 6227020800.0;
    p = fma(p, r, 1.0 / 479001600.0);
    p = fma(p, r, 1.0 / 39916800.0);
    p = fma(p, r, 1.0 / 3628800.0);
    p = fma(p, r, 1.0 / 362880.0);
    p = fma(p, r, 1.0 / 40320.0);
    p = fma(p, r, 1.0 / 5040.0);
    p = fma(p, r, 1.0 / 720.0);
    p = fma(p, r, 1.0 / 120.0);
    p = fma(p, r, 1.0 / 24.0);
    p = fma(p, r, 1.0 / 6.0);
    p = fma(p, r, 0.5);
    p = fma(p, r, 1.0);
    p = fma(p, r, 1.0);
    return ldexp(p, (int)k);
}
PT_HD bool is_nan(double v) { return v != v; }
PT_HD bool is_inf(double v) { return v == HUGE_VAL || v == -HUGE_VAL; }
PT_HD float qnan() { return (float)(HUGE_VAL - HUGE_VAL); }
PT_HD double copysign_d(double mag, double sgn) {
    double a = mag < 0 ? -mag : mag;
    // signbit via division keeps -0.0 handling exact: 1/-0 = -inf
    bool neg = (sgn < 0) || (sgn == 0 && (1.0 / sgn) < 0);
    return neg ? -a : a;
}

}  // namespace dmi

// -DPT_LIBM_MATH=1 (measurement builds only, tools/build_variants.sh): the DEVICE code calls CUDA's float functions
// instead — what the deterministic design costs (frames then differ from the oracle in the last ulp of some draws
// and are compared by relMSE, not by bits).  The product is built without it.
#ifndef PT_LIBM_MATH
#define PT_LIBM_MATH 0
#endif
#if PT_LIBM_MATH && defined(__CUDA_ARCH__)
#define PT_LIBM(expr) return (expr)
#else
#define PT_LIBM(expr)
#endif

PT_HD float dm_sin(float x) {
    PT_LIBM(sinf(x));
    double xd = x;
    if (dmi::is_nan(xd) || dmi::is_inf(xd)) return dmi::qnan();
    double s, c;
    dmi::sincos_d(xd, s, c);
    return (float)s;
}
PT_HD float dm_cos(float x) {
    PT_LIBM(cosf(x));
    double xd = x;
    if (dmi::is_nan(xd) || dmi::is_inf(xd)) return dmi::qnan();
    double s, c;
    dmi::sincos_d(xd, s, c);
    return (float)c;
}
PT_HD float dm_tan(float x) {
    PT_LIBM(tanf(x));
    double xd = x;
    if (dmi::is_nan(xd) || dmi::is_inf(xd)) return dmi::qnan();
    double s, c;
    dmi::sincos_d(xd, s, c);
    return (float)(s / c);
}
PT_HD float dm_atan(float x) {
    PT_LIBM(atanf(x));
    double xd = x;
    if (dmi::is_nan(xd)) return dmi::qnan();
    double a = dmi::atan_core(xd < 0 ? -xd : xd);
    return (float)dmi::copysign_d(a, xd);
}
PT_HD float dm_atan2(float y, float x) {
    PT_LIBM(atan2f(y, x));
    double yd = y, xd = x;
    if (dmi::is_nan(yd) || dmi::is_nan(xd)) return dmi::qnan();
    bool xneg = (xd < 0) || (xd == 0 && (1.0 / xd) < 0);
    if (yd == 0) return (float)dmi::copysign_d(xneg ? dmi::kPiD : 0.0, yd);
    if (xd == 0) return (float)dmi::copysign_d(dmi::kPio2Hi, yd);
    double a;
    if (dmi::is_inf(xd) && dmi::is_inf(yd))
        a = dmi::kPio4;
    else {
        double q = yd / xd;
        a = dmi::atan_core(q < 0 ? -q : q);
    }
    if (xneg) a = (dmi::kPiD - a) + 2.0 * dmi::kPio2Lo;
    return (float)dmi::copysign_d(a, yd);
}
PT_HD float dm_asin(float x) {
    PT_LIBM(asinf(x));
    double xd = x;
    if (dmi::is_nan(xd) || xd > 1.0 || xd < -1.0) return dmi::qnan();
    double c = sqrt((1.0 - xd) * (1.0 + xd));
    double ax = xd < 0 ? -xd : xd;
    double a = (c == 0.0) ? (dmi::kPio2Hi + dmi::kPio2Lo) : dmi::atan_core(ax / c);
    return (float)dmi::copysign_d(a, xd);
}
PT_HD float dm_log(float x) {
    PT_LIBM(logf(x));
    double xd = x;
    if (dmi::is_nan(xd) || xd < 0) return dmi::qnan();
    if (xd == 0) return (float)(-HUGE_VAL);
    if (dmi::is_inf(xd)) return (float)HUGE_VAL;
    return (float)dmi::log_core(xd);
}
// x >= 0 is the only domain the integrator uses (colours, alpha^2); x < 0 -> NaN.
PT_HD float dm_pow(float x, float y) {
    PT_LIBM(powf(x, y));
    double xd = x, yd = y;
    if (yd == 0 || xd == 1.0) return 1.0f;
    if (dmi::is_nan(xd) || dmi::is_nan(yd) || xd < 0) return dmi::qnan();
    if (xd == 0) return yd > 0 ? 0.0f : (float)HUGE_VAL;
    if (dmi::is_inf(xd)) return yd > 0 ? (float)HUGE_VAL : 0.0f;
    if (dmi::is_inf(yd)) {
        bool small = xd < 1.0;
        return (small == (yd > 0)) ? 0.0f : (float)HUGE_VAL;
    }
    return (float)dmi::exp_core(yd * dmi::log_core(xd));
}

}  // namespace pt
