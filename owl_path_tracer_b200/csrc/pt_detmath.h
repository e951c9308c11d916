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

// The binary64 routines are OUT OF LINE on the device.  Inlined, ptxas treated them as free arithmetic: k_shade
// evaluated the three pow() of the sheen tint (3 x log + exp, ~400 binary64 instructions) for EVERY shaded vertex, hoisted
// above the `sheen <= 0` test that skips them (ncu, round 2: 17 % of k_shade's instructions were binary64 on a scene
// without sheen).  A call is never speculated.
#if defined(__CUDACC__)
#define PT_HD_CALL __host__ __device__ __noinline__ inline
#else
#define PT_HD_CALL __attribute__((noinline)) inline
#endif

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

// ---------------------------------------------------------------------------- binary32 internals (build option)
// -DPT_TRIG_F32=1 evaluates sin / cos / tan / atan / atan2 / asin in binary32: Cody-Waite reduction and short minimax
// polynomials built from +, -, *, /, fmaf, sqrtf, floorf only — still the same bits from nvcc and g++, 1-3 ulp, the
// accuracy class of CUDA's own float functions.  NOT the default: measured on the headline frame (dragon 1080p, 1024 spp,
// 128x128 crop against the glibc-math oracle, profiles/r02f_trig_probe.jsonl) the binary64 route differs from glibc in
// 1.5 % of the pixels, per-pixel relMSE 3.8e-4 — inside north_star's 1e-3 — while the binary32 route differs in 15 %
// and reaches 4.3e-3 (a path that diverges by one ulp somewhere meets another firefly of this very noisy estimator).
// Speed: the same on the dragon, 3 % of a frame on the car (profiles/r02f_trig_ab.jsonl).  Parity wins.
// Coefficients: Chebyshev-node interpolation in 50-digit arithmetic (mpmath), rounded to binary32.
namespace dmf {
constexpr float kTwoOverPi = 0.636619747f;
constexpr float kPio2Hi = 1.57079637f;        // float(pi/2)
constexpr float kPio2Lo = -4.37113883e-8f;    // float(pi/2 - kPio2Hi)
constexpr float kPio2Lo2 = -1.7151245e-15f;   // what is left
constexpr float kPio4Hi = 0.785398185f, kPio4Lo = -2.18556941e-8f;
constexpr float kPiHi = 3.14159274f, kPiLo = -8.74227766e-8f;
constexpr float kTrigMax = 100.0f;            // beyond: the binary64 route (the integrator's arguments are below 7; for large
                                              // k the rounding of x - k pi/2 would cost accuracy next to the zeros)

PT_HD bool is_nan(float v) { return v != v; }
PT_HD bool is_inf(float v) { return v == (float)HUGE_VAL || v == -(float)HUGE_VAL; }
PT_HD bool sign_bit(float v) { return (v < 0.0f) || (v == 0.0f && (1.0f / v) < 0.0f); }
PT_HD float with_sign(float mag, float sgn) {
    const float a = mag < 0.0f ? -mag : mag;
    return sign_bit(sgn) ? -a : a;
}
// x = k pi/2 + r, |r| <= pi/4 (+ rounding); sin r and cos r; returns k mod 4
PT_HD int sincos_reduced(float x, float& sr, float& cr) {
    const float k = floorf(fmaf(x, kTwoOverPi, 0.5f));
    float r = fmaf(-k, kPio2Hi, x);
    r = fmaf(-k, kPio2Lo, r);
    r = fmaf(-k, kPio2Lo2, r);
    const float z = r * r;
    float ps = 2.724378191e-06f;
    ps = fmaf(ps, z, -1.984003902e-04f);
    ps = fmaf(ps, z, 8.333331905e-03f);
    ps = fmaf(ps, z, -1.666666716e-01f);
    sr = fmaf(r * z, ps, r);                         // r + r^3 S(r^2)
    float pc = -2.729583173e-07f;
    pc = fmaf(pc, z, 2.480055991e-05f);
    pc = fmaf(pc, z, -1.388888806e-03f);
    pc = fmaf(pc, z, 4.166666791e-02f);
    cr = fmaf(z * z, pc, fmaf(-0.5f, z, 1.0f));      // 1 - r^2/2 + r^4 C(r^2)
    return (int)k & 3;
}
PT_HD void sincos_f(float x, float& s, float& c) {
    float sr, cr;
    const int q = sincos_reduced(x, sr, cr);
    const float a = (q & 1) ? cr : sr, b = (q & 1) ? sr : cr;
    s = (q & 2) ? -a : a;                            // q: 0 (s, c)  1 (c, -s)  2 (-s, -c)  3 (-c, s)
    c = ((q + 1) & 2) ? -b : b;
}
PT_HD float atan_f(float a) {  // a >= 0 (may be +inf); result in [0, pi/2]
    const bool inv = a > 1.0f;
    const float t = inv ? 1.0f / a : a;
    const bool shift = t > 0.414213562f;  // tan(pi/8)
    const float u = shift ? (t - 1.0f) / (t + 1.0f) : t;
    const float z = u * u;
    float p = 5.044116825e-02f;
    p = fmaf(p, z, -8.623115718e-02f);
    p = fmaf(p, z, 1.107115597e-01f);
    p = fmaf(p, z, -1.428413987e-01f);
    p = fmaf(p, z, 1.999997646e-01f);
    p = fmaf(p, z, -3.333333433e-01f);
    float r = fmaf(u * z, p, u);                     // u + u^3 A(u^2)
    if (shift) r = kPio4Hi + (r + kPio4Lo);
    if (inv) r = (kPio2Hi - r) + kPio2Lo;
    return r;
}
}  // namespace dmf

// -DPT_LIBM_MATH=1 (measurement builds only, tools/build_variants.sh): the DEVICE code calls CUDA's float functions
// instead — what the deterministic design costs (frames then differ from the oracle in the last ulp of some draws
// and are compared by relMSE, not by bits).  The product is built without it.
#ifndef PT_TRIG_F32
#define PT_TRIG_F32 0
#endif
#ifndef PT_LIBM_MATH
#define PT_LIBM_MATH 0
#endif
#if PT_LIBM_MATH && defined(__CUDA_ARCH__)
#define PT_LIBM(expr) return (expr)
#else
#define PT_LIBM(expr)
#endif

// binary64 sin / cos / tan (which: 0, 1, 2), inlined: a handful of polynomial steps, cheap enough to sit in line
PT_HD float trig_d(float x, int which) {
    double s, c;
    dmi::sincos_d((double)x, s, c);
    return which == 0 ? (float)s : (which == 1 ? (float)c : (float)(s / c));
}
// |x| > dmf::kTrigMax with the binary32 route: the binary64 one, out of line
PT_HD_CALL float trig_big(float x, int which) {
    double s, c;
    dmi::sincos_d((double)x, s, c);
    return which == 0 ? (float)s : (which == 1 ? (float)c : (float)(s / c));
}

PT_HD float dm_sin(float x) {
    PT_LIBM(sinf(x));
    if (dmf::is_nan(x) || dmf::is_inf(x)) return dmi::qnan();
#if PT_TRIG_F32
    if (x <= dmf::kTrigMax && x >= -dmf::kTrigMax) {
        float sf, cf;
        dmf::sincos_f(x, sf, cf);
        return sf;
    }
    return trig_big(x, 0);
#else
    return trig_d(x, 0);
#endif
}
PT_HD float dm_cos(float x) {
    PT_LIBM(cosf(x));
    if (dmf::is_nan(x) || dmf::is_inf(x)) return dmi::qnan();
#if PT_TRIG_F32
    if (x <= dmf::kTrigMax && x >= -dmf::kTrigMax) {
        float sf, cf;
        dmf::sincos_f(x, sf, cf);
        return cf;
    }
    return trig_big(x, 1);
#else
    return trig_d(x, 1);
#endif
}
PT_HD float dm_tan(float x) {
    PT_LIBM(tanf(x));
    if (dmf::is_nan(x) || dmf::is_inf(x)) return dmi::qnan();
#if PT_TRIG_F32
    if (x <= dmf::kTrigMax && x >= -dmf::kTrigMax) {
        float sf, cf;
        dmf::sincos_f(x, sf, cf);
        return sf / cf;
    }
    return trig_big(x, 2);
#else
    return trig_d(x, 2);
#endif
}
PT_HD float dm_atan(float x) {
    PT_LIBM(atanf(x));
    if (dmf::is_nan(x)) return dmi::qnan();
#if PT_TRIG_F32
    return dmf::with_sign(dmf::atan_f(x < 0.0f ? -x : x), x);
#else
    const double xd = x;
    return (float)dmi::copysign_d(dmi::atan_core(xd < 0 ? -xd : xd), xd);
#endif
}
PT_HD float dm_atan2(float y, float x) {
    PT_LIBM(atan2f(y, x));
    if (dmf::is_nan(y) || dmf::is_nan(x)) return dmi::qnan();
#if !PT_TRIG_F32
    {
        const double yd = y, xd = x;
        const bool xn = (xd < 0) || (xd == 0 && (1.0 / xd) < 0);
        if (yd == 0) return (float)dmi::copysign_d(xn ? dmi::kPiD : 0.0, yd);
        if (xd == 0) return (float)dmi::copysign_d(dmi::kPio2Hi, yd);
        double a;
        if (dmi::is_inf(xd) && dmi::is_inf(yd))
            a = dmi::kPio4;
        else {
            const double q = yd / xd;
            a = dmi::atan_core(q < 0 ? -q : q);
        }
        if (xn) a = (dmi::kPiD - a) + 2.0 * dmi::kPio2Lo;
        return (float)dmi::copysign_d(a, yd);
    }
#endif
    const bool xneg = dmf::sign_bit(x);
    if (y == 0.0f) return dmf::with_sign(xneg ? dmf::kPiHi : 0.0f, y);
    if (x == 0.0f) return dmf::with_sign(dmf::kPio2Hi, y);
    float a;
    if (dmf::is_inf(x) && dmf::is_inf(y))
        a = dmf::kPio4Hi;
    else {
        const float q = y / x;   // an overflowing quotient gives pi/2, an underflowing one 0 (or pi): both within an ulp
        a = dmf::atan_f(q < 0.0f ? -q : q);
    }
    if (xneg) a = (dmf::kPiHi - a) + dmf::kPiLo;
    return dmf::with_sign(a, y);
}
PT_HD float dm_asin(float x) {
    PT_LIBM(asinf(x));
    if (dmf::is_nan(x) || x > 1.0f || x < -1.0f) return dmi::qnan();
#if !PT_TRIG_F32
    {
        const double xd = x;
        const double cd = sqrt((1.0 - xd) * (1.0 + xd));
        const double ax = xd < 0 ? -xd : xd;
        const double a = (cd == 0.0) ? (dmi::kPio2Hi + dmi::kPio2Lo) : dmi::atan_core(ax / cd);
        return (float)dmi::copysign_d(a, xd);
    }
#endif
    const float c = sqrtf((1.0f - x) * (1.0f + x));
    const float ax = x < 0.0f ? -x : x;
    const float a = (c == 0.0f) ? dmf::kPio2Hi : dmf::atan_f(ax / c);
    return dmf::with_sign(a, x);
}
PT_HD_CALL float dm_log(float x) {
    PT_LIBM(logf(x));
    double xd = x;
    if (dmi::is_nan(xd) || xd < 0) return dmi::qnan();
    if (xd == 0) return (float)(-HUGE_VAL);
    if (dmi::is_inf(xd)) return (float)HUGE_VAL;
    return (float)dmi::log_core(xd);
}
// x >= 0 is the only domain the integrator uses (colours, alpha^2); x < 0 -> NaN.
PT_HD_CALL float dm_pow(float x, float y) {
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
