// oracle.cpp — scalar CPU restatement of the reference's render path.  TEST INFRASTRUCTURE.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may build, link or call this file.  The product (libb200pt.so) never does.
//
// What it restates (all paths relative to /root/reference/path_tracer/src):
//   random.hpp:34-85            TEA seed + LCG draw                      -> Rng
//   camera.cpp:3-22             look-at camera                           -> oracle_camera
//   device/device.cu:220-254    __raygen__ray_gen                        -> render_pixel
//   device/device.cu:113-218    trace_path                               -> trace_path
//   device/device.cu:23-94      env lookup, mesh fetch, texture fetch    -> env_lookup, ...
//   device/disney/*.cuh         Disney lobes                             -> disney_*
//   math.hpp, sample_methods.hpp:19-65                                   -> helpers below
//   extern/owl/.../owl_device.h:79-91   make_8bit / make_rgba            -> quantise8
//   extern/owl/.../vec.h:330-355        dot, cross, normalize            -> V3 helpers
// What it cannot restate: the OptiX driver's BVH traversal and watertight triangle test
// (closed source, call site device.cu:133).  The closest hit is therefore DEFINED here
// as: smallest Moller-Trumbore t in (tmin, tmax), ties to the smaller global triangle
// id — independent of any acceleration structure.  Two finders implement it: brute force
// over all triangles, and a plain binary BVH (median split) with padded boxes; tests check
// they agree.
//
// PARITY STATUS: pinned by the reference itself, except at the traversal boundary.  oracle/ref/ compiles the
// reference's own device.cu (raygen, trace_path, hit / miss programs), device/disney/*.cuh, math.hpp,
// sample_methods.hpp, random.hpp and camera.cpp for the host (oracle/_ref/libref_device.so; one mechanical edit, see
// oracle/ref/patch.py) and tests/test_ref_device.py requires the glibc build of this file to equal it BIT FOR BIT:
// sample_disney on the 960-case grid + 262,144 random cases, RNG, camera, ONB, make_rgba, and whole frames of all six
// scenes (float mean, RGBA8, ray count).  tests/golden/ref_*.npz are vectors generated from that library
// (tools/make_golden_ref.py) so the pin also holds where /root/reference is absent.  Unpinned by nature: the closest
// hit itself (OptiX is closed source and the reference ships no vectors for it) — defined above.
//
// Elementary functions: default build uses glibc libm (independent of the product).
// -DORACLE_DETMATH switches to the product's deterministic functions
// (owl_path_tracer_b200/csrc/pt_detmath.h) so that whole images can be compared bit for
// bit with the GPU; tests/test_detmath.py checks those functions against float64.
//
// Build: see oracle/Makefile (g++ -O2 -ffp-contract=off -fno-fast-math).
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <atomic>
#include <thread>
#include <vector>

#ifdef ORACLE_DETMATH
#include "../owl_path_tracer_b200/csrc/pt_detmath.h"
static inline float m_sin(float x) { return pt::dm_sin(x); }
static inline float m_cos(float x) { return pt::dm_cos(x); }
static inline float m_tan(float x) { return pt::dm_tan(x); }
static inline float m_atan(float x) { return pt::dm_atan(x); }
static inline float m_atan2(float y, float x) { return pt::dm_atan2(y, x); }
static inline float m_asin(float x) { return pt::dm_asin(x); }
static inline float m_log(float x) { return pt::dm_log(x); }
static inline float m_pow(float x, float y) { return pt::dm_pow(x, y); }
#else
static inline float m_sin(float x) { return sinf(x); }
static inline float m_cos(float x) { return cosf(x); }
static inline float m_tan(float x) { return tanf(x); }
static inline float m_atan(float x) { return atanf(x); }
static inline float m_atan2(float y, float x) { return atan2f(y, x); }
static inline float m_asin(float x) { return asinf(x); }
static inline float m_log(float x) { return logf(x); }
static inline float m_pow(float x, float y) { return powf(x, y); }
#endif

namespace {

// ---------------------------------------------------------------- vectors (owl vec.h)
struct V3 {
    float x, y, z;
};
struct V2 {
    float x, y;
};
inline V3 v3(float a) { return {a, a, a}; }
inline V3 add(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 mul(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline V3 scale(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }   // vec * scalar
inline V3 scale(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }   // scalar * vec
inline V3 divs(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline V3 neg(V3 a) { return {-a.x, -a.y, -a.z}; }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }            // vec.h:346
inline V3 cross(V3 a, V3 b) {                                                          // vec.h:330
    return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y};
}
inline V3 normalize(V3 v) {  // vec.h:353: v * rsqrt(dot), rsqrt = 1.f/sqrtf (owl-common.h:263)
    float r = 1.f / sqrtf(dot(v, v));
    return scale(v, r);
}

// ---------------------------------------------------------------- constants (types.hpp:9-18)
const float pi = 3.14159265358979323f;
const float two_pi = 6.28318530717958648f;
const float pi_over_two = 1.57079632679489661f;
const float pi_over_four = 0.78539816339744830f;
const float inv_pi = 0.31830988618379067f;
const float t_min = 1E-3f;
const float t_max = 1E10f;
const float alpha_min = 0.001f;

// ---------------------------------------------------------------- math.hpp
inline float lerp(float a, float b, float t) { return a + (b - a) * t; }                 // :6
inline V3 lerp(V3 a, V3 b, float t) { return add(a, scale(sub(b, a), t)); }              // :10
inline float sqr(float v) { return v * v; }                                              // :16
inline float clampf(float x, float lo, float hi) { return fminf(hi, fmaxf(lo, x)); }     // disney_helper.cuh:14
inline float cos_theta(V3 w) { return w.z; }                                             // :22
inline float sin_theta(V3 w) { return sqrtf(fmaxf(0.0f, 1.0f - sqr(cos_theta(w)))); }    // :24
inline float tan_theta(V3 w) { return sin_theta(w) / cos_theta(w); }                     // :26
inline float cos_phi(V3 w) {                                                             // :28-32
    float s = sin_theta(w);
    return (s == 0) ? 1.0f : clampf(w.x / s, -1.0f, 1.0f);
}
inline float sin_phi(V3 w) {                                                             // :34-38
    float s = sin_theta(w);
    return (s == 0) ? 1.0f : clampf(w.y / s, -1.0f, 1.0f);
}
inline V3 sphere_dir(float theta, float phi) {                                           // :42-48
    float x = m_sin(theta) * m_cos(phi);
    float y = m_sin(theta) * m_sin(phi);
    float z = m_cos(theta);
    return {x, y, z};
}
inline V3 sphere_dir(float sin_t, float cos_t, float phi) {                              // :50-56
    return {sin_t * m_cos(phi), sin_t * m_sin(phi), cos_t};
}
inline V3 reflect(V3 w, V3 n) { return sub(scale(2.0f, scale(dot(w, n), n)), w); }       // :58-61
inline bool refract(V3 w, V3 n, float eta, V3& wi) {                                     // :63-77
    if (eta == 1.0f) {
        wi = neg(w);
        return true;
    }
    float cos_i = dot(w, n);
    float sin2_i = fmaxf(0.0f, 1.0f - sqr(cos_i));
    float sin2_t = eta * eta * sin2_i;
    if (sin2_t > 1.0f) return false;
    float cos_t = sqrtf(1.0f - sin2_t);
    wi = add(scale(eta, neg(w)), scale(eta * cos_i - cos_t, n));
    return true;
}
inline bool same_hemisphere(V3 a, V3 b) { return a.z * b.z > 0.0f; }                     // :79-82
inline void onb(V3 n, V3& t, V3& b) {                                                    // :86-95
    if (n.x != n.y || n.x != n.z)
        t = {n.z - n.y, n.x - n.z, n.y - n.x};
    else
        t = {n.z - n.y, n.x + n.z, -n.y - n.x};
    t = normalize(t);
    b = cross(n, t);
}
inline V3 to_local(V3 t, V3 b, V3 n, V3 w) { return normalize(V3{dot(w, t), dot(w, b), dot(w, n)}); }  // :98-101
inline V3 to_world(V3 t, V3 b, V3 n, V3 w) {                                             // :104-107
    return normalize(add(add(scale(w.x, t), scale(w.y, b)), scale(w.z, n)));
}

// ---------------------------------------------------------------- random.hpp:34-85
struct Rng {
    uint32_t state;
    void seed(uint32_t u, uint32_t v) {
        uint32_t s = 0;
        for (int n = 0; n < 4; n++) {
            s += 0x9e3779b9;
            u += ((v << 4) + 0xa341316c) ^ (v + s) ^ ((v >> 5) + 0xc8013ea4);
            v += ((u << 4) + 0xad90777d) ^ (u + s) ^ ((u >> 5) + 0x7e95761e);
        }
        state = u;
    }
    float next() {
        state = 16807u * state + 1013904223u;
        return ldexpf((float)state, -32);
    }
};

// ---------------------------------------------------------------- material_data (device_global.hpp:19-36)
struct Mat {
    float base_color[3];
    float subsurface, metallic, specular, specular_tint, roughness, anisotropic, sheen, sheen_tint,
        clearcoat, clearcoat_gloss, ior, specular_transmission, specular_transmission_roughness, emission;
};
static_assert(sizeof(Mat) == 68, "material_data is 17 floats");
inline V3 base_of(const Mat& m) { return {m.base_color[0], m.base_color[1], m.base_color[2]}; }

enum { LOBE_NONE = -1, LOBE_DIFFUSE = 0, LOBE_CLEARCOAT = 1, LOBE_METALLIC = 2, LOBE_GLASS = 3 };  // disney.cuh:9-13

// ---------------------------------------------------------------- disney_helper.cuh
inline float luminance(V3 c) { return dot(V3{0.2126f, 0.7152f, 0.0722f}, c); }           // :9-12
inline float schlick_weight(float c) {                                                   // :19-24
    float m = clampf(1.0f - c, 0.0f, 1.0f);
    float m2 = m * m;
    return m2 * m2 * m;
}
inline float roughness_to_alpha(float r) { return fmaxf(alpha_min, clampf(sqr(r), 0.0f, 1.0f)); }  // :39-42
inline V2 roughness_to_alpha(float r, float aniso) {                                     // :44-48
    float aspect = sqrtf(1.0f - 0.9f * aniso);
    return {fmaxf(alpha_min, sqr(r) / aspect), fmaxf(alpha_min, sqr(r) * aspect)};
}
inline float fresnel_equation(V3 i, V3 m, float eta_i, float eta_t) {                    // :51-60
    float c = fabsf(dot(i, m));
    float denom = sqr(eta_t / eta_i) - 1.0f + sqr(c);
    if (denom < 0.0f) return 1.0f;
    float g = sqrtf(denom);
    return 0.5f * sqr((g - c) / (g + c)) * (1.0f + sqr(c * (g + c) - 1.0f) / sqr(c * (g - c) + 1.0f));
}

// ---------------------------------------------------------------- sample_methods.hpp:19-57
inline V2 concentric_disk(V2 u) {
    float dx = 2.0f * u.x - 1;
    float dy = 2.0f * u.y - 1;
    if (dx == 0 && dy == 0) return {0.0f, 0.0f};
    float phi, r;
    if (fabsf(dx) > fabsf(dy)) {
        r = dx;
        phi = pi_over_four * (dy / dx);
    } else {
        r = dy;
        phi = pi_over_two - pi_over_four * (dx / dy);
    }
    return {r * m_cos(phi), r * m_sin(phi)};
}
inline V3 cosine_hemisphere(V2 u) {
    V2 p = concentric_disk(u);
    float c = sqrtf(fmaxf(0.0f, 1.0f - sqr(p.x) - sqr(p.y)));
    return {p.x, p.y, c};
}

// ---------------------------------------------------------------- disney_specular.cuh
float lambda(V3 w, float ax, float ay) {                                                 // :17-28
    float tt = tan_theta(w);
    if (isinf(tt)) return 0.0f;
    float alpha0 = sqrtf(sqr(cos_phi(w) * ax) + sqr(sin_phi(w) * ay));
    float a = 1.0f / (alpha0 * tt);
    return (-1.0f + sqrtf(1.0f + 1.0f / sqr(a))) / 2.0f;
}
float g1_smith(V3 w, float ax, float ay) { return 1.0f / (1.0f + lambda(w, ax, ay)); }   // :32-35
float g2_separable(V3 wo, V3 wi, float ax, float ay) { return g1_smith(wo, ax, ay) * g1_smith(wi, ax, ay); }  // :39-42
float g2_correlated(V3 wo, V3 wi, float ax, float ay) {                                  // :47-50
    return 1.0f / (1.0f + lambda(wo, ax, ay) + lambda(wi, ax, ay));
}
float d_gtr2(V3 wm, float ax, float ay) {                                                // :53-60
    float tan2 = sqr(tan_theta(wm));
    if (isinf(tan2)) return 0.0f;
    float cos4 = sqr(sqr(cos_theta(wm)));
    float e = 1.0f + tan2 * (sqr(cos_phi(wm)) / sqr(ax) + sqr(sin_phi(wm)) / sqr(ay));
    return 1.0f / (pi * ax * ay * cos4 * sqr(e));
}
V3 sample_gtr2_ndf(float ax, float ay, V2 u) {                                           // :64-81
    float phi = m_atan(ay / ax * m_tan(two_pi * u.y + inv_pi));
    if (u.y > .5f) phi += pi;
    float sin_p = m_sin(phi), cos_p = m_cos(phi);
    float ax2 = sqr(ax), ay2 = sqr(ay);
    float alpha2 = 1 / (sqr(cos_p) / ax2 + sqr(sin_p) / ay2);
    float tan2 = alpha2 * u.x / (1 - u.x);
    float cos_t = 1 / sqrtf(1 + tan2);
    float sin_t = sqrtf(fmaxf(0.f, 1.f - sqr(cos_t)));
    return normalize(sphere_dir(sin_t, cos_t, phi));
}
V3 eval_specular_brdf(const Mat& m, V3 wo, V3 wh, V3 wi, float& pdf) {                   // :125-149
    V3 base = base_of(m);
    float lum = luminance(base);
    V3 c_tint = lum > 0.0f ? divs(base, lum) : v3(1.0f);
    V3 c_spec = lerp(scale(.08f * m.specular, lerp(v3(1.0f), c_tint, m.specular_tint)), base, m.metallic);
    V2 a = roughness_to_alpha(m.roughness, m.anisotropic);
    float d = d_gtr2(wh, a.x, a.y);
    float g = g2_correlated(wo, wi, a.x, a.y);
    V3 f = lerp(c_spec, v3(1.0f), schlick_weight(dot(wi, wh)));
    pdf = d * g1_smith(wo, a.x, a.y) * fmaxf(0.0f, dot(wo, wh)) / (4.0f * cos_theta(wo));
    return divs(scale(d * g, f), 4.0f * fabsf(cos_theta(wo)));
}
V3 sample_specular_brdf(const Mat& m, V3 wo, Rng& rng, V3& wi, float& pdf) {             // :151-170
    V2 a = roughness_to_alpha(m.roughness, m.anisotropic);
    V2 u;
    u.x = rng.next();
    u.y = rng.next();
    V3 wh = sample_gtr2_ndf(a.x, a.y, u);
    if (dot(wo, wh) < 0.0f) wh = neg(wh);
    wi = reflect(wo, wh);
    if (cos_theta(wi) <= 0.0f) {
        pdf = 0.0f;
        return v3(0.0f);
    }
    return eval_specular_brdf(m, wo, wh, wi, pdf);
}
V3 eval_specular_bsdf(const Mat& m, V3 wo, V3 wh, V3 wi, float& pdf) {                   // :193-214
    float eta_i = cos_theta(wo) > 0.0f ? 1.0f : m.ior;                                   // disney_helper.cuh:31-37
    float eta_t = cos_theta(wo) > 0.0f ? m.ior : 1.0f;
    float eta = eta_i / eta_t;
    float R = fresnel_equation(wo, wh, eta_i, eta_t);
    float T = 1.0f - R;
    if (same_hemisphere(wo, wi)) {
        pdf = R / (R + T);
        return divs(scale(base_of(m), R), fabsf(cos_theta(wi)));
    }
    pdf = T / (R + T);
    V3 b = base_of(m);
    V3 sb = {sqrtf(b.x), sqrtf(b.y), sqrtf(b.z)};
    return divs(divs(scale(sb, T), fabsf(cos_theta(wi))), sqr(eta));
}
V3 sample_specular_bsdf(const Mat& m, V3 wo, Rng& rng, V3& wi, float& pdf) {             // :216-244
    V2 u;
    u.x = rng.next();
    u.y = rng.next();
    float a = roughness_to_alpha(m.specular_transmission_roughness);
    float theta = m_atan((a * sqrtf(u.x)) / sqrtf(1.0f - u.x));                          // :175-180
    float phi = two_pi * u.y;
    V3 wh = sphere_dir(theta, phi);
    if (cos_theta(wo) < 0.0f && !same_hemisphere(wo, wh)) wh = neg(wh);
    float eta_i = cos_theta(wo) > 0.0f ? 1.0f : m.ior;
    float eta_t = cos_theta(wo) > 0.0f ? m.ior : 1.0f;
    float eta = eta_i / eta_t;
    float R = fresnel_equation(wo, wh, eta_i, eta_t);
    float T = 1.0f - R;
    if (!refract(wo, wh, eta, wi) || rng.next() < R / (R + T)) {
        V2 al = roughness_to_alpha(m.roughness, m.anisotropic);
        V2 u2;
        u2.x = rng.next();
        u2.y = rng.next();
        wh = sample_gtr2_ndf(al.x, al.y, u2);
        wi = normalize(reflect(wo, wh));
    }
    return eval_specular_bsdf(m, wo, wh, wi, pdf);
}

// ---------------------------------------------------------------- disney_diffuse.cuh:26-62
V3 sample_diffuse(const Mat& m, V3 wo, Rng& rng, V3& wi, float& pdf) {
    V2 u;
    u.x = rng.next();
    u.y = rng.next();
    wi = cosine_hemisphere(u);
    float fo = schlick_weight(cos_theta(wo));
    float fi = schlick_weight(cos_theta(wi));
    V3 lambert = scale(base_of(m), inv_pi);
    float fd = (1.0f - 0.5f * fo) * (1.0f - 0.5f * fi);
    float rr = m.roughness * (dot(wo, wi) + 1.0f);
    float fr = rr * (fi + fo + fo * fi * (rr - 1.0f));
    pdf = fabsf(cos_theta(wi)) * inv_pi;                                                  // sample_methods.hpp:59-62
    return scale(lambert, fd + fr);
}

// ---------------------------------------------------------------- disney_clearcoat.cuh:13-78
V3 sample_clearcoat(const Mat& m, V3 wo, Rng& rng, V3& wi, float& pdf) {
    float a = lerp(0.1f, .001f, m.clearcoat_gloss);
    V2 u;
    u.x = rng.next();
    u.y = rng.next();
    float alpha2 = sqr(a);                                                                // :24-35
    float cos_t = sqrtf(fmaxf(0.0f, (1.0f - m_pow(alpha2, 1.0f - u.x)) / (1.0f - alpha2)));
    float sin_t = sqrtf(fmaxf(0.0f, 1.0f - sqr(cos_t)));
    float phi = two_pi * u.y;
    V3 wh = sphere_dir(sin_t, cos_t, phi);
    if (!same_hemisphere(wo, wh)) wh = neg(wh);
    if (dot(wh, wo) < 0.0f) wh = neg(wh);                                                 // :67
    wh = normalize(wh);
    wi = reflect(wo, wh);
    if (!same_hemisphere(wo, wi)) {
        pdf = 0.0f;
        return v3(0.0f);
    }
    if (m.clearcoat <= 0.0f) {                                                            // :47-51
        pdf = 0.0f;
        return v3(0.0f);
    }
    float d;                                                                              // :13-20
    if (a >= 1)
        d = inv_pi;
    else {
        float a2 = sqr(a);
        d = (a2 - 1.0f) / (pi * m_log(a2) * (1.0f + (a2 - 1.0f) * sqr(cos_theta(wh))));
    }
    float f = lerp(1.0f, schlick_weight(cos_theta(wi)), 0.04f);
    float g = g2_separable(wo, wi, .25f, .25f);
    pdf = d / (4.0f * dot(wh, wi));
    return v3(d * g * f / (4.0f * fabsf(cos_theta(wo)) * fabsf(cos_theta(wi))));
}

// ---------------------------------------------------------------- disney_sheen.cuh:15-37
V3 eval_sheen(const Mat& m, V3 wo, V3 wi) {
    if (m.sheen <= 0.0f) return v3(0.0f);
    V3 wh = add(wi, wo);
    if (wh.x == 0.0f && wh.y == 0.0f && wh.z == 0.0f) return v3(0.0f);
    wh = normalize(wh);
    V3 b = base_of(m);
    V3 lin = {m_pow(b.x, 2.2f), m_pow(b.y, 2.2f), m_pow(b.z, 2.2f)};                      // disney_helper.cuh:4-7
    float lum = luminance(lin);
    float cos_d = dot(wi, wh);
    V3 tint = (lum > 0.0f) ? divs(b, lum) : v3(1.0f);
    return scale(scale(lerp(v3(1.0f), tint, m.sheen_tint), m.sheen), schlick_weight(cos_d));
}

// ---------------------------------------------------------------- next-event estimation support
// Not part of the reference's integrator (its light sampling code is dead); an optional
// mode assembled from the reference's own eval_* functions and lights.hpp.  Mirrors
// owl_path_tracer_b200/csrc/pt_disney.h eval_mixture; see DESIGN.md "NEE mode" for the estimator.
V3 eval_diffuse(const Mat& m, V3 wo, V3 wi, float& pdf) {                                 // disney_diffuse.cuh:26-55
    float fo = schlick_weight(cos_theta(wo));
    float fi = schlick_weight(cos_theta(wi));
    V3 lambert = scale(base_of(m), inv_pi);
    float fd = (1.0f - 0.5f * fo) * (1.0f - 0.5f * fi);
    float rr = m.roughness * (dot(wo, wi) + 1.0f);
    float fr = rr * (fi + fo + fo * fi * (rr - 1.0f));
    pdf = fabsf(cos_theta(wi)) * inv_pi;
    return scale(lambert, fd + fr);
}
V3 eval_clearcoat(const Mat& m, V3 wo, V3 wh, V3 wi, float& pdf) {                       // disney_clearcoat.cuh:13-20,45-59
    float a = lerp(0.1f, .001f, m.clearcoat_gloss);
    float a2 = sqr(a);
    float d;
    if (a >= 1)
        d = inv_pi;
    else
        d = (a2 - 1.0f) / (pi * m_log(a2) * (1.0f + (a2 - 1.0f) * sqr(cos_theta(wh))));
    float f = lerp(1.0f, schlick_weight(cos_theta(wi)), 0.04f);
    float g = g2_separable(wo, wi, .25f, .25f);
    pdf = d / (4.0f * dot(wh, wi));
    return v3(d * g * f / (4.0f * fabsf(cos_theta(wo)) * fabsf(cos_theta(wi))));
}
// f_eff = sum_l p_l (f_l + sheen), pdf_eff = sum_l p_l pdf_l over metal, clearcoat, diffuse
void eval_mixture(const Mat& m, V3 wo, V3 wi, V3& f_eff, float& pdf_eff) {
    float diffuse_w = (1.0f - m.specular_transmission) * (1.0f - m.metallic);
    float metallic_w = m.metallic;
    float clearcoat_w = .25f * m.clearcoat;
    float glass_w = (1.0f - m.metallic) * m.specular_transmission;
    float factor = 1.0f / (metallic_w + glass_w + diffuse_w + clearcoat_w);
    float p_metallic = metallic_w * factor;
    float p_diffuse = diffuse_w * factor;
    float p_clearcoat = clearcoat_w * factor;
    f_eff = v3(0.0f);
    pdf_eff = 0.0f;
    V3 wh = add(wo, wi);
    const bool has_wh = !(wh.x == 0.0f && wh.y == 0.0f && wh.z == 0.0f);
    if (has_wh) wh = normalize(wh);
    const V3 sheen = eval_sheen(m, wo, wi);
    if (p_metallic > 0.0f && has_wh && cos_theta(wi) > 0.0f) {
        float pdf;
        V3 f = eval_specular_brdf(m, wo, wh, wi, pdf);
        f_eff = add(f_eff, scale(p_metallic, add(f, sheen)));
        pdf_eff = pdf_eff + p_metallic * pdf;
    }
    if (p_clearcoat > 0.0f && has_wh && same_hemisphere(wo, wi)) {
        float pdf;
        V3 f = eval_clearcoat(m, wo, wh, wi, pdf);
        f_eff = add(f_eff, scale(p_clearcoat, add(f, sheen)));
        pdf_eff = pdf_eff + p_clearcoat * pdf;
    }
    if (p_diffuse > 0.0f && cos_theta(wi) > 0.0f) {
        float pdf;
        V3 f = eval_diffuse(m, wo, wi, pdf);
        f_eff = add(f_eff, scale(p_diffuse, add(f, sheen)));
        pdf_eff = pdf_eff + p_diffuse * pdf;
    }
}
inline float power_heuristic(float pdf_f, float pdf_g) {                                  // lights.hpp:56-61, n_f = n_g = 1
    float f = pdf_f, g = pdf_g;
    return (f * f) / (f * f + g * g);
}
inline float pdf_area_to_solid_angle(float pdf_area, float dist_sqr, float cos_t) {     // lights.hpp:23-28
    float a = fabsf(cos_t);
    if (a < 1e-4f) return 0.0f;
    return pdf_area * dist_sqr / a;
}

// ---------------------------------------------------------------- disney.cuh:15-66
V3 sample_disney(const Mat& m, V3 wo, Rng& rng, V3& wi, float& pdf, int& sampled_lobe) {
    float diffuse_w = (1.0f - m.specular_transmission) * (1.0f - m.metallic);
    float metallic_w = m.metallic;
    float clearcoat_w = .25f * m.clearcoat;
    float glass_w = (1.0f - m.metallic) * m.specular_transmission;
    float factor = 1.0f / (metallic_w + glass_w + diffuse_w + clearcoat_w);
    float p_metallic = metallic_w * factor;
    float p_glass = glass_w * factor;
    float p_diffuse = diffuse_w * factor;
    float p_clearcoat = clearcoat_w * factor;

    bool force_btdf = cos_theta(wo) < 0.0f && sampled_lobe == LOBE_GLASS;
    float p = rng.next();
    V3 f = v3(0.0f);
    if (!force_btdf && p <= p_metallic) {
        f = sample_specular_brdf(m, wo, rng, wi, pdf);
        sampled_lobe = LOBE_METALLIC;
    } else if (!force_btdf && p > p_metallic && p <= (p_metallic + p_clearcoat)) {
        f = sample_clearcoat(m, wo, rng, wi, pdf);
        sampled_lobe = LOBE_CLEARCOAT;
    } else if (!force_btdf && p > p_metallic + p_clearcoat && p <= (p_metallic + p_clearcoat + p_diffuse)) {
        f = sample_diffuse(m, wo, rng, wi, pdf);
        sampled_lobe = LOBE_DIFFUSE;
    } else if (force_btdf || p_glass >= 0.0f) {
        f = sample_specular_bsdf(m, wo, rng, wi, pdf);
        sampled_lobe = LOBE_GLASS;
    }
    return add(f, eval_sheen(m, wo, wi));
}

// ---------------------------------------------------------------- scene
struct MeshRec {
    std::vector<V3> verts, normals;
    std::vector<V2> uvs;
    std::vector<int32_t> idx;  // 3 per triangle, local
    int material;
    std::vector<uint32_t> tex;
    int tex_w = 0, tex_h = 0;
    int first_tri = 0;
};
struct TriRef {
    V3 v0, e1, e2;
    uint32_t id;  // global triangle id
};
struct BNode {
    float lo[3], hi[3];
    int left, right;  // children, or (first,count) when leaf
    int first, count;
};
struct HitRec {
    int32_t mesh, prim;
    float t, u, v;
};

struct Scene {
    std::vector<MeshRec> meshes;
    std::vector<Mat> mats;
    std::vector<uint32_t> env;
    std::vector<float> env_f32;   // float RGB map (oracle_set_environment_f32) instead of env; same addressing
    int env_w = 0, env_h = 0;
    bool env_use = false, env_auto = false;
    V3 env_color{0, 0, 0};
    float env_intensity = 0;
    V3 cam_origin{}, cam_llc{}, cam_h{}, cam_v{};
    int width = 0, height = 0, max_samples = 1, max_depth = 1;
    // acceleration
    std::vector<TriRef> tris;       // BVH order
    std::vector<TriRef> tris_flat;  // id order
    std::vector<BNode> nodes;
    std::vector<int32_t> tri_mesh;  // global id -> mesh
    // next-event estimation mode (oracle_set_nee): emissive triangles, chosen by area * emission
    bool nee = false;
    int transfer = 0;      // output transfer of the RGBA8 frame (oracle_set_output)
    struct Light {
        int mesh, prim;
        float area, prob, emission;
    };
    std::vector<Light> lights;
    std::vector<float> light_cdf;
    std::vector<int> mesh_light_first;
    std::vector<float> env_cdf_rows, env_cdf_cols;  // NEE mode, environment map in use: luminance cdfs (rows, then columns per row)
    std::atomic<uint64_t> rays{0};
    std::atomic<uint64_t> shadow_rays{0};  // of the last render (NEE mode)
};

// Moller-Trumbore.  This expression order is the parity contract with the GPU
// (owl_path_tracer_b200/csrc/pt_bvh8.h ray_triangle): products accumulated with explicit fused
// multiply-adds, one IEEE rounding each (std::fmaf is exact on every host, with or without FMA hardware).
inline float dot_fma(V3 a, V3 b) { return std::fmaf(a.z, b.z, std::fmaf(a.y, b.y, a.x * b.x)); }
inline V3 cross_fma(V3 a, V3 b) {
    return V3{std::fmaf(a.y, b.z, -(a.z * b.y)), std::fmaf(a.z, b.x, -(a.x * b.z)), std::fmaf(a.x, b.y, -(a.y * b.x))};
}
inline bool ray_triangle(V3 o, V3 d, const TriRef& tr, float tmin, float tmax, float& t, float& u, float& v) {
    V3 p = cross_fma(d, tr.e2);
    float det = dot_fma(tr.e1, p);
    if (!(det != 0.0f)) return false;
    float inv = 1.0f / det;
    V3 s = sub(o, tr.v0);
    float uu = dot_fma(s, p) * inv;
    if (!(uu >= 0.0f && uu <= 1.0f)) return false;
    V3 q = cross_fma(s, tr.e1);
    float vv = dot_fma(d, q) * inv;
    if (!(vv >= 0.0f && uu + vv <= 1.0f)) return false;
    float tt = dot_fma(tr.e2, q) * inv;
    if (!(tt > tmin && tt < tmax)) return false;
    t = tt;
    u = uu;
    v = vv;
    return true;
}

struct Closest {
    float t, u, v;
    uint32_t id;
};
inline void consider(const TriRef& tr, V3 o, V3 d, float tmin, float tmax, Closest& c) {
    float t, u, v;
    if (ray_triangle(o, d, tr, tmin, tmax, t, u, v)) {
        if (t < c.t || (t == c.t && tr.id < c.id)) c = {t, u, v, tr.id};
    }
}

Closest closest_brute(const Scene& sc, V3 o, V3 d, float tmin, float tmax) {
    Closest c{tmax, 0, 0, 0xFFFFFFFFu};
    for (const TriRef& tr : sc.tris_flat) consider(tr, o, d, tmin, tmax, c);
    return c;
}

Closest closest_bvh(const Scene& sc, V3 o, V3 d, float tmin, float tmax) {
    Closest c{tmax, 0, 0, 0xFFFFFFFFu};
    if (sc.nodes.empty()) return c;
    // double-precision slab test with a relative guard band: never culls a box the
    // single-precision triangle test could still hit
    const double ox = o.x, oy = o.y, oz = o.z;
    const double ix = 1.0 / (double)d.x, iy = 1.0 / (double)d.y, iz = 1.0 / (double)d.z;
    int stack[128];
    int sp = 0;
    stack[sp++] = 0;
    while (sp) {
        const BNode& n = sc.nodes[stack[--sp]];
        double t0 = tmin, t1 = c.t;
        const double lo[3] = {n.lo[0], n.lo[1], n.lo[2]}, hi[3] = {n.hi[0], n.hi[1], n.hi[2]};
        const double oo[3] = {ox, oy, oz}, ii[3] = {ix, iy, iz};
        const float dd[3] = {d.x, d.y, d.z};
        bool miss = false;
        for (int a = 0; a < 3 && !miss; ++a) {
            if (dd[a] == 0.0f) {
                if (oo[a] < lo[a] || oo[a] > hi[a]) miss = true;
                continue;
            }
            double ta = (lo[a] - oo[a]) * ii[a], tb = (hi[a] - oo[a]) * ii[a];
            if (ta > tb) std::swap(ta, tb);
            ta -= fabs(ta) * 1e-6 + 1e-9;
            tb += fabs(tb) * 1e-6 + 1e-9;
            if (ta > t0) t0 = ta;
            if (tb < t1) t1 = tb;
            if (t0 > t1) miss = true;
        }
        if (miss) continue;
        if (n.count > 0) {
            for (int k = 0; k < n.count; ++k) consider(sc.tris[n.first + k], o, d, tmin, tmax, c);
        } else {
            stack[sp++] = n.left;
            stack[sp++] = n.right;
        }
    }
    return c;
}

void build_bvh(Scene& sc) {
    sc.tris_flat.clear();
    sc.tri_mesh.clear();
    int first = 0;
    float maxabs = 0.0f;
    for (size_t mi = 0; mi < sc.meshes.size(); ++mi) {
        MeshRec& m = sc.meshes[mi];
        m.first_tri = first;
        const int nt = (int)m.idx.size() / 3;
        for (int k = 0; k < nt; ++k) {
            V3 a = m.verts[m.idx[3 * k]], b = m.verts[m.idx[3 * k + 1]], c = m.verts[m.idx[3 * k + 2]];
            TriRef tr{a, sub(b, a), sub(c, a), (uint32_t)(first + k)};
            sc.tris_flat.push_back(tr);
            sc.tri_mesh.push_back((int32_t)mi);
        }
        for (const V3& v : m.verts) maxabs = fmaxf(maxabs, fmaxf(fabsf(v.x), fmaxf(fabsf(v.y), fabsf(v.z))));
        first += nt;
    }
    const int n = (int)sc.tris_flat.size();
    sc.tris = sc.tris_flat;
    sc.nodes.clear();
    if (n == 0) return;
    const float pad = maxabs * 1e-5f + 1e-30f;
    struct Box {
        float lo[3], hi[3], c[3];
    };
    std::vector<Box> boxes(n);
    auto box_of = [&](const TriRef& t) {
        Box b;
        V3 p[3] = {t.v0, add(t.v0, t.e1), add(t.v0, t.e2)};
        for (int a = 0; a < 3; ++a) {
            float q0 = (&p[0].x)[a], q1 = (&p[1].x)[a], q2 = (&p[2].x)[a];
            b.lo[a] = fminf(q0, fminf(q1, q2)) - pad;
            b.hi[a] = fmaxf(q0, fmaxf(q1, q2)) + pad;
            b.c[a] = 0.5f * (b.lo[a] + b.hi[a]);
        }
        return b;
    };
    std::vector<int> order(n);
    for (int i = 0; i < n; ++i) {
        order[i] = i;
        boxes[i] = box_of(sc.tris_flat[i]);
    }
    sc.nodes.reserve(2 * n);
    struct Job {
        int node, first, count;
    };
    std::vector<Job> jobs;
    sc.nodes.push_back(BNode{});
    jobs.push_back({0, 0, n});
    while (!jobs.empty()) {
        Job j = jobs.back();
        jobs.pop_back();
        BNode nd{};
        for (int a = 0; a < 3; ++a) {
            nd.lo[a] = 3e38f;
            nd.hi[a] = -3e38f;
        }
        float clo[3] = {3e38f, 3e38f, 3e38f}, chi[3] = {-3e38f, -3e38f, -3e38f};
        for (int k = 0; k < j.count; ++k) {
            const Box& b = boxes[order[j.first + k]];
            for (int a = 0; a < 3; ++a) {
                nd.lo[a] = fminf(nd.lo[a], b.lo[a]);
                nd.hi[a] = fmaxf(nd.hi[a], b.hi[a]);
                clo[a] = fminf(clo[a], b.c[a]);
                chi[a] = fmaxf(chi[a], b.c[a]);
            }
        }
        if (j.count <= 4) {
            nd.first = j.first;
            nd.count = j.count;
            sc.nodes[j.node] = nd;
            continue;
        }
        int axis = 0;
        if (chi[1] - clo[1] > chi[axis] - clo[axis]) axis = 1;
        if (chi[2] - clo[2] > chi[axis] - clo[axis]) axis = 2;
        const int mid = j.count / 2;
        std::nth_element(order.begin() + j.first, order.begin() + j.first + mid, order.begin() + j.first + j.count,
                         [&](int x, int y) { return boxes[x].c[axis] < boxes[y].c[axis]; });
        nd.count = 0;
        nd.left = (int)sc.nodes.size();
        nd.right = nd.left + 1;
        sc.nodes[j.node] = nd;
        sc.nodes.push_back(BNode{});
        sc.nodes.push_back(BNode{});
        jobs.push_back({nd.left, j.first, mid});
        jobs.push_back({nd.right, j.first + mid, j.count - mid});
    }
    for (int i = 0; i < n; ++i) sc.tris[i] = sc.tris_flat[order[i]];
}

// ---------------------------------------------------------------- device.cu helpers
// tex2D on an RGBA8 / point / clamp / normalised-coordinate texture (owl.hpp:248-257)
V3 fetch_rgba8(const uint32_t* img, int w, int h, float u, float v) {
    float fx = floorf(u * (float)w), fy = floorf(v * (float)h);
    int ix = (fx >= 0.0f) ? (fx < (float)w ? (int)fx : w - 1) : 0;
    int iy = (fy >= 0.0f) ? (fy < (float)h ? (int)fy : h - 1) : 0;
    uint32_t p = img[(size_t)iy * w + ix];
    return {(float)(p & 0xFF) / 255.0f, (float)((p >> 8) & 0xFF) / 255.0f, (float)((p >> 16) & 0xFF) / 255.0f};
}
V3 sample_environment(const Scene& sc, V3 dir) {                                         // device.cu:23-39
    float u = 0.5f + m_atan2(dir.x, dir.z) / (2.0f * pi);
    float v = 0.5f + m_asin(dir.y) / pi;
    if (!sc.env_f32.empty()) {  // float map: same point sampling, radiance not limited to [0, 1]
        float fx = floorf(u * (float)sc.env_w), fy = floorf(v * (float)sc.env_h);
        int ix = (fx >= 0.0f) ? (fx < (float)sc.env_w ? (int)fx : sc.env_w - 1) : 0;
        int iy = (fy >= 0.0f) ? (fy < (float)sc.env_h ? (int)fy : sc.env_h - 1) : 0;
        const float* p = &sc.env_f32[3 * ((size_t)iy * (size_t)sc.env_w + (size_t)ix)];
        return {p[0], p[1], p[2]};
    }
    return fetch_rgba8(sc.env.data(), sc.env_w, sc.env_h, u, v);
}
inline bool has_env_map(const Scene& sc) { return !sc.env.empty() || !sc.env_f32.empty(); }

// --- NEE mode, importance sampling of the environment map -------------------------------------
// P(texel) ~ Rec.709 luminance of the RGBA8 texel; sums in binary64, rounded once per entry.
void build_env_table(Scene& sc) {
    sc.env_cdf_rows.clear();
    sc.env_cdf_cols.clear();
    if (!sc.nee || !sc.env_use || !has_env_map(sc)) return;
    const int w = sc.env_w, h = sc.env_h;
    sc.env_cdf_rows.resize((size_t)h);
    sc.env_cdf_cols.resize((size_t)w * (size_t)h);
    std::vector<double> row_acc((size_t)h), col_acc((size_t)w);
    double total = 0.0;
    for (int j = 0; j < h; ++j) {
        double acc = 0.0;
        for (int i = 0; i < w; ++i) {
            const size_t k = (size_t)j * (size_t)w + (size_t)i;
            float lum;
            if (!sc.env_f32.empty()) {
                lum = 0.2126f * sc.env_f32[3 * k] + 0.7152f * sc.env_f32[3 * k + 1] + 0.0722f * sc.env_f32[3 * k + 2];
                lum = lum > 0.0f ? lum : 0.0f;
            } else {
                const uint32_t p = sc.env[k];
                const float r = (float)(p & 0xFFu) / 255.0f, g = (float)((p >> 8) & 0xFFu) / 255.0f, b = (float)((p >> 16) & 0xFFu) / 255.0f;
                lum = 0.2126f * r + 0.7152f * g + 0.0722f * b;
            }
            acc += (double)lum;
            col_acc[(size_t)i] = acc;
        }
        float* row = sc.env_cdf_cols.data() + (size_t)j * (size_t)w;
        for (int i = 0; i < w; ++i) row[i] = acc > 0.0 ? (float)(col_acc[(size_t)i] / acc) : (float)(i + 1) / (float)w;
        row[w - 1] = 1.0f;
        total += acc;
        row_acc[(size_t)j] = total;
    }
    if (!(total > 0.0)) {
        sc.env_cdf_rows.clear();
        sc.env_cdf_cols.clear();
        return;
    }
    for (int j = 0; j < h; ++j) sc.env_cdf_rows[(size_t)j] = (float)(row_acc[(size_t)j] / total);
    sc.env_cdf_rows[(size_t)h - 1] = 1.0f;
}
// density (per solid angle) with which the map sampler produces `dir`: probability of the texel the miss
// branch reads for dir, over the texel's solid angle (du dv = dlon dlat / 2 pi^2, d omega = cos(lat) dlon dlat)
float env_pdf(const Scene& sc, V3 dir) {
    const float u = 0.5f + m_atan2(dir.x, dir.z) / (2.0f * pi);
    const float v = 0.5f + m_asin(dir.y) / pi;
    const float fx = floorf(u * (float)sc.env_w), fy = floorf(v * (float)sc.env_h);
    const int ix = (fx >= 0.0f) ? (fx < (float)sc.env_w ? (int)fx : sc.env_w - 1) : 0;
    const int iy = (fy >= 0.0f) ? (fy < (float)sc.env_h ? (int)fy : sc.env_h - 1) : 0;
    const float prow = sc.env_cdf_rows[iy] - (iy ? sc.env_cdf_rows[iy - 1] : 0.0f);
    const float* row = sc.env_cdf_cols.data() + (size_t)iy * (size_t)sc.env_w;
    const float pcol = row[ix] - (ix ? row[ix - 1] : 0.0f);
    const float cl = sqrtf(fmaxf(0.0f, 1.0f - dir.y * dir.y));
    if (cl < 1e-4f) return 0.0f;
    return prow * pcol * ((float)sc.env_w * (float)sc.env_h) / (2.0f * pi * pi * cl);
}
int cdf_find(const float* cdf, int n, float u) {  // first entry with cdf >= u
    int lo = 0, hi = n - 1;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (u <= cdf[mid]) hi = mid;
        else lo = mid + 1;
    }
    return lo;
}

typedef Closest (*Finder)(const Scene&, V3, V3, float, float);

// light table of the NEE mode: every triangle of every mesh whose material emits, in mesh order;
// weight = triangle_area (lights.hpp:15-20) * emission, cdf accumulated in binary64
void build_lights(Scene& sc) {
    sc.lights.clear();
    sc.light_cdf.clear();
    sc.mesh_light_first.assign(sc.meshes.size(), -1);
    if (!sc.nee) return;
    std::vector<double> acc;
    double total = 0.0;
    for (size_t mi = 0; mi < sc.meshes.size(); ++mi) {
        const MeshRec& mesh = sc.meshes[mi];
        const float emission = sc.mats[mesh.material].emission;
        if (!(emission > 0.0f)) continue;
        sc.mesh_light_first[mi] = (int)sc.lights.size();
        for (int prim = 0; prim < (int)(mesh.idx.size() / 3); ++prim) {
            const V3 a = mesh.verts[mesh.idx[3 * prim]], b = mesh.verts[mesh.idx[3 * prim + 1]], c = mesh.verts[mesh.idx[3 * prim + 2]];
            const V3 n = cross(sub(b, a), sub(c, a));
            const float area = 0.5f * sqrtf(dot(n, n));
            const float w = area * emission;
            total += (double)w;
            acc.push_back(total);
            sc.lights.push_back(Scene::Light{(int)mi, prim, area, 0.0f, emission});
        }
    }
    if (!(total > 0.0)) {
        sc.lights.clear();
        sc.mesh_light_first.assign(sc.meshes.size(), -1);
        return;
    }
    sc.light_cdf.resize(sc.lights.size());
    for (size_t i = 0; i < sc.lights.size(); ++i) sc.light_cdf[i] = (float)(acc[i] / total);
    sc.light_cdf.back() = 1.0f;
    for (size_t i = 0; i < sc.lights.size(); ++i) sc.lights[i].prob = sc.light_cdf[i] - (i ? sc.light_cdf[i - 1] : 0.0f);
}

// ---------------------------------------------------------------- device.cu:113-218
V3 trace_path(Scene& sc, V3 org, V3 dir, Rng& rng, Finder find, uint64_t& nrays, uint64_t& nshadow) {
    V3 radiance = v3(0.0f);
    V3 throughput = v3(1.0f);
    int sampled_lobe = LOBE_NONE;
    int nan_retries = 0;
    // next-event estimation mode: light-sampled radiance gathered along the path, and the density
    // with which the BSDF generated the current ray (<= 0: an emitter it reaches counts fully)
    // light sampling draws from the emissive triangles, from the environment map, or (one extra draw) from
    // either with probability 1/2 when the scene has both
    const bool tri_l = sc.nee && !sc.lights.empty(), env_l = sc.nee && !sc.env_cdf_rows.empty();
    const bool nee = tri_l || env_l;
    const float p_sel = (tri_l && env_l) ? 0.5f : 1.0f;
    V3 gathered = v3(0.0f);
    float prev_pdf = -1.0f;
    for (int depth = 0; depth < sc.max_depth; ++depth) {
        Closest h = find(sc, org, dir, t_min, t_max);
        ++nrays;
        if (h.id == 0xFFFFFFFFu) {                                                       // :136-148
            if (sc.env_use && has_env_map(sc))
                radiance = add(radiance, sample_environment(sc, dir));
            else if (sc.env_auto)
                radiance = add(radiance, lerp(v3(1.0f), V3{0.5f, 0.7f, 1.0f}, 0.5f * (dir.y + 1.0f)));
            else
                radiance = add(radiance, sc.env_color);
            radiance = scale(radiance, sc.env_intensity);
            if (env_l && prev_pdf > 0.0f) radiance = scale(radiance, power_heuristic(prev_pdf, p_sel * env_pdf(sc, dir)));
            break;
        }
        const int mesh_index = sc.tri_mesh[h.id];
        const MeshRec& mesh = sc.meshes[mesh_index];
        const int prim = (int)h.id - mesh.first_tri;
        Mat material = sc.mats[mesh.material];                                            // :150-154
        const int i0 = mesh.idx[3 * prim], i1 = mesh.idx[3 * prim + 1], i2 = mesh.idx[3 * prim + 2];
        const float bx = h.u, by = h.v;
        const float bw = 1 - bx - by;
        if (material.emission > 0.0f) {                                                   // :157-161
            if (tri_l && prev_pdf > 0.0f && sc.mesh_light_first[mesh_index] >= 0) {
                const Scene::Light& lt = sc.lights[sc.mesh_light_first[mesh_index] + prim];
                // the area -> solid-angle density needs the GEOMETRIC unit normal of the emitter (the reference's unused
                // sample_triangle, lights.hpp:30-48, takes the interpolated shading normal: biased for smooth emitters)
                V3 nl = normalize(cross(sub(mesh.verts[i1], mesh.verts[i0]), sub(mesh.verts[i2], mesh.verts[i0])));
                float pdf_light = p_sel * (lt.prob * pdf_area_to_solid_angle(1.0f / lt.area, h.t * h.t, dot(neg(dir), nl)));
                radiance = v3(material.emission * power_heuristic(prev_pdf, pdf_light));
            } else {
                radiance = v3(material.emission);
            }
            break;
        }
        V3 v_p = add(add(scale(bw, mesh.verts[i0]), scale(bx, mesh.verts[i1])), scale(by, mesh.verts[i2]));  // :59
        V3 v_n = normalize(add(add(scale(bw, mesh.normals[i0]), scale(bx, mesh.normals[i1])), scale(by, mesh.normals[i2])));  // :72
        if (!mesh.tex.empty()) {                                                          // :75-94
            V2 t0 = mesh.uvs[i0], t1 = mesh.uvs[i1], t2 = mesh.uvs[i2];
            float tu = bw * t0.x + bx * t1.x + by * t2.x;
            float tv = bw * t0.y + bx * t1.y + by * t2.y;
            V3 c = fetch_rgba8(mesh.tex.data(), mesh.tex_w, mesh.tex_h, tu, tv);
            material.base_color[0] = c.x;
            material.base_color[1] = c.y;
            material.base_color[2] = c.z;
        }
        V3 wo = neg(normalize(dir));                                                      // :267-268
        V3 T, B;
        onb(v_n, T, B);
        V3 local_wo = to_local(T, B, v_n, wo);
        if (nee && nan_retries == 0) {
            bool use_env = env_l;
            if (tri_l && env_l) use_env = rng.next() < 0.5f;
            V3 ldir = v3(0.0f), emitted = v3(0.0f);
            float pdf_light = 0.0f, tmax = 0.0f;
            if (use_env) {
                // one texel by luminance, a uniform point in it; density and radiance re-read through the
                // direction -> texel mapping of the miss branch
                const float u0 = rng.next(), u1 = rng.next(), u2 = rng.next(), u3 = rng.next();
                const int iy = cdf_find(sc.env_cdf_rows.data(), sc.env_h, u0);
                const int ix = cdf_find(sc.env_cdf_cols.data() + (size_t)iy * (size_t)sc.env_w, sc.env_w, u1);
                const float lon = (((float)ix + u2) / (float)sc.env_w - 0.5f) * (2.0f * pi);
                const float lat = (((float)iy + u3) / (float)sc.env_h - 0.5f) * pi;
                const float cl = m_cos(lat);
                ldir = normalize(V3{m_sin(lon) * cl, m_sin(lat), m_cos(lon) * cl});
                pdf_light = p_sel * env_pdf(sc, ldir);
                emitted = scale(add(v3(0.0f), sample_environment(sc, ldir)), sc.env_intensity);
                tmax = t_max;
            } else {
                // one emissive triangle by power, one point on it by area (lights.hpp:30-48), a shadow
                // ray over (t_min, distance - t_min) (visibiliy_test, device.cu:97-110)
                const float us = rng.next(), u0 = rng.next(), u1 = rng.next();
                const int lo = cdf_find(sc.light_cdf.data(), (int)sc.lights.size(), us);
                const Scene::Light& lt = sc.lights[lo];
                const MeshRec& lm = sc.meshes[lt.mesh];
                const int j0 = lm.idx[3 * lt.prim], j1 = lm.idx[3 * lt.prim + 1], j2 = lm.idx[3 * lt.prim + 2];
                const float su0 = sqrtf(u0);                                              // sample_methods.hpp:82-86
                const float cx = 1.0f - su0, cy = u1 * su0;
                const float cz = 1.0f - cx - cy;
                V3 pos = add(add(scale(cz, lm.verts[j0]), scale(cx, lm.verts[j1])), scale(cy, lm.verts[j2]));
                V3 nl = normalize(cross(sub(lm.verts[j1], lm.verts[j0]), sub(lm.verts[j2], lm.verts[j0])));   // geometric unit normal
                ldir = sub(pos, v_p);
                const float dist2 = dot(ldir, ldir);
                const float dist = sqrtf(dist2);
                ldir = divs(ldir, dist);
                pdf_light = p_sel * (lt.prob * pdf_area_to_solid_angle(1.0f / lt.area, dist2, dot(neg(ldir), nl)));
                if (!(dist > 2.0f * t_min)) pdf_light = 0.0f;
                emitted = v3(lt.emission);
                tmax = dist - t_min;
            }
            if (pdf_light > 0.0f) {
                V3 lwi = normalize(V3{dot(ldir, T), dot(ldir, B), dot(ldir, v_n)});
                V3 f_eff;
                float pdf_eff;
                eval_mixture(material, local_wo, lwi, f_eff, pdf_eff);
                if (pdf_eff > 0.0f) {
                    const float w = power_heuristic(pdf_light, pdf_eff);
                    V3 c = mul(throughput, mul(f_eff, scale(emitted, fabsf(cos_theta(lwi)) * w / pdf_light)));
                    const bool bad = isinf(c.x) || isinf(c.y) || isinf(c.z) || isnan(c.x) || isnan(c.y) || isnan(c.z);
                    if (!bad && (c.x != 0.0f || c.y != 0.0f || c.z != 0.0f)) {
                        Closest blocker = find(sc, v_p, ldir, t_min, tmax);
                        ++nshadow;
                        if (blocker.id == 0xFFFFFFFFu) gathered = add(gathered, c);
                    }
                }
            }
        }
        V3 local_wi = v3(0.0f);
        float pdf = 0.0f;
        V3 f = sample_disney(material, local_wo, rng, local_wi, pdf, sampled_lobe);
        V3 wi = to_world(T, B, v_n, local_wi);
        if (pdf < 1E-5f) break;                                                           // :193-194
        if (isinf(f.x) || isinf(f.y) || isinf(f.z) || isnan(f.x) || isnan(f.y) || isnan(f.z)) {  // :196-201
            if (++nan_retries >= 64) break;  // the reference would spin; the product gives up after 64 too
            --depth;
            --nrays;  // the product does not re-trace the identical ray
            continue;
        }
        nan_retries = 0;
        throughput = mul(throughput, divs(scale(f, fabsf(cos_theta(local_wi))), pdf));    // :204
        org = v_p;
        dir = wi;
        if (nee) {
            V3 fe;
            float pe;
            eval_mixture(material, local_wo, local_wi, fe, pe);
            prev_pdf = sampled_lobe == LOBE_GLASS ? -1.0f : pe;
        }
        float beta_max = fmaxf(throughput.x, fmaxf(throughput.y, throughput.z));
        if (sampled_lobe != LOBE_GLASS && depth > 3) {                                    // :209-214
            float q = fmaxf(.05f, 1 - beta_max);
            if (rng.next() > q) break;
        }
    }
    if (nee) return add(gathered, mul(radiance, throughput));
    return mul(radiance, throughput);
}

uint32_t quantise8(float f) {  // owl_device.h:79-82; NaN -> 0 as on the device
    float s = f * 256.f;
    if (!(s == s)) return 0;
    if (s >= 255.0f) return 255;
    if (s <= 0.0f) return 0;
    return (uint32_t)(int)s;
}

// device.cu:220-254 for one pixel; returns the SUM over samples (caller scales)
// lum2 (not in the reference): sum over the samples of the squared Rec.709 luminance of the sample value, the second
// moment behind the progressive mode's variance estimate (b200pt_pass_info::rel_variance)
V3 render_pixel(Scene& sc, int x, int y, int seed_row_offset, int nsamples, Finder find, uint64_t& nrays, uint64_t& nshadow,
                float* lum2 = nullptr) {
    Rng rng;
    rng.seed((uint32_t)x, (uint32_t)(y + seed_row_offset));
    V3 color = v3(0.0f);
    float second = 0.0f;
    for (int s = 0; s < nsamples; ++s) {
        float rx = rng.next();
        float ry = rng.next();
        float su = ((float)x + rx) / (float)sc.width;
        float sv = ((float)y + ry) / (float)sc.height;
        V3 dir = normalize(sub(add(add(sc.cam_llc, scale(su, sc.cam_h)), scale(sv, sc.cam_v)), sc.cam_origin));
        const V3 value = trace_path(sc, sc.cam_origin, dir, rng, find, nrays, nshadow);
        color = add(color, value);
        const float l = luminance(value);
        second += l * l;
    }
    if (lum2) *lum2 = second;
    return color;
}

}  // namespace

// ==================================================================== C interface
extern "C" {

struct oracle_scene {
    Scene s;
};

oracle_scene* oracle_create() { return new oracle_scene(); }
void oracle_destroy(oracle_scene* o) { delete o; }
int oracle_uses_detmath() {
#ifdef ORACLE_DETMATH
    return 1;
#else
    return 0;
#endif
}

int oracle_add_mesh(oracle_scene* o, const float* verts, const float* normals, const float* uvs, int nv,
                    const int32_t* idx, int nt, int material, const uint8_t* tex, int tw, int th) {
    MeshRec m;
    m.verts.resize(nv);
    m.normals.resize(nv);
    for (int i = 0; i < nv; ++i) {
        m.verts[i] = {verts[3 * i], verts[3 * i + 1], verts[3 * i + 2]};
        m.normals[i] = {normals[3 * i], normals[3 * i + 1], normals[3 * i + 2]};
    }
    if (uvs) {
        m.uvs.resize(nv);
        for (int i = 0; i < nv; ++i) m.uvs[i] = {uvs[2 * i], uvs[2 * i + 1]};
    }
    m.idx.assign(idx, idx + 3 * (size_t)nt);
    for (int32_t v : m.idx)
        if (v < 0 || v >= nv) return -1;
    m.material = material;
    if (tex && tw > 0 && th > 0) {
        if (!uvs) return -1;
        m.tex.resize((size_t)tw * th);
        memcpy(m.tex.data(), tex, (size_t)tw * th * 4);
        m.tex_w = tw;
        m.tex_h = th;
    }
    o->s.meshes.push_back(std::move(m));
    return (int)o->s.meshes.size() - 1;
}
void oracle_set_materials(oracle_scene* o, const float* m, int nm) {
    o->s.mats.resize(nm);
    memcpy(o->s.mats.data(), m, sizeof(Mat) * (size_t)nm);
}
void oracle_set_environment(oracle_scene* o, const uint8_t* rgba, int w, int h, int use, int auto_,
                            const float* color, float intensity) {
    o->s.env.clear();
    o->s.env_f32.clear();
    o->s.env_w = o->s.env_h = 0;
    if (rgba && w > 0 && h > 0) {
        o->s.env.resize((size_t)w * h);
        memcpy(o->s.env.data(), rgba, (size_t)w * h * 4);
        o->s.env_w = w;
        o->s.env_h = h;
    }
    o->s.env_use = use != 0;
    o->s.env_auto = auto_ != 0;
    o->s.env_color = {color[0], color[1], color[2]};
    o->s.env_intensity = intensity;
}
void oracle_set_environment_f32(oracle_scene* o, const float* rgb, int w, int h, int use, int auto_, const float* color,
                                float intensity) {
    o->s.env.clear();
    o->s.env_f32.clear();
    o->s.env_w = o->s.env_h = 0;
    if (rgb && w > 0 && h > 0) {
        o->s.env_f32.assign(rgb, rgb + (size_t)w * h * 3);
        o->s.env_w = w;
        o->s.env_h = h;
    }
    o->s.env_use = use != 0;
    o->s.env_auto = auto_ != 0;
    o->s.env_color = {color[0], color[1], color[2]};
    o->s.env_intensity = intensity;
}
void oracle_set_camera(oracle_scene* o, const float* origin, const float* llc, const float* hor, const float* ver) {
    o->s.cam_origin = {origin[0], origin[1], origin[2]};
    o->s.cam_llc = {llc[0], llc[1], llc[2]};
    o->s.cam_h = {hor[0], hor[1], hor[2]};
    o->s.cam_v = {ver[0], ver[1], ver[2]};
}
void oracle_set_render(oracle_scene* o, int w, int h, int spp, int depth) {
    o->s.width = w;
    o->s.height = h;
    o->s.max_samples = spp;
    o->s.max_depth = depth;
}
void oracle_build(oracle_scene* o) { build_bvh(o->s); }
// next-event estimation mode on/off (see trace_path); shadow rays of the last render
void oracle_set_nee(oracle_scene* o, int on) { o->s.nee = on != 0; }
// 0: linear (the reference as shipped); 1: color = saturate(pow(color, 1/2.2)), device.cu:248 (commented out there)
void oracle_set_output(oracle_scene* o, int transfer) { o->s.transfer = transfer; }
uint64_t oracle_shadow_rays(oracle_scene* o) { return o->s.shadow_rays.load(); }

// camera.cpp:3-22.  out: origin, llc, horizontal, vertical (12 floats)
void oracle_camera(const float* from, const float* at, const float* up, float vfov, int w, int h, float* out) {
    float aspect = (float)w / (float)h;
    float theta = vfov * pi / 180.0f;
    float hh = m_tan(theta / 2);
    float vh = 2.0f * hh;
    float vw = aspect * vh;
    V3 origin = {from[0], from[1], from[2]};
    V3 wv = normalize(sub(origin, V3{at[0], at[1], at[2]}));
    V3 u = normalize(cross(V3{up[0], up[1], up[2]}, wv));
    V3 v = normalize(cross(wv, u));
    V3 horizontal = scale(vw, u);
    V3 vertical = scale(vh, v);
    V3 llc = sub(sub(sub(origin, divs(horizontal, 2.0f)), divs(vertical, 2.0f)), wv);
    float r[12] = {origin.x, origin.y, origin.z, llc.x, llc.y, llc.z, horizontal.x, horizontal.y, horizontal.z,
                   vertical.x, vertical.y, vertical.z};
    memcpy(out, r, sizeof(r));
}

// Render rows [y0, y1) x columns [x0, x1) of the frame (full-frame coordinates).
//   rank/world/shard_mode: same meaning as b200pt_config (include/b200pt.h).
//   sum_out:  W*H*4 floats, un-normalised sample sums of THIS rank (index x + W*y, y as launched)
//   mean_out: W*H*4 floats, sum * (1/total_spp), row-flipped like the framebuffer (may be NULL)
//   rgba_out: W*H uint32 RGBA8, row-flipped (may be NULL)  [only meaningful for world==1]
// brute != 0 uses the brute-force finder.  Returns radiance segments traced.
uint64_t oracle_render(oracle_scene* o, int rank, int world, int shard_mode, int x0, int y0, int x1, int y1,
                       int brute, int nthreads, float* sum_out, float* mean_out, uint32_t* rgba_out) {
    Scene& sc = o->s;
    const int W = sc.width, H = sc.height;
    Finder find = brute ? closest_brute : closest_bvh;
    int nsamples = sc.max_samples, seed_off = 0, row_first = 0, row_step = 1;
    if (world > 1) {
        if (shard_mode == 0) {
            nsamples = sc.max_samples / world + (rank < sc.max_samples % world ? 1 : 0);
            seed_off = rank * H;
        } else {
            row_first = rank;
            row_step = world;
        }
    }
    if (nthreads < 1) nthreads = 1;
    build_lights(sc);
    build_env_table(sc);
    sc.shadow_rays = 0;
    std::atomic<int> next_row{y0};
    std::atomic<uint64_t> total{0};
    auto worker = [&]() {
        uint64_t nr = 0, ns = 0;
        for (;;) {
            int y = next_row.fetch_add(1);
            if (y >= y1) break;
            if (y < row_first || (y - row_first) % row_step != 0) continue;
            for (int x = x0; x < x1; ++x) {
                float second = 0.0f;
                V3 c = render_pixel(sc, x, y, seed_off, nsamples, find, nr, ns, &second);
                float* d = sum_out + 4 * ((size_t)x + (size_t)W * y);
                d[0] = c.x;
                d[1] = c.y;
                d[2] = c.z;
                d[3] = second;
            }
        }
        total += nr;
        sc.shadow_rays += ns;
    };
    std::vector<std::thread> th;
    for (int i = 1; i < nthreads; ++i) th.emplace_back(worker);
    worker();
    for (auto& t : th) t.join();
    if (mean_out || rgba_out) {
        const float inv = 1.0f / (float)sc.max_samples;                                   // device.cu:247
        for (int y = y0; y < y1; ++y)
            for (int x = x0; x < x1; ++x) {
                const float* s = sum_out + 4 * ((size_t)x + (size_t)W * y);
                V3 c = {s[0] * inv, s[1] * inv, s[2] * inv};
                size_t ofs = (size_t)x + (size_t)W * (H - 1 - y);                         // device.cu:251
                if (mean_out) {
                    mean_out[4 * ofs] = c.x;
                    mean_out[4 * ofs + 1] = c.y;
                    mean_out[4 * ofs + 2] = c.z;
                    mean_out[4 * ofs + 3] = 1.0f;
                }
                if (rgba_out) {
                    if (sc.transfer == 1) {  // the reference's commented-out line, device.cu:248
                        const float g = 1.0f / 2.2f;
                        c = {fminf(fmaxf(m_pow(c.x, g), 0.0f), 1.0f), fminf(fmaxf(m_pow(c.y, g), 0.0f), 1.0f),
                             fminf(fmaxf(m_pow(c.z, g), 0.0f), 1.0f)};
                    }
                    rgba_out[ofs] = (quantise8(c.x) << 0) + (quantise8(c.y) << 8) + (quantise8(c.z) << 16) + (0xffU << 24);
                }
            }
    }
    return total.load();
}

// closest hit of sample 0's camera ray of every pixel (index x + W*y)
void oracle_primary_hits(oracle_scene* o, int brute, int nthreads, HitRec* out) {
    Scene& sc = o->s;
    const int W = sc.width, H = sc.height;
    Finder find = brute ? closest_brute : closest_bvh;
    std::atomic<int> next_row{0};
    auto worker = [&]() {
        for (;;) {
            int y = next_row.fetch_add(1);
            if (y >= H) break;
            for (int x = 0; x < W; ++x) {
                Rng rng;
                rng.seed((uint32_t)x, (uint32_t)y);
                float rx = rng.next(), ry = rng.next();
                float su = ((float)x + rx) / (float)W, sv = ((float)y + ry) / (float)H;
                V3 dir = normalize(sub(add(add(sc.cam_llc, scale(su, sc.cam_h)), scale(sv, sc.cam_v)), sc.cam_origin));
                Closest c = find(sc, sc.cam_origin, dir, t_min, t_max);
                HitRec& h = out[(size_t)x + (size_t)W * y];
                if (c.id == 0xFFFFFFFFu)
                    h = {-1, -1, t_max, 0, 0};
                else {
                    int mi = sc.tri_mesh[c.id];
                    h = {mi, (int)c.id - sc.meshes[mi].first_tri, c.t, c.u, c.v};
                }
            }
        }
    };
    std::vector<std::thread> th;
    for (int i = 1; i < nthreads; ++i) th.emplace_back(worker);
    worker();
    for (auto& t : th) t.join();
}

void oracle_trace(oracle_scene* o, const float* org, const float* dir, int n, float tmin, float tmax, int brute,
                  HitRec* out) {
    Scene& sc = o->s;
    Finder find = brute ? closest_brute : closest_bvh;
    for (int i = 0; i < n; ++i) {
        Closest c = find(sc, V3{org[3 * i], org[3 * i + 1], org[3 * i + 2]}, V3{dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]},
                         tmin, tmax);
        if (c.id == 0xFFFFFFFFu)
            out[i] = {-1, -1, tmax, 0, 0};
        else {
            int mi = sc.tri_mesh[c.id];
            out[i] = {mi, (int)c.id - sc.meshes[mi].first_tri, c.t, c.u, c.v};
        }
    }
}

// same signature as b200pt_sample_disney (include/b200pt.h); materials passed directly
void oracle_sample_disney(const float* mats, int n, const int32_t* material, const float* wo, const int32_t* lobe_in,
                          const uint32_t* rng_in, float* f, float* wi, float* pdf, int32_t* lobe_out, uint32_t* rng_out) {
    const Mat* mm = (const Mat*)mats;
    for (int i = 0; i < n; ++i) {
        Rng r;
        r.state = rng_in[i];
        int lobe = lobe_in[i];
        V3 w_i = v3(0.0f);
        float p = 0.0f;
        V3 fr = sample_disney(mm[material[i]], V3{wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]}, r, w_i, p, lobe);
        f[3 * i] = fr.x;
        f[3 * i + 1] = fr.y;
        f[3 * i + 2] = fr.z;
        wi[3 * i] = w_i.x;
        wi[3 * i + 1] = w_i.y;
        wi[3 * i + 2] = w_i.z;
        pdf[i] = p;
        lobe_out[i] = lobe;
        rng_out[i] = r.state;
    }
}

// RNG known-answer hook: first n draws of pixel (x,y)
void oracle_rng(uint32_t x, uint32_t y, int n, float* out, uint32_t* state_out) {
    Rng r;
    r.seed(x, y);
    for (int i = 0; i < n; ++i) out[i] = r.next();
    if (state_out) *state_out = r.state;
}

// elementary function hook (fn: 0 sin 1 cos 2 tan 3 atan 4 asin 5 log 6 atan2(x,y) 7 pow(x,y))
void oracle_math(int fn, int n, const float* x, const float* y, float* out) {
    for (int i = 0; i < n; ++i) {
        switch (fn) {
            case 0: out[i] = m_sin(x[i]); break;
            case 1: out[i] = m_cos(x[i]); break;
            case 2: out[i] = m_tan(x[i]); break;
            case 3: out[i] = m_atan(x[i]); break;
            case 4: out[i] = m_asin(x[i]); break;
            case 5: out[i] = m_log(x[i]); break;
            case 6: out[i] = m_atan2(x[i], y[i]); break;
            case 7: out[i] = m_pow(x[i], y[i]); break;
            default: out[i] = 0;
        }
    }
}

// math.hpp:86-107 hook: onb of n, then to_local(w) and to_world(w); out = t b local world per case (12 floats)
void oracle_onb(const float* n_in, const float* w_in, int n, float* out) {
    for (int i = 0; i < n; ++i) {
        V3 nn{n_in[3 * i], n_in[3 * i + 1], n_in[3 * i + 2]}, w{w_in[3 * i], w_in[3 * i + 1], w_in[3 * i + 2]};
        V3 t{}, b{};
        onb(nn, t, b);
        V3 v[4] = {t, b, to_local(t, b, nn, w), to_world(t, b, nn, w)};
        for (int k = 0; k < 4; ++k) {
            out[12 * i + 3 * k] = v[k].x;
            out[12 * i + 3 * k + 1] = v[k].y;
            out[12 * i + 3 * k + 2] = v[k].z;
        }
    }
}

// lights.hpp:23-28, :56-61 hooks (dead code in the reference; the NEE mode builds on them)
float oracle_pdf_area_to_solid_angle(float pdf_area, float dist_sqr, float cos_t) {
    return pdf_area_to_solid_angle(pdf_area, dist_sqr, cos_t);
}
float oracle_power_heuristic(float pdf_f, float pdf_g) { return power_heuristic(pdf_f, pdf_g); }

uint32_t oracle_quantise(float r, float g, float b) {
    return (quantise8(r) << 0) + (quantise8(g) << 8) + (quantise8(b) << 16) + (0xffU << 24);
}

}  // extern "C"
