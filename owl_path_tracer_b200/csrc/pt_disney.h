// pt_disney.h — Disney BSDF sampling as the reference integrator does it, written for
// the shade kernel.  All vectors are in the local shading frame (normal = +z).
//
// Behaviour follows path_tracer/src/device/disney/*.cuh of the reference exactly,
// including its quirks: lobe-selection probability is
// not part of the pdf, the metal lobe samples the NDF but reports the VNDF pdf, the
// clearcoat weight only gates, the glass pdf is the Fresnel split, sheen is added
// after every lobe.  Expression order is kept so that results are bit-identical to
// the scalar oracle under the floating point contract of pt_defs.h.
#pragma once
#include "pt_defs.h"
#include "pt_detmath.h"
#include "pt_rng.h"

namespace pt {

// math.hpp:22-38 — local-frame trigonometry
PT_HD float cosT(f3 w) { return w.z; }
PT_HD float sinT(f3 w) { return sqrtf(fmaxf(0.0f, 1.0f - sqr(w.z))); }
PT_HD float tanT(f3 w) { return sinT(w) / w.z; }
PT_HD float cosP(f3 w) {
    float s = sinT(w);
    return (s == 0) ? 1.0f : clampf(w.x / s, -1.0f, 1.0f);
}
PT_HD float sinP(f3 w) {
    float s = sinT(w);
    return (s == 0) ? 1.0f : clampf(w.y / s, -1.0f, 1.0f);
}
PT_HD bool same_side(f3 a, f3 b) { return a.z * b.z > 0.0f; }
// math.hpp:58-61
PT_HD f3 mirror(f3 w, f3 n) { return 2.0f * (dot(w, n) * n) - w; }
// math.hpp:63-77
PT_HD bool bend(f3 w, f3 n, float eta, f3& wi) {
    if (eta == 1.0f) {
        wi = -w;
        return true;
    }
    float ci = dot(w, n);
    float s2i = fmaxf(0.0f, 1.0f - sqr(ci));
    float s2t = eta * eta * s2i;
    if (s2t > 1.0f) return false;
    float ct = sqrtf(1.0f - s2t);
    wi = eta * (-w) + (eta * ci - ct) * n;
    return true;
}
// math.hpp:51-56
PT_HD f3 sph(float st, float ct, float phi) { return f3{st * dm_cos(phi), st * dm_sin(phi), ct}; }

// disney_helper.cuh:19-24
PT_HD float schlick_w(float c) {
    float m = clampf(1.0f - c, 0.0f, 1.0f);
    float m2 = m * m;
    return m2 * m2 * m;
}
PT_HD float lum709(f3 c) { return dot(f3{0.2126f, 0.7152f, 0.0722f}, c); }
// disney_helper.cuh:39-48
PT_HD float alpha_iso(float r) { return fmaxf(kAlphaMin, clampf(sqr(r), 0.0f, 1.0f)); }
PT_HD f2 alpha_aniso(float r, float aniso) {
    float aspect = sqrtf(1.0f - 0.9f * aniso);
    return f2{fmaxf(kAlphaMin, sqr(r) / aspect), fmaxf(kAlphaMin, sqr(r) * aspect)};
}
// disney_helper.cuh:51-60 (dielectric Fresnel, TIR -> 1)
PT_HD float fresnel_diel(f3 i, f3 m, float eta_i, float eta_t) {
    float c = fabsf(dot(i, m));
    float den = sqr(eta_t / eta_i) - 1.0f + sqr(c);
    if (den < 0.0f) return 1.0f;
    float g = sqrtf(den);
    return 0.5f * sqr((g - c) / (g + c)) * (1.0f + sqr(c * (g + c) - 1.0f) / sqr(c * (g - c) + 1.0f));
}

// disney_specular.cuh:17-28
PT_HD float smith_lambda(f3 w, float ax, float ay) {
    float tt = tanT(w);
    if (isinf(tt)) return 0.0f;
    float a0 = sqrtf(sqr(cosP(w) * ax) + sqr(sinP(w) * ay));
    float a = 1.0f / (a0 * tt);
    return (-1.0f + sqrtf(1.0f + 1.0f / sqr(a))) / 2.0f;
}
PT_HD float smith_g1(f3 w, float ax, float ay) { return 1.0f / (1.0f + smith_lambda(w, ax, ay)); }
// disney_specular.cuh:53-60
PT_HD float ggx_d(f3 wm, float ax, float ay) {
    float t2 = sqr(tanT(wm));
    if (isinf(t2)) return 0.0f;
    float c4 = sqr(sqr(wm.z));
    float e = 1.0f + t2 * (sqr(cosP(wm)) / sqr(ax) + sqr(sinP(wm)) / sqr(ay));
    return 1.0f / (kPi * ax * ay * c4 * sqr(e));
}
// disney_specular.cuh:64-81 (note the "+ inv_pi" phase, Q9)
PT_HD f3 ggx_sample_ndf(float ax, float ay, float u0, float u1) {
    float phi = dm_atan(ay / ax * dm_tan(kTwoPi * u1 + kInvPi));
    if (u1 > .5f) phi += kPi;
    float sp = dm_sin(phi), cp = dm_cos(phi);
    float ax2 = sqr(ax), ay2 = sqr(ay);
    float a2 = 1.0f / (sqr(cp) / ax2 + sqr(sp) / ay2);
    float tan2 = a2 * u0 / (1.0f - u0);
    float ct = 1.0f / sqrtf(1.0f + tan2);
    float st = sqrtf(fmaxf(0.f, 1.f - sqr(ct)));
    return normalize(f3{st * cp, st * sp, ct});
}

struct BsdfSample {
    f3 f;
    f3 wi;
    float pdf;
};

// eval_disney_diffuse, disney_diffuse.cuh:26-55 (pdf_cosine_hemisphere, sample_methods.hpp:59-62)
PT_HD f3 eval_diffuse(const Material& m, f3 wo, f3 wi, float& pdf) {
    float fo = schlick_w(wo.z);
    float fi = schlick_w(wi.z);
    f3 lambert = f3{m.base_color[0], m.base_color[1], m.base_color[2]} * kInvPi;
    float fd = (1.0f - 0.5f * fo) * (1.0f - 0.5f * fi);
    float rr = m.roughness * (dot(wo, wi) + 1.0f);
    float fr = rr * (fi + fo + fo * fi * (rr - 1.0f));
    pdf = fabsf(wi.z) * kInvPi;
    return lambert * (fd + fr);
}

// disney_diffuse.cuh:57-62 with sample_methods.hpp:19-57
PT_HD void lobe_diffuse(const Material& m, f3 wo, uint32_t& rng, BsdfSample& s) {
    float ux = rng_next(rng);
    float uy = rng_next(rng);
    float dx = 2.0f * ux - 1.0f;
    float dy = 2.0f * uy - 1.0f;
    float px = 0.0f, py = 0.0f;
    if (!(dx == 0 && dy == 0)) {
        float phi, r;
        if (fabsf(dx) > fabsf(dy)) {
            r = dx;
            phi = kPiOverFour * (dy / dx);
        } else {
            r = dy;
            phi = kPiOverTwo - kPiOverFour * (dx / dy);
        }
        px = r * dm_cos(phi);
        py = r * dm_sin(phi);
    }
    f3 wi{px, py, sqrtf(fmaxf(0.0f, 1.0f - sqr(px) - sqr(py)))};
    s.wi = wi;
    s.f = eval_diffuse(m, wo, wi, s.pdf);
}

// eval_disney_specular_brdf, disney_specular.cuh:125-149
PT_HD f3 eval_metal(const Material& m, f3 wo, f3 wh, f3 wi, float& pdf) {
    f2 a = alpha_aniso(m.roughness, m.anisotropic);
    f3 base{m.base_color[0], m.base_color[1], m.base_color[2]};
    float lum = lum709(base);
    f3 tint = lum > 0.0f ? base / lum : mk3(1.0f);
    f3 cspec = lerp3(.08f * m.specular * lerp3(mk3(1.0f), tint, m.specular_tint), base, m.metallic);
    float d = ggx_d(wh, a.x, a.y);
    float g = 1.0f / (1.0f + smith_lambda(wo, a.x, a.y) + smith_lambda(wi, a.x, a.y));
    f3 fr = lerp3(cspec, mk3(1.0f), schlick_w(dot(wi, wh)));
    pdf = d * smith_g1(wo, a.x, a.y) * fmaxf(0.0f, dot(wo, wh)) / (4.0f * wo.z);
    return d * g * fr / (4.0f * fabsf(wo.z));
}

// disney_specular.cuh:151-170
PT_HD void lobe_metal(const Material& m, f3 wo, uint32_t& rng, BsdfSample& s) {
    f2 a = alpha_aniso(m.roughness, m.anisotropic);
    float u0 = rng_next(rng);
    float u1 = rng_next(rng);
    f3 wh = ggx_sample_ndf(a.x, a.y, u0, u1);
    if (dot(wo, wh) < 0.0f) wh = -wh;
    f3 wi = mirror(wo, wh);
    s.wi = wi;
    if (wi.z <= 0.0f) {
        s.pdf = 0.0f;
        s.f = mk3(0.0f);
        return;
    }
    s.f = eval_metal(m, wo, wh, wi, s.pdf);
}

// eval_disney_clearcoat (m.clearcoat > 0 checked by the caller), disney_clearcoat.cuh:13-20,45-59
PT_HD f3 eval_clearcoat(const Material& m, f3 wo, f3 wh, f3 wi, float& pdf) {
    float a = lerpf(0.1f, .001f, m.clearcoat_gloss);
    float a2 = sqr(a);
    float d;
    if (a >= 1)
        d = kInvPi;
    else
        d = (a2 - 1.0f) / (kPi * dm_log(a2) * (1.0f + (a2 - 1.0f) * sqr(wh.z)));
    float fr = lerpf(1.0f, schlick_w(wi.z), 0.04f);
    float g = smith_g1(wo, .25f, .25f) * smith_g1(wi, .25f, .25f);
    pdf = d / (4.0f * dot(wh, wi));
    return mk3(d * g * fr / (4.0f * fabsf(wo.z) * fabsf(wi.z)));
}

// disney_clearcoat.cuh:22-35,61-78
PT_HD void lobe_clearcoat(const Material& m, f3 wo, uint32_t& rng, BsdfSample& s) {
    float a = lerpf(0.1f, .001f, m.clearcoat_gloss);
    float u0 = rng_next(rng);
    float u1 = rng_next(rng);
    float a2 = sqr(a);
    float ct = sqrtf(fmaxf(0.0f, (1.0f - dm_pow(a2, 1.0f - u0)) / (1.0f - a2)));
    float st = sqrtf(fmaxf(0.0f, 1.0f - sqr(ct)));
    float phi = kTwoPi * u1;
    f3 wh = sph(st, ct, phi);
    if (!same_side(wo, wh)) wh = -wh;
    if (dot(wh, wo) < 0.0f) wh = -wh;
    wh = normalize(wh);
    f3 wi = mirror(wo, wh);
    s.wi = wi;
    if (!same_side(wo, wi) || m.clearcoat <= 0.0f) {
        s.pdf = 0.0f;
        s.f = mk3(0.0f);
        return;
    }
    s.f = eval_clearcoat(m, wo, wh, wi, s.pdf);
}

// disney_specular.cuh:175-244
PT_HD void lobe_glass(const Material& m, f3 wo, uint32_t& rng, BsdfSample& s) {
    float a = alpha_iso(m.specular_transmission_roughness);
    float u0 = rng_next(rng);
    float u1 = rng_next(rng);
    float theta = dm_atan((a * sqrtf(u0)) / sqrtf(1.0f - u0));
    float phi = kTwoPi * u1;
    f3 wh{dm_sin(theta) * dm_cos(phi), dm_sin(theta) * dm_sin(phi), dm_cos(theta)};
    if (wo.z < 0.0f && !same_side(wo, wh)) wh = -wh;

    float eta_i = wo.z > 0.0f ? 1.0f : m.ior;
    float eta_t = wo.z > 0.0f ? m.ior : 1.0f;
    float eta = eta_i / eta_t;
    float R = fresnel_diel(wo, wh, eta_i, eta_t);
    float T = 1.0f - R;
    f3 wi = s.wi;
    bool reflect_it = !bend(wo, wh, eta, wi);
    if (!reflect_it) reflect_it = rng_next(rng) < R / (R + T);
    if (reflect_it) {
        f2 aa = alpha_aniso(m.roughness, m.anisotropic);
        float v0 = rng_next(rng);
        float v1 = rng_next(rng);
        wh = ggx_sample_ndf(aa.x, aa.y, v0, v1);
        wi = normalize(mirror(wo, wh));
    }
    // eval_disney_specular_bsdf with the (possibly new) half vector
    R = fresnel_diel(wo, wh, eta_i, eta_t);
    T = 1.0f - R;
    f3 base{m.base_color[0], m.base_color[1], m.base_color[2]};
    s.wi = wi;
    if (same_side(wo, wi)) {
        s.pdf = R / (R + T);
        s.f = base * R / fabsf(wi.z);
    } else {
        s.pdf = T / (R + T);
        s.f = (f3{sqrtf(base.x), sqrtf(base.y), sqrtf(base.z)} * T / fabsf(wi.z)) / sqr(eta);
    }
}

// disney_sheen.cuh:15-37
PT_HD f3 sheen_term(const Material& m, f3 wo, f3 wi) {
    if (m.sheen <= 0.0f) return mk3(0.0f);
    f3 wh = wi + wo;
    if (wh.x == 0.0f && wh.y == 0.0f && wh.z == 0.0f) return mk3(0.0f);
    wh = normalize(wh);
    f3 base{m.base_color[0], m.base_color[1], m.base_color[2]};
    float lum = lum709(f3{dm_pow(base.x, 2.2f), dm_pow(base.y, 2.2f), dm_pow(base.z, 2.2f)});
    float cd = dot(wi, wh);
    f3 tint = (lum > 0.0f) ? base / lum : mk3(1.0f);
    return lerp3(mk3(1.0f), tint, m.sheen_tint) * m.sheen * schlick_w(cd);
}

// disney.cuh:15-66.  `lobe` is read (force_btdf) and updated.  s.wi must come in as
// the caller's initial value (zero vector, device.cu:181) — it is what survives when
// no branch runs (all weights NaN).
PT_HD void sample_disney(const Material& m, f3 wo, uint32_t& rng, int& lobe, BsdfSample& s) {
    float dw = (1.0f - m.specular_transmission) * (1.0f - m.metallic);
    float mw = m.metallic;
    float cw = .25f * m.clearcoat;
    float gw = (1.0f - m.metallic) * m.specular_transmission;
    float factor = 1.0f / (mw + gw + dw + cw);
    float pm = mw * factor;
    float pg = gw * factor;
    float pd = dw * factor;
    float pc = cw * factor;

    bool force = wo.z < 0.0f && lobe == LOBE_GLASS;
    float p = rng_next(rng);
    s.f = mk3(0.0f);
    if (!force && p <= pm) {
        lobe_metal(m, wo, rng, s);
        lobe = LOBE_METALLIC;
    } else if (!force && p > pm && p <= (pm + pc)) {
        lobe_clearcoat(m, wo, rng, s);
        lobe = LOBE_CLEARCOAT;
    } else if (!force && p > pm + pc && p <= (pm + pc + pd)) {
        lobe_diffuse(m, wo, rng, s);
        lobe = LOBE_DIFFUSE;
    } else if (force || pg >= 0.0f) {
        lobe_glass(m, wo, rng, s);
        lobe = LOBE_GLASS;
    }
    s.f = s.f + sheen_term(m, wo, s.wi);
}

// ---- next-event estimation support (B200PT_FLAG_NEE; no counterpart in the reference's integrator,
// built from its eval_* functions).  The reference's estimator picks ONE lobe l with probability
// p_l and weights (f_l + sheen) / pdf_l, so in expectation it integrates
//     f_eff(wo, wi) = sum_l p_l (f_l(wo, wi) + sheen(wo, wi))
// and generates directions with density pdf_eff = sum_l p_l pdf_l.  eval_mixture returns both for a
// GIVEN wi over the three lobes a light sample can reach (metal, clearcoat, diffuse); the glass lobe
// reports a probability instead of a density (disney_specular.cuh:204-213) and is treated as
// specular: it is left out here and emitters found through it count fully.
// The hemisphere conditions are those under which each sampler returns a non-zero sample.
PT_HD void eval_mixture(const Material& m, f3 wo, f3 wi, f3& f_eff, float& pdf_eff) {
    float dw = (1.0f - m.specular_transmission) * (1.0f - m.metallic);
    float mw = m.metallic;
    float cw = .25f * m.clearcoat;
    float gw = (1.0f - m.metallic) * m.specular_transmission;
    float factor = 1.0f / (mw + gw + dw + cw);
    float pm = mw * factor;
    float pd = dw * factor;
    float pc = cw * factor;
    f_eff = mk3(0.0f);
    pdf_eff = 0.0f;
    f3 wh = wo + wi;
    const bool has_wh = !(wh.x == 0.0f && wh.y == 0.0f && wh.z == 0.0f);
    if (has_wh) wh = normalize(wh);
    const f3 sheen = sheen_term(m, wo, wi);
    if (pm > 0.0f && has_wh && wi.z > 0.0f) {          // lobe_metal: rejects cos(wi) <= 0
        float pdf;
        f3 f = eval_metal(m, wo, wh, wi, pdf);
        f_eff = f_eff + pm * (f + sheen);
        pdf_eff = pdf_eff + pm * pdf;
    }
    if (pc > 0.0f && has_wh && same_side(wo, wi)) {    // lobe_clearcoat: rejects the other hemisphere
        float pdf;
        f3 f = eval_clearcoat(m, wo, wh, wi, pdf);
        f_eff = f_eff + pc * (f + sheen);
        pdf_eff = pdf_eff + pc * pdf;
    }
    if (pd > 0.0f && wi.z > 0.0f) {                    // cosine hemisphere: only generates cos(wi) >= 0
        float pdf;
        f3 f = eval_diffuse(m, wo, wi, pdf);
        f_eff = f_eff + pd * (f + sheen);
        pdf_eff = pdf_eff + pd * pdf;
    }
}

// power_heuristic(1, a, 1, b), lights.hpp:56-61
PT_HD float power_heuristic(float pdf_f, float pdf_g) {
    float f = pdf_f, g = pdf_g;
    return (f * f) / (f * f + g * g);
}

}  // namespace pt
