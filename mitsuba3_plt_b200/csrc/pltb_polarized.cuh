// pltb_polarized.cuh -- the `*_spectral_polarized` variants of the plt path: Mueller-matrix arithmetic and the
// polarized branches of the BSDFs, per wavelength.
//
// In these variants Spectrum = MuellerMatrix<Spectrum<float, 4>> (include/mitsuba/core/spectrum.h:60-95): every BSDF weight is a
// 4x4 matrix per wavelength.  All matrix operations of the reference act independently per wavelength, so the kernels work on
// one wavelength at a time with a 4x4 float matrix (F44) instead of holding 64 floats; products follow Dr.Jit's fmadd order
// (ext/drjit/include/drjit/matrix.h:105-121, src/python/base.cpp:774-786), including the structurally zero terms, so that the
// arithmetic (and its 0 * inf behaviour) is the reference's.  Complex arithmetic: ext/drjit/include/drjit/complex.h:61-185;
// fresnel_polarized: include/mitsuba/render/fresnel.h:147-275; mueller.h: include/mitsuba/render/mueller.h:36-405.
// The integrator only ever reads the FIRST COLUMN of the path matrices (stokes.cpp:108-121 reads spec.entry(i, 0), the film
// reads spec(0, 0)), and column 0 of a product is the product with column 0: the wavefront state carries 4-vectors (C4), only
// the per-vertex replay weights W_i are stored as full matrices.
#pragma once
#include "pltb_bsdf.cuh"

namespace pltb {

struct Cx { float re, im; };
PLTB_HD Cx cx(float re, float im = 0.f) { Cx r; r.re = re; r.im = im; return r; }
PLTB_HD Cx cx_mul(Cx a, Cx b) { return cx(fmsub(a.re, b.re, a.im * b.im), fmadd(a.re, b.im, a.im * b.re)); }
PLTB_HD Cx cx_scale(Cx a, float s) { return cx(a.re * s, a.im * s); }
PLTB_HD Cx cx_conj(Cx a) { return cx(a.re, -a.im); }
PLTB_HD float cx_sqnorm(Cx a) { return fmadd(a.im, a.im, a.re * a.re); }
PLTB_HD float cx_abs(Cx a) { return sqrtf(cx_sqnorm(a)); }
PLTB_HD Cx cx_rcp(Cx a) { return cx_scale(cx_conj(a), rcp(cx_sqnorm(a))); }
PLTB_HD Cx cx_div(Cx a, Cx b) { return cx_mul(a, cx_rcp(b)); }
PLTB_HD Cx cx_add(Cx a, Cx b) { return cx(a.re + b.re, a.im + b.im); }
PLTB_HD Cx cx_sub(Cx a, Cx b) { return cx(a.re - b.re, a.im - b.im); }
PLTB_HD Cx cx_sqrt(Cx z) {
    const float n = cx_abs(z), t1 = sqrtf(.5f * (n + fabsf(z.re))), t2 = .5f * z.im / t1;
    const bool zero = n == 0.f, m = z.re >= 0.f;
    return cx(m ? t1 : fabsf(t2), zero ? 0.f : (m ? t2 : copysignf(t1, z.im)));
}

struct F44 { float m[4][4]; };
struct C4 { float v[4]; };   // one wavelength of a Stokes-like column
PLTB_HD F44 f44_zero() { F44 r; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i][j] = 0.f; return r; }
PLTB_HD F44 f44_diag(float v) { F44 r = f44_zero(); for (int i = 0; i < 4; ++i) r.m[i][i] = v; return r; }
PLTB_HD F44 f44_depolarizer(float v) { F44 r = f44_zero(); r.m[0][0] = v; return r; }
PLTB_HD F44 f44_mul(const F44 &a, const F44 &b) {
    F44 r;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            float v = a.m[i][0] * b.m[0][c];
#pragma unroll
            for (int j = 1; j < 4; ++j) v = fmadd(a.m[i][j], b.m[j][c], v);
            r.m[i][c] = v;
        }
    return r;
}
PLTB_HD C4 f44_apply(const F44 &a, const C4 &x) {   // column 0 of a * X
    C4 r;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float v = a.m[i][0] * x.v[0];
#pragma unroll
        for (int j = 1; j < 4; ++j) v = fmadd(a.m[i][j], x.v[j], v);
        r.v[i] = v;
    }
    return r;
}
PLTB_HD F44 f44_scale(const F44 &a, float s) { F44 r; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i][j] = a.m[i][j] * s; return r; }
PLTB_HD F44 f44_transpose(const F44 &a) { F44 r; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i][j] = a.m[j][i]; return r; }

struct FresnelPol { Cx a_s, a_p; float cos_theta_t, eta_it, eta_ti; };
PLTB_HD FresnelPol fresnel_polarized(float cos_theta_i, float eta) {   // fresnel.h:147-199
    const bool outside = cos_theta_i >= 0.f;
    const float rcp_eta = rcp(eta), eta_it = outside ? eta : rcp_eta, eta_ti = outside ? rcp_eta : eta;
    const float cos_theta_t_sqr = fnmadd(fnmadd(cos_theta_i, cos_theta_i, 1.f), eta_ti * eta_ti, 1.f);
    const float cos_theta_i_abs = fabsf(cos_theta_i);
    Cx cos_theta_t = cx_sqrt(cx(cos_theta_t_sqr, 0.f));
    cos_theta_t = cx(mulsign(cos_theta_t.re, cos_theta_t_sqr), mulsign(cos_theta_t.im, cos_theta_t_sqr));
    const Cx ec = cx_scale(cos_theta_t, eta_it);
    FresnelPol f;
    f.a_s = cx_div(cx_sub(cx(cos_theta_i_abs), ec), cx_add(cx(cos_theta_i_abs), ec));
    f.a_p = cx_div(cx_sub(cx(eta_it * cos_theta_i_abs), cos_theta_t), cx_add(cx(eta_it * cos_theta_i_abs), cos_theta_t));
    if (eta == 1.f || eta == 0.f) { f.a_s = cx(0.f); f.a_p = cx(0.f); }
    f.cos_theta_t = cos_theta_t_sqr >= 0.f ? mulsign_neg(cos_theta_t.re, cos_theta_i) : 0.f;
    f.eta_it = eta_it; f.eta_ti = eta_ti;
    return f;
}
PLTB_HD FresnelPol fresnel_polarized(float cos_theta_i, Cx eta) {      // fresnel.h:201-250
    const bool outside = cos_theta_i >= 0.f;
    if (eta.im > 0.f) eta = cx_conj(eta);
    const Cx rcp_eta = cx_rcp(eta), eta_it = outside ? eta : rcp_eta, eta_ti = outside ? rcp_eta : eta;
    const Cx cos_theta_t_sqr = cx_sub(cx(1.f), cx_scale(cx_mul(eta_ti, eta_ti), fnmadd(cos_theta_i, cos_theta_i, 1.f)));
    const float cos_theta_i_abs = fabsf(cos_theta_i);
    Cx cos_theta_t = cx_sqrt(cos_theta_t_sqr);
    if (cos_theta_t.im > 0.f) cos_theta_t = cx_conj(cos_theta_t);
    const Cx ec = cx_mul(eta_it, cos_theta_t), ei = cx_scale(eta_it, cos_theta_i_abs);
    FresnelPol f;
    f.a_s = cx_div(cx_sub(cx(cos_theta_i_abs), ec), cx_add(cx(cos_theta_i_abs), ec));
    f.a_p = cx_div(cx_sub(ei, cos_theta_t), cx_add(ei, cos_theta_t));
    const bool index_matched = cx_sqnorm(eta) == 1.f && eta.im == 0.f, invalid = cx_sqnorm(eta) == 0.f;
    if (index_matched || invalid) { f.a_s = cx(0.f); f.a_p = cx(0.f); }
    f.cos_theta_t = cos_theta_t_sqr.re >= 0.f ? mulsign_neg(cos_theta_t.re, cos_theta_i) : 0.f;
    f.eta_it = f.eta_ti = 0.f;
    return f;
}

PLTB_HD F44 mueller_reflection(const FresnelPol &f) {   // mueller::specular_reflection, mueller.h:197-226
    const float normalization = rsqrt_(cx_sqnorm(f.a_p) * cx_sqnorm(f.a_s));   // sincos_arg_diff, complex.h:159-164
    const Cx v = cx_scale(cx_mul(f.a_p, cx_conj(f.a_s)), normalization);
    float sin_delta = v.im, cos_delta = v.re;
    const float r_s = cx_abs(cx_mul(f.a_s, f.a_s)), r_p = cx_abs(cx_mul(f.a_p, f.a_p));
    const float a = 0.5f * (r_s + r_p), b = 0.5f * (r_s - r_p), c = sqrtf(r_s * r_p);
    if (c == 0.f) { sin_delta = 0.f; cos_delta = 0.f; }
    F44 r = f44_zero();
    r.m[0][0] = a; r.m[0][1] = b; r.m[1][0] = b; r.m[1][1] = a;
    r.m[2][2] = c * cos_delta; r.m[2][3] = -c * sin_delta; r.m[3][2] = c * sin_delta; r.m[3][3] = c * cos_delta;
    return r;
}
PLTB_HD F44 mueller_transmission(float cos_theta_i, float eta) {   // mueller::specular_transmission, mueller.h:237-266
    const FresnelPol f = fresnel_polarized(cos_theta_i, eta);
    const float factor = -f.eta_it * (fabsf(cos_theta_i) > 1e-8f ? f.cos_theta_t / cos_theta_i : 0.f);
    const float a_s_r = 1.f + f.a_s.re, a_p_r = (1.f + f.a_p.re) * f.eta_ti;
    const float t_s = sqr(a_s_r), t_p = sqr(a_p_r);
    const float a = 0.5f * factor * (t_s + t_p), b = 0.5f * factor * (t_s - t_p), c = factor * sqrtf(t_s * t_p);
    F44 r = f44_zero();
    r.m[0][0] = a; r.m[0][1] = b; r.m[1][0] = b; r.m[1][1] = a; r.m[2][2] = c; r.m[3][3] = c;
    return r;
}
PLTB_HD float pdf_rgb_spectrum(float wl) {   // include/mitsuba/core/spectrum.h:463-467
    const float tmp = rcp(coshf(0.0072f * (wl - 538.f)));
    return (wl >= 360.f && wl <= 830.f) ? 0.003939804229326285f * tmp * tmp : 0.f;
}
PLTB_HD V3 stokes_basis(V3 forward) { V3 s, t; coordinate_system(forward, s, t); return s; }   // mueller.h:284-287
// mueller::rotator (mueller.h:163-172) of the angle of mueller::rotate_stokes_basis (mueller.h:312-323) is held as (sin 2 theta,
// cos 2 theta).  The frame alignment around a reflection / transmission matrix (conductor.cpp:285-300, dielectric.cpp:339-357, roughconductor.cpp:
// 285-300, roughgrating.cpp:541-560): the two rotators of mueller::rotate_mueller_basis (mueller.h:358-371) do not depend on the
// wavelength, so they are built once per vertex; align() then forms R_out * M * transpose(R_in) for one wavelength.
struct StokesFrames {
    float s_out, c_out, s_in, c_in;   // sin / cos of twice the two rotation angles
    PLTB_HD static void angle(V3 forward, V3 basis_current, V3 basis_target, float &s, float &c) {   // rotate_stokes_basis + rotator
        float theta = unit_angle(normalize(basis_current), normalize(basis_target));
        if (dot(forward, cross(basis_current, basis_target)) < 0) theta *= -1.f;
        s = sinf(2.f * theta); c = cosf(2.f * theta);
    }
    PLTB_HD void init(V3 n, V3 wo_hat, V3 wi_hat) {
        V3 s_axis_in = cross(n, -wo_hat), s_axis_out = cross(n, wi_hat);
        const bool collinear = s_axis_in.x == 0.f && s_axis_in.y == 0.f && s_axis_in.z == 0.f;
        s_axis_in = collinear ? mk3(1, 0, 0) : normalize(s_axis_in);
        s_axis_out = collinear ? mk3(1, 0, 0) : normalize(s_axis_out);
        angle(-wo_hat, s_axis_in, stokes_basis(-wo_hat), s_in, c_in);
        angle(wi_hat, s_axis_out, stokes_basis(wi_hat), s_out, c_out);
    }
    PLTB_HD static F44 rotator(float s, float c) { F44 r = f44_diag(1.f); r.m[1][1] = c; r.m[1][2] = s; r.m[2][1] = -s; r.m[2][2] = c; return r; }
    PLTB_HD F44 align(const F44 &m) const { return f44_mul(f44_mul(rotator(s_out, c_out), m), f44_transpose(rotator(s_in, c_in))); }
};

// ---- polarized BSDF weights, one wavelength at a time ----------------------------------------------------------------------------
// Everything that does not depend on the Spectrum type (sampled direction, pdf, flags, the unpolarized weight of the
// depolarizing BSDFs) comes from the unpolarized code; these builders add the Mueller structure.
struct PolVertex {
    // prepared once per vertex by pol_prepare(), then queried per wavelength
    uint32_t type;
    bool flipped;          // twosided back side: wi / wo were flipped before the nested BSDF saw them
    V3 wi, wo;             // as the nested BSDF sees them
    bool sample_active;    // the sampled weight is non-zero (same predicate as the unpolarized code)
    S4 eta, k, refl, trans;   // conductor eta / k, specular_reflectance, specular_transmittance (1 when absent)
    StokesFrames fs, fw;   // frames of the sampled weight / of the replay weight
    float cos_s, cos_w;    // cosines handed to specular_reflection for the two
    S4 w_scalar;           // unpolarized factor of the sampled weight (roughconductor / roughgrating: G1 [* specular_reflectance])
    float intensity, multiplier;   // roughgrating: sampled lobe intensity and the `multiplier` parameter
    S4 u_weight;           // unpolarized sampled weight (diffuse, principled)
    float sample_pdf;      // dielectric: bs.pdf
    bool selected_r;       // dielectric
    float eta_ti2;         // dielectric: square(eta_ti) of the unpolarized Fresnel term
    // replay weight (wbsdf_weight)
    bool weight_active; float weight_pdf; S4 weight_val;   // rough*: eval_pdf ratio pieces; principled: eval, pdf
    float wpdf;            // dielectric
};

PLTB_HD F44 pol_conductor_matrix(const PolVertex &p, const StokesFrames &f, float cos_theta, int k) {
    return f.align(mueller_reflection(fresnel_polarized(cos_theta, cx(p.eta.v[k], p.k.v[k]))));
}

// BSDF::sample weight for wavelength k (diffuse.cpp:122-124, conductor.cpp:273-302, dielectric.cpp:308-396, roughconductor.cpp:
// 270-310, roughgrating.cpp:523-593, principled.cpp:486,708)
PLTB_HD F44 pol_sample_weight(const PolVertex &p, int k) {
    switch (p.type) {
        case PLTB_BSDF_DIFFUSE:
        case PLTB_BSDF_PRINCIPLED: return f44_depolarizer(p.u_weight.v[k]);
        case PLTB_BSDF_CONDUCTOR:
            if (!p.sample_active) return f44_zero();
            return f44_mul(pol_conductor_matrix(p, p.fs, p.cos_s, k), f44_diag(p.refl.v[k]));
        case PLTB_BSDF_DIELECTRIC: {
            F44 w = p.selected_r ? mueller_reflection(fresnel_polarized(p.cos_s, p.eta.v[0])) : mueller_transmission(p.cos_s, p.eta.v[0]);
            w = f44_scale(w, rcp(p.sample_pdf));
            w = p.fs.align(w);
            w = f44_mul(w, f44_diag(p.selected_r ? p.refl.v[k] : p.trans.v[k]));
            if (!p.selected_r) w = f44_scale(w, p.eta_ti2);
            return w;
        }
        case PLTB_BSDF_ROUGHCONDUCTOR:
            if (!p.sample_active) return f44_zero();
            return f44_scale(pol_conductor_matrix(p, p.fs, p.cos_s, k), p.w_scalar.v[k]);
        case PLTB_BSDF_ROUGHGRATING:   // (F * weight * intensity * multiplier): three successive elementwise products (roughgrating.cpp:583)
            if (!p.sample_active) return f44_zero();
            return f44_scale(f44_scale(f44_scale(pol_conductor_matrix(p, p.fs, p.cos_s, k), p.w_scalar.v[k]), p.intensity), p.multiplier);
        default: return f44_zero();
    }
}

// wbsdf_weight for wavelength k (diffuse.cpp:182-200: a *scaled identity*; conductor.cpp:320-376; dielectric.cpp:522-565 with its
// swapped directions and the "nan check" select against the identity; base class: eval_pdf ratio, bsdf.cpp:85-97)
PLTB_HD F44 pol_replay_weight(const PolVertex &p, int k) {
    switch (p.type) {
        case PLTB_BSDF_DIFFUSE: return f44_diag(p.refl.v[k]);
        case PLTB_BSDF_CONDUCTOR:
            if (!(p.wi.z > 0.f)) return f44_zero();
            return f44_mul(pol_conductor_matrix(p, p.fw, p.cos_w, k), f44_diag(p.refl.v[k]));
        case PLTB_BSDF_DIELECTRIC: {
            F44 w = p.wo.z > 0.0f ? mueller_reflection(fresnel_polarized(p.cos_w, p.eta.v[0])) : mueller_transmission(p.cos_w, p.eta.v[0]);
            w = f44_scale(w, rcp(p.wpdf));
            // dr::select(weight > 1.0f, weight, 1.0): both scalars convert to the identity matrix (matrix.h:68-73)
            for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { const float id = i == j ? 1.0f : 0.0f; w.m[i][j] = w.m[i][j] > id ? w.m[i][j] : id; }
            w = p.fw.align(w);
            return f44_mul(w, f44_diag(p.wo.z > 0.0f ? p.refl.v[k] : p.trans.v[k]));
        }
        case PLTB_BSDF_PRINCIPLED:
            return p.weight_pdf > 0.0f ? f44_scale(f44_depolarizer(p.weight_val.v[k]), rcp(p.weight_pdf)) : f44_zero();
        case PLTB_BSDF_ROUGHCONDUCTOR:
        case PLTB_BSDF_ROUGHGRATING:
            if (!(p.weight_active && p.weight_pdf > 0.0f)) return f44_zero();
            return f44_scale(f44_scale(pol_conductor_matrix(p, p.fw, p.cos_w, k), p.weight_val.v[k]), rcp(p.weight_pdf));
        default: return f44_zero();
    }
}

// Prepares both weights of a vertex.  `so` is the unpolarized wbsdf_sample result for the same inputs (direction, pdf, flags and
// -- for the depolarizing BSDFs -- the weight itself); s2 re-derives the microfacet normal of the rough BSDFs.
PLTB_NI void pol_prepare(const DeviceScene &sc, const pltb_bsdf *b, V3 wi, V2 uv, const S4 &wl, V2 s2, const BSDFSampleOut &so,
                         PolVertex &p) {
    p.type = b ? b->type : 0u;
    p.sample_active = false; p.weight_active = false; p.weight_pdf = 0.f; p.flipped = false;
    if (!b) return;
    V3 wo = so.wo;
    if (b->flags & PLTB_BSDF_FLAG_TWOSIDED) { p.flipped = wi.z < 0.f; wo.z = mulsign(wo.z, wi.z); wi.z = fabsf(wi.z); }
    p.wi = wi; p.wo = wo;
    p.u_weight = so.weight;
    p.refl = s4(1.f); p.trans = s4(1.f); p.eta = s4(0.f); p.k = s4(0.f);
    const float cos_theta_i = wi.z;
    switch (p.type) {
        case PLTB_BSDF_DIFFUSE:
            p.refl = tex_eval(sc, b->tex[0], wl, uv);   // replay: diag(reflectance); sample: depolarizer(so.weight)
            return;
        case PLTB_BSDF_PRINCIPLED: {
            p.weight_pdf = principled_pdf(*b, wi, wo);
            p.weight_val = p.weight_pdf > 0.0f ? principled_eval(sc, *b, wi, uv, wl, wo) : s4(0.f);
            return;
        }
        case PLTB_BSDF_CONDUCTOR:
            p.eta = tex_eval(sc, b->tex[0], wl, uv); p.k = tex_eval(sc, b->tex[1], wl, uv); p.refl = tex_eval(sc, b->tex[2], wl, uv);
            p.sample_active = cos_theta_i > 0.f;
            p.fs.init(mk3(0, 0, 1), wo, wi); p.cos_s = wo.z;
            p.fw = p.fs; p.cos_w = wo.z;
            return;
        case PLTB_BSDF_DIELECTRIC: {
            p.eta = s4(b->f[0]);
            if (b->tex[2] >= 0) p.refl = tex_eval(sc, b->tex[2], wl, uv);
            if (b->tex[3] >= 0) p.trans = tex_eval(sc, b->tex[3], wl, uv);
            const FresnelResult f = fresnel(cos_theta_i, b->f[0]);
            p.selected_r = (so.type & F_DeltaReflection) != 0;
            p.sample_pdf = so.pdf; p.eta_ti2 = sqr(f.eta_ti);
            p.fs.init(mk3(0, 0, 1), wo, wi); p.cos_s = wo.z;
            p.wpdf = wo.z > 0.0f ? f.r : 1.f - f.r;
            p.fw.init(mk3(0, 0, 1), wi, wo); p.cos_w = wi.z;   // dielectric.cpp:522-524: wo_hat = si.wi, wi_hat = wo
            return;
        }
        case PLTB_BSDF_ROUGHCONDUCTOR:
        case PLTB_BSDF_ROUGHGRATING: {
            p.eta = tex_eval(sc, b->tex[0], wl, uv); p.k = tex_eval(sc, b->tex[1], wl, uv);
            const S4 refl = b->tex[2] >= 0 ? tex_eval(sc, b->tex[2], wl, uv) : s4(1.f);
            Microfacet distr = make_mf(*b);
            if (cos_theta_i > 0.f) {
                float pdf; const V3 m = distr.sample(wi, s2, pdf);
                const V3 wo_hat = p.type == PLTB_BSDF_ROUGHGRATING ? reflect(wi, m) : wo;   // roughgrating.cpp:534-540: the reflected direction
                p.sample_active = (pdf != 0.f) && wo.z > 0.f;
                const float weight = distr.sample_visible ? distr.smith_g1(wo_hat, m)
                                                          : distr.G(wi, wo_hat, m) * dot(wi, m) / (cos_theta_i * m.z);
                p.w_scalar = s4(weight) * refl;
                p.fs.init(m, wo_hat, wi); p.cos_s = dot(wo_hat, m);
            }
            // replay weight: eval_pdf of the rough conductor lobe (roughconductor.cpp:432-516, roughgrating.cpp:1036-1127)
            const V3 H = normalize(wo + wi);
            bool active = cos_theta_i > 0.f && wo.z > 0.f && dot(wi, H) > 0.f && dot(wo, H) > 0.f;
            if (active) {
                const float D = distr.eval(H);
                active = D != 0.f;
                const float g1 = distr.smith_g1(wi, H), G = g1 * distr.smith_g1(wo, H);
                const float val = D * G / (4.f * cos_theta_i);
                p.weight_val = p.type == PLTB_BSDF_ROUGHCONDUCTOR ? s4(val) * refl : s4(val);
                p.weight_pdf = distr.sample_visible ? D * g1 / (4.f * cos_theta_i) : distr.pdf(wi, H) / (4.f * dot(wo, H));
                p.weight_active = active;
                p.fw.init(H, wo, wi); p.cos_w = dot(wo, H);
            }
            p.intensity = so.aux; p.multiplier = b->f[7];
            return;
        }
        default: return;
    }
}

// First row of wbsdf_eval in a polarized variant for wavelength k (the only part of `em_weight * mis_em * bsdf_val * alpha` that
// survives the depolarizing emitter weight on its left, plt.py:300): diffuse / principled: depolarizer(value); rough conductor
// and grating: F(H) * value (roughconductor.cpp:351-389, roughgrating.cpp:924-968).  `value` is the unpolarized factor.
struct PolEval { uint32_t type; bool zero; S4 eta, k; StokesFrames f; float cos_h; };
PLTB_NI void pol_eval_prepare(const DeviceScene &sc, const pltb_bsdf *b, V3 wi, V2 uv, const S4 &wl, V3 wo, PolEval &e) {
    e.type = b ? b->type : 0u; e.zero = true;
    if (!b) return;
    if (b->flags & PLTB_BSDF_FLAG_TWOSIDED) { wo.z = mulsign(wo.z, wi.z); wi.z = fabsf(wi.z); }
    if (e.type == PLTB_BSDF_ROUGHCONDUCTOR || e.type == PLTB_BSDF_ROUGHGRATING) {
        e.eta = tex_eval(sc, b->tex[0], wl, uv); e.k = tex_eval(sc, b->tex[1], wl, uv);
        const V3 H = normalize(wo + wi);
        e.f.init(H, wo, wi); e.cos_h = dot(wo, H);
        e.zero = false;
    } else if (e.type == PLTB_BSDF_DIFFUSE || e.type == PLTB_BSDF_PRINCIPLED) e.zero = false;
}
PLTB_HD void pol_eval_row0(const PolEval &e, float value_k, int k, float row[4]) {
    if (e.zero) { row[0] = row[1] = row[2] = row[3] = 0.f; return; }
    if (e.type == PLTB_BSDF_DIFFUSE || e.type == PLTB_BSDF_PRINCIPLED) { row[0] = value_k; row[1] = row[2] = row[3] = 0.f; return; }
    const F44 F = e.f.align(mueller_reflection(fresnel_polarized(e.cos_h, cx(e.eta.v[k], e.k.v[k]))));
    for (int j = 0; j < 4; ++j) row[j] = F.m[0][j] * value_k;
}

}  // namespace pltb
