// pltb_principled.cuh -- the `principled` BSDF (src/bsdfs/principled.cpp, principledhelpers.h) without its transmission
// lobe (spec_trans absent or 0; the host loader rejects anything else): diffuse + retro-reflection + fake subsurface +
// sheen, clearcoat (GTR1), main specular (anisotropic GGX with the tinted / metallic Schlick-Fresnel blend).
// Used by scenes/gratingcap; wbsdf_sample / eval / pdf / weight are the BSDF base-class defaults.
#pragma once
#include "pltb_shading.cuh"

namespace pltb {

PLTB_HD float dr_lerp(float a, float b, float t) { return fmadd(b, t, fnmadd(a, t, a)); }   // drjit array_router.h:350
PLTB_HD float schlick_weight(float cos_i) {   // principledhelpers.h:108-112
    float m = fminf(fmaxf(1.0f - cos_i, 0.0f), 1.0f);
    return sqr(sqr(m)) * m;
}
PLTB_HD float schlick_R0_eta(float eta) { return sqr((eta - 1.0f) / (eta + 1.0f)); }
PLTB_HD float calc_schlick(float R0, float cos_theta_i, float eta) {   // principledhelpers.h:114-128
    bool outside = cos_theta_i >= 0.0f;
    float rcp_eta = rcp(eta), eta_it = outside ? eta : rcp_eta, eta_ti = outside ? rcp_eta : eta;
    float cos_theta_t_sqr = fnmadd(fnmadd(cos_theta_i, cos_theta_i, 1.0f), sqr(eta_ti), 1.0f);
    float cos_theta_t = safe_sqrt(cos_theta_t_sqr);
    return eta_it > 1.0f ? dr_lerp(schlick_weight(fabsf(cos_theta_i)), 1.0f, R0) : dr_lerp(schlick_weight(cos_theta_t), 1.0f, R0);
}
PLTB_HD bool mac_mic_compatibility(V3 m, V3 wi, V3 wo, float cos_theta_i, bool reflection) {   // principledhelpers.h:134-146
    V3 ms = mk3(mulsign(m.x, cos_theta_i), mulsign(m.y, cos_theta_i), mulsign(m.z, cos_theta_i));
    if (reflection) return dot(wi, ms) > 0.0f && dot(wo, ms) > 0.0f;
    V3 mn = mk3(mulsign_neg(m.x, cos_theta_i), mulsign_neg(m.y, cos_theta_i), mulsign_neg(m.z, cos_theta_i));
    return dot(wi, ms) > 0.0f && dot(wo, mn) > 0.0f;
}
// mitsuba::luminance(value, wavelengths), spectral mode: mean(cie1931_y(wl) * value) (spectrum.h:272-295, 414-429)
PLTB_HD float spectral_luminance(const DeviceScene &sc, const S4 &value, const S4 &wl) {
    if (sc.rgb) return value.v[0] * 0.212671f + value.v[1] * 0.715160f + value.v[2] * 0.072169f;   // luminance(Color3f), spectrum.h:431-434
    const float *y = sc.cie_srgb + 3 * 95;
    float acc[4];
    for (int i = 0; i < 4; ++i) {
        float t = (wl.v[i] - 360.f) * (94.f / (830.f - 360.f));
        bool active = wl.v[i] >= 360.f && wl.v[i] <= 830.f;
        int32_t ti = (int32_t) t;
        uint32_t i0 = (uint32_t) (ti < 0 ? 0 : (ti > 93 ? 93 : ti));
        float w1 = t - (float) i0, w0 = 1.f - w1;
        acc[i] = (active ? fmadd(w0, PLTB_LDG(y + i0), w1 * PLTB_LDG(y + i0 + 1)) : 0.f) * value.v[i];
    }
    return (acc[0] + acc[1] + acc[2] + acc[3]) * 0.25f;
}

struct Principled {
    float eta, roughness, anisotropic, metallic, clearcoat, clearcoat_gloss, sheen, sheen_tint, flatness, spec_tint, spec_trans;
    float spec_srate, clearcoat_srate, diff_refl_srate;
    bool has_anisotropic, has_metallic, has_clearcoat, has_sheen, has_sheen_tint, has_flatness, has_spec_tint, has_spec_trans;
    PLTB_HD void init(const pltb_bsdf &b) {
        const uint32_t h = b.u[0];
        has_anisotropic = (h & PLTB_PRINCIPLED_HAS_ANISOTROPIC) != 0; has_metallic = (h & PLTB_PRINCIPLED_HAS_METALLIC) != 0;
        has_clearcoat = (h & PLTB_PRINCIPLED_HAS_CLEARCOAT) != 0; has_sheen = (h & PLTB_PRINCIPLED_HAS_SHEEN) != 0;
        has_sheen_tint = (h & PLTB_PRINCIPLED_HAS_SHEEN_TINT) != 0; has_flatness = (h & PLTB_PRINCIPLED_HAS_FLATNESS) != 0;
        has_spec_tint = (h & PLTB_PRINCIPLED_HAS_SPEC_TINT) != 0; has_spec_trans = (h & PLTB_PRINCIPLED_HAS_SPEC_TRANS) != 0;
        eta = b.f[0]; roughness = b.f[1];
        anisotropic = has_anisotropic ? b.f[2] : 0.f; metallic = has_metallic ? b.f[3] : 0.f;
        clearcoat = has_clearcoat ? b.f[4] : 0.f; clearcoat_gloss = b.f[5];
        sheen = has_sheen ? b.f[6] : 0.f; sheen_tint = b.f[7];
        flatness = has_flatness ? b.f[8] : 0.f; spec_tint = has_spec_tint ? b.f[9] : 0.f;
        spec_trans = has_spec_trans ? b.f[10] : 0.f;
        spec_srate = u2f(b.u[1]); clearcoat_srate = u2f(b.u[2]); diff_refl_srate = u2f(b.u[3]);
    }
    PLTB_HD Microfacet spec_distr() const {   // calc_dist_params (principledhelpers.h:214-226); GGX, sample_visible
        Microfacet m; m.type = PLTB_MICROFACET_GGX; m.sample_visible = true;
        float r2 = sqr(roughness);
        if (!has_anisotropic) { m.au = m.av = fmaxf(0.001f, r2); }
        else {
            float aspect = sqrtf(1.0f - 0.9f * anisotropic);
            m.au = fmaxf(0.001f, r2 / aspect); m.av = fmaxf(0.001f, r2 * aspect);
        }
        m.au = fmaxf(m.au, 1e-4f); m.av = fmaxf(m.av, 1e-4f);   // MicrofacetDistribution::configure (microfacet.h:169-173)
        return m;
    }
    // lobe selection probabilities (principled.cpp:376-399 == :760-786)
    PLTB_HD void probs(float F_spec, bool front, float &p_spec, float &p_cc, float &p_diff, float &p_trans) const {
        float brdf = (1.0f - metallic) * (1.0f - spec_trans), bsdf = has_spec_trans ? (1.0f - metallic) * spec_trans : 0.0f;
        float ps = front ? spec_srate * (1.0f - bsdf * (1.0f - F_spec)) : F_spec;
        float pt = has_spec_trans ? (front ? spec_srate * bsdf * (1.0f - F_spec) : (1.0f - F_spec)) : 0.0f;
        float pc = has_clearcoat ? (front ? 0.25f * clearcoat * clearcoat_srate : 0.0f) : 0.0f;
        float pd = front ? brdf * diff_refl_srate : 0.f;
        float rcp_tot = rcp(ps + pt + pc + pd);
        p_spec = ps * rcp_tot; p_cc = pc * rcp_tot; p_diff = pd * rcp_tot; p_trans = pt * rcp_tot;
    }
};

// GTR1Isotropic (principledhelpers.h:17-56)
struct GTR1 {
    float alpha;
    PLTB_HD float eval(V3 m) const {
        float cos_theta2 = sqr(m.z), alpha2 = sqr(alpha);
        float result = (alpha2 - 1.f) / (kPi * logf(alpha2) * (1.f + (alpha2 - 1.f) * cos_theta2));
        return result * m.z > 1e-20f ? result : 0.f;
    }
    PLTB_HD float pdf(V3 m) const { return m.z < 0.f ? 0.f : m.z * eval(m); }
    PLTB_HD V3 sample(V2 s) const {
        float ang = (2.f * kPi) * s.x, sin_phi = sinf(ang), cos_phi = cosf(ang), alpha2 = sqr(alpha);
        float cos_theta2 = (1.f - powf(alpha2, 1.f - s.y)) / (1.f - alpha2);
        float sin_theta = sqrtf(fmaxf(0.f, 1.f - cos_theta2)), cos_theta = sqrtf(fmaxf(0.f, cos_theta2));
        return mk3(cos_phi * sin_theta, sin_phi * sin_theta, cos_theta);
    }
};
PLTB_HD float smith_ggx1(V3 v, V3 wh, float alpha) {   // principledhelpers.h:73-93
    float alpha_2 = sqr(alpha), cos_theta = fabsf(v.z), cos_theta_2 = sqr(cos_theta), tan_theta_2 = (1.0f - cos_theta_2) / cos_theta_2;
    float result = 2.0f * rcp(1.0f + sqrtf(1.0f + alpha_2 * tan_theta_2));
    if (v.z == 1.f) result = 1.f;
    if (dot(v, wh) * v.z <= 0.f) result = 0.f;
    return result;
}

// Principled::eval (principled.cpp:488-712)
PLTB_NI S4 principled_eval(const DeviceScene &sc, const pltb_bsdf &b, V3 wi, V2 uv, const S4 &wl, V3 wo) {
    Principled P; P.init(b);
    const float cos_theta_i = wi.z;
    if (!(cos_theta_i != 0.0f)) return s4(0.f);
    const S4 base_color = tex_eval(sc, b.tex[0], wl, uv);
    const float brdf = (1.0f - P.metallic) * (1.0f - P.spec_trans), bsdf = (1.0f - P.metallic) * P.spec_trans;
    const float cos_theta_o = wo.z;
    const bool reflect_ = cos_theta_i * cos_theta_o > 0.0f, refract_ = cos_theta_i * cos_theta_o < 0.0f, front_side = cos_theta_i > 0.0f;
    const float inv_eta = rcp(P.eta), eta_path = front_side ? P.eta : inv_eta, inv_eta_path = front_side ? inv_eta : P.eta;
    const Microfacet spec_dist = P.spec_distr();
    V3 wh = normalize(wi + wo * (reflect_ ? 1.0f : eta_path));
    wh = mk3(mulsign(wh.x, wh.z), mulsign(wh.y, wh.z), mulsign(wh.z, wh.z));
    const FresnelResult fr = fresnel(dot(wi, wh), P.eta);
    const float F_spec_dielectric = fr.r;
    const bool reflection_compat = mac_mic_compatibility(wh, wi, wo, cos_theta_i, true);
    const bool spec_reflect_active = reflect_ && reflection_compat && F_spec_dielectric > 0.0f;
    const bool clearcoat_active = P.has_clearcoat && P.clearcoat > 0.0f && reflect_ && reflection_compat && front_side;
    const bool spec_trans_active = P.has_spec_trans && bsdf > 0.0f && refract_ && mac_mic_compatibility(wh, wi, wo, cos_theta_i, false) &&
                                   F_spec_dielectric < 1.0f;
    const bool diffuse_active = brdf > 0.0f && reflect_ && front_side;
    const bool sheen_active = P.has_sheen && P.sheen > 0.0f && reflect_ && (1.0f - P.metallic > 0.0f) && front_side;
    const float D = spec_dist.eval(wh), G = spec_dist.G(wi, wo, wh);
    S4 value = s4(0.f);
    if (spec_reflect_active) {
        const float lum = P.has_spec_tint ? spectral_luminance(sc, base_color, wl) : 1.0f;
        // principled_fresnel (principledhelpers.h:148-186)
        const float cti = dot(wi, wh);
        const float eta_it = cti >= 0.0f ? P.eta : rcp(P.eta);
        for (int i = 0; i < 4; ++i) {
            float F_schlick = 0.f;
            if (P.has_metallic) F_schlick += P.metallic * calc_schlick(base_color.v[i], cti, P.eta);
            if (P.has_spec_tint) {
                float c_tint = lum > 0.0f ? base_color.v[i] / lum : 1.0f;
                float F0 = c_tint * schlick_R0_eta(eta_it);
                F_schlick += (1.0f - P.metallic) * P.spec_tint * calc_schlick(F0, cti, P.eta);
            }
            float F_front = (1.0f - P.metallic) * (1.0f - P.spec_tint) * F_spec_dielectric + F_schlick;
            float Fp = front_side ? F_front : bsdf * F_spec_dielectric;
            value.v[i] += Fp * D * G / (4.0f * fabsf(cos_theta_i));
        }
    }
    if (spec_trans_active) {   // main specular transmission (principled.cpp:604-621), TransportMode::Radiance: scale = inv_eta_path^2
        const float scale = sqr(inv_eta_path), dwi = dot(wi, wh), dwo = dot(wo, wh);
        const float t = fabsf((scale * (1.0f - F_spec_dielectric) * D * G * eta_path * eta_path * dwi * dwo) /
                              (cos_theta_i * sqr(dwi + eta_path * dwo)));
        for (int i = 0; i < 4; ++i) value.v[i] += sqrtf(base_color.v[i]) * bsdf * t;
    }
    if (clearcoat_active) {
        const float Fcc = calc_schlick(0.04f, dot(wi, wh), P.eta);
        GTR1 mf; mf.alpha = dr_lerp(0.1f, 0.001f, P.clearcoat_gloss);
        const float Dcc = mf.eval(wh);
        const float G_cc = smith_ggx1(wi, wh, 0.25f) * smith_ggx1(wo, wh, 0.25f);
        const float add = (P.clearcoat * 0.25f) * Fcc * Dcc * G_cc * fabsf(cos_theta_o);
        for (int i = 0; i < 4; ++i) value.v[i] += add;
    }
    if (diffuse_active) {
        const float Fo = schlick_weight(fabsf(cos_theta_o)), Fi = schlick_weight(fabsf(cos_theta_i));
        const float f_diff = (1.0f - 0.5f * Fi) * (1.0f - 0.5f * Fo);
        const float cos_theta_d = dot(wh, wo);
        const float Rr = 2.0f * P.roughness * sqr(cos_theta_d);
        const float f_retro = Rr * (Fo + Fi + Fo * Fi * (Rr - 1.0f));
        float lobe;
        if (P.has_flatness) {
            const float Fss90 = Rr / 2.0f;
            const float Fss = dr_lerp(1.0f, Fss90, Fo) * dr_lerp(1.0f, Fss90, Fi);
            const float f_ss = 1.25f * (Fss * (1.0f / (fabsf(cos_theta_o) + fabsf(cos_theta_i)) - 0.5f) + 0.5f);
            lobe = dr_lerp(f_diff + f_retro, f_ss, P.flatness);
        } else lobe = f_diff + f_retro;
        for (int i = 0; i < 4; ++i) value.v[i] += brdf * fabsf(cos_theta_o) * base_color.v[i] * kInvPi * lobe;
        if (sheen_active) {
            const float Fd = schlick_weight(fabsf(cos_theta_d));
            if (P.has_sheen_tint) {
                const float lum = spectral_luminance(sc, base_color, wl);
                for (int i = 0; i < 4; ++i) {
                    float c_tint = lum > 0.0f ? base_color.v[i] / lum : 1.0f;
                    float c_sheen = dr_lerp(1.0f, c_tint, P.sheen_tint);
                    value.v[i] += P.sheen * (1.0f - P.metallic) * Fd * c_sheen * fabsf(cos_theta_o);
                }
            } else {
                const float add = P.sheen * (1.0f - P.metallic) * Fd * fabsf(cos_theta_o);
                for (int i = 0; i < 4; ++i) value.v[i] += add;
            }
        }
    }
    return value;
}

// Principled::pdf (principled.cpp:714-834)
PLTB_NI float principled_pdf(const pltb_bsdf &b, V3 wi, V3 wo) {
    Principled P; P.init(b);
    const float cos_theta_i = wi.z;
    if (!(cos_theta_i != 0.0f)) return 0.0f;
    const bool front_side = cos_theta_i > 0.0f;
    const float eta_path = front_side ? P.eta : rcp(P.eta), cos_theta_o = wo.z;
    const bool reflect_ = cos_theta_i * cos_theta_o > 0.0f, refract_ = cos_theta_i * cos_theta_o < 0.0f;
    V3 wh = normalize(wi + wo * (reflect_ ? 1.0f : eta_path));
    wh = mk3(mulsign(wh.x, wh.z), mulsign(wh.y, wh.z), mulsign(wh.z, wh.z));
    const Microfacet spec_distr = P.spec_distr();
    const FresnelResult fr = fresnel(dot(wi, wh), P.eta);
    float p_spec, p_cc, p_diff, p_trans;
    P.probs(fr.r, front_side, p_spec, p_cc, p_diff, p_trans);
    float dwh_dwo_abs;   // different for reflection and transmission (principled.cpp:789-801)
    if (P.has_spec_trans) {
        const float dot_wi_h = dot(wi, wh), dot_wo_h = dot(wo, wh);
        dwh_dwo_abs = fabsf(reflect_ ? rcp(4.0f * dot_wo_h) : (sqr(eta_path) * dot_wo_h) / sqr(dot_wi_h + eta_path * dot_wo_h));
    } else {
        dwh_dwo_abs = fabsf(rcp(4.0f * dot(wo, wh)));
    }
    float pdf = 0.0f;
    const bool macmic = mac_mic_compatibility(wh, wi, wo, cos_theta_i, true) && reflect_;
    const V3 wi_s = mk3(mulsign(wi.x, cos_theta_i), mulsign(wi.y, cos_theta_i), mulsign(wi.z, cos_theta_i));
    if (macmic) pdf += p_spec * spec_distr.pdf(wi_s, wh) * dwh_dwo_abs;
    if (reflect_) pdf += p_diff * (kInvPi * wo.z);
    if (P.has_spec_trans && mac_mic_compatibility(wh, wi, wo, cos_theta_i, false) && refract_)
        pdf += p_trans * spec_distr.pdf(wi_s, wh) * dwh_dwo_abs;
    if (P.has_clearcoat) {
        GTR1 cc; cc.alpha = dr_lerp(0.1f, 0.001f, P.clearcoat_gloss);
        if (macmic) pdf += p_cc * cc.pdf(wh) * dwh_dwo_abs;
    }
    return pdf;
}

// Principled::sample (principled.cpp:330-486)
PLTB_HD void principled_sample(const DeviceScene &sc, const pltb_bsdf &b, V3 wi, V2 uv, const S4 &wl, float s1, V2 s2,
                               V3 &out_wo, float &out_pdf, float &out_eta, uint32_t &out_type, S4 &out_weight) {
    const float cos_theta_i = wi.z;
    if (!(cos_theta_i != 0.0f)) return;
    Principled P; P.init(b);
    const bool front_side = cos_theta_i > 0.0f;
    const Microfacet spec_distr = P.spec_distr();
    const V3 wi_s = mk3(mulsign(wi.x, cos_theta_i), mulsign(wi.y, cos_theta_i), mulsign(wi.z, cos_theta_i));
    float mpdf; const V3 m_spec = spec_distr.sample(wi_s, s2, mpdf);
    const FresnelResult fr = fresnel(dot(wi, m_spec), P.eta);
    const float bsdf = P.has_spec_trans ? (1.0f - P.metallic) * P.spec_trans : 0.0f;
    bool active = front_side || bsdf > 0.0f;   // the inside case is only sampled when the BSDF major lobe is on
    float p_spec, p_cc, p_diff, p_trans;
    P.probs(fr.r, front_side, p_spec, p_cc, p_diff, p_trans);
    float curr = 0.0f;
    const bool sample_diffuse = active && s1 < p_diff;
    curr += p_diff;
    const bool sample_clearcoat = P.has_clearcoat && active && s1 >= curr && s1 < curr + p_cc;
    curr += p_cc;
    const bool sample_spec_trans = P.has_spec_trans && active && s1 >= curr && s1 < curr + p_trans;
    curr += p_trans;
    const bool sample_spec_reflect = active && s1 >= curr;
    out_eta = 1.0f;
    if (sample_spec_reflect) {
        V3 wo = reflect(wi, m_spec);
        out_wo = wo; out_type = F_GlossyReflection;
        active = active && mac_mic_compatibility(m_spec, wi, wo, cos_theta_i, true) && cos_theta_i * wo.z > 0.0f;
    }
    if (sample_spec_trans) {   // refract(wi, m, cos_theta_t, eta_ti) (fresnel.h:310-314)
        const float c = fmadd(dot(wi, m_spec), fr.eta_ti, fr.cos_theta_t);
        V3 wo = mk3(fmsub(m_spec.x, c, wi.x * fr.eta_ti), fmsub(m_spec.y, c, wi.y * fr.eta_ti), fmsub(m_spec.z, c, wi.z * fr.eta_ti));
        out_wo = wo; out_type = F_GlossyTransmission; out_eta = fr.eta_it;
        active = active && mac_mic_compatibility(m_spec, wi, wo, cos_theta_i, false) && cos_theta_i * wo.z < 0.0f;
    }
    if (sample_clearcoat) {
        GTR1 cc; cc.alpha = dr_lerp(0.1f, 0.001f, P.clearcoat_gloss);
        V3 m_cc = cc.sample(s2);
        V3 wo = reflect(wi, m_cc);
        out_wo = wo; out_type = F_GlossyReflection;
        active = active && mac_mic_compatibility(m_cc, wi, wo, cos_theta_i, true) && cos_theta_i * wo.z > 0.0f;
    }
    if (sample_diffuse) {
        V3 wo = square_to_cosine_hemisphere(s2);
        out_wo = wo; out_type = F_DiffuseReflection;
        active = active && cos_theta_i * wo.z > 0.0f;
    }
    out_pdf = active ? principled_pdf(b, wi, out_wo) : 0.0f;
    active = active && out_pdf > 0.0f;
    if (active) out_weight = principled_eval(sc, b, wi, uv, wl, out_wo) / out_pdf;
}

}  // namespace pltb
