// pltb_bsdf.cuh -- the wbsdf_sample / wbsdf_eval / wbsdf_pdf / wbsdf_weight family for the BSDFs on the plt path,
// including the TwoSidedBRDF wrapper.  JIT-variant semantics (see oracle/plt_oracle.cpp header).
#pragma once
#include "pltb_shading.cuh"
#include "pltb_principled.cuh"

namespace pltb {

struct BSDFSampleOut { V3 wo; float pdf, eta; uint32_t type; S4 weight; float aux; /* roughgrating: sampled lobe intensity */ };

PLTB_HD S4 conductor_F(const DeviceScene &sc, const pltb_bsdf &b, const S4 &wl, V2 uv, float cos_theta) {
    if (b.tex[0] >= 0 && b.tex[1] >= 0) {
        // spectrally and spatially flat eta and k (`uniform` spectra, the plugins' defaults 0 / 1 included): the four lanes would run
        // the same operations on the same inputs -- one Fresnel evaluation instead of four
        const pltb_texture &te = sc.textures[b.tex[0]], &tk = sc.textures[b.tex[1]];
        if (te.type == PLTB_TEX_UNIFORM && tk.type == PLTB_TEX_UNIFORM) return s4(fresnel_conductor(cos_theta, te.p[0], tk.p[0]));
    }
    S4 eta = tex_eval(sc, b.tex[0], wl, uv), k = tex_eval(sc, b.tex[1], wl, uv), F;
    for (int i = 0; i < 4; ++i) F.v[i] = fresnel_conductor(cos_theta, eta.v[i], k.v[i]);
    return F;
}

// (Measured and rejected: pooling the zero-to-two precise Bessel evaluations per lane of RoughGrating::wbsdf_sample across the
// warp -- ballots to number the (a, order) requests, one nearly full round of bessel_j, shuffles to hand the squares back --
// gave bit-identical results but k_shade<grating> went from 133 to 159 ms per 2^30 samples: the bookkeeping and the fragmented
// __activemask() groups cost more than the half-empty rounds they replace.)

// wi already has the twosided flip applied.
PLTB_HD BSDFSampleOut bsdf_sample_inner(const DeviceScene &sc, const pltb_bsdf &b, uint32_t type, V3 wi, V2 uv, const S4 &wl,
                                        float s1, V2 s2, V2 lobe_s2) {
    BSDFSampleOut o;
    o.wo = mk3(0, 0, 0); o.pdf = 0.f; o.eta = 0.f; o.type = 0; o.weight = s4(0.f); o.aux = 0.f;
    const float cos_theta_i = wi.z;
    switch (type) {
        case PLTB_BSDF_DIFFUSE: {  // src/bsdfs/diffuse.cpp:101-125
            if (!(cos_theta_i > 0.f)) return o;
            o.wo = square_to_cosine_hemisphere(s2);
            o.pdf = kInvPi * o.wo.z;
            o.eta = 1.f; o.type = F_DiffuseReflection;
            if (o.pdf > 0.f) o.weight = tex_eval(sc, b.tex[0], wl, uv);
            return o;
        }
        case PLTB_BSDF_CONDUCTOR: {  // src/bsdfs/conductor.cpp:248-308
            if (!(cos_theta_i > 0.f)) return o;
            o.type = F_DeltaReflection; o.wo = reflect(wi); o.eta = 1.f; o.pdf = 1.f;
            o.weight = tex_eval(sc, b.tex[2], wl, uv) * conductor_F(sc, b, wl, uv, cos_theta_i);
            return o;
        }
        case PLTB_BSDF_DIELECTRIC: {  // src/bsdfs/dielectric.cpp:265-400
            FresnelResult f = fresnel(cos_theta_i, b.f[0]);
            float t_i = 1.f - f.r;
            bool selected_r = s1 <= f.r;
            o.pdf = selected_r ? f.r : t_i;
            o.type = selected_r ? F_DeltaReflection : F_DeltaTransmission;
            o.wo = selected_r ? reflect(wi) : refract(wi, f.cos_theta_t, f.eta_ti);
            o.eta = selected_r ? 1.f : f.eta_it;
            S4 w = s4(1.f);
            if (selected_r) { if (b.tex[2] >= 0) w = w * tex_eval(sc, b.tex[2], wl, uv); }
            else {
                if (b.tex[3] >= 0) w = w * tex_eval(sc, b.tex[3], wl, uv);
                w = w * sqr(f.eta_ti);
            }
            o.weight = w;
            return o;
        }
        case PLTB_BSDF_PRINCIPLED:
            principled_sample(sc, b, wi, uv, wl, s1, s2, o.wo, o.pdf, o.eta, o.type, o.weight);
            return o;
        case PLTB_BSDF_ROUGHCONDUCTOR: {  // src/bsdfs/roughconductor.cpp:225-311
            if (!(cos_theta_i > 0.f)) return o;
            Microfacet distr = make_mf(b);
            float pdf; V3 m = distr.sample(wi, s2, pdf);
            o.pdf = pdf; o.wo = reflect(wi, m); o.eta = 1.f; o.type = F_GlossyReflection;
            bool active = (o.pdf != 0.f) && o.wo.z > 0.f;
            float weight = distr.sample_visible ? distr.smith_g1(o.wo, m)
                                                : distr.G(wi, o.wo, m) * dot(wi, m) / (cos_theta_i * m.z);
            o.pdf /= 4.f * dot(o.wo, m);
            S4 F = conductor_F(sc, b, wl, uv, dot(wi, m));
            S4 w = s4(weight);
            if (b.tex[2] >= 0) w = w * tex_eval(sc, b.tex[2], wl, uv);
            if (active) o.weight = F * w;
            return o;
        }
        case PLTB_BSDF_ROUGHGRATING: {  // src/bsdfs/roughgrating.cpp:449-595, sample_diffract :414-447
            if (!(cos_theta_i > 0.f)) return o;
            Microfacet distr = make_mf(b);
            float pdf; V3 m = distr.sample(wi, s2, pdf);
            o.pdf = pdf;
            V3 reflection_dir = reflect(wi, m);
            // spectral: the hero wavelength; RGB variants: "Use sample 1 for random wavelength sample" (roughgrating.cpp:499-506)
            float wavelength = sc.rgb ? s1 * (830.f - 150.0f - 360.f) + 360.f : wl.v[0];
            Frame nf = frame_from_normal(m);
            V3 wi_local = nf.to_local(wi);
            Grating g = make_grating(b, uv);
            V2 pdf_xy; int lx, ly;
            g.sample_lobe(sc.bessel, lobe_s2, wi_local, wavelength * 1e-3f, lx, ly, pdf_xy);
            float intensity = g.lobe_intensity_sampled(sc.bessel, lx, ly, wi_local, wavelength);  // nm as written (N2)
            bool valid; V3 wo_local = g.diffract(wi_local, lx, ly, wavelength * 1e-3f, valid);
            o.wo = nf.to_world(wo_local);
            float grating_pdf = pdf_xy.x * pdf_xy.y;
            o.eta = 1.f; o.type = F_GlossyReflection;
            bool active = (o.pdf != 0.f) && o.wo.z > 0.f;
            o.pdf *= grating_pdf;
            float weight = distr.sample_visible ? distr.smith_g1(reflection_dir, m)
                                                : distr.G(wi, reflection_dir, m) * dot(wi, m) / (cos_theta_i * m.z);
            o.pdf /= 4.f * dot(reflection_dir, m);
            S4 F = conductor_F(sc, b, wl, uv, dot(wi, m));
            S4 w = s4(weight);
            if (b.tex[2] >= 0) w = w * tex_eval(sc, b.tex[2], wl, uv);
            if (active) o.weight = ((F * w) * intensity) * b.f[7];
            o.aux = intensity;
            return o;
        }
        default: return o;
    }
}

// conventional rough-conductor eval_pdf (roughconductor.cpp:432-516; RoughGrating::eval_pdf roughgrating.cpp:1036-1127)
PLTB_HD void rough_eval_pdf(const DeviceScene &sc, const pltb_bsdf &b, V3 wi, V2 uv, const S4 &wl, V3 wo, bool apply_refl,
                            S4 &value, float &pdf) {
    value = s4(0.f); pdf = 0.f;
    float cos_theta_i = wi.z, cos_theta_o = wo.z;
    V3 H = normalize(wo + wi);
    bool active = cos_theta_i > 0.f && cos_theta_o > 0.f && dot(wi, H) > 0.f && dot(wo, H) > 0.f;
    if (!active) return;
    Microfacet distr = make_mf(b);
    float D = distr.eval(H);
    if (!(D != 0.f)) return;   // value and pdf stay 0: no need for G and the Fresnel factor (half of the diffracted directions)
    float g1 = distr.smith_g1(wi, H), G = g1 * distr.smith_g1(wo, H);
    float val = D * G / (4.f * cos_theta_i);
    S4 F = conductor_F(sc, b, wl, uv, dot(wi, H));
    S4 v = s4(val);
    if (apply_refl && b.tex[2] >= 0) v = v * tex_eval(sc, b.tex[2], wl, uv);
    float p = distr.sample_visible ? D * g1 / (4.f * cos_theta_i) : distr.pdf(wi, H) / (4.f * dot(wo, H));
    if (active) { value = F * v; pdf = p; }
}

// RoughGrating::wbsdf_eval, spectral branch (roughgrating.cpp:841-969), in two steps with the reference's arithmetic and
// summation order:
//   grating_scan   walks the (lx, ly, lambda) loops and lists the lobes whose window contains wo.  Invalid lobes (|sin_o| > 1)
//                  add exactly 0 and are skipped; the window test |unit_angle(c, wo)| < a_win (drjit sphere.h:53-60) is done on
//                  the squared chord (for dot >= 0 the angle is 2 asin(|wo - c| / 2), monotone in |wo - c|^2; dot < 0 means an
//                  angle >= pi/2, which only passes for a_win > pi/2, i.e. alpha > 0.6: those BSDFs take the exact test); and a
//                  conservative pre-filter in the grating's sine space rejects lobes before their direction is reconstructed.
//   grating_finish evaluates the per-order terms J_l(a)^2 the listed lobes need -- in one uniform (lambda, order) loop, so the
//                  lanes of a warp that need the same term evaluate it together -- and accumulates the listed lobes in loop
//                  order.  On cbox-grating 9 % of the NEE directions hit any window, so 12 Bessel evaluations per item (392 as
//                  written in the reference) become ~0.4.
struct GratingHits {
    static constexpr int kMax = 12;
    uint32_t need;             // bit (8 * i + l): term of order l at wavelength i is needed
    int n;                     // lobes inside their window (may exceed kMax: then `hit` is incomplete and finish() rescans)
    uint16_t hit[kMax];        // (lx + 4) | (ly + 4) << 4 | i << 8, in loop order
};

struct GratingLoop {
    Grating g;
    V2 sin_i;
    float wl_um[4];
    float a_win, chord_sq, awo, bwo, tolx, toly;
    int half, lx0, lx1, ly0, ly1;   // lobe ranges that can pass the pre-filter (whole [-half, half] when it is off)
    bool small_win;

    PLTB_HD void init(const pltb_bsdf &b, V3 wi, V2 uv, const S4 &wl, V3 wo) {
        g = make_grating(b, uv);
        a_win = 2 * sqrtf(b.f[1] * b.f[2]);   // Texture::eval_1 of alpha_u / alpha_v: unclamped (roughgrating.cpp:871-874)
        half = (int) g.lobes / 2;
        for (int i = 0; i < 4; ++i) wl_um[i] = wl.v[i] * 1e-3f;
        sin_i = g.sin_i(wi);
        const float half_win = fminf(0.5f * a_win, 0.5f * kPi), chord = 2.f * sinf(half_win);
        chord_sq = chord * chord;
        // Pre-filter: for a valid lobe the direction c is a unit vector with c.x / |c.xz| == da and c.y / |c.yz| == db
        // (diffractiongrating.h:211-225).  v -> v.x / |v.xz| has gradient <= 1 / |v.xz|, so |c - wo| < chord implies
        // |da - wo.x / |wo.xz|| <= chord / (|wo.xz| - chord) (likewise in y); a 1.5x margin covers rounding.
        small_win = a_win <= 0.5f * kPi;
        const float pwx = sqrtf(sqr(wo.x) + sqr(wo.z)), pwy = sqrtf(sqr(wo.y) + sqr(wo.z));
        const bool fx = small_win && pwx > 2.f * chord, fy = small_win && pwy > 2.f * chord;
        awo = fx ? wo.x / pwx : 0.f; bwo = fy ? wo.y / pwy : 0.f;
        tolx = fx ? 1.5f * chord / (pwx - chord) + 1e-5f : inf_f();
        toly = fy ? 1.5f * chord / (pwy - chord) + 1e-5f : inf_f();
        // The same two tests bound the lobe indices: per wavelength the rotated lobe (lrx, lry) must lie in a box around
        // ((awo + sin_i.x) / (wl ipx), (bwo + sin_i.y) / (wl ipy)); rotating the union of the four boxes back gives integer
        // ranges for lx, ly.  The loops below still apply the tests themselves, so this only skips iterations that fail them
        // (196 -> ~10 per item on cbox-grating); margins cover the rounding of this float arithmetic.
        lx0 = ly0 = -half; lx1 = ly1 = half;
        if (fx && fy && g.inv_period.x > 1e-6f && g.inv_period.y > 1e-6f) {
            float ax0 = inf_f(), ax1 = -inf_f(), ay0 = inf_f(), ay1 = -inf_f();
            const float cx0 = fmaxf(awo - tolx, -1.f) + sin_i.x, cx1 = fminf(awo + tolx, 1.f) + sin_i.x;
            const float cy0 = fmaxf(bwo - toly, -1.f) + sin_i.y, cy1 = fminf(bwo + toly, 1.f) + sin_i.y;
            for (int i = 0; i < 4; ++i) {
                const float rx = 1.f / (wl_um[i] * g.inv_period.x), ry = 1.f / (wl_um[i] * g.inv_period.y);
                ax0 = fminf(ax0, fminf(cx0 * rx, cx1 * rx)); ax1 = fmaxf(ax1, fmaxf(cx0 * rx, cx1 * rx));
                ay0 = fminf(ay0, fminf(cy0 * ry, cy1 * ry)); ay1 = fmaxf(ay1, fmaxf(cy0 * ry, cy1 * ry));
            }
            if (ax0 <= ax1 && ay0 <= ay1 && finite_(ax0 + ax1 + ay0 + ay1)) {
                const float mx = 1e-3f * (1.f + fmaxf(fabsf(ax0), fabsf(ax1))), my = 1e-3f * (1.f + fmaxf(fabsf(ay0), fabsf(ay1)));
                ax0 -= mx; ax1 += mx; ay0 -= my; ay1 += my;
                const float cg = g.dir.x, sg = g.dir.y;   // l = R^T lr
                const float xa = cg * ax0, xb = cg * ax1, xc = sg * ay0, xd = sg * ay1;
                const float ya = -sg * ax0, yb = -sg * ax1, yc = cg * ay0, yd = cg * ay1;
                const float fl0 = fminf(xa, xb) + fminf(xc, xd) - 0.01f, fl1 = fmaxf(xa, xb) + fmaxf(xc, xd) + 0.01f;
                const float gl0 = fminf(ya, yb) + fminf(yc, yd) - 0.01f, gl1 = fmaxf(ya, yb) + fmaxf(yc, yd) + 0.01f;
                const float fh = (float) half;
                lx0 = (int) ceilf(fmaxf(fl0, -fh)); lx1 = (int) floorf(fminf(fl1, fh));
                ly0 = (int) ceilf(fmaxf(gl0, -fh)); ly1 = (int) floorf(fminf(gl1, fh));
            } else if (!(ax0 <= ax1 && ay0 <= ay1)) {
                lx0 = 1; lx1 = 0;   // the box is empty (NaN-free): nothing can pass
            }
        }
    }
    // calls f(lx, ly, i) for every lobe whose window contains wo, in the reference's loop order
    template <class F> PLTB_HD void for_each_inside(V3 wo, F &&f) const {
        const float cos_g = g.dir.x, sin_g = g.dir.y;
        for (int lx = lx0; lx <= lx1; ++lx) {
#pragma unroll 1
            for (int ly = ly0; ly <= ly1; ++ly) {
                const float lrx = cos_g * (float) lx - sin_g * (float) ly, lry = sin_g * (float) lx + cos_g * (float) ly;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float da = wl_um[i] * lrx * g.inv_period.x - sin_i.x, db = wl_um[i] * lry * g.inv_period.y - sin_i.y;
                    if (!(fabsf(da) <= 1.0f && fabsf(db) <= 1.0f)) continue;   // lobe_active == false: adds 0
                    if (fabsf(da - awo) > tolx || fabsf(db - bwo) > toly) continue;   // cannot be inside the window
                    const float m = (sqr(da) - 1) / (sqr(da * db) - 1), qq = 1 - sqr(db) * m;
                    const V3 c = mk3(da * sqrtf(fmaxf(0.0f, qq)), db * sqrtf(m), sqrtf(fmaxf(0.0f, 1 - sqr(da) * qq - sqr(db) * m)));
                    bool inside;
                    if (small_win) inside = dot(c, wo) >= 0.f && sqnorm(wo - c) < chord_sq;
                    else inside = fabsf(unit_angle(c, wo)) < a_win;
                    if (inside) f(lx, ly, i);
                }
            }
        }
    }
};

PLTB_HD void grating_scan(const GratingLoop &gl, V3 wo, GratingHits &h) {
    h.need = 0; h.n = 0;
    const bool one_d = gl.g.is_1d();
    gl.for_each_inside(wo, [&](int lx, int ly, int i) {
        const int alx = lx < 0 ? -lx : lx, aly = ly < 0 ? -ly : ly;
        if (alx) h.need |= 1u << (8 * i + alx);
        if (aly && !one_d) h.need |= 1u << (8 * i + aly);
        if (h.n < GratingHits::kMax) h.hit[h.n] = (uint16_t) ((lx + 4) | ((ly + 4) << 4) | (i << 8));
        h.n++;
    });
}

// the lobe sum of grating_finish without the Fresnel factor and specular_reflectance (the polarized variants multiply it with a
// Mueller matrix instead, roughgrating.cpp:924-968)
PLTB_HD S4 grating_lobe_sum(const DeviceScene &sc, const pltb_bsdf &b, const GratingLoop &gl, V3 wi, V3 wo, const GratingHits &h) {
    const Grating &g = gl.g;
    S4 result = s4(0.f);
    if (h.n > 0) {
        const bool one_d = g.is_1d(), rescan = h.n > GratingHits::kMax;
        float term[4][PLTB_MAX_LOBES / 2 + 1], coh[4];
        const float angle_offset = fabsf(unit_angle(reflect(wi), wo));
        for (int i = 0; i < 4; ++i) {
            const float a = kFourPi * g.q / (gl.wl_um[i] * fabsf(wi.z));
            term[i][0] = 1.0f;
            for (int l = 1; l <= gl.half; ++l)
                if (rescan || (h.need >> (8 * i + l)) & 1u) term[i][l] = g.axis_term(sc.bessel, a, l);
            // Coherence(m_coherence, 1.0).inv_coherence_det(k), include/mitsuba/plt/plt.h:57-100
            float k = kTwoPi / gl.wl_um[i];
            float s = k / (kTwoPi * (1.0f * 1e+3f));
            float d00 = s * b.f[8], d01 = s * 0.f;
            float det = fmsub(d00, d00, d01 * d01);
            float inv_det = isnan_(det) ? 0.0f : det;
            float ac = expf(-0.5f * angle_offset * inv_det * angle_offset);
            coh[i] = isnan_(ac) ? 0.0f : ac;
        }
        auto add = [&](int lx, int ly, int i) {
            const int alx = lx < 0 ? -lx : lx, aly = ly < 0 ? -ly : ly;
            float ix = term[i][alx], iy = one_d ? ix : term[i][aly];
            float lobe_intensity = g.multiplier * ix * iy;
            if (lx == 0 && ly == 0) result.v[i] += lobe_intensity;
            else result.v[i] += lobe_intensity * coh[i];
        };
        if (rescan) gl.for_each_inside(wo, add);
        else
            for (int k = 0; k < h.n; ++k) add((int) (h.hit[k] & 15) - 4, (int) ((h.hit[k] >> 4) & 15) - 4, (int) (h.hit[k] >> 8));
    }
    return result;
}

// cie1931_xyz (include/mitsuba/core/spectrum.h:236-264) followed by xyz_to_srgb (:406-412; Matrix * vector = fmadd chain over the
// columns); rows 4, 3, 5 of the scene's colour table are X, Y, Z
PLTB_HD void wavelength_to_srgb(const float *cie, float wl, float rgb[3]) {
    const float t = (wl - 360.f) * (94 / (830.f - 360.f));
    const bool active = wl >= 360.f && wl <= 830.f;
    uint32_t i0 = (uint32_t) fmaxf(t, 0.f);
    if (i0 > 93u) i0 = 93u;
    const float w1 = t - (float) i0, w0 = 1.f - w1;
    const float *X = cie + 4 * 95, *Y = cie + 3 * 95, *Z = cie + 5 * 95;
    const float x = active ? fmadd(w0, PLTB_LDG(X + i0), w1 * PLTB_LDG(X + i0 + 1)) : 0.f,
                y = active ? fmadd(w0, PLTB_LDG(Y + i0), w1 * PLTB_LDG(Y + i0 + 1)) : 0.f,
                z = active ? fmadd(w0, PLTB_LDG(Z + i0), w1 * PLTB_LDG(Z + i0 + 1)) : 0.f;
    rgb[0] = fmadd(-0.498535f, z, fmadd(-1.537150f, y, 3.240479f * x));
    rgb[1] = fmadd(0.041556f, z, fmadd(1.875991f, y, -0.969256f * x));
    rgb[2] = fmadd(1.057311f, z, fmadd(-0.204043f, y, 0.055648f * x));
}

// The RGB fallback of RoughGrating::wbsdf_eval (roughgrating.cpp:716-790) before the Fresnel factor and specular_reflectance:
// exhaustive search over the lobes at the path's three sampling wavelengths (wl lanes 0..2), each lobe coloured with
// xyz_to_srgb(cie1931_xyz(wl)); every lobe carries the angular coherence factor.  A lobe outside its window adds exactly 0 (the
// select discards its intensity; colour and coherence are finite), so its Bessel terms are not evaluated.
PLTB_NI S4 grating_lobe_sum_rgb(const DeviceScene &sc, const pltb_bsdf &b, V3 wi, V2 uv, const S4 &wl, V3 wo) {
    const Grating g = make_grating(b, uv);
    const V3 reflection_dir = reflect(wi);
    const V2 si_ = g.sin_i(wi);
    float result[3] = { 0.f, 0.f, 0.f };
    const float a = 2 * sqrtf(b.f[1] * b.f[2]);
    const int half = (int) g.lobes / 2;
    const float angle_offset = fabsf(unit_angle(reflection_dir, wo));
    float colour[3][3], coh[3];
    for (int i = 0; i < 3; ++i) {
        wavelength_to_srgb(sc.cie_srgb, wl.v[i], colour[i]);
        const float k = kTwoPi / (wl.v[i] * 1e-3f);
        const float s = k / (kTwoPi * (1.0f * 1e+3f));   // Coherence(m_coherence, 1.0).inv_coherence_det(k), plt.h:57-100
        const float d00 = s * b.f[8], d01 = s * 0.f;
        const float det = fmsub(d00, d00, d01 * d01);
        const float inv_det = isnan_(det) ? 0.0f : det;
        const float ac = expf(-0.5f * angle_offset * inv_det * angle_offset);
        coh[i] = isnan_(ac) ? 0.0f : ac;
    }
    for (int lx = -half; lx < half + 1; lx++) {
#pragma unroll 1
        for (int ly = -half; ly < half + 1; ly++) {
#pragma unroll 1
            for (int i = 0; i < 3; i++) {
                const float wl_um = wl.v[i] * 1e-3f;
                bool lobe_active;
                const V3 center_dir = g.diffract_s(si_, lx, ly, wl_um, lobe_active);
                if (!(lobe_active && fabsf(unit_angle(center_dir, wo)) < a)) continue;
                const float lobes_result = g.lobe_intensity(sc.bessel, lx, ly, wi, wl_um);
                for (int c = 0; c < 3; ++c) result[c] += (lobes_result * colour[i][c]) * coh[i];
            }
        }
    }
    S4 r; r.v[0] = result[0]; r.v[1] = result[1]; r.v[2] = r.v[3] = result[2];
    return r;
}

PLTB_HD S4 grating_finish(const DeviceScene &sc, const pltb_bsdf &b, const GratingLoop &gl, V3 wi, V2 uv, const S4 &wl, V3 wo,
                          const GratingHits &h) {
    S4 result = grating_lobe_sum(sc, b, gl, wi, wo, h);
    V3 H = normalize(wo + wi);
    S4 F = conductor_F(sc, b, wl, uv, dot(wi, H));
    if (b.tex[2] >= 0) result = result * tex_eval(sc, b.tex[2], wl, uv);
    return F * result;
}

PLTB_HD S4 grating_eval_rgb(const DeviceScene &sc, const pltb_bsdf &b, V3 wi, V2 uv, const S4 &wl, V3 wo) {   // roughgrating.cpp:716-839
    S4 result = grating_lobe_sum_rgb(sc, b, wi, uv, wl, wo);
    V3 H = normalize(wo + wi);
    S4 F = conductor_F(sc, b, wl, uv, dot(wi, H));
    if (b.tex[2] >= 0) result = result * tex_eval(sc, b.tex[2], wl, uv);
    return F * result;
}
PLTB_HD S4 grating_eval(const DeviceScene &sc, const pltb_bsdf &b, V3 wi, V2 uv, const S4 &wl, V3 wo) {
    GratingLoop gl; gl.init(b, wi, uv, wl, wo);
    GratingHits h; grating_scan(gl, wo, h);
    return grating_finish(sc, b, gl, wi, uv, wl, wo, h);
}

// Rgb: the RGB variants (DeviceScene::rgb) -- a template parameter because only roughgrating's evaluation differs, and the spectral
// instantiations should not carry the fallback's registers (k_nee_post<grating> 96 -> 128 with a run-time branch)
template <bool Rgb = false>
PLTB_HD S4 bsdf_eval_inner(const DeviceScene &sc, const pltb_bsdf &b, uint32_t type, V3 wi, V2 uv, const S4 &wl, V3 wo) {
    float cos_theta_i = wi.z, cos_theta_o = wo.z;
    switch (type) {
        case PLTB_BSDF_DIFFUSE:  // diffuse.cpp:127-142
            if (!(cos_theta_i > 0.f && cos_theta_o > 0.f)) return s4(0.f);
            return (tex_eval(sc, b.tex[0], wl, uv) * kInvPi) * cos_theta_o;
        case PLTB_BSDF_PRINCIPLED: return principled_eval(sc, b, wi, uv, wl, wo);
        case PLTB_BSDF_ROUGHCONDUCTOR: {  // roughconductor.cpp:313-390
            if (!(cos_theta_i > 0.f && cos_theta_o > 0.f)) return s4(0.f);
            V3 H = normalize(wo + wi);
            Microfacet distr = make_mf(b);
            float D = distr.eval(H);
            float G = distr.G(wi, wo, H);
            S4 result = s4(D * G / (4.f * cos_theta_i));
            S4 F = conductor_F(sc, b, wl, uv, dot(wi, H));
            if (b.tex[2] >= 0) result = result * tex_eval(sc, b.tex[2], wl, uv);
            return D != 0.f ? F * result : s4(0.f);
        }
        case PLTB_BSDF_ROUGHGRATING: return Rgb ? grating_eval_rgb(sc, b, wi, uv, wl, wo) : grating_eval(sc, b, wi, uv, wl, wo);
        default: return s4(0.f);
    }
}

// wbsdf_eval without its Fresnel factor (wi / wo already twosided-flipped): what the polarized variants scale a Mueller matrix
// with -- the whole value for the depolarizing BSDFs (diffuse, principled), D G / (4 cos) [* specular_reflectance] for the rough
// conductor (0 where the unpolarized eval is masked), the lobe sum [* specular_reflectance] for the grating
template <bool Rgb = false>
PLTB_HD S4 bsdf_eval_nofresnel(const DeviceScene &sc, const pltb_bsdf &b, uint32_t type, V3 wi, V2 uv, const S4 &wl, V3 wo) {
    switch (type) {
        case PLTB_BSDF_DIFFUSE:
        case PLTB_BSDF_PRINCIPLED: return bsdf_eval_inner(sc, b, type, wi, uv, wl, wo);
        case PLTB_BSDF_ROUGHCONDUCTOR: {
            if (!(wi.z > 0.f && wo.z > 0.f)) return s4(0.f);
            V3 H = normalize(wo + wi);
            Microfacet distr = make_mf(b);
            float D = distr.eval(H), G = distr.G(wi, wo, H);
            S4 result = s4(D * G / (4.f * wi.z));
            if (b.tex[2] >= 0) result = result * tex_eval(sc, b.tex[2], wl, uv);
            return D != 0.f ? result : s4(0.f);
        }
        case PLTB_BSDF_ROUGHGRATING: {
            if (Rgb) {
                S4 result = grating_lobe_sum_rgb(sc, b, wi, uv, wl, wo);
                if (b.tex[2] >= 0) result = result * tex_eval(sc, b.tex[2], wl, uv);
                return result;
            }
            GratingLoop gl; gl.init(b, wi, uv, wl, wo);
            GratingHits h; grating_scan(gl, wo, h);
            S4 result = grating_lobe_sum(sc, b, gl, wi, wo, h);
            if (b.tex[2] >= 0) result = result * tex_eval(sc, b.tex[2], wl, uv);
            return result;
        }
        default: return s4(0.f);
    }
}

PLTB_HD float bsdf_pdf_inner(const DeviceScene &sc, const pltb_bsdf &b, uint32_t type, V3 wi, V2 uv, const S4 &wl, V3 wo) {
    float cos_theta_i = wi.z, cos_theta_o = wo.z;
    switch (type) {
        case PLTB_BSDF_DIFFUSE: return (cos_theta_i > 0.f && cos_theta_o > 0.f) ? kInvPi * wo.z : 0.f;
        case PLTB_BSDF_PRINCIPLED: return principled_pdf(b, wi, wo);
        case PLTB_BSDF_ROUGHCONDUCTOR: {  // roughconductor.cpp:392-430
            V3 m = normalize(wo + wi);
            bool active = cos_theta_i > 0.f && cos_theta_o > 0.f && dot(wi, m) > 0.f && dot(wo, m) > 0.f;
            if (!active) return 0.f;
            Microfacet distr = make_mf(b);
            return distr.sample_visible ? distr.eval(m) * distr.smith_g1(wi, m) / (4.f * cos_theta_i)
                                        : distr.pdf(wi, m) / (4.f * dot(wo, m));
        }
        case PLTB_BSDF_ROUGHGRATING: {  // roughgrating.cpp:1009-1034 (RGB variants: wl.v[0] is sd.sampling_wavelengths[0], :1027-1031)
            float k = kTwoPi / (wl.v[0] * 1e-3f);
            return expf(-sqr(wi.z * b.f[6] * k));
        }
        default: return 0.f;
    }
}

PLTB_HD S4 bsdf_weight_inner(const DeviceScene &sc, const pltb_bsdf &b, uint32_t type, V3 wi, V2 uv, const S4 &wl, V3 wo) {
    switch (type) {
        case PLTB_BSDF_DIFFUSE: return tex_eval(sc, b.tex[0], wl, uv);  // diffuse.cpp:182-200
        case PLTB_BSDF_CONDUCTOR:  // conductor.cpp:320-376
            if (!(wi.z > 0.f)) return s4(0.f);
            return tex_eval(sc, b.tex[2], wl, uv) * conductor_F(sc, b, wl, uv, wi.z);
        case PLTB_BSDF_DIELECTRIC: {  // dielectric.cpp:404-590
            S4 w = s4(1.f);
            if (wo.z > 0.0f) { if (b.tex[2] >= 0) w = w * tex_eval(sc, b.tex[2], wl, uv); }
            else { if (b.tex[3] >= 0) w = w * tex_eval(sc, b.tex[3], wl, uv); }
            return w;
        }
        case PLTB_BSDF_PRINCIPLED: {  // base wbsdf_weight: eval / pdf (bsdf.cpp:85-97)
            float p = principled_pdf(b, wi, wo);
            return p > 0.0f ? principled_eval(sc, b, wi, uv, wl, wo) / p : s4(0.f);
        }
        case PLTB_BSDF_ROUGHCONDUCTOR:
        case PLTB_BSDF_ROUGHGRATING: {  // base wbsdf_weight, src/render/bsdf.cpp:85-97
            S4 e; float p;
            rough_eval_pdf(sc, b, wi, uv, wl, wo, type == PLTB_BSDF_ROUGHCONDUCTOR, e, p);
            return p > 0.0f ? e / p : s4(0.f);
        }
        default: return s4(0.f);
    }
}

// TwoSidedBRDF with a single nested BRDF (src/bsdfs/twosided.cpp:262-420)
PLTB_HD const pltb_bsdf *si_bsdf(const DeviceScene &sc, int32_t shape) {
    if (shape < 0) return nullptr;
    int32_t id = sc.shapes[shape].bsdf;
    return id < 0 ? nullptr : sc.bsdfs + id;
}
// the material class of a queue is the BSDF plugin type (one exception below): a compile-time constant inside k_shade<Class> / k_nee_post<Class>, so the
// switch statements of the *_inner functions keep one case; Class < 0: generic (type read from the record)
template <int Class> PLTB_HD uint32_t class_type(const pltb_bsdf *b) {
#ifndef PLTB_CLASS_PER_TYPE
    if (Class == PLTB_BSDF_CONDUCTOR) {   // the shared class of the delta BSDFs (pltb_bvh.h: material_class)
        // a two-valued expression rather than b->type itself: the switch statements of the *_inner functions then keep two cases
        return b->type == PLTB_BSDF_DIELECTRIC ? (uint32_t) PLTB_BSDF_DIELECTRIC : (uint32_t) PLTB_BSDF_CONDUCTOR;
    }
#endif
    if (Class >= 0) return (uint32_t) Class;
    return b->type;
}
template <int Class = -1>
PLTB_HD BSDFSampleOut wbsdf_sample(const DeviceScene &sc, const pltb_bsdf *b, V3 wi, V2 uv, const S4 &wl, float s1, V2 s2,
                                   V2 lobe_s2) {
    if (!b) { BSDFSampleOut o; o.wo = mk3(0, 0, 0); o.pdf = 0.f; o.eta = 0.f; o.type = 0; o.weight = s4(0.f); o.aux = 0.f; return o; }
    const uint32_t type = class_type<Class>(b);
    if (b->flags & PLTB_BSDF_FLAG_TWOSIDED) {
        float wz = wi.z;
        wi.z = fabsf(wi.z);
        BSDFSampleOut o = bsdf_sample_inner(sc, *b, type, wi, uv, wl, s1, s2, lobe_s2);
        o.wo.z = mulsign(o.wo.z, wz);
        return o;
    }
    return bsdf_sample_inner(sc, *b, type, wi, uv, wl, s1, s2, lobe_s2);
}
PLTB_HD void twosided_flip(const pltb_bsdf *b, V3 &wi, V3 &wo) {
    if (b->flags & PLTB_BSDF_FLAG_TWOSIDED) { wo.z = mulsign(wo.z, wi.z); wi.z = fabsf(wi.z); }
}
template <int Class = -1, bool Rgb = false>
PLTB_HD S4 wbsdf_eval(const DeviceScene &sc, const pltb_bsdf *b, V3 wi, V2 uv, const S4 &wl, V3 wo) {
    if (!b) return s4(0.f);
    twosided_flip(b, wi, wo);
    return bsdf_eval_inner<Rgb>(sc, *b, class_type<Class>(b), wi, uv, wl, wo);
}
template <int Class = -1>
PLTB_HD float wbsdf_pdf(const DeviceScene &sc, const pltb_bsdf *b, V3 wi, V2 uv, const S4 &wl, V3 wo) {
    if (!b) return 0.f;
    twosided_flip(b, wi, wo);
    return bsdf_pdf_inner(sc, *b, class_type<Class>(b), wi, uv, wl, wo);
}
template <int Class = -1>
PLTB_HD S4 wbsdf_weight(const DeviceScene &sc, const pltb_bsdf *b, V3 wi, V2 uv, const S4 &wl, V3 wo) {
    if (!b) return s4(0.f);
    twosided_flip(b, wi, wo);
    return bsdf_weight_inner(sc, *b, class_type<Class>(b), wi, uv, wl, wo);
}

}  // namespace pltb
