// pltb_shading.cuh -- surface interaction reconstruction, microfacet / grating models and the wbsdf_* family.
// Reference (paths relative to the reference tree): src/render/mesh.cpp:1379-1551, include/mitsuba/render/
// interaction.h:140-170,278-289,497-517, include/mitsuba/render/microfacet.h:100-431, include/mitsuba/plt/
// diffractiongrating.h:32-290, src/bsdfs/{diffuse,conductor,dielectric,roughconductor,roughgrating,twosided}.cpp,
// src/render/bsdf.cpp:22-97.
#pragma once
#include "pltb_scene.cuh"

namespace pltb {

enum : uint32_t {
    F_Null = 0x1, F_DiffuseReflection = 0x2, F_GlossyReflection = 0x8, F_GlossyTransmission = 0x10, F_DeltaReflection = 0x20,
    F_DeltaTransmission = 0x40, F_Smooth = 0x2 | 0x4 | 0x8 | 0x10, F_Delta = 0x1 | 0x20 | 0x40
};

struct SI {
    float t;
    V3 p, n;
    V2 uv;
    Frame sh;
    V3 wi;
    int32_t shape;
    PLTB_HD bool valid() const { return t != inf_f(); }
    PLTB_HD V3 offset_p(V3 d) const {  // interaction.h:165-170
        float mag = (1.f + fmaxf(fmaxf(fabsf(p.x), fabsf(p.y)), fabsf(p.z))) * kRayEps;
        mag = mulsign(mag, dot(n, d));
        return fmadd3(n, mag, p);
    }
    PLTB_HD Ray spawn_ray(V3 d) const { Ray r; r.o = offset_p(d); r.d = d; r.maxt = kLargest; return r; }
    PLTB_HD Ray spawn_ray_to(V3 target) const {  // interaction.h:145-152
        Ray r;
        r.o = offset_p(target - p);
        V3 d = target - r.o;
        float dist = norm(d);
        r.d = d / dist;
        r.maxt = dist * (1.f - kShadowEps);
        return r;
    }
};

PLTB_NI SI make_si(const DeviceScene &sc, const Ray &ray, const Hit &h) {
    SI si;
    si.t = inf_f(); si.shape = -1;
    si.p = si.n = mk3(0, 0, 0); si.uv = mk2(0, 0);
    si.sh.s = si.sh.t = si.sh.n = mk3(0, 0, 0);
    if (h.shape < 0) { si.wi = -ray.d; return si; }
    const pltb_shape &sh = sc.shapes[h.shape];
    const uint32_t flags = sh.flags;
    si.shape = h.shape; si.t = h.t;
    V3 dp_du, dp_dv;
    if (sh.type == PLTB_SHAPE_MESH) {
        uint32_t f0 = PLTB_LDG(sc.faces + 3 * h.prim), f1 = PLTB_LDG(sc.faces + 3 * h.prim + 1),
                 f2 = PLTB_LDG(sc.faces + 3 * h.prim + 2);
        V3 p0 = ld3(sc.positions, f0), p1 = ld3(sc.positions, f1), p2 = ld3(sc.positions, f2);
        float b1 = h.u, b2 = h.v, b0 = 1.f - b1 - b2;
        si.p = fmadd3(p0, b0, fmadd3(p1, b1, p2 * b2));
        V3 dp0 = p1 - p0, dp1 = p2 - p0;
        si.n = normalize(cross(dp0, dp1));
        si.uv = mk2(b1, b2);
        coordinate_system(si.n, dp_du, dp_dv);
        if (flags & PLTB_SHAPE_HAS_UV) {
            V2 uv0 = ld2(sc.texcoords, f0), uv1 = ld2(sc.texcoords, f1), uv2 = ld2(sc.texcoords, f2);
            si.uv = mk2(fmadd(uv2.x, b2, fmadd(uv1.x, b1, uv0.x * b0)), fmadd(uv2.y, b2, fmadd(uv1.y, b1, uv0.y * b0)));
            V2 duv0 = mk2(uv1.x - uv0.x, uv1.y - uv0.y), duv1 = mk2(uv2.x - uv0.x, uv2.y - uv0.y);
            float det = fmsub(duv0.x, duv1.y, duv0.y * duv1.x), inv_det = rcp(det);
            if (det != 0.f) {
                dp_du = mk3(fmsub(duv1.y, dp0.x, duv0.y * dp1.x), fmsub(duv1.y, dp0.y, duv0.y * dp1.y),
                            fmsub(duv1.y, dp0.z, duv0.y * dp1.z)) * inv_det;
            }
        }
        if (flags & PLTB_SHAPE_HAS_NORMALS) {
            V3 n0 = ld3(sc.normals, f0), n1 = ld3(sc.normals, f1), n2 = ld3(sc.normals, f2);
            V3 n = fmadd3(n2, b2, fmadd3(n1, b1, n0 * b0));
            si.sh.n = n * rsqrt_(sqnorm(n));
        } else {
            si.sh.n = si.n;
        }
        if (flags & PLTB_SHAPE_FLIP_NORMALS) { si.n = -si.n; si.sh.n = -si.sh.n; }
    } else if (sh.type == PLTB_SHAPE_SPHERE) {  // src/shapes/sphere.cpp:438-505
        V3 c = mk3(sh.center[0], sh.center[1], sh.center[2]);
        si.sh.n = normalize(fmadd3(ray.d, h.t, ray.o) - c);
        si.p = fmadd3(si.sh.n, sh.radius, c);
        V3 local = xf_point(sh.to_object, si.p);
        float theta = acosf(fminf(fmaxf(local.z, -1.f), 1.f)), phi = atan2f(local.y, local.x);
        if (phi < 0.f) phi += 2.f * kPi;
        si.uv = mk2(phi * (0.5f * kInvPi), theta * kInvPi);
        dp_du = xf_vector(sh.to_world, mk3(-local.y, local.x, 0.f) * (2.f * kPi));
        if (flags & PLTB_SHAPE_FLIP_NORMALS) si.sh.n = -si.sh.n;
        si.n = si.sh.n;
    } else {  // src/shapes/rectangle.cpp:297-330
        si.p = xf_point(sh.to_world, mk3(h.u, h.v, 0.f));
        si.n = mk3(sh.frame_n[0], sh.frame_n[1], sh.frame_n[2]);
        si.sh.n = si.n;
        dp_du = xf_vector(sh.to_world, mk3(2.f, 0.f, 0.f));
        si.uv = mk2(fmadd(h.u, .5f, .5f), fmadd(h.v, .5f, .5f));
        if (flags & PLTB_SHAPE_FLIP_NORMALS) { si.n = -si.n; si.sh.n = -si.sh.n; }
    }
    // initialize_sh_frame (interaction.h:278-289)
    float nd = dot(si.sh.n, dp_du);
    si.sh.s = normalize(fmadd3(si.sh.n, -nd, dp_du));
    if (dp_du.x == 0.f && dp_du.y == 0.f && dp_du.z == 0.f) { V3 tt; coordinate_system(si.sh.n, si.sh.s, tt); }
    si.sh.t = cross(si.sh.n, si.sh.s);
    si.wi = si.sh.to_local(-ray.d);
    return si;
}

// ---- MicrofacetDistribution ------------------------------------------------------------------------------------------------
struct Microfacet {
    uint32_t type; float au, av; bool sample_visible;
    PLTB_HD float eval(V3 m) const {
        float alpha_uv = au * av, cos_theta = m.z, cos_theta_2 = sqr(cos_theta), result;
        if (type == PLTB_MICROFACET_BECKMANN)
            result = expf(-(sqr(m.x / au) + sqr(m.y / av)) / cos_theta_2) / (kPi * alpha_uv * sqr(cos_theta_2));
        else
            result = rcp(kPi * alpha_uv * sqr(sqr(m.x / au) + sqr(m.y / av) + sqr(m.z)));
        return result * cos_theta > 1e-20f ? result : 0.f;
    }
    PLTB_HD float smith_g1(V3 v, V3 m) const {
        float xy_alpha_2 = sqr(au * v.x) + sqr(av * v.y), tan_theta_alpha_2 = xy_alpha_2 / sqr(v.z), result;
        if (type == PLTB_MICROFACET_BECKMANN) {
            float a = rsqrt_(tan_theta_alpha_2), a_sqr = sqr(a);
            result = a >= 1.6f ? 1.f : (3.535f * a + 2.181f * a_sqr) / (1.f + 2.276f * a + 2.577f * a_sqr);
        } else {
            result = 2.f / (1.f + sqrtf(1.f + tan_theta_alpha_2));
        }
        if (xy_alpha_2 == 0.f) result = 1.f;
        if (dot(v, m) * v.z <= 0.f) result = 0.f;
        return result;
    }
    PLTB_HD float G(V3 wi, V3 wo, V3 m) const { return smith_g1(wi, m) * smith_g1(wo, m); }
    PLTB_HD float pdf(V3 wi, V3 m) const {
        float result = eval(m);
        if (sample_visible) result *= smith_g1(wi, m) * fabsf(dot(wi, m)) / wi.z;
        else result *= m.z;
        return result;
    }
    PLTB_HD V2 sample_visible_11(float cos_theta_i, V2 sample) const {
        if (type == PLTB_MICROFACET_BECKMANN) {
            float tan_theta_i = safe_sqrt(fnmadd(cos_theta_i, cos_theta_i, 1.f)) / cos_theta_i;
            float cot_theta_i = rcp(tan_theta_i);
            float maxval = dr_erf(cot_theta_i);
            sample.x = fmaxf(fminf(sample.x, 1.f - 1e-6f), 1e-6f);
            sample.y = fmaxf(fminf(sample.y, 1.f - 1e-6f), 1e-6f);
            float x = maxval - (maxval + 1.f) * dr_erf(sqrtf(-logf(sample.x)));
            sample.x *= 1.f + maxval + kInvSqrtPi * tan_theta_i * expf(-sqr(cot_theta_i));
            for (int i = 0; i < 3; ++i) {
                float slope = dr_erfinv(x),
                      value = 1.f + x + kInvSqrtPi * tan_theta_i * expf(-sqr(slope)) - sample.x,
                      derivative = 1.f - slope * tan_theta_i;
                x -= value / derivative;
            }
            return mk2(dr_erfinv(x), dr_erfinv(fmsub(2.f, sample.y, 1.f)));
        } else {
            V2 p = square_to_uniform_disk_concentric(sample);
            float s = 0.5f * (1.f + cos_theta_i);
            float la = safe_sqrt(1.f - sqr(p.x));
            p.y = fmadd(p.y, s, fnmadd(la, s, la));
            float x = p.x, y = p.y, z = safe_sqrt(1.f - fmadd(p.y, p.y, p.x * p.x));
            float sin_theta_i = safe_sqrt(1.f - sqr(cos_theta_i));
            float nrm = rcp(fmadd(sin_theta_i, y, cos_theta_i * z));
            return mk2(fmsub(cos_theta_i, y, sin_theta_i * z) * nrm, x * nrm);
        }
    }
    PLTB_NI V3 sample(V3 wi, V2 sample, float &pdf_out) const {
        if (!sample_visible) {
            float sin_phi, cos_phi, cos_theta, cos_theta_2, alpha_2;
            if (au == av) {
                float ang = (2.f * kPi) * sample.y;
                sin_phi = sinf(ang); cos_phi = cosf(ang);
                alpha_2 = au * au;
            } else {
                float ratio = av / au, tmp = ratio * tanf((2.f * kPi) * sample.y);
                cos_phi = rsqrt_(fmadd(tmp, tmp, 1.f));
                cos_phi = mulsign(cos_phi, fabsf(sample.y - .5f) - .25f);
                sin_phi = cos_phi * tmp;
                alpha_2 = rcp(sqr(cos_phi / au) + sqr(sin_phi / av));
            }
            if (type == PLTB_MICROFACET_BECKMANN) {
                cos_theta = rsqrt_(fnmadd(alpha_2, logf(1.f - sample.x), 1.f));
                cos_theta_2 = sqr(cos_theta);
                float cos_theta_3 = fmaxf(cos_theta_2 * cos_theta, 1e-20f);
                pdf_out = (1.f - sample.x) / (kPi * au * av * cos_theta_3);
            } else {
                float tan_theta_m_2 = alpha_2 * sample.x / (1.f - sample.x);
                cos_theta = rsqrt_(1.f + tan_theta_m_2);
                cos_theta_2 = sqr(cos_theta);
                float temp = 1.f + tan_theta_m_2 / alpha_2, cos_theta_3 = fmaxf(cos_theta_2 * cos_theta, 1e-20f);
                pdf_out = rcp(kPi * au * av * cos_theta_3 * sqr(temp));
            }
            float sin_theta = sqrtf(1.f - cos_theta_2);
            return mk3(cos_phi * sin_theta, sin_phi * sin_theta, cos_theta);
        }
        V3 wi_p = normalize(mk3(au * wi.x, av * wi.y, wi.z));
        float sin_theta_2 = fmadd(wi_p.x, wi_p.x, sqr(wi_p.y)), inv_sin_theta = rsqrt_(sin_theta_2);
        float cos_phi, sin_phi;
        if (fabsf(sin_theta_2) <= 4.f * kEps) { cos_phi = 1.f; sin_phi = 0.f; }
        else {
            cos_phi = fminf(fmaxf(wi_p.x * inv_sin_theta, -1.f), 1.f);
            sin_phi = fminf(fmaxf(wi_p.y * inv_sin_theta, -1.f), 1.f);
        }
        V2 slope = sample_visible_11(wi_p.z, sample);
        slope = mk2(fmsub(cos_phi, slope.x, sin_phi * slope.y) * au, fmadd(sin_phi, slope.x, cos_phi * slope.y) * av);
        V3 m = normalize(mk3(-slope.x, -slope.y, 1.f));
        pdf_out = eval(m) * smith_g1(wi, m) * fabsf(dot(wi, m)) / wi.z;
        return m;
    }
};
PLTB_HD Microfacet make_mf(const pltb_bsdf &b) {
    Microfacet m; m.type = b.u[0]; m.au = fmaxf(b.f[1], 1e-4f); m.av = fmaxf(b.f[2], 1e-4f); m.sample_visible = b.u[1] != 0;
    return m;
}

// ---- DiffractionGrating -------------------------------------------------------------------------------------------------------
struct Grating {
    V2 dir, inv_period; float q, multiplier; uint32_t lobes, type;
    PLTB_HD bool is_1d() const { return inv_period.y < kEps; }
    PLTB_HD float alpha(V3 wi, float k) const { return expf(-sqr(wi.z * q * k)); }
    // squared order-l term of lobe_intensity for one axis given a = 4 pi q / (wl |wi.z|)  (:228-272)
    PLTB_HD float axis_term(const BesselTables &bt, float a, int l) const {
        if (l == 0) return 1.0f;
        uint32_t gtype = type & 0xFu;
        float fl = (float) l;
        if (gtype == 0) return sqr(bessel_j(bt, a, l));
        if (gtype == 1) return sinf(a / 2.0f) * sinc_(kPi * fl / 2.0f);
        return 1.f / sqrtf(fabsf(fl));
    }
    // the same for sample_lobe's selection weights (sinusoidal gratings: reduced-precision Bessel, see bessel_j_fast)
    PLTB_HD float axis_term_select(const BesselTables &bt, float a, int l) const {
#ifndef PLTB_EXACT_SELECT
        if (l != 0 && (type & 0xFu) == 0) {
            float j;
            if (bessel_j_tab(bt, a, l, j)) return sqr(j);   // tabulated 20-node rule (BesselTables::tab)
            return sqr(bessel_j_fast(bt, a, l));
        }
#endif
        return axis_term(bt, a, l);
    }
    // lobe_intensity for the lobe RoughGrating::wbsdf_sample has just drawn (roughgrating.cpp:440, the nanometre-unit call N2): the
    // value scales the sample-phase throughput only, never the replayed one, so sinusoidal gratings read the tabulated rule
    PLTB_HD float lobe_intensity_sampled(const BesselTables &bt, int lx, int ly, V3 wi, float wl) const {
#ifndef PLTB_EXACT_SELECT
        if ((type & 0xFu) == 0) {
            const float a = kFourPi * q / (wl * fabsf(wi.z));
            float jx = 1.f, jy = 1.f;
            const bool okx = lx == 0 || bessel_j_tab(bt, a, lx, jx), oky = ly == 0 || is_1d() || bessel_j_tab(bt, a, ly, jy);
            if (okx && oky) {
                const float ix = lx == 0 ? 1.0f : sqr(jx);
                const float iy = is_1d() ? ix : (ly == 0 ? 1.0f : sqr(jy));
                return multiplier * ix * iy;
            }
        }
#endif
        return lobe_intensity(bt, lx, ly, wi, wl);
    }
    PLTB_HD float lobe_intensity(const BesselTables &bt, int lx, int ly, V3 wi, float wl) const {
        const float a = kFourPi * q / (wl * fabsf(wi.z));
#if defined(__CUDA_ARCH__) && !defined(PLTB_BESSEL_SEQUENTIAL)
        if ((type & 0xFu) == 0 && !is_1d()) {   // sinusoidal, 2-D: both orders in one pass (same values as the two calls below)
            if (lx == 0 && ly == 0) return multiplier * 1.0f * 1.0f;
            float jx, jy;
            bessel_j_pair(bt, a, lx, ly, jx, jy);
            const float ix = lx == 0 ? 1.0f : sqr(jx), iy = ly == 0 ? 1.0f : sqr(jy);
            return multiplier * ix * iy;
        }
#endif
        float ix = axis_term(bt, a, lx);
        float iy = is_1d() ? ix : axis_term(bt, a, ly);
        return multiplier * ix * iy;
    }
    PLTB_HD void sample_lobe(const BesselTables &bt, V2 sample2, V3 wi, float wl, int &lobe_x, int &lobe_y, V2 &pdf) const {
        float intensities[PLTB_MAX_LOBES];
        float total = 0.0f;
        const uint32_t n = lobes / 2 + 1;
        for (uint32_t l = 0; l < n; ++l) {
            // lobe_intensity((l, 0)): the y term is order 0 (== 1) unless the grating is 1-D, where it repeats the x term
            const float a = kFourPi * q / (wl * fabsf(wi.z)), ix = axis_term_select(bt, a, (int) l);
            float li = multiplier * ix * (is_1d() ? ix : 1.0f);
            if (l == 0) li /= 2.0f;
            total += li;
            intensities[l] = li;
        }
        float cdf = 0.0f, lx = 0.f, ly = 0.f;
        pdf = mk2(0.f, 0.f);
        V2 rn = mk2((sample2.x - 0.5f) * 2.0f, (sample2.y - 0.5f) * 2.0f);
        for (uint32_t l = 0; l < n; ++l) {
            float p = intensities[l] / total;
            bool s1 = fabsf(rn.x) > cdf, s2 = fabsf(rn.y) > cdf;
            if (s1) { pdf.x = l == 0 ? p : p / 2.0f; lx = (float) l; }
            if (s2) { pdf.y = l == 0 ? p : p / 2.0f; ly = (float) l; }
            cdf += p;
        }
        lobe_x = (int) (lx * sign1(rn.x));
        lobe_y = (int) (ly * sign1(rn.y));
    }
    // direction-independent part of diffract(): sin of the incident angles (:204-209)
    PLTB_HD V2 sin_i(V3 wi) const {
        float px = sqrtf(sqr(wi.x) + sqr(wi.z)), py = sqrtf(sqr(wi.y) + sqr(wi.z));
        return mk2(px > kEps ? wi.x / px : 0.0f, py > kEps ? wi.y / py : 0.0f);
    }
    PLTB_HD V3 diffract_s(V2 si_, int lx, int ly, float wl, bool &valid) const {  // (:211-225)
        float cos_g = dir.x, sin_g = dir.y;
        float lrx = cos_g * (float) lx - sin_g * (float) ly, lry = sin_g * (float) lx + cos_g * (float) ly;
        float a = wl * lrx * inv_period.x - si_.x, b = wl * lry * inv_period.y - si_.y;
        float m = (sqr(a) - 1) / (sqr(a * b) - 1);
        float qq = 1 - sqr(b) * m;
        valid = fabsf(a) <= 1.0f && fabsf(b) <= 1.0f;
        return mk3(a * sqrtf(fmaxf(0.0f, qq)), b * sqrtf(m), sqrtf(fmaxf(0.0f, 1 - sqr(a) * qq - sqr(b) * m)));
    }
    PLTB_HD V3 diffract(V3 wi, int lx, int ly, float wl, bool &valid) const { return diffract_s(sin_i(wi), lx, ly, wl, valid); }
};
PLTB_HD Grating make_grating(const pltb_bsdf &b, V2 uv) {
    Grating g;
    g.inv_period = mk2(b.f[4], b.f[5]); g.q = b.f[6]; g.multiplier = b.f[7]; g.lobes = b.u[2]; g.type = b.u[3];
    float angle = b.f[3];
    if ((g.type & 0x10u) == 0) {
        g.dir = mk2(cosf(angle), sinf(angle));
    } else {
        V2 r = mk2(uv.x - 0.5f, uv.y - 0.5f);
        float il = rsqrt_(fmadd(r.y, r.y, r.x * r.x));
        r = mk2(r.x * il, r.y * il);
        float c = cosf(angle), s = sinf(angle), vx = r.x, vy = -r.y;
        g.dir = mk2(fmadd(s, vy, c * vx), fmadd(c, vy, -s * vx));
    }
    return g;
}

}  // namespace pltb
