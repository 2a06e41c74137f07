// pltb_envmap.cuh -- EnvironmentMapEmitter (src/emitters/envmap.cpp:343-580) and the Hierarchical2D<Float, 0> warp it samples
// (include/mitsuba/core/distr_2d.h:475-700), over the flattened block built at load time (mitsuba3_plt_b200/distr2d.py).
#pragma once
#include "pltb_scene.cuh"

namespace pltb {

PLTB_HD float clip01(float x) { return fminf(fmaxf(x, 0.f), 1.f); }
PLTB_HD float lerp_(float a, float b, float t) { return fmadd(b, t, fnmadd(a, t, a)); }   // dr::lerp
// warp::interval_to_linear (include/mitsuba/core/warp.h:447-453)
PLTB_HD float interval_to_linear(float v0, float v1, float sample) {
    if (fabsf(v0 - v1) > 1e-4f * (v0 + v1)) return (v0 - safe_sqrt(lerp_(sqr(v0), sqr(v1), sample))) / (v0 - v1);
    return sample;
}

struct Hier2D {
    const float *blk;
    uint32_t n_levels, res_x, res_y;
    PLTB_HD explicit Hier2D(const float *b) : blk(b), n_levels(f2u(PLTB_LDG(b))), res_x(f2u(PLTB_LDG(b + 1))), res_y(f2u(PLTB_LDG(b + 2))) {}
    PLTB_HD const float *level(uint32_t l) const { return blk + f2u(PLTB_LDG(blk + 4 + 2 * l)); }
    PLTB_HD uint32_t width(uint32_t l) const { return f2u(PLTB_LDG(blk + 5 + 2 * l)); }
    PLTB_HD static uint32_t index(uint32_t x, uint32_t y, uint32_t w) {   // Level::index (distr_2d.h:772-776)
        return ((x & 1u) | (((x & ~1u) | (y & 1u)) << 1)) + (y & ~1u) * w;
    }
    // sample() (distr_2d.h:481-563)
    PLTB_HD void sample(V2 s, V2 &out, float &pdf) const {
        s.x = clip01(s.x); s.y = clip01(s.y);
        uint32_t ox = 0, oy = 0;
        for (int l = (int) n_levels - 2; l > 0; --l) {
            const float *lv = level((uint32_t) l);
            ox <<= 1; oy <<= 1;
            const uint32_t i = index(ox, oy, width((uint32_t) l));
            const float v00 = PLTB_LDG(lv + i), v10 = PLTB_LDG(lv + i + 1), v01 = PLTB_LDG(lv + i + 2), v11 = PLTB_LDG(lv + i + 3);
            s.x = clip01(s.x); s.y = clip01(s.y);
            const float r0 = v00 + v10, r1 = v01 + v11;
            s.y *= r0 + r1;
            bool m = s.y > r0;
            if (m) { oy += 1; s.y -= r0; }
            s.y /= m ? r1 : r0;
            const float c0 = m ? v01 : v00, c1 = m ? v11 : v10;
            s.x *= c0 + c1;
            m = s.x > c0;
            if (m) s.x -= c0;
            s.x /= m ? c1 : c0;
            if (m) ox += 1;
        }
        const float *l0 = level(0);
        const uint32_t w = width(0), i = ox + oy * w;
        const float v00 = PLTB_LDG(l0 + i), v10 = PLTB_LDG(l0 + i + 1), v01 = PLTB_LDG(l0 + i + w), v11 = PLTB_LDG(l0 + i + w + 1);
        // warp::square_to_bilinear (warp.h:480-494)
        const float r0 = v00 + v10, r1 = v01 + v11;
        s.y = interval_to_linear(r0, r1, s.y);
        const float c0 = lerp_(v00, v01, s.y), c1 = lerp_(v10, v11, s.y);
        s.x = interval_to_linear(c0, c1, s.x);
        pdf = lerp_(c0, c1, s.x);
        const float psx = 1.f / (float) (res_x - 1), psy = 1.f / (float) (res_y - 1);   // m_patch_size
        out.x = ((float) (int32_t) ox + s.x) * psx;
        out.y = ((float) (int32_t) oy + s.y) * psy;
    }
    // eval() (distr_2d.h:668-699)
    PLTB_HD float eval(V2 pos) const {
        pos.x = clip01(pos.x) * (float) (res_x - 1); pos.y = clip01(pos.y) * (float) (res_y - 1);
        uint32_t ox = (uint32_t) (int32_t) pos.x, oy = (uint32_t) (int32_t) pos.y;
        if (ox > res_x - 2) ox = res_x - 2;
        if (oy > res_y - 2) oy = res_y - 2;
        pos.x -= (float) (int32_t) ox; pos.y -= (float) (int32_t) oy;
        const float *l0 = level(0);
        const uint32_t w = width(0), i = ox + oy * w;
        const float v00 = PLTB_LDG(l0 + i), v10 = PLTB_LDG(l0 + i + 1), v01 = PLTB_LDG(l0 + i + w), v11 = PLTB_LDG(l0 + i + w + 1);
        return lerp_(lerp_(v00, v10, pos.x), lerp_(v01, v11, pos.x), pos.y);   // square_to_bilinear_pdf (warp.h:516-521)
    }
};

PLTB_HD float safe_acos(float x) { return acosf(fminf(fmaxf(x, -1.f), 1.f)); }
PLTB_HD V2 envmap_dir_to_uv(V3 v) {   // envmap.cpp:348-350
    return mk2(atan2f(v.x, -v.z) * (0.5f * kInvPi), safe_acos(v.y) * kInvPi);
}
PLTB_HD float envmap_inv_sin_theta(V3 d) {   // safe_rsqrt(max(d.x^2 + d.z^2, Epsilon^2)) (envmap.cpp:443-444)
    const float v = fmaxf(sqr(d.x) + sqr(d.z), sqr(kEps));
    return 1.f / sqrtf(fmaxf(v, 0.f));
}
// EnvironmentMapEmitter::eval_spectrum (envmap.cpp:524-580), include_whitepoint = true
PLTB_NI S4 envmap_eval_spectrum(const DeviceScene &sc, const pltb_emitter &em, V2 uv, const S4 &wl) {
    const uint32_t rx = em.env_res[0], ry = em.env_res[1];
    uv.x -= .5f / (float) (rx - 1u);
    uv.x -= floorf(uv.x); uv.y -= floorf(uv.y);
    uv.x *= (float) (rx - 1u); uv.y *= (float) (ry - 1u);
    uint32_t px = (uint32_t) uv.x, py = (uint32_t) uv.y;
    if (px > rx - 2u) px = rx - 2u;
    if (py > ry - 2u) py = ry - 2u;
    const float w1x = uv.x - (float) px, w1y = uv.y - (float) py, w0x = 1.f - w1x, w0y = 1.f - w1y;
    // texel records stay in global memory: srgb_model_eval reads its coefficients through the read-only path
    const float *v00 = sc.tex_data + em.env_data + 4 * (size_t) (py * rx + px), *v10 = v00 + 4, *v01 = v00 + 4 * (size_t) rx,
                *v11 = v01 + 4;
    if (sc.rgb) {   // envmap.cpp:569-579: texels are (r, g, b, unused); no whitepoint
        S4 result;
        for (int c = 0; c < 3; ++c) {
            const float v0 = fmadd(w0x, PLTB_LDG(v00 + c), w1x * PLTB_LDG(v10 + c)), v1 = fmadd(w0x, PLTB_LDG(v01 + c), w1x * PLTB_LDG(v11 + c));
            result.v[c] = fmadd(w0y, v0, w1y * v1) * em.env_scale;
        }
        result.v[3] = result.v[2];
        return result;
    }
    const float q00 = PLTB_LDG(v00 + 3), q10 = PLTB_LDG(v10 + 3), q01 = PLTB_LDG(v01 + 3), q11 = PLTB_LDG(v11 + 3);
    const S4 white = tex_eval(sc, em.radiance, wl, mk2(0.f, 0.f));
    const float f0 = fmadd(w0x, q00, w1x * q10), f1 = fmadd(w0x, q01, w1x * q11), f = fmadd(w0y, f0, w1y * f1);
    S4 result;
    for (int i = 0; i < 4; ++i) {
        const float s00 = srgb_model_eval(v00, wl.v[i]), s10 = srgb_model_eval(v10, wl.v[i]);
        const float s01 = srgb_model_eval(v01, wl.v[i]), s11 = srgb_model_eval(v11, wl.v[i]);
        const float s0 = fmadd(w0x, s00, w1x * s10), s1 = fmadd(w0x, s01, w1x * s11), sv = fmadd(w0y, s0, w1y * s1);
        result.v[i] = ((sv * f) * em.env_scale) * white.v[i];
    }
    return result;
}
// ::pdf_direction (envmap.cpp:468-484)
PLTB_NI float envmap_pdf_direction(const DeviceScene &sc, const pltb_emitter &em, V3 d_world) {
    const V3 d = xf_vector(em.to_local, d_world);
    V2 uv = envmap_dir_to_uv(d);
    uv.x -= .5f / (float) (em.env_res[0] - 1u);
    uv.x -= floorf(uv.x); uv.y -= floorf(uv.y);
    const Hier2D warp(sc.tex_data + em.env_warp);
    return warp.eval(uv) * envmap_inv_sin_theta(d) * (1.f / (2.f * sqr(kPi)));
}
// ::sample_direction (envmap.cpp:421-466) without the spectrum: direction, distance, uv and solid-angle density
PLTB_NI void envmap_sample_direction(const DeviceScene &sc, const pltb_emitter &em, V3 p, V2 sample, V3 &d_world, float &dist,
                                     V2 &uv, float &pdf_out) {
    const Hier2D warp(sc.tex_data + em.env_warp);
    float pdf;
    warp.sample(sample, uv, pdf);
    uv.x += .5f / (float) (em.env_res[0] - 1u);
    const bool active = pdf > 0.f;
    const float theta = uv.y * kPi, phi = uv.x * (2.f * kPi);
    const float st = sinf(theta), ct = cosf(theta), sp = sinf(phi), cp = cosf(phi);
    const V3 dl = mk3(sp * st, ct, -(cp * st));   // dr::sphdir(theta, phi) = (cp st, sp st, ct) -> (d.y, d.z, -d.x)
    const V3 c = mk3(em.bsphere_center[0], em.bsphere_center[1], em.bsphere_center[2]);
    dist = 2.f * fmaxf(em.bsphere_radius, norm(p - c));
    const float inv_sin_theta = envmap_inv_sin_theta(dl);
    d_world = xf_vector(em.to_world, dl);
    pdf_out = active ? pdf * inv_sin_theta * (1.f / (2.f * sqr(kPi))) : 0.f;
}

}  // namespace pltb
