// pltb_stages_pol.cuh -- wavefront stages of the polarized variants (PLTB_RENDER_POLARIZED): the same sample-phase
// loop and solve phase as pltb_stages.cuh with Spectrum = Mueller matrix.  Ray generation, traversal, classification, emitter
// sampling and shadow rays are shared with the unpolarized path (none of them touches the Spectrum type); what changes is
//   stage_shade_pol     BSDF weights are 4x4 matrices per wavelength; throughputs are tracked by their first column
//                       (alpha = W * alpha, plt.py:123; replay alpha_i = W_0 (W_1 (... e_0)), plt.py:440-470)
//   stage_nee_post_pol  em_weight * mis_em * bsdf_val * alpha (plt.py:300) reduces to the (0, 0) entry: row 0 of bsdf_val times
//                       the first column of alpha_i
//   gather / stokes     first column of L, rotated to the sensor frame, 12 Stokes AOVs (stokes.cpp:88-129)
#pragma once
#include "pltb_stages.cuh"
#include "pltb_polarized.cuh"

namespace pltb {

PLTB_HD void pol_store_col(float4 *base, size_t stride, size_t r, const C4 a[4]) {   // a[wavelength].v[component] -> [component] S4
    for (int c = 0; c < 4; ++c) base[(size_t) c * stride + r] = mkf4(a[0].v[c], a[1].v[c], a[2].v[c], a[3].v[c]);
}
PLTB_HD void pol_load_col(const float4 *base, size_t stride, size_t r, C4 a[4]) {
    for (int c = 0; c < 4; ++c) {
        const float4 v = base[(size_t) c * stride + r];
        a[0].v[c] = v.x; a[1].v[c] = v.y; a[2].v[c] = v.z; a[3].v[c] = v.w;
    }
}

// The Stokes-frame rotation StokesIntegrator::sample applies to the result (stokes.cpp:97-106): from the implicit basis of -ray.d
// to cross(ray.d, sensor "up"); only the angle depends on the sample, so it is computed with the primary ray and kept.
PLTB_HD float stokes_rotation_angle(const pltb_sensor &sn, V3 ray_d, bool batch) {
    const V3 current_basis = stokes_basis(-ray_d);
    const V3 vertical = batch ? mk3(0.f, 1.f, 0.f) : xf_vector(sn.to_world, mk3(0.f, 1.f, 0.f));
    const V3 target_basis = cross(ray_d, vertical);
    float theta = unit_angle(normalize(current_basis), normalize(target_basis));
    if (dot(-ray_d, cross(current_basis, target_basis)) < 0) theta *= -1.f;
    return theta;
}

// first column of W_0 (W_1 (... (W_{i-1} I))) from the stored replay weights (replay_path, plt.py:440-470).
// `flags` = p_flags of the path: bit 1 + j says that W_j is a scaled identity (the replay weight of `diffuse`, diffuse.cpp:182-200),
// stored as ONE float4 of the four per-wavelength factors instead of sixteen: for finite operands diag(s) * a is s * a component by
// component (the fmadd chain of f44_apply only adds exact zeros), and a closed box is mostly diffuse walls -- the O(depth^2) re-read
// of the weights was the largest stream of the polarized shading kernel (256 B per earlier vertex).
PLTB_HD void pol_replay_column(const PassState &st, uint32_t local, uint32_t i, uint32_t flags, C4 ai[4]) {
    for (int k = 0; k < 4; ++k) { ai[k].v[0] = 1.f; ai[k].v[1] = ai[k].v[2] = ai[k].v[3] = 0.f; }
    for (int j = (int) i - 1; j >= 0; --j) {
        const size_t rj = w_index(st, local, (uint32_t) j);
        if (flags >> (1 + j) & 1u) {
            const float4 s = st.pv_w[rj];
            const float sv[4] = { s.x, s.y, s.z, s.w };
            float fsum = 0.f;
            for (int k = 0; k < 4; ++k) fsum += (ai[k].v[0] + ai[k].v[1]) + (ai[k].v[2] + ai[k].v[3]);
            if (finite_(fsum)) {
                for (int k = 0; k < 4; ++k) for (int c = 0; c < 4; ++c) ai[k].v[c] = sv[k] * ai[k].v[c];
            } else {   // 0 * inf / NaN terms of the full product must show up as they do in the reference
                for (int k = 0; k < 4; ++k) ai[k] = f44_apply(f44_diag(sv[k]), ai[k]);
            }
            continue;
        }
        for (int k = 0; k < 4; ++k) {
            F44 W;
            for (int row = 0; row < 4; ++row) {
                const float4 v = st.pv_w[(size_t) (4 * k + row) * st.nv + rj];
                W.m[row][0] = v.x; W.m[row][1] = v.y; W.m[row][2] = v.z; W.m[row][3] = v.w;
            }
            ai[k] = f44_apply(W, ai[k]);
        }
    }
}

// bounce record of a path that left the scene at bounce i: nothing, or the environment's emission alpha_i * depolarizer(mis * Le)
PLTB_NI void pol_miss_record(const DeviceScene &sc, const PassState &st, uint32_t local, uint32_t i) {
    const size_t r = (size_t) i * st.chunk + local;   // (only called when the scene has an environment emitter)
    C4 em[4];
    const S4 Lem = env_miss_lem(sc, st, local, i);
    C4 ai[4];
    pol_replay_column(st, local, i, i > 0 ? (uint32_t) st.p_flags[local] : 0u, ai);
    for (int k = 0; k < 4; ++k) for (int c = 0; c < 4; ++c) em[k].v[c] = ai[k].v[c] * Lem.v[k];
    pol_store_col(st.pv_em, st.nv, r, em);
}

template <int Class, class Push>
PLTB_HD void stage_shade_pol(const DeviceScene &sc, const PassState &st, uint32_t local, uint32_t i, uint32_t slot, uint32_t live_pos, Push &push, uint32_t cnt[4]) {
    const float4 o4 = st.p_ray_o[local], d4 = st.p_ray_d[local], h4 = st.p_hit[local];
    Ray ray; ray.o = mk3(o4.x, o4.y, o4.z); ray.d = mk3(d4.x, d4.y, d4.z); ray.maxt = o4.w;
    Hit h; h.t = h4.x; h.u = h4.y; h.v = h4.z; h.prim = f2u(h4.w) & kHitPrimMask; h.cls = (uint32_t) Class;
    h.shape = h.prim < sc.n_faces ? (int32_t) PLTB_LDG(sc.face_shape + h.prim) : (int32_t) (h.prim - sc.n_faces);
    const S4 wl = f4_to_s4(st.wl[local]);
    PCG32 rng; rng.state = st.rng_state[local]; rng.inc = st.rng_inc[local];
    C4 alpha[4];
    for (int k = 0; k < 4; ++k) { alpha[k].v[0] = 1.f; alpha[k].v[1] = alpha[k].v[2] = alpha[k].v[3] = 0.f; }   // Spectrum(1): identity
    float eta = 1.f, last_nd_pdf = 1.f;
    bool prev_bsdf_delta = true;
    V3 prev_p = mk3(0.f, 0.f, 0.f);
    uint64_t flags = 0ull;   // PassState::p_flags: bit 0 prev_bsdf_delta, bits 1 + j scaled-identity replay weights (pol_replay_column), presence masks
    if (i > 0) {
        pol_load_col(st.pp_alpha, st.chunk, local, alpha);
        float4 pv = st.p_prev[local];
        prev_p = mk3(pv.x, pv.y, pv.z); last_nd_pdf = pv.w;
        eta = d4.w;
        flags = st.p_flags[local];
        prev_bsdf_delta = (flags & 1ull) != 0;
    }
    SI si = make_si(sc, ray, h);
    bool active_next = (i + 1 < st.max_depth);
    float sample_1 = rng.next_float32();
    V2 sample_2, lobe_sample_2;
    sample_2.x = rng.next_float32(); sample_2.y = rng.next_float32();
    lobe_sample_2.x = rng.next_float32(); lobe_sample_2.y = rng.next_float32();
    const pltb_bsdf *b = si_bsdf(sc, si.shape);
    BSDFSampleOut so = wbsdf_sample<Class>(sc, b, si.wi, si.uv, wl, sample_1, sample_2, lobe_sample_2);
    PolVertex pv;
    pol_prepare(sc, b, si.wi, si.uv, wl, sample_2, so, pv);

    // replay_path: first column of W_0 (W_1 (... (W_{i-1} I)))
    C4 ai[4];
    pol_replay_column(st, local, i, (uint32_t) flags, ai);
    // solve_replay_emissive: Li = alpha * (mis_bsdf * depolarizer(Le)): first column = alpha[:, 0] * (mis * Le)
    const size_t r = (size_t) i * st.chunk + local;
    C4 em[4];
    for (int k = 0; k < 4; ++k) for (int c = 0; c < 4; ++c) em[k].v[c] = 0.f;
    const int32_t em_id = sc.shapes[si.shape].emitter;
    if (em_id >= 0) {
        const pltb_emitter &e = sc.emitters[em_id];
        float em_pdf = 0.f;
        if (!prev_bsdf_delta) {
            V3 rel = si.p - prev_p;
            float dist = norm(rel);
            V3 d = rel / dist;
            if (dot(d, si.sh.n) < 0.f) em_pdf = shape_pdf_direction(sc.shapes[e.shape], prev_p, d, si.sh.n, dist);
            em_pdf *= sc.emitter_pmf;
        }
        const float mis_bsdf = mis_weight(last_nd_pdf, em_pdf);
        const S4 Lem = (si.wi.z > 0.f ? tex_eval(sc, e.radiance, wl, si.uv) : s4(0.f)) * mis_bsdf;
        for (int k = 0; k < 4; ++k) {   // (alpha * D)(c, 0) = alpha(c, 0) * Lem, then three fmadds with the zero entries of D
            for (int c = 0; c < 4; ++c) {
                float v = ai[k].v[c] * Lem.v[k];
                em[k].v[c] = v;
            }
        }
    }
    if (em_id >= 0) { pol_store_col(st.pv_em, st.nv, r, em); flags |= 1ull << (kFlagEm + i); }   // other vertices have no emissive term
    if (active_next) {
        const size_t rw = w_index(st, local, i);
        if (pv.type == PLTB_BSDF_DIFFUSE) {   // diag(reflectance): the four factors, flagged in p_flags
            st.pv_w[rw] = s4_to_f4(pv.refl);
            flags |= 1ull << (1 + i);
        } else {
            for (int k = 0; k < 4; ++k) {
                const F44 W = pol_replay_weight(pv, k);
                for (int row = 0; row < 4; ++row)
                    st.pv_w[(size_t) (4 * k + row) * st.nv + rw] = mkf4(W.m[row][0], W.m[row][1], W.m[row][2], W.m[row][3]);
            }
        }
    }
    cnt[2]++;
    if (Class == PLTB_BSDF_ROUGHGRATING) cnt[3]++;

    Ray next = si.spawn_ray(si.sh.to_world(so.wo));
    if (active_next) {
        for (int k = 0; k < 4; ++k) alpha[k] = f44_apply(pol_sample_weight(pv, k), alpha[k]);   // bsdf_weight.L * alpha
        eta = so.eta;
    }
    float alpha_max = fmaxf(fmaxf(alpha[0].v[0], alpha[1].v[0]), fmaxf(alpha[2].v[0], alpha[3].v[0]));   // max(unpolarized_spectrum(alpha))
    active_next = active_next && alpha_max != 0.f;
    float rr_prob = fminf(alpha_max * sqr(eta), 0.95f);
    bool rr_active = (i + 1) >= st.rr_depth;
    if (rr_active) { const float rcp_thp = rcp(rr_prob); for (int k = 0; k < 4; ++k) for (int c = 0; c < 4; ++c) alpha[k].v[c] *= rcp_thp; }
    bool rr_continue = rng.next_float32() < rr_prob;
    active_next = active_next && (!rr_active || rr_continue);

    constexpr bool kSmoothClass = (kSmoothClassMask >> Class & 1u) != 0;
    const bool nee = kSmoothClass && (so.type & F_Smooth) != 0 && sc.n_emitters > 0;
    if (nee) {
        flags |= 1ull << (kFlagNee + i);
        st.v_p_u[slot] = mkf4(si.p.x, si.p.y, si.p.z, si.uv.x);
        st.v_n_v[slot] = mkf4(si.n.x, si.n.y, si.n.z, si.uv.y);
        st.v_s_wix[slot] = mkf4(si.sh.s.x, si.sh.s.y, si.sh.s.z, si.wi.x);
        st.v_shn_wiy[slot] = mkf4(si.sh.n.x, si.sh.n.y, si.sh.n.z, si.wi.y);
        pol_store_col(st.pv_alpha, st.nv, slot, ai);
        float2 ws; ws.x = si.wi.z; ws.y = u2f((uint32_t) si.shape);
        st.v_wiz_shape[slot] = ws;
        st.rec_entry[slot] = local | (i << 27);
    } else if (kSmoothClass && sc.n_emitters > 0) {
        st.rec_entry[slot] = kNoRecord;
    }
    st.rng_state[local] = rng.state;
    if (active_next) {
        const bool delta = (so.type & F_Delta) != 0;
        st.p_ray_o[local] = mkf4(next.o.x, next.o.y, next.o.z, next.maxt);
        st.p_ray_d[local] = mkf4(next.d.x, next.d.y, next.d.z, eta);
        pol_store_col(st.pp_alpha, st.chunk, local, alpha);
        st.p_prev[local] = mkf4(si.p.x, si.p.y, si.p.z, delta ? last_nd_pdf : so.pdf);
        st.p_flags[local] = (flags & ~1ull) | (delta ? 1ull : 0ull);
    } else {
        st.v_mask[local] = pack_mask(flags);
        st.n_vertices[local] = i + 1;
    }
    if (i + 1 < st.max_depth) st.live[live_pos] = active_next ? local : kNoPath;
}

// solve_replay_NEE's BSDF half in a polarized variant.  em_weight is depolarizer(e), so ((e mis) * B) has only row 0,
// (e mis) B(0, j), and its product with alpha contributes only to L(0, 0): sum_j X(0, j) alpha(j, 0) in fmadd order.
template <int Class, bool Rgb>
PLTB_HD void stage_nee_post_pol(const DeviceScene &sc, const PassState &st, uint32_t slot) {   // Rgb: see bsdf_eval_inner
    const uint32_t entry = st.rec_entry[slot], local = entry & 0x7FFFFFFu, depth = entry >> 27;
    const size_t r = slot, rv = (size_t) depth * st.chunk + local;   // record slot / vertex index
    const float4 a0 = st.v_p_u[r], a1 = st.v_n_v[r], a2 = st.v_s_wix[r], a3 = st.v_shn_wiy[r], w4 = st.v_w[r];
    const float2 ws = st.v_wiz_shape[r];
    SI si;
    si.t = 0.f;
    si.p = mk3(a0.x, a0.y, a0.z); si.n = mk3(a1.x, a1.y, a1.z); si.uv = mk2(a0.w, a1.w);
    si.sh.s = mk3(a2.x, a2.y, a2.z); si.sh.n = mk3(a3.x, a3.y, a3.z); si.sh.t = cross(si.sh.n, si.sh.s);
    si.wi = mk3(a2.w, a3.w, ws.x);
    si.shape = (int32_t) f2u(ws.y);
    const S4 wl = f4_to_s4(st.wl[local]);
    DirectionSample ds;
    ds.p = mk3(w4.x, w4.y, w4.z);
    ds.delta = signbit_(w4.w);
    ds.pdf = fabsf(w4.w);            // 0 when the shadow ray was occluded
    ds.d = normalize(ds.p - si.p);
    const S4 em_weight = ds.pdf != 0.0f ? f4_to_s4(st.v_emw[r]) : s4(0.f);
    const V3 wo = si.sh.to_local(ds.d);
    const pltb_bsdf *b = si_bsdf(sc, si.shape);
    V3 wi_f = si.wi, wo_f = wo;
    if (b) twosided_flip(b, wi_f, wo_f);
    const S4 value = b ? bsdf_eval_nofresnel<Rgb>(sc, *b, class_type<Class>(b), wi_f, si.uv, wl, wo_f) : s4(0.f);
    const float bsdf_pdf = wbsdf_pdf<Class>(sc, b, si.wi, si.uv, wl, wo);
    const float mis_em = ds.delta ? 1.0f : mis_weight(ds.pdf, bsdf_pdf);
    PolEval pe;
    pol_eval_prepare(sc, b, si.wi, si.uv, wl, wo, pe);
    C4 ai[4];
    pol_load_col(st.pv_alpha, st.nv, r, ai);
    S4 out;
    for (int k = 0; k < 4; ++k) {
        float row[4];
        pol_eval_row0(pe, value.v[k], k, row);
        const float e = em_weight.v[k] * mis_em;
        float v = (e * row[0]) * ai[k].v[0];
        for (int j = 1; j < 4; ++j) v = fmadd(e * row[j], ai[k].v[j], v);
        out.v[k] = v;
    }
    st.v_nee[rv] = s4_to_f4(out);
}

}  // namespace pltb
