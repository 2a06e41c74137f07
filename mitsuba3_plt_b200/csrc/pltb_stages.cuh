// pltb_stages.cuh -- bodies of the wavefront stages (see pltb_integrator.cuh), templated over the list-append policy and the
// material class; instantiated by the CUDA kernels of pltb_kernels.cu.
#pragma once
#include "pltb_integrator.cuh"
#include "pltb_envmap.cuh"

namespace pltb {

PLTB_HD float4 mkf4(float x, float y, float z, float w) { float4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
PLTB_HD float4 s4_to_f4(const S4 &s) { return mkf4(s.v[0], s.v[1], s.v[2], s.v[3]); }
PLTB_HD S4 f4_to_s4(const float4 &f) { S4 s; s.v[0] = f.x; s.v[1] = f.y; s.v[2] = f.z; s.v[3] = f.w; return s; }

// Work order of a render: sample-major, w = s_rel * n_pix + pixel (pixel fastest).  Adjacent threads work on adjacent pixels
// of the same sample index, so primary rays stay coherent and the 5x5 film splats of a warp go to 32 neighbouring pixels
// (coalesced vector atomics to distinct addresses) instead of 32 times to the same 25 pixels.
// work index -> window pixel index / sample-in-window
PLTB_HD void split_work(const PassParams &pp, uint64_t w, uint32_t &pix, uint32_t &s_rel) {
    s_rel = (uint32_t) (w / pp.n_pix);
    pix = (uint32_t) (w - (uint64_t) s_rel * pp.n_pix);
}
// -> pixel / sample-in-pixel / reference lane index (common.py:256-266)
PLTB_HD void decode_sample(const PassParams &pp, uint64_t w, uint32_t &x, uint32_t &y, uint32_t &s, uint32_t &idx) {
    uint32_t pix, s_rel;
    split_work(pp, w, pix, s_rel);
    s = s_rel + pp.spp_begin;
    uint32_t row = pix / pp.crop_w;
    x = pp.crop_x + (pix - row * pp.crop_w);
    y = pp.crop_y + row;
    idx = (y * pp.film_w + x) * pp.spp + s;
}

PLTB_HD float stokes_rotation_angle(const pltb_sensor &sn, V3 ray_d, bool batch);   // pltb_stages_pol.cuh
struct PassState;
PLTB_NI void pol_miss_record(const DeviceScene &sc, const PassState &st, uint32_t local, uint32_t i);   // pltb_stages_pol.cuh

// Closest-hit bookkeeping shared by k_raygen (bounce 0) and k_trace (bounce >= 1): a hit is routed to the shade queue of
// its material class; a miss finalises the path in place (plt.py:93-168 with si invalid: the five BSDF draws and the RR
// draw are still consumed, the bounce record contributes nothing, the loop ends).
// solve_replay_emissive (plt.py:315-405) for a ray that left the scene at bounce i when the scene has an environment emitter:
// si.emitter(scene) of the invalid interaction is the environment (interaction.h:343-346), DirectionSample3f(scene, si, prev_si)
// takes d = -si.wi = ray.d (records.h:173-180).  Constant background: pdf_direction = 1 / (4 pi) (constant.cpp:152-157),
// eval = radiance (:92-96).  The replayed throughput is the prefix product of the recorded weights, as in stage_shade.
// the MIS-weighted environment radiance mis_bsdf * Le seen by the escaped ray (before the replayed throughput)
PLTB_NI S4 env_miss_lem(const DeviceScene &sc, const PassState &st, uint32_t local, uint32_t i) {
    const pltb_emitter &e = sc.emitters[sc.env];
    const S4 wl = f4_to_s4(st.wl[local]);
    float last_nd_pdf = 1.f;
    bool prev_bsdf_delta = true;
    if (i > 0) {
        last_nd_pdf = st.p_prev[local].w;
        prev_bsdf_delta = (st.p_flags[local] & 1u) != 0;
    }
    float em_pdf = 0.f;
    S4 Le;
    if (e.type == PLTB_EMITTER_ENVMAP) {   // EnvironmentMapEmitter::pdf_direction / ::eval (envmap.cpp:468-484, 343-353)
        const float4 d4 = st.p_ray_d[local];
        const V3 d = mk3(d4.x, d4.y, d4.z);   // -si.wi of the invalid interaction
        if (!prev_bsdf_delta) em_pdf = envmap_pdf_direction(sc, e, d) * sc.emitter_pmf;
        Le = envmap_eval_spectrum(sc, e, envmap_dir_to_uv(xf_vector(e.to_local, d)), wl);
    } else {
        if (!prev_bsdf_delta) em_pdf = kInvFourPi * sc.emitter_pmf;
        Le = tex_eval(sc, e.radiance, wl, mk2(0.f, 0.f));
    }
    const float mis_bsdf = mis_weight(last_nd_pdf, em_pdf);
    return Le * mis_bsdf;
}
PLTB_NI S4 env_miss_emission(const DeviceScene &sc, const PassState &st, uint32_t local, uint32_t i) {
    S4 ai = s4(1.f);
    for (int j = (int) i - 1; j >= 0; --j) ai = f4_to_s4(st.v_w[w_index(st, local, (uint32_t) j)]) * ai;
    return ai * env_miss_lem(sc, st, local, i);
}

template <class Push>
PLTB_HD void route_hit(const DeviceScene &sc, const PassState &st, uint32_t local, uint32_t i, const Hit &h, PCG32 &rng,
                       bool rng_dirty, Push &push) {
    if (h.shape >= 0) {
        if (rng_dirty) {   // bounce 0 (k_raygen); k_trace has already written the record
            st.p_hit[local] = mkf4(h.t, h.u, h.v, u2f(h.prim | (h.cls << kHitClsShift)));
            st.rng_state[local] = rng.state; st.rng_inc[local] = rng.inc;
        }
        PLTB_ASSERT(h.cls < (uint32_t) kClasses && st.shade_q[h.cls] != nullptr && local < st.chunk && i < st.max_depth);
        push(st.shade_q[h.cls], st.counts + cnt_shade(i, h.cls), local);
    } else {
        for (int k = 0; k < 6; ++k) rng.state = rng.state * 0x5851f42d4c957f2dULL + rng.inc;
        st.rng_state[local] = rng.state; st.rng_inc[local] = rng.inc;
        const size_t r = (size_t) i * st.chunk + local;
        uint64_t flags = i > 0 ? st.p_flags[local] : 0ull;
        if (sc.env >= 0) {   // the escaped ray sees the environment: an emissive term at this vertex
            if (st.pv_em) pol_miss_record(sc, st, local, i);
            else st.v_em[r] = s4_to_f4(env_miss_emission(sc, st, local, i));
            flags |= 1ull << (kFlagEm + i);
        }
        st.v_mask[local] = pack_mask(flags);
        st.n_vertices[local] = i + 1;
    }
}

// ADIntegrator.sample_rays (common.py:231-309) + the first iteration's scene.ray_intersect (plt.py:96)
template <class Push>
PLTB_HD int stage_raygen(const DeviceScene &sc, const PassState &st, const PassParams &pp, uint32_t local, Push &push,
                         uint32_t cnt[4]) {   // returns the material class of what the primary ray hit, -1 for a miss
    uint32_t x, y, s, idx;
    decode_sample(pp, pp.first + local, x, y, s, idx);
    // PCG32Sampler::seed (src/render/sampler.cpp:125-145)
    uint32_t v0, v1;
    sample_tea_32(pp.seed, idx, v0, v1);
    PCG32 rng; rng.seed(v0, v1);
    float jx = rng.next_float32(), jy = rng.next_float32();
    V2 pos_f = mk2((float) x + jx, (float) y + jy);
    const float wav_sample = sc.rgb ? 0.f : rng.next_float32();   // only drawn in spectral variants (common.py:297-299)
    const uint32_t fw = sc.film.width;
    uint32_t sidx = x / fw;   // BatchSensor: film split in x (src/sensors/batch.cpp:131-190)
    const pltb_sensor &sn = sc.sensors[sidx];
    V2 pos_adj = mk2((pos_f.x - (float) (sidx * fw)) * rcp((float) fw), pos_f.y * rcp((float) pp.film_h));
    S4 wl = s4(0.f), wgt = s4(1.f);   // RGB variants: no wavelengths, ray weight 1 (sensor.cpp:113-127)
    if (!sc.rgb) sample_wavelengths(sc, sn, wav_sample, wl, wgt);
    Ray ray = sensor_ray(sn, pos_adj);
    if (!sc.rgb) st.wl[local] = s4_to_f4(wl);
    st.sensor_w[local] = s4_to_f4(wgt);
    float2 fp; fp.x = pos_f.x; fp.y = pos_f.y;
    st.film_pos[local] = fp;
    if (st.n_shadow) st.n_shadow[local] = 0;
    if (st.stokes_theta) st.stokes_theta[local] = stokes_rotation_angle(sc.sensors[0], ray.d, sc.n_sensors > 1);
    // sampling_wavelengths (plt.py:64-70): four draws times scale_wl = 680; dead in spectral variants (N1), in RGB variants the
    // first three are what roughgrating's RGB fallback evaluates (kept in the wavelength slot, third one repeated)
    {
        float u[4];
        for (int k = 0; k < 4; ++k) u[k] = rng.next_float32();
        if (sc.rgb) st.wl[local] = mkf4(u[0] * 680.f, u[1] * 680.f, u[2] * 680.f, u[2] * 680.f);
    }
    Hit h = bvh_intersect<false>(sc, ray);
    cnt[0]++;
    if (h.shape >= 0 || sc.env >= 0) {   // a primary ray that leaves a scene without environment is never looked at again
        st.p_ray_o[local] = mkf4(ray.o.x, ray.o.y, ray.o.z, ray.maxt);
        st.p_ray_d[local] = mkf4(ray.d.x, ray.d.y, ray.d.z, 1.f);
    }
    route_hit(sc, st, local, 0, h, rng, true, push);
    return h.shape >= 0 ? (int) h.cls : -1;
}

// One iteration of the sample-phase loop (plt.py:93-168) for a path that hit a surface of material class `Class`, fused with
// the parts of the solve phase that only depend on vertices 0..i (replay_path :408-472, solve_replay_emissive :315-405).
template <int Class, class Push>
PLTB_HD void stage_shade(const DeviceScene &sc, const PassState &st, uint32_t local, uint32_t i, uint32_t slot, uint32_t live_pos, Push &push, uint32_t cnt[4]) {
    PLTB_ASSERT(local < st.chunk && i < st.max_depth);
    const float4 o4 = st.p_ray_o[local], d4 = st.p_ray_d[local], h4 = st.p_hit[local];
    Ray ray; ray.o = mk3(o4.x, o4.y, o4.z); ray.d = mk3(d4.x, d4.y, d4.z); ray.maxt = o4.w;
    Hit h; h.t = h4.x; h.u = h4.y; h.v = h4.z; h.prim = f2u(h4.w) & kHitPrimMask; h.cls = (uint32_t) Class;
    PLTB_ASSERT(f2u(h4.w) != kHitMiss && ((f2u(h4.w) >> kHitClsShift) & 7u) == (uint32_t) Class && h.prim < sc.n_faces + sc.n_shapes);
    h.shape = h.prim < sc.n_faces ? (int32_t) PLTB_LDG(sc.face_shape + h.prim) : (int32_t) (h.prim - sc.n_faces);
    const S4 wl = f4_to_s4(st.wl[local]);
    PCG32 rng; rng.state = st.rng_state[local]; rng.inc = st.rng_inc[local];
    S4 alpha = s4(1.f);
    float eta = 1.f, last_nd_pdf = 1.f;
    bool prev_bsdf_delta = true;
    V3 prev_p = mk3(0.f, 0.f, 0.f);
    uint64_t flags = 0ull;   // PassState::p_flags
    if (i > 0) {
        alpha = f4_to_s4(st.p_alpha[local]);
        float4 pv = st.p_prev[local];
        prev_p = mk3(pv.x, pv.y, pv.z); last_nd_pdf = pv.w;
        eta = d4.w;
        flags = st.p_flags[local];
        prev_bsdf_delta = (flags & 1ull) != 0;
    }
    SI si = make_si(sc, ray, h);
    bool active_next = (i + 1 < st.max_depth);   // si is valid here; depth == i on entry
    float sample_1 = rng.next_float32();
    V2 sample_2, lobe_sample_2;
    sample_2.x = rng.next_float32(); sample_2.y = rng.next_float32();
    lobe_sample_2.x = rng.next_float32(); lobe_sample_2.y = rng.next_float32();
    const pltb_bsdf *b = si_bsdf(sc, si.shape);
    BSDFSampleOut so = wbsdf_sample<Class>(sc, b, si.wi, si.uv, wl, sample_1, sample_2, lobe_sample_2);

    // replay_path for this vertex: alpha_i = w_0 * (w_1 * (... * (w_{i-1} * 1)))
    S4 ai = s4(1.f);
    for (int j = (int) i - 1; j >= 0; --j) ai = f4_to_s4(st.v_w[w_index(st, local, (uint32_t) j)]) * ai;
    // solve_replay_emissive
    S4 em = s4(0.f);
    const int32_t em_id = sc.shapes[si.shape].emitter;
    if (em_id >= 0) {
        const pltb_emitter &e = sc.emitters[em_id];
        float em_pdf = 0.f;
        if (!prev_bsdf_delta) {   // Scene::pdf_emitter_direction -> AreaLight::pdf_direction (area.cpp:170-196)
            V3 rel = si.p - prev_p;
            float dist = norm(rel);
            V3 d = rel / dist;
            if (dot(d, si.sh.n) < 0.f) em_pdf = shape_pdf_direction(sc.shapes[e.shape], prev_p, d, si.sh.n, dist);
            em_pdf *= sc.emitter_pmf;
        }
        float mis_bsdf = mis_weight(last_nd_pdf, em_pdf);
        S4 Le = si.wi.z > 0.f ? tex_eval(sc, e.radiance, wl, si.uv) : s4(0.f);   // AreaLight::eval (area.cpp:82-89)
        em = ai * (Le * mis_bsdf);
    }
    const size_t r = (size_t) i * st.chunk + local;
    if (active_next) st.v_w[w_index(st, local, i)] = s4_to_f4(wbsdf_weight<Class>(sc, b, si.wi, si.uv, wl, so.wo));
    cnt[2]++;
    if (Class == PLTB_BSDF_ROUGHGRATING) cnt[3]++;

    Ray next = si.spawn_ray(si.sh.to_world(so.wo));
    if (active_next) { alpha = so.weight * alpha; eta = so.eta; }
    float alpha_max = smax(alpha);
    active_next = active_next && alpha_max != 0.f;
    float rr_prob = fminf(alpha_max * sqr(eta), 0.95f);
    bool rr_active = (i + 1) >= st.rr_depth;
    if (rr_active) alpha = alpha * rcp(rr_prob);
    bool rr_continue = rng.next_float32() < rr_prob;
    active_next = active_next && (!rr_active || rr_continue);

    // bounce record
    constexpr bool kSmoothClass = (kSmoothClassMask >> Class & 1u) != 0;
    PLTB_ASSERT(kSmoothClass || (so.type & F_Smooth) == 0);
    const bool nee = kSmoothClass && (so.type & F_Smooth) != 0 && sc.n_emitters > 0;
    if (em_id >= 0) { st.v_em[r] = s4_to_f4(em); flags |= 1ull << (kFlagEm + i); }   // other vertices have no emissive term
    if (nee) flags |= 1ull << (kFlagNee + i);   // the NEE steps write v_nee[r]
    st.rng_state[local] = rng.state;
    if (active_next) {
        const bool delta = (so.type & F_Delta) != 0;
        st.p_ray_o[local] = mkf4(next.o.x, next.o.y, next.o.z, next.maxt);
        st.p_ray_d[local] = mkf4(next.d.x, next.d.y, next.d.z, eta);
        st.p_alpha[local] = s4_to_f4(alpha);
        st.p_prev[local] = mkf4(si.p.x, si.p.y, si.p.z, delta ? last_nd_pdf : so.pdf);
        st.p_flags[local] = (flags & ~1ull) | (delta ? 1ull : 0ull);
    } else {
        st.v_mask[local] = pack_mask(flags);
        st.n_vertices[local] = i + 1;
    }
    if (nee) {   // NEE record, dense (PassState)
        st.v_p_u[slot] = mkf4(si.p.x, si.p.y, si.p.z, si.uv.x);
        st.v_n_v[slot] = mkf4(si.n.x, si.n.y, si.n.z, si.uv.y);
        st.v_s_wix[slot] = mkf4(si.sh.s.x, si.sh.s.y, si.sh.s.z, si.wi.x);
        st.v_shn_wiy[slot] = mkf4(si.sh.n.x, si.sh.n.y, si.sh.n.z, si.wi.y);
        st.v_alpha[slot] = s4_to_f4(ai);
        float2 ws; ws.x = si.wi.z; ws.y = u2f((uint32_t) si.shape);
        st.v_wiz_shape[slot] = ws;
        st.rec_entry[slot] = local | (i << 27);
    } else if (kSmoothClass && sc.n_emitters > 0) {
        st.rec_entry[slot] = kNoRecord;
    }
    if (i + 1 < st.max_depth) st.live[live_pos] = active_next ? local : kNoPath;   // the next bounce's ray list (PassState::live)
}

// solve_replay_NEE (plt.py:221-300) for one queued vertex, in three steps so that every step runs in dense warps:
//   stage_nee_light  emitter sampling (Scene::sample_emitter_direction without the visibility test) and the emitter weight of
//                    the unoccluded sample; vertices with a non-zero pdf are appended to the shadow-ray list
//   k_trace<true>    any-hit traversal of the shadow rays by persistent warps; an occluded sample gets ds.pdf = 0, which is
//                    what scene.cpp:330-340 does
//   stage_nee_post   wbsdf_eval / wbsdf_pdf, MIS weight, times the replayed throughput
// Between the steps the record slot carries {ds.p, +-ds.pdf (sign bit: ds.delta)} in v_w (free once the sample phase is
// over) and the emitter weight in v_emw.
template <class Push>
PLTB_HD void stage_nee_light(const DeviceScene &sc, const PassState &st, uint32_t slot, Push &push, uint32_t cnt[4]) {
    const uint32_t entry = st.rec_entry[slot], local = entry & 0x7FFFFFFu, depth = entry >> 27;
    PLTB_ASSERT(local < st.chunk && depth < st.max_depth && slot < st.nv);
    const size_t r = slot;   // NEE records and the scratch of the NEE steps are indexed by record slot
    const float4 a0 = st.v_p_u[r];
    const V3 p = mk3(a0.x, a0.y, a0.z);
    const S4 wl = f4_to_s4(st.wl[local]);
    PCG32 rng; rng.state = st.rng_state[local]; rng.inc = st.rng_inc[local];
    for (uint32_t k = 0; k < 2 * depth; ++k) rng.state = rng.state * 0x5851f42d4c957f2dULL + rng.inc;
    V2 sample; sample.x = rng.next_float32(); sample.y = rng.next_float32();

    // Scene::sample_emitter_direction (src/render/scene.cpp:294-348) + sample_emitter (:222-245)
    uint32_t index = 0;
    if (sc.n_emitters > 1) {
        float cf = (float) sc.n_emitters, scaled = sample.x * cf;
        index = (uint32_t) scaled;
        if (index > sc.n_emitters - 1u) index = sc.n_emitters - 1u;
        sample.x = scaled - (float) index;
    }
    const pltb_emitter &em = sc.emitters[index];
    DirectionSample ds;
    if (em.type == PLTB_EMITTER_AREA) {
        ds = shape_sample_direction(sc, sc.shapes[em.shape], p, sample);   // AreaLight::sample_direction (area.cpp:116-168)
    } else if (em.type == PLTB_EMITTER_CONSTANT) {   // ConstantBackgroundEmitter::sample_direction (constant.cpp:121-150)
        const float z = fnmadd(2.f, sample.y, 1.f), rr = safe_sqrt(fnmadd(z, z, 1.f)), a = 2.f * kPi * sample.x;
        const V3 dir = mk3(rr * cosf(a), rr * sinf(a), z);   // warp::square_to_uniform_sphere (warp.h:250-255)
        const V3 c = mk3(em.bsphere_center[0], em.bsphere_center[1], em.bsphere_center[2]);
        const float dist = 2.f * fmaxf(em.bsphere_radius, norm(p - c));
        ds.p = fmadd3(dir, dist, p);
        ds.n = -dir; ds.uv = sample; ds.pdf = kInvFourPi; ds.delta = false; ds.d = dir; ds.dist = dist;
    } else if (em.type == PLTB_EMITTER_ENVMAP) {   // EnvironmentMapEmitter::sample_direction (envmap.cpp:421-466)
        V3 dw;
        envmap_sample_direction(sc, em, p, sample, dw, ds.dist, ds.uv, ds.pdf);
        ds.p = p + dw * ds.dist;
        ds.n = -dw; ds.delta = false; ds.d = dw;
    } else if (em.type == PLTB_EMITTER_DIRECTIONALSPOT) {   // DirectionalSpreadEmitter::sample_direction (directionalspot.cpp:161-192)
        const float rr = sqrtf(sample.y), ang = kTwoPi * sample.x;   // warp::square_to_uniform_disk (warp.h:25-30) * sin(spread)
        const float sn = sinf(ang), cs = cosf(ang);
        const V3 dd = xf_vector(em.to_world, normalize(mk3((cs * rr) * em.env_pad[0], (sn * rr) * em.env_pad[0], 1.f)));
        const V3 c = mk3(em.bsphere_center[0], em.bsphere_center[1], em.bsphere_center[2]);
        const float dist = 2.f * fmaxf(em.bsphere_radius, norm(p - c));
        ds.p = p - dd * dist;
        ds.n = dd; ds.uv = mk2(0.f, 0.f); ds.pdf = 1.f; ds.delta = true; ds.d = -dd; ds.dist = dist;
    } else {   // DirectionalEmitter::sample_direction (src/emitters/directional.cpp:149-176)
        V3 dirn = mk3(em.direction[0], em.direction[1], em.direction[2]);
        ds.d = -dirn;
        float dist = 2.f * em.bsphere_radius;
        ds.p = fmadd3(ds.d, dist, p);
        ds.n = dirn; ds.pdf = 1.f; ds.delta = true; ds.dist = dist; ds.uv = mk2(0.f, 0.f);
    }
    ds.pdf *= sc.emitter_pmf;
    ds.d = normalize(ds.p - p);   // plt.py:264-269
    S4 em_val = s4(0.f);
    if (em.type == PLTB_EMITTER_AREA) { if (dot(ds.d, ds.n) < 0.f) em_val = tex_eval(sc, em.radiance, wl, ds.uv); }
    else if (em.type == PLTB_EMITTER_ENVMAP) em_val = envmap_eval_spectrum(sc, em, ds.uv, wl);   // ::eval_direction (:486-491)
    else em_val = tex_eval(sc, em.radiance, wl, ds.uv);
    const S4 em_weight = ds.pdf != 0.0f ? em_val / ds.pdf : s4(0.f);
    st.v_emw[r] = s4_to_f4(em_weight);
    st.v_w[r] = mkf4(ds.p.x, ds.p.y, ds.p.z, ds.delta ? -ds.pdf : ds.pdf);
    if (ds.pdf != 0.f) {
        cnt[1]++;
#if defined(__CUDA_ARCH__)
        if (st.n_shadow) atomicAdd(st.n_shadow + local, 1u);
#endif
        push(st.shadow_q, st.counts + kCntShadow, slot);
    }
}

// The shadow ray of a queued vertex (SurfaceInteraction::spawn_ray_to, interaction.h:145-152), rebuilt from the record.
PLTB_HD Ray nee_shadow_ray(const PassState &st, size_t r) {
    const float4 a0 = st.v_p_u[r], a1 = st.v_n_v[r], w4 = st.v_w[r];
    SI si;
    si.p = mk3(a0.x, a0.y, a0.z); si.n = mk3(a1.x, a1.y, a1.z);
    return si.spawn_ray_to(mk3(w4.x, w4.y, w4.z));
}

// does the light direction fall inside any lobe window of the grating?  (the divergent part of stage_nee_post<roughgrating>;
// k_nee_post substitutes a functor that hands the lit items of a block to dense threads -- same function, same inputs, other thread)
struct GratingInsideDirect {
    PLTB_HD bool operator()(bool want, const pltb_bsdf *b, V3 wi_f, V3 wo_f, V2 uv, const S4 &wl) const {
        if (!want) return false;
        GratingLoop gl; gl.init(*b, wi_f, uv, wl, wo_f);
        GratingHits h; grating_scan(gl, wo_f, h);
        return h.n > 0;
    }
};

// Mode (roughgrating, spectral variants): the queue is processed by two kernels running this one body.
//   kNeeFilter  k_nee_post<grating>: loads the record, finishes the unlit well-formed items (contribution exactly 0) and the
//               malformed ones (non-finite or non-unit inputs, the reference's quirks of DESIGN.md 5: evaluated as written), and
//               lists the lit ones in the scan list (shadow_q, dead after the shadow trace)
//   kNeeScan    k_nee_scan: the lobe-window scan of the lit items, one per lane, all lanes busy (inside one kernel, one item per
//               thread ran the scan at 10 of 32 lanes; a block-level hand-over to dense threads left 29 % of the warp samples at its
//               barrier: 53.6 / 47.2 / 45.4 ms per 2^30 samples on cbox-grating.  Moving the malformed items to a third kernel made
//               the filter 40 registers wide but no faster, and the handful of exhaustive evaluations cost 0.2 ms per pass alone.)
//   kNeeFull    everything in one call: the other material classes and the RGB variants
enum { kNeeFull = 0, kNeeFilter = 1, kNeeScan = 2 };
template <int Class, bool Rgb, int Mode, class Push, class Inside>
PLTB_HD void stage_nee_post(const DeviceScene &sc, const PassState &st, uint32_t slot, bool valid, Push &push, Inside &&inside_fn) {
    const uint32_t entry = valid ? st.rec_entry[slot] : 0u, local = entry & 0x7FFFFFFu, depth = entry >> 27;
    const size_t r = slot, rv = (size_t) depth * st.chunk + local;   // record slot / vertex index (where the result goes)
    SI si;
    si.t = 0.f; si.uv = mk2(0.f, 0.f); si.wi = mk3(0.f, 0.f, 1.f);
    S4 wl = s4(0.f), em_weight = s4(0.f), a = s4(0.f);
    DirectionSample ds;
    ds.delta = false; ds.pdf = 0.f;
    V3 wo = mk3(0.f, 0.f, 1.f);
    const pltb_bsdf *b = nullptr;
    if (valid) {
        const float4 a0 = st.v_p_u[r], a1 = st.v_n_v[r], a2 = st.v_s_wix[r], a3 = st.v_shn_wiy[r], w4 = st.v_w[r];
        const float2 ws = st.v_wiz_shape[r];
        si.p = mk3(a0.x, a0.y, a0.z); si.n = mk3(a1.x, a1.y, a1.z); si.uv = mk2(a0.w, a1.w);
        si.sh.s = mk3(a2.x, a2.y, a2.z); si.sh.n = mk3(a3.x, a3.y, a3.z); si.sh.t = cross(si.sh.n, si.sh.s);
        si.wi = mk3(a2.w, a3.w, ws.x);
        si.shape = (int32_t) f2u(ws.y);
        wl = f4_to_s4(st.wl[local]);
        ds.p = mk3(w4.x, w4.y, w4.z);
        ds.delta = signbit_(w4.w);
        ds.pdf = fabsf(w4.w);            // 0 when the shadow ray was occluded
        ds.d = normalize(ds.p - si.p);
        em_weight = ds.pdf != 0.0f ? f4_to_s4(st.v_emw[r]) : s4(0.f);
        wo = si.sh.to_local(ds.d);
        b = si_bsdf(sc, si.shape);
        a = f4_to_s4(st.v_alpha[r]);
    }
    if (Class == PLTB_BSDF_ROUGHGRATING && !Rgb) {
        // RoughGrating::wbsdf_eval is only non-zero inside the few-degree windows around the diffraction lobes: 9 % of the NEE
        // directions on cbox-grating, 6 % once unlit samples are excluded.  For well-formed inputs (finite, unit wi / wo,
        // finite throughput: then every factor of the product is finite) an unlit sample or one outside every window
        // contributes exactly 0, as in the reference; the rest is compacted into `queue_eval` so that the Bessel terms are
        // evaluated by dense warps (k_nee_eval).  Anything else takes the generic path below.
        V3 wi_f = si.wi, wo_f = wo;
        if (valid) twosided_flip(b, wi_f, wo_f);
        const float fsum = wi_f.x + wi_f.y + wi_f.z + wo_f.x + wo_f.y + wo_f.z + a.v[0] + a.v[1] + a.v[2] + a.v[3] +
                           em_weight.v[0] + em_weight.v[1] + em_weight.v[2] + em_weight.v[3];
        const bool sane = finite_(fsum) && wi_f.z != 0.f && fabsf(sqnorm(wi_f) - 1.f) < 1e-3f && fabsf(sqnorm(wo_f) - 1.f) < 1e-3f;
        const bool lit = em_weight.v[0] != 0.f || em_weight.v[1] != 0.f || em_weight.v[2] != 0.f || em_weight.v[3] != 0.f;
        // (RGB variants, Rgb = true: the fallback's lobe search differs -- other wavelengths, a coherence factor on every lobe --
        // and takes the generic path below)
        if (Mode == kNeeFilter && valid && sane && lit) { push(st.shadow_q, st.counts + kCntScan, slot); return; }
        const bool inside = Mode != kNeeFilter && inside_fn(valid && sane && lit, b, wi_f, wo_f, si.uv, wl);
        if (!valid) return;
        if (sane || Mode == kNeeScan) {
            if (!inside) { st.v_nee[rv] = mkf4(0.f, 0.f, 0.f, 0.f); return; }
            float bsdf_pdf = wbsdf_pdf<Class>(sc, b, si.wi, si.uv, wl, wo);
            float mis_em = ds.delta ? 1.0f : mis_weight(ds.pdf, bsdf_pdf);
            st.v_emw[r] = s4_to_f4(em_weight * mis_em);    // finished by k_nee_eval
            st.v_w[r] = mkf4(wo.x, wo.y, wo.z, 0.f);
            push(st.queue_eval, st.counts + kCntEval, slot);
            return;
        }
    }
    if (!valid) return;
    S4 bsdf_val = wbsdf_eval<Class, Rgb>(sc, b, si.wi, si.uv, wl, wo);
    float bsdf_pdf = wbsdf_pdf<Class>(sc, b, si.wi, si.uv, wl, wo);
    float mis_em = ds.delta ? 1.0f : mis_weight(ds.pdf, bsdf_pdf);
    st.v_nee[rv] = s4_to_f4(((em_weight * mis_em) * bsdf_val) * a);
}

// second half of solve_replay_NEE for the roughgrating vertices whose light direction falls inside a lobe window
PLTB_HD void stage_nee_eval(const DeviceScene &sc, const PassState &st, uint32_t slot) {
    const uint32_t entry = st.rec_entry[slot], local = entry & 0x7FFFFFFu, depth = entry >> 27;
    const size_t r = slot, rv = (size_t) depth * st.chunk + local;
    const float4 a0 = st.v_p_u[r], a1 = st.v_n_v[r], a2 = st.v_s_wix[r], a3 = st.v_shn_wiy[r], w4 = st.v_w[r];
    const float2 ws = st.v_wiz_shape[r];
    const V3 wi = mk3(a2.w, a3.w, ws.x), wo = mk3(w4.x, w4.y, w4.z);
    const V2 uv = mk2(a0.w, a1.w);
    const S4 wl = f4_to_s4(st.wl[local]);
    const pltb_bsdf *b = si_bsdf(sc, (int32_t) f2u(ws.y));
    S4 bsdf_val = wbsdf_eval<PLTB_BSDF_ROUGHGRATING>(sc, b, wi, uv, wl, wo);
    st.v_nee[rv] = s4_to_f4((f4_to_s4(st.v_emw[r]) * bsdf_val) * f4_to_s4(st.v_alpha[r]));
}

// plt_solve_phase accumulation order (plt.py:191-216): emissive and NEE terms of every recorded vertex, in depth order
PLTB_HD S4 splat_gather(const PassState &st, uint32_t local, uint32_t n) {
    S4 L = s4(0.f);
    const uint32_t m = st.v_mask[local];   // terms that exist; the others are exact zeros
    for (uint32_t i = 0; i < n; ++i) {
        const size_t r = (size_t) i * st.chunk + local;
        if (m >> i & 1u) L = L + f4_to_s4(st.v_em[r]);
        if (m >> (16 + i) & 1u) L = L + f4_to_s4(st.v_nee[r]);
    }
    return L;
}
PLTB_HD S4 splat_gather(const PassState &st, uint32_t local) { return splat_gather(st, local, st.n_vertices[local]); }
// _splat_to_block (common.py:370-397): sensor weight, spectrum -> sRGB, channels [R, G, B, W = 1]
PLTB_HD void splat_values(const DeviceScene &sc, const S4 &L, const S4 &wl, const S4 &sensor_w, float values[4]) {
    if (sc.rgb) { const S4 v = L * sensor_w; values[0] = v.v[0]; values[1] = v.v[1]; values[2] = v.v[2]; }   // rgb = value (common.py:386-391)
    else spectrum_to_srgb(sc.cie_srgb, L * sensor_w, wl, values);
    values[3] = 1.f;
}
PLTB_HD void splat_values(const DeviceScene &sc, const PassState &st, uint32_t local, const S4 &L, float values[4]) {
    splat_values(sc, L, f4_to_s4(st.wl[local]), f4_to_s4(st.sensor_w[local]), values);
}
// ImageBlock::put (src/render/imageblock.cpp:429-500), one atomic per tap straight into the film
template <class Film>
PLTB_HD void splat_taps(const DeviceScene &sc, const PassParams &pp, const float values[4], float2 pos, Film &film) {
    const int W = (int) pp.film_w, H = (int) pp.film_h;
    if (sc.film.rfilter == PLTB_RFILTER_BOX) {
        int px = (int) floorf(pos.x), py = (int) floorf(pos.y);
        if (px >= 0 && py >= 0 && px < W && py < H) film.add((size_t) py * W + px, values, 1.f);
        return;
    }
    const int nr = (int) ceilf(sc.film.radius - .5f), count = 2 * nr + 1;
    const int px = (int) floorf(pos.x) - nr, py = (int) floorf(pos.y) - nr;
    float relx = ((float) px + .5f) - pos.x, rely = ((float) py + .5f) - pos.y;
    float wx[9], wy[9];
    for (int k = 0; k < count && k < 9; ++k) {
        wx[k] = rfilter_eval(sc.film, relx); wy[k] = rfilter_eval(sc.film, rely);
        relx += 1; rely += 1;
    }
    for (int ys = 0; ys < count && ys < 9; ++ys) {
        int yy = py + ys;
        if (yy < 0 || yy >= H) continue;
        for (int xs = 0; xs < count && xs < 9; ++xs) {
            int xx = px + xs;
            if (xx < 0 || xx >= W) continue;
            film.add((size_t) yy * W + xx, values, wy[ys] * wx[xs]);
        }
    }
}
// stage 3 for one sample; returns L
template <class Film>
PLTB_HD S4 stage_splat(const DeviceScene &sc, const PassState &st, const PassParams &pp, uint32_t local, Film &film, bool do_splat) {
    S4 L = splat_gather(st, local);
    if (!do_splat) return L;
    float values[4];
    splat_values(sc, st, local, L, values);
    splat_taps(sc, pp, values, st.film_pos[local], film);
    return L;
}

}  // namespace pltb
