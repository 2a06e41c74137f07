// pltb_integrator.cuh -- emitter / sensor / film helpers and the wavefront state of one render pass (DESIGN.md "Kernels"):
//   k_raygen          one thread per Monte-Carlo sample: PCG32/TEA seeding, pixel jitter, wavelength sampling, sensor ray,
//                     then the bounce-0 closest-hit traversal; hits are appended to the shade queue of their material class.
//   k_shade<class>(i) one thread per hit of bounce i (dense, one material class per launch): surface interaction,
//                     wbsdf_sample, and -- because every replay weight depends only on the vertex itself -- the
//                     replay weight w_i, the replayed throughput alpha_i and the MIS-weighted emissive term; RR; bounce
//                     record; NEE work item; continuing paths are appended to the live list.
//   k_trace(i)        one thread per live path: BVH closest hit for bounce i >= 1; misses are finalised in place.
//   k_nee<class>      one thread per queued vertex.  solve_replay_NEE: emitter sampling, shadow ray, wbsdf_eval, wbsdf_pdf.
//   k_splat           one thread per sample.  Sums the contributions in the reference's order, spectrum -> sRGB, film splat.
// Reference: scripts/rendering/integrators/plt.py:50-529, src/python/python/ad/integrators/common.py:231-397.
#pragma once
#include "pltb_bsdf.cuh"

namespace pltb {

// ---- emitters ---------------------------------------------------------------------------------------------------------------
struct DirectionSample { V3 p, n, d; V2 uv; float pdf, dist; bool delta; };

// Mesh::sample_position (src/render/mesh.cpp:751-805), DiscreteDistribution::sample_reuse (distr_1d.h:116-182)
PLTB_HD void mesh_sample_position(const DeviceScene &sc, const pltb_shape &sh, V2 sample, V3 &p, V3 &n, V2 &uv, float &pdf) {
    const float *cdf = sc.area_cdf + sh.cdf_offset;
    float sum = sh.area_sum, value = sample.y * sum;
    uint32_t lo = 0, hi = sh.n_faces - 1;
    while (lo < hi) {
        uint32_t mid = (lo + hi) / 2;
        float c = PLTB_LDG(cdf + mid);
        bool pred = ((c < value) || c == 0.f) && c != sum;
        if (pred) lo = mid + 1; else hi = mid;
    }
    uint32_t face = lo;
    float c_hi = PLTB_LDG(cdf + face), c_lo = face > 0 ? PLTB_LDG(cdf + face - 1) : 0.f;
    float pmf = (c_hi - c_lo) * sh.area_norm;
    float cdf_prev = face > 0 ? c_lo * sh.area_norm : 0.f;
    sample.y = (sample.y - cdf_prev) / pmf;
    uint32_t fi = sh.first_face + face;
    uint32_t f0 = PLTB_LDG(sc.faces + 3 * fi), f1 = PLTB_LDG(sc.faces + 3 * fi + 1), f2 = PLTB_LDG(sc.faces + 3 * fi + 2);
    V3 p0 = ld3(sc.positions, f0), p1 = ld3(sc.positions, f1), p2 = ld3(sc.positions, f2);
    V3 e0 = p1 - p0, e1 = p2 - p0;
    V2 b = square_to_uniform_triangle(sample);
    p = fmadd3(e0, b.x, fmadd3(e1, b.y, p0));
    pdf = sh.area_norm;
    if (sh.flags & PLTB_SHAPE_HAS_UV) {
        V2 uv0 = ld2(sc.texcoords, f0), uv1 = ld2(sc.texcoords, f1), uv2 = ld2(sc.texcoords, f2);
        float b0 = 1.f - b.x - b.y;
        uv = mk2(fmadd(uv0.x, b0, fmadd(uv1.x, b.x, uv2.x * b.y)), fmadd(uv0.y, b0, fmadd(uv1.y, b.x, uv2.y * b.y)));
    } else uv = b;
    if (sh.flags & PLTB_SHAPE_HAS_NORMALS) {
        V3 n0 = ld3(sc.normals, f0), n1 = ld3(sc.normals, f1), n2 = ld3(sc.normals, f2);
        n = fmadd3(n0, 1.f - b.x - b.y, fmadd3(n1, b.x, n2 * b.y));
    } else n = cross(e0, e1);
    n = normalize(n);
    if (sh.flags & PLTB_SHAPE_FLIP_NORMALS) n = -n;
}

// Shape::sample_direction (src/render/shape.cpp:356-373), Sphere::sample_direction (src/shapes/sphere.cpp:234-309)
PLTB_HD DirectionSample shape_sample_direction(const DeviceScene &sc, const pltb_shape &sh, V3 ref_p, V2 sample) {
    DirectionSample ds;
    ds.uv = mk2(0, 0); ds.delta = false; ds.pdf = 0.f;
    if (sh.type == PLTB_SHAPE_SPHERE) {
        V3 center = mk3(sh.center[0], sh.center[1], sh.center[2]);
        V3 dc_v = center - ref_p;
        float dc_2 = sqnorm(dc_v), radius = sh.radius;
        float radius_adj = radius * ((sh.flags & PLTB_SHAPE_FLIP_NORMALS) ? (1.f + kRayEps) : (1.f - kRayEps));
        if (dc_2 > sqr(radius_adj)) {
            float inv_dc = rsqrt_(dc_2), sin_theta_max = radius * inv_dc, sin_theta_max_2 = sqr(sin_theta_max),
                  inv_sin_theta_max = rcp(sin_theta_max), cos_theta_max = safe_sqrt(1.f - sin_theta_max_2);
            float sin_theta_2 = sin_theta_max_2 > 0.00068523f ? 1.f - sqr(fmadd(cos_theta_max - 1.f, sample.x, 1.f))
                                                               : sin_theta_max_2 * sample.x,
                  cos_theta = safe_sqrt(1.f - sin_theta_2);
            float cos_alpha = sin_theta_2 * inv_sin_theta_max +
                              cos_theta * safe_sqrt(fnmadd(sin_theta_2, sqr(inv_sin_theta_max), 1.f)),
                  sin_alpha = safe_sqrt(fnmadd(cos_alpha, cos_alpha, 1.f));
            float ang = sample.y * (2.f * kPi);
            float sin_phi = sinf(ang), cos_phi = cosf(ang);
            Frame fr = frame_from_normal(dc_v * -inv_dc);
            V3 dd = fr.to_world(mk3(cos_phi * sin_alpha, sin_phi * sin_alpha, cos_alpha));
            ds.p = fmadd3(dd, radius, center);
            ds.n = dd;
            ds.d = ds.p - ref_p;
            float dist2 = sqnorm(ds.d);
            ds.dist = sqrtf(dist2);
            ds.d = ds.d / ds.dist;
            ds.pdf = (0.5f * kInvPi) / (1.f - cos_theta_max);
            if (ds.dist == 0.f) ds.pdf = 0.f;
        } else {
            float z = fnmadd(2.f, sample.y, 1.f), r = safe_sqrt(fnmadd(z, z, 1.f));
            float ang = 2.f * kPi * sample.x;
            V3 dd = mk3(r * cosf(ang), r * sinf(ang), z);
            ds.p = fmadd3(dd, radius, center);
            ds.n = dd;
            ds.d = ds.p - ref_p;
            float dist2 = sqnorm(ds.d);
            ds.dist = sqrtf(dist2);
            ds.d = ds.d / ds.dist;
            ds.pdf = (kInvPi * 0.25f / sqr(radius)) * dist2 / fabsf(dot(ds.d, ds.n));
            if (!finite_(ds.pdf)) ds.pdf = 0.f;
        }
        if (sh.flags & PLTB_SHAPE_FLIP_NORMALS) ds.n = -ds.n;
        return ds;
    }
    float pdf;
    if (sh.type == PLTB_SHAPE_MESH) {
        mesh_sample_position(sc, sh, sample, ds.p, ds.n, ds.uv, pdf);
    } else {  // Rectangle::sample_position, src/shapes/rectangle.cpp:186-204
        ds.p = xf_point(sh.to_world, mk3(fmsub(sample.x, 2.f, 1.f), fmsub(sample.y, 2.f, 1.f), 0.f));
        ds.n = mk3(sh.frame_n[0], sh.frame_n[1], sh.frame_n[2]);
        if (sh.flags & PLTB_SHAPE_FLIP_NORMALS) ds.n = -ds.n;
        ds.uv = sample;
        pdf = sh.inv_area;
    }
    ds.pdf = pdf;
    ds.d = ds.p - ref_p;
    float dist_squared = sqnorm(ds.d);
    ds.dist = sqrtf(dist_squared);
    ds.d = ds.d / ds.dist;
    float dp = fabsf(dot(ds.d, ds.n));
    float x = dist_squared / dp;
    ds.pdf *= finite_(x) ? x : 0.f;
    return ds;
}

// Shape::pdf_direction (shape.cpp:375-386), Sphere::pdf_direction (sphere.cpp:311-330)
PLTB_HD float shape_pdf_direction(const pltb_shape &sh, V3 ref_p, V3 d, V3 n, float dist) {
    if (sh.type == PLTB_SHAPE_SPHERE) {
        V3 center = mk3(sh.center[0], sh.center[1], sh.center[2]);
        float sin_alpha = sh.radius * rcp(norm(center - ref_p)), cos_alpha = safe_sqrt(1.f - sin_alpha * sin_alpha);
        if (sin_alpha < 1.f - kEps) return (0.5f * kInvPi) / (1.f - cos_alpha);
        return (kInvPi * 0.25f / sqr(sh.radius)) * sqr(dist) / fabsf(dot(d, n));
    }
    float pdf = sh.type == PLTB_SHAPE_MESH ? sh.area_norm : sh.inv_area;
    float dp = fabsf(dot(d, n));
    pdf *= dp != 0.f ? (dist * dist) / dp : 0.f;
    return pdf;
}

// ---- sensor -------------------------------------------------------------------------------------------------------------------
// sample_wavelength (include/mitsuba/core/spectrum.h:446-480), math::sample_shifted (math.h:554-577),
// Sensor::sample_wavelengths with an SRF (src/render/sensor.cpp:113-127)
PLTB_HD void sample_wavelengths(const DeviceScene &sc, const pltb_sensor &sn, float sample, S4 &wl, S4 &weight) {
    for (int i = 0; i < 4; ++i) {
        float u = sample + (float) i / 4.f;
        if (u > 1.f) u -= 1.f;
        if (sn.srf >= 0) {
            const pltb_texture &t = sc.textures[sn.srf];
            const float *data = sc.tex_data + t.offset;
            const uint32_t n = t.count;
            const float *nodes = nullptr, *vals, *cdf;
            if (t.type == PLTB_TEX_IRREGULAR) { nodes = data; vals = data + n; cdf = data + 2 * n; }
            else { vals = data; cdf = data + n; }
            float integral = cdf[n - 2], s = u * integral;
            uint32_t a = 0, bnd = n - 2;
            while (a < bnd) { uint32_t m = (a + bnd) / 2; if (cdf[m] < s) a = m + 1; else bnd = m; }
            uint32_t index = a;
            float y0 = vals[index], y1 = vals[index + 1], c0 = index > 0 ? cdf[index - 1] : 0.f, x0, x1;
            if (nodes) { x0 = nodes[index]; x1 = nodes[index + 1]; }
            else { float isz = (t.p[4] - t.p[3]) / (float) (n - 1); x0 = fmadd((float) index, isz, t.p[3]); x1 = x0 + isz; }
            float w = x1 - x0, ss = (s - c0) / w;
            float t_lin = y0 != y1 ? (y0 - safe_sqrt(fmadd(y0, y0, 2.f * ss * (y1 - y0)))) / (y0 - y1) : ss / y0;
            wl.v[i] = fmadd(t_lin, w, x0);
            weight.v[i] = integral;
        } else {
            wl.v[i] = 538.f - atanhf(0.8569106254698279f - 1.8275019724092267f * u) * 138.88888888888889f;
            float tmp = coshf(0.0072f * (wl.v[i] - 538.f));
            weight.v[i] = 253.82f * tmp * tmp;
        }
    }
}

// PerspectiveCamera::sample_ray_differential (src/sensors/perspective.cpp:240-280), OrthographicCamera::sample_ray
PLTB_HD Ray sensor_ray(const pltb_sensor &sn, V2 pos) {
    V3 near_p = xf_point_proj(sn.sample_to_camera, mk3(pos.x + sn.pp_offset[0], pos.y + sn.pp_offset[1], 0.f));
    Ray ray;
    if (sn.type == PLTB_SENSOR_PERSPECTIVE) {
        V3 d = normalize(near_p);
        ray.o = mk3(sn.to_world[3], sn.to_world[7], sn.to_world[11]);
        ray.d = xf_vector(sn.to_world, d);
        float inv_z = rcp(d.z), near_t = sn.near_clip * inv_z, far_t = sn.far_clip * inv_z;
        ray.o = fmadd3(ray.d, near_t, ray.o);
        ray.maxt = far_t - near_t;
    } else {
        ray.o = xf_point(sn.to_world, near_p);
        ray.d = normalize(xf_vector(sn.to_world, mk3(0.f, 0.f, 1.f)));
        ray.maxt = sn.far_clip - sn.near_clip;
    }
    return ray;
}

// linear_rgb_rec + spectrum_to_srgb (include/mitsuba/core/spectrum.h:330-392)
PLTB_HD void spectrum_to_srgb(const float *cie_srgb, const S4 &value, const S4 &wl, float rgb[3]) {
    float acc[3] = { 0.f, 0.f, 0.f };
    for (int i = 0; i < 4; ++i) {
        float t = (wl.v[i] - 360.f) * (94 / (830.f - 360.f));
        bool active = wl.v[i] >= 360.f && wl.v[i] <= 830.f;
        uint32_t i0 = (uint32_t) fmaxf(t, 0.f);
        if (i0 > 93u) i0 = 93u;
        float w1 = t - (float) i0, w0 = 1.f - w1;
        for (int c = 0; c < 3; ++c) {
            float v = active ? fmadd(w0, PLTB_LDG(cie_srgb + 95 * c + i0), w1 * PLTB_LDG(cie_srgb + 95 * c + i0 + 1)) : 0.f;
            float prod = v * value.v[i];
            acc[c] = i == 0 ? prod : acc[c] + prod;
        }
    }
    for (int c = 0; c < 3; ++c) rgb[c] = (acc[c] * 0.25f) * (float) (1.0 / 106.7502593994140625);
}

PLTB_HD float rfilter_eval(const pltb_film &f, float x) {  // gaussian.cpp:94-103 (CUDA form), tent.cpp:24-26
    if (f.rfilter == PLTB_RFILTER_GAUSSIAN) {
        float alpha = -1.f / (2.f * sqr(f.stddev));
        float bias = expf(alpha * sqr(f.radius));
        return fmaxf(0.f, exp2f((1.4426950408889634074f * alpha) * sqr(x)) - bias);
    }
    return fmaxf(0.f, 1.f - fabsf(x * (1.f / f.radius)));
}

// ---- wavefront state ------------------------------------------------------------------------------------------------------------
// All arrays are SoA.  Per-sample and per-path arrays are indexed by the sample's local id in the pass; vertex records are
// depth-major (index = depth * chunk + sample_local) so that accesses of a warp are coalesced float4s.
// Material classes: one shade / NEE queue per BSDF plugin type (indexed by PLTB_BSDF_*; conductor, dielectric and "no BSDF" share
// one, pltb_bvh.h: material_class), so that a warp of k_shade / k_nee_post runs one plugin's code (the class is a template
// parameter).  With two classes (grating / everything else) the "everything else" kernel ran at 14-17 of 32 lanes on cbox and
// veach-mis and missed the instruction cache (profiles/r02_ncu_kernels_c{1,4}_2class.txt).
constexpr int kClasses = 7;

// Per-path streams that the list-driven kernels (k_shade, k_trace, the NEE steps) read together are INTERLEAVED so that a path's
// values share 32-byte sectors: those kernels gather by path id, and a gathered 16-byte (8-byte) element uses half (a quarter) of
// the sector it pulls from DRAM.  `Strided<T, N>` is stream k of N interleaved ones: element i sits at base[N i] (base = buffer + k).
template <class T, int N> struct Strided {
    T *base;
    PLTB_HD T &operator[](size_t i) const { return base[(size_t) N * i]; }
};
struct PassState {
    uint32_t chunk;            // samples per pass (capacity of every per-sample array)
    uint32_t max_depth, rr_depth;
    // per sample
    float4 *wl;                // sampled wavelengths
    float4 *sensor_w;          // sensor spectral weight
    float2 *film_pos;          // splat position
    Strided<uint64_t, 4> rng_state, rng_inc;   // PCG32 state (advanced bounce by bounce; NEE draws continue from the final state);
                                               // one 32-byte record per sample with p_flags: {state, inc, flags, -}
    uint32_t *n_vertices;      // bounce records written (== closest-hit rays traced: one per loop iteration)
    uint32_t *n_shadow;        // shadow rays traced per sample; nullptr unless the caller asked for pltb_sample_dump.n_rays
    // per path, carried from bounce to bounce
    Strided<float4, 2> p_ray_o;   // {ray.o, ray.maxt}   } one 32-byte record per path
    Strided<float4, 2> p_ray_d;   // {ray.d, eta}        }
    float4 *p_hit;             // {t, u, v, prim id bits} of the closest hit found by k_trace
    Strided<float4, 2> p_alpha;   // sample-phase throughput alpha   } one 32-byte record per path
    Strided<float4, 2> p_prev;    // {prev_si.p, last_nd_pdf}        }
    Strided<uint64_t, 4> p_flags;       // bit 0: prev_bsdf_delta; bits 1 + j: (polarized) the replay weight of vertex j is a scaled identity;
                               // bits 17 + j / 33 + j: vertex j has an emissive / an NEE term (kFlagEm / kFlagNee)
    uint32_t *v_mask;          // per sample, written when the path ends: bit j: vertex j has an emissive term (v_em / pv_em written),
                               // bit 16 + j: an NEE term (v_nee written).  Most terms are exact zeros (only emitter hits have an
                               // emissive term, only smooth BSDFs queue NEE): they are neither stored nor summed
    // per vertex (depth-major)
    float4 *v_w;               // replay weight w_i = wbsdf_weight(si_i, wo_i); after the sample phase: [slot] {ds.p, +-ds.pdf} of the NEE steps
    // NEE records are stored DENSELY, indexed by record slot: the vertex k_shade<c> shades at bounce i from position e of
    // shade_q[c] owns slot  slot_base(i, c) + e,  slot_base = the number of vertices shaded by earlier launches of the pass (the sum
    // of the shade-queue counts of the (bounce, class) pairs before (i, c), warp_slot_prefix in pltb_kernels.cu).  No counter is
    // involved: a slot costs k_shade no atomic, and the NEE work list of a class is implicit -- its kernels walk the slot runs of
    // that class, bounce by bounce.  Vertices that queue no NEE item (delta samples) leave rec_entry[slot] = kNoRecord.
    // (Indexed by depth * chunk + sample the records touched half-used sectors: 304 B of DRAM per roughgrating item for 152 B of
    // record; slots handed out by one atomic counter plus one per-class queue counter cost k_shade three same-line atomics per
    // warp: 39 % of its stall samples at bounce 0 of veach-mis, profiles/r02_ab_session17.txt.)
    float4 *v_p_u, *v_n_v, *v_s_wix, *v_shn_wiy;   // [slot] {p, uv.x} {n, uv.y} {sh.s, wi.x} {sh.n, wi.y}
    float4 *v_alpha;           // [slot] replayed throughput alpha_i
    float4 *v_emw;             // [slot] emitter weight between the NEE steps (k_nee_light -> k_nee_post -> k_nee_eval)
    uint32_t *rec_entry;       // [slot] sample_local | depth << 27 of the vertex, or kNoRecord
    float4 *L_out[4];          // where k_gather leaves a sample's sum for the splat (dense, [sample]): row 0 of v_em -- polarized:
                               // of the four pv_em components -- which the gathering thread has consumed by then
    float4 *v_em;              // MIS-weighted emissive contribution alpha_i * Lem_i
    float4 *v_nee;             // NEE contribution (k_nee)
    float2 *v_wiz_shape;       // [slot] {wi.z, shape id bits}
    // polarized variants only (PLTB_RENDER_POLARIZED; nullptr otherwise).  Component-major: [c][...] with c = Stokes component.
    float4 *pp_alpha;          // [4][chunk]   first column of the sample-phase throughput matrix
    float4 *pv_w;              // [16][nv]     replay weight W_i as 4 rows x 4 wavelengths: index (4 * wavelength + row) * nv + r
    float4 *pv_alpha;          // [4][nv]      first column of the replayed throughput alpha_i, by record slot (the NEE step reads it)
    float4 *pv_em;             // [4][nv]      first column of alpha_i * Lem_i
    float *stokes_theta;       // [chunk]      angle of the final Stokes-frame rotation to the sensor's x axis (stokes.cpp:100-106)
    size_t nv;                 // chunk * max_depth slots
    // work lists (path ids)
    uint32_t *live;            // paths whose next ray must be traced (written by k_shade of bounce i, read by k_trace i+1).  Like a
                               // record slot, the position is implicit: the vertex at position e of shade_q[c] writes position
                               // (number of vertices of the classes below c at this bounce) + e -- its path id, or kNoPath if the
                               // path ended; the list of bounce i+1 has as many positions as bounce i shaded vertices
                               // (live_count).  The other lists are filled with warp-aggregated atomics.
    uint8_t *hit_cls;          // [position in live] material class of what the ray hit, 0xFF = miss: k_trace -> k_classify (dense, so
                               // the routing pass does not fetch the 16-byte hit records through the list)
    uint32_t *shade_q[kClasses];   // hits of the current bounce by material class (== BSDF plugin type, PLTB_BSDF_*); nullptr for
                                   // classes the scene does not contain
    uint32_t *queue_eval;      // roughgrating NEE items that reached a lobe window (k_nee_post<1> -> k_nee_eval)
    uint32_t *shadow_q;        // NEE items of both classes whose light sample has a non-zero pdf (k_nee_light -> k_trace<true>); after
                               // the shadow trace the array is reused as the scan list of k_nee_post<grating> (counts[kCntScan])
    // counters, zeroed once per pass (cnt_* below index them)
    uint32_t *counts;
    // statistics
    unsigned long long *counters;  // closest, shadow, vertices, grating vertices, nee_other, nee_grating
};
// layout of PassState::counts: per bounce i a block of kCountsStride counters {unused, trace cursor, shade_q[c] counts}; then (after
// kClasses unused ones) the queue_eval count, the shadow_q count, the shadow-ray cursor and the scan-list count
// Every counter sits in its own 256-byte block (kCntPad words apart; 128 B apart they still collide in pairs: k_classify 21.8 ->
// 19.0 ms per 2^30 samples on cbox-grating): atomics on words of one line queue up in the same L2 slice
// (veach-mis 917 -> 963 Msamples/s when k_shade still appended to three lists per item, profiles/r02_ab_session17.txt).
#ifndef PLTB_CNT_PAD
#define PLTB_CNT_PAD 64
#endif
constexpr int kCntPad = PLTB_CNT_PAD;
constexpr int kCountsStride = 2 + kClasses + 3;   // 12
constexpr int kCountsNee = kCountsStride * (PLTB_MAX_DEPTH + 1);
constexpr int kCountsTotal = (kCountsNee + kClasses + 5) * kCntPad;
PLTB_HD int cnt_cursor(uint32_t i) { return kCntPad * (kCountsStride * (int) i + 1); }
PLTB_HD int cnt_shade(uint32_t i, uint32_t c) { return kCntPad * (kCountsStride * (int) i + 2 + (int) c); }
constexpr int kCntEval = kCntPad * (kCountsNee + kClasses), kCntShadow = kCntPad * (kCountsNee + kClasses + 1),
              kCntShadowCursor = kCntPad * (kCountsNee + kClasses + 2);
constexpr int kFlagEm = 17, kFlagNee = 33;   // first bit of the two presence masks in PassState::p_flags
PLTB_HD uint32_t pack_mask(uint64_t flags) { return (uint32_t) ((flags >> kFlagEm) & 0xFFFFull) | ((uint32_t) ((flags >> kFlagNee) & 0xFFFFull) << 16); }
// index of the replay weight of vertex j of a path in v_w / pv_w during the sample phase: depth-major.  (Path-major -- a path's
// weights consecutive, so that the replay product reads whole sectors -- measured slower: the per-vertex store becomes a strided
// partial-sector write; veach-mis 1127 -> 1106, C2 2234 -> 2211, C2pol 746 -> 665 Msamples/s, profiles/r02_ab_session17.txt)
PLTB_HD size_t w_index(const PassState &st, uint32_t local, uint32_t j) { return (size_t) j * st.chunk + local; }
constexpr uint32_t kNoRecord = 0xFFFFFFFFu;   // rec_entry of a vertex without NEE item
constexpr uint32_t kNoPath = 0xFFFFFFFFu;     // live-list position of a path that ended at the previous bounce
// material classes whose samples can be smooth (diffuse, roughconductor, roughgrating, principled): only they run the NEE kernels
constexpr uint32_t kSmoothClassMask = (1u << PLTB_BSDF_DIFFUSE) | (1u << PLTB_BSDF_ROUGHCONDUCTOR) | (1u << PLTB_BSDF_ROUGHGRATING) |
                                      (1u << PLTB_BSDF_PRINCIPLED);
constexpr int kCntScan = kCntPad * (kCountsNee + kClasses + 3);   // roughgrating NEE items whose lobe windows have to be scanned (k_nee_post<grating> -> k_nee_scan)

struct PassParams {
    uint32_t seed, spp, spp_begin, spp_count;
    uint32_t crop_x, crop_y, crop_w, crop_h;
    uint32_t film_w, film_h;       // whole film (n_sensors * film.width, film.height)
    uint32_t n_pix;                // crop_w * crop_h
    uint64_t first;                // first sample (in window order) of this pass
    uint32_t count;                // samples in this pass
};

}  // namespace pltb
