// pltb_scene.cuh -- device-resident flattened scene: geometry SoA, 4-wide BVH, plugin tables, texture evaluation.
// Layout in HBM (all tables are small and stay L1/L2-resident for the config scenes):
//   bvh_nodes : float4[8] per 4-wide node  lo.x[4] lo.y[4] lo.z[4] hi.x[4] hi.y[4] hi.z[4] child[4] (int bits) pad[4]
//   tri_data  : float4[3] per leaf-ordered primitive {p0.xyz,e1.x}{e1.yz,e2.xy}{e2.z, prim id, shape id, material class}
//   vertex arrays / faces as uploaded from pltb_scene_desc; plugin tables as arrays of the ABI structs.
#pragma once
#include "../../include/pltb.h"
#include "pltb_math.cuh"

#if defined(__CUDA_ARCH__)
#  define PLTB_LDG(p) __ldg(p)
#else
#  define PLTB_LDG(p) (*(p))
#endif
#if !defined(__CUDACC__)
struct float4 { float x, y, z, w; };
struct float2 { float x, y; };
#endif

namespace pltb {

struct DeviceScene {
    const float4 *bvh_nodes;
    const float4 *tri_data;
    uint32_t n_nodes, n_prims;
    const float *positions, *normals, *texcoords;
    const uint32_t *faces, *face_shape;
    const pltb_shape *shapes;
    const pltb_bsdf *bsdfs;
    const pltb_texture *textures;
    const float *tex_data;
    const pltb_emitter *emitters;
    const float *area_cdf;
    const pltb_sensor *sensors;
    const float *cie_srgb;
    pltb_film film;
    uint32_t n_faces, n_shapes, n_emitters, n_sensors;
    float emitter_pmf;
    int32_t env;   // index of the environment emitter (Scene::environment(), src/render/scene.cpp:60-67), -1 if none
    uint32_t rgb;  // PLTB_RENDER_RGB (set per render call): Spectrum = Color3f in lanes 0..2 of S4, lane 3 repeats lane 2 (so per-lane
                   // arithmetic and max() are those of a 3-vector); the per-sample wavelength slot then holds the path's
                   // `sampling_wavelengths` (plt.py:64-70,158), which only roughgrating's RGB fallback reads
    BesselTables bessel;
};

struct Ray { V3 o, d; float maxt; };
struct Hit { float t, u, v; uint32_t prim; int32_t shape; uint32_t cls; };   // cls: material class of the shape (tri_data)
// closest hit as stored per path: {t, u, v, prim | cls << 28}; prim < 2^24 (pltb_scene_create), cls < 8; 0xFFFFFFFF = miss
constexpr uint32_t kHitPrimMask = 0x00FFFFFFu, kHitMiss = 0xFFFFFFFFu;
constexpr int kHitClsShift = 28;

PLTB_HD V3 ld3(const float *p, uint32_t i) { return mk3(PLTB_LDG(p + 3 * i), PLTB_LDG(p + 3 * i + 1), PLTB_LDG(p + 3 * i + 2)); }
PLTB_HD V2 ld2(const float *p, uint32_t i) { return mk2(PLTB_LDG(p + 2 * i), PLTB_LDG(p + 2 * i + 1)); }

// Sphere / rectangle intersection (src/shapes/sphere.cpp:393-436 -- double precision as in the reference;
// src/shapes/rectangle.cpp:262-283)
PLTB_HD bool sphere_intersect(const pltb_shape &sh, const Ray &ray, float &t_out) {
    double ox = (double) ray.o.x - (double) sh.center[0], oy = (double) ray.o.y - (double) sh.center[1],
           oz = (double) ray.o.z - (double) sh.center[2];
    double dx = ray.d.x, dy = ray.d.y, dz = ray.d.z, radius = sh.radius, maxt = ray.maxt;
    double A = dx * dx + dy * dy + dz * dz, B = 2.0 * (ox * dx + oy * dy + oz * dz),
           Cc = ox * ox + oy * oy + oz * oz - radius * radius;
    if (A == 0.0) return false;
    double discrim = B * B - 4.0 * A * Cc;
    if (discrim < 0) return false;
    double temp = -0.5 * (B + copysign(sqrt(discrim), B));
    double x0p = temp / A, x1p = Cc / temp;
    double near_t = fmin(x0p, x1p), far_t = fmax(x0p, x1p);
    bool out_bounds = !(near_t <= maxt && far_t >= 0.0), in_bounds = near_t < 0.0 && far_t > maxt;
    if (out_bounds || in_bounds) return false;
    t_out = (float) (near_t < 0.0 ? far_t : near_t);
    return true;
}
PLTB_HD bool rectangle_intersect(const pltb_shape &sh, const Ray &ray, float &t_out, float &u, float &v) {
    V3 o = xf_point(sh.to_object, ray.o), d = xf_vector(sh.to_object, ray.d);
    float t = -o.z / d.z;
    V3 local = fmadd3(d, t, o);
    bool active = t >= 0.f && t <= ray.maxt && fabsf(local.x) <= 1.f && fabsf(local.y) <= 1.f;
    if (!active) return false;
    t_out = t; u = local.x; v = local.y;
    return true;
}

// Scene::ray_intersect / ray_test (src/render/scene.cpp:176-205) over our own 4-wide BVH: ordered descent with a small stack of
// (entry distance, child) pairs so that subtrees behind the closest hit found so far are dropped when they are popped, without
// fetching their node; Moeller-Trumbore exactly as Mesh::moeller_trumbore (include/mitsuba/render/mesh.h:460-481).
// A node is one 128-byte line (pltb_bvh.h): seven 16-byte loads, four slab tests, a 5-exchange sorting network on keys
// (distance bits | child slot).  The traversal is a resumable state machine (one inner node or one leaf per step) so that
// k_trace can keep the lanes of a warp busy by handing finished lanes a new ray while the others are still walking the tree.
constexpr int kBvhStack = 64;   // entries; pltb_scene_create refuses scenes whose tree could need more (HostBVH::stack_need)

struct Traversal {
    Ray ray;
    float ix, iy, iz;
    int32_t node;   // >= 0: inner node index; < 0: leaf ~(first | count << 24)
    int sp;
    bool done;
    Hit best;
    // key << 32 | child reference.  The array lives outside the struct (the caller's local array): a dynamically indexed member
    // keeps the whole struct -- ray, node, best hit -- in local memory, and k_trace is bound by the L1 data pipe.
    unsigned long long *stack;

    PLTB_HD void push(unsigned long long e) { stack[sp++] = e; }
    PLTB_HD unsigned long long top_pop() { return stack[--sp]; }
    PLTB_HD void begin(const DeviceScene &sc, const Ray &r, unsigned long long *stack_mem) {
        ray = r; stack = stack_mem;
        best.t = inf_f(); best.u = best.v = 0.f; best.prim = 0; best.shape = -1; best.cls = 0;
        sp = 0; node = 0;
        // A ray with a NaN component (the reference lets paths continue with alpha = inf * 0, DESIGN.md "quirks") fails every
        // triangle / shape test (NaN comparisons are false) but would pass every slab test and walk the whole tree: report
        // the miss right away.
        done = sc.n_nodes == 0 || isnan_(r.o.x + r.o.y + r.o.z + r.d.x + r.d.y + r.d.z);
        ix = 1.f / r.d.x; iy = 1.f / r.d.y; iz = 1.f / r.d.z;
    }
    PLTB_HD void pop() {
        while (sp > 0) {
            const unsigned long long e = top_pop();
            // the key holds the entry distance rounded down: a subtree that starts behind the closest hit cannot improve it
            if (u2f((uint32_t) (e >> 32) & ~3u) <= ray.maxt) { node = (int32_t) (uint32_t) e; return; }
        }
        done = true;
    }
    PLTB_HD static uint32_t slab_key(float lox, float loy, float loz, float hix, float hiy, float hiz, float ox, float oy, float oz,
                                     float ix, float iy, float iz, float maxt, uint32_t slot) {
        const float t0x = (lox - ox) * ix, t1x = (hix - ox) * ix;
        const float t0y = (loy - oy) * iy, t1y = (hiy - oy) * iy;
        const float t0z = (loz - oz) * iz, t1z = (hiz - oz) * iz;
        const float n = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fmaxf(fminf(t0z, t1z), 0.f));
        const float f = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fminf(fmaxf(t0z, t1z), maxt));
        // boxes are padded at build time; the 1 + 4 ulp factor keeps the slab test conservative
        return n <= f * 1.0000005f ? ((f2u(n) & ~3u) | slot) : 0xFFFFFFFFu;
    }
    PLTB_HD void inner_step(const DeviceScene &sc) {
        const float4 *n = sc.bvh_nodes + 8 * (size_t) node;
        const float4 lx = PLTB_LDG(n), ly = PLTB_LDG(n + 1), lz = PLTB_LDG(n + 2), hx = PLTB_LDG(n + 3), hy = PLTB_LDG(n + 4),
                     hz = PLTB_LDG(n + 5), ch = PLTB_LDG(n + 6);
        const float ox = ray.o.x, oy = ray.o.y, oz = ray.o.z, mt = ray.maxt;
        uint32_t k0 = slab_key(lx.x, ly.x, lz.x, hx.x, hy.x, hz.x, ox, oy, oz, ix, iy, iz, mt, 0u);
        uint32_t k1 = slab_key(lx.y, ly.y, lz.y, hx.y, hy.y, hz.y, ox, oy, oz, ix, iy, iz, mt, 1u);
        uint32_t k2 = slab_key(lx.z, ly.z, lz.z, hx.z, hy.z, hz.z, ox, oy, oz, ix, iy, iz, mt, 2u);
        uint32_t k3 = slab_key(lx.w, ly.w, lz.w, hx.w, hy.w, hz.w, ox, oy, oz, ix, iy, iz, mt, 3u);
        // sorting network (0,1)(2,3)(0,2)(1,3)(1,2): ascending entry distance, misses (0xFFFFFFFF) last
        uint32_t t;
        t = umin_(k0, k1); k1 = umax_(k0, k1); k0 = t;
        t = umin_(k2, k3); k3 = umax_(k2, k3); k2 = t;
        t = umin_(k0, k2); k2 = umax_(k0, k2); k0 = t;
        t = umin_(k1, k3); k3 = umax_(k1, k3); k1 = t;
        t = umin_(k1, k2); k2 = umax_(k1, k2); k1 = t;
        if (k0 == 0xFFFFFFFFu) { pop(); return; }
        PLTB_ASSERT(sp + 3 <= kBvhStack);
        if (k3 != 0xFFFFFFFFu) push(((unsigned long long) k3 << 32) | child_ref(ch, k3));
        if (k2 != 0xFFFFFFFFu) push(((unsigned long long) k2 << 32) | child_ref(ch, k2));
        if (k1 != 0xFFFFFFFFu) push(((unsigned long long) k1 << 32) | child_ref(ch, k1));
        node = (int32_t) child_ref(ch, k0);
    }
    PLTB_HD static uint32_t child_ref(const float4 &ch, uint32_t key) {
        const uint32_t s = key & 3u;
        return f2u(s == 0 ? ch.x : s == 1 ? ch.y : s == 2 ? ch.z : ch.w);
    }
    template <bool AnyHit> PLTB_HD void leaf_step(const DeviceScene &sc) {
        uint32_t leaf = ~(uint32_t) node, first = leaf & 0xFFFFFFu, count = leaf >> 24;
        for (uint32_t i = first; i < first + count; ++i) {
            const float4 *tp = sc.tri_data + 3 * (size_t) i;
            float4 q0 = PLTB_LDG(tp), q1 = PLTB_LDG(tp + 1), q2 = PLTB_LDG(tp + 2);
            uint32_t prim = f2u(q2.y);
            int32_t shape = (int32_t) f2u(q2.z);
            PLTB_ASSERT(i < sc.n_prims && shape >= 0 && (uint32_t) shape < sc.n_shapes && f2u(q2.w) < 8u);
            float t, u = 0.f, v = 0.f;
            bool hit;
            if (prim < sc.n_faces) {
                V3 p0 = mk3(q0.x, q0.y, q0.z), e1 = mk3(q0.w, q1.x, q1.y), e2 = mk3(q1.z, q1.w, q2.x);
                V3 pvec = cross(ray.d, e2);
                float inv_det = rcp(dot(e1, pvec));
                V3 tvec = ray.o - p0;
                u = dot(tvec, pvec) * inv_det;
                hit = u >= 0.f && u <= 1.f;
                V3 qvec = cross(tvec, e1);
                v = dot(ray.d, qvec) * inv_det;
                hit &= v >= 0.f && u + v <= 1.f;
                t = dot(e2, qvec) * inv_det;
                hit &= t >= 0.f && t <= ray.maxt;
            } else {
                const pltb_shape &sh = sc.shapes[shape];
                if (sh.type == PLTB_SHAPE_SPHERE) hit = sphere_intersect(sh, ray, t);
                else hit = rectangle_intersect(sh, ray, t, u, v);
            }
            if (hit && t < best.t) {
                best.t = t; best.u = u; best.v = v; best.prim = prim; best.shape = shape; best.cls = f2u(q2.w);
                if (AnyHit) { done = true; return; }
                ray.maxt = t;
            }
        }
        pop();
    }
    template <bool AnyHit> PLTB_HD void step(const DeviceScene &sc) {
        if (node >= 0) inner_step(sc);
        else leaf_step<AnyHit>(sc);
    }
};

// one-ray-per-thread form (coherent primary rays in k_raygen): the same steps, run to completion
template <bool AnyHit> PLTB_NI Hit bvh_intersect(const DeviceScene &sc, Ray ray) {
    unsigned long long stack_mem[kBvhStack];
    Traversal tr;
    tr.begin(sc, ray, stack_mem);
    // while-while: all lanes walk inner nodes until each sits at a leaf (or is done), then the leaves are tested together
    // (one inner-or-leaf step per iteration ran both bodies at partial occupancy: k_raygen 78.8 -> 73.3 ms per 2^30 samples)
    while (!tr.done) {
        while (!tr.done && tr.node >= 0) tr.inner_step(sc);
        if (!tr.done) tr.template leaf_step<AnyHit>(sc);
    }
    return tr.best;
}

// ---- textures / spectra -------------------------------------------------------------------------------------------------
PLTB_HD float srgb_model_eval(const float *c, float wl) {  // include/mitsuba/render/srgb.h:9-23
    float c0 = PLTB_LDG(c), c1 = PLTB_LDG(c + 1), c2 = PLTB_LDG(c + 2);
    float v = fmadd(fmadd(c0, wl, c1), wl, c2);
    if (isinf_(c2)) return fmadd(sign1(c2), .5f, .5f);
    return fmaxf(0.f, fmadd(.5f * v, rsqrt_smooth(fmadd(v, v, 1.f)), .5f));
}
PLTB_HD float regular_eval(const float *vals, uint32_t count, float lo, float hi, float x) {  // distr_1d.h:364-379
    if (!(x >= lo && x <= hi)) return 0.f;
    float inv_interval = rcp((hi - lo) / (float) (count - 1));
    x = (x - lo) * inv_interval;
    uint32_t index = (uint32_t) fmaxf(x, 0.f);
    if (index > count - 2) index = count - 2;
    float y0 = PLTB_LDG(vals + index), y1 = PLTB_LDG(vals + index + 1), w1 = x - (float) index, w0 = 1.f - w1;
    return fmadd(w0, y0, w1 * y1);
}
PLTB_HD float irregular_eval(const float *nodes, const float *vals, uint32_t count, float x) {  // distr_1d.h:705-730
    if (!(x >= nodes[0] && x <= nodes[count - 1])) return 0.f;
    uint32_t a = 0, b = count;
    while (a < b) { uint32_t m = (a + b) / 2; if (nodes[m] < x) a = m + 1; else b = m; }
    uint32_t i0 = (a > 1u ? a : 1u) - 1u;
    if (i0 > count - 2) i0 = count - 2;
    float x0 = nodes[i0], x1 = nodes[i0 + 1], y0 = vals[i0], y1 = vals[i0 + 1];
    float t = (x - x0) / (x1 - x0);
    return fmadd(y1 - y0, t, y0);
}
PLTB_HD float blackbody_eval(float temperature, float wl) {  // src/spectra/blackbody.cpp:128-150
    const float c = 2.99792458e+8f, h = 6.62607004e-34f, k = 1.38064852e-23f;
    const float c0 = 2 * h * c * c, c1 = h * c / k;
    float lambda = wl * 1e-9f, lambda2 = lambda * lambda, lambda5 = lambda2 * lambda2 * lambda;
    return 1e-9f * c0 / (lambda5 * (expf(c1 / (lambda * temperature)) - 1.f));
}
PLTB_HD int wrap_coord(int pos, int shape, uint32_t mode) {  // ext/drjit/include/drjit/texture.h:1222-1250
    if (mode == PLTB_WRAP_CLAMP) return pos < 0 ? 0 : (pos > shape - 1 ? shape - 1 : pos);
    int shift_neg = pos < 0 ? pos + 1 : pos;
    int div = shift_neg / shape;
    int mod = pos - div * shape;
    if (mod < 0) mod += shape;
    if (mode == PLTB_WRAP_MIRROR && !(((div & 1) == 0) ^ (pos < 0))) mod = shape - 1 - mod;
    return mod;
}

PLTB_NI S4 tex_eval(const DeviceScene &sc, int32_t id, const S4 &wl, V2 uv) {
    if (id < 0) return s4(1.f);
    const pltb_texture &t = sc.textures[id];
    const float *data = sc.tex_data + t.offset;
    S4 r = s4(0.f);
    switch (t.type) {
        case PLTB_TEX_UNIFORM: return s4(t.p[0]);
        case PLTB_TEX_RGB: r.v[0] = t.p[0]; r.v[1] = t.p[1]; r.v[2] = r.v[3] = t.p[2]; return r;   // srgb.cpp:56-57 in an RGB variant
        case PLTB_TEX_SRGB:
            for (int i = 0; i < 4; ++i) r.v[i] = srgb_model_eval(t.p, wl.v[i]);
            return r;
        case PLTB_TEX_D65:
            for (int i = 0; i < 4; ++i) {
                float dv = regular_eval(data, t.count, t.p[3], t.p[4], wl.v[i]);
                r.v[i] = (t.flags & 1u) ? dv * srgb_model_eval(t.p, wl.v[i]) : dv;
            }
            return r;
        case PLTB_TEX_REGULAR:
            for (int i = 0; i < 4; ++i) r.v[i] = regular_eval(data, t.count, t.p[3], t.p[4], wl.v[i]);
            return r;
        case PLTB_TEX_IRREGULAR:
            for (int i = 0; i < 4; ++i) r.v[i] = irregular_eval(data, data + t.count, t.count, wl.v[i]);
            return r;
        case PLTB_TEX_BLACKBODY:
            for (int i = 0; i < 4; ++i)
                r.v[i] = (wl.v[i] >= t.p[3] && wl.v[i] <= t.p[4]) ? blackbody_eval(t.p[0], wl.v[i]) : 0.f;
            return r;
        case PLTB_TEX_BITMAP: {  // src/textures/bitmap.cpp:700-737
            float u = fmadd(t.to_uv[1], uv.y, fmadd(t.to_uv[0], uv.x, t.to_uv[2])),
                  v = fmadd(t.to_uv[4], uv.y, fmadd(t.to_uv[3], uv.x, t.to_uv[5]));
            float pu = fmadd(u, (float) t.width, -.5f), pv = fmadd(v, (float) t.height, -.5f);
            int iu = (int) floorf(pu), iv = (int) floorf(pv);
            int x0 = wrap_coord(iu, (int) t.width, t.wrap), x1 = wrap_coord(iu + 1, (int) t.width, t.wrap),
                y0 = wrap_coord(iv, (int) t.height, t.wrap), y1 = wrap_coord(iv + 1, (int) t.height, t.wrap);
            const float *v00 = data + 3 * ((size_t) y0 * t.width + x0), *v10 = data + 3 * ((size_t) y0 * t.width + x1),
                        *v01 = data + 3 * ((size_t) y1 * t.width + x0), *v11 = data + 3 * ((size_t) y1 * t.width + x1);
            float w1x = pu - (float) iu, w0x = 1.f - w1x, w1y = pv - (float) iv, w0y = 1.f - w1y;
            if (t.flags & PLTB_TEXFLAG_RGB) {   // RGB variants: linear RGB texels, dr::Texture::eval (drjit/texture.h:386-419)
                for (int c = 0; c < 3; ++c) {
                    float acc = fmadd(PLTB_LDG(v00 + c), w0x * w0y, 0.f);
                    acc = fmadd(PLTB_LDG(v10 + c), w1x * w0y, acc);
                    acc = fmadd(PLTB_LDG(v01 + c), w0x * w1y, acc);
                    r.v[c] = fmadd(PLTB_LDG(v11 + c), w1x * w1y, acc);
                }
                r.v[3] = r.v[2];
                return r;
            }
            for (int i = 0; i < 4; ++i) {
                float c00 = srgb_model_eval(v00, wl.v[i]), c10 = srgb_model_eval(v10, wl.v[i]),
                      c01 = srgb_model_eval(v01, wl.v[i]), c11 = srgb_model_eval(v11, wl.v[i]);
                float c0 = fmadd(w0x, c00, w1x * c10), c1 = fmadd(w0x, c01, w1x * c11);
                r.v[i] = fmadd(w0y, c0, w1y * c1);
            }
            return r;
        }
        default: return r;
    }
}

}  // namespace pltb
