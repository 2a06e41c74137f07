// pltb_bvh.h -- host-side binned-SAH BVH builder producing the 2-wide node layout traversed by bvh_intersect()
// (pltb_scene.cuh).  Replaces the reference's kd-tree / Embree / OptiX acceleration structures
// (include/mitsuba/render/kdtree.h, src/render/scene_embree.inl, scene_optix.inl); the config scenes have <= 3k
// primitives, so a host build takes well under a millisecond per thousand triangles.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>

#include "../../include/pltb.h"

namespace pltb {

struct HostBVH {
    std::vector<float> nodes;     // 16 floats per node
    std::vector<float> tri_data;  // 12 floats per leaf-ordered primitive
    uint32_t n_nodes = 0, n_prims = 0;
};

namespace detail {
struct Box {
    float lo[3], hi[3];
    Box() { for (int k = 0; k < 3; ++k) { lo[k] = std::numeric_limits<float>::infinity(); hi[k] = -lo[k]; } }
    void grow(const Box &b) { for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], b.lo[k]); hi[k] = std::max(hi[k], b.hi[k]); } }
    void grow(const float *p) { for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], p[k]); hi[k] = std::max(hi[k], p[k]); } }
    float area() const {
        float d[3] = { hi[0] - lo[0], hi[1] - lo[1], hi[2] - lo[2] };
        if (d[0] < 0) return 0.f;
        return 2.f * (d[0] * d[1] + d[1] * d[2] + d[2] * d[0]);
    }
};
inline uint32_t fbits(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
inline float bitsf(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

struct Builder {
    const pltb_scene_desc *d;
    std::vector<uint32_t> prims;   // < n_faces: triangle id, else n_faces + shape id
    std::vector<Box> boxes;        // per prims index (by primitive id slot)
    std::vector<float> cent;       // 3 per slot
    HostBVH *out;
#ifndef PLTB_BVH_LEAF
#define PLTB_BVH_LEAF 4   // primitives per leaf at which splitting stops
#endif
    static constexpr int kBins = 16, kLeaf = PLTB_BVH_LEAF;

    int32_t make_leaf(uint32_t first, uint32_t count) {
        uint32_t base = out->n_prims;
        for (uint32_t i = first; i < first + count; ++i) {
            uint32_t p = prims[i];
            float rec[12] = { 0 };
            if (p < d->n_faces) {
                const float *p0 = d->positions + 3 * d->faces[3 * p], *p1 = d->positions + 3 * d->faces[3 * p + 1],
                            *p2 = d->positions + 3 * d->faces[3 * p + 2];
                for (int k = 0; k < 3; ++k) { rec[k] = p0[k]; rec[3 + k] = p1[k] - p0[k]; rec[6 + k] = p2[k] - p0[k]; }
                rec[9] = bitsf(p); rec[10] = bitsf(d->face_shape[p]);
            } else {
                rec[9] = bitsf(p); rec[10] = bitsf(p - d->n_faces);
            }
            out->tri_data.insert(out->tri_data.end(), rec, rec + 12);
            out->n_prims++;
        }
        return (int32_t) ~(base | (count << 24));
    }

    // returns child reference (>= 0 inner node index, < 0 leaf) and its bounds
    int32_t build(uint32_t first, uint32_t count, Box &bounds) {
        bounds = Box();
        Box cb;
        for (uint32_t i = first; i < first + count; ++i) { bounds.grow(boxes[prims[i]]); cb.grow(&cent[3 * prims[i]]); }
        if (count <= (uint32_t) kLeaf) return make_leaf(first, count);
        int best_axis = -1, best_bin = -1;
        float best_cost = std::numeric_limits<float>::infinity();
        for (int axis = 0; axis < 3; ++axis) {
            float ext = cb.hi[axis] - cb.lo[axis];
            if (!(ext > 0)) continue;
            Box bb[kBins]; uint32_t bc[kBins] = { 0 };
            float scale = kBins / ext;
            for (uint32_t i = first; i < first + count; ++i) {
                int b = std::min(kBins - 1, (int) ((cent[3 * prims[i] + axis] - cb.lo[axis]) * scale));
                bb[b].grow(boxes[prims[i]]); bc[b]++;
            }
            float la[kBins], ra[kBins]; uint32_t lc[kBins], rc[kBins];
            Box acc; uint32_t n = 0;
            for (int b = 0; b < kBins; ++b) { acc.grow(bb[b]); n += bc[b]; la[b] = acc.area(); lc[b] = n; }
            acc = Box(); n = 0;
            for (int b = kBins - 1; b >= 0; --b) { acc.grow(bb[b]); n += bc[b]; ra[b] = acc.area(); rc[b] = n; }
            for (int b = 0; b < kBins - 1; ++b) {
                if (lc[b] == 0 || rc[b + 1] == 0) continue;
                float cost = la[b] * lc[b] + ra[b + 1] * rc[b + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = axis; best_bin = b; }
            }
        }
        uint32_t mid;
        if (best_axis < 0) {
            mid = first + count / 2;   // all centroids coincide
        } else {
            float ext = cb.hi[best_axis] - cb.lo[best_axis], scale = kBins / ext, lo = cb.lo[best_axis];
            auto it = std::partition(prims.begin() + first, prims.begin() + first + count, [&](uint32_t p) {
                return std::min(kBins - 1, (int) ((cent[3 * p + best_axis] - lo) * scale)) <= best_bin;
            });
            mid = (uint32_t) (it - prims.begin());
            if (mid == first || mid == first + count) mid = first + count / 2;
        }
        uint32_t idx = out->n_nodes++;
        out->nodes.resize(16 * (size_t) out->n_nodes, 0.f);
        Box b0, b1;
        int32_t c0 = build(first, mid - first, b0);
        int32_t c1 = build(mid, first + count - mid, b1);
        float *n = &out->nodes[16 * (size_t) idx];
        // pad the child boxes so that the device slab test is conservative w.r.t. the exact triangle test
        auto pad = [](Box &b) {
            for (int k = 0; k < 3; ++k) {
                float e = 1e-5f * std::max(std::fabs(b.lo[k]), std::fabs(b.hi[k])) + 1e-6f * (b.hi[k] - b.lo[k]) + 1e-7f;
                b.lo[k] -= e; b.hi[k] += e;
            }
        };
        pad(b0); pad(b1);
        n[0] = b0.lo[0]; n[1] = b0.lo[1]; n[2] = b0.lo[2]; n[3] = b0.hi[0];
        n[4] = b0.hi[1]; n[5] = b0.hi[2]; n[6] = b1.lo[0]; n[7] = b1.lo[1];
        n[8] = b1.lo[2]; n[9] = b1.hi[0]; n[10] = b1.hi[1]; n[11] = b1.hi[2];
        n[12] = bitsf((uint32_t) c0); n[13] = bitsf((uint32_t) c1);
        return (int32_t) idx;
    }
};
}  // namespace detail

inline void build_bvh(const pltb_scene_desc *d, HostBVH &out) {
    using namespace detail;
    Builder b;
    b.d = d; b.out = &out;
    const uint32_t slots = d->n_faces + d->n_shapes;
    b.boxes.resize(slots); b.cent.assign(3 * (size_t) slots, 0.f);
    for (uint32_t f = 0; f < d->n_faces; ++f) {
        for (int c = 0; c < 3; ++c) b.boxes[f].grow(d->positions + 3 * d->faces[3 * f + c]);
        b.prims.push_back(f);
    }
    for (uint32_t s = 0; s < d->n_shapes; ++s) {
        const pltb_shape &sh = d->shapes[s];
        if (sh.type == PLTB_SHAPE_MESH) continue;
        Box &bx = b.boxes[d->n_faces + s];
        if (sh.type == PLTB_SHAPE_SPHERE) {
            for (int k = 0; k < 3; ++k) { bx.lo[k] = sh.center[k] - sh.radius * 1.0001f; bx.hi[k] = sh.center[k] + sh.radius * 1.0001f; }
        } else {
            for (int c = 0; c < 4; ++c) {
                float lx = (c & 1) ? 1.f : -1.f, ly = (c & 2) ? 1.f : -1.f, p[3];
                for (int k = 0; k < 3; ++k) p[k] = sh.to_world[4 * k] * lx + sh.to_world[4 * k + 1] * ly + sh.to_world[4 * k + 3];
                bx.grow(p);
            }
        }
        b.prims.push_back(d->n_faces + s);
    }
    for (uint32_t p : b.prims)
        for (int k = 0; k < 3; ++k) b.cent[3 * p + k] = 0.5f * (b.boxes[p].lo[k] + b.boxes[p].hi[k]);
    out = HostBVH();
    if (b.prims.empty()) return;
    Box root;
    int32_t r = b.build(0, (uint32_t) b.prims.size(), root);
    if (r < 0) {
        // a single leaf: wrap it in one inner node whose second child is an empty leaf
        out.n_nodes = 1;
        out.nodes.assign(16, 0.f);
        float *n = out.nodes.data();
        n[0] = root.lo[0] - 1e-4f; n[1] = root.lo[1] - 1e-4f; n[2] = root.lo[2] - 1e-4f; n[3] = root.hi[0] + 1e-4f;
        n[4] = root.hi[1] + 1e-4f; n[5] = root.hi[2] + 1e-4f;
        for (int k = 0; k < 6; ++k) n[6 + k] = n[k];
        n[12] = bitsf((uint32_t) r); n[13] = bitsf(~0u);   // second child: empty leaf
    }
}

}  // namespace pltb
