// pltb_bvh.h -- host-side binned-SAH BVH builder producing the 4-wide node layout traversed by Traversal / bvh_intersect()
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
    std::vector<float> nodes;     // 32 floats per 4-wide node
    std::vector<float> tri_data;  // 12 floats per leaf-ordered primitive
    uint32_t n_nodes = 0, n_prims = 0;
    uint32_t stack_need = 0;      // traversal-stack entries the deepest path can require (checked against kBvhStack)
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
inline uint32_t f2u_(float f) { return fbits(f); }
}  // namespace detail
// Material class of a BSDF plugin type (the shade / NEE queue its vertices go to): its own type, except that the delta BSDFs
// (conductor, dielectric) and shapes without a BSDF share the conductor's class -- their code is small, and on the closed cbox
// every further queue makes the per-path state accesses of all queues sparser (measured: one queue per type cost 17 % there).
inline uint32_t material_class(uint32_t bsdf_type) {
#ifdef PLTB_CLASS_PER_TYPE
    return bsdf_type;
#else
    return (bsdf_type == PLTB_BSDF_NULL || bsdf_type == PLTB_BSDF_DIELECTRIC) ? (uint32_t) PLTB_BSDF_CONDUCTOR : bsdf_type;
#endif
}
namespace detail {

struct Builder {
    const pltb_scene_desc *d;
    std::vector<uint32_t> prims;   // < n_faces: triangle id, else n_faces + shape id
    std::vector<Box> boxes;        // per prims index (by primitive id slot)
    std::vector<float> cent;       // 3 per slot
    HostBVH *out;
#ifndef PLTB_BVH_LEAF
#define PLTB_BVH_LEAF 4   // primitives per leaf at which splitting stops
#endif
    static constexpr int kBins = 16, kLeaf = PLTB_BVH_LEAF, kMaxDepth = 48;

    // binary tree of the binned-SAH build; collapsed into 4-wide nodes by emit()
    struct TNode { Box box; int32_t left = -1, right = -1; uint32_t first = 0, count = 0; };
    std::vector<TNode> tree;

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
            // material class of the owning shape (the shade queue its hits are routed to): the BSDF plugin type, 0 without a BSDF
            const int32_t bid = d->shapes[f2u_(rec[10])].bsdf;
            rec[11] = bitsf(material_class(bid >= 0 ? d->bsdfs[bid].type : 0u));
            out->tri_data.insert(out->tri_data.end(), rec, rec + 12);
            out->n_prims++;
        }
        return (int32_t) ~(base | (count << 24));
    }

    // binned SAH split of prims[first, first + count); past kMaxDepth (degenerate input) median splits bound the depth
    int32_t build(uint32_t first, uint32_t count, int depth) {
        const int32_t idx = (int32_t) tree.size();
        tree.emplace_back();
        Box bounds, cb;
        for (uint32_t i = first; i < first + count; ++i) { bounds.grow(boxes[prims[i]]); cb.grow(&cent[3 * prims[i]]); }
        tree[idx].box = bounds; tree[idx].first = first; tree[idx].count = count;
        if (count <= (uint32_t) kLeaf) return idx;
        int best_axis = -1, best_bin = -1;
        float best_cost = std::numeric_limits<float>::infinity();
        for (int axis = 0; axis < 3 && depth < kMaxDepth; ++axis) {
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
            mid = first + count / 2;   // all centroids coincide, or the depth bound was reached
        } else {
            float ext = cb.hi[best_axis] - cb.lo[best_axis], scale = kBins / ext, lo = cb.lo[best_axis];
            auto it = std::partition(prims.begin() + first, prims.begin() + first + count, [&](uint32_t p) {
                return std::min(kBins - 1, (int) ((cent[3 * p + best_axis] - lo) * scale)) <= best_bin;
            });
            mid = (uint32_t) (it - prims.begin());
            if (mid == first || mid == first + count) mid = first + count / 2;
        }
        const int32_t l = build(first, mid - first, depth + 1);
        const int32_t r = build(mid, first + count - mid, depth + 1);
        tree[idx].left = l; tree[idx].right = r;
        return idx;
    }

    // Collapse the binary subtree at `t` (an inner node) into one 4-wide node: the children of `t`, with the inner child of
    // the largest surface area replaced by its own two children until there are four (or only leaves are left).
    // Node layout (8 float4 = one 128-byte line): lo.x[4] lo.y[4] lo.z[4] hi.x[4] hi.y[4] hi.z[4] child[4] pad[4]; empty slots
    // are far-away point boxes and reference an empty leaf.  Returns the node index; `need` receives the traversal-stack
    // entries the subtree can require (3 pushed siblings per level).
    int32_t emit(int32_t t, uint32_t &need) {
        int32_t kids[4]; int n = 0;
        kids[n++] = tree[t].left; kids[n++] = tree[t].right;
        while (n < 4) {
            int pick = -1; float best = -1.f;
            for (int i = 0; i < n; ++i)
                if (tree[kids[i]].left >= 0 && tree[kids[i]].box.area() > best) { best = tree[kids[i]].box.area(); pick = i; }
            if (pick < 0) break;
            const int32_t k = kids[pick];
            kids[pick] = tree[k].left; kids[n++] = tree[k].right;
        }
        const uint32_t idx = out->n_nodes++;
        out->nodes.resize(32 * (size_t) out->n_nodes, 0.f);
        int32_t ref[4]; Box bx[4]; uint32_t sub = 0; bool empty[4];
        for (int i = 0; i < 4; ++i) {
            empty[i] = i >= n || (tree[kids[i]].left < 0 && tree[kids[i]].count == 0);
            if (empty[i]) { ref[i] = (int32_t) ~0u; continue; }
            const TNode &k = tree[kids[i]];
            bx[i] = k.box;
            if (k.left < 0) ref[i] = make_leaf(k.first, k.count);
            else { uint32_t s = 0; ref[i] = emit(kids[i], s); sub = std::max(sub, s); }
            // pad the child boxes so that the device slab test is conservative w.r.t. the exact triangle test
            for (int a = 0; a < 3; ++a) {
                float e = 1e-5f * std::max(std::fabs(bx[i].lo[a]), std::fabs(bx[i].hi[a])) + 1e-6f * (bx[i].hi[a] - bx[i].lo[a]) + 1e-7f;
                bx[i].lo[a] -= e; bx[i].hi[a] += e;
            }
        }
        float *nd = &out->nodes[32 * (size_t) idx];
        // an empty slot is a far-away point box (an infinite one would pass the slab test of rays with maxt = FLT_MAX, whose
        // padded far distance overflows); should a ray ever aim at it, it finds an empty leaf
        const float far_away = 1e30f;
        for (int i = 0; i < 4; ++i) {
            for (int a = 0; a < 3; ++a) {
                nd[4 * a + i] = empty[i] ? far_away : bx[i].lo[a];
                nd[12 + 4 * a + i] = empty[i] ? far_away : bx[i].hi[a];
            }
            nd[24 + i] = bitsf((uint32_t) ref[i]);
        }
        need = 3u + sub;
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
    int32_t root = b.build(0, (uint32_t) b.prims.size(), 0);
    if (b.tree[root].left < 0) {
        // a single leaf: wrap it in one node whose other slots are empty
        Builder::TNode leaf = b.tree[root];
        b.tree.push_back(leaf); b.tree.push_back(Builder::TNode());
        b.tree[root].left = (int32_t) b.tree.size() - 2; b.tree[root].right = (int32_t) b.tree.size() - 1;
    }
    b.emit(root, out.stack_need);
}

}  // namespace pltb
