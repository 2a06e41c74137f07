// Host-side check of the 4-wide BVH (csrc/pltb_bvh.h builder + the Traversal state machine of csrc/pltb_scene.cuh compiled as
// plain C++): closest hits and occlusion of random rays against brute force over all primitives.  Test infrastructure only.
#include <stdio.h>
#include <stdlib.h>
#include <random>
#include <vector>
#include "pltb_scene.cuh"
#include "pltb_bvh.h"
using namespace pltb;

int main(int argc, char **argv) {
    const int n_tris = argc > 1 ? atoi(argv[1]) : 3000, n_rays = argc > 2 ? atoi(argv[2]) : 20000;
    const int mode = argc > 3 ? atoi(argv[3]) : 0;   // 0: random soup, 1: long thin slivers along x (deep, skewed trees)
    std::mt19937 rng(11);
    std::uniform_real_distribution<float> U(0.f, 1.f);
    std::vector<float> pos; std::vector<uint32_t> faces, face_shape;
    for (int t = 0; t < n_tris; ++t) {
        float c[3] = { 2 * U(rng) - 1, 2 * U(rng) - 1, 2 * U(rng) - 1 };
        if (mode == 1) { c[0] = (float) t / n_tris; c[1] *= 1e-3f; c[2] *= 1e-3f; }
        for (int v = 0; v < 3; ++v)
            for (int k = 0; k < 3; ++k) pos.push_back(c[k] + (mode == 1 && k == 0 ? 0.5f : 0.15f) * (U(rng) - .5f));
        for (int v = 0; v < 3; ++v) faces.push_back(3 * t + v);
        face_shape.push_back(0);
    }
    pltb_shape shapes[2] = {};
    shapes[0].type = PLTB_SHAPE_MESH; shapes[0].n_faces = n_tris; shapes[0].bsdf = shapes[1].bsdf = -1;
    shapes[1].type = PLTB_SHAPE_SPHERE; shapes[1].center[0] = 0.2f; shapes[1].center[1] = -0.1f; shapes[1].center[2] = 0.3f; shapes[1].radius = 0.25f;
    pltb_scene_desc d = {};
    d.n_vertices = 3 * n_tris; d.n_faces = n_tris; d.n_shapes = 2;
    d.positions = pos.data(); d.faces = faces.data(); d.face_shape = face_shape.data(); d.shapes = shapes;
    HostBVH bvh;
    build_bvh(&d, bvh);
    DeviceScene sc = {};
    sc.bvh_nodes = (const float4 *) bvh.nodes.data(); sc.tri_data = (const float4 *) bvh.tri_data.data();
    sc.n_nodes = bvh.n_nodes; sc.n_prims = bvh.n_prims; sc.n_faces = n_tris; sc.shapes = shapes; sc.n_shapes = 2;
    long bad = 0, hits = 0, steps = 0; int max_sp = 0;
    for (int r = 0; r < n_rays; ++r) {
        Ray ray;
        ray.o = mk3(3 * U(rng) - 1.5f, 3 * U(rng) - 1.5f, 3 * U(rng) - 1.5f);
        V3 tgt = mk3(2 * U(rng) - 1, 2 * U(rng) - 1, 2 * U(rng) - 1);
        ray.d = normalize(tgt - ray.o);
        if (r % 7 == 0) ray.d.y = 0.f;   // axis-aligned components: infinite reciprocals
        ray.maxt = r % 3 ? kLargest : 1.5f;
        // brute force over the leaf-ordered records with the same primitive test
        unsigned long long stk_bf[kBvhStack], stk_tr[kBvhStack], stk_sh[kBvhStack];
        Traversal bf; bf.begin(sc, ray, stk_bf); bf.done = false;
        for (uint32_t i = 0; i < bvh.n_prims; ++i) { bf.node = (int32_t) ~(i | (1u << 24)); bf.sp = 0; bf.leaf_step<false>(sc); bf.done = false; }
        Traversal tr; tr.begin(sc, ray, stk_tr);
        while (!tr.done) { tr.step<false>(sc); steps++; if (tr.sp > max_sp) max_sp = tr.sp; }
        Traversal sh; sh.begin(sc, ray, stk_sh);
        while (!sh.done) sh.step<true>(sc);
        const bool same = tr.best.shape == bf.best.shape && (bf.best.shape < 0 || (tr.best.t == bf.best.t && tr.best.prim == bf.best.prim));
        const bool occl = (sh.best.shape >= 0) == (bf.best.shape >= 0);
        if (!same || !occl) { if (bad++ < 5) printf("MISMATCH ray %d: bvh t %g prim %u, brute t %g prim %u\n", r, tr.best.t, tr.best.prim, bf.best.t, bf.best.prim); }
        hits += bf.best.shape >= 0;
    }
    printf("{\"rays\": %d, \"hits\": %ld, \"mismatches\": %ld, \"nodes\": %u, \"prims\": %u, \"stack_need\": %u, \"max_sp\": %d, \"steps_per_ray\": %.2f}\n",
           n_rays, hits, bad, bvh.n_nodes, bvh.n_prims, bvh.stack_need, max_sp, (double) steps / n_rays);
    return bad || max_sp > (int) bvh.stack_need ? 1 : 0;
}
