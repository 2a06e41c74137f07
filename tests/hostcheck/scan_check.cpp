// Host-side check of GratingLoop's lobe-index ranges (csrc/pltb_bsdf.cuh): the ranged (lx, ly) loops must list exactly the
// lobes the full [-half, half]^2 loops list.  Test infrastructure only: compiles the device headers as plain C++.
#include <stdio.h>
#include <stdlib.h>
#include <random>
#include "pltb_bsdf.cuh"
using namespace pltb;

int main(int argc, char **argv) {
    const long N = argc > 1 ? atol(argv[1]) : 2000000;
    std::mt19937 rng(7);
    std::uniform_real_distribution<float> U(0.f, 1.f);
    long hits = 0, iters_full = 0, iters_ranged = 0, bad = 0;
    for (long n = 0; n < N; ++n) {
        pltb_bsdf b{};
        b.type = PLTB_BSDF_ROUGHGRATING;
        const int cfg = n % 4;
        b.f[1] = b.f[2] = cfg == 0 ? 0.02f : cfg == 1 ? 0.03f : cfg == 2 ? 0.01f : 0.2f * U(rng) + 0.005f;
        b.f[3] = cfg == 0 ? 1.509f : cfg == 1 ? 0.f : 6.28f * U(rng);
        b.f[4] = cfg == 0 ? 0.7f : cfg == 1 ? 0.7f : 0.9f * U(rng) + 0.05f;
        b.f[5] = cfg == 0 ? 0.4f : cfg == 1 ? 0.7f : (cfg == 2 ? 0.9f : (U(rng) < 0.2f ? 0.f : U(rng)));
        b.f[6] = 0.04f; b.f[7] = 10.f;
        b.u[2] = cfg == 1 ? 5 : (cfg == 3 ? 1 + 2 * (int) (U(rng) * 4.99f) : 7);
        b.u[3] = 0 | ((cfg == 0 || (cfg == 3 && U(rng) < 0.5f)) ? 0x10 : 0);
        auto dir = [&](bool up) {
            float z = up ? 0.02f + 0.98f * U(rng) : 2 * U(rng) - 1, ph = kTwoPi * U(rng), r = sqrtf(fmaxf(0.f, 1 - z * z));
            return mk3(r * cosf(ph), r * sinf(ph), z);
        };
        V3 wi = dir(true), wo;
        V2 uv = mk2(U(rng), U(rng));
        S4 wl; for (int i = 0; i < 4; ++i) wl.v[i] = 360.f + 470.f * U(rng);
        if (U(rng) < 0.7f) {   // aim near an actual lobe so that windows are hit often
            Grating g = make_grating(b, uv);
            int half = (int) g.lobes / 2;
            int lx = (int) (U(rng) * (2 * half + 1)) - half, ly = (int) (U(rng) * (2 * half + 1)) - half;
            bool valid; V3 c = g.diffract(wi, lx, ly, wl.v[(int) (U(rng) * 3.99f)] * 1e-3f, valid);
            V3 j = mk3(U(rng) - .5f, U(rng) - .5f, U(rng) - .5f) * (0.1f * b.f[1] * 4);
            wo = normalize(c + j);
            if (!valid || !finite_(wo.x + wo.y + wo.z)) wo = dir(false);
        } else wo = dir(false);
        GratingLoop gl; gl.init(b, wi, uv, wl, wo);
        GratingHits a, f;
        grating_scan(gl, wo, a);
        iters_ranged += (long) (gl.lx1 >= gl.lx0 ? gl.lx1 - gl.lx0 + 1 : 0) * (gl.ly1 >= gl.ly0 ? gl.ly1 - gl.ly0 + 1 : 0);
        gl.lx0 = gl.ly0 = -gl.half; gl.lx1 = gl.ly1 = gl.half;
        grating_scan(gl, wo, f);
        iters_full += (long) (2 * gl.half + 1) * (2 * gl.half + 1);
        hits += f.n > 0;
        bool same = a.n == f.n && a.need == f.need;
        for (int k = 0; same && k < f.n && k < GratingHits::kMax; ++k) same = a.hit[k] == f.hit[k];
        if (!same) { if (bad++ < 5) printf("MISMATCH n=%ld ranged %d full %d\n", n, a.n, f.n); }
    }
    printf("{\"items\": %ld, \"with_hits\": %ld, \"mismatches\": %ld, \"lobe_iters_full\": %ld, \"lobe_iters_ranged\": %ld}\n",
           N, hits, bad, iters_full, iters_ranged);
    return bad ? 1 : 0;
}
