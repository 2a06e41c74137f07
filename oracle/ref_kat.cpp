// ref_kat.cpp -- known-answer generator.  Compiled ONLY against the reference's own headers (oracle/ref_overlay.sh);
// instantiates the reference's scalar_spectral types and prints JSON known answers that tests/test_oracle_kat.py holds
// the CPU oracle (and through it the CUDA kernels) to.  No reference code is restated here: this file only calls it.
#include <mitsuba/core/math.h>
#include <mitsuba/core/random.h>
#include <mitsuba/core/spectrum.h>
#include <mitsuba/core/warp.h>
#include <mitsuba/plt/diffractiongrating.h>
#include <mitsuba/plt/plt.h>
#include <mitsuba/render/fresnel.h>
#include <mitsuba/render/microfacet.h>
#include <drjit/random.h>
#include <cstdarg>
#include <cstdio>

using namespace mitsuba;
using Float = float;
using Spec = Spectrum<float, 4>;
using Grating = DiffractionGrating<Float, Spec>;
using V3 = Vector<float, 3>;
using V2 = Vector<float, 2>;
using P2 = Point<float, 2>;
using V2i = Vector<int32_t, 2>;

static bool first_item;
static void open_arr(const char *name, bool first = false) { printf("%s\"%s\": [", first ? "" : ",\n", name); first_item = true; }
static void close_arr() { printf("]"); }
static void item(const char *fmt, ...) {
    va_list ap; va_start(ap, fmt);
    if (!first_item) printf(", ");
    first_item = false;
    vprintf(fmt, ap); va_end(ap);
}

int main() {
    printf("{\n");
    // ---- math::bessel_j (include/mitsuba/core/math.h) ----
    open_arr("bessel_j", true);
    const float xs[] = { 0.f, 0.05f, 0.5f, 0.9139f, 1.7f, 3.7f, 7.25f, 12.5f, 30.f, 44.f, 46.5f, 50.f, 80.f, -2.3f };
    for (float x : xs) for (int nu = 0; nu <= 4; ++nu)
        item("[%.9g, %d, %.9g]", x, nu, math::bessel_j<float>(x, nu));
    close_arr();
    // ---- DiffractionGrating (include/mitsuba/plt/diffractiongrating.h) ----
    struct G { float angle, ipx, ipy, q; uint32_t lobes; uint32_t type; bool radial; float mult, u, v; };
    const G gs[] = {
        { 1.509f, 0.7f, 0.4f, 0.04f, 7, 0, false, 10.f, 0.3f, 0.6f },   // cbox-grating parameters, non radial
        { 1.509f, 0.7f, 0.4f, 0.04f, 7, 0, true, 10.f, 0.3f, 0.6f },    // radial
        { 0.0f, 0.7f, 0.7f, 0.05f, 5, 0, false, 10.f, 0.5f, 0.25f },    // spectrograph parameters
        { 0.4f, 1.2f, 0.0f, 0.3f, 5, 1, false, 1.f, 0.1f, 0.9f },       // rectangular, 1-D
        { 0.9f, 0.3f, 0.5f, 0.1f, 9, 2, true, 2.f, 0.8f, 0.2f },        // linear, radial, max lobes
    };
    const V3 wis[] = { dr::normalize(V3(0.3f, -0.2f, 0.9f)), dr::normalize(V3(-0.6f, 0.1f, 0.3f)), V3(0.f, 0.f, 1.f) };
    const float wls[] = { 0.38f, 0.55f, 0.72f };
    open_arr("grating");
    int gi = 0;
    for (const G &g : gs) {
        uint32_t t = (g.type == 0 ? (uint32_t) DiffractionGratingType::Sinusoidal
                     : g.type == 1 ? (uint32_t) DiffractionGratingType::Rectangular
                                   : (uint32_t) DiffractionGratingType::Linear);
        if (g.radial) t |= (uint32_t) DiffractionGratingType::Radial;
        Grating gr(g.angle, V2(g.ipx, g.ipy), g.q, g.lobes, (DiffractionGratingType) t, g.mult, P2(g.u, g.v));
        for (const V3 &wi : wis) for (float wl : wls) {
            int half = (int) g.lobes / 2;
            for (int lx = -half; lx <= half; ++lx) for (int ly = -1; ly <= 1; ++ly) {
                V2i lobe(lx, ly);
                float I = gr.lobe_intensity(lobe, wi, wl);
                float pdf = gr.lobe_pdf(lobe, wi, wl);
                auto [d, valid] = gr.diffract(wi, lobe, wl);
                item("{\"g\": %d, \"wi\": [%.9g, %.9g, %.9g], \"wl\": %.9g, \"lobe\": [%d, %d], \"intensity\": %.9g, "
                     "\"pdf\": %.9g, \"dir\": [%.9g, %.9g, %.9g], \"valid\": %d}", gi, wi.x(), wi.y(), wi.z(), wl, lx, ly, I,
                     pdf, d.x(), d.y(), d.z(), (int) valid);
            }
        }
        ++gi;
    }
    close_arr();
    open_arr("grating_sample");
    gi = 0;
    for (const G &g : gs) {
        uint32_t t = (g.type == 0 ? (uint32_t) DiffractionGratingType::Sinusoidal
                     : g.type == 1 ? (uint32_t) DiffractionGratingType::Rectangular
                                   : (uint32_t) DiffractionGratingType::Linear);
        if (g.radial) t |= (uint32_t) DiffractionGratingType::Radial;
        Grating gr(g.angle, V2(g.ipx, g.ipy), g.q, g.lobes, (DiffractionGratingType) t, g.mult, P2(g.u, g.v));
        const float us[][2] = { { 0.83f, 0.21f }, { 0.02f, 0.97f }, { 0.5f, 0.5f }, { 0.61f, 0.38f }, { 0.999f, 0.001f } };
        for (auto &u : us) {
            auto [lobe, pdf] = gr.sample_lobe(V2(u[0], u[1]), wis[0], 0.55f);
            float al = gr.alpha(wis[0], 2.f * dr::Pi<float> / 0.55f);
            item("{\"g\": %d, \"u\": [%.9g, %.9g], \"lobe\": [%d, %d], \"pdf\": [%.9g, %.9g], \"alpha\": %.9g}", gi, u[0], u[1],
                 lobe.x(), lobe.y(), pdf.x(), pdf.y(), al);
        }
        ++gi;
    }
    close_arr();
    printf(",\n\"grating_params\": [");
    for (size_t i = 0; i < sizeof(gs) / sizeof(gs[0]); ++i)
        printf("%s[%.9g, %.9g, %.9g, %.9g, %u, %u, %d, %.9g, %.9g, %.9g]", i ? ", " : "", gs[i].angle, gs[i].ipx, gs[i].ipy, gs[i].q,
               gs[i].lobes, gs[i].type, (int) gs[i].radial, gs[i].mult, gs[i].u, gs[i].v);
    printf("]");
    // ---- Fresnel (include/mitsuba/render/fresnel.h) ----
    open_arr("fresnel");
    for (float c : { 1.f, 0.7f, 0.2f, 0.f, -0.3f, -1.f }) for (float eta : { 1.5f, 1.4f, 1.f / 1.4f }) {
        auto [r, ct, eit, eti] = fresnel<float>(c, eta);
        item("[%.9g, %.9g, %.9g, %.9g, %.9g, %.9g]", c, eta, r, ct, eit, eti);
    }
    close_arr();
    open_arr("fresnel_conductor");
    for (float c : { 1.f, 0.7f, 0.2f, 0.01f }) for (auto ek : { std::pair<float, float>{ 0.f, 1.f }, { 0.2f, 3.9f }, { 1.657f, 9.224f } })
        item("[%.9g, %.9g, %.9g, %.9g]", c, ek.first, ek.second, fresnel_conductor<float>(c, dr::Complex<float>(ek.first, ek.second)));
    close_arr();
    // ---- MicrofacetDistribution (include/mitsuba/render/microfacet.h) ----
    open_arr("microfacet");
    for (int type = 0; type < 2; ++type) for (auto al : { std::pair<float, float>{ 0.02f, 0.02f }, { 0.1f, 0.1f }, { 0.3f, 0.05f } }) for (int vis = 0; vis < 2; ++vis) {
        MicrofacetDistribution<float, Spec> distr(type == 0 ? MicrofacetType::Beckmann : MicrofacetType::GGX, al.first, al.second, vis != 0);
        for (const V3 &wi : wis) for (auto u : { std::pair<float, float>{ 0.3f, 0.8f }, { 0.95f, 0.05f }, { 0.5f, 0.5f } }) {
            auto [m, pdf] = distr.sample(wi, P2(u.first, u.second));
            float D = distr.eval(m), g1 = distr.smith_g1(wi, m);
            item("{\"type\": %d, \"au\": %.9g, \"av\": %.9g, \"vis\": %d, \"wi\": [%.9g, %.9g, %.9g], \"u\": [%.9g, %.9g], "
                 "\"m\": [%.9g, %.9g, %.9g], \"pdf\": %.9g, \"D\": %.9g, \"G1\": %.9g}", type, al.first, al.second, vis, wi.x(), wi.y(),
                 wi.z(), u.first, u.second, m.x(), m.y(), m.z(), pdf, D, g1);
        }
    }
    close_arr();
    // ---- RNG (include/mitsuba/core/random.h, drjit/random.h) ----
    open_arr("tea32");
    for (uint32_t v0 : { 0u, 1u, 12345u }) for (uint32_t v1 : { 0u, 1u, 4294967295u }) {
        auto [a, b] = sample_tea_32<uint32_t>(v0, v1, 4);
        item("[%u, %u, %u, %u]", v0, v1, a, b);
    }
    close_arr();
    open_arr("pcg32");
    for (auto ss : { std::pair<uint64_t, uint64_t>{ 0x853c49e6748fea9bULL, 0xda3e39cb94b95bdbULL }, { 42, 54 }, { 0, 1 } }) {
        dr::PCG32<uint32_t> rng(1, ss.first, ss.second);
        float f0 = rng.next_float32(), f1 = rng.next_float32(), f2 = rng.next_float32(), f3 = rng.next_float32();
        item("[%llu, %llu, %.9g, %.9g, %.9g, %.9g]", (unsigned long long) ss.first, (unsigned long long) ss.second, f0, f1, f2, f3);
    }
    close_arr();
    // ---- wavelength sampling (include/mitsuba/core/spectrum.h) ----
    open_arr("sample_rgb_spectrum");
    for (float u : { 0.f, 0.1f, 0.37f, 0.5f, 0.9f, 0.999f }) {
        auto [wl, w] = sample_rgb_spectrum<float>(u);
        item("[%.9g, %.9g, %.9g]", u, wl, w);
    }
    close_arr();
    printf("\n}\n");
    return 0;
}
