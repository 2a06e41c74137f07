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
