// ref_kat_pol.cpp -- second known-answer generator, compiled ONLY against the reference's own headers (oracle/ref_overlay.sh):
// the polarized building blocks (include/mitsuba/render/mueller.h, fresnel_polarized of fresnel.h, Dr.Jit's Matrix / Complex
// arithmetic) and further header-only functions the oracle restates (array / scalar division, warp.h, spectrum.h colour
// conversion, srgb.h, plt.h Coherence).  No reference code is restated here: this file only calls it.  Output:
// tests/golden/ref_kat_pol.json (oracle/gen_golden.py), checked by tests/test_oracle_kat.py.
#include <mitsuba/core/math.h>
#include <mitsuba/core/spectrum.h>
#include <mitsuba/core/vector.h>
#include <mitsuba/core/warp.h>
#include <mitsuba/plt/plt.h>
#include <mitsuba/render/fresnel.h>
#include <mitsuba/render/mueller.h>
#include <mitsuba/render/srgb.h>
#include <cstdarg>
#include <cstdio>

using namespace mitsuba;
using Float = float;
using Spec = Spectrum<float, 4>;
using Mueller = MuellerMatrix<Spec>;
using V3 = Vector<float, 3>;
using V2 = Vector<float, 2>;
using P2 = Point<float, 2>;
using Cx = dr::Complex<float>;
using CxS = dr::Complex<Spec>;

static bool first_item;
static void open_arr(const char *name, bool first = false) { printf("%s\"%s\": [", first ? "" : ",\n", name); first_item = true; }
static void close_arr() { printf("]"); }
static void item(const char *fmt, ...) {
    va_list ap; va_start(ap, fmt);
    if (!first_item) printf(", ");
    first_item = false;
    vprintf(fmt, ap); va_end(ap);
}
static void print_mueller(const Mueller &M) {   // 16 x 4 floats, row-major, as part of the current item
    printf("[");
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) for (int k = 0; k < 4; ++k)
        printf("%s%.9g", (i || j || k) ? ", " : "", M(i, j)[k]);
    printf("]");
}
static Mueller test_matrix(int seed) {
    Mueller M;
    uint32_t s = 1234567u + 977u * seed;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) for (int k = 0; k < 4; ++k) {
        s = s * 1664525u + 1013904223u;
        M(i, j)[k] = ((s >> 8) & 0xFFFF) / 65536.f * 2.f - 1.f;
    }
    return M;
}

int main() {
    printf("{\n");
    // ---- Dr.Jit: array / scalar is array * rcp(scalar) (ext/drjit/include/drjit/array_router.h:218-232) ----
    open_arr("array_div_scalar", true);
    for (float s : { 3.f, 0.7f, 1.3370001f, 123.456f, 1e-3f }) {
        V3 v = V3(0.3f, -1.7f, 2.9f) / s;
        Spec sp = Spec(0.25f, 1.1f, 7.3f, 0.0123f) / s;
        item("[%.9g, %.9g, %.9g, %.9g, %.9g, %.9g, %.9g, %.9g]", s, v.x(), v.y(), v.z(), sp[0], sp[1], sp[2], sp[3]);
    }
    close_arr();
    // ---- Dr.Jit Matrix: product order, scaled identity from a value, matrix * value, matrix / value ----
    open_arr("matrix_ops");
    {
        Mueller A = test_matrix(0), B = test_matrix(1);
        printf("{\"A\": "); print_mueller(A); printf(", \"B\": "); print_mueller(B);
        printf(", \"AB\": "); print_mueller(A * B);
        printf(", \"from_value\": "); print_mueller(Mueller(Spec(0.5f, 0.25f, 2.f, 3.f)));
        printf(", \"A_times_spec\": "); print_mueller(A * Spec(0.5f, 0.25f, 2.f, 3.f));
        printf(", \"A_div_float\": "); print_mueller(A / 0.7f);
        printf(", \"depolarizer\": "); print_mueller(mueller::depolarizer<Spec>(Spec(0.5f, 0.25f, 2.f, 3.f)));
        printf(", \"A_times_absorber\": "); print_mueller(A * mueller::absorber<Spec>(Spec(0.5f, 0.25f, 2.f, 3.f)));
        printf(", \"transpose\": "); print_mueller(dr::transpose(A));
        // dielectric.cpp:533-536 "hacky nan check": dr::select(weight > 1.0f, weight, 1.0) on a Mueller matrix
        Mueller W = A * 3.f;
        printf(", \"select_gt_one\": "); print_mueller(dr::select(W > 1.0f, W, 1.0));
        printf("}");
        first_item = false;
    }
    close_arr();
    // ---- fresnel_polarized (include/mitsuba/render/fresnel.h:147-275) ----
    open_arr("fresnel_polarized_real");
    for (float c : { 1.f, 0.7f, 0.2f, 0.05f, -0.3f, -0.9f, 1.7f, 6.5f }) for (float eta : { 1.5f, 1.4f, 1.f / 1.4f, 1.f }) {
        auto [a_s, a_p, ct, eit, eti] = fresnel_polarized<float>(c, eta);
        item("[%.9g, %.9g, %.9g, %.9g, %.9g, %.9g, %.9g, %.9g, %.9g]", c, eta, dr::real(a_s), dr::imag(a_s), dr::real(a_p), dr::imag(a_p), ct, eit, eti);
    }
    close_arr();
    open_arr("fresnel_polarized_complex");
    for (float c : { 1.f, 0.7f, 0.2f, 0.01f, 1.7f, 6.5f, 27.5f }) for (auto ek : { std::pair<float, float>{ 0.f, 1.f }, { 0.2f, 3.9f }, { 1.657f, 9.224f } }) {
        auto [a_s, a_p, ct, eit, eti] = fresnel_polarized<float>(c, Cx(ek.first, ek.second));
        item("[%.9g, %.9g, %.9g, %.9g, %.9g, %.9g, %.9g, %.9g]", c, ek.first, ek.second, dr::real(a_s), dr::imag(a_s), dr::real(a_p), dr::imag(a_p), ct);
    }
    close_arr();
    // ---- mueller.h ----
    open_arr("specular_reflection_complex");
    for (float c : { 1.f, 0.7f, 0.2f, 0.01f, 1.7f, 27.5f }) for (auto ek : { std::pair<float, float>{ 0.f, 1.f }, { 0.2f, 3.9f }, { 1.657f, 9.224f } }) {
        Mueller M = mueller::specular_reflection(Spec(c), CxS(Spec(ek.first), Spec(ek.second)));
        printf("%s{\"c\": %.9g, \"eta\": %.9g, \"k\": %.9g, \"M\": ", first_item ? "" : ", ", c, ek.first, ek.second); print_mueller(M); printf("}");
        first_item = false;
    }
    close_arr();
    open_arr("specular_reflection_real");
    for (float c : { 1.f, 0.7f, 0.2f, -0.3f, -0.9f }) for (float eta : { 1.4f, 1.f / 1.4f }) {
        Mueller M = mueller::specular_reflection(Spec(c), Spec(eta));
        printf("%s{\"c\": %.9g, \"eta\": %.9g, \"M\": ", first_item ? "" : ", ", c, eta); print_mueller(M); printf("}");
        first_item = false;
    }
    close_arr();
    open_arr("specular_transmission");
    for (float c : { 1.f, 0.7f, 0.2f, -0.3f, -0.9f }) for (float eta : { 1.4f, 1.f / 1.4f }) {
        Mueller M = mueller::specular_transmission(Spec(c), Spec(eta));
        printf("%s{\"c\": %.9g, \"eta\": %.9g, \"M\": ", first_item ? "" : ", ", c, eta); print_mueller(M); printf("}");
        first_item = false;
    }
    close_arr();
    const V3 dirs[] = { dr::normalize(V3(0.3f, -0.2f, 0.9f)), dr::normalize(V3(-0.6f, 0.1f, 0.3f)), V3(0.f, 0.f, 1.f),
                        dr::normalize(V3(0.1f, 0.8f, -0.4f)), V3(0.f, 0.f, -1.f) };
    open_arr("stokes_basis");
    for (const V3 &d : dirs) { V3 b = mueller::stokes_basis(d); item("[%.9g, %.9g, %.9g, %.9g, %.9g, %.9g]", d.x(), d.y(), d.z(), b.x(), b.y(), b.z()); }
    close_arr();
    open_arr("rotate_stokes_basis");
    for (const V3 &f : dirs) for (const V3 &t : dirs) {
        V3 cur = mueller::stokes_basis(f), tgt = dr::cross(f, t);
        if (dr::squared_norm(tgt) < 1e-6f) continue;
        Mueller R = mueller::rotate_stokes_basis<V3, Spec>(f, cur, tgt);
        printf("%s{\"forward\": [%.9g, %.9g, %.9g], \"current\": [%.9g, %.9g, %.9g], \"target\": [%.9g, %.9g, %.9g], \"M\": ", first_item ? "" : ", ",
               f.x(), f.y(), f.z(), cur.x(), cur.y(), cur.z(), tgt.x(), tgt.y(), tgt.z());
        print_mueller(R); printf("}");
        first_item = false;
    }
    close_arr();
    open_arr("rotate_mueller_basis");
    {
        int k = 0;
        for (const V3 &wi : dirs) for (const V3 &wo : dirs) {
            if (wi.z() <= 0.f || wo.z() <= 0.f) continue;
            // the frame juggling of conductor.cpp:270-300 around an arbitrary matrix
            V3 n(0, 0, 1), wo_hat = wo, wi_hat = wi;
            V3 s_in = dr::cross(n, -wo_hat), s_out = dr::cross(n, wi_hat);
            bool collinear = dr::all(s_in == V3(0));
            s_in = collinear ? V3(1, 0, 0) : dr::normalize(s_in);
            s_out = collinear ? V3(1, 0, 0) : dr::normalize(s_out);
            Mueller M = test_matrix(2 + k);
            Mueller R = mueller::rotate_mueller_basis<V3, Spec>(M, -wo_hat, s_in, mueller::stokes_basis(-wo_hat), wi_hat, s_out, mueller::stokes_basis(wi_hat));
            printf("%s{\"wi\": [%.9g, %.9g, %.9g], \"wo\": [%.9g, %.9g, %.9g], \"seed\": %d, \"M\": ", first_item ? "" : ", ", wi.x(), wi.y(), wi.z(),
                   wo.x(), wo.y(), wo.z(), 2 + k);
            print_mueller(R); printf("}");
            first_item = false; ++k;
        }
    }
    close_arr();
    // ---- spectrum.h: pdf_rgb_spectrum, colour conversion; srgb.h ----
    open_arr("pdf_rgb_spectrum");
    for (float wl : { 350.f, 360.f, 400.f, 538.f, 600.f, 829.f, 831.f }) item("[%.9g, %.9g]", wl, pdf_rgb_spectrum<float>(wl));
    close_arr();
    // (spectrum_to_srgb / spectrum_to_xyz need the CIE tables of src/core/spectrum.cpp, which is not header-only: not pinned here)
    open_arr("srgb_model_eval");
    for (auto co : { dr::Array<float, 3>(0.f, 0.f, 0.f), dr::Array<float, 3>(1e-4f, -0.1f, 20.f), dr::Array<float, 3>(-2e-5f, 0.02f, -4.f),
                     dr::Array<float, 3>(0.f, 0.f, dr::Infinity<float>), dr::Array<float, 3>(0.f, 0.f, -dr::Infinity<float>) })
        for (float wl : { 380.f, 550.f, 720.f }) {
            Spec r = srgb_model_eval<Spec>(co, Spec(wl));
            item("[%.9g, %.9g, %.9g, %.9g, %.9g]", co[0], co[1], (double) co[2], wl, r[0]);
        }
    close_arr();
    // ---- warp.h ----
    open_arr("warp");
    for (auto u : { std::pair<float, float>{ 0.3f, 0.8f }, { 0.95f, 0.05f }, { 0.5f, 0.5f }, { 0.f, 0.f }, { 0.999f, 0.25f } }) {
        P2 s(u.first, u.second);
        P2 dk = warp::square_to_uniform_disk_concentric(s);
        V3 ch = warp::square_to_cosine_hemisphere(s), us = warp::square_to_uniform_sphere(s), uc = warp::square_to_uniform_cone(s, 0.8f);
        P2 tr = warp::square_to_uniform_triangle(s);
        item("{\"u\": [%.9g, %.9g], \"disk_concentric\": [%.9g, %.9g], \"cosine_hemisphere\": [%.9g, %.9g, %.9g], \"uniform_sphere\": [%.9g, %.9g, %.9g], "
             "\"uniform_cone_0.8\": [%.9g, %.9g, %.9g], \"uniform_triangle\": [%.9g, %.9g]}", u.first, u.second, dk.x(), dk.y(), ch.x(), ch.y(), ch.z(),
             us.x(), us.y(), us.z(), uc.x(), uc.y(), uc.z(), tr.x(), tr.y());
    }
    close_arr();
    // ---- vector.h: coordinate_system; drjit sphere.h unit_angle ----
    open_arr("coordinate_system");
    for (const V3 &d : dirs) { auto [s, t] = coordinate_system(d); item("[%.9g, %.9g, %.9g, %.9g, %.9g, %.9g, %.9g, %.9g, %.9g]", d.x(), d.y(), d.z(), s.x(), s.y(), s.z(), t.x(), t.y(), t.z()); }
    close_arr();
    open_arr("unit_angle");
    for (const V3 &a : dirs) for (const V3 &b : dirs) item("[%.9g, %.9g, %.9g, %.9g, %.9g, %.9g, %.9g]", a.x(), a.y(), a.z(), b.x(), b.y(), b.z(), dr::unit_angle(a, b));
    close_arr();
    // ---- plt.h: Coherence ----
    open_arr("coherence");
    for (float k : { 2.f * dr::Pi<float> / 0.38f, 2.f * dr::Pi<float> / 0.55f, 2.f * dr::Pi<float> / 0.72f }) {
        Coherence<float, Spec> c(1e-18f, 1.0f);
        float det0 = c.inv_coherence_det(k);
        c.propagate(2.5f, true);
        float det1 = c.inv_coherence_det(k);
        item("[%.9g, %.9g, %.9g, %.9g]", k, det0, det1, c.opl);
    }
    close_arr();
    printf("\n}\n");
    return 0;
}
