/*
 * plt_oracle.cpp -- CPU ORACLE (test infrastructure, NOT the product).
 *
 * A scalar float32 C++17 restatement of the reference's `plt` forward render path, one Monte-Carlo sample at a
 * time, written function by function after the reference files cited below (paths relative to the reference
 * tree, ZenithGD/mitsuba3-plt @ bc454f3d).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` legs may load this library; the product path (libpltb200.so) never does.
 *
 * PARITY STATUS: the reference has no test or golden vector for plt.py / roughgrating.cpp / include/mitsuba/plt
 *  and the fork cannot be built here, so integrator-level parity is UNPINNED.  Component level
 * pins: TEA / PCG32 (src/core/tests/test_random.py, src/samplers/tests/test_independent.py), Fresnel and
 * microfacet values (src/render/tests/test_fresnel.py, test_microfacet.py) and KATs produced by compiling the
 * reference's own PLT headers (oracle/ref_kat.cpp -> tests/golden/ref_kat.json); see tests/test_oracle_kat.py.
 *
 * Semantics follow the JIT (llvm_/cuda_) variants, the only ones in which plt.py runs:
 * masked lanes of a virtual call return zero, and `active &=` inside a callee does not mask a result that the
 * callee returns unmasked (e.g. RoughGrating::wbsdf_eval, roughgrating.cpp:967).
 *
 * Numerics contract shared with the CUDA kernels (DESIGN.md "Numerics contract"): float32 everywhere except
 * the PCG32 state; no implicit FMA contraction (-ffp-contract=off); fmadd/fmsub/fnmadd are written only where
 * the reference writes dr::fmadd & co; dot(a,b) = fma(a.z,b.z, fma(a.y,b.y, a.x*b.x)) (drjit array_base.h:597);
 * cross = fmsub form (array_router.h:665); rcp = 1/x, rsqrt = 1/sqrt(x) (array_utils.h:213-219).
 */
#include "../include/pltb.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

namespace {

constexpr float kPi = 3.14159265358979323846f;
constexpr float kInvPi = 0.31830988618379067154f;
constexpr float kInvFourPi = 0.07957747154594766788f;
constexpr float kTwoPi = 6.28318530717958647692f;
constexpr float kFourPi = 12.56637061435917295385f;
constexpr float kInvSqrtPi = 0.56418958354775628695f;
constexpr float kEps = 5.9604644775390625e-8f;   /* dr::Epsilon<float> = 2^-24 */
constexpr float kRayEps = kEps * 1500.f;         /* include/mitsuba/core/math.h:20 */
constexpr float kShadowEps = kRayEps * 10.f;     /* math.h:22 */
constexpr float kInf = std::numeric_limits<float>::infinity();
constexpr float kLargest = std::numeric_limits<float>::max();

/* BSDFFlags, include/mitsuba/render/bsdf.h:60-120 */
enum : uint32_t {
    F_Null = 0x1, F_DiffuseReflection = 0x2, F_DiffuseTransmission = 0x4, F_GlossyReflection = 0x8,
    F_GlossyTransmission = 0x10, F_DeltaReflection = 0x20, F_DeltaTransmission = 0x40,
    F_Smooth = 0x2 | 0x4 | 0x8 | 0x10, F_Delta = 0x1 | 0x20 | 0x40
};

inline float fmadd(float a, float b, float c) { return std::fmaf(a, b, c); }
inline float fmsub(float a, float b, float c) { return std::fmaf(a, b, -c); }
inline float fnmadd(float a, float b, float c) { return std::fmaf(-a, b, c); }
inline float sqr(float x) { return x * x; }
inline float rcp(float x) { return 1.f / x; }
inline float rsqrt(float x) { return 1.f / std::sqrt(x); }
inline float safe_sqrt(float x) { return std::sqrt(std::max(x, 0.f)); }
inline float mulsign(float a, float b) { return std::signbit(b) ? -a : a; }
inline float mulsign_neg(float a, float b) { return std::signbit(b) ? a : -a; }
inline float sign1(float v) { return std::copysign(1.f, v); }

struct V2 { float x = 0, y = 0; };
struct V3 {
    float x = 0, y = 0, z = 0;
    V3() = default;
    V3(float a, float b, float c) : x(a), y(b), z(c) {}
};
inline V3 operator+(V3 a, V3 b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
inline V3 operator-(V3 a, V3 b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
inline V3 operator-(V3 a) { return { -a.x, -a.y, -a.z }; }
inline V3 operator*(V3 a, float s) { return { a.x * s, a.y * s, a.z * s }; }
/* Dr.Jit: array / scalar is array * rcp(scalar) ("reciprocal approximation", ext/drjit/include/drjit/array_router.h:218-232;
 * Python: src/python/base.cpp:220-237); rcp of a scalar float is 1 / x.  Pinned by tests/golden/ref_kat_pol.json. */
inline V3 operator/(V3 a, float s) { const float r = 1.f / s; return { a.x * r, a.y * r, a.z * r }; }
inline float dot(V3 a, V3 b) { return fmadd(a.z, b.z, fmadd(a.y, b.y, a.x * b.x)); }
inline float sqnorm(V3 a) { return fmadd(a.z, a.z, fmadd(a.y, a.y, a.x * a.x)); }
inline float norm(V3 a) { return std::sqrt(sqnorm(a)); }
inline V3 normalize(V3 a) { return a * rsqrt(sqnorm(a)); }
inline V3 cross(V3 a, V3 b) {
    return { fmsub(a.y, b.z, a.z * b.y), fmsub(a.z, b.x, a.x * b.z), fmsub(a.x, b.y, a.y * b.x) };
}
inline V3 fmadd(V3 a, float s, V3 c) { return { fmadd(a.x, s, c.x), fmadd(a.y, s, c.y), fmadd(a.z, s, c.z) }; }

struct S4 {
    float v[4] = { 0, 0, 0, 0 };
    S4() = default;
    explicit S4(float a) { v[0] = v[1] = v[2] = v[3] = a; }
    float &operator[](int i) { return v[i]; }
    float operator[](int i) const { return v[i]; }
};
inline S4 operator*(S4 a, S4 b) { S4 r; for (int i = 0; i < 4; ++i) r[i] = a[i] * b[i]; return r; }
inline S4 operator*(S4 a, float b) { S4 r; for (int i = 0; i < 4; ++i) r[i] = a[i] * b; return r; }
inline S4 operator+(S4 a, S4 b) { S4 r; for (int i = 0; i < 4; ++i) r[i] = a[i] + b[i]; return r; }
inline S4 operator/(S4 a, float b) { S4 r; const float rb = 1.f / b; for (int i = 0; i < 4; ++i) r[i] = a[i] * rb; return r; }
inline float smax(S4 a) { return std::max(std::max(a[0], a[1]), std::max(a[2], a[3])); }

/* include/mitsuba/core/vector.h:116-136 */
inline void coordinate_system(V3 n, V3 &s, V3 &t) {
    float sign = sign1(n.z), a = -rcp(sign + n.z), b = n.x * n.y * a;
    s = V3(mulsign(sqr(n.x) * a, n.z) + 1.f, mulsign(b, n.z), mulsign_neg(n.x, n.z));
    t = V3(b, fmadd(n.y, n.y * a, sign), -n.y);
}

/* include/mitsuba/core/frame.h */
struct Frame {
    V3 s, t, n;
    Frame() = default;
    explicit Frame(V3 v) : n(v) { coordinate_system(v, s, t); }
    V3 to_local(V3 v) const { return { dot(v, s), dot(v, t), dot(v, n) }; }
    V3 to_world(V3 v) const { return fmadd(n, v.z, fmadd(t, v.y, s * v.x)); }
};

/* ------------------------------------------------------------------------------------------------------------
 * RNG: sample_tea_32 (include/mitsuba/core/random.h:77-90), PCG32 (ext/drjit/include/drjit/random.h:57-138),
 * PCG32Sampler::seed (src/render/sampler.cpp:125-145)
 * ---------------------------------------------------------------------------------------------------------- */
inline void sample_tea_32(uint32_t v0, uint32_t v1, uint32_t &o0, uint32_t &o1, int rounds = 4) {
    uint32_t sum = 0;
    for (int i = 0; i < rounds; ++i) {
        sum += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    o0 = v0; o1 = v1;
}

struct PCG32 {
    uint64_t state = 0, inc = 0;
    void seed(uint64_t initstate, uint64_t initseq) {
        state = 0;
        inc = (initseq << 1) | 1u;
        next_uint32();
        state += initstate;
        next_uint32();
    }
    uint32_t next_uint32() {
        uint64_t old = state;
        state = old * 0x5851f42d4c957f2dULL + inc;
        uint32_t xorshifted = (uint32_t) (((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t) (old >> 59u);
        return (xorshifted >> rot) | (xorshifted << ((-(int32_t) rot) & 31));
    }
    float next_float32() {
        uint32_t u = (next_uint32() >> 9) | 0x3f800000u;
        float f;
        std::memcpy(&f, &u, 4);
        return f - 1.f;
    }
};

/* ------------------------------------------------------------------------------------------------------------
 * Special functions
 * ---------------------------------------------------------------------------------------------------------- */

/* drjit::erf, float path (ext/drjit/include/drjit/math.h:1497-1553); estrin evaluation order as
 * drjit::detail::estrin_impl (pairs combined with fmadd, powers by squaring). */
inline float estrin(float x, const float *c, int n) {
    float tmp[16];
    int m = (n + 1) / 2;
    for (int i = 0; i < n / 2; ++i)
        tmp[i] = fmadd(x, c[2 * i + 1], c[2 * i]);
    if (n & 1)
        tmp[m - 1] = c[n - 1];
    if (m == 1)
        return tmp[0];
    return estrin(x * x, tmp, m);
}

inline float dr_erf(float x) {
    float xa = std::fabs(x), x2 = x * x;
    static const float c0c[] = { 0x1.20dd74p+0f, -0x1.812672p-2f, 0x1.ce0934p-4f,
                                 -0x1.b5a334p-6f, 0x1.4246b4p-8f, -0x1.273facp-11f };
    static const float c1c[] = { -0x1.a0d71ap+0f, -0x1.d51e3ap-1f, -0x1.3a904cp-3f, 0x1.1c395cp-5f,
                                 -0x1.6856bep-8f, 0x1.180f1ep-11f, -0x1.8ca9f6p-16f };
    float c0 = estrin(x2, c0c, 6), c1 = estrin(xa, c1c, 7);
    float xb = 1.f - std::exp2(c1 * xa);
    if (xa < 1.f)
        return x * c0;
    return std::copysign(std::isfinite(xb) ? xb : 1.f, x);
}

/* drjit::erfinv (math.h:1557-1580) */
inline float dr_erfinv(float x) {
    float w = -std::log((1.f - x) * (1.f + x));
    float w1 = w - 2.5f, w2 = std::sqrt(w) - 3.f;
    static const float p1c[] = { 1.50140941f, 0.246640727f, -0.00417768164f, -0.00125372503f, 0.00021858087f,
                                 -4.39150654e-06f, -3.5233877e-06f, 3.43273939e-07f, 2.81022636e-08f };
    static const float p2c[] = { 2.83297682f, 1.00167406f, 0.00943887047f, -0.0076224613f, 0.00573950773f,
                                 -0.00367342844f, 0.00134934322f, 0.000100950558f, -0.000200214257f };
    float p1 = estrin(w1, p1c, 9), p2 = estrin(w2, p2c, 9);
    return (w < 5.f ? p1 : p2) * x;
}

/* math::bessel_j and helpers, include/mitsuba/core/math.h:204-347 */
inline float bessel_gm(float x, int nu, float t) { return std::cos((float) nu * t - x * std::sin(t)); }

inline float bessel_j_small(float x, int nu, uint32_t points = 20) {
    /* integrate_simpson(gm, 0, pi, points), math.h:204-223 */
    const float a = 0.f, b = kPi;
    const float h = (b - a) / (float) (points - 1);
    float I_odd = 0.f, I_even = 0.f;
    for (uint32_t i = 1; i < points - 1; i += 2)
        I_odd += bessel_gm(x, nu, a + h * (float) i);
    for (uint32_t i = 2; i < points - 1; i += 2)
        I_even += bessel_gm(x, nu, a + h * (float) i);
    float integ = (h / 3.0f) * (bessel_gm(x, nu, a) + 4.0f * I_odd + 2.0f * I_even + bessel_gm(x, nu, b));
    return kInvPi * integ;
}

inline float bessel_j_asymp(float x, int nu) {
    const float sign = ((x >= 0) || (nu % 2 != 0)) ? 1.0f : -1.0f;
    const float x_abs = std::fabs(x);
    if (x_abs > 10 * kEps)
        return sign * std::sqrt(2.0f / (kPi * x_abs)) * std::cos(x_abs - kPi * (float) nu / 2.0f - kPi / 4.0f);
    return nu == 0 ? 1.0f : 0.0f;
}

inline float bessel_j(float x, int nu, uint32_t points = 20, float c = 45, float f = 1.5f) {
    const float nu_abs = (float) std::abs(nu);
    const float nsign = sign1((float) nu);
    float center = c + nu_abs;
    float weight_fn = 1.0f / (1.0f + std::exp(-f * (std::fabs(x) - center)));
    return nsign * ((1.0f - weight_fn) * bessel_j_small(x, (int) nu_abs, points) +
                    weight_fn * bessel_j_asymp(x, (int) nu_abs));
}

/* math::sinc, math.h:239-242 */
inline float sinc(float x) {
    float xx = std::fabs(x);
    return xx < 1e-5f ? 1.0f : std::sin(xx) / xx;
}

/* drjit unit_angle, ext/drjit/include/drjit/sphere.h:53-60 */
inline float unit_angle(V3 a, V3 b) {
    float dot_uv = dot(a, b);
    V3 am(mulsign(a.x, dot_uv), mulsign(a.y, dot_uv), mulsign(a.z, dot_uv));
    float temp = 2.f * std::asin(.5f * norm(b - am));
    return dot_uv >= 0 ? temp : kPi - temp;
}

/* common.py:1304-1312 */
inline float mis_weight(float pdf_a, float pdf_b) {
    float a2 = sqr(pdf_a), b2 = sqr(pdf_b);
    float w = a2 / (a2 + b2);
    return std::isfinite(w) ? w : 0.f;
}

/* ------------------------------------------------------------------------------------------------------------
 * Warps, include/mitsuba/core/warp.h
 * ---------------------------------------------------------------------------------------------------------- */
inline V2 square_to_uniform_disk_concentric(V2 sample) { /* warp.h:54-90 */
    float x = fmsub(2.f, sample.x, 1.f), y = fmsub(2.f, sample.y, 1.f);
    bool is_zero = (x == 0.f) && (y == 0.f), quadrant_1_or_3 = std::fabs(x) < std::fabs(y);
    float r = quadrant_1_or_3 ? y : x, rp = quadrant_1_or_3 ? x : y;
    float phi = 0.25f * kPi * rp / r;
    if (quadrant_1_or_3) phi = 0.5f * kPi - phi;
    if (is_zero) phi = 0.f;
    float s = std::sin(phi), c = std::cos(phi);
    return { r * c, r * s };
}
inline V3 square_to_cosine_hemisphere(V2 sample) { /* warp.h:412-420 */
    V2 p = square_to_uniform_disk_concentric(sample);
    float z = safe_sqrt(1.f - fmadd(p.y, p.y, p.x * p.x));
    return { p.x, p.y, z };
}
inline V2 square_to_uniform_triangle(V2 sample) { /* warp.h:153-156 */
    float t = safe_sqrt(1.f - sample.x);
    return { 1.f - t, t * sample.y };
}

/* ------------------------------------------------------------------------------------------------------------
 * Fresnel, include/mitsuba/render/fresnel.h
 * ---------------------------------------------------------------------------------------------------------- */
struct FresnelResult { float r, cos_theta_t, eta_it, eta_ti; };
inline FresnelResult fresnel(float cos_theta_i, float eta) { /* fresnel.h:35-91 */
    bool outside = cos_theta_i >= 0.f;
    float rcp_eta = rcp(eta), eta_it = outside ? eta : rcp_eta, eta_ti = outside ? rcp_eta : eta;
    float cos_theta_t_sqr = fnmadd(fnmadd(cos_theta_i, cos_theta_i, 1.f), eta_ti * eta_ti, 1.f);
    float cos_theta_i_abs = std::fabs(cos_theta_i), cos_theta_t_abs = safe_sqrt(cos_theta_t_sqr);
    bool index_matched = eta == 1.f, special_case = index_matched || (cos_theta_i_abs == 0.f);
    float r_sc = index_matched ? 0.f : 1.f;
    float a_s = fnmadd(eta_it, cos_theta_t_abs, cos_theta_i_abs) / fmadd(eta_it, cos_theta_t_abs, cos_theta_i_abs);
    float a_p = fnmadd(eta_it, cos_theta_i_abs, cos_theta_t_abs) / fmadd(eta_it, cos_theta_i_abs, cos_theta_t_abs);
    float r = 0.5f * (sqr(a_s) + sqr(a_p));
    if (special_case) r = r_sc;
    float cos_theta_t = mulsign_neg(cos_theta_t_abs, cos_theta_i);
    return { r, cos_theta_t, eta_it, eta_ti };
}
inline float fresnel_conductor(float cos_theta_i, float eta_r, float eta_i) { /* fresnel.h:93-124 */
    float cos_theta_i_2 = cos_theta_i * cos_theta_i, sin_theta_i_2 = 1.f - cos_theta_i_2,
          sin_theta_i_4 = sin_theta_i_2 * sin_theta_i_2;
    float temp_1 = eta_r * eta_r - eta_i * eta_i - sin_theta_i_2,
          a_2_pb_2 = safe_sqrt(temp_1 * temp_1 + 4.f * eta_i * eta_i * eta_r * eta_r),
          a = safe_sqrt(.5f * (a_2_pb_2 + temp_1));
    float term_1 = a_2_pb_2 + cos_theta_i_2, term_2 = 2.f * cos_theta_i * a;
    float r_s = (term_1 - term_2) / (term_1 + term_2);
    float term_3 = a_2_pb_2 * cos_theta_i_2 + sin_theta_i_4, term_4 = term_2 * sin_theta_i_2;
    float r_p = r_s * (term_3 - term_4) / (term_3 + term_4);
    return 0.5f * (r_s + r_p);
}
inline V3 reflect(V3 wi) { return { -wi.x, -wi.y, wi.z }; }                       /* fresnel.h:276-279 */
inline V3 reflect(V3 wi, V3 m) {                                                  /* fresnel.h:281-285 */
    float s = 2.f * dot(wi, m);
    return { fmsub(m.x, s, wi.x), fmsub(m.y, s, wi.y), fmsub(m.z, s, wi.z) };
}
inline V3 refract(V3 wi, float cos_theta_t, float eta_ti) { return { -eta_ti * wi.x, -eta_ti * wi.y, cos_theta_t }; }

/* ------------------------------------------------------------------------------------------------------------
 * MicrofacetDistribution, include/mitsuba/render/microfacet.h:100-431
 * ---------------------------------------------------------------------------------------------------------- */
struct Microfacet {
    uint32_t type;
    float au, av;
    bool sample_visible;
    Microfacet(uint32_t t, float a_u, float a_v, bool sv)
        : type(t), au(std::max(a_u, 1e-4f)), av(std::max(a_v, 1e-4f)), sample_visible(sv) {}

    float eval(V3 m) const { /* :185-206 */
        float alpha_uv = au * av, cos_theta = m.z, cos_theta_2 = sqr(cos_theta), result;
        if (type == PLTB_MICROFACET_BECKMANN)
            result = std::exp(-(sqr(m.x / au) + sqr(m.y / av)) / cos_theta_2) / (kPi * alpha_uv * sqr(cos_theta_2));
        else
            result = rcp(kPi * alpha_uv * sqr(sqr(m.x / au) + sqr(m.y / av) + sqr(m.z)));
        return result * cos_theta > 1e-20f ? result : 0.f;
    }
    float smith_g1(V3 v, V3 m) const { /* :335-366 */
        float xy_alpha_2 = sqr(au * v.x) + sqr(av * v.y), tan_theta_alpha_2 = xy_alpha_2 / sqr(v.z), result;
        if (type == PLTB_MICROFACET_BECKMANN) {
            float a = rsqrt(tan_theta_alpha_2), a_sqr = sqr(a);
            result = a >= 1.6f ? 1.f : (3.535f * a + 2.181f * a_sqr) / (1.f + 2.276f * a + 2.577f * a_sqr);
        } else {
            result = 2.f / (1.f + std::sqrt(1.f + tan_theta_alpha_2));
        }
        if (xy_alpha_2 == 0.f) result = 1.f;
        if (dot(v, m) * v.z <= 0.f) result = 0.f;
        return result;
    }
    float G(V3 wi, V3 wo, V3 m) const { return smith_g1(wi, m) * smith_g1(wo, m); }
    float pdf(V3 wi, V3 m) const { /* :208-216 */
        float result = eval(m);
        if (sample_visible)
            result *= smith_g1(wi, m) * std::fabs(dot(wi, m)) / wi.z;
        else
            result *= m.z;
        return result;
    }
    V2 sample_visible_11(float cos_theta_i, V2 sample) const { /* :368-429 */
        if (type == PLTB_MICROFACET_BECKMANN) {
            float tan_theta_i = safe_sqrt(fnmadd(cos_theta_i, cos_theta_i, 1.f)) / cos_theta_i;
            float cot_theta_i = rcp(tan_theta_i);
            float maxval = dr_erf(cot_theta_i);
            sample.x = std::max(std::min(sample.x, 1.f - 1e-6f), 1e-6f);
            sample.y = std::max(std::min(sample.y, 1.f - 1e-6f), 1e-6f);
            float x = maxval - (maxval + 1.f) * dr_erf(std::sqrt(-std::log(sample.x)));
            sample.x *= 1.f + maxval + kInvSqrtPi * tan_theta_i * std::exp(-sqr(cot_theta_i));
            for (int i = 0; i < 3; ++i) {
                float slope = dr_erfinv(x),
                      value = 1.f + x + kInvSqrtPi * tan_theta_i * std::exp(-sqr(slope)) - sample.x,
                      derivative = 1.f - slope * tan_theta_i;
                x -= value / derivative;
            }
            return { dr_erfinv(x), dr_erfinv(fmsub(2.f, sample.y, 1.f)) };
        } else {
            V2 p = square_to_uniform_disk_concentric(sample);
            float s = 0.5f * (1.f + cos_theta_i);
            /* dr::lerp(a, b, t) = fmadd(b, t, fnmadd(a, t, a)) (drjit array_router.h) */
            float la = safe_sqrt(1.f - sqr(p.x));
            p.y = fmadd(p.y, s, fnmadd(la, s, la));
            float x = p.x, y = p.y, z = safe_sqrt(1.f - fmadd(p.y, p.y, p.x * p.x));
            float sin_theta_i = safe_sqrt(1.f - sqr(cos_theta_i));
            float nrm = rcp(fmadd(sin_theta_i, y, cos_theta_i * z));
            return { fmsub(cos_theta_i, y, sin_theta_i * z) * nrm, x * nrm };
        }
    }
    /* returns m, pdf (:244-327) */
    V3 sample(V3 wi, V2 sample, float &pdf_out) const {
        if (!sample_visible) {
            float sin_phi, cos_phi, cos_theta, cos_theta_2, alpha_2;
            if (au == av) {
                float ang = (2.f * kPi) * sample.y;
                sin_phi = std::sin(ang); cos_phi = std::cos(ang);
                alpha_2 = au * au;
            } else {
                float ratio = av / au, tmp = ratio * std::tan((2.f * kPi) * sample.y);
                cos_phi = rsqrt(fmadd(tmp, tmp, 1.f));
                cos_phi = mulsign(cos_phi, std::fabs(sample.y - .5f) - .25f);
                sin_phi = cos_phi * tmp;
                alpha_2 = rcp(sqr(cos_phi / au) + sqr(sin_phi / av));
            }
            if (type == PLTB_MICROFACET_BECKMANN) {
                cos_theta = rsqrt(fnmadd(alpha_2, std::log(1.f - sample.x), 1.f));
                cos_theta_2 = sqr(cos_theta);
                float cos_theta_3 = std::max(cos_theta_2 * cos_theta, 1e-20f);
                pdf_out = (1.f - sample.x) / (kPi * au * av * cos_theta_3);
            } else {
                float tan_theta_m_2 = alpha_2 * sample.x / (1.f - sample.x);
                cos_theta = rsqrt(1.f + tan_theta_m_2);
                cos_theta_2 = sqr(cos_theta);
                float temp = 1.f + tan_theta_m_2 / alpha_2, cos_theta_3 = std::max(cos_theta_2 * cos_theta, 1e-20f);
                pdf_out = rcp(kPi * au * av * cos_theta_3 * sqr(temp));
            }
            float sin_theta = std::sqrt(1.f - cos_theta_2);
            return { cos_phi * sin_theta, sin_phi * sin_theta, cos_theta };
        }
        V3 wi_p = normalize(V3(au * wi.x, av * wi.y, wi.z));
        /* Frame3f::sincos_phi, include/mitsuba/core/frame.h:111-122 */
        float sin_theta_2 = fmadd(wi_p.x, wi_p.x, sqr(wi_p.y)), inv_sin_theta = rsqrt(sin_theta_2);
        float rx = wi_p.x * inv_sin_theta, ry = wi_p.y * inv_sin_theta;
        float cos_phi, sin_phi;
        if (std::fabs(sin_theta_2) <= 4.f * kEps) { cos_phi = 1.f; sin_phi = 0.f; }
        else { cos_phi = std::min(std::max(rx, -1.f), 1.f); sin_phi = std::min(std::max(ry, -1.f), 1.f); }
        float cos_theta = wi_p.z;
        V2 slope = sample_visible_11(cos_theta, sample);
        slope = { fmsub(cos_phi, slope.x, sin_phi * slope.y) * au, fmadd(sin_phi, slope.x, cos_phi * slope.y) * av };
        V3 m = normalize(V3(-slope.x, -slope.y, 1.f));
        pdf_out = eval(m) * smith_g1(wi, m) * std::fabs(dot(wi, m)) / wi.z;
        return m;
    }
};

/* ------------------------------------------------------------------------------------------------------------
 * DiffractionGrating, include/mitsuba/plt/diffractiongrating.h:32-290
 * ---------------------------------------------------------------------------------------------------------- */
struct Grating {
    V2 dir, inv_period;
    float q, multiplier;
    uint32_t lobes, type;
    Grating(float angle, V2 inv_p, float q_, uint32_t lobes_, uint32_t type_, float mult, V2 uv)
        : inv_period(inv_p), q(q_), multiplier(mult), lobes(lobes_), type(type_) {
        if ((type & 0x10u) == 0) {
            dir = { std::cos(angle), std::sin(angle) };
        } else {
            V2 r = { uv.x - 0.5f, uv.y - 0.5f };
            float il = rsqrt(fmadd(r.y, r.y, r.x * r.x));
            r = { r.x * il, r.y * il };
            float c = std::cos(angle), s = std::sin(angle);
            /* Matrix2f(c, s, -s, c) * Vector2f(r.x, -r.y): drjit matrix*vector = fmadd over columns
               (ext/drjit/include/drjit/matrix.h): col0 * v.x, then fmadd(col1, v.y, .) */
            float vx = r.x, vy = -r.y;
            dir = { fmadd(s, vy, c * vx), fmadd(c, vy, -s * vx) };
        }
    }
    bool is_1d() const { return inv_period.y < kEps; }
    float alpha(V3 wi, float k) const { return std::exp(-sqr(wi.z * q * k)); } /* :78-83 */

    float lobe_intensity(int lx_i, int ly_i, V3 wi, float wl) const { /* :228-272 */
        uint32_t gtype = type & 0xFu;
        const float a = kFourPi * q / (wl * std::fabs(wi.z));
        float ix, iy, lx = (float) lx_i, ly = (float) ly_i;
        switch (gtype) {
            case 0: /* sinusoidal */
                ix = lx_i == 0 ? 1.0f : sqr(bessel_j(a, lx_i));
                iy = is_1d() ? ix : (ly_i == 0 ? 1.0f : sqr(bessel_j(a, ly_i)));
                break;
            case 1: /* rectangular */
                ix = lx_i == 0 ? 1.0f : std::sin(a / 2.0f) * sinc(kPi * lx / 2.0f);
                iy = is_1d() ? ix : (ly_i == 0 ? 1.0f : std::sin(a / 2.0f) * sinc(kPi * ly / 2.0f));
                break;
            default:
                ix = lx_i == 0 ? 1.f : 1.f / std::sqrt(std::fabs(lx));
                iy = is_1d() ? ix : (ly_i == 0 ? 1.f : 1.f / std::sqrt(std::fabs(ly)));
                break;
        }
        return multiplier * ix * iy;
    }

    void sample_lobe(V2 sample2, V3 wi, float wl, int lobe[2], V2 &pdf) const { /* :105-151 */
        float intensities[PLTB_MAX_LOBES];
        float total = 0.0f;
        for (uint32_t l = 0; l < lobes / 2 + 1; ++l) {
            float li = lobe_intensity((int) l, 0, wi, wl);
            if (l == 0) li /= 2.0f;
            total += li;
            intensities[l] = li;
        }
        float cdf = 0.0f;
        pdf = { 0.f, 0.f };
        float lobe_x = 0, lobe_y = 0;
        V2 rn = { (sample2.x - 0.5f) * 2.0f, (sample2.y - 0.5f) * 2.0f };
        V2 rnd_sign = { sign1(rn.x), sign1(rn.y) };
        for (uint32_t l = 0; l < lobes / 2 + 1; ++l) {
            float p = intensities[l] / total;
            bool s1 = std::fabs(rn.x) > cdf, s2 = std::fabs(rn.y) > cdf;
            if (s1) { pdf.x = l == 0 ? p : p / 2.0f; lobe_x = (float) l; }
            if (s2) { pdf.y = l == 0 ? p : p / 2.0f; lobe_y = (float) l; }
            cdf += p;
        }
        lobe[0] = (int) (lobe_x * rnd_sign.x);
        lobe[1] = (int) (lobe_y * rnd_sign.y);
    }

    float lobe_pdf(int lx, int ly, V3 wi, float wl) const { /* :164-190 */
        float total = 0.f, ix = 0.f, iy = 0.f;
        for (uint32_t l = 0; l <= lobes / 2; ++l) {
            float li = lobe_intensity((int) l, 0, wi, wl);
            if (l == 0) li *= 0.5f;
            total += li;
            if ((uint32_t) std::abs(lx) == l) ix = li;
            if ((uint32_t) std::abs(ly) == l) iy = li;
        }
        return (ix / total) * (iy / total);
    }

    V3 diffract(V3 wi, int lx, int ly, float wl, bool &valid) const { /* :201-226 */
        float px = std::sqrt(sqr(wi.x) + sqr(wi.z)), py = std::sqrt(sqr(wi.y) + sqr(wi.z));
        float sin_ix = px > kEps ? wi.x / px : 0.0f, sin_iy = py > kEps ? wi.y / py : 0.0f;
        float cos_g = dir.x, sin_g = dir.y;
        float lrx = cos_g * (float) lx - sin_g * (float) ly, lry = sin_g * (float) lx + cos_g * (float) ly;
        float a = wl * lrx * inv_period.x - sin_ix, b = wl * lry * inv_period.y - sin_iy;
        float m = (sqr(a) - 1) / (sqr(a * b) - 1);
        float qq = 1 - sqr(b) * m;
        V3 diff(a * std::sqrt(std::max(0.0f, qq)), b * std::sqrt(m),
                std::sqrt(std::max(0.0f, 1 - sqr(a) * qq - sqr(b) * m)));
        valid = std::fabs(a) <= 1.0f && std::fabs(b) <= 1.0f;
        return diff;
    }
};

/* ------------------------------------------------------------------------------------------------------------
 * Scene access
 * ---------------------------------------------------------------------------------------------------------- */
struct BVHNode { float lo[3], hi[3]; uint32_t left, right, first, count; };

struct Scene {
    const pltb_scene_desc *d;
    std::vector<BVHNode> nodes;
    std::vector<uint32_t> prims; /* < n_faces: triangle, else analytic shape id + n_faces */
    float emitter_pmf = 0.f;
    int32_t env = -1; /* index of the environment emitter (Scene::environment(), scene.cpp:60-67), -1 if none */
    bool rgb = false; /* PLTB_RENDER_RGB: the `*_rgb` variants -- Spectrum = Color3f in lanes 0..2 of S4 (lane 3 repeats lane 2, so
                         per-lane arithmetic and max() are those of a 3-vector); SI::wavelengths then carries the path's
                         `sampling_wavelengths` (plt.py:64-70,158) instead of sensor wavelengths, which do not exist */
    V3 pos(uint32_t i) const { return { d->positions[3 * i], d->positions[3 * i + 1], d->positions[3 * i + 2] }; }
    V3 nrm(uint32_t i) const { return { d->normals[3 * i], d->normals[3 * i + 1], d->normals[3 * i + 2] }; }
    V2 uv(uint32_t i) const { return { d->texcoords[2 * i], d->texcoords[2 * i + 1] }; }
};

struct Ray { V3 o, d; float maxt; };

/* Mesh::moeller_trumbore, include/mitsuba/render/mesh.h:460-481 */
inline bool moeller_trumbore(const Ray &ray, V3 p0, V3 p1, V3 p2, float &t, float &u, float &v) {
    V3 e1 = p1 - p0, e2 = p2 - p0;
    V3 pvec = cross(ray.d, e2);
    float inv_det = rcp(dot(e1, pvec));
    V3 tvec = ray.o - p0;
    u = dot(tvec, pvec) * inv_det;
    bool active = u >= 0.f && u <= 1.f;
    V3 qvec = cross(tvec, e1);
    v = dot(ray.d, qvec) * inv_det;
    active &= v >= 0.f && u + v <= 1.f;
    t = dot(e2, qvec) * inv_det;
    active &= t >= 0.f && t <= ray.maxt;
    return active;
}

inline V3 xf_point(const float *m, V3 p) { /* Transform::transform_affine, transform.h:119-127 */
    return { fmadd(m[2], p.z, fmadd(m[1], p.y, fmadd(m[0], p.x, m[3]))),
             fmadd(m[6], p.z, fmadd(m[5], p.y, fmadd(m[4], p.x, m[7]))),
             fmadd(m[10], p.z, fmadd(m[9], p.y, fmadd(m[8], p.x, m[11]))) };
}
inline V3 xf_point_proj(const float *m, V3 p) { /* Transform::operator*(Point), transform.h:134-142 */
    float x = fmadd(m[2], p.z, fmadd(m[1], p.y, fmadd(m[0], p.x, m[3]))),
          y = fmadd(m[6], p.z, fmadd(m[5], p.y, fmadd(m[4], p.x, m[7]))),
          z = fmadd(m[10], p.z, fmadd(m[9], p.y, fmadd(m[8], p.x, m[11]))),
          w = fmadd(m[14], p.z, fmadd(m[13], p.y, fmadd(m[12], p.x, m[15])));
    return { x / w, y / w, z / w };
}
inline V3 xf_vector(const float *m, V3 v) { /* Transform::operator*(Vector), transform.h:149-158 */
    return { fmadd(m[2], v.z, fmadd(m[1], v.y, m[0] * v.x)), fmadd(m[6], v.z, fmadd(m[5], v.y, m[4] * v.x)),
             fmadd(m[10], v.z, fmadd(m[9], v.y, m[8] * v.x)) };
}

/* Analytic shapes.  Sphere::ray_intersect_preliminary_impl (src/shapes/sphere.cpp:393-436) is evaluated in
 * double precision by the reference; Rectangle::ray_intersect_preliminary_impl (src/shapes/rectangle.cpp:262-283). */
inline bool sphere_intersect(const pltb_shape &sh, const Ray &ray, float &t_out) {
    double ox = (double) ray.o.x - (double) sh.center[0], oy = (double) ray.o.y - (double) sh.center[1],
           oz = (double) ray.o.z - (double) sh.center[2];
    double dx = ray.d.x, dy = ray.d.y, dz = ray.d.z, radius = sh.radius, maxt = ray.maxt;
    double A = dx * dx + dy * dy + dz * dz, B = 2.0 * (ox * dx + oy * dy + oz * dz),
           C = ox * ox + oy * oy + oz * oz - radius * radius;
    /* math::solve_quadratic, include/mitsuba/core/math.h:440-490 */
    bool linear_case = A == 0.0;
    if (linear_case) return false; /* d != 0 always */
    double discrim = B * B - 4.0 * A * C;
    if (discrim < 0) return false;
    double temp = -0.5 * (B + std::copysign(std::sqrt(discrim), B));
    double x0p = temp / A, x1p = C / temp;
    double near_t = std::min(x0p, x1p), far_t = std::max(x0p, x1p);
    bool out_bounds = !(near_t <= maxt && far_t >= 0.0), in_bounds = near_t < 0.0 && far_t > maxt;
    if (out_bounds || in_bounds) return false;
    t_out = (float) (near_t < 0.0 ? far_t : near_t);
    return true;
}
inline bool rectangle_intersect(const pltb_shape &sh, const Ray &ray, float &t_out, V2 &uv) {
    V3 o = xf_point(sh.to_object, ray.o), d = xf_vector(sh.to_object, ray.d);
    float t = -o.z / d.z;
    V3 local = fmadd(d, t, o);
    bool active = t >= 0.f && t <= ray.maxt && std::fabs(local.x) <= 1.f && std::fabs(local.y) <= 1.f;
    if (!active) return false;
    t_out = t; uv = { local.x, local.y };
    return true;
}

void prim_bounds(const Scene &sc, uint32_t prim, float lo[3], float hi[3]) {
    const pltb_scene_desc *d = sc.d;
    if (prim < d->n_faces) {
        for (int k = 0; k < 3; ++k) { lo[k] = kInf; hi[k] = -kInf; }
        for (int c = 0; c < 3; ++c) {
            V3 p = sc.pos(d->faces[3 * prim + c]);
            float pv[3] = { p.x, p.y, p.z };
            for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], pv[k]); hi[k] = std::max(hi[k], pv[k]); }
        }
    } else {
        const pltb_shape &sh = d->shapes[prim - d->n_faces];
        if (sh.type == PLTB_SHAPE_SPHERE) {
            for (int k = 0; k < 3; ++k) { lo[k] = sh.center[k] - sh.radius; hi[k] = sh.center[k] + sh.radius; }
        } else {
            for (int k = 0; k < 3; ++k) { lo[k] = kInf; hi[k] = -kInf; }
            for (int c = 0; c < 4; ++c) {
                V3 p = xf_point(sh.to_world, V3((c & 1) ? 1.f : -1.f, (c & 2) ? 1.f : -1.f, 0.f));
                float pv[3] = { p.x, p.y, p.z };
                for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], pv[k]); hi[k] = std::max(hi[k], pv[k]); }
            }
        }
    }
}

/* A plain median-split BVH: the oracle's own accel (closest-hit results do not depend on the accel except for
 * exact ties; the reference uses a kd-tree / Embree / OptiX). */
uint32_t build_bvh(Scene &sc, std::vector<float> &lo, std::vector<float> &hi, uint32_t first, uint32_t count) {
    uint32_t idx = (uint32_t) sc.nodes.size();
    sc.nodes.push_back({});
    BVHNode n{};
    for (int k = 0; k < 3; ++k) { n.lo[k] = kInf; n.hi[k] = -kInf; }
    float clo[3] = { kInf, kInf, kInf }, chi[3] = { -kInf, -kInf, -kInf };
    for (uint32_t i = first; i < first + count; ++i) {
        uint32_t p = sc.prims[i];
        for (int k = 0; k < 3; ++k) {
            n.lo[k] = std::min(n.lo[k], lo[3 * p + k]); n.hi[k] = std::max(n.hi[k], hi[3 * p + k]);
            float c = 0.5f * (lo[3 * p + k] + hi[3 * p + k]);
            clo[k] = std::min(clo[k], c); chi[k] = std::max(chi[k], c);
        }
    }
    if (count <= 4) {
        n.first = first; n.count = count;
    } else {
        int axis = 0;
        for (int k = 1; k < 3; ++k) if (chi[k] - clo[k] > chi[axis] - clo[axis]) axis = k;
        uint32_t mid = first + count / 2;
        std::nth_element(sc.prims.begin() + first, sc.prims.begin() + mid, sc.prims.begin() + first + count,
                         [&](uint32_t a, uint32_t b) {
                             return lo[3 * a + axis] + hi[3 * a + axis] < lo[3 * b + axis] + hi[3 * b + axis];
                         });
        n.count = 0;
        n.left = build_bvh(sc, lo, hi, first, mid - first);
        n.right = build_bvh(sc, lo, hi, mid, first + count - mid);
    }
    sc.nodes[idx] = n;
    return idx;
}

struct Hit { float t = kInf; float u = 0, v = 0; uint32_t prim = 0; int32_t shape = -1; };

inline bool slab(const BVHNode &n, const Ray &r, const float inv[3], float tmax) {
    float t0 = 0.f, t1 = tmax;
    const float o[3] = { r.o.x, r.o.y, r.o.z };
    for (int k = 0; k < 3; ++k) {
        float a = (n.lo[k] - o[k]) * inv[k], b = (n.hi[k] - o[k]) * inv[k];
        if (a > b) std::swap(a, b);
        /* conservative: widen by a few ulps so the box test never rejects a hit the triangle test accepts */
        a -= std::fabs(a) * 4e-7f; b += std::fabs(b) * 4e-7f;
        if (!(a <= t1 && b >= t0)) { if (!std::isnan(a) && !std::isnan(b)) return false; }
        if (a > t0) t0 = a;
        if (b < t1) t1 = b;
    }
    return true;
}

/* Scene::ray_intersect / ray_test, src/render/scene.cpp:176-205 */
Hit intersect(const Scene &sc, const Ray &ray, bool any_hit) {
    Hit best;
    float inv[3] = { 1.f / ray.d.x, 1.f / ray.d.y, 1.f / ray.d.z };
    Ray r = ray;
    uint32_t stack[64]; int sp = 0;
    stack[sp++] = 0;
    const pltb_scene_desc *d = sc.d;
    while (sp) {
        const BVHNode &n = sc.nodes[stack[--sp]];
        if (!slab(n, r, inv, r.maxt)) continue;
        if (n.count == 0) { stack[sp++] = n.left; stack[sp++] = n.right; continue; }
        for (uint32_t i = n.first; i < n.first + n.count; ++i) {
            uint32_t p = sc.prims[i];
            float t, u = 0, v = 0; bool hit; int32_t shape;
            if (p < d->n_faces) {
                hit = moeller_trumbore(r, sc.pos(d->faces[3 * p]), sc.pos(d->faces[3 * p + 1]), sc.pos(d->faces[3 * p + 2]), t, u, v);
                shape = (int32_t) d->face_shape[p];
            } else {
                shape = (int32_t) (p - d->n_faces);
                const pltb_shape &sh = d->shapes[shape];
                V2 uv;
                if (sh.type == PLTB_SHAPE_SPHERE) hit = sphere_intersect(sh, r, t);
                else { hit = rectangle_intersect(sh, r, t, uv); u = uv.x; v = uv.y; }
            }
            if (hit && t < best.t) {
                best.t = t; best.u = u; best.v = v; best.prim = p; best.shape = shape;
                if (any_hit) return best;
                r.maxt = t;
            }
        }
    }
    return best;
}

/* SurfaceInteraction (the fields the path reads), include/mitsuba/render/interaction.h */
struct SI {
    float t = kInf;
    V3 p, n;
    V2 uv;
    Frame sh;
    V3 dp_du, dp_dv;
    V3 wi;
    S4 wavelengths;
    int32_t shape = -1;
    uint32_t prim = 0;
    bool valid() const { return t != kInf; }
    V3 to_world(V3 v) const { return sh.to_world(v); }
    V3 to_local(V3 v) const { return sh.to_local(v); }
    V3 offset_p(V3 d) const { /* interaction.h:165-170 */
        float mag = (1.f + std::max(std::max(std::fabs(p.x), std::fabs(p.y)), std::fabs(p.z))) * kRayEps;
        mag = mulsign(mag, dot(n, d));
        return fmadd(n, mag, p);   /* fmadd(mag, n, p) */
    }
    Ray spawn_ray(V3 d) const { return { offset_p(d), d, kLargest }; }
    Ray spawn_ray_to(V3 target) const { /* interaction.h:145-152 */
        V3 o = offset_p(target - p);
        V3 d = target - o;
        float dist = norm(d);
        d = d / dist;
        return { o, d, dist * (1.f - kShadowEps) };
    }
};

/* Mesh::compute_surface_interaction (src/render/mesh.cpp:1379-1551) + finalize_surface_interaction
 * (interaction.h:497-517) + initialize_sh_frame (:278-289); sphere.cpp:438-505, rectangle.cpp:297-330. */
SI make_si(const Scene &sc, const Ray &ray, const Hit &h, const S4 &wavelengths) {
    SI si;
    si.wavelengths = wavelengths;
    if (h.shape < 0) { /* miss: t = inf, wi = -ray.d */
        si.wi = -ray.d;
        return si;
    }
    const pltb_scene_desc *d = sc.d;
    const pltb_shape &sh = d->shapes[h.shape];
    si.shape = h.shape; si.prim = h.prim; si.t = h.t;
    if (sh.type == PLTB_SHAPE_MESH) {
        uint32_t f0 = d->faces[3 * h.prim], f1 = d->faces[3 * h.prim + 1], f2 = d->faces[3 * h.prim + 2];
        V3 p0 = sc.pos(f0), p1 = sc.pos(f1), p2 = sc.pos(f2);
        float b1 = h.u, b2 = h.v, b0 = 1.f - b1 - b2;
        si.p = fmadd(p0, b0, fmadd(p1, b1, p2 * b2));
        si.n = normalize(cross(p1 - p0, p2 - p0));
        si.uv = { b1, b2 };
        coordinate_system(si.n, si.dp_du, si.dp_dv);
        V3 dp0 = p1 - p0, dp1 = p2 - p0;
        if (sh.flags & PLTB_SHAPE_HAS_UV) {
            V2 uv0 = sc.uv(f0), uv1 = sc.uv(f1), uv2 = sc.uv(f2);
            si.uv = { fmadd(uv2.x, b2, fmadd(uv1.x, b1, uv0.x * b0)), fmadd(uv2.y, b2, fmadd(uv1.y, b1, uv0.y * b0)) };
            V2 duv0 = { uv1.x - uv0.x, uv1.y - uv0.y }, duv1 = { uv2.x - uv0.x, uv2.y - uv0.y };
            float det = fmsub(duv0.x, duv1.y, duv0.y * duv1.x), inv_det = rcp(det);
            if (det != 0.f) {
                si.dp_du = V3(fmsub(duv1.y, dp0.x, duv0.y * dp1.x), fmsub(duv1.y, dp0.y, duv0.y * dp1.y),
                              fmsub(duv1.y, dp0.z, duv0.y * dp1.z)) * inv_det;
                si.dp_dv = V3(fnmadd(duv1.x, dp0.x, duv0.x * dp1.x), fnmadd(duv1.x, dp0.y, duv0.x * dp1.y),
                              fnmadd(duv1.x, dp0.z, duv0.x * dp1.z)) * inv_det;
            }
        }
        if (sh.flags & PLTB_SHAPE_HAS_NORMALS) {
            V3 n0 = sc.nrm(f0), n1 = sc.nrm(f1), n2 = sc.nrm(f2);
            V3 n = fmadd(n2, b2, fmadd(n1, b1, n0 * b0));
            float il = rsqrt(sqnorm(n));
            si.sh.n = n * il;
        } else {
            si.sh.n = si.n;
        }
        if (sh.flags & PLTB_SHAPE_FLIP_NORMALS) { si.n = -si.n; si.sh.n = -si.sh.n; }
    } else if (sh.type == PLTB_SHAPE_SPHERE) {
        /* src/shapes/sphere.cpp:438-505 */
        V3 c(sh.center[0], sh.center[1], sh.center[2]);
        si.sh.n = normalize(fmadd(ray.d, h.t, ray.o) - c);
        si.p = fmadd(si.sh.n, sh.radius, c);
        V3 local = xf_point(sh.to_object, si.p);
        float rd_2 = sqr(local.x) + sqr(local.y), theta = std::acos(std::min(std::max(local.z, -1.f), 1.f) /* unit sphere in object space */),
              phi = std::atan2(local.y, local.x);
        if (phi < 0.f) phi += 2.f * kPi;
        si.uv = { phi * (0.5f * kInvPi), theta * kInvPi };
        si.dp_du = xf_vector(sh.to_world, V3(-local.y, local.x, 0.f) * (2.f * kPi));
        float rd = std::sqrt(rd_2), inv_rd = rcp(rd), cos_phi = local.x * inv_rd, sin_phi = local.y * inv_rd;
        V3 dpdv = xf_vector(sh.to_world, V3(local.z * cos_phi, local.z * sin_phi, -rd) * kPi);
        if (rd == 0.f) dpdv = V3(1.f, 0.f, 0.f);
        si.dp_dv = dpdv;
        if (sh.flags & PLTB_SHAPE_FLIP_NORMALS) si.sh.n = -si.sh.n;
        si.n = si.sh.n;
    } else {
        /* src/shapes/rectangle.cpp:297-330 */
        si.p = xf_point(sh.to_world, V3(h.u, h.v, 0.f));
        si.n = V3(sh.frame_n[0], sh.frame_n[1], sh.frame_n[2]);
        si.sh.n = si.n;
        si.dp_du = xf_vector(sh.to_world, V3(2.f, 0.f, 0.f));
        si.dp_dv = xf_vector(sh.to_world, V3(0.f, 2.f, 0.f));
        si.uv = { fmadd(h.u, .5f, .5f), fmadd(h.v, .5f, .5f) };
        if (sh.flags & PLTB_SHAPE_FLIP_NORMALS) { si.n = -si.n; si.sh.n = -si.sh.n; }
    }
    /* initialize_sh_frame */
    float nd = dot(si.sh.n, si.dp_du);
    si.sh.s = normalize(fmadd(si.sh.n, -nd, si.dp_du));
    if (si.dp_du.x == 0.f && si.dp_du.y == 0.f && si.dp_du.z == 0.f) {
        V3 tt; coordinate_system(si.sh.n, si.sh.s, tt);
    }
    si.sh.t = cross(si.sh.n, si.sh.s);
    si.wi = si.to_local(-ray.d);
    return si;
}

/* ------------------------------------------------------------------------------------------------------------
 * Textures / spectra
 * ---------------------------------------------------------------------------------------------------------- */
inline float srgb_model_eval(const float *c, float wl) { /* include/mitsuba/render/srgb.h:9-23 */
    float v = fmadd(fmadd(c[0], wl, c[1]), wl, c[2]);
    if (std::isinf(c[2])) return fmadd(sign1(c[2]), .5f, .5f);
    return std::max(0.f, fmadd(.5f * v, rsqrt(fmadd(v, v, 1.f)), .5f));
}
inline float regular_eval(const float *vals, uint32_t count, float lo, float hi, float x) {
    /* ContinuousDistribution::eval_pdf, include/mitsuba/core/distr_1d.h:364-379 */
    if (!(x >= lo && x <= hi)) return 0.f;
    float inv_interval = rcp((hi - lo) / (float) (count - 1));
    x = (x - lo) * inv_interval;
    uint32_t index = std::min((uint32_t) std::max(x, 0.f), count - 2);
    float y0 = vals[index], y1 = vals[index + 1], w1 = x - (float) index, w0 = 1.f - w1;
    return fmadd(w0, y0, w1 * y1);
}
inline float irregular_eval(const float *nodes, const float *vals, uint32_t count, float x) {
    /* IrregularContinuousDistribution::eval_pdf, distr_1d.h:705-730 */
    if (!(x >= nodes[0] && x <= nodes[count - 1])) return 0.f;
    uint32_t lo = 0, hi = count - 1; /* binary_search: first index with nodes[i] >= x ... */
    uint32_t idx = 0;
    { uint32_t a = 0, b = count; while (a < b) { uint32_t m = (a + b) / 2; if (nodes[m] < x) a = m + 1; else b = m; } idx = a; }
    (void) lo; (void) hi;
    uint32_t i0 = std::min(std::max(idx, 1u) - 1u, count - 2), i1 = i0 + 1;
    float x0 = nodes[i0], x1 = nodes[i1], y0 = vals[i0], y1 = vals[i1];
    float t = (x - x0) / (x1 - x0);
    return fmadd(y1 - y0, t, y0);
}
inline float blackbody_eval(float temperature, float wl) { /* src/spectra/blackbody.cpp:128-150 */
    const float c = 2.99792458e+8f, h = 6.62607004e-34f, k = 1.38064852e-23f;
    const float c0 = 2 * h * c * c, c1 = h * c / k;
    float lambda = wl * 1e-9f, lambda2 = lambda * lambda, lambda5 = lambda2 * lambda2 * lambda;
    float P = 1e-9f * c0 / (lambda5 * (std::exp(c1 / (lambda * temperature)) - 1.f));
    return P;
}
inline int wrap_coord(int pos, int shape, uint32_t mode) { /* ext/drjit/include/drjit/texture.h:1222-1250 */
    if (mode == PLTB_WRAP_CLAMP) return std::min(std::max(pos, 0), shape - 1);
    int shift_neg = pos < 0 ? pos + 1 : pos;
    int div = shift_neg / shape;
    int mod = pos - div * shape;
    if (mod < 0) mod += shape;
    if (mode == PLTB_WRAP_MIRROR && !(((div & 1) == 0) ^ (pos < 0))) mod = shape - 1 - mod;
    return mod;
}

S4 tex_eval(const Scene &sc, int32_t id, const S4 &wl, V2 uv) {
    S4 r;
    if (id < 0) return S4(1.f);
    const pltb_texture &t = sc.d->textures[id];
    const float *data = sc.d->tex_data + t.offset;
    switch (t.type) {
        case PLTB_TEX_UNIFORM: return S4(t.p[0]);
        case PLTB_TEX_RGB: r[0] = t.p[0]; r[1] = t.p[1]; r[2] = r[3] = t.p[2]; return r;   /* srgb.cpp:56-57, eval :77-98 in RGB mode */
        case PLTB_TEX_SRGB: for (int i = 0; i < 4; ++i) r[i] = srgb_model_eval(t.p, wl[i]); return r;
        case PLTB_TEX_D65: /* src/spectra/d65.cpp:148-162 */
            for (int i = 0; i < 4; ++i) {
                float dv = regular_eval(data, t.count, t.p[3], t.p[4], wl[i]);
                r[i] = (t.flags & 1u) ? dv * srgb_model_eval(t.p, wl[i]) : dv;
            }
            return r;
        case PLTB_TEX_REGULAR: for (int i = 0; i < 4; ++i) r[i] = regular_eval(data, t.count, t.p[3], t.p[4], wl[i]); return r;
        case PLTB_TEX_IRREGULAR: for (int i = 0; i < 4; ++i) r[i] = irregular_eval(data, data + t.count, t.count, wl[i]); return r;
        case PLTB_TEX_BLACKBODY:
            for (int i = 0; i < 4; ++i) r[i] = (wl[i] >= t.p[3] && wl[i] <= t.p[4]) ? blackbody_eval(t.p[0], wl[i]) : 0.f;
            return r;
        case PLTB_TEX_BITMAP: { /* src/textures/bitmap.cpp:700-737 (interpolate_spectral, linear filter) */
            float u = fmadd(t.to_uv[1], uv.y, fmadd(t.to_uv[0], uv.x, t.to_uv[2])),
                  v = fmadd(t.to_uv[4], uv.y, fmadd(t.to_uv[3], uv.x, t.to_uv[5]));
            float pu = fmadd(u, (float) t.width, -.5f), pv = fmadd(v, (float) t.height, -.5f);
            int iu = (int) std::floor(pu), iv = (int) std::floor(pv);
            int x0 = wrap_coord(iu, (int) t.width, t.wrap), x1 = wrap_coord(iu + 1, (int) t.width, t.wrap),
                y0 = wrap_coord(iv, (int) t.height, t.wrap), y1 = wrap_coord(iv + 1, (int) t.height, t.wrap);
            const float *v00 = data + 3 * ((size_t) y0 * t.width + x0), *v10 = data + 3 * ((size_t) y0 * t.width + x1),
                        *v01 = data + 3 * ((size_t) y1 * t.width + x0), *v11 = data + 3 * ((size_t) y1 * t.width + x1);
            float w1x = pu - (float) iu, w0x = 1.f - w1x, w1y = pv - (float) iv, w0y = 1.f - w1y;
            if (t.flags & PLTB_TEXFLAG_RGB) {   /* RGB variants: texels are linear RGB, dr::Texture::eval with FilterMode::Linear
                                                   (ext/drjit/include/drjit/texture.h:386-419): four fmadd accumulations */
                for (int c = 0; c < 3; ++c) {
                    float acc = fmadd(v00[c], w0x * w0y, 0.f);
                    acc = fmadd(v10[c], w1x * w0y, acc);
                    acc = fmadd(v01[c], w0x * w1y, acc);
                    r[c] = fmadd(v11[c], w1x * w1y, acc);
                }
                r[3] = r[2];
                return r;
            }
            for (int i = 0; i < 4; ++i) {
                float c00 = srgb_model_eval(v00, wl[i]), c10 = srgb_model_eval(v10, wl[i]),
                      c01 = srgb_model_eval(v01, wl[i]), c11 = srgb_model_eval(v11, wl[i]);
                float c0 = fmadd(w0x, c00, w1x * c10), c1 = fmadd(w0x, c01, w1x * c11);
                r[i] = fmadd(w0y, c0, w1y * c1);
            }
            return r;
        }
    }
    return r;
}

/* ------------------------------------------------------------------------------------------------------------
 * BSDFs
 * ---------------------------------------------------------------------------------------------------------- */
struct BSDFSample { V3 wo; float pdf = 0.f, eta = 0.f; uint32_t type = 0; };
struct SampleOut { BSDFSample bs; S4 weight; int lobe[2] = { 0, 0 }; bool lobe_valid = true; /* diffract()'s mask (unused by the reference) */ };

inline Grating make_grating(const pltb_bsdf &b, V2 uv) {
    return Grating(b.f[3], V2{ b.f[4], b.f[5] }, b.f[6], b.u[2], b.u[3], b.f[7], uv);
}
inline S4 conductor_F(const Scene &sc, const pltb_bsdf &b, const SI &si, float cos_theta) {
    S4 eta = tex_eval(sc, b.tex[0], si.wavelengths, si.uv), k = tex_eval(sc, b.tex[1], si.wavelengths, si.uv), F;
    for (int i = 0; i < 4; ++i) F[i] = fresnel_conductor(cos_theta, eta[i], k[i]);
    return F;
}

/* ---- principled BSDF (src/bsdfs/principled.cpp, principledhelpers.h) ------------------ */
inline float dr_lerp(float a, float b, float t) { return fmadd(b, t, fnmadd(a, t, a)); }
inline float as_float(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
inline float schlick_weight(float cos_i) { /* principledhelpers.h:108-112 */
    float m = std::min(std::max(1.0f - cos_i, 0.0f), 1.0f);
    return sqr(sqr(m)) * m;
}
inline float schlick_R0_eta(float eta) { return sqr((eta - 1.0f) / (eta + 1.0f)); }
inline float calc_schlick(float R0, float cos_theta_i, float eta) { /* principledhelpers.h:114-128 */
    bool outside = cos_theta_i >= 0.0f;
    float rcp_eta = rcp(eta), eta_it = outside ? eta : rcp_eta, eta_ti = outside ? rcp_eta : eta;
    float cos_theta_t_sqr = fnmadd(fnmadd(cos_theta_i, cos_theta_i, 1.0f), sqr(eta_ti), 1.0f);
    float cos_theta_t = safe_sqrt(cos_theta_t_sqr);
    return eta_it > 1.0f ? dr_lerp(schlick_weight(std::fabs(cos_theta_i)), 1.0f, R0) : dr_lerp(schlick_weight(cos_theta_t), 1.0f, R0);
}
inline bool mac_mic_compatibility(V3 m, V3 wi, V3 wo, float cos_theta_i, bool reflection) { /* :134-146 */
    V3 ms(mulsign(m.x, cos_theta_i), mulsign(m.y, cos_theta_i), mulsign(m.z, cos_theta_i));
    if (reflection) return dot(wi, ms) > 0.0f && dot(wo, ms) > 0.0f;
    V3 mn(mulsign_neg(m.x, cos_theta_i), mulsign_neg(m.y, cos_theta_i), mulsign_neg(m.z, cos_theta_i));
    return dot(wi, ms) > 0.0f && dot(wo, mn) > 0.0f;
}
/* mitsuba::luminance(value, wavelengths) in spectral mode: mean(cie1931_y(wl) * value) (spectrum.h:272-295,414-429) */
inline float spectral_luminance(const Scene &sc, const S4 &value, const S4 &wl) {
    if (sc.rgb) return value[0] * 0.212671f + value[1] * 0.715160f + value[2] * 0.072169f;   /* luminance(Color3f), spectrum.h:431-434 */
    const float *y = sc.d->cie_srgb + 3 * 95;
    float acc[4];
    for (int i = 0; i < 4; ++i) {
        float t = (wl[i] - 360.f) * (94.f / (830.f - 360.f));
        bool active = wl[i] >= 360.f && wl[i] <= 830.f;
        uint32_t i0 = (uint32_t) std::min(std::max((int32_t) t, 0), 93), i1 = i0 + 1;
        float w1 = t - (float) i0, w0 = 1.f - w1;
        acc[i] = (active ? fmadd(w0, y[i0], w1 * y[i1]) : 0.f) * value[i];
    }
    return (acc[0] + acc[1] + acc[2] + acc[3]) * 0.25f;
}
struct Principled {
    float eta, roughness, anisotropic, metallic, clearcoat, clearcoat_gloss, sheen, sheen_tint, flatness, spec_tint, spec_trans;
    float spec_srate, clearcoat_srate, diff_refl_srate;
    bool has_anisotropic, has_metallic, has_clearcoat, has_sheen, has_sheen_tint, has_flatness, has_spec_tint, has_spec_trans;
    explicit Principled(const pltb_bsdf &b) {
        const uint32_t h = b.u[0];
        has_anisotropic = h & PLTB_PRINCIPLED_HAS_ANISOTROPIC; has_metallic = h & PLTB_PRINCIPLED_HAS_METALLIC;
        has_clearcoat = h & PLTB_PRINCIPLED_HAS_CLEARCOAT; has_sheen = h & PLTB_PRINCIPLED_HAS_SHEEN;
        has_sheen_tint = h & PLTB_PRINCIPLED_HAS_SHEEN_TINT; has_flatness = h & PLTB_PRINCIPLED_HAS_FLATNESS;
        has_spec_tint = h & PLTB_PRINCIPLED_HAS_SPEC_TINT; has_spec_trans = h & PLTB_PRINCIPLED_HAS_SPEC_TRANS;
        eta = b.f[0]; roughness = b.f[1];
        anisotropic = has_anisotropic ? b.f[2] : 0.f; metallic = has_metallic ? b.f[3] : 0.f;
        clearcoat = has_clearcoat ? b.f[4] : 0.f; clearcoat_gloss = b.f[5];
        sheen = has_sheen ? b.f[6] : 0.f; sheen_tint = b.f[7];
        flatness = has_flatness ? b.f[8] : 0.f; spec_tint = has_spec_tint ? b.f[9] : 0.f;
        spec_trans = has_spec_trans ? b.f[10] : 0.f;
        spec_srate = as_float(b.u[1]); clearcoat_srate = as_float(b.u[2]); diff_refl_srate = as_float(b.u[3]);
    }
    Microfacet spec_distr() const { /* calc_dist_params (principledhelpers.h:214-226); GGX, sample_visible */
        float r2 = sqr(roughness);
        if (!has_anisotropic) { float a = std::max(0.001f, r2); return Microfacet(PLTB_MICROFACET_GGX, a, a, true); }
        float aspect = std::sqrt(1.0f - 0.9f * anisotropic);
        return Microfacet(PLTB_MICROFACET_GGX, std::max(0.001f, r2 / aspect), std::max(0.001f, r2 * aspect), true);
    }
    /* lobe selection probabilities, front side, no transmission (principled.cpp:376-399 == :760-786) */
    void probs(float F_spec, float &p_spec, float &p_cc, float &p_diff, bool front, float &p_trans) const {
        float brdf = (1.0f - metallic) * (1.0f - spec_trans), bsdf = has_spec_trans ? (1.0f - metallic) * spec_trans : 0.0f;
        float ps = front ? spec_srate * (1.0f - bsdf * (1.0f - F_spec)) : F_spec;
        float pt = has_spec_trans ? (front ? spec_srate * bsdf * (1.0f - F_spec) : (1.0f - F_spec)) : 0.0f;
        float pc = has_clearcoat ? (front ? 0.25f * clearcoat * clearcoat_srate : 0.0f) : 0.0f;
        float pd = front ? brdf * diff_refl_srate : 0.f;
        float rcp_tot = rcp(ps + pt + pc + pd);
        p_spec = ps * rcp_tot; p_cc = pc * rcp_tot; p_diff = pd * rcp_tot; p_trans = pt * rcp_tot;
    }
};
/* GTR1Isotropic (principledhelpers.h:17-56) */
struct GTR1 {
    float alpha;
    float eval(V3 m) const {
        float cos_theta2 = sqr(m.z), alpha2 = sqr(alpha);
        float result = (alpha2 - 1.f) / (kPi * std::log(alpha2) * (1.f + (alpha2 - 1.f) * cos_theta2));
        return result * m.z > 1e-20f ? result : 0.f;
    }
    float pdf(V3 m) const { return m.z < 0.f ? 0.f : m.z * eval(m); }
    V3 sample(V2 s) const {
        float ang = (2.f * kPi) * s.x, sin_phi = std::sin(ang), cos_phi = std::cos(ang), alpha2 = sqr(alpha);
        float cos_theta2 = (1.f - std::pow(alpha2, 1.f - s.y)) / (1.f - alpha2);
        float sin_theta = std::sqrt(std::max(0.f, 1.f - cos_theta2)), cos_theta = std::sqrt(std::max(0.f, cos_theta2));
        return V3(cos_phi * sin_theta, sin_phi * sin_theta, cos_theta);
    }
};
inline float smith_ggx1(V3 v, V3 wh, float alpha) { /* principledhelpers.h:73-93 */
    float alpha_2 = sqr(alpha), cos_theta = std::fabs(v.z), cos_theta_2 = sqr(cos_theta), tan_theta_2 = (1.0f - cos_theta_2) / cos_theta_2;
    float result = 2.0f * rcp(1.0f + std::sqrt(1.0f + alpha_2 * tan_theta_2));
    if (v.z == 1.f) result = 1.f;
    if (dot(v, wh) * v.z <= 0.f) result = 0.f;
    return result;
}

/* Principled::eval (principled.cpp:488-712) */
S4 principled_eval(const Scene &sc, const pltb_bsdf &b, const SI &si, V3 wo) {
    const Principled P(b);
    float cos_theta_i = si.wi.z;
    if (!(cos_theta_i != 0.0f)) return S4(0.f);
    S4 base_color = tex_eval(sc, b.tex[0], si.wavelengths, si.uv);
    float brdf = (1.0f - P.metallic) * (1.0f - P.spec_trans), bsdf = (1.0f - P.metallic) * P.spec_trans;
    float cos_theta_o = wo.z;
    bool reflect = cos_theta_i * cos_theta_o > 0.0f, refract_ = cos_theta_i * cos_theta_o < 0.0f;
    bool front_side = cos_theta_i > 0.0f;
    float inv_eta = rcp(P.eta), eta_path = front_side ? P.eta : inv_eta, inv_eta_path = front_side ? inv_eta : P.eta;
    Microfacet spec_dist = P.spec_distr();
    V3 wh = normalize(si.wi + wo * (reflect ? 1.0f : eta_path));
    wh = V3(mulsign(wh.x, wh.z), mulsign(wh.y, wh.z), mulsign(wh.z, wh.z));
    FresnelResult fr = fresnel(dot(si.wi, wh), P.eta);
    float F_spec_dielectric = fr.r;
    bool reflection_compat = mac_mic_compatibility(wh, si.wi, wo, cos_theta_i, true);
    bool spec_reflect_active = reflect && reflection_compat && F_spec_dielectric > 0.0f;
    bool clearcoat_active = P.has_clearcoat && P.clearcoat > 0.0f && reflect && reflection_compat && front_side;
    bool spec_trans_active = P.has_spec_trans && bsdf > 0.0f && refract_ && mac_mic_compatibility(wh, si.wi, wo, cos_theta_i, false) &&
                             F_spec_dielectric < 1.0f;   /* principled.cpp:563-565 */
    bool diffuse_active = brdf > 0.0f && reflect && front_side;
    bool sheen_active = P.has_sheen && P.sheen > 0.0f && reflect && (1.0f - P.metallic > 0.0f) && front_side;
    float D = spec_dist.eval(wh), G = spec_dist.G(si.wi, wo, wh);
    S4 value(0.f);
    if (spec_reflect_active) {
        float lum = P.has_spec_tint ? spectral_luminance(sc, base_color, si.wavelengths) : 1.0f;
        /* principled_fresnel (principledhelpers.h:148-186) */
        float cti = dot(si.wi, wh);
        bool outside = cti >= 0.0f;
        float eta_it = outside ? P.eta : rcp(P.eta);
        S4 F_schlick(0.f);
        if (P.has_metallic)
            for (int i = 0; i < 4; ++i) F_schlick[i] += P.metallic * calc_schlick(base_color[i], cti, P.eta);
        if (P.has_spec_tint)
            for (int i = 0; i < 4; ++i) {
                float c_tint = lum > 0.0f ? base_color[i] / lum : 1.0f;
                float F0 = c_tint * schlick_R0_eta(eta_it);
                F_schlick[i] += (1.0f - P.metallic) * P.spec_tint * calc_schlick(F0, cti, P.eta);
            }
        for (int i = 0; i < 4; ++i) {
            float F_front = (1.0f - P.metallic) * (1.0f - P.spec_tint) * F_spec_dielectric + F_schlick[i];
            float Fp = front_side ? F_front : bsdf * F_spec_dielectric;
            value[i] += Fp * D * G / (4.0f * std::fabs(cos_theta_i));
        }
    }
    if (spec_trans_active) { /* main specular transmission, principled.cpp:604-621; ctx.mode == Radiance: scale = inv_eta_path^2 */
        float scale = sqr(inv_eta_path), dwi = dot(si.wi, wh), dwo = dot(wo, wh);
        float t = std::fabs((scale * (1.0f - F_spec_dielectric) * D * G * eta_path * eta_path * dwi * dwo) /
                            (cos_theta_i * sqr(dwi + eta_path * dwo)));
        for (int i = 0; i < 4; ++i) value[i] += std::sqrt(base_color[i]) * bsdf * t;
    }
    if (clearcoat_active) {
        float Fcc = calc_schlick(0.04f, dot(si.wi, wh), P.eta);
        GTR1 mf{ dr_lerp(0.1f, 0.001f, P.clearcoat_gloss) };
        float Dcc = mf.eval(wh);
        float G_cc = smith_ggx1(si.wi, wh, 0.25f) * smith_ggx1(wo, wh, 0.25f);
        float add = (P.clearcoat * 0.25f) * Fcc * Dcc * G_cc * std::fabs(cos_theta_o);
        for (int i = 0; i < 4; ++i) value[i] += add;
    }
    if (diffuse_active) {
        float Fo = schlick_weight(std::fabs(cos_theta_o)), Fi = schlick_weight(std::fabs(cos_theta_i));
        float f_diff = (1.0f - 0.5f * Fi) * (1.0f - 0.5f * Fo);
        float cos_theta_d = dot(wh, wo);
        float Rr = 2.0f * P.roughness * sqr(cos_theta_d);
        float f_retro = Rr * (Fo + Fi + Fo * Fi * (Rr - 1.0f));
        float lobe;
        if (P.has_flatness) {
            float Fss90 = Rr / 2.0f;
            float Fss = dr_lerp(1.0f, Fss90, Fo) * dr_lerp(1.0f, Fss90, Fi);
            float f_ss = 1.25f * (Fss * (1.0f / (std::fabs(cos_theta_o) + std::fabs(cos_theta_i)) - 0.5f) + 0.5f);
            lobe = dr_lerp(f_diff + f_retro, f_ss, P.flatness);
        } else lobe = f_diff + f_retro;
        for (int i = 0; i < 4; ++i) value[i] += brdf * std::fabs(cos_theta_o) * base_color[i] * kInvPi * lobe;
        if (sheen_active) {
            float Fd = schlick_weight(std::fabs(cos_theta_d));
            if (P.has_sheen_tint) {
                float lum = spectral_luminance(sc, base_color, si.wavelengths);
                for (int i = 0; i < 4; ++i) {
                    float c_tint = lum > 0.0f ? base_color[i] / lum : 1.0f;
                    float c_sheen = dr_lerp(1.0f, c_tint, P.sheen_tint);
                    value[i] += P.sheen * (1.0f - P.metallic) * Fd * c_sheen * std::fabs(cos_theta_o);
                }
            } else {
                float add = P.sheen * (1.0f - P.metallic) * Fd * std::fabs(cos_theta_o);
                for (int i = 0; i < 4; ++i) value[i] += add;
            }
        }
    }
    return value;
}

/* Principled::pdf (principled.cpp:714-834) */
float principled_pdf(const pltb_bsdf &b, const SI &si, V3 wo) {
    const Principled P(b);
    float cos_theta_i = si.wi.z;
    if (!(cos_theta_i != 0.0f)) return 0.0f;
    bool front_side = cos_theta_i > 0.0f;
    float eta_path = front_side ? P.eta : rcp(P.eta), cos_theta_o = wo.z;
    bool reflect = cos_theta_i * cos_theta_o > 0.0f, refract_ = cos_theta_i * cos_theta_o < 0.0f;
    V3 wh = normalize(si.wi + wo * (reflect ? 1.0f : eta_path));
    wh = V3(mulsign(wh.x, wh.z), mulsign(wh.y, wh.z), mulsign(wh.z, wh.z));
    Microfacet spec_distr = P.spec_distr();
    FresnelResult fr = fresnel(dot(si.wi, wh), P.eta);
    float p_spec, p_cc, p_diff, p_trans;
    P.probs(fr.r, p_spec, p_cc, p_diff, front_side, p_trans);
    float dwh_dwo_abs; /* principled.cpp:789-801 */
    if (P.has_spec_trans) {
        float dot_wi_h = dot(si.wi, wh), dot_wo_h = dot(wo, wh);
        dwh_dwo_abs = std::fabs(reflect ? rcp(4.0f * dot_wo_h) : (sqr(eta_path) * dot_wo_h) / sqr(dot_wi_h + eta_path * dot_wo_h));
    } else {
        dwh_dwo_abs = std::fabs(rcp(4.0f * dot(wo, wh)));
    }
    float pdf = 0.0f;
    bool macmic = mac_mic_compatibility(wh, si.wi, wo, cos_theta_i, true) && reflect;
    V3 wi_s(mulsign(si.wi.x, cos_theta_i), mulsign(si.wi.y, cos_theta_i), mulsign(si.wi.z, cos_theta_i));
    if (macmic) pdf += p_spec * spec_distr.pdf(wi_s, wh) * dwh_dwo_abs;
    if (reflect) pdf += p_diff * (kInvPi * wo.z);
    if (P.has_spec_trans && mac_mic_compatibility(wh, si.wi, wo, cos_theta_i, false) && refract_) /* :816-828 */
        pdf += p_trans * spec_distr.pdf(wi_s, wh) * dwh_dwo_abs;
    if (P.has_clearcoat) {
        GTR1 cc{ dr_lerp(0.1f, 0.001f, P.clearcoat_gloss) };
        if (macmic) pdf += p_cc * cc.pdf(wh) * dwh_dwo_abs;
    }
    return pdf;
}

/* `si` already has the twosided flip applied (wi.z = |wi.z|). */
SampleOut bsdf_sample_inner(const Scene &sc, const pltb_bsdf &b, const SI &si, float s1, V2 s2, V2 lobe_s2) {
    SampleOut o;
    float cos_theta_i = si.wi.z;
    switch (b.type) {
        case PLTB_BSDF_DIFFUSE: { /* src/bsdfs/diffuse.cpp:101-125 */
            if (!(cos_theta_i > 0.f)) return o;
            o.bs.wo = square_to_cosine_hemisphere(s2);
            o.bs.pdf = kInvPi * o.bs.wo.z;
            o.bs.eta = 1.f; o.bs.type = F_DiffuseReflection;
            if (o.bs.pdf > 0.f) o.weight = tex_eval(sc, b.tex[0], si.wavelengths, si.uv);
            return o;
        }
        case PLTB_BSDF_CONDUCTOR: { /* src/bsdfs/conductor.cpp:248-308 */
            if (!(cos_theta_i > 0.f)) return o;
            o.bs.type = F_DeltaReflection; o.bs.wo = reflect(si.wi); o.bs.eta = 1.f; o.bs.pdf = 1.f;
            o.weight = tex_eval(sc, b.tex[2], si.wavelengths, si.uv) * conductor_F(sc, b, si, cos_theta_i);
            return o;
        }
        case PLTB_BSDF_DIELECTRIC: { /* src/bsdfs/dielectric.cpp:265-400 */
            FresnelResult f = fresnel(cos_theta_i, b.f[0]);
            float t_i = 1.f - f.r;
            bool selected_r = s1 <= f.r;
            o.bs.pdf = selected_r ? f.r : t_i;
            o.bs.type = selected_r ? F_DeltaReflection : F_DeltaTransmission;
            o.bs.wo = selected_r ? reflect(si.wi) : refract(si.wi, f.cos_theta_t, f.eta_ti);
            o.bs.eta = selected_r ? 1.f : f.eta_it;
            S4 w(1.f);
            if (selected_r) { if (b.tex[2] >= 0) w = w * tex_eval(sc, b.tex[2], si.wavelengths, si.uv); }
            else {
                if (b.tex[3] >= 0) w = w * tex_eval(sc, b.tex[3], si.wavelengths, si.uv);
                w = w * sqr(f.eta_ti);
            }
            o.weight = w;
            return o;
        }
        case PLTB_BSDF_PRINCIPLED: { /* src/bsdfs/principled.cpp:330-486 */
            if (!(cos_theta_i != 0.0f)) return o;
            const Principled P(b);
            bool front_side = cos_theta_i > 0.0f;
            Microfacet spec_distr = P.spec_distr();
            V3 wi_s(mulsign(si.wi.x, cos_theta_i), mulsign(si.wi.y, cos_theta_i), mulsign(si.wi.z, cos_theta_i));
            float mpdf; V3 m_spec = spec_distr.sample(wi_s, s2, mpdf);
            FresnelResult fr = fresnel(dot(si.wi, m_spec), P.eta);
            float bsdf = P.has_spec_trans ? (1.0f - P.metallic) * P.spec_trans : 0.0f;
            bool active = front_side || bsdf > 0.0f;   /* principled.cpp:370-372 */
            float p_spec, p_cc, p_diff, p_trans;
            P.probs(fr.r, p_spec, p_cc, p_diff, front_side, p_trans);
            float curr = 0.0f;
            bool sample_diffuse = active && s1 < p_diff;
            curr += p_diff;
            bool sample_clearcoat = P.has_clearcoat && active && s1 >= curr && s1 < curr + p_cc;
            curr += p_cc;
            bool sample_spec_trans = P.has_spec_trans && active && s1 >= curr && s1 < curr + p_trans;
            curr += p_trans;
            bool sample_spec_reflect = active && s1 >= curr;
            o.bs.eta = 1.0f;
            if (sample_spec_reflect) {
                V3 wo = reflect(si.wi, m_spec);
                o.bs.wo = wo; o.bs.type = F_GlossyReflection;
                active = active && mac_mic_compatibility(m_spec, si.wi, wo, cos_theta_i, true) && cos_theta_i * wo.z > 0.0f;
            }
            if (sample_spec_trans) { /* :437-453; refract(wi, m, cos_theta_t, eta_ti), fresnel.h:310-314 */
                float c = fmadd(dot(si.wi, m_spec), fr.eta_ti, fr.cos_theta_t);
                V3 wo(fmsub(m_spec.x, c, si.wi.x * fr.eta_ti), fmsub(m_spec.y, c, si.wi.y * fr.eta_ti), fmsub(m_spec.z, c, si.wi.z * fr.eta_ti));
                o.bs.wo = wo; o.bs.type = F_GlossyTransmission; o.bs.eta = fr.eta_it;
                active = active && mac_mic_compatibility(m_spec, si.wi, wo, cos_theta_i, false) && cos_theta_i * wo.z < 0.0f;
            }
            if (sample_clearcoat) {
                GTR1 cc{ dr_lerp(0.1f, 0.001f, P.clearcoat_gloss) };
                V3 m_cc = cc.sample(s2);
                V3 wo = reflect(si.wi, m_cc);
                o.bs.wo = wo; o.bs.type = F_GlossyReflection;
                active = active && mac_mic_compatibility(m_cc, si.wi, wo, cos_theta_i, true) && cos_theta_i * wo.z > 0.0f;
            }
            if (sample_diffuse) {
                V3 wo = square_to_cosine_hemisphere(s2);
                o.bs.wo = wo; o.bs.type = F_DiffuseReflection;
                active = active && cos_theta_i * wo.z > 0.0f;
            }
            o.bs.pdf = active ? principled_pdf(b, si, o.bs.wo) : 0.0f;
            active = active && o.bs.pdf > 0.0f;
            if (active) o.weight = principled_eval(sc, b, si, o.bs.wo) / o.bs.pdf;
            return o;
        }
        case PLTB_BSDF_ROUGHCONDUCTOR: { /* src/bsdfs/roughconductor.cpp:225-311 */
            if (!(cos_theta_i > 0.f)) return o;
            Microfacet distr(b.u[0], b.f[1], b.f[2], b.u[1] != 0);
            float pdf; V3 m = distr.sample(si.wi, s2, pdf);
            o.bs.pdf = pdf; o.bs.wo = reflect(si.wi, m); o.bs.eta = 1.f; o.bs.type = F_GlossyReflection;
            bool active = (o.bs.pdf != 0.f) && o.bs.wo.z > 0.f;
            float weight = distr.sample_visible ? distr.smith_g1(o.bs.wo, m)
                                                : distr.G(si.wi, o.bs.wo, m) * dot(si.wi, m) / (cos_theta_i * m.z);
            o.bs.pdf /= 4.f * dot(o.bs.wo, m);
            S4 F = conductor_F(sc, b, si, dot(si.wi, m));
            S4 w(weight);
            if (b.tex[2] >= 0) w = w * tex_eval(sc, b.tex[2], si.wavelengths, si.uv);
            if (active) o.weight = F * w;
            return o;
        }
        case PLTB_BSDF_ROUGHGRATING: { /* src/bsdfs/roughgrating.cpp:449-595 + sample_diffract :414-447 */
            if (!(cos_theta_i > 0.f)) return o;
            Microfacet distr(b.u[0], b.f[1], b.f[2], b.u[1] != 0);
            float pdf; V3 m = distr.sample(si.wi, s2, pdf);
            o.bs.pdf = pdf;
            V3 reflection_dir = reflect(si.wi, m);
            /* spectral: the hero wavelength; RGB variants: "Use sample 1 for random wavelength sample" (roughgrating.cpp:499-506) */
            float wavelength = sc.rgb ? s1 * (830.f - 150.0f - 360.f) + 360.f : si.wavelengths[0];
            /* sample_diffract */
            Frame nf(m);
            V3 wi_local = nf.to_local(si.wi);
            Grating g = make_grating(b, si.uv);
            V2 pdf_xy;
            g.sample_lobe(lobe_s2, wi_local, wavelength * 1e-3f, o.lobe, pdf_xy);
            float intensity = g.lobe_intensity(o.lobe[0], o.lobe[1], wi_local, wavelength); /* nm, as written (N2) */
            bool valid; V3 wo_local = g.diffract(wi_local, o.lobe[0], o.lobe[1], wavelength * 1e-3f, valid);
            o.bs.wo = nf.to_world(wo_local);
            o.lobe_valid = valid;   /* returned as active_diffracted and never read (roughgrating.cpp:514-521); kept for the trace */
            float grating_pdf = pdf_xy.x * pdf_xy.y;
            o.bs.eta = 1.f; o.bs.type = F_GlossyReflection;
            bool active = (o.bs.pdf != 0.f) && o.bs.wo.z > 0.f;
            o.bs.pdf *= grating_pdf;
            float weight = distr.sample_visible ? distr.smith_g1(reflection_dir, m)
                                                : distr.G(si.wi, reflection_dir, m) * dot(si.wi, m) / (cos_theta_i * m.z);
            o.bs.pdf /= 4.f * dot(reflection_dir, m);
            S4 F = conductor_F(sc, b, si, dot(si.wi, m));
            S4 w(weight);
            if (b.tex[2] >= 0) w = w * tex_eval(sc, b.tex[2], si.wavelengths, si.uv);
            if (active) o.weight = ((F * w) * intensity) * b.f[7];
            return o;
        }
        default: return o;
    }
}

/* eval_pdf of the conventional rough-conductor lobe; used by roughconductor (src/bsdfs/roughconductor.cpp:432-516)
 * and by RoughGrating::eval_pdf (roughgrating.cpp:1036-1127; there specular_reflectance is NOT applied). */
void rough_eval_pdf(const Scene &sc, const pltb_bsdf &b, const SI &si, V3 wo, bool apply_refl, S4 &value, float &pdf) {
    value = S4(0.f); pdf = 0.f;
    float cos_theta_i = si.wi.z, cos_theta_o = wo.z;
    V3 H = normalize(wo + si.wi);
    bool active = cos_theta_i > 0.f && cos_theta_o > 0.f && dot(si.wi, H) > 0.f && dot(wo, H) > 0.f;
    if (!active) return;
    Microfacet distr(b.u[0], b.f[1], b.f[2], b.u[1] != 0);
    float D = distr.eval(H);
    active &= D != 0.f;
    float smith_g1_wi = distr.smith_g1(si.wi, H), G = smith_g1_wi * distr.smith_g1(wo, H);
    float val = D * G / (4.f * cos_theta_i);
    S4 F = conductor_F(sc, b, si, dot(si.wi, H));
    S4 v(val);
    if (apply_refl && b.tex[2] >= 0) v = v * tex_eval(sc, b.tex[2], si.wavelengths, si.uv);
    float p = distr.sample_visible ? D * smith_g1_wi / (4.f * cos_theta_i) : distr.pdf(si.wi, H) / (4.f * dot(wo, H));
    if (active) { value = F * v; pdf = p; }
}

/* Coherence(diffusivity, opl).inv_coherence_det(k) (include/mitsuba/plt/plt.h:27-59): det((k / (2 pi opl 1e3)) * diag(diffusivity)) */
inline float inv_coherence_det(float k, float diffusivity, float opl) {
    float rmm = opl * 1e+3f, s = k / (kTwoPi * rmm);
    float d00 = s * diffusivity, d11 = s * diffusivity, d01 = s * 0.f, d10 = s * 0.f;
    return fmsub(d00, d11, d01 * d10);
}

/* The exhaustive lobe search of RoughGrating::wbsdf_eval (roughgrating.cpp:845-896): sum over (lx, ly, wavelength) of the lobe
 * intensities whose window contains wo, before the Fresnel factor (a Spectrum in unpolarized, a Mueller matrix in polarized
 * variants) and specular_reflectance are applied. */
S4 grating_lobe_sum(const Scene &sc, const pltb_bsdf &b, const SI &si, V3 wo) {
    (void) sc;
    Grating g = make_grating(b, si.uv);
    V3 reflection_dir = reflect(si.wi);
    S4 result(0.f);
    float a = 2 * std::sqrt(b.f[1] * b.f[2]); /* texture eval_1 of alpha_u/alpha_v: unclamped */
    int half = (int) b.u[2] / 2;
    for (int lx = -half; lx < half + 1; lx++) {
        for (int ly = -half; ly < half + 1; ly++) {
            for (int i = 0; i < 4; i++) {
                float wl = si.wavelengths[i];
                float k = kTwoPi / (wl * 1e-3f);
                float lobe_intensity = g.lobe_intensity(lx, ly, si.wi, wl * 1e-3f);
                bool lobe_active; V3 center_dir = g.diffract(si.wi, lx, ly, wl * 1e-3f, lobe_active);
                if (lx == 0 && ly == 0) {
                    result[i] += (lobe_active && std::fabs(unit_angle(center_dir, wo)) < a) ? lobe_intensity : 0.0f;
                } else {
                    float det = inv_coherence_det(k, b.f[8], 1.0f);   /* Coherence3f coh(m_coherence, 1.0f), roughgrating.cpp:879 */
                    float angle_offset = std::fabs(unit_angle(reflection_dir, wo));
                    float inv_det = std::isnan(det) ? 0.0f : det;
                    float angular_coh = std::exp(-0.5f * angle_offset * inv_det * angle_offset);
                    if (std::isnan(angular_coh)) angular_coh = 0.0f;
                    result[i] += (lobe_active && std::fabs(unit_angle(center_dir, wo)) < a) ? lobe_intensity * angular_coh : 0.0f;
                }
            }
        }
    }
    return result;
}

/* cie1931_xyz (include/mitsuba/core/spectrum.h:236-264) followed by xyz_to_srgb (:406-412; Matrix * vector = fmadd chain over the
 * columns, ext/drjit/include/drjit/matrix.h:142-150).  cie = the 6 x 95 table of pltb_scene_desc: rows 4, 3, 5 are X, Y, Z. */
inline void wavelength_to_srgb(const float *cie, float wl, float rgb[3]) {
    float t = (wl - 360.f) * (94 / (830.f - 360.f));
    bool active = wl >= 360.f && wl <= 830.f;
    uint32_t i0 = std::min((uint32_t) std::max(t, 0.f), 93u), i1 = i0 + 1;   /* clip(UInt32(t), 0, 93) */
    float w1 = t - (float) i0, w0 = 1.f - w1;
    const float *X = cie + 4 * 95, *Y = cie + 3 * 95, *Z = cie + 5 * 95;
    float x = active ? fmadd(w0, X[i0], w1 * X[i1]) : 0.f, y = active ? fmadd(w0, Y[i0], w1 * Y[i1]) : 0.f,
          z = active ? fmadd(w0, Z[i0], w1 * Z[i1]) : 0.f;
    rgb[0] = fmadd(-0.498535f, z, fmadd(-1.537150f, y, 3.240479f * x));
    rgb[1] = fmadd(0.041556f, z, fmadd(1.875991f, y, -0.969256f * x));
    rgb[2] = fmadd(1.057311f, z, fmadd(-0.204043f, y, 0.055648f * x));
}

/* The RGB fallback of RoughGrating::wbsdf_eval (roughgrating.cpp:716-790), before the Fresnel factor and specular_reflectance:
 * exhaustive search over the lobes at the three `sampling_wavelengths` of the path (si.wavelengths here), each lobe coloured with
 * xyz_to_srgb(cie1931_xyz(wl)); every lobe (the (0, 0) one too) carries the angular coherence factor. */
S4 grating_lobe_sum_rgb(const Scene &sc, const pltb_bsdf &b, const SI &si, V3 wo) {
    Grating g = make_grating(b, si.uv);
    V3 reflection_dir = reflect(si.wi);
    float result[3] = { 0.f, 0.f, 0.f };
    float a = 2 * std::sqrt(b.f[1] * b.f[2]);
    int half = (int) b.u[2] / 2;
    for (int lx = -half; lx < half + 1; lx++) {
        for (int ly = -half; ly < half + 1; ly++) {
            for (int i = 0; i < 3; i++) {
                float wl = si.wavelengths[i];
                float k = kTwoPi / (wl * 1e-3f);
                float lobe_intensity = g.lobe_intensity(lx, ly, si.wi, wl * 1e-3f);
                bool lobe_active; V3 center_dir = g.diffract(si.wi, lx, ly, wl * 1e-3f, lobe_active);
                float colour[3]; wavelength_to_srgb(sc.d->cie_srgb, wl, colour);
                float lobes_result = (lobe_active && std::fabs(unit_angle(center_dir, wo)) < a) ? lobe_intensity : 0.0f;
                float det = inv_coherence_det(k, b.f[8], 1.0f);
                float angle_offset = std::fabs(unit_angle(reflection_dir, wo));
                float inv_det = std::isnan(det) ? 0.0f : det;
                float angular_coh = std::exp(-0.5f * angle_offset * inv_det * angle_offset);
                if (std::isnan(angular_coh)) angular_coh = 0.0f;
                for (int c = 0; c < 3; ++c) result[c] += (lobes_result * colour[c]) * angular_coh;   /* lobes_result * colour * angular_coh */
            }
        }
    }
    S4 r; r[0] = result[0]; r[1] = result[1]; r[2] = r[3] = result[2];
    return r;
}
inline S4 grating_lobe_sum_any(const Scene &sc, const pltb_bsdf &b, const SI &si, V3 wo) {
    return sc.rgb ? grating_lobe_sum_rgb(sc, b, si, wo) : grating_lobe_sum(sc, b, si, wo);
}

/* wbsdf_eval: base -> eval (src/render/bsdf.cpp:59-71) */
S4 bsdf_eval_inner(const Scene &sc, const pltb_bsdf &b, const SI &si, V3 wo) {
    float cos_theta_i = si.wi.z, cos_theta_o = wo.z;
    switch (b.type) {
        case PLTB_BSDF_DIFFUSE: { /* diffuse.cpp:127-142 */
            if (!(cos_theta_i > 0.f && cos_theta_o > 0.f)) return S4(0.f);
            return (tex_eval(sc, b.tex[0], si.wavelengths, si.uv) * kInvPi) * cos_theta_o;
        }
        case PLTB_BSDF_PRINCIPLED: return principled_eval(sc, b, si, wo);
        case PLTB_BSDF_ROUGHCONDUCTOR: { /* roughconductor.cpp:313-390 */
            if (!(cos_theta_i > 0.f && cos_theta_o > 0.f)) return S4(0.f);
            V3 H = normalize(wo + si.wi);
            Microfacet distr(b.u[0], b.f[1], b.f[2], b.u[1] != 0);
            float D = distr.eval(H);
            bool active = D != 0.f;
            float G = distr.G(si.wi, wo, H);
            S4 result(D * G / (4.f * cos_theta_i));
            S4 F = conductor_F(sc, b, si, dot(si.wi, H));
            if (b.tex[2] >= 0) result = result * tex_eval(sc, b.tex[2], si.wavelengths, si.uv);
            return active ? F * result : S4(0.f);
        }
        case PLTB_BSDF_ROUGHGRATING: { /* roughgrating.cpp:676-970, spectral branch :841-969 (JIT semantics: the
                                          cos_theta tests do not mask the returned value) */
            V3 H = normalize(wo + si.wi);
            S4 result = grating_lobe_sum_any(sc, b, si, wo);   /* RGB variants: the fallback of :716-790 */
            S4 F = conductor_F(sc, b, si, dot(si.wi, H));
            if (b.tex[2] >= 0) result = result * tex_eval(sc, b.tex[2], si.wavelengths, si.uv);
            return F * result;
        }
        default: return S4(0.f); /* conductor / dielectric: eval == 0 */
    }
}

/* wbsdf_pdf: base -> pdf (bsdf.cpp:73-83); RoughGrating::wbsdf_pdf (roughgrating.cpp:1009-1034) */
float bsdf_pdf_inner(const Scene &sc, const pltb_bsdf &b, const SI &si, V3 wo) {
    float cos_theta_i = si.wi.z, cos_theta_o = wo.z;
    switch (b.type) {
        case PLTB_BSDF_DIFFUSE: return (cos_theta_i > 0.f && cos_theta_o > 0.f) ? kInvPi * wo.z : 0.f;
        case PLTB_BSDF_PRINCIPLED: return principled_pdf(b, si, wo);
        case PLTB_BSDF_ROUGHCONDUCTOR: { /* roughconductor.cpp:392-430 */
            V3 m = normalize(wo + si.wi);
            bool active = cos_theta_i > 0.f && cos_theta_o > 0.f && dot(si.wi, m) > 0.f && dot(wo, m) > 0.f;
            if (!active) return 0.f;
            Microfacet distr(b.u[0], b.f[1], b.f[2], b.u[1] != 0);
            return distr.sample_visible ? distr.eval(m) * distr.smith_g1(si.wi, m) / (4.f * cos_theta_i)
                                        : distr.pdf(si.wi, m) / (4.f * dot(wo, m));
        }
        case PLTB_BSDF_ROUGHGRATING: {
            Grating g = make_grating(b, si.uv);
            float k = kTwoPi / (si.wavelengths[0] * 1e-3f);   /* RGB variants: sd.sampling_wavelengths[0] (:1027-1031) == si.wavelengths[0] here */
            return g.alpha(si.wi, k);
        }
        default: (void) sc; return 0.f;
    }
}

/* wbsdf_weight: base (bsdf.cpp:85-97) and overrides (diffuse.cpp:182-200, conductor.cpp:320-376,
 * dielectric.cpp:404-590) */
S4 bsdf_weight_inner(const Scene &sc, const pltb_bsdf &b, const SI &si, V3 wo) {
    switch (b.type) {
        case PLTB_BSDF_DIFFUSE: return tex_eval(sc, b.tex[0], si.wavelengths, si.uv);
        case PLTB_BSDF_CONDUCTOR: {
            if (!(si.wi.z > 0.f)) return S4(0.f);
            return tex_eval(sc, b.tex[2], si.wavelengths, si.uv) * conductor_F(sc, b, si, si.wi.z);
        }
        case PLTB_BSDF_DIELECTRIC: {
            bool selected_r = wo.z > 0.0f;
            S4 w(1.f);
            if (selected_r) { if (b.tex[2] >= 0) w = w * tex_eval(sc, b.tex[2], si.wavelengths, si.uv); }
            else { if (b.tex[3] >= 0) w = w * tex_eval(sc, b.tex[3], si.wavelengths, si.uv); }
            return w;
        }
        case PLTB_BSDF_PRINCIPLED: { /* base wbsdf_weight: eval / pdf (bsdf.cpp:85-97) */
            float p = principled_pdf(b, si, wo);
            return p > 0.0f ? principled_eval(sc, b, si, wo) / p : S4(0.f);
        }
        case PLTB_BSDF_ROUGHCONDUCTOR:
        case PLTB_BSDF_ROUGHGRATING: {
            S4 e; float p;
            rough_eval_pdf(sc, b, si, wo, b.type == PLTB_BSDF_ROUGHCONDUCTOR, e, p);
            return p > 0.0f ? e / p : S4(0.f);
        }
        default: return S4(0.f);
    }
}

/* TwoSidedBRDF wrappers with m_brdf[0] == m_brdf[1] (src/bsdfs/twosided.cpp:262-420) */
inline const pltb_bsdf *si_bsdf(const Scene &sc, const SI &si) {
    if (si.shape < 0) return nullptr;
    int32_t id = sc.d->shapes[si.shape].bsdf;
    return id < 0 ? nullptr : &sc.d->bsdfs[id];
}
SampleOut wbsdf_sample(const Scene &sc, const SI &si_, float s1, V2 s2, V2 lobe_s2) {
    const pltb_bsdf *b = si_bsdf(sc, si_);
    if (!b) return SampleOut();
    if (b->flags & PLTB_BSDF_FLAG_TWOSIDED) {
        SI si = si_;
        si.wi.z = std::fabs(si.wi.z);
        SampleOut o = bsdf_sample_inner(sc, *b, si, s1, s2, lobe_s2);
        o.bs.wo.z = mulsign(o.bs.wo.z, si_.wi.z);
        return o;
    }
    return bsdf_sample_inner(sc, *b, si_, s1, s2, lobe_s2);
}
template <typename R, typename Fn> R twosided_call(const Scene &sc, const SI &si_, V3 wo, Fn fn, R zero) {
    const pltb_bsdf *b = si_bsdf(sc, si_);
    if (!b) return zero;
    if (b->flags & PLTB_BSDF_FLAG_TWOSIDED) {
        SI si = si_;
        wo.z = mulsign(wo.z, si.wi.z);
        si.wi.z = std::fabs(si.wi.z);
        return fn(sc, *b, si, wo);
    }
    return fn(sc, *b, si_, wo);
}
S4 wbsdf_eval(const Scene &sc, const SI &si, V3 wo) { return twosided_call<S4>(sc, si, wo, bsdf_eval_inner, S4(0.f)); }
float wbsdf_pdf(const Scene &sc, const SI &si, V3 wo) { return twosided_call<float>(sc, si, wo, bsdf_pdf_inner, 0.f); }
S4 wbsdf_weight(const Scene &sc, const SI &si, V3 wo) { return twosided_call<S4>(sc, si, wo, bsdf_weight_inner, S4(0.f)); }

/* ------------------------------------------------------------------------------------------------------------
 * Emitters
 * ---------------------------------------------------------------------------------------------------------- */
/* ------------------------------------------------------------------------------------------------------------
 * Hierarchical2D<Float, 0> (include/mitsuba/core/distr_2d.h:475-700) over the flattened block built by
 * mitsuba3_plt_b200/distr2d.py, and EnvironmentMapEmitter (src/emitters/envmap.cpp:343-560)
 * ---------------------------------------------------------------------------------------------------------- */
inline uint32_t as_u32(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
inline float clip01(float x) { return std::min(std::max(x, 0.f), 1.f); }
inline float lerp_(float a, float b, float t) { return fmadd(b, t, fnmadd(a, t, a)); } /* dr::lerp */
/* warp::interval_to_linear (warp.h:447-453) */
inline float interval_to_linear(float v0, float v1, float sample) {
    if (std::fabs(v0 - v1) > 1e-4f * (v0 + v1))
        return (v0 - safe_sqrt(lerp_(sqr(v0), sqr(v1), sample))) / (v0 - v1);
    return sample;
}
struct Hier2D {
    const float *blk; uint32_t n_levels, res_x, res_y;
    explicit Hier2D(const float *b) : blk(b), n_levels(as_u32(b[0])), res_x(as_u32(b[1])), res_y(as_u32(b[2])) {}
    const float *level(uint32_t l) const { return blk + as_u32(blk[4 + 2 * l]); }
    uint32_t width(uint32_t l) const { return as_u32(blk[5 + 2 * l]); }
    static uint32_t index(uint32_t x, uint32_t y, uint32_t w) { return ((x & 1u) | (((x & ~1u) | (y & 1u)) << 1)) + (y & ~1u) * w; }
    /* sample(), distr_2d.h:481-563 */
    void sample(V2 s, V2 &out, float &pdf) const {
        s.x = clip01(s.x); s.y = clip01(s.y);
        uint32_t ox = 0, oy = 0;
        for (int l = (int) n_levels - 2; l > 0; --l) {
            const float *lv = level((uint32_t) l);
            ox <<= 1; oy <<= 1;
            uint32_t i = index(ox, oy, width((uint32_t) l));
            float v00 = lv[i], v10 = lv[i + 1], v01 = lv[i + 2], v11 = lv[i + 3];
            s.x = clip01(s.x); s.y = clip01(s.y);
            float r0 = v00 + v10, r1 = v01 + v11;
            s.y *= r0 + r1;
            bool m = s.y > r0;
            if (m) { oy += 1; s.y -= r0; }
            s.y /= m ? r1 : r0;
            float c0 = m ? v01 : v00, c1 = m ? v11 : v10;
            s.x *= c0 + c1;
            m = s.x > c0;
            if (m) s.x -= c0;
            s.x /= m ? c1 : c0;
            if (m) ox += 1;
        }
        const float *l0 = level(0);
        uint32_t w = width(0), i = ox + oy * w;
        float v00 = l0[i], v10 = l0[i + 1], v01 = l0[i + w], v11 = l0[i + w + 1];
        /* warp::square_to_bilinear, warp.h:480-494 */
        float r0 = v00 + v10, r1 = v01 + v11;
        s.y = interval_to_linear(r0, r1, s.y);
        float c0 = lerp_(v00, v01, s.y), c1 = lerp_(v10, v11, s.y);
        s.x = interval_to_linear(c0, c1, s.x);
        pdf = lerp_(c0, c1, s.x);
        float psx = 1.f / (float) (res_x - 1), psy = 1.f / (float) (res_y - 1); /* m_patch_size */
        out.x = ((float) (int32_t) ox + s.x) * psx;
        out.y = ((float) (int32_t) oy + s.y) * psy;
    }
    /* eval(), distr_2d.h:668-699 */
    float eval(V2 pos) const {
        pos.x = clip01(pos.x) * (float) (res_x - 1); pos.y = clip01(pos.y) * (float) (res_y - 1);
        uint32_t ox = std::min((uint32_t) (int32_t) pos.x, res_x - 2), oy = std::min((uint32_t) (int32_t) pos.y, res_y - 2);
        pos.x -= (float) (int32_t) ox; pos.y -= (float) (int32_t) oy;
        const float *l0 = level(0);
        uint32_t w = width(0), i = ox + oy * w;
        float v00 = l0[i], v10 = l0[i + 1], v01 = l0[i + w], v11 = l0[i + w + 1];
        return lerp_(lerp_(v00, v10, pos.x), lerp_(v01, v11, pos.x), pos.y); /* square_to_bilinear_pdf, warp.h:516-521 */
    }
};
inline float safe_acos(float x) { return std::acos(std::min(std::max(x, -1.f), 1.f)); }
/* EnvironmentMapEmitter::eval_spectrum (envmap.cpp:524-580), include_whitepoint = true */
S4 tex_eval(const Scene &sc, int32_t id, const S4 &wl, V2 uv);
inline float srgb_model_eval(const float *c, float wl);
S4 envmap_eval_spectrum(const Scene &sc, const pltb_emitter &em, V2 uv, const S4 &wl) {
    const uint32_t rx = em.env_res[0], ry = em.env_res[1];
    uv.x -= .5f / (float) (rx - 1u);
    uv.x -= std::floor(uv.x); uv.y -= std::floor(uv.y);
    uv.x *= (float) (rx - 1u); uv.y *= (float) (ry - 1u);
    uint32_t px = std::min((uint32_t) uv.x, rx - 2u), py = std::min((uint32_t) uv.y, ry - 2u);
    float w1x = uv.x - (float) px, w1y = uv.y - (float) py, w0x = 1.f - w1x, w0y = 1.f - w1y;
    const float *data = sc.d->tex_data + em.env_data;
    uint32_t index = py * rx + px;
    const float *v00 = data + 4 * (size_t) index, *v10 = v00 + 4, *v01 = v00 + 4 * (size_t) rx, *v11 = v01 + 4;
    if (sc.rgb) {   /* envmap.cpp:569-579: texels are (r, g, b, unused); no whitepoint */
        S4 result;
        for (int c = 0; c < 3; ++c) {
            float v0 = fmadd(w0x, v00[c], w1x * v10[c]), v1 = fmadd(w0x, v01[c], w1x * v11[c]);
            result[c] = fmadd(w0y, v0, w1y * v1) * em.env_scale;
        }
        result[3] = result[2];
        return result;
    }
    S4 white = tex_eval(sc, em.radiance, wl, V2{ 0.f, 0.f });
    S4 result;
    float f0 = fmadd(w0x, v00[3], w1x * v10[3]), f1 = fmadd(w0x, v01[3], w1x * v11[3]), f = fmadd(w0y, f0, w1y * f1);
    for (int i = 0; i < 4; ++i) {
        float s00 = srgb_model_eval(v00, wl[i]), s10 = srgb_model_eval(v10, wl[i]);
        float s01 = srgb_model_eval(v01, wl[i]), s11 = srgb_model_eval(v11, wl[i]);
        float s0 = fmadd(w0x, s00, w1x * s10), s1 = fmadd(w0x, s01, w1x * s11), sv = fmadd(w0y, s0, w1y * s1);
        result[i] = ((sv * f) * em.env_scale) * white[i];
    }
    return result;
}
inline V2 envmap_dir_to_uv(V3 v) { /* envmap.cpp:348-350 */
    return V2{ std::atan2(v.x, -v.z) * (0.5f * kInvPi), safe_acos(v.y) * kInvPi };
}
inline float envmap_inv_sin_theta(V3 d) { /* safe_rsqrt(max(d.x^2 + d.z^2, Epsilon^2)), envmap.cpp:443-444 */
    float v = std::max(sqr(d.x) + sqr(d.z), sqr(kEps));
    return 1.f / std::sqrt(std::max(v, 0.f));
}
/* ::pdf_direction (envmap.cpp:468-484) */
float envmap_pdf_direction(const Scene &sc, const pltb_emitter &em, V3 d_world) {
    V3 d = xf_vector(em.to_local, d_world);
    V2 uv = envmap_dir_to_uv(d);
    uv.x -= .5f / (float) (em.env_res[0] - 1u);
    uv.x -= std::floor(uv.x); uv.y -= std::floor(uv.y);
    Hier2D warp(sc.d->tex_data + em.env_warp);
    return warp.eval(uv) * envmap_inv_sin_theta(d) * (1.f / (2.f * sqr(kPi)));
}
/* warp::square_to_uniform_sphere (include/mitsuba/core/warp.h:250-255) */
inline V3 square_to_uniform_sphere(V2 sample) {
    float z = fnmadd(2.f, sample.y, 1.f), r = safe_sqrt(fnmadd(z, z, 1.f));
    float a = 2.f * kPi * sample.x;
    return V3(r * std::cos(a), r * std::sin(a), z);
}
struct DirectionSample { V3 p, n, d; V2 uv; float pdf = 0.f, dist = 0.f; bool delta = false; int32_t emitter = -1; };

/* Mesh::sample_position (src/render/mesh.cpp:751-805) with DiscreteDistribution::sample_reuse
 * (include/mitsuba/core/distr_1d.h:116-182, JIT predicate) */
void mesh_sample_position(const Scene &sc, const pltb_shape &sh, V2 sample, V3 &p, V3 &n, V2 &uv, float &pdf) {
    const pltb_scene_desc *d = sc.d;
    const float *cdf = d->area_cdf + sh.cdf_offset;
    float sum = sh.area_sum, value = sample.y * sum;
    uint32_t lo = 0, hi = sh.n_faces - 1; /* first index in [0, n-1] whose predicate is false */
    while (lo < hi) {
        uint32_t mid = (lo + hi) / 2;
        float c = cdf[mid];
        bool pred = ((c < value) || c == 0.f) && c != sum;
        if (pred) lo = mid + 1; else hi = mid;
    }
    uint32_t face = lo;
    float pmf = (cdf[face] - (face > 0 ? cdf[face - 1] : 0.f)) * sh.area_norm;
    float cdf_prev = face > 0 ? cdf[face - 1] * sh.area_norm : 0.f;
    sample.y = (sample.y - cdf_prev) / pmf;
    uint32_t fi = sh.first_face + face;
    uint32_t f0 = d->faces[3 * fi], f1 = d->faces[3 * fi + 1], f2 = d->faces[3 * fi + 2];
    V3 p0 = sc.pos(f0), p1 = sc.pos(f1), p2 = sc.pos(f2);
    V3 e0 = p1 - p0, e1 = p2 - p0;
    V2 b = square_to_uniform_triangle(sample);
    p = fmadd(e0, b.x, fmadd(e1, b.y, p0));
    pdf = sh.area_norm;
    if (sh.flags & PLTB_SHAPE_HAS_UV) {
        V2 uv0 = sc.uv(f0), uv1 = sc.uv(f1), uv2 = sc.uv(f2);
        float b0 = 1.f - b.x - b.y;
        uv = { fmadd(uv0.x, b0, fmadd(uv1.x, b.x, uv2.x * b.y)), fmadd(uv0.y, b0, fmadd(uv1.y, b.x, uv2.y * b.y)) };
    } else uv = b;
    if (sh.flags & PLTB_SHAPE_HAS_NORMALS) {
        V3 n0 = sc.nrm(f0), n1 = sc.nrm(f1), n2 = sc.nrm(f2);
        n = fmadd(n0, 1.f - b.x - b.y, fmadd(n1, b.x, n2 * b.y));
    } else n = cross(e0, e1);
    n = normalize(n);
    if (sh.flags & PLTB_SHAPE_FLIP_NORMALS) n = -n;
}

/* Shape::sample_direction (src/render/shape.cpp:356-373) for meshes and rectangles (rectangle.cpp:206-240 has its
 * own but equivalent area-based version); Sphere::sample_direction (src/shapes/sphere.cpp:234-309). */
DirectionSample shape_sample_direction(const Scene &sc, const pltb_shape &sh, V3 ref_p, V2 sample) {
    DirectionSample ds;
    if (sh.type == PLTB_SHAPE_SPHERE) {
        V3 center(sh.center[0], sh.center[1], sh.center[2]);
        V3 dc_v = center - ref_p;
        float dc_2 = sqnorm(dc_v);
        float radius = sh.radius;
        float radius_adj = radius * ((sh.flags & PLTB_SHAPE_FLIP_NORMALS) ? (1.f + kRayEps) : (1.f - kRayEps));
        bool outside = dc_2 > sqr(radius_adj);
        if (outside) {
            float inv_dc = rsqrt(dc_2), sin_theta_max = radius * inv_dc, sin_theta_max_2 = sqr(sin_theta_max),
                  inv_sin_theta_max = rcp(sin_theta_max), cos_theta_max = safe_sqrt(1.f - sin_theta_max_2);
            float sin_theta_2 = sin_theta_max_2 > 0.00068523f ? 1.f - sqr(fmadd(cos_theta_max - 1.f, sample.x, 1.f))
                                                               : sin_theta_max_2 * sample.x,
                  cos_theta = safe_sqrt(1.f - sin_theta_2);
            float cos_alpha = sin_theta_2 * inv_sin_theta_max +
                              cos_theta * safe_sqrt(fnmadd(sin_theta_2, sqr(inv_sin_theta_max), 1.f)),
                  sin_alpha = safe_sqrt(fnmadd(cos_alpha, cos_alpha, 1.f));
            float ang = sample.y * (2.f * kPi);
            float sin_phi = std::sin(ang), cos_phi = std::cos(ang);
            Frame fr(dc_v * -inv_dc);
            V3 dd = fr.to_world(V3(cos_phi * sin_alpha, sin_phi * sin_alpha, cos_alpha));
            ds.p = fmadd(dd, radius, center);
            ds.n = dd;
            ds.d = ds.p - ref_p;
            float dist2 = sqnorm(ds.d);
            ds.dist = std::sqrt(dist2);
            ds.d = ds.d / ds.dist;
            /* warp::square_to_uniform_cone_pdf, warp.h:330-340 */
            ds.pdf = (0.5f * kInvPi) / (1.f - cos_theta_max);
            if (ds.dist == 0.f) ds.pdf = 0.f;   /* sphere.cpp: masked(ds.pdf, ds.dist == 0) = 0 */
        } else {
            /* inside: uniform sphere sampling */
            float z = fnmadd(2.f, sample.y, 1.f), r = safe_sqrt(fnmadd(z, z, 1.f));
            float ang = 2.f * kPi * sample.x;
            V3 dd(r * std::cos(ang), r * std::sin(ang), z);
            ds.p = fmadd(dd, radius, center);
            ds.n = dd;
            ds.d = ds.p - ref_p;
            float dist2 = sqnorm(ds.d);
            ds.dist = std::sqrt(dist2);
            ds.d = ds.d / ds.dist;
            ds.pdf = (kInvPi * 0.25f / sqr(radius)) * dist2 / std::fabs(dot(ds.d, ds.n));
            if (!std::isfinite(ds.pdf)) ds.pdf = 0.f;
        }
        if (sh.flags & PLTB_SHAPE_FLIP_NORMALS) ds.n = -ds.n;
        return ds;
    }
    float pdf;
    if (sh.type == PLTB_SHAPE_MESH) {
        mesh_sample_position(sc, sh, sample, ds.p, ds.n, ds.uv, pdf);
    } else { /* Rectangle::sample_position, rectangle.cpp:186-204 */
        ds.p = xf_point(sh.to_world, V3(fmsub(sample.x, 2.f, 1.f), fmsub(sample.y, 2.f, 1.f), 0.f));
        ds.n = V3(sh.frame_n[0], sh.frame_n[1], sh.frame_n[2]);
        if (sh.flags & PLTB_SHAPE_FLIP_NORMALS) ds.n = -ds.n;
        ds.uv = sample;
        pdf = sh.inv_area;
    }
    ds.pdf = pdf;
    ds.d = ds.p - ref_p;
    float dist_squared = sqnorm(ds.d);
    ds.dist = std::sqrt(dist_squared);
    ds.d = ds.d / ds.dist;
    float dp = std::fabs(dot(ds.d, ds.n));
    float x = dist_squared / dp;
    ds.pdf *= std::isfinite(x) ? x : 0.f;
    return ds;
}

/* Shape::pdf_direction (shape.cpp:375-386); Sphere::pdf_direction (sphere.cpp:311-330) */
float shape_pdf_direction(const pltb_shape &sh, V3 ref_p, const DirectionSample &ds) {
    if (sh.type == PLTB_SHAPE_SPHERE) {
        V3 center(sh.center[0], sh.center[1], sh.center[2]);
        float sin_alpha = sh.radius * rcp(norm(center - ref_p)), cos_alpha = safe_sqrt(1.f - sin_alpha * sin_alpha);
        if (sin_alpha < 1.f - kEps /* DeltaEpsilon? use OneMinusEpsilon */)
            return (0.5f * kInvPi) / (1.f - cos_alpha);
        return (kInvPi * 0.25f / sqr(sh.radius)) * sqr(ds.dist) / std::fabs(dot(ds.d, ds.n));
    }
    float pdf = sh.type == PLTB_SHAPE_MESH ? sh.area_norm : sh.inv_area;
    float dp = std::fabs(dot(ds.d, ds.n));
    pdf *= dp != 0.f ? (ds.dist * ds.dist) / dp : 0.f;
    return pdf;
}

/* ------------------------------------------------------------------------------------------------------------
 * The integrator: PLTIntegrator.sample (scripts/rendering/integrators/plt.py:50-529)
 * ---------------------------------------------------------------------------------------------------------- */
struct Bounce {
    SI si;
    V3 wo;
    uint32_t flags = 0;
    float last_nd_pdf = 0.f;
    int lobe[2] = { 0, 0 };
    bool active = false;
    bool lobe_valid = true;   /* trace only */
};

/* Non-finite trace (tools/nan_trace.py): with PLT_ORACLE_NAN_TRACE=<file> every non-finite term of plt_solve_phase is written as
 * one JSON line naming the factor that is not finite and the state of the path that led there (per bounce: BSDF type, |wo|,
 * sampled lobe and whether diffract() flagged it invalid, |wi| of the vertex).  Test infrastructure, single-threaded use. */
static FILE *nan_trace_file() {
    static FILE *f = [] { const char *p = getenv("PLT_ORACLE_NAN_TRACE"); return p ? fopen(p, "a") : (FILE *) nullptr; }();
    return f;
}
static bool s4_finite(const S4 &s) { return std::isfinite(s[0]) && std::isfinite(s[1]) && std::isfinite(s[2]) && std::isfinite(s[3]); }

struct Counters { uint64_t closest = 0, shadow = 0, vertices = 0, grating = 0; };

inline S4 emitter_radiance(const Scene &sc, const pltb_emitter &em, const S4 &wl, V2 uv) { return tex_eval(sc, em.radiance, wl, uv); }

/* replay_path, plt.py:408-472 */
S4 replay_path(const Scene &sc, const Bounce *bb, int bounce_idx) {
    S4 alpha(1.f);
    for (int i = bounce_idx - 1; i >= 0; --i) {
        const Bounce &b = bb[i];
        /* coherence.propagate(t): carried but inert */
        if (b.active) alpha = wbsdf_weight(sc, b.si, b.wo) * alpha;
    }
    return alpha;
}

/* MIS-weighted emitted radiance seen at bounce `bi` (solve_replay_emissive, plt.py:315-405, before the replayed throughput is
 * applied): mis_weight(last_nd_pdf, pdf_emitter_direction) * emitter.eval.  False when the vertex sees no emitter. */
bool emissive_term(const Scene &sc, const Bounce *bb, uint32_t bi, S4 &Lem) {
    const pltb_scene_desc *d = sc.d;
    const Bounce &b = bb[bi];
    if (b.si.valid()) {
        int32_t em_id = d->shapes[b.si.shape].emitter;
        if (em_id < 0) return false;
        const pltb_emitter &em = d->emitters[em_id];
        bool prev_delta = bi > 0 ? (bb[bi - 1].flags & F_Delta) != 0 : true;
        V3 prev_p = bi > 0 ? bb[bi - 1].si.p : V3(0, 0, 0);
        /* DirectionSample3f(scene, si, prev_si), records.h:173-180 */
        DirectionSample ds;
        ds.p = b.si.p; ds.n = b.si.sh.n; ds.uv = b.si.uv;
        V3 rel = b.si.p - prev_p;
        ds.dist = norm(rel);
        ds.d = rel / ds.dist;
        /* Scene::pdf_emitter_direction (scene.cpp:350-361) -> AreaLight::pdf_direction (area.cpp:170-196) */
        float em_pdf = 0.f;
        if (!prev_delta) {
            float dp = dot(ds.d, ds.n);
            if (dp < 0.f) em_pdf = shape_pdf_direction(d->shapes[em.shape], prev_p, ds);
            em_pdf *= sc.emitter_pmf;
        }
        float mis_bsdf = mis_weight(b.last_nd_pdf, em_pdf);
        /* AreaLight::eval, area.cpp:82-89 */
        S4 Le = b.si.wi.z > 0.f ? emitter_radiance(sc, em, b.si.wavelengths, b.si.uv) : S4(0.f);
        Lem = Le * mis_bsdf;
        return true;
    }
    if (sc.env < 0) return false;
    /* the ray escaped: si.emitter(scene) is the environment (interaction.h:343-346); DirectionSample3f(scene, si, prev_si)
     * takes d = -si.wi (records.h:173-180) */
    const pltb_emitter &em = d->emitters[sc.env];
    bool prev_delta = bi > 0 ? (bb[bi - 1].flags & F_Delta) != 0 : true;
    /* ConstantBackgroundEmitter::pdf_direction (constant.cpp:152-157) / EnvironmentMapEmitter (envmap.cpp:468-484) */
    float em_pdf = 0.f;
    if (!prev_delta)
        em_pdf = (em.type == PLTB_EMITTER_ENVMAP ? envmap_pdf_direction(sc, em, -b.si.wi) : kInvFourPi) * sc.emitter_pmf;
    float mis_bsdf = mis_weight(b.last_nd_pdf, em_pdf);
    S4 Le;
    if (em.type == PLTB_EMITTER_ENVMAP) /* ::eval, envmap.cpp:343-353 */
        Le = envmap_eval_spectrum(sc, em, envmap_dir_to_uv(xf_vector(em.to_local, -b.si.wi)), b.si.wavelengths);
    else
        Le = emitter_radiance(sc, em, b.si.wavelengths, V2{ 0.f, 0.f }); /* ::eval, constant.cpp:92-96 */
    Lem = Le * mis_bsdf;
    return true;
}

/* The emitter half of solve_replay_NEE (plt.py:221-269) for one recorded bounce: Scene::sample_emitter_direction with its
 * shadow ray, the re-evaluated emitter value and em_weight = em_val / ds.pdf.  False when NEE is inactive for the vertex. */
struct NeeSample { S4 em_weight; V3 wo; float pdf = 0.f; bool delta = false; };
bool nee_light_sample(const Scene &sc, const Bounce &b, V2 u2, Counters &cnt, NeeSample &out) {
    const pltb_scene_desc *d = sc.d;
    bool active_em = b.active && (b.flags & F_Smooth) != 0;
    if (!(active_em && d->n_emitters > 0)) return false;
    const SI &si = b.si;
    /* Scene::sample_emitter_direction (scene.cpp:294-348) + sample_emitter (:222-245) */
    uint32_t index = 0;
    V2 sample = u2;
    if (d->n_emitters > 1) {
        float cf = (float) d->n_emitters, scaled = sample.x * cf;
        index = std::min((uint32_t) scaled, d->n_emitters - 1u);
        sample.x = scaled - (float) index;
    }
    const pltb_emitter &em = d->emitters[index];
    DirectionSample ds;
    if (em.type == PLTB_EMITTER_AREA) {
        ds = shape_sample_direction(sc, d->shapes[em.shape], si.p, sample); /* area.cpp:116-168 */
    } else if (em.type == PLTB_EMITTER_CONSTANT) { /* ::sample_direction, constant.cpp:121-150 */
        V3 dir = square_to_uniform_sphere(sample);
        V3 c(em.bsphere_center[0], em.bsphere_center[1], em.bsphere_center[2]);
        float radius = std::max(em.bsphere_radius, norm(si.p - c)), dist = 2.f * radius;
        ds.p = fmadd(dir, dist, si.p);
        ds.n = -dir; ds.uv = sample; ds.pdf = kInvFourPi; ds.delta = false; ds.d = dir; ds.dist = dist;
    } else if (em.type == PLTB_EMITTER_ENVMAP) { /* ::sample_direction, envmap.cpp:421-466 */
        Hier2D warp(d->tex_data + em.env_warp);
        V2 uv; float pdf;
        warp.sample(sample, uv, pdf);
        uv.x += .5f / (float) (em.env_res[0] - 1u);
        bool act_s = pdf > 0.f;
        float theta = uv.y * kPi, phi = uv.x * (2.f * kPi);
        float st = std::sin(theta), ct = std::cos(theta), sp = std::sin(phi), cp = std::cos(phi);
        V3 sd(cp * st, sp * st, ct); /* dr::sphdir */
        V3 dl(sd.y, sd.z, -sd.x);
        V3 c(em.bsphere_center[0], em.bsphere_center[1], em.bsphere_center[2]);
        float radius = std::max(em.bsphere_radius, norm(si.p - c)), dist = 2.f * radius;
        float inv_sin_theta = envmap_inv_sin_theta(dl);
        V3 dw = xf_vector(em.to_world, dl);
        ds.p = si.p + dw * dist;
        ds.n = -dw; ds.uv = uv; ds.delta = false; ds.d = dw; ds.dist = dist;
        ds.pdf = act_s ? pdf * inv_sin_theta * (1.f / (2.f * sqr(kPi))) : 0.f;
    } else if (em.type == PLTB_EMITTER_DIRECTIONALSPOT) { /* DirectionalSpreadEmitter::sample_direction, src/emitters/directionalspot.cpp:161-192 */
        /* warp::square_to_uniform_disk (warp.h:37-42) * sin(spread_angle) */
        float r = std::sqrt(sample.y), ang = kTwoPi * sample.x;
        float sn = std::sin(ang), cs = std::cos(ang);
        V2 angular = { (cs * r) * em.env_pad[0], (sn * r) * em.env_pad[0] };
        V3 dd = xf_vector(em.to_world, normalize(V3(angular.x, angular.y, 1.f)));
        V3 c(em.bsphere_center[0], em.bsphere_center[1], em.bsphere_center[2]);
        float radius = std::max(em.bsphere_radius, norm(si.p - c)), dist = 2.f * radius;
        ds.p = si.p - dd * dist;
        ds.n = dd; ds.uv = V2{ 0.f, 0.f }; ds.pdf = 1.f; ds.delta = true; ds.d = -dd; ds.dist = dist;
    } else { /* DirectionalEmitter::sample_direction, src/emitters/directional.cpp:149-176 */
        V3 dirn(em.direction[0], em.direction[1], em.direction[2]);
        ds.d = -dirn;
        float dist = 2.f * em.bsphere_radius;
        ds.p = fmadd(ds.d, dist, si.p);
        ds.n = dirn; ds.pdf = 1.f; ds.delta = true; ds.dist = dist;
    }
    ds.emitter = (int32_t) index;
    ds.pdf *= sc.emitter_pmf;
    if (ds.pdf != 0.f) {
        cnt.shadow++;
        Ray sr = si.spawn_ray_to(ds.p);
        Hit sh = intersect(sc, sr, true);
        if (sh.shape >= 0) ds.pdf = 0.f;
    }
    /* back in plt.py:264-269 */
    ds.d = normalize(ds.p - si.p);
    S4 em_val(0.f);
    if (em.type == PLTB_EMITTER_AREA) { /* AreaLight::eval_direction, area.cpp:198-208 */
        if (dot(ds.d, ds.n) < 0.f) em_val = emitter_radiance(sc, em, si.wavelengths, ds.uv);
    } else if (em.type == PLTB_EMITTER_ENVMAP) { /* ::eval_direction, envmap.cpp:486-491 */
        em_val = envmap_eval_spectrum(sc, em, ds.uv, si.wavelengths);
    } else { /* DirectionalEmitter / ConstantBackgroundEmitter::eval_direction (directional.cpp:189-195) */
        em_val = emitter_radiance(sc, em, si.wavelengths, ds.uv);
    }
    out.em_weight = ds.pdf != 0.0f ? em_val / ds.pdf : S4(0.f);
    out.wo = si.to_local(ds.d);
    out.pdf = ds.pdf; out.delta = ds.delta;
    return true;
}

static void trace_nonfinite(const Scene &sc, const Bounce *bb, uint32_t n_written, uint32_t bi, const char *term, const S4 &mis,
                            const S4 &emitter, const S4 &bsdf_val, const S4 &alpha) {
    FILE *f = nan_trace_file();
    const char *first = !s4_finite(emitter) ? "emitter" : !s4_finite(mis) ? "mis" : !s4_finite(bsdf_val) ? "bsdf_val"
                        : !s4_finite(alpha) ? "replay_alpha" : "product(0*inf)";
    fprintf(f, "{\"term\": \"%s\", \"bounce\": %u, \"first_nonfinite_factor\": \"%s\", \"emitter\": %g, \"mis\": %g, \"bsdf_val\": %g, "
               "\"alpha\": %g, \"path\": [", term, bi, first, emitter[0], mis[0], bsdf_val[0], alpha[0]);
    for (uint32_t j = 0; j < n_written && j <= bi; ++j) {
        const Bounce &b = bb[j];
        const pltb_bsdf *bs = b.si.valid() ? si_bsdf(sc, b.si) : nullptr;
        S4 w = (b.si.valid() && j < bi) ? wbsdf_weight(sc, b.si, b.wo) : S4(0.f);
        fprintf(f, "%s{\"bsdf\": %d, \"wi_norm\": %g, \"wo_norm\": %g, \"lobe\": [%d, %d], \"lobe_valid\": %s, \"weight\": %g}", j ? ", " : "",
                bs ? (int) bs->type : -1, norm(b.si.wi), norm(b.wo), b.lobe[0], b.lobe[1], b.lobe_valid ? "true" : "false", w[0]);
    }
    fprintf(f, "]}\n");
    fflush(f);
}

/* plt.py:64-70: four draws scaled by scale_wl = (MI_CIE_MAX - 150 - MI_CIE_MIN) + MI_CIE_MIN = 680.  Dead in the spectral
 * variants (N1); in the RGB variants the first three are the path's `sampling_wavelengths` (BounceData / PLTSamplePhaseData,
 * :158,280), which the RGB fallback of roughgrating reads -- they replace `wavelengths` (empty in RGB mode) from here on. */
inline void draw_sampling_wavelengths(const Scene &sc, PCG32 &rng, S4 &wavelengths) {
    float u[4];
    for (int i = 0; i < 4; ++i) u[i] = rng.next_float32();
    if (sc.rgb) { wavelengths[0] = u[0] * 680.f; wavelengths[1] = u[1] * 680.f; wavelengths[2] = wavelengths[3] = u[2] * 680.f; }
}

S4 plt_sample(const Scene &sc, PCG32 &rng, Ray ray, S4 &wavelengths, uint32_t max_depth, uint32_t rr_depth,
              Counters &cnt) {
    Bounce bb[PLTB_MAX_DEPTH];

    /* ---- plt_sample_phase, plt.py:50-171 ---- */
    draw_sampling_wavelengths(sc, rng, wavelengths);
    S4 alpha(1.f);
    float eta = 1.f;
    uint32_t depth = 0, i = 0;
    bool active = true, prev_bsdf_delta = true;
    float last_nd_pdf = 1.f;
    while (active) {
        Hit h = intersect(sc, ray, false);
        cnt.closest++;
        SI si = make_si(sc, ray, h, wavelengths);
        bool active_next = (depth + 1 < max_depth) && si.valid();
        float sample_1 = rng.next_float32();
        V2 sample_2, lobe_sample_2;
        sample_2.x = rng.next_float32(); sample_2.y = rng.next_float32();
        lobe_sample_2.x = rng.next_float32(); lobe_sample_2.y = rng.next_float32();
        SampleOut so = wbsdf_sample(sc, si, sample_1, sample_2, lobe_sample_2);
        ray = si.spawn_ray(si.to_world(so.bs.wo));
        float pdf = so.bs.pdf;
        if (active_next) { alpha = so.weight * alpha; eta = so.bs.eta; }
        prev_bsdf_delta = (so.bs.type & F_Delta) != 0;
        if (si.valid()) depth += 1;
        float alpha_max = smax(alpha);
        active_next &= alpha_max != 0;
        float rr_prob = std::min(alpha_max * sqr(eta), 0.95f);
        bool rr_active = depth >= rr_depth;
        float rcp_thp = rcp(rr_prob);
        if (rr_active) alpha = alpha * rcp_thp;
        bool rr_continue = rng.next_float32() < rr_prob;
        active_next &= !rr_active || rr_continue;
        Bounce &b = bb[i];
        b.si = si; b.wo = so.bs.wo; b.flags = so.bs.type; b.last_nd_pdf = last_nd_pdf;
        b.lobe[0] = so.lobe[0]; b.lobe[1] = so.lobe[1]; b.active = true; b.lobe_valid = so.lobe_valid;
        if (si.valid()) {
            cnt.vertices++;
            const pltb_bsdf *bs = si_bsdf(sc, si);
            if (bs && bs->type == PLTB_BSDF_ROUGHGRATING) cnt.grating++;
        }
        last_nd_pdf = prev_bsdf_delta ? last_nd_pdf : pdf;
        active = active_next;
        i += 1;
    }
    uint32_t n_written = i;

    /* ---- plt_solve_phase, plt.py:174-218 ---- */
    S4 L(0.f);
    for (uint32_t bi = 0; bi < max_depth; ++bi) {
        bool written = bi < n_written; /* unwritten slots: all-zero, active = false (N4) */
        /* solve_replay_emissive, :315-405 */
        S4 Lem;
        if (written && emissive_term(sc, bb, bi, Lem)) {
            S4 a = replay_path(sc, bb, (int) bi);
            L = L + a * Lem;
            if (nan_trace_file() && !s4_finite(a * Lem))
                trace_nonfinite(sc, bb, n_written, bi, "emissive", S4(1.f), Lem, S4(1.f), a);
        }
        /* solve_replay_NEE, :221-300 */
        V2 u2; u2.x = rng.next_float32(); u2.y = rng.next_float32(); /* drawn for every i, every lane */
        NeeSample ns;
        if (written && nee_light_sample(sc, bb[bi], u2, cnt, ns)) {
            const SI &si = bb[bi].si;
            S4 bsdf_val = wbsdf_eval(sc, si, ns.wo);
            float bsdf_pdf = wbsdf_pdf(sc, si, ns.wo);
            float mis_em = ns.delta ? 1.0f : mis_weight(ns.pdf, bsdf_pdf);
            S4 a = replay_path(sc, bb, (int) bi);
            L = L + ((ns.em_weight * mis_em) * bsdf_val) * a;
            if (nan_trace_file() && !s4_finite(((ns.em_weight * mis_em) * bsdf_val) * a))
                trace_nonfinite(sc, bb, n_written, bi, "nee", S4(mis_em), ns.em_weight, bsdf_val, a);
        }
    }
    return L;
}

void spectrum_to_srgb(const float *cie_srgb, const S4 &value, const S4 &wl, float rgb[3]);
#include "plt_oracle_pol.inc"

/* ------------------------------------------------------------------------------------------------------------
 * Sensor + film
 * ---------------------------------------------------------------------------------------------------------- */
/* sample_wavelength (include/mitsuba/core/spectrum.h:446-480) with math::sample_shifted (math.h:554-577);
 * Sensor::sample_wavelengths with an SRF (src/render/sensor.cpp:113-127) */
void sample_wavelengths(const Scene &sc, const pltb_sensor &sn, float sample, S4 &wl, S4 &weight) {
    S4 u;
    for (int i = 0; i < 4; ++i) {
        float v = sample + (float) i / 4.f;
        if (v > 1.f) v -= 1.f;
        u[i] = v;
    }
    if (sn.srf >= 0) {
        const pltb_texture &t = sc.d->textures[sn.srf];
        const float *data = sc.d->tex_data + t.offset;
        /* regular / irregular sample_spectrum: ContinuousDistribution::sample_pdf inversion.  The CDF (t.count
           floats for regular: count-1 entries + pad; irregular likewise) follows the values in tex_data. */
        for (int i = 0; i < 4; ++i) {
            const float *nodes = nullptr, *vals, *cdf;
            uint32_t n = t.count;
            if (t.type == PLTB_TEX_IRREGULAR) { nodes = data; vals = data + n; cdf = data + 2 * n; }
            else { vals = data; cdf = data + n; }
            float integral = cdf[n - 2];
            float s = u[i] * integral;
            uint32_t a = 0, bnd = n - 2; /* first index with cdf[i] >= s */
            while (a < bnd) { uint32_t m = (a + bnd) / 2; if (cdf[m] < s) a = m + 1; else bnd = m; }
            uint32_t index = a;
            float y0 = vals[index], y1 = vals[index + 1], c0 = index > 0 ? cdf[index - 1] : 0.f;
            float x0, x1;
            if (nodes) { x0 = nodes[index]; x1 = nodes[index + 1]; }
            else { float isz = (t.p[4] - t.p[3]) / (float) (n - 1); x0 = fmadd((float) index, isz, t.p[3]); x1 = x0 + isz; }
            float w = x1 - x0;
            float ss = (s - c0) / w; /* sample in [0, area/w] */
            float t_lin = y0 != y1 ? (y0 - safe_sqrt(fmadd(y0, y0, 2.f * ss * (y1 - y0)))) / (y0 - y1) : ss / y0;
            wl[i] = fmadd(t_lin, w, x0);
            weight[i] = integral; /* value / pdf = integral */
        }
        return;
    }
    for (int i = 0; i < 4; ++i) {
        wl[i] = 538.f - std::atanh(0.8569106254698279f - 1.8275019724092267f * u[i]) * 138.88888888888889f;
        float tmp = std::cosh(0.0072f * (wl[i] - 538.f));
        weight[i] = 253.82f * tmp * tmp;
    }
}

/* PerspectiveCamera::sample_ray_differential (src/sensors/perspective.cpp:240-280, differentials unused);
 * OrthographicCamera::sample_ray (src/sensors/orthographic.cpp:144-170) */
Ray sensor_ray(const pltb_sensor &sn, V2 pos) {
    V3 near_p = xf_point_proj(sn.sample_to_camera, V3(pos.x + sn.pp_offset[0], pos.y + sn.pp_offset[1], 0.f));
    Ray ray;
    if (sn.type == PLTB_SENSOR_PERSPECTIVE) {
        V3 d = normalize(near_p);
        ray.o = V3(sn.to_world[3], sn.to_world[7], sn.to_world[11]);
        ray.d = xf_vector(sn.to_world, d);
        float inv_z = rcp(d.z), near_t = sn.near_clip * inv_z, far_t = sn.far_clip * inv_z;
        ray.o = fmadd(ray.d, near_t, ray.o);
        ray.maxt = far_t - near_t;
    } else {
        ray.o = xf_point(sn.to_world, near_p);
        ray.d = normalize(xf_vector(sn.to_world, V3(0.f, 0.f, 1.f)));
        ray.maxt = sn.far_clip - sn.near_clip;
    }
    return ray;
}

/* linear_rgb_rec + spectrum_to_srgb, include/mitsuba/core/spectrum.h:330-392 */
void spectrum_to_srgb(const float *cie_srgb, const S4 &value, const S4 &wl, float rgb[3]) {
    float acc[3] = { 0, 0, 0 };
    for (int i = 0; i < 4; ++i) {
        float t = (wl[i] - 360.f) * (94 / (830.f - 360.f));
        bool active = wl[i] >= 360.f && wl[i] <= 830.f;
        uint32_t i0 = std::min((uint32_t) std::max(t, 0.f), 93u), i1 = i0 + 1;
        float w1 = t - (float) i0, w0 = 1.f - w1;
        for (int c = 0; c < 3; ++c) {
            float v = active ? fmadd(w0, cie_srgb[95 * c + i0], w1 * cie_srgb[95 * c + i1]) : 0.f;
            float prod = v * value[i];
            acc[c] = i == 0 ? prod : acc[c] + prod;
        }
    }
    for (int c = 0; c < 3; ++c) rgb[c] = (acc[c] * 0.25f) * (float) (1.0 / 106.7502593994140625);
}

/* Reconstruction filters: gaussian in its CUDA-variant form (src/rfilters/gaussian.cpp:94-103), tent
 * (src/rfilters/tent.cpp:24-26) */
inline float rfilter_eval(const pltb_film &f, float x) {
    if (f.rfilter == PLTB_RFILTER_GAUSSIAN) {
        float alpha = -1.f / (2.f * sqr(f.stddev));
        float bias = std::exp(alpha * sqr(f.radius));
        return std::max(0.f, std::exp2((1.4426950408889634074f * alpha) * sqr(x)) - bias);
    }
    return std::max(0.f, 1.f - std::fabs(x * (1.f / f.radius)));
}

/* ImageBlock::put, coalesced path (src/render/imageblock.cpp:429-500); film has no border. */
void splat(const pltb_film &f, uint32_t W, uint32_t H, float *film, V2 pos, const float *values, int nch = 4) {
    if (f.rfilter == PLTB_RFILTER_BOX) {
        int x = (int) std::floor(pos.x), y = (int) std::floor(pos.y);
        if (x < 0 || y < 0 || x >= (int) W || y >= (int) H) return;
        for (int k = 0; k < nch; ++k) film[nch * ((size_t) y * W + x) + k] += values[k];
        return;
    }
    int n = (int) std::ceil(f.radius - .5f), count = 2 * n + 1;
    int px = (int) std::floor(pos.x) - n, py = (int) std::floor(pos.y) - n;
    float relx = ((float) px + .5f) - pos.x, rely = ((float) py + .5f) - pos.y;
    float wx[16], wy[16];
    for (int i = 0; i < count; ++i) {
        wx[i] = rfilter_eval(f, relx); wy[i] = rfilter_eval(f, rely);
        relx += 1; rely += 1;
    }
    for (int ys = 0; ys < count; ++ys) {
        int y = py + ys;
        if (y < 0 || y >= (int) H) continue;
        for (int xs = 0; xs < count; ++xs) {
            int x = px + xs;
            if (x < 0 || x >= (int) W) continue;
            float w = wy[ys] * wx[xs];
            float *dst = film + nch * ((size_t) y * W + x);
            for (int k = 0; k < nch; ++k) dst[k] += values[k] * w;
        }
    }
}

void prepare_scene(Scene &sc, const pltb_scene_desc *d) {
    sc.d = d;
    uint32_t n_prims = d->n_faces;
    for (uint32_t s = 0; s < d->n_shapes; ++s) if (d->shapes[s].type != PLTB_SHAPE_MESH) n_prims++;
    sc.prims.clear();
    for (uint32_t i = 0; i < d->n_faces; ++i) sc.prims.push_back(i);
    for (uint32_t s = 0; s < d->n_shapes; ++s) if (d->shapes[s].type != PLTB_SHAPE_MESH) sc.prims.push_back(d->n_faces + s);
    uint32_t total = d->n_faces + d->n_shapes;
    std::vector<float> lo(3 * total, 0.f), hi(3 * total, 0.f);
    for (uint32_t p : sc.prims) prim_bounds(sc, p, &lo[3 * p], &hi[3 * p]);
    sc.nodes.clear();
    if (!sc.prims.empty()) build_bvh(sc, lo, hi, 0, (uint32_t) sc.prims.size());
    else { BVHNode n{}; for (int k = 0; k < 3; ++k) { n.lo[k] = kInf; n.hi[k] = -kInf; } n.count = 0; n.left = n.right = 0; sc.nodes.push_back(n); sc.prims.clear(); sc.nodes[0].first = 0; }
    sc.emitter_pmf = d->n_emitters ? 1.f / (float) d->n_emitters : 0.f;
    for (uint32_t i = 0; i < d->n_emitters; ++i)
        if (d->emitters[i].type == PLTB_EMITTER_CONSTANT || d->emitters[i].type == PLTB_EMITTER_ENVMAP) sc.env = (int32_t) i;
}

} // namespace

/* ================================================================================================================
 * C entry points (loaded by tests/, smoke() and bench.py's CPU legs only)
 * ============================================================================================================== */
extern "C" {

/* ADIntegrator.render (src/python/python/ad/integrators/common.py:46-110) for the sample window described by
 * `params`; film_rgbw (W*H*4 raw accumulator, W = n_sensors * film.width) may be NULL. Returns seconds spent. */
double plt_oracle_render(const pltb_scene_desc *desc, const pltb_render_params *prm, float *film_rgbw,
                         pltb_sample_dump *dump, pltb_stats *stats, int n_threads) {
    Scene sc;
    prepare_scene(sc, desc);
    sc.rgb = (prm->flags & PLTB_RENDER_RGB) != 0;
    const uint32_t fw = desc->film.width, W = fw * desc->n_sensors, H = desc->film.height;
    uint32_t cx = prm->crop_x, cy = prm->crop_y, cw = prm->crop_w ? prm->crop_w : W, ch = prm->crop_h ? prm->crop_h : H;
    if (n_threads <= 0) n_threads = (int) std::thread::hardware_concurrency();
    if (n_threads <= 0) n_threads = 1;
    const uint32_t max_depth = std::min<uint32_t>(prm->max_depth, PLTB_MAX_DEPTH);
    std::vector<std::vector<float>> films;
    const bool polarized = (prm->flags & PLTB_RENDER_POLARIZED) != 0, stokes = polarized && (prm->flags & PLTB_RENDER_STOKES) != 0;
    const int nch = PLTB_FILM_CHANNELS(stokes ? PLTB_RENDER_STOKES : 0);
    if (film_rgbw && !(prm->flags & PLTB_RENDER_NO_SPLAT)) films.assign(n_threads, std::vector<float>((size_t) W * H * nch, 0.f));
    std::vector<Counters> counters(n_threads);
    std::atomic<uint32_t> next_row{ 0 };
    auto t0 = std::chrono::steady_clock::now();
    auto worker = [&](int tid) {
        Counters &cnt = counters[tid];
        for (;;) {
            uint32_t row = next_row.fetch_add(1);
            if (row >= ch) break;
            uint32_t y = cy + row;
            for (uint32_t x = cx; x < cx + cw; ++x) {
                uint32_t pixel = y * W + x;
                for (uint32_t s = prm->spp_begin; s < prm->spp_begin + prm->spp_count; ++s) {
                    uint32_t idx = pixel * prm->spp + s; /* lane index of the reference wavefront, common.py:259-266 */
                    uint32_t v0, v1;
                    sample_tea_32(prm->seed, idx, v0, v1);
                    PCG32 rng; rng.seed(v0, v1);
                    V2 jitter; jitter.x = rng.next_float32(); jitter.y = rng.next_float32();
                    V2 pos_f = { (float) x + jitter.x, (float) y + jitter.y };
                    /* wavelength_sample is only drawn in spectral variants (common.py:297-299) */
                    float wav_sample = sc.rgb ? 0.f : rng.next_float32();
                    /* BatchSensor (src/sensors/batch.cpp:131-190): the film is split in x between the sensors */
                    uint32_t sidx = x / fw;
                    const pltb_sensor &sn = desc->sensors[sidx];
                    V2 pos_adj = { (pos_f.x - (float) (sidx * fw)) * rcp((float) fw), pos_f.y * rcp((float) H) };
                    S4 wl, wgt(1.f);   /* RGB variants: no wavelengths, ray weight 1 (sensor.cpp:113-127) */
                    if (!sc.rgb) sample_wavelengths(sc, sn, wav_sample, wl, wgt);
                    Ray ray = sensor_ray(sn, pos_adj);
                    Counters c0 = cnt;
                    S4 L;
                    float values[16];
                    if (polarized) {
                        M4 Lm = plt_sample_pol(sc, rng, ray, wl, max_depth, prm->rr_depth, cnt);
                        /* scene->sensors()[0]->world_transform() (stokes.cpp:100-102); a `batch` sensor's own transform is the
                         * identity in every config that uses one */
                        static const float ident[16] = { 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1 };
                        if (stokes) Lm = stokes_aovs(sc, Lm, ray, wl, desc->n_sensors > 1 ? ident : desc->sensors[0].to_world, values + 4);
                        L = Lm.e[0][0];   /* unpolarized_spectrum(spec), integrator.cpp:456 */
                        if (dump && dump->stokes) {
                            uint64_t di = ((uint64_t) row * cw + (x - cx)) * prm->spp_count + (s - prm->spp_begin);
                            if (di < dump->capacity)
                                for (int c = 0; c < 4; ++c) for (int k = 0; k < 4; ++k) dump->stokes[16 * di + 4 * c + k] = Lm.e[c][0][k];
                        }
                    } else {
                        L = plt_sample(sc, rng, ray, wl, max_depth, prm->rr_depth, cnt);
                    }
                    if (dump) {
                        uint64_t di = ((uint64_t) row * cw + (x - cx)) * prm->spp_count + (s - prm->spp_begin);
                        if (di < dump->capacity) {
                            if (dump->wavelengths) for (int k = 0; k < 4; ++k) dump->wavelengths[4 * di + k] = wl[k];
                            if (dump->radiance) for (int k = 0; k < 4; ++k) dump->radiance[4 * di + k] = L[k];
                            if (dump->n_rays) {
                                dump->n_rays[2 * di] = (uint32_t) (cnt.closest - c0.closest);
                                dump->n_rays[2 * di + 1] = (uint32_t) (cnt.shadow - c0.shadow);
                            }
                        }
                    }
                    if (!films.empty()) {
                        if (sc.rgb) { S4 v = L * wgt; values[0] = v[0]; values[1] = v[1]; values[2] = v[2]; }   /* rgb = value, common.py:386-391 */
                        else spectrum_to_srgb(desc->cie_srgb, L * wgt, wl, values);
                        values[3] = 1.f;
                        splat(desc->film, W, H, films[tid].data(), pos_f, values, nch);
                    }
                }
            }
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < n_threads; ++t) th.emplace_back(worker, t);
    worker(0);
    for (auto &t : th) t.join();
    double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (!films.empty()) {
        bool accumulate = (prm->flags & PLTB_RENDER_ACCUMULATE) != 0;
        for (size_t i = 0; i < (size_t) W * H * nch; ++i) {
            float acc = accumulate ? film_rgbw[i] : 0.f;
            for (int t = 0; t < n_threads; ++t) acc += films[t][i];
            film_rgbw[i] = acc;
        }
    }
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        stats->samples = (uint64_t) cw * ch * prm->spp_count;
        for (auto &c : counters) {
            stats->closest_rays += c.closest; stats->shadow_rays += c.shadow;
            stats->vertices += c.vertices; stats->grating_vertices += c.grating;
        }
        stats->ms_total = (float) (secs * 1e3);
    }
    return secs;
}

/* HDRFilm::develop (src/films/hdrfilm.cpp:304-400) with AOV channels: every channel / (w == 0 ? 1 : w), W dropped */
void plt_oracle_develop_n(const float *raw, uint64_t n_pixels, int nch, float *out) {
    for (uint64_t i = 0; i < n_pixels; ++i) {
        float w = raw[nch * i + 3];
        float d = w == 0.f ? 1.f : w;
        int o = 0;
        for (int c = 0; c < nch; ++c) if (c != 3) out[(nch - 1) * i + o++] = raw[nch * i + c] / d;
    }
}

/* HDRFilm::develop (src/films/hdrfilm.cpp:304-400): rgb / (w == 0 ? 1 : w) */
void plt_oracle_develop(const float *rgbw, uint64_t n_pixels, float *rgb) {
    for (uint64_t i = 0; i < n_pixels; ++i) {
        float w = rgbw[4 * i + 3];
        float d = w == 0.f ? 1.f : w;
        for (int c = 0; c < 3; ++c) rgb[3 * i + c] = rgbw[4 * i + c] / d;
    }
}

/* ---- component entry points for known-answer tests ---- */
void plt_oracle_tea32(uint32_t v0, uint32_t v1, int rounds, uint32_t *out) { sample_tea_32(v0, v1, out[0], out[1], rounds); }
void plt_oracle_pcg32_floats(uint64_t initstate, uint64_t initseq, int n, float *out, uint32_t *out_u32) {
    PCG32 r; r.seed(initstate, initseq);
    for (int i = 0; i < n; ++i) {
        if (out_u32) out_u32[i] = r.next_uint32(); else out[i] = r.next_float32();
    }
}
void plt_oracle_sampler_floats(uint32_t seed, uint32_t idx, int n, float *out) {
    uint32_t v0, v1; sample_tea_32(seed, idx, v0, v1);
    PCG32 r; r.seed(v0, v1);
    for (int i = 0; i < n; ++i) out[i] = r.next_float32();
}
float plt_oracle_bessel_j(float x, int nu) { return bessel_j(x, nu); }
float plt_oracle_erf(float x) { return dr_erf(x); }
float plt_oracle_erfinv(float x) { return dr_erfinv(x); }
float plt_oracle_fresnel_conductor(float c, float eta, float k) { return fresnel_conductor(c, eta, k); }
void plt_oracle_fresnel(float c, float eta, float *out) { FresnelResult f = fresnel(c, eta); out[0] = f.r; out[1] = f.cos_theta_t; out[2] = f.eta_it; out[3] = f.eta_ti; }
float plt_oracle_mf_eval(uint32_t type, float au, float av, const float *m) { return Microfacet(type, au, av, true).eval(V3(m[0], m[1], m[2])); }
float plt_oracle_mf_g1(uint32_t type, float au, float av, const float *v, const float *m) { return Microfacet(type, au, av, true).smith_g1(V3(v[0], v[1], v[2]), V3(m[0], m[1], m[2])); }
void plt_oracle_mf_sample(uint32_t type, float au, float av, int visible, const float *wi, const float *u, float *out) {
    float pdf; V3 m = Microfacet(type, au, av, visible != 0).sample(V3(wi[0], wi[1], wi[2]), V2{ u[0], u[1] }, pdf);
    out[0] = m.x; out[1] = m.y; out[2] = m.z; out[3] = pdf;
}
/* grating: params = angle, inv_px, inv_py, q, multiplier, uv.x, uv.y */
float plt_oracle_grating_intensity(const float *p, uint32_t lobes, uint32_t type, int lx, int ly, const float *wi, float wl) {
    return Grating(p[0], V2{ p[1], p[2] }, p[3], lobes, type, p[4], V2{ p[5], p[6] }).lobe_intensity(lx, ly, V3(wi[0], wi[1], wi[2]), wl);
}
int plt_oracle_grating_diffract(const float *p, uint32_t lobes, uint32_t type, int lx, int ly, const float *wi, float wl, float *out) {
    bool valid; V3 r = Grating(p[0], V2{ p[1], p[2] }, p[3], lobes, type, p[4], V2{ p[5], p[6] }).diffract(V3(wi[0], wi[1], wi[2]), lx, ly, wl, valid);
    out[0] = r.x; out[1] = r.y; out[2] = r.z; return valid ? 1 : 0;
}
float plt_oracle_grating_lobe_pdf(const float *p, uint32_t lobes, uint32_t type, int lx, int ly, const float *wi, float wl) {
    return Grating(p[0], V2{ p[1], p[2] }, p[3], lobes, type, p[4], V2{ p[5], p[6] }).lobe_pdf(lx, ly, V3(wi[0], wi[1], wi[2]), wl);
}
void plt_oracle_grating_sample_lobe(const float *p, uint32_t lobes, uint32_t type, const float *u, const float *wi, float wl, int *lobe, float *pdf) {
    V2 pd; Grating(p[0], V2{ p[1], p[2] }, p[3], lobes, type, p[4], V2{ p[5], p[6] }).sample_lobe(V2{ u[0], u[1] }, V3(wi[0], wi[1], wi[2]), wl, lobe, pd);
    pdf[0] = pd.x; pdf[1] = pd.y;
}
float plt_oracle_grating_alpha(const float *p, uint32_t lobes, uint32_t type, const float *wi, float k) {
    return Grating(p[0], V2{ p[1], p[2] }, p[3], lobes, type, p[4], V2{ p[5], p[6] }).alpha(V3(wi[0], wi[1], wi[2]), k);
}
void plt_oracle_sample_wavelengths(float sample, float *wl, float *w) {
    Scene sc; pltb_sensor sn{}; sn.srf = -1; S4 a, b; sample_wavelengths(sc, sn, sample, a, b);
    for (int i = 0; i < 4; ++i) { wl[i] = a[i]; w[i] = b[i]; }
}
void plt_oracle_spectrum_to_srgb(const float *cie_srgb, const float *value, const float *wl, float *rgb) {
    S4 v, w; for (int i = 0; i < 4; ++i) { v[i] = value[i]; w[i] = wl[i]; }
    spectrum_to_srgb(cie_srgb, v, w, rgb);
}
float plt_oracle_srgb_model_eval(const float *c, float wl) { return srgb_model_eval(c, wl); }
/* xyz_to_srgb(cie1931_xyz(wl)), the lobe colour of roughgrating's RGB fallback; cie = the 6 x 95 table of pltb_scene_desc */
void plt_oracle_wavelength_to_srgb(const float *cie, float wl, float *rgb) { wavelength_to_srgb(cie, wl, rgb); }
/* RoughGrating::wbsdf_eval / wbsdf_pdf in an RGB variant for the BSDF of shape `shape`: si.wavelengths = the path's sampling wavelengths.
 * out = { eval[3], pdf, lobe sum before Fresnel[3] } */
void plt_oracle_bsdf_probe_rgb(const pltb_scene_desc *d, int32_t shape, const float *wi, const float *uv, const float *swl3, const float *wo, float *out) {
    Scene sc; prepare_scene(sc, d); sc.rgb = true;
    SI si; si.t = 1.f; si.shape = shape; si.wi = V3(wi[0], wi[1], wi[2]); si.uv = V2{ uv[0], uv[1] };
    si.n = V3(0.f, 0.f, 1.f); si.sh.n = si.n; si.sh.s = V3(1.f, 0.f, 0.f); si.sh.t = V3(0.f, 1.f, 0.f);
    si.wavelengths[0] = swl3[0]; si.wavelengths[1] = swl3[1]; si.wavelengths[2] = si.wavelengths[3] = swl3[2];
    const V3 w(wo[0], wo[1], wo[2]);
    const S4 e = wbsdf_eval(sc, si, w);
    out[0] = e[0]; out[1] = e[1]; out[2] = e[2]; out[3] = wbsdf_pdf(sc, si, w);
    const pltb_bsdf *b = si_bsdf(sc, si);
    S4 ls = (b && b->type == PLTB_BSDF_ROUGHGRATING) ? grating_lobe_sum_rgb(sc, *b, si, w) : S4(0.f);
    out[4] = ls[0]; out[5] = ls[1]; out[6] = ls[2];
}
/* Principled::eval / ::pdf for a spatially uniform base colour (KAT against src/bsdfs/tests/test_principled.py:124-200);
 * cie is the 4 x 95 table of pltb_scene_desc (its Y row defines luminance()). out = { eval[4], pdf } */
void plt_oracle_principled_eval_pdf(const pltb_bsdf *b_in, float base_color, const float *cie, const float *wi, const float *wo,
                                    const float *wl, float *out) {
    pltb_texture t{}; t.type = PLTB_TEX_UNIFORM; t.p[0] = base_color;
    pltb_bsdf b = *b_in; b.tex[0] = 0;
    pltb_scene_desc d{}; d.textures = &t; d.n_textures = 1; d.cie_srgb = cie;
    Scene sc; sc.d = &d;
    SI si; si.wi = V3(wi[0], wi[1], wi[2]); si.uv = V2{ 0.f, 0.f };
    for (int i = 0; i < 4; ++i) si.wavelengths[i] = wl[i];
    V3 w(wo[0], wo[1], wo[2]);
    S4 e = principled_eval(sc, b, si, w);
    for (int i = 0; i < 4; ++i) out[i] = e[i];
    out[4] = principled_pdf(b, si, w);
}

void plt_oracle_hier2d_sample(const float *block, const float *u, float *out) { /* out: x, y, pdf */
    V2 o; float pdf; Hier2D(block).sample(V2{ u[0], u[1] }, o, pdf); out[0] = o.x; out[1] = o.y; out[2] = pdf;
}
/* envmap consistency probe (the reference's own check, src/emitters/tests/test_envmap.py:43-70): for a uniform sample u,
 * out = { ds.pdf from sample_direction, pdf_direction(ds.d), eval(si.wi = -ds.d)[0], eval_direction(ds.uv)[0], d.x, d.y, d.z } */
void plt_oracle_envmap_probe(const pltb_scene_desc *d, const float *u, const float *wl4, float *out) {
    Scene sc; sc.d = d;
    const pltb_emitter *emp = nullptr;
    for (uint32_t i = 0; i < d->n_emitters; ++i) if (d->emitters[i].type == PLTB_EMITTER_ENVMAP) emp = &d->emitters[i];
    if (!emp) { out[0] = -1.f; return; }
    const pltb_emitter &em = *emp;
    S4 wl; for (int i = 0; i < 4; ++i) wl[i] = wl4[i];
    Hier2D warp(d->tex_data + em.env_warp);
    V2 uv; float pdf;
    warp.sample(V2{ u[0], u[1] }, uv, pdf);
    uv.x += .5f / (float) (em.env_res[0] - 1u);
    float theta = uv.y * kPi, phi = uv.x * (2.f * kPi);
    V3 sd(std::cos(phi) * std::sin(theta), std::sin(phi) * std::sin(theta), std::cos(theta));
    V3 dl(sd.y, sd.z, -sd.x);
    V3 dw = xf_vector(em.to_world, dl);
    out[0] = pdf > 0.f ? pdf * envmap_inv_sin_theta(dl) * (1.f / (2.f * sqr(kPi))) : 0.f;
    out[1] = envmap_pdf_direction(sc, em, dw);
    out[2] = envmap_eval_spectrum(sc, em, envmap_dir_to_uv(xf_vector(em.to_local, dw)), wl)[0];
    out[3] = envmap_eval_spectrum(sc, em, uv, wl)[0];
    out[4] = dw.x; out[5] = dw.y; out[6] = dw.z;
}
/* ---- plugin-level probes for the reference's own plugin tests (tests/test_oracle_ref_tests.py) ---- */
/* BSDF of shape `shape`: out = { eval[4], pdf, weight[4], sample: wo[3], pdf, weight[4], eta } (src/render/bsdf.cpp:22-97 entry points) */
void plt_oracle_bsdf_probe(const pltb_scene_desc *d, int32_t shape, const float *wi, const float *uv, const float *wl4, const float *wo,
                           const float *u5, float *out) {
    Scene sc; prepare_scene(sc, d);
    SI si; si.t = 1.f; si.shape = shape; si.wi = V3(wi[0], wi[1], wi[2]); si.uv = V2{ uv[0], uv[1] };
    si.n = V3(0.f, 0.f, 1.f); si.sh.n = si.n; si.sh.s = V3(1.f, 0.f, 0.f); si.sh.t = V3(0.f, 1.f, 0.f);
    for (int k = 0; k < 4; ++k) si.wavelengths[k] = wl4[k];
    const V3 w(wo[0], wo[1], wo[2]);
    const S4 e = wbsdf_eval(sc, si, w), wt = wbsdf_weight(sc, si, w);
    for (int k = 0; k < 4; ++k) { out[k] = e[k]; out[5 + k] = wt[k]; }
    out[4] = wbsdf_pdf(sc, si, w);
    const SampleOut so = wbsdf_sample(sc, si, u5[0], V2{ u5[1], u5[2] }, V2{ u5[3], u5[4] });
    out[9] = so.bs.wo.x; out[10] = so.bs.wo.y; out[11] = so.bs.wo.z; out[12] = so.bs.pdf;
    for (int k = 0; k < 4; ++k) out[13 + k] = so.weight[k];
    out[17] = so.bs.eta;
}
/* batched forms of the probe for statistical tests (the reference's chi^2 tests, src/bsdfs/tests/test_principled.py:60-121):
 * sample: u = n x {sample1, sample2.x, sample2.y} -> out = n x {wo[3], pdf, weight[0]};  pdf: wo = n x 3 -> out = n x {pdf, eval[0]} */
static SI probe_si(int32_t shape, const float *wi, const float *uv, const float *wl4) {
    SI si; si.t = 1.f; si.shape = shape; si.wi = V3(wi[0], wi[1], wi[2]); si.uv = V2{ uv[0], uv[1] };
    si.n = V3(0.f, 0.f, 1.f); si.sh.n = si.n; si.sh.s = V3(1.f, 0.f, 0.f); si.sh.t = V3(0.f, 1.f, 0.f);
    for (int k = 0; k < 4; ++k) si.wavelengths[k] = wl4[k];
    return si;
}
void plt_oracle_bsdf_sample_batch(const pltb_scene_desc *d, int32_t shape, const float *wi, const float *uv, const float *wl4, uint64_t n,
                                  const float *u, float *out) {
    Scene sc; prepare_scene(sc, d);
    const SI si = probe_si(shape, wi, uv, wl4);
    for (uint64_t i = 0; i < n; ++i) {
        const SampleOut so = wbsdf_sample(sc, si, u[3 * i], V2{ u[3 * i + 1], u[3 * i + 2] }, V2{ 0.5f, 0.5f });
        out[5 * i] = so.bs.wo.x; out[5 * i + 1] = so.bs.wo.y; out[5 * i + 2] = so.bs.wo.z; out[5 * i + 3] = so.bs.pdf; out[5 * i + 4] = so.weight[0];
    }
}
void plt_oracle_bsdf_pdf_batch(const pltb_scene_desc *d, int32_t shape, const float *wi, const float *uv, const float *wl4, uint64_t n,
                               const float *wo, float *out) {
    Scene sc; prepare_scene(sc, d);
    const SI si = probe_si(shape, wi, uv, wl4);
    for (uint64_t i = 0; i < n; ++i) {
        const V3 w(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]);
        out[2 * i] = wbsdf_pdf(sc, si, w); out[2 * i + 1] = wbsdf_eval(sc, si, w)[0];
    }
}
/* Emitter sampling from a reference point (Scene::sample_emitter_direction without the visibility test, area.cpp:116-208):
 * out = { ds.d[3], ds.pdf, ds.dist, ds.p[3], ds.n[3], pdf_direction(ds), eval_direction[4], eval(wi = -d)[4], eval(wi = +d)[4] } */
void plt_oracle_area_emitter_probe(const pltb_scene_desc *d, uint32_t emitter, const float *p, const float *u, const float *wl4, float *out) {
    Scene sc; prepare_scene(sc, d);
    const pltb_emitter &em = d->emitters[emitter];
    const pltb_shape &sh = d->shapes[em.shape];
    S4 wl; for (int k = 0; k < 4; ++k) wl[k] = wl4[k];
    const V3 ref(p[0], p[1], p[2]);
    DirectionSample ds = shape_sample_direction(sc, sh, ref, V2{ u[0], u[1] });
    out[0] = ds.d.x; out[1] = ds.d.y; out[2] = ds.d.z; out[3] = ds.pdf; out[4] = ds.dist;
    out[5] = ds.p.x; out[6] = ds.p.y; out[7] = ds.p.z; out[8] = ds.n.x; out[9] = ds.n.y; out[10] = ds.n.z;
    out[11] = shape_pdf_direction(sh, ref, ds);
    const S4 rad = emitter_radiance(sc, em, wl, ds.uv);
    for (int k = 0; k < 4; ++k) {
        out[12 + k] = dot(ds.d, ds.n) < 0.f ? rad[k] : 0.f;   /* AreaLight::eval_direction, area.cpp:198-208 */
        out[16 + k] = rad[k];                                  /* AreaLight::eval with cos_theta(si.wi) > 0, area.cpp:83-91 */
        out[20 + k] = 0.f;                                     /* ... and with si.wi pointing into the shape */
    }
}
/* Sensor::sample_ray for film position `pos` in [0,1]^2: out = { o[3], d[3] } */
void plt_oracle_sensor_ray(const pltb_scene_desc *d, uint32_t sensor, const float *pos, float *out) {
    const Ray r = sensor_ray(d->sensors[sensor], V2{ pos[0], pos[1] });
    out[0] = r.o.x; out[1] = r.o.y; out[2] = r.o.z; out[3] = r.d.x; out[4] = r.d.y; out[5] = r.d.z;
}
/* Texture::eval of spectrum / texture `id` */
void plt_oracle_tex_probe(const pltb_scene_desc *d, int32_t id, const float *wl4, const float *uv, float *out) {
    Scene sc; prepare_scene(sc, d);
    S4 wl; for (int k = 0; k < 4; ++k) wl[k] = wl4[k];
    const S4 v = tex_eval(sc, id, wl, V2{ uv[0], uv[1] });
    for (int k = 0; k < 4; ++k) out[k] = v[k];
}
/* ---- polarized building blocks and further header-level functions (pinned on tests/golden/ref_kat_pol.json) ---- */
static void m4_store(const M4 &m, float *out) { for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) for (int k = 0; k < 4; ++k) out[16 * i + 4 * j + k] = m.e[i][j][k]; }
static M4 m4_load(const float *in) { M4 m; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) for (int k = 0; k < 4; ++k) m.e[i][j][k] = in[16 * i + 4 * j + k]; return m; }
void plt_oracle_array_div(const float *v3, const float *s4v, float s, float *out) {
    V3 v = V3(v3[0], v3[1], v3[2]) / s; S4 sp; for (int k = 0; k < 4; ++k) sp[k] = s4v[k]; sp = sp / s;
    out[0] = v.x; out[1] = v.y; out[2] = v.z; for (int k = 0; k < 4; ++k) out[3 + k] = sp[k];
}
void plt_oracle_m4_ops(const float *a, const float *b, const float *s4v, float f, float *out) {   /* out: 8 matrices of 64 floats */
    M4 A = m4_load(a), B = m4_load(b); S4 s; for (int k = 0; k < 4; ++k) s[k] = s4v[k];
    m4_store(A * B, out); m4_store(m4_diag(s), out + 64); m4_store(m4_scale(A, s), out + 128); m4_store(m4_div(A, f), out + 192);
    m4_store(m4_depolarizer(s), out + 256); m4_store(A * m4_absorber(s), out + 320); m4_store(m4_transpose(A), out + 384);
    m4_store(m4_select_gt_one(m4_scale(A, 3.f)), out + 448);
}
void plt_oracle_fresnel_polarized_real(float c, float eta, float *out) {
    FresnelPol f = fresnel_polarized(c, eta);
    out[0] = f.a_s.re; out[1] = f.a_s.im; out[2] = f.a_p.re; out[3] = f.a_p.im; out[4] = f.cos_theta_t; out[5] = f.eta_it; out[6] = f.eta_ti;
}
void plt_oracle_fresnel_polarized_complex(float c, float eta, float k, float *out) {
    FresnelPol f = fresnel_polarized(c, cx(eta, k));
    out[0] = f.a_s.re; out[1] = f.a_s.im; out[2] = f.a_p.re; out[3] = f.a_p.im; out[4] = f.cos_theta_t;
}
void plt_oracle_specular_reflection_complex(float c, float eta, float k, float *out) { m4_store(specular_reflection(S4(c), S4(eta), S4(k)), out); }
void plt_oracle_specular_reflection_real(float c, float eta, float *out) { m4_store(specular_reflection(c, eta), out); }
void plt_oracle_specular_transmission(float c, float eta, float *out) { m4_store(specular_transmission(c, eta), out); }
void plt_oracle_stokes_basis(const float *d, float *out) { V3 b = stokes_basis(V3(d[0], d[1], d[2])); out[0] = b.x; out[1] = b.y; out[2] = b.z; }
void plt_oracle_rotate_stokes_basis(const float *f, const float *c, const float *t, float *out) {
    m4_store(rotate_stokes_basis(V3(f[0], f[1], f[2]), V3(c[0], c[1], c[2]), V3(t[0], t[1], t[2])), out);
}
void plt_oracle_align_stokes_frames(const float *m, const float *n, const float *wo_hat, const float *wi_hat, float *out) {
    m4_store(align_stokes_frames(m4_load(m), V3(n[0], n[1], n[2]), V3(wo_hat[0], wo_hat[1], wo_hat[2]), V3(wi_hat[0], wi_hat[1], wi_hat[2])), out);
}
float plt_oracle_pdf_rgb_spectrum(float wl) { return pdf_rgb_spectrum(wl); }
void plt_oracle_warp(const float *u, float *out) {   /* disk_concentric[2] cosine_hemisphere[3] uniform_sphere[3] uniform_triangle[2] */
    V2 s{ u[0], u[1] };
    V2 d = square_to_uniform_disk_concentric(s); V3 c = square_to_cosine_hemisphere(s), sp = square_to_uniform_sphere(s); V2 t = square_to_uniform_triangle(s);
    out[0] = d.x; out[1] = d.y; out[2] = c.x; out[3] = c.y; out[4] = c.z; out[5] = sp.x; out[6] = sp.y; out[7] = sp.z; out[8] = t.x; out[9] = t.y;
}
void plt_oracle_coordinate_system(const float *n, float *out) {
    V3 s, t; coordinate_system(V3(n[0], n[1], n[2]), s, t);
    out[0] = s.x; out[1] = s.y; out[2] = s.z; out[3] = t.x; out[4] = t.y; out[5] = t.z;
}
float plt_oracle_unit_angle(const float *a, const float *b) { return unit_angle(V3(a[0], a[1], a[2]), V3(b[0], b[1], b[2])); }
float plt_oracle_inv_coherence_det(float k, float diffusivity, float opl) { return inv_coherence_det(k, diffusivity, opl); }

float plt_oracle_rfilter_eval(uint32_t type, float radius, float stddev, float x) {
    pltb_film f{}; f.rfilter = type; f.radius = radius; f.stddev = stddev;
    return rfilter_eval(f, x);
}
/* ImageBlock::put of one sample of weight 1 into a W x H single-channel block without border (imageblock.cpp:429-500) */
void plt_oracle_splat(uint32_t type, float radius, float stddev, uint32_t W, uint32_t H, const float *pos, float *film) {
    pltb_film f{}; f.rfilter = type; f.radius = radius; f.stddev = stddev;
    const float one = 1.f;
    splat(f, W, H, film, V2{ pos[0], pos[1] }, &one, 1);
}
float plt_oracle_hier2d_eval(const float *block, const float *pos) { return Hier2D(block).eval(V2{ pos[0], pos[1] }); }

} // extern "C"
