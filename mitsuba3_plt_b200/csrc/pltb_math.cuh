// pltb_math.cuh -- scalar/vector math of the plt path under the numerics contract:
// float32, no implicit contraction (-fmad=false), explicit fmaf exactly where the reference writes dr::fmadd & co.
// Compiles for the device (nvcc) and, for host-side emulation in tests, as plain C++ (g++ -ffp-contract=off).
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#  define PLTB_HD __host__ __device__ __forceinline__
#  define PLTB_D __device__ __forceinline__
#  define PLTB_NI inline __host__ __device__ __noinline__   /* big, multiply-used bodies: one copy per kernel keeps the I-cache footprint down */
#else
#  define PLTB_HD inline
#  define PLTB_D inline
#  define PLTB_NI inline
#endif

// -DPLTB_DEBUG_CHECKS: device-side assertions on the invariants the kernels rely on (traversal stack bound, queue routing, record
// indices).  compute-sanitizer is closed on the GPU pool this was developed on, so `tools/gpu.sh <tag> debugchecks` runs the GPU
// test-suite against this build instead (profiles/r02_debugchecks_7e.txt).
#if defined(PLTB_DEBUG_CHECKS) && defined(__CUDA_ARCH__)
#  include <assert.h>
#  define PLTB_ASSERT(x) assert(x)
#else
#  define PLTB_ASSERT(x) ((void) 0)
#endif

namespace pltb {

constexpr float kPi = 3.14159265358979323846f;
constexpr float kInvPi = 0.31830988618379067154f;
constexpr float kInvFourPi = 0.07957747154594766788f;
constexpr float kTwoPi = 6.28318530717958647692f;
constexpr float kFourPi = 12.56637061435917295385f;
constexpr float kInvSqrtPi = 0.56418958354775628695f;
constexpr float kEps = 5.9604644775390625e-8f;  // dr::Epsilon<float>
constexpr float kRayEps = kEps * 1500.f;        // include/mitsuba/core/math.h:20
constexpr float kShadowEps = kRayEps * 10.f;    // math.h:22
constexpr float kLargest = 3.402823466e+38f;

PLTB_HD float inf_f() { return __builtin_huge_valf(); }
PLTB_HD float fmadd(float a, float b, float c) { return fmaf(a, b, c); }
PLTB_HD float fmsub(float a, float b, float c) { return fmaf(a, b, -c); }
PLTB_HD float fnmadd(float a, float b, float c) { return fmaf(-a, b, c); }
PLTB_HD float sqr(float x) { return x * x; }
#if defined(__CUDA_ARCH__) && defined(PLTB_FAST_RCP)
PLTB_HD float rcp(float x) { return __frcp_rn(x); }   // == 1.f / x (correctly rounded) without the division slow path
#else
PLTB_HD float rcp(float x) { return 1.f / x; }
#endif
// dr::rsqrt.  Geometry and BSDF code (normalisations, microfacet sampling, the grating direction) gets the value of Dr.Jit's
// scalar fallback 1 / sqrt(x) -- two correctly rounded IEEE operations, spelled so that no division subroutine is called: a
// 2-ulp MUFU.RSQ there moves diffraction / microfacet branches and costs 5-10x more divergent samples than everything else
// together (profiles/r02_ab_numerics.txt).  Smooth colour arithmetic (srgb_model_eval: 16 calls per bitmap lookup) may use
// MUFU.RSQ (-DPLTB_FAST_RSQRT, the default; what Dr.Jit's CUDA backend emits for dr::rsqrt everywhere): no branch depends on it.
#if defined(__CUDA_ARCH__) && defined(PLTB_FAST_RSQRT_ALL)
PLTB_HD float rsqrt_(float x) { return rsqrtf(x); }
#elif defined(__CUDA_ARCH__)
PLTB_HD float rsqrt_(float x) { return __frcp_rn(__fsqrt_rn(x)); }
#else
PLTB_HD float rsqrt_(float x) { return 1.f / sqrtf(x); }
#endif
#if defined(__CUDA_ARCH__) && (defined(PLTB_FAST_RSQRT) || defined(PLTB_FAST_RSQRT_ALL))
PLTB_HD float rsqrt_smooth(float x) { return rsqrtf(x); }
#else
PLTB_HD float rsqrt_smooth(float x) { return rsqrt_(x); }
#endif
PLTB_HD float safe_sqrt(float x) { return sqrtf(fmaxf(x, 0.f)); }
PLTB_HD uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
PLTB_HD float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
PLTB_HD bool signbit_(float b) { return (f2u(b) >> 31) != 0; }
PLTB_HD float mulsign(float a, float b) { return u2f(f2u(a) ^ (f2u(b) & 0x80000000u)); }
PLTB_HD float mulsign_neg(float a, float b) { return u2f(f2u(a) ^ (~f2u(b) & 0x80000000u)); }
PLTB_HD float sign1(float v) { return u2f(0x3f800000u | (f2u(v) & 0x80000000u)); }
PLTB_HD uint32_t umin_(uint32_t a, uint32_t b) { return a < b ? a : b; }
PLTB_HD uint32_t umax_(uint32_t a, uint32_t b) { return a > b ? a : b; }
PLTB_HD bool finite_(float x) { return (f2u(x) & 0x7f800000u) != 0x7f800000u; }
PLTB_HD bool isnan_(float x) { return x != x; }
PLTB_HD bool isinf_(float x) { return (f2u(x) & 0x7fffffffu) == 0x7f800000u; }

struct V2 { float x, y; };
struct V3 { float x, y, z; };
PLTB_HD V2 mk2(float x, float y) { V2 r; r.x = x; r.y = y; return r; }
PLTB_HD V3 mk3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
PLTB_HD V3 operator+(V3 a, V3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
PLTB_HD V3 operator-(V3 a, V3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
PLTB_HD V3 operator-(V3 a) { return mk3(-a.x, -a.y, -a.z); }
PLTB_HD V3 operator*(V3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
// Dr.Jit: array / scalar is array * rcp(scalar) (ext/drjit/include/drjit/array_router.h:218-232)
PLTB_HD V3 operator/(V3 a, float s) { const float r = rcp(s); return mk3(a.x * r, a.y * r, a.z * r); }
PLTB_HD float dot(V3 a, V3 b) { return fmadd(a.z, b.z, fmadd(a.y, b.y, a.x * b.x)); }
PLTB_HD float sqnorm(V3 a) { return fmadd(a.z, a.z, fmadd(a.y, a.y, a.x * a.x)); }
PLTB_HD float norm(V3 a) { return sqrtf(sqnorm(a)); }
PLTB_HD V3 normalize(V3 a) { return a * rsqrt_(sqnorm(a)); }
PLTB_HD V3 cross(V3 a, V3 b) {
    return mk3(fmsub(a.y, b.z, a.z * b.y), fmsub(a.z, b.x, a.x * b.z), fmsub(a.x, b.y, a.y * b.x));
}
PLTB_HD V3 fmadd3(V3 a, float s, V3 c) { return mk3(fmadd(a.x, s, c.x), fmadd(a.y, s, c.y), fmadd(a.z, s, c.z)); }

// 4-wavelength spectrum
struct S4 { float v[4]; };
PLTB_HD S4 s4(float a) { S4 r; r.v[0] = r.v[1] = r.v[2] = r.v[3] = a; return r; }
PLTB_HD S4 operator*(S4 a, S4 b) { S4 r; for (int i = 0; i < 4; ++i) r.v[i] = a.v[i] * b.v[i]; return r; }
PLTB_HD S4 operator*(S4 a, float b) { S4 r; for (int i = 0; i < 4; ++i) r.v[i] = a.v[i] * b; return r; }
PLTB_HD S4 operator+(S4 a, S4 b) { S4 r; for (int i = 0; i < 4; ++i) r.v[i] = a.v[i] + b.v[i]; return r; }
PLTB_HD S4 operator/(S4 a, float b) { S4 r; const float rb = rcp(b); for (int i = 0; i < 4; ++i) r.v[i] = a.v[i] * rb; return r; }
PLTB_HD float smax(S4 a) { return fmaxf(fmaxf(a.v[0], a.v[1]), fmaxf(a.v[2], a.v[3])); }

// include/mitsuba/core/vector.h:116-136
PLTB_HD void coordinate_system(V3 n, V3 &s, V3 &t) {
    float sign = sign1(n.z), a = -rcp(sign + n.z), b = n.x * n.y * a;
    s = mk3(mulsign(sqr(n.x) * a, n.z) + 1.f, mulsign(b, n.z), mulsign_neg(n.x, n.z));
    t = mk3(b, fmadd(n.y, n.y * a, sign), -n.y);
}

struct Frame {
    V3 s, t, n;
    PLTB_HD V3 to_local(V3 v) const { return mk3(dot(v, s), dot(v, t), dot(v, n)); }
    PLTB_HD V3 to_world(V3 v) const { return fmadd3(n, v.z, fmadd3(t, v.y, s * v.x)); }
};
PLTB_HD Frame frame_from_normal(V3 n) { Frame f; f.n = n; coordinate_system(n, f.s, f.t); return f; }

// ---- RNG: TEA (include/mitsuba/core/random.h:77-90) + PCG32 (ext/drjit/include/drjit/random.h:57-138) --------------
PLTB_HD void sample_tea_32(uint32_t v0, uint32_t v1, uint32_t &o0, uint32_t &o1) {
    uint32_t sum = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        sum += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    o0 = v0; o1 = v1;
}

struct PCG32 {
    uint64_t state, inc;
    PLTB_HD uint32_t next_uint32() {
        uint64_t old = state;
        state = old * 0x5851f42d4c957f2dULL + inc;
        uint32_t xorshifted = (uint32_t) (((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t) (old >> 59u);
        return (xorshifted >> rot) | (xorshifted << ((0u - rot) & 31u));
    }
    PLTB_HD void seed(uint64_t initstate, uint64_t initseq) {
        state = 0; inc = (initseq << 1) | 1u;
        next_uint32(); state += initstate; next_uint32();
    }
    PLTB_HD float next_float32() { return u2f((next_uint32() >> 9) | 0x3f800000u) - 1.f; }
};

// ---- special functions ------------------------------------------------------------------------------------------------
// drjit::detail::estrin_impl (ext/drjit/include/drjit/math.h:26-41), unrolled for the two polynomial sizes used
PLTB_HD float estrin6(float x, const float *c) {
    float x2 = x * x, x4 = x2 * x2;
    float a = fmadd(x, c[1], c[0]), b = fmadd(x, c[3], c[2]), d = fmadd(x, c[5], c[4]);
    float e = fmadd(x2, b, a);
    return fmadd(x4, d, e);
}
PLTB_HD float estrin7(float x, const float *c) {
    float x2 = x * x, x4 = x2 * x2;
    float a = fmadd(x, c[1], c[0]), b = fmadd(x, c[3], c[2]), d = fmadd(x, c[5], c[4]), g = c[6];
    float e = fmadd(x2, b, a), f = fmadd(x2, g, d);
    return fmadd(x4, f, e);
}
PLTB_HD float estrin9(float x, const float *c) {
    float x2 = x * x, x4 = x2 * x2, x8 = x4 * x4;
    float a = fmadd(x, c[1], c[0]), b = fmadd(x, c[3], c[2]), d = fmadd(x, c[5], c[4]), g = fmadd(x, c[7], c[6]), h = c[8];
    float e = fmadd(x2, b, a), f = fmadd(x2, g, d);
    float k = fmadd(x4, f, e);
    return fmadd(x8, h, k);
}

// drjit::erf float path (math.h:1497-1553)
PLTB_HD float dr_erf(float x) {
    const float c0c[6] = { 0x1.20dd74p+0f, -0x1.812672p-2f, 0x1.ce0934p-4f, -0x1.b5a334p-6f, 0x1.4246b4p-8f, -0x1.273facp-11f };
    const float c1c[7] = { -0x1.a0d71ap+0f, -0x1.d51e3ap-1f, -0x1.3a904cp-3f, 0x1.1c395cp-5f,
                           -0x1.6856bep-8f, 0x1.180f1ep-11f, -0x1.8ca9f6p-16f };
    float xa = fabsf(x), x2 = x * x;
    float c0 = estrin6(x2, c0c), c1 = estrin7(xa, c1c);
    float xb = 1.f - exp2f(c1 * xa);
    if (xa < 1.f) return x * c0;
    return copysignf(finite_(xb) ? xb : 1.f, x);
}
// drjit::erfinv (math.h:1557-1580)
PLTB_HD float dr_erfinv(float x) {
    const float p1c[9] = { 1.50140941f, 0.246640727f, -0.00417768164f, -0.00125372503f, 0.00021858087f,
                           -4.39150654e-06f, -3.5233877e-06f, 3.43273939e-07f, 2.81022636e-08f };
    const float p2c[9] = { 2.83297682f, 1.00167406f, 0.00943887047f, -0.0076224613f, 0.00573950773f,
                           -0.00367342844f, 0.00134934322f, 0.000100950558f, -0.000200214257f };
    float w = -logf((1.f - x) * (1.f + x));
    float w1 = w - 2.5f, w2 = sqrtf(w) - 3.f;
    float p1 = estrin9(w1, p1c), p2 = estrin9(w2, p2c);
    return (w < 5.f ? p1 : p2) * x;
}

// math::bessel_j (include/mitsuba/core/math.h:204-347).  The 20 Simpson nodes t_i = h*i and sin(t_i) do not depend
// on the argument; they are tabulated once on the host with the same float ops (see pltb_bessel_tables()).
// `tab` (device memory, nullptr in host builds): math::bessel_j(x, l) for l = 1..kBesselTabOrders on [0, kBesselTabMax) as cubic
// Hermite segments {f(x_i), h f'(x_i), f(x_i+1), h f'(x_i+1)} with h = 1 / kBesselTabInvH, evaluated in double precision from the
// same 20-node rule (blend with the asymptotic form included).  One 16-byte load and 10 flops replace 20 cosines.  It stands in
// for the quadrature ONLY where the value does not scale radiance: DiffractionGrating::sample_lobe's selection weights and the
// sampled lobe's intensity in RoughGrating::wbsdf_sample, which enters the sample-phase throughput (Russian roulette, the
// `alpha != 0` test) but never the replayed one.  |table - float32 rule| <~ 2e-6 absolute (the rule's own float32 rounding).
#ifndef PLTB_BESSEL_INVH
#define PLTB_BESSEL_INVH 16   // segments per unit of x: interpolation error (h^4 / 384) |f''''| <= 4e-8; 64 KB, L1-resident
#endif
constexpr int kBesselTabOrders = 4, kBesselTabInvH = PLTB_BESSEL_INVH, kBesselTabN = 64 * PLTB_BESSEL_INVH;
constexpr float kBesselTabMax = (float) kBesselTabN / (float) kBesselTabInvH;   // 64
#if !defined(__CUDACC__)
struct pltb_host_float4;                      // host builds (tests/hostcheck) never have a table
typedef pltb_host_float4 pltb_tab4;
#else
typedef float4 pltb_tab4;
#endif
struct BesselTables { float t[20]; float sin_t[20]; float h_over_3; const pltb_tab4 *tab; };

PLTB_HD float bessel_j_small(const BesselTables &bt, float x, int nu) {
    float fnu = (float) nu;
    float I_odd = 0.f, I_even = 0.f;
#pragma unroll 1
    for (int i = 1; i < 19; i += 2) I_odd += cosf(fnu * bt.t[i] - x * bt.sin_t[i]);
#pragma unroll 1
    for (int i = 2; i < 19; i += 2) I_even += cosf(fnu * bt.t[i] - x * bt.sin_t[i]);
    float fa = cosf(fnu * bt.t[0] - x * bt.sin_t[0]), fb = cosf(fnu * bt.t[19] - x * bt.sin_t[19]);
    return kInvPi * (bt.h_over_3 * (fa + 4.0f * I_odd + 2.0f * I_even + fb));
}
PLTB_HD float bessel_j_asymp(float x, int nu) {
    const float sign = ((x >= 0) || (nu % 2 != 0)) ? 1.0f : -1.0f;
    const float x_abs = fabsf(x);
    if (x_abs > 10 * kEps)
        return sign * sqrtf(2.0f / (kPi * x_abs)) * cosf(x_abs - kPi * (float) nu / 2.0f - kPi / 4.0f);
    return nu == 0 ? 1.0f : 0.0f;
}
PLTB_NI float bessel_j(const BesselTables &bt, float x, int nu) {
    int nua = nu < 0 ? -nu : nu;
    const float nu_abs = (float) nua, nsign = nu < 0 ? -1.f : 1.f;
    float center = 45.f + nu_abs;
    float weight_fn = 1.0f / (1.0f + expf(-1.5f * (fabsf(x) - center)));
    return nsign * ((1.0f - weight_fn) * bessel_j_small(bt, x, nua) + weight_fn * bessel_j_asymp(x, nua));
}
// J_nu1(x) and J_nu2(x) in ONE pass over the Simpson nodes: the same operations per value as two bessel_j calls (bit-identical
// results), but the two independent cosf chains overlap and a warp whose lanes need different orders makes one trip through the
// loop instead of two half-empty ones (RoughGrating::wbsdf_sample evaluates lobe_intensity((lx, ly)) with lx, ly in -3..3).
PLTB_NI void bessel_j_pair(const BesselTables &bt, float x, int nu1, int nu2, float &j1, float &j2) {
    const int n1 = nu1 < 0 ? -nu1 : nu1, n2 = nu2 < 0 ? -nu2 : nu2;
    const float f1 = (float) n1, f2 = (float) n2;
    float o1 = 0.f, e1 = 0.f, o2 = 0.f, e2 = 0.f;
#pragma unroll 1
    for (int i = 1; i < 19; i += 2) { o1 += cosf(f1 * bt.t[i] - x * bt.sin_t[i]); o2 += cosf(f2 * bt.t[i] - x * bt.sin_t[i]); }
#pragma unroll 1
    for (int i = 2; i < 19; i += 2) { e1 += cosf(f1 * bt.t[i] - x * bt.sin_t[i]); e2 += cosf(f2 * bt.t[i] - x * bt.sin_t[i]); }
    const float a1 = cosf(f1 * bt.t[0] - x * bt.sin_t[0]), b1 = cosf(f1 * bt.t[19] - x * bt.sin_t[19]);
    const float a2 = cosf(f2 * bt.t[0] - x * bt.sin_t[0]), b2 = cosf(f2 * bt.t[19] - x * bt.sin_t[19]);
    const float s1 = kInvPi * (bt.h_over_3 * (a1 + 4.0f * o1 + 2.0f * e1 + b1)), s2 = kInvPi * (bt.h_over_3 * (a2 + 4.0f * o2 + 2.0f * e2 + b2));
    const float w1 = 1.0f / (1.0f + expf(-1.5f * (fabsf(x) - (45.f + f1)))), w2 = 1.0f / (1.0f + expf(-1.5f * (fabsf(x) - (45.f + f2))));
    j1 = (nu1 < 0 ? -1.f : 1.f) * ((1.0f - w1) * s1 + w1 * bessel_j_asymp(x, n1));
    j2 = (nu2 < 0 ? -1.f : 1.f) * ((1.0f - w2) * s2 + w2 * bessel_j_asymp(x, n2));
}
// Reduced-precision J_nu for *sampling decisions only* (DiffractionGrating::sample_lobe: which lobe is drawn and its discrete
// pdf): the Simpson terms use MUFU.COS after a two-constant reduction to [-pi, pi] (absolute error ~5e-7 per term, ~1e-7 in
// J).  Everything that scales radiance (lobe_intensity in wbsdf_sample / wbsdf_eval) keeps the precise cosf above: with the
// nanometre-unit call of roughgrating.cpp:440 (N2) J_1 is ~4e-4, where 1e-7 would already be a 2e-4 relative error.
PLTB_HD float cos_fast(float a) {
#if defined(__CUDA_ARCH__)
    const float k = rintf(a * 0.15915494309189535f);
    float r = fmaf(-k, 6.2831854820251465f, a);
    r = fmaf(-k, -1.7484555e-7f, r);
    return __cosf(r);
#else
    return cosf(a);
#endif
}
PLTB_NI float bessel_j_fast(const BesselTables &bt, float x, int nu) {   // nu >= 0
    const float fnu = (float) nu;
    float I_odd = 0.f, I_even = 0.f;
#pragma unroll 1
    for (int i = 1; i < 19; i += 2) {
        I_odd += cos_fast(fnu * bt.t[i] - x * bt.sin_t[i]);
        I_even += cos_fast(fnu * bt.t[i + 1] - x * bt.sin_t[i + 1]);
    }
    const float fa = cos_fast(fnu * bt.t[0] - x * bt.sin_t[0]), fb = cos_fast(fnu * bt.t[19] - x * bt.sin_t[19]);
    const float small_ = kInvPi * (bt.h_over_3 * (fa + 4.0f * I_odd + 2.0f * I_even + fb));
    const float weight_fn = 1.0f / (1.0f + expf(-1.5f * (fabsf(x) - (45.f + fnu))));
    return (1.0f - weight_fn) * small_ + weight_fn * bessel_j_asymp(x, nu);
}
// table lookup of bessel_j(x, nu) (see BesselTables::tab); false when the argument is outside the table or there is none
PLTB_HD bool bessel_j_tab(const BesselTables &bt, float x, int nu, float &j) {
#if defined(__CUDA_ARCH__) && !defined(PLTB_NO_BESSEL_TABLE)
    const int n = nu < 0 ? -nu : nu;
    if (bt.tab == nullptr || n < 1 || n > kBesselTabOrders || !(x >= 0.f && x < kBesselTabMax)) return false;
    const float u = x * (float) kBesselTabInvH;
    int i = (int) u;
    if (i > kBesselTabN - 1) i = kBesselTabN - 1;
    const float s = u - (float) i, s2 = s * s, s3 = s2 * s;
    const float4 c = __ldg(bt.tab + (n - 1) * kBesselTabN + i);
    const float h00 = 2.f * s3 - 3.f * s2 + 1.f, h10 = s3 - 2.f * s2 + s, h01 = 3.f * s2 - 2.f * s3, h11 = s3 - s2;
    const float v = h00 * c.x + h10 * c.y + h01 * c.z + h11 * c.w;
    j = nu < 0 ? -v : v;   // bessel_j: sign(nu) * (value at |nu|)
    return true;
#else
    (void) bt; (void) x; (void) nu; (void) j;
    return false;
#endif
}
PLTB_HD float sinc_(float x) { float xx = fabsf(x); return xx < 1e-5f ? 1.0f : sinf(xx) / xx; }

// ext/drjit/include/drjit/sphere.h:53-60
PLTB_HD float unit_angle(V3 a, V3 b) {
    float dot_uv = dot(a, b);
    V3 am = mk3(mulsign(a.x, dot_uv), mulsign(a.y, dot_uv), mulsign(a.z, dot_uv));
    float temp = 2.f * asinf(.5f * norm(b - am));
    return dot_uv >= 0 ? temp : kPi - temp;
}
// src/python/python/ad/integrators/common.py:1304-1312
PLTB_HD float mis_weight(float pdf_a, float pdf_b) {
    float a2 = sqr(pdf_a), b2 = sqr(pdf_b), w = a2 / (a2 + b2);
    return finite_(w) ? w : 0.f;
}

// ---- warps (include/mitsuba/core/warp.h) ---------------------------------------------------------------------------------
PLTB_HD V2 square_to_uniform_disk_concentric(V2 sample) {
    float x = fmsub(2.f, sample.x, 1.f), y = fmsub(2.f, sample.y, 1.f);
    bool is_zero = (x == 0.f) && (y == 0.f), q13 = fabsf(x) < fabsf(y);
    float r = q13 ? y : x, rp = q13 ? x : y;
    float phi = 0.25f * kPi * rp / r;
    if (q13) phi = 0.5f * kPi - phi;
    if (is_zero) phi = 0.f;
    float s = sinf(phi), c = cosf(phi);
    return mk2(r * c, r * s);
}
PLTB_HD V3 square_to_cosine_hemisphere(V2 sample) {
    V2 p = square_to_uniform_disk_concentric(sample);
    float z = safe_sqrt(1.f - fmadd(p.y, p.y, p.x * p.x));
    return mk3(p.x, p.y, z);
}
PLTB_HD V2 square_to_uniform_triangle(V2 sample) {
    float t = safe_sqrt(1.f - sample.x);
    return mk2(1.f - t, t * sample.y);
}

// ---- Fresnel (include/mitsuba/render/fresnel.h) --------------------------------------------------------------------------
struct FresnelResult { float r, cos_theta_t, eta_it, eta_ti; };
PLTB_HD FresnelResult fresnel(float cos_theta_i, float eta) {
    bool outside = cos_theta_i >= 0.f;
    float rcp_eta = rcp(eta), eta_it = outside ? eta : rcp_eta, eta_ti = outside ? rcp_eta : eta;
    float cos_theta_t_sqr = fnmadd(fnmadd(cos_theta_i, cos_theta_i, 1.f), eta_ti * eta_ti, 1.f);
    float ci = fabsf(cos_theta_i), ct = safe_sqrt(cos_theta_t_sqr);
    bool index_matched = eta == 1.f, special = index_matched || (ci == 0.f);
    float a_s = fnmadd(eta_it, ct, ci) / fmadd(eta_it, ct, ci);
    float a_p = fnmadd(eta_it, ci, ct) / fmadd(eta_it, ci, ct);
    float r = 0.5f * (sqr(a_s) + sqr(a_p));
    if (special) r = index_matched ? 0.f : 1.f;
    FresnelResult f; f.r = r; f.cos_theta_t = mulsign_neg(ct, cos_theta_i); f.eta_it = eta_it; f.eta_ti = eta_ti;
    return f;
}
PLTB_HD float fresnel_conductor(float cos_theta_i, float eta_r, float eta_i) {
    float c2 = cos_theta_i * cos_theta_i, s2 = 1.f - c2, s4_ = s2 * s2;
    float temp_1 = eta_r * eta_r - eta_i * eta_i - s2,
          a_2_pb_2 = safe_sqrt(temp_1 * temp_1 + 4.f * eta_i * eta_i * eta_r * eta_r),
          a = safe_sqrt(.5f * (a_2_pb_2 + temp_1));
    float term_1 = a_2_pb_2 + c2, term_2 = 2.f * cos_theta_i * a;
    float r_s = (term_1 - term_2) / (term_1 + term_2);
    float term_3 = a_2_pb_2 * c2 + s4_, term_4 = term_2 * s2;
    float r_p = r_s * (term_3 - term_4) / (term_3 + term_4);
    return 0.5f * (r_s + r_p);
}
PLTB_HD V3 reflect(V3 wi) { return mk3(-wi.x, -wi.y, wi.z); }
PLTB_HD V3 reflect(V3 wi, V3 m) {
    float s = 2.f * dot(wi, m);
    return mk3(fmsub(m.x, s, wi.x), fmsub(m.y, s, wi.y), fmsub(m.z, s, wi.z));
}
PLTB_HD V3 refract(V3 wi, float cos_theta_t, float eta_ti) { return mk3(-eta_ti * wi.x, -eta_ti * wi.y, cos_theta_t); }

// ---- transforms (include/mitsuba/core/transform.h:119-158), row-major 4x4 ------------------------------------------------
PLTB_HD V3 xf_point(const float *m, V3 p) {
    return mk3(fmadd(m[2], p.z, fmadd(m[1], p.y, fmadd(m[0], p.x, m[3]))),
               fmadd(m[6], p.z, fmadd(m[5], p.y, fmadd(m[4], p.x, m[7]))),
               fmadd(m[10], p.z, fmadd(m[9], p.y, fmadd(m[8], p.x, m[11]))));
}
PLTB_HD V3 xf_point_proj(const float *m, V3 p) {
    float x = fmadd(m[2], p.z, fmadd(m[1], p.y, fmadd(m[0], p.x, m[3]))),
          y = fmadd(m[6], p.z, fmadd(m[5], p.y, fmadd(m[4], p.x, m[7]))),
          z = fmadd(m[10], p.z, fmadd(m[9], p.y, fmadd(m[8], p.x, m[11]))),
          w = fmadd(m[14], p.z, fmadd(m[13], p.y, fmadd(m[12], p.x, m[15])));
    return mk3(x / w, y / w, z / w);
}
PLTB_HD V3 xf_vector(const float *m, V3 v) {
    return mk3(fmadd(m[2], v.z, fmadd(m[1], v.y, m[0] * v.x)), fmadd(m[6], v.z, fmadd(m[5], v.y, m[4] * v.x)),
               fmadd(m[10], v.z, fmadd(m[9], v.y, m[8] * v.x)));
}

}  // namespace pltb
