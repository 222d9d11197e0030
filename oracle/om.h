// ORACLE (test infrastructure, NOT product code): scalar fp32 math helpers.
//
// Restates the Dr.Jit / Mitsuba scalar semantics the reference's llvm_ad_rgb kernels execute:
//   dot/cross/normalize  : drjit/include/drjit/array_base.h:597-625, array_router.h:651-682
//   coordinate_system    : mitsuba3/include/mitsuba/core/vector.h:118-138
//   Frame to_local/world : mitsuba3/include/mitsuba/core/frame.h:34-41
//   TEA                  : mitsuba3/include/mitsuba/core/random.h:77-90
//   PCG32                : drjit/include/drjit/random.h:60-79,211-213
//   warps                : mitsuba3/include/mitsuba/core/warp.h:54-90,412-436,447-494,516-521
//   erfinv (Giles)       : drjit/include/drjit/math.h:1556-1578
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline may use anything in oracle/.
// Build with -ffp-contract=off: fused operations appear only where the reference writes fmadd.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>

namespace orc {

static const float Pi = 3.14159265358979323846f;
static const float InvPi = 0.31830988618379067154f;
static const float InvTwoPi = 0.15915494309189533577f;
static const float TwoPi = 6.28318530717958647692f;
static const float InvSqrtPi = 0.56418958354775628695f;
static const float Eps = 1.1920929e-07f * 0.5f;           // dr::Epsilon<float> = 2^-24
static const float RayEpsilon = Eps * 1500.f;             // math.h:18-20
static const float ShadowEpsilon = RayEpsilon * 10.f;     // math.h:22
static const float Inf = std::numeric_limits<float>::infinity();
static const float Largest = std::numeric_limits<float>::max();

struct V2 { float x, y; };
struct V3 {
    float x, y, z;
    V3() : x(0), y(0), z(0) {}
    V3(float a) : x(a), y(a), z(a) {}
    V3(float a, float b, float c) : x(a), y(b), z(c) {}
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
static inline V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline V3 operator*(V3 a, V3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline V3 operator/(V3 a, V3 b) { return V3(a.x / b.x, a.y / b.y, a.z / b.z); }
static inline V3 operator*(V3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
static inline V3 operator*(float s, V3 a) { return V3(a.x * s, a.y * s, a.z * s); }
static inline V3 operator/(V3 a, float s) { return V3(a.x / s, a.y / s, a.z / s); }
static inline V3 operator-(V3 a) { return V3(-a.x, -a.y, -a.z); }
static inline V3 &operator+=(V3 &a, V3 b) { a = a + b; return a; }
static inline V3 &operator*=(V3 &a, V3 b) { a = a * b; return a; }
static inline V3 &operator*=(V3 &a, float b) { a = a * b; return a; }
static inline V3 fmadd(V3 a, V3 b, V3 c) { return V3(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y), fmaf(a.z, b.z, c.z)); }
static inline V3 fmadd(V3 a, float b, V3 c) { return V3(fmaf(a.x, b, c.x), fmaf(a.y, b, c.y), fmaf(a.z, b, c.z)); }
static inline V3 fmadd(float a, V3 b, V3 c) { return fmadd(b, a, c); }
static inline float sqr(float x) { return x * x; }
static inline float dot(V3 a, V3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
static inline float squared_norm(V3 a) { return fmaf(a.z, a.z, fmaf(a.y, a.y, a.x * a.x)); }
static inline float norm(V3 a) { return sqrtf(squared_norm(a)); }
static inline float rcp(float x) { return 1.f / x; }
static inline float rsqrt(float x) { return 1.f / sqrtf(x); }
static inline float safe_sqrt(float x) { return sqrtf(fmaxf(x, 0.f)); }
static inline float safe_rsqrt(float x) { return rsqrt(fmaxf(x, 0.f)); }
static inline float clipf(float v, float lo, float hi) { return fmaxf(fminf(v, hi), lo); }
static inline float safe_acos(float x) { return acosf(clipf(x, -1.f, 1.f)); }
static inline V3 normalize(V3 a) { return a * rsqrt(squared_norm(a)); }
static inline V3 cross(V3 a, V3 b) {
    // fmsub(a.yzx, b.zxy, a.zxy * b.yzx)
    return V3(fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x)));
}
static inline float hmax(V3 a) { return fmaxf(fmaxf(a.x, a.y), a.z); }
static inline float mulsign(float a, float b) { return std::signbit(b) ? -a : a; }
static inline float mulsign_neg(float a, float b) { return std::signbit(b) ? a : -a; }
static inline float signf(float x) { return std::signbit(x) ? -1.f : 1.f; }
static inline float lerp(float a, float b, float t) { return fmaf(b, t, fmaf(-a, t, a)); } // array_router.h:350-352
static inline bool finite(float x) { return std::isfinite(x); }

// vector.h:118-138
static inline void coordinate_system(V3 n, V3 &s, V3 &t) {
    float sign = signf(n.z), a = -rcp(sign + n.z), b = n.x * n.y * a;
    s = V3(mulsign(sqr(n.x) * a, n.z) + 1.f, mulsign(b, n.z), mulsign_neg(n.x, n.z));
    t = V3(b, fmaf(n.y, n.y * a, sign), -n.y);
}

struct Frame {
    V3 s, t, n;
    V3 to_local(V3 v) const { return V3(dot(v, s), dot(v, t), dot(v, n)); }
    V3 to_world(V3 v) const { return fmadd(n, v.z, fmadd(t, v.y, s * v.x)); }
};

// row-major 4x4 helpers (Transform::operator*, transform.h:112-160)
static inline V3 xf_point_affine(const float *m, V3 p) {
    // result = col3; result = fmadd(col_i, p_i, result) for i = 0..2
    V3 r(m[3], m[7], m[11]);
    r = fmadd(V3(m[0], m[4], m[8]), p.x, r);
    r = fmadd(V3(m[1], m[5], m[9]), p.y, r);
    r = fmadd(V3(m[2], m[6], m[10]), p.z, r);
    return r;
}
static inline V3 xf_point_persp(const float *m, V3 p) {
    float r[4] = { m[3], m[7], m[11], m[15] };
    for (int i = 0; i < 3; ++i) {
        float a = p[i];
        for (int j = 0; j < 4; ++j) r[j] = fmaf(m[j * 4 + i], a, r[j]);
    }
    return V3(r[0] / r[3], r[1] / r[3], r[2] / r[3]);
}
static inline V3 xf_vector(const float *m, V3 v) {
    V3 r = V3(m[0], m[4], m[8]) * v.x;      // transform.h:148-156
    r = fmadd(V3(m[1], m[5], m[9]), v.y, r);
    r = fmadd(V3(m[2], m[6], m[10]), v.z, r);
    return r;
}

// --------------------------------------------------------------------- RNG
static inline void sample_tea_32(uint32_t v0, uint32_t v1, uint32_t &o0, uint32_t &o1, int rounds = 4) {
    uint32_t sum = 0;
    for (int i = 0; i < rounds; ++i) {
        sum += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    o0 = v0; o1 = v1;
}
static inline float u32_to_float01(uint32_t u) {
    uint32_t bits = (u >> 9) | 0x3f800000u;
    float f; std::memcpy(&f, &bits, 4);
    return f - 1.f;
}
struct PCG32 {
    uint64_t state, inc;
    static constexpr uint64_t MULT = 0x5851f42d4c957f2dULL;
    void seed(uint64_t initstate = 0x853c49e6748fea9bULL, uint64_t initseq = 0xda3e39cb94b95bdbULL) {
        state = 0; inc = (initseq << 1) | 1u;
        next_uint32(); state += initstate; next_uint32();
    }
    uint32_t next_uint32() {
        uint64_t old = state;
        state = old * MULT + inc;
        uint32_t xorshifted = (uint32_t) (((old >> 18) ^ old) >> 27), rot = (uint32_t) (old >> 59);
        return (xorshifted >> rot) | (xorshifted << ((-(int32_t) rot) & 31));
    }
    float next_float32() { return u32_to_float01(next_uint32()); }
};
// PCG32Sampler::seed for JIT variants (sampler.cpp:129-148): lane idx scrambled through TEA
static inline PCG32 sampler_seed(uint32_t base_seed_plus_seed, uint32_t lane) {
    uint32_t v0, v1; sample_tea_32(base_seed_plus_seed, lane, v0, v1);
    PCG32 r; r.seed((uint64_t) v0, (uint64_t) v1); return r;
}

// ------------------------------------------------------------------- warps
static inline V2 square_to_uniform_disk_concentric(V2 s) {
    float x = fmaf(2.f, s.x, -1.f), y = fmaf(2.f, s.y, -1.f);
    bool is_zero = (x == 0.f) && (y == 0.f), q13 = fabsf(x) < fabsf(y);
    float r = q13 ? y : x, rp = q13 ? x : y;
    float phi = 0.25f * Pi * rp / r;
    if (q13) phi = 0.5f * Pi - phi;
    if (is_zero) phi = 0.f;
    float sn = sinf(phi), cs = cosf(phi);
    return V2{ r * cs, r * sn };
}
static inline V3 square_to_cosine_hemisphere(V2 s) {
    V2 p = square_to_uniform_disk_concentric(s);
    float z = safe_sqrt(1.f - fmaf(p.y, p.y, p.x * p.x));
    return V3(p.x, p.y, z);
}
static inline float square_to_cosine_hemisphere_pdf(V3 v) { return InvPi * v.z; }
static inline V2 square_to_uniform_triangle(V2 s) { // warp.h:153-156
    float t = safe_sqrt(1.f - s.x);
    return V2{ 1.f - t, t * s.y };
}
static inline float interval_to_linear(float v0, float v1, float sample) {
    if (fabsf(v0 - v1) > 1e-4f * (v0 + v1))
        return (v0 - safe_sqrt(lerp(sqr(v0), sqr(v1), sample))) / (v0 - v1);
    return sample;
}
static inline V2 square_to_bilinear(float v00, float v10, float v01, float v11, V2 s, float &pdf) {
    float r0 = v00 + v10, r1 = v01 + v11;
    s.y = interval_to_linear(r0, r1, s.y);
    float c0 = lerp(v00, v01, s.y), c1 = lerp(v10, v11, s.y);
    s.x = interval_to_linear(c0, c1, s.x);
    pdf = lerp(c0, c1, s.x);
    return s;
}
static inline float square_to_bilinear_pdf(float v00, float v10, float v01, float v11, V2 s) {
    return lerp(lerp(v00, v10, s.x), lerp(v01, v11, s.x), s.y);
}

// Estrin evaluation, n coefficients (drjit/math.h:26-41)
template <int n> static inline float estrin(float x, const float (&c)[n]) {
    constexpr int n_rec = (n - 1) / 2, n_fma = n / 2;
    float r[n_rec + 1];
    for (int i = 0; i < n_fma; ++i) r[i] = fmaf(x, c[2 * i + 1], c[2 * i]);
    if (n_rec == n_fma) r[n_rec] = c[n - 1];
    if constexpr (n_rec == 0) return r[0];
    else return estrin<n_rec + 1>(x * x, r);
}
static inline float erfinv(float x) {
    float w = -logf((1.f - x) * (1.f + x));
    float w1 = w - 2.5f, w2 = sqrtf(w) - 3.f;
    const float c1[9] = { 1.50140941f, 0.246640727f, -0.00417768164f, -0.00125372503f, 0.00021858087f,
                          -4.39150654e-06f, -3.5233877e-06f, 3.43273939e-07f, 2.81022636e-08f };
    const float c2[9] = { 2.83297682f, 1.00167406f, 0.00943887047f, -0.0076224613f, 0.00573950773f,
                          -0.00367342844f, 0.00134934322f, 0.000100950558f, -0.000200214257f };
    float p1 = estrin<9>(w1, c1), p2 = estrin<9>(w2, c2);
    return (w < 5.f ? p1 : p2) * x;
}

} // namespace orc
