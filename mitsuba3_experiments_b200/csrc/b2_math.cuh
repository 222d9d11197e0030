// b2_math.cuh -- fp32 device math for the b2path kernels (sm_100a).
//
// The translation unit is compiled with -fmad=false: a fused multiply-add appears ONLY where the
// reference writes dr::fmadd / fmsub / fnmadd (dot, cross, Frame::to_world, Transform, lerp, ...), so that
// the per-sample arithmetic follows the llvm_ad_rgb semantics of the reference:
//   dot / cross / normalize   drjit/include/drjit/array_base.h:597-625, array_router.h:651-682
//   coordinate_system         mitsuba3/include/mitsuba/core/vector.h:118-138
//   Frame                     mitsuba3/include/mitsuba/core/frame.h:34-41
//   TEA / PCG32               mitsuba3/include/mitsuba/core/random.h:77-90, drjit/include/drjit/random.h:60-79,211-213
//   warps                     mitsuba3/include/mitsuba/core/warp.h:54-90,153-156,412-436,447-521
// Division and sqrt are IEEE (nvcc default -prec-div/-prec-sqrt); no fast-math intrinsics on the parity path.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define B2_DEV __device__ __forceinline__
#define B2_HD __host__ __device__ __forceinline__

namespace b2 {

constexpr float kPi = 3.14159265358979323846f;
constexpr float kInvPi = 0.31830988618379067154f;
constexpr float kInvTwoPi = 0.15915494309189533577f;
constexpr float kTwoPi = 6.28318530717958647692f;
constexpr float kInvSqrtPi = 0.56418958354775628695f;
constexpr float kEps = 5.9604645e-08f;              // dr::Epsilon<float> = 2^-24
constexpr float kRayEps = kEps * 1500.f;            // math.h:18-20
constexpr float kShadowEps = kRayEps * 10.f;        // math.h:22
constexpr float kLargest = 3.402823466e+38f;

struct f2 { float x, y; };
struct f3 {
    float x, y, z;
    B2_HD f3() : x(0.f), y(0.f), z(0.f) {}
    B2_HD explicit f3(float a) : x(a), y(a), z(a) {}
    B2_HD f3(float a, float b, float c) : x(a), y(b), z(c) {}
};

B2_DEV f3 operator+(f3 a, f3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
B2_DEV f3 operator-(f3 a, f3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
B2_DEV f3 operator*(f3 a, f3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
B2_DEV f3 operator/(f3 a, f3 b) { return f3(a.x / b.x, a.y / b.y, a.z / b.z); }
B2_DEV f3 operator*(f3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
B2_DEV f3 operator*(float s, f3 a) { return f3(a.x * s, a.y * s, a.z * s); }
B2_DEV f3 operator/(f3 a, float s) { return f3(a.x / s, a.y / s, a.z / s); }
B2_DEV f3 operator-(f3 a) { return f3(-a.x, -a.y, -a.z); }
B2_DEV f3 &operator+=(f3 &a, f3 b) { a = a + b; return a; }
B2_DEV f3 &operator*=(f3 &a, f3 b) { a = a * b; return a; }
B2_DEV f3 &operator*=(f3 &a, float b) { a = a * b; return a; }

B2_DEV float fma_(float a, float b, float c) { return __fmaf_rn(a, b, c); }
B2_DEV f3 fmadd(f3 a, f3 b, f3 c) { return f3(fma_(a.x, b.x, c.x), fma_(a.y, b.y, c.y), fma_(a.z, b.z, c.z)); }
B2_DEV f3 fmadd(f3 a, float b, f3 c) { return f3(fma_(a.x, b, c.x), fma_(a.y, b, c.y), fma_(a.z, b, c.z)); }
B2_DEV float sqr(float x) { return x * x; }
B2_DEV float dot(f3 a, f3 b) { return fma_(a.z, b.z, fma_(a.y, b.y, a.x * b.x)); }
B2_DEV float squared_norm(f3 a) { return dot(a, a); }
B2_DEV float norm(f3 a) { return sqrtf(squared_norm(a)); }
B2_DEV float rcp(float x) { return 1.f / x; }
B2_DEV float rsqrt_(float x) { return 1.f / sqrtf(x); }
B2_DEV float safe_sqrt(float x) { return sqrtf(fmaxf(x, 0.f)); }
B2_DEV float safe_rsqrt(float x) { return rsqrt_(fmaxf(x, 0.f)); }
B2_DEV float clipf(float v, float lo, float hi) { return fmaxf(fminf(v, hi), lo); }
B2_DEV float safe_acos(float x) { return acosf(clipf(x, -1.f, 1.f)); }
B2_DEV f3 normalize(f3 a) { return a * rsqrt_(squared_norm(a)); }
B2_DEV f3 cross(f3 a, f3 b) {
    return f3(fma_(a.y, b.z, -(a.z * b.y)), fma_(a.z, b.x, -(a.x * b.z)), fma_(a.x, b.y, -(a.y * b.x)));
}
B2_DEV float hmax(f3 a) { return fmaxf(fmaxf(a.x, a.y), a.z); }
B2_DEV bool signbit_(float x) { return (__float_as_uint(x) >> 31) != 0u; }
B2_DEV float mulsign(float a, float b) { return signbit_(b) ? -a : a; }
B2_DEV float mulsign_neg(float a, float b) { return signbit_(b) ? a : -a; }
B2_DEV float lerp_(float a, float b, float t) { return fma_(b, t, fma_(-a, t, a)); }   // array_router.h:350-352
B2_DEV bool finite_(float x) { return (__float_as_uint(x) & 0x7f800000u) != 0x7f800000u; }
B2_DEV bool is_zero(f3 a) { return a.x == 0.f && a.y == 0.f && a.z == 0.f; }

// vector.h:118-138
B2_DEV void coordinate_system(f3 n, f3 &s, f3 &t) {
    float sign = signbit_(n.z) ? -1.f : 1.f, a = -rcp(sign + n.z), b = n.x * n.y * a;
    s = f3(mulsign(sqr(n.x) * a, n.z) + 1.f, mulsign(b, n.z), mulsign_neg(n.x, n.z));
    t = f3(b, fma_(n.y, n.y * a, sign), -n.y);
}

struct Frame {
    f3 s, t, n;
    B2_DEV f3 to_local(f3 v) const { return f3(dot(v, s), dot(v, t), dot(v, n)); }
    B2_DEV f3 to_world(f3 v) const { return fmadd(n, v.z, fmadd(t, v.y, s * v.x)); }
};

// Row-major 4x4 transforms (Transform::operator*, transform.h:112-160)
B2_DEV f3 xf_point_affine(const float *m, f3 p) {
    f3 r(m[3], m[7], m[11]);
    r = fmadd(f3(m[0], m[4], m[8]), p.x, r);
    r = fmadd(f3(m[1], m[5], m[9]), p.y, r);
    r = fmadd(f3(m[2], m[6], m[10]), p.z, r);
    return r;
}
B2_DEV f3 xf_point_persp(const float *m, f3 p) {
    float r0 = m[3], r1 = m[7], r2 = m[11], r3 = m[15];
    r0 = fma_(m[0], p.x, r0); r1 = fma_(m[4], p.x, r1); r2 = fma_(m[8], p.x, r2); r3 = fma_(m[12], p.x, r3);
    r0 = fma_(m[1], p.y, r0); r1 = fma_(m[5], p.y, r1); r2 = fma_(m[9], p.y, r2); r3 = fma_(m[13], p.y, r3);
    r0 = fma_(m[2], p.z, r0); r1 = fma_(m[6], p.z, r1); r2 = fma_(m[10], p.z, r2); r3 = fma_(m[14], p.z, r3);
    return f3(r0 / r3, r1 / r3, r2 / r3);
}
B2_DEV f3 xf_vector(const float *m, f3 v) {
    f3 r = f3(m[0], m[4], m[8]) * v.x;
    r = fmadd(f3(m[1], m[5], m[9]), v.y, r);
    r = fmadd(f3(m[2], m[6], m[10]), v.z, r);
    return r;
}

// ------------------------------------------------------------------------------------------- RNG
// sample_tea_32 (core/random.h:77-90), 4 rounds
B2_HD void tea32(uint32_t v0, uint32_t v1, uint32_t &o0, uint32_t &o1) {
    uint32_t sum = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        sum += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    o0 = v0; o1 = v1;
}

// PCG32 (drjit/random.h:60-79,211-213) seeded as PCG32Sampler::seed does for JIT variants (sampler.cpp:129-148):
// (v0, v1) = tea(base_seed + seed, lane); state = 0; inc = (v1 << 1) | 1; step; state += v0; step.
// The stream of a lane is a pure function of (seed, lane): nothing is ever stored in global memory.
struct Pcg32 {
    uint64_t state, inc;
    static constexpr uint64_t kMult = 0x5851f42d4c957f2dULL;
    B2_HD void step() { state = state * kMult + inc; }
    B2_HD void seed_lane(uint32_t seed, uint32_t lane) {
        uint32_t v0, v1; tea32(seed, lane, v0, v1);
        state = 0; inc = ((uint64_t) v1 << 1) | 1u;
        step(); state += (uint64_t) v0; step();
    }
    B2_HD uint32_t next_u32() {
        uint64_t old = state; step();
        uint32_t xs = (uint32_t) (((old >> 18) ^ old) >> 27), rot = (uint32_t) (old >> 59);
        return (xs >> rot) | (xs << ((32u - rot) & 31u));
    }
    B2_DEV float next_1d() { return __uint_as_float((next_u32() >> 9) | 0x3f800000u) - 1.f; }
    B2_DEV f2 next_2d() { float a = next_1d(); float b = next_1d(); return f2{ a, b }; }
    B2_HD void skip(int n) { for (int i = 0; i < n; ++i) step(); }
};

// ------------------------------------------------------------------------------------------- warps
B2_DEV f2 square_to_uniform_disk_concentric(f2 s) {
    float x = fma_(2.f, s.x, -1.f), y = fma_(2.f, s.y, -1.f);
    bool zero = (x == 0.f) && (y == 0.f), q13 = fabsf(x) < fabsf(y);
    float r = q13 ? y : x, rp = q13 ? x : y;
    float phi = 0.25f * kPi * rp / r;
    if (q13) phi = 0.5f * kPi - phi;
    if (zero) phi = 0.f;
    float sn, cs; sincosf(phi, &sn, &cs);
    return f2{ r * cs, r * sn };
}
B2_DEV f3 square_to_cosine_hemisphere(f2 s) {
    f2 p = square_to_uniform_disk_concentric(s);
    float z = safe_sqrt(1.f - fma_(p.y, p.y, p.x * p.x));
    return f3(p.x, p.y, z);
}
B2_DEV f2 square_to_uniform_triangle(f2 s) {
    float t = safe_sqrt(1.f - s.x);
    return f2{ 1.f - t, t * s.y };
}
B2_DEV float interval_to_linear(float v0, float v1, float sample) {
    if (fabsf(v0 - v1) > 1e-4f * (v0 + v1))
        return (v0 - safe_sqrt(lerp_(sqr(v0), sqr(v1), sample))) / (v0 - v1);
    return sample;
}
B2_DEV f2 square_to_bilinear(float v00, float v10, float v01, float v11, f2 s, float &pdf) {
    float r0 = v00 + v10, r1 = v01 + v11;
    s.y = interval_to_linear(r0, r1, s.y);
    float c0 = lerp_(v00, v01, s.y), c1 = lerp_(v10, v11, s.y);
    s.x = interval_to_linear(c0, c1, s.x);
    pdf = lerp_(c0, c1, s.x);
    return s;
}
B2_DEV float square_to_bilinear_pdf(float v00, float v10, float v01, float v11, f2 s) {
    return lerp_(lerp_(v00, v10, s.x), lerp_(v01, v11, s.x), s.y);
}

// Estrin's scheme exactly as dr::estrin expands it (drjit/math.h:26-41) for 10 and 9 coefficients
B2_DEV float estrin10(float x, const float *c) {
    float x2 = x * x, x4 = x2 * x2, x8 = x4 * x4;
    float a0 = fma_(x, c[1], c[0]), a1 = fma_(x, c[3], c[2]), a2 = fma_(x, c[5], c[4]), a3 = fma_(x, c[7], c[6]), a4 = fma_(x, c[9], c[8]);
    float b0 = fma_(x2, a1, a0), b1 = fma_(x2, a3, a2), b2 = a4;
    float c0 = fma_(x4, b1, b0), c1 = b2;
    return fma_(x8, c1, c0);
}
B2_DEV float estrin9(float x, const float *c) {
    float x2 = x * x, x4 = x2 * x2, x8 = x4 * x4;
    float a0 = fma_(x, c[1], c[0]), a1 = fma_(x, c[3], c[2]), a2 = fma_(x, c[5], c[4]), a3 = fma_(x, c[7], c[6]), a4 = c[8];
    float b0 = fma_(x2, a1, a0), b1 = fma_(x2, a3, a2), b2 = a4;
    float c0 = fma_(x4, b1, b0), c1 = b2;
    return fma_(x8, c1, c0);
}
// dr::erfinv (drjit/math.h:1556-1578, Giles' single-precision approximation)
B2_DEV float erfinv_(float x) {
    const float c1[9] = { 1.50140941f, 0.246640727f, -0.00417768164f, -0.00125372503f, 0.00021858087f,
                          -4.39150654e-06f, -3.5233877e-06f, 3.43273939e-07f, 2.81022636e-08f };
    const float c2[9] = { 2.83297682f, 1.00167406f, 0.00943887047f, -0.0076224613f, 0.00573950773f,
                          -0.00367342844f, 0.00134934322f, 0.000100950558f, -0.000200214257f };
    float w = -logf((1.f - x) * (1.f + x));
    float p1 = estrin9(w - 2.5f, c1), p2 = estrin9(sqrtf(w) - 3.f, c2);
    return (w < 5.f ? p1 : p2) * x;
}

}  // namespace b2
