// oracle/_ref/libref_pin.so: the REFERENCE's own header-only math, compiled in scalar float mode from
// the sources where they lie under /root/reference (nothing is copied into this repository).
// Used only by tests/ and tests/golden/make_golden.py to pin the oracle restatement.
//   drjit/include/drjit/random.h        PCG32
//   mitsuba/core/random.h               sample_tea_32 / sample_tea_float32
//   mitsuba/core/warp.h                 square_to_* warps, interval_to_linear, square_to_bilinear
//   mitsuba/render/fresnel.h            fresnel, fresnel_conductor
//   mitsuba/render/microfacet.h         MicrofacetDistribution, eval_transmittance / eval_reflectance (+ core/quad.h)
//   mitsuba/core/distr_2d.h             Hierarchical2D<float, 0>
//   mitsuba/core/vector.h               coordinate_system
#include <vector>
#include <string>
#include <array>
#include <tuple>
#include <memory>
#include <sstream>
#include <mitsuba/mitsuba.h>
#include <mitsuba/core/fwd.h>
#include <mitsuba/core/object.h>
#include <mitsuba/core/random.h>
#include <mitsuba/core/warp.h>
#include <mitsuba/core/distr_2d.h>
#include <mitsuba/render/fresnel.h>
#include <mitsuba/render/microfacet.h>
#include <drjit/random.h>
#include <drjit/math.h>
#include <cstdarg>
#include <cstdio>
#include <stdexcept>

// link stubs for symbols normally provided by libdrjit-core / libmitsuba
extern "C" int jit_flag(JitFlag) { return 0; }
extern "C" void jit_log(LogLevel, const char *, ...) {}
#undef Throw
namespace mitsuba { namespace detail {
void Throw(LogLevel, const char *, const char *file, int line, std::string_view msg) {
    throw std::runtime_error(std::string(file) + ":" + std::to_string(line) + " " + std::string(msg));
}
} }

using namespace mitsuba;
namespace dr = drjit;
using Float = float;
using Vector3f = Vector<float, 3>;
using Point2f = Point<float, 2>;
using MF = MicrofacetDistribution<float, Color<float, 3>>;

extern "C" {
void ref_tea32(uint32_t v0, uint32_t v1, int rounds, uint32_t *out) { auto r = sample_tea_32<uint32_t>(v0, v1, rounds); out[0] = r.first; out[1] = r.second; }
float ref_tea_float32(uint32_t v0, uint32_t v1, int rounds) { return sample_tea_float32<uint32_t>(v0, v1, rounds); }
void ref_pcg32_floats(uint64_t initstate, uint64_t initseq, int n, float *out) {
    dr::PCG32<float> rng; rng.seed(initstate, initseq);
    for (int i = 0; i < n; ++i) out[i] = rng.next_float32();
}
void ref_pcg32_default_floats(int n, float *out) { dr::PCG32<float> rng; for (int i = 0; i < n; ++i) out[i] = rng.next_float32(); }
void ref_warp(int which, const float *in, float *out) {
    if (which == 0) { auto r = warp::square_to_uniform_disk_concentric(Point2f(in[0], in[1])); out[0] = r.x(); out[1] = r.y(); }
    else if (which == 1) { auto r = warp::square_to_cosine_hemisphere(Point2f(in[0], in[1])); out[0] = r.x(); out[1] = r.y(); out[2] = r.z(); }
    else if (which == 2) { auto r = warp::square_to_uniform_triangle(Point2f(in[0], in[1])); out[0] = r.x(); out[1] = r.y(); }
    else if (which == 3) out[0] = warp::interval_to_linear(in[0], in[1], in[2]);
    else if (which == 4) out[0] = dr::erfinv(in[0]);
    else if (which == 5) { auto r = coordinate_system(Vector3f(in[0], in[1], in[2])); for (int i = 0; i < 3; ++i) { out[i] = r.first[i]; out[3 + i] = r.second[i]; } }
}
void ref_fresnel(float cos_i, float eta, float *out4) { auto r = fresnel(cos_i, eta); out4[0] = std::get<0>(r); out4[1] = std::get<1>(r); out4[2] = std::get<2>(r); out4[3] = std::get<3>(r); }
float ref_fresnel_conductor(float cos_i, float eta, float k) { return fresnel_conductor(cos_i, dr::Complex<float>(eta, k)); }
void ref_microfacet(int type, float au, float av, int visible, int op, const float *in, float *out) {
    MF d(type == 0 ? MicrofacetType::Beckmann : MicrofacetType::GGX, au, av, visible != 0);
    if (op == 0) { auto r = d.sample(Vector3f(in[0], in[1], in[2]), Point2f(in[3], in[4])); out[0] = r.first.x(); out[1] = r.first.y(); out[2] = r.first.z(); out[3] = r.second; }
    else if (op == 1) out[0] = d.eval(Vector3f(in[0], in[1], in[2]));
    else if (op == 2) out[0] = d.smith_g1(Vector3f(in[0], in[1], in[2]), Vector3f(in[3], in[4], in[5]));
    else if (op == 3) out[0] = d.pdf(Vector3f(in[0], in[1], in[2]), Vector3f(in[3], in[4], in[5]));
}
// the rough-plastic precomputation exactly as RoughPlastic::parameters_changed runs it (roughplastic.cpp:232-262)
void ref_rough_tables(int type, float alpha, float eta, float *trans64, float *internal_reflectance) {
    using FloatX = dr::DynamicArray<float>;
    using Vector3fX = Vector<FloatX, 3>;
    using FloatP = dr::Packet<float>;
    MicrofacetDistribution<FloatP, Color<float, 3>> distr(type == 0 ? MicrofacetType::Beckmann : MicrofacetType::GGX, alpha);
    FloatX mu = dr::maximum(1e-6f, dr::linspace<FloatX>(0, 1, 64));
    FloatX zero = dr::zeros<FloatX>(64);
    Vector3fX wi = Vector3fX(dr::sqrt(1 - mu * mu), zero, mu);
    auto t = eval_transmittance(distr, wi, eta);
    for (int i = 0; i < 64; ++i) trans64[i] = t[i];
    *internal_reflectance = dr::mean(eval_reflectance(distr, wi, 1.f / eta) * wi.z()) * 2.f;
}
// Hierarchical2D<float,0> over data [h][w]
void *ref_h2d_create(const float *data, uint32_t w, uint32_t h) { return new Hierarchical2D<float, 0>(data, Vector<uint32_t, 2>(w, h)); }
void ref_h2d_destroy(void *p) { delete (Hierarchical2D<float, 0> *) p; }
void ref_h2d_sample(void *p, float u, float v, float *out3) { auto r = ((Hierarchical2D<float, 0> *) p)->sample(Point2f(u, v)); out3[0] = r.first.x(); out3[1] = r.first.y(); out3[2] = r.second; }
float ref_h2d_eval(void *p, float u, float v) { return ((Hierarchical2D<float, 0> *) p)->eval(Point2f(u, v)); }
}
