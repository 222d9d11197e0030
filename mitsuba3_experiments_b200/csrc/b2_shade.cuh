// b2_shade.cuh -- surface interaction, textures, BSDFs and emitters (device code).
//
// Follows, operation for operation, the llvm_ad_rgb arithmetic of the reference plugins (file:line relative to
// /root/reference/mitsuba3):
//   surface interaction  src/render/mesh.cpp:1386-1560, include/mitsuba/render/interaction.h:140-169,278-288,497-516
//   bitmap texture       src/textures/bitmap.cpp:419-451,783-800 -> ext/drjit/include/drjit/texture.h:478-551 (fp32 software bilinear)
//   diffuse              src/bsdfs/diffuse.cpp:100-179
//   rough conductor      src/bsdfs/roughconductor.cpp:227-518, include/mitsuba/render/microfacet.h:185-421, fresnel.h:92-117
//   dielectric           src/bsdfs/dielectric.cpp:245-370, fresnel.h:34-71,276-295
//   twosided             src/bsdfs/twosided.cpp:110-260
//   area light           src/emitters/area.cpp:83-197, src/render/shape.cpp:354-384, src/shapes/rectangle.cpp:182-200,
//                        src/render/mesh.cpp:758-812
//   envmap               src/emitters/envmap.cpp:353-363,414-476,528-586, include/mitsuba/core/distr_2d.h:517-602,695-726
//   constant             src/emitters/constant.cpp
//   emitter selection    src/render/scene.cpp:252-392, include/mitsuba/render/records.h:173-180
#pragma once
#include "b2_bvh.cuh"
#include <cuda_fp16.h>

namespace b2 {

enum { LOBE_DIFFUSE = 1, LOBE_GLOSSY = 2, LOBE_DELTA_R = 4, LOBE_DELTA_T = 8, LOBE_NULL = 16,
       LOBE_SMOOTH = LOBE_DIFFUSE | LOBE_GLOSSY, LOBE_DELTA = LOBE_DELTA_R | LOBE_DELTA_T | LOBE_NULL };   // BSDFFlags::Delta includes Null

// Shading classes: the wavefront sorts the hits of a bounce by class and runs one specialised shade kernel per class
// (CLS >= 0 turns the BSDF-type switch and the textured/constant test into compile-time constants); CLS_GENERIC keeps
// both at run time (megakernel, adjoint replay).
// CLS_CONDUCTOR = rough conductor, CLS_PLASTIC = rough plastic, CLS_ROUGHDIEL = rough dielectric (one kernel each: the single "glossy"
// kernel that switched between the three at run time was 15 520 SASS instructions and spent 86 % of its cycles waiting for
// instruction fetch, profiles/README.md round 2), CLS_DIELECTRIC = the delta BSDFs (smooth dielectric, smooth conductor).
enum { CLS_GENERIC = -1, CLS_DIFFUSE = 0, CLS_DIFFUSE_TEX = 1, CLS_CONDUCTOR = 2, CLS_DIELECTRIC = 3, CLS_MISS = 4, CLS_PLASTIC = 5, CLS_ROUGHDIEL = 6, CLS_COUNT = 7 };
template <int CLS> B2_DEV int cls_type(int runtime_type) {
    return (CLS == CLS_DIFFUSE || CLS == CLS_DIFFUSE_TEX) ? B2_BSDF_DIFFUSE
           : CLS == CLS_CONDUCTOR ? B2_BSDF_ROUGHCONDUCTOR : CLS == CLS_PLASTIC ? B2_BSDF_ROUGHPLASTIC : CLS == CLS_ROUGHDIEL ? B2_BSDF_ROUGHDIELECTRIC
           : CLS == CLS_DIELECTRIC ? (runtime_type == B2_BSDF_CONDUCTOR ? B2_BSDF_CONDUCTOR : B2_BSDF_DIELECTRIC)   // the delta class
           : runtime_type;
}
template <int CLS> B2_DEV bool cls_textured(int tex) { return CLS == CLS_DIFFUSE ? false : CLS == CLS_DIFFUSE_TEX ? true : tex >= 0; }

struct SI {
    bool valid;
    f3 p, n, wi;
    f2 uv;
    Frame sh;
    int mesh, bsdf, emitter;         // emitter: index into the emitter table (-1 none; env emitter when !valid)
};
struct DS { f3 p, n, d; float pdf, dist; bool delta; int emitter; };
struct BS { f3 wo; float pdf, eta; uint32_t type; };

struct Ray { f3 o, d; float maxt; };

// -------------------------------------------------------------------------------- surface interaction
B2_DEV f3 load3(const float *p, uint32_t i) { return f3(__ldg(p + 3 * (size_t) i), __ldg(p + 3 * (size_t) i + 1), __ldg(p + 3 * (size_t) i + 2)); }
B2_DEV f2 load2(const float *p, uint32_t i) { return f2{ __ldg(p + 2 * (size_t) i), __ldg(p + 2 * (size_t) i + 1) }; }

template <bool COUNT>
B2_DEV void make_si(const DScene &sc, const Ray &ray, const Hit &h, SI &si, Counters<COUNT> &cnt) {
    si.valid = h.valid();
    si.uv = f2{ 0.f, 0.f };
    if (!si.valid) { si.wi = -ray.d; si.mesh = -1; si.bsdf = -1; si.emitter = sc.env_emitter; si.p = f3(); si.n = f3(); return; }
    cnt.add(C_VERTICES);
    const uint32_t f = h.face;
    const DMesh m = sc.meshes[h.mesh];
    si.mesh = (int) h.mesh; si.bsdf = m.bsdf; si.emitter = m.emitter;
    const float4 *rec = sc.shade + (size_t) f * 6;
    const float4 r0 = __ldg(rec), r1 = __ldg(rec + 1), r2 = __ldg(rec + 2);
    const f3 p0(r0.x, r0.y, r0.z), p1(r0.w, r1.x, r1.y), p2(r1.z, r1.w, r2.x);
    const float b1 = h.u, b2 = h.v, b0 = 1.f - b1 - b2;
    si.p = fmadd(p0, b0, fmadd(p1, b1, p2 * b2));
    const f3 dp0 = p1 - p0, dp1 = p2 - p0;
    si.n = normalize(cross(dp0, dp1));
    si.uv = f2{ b1, b2 };
    f3 dp_du, dp_dv;
    coordinate_system(si.n, dp_du, dp_dv);
    if (m.flags & B2_MESH_HAS_UV) {
        const float4 r4 = __ldg(rec + 4), r5 = __ldg(rec + 5);
        const f2 uv0{ r4.z, r4.w }, uv1{ r5.x, r5.y }, uv2{ r5.z, r5.w };
        si.uv.x = fma_(uv2.x, b2, fma_(uv1.x, b1, uv0.x * b0));
        si.uv.y = fma_(uv2.y, b2, fma_(uv1.y, b1, uv0.y * b0));
        const float du0x = uv1.x - uv0.x, du0y = uv1.y - uv0.y, du1x = uv2.x - uv0.x, du1y = uv2.y - uv0.y;
        const float det = fma_(du0x, du1y, -(du0y * du1x)), inv_det = rcp(det);
        if (det != 0.f) {
            dp_du = fmadd(f3(du1y), dp0, -(dp1 * du0y)) * inv_det;
            dp_dv = fmadd(f3(-du1x), dp0, dp1 * du0x) * inv_det;
        }
    }
    if (m.flags & B2_MESH_HAS_NORMALS) {
        const float4 r3 = __ldg(rec + 3), r4n = __ldg(rec + 4);
        const f3 n0(r2.y, r2.z, r2.w), n1(r3.x, r3.y, r3.z), n2(r3.w, r4n.x, r4n.y);
        const f3 n = fmadd(n2, b2, fmadd(n1, b1, n0 * b0));
        si.sh.n = n * rsqrt_(squared_norm(n));
    } else si.sh.n = si.n;
    if (m.flags & B2_MESH_FLIP_NORMALS) { si.n = -si.n; si.sh.n = -si.sh.n; }
    si.sh.s = normalize(fmadd(si.sh.n, -dot(si.sh.n, dp_du), dp_du));
    if (is_zero(dp_du)) { f3 t; coordinate_system(si.sh.n, si.sh.s, t); }
    si.sh.t = cross(si.sh.n, si.sh.s);
    si.wi = si.sh.to_local(-ray.d);
}

// Interaction::offset_p / spawn_ray / spawn_ray_to (interaction.h:140-169)
B2_DEV f3 offset_p(f3 p, f3 n, f3 dir) {
    float mag = (1.f + hmax(f3(fabsf(p.x), fabsf(p.y), fabsf(p.z)))) * kRayEps;
    mag = mulsign(mag, dot(n, dir));
    return fmadd(f3(mag), n, p);
}
B2_DEV Ray spawn_ray(f3 p, f3 n, f3 dir) { return Ray{ offset_p(p, n, dir), dir, kLargest }; }
B2_DEV Ray spawn_ray_to(f3 p, f3 n, f3 target) {
    f3 o = offset_p(p, n, target - p);
    f3 dd = target - o; float dist = norm(dd); dd = dd / dist;
    return Ray{ o, dd, dist * (1.f - kShadowEps) };
}

// -------------------------------------------------------------------------------- textures
B2_DEV int wrap_coord(int pos, int res, int mode) {
    if (mode == B2_WRAP_CLAMP) return min(max(pos, 0), res - 1);
    int shift = pos < 0 ? pos + 1 : pos, div = shift / res, mod = pos - div * res;
    if (mod < 0) mod += res;
    if (mode == B2_WRAP_MIRROR) mod = (((div & 1) == 0) != (pos < 0)) ? mod : res - 1 - mod;
    return mod;
}
B2_DEV f3 texel(const DTexture &t, uint32_t idx) {
    if (t.format == B2_TEX_F16) {
        const __half *p = (const __half *) t.data + 3 * (size_t) idx;
        return f3(__half2float(p[0]), __half2float(p[1]), __half2float(p[2]));
    }
    const float *p = (const float *) t.data + 3 * (size_t) idx;
    return f3(__ldg(p), __ldg(p + 1), __ldg(p + 2));
}
// returns the interpolated value; idx/w receive the 4 texel indices and weights (adjoint scatter)
template <bool COUNT>
B2_DEV f3 tex_eval(const DScene &sc, int ti, f2 uv_, uint32_t *idx, float *w, Counters<COUNT> &cnt) {
    const DTexture t = sc.textures[ti]; cnt.add(C_TEX);
    const float *m = t.to_uv;
    const f2 uv{ fma_(m[1], uv_.y, fma_(m[0], uv_.x, m[2])), fma_(m[4], uv_.y, fma_(m[3], uv_.x, m[5])) };
    if (t.filter == B2_FILTER_NEAREST) {
        int x = wrap_coord((int) floorf(uv.x * (float) t.width), t.width, t.wrap), y = wrap_coord((int) floorf(uv.y * (float) t.height), t.height, t.wrap);
        idx[0] = idx[1] = idx[2] = idx[3] = (uint32_t) y * t.width + x; w[0] = 1.f; w[1] = w[2] = w[3] = 0.f;
        return texel(t, idx[0]);
    }
    const float px = fma_(uv.x, (float) t.width, -.5f), py = fma_(uv.y, (float) t.height, -.5f);
    const int ix = (int) floorf(px), iy = (int) floorf(py);
    const float w1x = px - (float) ix, w1y = py - (float) iy, w0x = 1.f - w1x, w0y = 1.f - w1y;
    const int x0 = wrap_coord(ix, t.width, t.wrap), x1 = wrap_coord(ix + 1, t.width, t.wrap);
    const int y0 = wrap_coord(iy, t.height, t.wrap), y1 = wrap_coord(iy + 1, t.height, t.wrap);
    idx[0] = y0 * t.width + x0; idx[1] = y0 * t.width + x1; idx[2] = y1 * t.width + x0; idx[3] = y1 * t.width + x1;
    w[0] = w0x * w0y; w[1] = w1x * w0y; w[2] = w0x * w1y; w[3] = w1x * w1y;
    f3 out(0.f);
#pragma unroll
    for (int i = 0; i < 4; ++i) out = fmadd(texel(t, idx[i]), w[i], out);
    return out;
}
// Reflectance of a BSDF at a vertex.  The texture is fetched once per vertex and reused by eval_pdf / sample (the
// reference's JIT merges the two identical lookups the same way); every logical evaluation is still counted.
struct ReflCache { bool have = false; f3 v; };
template <int CLS = CLS_GENERIC, bool COUNT>
B2_DEV f3 reflectance(const DScene &sc, const DBsdf &b, const SI &si, ReflCache &rc, Counters<COUNT> &cnt) {
    if (!cls_textured<CLS>(b.tex)) return f3(b.refl[0], b.refl[1], b.refl[2]);
    if (!rc.have) { uint32_t idx[4]; float w[4]; rc.v = tex_eval(sc, b.tex, si.uv, idx, w, cnt); rc.have = true; }
    else cnt.add(C_TEX);
    return rc.v;
}

// -------------------------------------------------------------------------------- Fresnel
B2_DEV void fresnel(float cos_i, float eta, float &r, float &cos_t, float &eta_it, float &eta_ti) {
    const bool outside = cos_i >= 0.f;
    const float rcp_eta = rcp(eta); eta_it = outside ? eta : rcp_eta; eta_ti = outside ? rcp_eta : eta;
    const float cos_t_sqr = fma_(-fma_(-cos_i, cos_i, 1.f), eta_ti * eta_ti, 1.f);
    const float ci = fabsf(cos_i), ct = safe_sqrt(cos_t_sqr);
    const bool index_matched = eta == 1.f, special = index_matched || ci == 0.f;
    const float a_s = fma_(-eta_it, ct, ci) / fma_(eta_it, ct, ci);
    const float a_p = fma_(-eta_it, ci, ct) / fma_(eta_it, ci, ct);
    r = 0.5f * (sqr(a_s) + sqr(a_p));
    if (special) r = index_matched ? 0.f : 1.f;
    cos_t = mulsign_neg(ct, cos_i);
}
B2_DEV float fresnel_conductor(float cos_i, float eta_r, float eta_i) {
    const float c2 = cos_i * cos_i, s2 = 1.f - c2, s4 = s2 * s2;
    const float temp_1 = eta_r * eta_r - eta_i * eta_i - s2;
    const float a_2_pb_2 = safe_sqrt(temp_1 * temp_1 + 4.f * eta_i * eta_i * eta_r * eta_r);
    const float a = safe_sqrt(.5f * (a_2_pb_2 + temp_1));
    const float term_1 = a_2_pb_2 + c2, term_2 = 2.f * cos_i * a;
    const float r_s = (term_1 - term_2) / (term_1 + term_2);
    const float term_3 = a_2_pb_2 * c2 + s4, term_4 = term_2 * s2;
    const float r_p = r_s * (term_3 - term_4) / (term_3 + term_4);
    return 0.5f * (r_s + r_p);
}

// -------------------------------------------------------------------------------- microfacet distribution
struct Microfacet {
    int type; float au, av; bool visible;
    B2_DEV Microfacet(int t, float a_u, float a_v, bool vis) : type(t), au(fmaxf(a_u, 1e-4f)), av(fmaxf(a_v, 1e-4f)), visible(vis) {}
    B2_DEV void scale_alpha(float v) { au *= v; av *= v; }      // microfacet.h:174-177
    B2_DEV float eval(f3 m) const {
        const float alpha_uv = au * av, ct = m.z, ct2 = sqr(ct);
        float result;
        if (type == B2_MF_BECKMANN) result = expf(-(sqr(m.x / au) + sqr(m.y / av)) / ct2) / (kPi * alpha_uv * sqr(ct2));
        else result = rcp(kPi * alpha_uv * sqr(sqr(m.x / au) + sqr(m.y / av) + sqr(m.z)));
        return result * ct > 1e-20f ? result : 0.f;
    }
    B2_DEV float smith_g1(f3 v, f3 m) const {
        const float xy_alpha_2 = sqr(au * v.x) + sqr(av * v.y), tan_theta_alpha_2 = xy_alpha_2 / sqr(v.z);
        float result;
        if (type == B2_MF_BECKMANN) {
            const float a = rsqrt_(tan_theta_alpha_2), a_sqr = sqr(a);
            result = a >= 1.6f ? 1.f : (3.535f * a + 2.181f * a_sqr) / (1.f + 2.276f * a + 2.577f * a_sqr);
        } else result = 2.f / (1.f + sqrtf(1.f + tan_theta_alpha_2));
        if (xy_alpha_2 == 0.f) result = 1.f;
        if (dot(v, m) * v.z <= 0.f) result = 0.f;
        return result;
    }
    B2_DEV float G(f3 wi, f3 wo, f3 m) const { return smith_g1(wi, m) * smith_g1(wo, m); }
    B2_DEV float pdf(f3 wi, f3 m) const {
        float result = eval(m);
        if (visible) result *= smith_g1(wi, m) * fabsf(dot(wi, m)) / wi.z; else result *= m.z;
        return result;
    }
    B2_DEV f2 sample_visible_11(float cos_theta_i, f2 sample) const {
        if (type == B2_MF_BECKMANN) {
            const float tan_theta_i = safe_sqrt(fma_(-cos_theta_i, cos_theta_i, 1.f)) / cos_theta_i, cot_theta_i = rcp(tan_theta_i);
            const float maxval = erff(cot_theta_i);
            sample.x = fmaxf(fminf(sample.x, 1.f - 1e-6f), 1e-6f); sample.y = fmaxf(fminf(sample.y, 1.f - 1e-6f), 1e-6f);
            float x = maxval - (maxval + 1.f) * erff(sqrtf(-logf(sample.x)));
            sample.x *= 1.f + maxval + kInvSqrtPi * tan_theta_i * expf(-sqr(cot_theta_i));
#pragma unroll 1
            for (int i = 0; i < 3; ++i) {
                const float slope = erfinv_(x), value = 1.f + x + kInvSqrtPi * tan_theta_i * expf(-sqr(slope)) - sample.x,
                            derivative = 1.f - slope * tan_theta_i;
                x -= value / derivative;
            }
            return f2{ erfinv_(x), erfinv_(fma_(2.f, sample.y, -1.f)) };
        }
        f2 p = square_to_uniform_disk_concentric(sample);
        const float s = 0.5f * (1.f + cos_theta_i);
        p.y = lerp_(safe_sqrt(1.f - sqr(p.x)), p.y, s);
        const float x = p.x, y = p.y, z = safe_sqrt(1.f - fma_(p.y, p.y, p.x * p.x));
        const float sin_theta_i = safe_sqrt(1.f - sqr(cos_theta_i));
        const float nrm = rcp(fma_(sin_theta_i, y, cos_theta_i * z));
        return f2{ fma_(cos_theta_i, y, -(sin_theta_i * z)) * nrm, x * nrm };
    }
    B2_DEV f3 sample(f3 wi, f2 sample, float &pdf_out) const {
        if (!visible) {
            float sin_phi, cos_phi, cos_theta, cos_theta_2, alpha_2, pdfv;
            if (au == av) { sincosf(kTwoPi * sample.y, &sin_phi, &cos_phi); alpha_2 = au * au; }
            else {
                const float ratio = av / au, tmp = ratio * tanf(kTwoPi * sample.y);
                cos_phi = rsqrt_(fma_(tmp, tmp, 1.f)); cos_phi = mulsign(cos_phi, fabsf(sample.y - .5f) - .25f);
                sin_phi = cos_phi * tmp; alpha_2 = rcp(sqr(cos_phi / au) + sqr(sin_phi / av));
            }
            if (type == B2_MF_BECKMANN) {
                cos_theta = rsqrt_(fma_(-alpha_2, logf(1.f - sample.x), 1.f)); cos_theta_2 = sqr(cos_theta);
                const float c3 = fmaxf(cos_theta_2 * cos_theta, 1e-20f);
                pdfv = (1.f - sample.x) / (kPi * au * av * c3);
            } else {
                const float tan2 = alpha_2 * sample.x / (1.f - sample.x);
                cos_theta = rsqrt_(1.f + tan2); cos_theta_2 = sqr(cos_theta);
                const float temp = 1.f + tan2 / alpha_2, c3 = fmaxf(cos_theta_2 * cos_theta, 1e-20f);
                pdfv = rcp(kPi * au * av * c3 * sqr(temp));
            }
            const float sin_theta = sqrtf(1.f - cos_theta_2);
            pdf_out = pdfv; return f3(cos_phi * sin_theta, sin_phi * sin_theta, cos_theta);
        }
        const f3 wi_p = normalize(f3(au * wi.x, av * wi.y, wi.z));
        float sin_phi, cos_phi;                                   // Frame::sincos_phi (frame.h:111-122)
        {
            const float st2 = fma_(wi_p.x, wi_p.x, sqr(wi_p.y)), inv = rsqrt_(st2);
            if (fabsf(st2) <= 4.f * kEps) { cos_phi = 1.f; sin_phi = 0.f; }
            else { cos_phi = clipf(wi_p.x * inv, -1.f, 1.f); sin_phi = clipf(wi_p.y * inv, -1.f, 1.f); }
        }
        f2 slope = sample_visible_11(wi_p.z, sample);
        slope = f2{ fma_(cos_phi, slope.x, -(sin_phi * slope.y)) * au, fma_(sin_phi, slope.x, cos_phi * slope.y) * av };
        const f3 m = normalize(f3(-slope.x, -slope.y, 1.f));
        pdf_out = eval(m) * smith_g1(wi, m) * fabsf(dot(wi, m)) / wi.z;
        return m;
    }
};

// -------------------------------------------------------------------------------- BSDFs
template <int CLS = CLS_GENERIC>
B2_DEV uint32_t bsdf_lobes(const DBsdf &b) {
    const int type = cls_type<CLS>(b.type);
    const uint32_t l = type == B2_BSDF_DIFFUSE ? LOBE_DIFFUSE : ((type == B2_BSDF_ROUGHCONDUCTOR || type == B2_BSDF_ROUGHDIELECTRIC) ? LOBE_GLOSSY
                       : (type == B2_BSDF_ROUGHPLASTIC ? (LOBE_GLOSSY | LOBE_DIFFUSE) : (LOBE_DELTA_R | LOBE_DELTA_T)));
    return (b.flags & B2_BSDF_FLAG_MASK) ? (l | LOBE_NULL) : l;      // mask.cpp:96-110
}
B2_DEV f3 fresnel_conductor3(float c, const DBsdf &b) {
    return f3(fresnel_conductor(c, b.eta[0], b.k[0]), fresnel_conductor(c, b.eta[1], b.k[1]), fresnel_conductor(c, b.eta[2], b.k[2]));
}

// RoughPlastic::lerp_gather (roughplastic.cpp:343-351)
B2_DEV float lerp_gather(const float *__restrict__ data, float x, int size) {
    x *= (float) (size - 1);
    const uint32_t index = min((uint32_t) x, (uint32_t) (size - 2));
    const float v0 = __ldg(data + index), v1 = __ldg(data + index + 1);
    return lerp_(v0, v1, x - (float) index);
}
// RoughPlastic::eval / pdf (roughplastic.cpp:301-341, 353-393); wi.z > 0 and wo.z > 0 are the caller's precondition
template <class RF>
B2_DEV void roughplastic_eval_pdf(const DBsdf &b, f3 wi, f3 wo, RF &&refl, f3 &val, float &pdf, bool deriv = false) {
    const float ci = wi.z, co = wo.z;
    const Microfacet distr(b.distribution, b.alpha_u, b.alpha_u, b.sample_visible != 0);
    const f3 H = normalize(wo + wi);
    const float D = distr.eval(H);
    float F, cos_t, eta_it, eta_ti; fresnel(dot(wi, H), b.ior, F, cos_t, eta_it, eta_ti);
    const float G = distr.G(wi, wo, H);
    const f3 value = f3(F * D * G / (4.f * ci)) * f3(b.trans[0], b.trans[1], b.trans[2]);      // specular_reflectance
    const float t_i = lerp_gather(b.rough_table, ci, 64), t_o = lerp_gather(b.rough_table, co, 64);
    f3 diff = refl();
    if (deriv) {
        // d/d rho of rho / (1 - rho r) = 1 / (1 - rho r)^2 (nonlinear), of rho / (1 - r) = 1 / (1 - r); the specular lobe does not depend on rho
        const f3 den = f3(1.f) - (b.nonlinear ? diff * b.int_refl : f3(b.int_refl));
        diff = b.nonlinear ? f3(1.f) / (den * den) : f3(1.f) / den;
    } else diff = diff / (f3(1.f) - (b.nonlinear ? diff * b.int_refl : f3(b.int_refl)));
    const float inv_eta_2 = 1.f / (b.ior * b.ior);
    val = (deriv ? f3(0.f) : value) + diff * (kInvPi * inv_eta_2 * co * t_i * t_o);
    float prob_specular = (1.f - t_i) * b.spec_weight, prob_diffuse = t_i * (1.f - b.spec_weight);
    prob_specular = prob_specular / (prob_specular + prob_diffuse); prob_diffuse = 1.f - prob_specular;
    float r = b.sample_visible ? D * distr.smith_g1(wi, H) / (4.f * ci) : distr.pdf(wi, H) / (4.f * dot(wo, H));
    r *= prob_specular;
    r += prob_diffuse * (kInvPi * co);
    pdf = r;
}

// eval_pdf.  `refl()` yields the (possibly textured) reflectance at si and is invoked exactly where the reference
// evaluates it.  deriv = true returns d value / d reflectance per channel instead of the value (PRB adjoint, prb.py:202-250): the
// value at reflectance := 1 for the BSDFs that are linear in it (diffuse, rough conductor, the reflection lobe of the rough
// dielectric -- its transmission lobe does not depend on specular_reflectance), the closed form for rough plastic.
template <int CLS = CLS_GENERIC, class RF>
B2_DEV void bsdf_eval_pdf_nested(const DBsdf &b, f3 wi, f3 wo, RF &&refl, f3 &val, float &pdf, bool deriv = false) {
    val = f3(0.f); pdf = 0.f;
    if (b.flags & B2_BSDF_FLAG_TWOSIDED) { wo.z = mulsign(wo.z, wi.z); wi.z = fabsf(wi.z); }
    const float ci = wi.z, co = wo.z;
    const int type = cls_type<CLS>(b.type);
    if (type == B2_BSDF_DIFFUSE) {
        if (!(ci > 0.f && co > 0.f)) return;
        val = (deriv ? f3(1.f) : refl()) * kInvPi * co;
        pdf = kInvPi * co;
    } else if (type == B2_BSDF_ROUGHCONDUCTOR) {
        const f3 H = normalize(wo + wi);
        if (!(ci > 0.f && co > 0.f && dot(wi, H) > 0.f && dot(wo, H) > 0.f)) return;
        const Microfacet distr(b.distribution, b.alpha_u, b.alpha_v, b.sample_visible != 0);
        const float D = distr.eval(H);
        const float g1_wi = distr.smith_g1(wi, H), G = g1_wi * distr.smith_g1(wo, H);
        const float value = D * G / (4.f * ci);
        const f3 F = fresnel_conductor3(dot(wi, H), b);
        const f3 v = f3(value) * (deriv ? f3(1.f) : refl());
        const float p = b.sample_visible ? D * g1_wi / (4.f * ci) : distr.pdf(wi, H) / (4.f * dot(wo, H));
        if (D != 0.f) { val = F * v; pdf = p; }
    } else if (type == B2_BSDF_ROUGHPLASTIC) {
        if (!(ci > 0.f && co > 0.f)) return;
        roughplastic_eval_pdf(b, wi, wo, refl, val, pdf, deriv);
    } else if (type == B2_BSDF_ROUGHDIELECTRIC) {
        // RoughDielectric::eval_pdf (roughdielectric.cpp:503-605)
        if (ci == 0.f) return;
        const bool reflect = ci * co > 0.f;
        const float eta = ci > 0.f ? b.ior : rcp(b.ior), inv_eta = ci > 0.f ? rcp(b.ior) : b.ior;
        f3 m = normalize(wi + wo * (reflect ? 1.f : eta));
        m = f3(mulsign(m.x, m.z), mulsign(m.y, m.z), mulsign(m.z, m.z));
        const float dwi = dot(wi, m), dwo = dot(wo, m);
        if (!(dwi * ci > 0.f && dwo * co > 0.f)) return;
        Microfacet distr(b.distribution, b.alpha_u, b.alpha_v, b.sample_visible != 0);
        const float D = distr.eval(m);
        float F, cos_t, eta_it, eta_ti; fresnel(dwi, b.ior, F, cos_t, eta_it, eta_ti);
        const float G = distr.G(wi, wo, m);
        if (reflect) {
            val = f3(F * D * G / (4.f * fabsf(ci)));
            val = val * (deriv ? f3(1.f) : refl());
        } else {
            const float scale = sqr(inv_eta);
            val = f3(fabsf((scale * (1.f - F) * D * G * eta * eta * dwi * dwo) / (ci * sqr(dwi + eta * dwo))));
            val = deriv ? f3(0.f) : val * f3(b.trans[0], b.trans[1], b.trans[2]);
        }
        if (!b.sample_visible) distr.scale_alpha(1.2f - .2f * sqrtf(fabsf(ci)));
        float p = distr.pdf(ci < 0.f ? -wi : wi, m);
        p *= reflect ? F : 1.f - F;
        const float dwh_dwo = reflect ? rcp(4.f * dwo) : (eta * eta * dwo) / sqr(dwi + eta * dwo);
        pdf = p * fabsf(dwh_dwo);
    }
}

// OpacityMask::eval_pdf (mask.cpp:196-215): nested value and pdf scaled by the opacity
template <int CLS = CLS_GENERIC, class RF>
B2_DEV void bsdf_eval_pdf(const DBsdf &b, f3 wi, f3 wo, RF &&refl, f3 &val, float &pdf, bool deriv = false) {
    bsdf_eval_pdf_nested<CLS>(b, wi, wo, refl, val, pdf, deriv);
    if (b.flags & B2_BSDF_FLAG_MASK) { val = val * b.opacity; pdf *= b.opacity; }
}

// sample.  Same lazy `refl()` convention as bsdf_eval_pdf.
template <int CLS = CLS_GENERIC, class RF>
B2_DEV void bsdf_sample(const DBsdf &b, f3 wi_in, float s1, f2 s2, RF &&refl, BS &bs, f3 &weight) {
    const int type = cls_type<CLS>(b.type);
    f3 wi = wi_in;
    bs.wo = f3(); bs.pdf = 0.f; bs.eta = 0.f; bs.type = 0u; weight = f3(0.f);
    const bool masked = (b.flags & B2_BSDF_FLAG_MASK) != 0;
    if (masked) {           // OpacityMask::sample (mask.cpp:123-171)
        if (!(s1 < b.opacity)) { bs.wo = -wi_in; bs.eta = 1.f; bs.type = LOBE_NULL; bs.pdf = 1.f - b.opacity; weight = f3(1.f); return; }
        s1 /= b.opacity;
    }
    const bool two = (b.flags & B2_BSDF_FLAG_TWOSIDED) != 0;
    if (two) wi.z = fabsf(wi.z);
    const float ci = wi.z;
    if (type == B2_BSDF_DIFFUSE) {
        if (ci > 0.f) {
            bs.wo = square_to_cosine_hemisphere(s2);
            bs.pdf = kInvPi * bs.wo.z;
            bs.eta = 1.f; bs.type = LOBE_DIFFUSE;
            if (bs.pdf > 0.f) weight = refl();
        }
    } else if (type == B2_BSDF_ROUGHCONDUCTOR) {
        if (ci > 0.f) {
            const Microfacet distr(b.distribution, b.alpha_u, b.alpha_v, b.sample_visible != 0);
            const f3 m = distr.sample(wi, s2, bs.pdf);
            bs.wo = fmadd(m, 2.f * dot(wi, m), -wi);
            bs.eta = 1.f; bs.type = LOBE_GLOSSY;
            const bool active = (bs.pdf != 0.f) && bs.wo.z > 0.f;
            const float w = b.sample_visible ? distr.smith_g1(bs.wo, m) : distr.G(wi, bs.wo, m) * dot(wi, m) / (ci * m.z);
            bs.pdf /= 4.f * dot(bs.wo, m);
            const f3 F = fresnel_conductor3(dot(wi, m), b);
            const f3 wv = f3(w) * refl();
            if (active) weight = F * wv;
        }
    } else if (type == B2_BSDF_ROUGHPLASTIC) {
        // RoughPlastic::sample (roughplastic.cpp:264-299)
        if (ci > 0.f) {
            const float t_i = lerp_gather(b.rough_table, ci, 64);
            float prob_specular = (1.f - t_i) * b.spec_weight; const float prob_diffuse = t_i * (1.f - b.spec_weight);
            prob_specular = prob_specular / (prob_specular + prob_diffuse);
            bs.eta = 1.f;
            if (s1 < prob_specular) {
                const Microfacet distr(b.distribution, b.alpha_u, b.alpha_u, b.sample_visible != 0);
                float unused; const f3 m = distr.sample(wi, s2, unused);
                bs.wo = fmadd(m, 2.f * dot(wi, m), -wi); bs.type = LOBE_GLOSSY;
            } else { bs.wo = square_to_cosine_hemisphere(s2); bs.type = LOBE_DIFFUSE; }
            f3 value(0.f); float p = 0.f;
            if (bs.wo.z > 0.f) roughplastic_eval_pdf(b, wi, bs.wo, refl, value, p);
            bs.pdf = p;
            if (p > 0.f) weight = value / f3(p);
        }
    } else if (type == B2_BSDF_ROUGHDIELECTRIC) {
        // RoughDielectric::sample (roughdielectric.cpp:242-357)
        if (ci != 0.f) {
            Microfacet distr(b.distribution, b.alpha_u, b.alpha_v, b.sample_visible != 0), sample_distr = distr;
            if (!b.sample_visible) sample_distr.scale_alpha(1.2f - .2f * sqrtf(fabsf(ci)));
            const f3 m = sample_distr.sample(ci < 0.f ? -wi : wi, s2, bs.pdf);
            const bool active = bs.pdf != 0.f;
            float F, cos_t, eta_it, eta_ti; fresnel(dot(wi, m), b.ior, F, cos_t, eta_it, eta_ti);
            const bool sel_r = s1 <= F && active, sel_t = !sel_r && active;
            bs.pdf *= sel_r ? F : 1.f - F;
            bs.eta = sel_r ? 1.f : eta_it; bs.type = LOBE_GLOSSY;
            f3 w(1.f); float dwh_dwo = 0.f;
            if (sel_r) {
                bs.wo = fmadd(m, 2.f * dot(wi, m), -wi);                             // reflect(wi, m)
                w = w * refl();
                dwh_dwo = rcp(4.f * dot(bs.wo, m));
            }
            if (sel_t) {
                bs.wo = fmadd(m, fma_(dot(wi, m), eta_ti, cos_t), -(wi * eta_ti));   // refract(wi, m, cos_t, eta_ti)
                w = w * (f3(b.trans[0], b.trans[1], b.trans[2]) * sqr(eta_ti));
                dwh_dwo = (sqr(bs.eta) * dot(bs.wo, m)) / sqr(dot(wi, m) + bs.eta * dot(bs.wo, m));
            }
            if (active) {
                w = w * (b.sample_visible ? distr.smith_g1(bs.wo, m) : distr.G(wi, bs.wo, m) * dot(wi, m) / (ci * m.z));
                weight = w;
            }
            bs.pdf *= fabsf(dwh_dwo);
        }
    } else if (type == B2_BSDF_CONDUCTOR) {
        // SmoothConductor::sample (conductor.cpp:249-307)
        if (ci > 0.f) {
            bs.wo = f3(-wi.x, -wi.y, wi.z);
            bs.pdf = 1.f; bs.eta = 1.f; bs.type = LOBE_DELTA_R;
            weight = refl() * fresnel_conductor3(ci, b);
        }
    } else {
        float r_i, cos_t, eta_it, eta_ti; fresnel(ci, b.ior, r_i, cos_t, eta_it, eta_ti);
        const float t_i = 1.f - r_i;
        const bool sel_r = s1 <= r_i;
        bs.pdf = sel_r ? r_i : t_i;
        bs.type = sel_r ? LOBE_DELTA_R : LOBE_DELTA_T;
        bs.wo = sel_r ? f3(-wi.x, -wi.y, wi.z) : f3(-eta_ti * wi.x, -eta_ti * wi.y, cos_t);
        bs.eta = sel_r ? 1.f : eta_it;
        weight = f3(1.f);
        if (sel_r) weight *= f3(b.refl[0], b.refl[1], b.refl[2]);
        else { weight *= f3(b.trans[0], b.trans[1], b.trans[2]); weight *= sqr(eta_ti); }
    }
    if (two) bs.wo.z = mulsign(bs.wo.z, wi_in.z);
    if (masked) bs.pdf *= b.opacity;
}

// -------------------------------------------------------------------------------- environment map
B2_DEV f3 env_texel(const DEnvmap &e, uint32_t i) { const float *p = e.data + 3 * (size_t) i; return f3(__ldg(p), __ldg(p + 1), __ldg(p + 2)); }
B2_DEV f3 env_eval_uv(const DEnvmap &e, f2 uv) {
    const uint32_t rx = e.width, ry = e.height;
    uv.x -= .5f / (float) (rx - 1u);
    uv.x -= floorf(uv.x); uv.y -= floorf(uv.y);
    uv.x *= (float) (rx - 1u); uv.y *= (float) (ry - 1u);
    const uint32_t px = min((uint32_t) uv.x, rx - 2u), py = min((uint32_t) uv.y, ry - 2u);
    const float w1x = uv.x - (float) px, w1y = uv.y - (float) py, w0x = 1.f - w1x, w0y = 1.f - w1y;
    const uint32_t index = py * rx + px;
    const f3 v00 = env_texel(e, index), v10 = env_texel(e, index + 1), v01 = env_texel(e, index + rx), v11 = env_texel(e, index + rx + 1);
    const f3 v0 = fmadd(f3(w0x), v00, v10 * w1x), v1 = fmadd(f3(w0x), v01, v11 * w1x);
    const f3 v = fmadd(f3(w0y), v0, v1 * w1y);
    return v * e.scale;
}
B2_DEV f3 xf3x4_vector(const float *m, f3 v) {
    f3 r = f3(m[0], m[4], m[8]) * v.x;
    r = fmadd(f3(m[1], m[5], m[9]), v.y, r);
    r = fmadd(f3(m[2], m[6], m[10]), v.z, r);
    return r;
}
B2_DEV f3 env_eval_dir(const DEnvmap &e, f3 dir_world) {
    const f3 v = xf3x4_vector(e.from_world, dir_world);
    const f2 uv{ atan2f(v.x, -v.z) * kInvTwoPi, safe_acos(v.y) * kInvPi };
    return env_eval_uv(e, uv);
}
B2_DEV uint32_t warp_index(uint32_t x, uint32_t y, uint32_t width) {
    return ((x & 1u) | (((x & ~1u) | (y & 1u)) << 1)) + ((y & ~1u) * width);
}
B2_DEV f2 warp_sample(const DEnvmap &e, f2 sample, float &pdf) {
    sample.x = clipf(sample.x, 0.f, 1.f); sample.y = clipf(sample.y, 0.f, 1.f);
    uint32_t ox = 0, oy = 0;
    for (int l = e.n_levels - 2; l > 0; --l) {
        const float *lv = e.warp + e.level_offset[l];
        ox <<= 1; oy <<= 1;
        const uint32_t i = warp_index(ox, oy, e.level_width[l]);
        const float4 q = __ldg((const float4 *) (lv + i));           // 2x2 block, 16-byte aligned by construction
        const float v00 = q.x, v10 = q.y, v01 = q.z, v11 = q.w;
        sample.x = clipf(sample.x, 0.f, 1.f); sample.y = clipf(sample.y, 0.f, 1.f);
        const float r0 = v00 + v10, r1 = v01 + v11;
        sample.y *= r0 + r1;
        bool mask = sample.y > r0;
        if (mask) { oy += 1; sample.y -= r0; }
        sample.y /= mask ? r1 : r0;
        const float c0 = mask ? v01 : v00, c1 = mask ? v11 : v10;
        sample.x *= c0 + c1;
        mask = sample.x > c0;
        if (mask) sample.x -= c0;
        sample.x /= mask ? c1 : c0;
        if (mask) ox += 1;
    }
    const float *l0 = e.warp + e.level_offset[0]; const uint32_t w0 = e.level_width[0];
    const uint32_t i = ox + oy * w0;
    sample = square_to_bilinear(__ldg(l0 + i), __ldg(l0 + i + 1), __ldg(l0 + i + w0), __ldg(l0 + i + w0 + 1), sample, pdf);
    const float psx = 1.f / (float) (e.width - 1), psy = 1.f / (float) (e.height - 1);
    return f2{ ((float) (int) ox + sample.x) * psx, ((float) (int) oy + sample.y) * psy };
}
// Fast mode (B2_ENV_WARP_ALIAS): the bilinear patch is drawn from a Vose alias table over the patch masses (mean of the
// four corner values = the integral of the interpolant over the patch) -- two dependent 4-byte loads instead of the
// ten dependent 2x2-block loads of the hierarchical descent -- and the position inside the patch by the same
// square_to_bilinear.  The density is the same function warp_eval() returns, so MIS stays consistent; only the mapping
// (u1, u2) -> direction differs from the reference, i.e. parity holds in expectation, not per sample.
B2_DEV f2 alias_sample(const DEnvmap &e, f2 sample, float &pdf) {
    const uint32_t pw = (uint32_t) e.width - 1u, ph = (uint32_t) e.height - 1u, n = pw * ph;
    sample.x = clipf(sample.x, 0.f, 1.f); sample.y = clipf(sample.y, 0.f, 1.f);
    const float scaled = sample.x * (float) n;
    uint32_t bin = min((uint32_t) scaled, n - 1u);
    float u = scaled - (float) bin;
    const float q = __ldg(e.alias_prob + bin);
    if (u < q) u = u / q;
    else { u = (u - q) / (1.f - q); bin = __ldg(e.alias_idx + bin); }
    u = clipf(u, 0.f, 1.f);
    const uint32_t oy = bin / pw, ox = bin - oy * pw;
    const float *l0 = e.warp + e.level_offset[0]; const uint32_t w0 = e.level_width[0];
    const uint32_t i = ox + oy * w0;
    f2 s = square_to_bilinear(__ldg(l0 + i), __ldg(l0 + i + 1), __ldg(l0 + i + w0), __ldg(l0 + i + w0 + 1), f2{ u, sample.y }, pdf);
    const float psx = 1.f / (float) pw, psy = 1.f / (float) ph;
    return f2{ ((float) (int) ox + s.x) * psx, ((float) (int) oy + s.y) * psy };
}
B2_DEV float warp_eval(const DEnvmap &e, f2 pos) {
    pos.x = clipf(pos.x, 0.f, 1.f); pos.y = clipf(pos.y, 0.f, 1.f);
    pos.x *= (float) (e.width - 1); pos.y *= (float) (e.height - 1);
    const uint32_t ox = min((uint32_t) (int) pos.x, (uint32_t) e.width - 2u), oy = min((uint32_t) (int) pos.y, (uint32_t) e.height - 2u);
    pos.x -= (float) (int) ox; pos.y -= (float) (int) oy;
    const float *l0 = e.warp + e.level_offset[0]; const uint32_t w0 = e.level_width[0];
    const uint32_t i = ox + oy * w0;
    return square_to_bilinear_pdf(__ldg(l0 + i), __ldg(l0 + i + 1), __ldg(l0 + i + w0), __ldg(l0 + i + w0 + 1), pos);
}

// -------------------------------------------------------------------------------- emitters
template <bool COUNT>
B2_DEV f3 emitter_eval(const DScene &sc, int em, const SI &si, Counters<COUNT> &cnt) {
    if (em < 0) return f3(0.f);
    const DEmitter &e = sc.emitters[em];
    if (e.type == B2_EMITTER_AREA) return si.wi.z > 0.f ? f3(e.radiance[0], e.radiance[1], e.radiance[2]) : f3(0.f);
    cnt.add(C_ESCAPES);
    if (e.type == B2_EMITTER_CONSTANT) return f3(e.radiance[0], e.radiance[1], e.radiance[2]);
    return env_eval_dir(sc.env, -si.wi);
}

// Emitter::sample_direction without the visibility test
template <bool COUNT>
B2_DEV void emitter_sample_direction(const DScene &sc, int em, f3 ref_p, f2 sample, int env_warp, DS &ds, f3 &spec, Counters<COUNT> &cnt) {
    const DEmitter &e = sc.emitters[em];
    ds.emitter = em; ds.delta = false; ds.pdf = 0.f; ds.dist = 0.f; spec = f3(0.f);
    if (e.type == B2_EMITTER_AREA) {
        cnt.add(C_NEE_AREA);
        if (e.shape_kind == B2_SHAPE_RECTANGLE) {
            const float *m = e.to_world; const f3 q(fma_(sample.x, 2.f, -1.f), fma_(sample.y, 2.f, -1.f), 0.f);
            f3 r(m[3], m[7], m[11]);
            r = fmadd(f3(m[0], m[4], m[8]), q.x, r); r = fmadd(f3(m[1], m[5], m[9]), q.y, r); r = fmadd(f3(m[2], m[6], m[10]), q.z, r);
            ds.p = r; ds.n = f3(e.normal[0], e.normal[1], e.normal[2]);
        } else {
            const DMesh m = sc.meshes[e.mesh];
            const float *cdf = sc.area_cdf + e.cdf_offset; const uint32_t nf = m.n_faces;
            const float total = __ldg(cdf + nf - 1), value = sample.y * total, nrmz = rcp(total);
            uint32_t lo = 0, hi = nf - 1;
            while (lo < hi) { const uint32_t mid = (lo + hi) / 2; if (__ldg(cdf + mid) < value) lo = mid + 1; else hi = mid; }
            const uint32_t fi = lo; const float prev = fi ? __ldg(cdf + fi - 1) : 0.f;
            const float sy = (sample.y - prev * nrmz) / ((__ldg(cdf + fi) - prev) * nrmz);
            const uint32_t f = m.first_face + fi;
            const uint32_t i0 = __ldg(sc.faces + 3 * (size_t) f), i1 = __ldg(sc.faces + 3 * (size_t) f + 1), i2 = __ldg(sc.faces + 3 * (size_t) f + 2);
            const f3 p0 = load3(sc.pos, i0), p1 = load3(sc.pos, i1), p2 = load3(sc.pos, i2);
            const f3 e0 = p1 - p0, e1 = p2 - p0;
            const f2 b = square_to_uniform_triangle(f2{ sample.x, sy });
            ds.p = fmadd(e0, b.x, fmadd(e1, b.y, p0));
            ds.n = normalize(cross(e0, e1));
            if (m.flags & B2_MESH_HAS_NORMALS) {
                const f3 n0 = load3(sc.nrm, i0), n1 = load3(sc.nrm, i1), n2 = load3(sc.nrm, i2);
                ds.n = normalize(fmadd(n0, 1.f - b.x - b.y, fmadd(n1, b.x, n2 * b.y)));
            }
            if (m.flags & B2_MESH_FLIP_NORMALS) ds.n = -ds.n;
        }
        ds.d = ds.p - ref_p;
        const float dist2 = squared_norm(ds.d); ds.dist = sqrtf(dist2); ds.d = ds.d / ds.dist;
        const float dn = dot(ds.d, ds.n), dp = fabsf(dn), x = dist2 / dp;
        ds.pdf = e.inv_area * (finite_(x) ? x : 0.f);
        if (dn < 0.f && ds.pdf != 0.f) spec = f3(e.radiance[0], e.radiance[1], e.radiance[2]) / ds.pdf;
    } else if (e.type == B2_EMITTER_ENVMAP) {
        cnt.add(C_NEE_ENV);
        const DEnvmap &env = sc.env;
        float pdf; f2 uv = (env_warp == B2_ENV_WARP_ALIAS && env.alias_prob) ? alias_sample(env, sample, pdf) : warp_sample(env, sample, pdf);
        uv.x += .5f / (float) (env.width - 1);
        const bool active = pdf > 0.f;
        const float theta = uv.y * kPi, phi = uv.x * kTwoPi;
        float st, ct, sp, cp; sincosf(theta, &st, &ct); sincosf(phi, &sp, &cp);
        const f3 dsph(cp * st, sp * st, ct);
        f3 dir(dsph.y, dsph.z, -dsph.x);
        const f3 bc(env.bsphere_center[0], env.bsphere_center[1], env.bsphere_center[2]);
        const float radius = fmaxf(env.bsphere_radius, norm(ref_p - bc)), dist = 2.f * radius;
        const float inv_sin_theta = safe_rsqrt(fmaxf(sqr(dir.x) + sqr(dir.z), sqr(kEps)));
        dir = xf3x4_vector(env.to_world, dir);
        ds.p = ref_p + dir * dist; ds.n = -dir;
        ds.pdf = active ? pdf * inv_sin_theta * (1.f / (2.f * sqr(kPi))) : 0.f;
        ds.d = dir; ds.dist = dist;
        if (active) spec = env_eval_uv(env, uv) / ds.pdf;
    } else if (e.type == B2_EMITTER_POINT) {
        // PointLight::sample_direction (emitters/point.cpp:119-148): delta position, pdf 1, intensity / dist^2
        cnt.add(C_NEE_AREA);
        ds.p = f3(e.to_world[3], e.to_world[7], e.to_world[11]); ds.n = f3(0.f); ds.pdf = 1.f; ds.delta = true;
        ds.d = ds.p - ref_p;
        const float dist2 = dot(ds.d, ds.d), inv_dist = 1.f / sqrtf(dist2);
        ds.dist = sqrtf(dist2); ds.d = ds.d * inv_dist;
        spec = f3(e.radiance[0], e.radiance[1], e.radiance[2]) * sqr(inv_dist);
    } else {
        cnt.add(C_NEE_ENV);
        const DEnvmap &env = sc.env;
        const float z = fma_(-2.f, sample.y, 1.f), r = safe_sqrt(fma_(-z, z, 1.f));
        float sa, ca; sincosf(kTwoPi * sample.x, &sa, &ca);
        const f3 dir(r * ca, r * sa, z);
        const f3 bc(env.bsphere_center[0], env.bsphere_center[1], env.bsphere_center[2]);
        const float radius = fmaxf(env.bsphere_radius, norm(ref_p - bc)), dist = 2.f * radius;
        ds.p = ref_p + dir * dist; ds.n = -dir; ds.pdf = 0.25f * kInvPi; ds.d = dir; ds.dist = dist;
        spec = f3(e.radiance[0], e.radiance[1], e.radiance[2]) / ds.pdf;
    }
}

// DirectionSample3f(scene, si, ref) (records.h:173-180)
B2_DEV void direction_sample(const SI &si, f3 ref_p, DS &ds) {
    ds.p = si.p; ds.n = si.sh.n; ds.pdf = 0.f; ds.delta = false;
    const f3 rel = si.p - ref_p; ds.dist = norm(rel);
    ds.d = si.valid ? rel / ds.dist : -si.wi;
    ds.emitter = si.emitter;
}
// Scene::pdf_emitter_direction (scene.cpp:381-392)
B2_DEV float pdf_emitter_direction(const DScene &sc, const DS &ds, bool active) {
    if (!active || ds.emitter < 0) return 0.f;
    const DEmitter &e = sc.emitters[ds.emitter];
    const float pmf = 1.f / (float) sc.n_emitters;
    float value;
    if (e.type == B2_EMITTER_AREA) {
        const float dp = dot(ds.d, ds.n);
        if (!(dp < 0.f)) return 0.f;
        const float adp = fabsf(dp);
        value = e.inv_area * (adp != 0.f ? (ds.dist * ds.dist) / adp : 0.f);
    } else if (e.type == B2_EMITTER_ENVMAP) {
        const DEnvmap &env = sc.env;
        const f3 dd = xf3x4_vector(env.from_world, ds.d);
        f2 uv{ atan2f(dd.x, -dd.z) * kInvTwoPi, safe_acos(dd.y) * kInvPi };
        uv.x -= .5f / (float) (env.width - 1u);
        uv.x -= floorf(uv.x); uv.y -= floorf(uv.y);
        const float inv_sin_theta = safe_rsqrt(fmaxf(sqr(dd.x) + sqr(dd.z), sqr(kEps)));
        value = warp_eval(env, uv) * inv_sin_theta * (1.f / (2.f * sqr(kPi)));
    } else value = 0.25f * kInvPi;
    return value * pmf;
}

// Scene::sample_emitter_direction (scene.cpp:319-369): emitter choice (:252-275), sample, shadow ray
// `occluded(ray)` answers the visibility query: the megakernel traces the shadow ray on the spot, the wavefront shade
// kernel records it for the next trace launch and answers "visible" (the contribution is then added by the trace
// kernel only if the ray arrives; an occluded sample contributes exactly zero in either formulation).
template <bool COUNT, class OCC>
B2_DEV void sample_emitter_direction(const DScene &sc, const SI &ref, f2 sample, bool active, int env_warp, DS &ds, f3 &spec, Counters<COUNT> &cnt, OCC &&occluded) {
    ds.p = f3(); ds.n = f3(); ds.d = f3(); ds.pdf = 0.f; ds.dist = 0.f; ds.delta = false; ds.emitter = -1; spec = f3(0.f);
    const uint32_t n = sc.n_emitters;
    if (!active || n == 0) return;
    uint32_t index = 0; float weight = 1.f;
    if (n >= 2) {
        const float nf = (float) n, scaled = sample.x * nf;
        index = min((uint32_t) scaled, n - 1u);
        weight = nf; sample.x = scaled - (float) index;
    }
    emitter_sample_direction(sc, (int) index, ref.p, sample, env_warp, ds, spec, cnt);
    ds.pdf *= 1.f / (float) n;
    spec *= weight;
    if (ds.pdf != 0.f) {
        const Ray sr = spawn_ray_to(ref.p, ref.n, ds.p);
        if (occluded(sr)) { spec = f3(0.f); ds.pdf = 0.f; }
    }
}

}  // namespace b2
