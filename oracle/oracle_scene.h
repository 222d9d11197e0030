// ORACLE (test infrastructure, NOT product code): scene access, CPU BVH, surface interaction,
// textures, BSDFs and emitters -- a scalar CPU restatement of the reference plugins.
// Each function cites the reference file:line (relative to /root/reference/mitsuba3) it follows.
#pragma once
#include "../include/b2path.h"
#include "om.h"
#include <algorithm>
#include <vector>

namespace orc {

struct Ray { V3 o, d; float maxt; };
struct Hit { float t, u, v; uint32_t face; bool valid() const { return face != 0xffffffffu; } };

enum { LOBE_DIFFUSE = 1, LOBE_GLOSSY = 2, LOBE_DELTA_R = 4, LOBE_DELTA_T = 8, LOBE_NULL = 16,
       LOBE_SMOOTH = LOBE_DIFFUSE | LOBE_GLOSSY, LOBE_DELTA = LOBE_DELTA_R | LOBE_DELTA_T | LOBE_NULL };   // BSDFFlags::Delta includes Null

struct SI {                 // SurfaceInteraction3f subset (interaction.h)
    bool valid = false;
    float t = Inf;
    V3 p, n, wi, dp_du, dp_dv;
    V2 uv{0, 0};
    Frame sh;
    uint32_t face = 0xffffffffu;
    int mesh = -1;
};
struct DS {                 // DirectionSample3f (records.h)
    V3 p, n, d; V2 uv{0, 0};
    float pdf = 0, dist = 0; bool delta = false; int emitter = -1;
};
struct BS { V3 wo; float pdf = 0, eta = 0; uint32_t type = 0; };

struct Counters {
    uint64_t samples = 0, closest = 0, shadow = 0, vertices = 0, tex = 0, nee_area = 0, nee_env = 0,
             escapes = 0, taps = 0, grad_atomics = 0, nodes = 0, tris = 0;
    void add(const Counters &o) {
        samples += o.samples; closest += o.closest; shadow += o.shadow; vertices += o.vertices; tex += o.tex;
        nee_area += o.nee_area; nee_env += o.nee_env; escapes += o.escapes; taps += o.taps;
        grad_atomics += o.grad_atomics; nodes += o.nodes; tris += o.tris;
    }
};

// ---------------------------------------------------------------------------------------------
// Binary BVH (binned SAH).  Stands in for Embree (scene_embree.inl:98-118,245-433): any exact
// closest-hit structure returns the same (t,u,v) as long as the triangle test is the same.
// ---------------------------------------------------------------------------------------------
struct BVHNode { float lo[3], hi[3]; uint32_t left, count; /* count>0: leaf, left = first prim */ };

// debug recorder (tools/diag_*.py): when set, every closest-hit query of this thread appends ray(7) + hit(4)
struct TraceLog { float *out; int max_records, n; };
static thread_local TraceLog *g_trace_log = nullptr;

static inline void rough_tables(int type, float alpha, float eta, float *trans, float &internal_reflectance);
struct RoughTab { int type; float alpha, eta; float trans[64]; float internal_reflectance; };

struct Scene {
    const b2_scene_desc *d = nullptr;
    mutable std::vector<RoughTab> rough;  // rough-plastic tables by (distribution, alpha, eta); filled in init()
    const RoughTab &rough_of(const b2_bsdf &b) const {
        for (const RoughTab &t : rough) if (t.type == b.distribution && t.alpha == b.alpha_u && t.eta == b.ior) return t;
        RoughTab t; t.type = b.distribution; t.alpha = b.alpha_u; t.eta = b.ior;     // (single-threaded callers only: init, test hooks)
        rough_tables(t.type, t.alpha, t.eta, t.trans, t.internal_reflectance);
        rough.push_back(t);
        return rough.back();
    }
    std::vector<BVHNode> nodes;
    std::vector<uint32_t> prims;          // face indices in leaf order
    std::vector<int> face_mesh;           // face -> mesh index
    int env_emitter = -1;                 // index of the environment emitter in the emitter list

    V3 pos(uint32_t v) const { return V3(d->positions[3 * v], d->positions[3 * v + 1], d->positions[3 * v + 2]); }
    V3 nrm(uint32_t v) const { return V3(d->normals[3 * v], d->normals[3 * v + 1], d->normals[3 * v + 2]); }
    V2 tc(uint32_t v) const { return V2{ d->texcoords[2 * v], d->texcoords[2 * v + 1] }; }

    void init(const b2_scene_desc *desc) {
        d = desc;
        face_mesh.assign(d->n_faces, -1);
        for (uint32_t m = 0; m < d->n_meshes; ++m)
            for (uint32_t f = 0; f < d->meshes[m].n_faces; ++f) face_mesh[d->meshes[m].first_face + f] = (int) m;
        for (uint32_t e = 0; e < d->n_emitters; ++e)
            if (d->emitters[e].type == B2_EMITTER_ENVMAP || d->emitters[e].type == B2_EMITTER_CONSTANT) env_emitter = (int) e;
        rough.reserve(d->n_bsdfs + 1);
        for (uint32_t i = 0; i < d->n_bsdfs; ++i) if (d->bsdfs[i].type == B2_BSDF_ROUGHPLASTIC) rough_of(d->bsdfs[i]);
        build();
    }

    // ---- build
    struct PrimInfo { float lo[3], hi[3], c[3]; uint32_t face; };
    void build() {
        nodes.clear(); prims.clear();
        uint32_t n = d->n_faces;
        if (n == 0) return;
        std::vector<PrimInfo> info(n);
        for (uint32_t f = 0; f < n; ++f) {
            PrimInfo &pi = info[f]; pi.face = f;
            for (int k = 0; k < 3; ++k) { pi.lo[k] = Inf; pi.hi[k] = -Inf; }
            for (int j = 0; j < 3; ++j) {
                V3 p = pos(d->faces[3 * f + j]);
                for (int k = 0; k < 3; ++k) { pi.lo[k] = fminf(pi.lo[k], p[k]); pi.hi[k] = fmaxf(pi.hi[k], p[k]); }
            }
            for (int k = 0; k < 3; ++k) pi.c[k] = 0.5f * (pi.lo[k] + pi.hi[k]);
        }
        nodes.reserve(2 * n);
        nodes.push_back(BVHNode());
        build_rec(0, info, 0, n);
        prims.resize(n);
        for (uint32_t i = 0; i < n; ++i) prims[i] = info[i].face;
    }
    static float half_area(const float *lo, const float *hi) {
        float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        return dx * dy + dy * dz + dz * dx;
    }
    void build_rec(uint32_t ni, std::vector<PrimInfo> &info, uint32_t b, uint32_t e) {
        float lo[3] = { Inf, Inf, Inf }, hi[3] = { -Inf, -Inf, -Inf }, clo[3] = { Inf, Inf, Inf }, chi[3] = { -Inf, -Inf, -Inf };
        for (uint32_t i = b; i < e; ++i)
            for (int k = 0; k < 3; ++k) {
                lo[k] = fminf(lo[k], info[i].lo[k]); hi[k] = fmaxf(hi[k], info[i].hi[k]);
                clo[k] = fminf(clo[k], info[i].c[k]); chi[k] = fmaxf(chi[k], info[i].c[k]);
            }
        for (int k = 0; k < 3; ++k) { nodes[ni].lo[k] = lo[k]; nodes[ni].hi[k] = hi[k]; }
        uint32_t n = e - b;
        int axis = 0;
        for (int k = 1; k < 3; ++k) if (chi[k] - clo[k] > chi[axis] - clo[axis]) axis = k;
        if (n <= 2 || !(chi[axis] > clo[axis])) {
            if (n <= 4) { nodes[ni].left = b; nodes[ni].count = n; return; }
            // degenerate centroids: split in the middle
            uint32_t mid = b + n / 2;
            uint32_t l = (uint32_t) nodes.size(); nodes.push_back(BVHNode()); nodes.push_back(BVHNode());
            nodes[ni].left = l; nodes[ni].count = 0;
            build_rec(l, info, b, mid); build_rec(l + 1, info, mid, e);
            return;
        }
        const int NB = 16;
        struct Bin { float lo[3], hi[3]; uint32_t n; } bins[NB];
        for (auto &bn : bins) { bn.n = 0; for (int k = 0; k < 3; ++k) { bn.lo[k] = Inf; bn.hi[k] = -Inf; } }
        float scale = NB / (chi[axis] - clo[axis]);
        auto bin_of = [&](const PrimInfo &p) { int i = (int) ((p.c[axis] - clo[axis]) * scale); return std::min(std::max(i, 0), NB - 1); };
        for (uint32_t i = b; i < e; ++i) {
            Bin &bn = bins[bin_of(info[i])]; bn.n++;
            for (int k = 0; k < 3; ++k) { bn.lo[k] = fminf(bn.lo[k], info[i].lo[k]); bn.hi[k] = fmaxf(bn.hi[k], info[i].hi[k]); }
        }
        float right_area[NB]; uint32_t right_n[NB];
        { float l2[3] = { Inf, Inf, Inf }, h2[3] = { -Inf, -Inf, -Inf }; uint32_t c = 0;
          for (int i = NB - 1; i > 0; --i) { c += bins[i].n; for (int k = 0; k < 3; ++k) { l2[k] = fminf(l2[k], bins[i].lo[k]); h2[k] = fmaxf(h2[k], bins[i].hi[k]); }
              right_area[i] = c ? half_area(l2, h2) : 0.f; right_n[i] = c; } }
        float best = Inf; int best_split = -1;
        { float l2[3] = { Inf, Inf, Inf }, h2[3] = { -Inf, -Inf, -Inf }; uint32_t c = 0;
          for (int i = 0; i < NB - 1; ++i) { c += bins[i].n; for (int k = 0; k < 3; ++k) { l2[k] = fminf(l2[k], bins[i].lo[k]); h2[k] = fmaxf(h2[k], bins[i].hi[k]); }
              if (c == 0 || right_n[i + 1] == 0) continue;
              float cost = half_area(l2, h2) * c + right_area[i + 1] * right_n[i + 1];
              if (cost < best) { best = cost; best_split = i; } } }
        uint32_t mid;
        if (best_split < 0) mid = b + n / 2;
        else {
            auto it = std::partition(info.begin() + b, info.begin() + e, [&](const PrimInfo &p) { return bin_of(p) <= best_split; });
            mid = (uint32_t) (it - info.begin());
            if (mid == b || mid == e) mid = b + n / 2;
        }
        if (n <= 4 && best >= half_area(lo, hi) * n) { nodes[ni].left = b; nodes[ni].count = n; return; }
        uint32_t l = (uint32_t) nodes.size(); nodes.push_back(BVHNode()); nodes.push_back(BVHNode());
        nodes[ni].left = l; nodes[ni].count = 0;
        build_rec(l, info, b, mid); build_rec(l + 1, info, mid, e);
    }

    // ---- Mesh::moeller_trumbore (include/mitsuba/render/mesh.h:469-492)
    bool tri_test(const Ray &ray, uint32_t f, float maxt, float &t, float &u, float &v) const {
        V3 p0 = pos(d->faces[3 * f]), p1 = pos(d->faces[3 * f + 1]), p2 = pos(d->faces[3 * f + 2]);
        V3 e1 = p1 - p0, e2 = p2 - p0;
        V3 pvec = cross(ray.d, e2);
        float inv_det = rcp(dot(e1, pvec));
        V3 tvec = ray.o - p0;
        u = dot(tvec, pvec) * inv_det;
        bool ok = u >= 0.f && u <= 1.f;
        V3 qvec = cross(tvec, e1);
        v = dot(ray.d, qvec) * inv_det;
        ok = ok && v >= 0.f && u + v <= 1.f;
        t = dot(e2, qvec) * inv_det;
        ok = ok && t >= 0.f && t <= maxt;
        return ok;
    }
    static bool box_test(const BVHNode &n, const Ray &r, const float *inv, float tmax) {
        float t0 = 0.f, t1 = tmax;
        for (int k = 0; k < 3; ++k) {
            float a = (n.lo[k] - r.o[k]) * inv[k], b = (n.hi[k] - r.o[k]) * inv[k];
            float tn = fminf(a, b), tf = fmaxf(a, b);       // fmin/fmax drop NaNs (0 * inf)
            t0 = fmaxf(t0, tn); t1 = fminf(t1, tf * 1.0000004f);
        }
        return t0 <= t1;
    }
    // Scene::ray_intersect_preliminary (scene.cpp:209-224): closest hit in [0, maxt]
    Hit intersect(const Ray &ray, Counters &c) const {
        c.closest++;
        Hit best{ ray.maxt, 0, 0, 0xffffffffu };
        if (nodes.empty()) return best;
        float inv[3] = { 1.f / ray.d.x, 1.f / ray.d.y, 1.f / ray.d.z };
        uint32_t stack[128]; int sp = 0; stack[sp++] = 0;
        while (sp) {
            const BVHNode &n = nodes[stack[--sp]]; c.nodes++;
            if (!box_test(n, ray, inv, best.t)) continue;
            if (n.count) {
                for (uint32_t i = 0; i < n.count; ++i) {
                    uint32_t f = prims[n.left + i]; float t, u, v; c.tris++;
                    // deterministic tie rule: equal t -> lower face index wins, independent of traversal order
                    if (tri_test(ray, f, ray.maxt, t, u, v) && (!best.valid() || t < best.t || (t == best.t && f < best.face))) best = Hit{ t, u, v, f };
                }
            } else { stack[sp++] = n.left; stack[sp++] = n.left + 1; }
        }
        if (g_trace_log && g_trace_log->n < g_trace_log->max_records) {
            float *o = g_trace_log->out + 11 * g_trace_log->n++;
            o[0] = ray.o.x; o[1] = ray.o.y; o[2] = ray.o.z; o[3] = ray.d.x; o[4] = ray.d.y; o[5] = ray.d.z; o[6] = ray.maxt;
            o[7] = best.t; o[8] = best.u; o[9] = best.v; std::memcpy(o + 10, &best.face, 4);
        }
        return best;
    }
    // Scene::ray_test (scene.cpp:226-235)
    bool occluded(const Ray &ray, Counters &c) const {
        c.shadow++;
        if (nodes.empty()) return false;
        float inv[3] = { 1.f / ray.d.x, 1.f / ray.d.y, 1.f / ray.d.z };
        uint32_t stack[128]; int sp = 0; stack[sp++] = 0;
        while (sp) {
            const BVHNode &n = nodes[stack[--sp]]; c.nodes++;
            if (!box_test(n, ray, inv, ray.maxt)) continue;
            if (n.count) {
                for (uint32_t i = 0; i < n.count; ++i) { float t, u, v; c.tris++; if (tri_test(ray, prims[n.left + i], ray.maxt, t, u, v)) return true; }
            } else { stack[sp++] = n.left; stack[sp++] = n.left + 1; }
        }
        return false;
    }

    // ---- PreliminaryIntersection::compute_surface_interaction (interaction.h:712-738)
    //      -> Mesh::compute_surface_interaction (mesh.cpp:1386-1560) -> finalize (interaction.h:497-516)
    SI make_si(const Ray &ray, const Hit &h, Counters &c) const {
        SI si;
        if (!h.valid()) { si.wi = -ray.d; return si; }
        c.vertices++;
        uint32_t f = h.face; const b2_mesh &m = d->meshes[face_mesh[f]];
        uint32_t i0 = d->faces[3 * f], i1 = d->faces[3 * f + 1], i2 = d->faces[3 * f + 2];
        V3 p0 = pos(i0), p1 = pos(i1), p2 = pos(i2);
        float b1 = h.u, b2 = h.v, b0 = 1.f - b1 - b2;
        si.valid = true; si.face = f; si.mesh = face_mesh[f]; si.t = h.t;
        si.p = fmadd(p0, b0, fmadd(p1, b1, p2 * b2));
        si.n = normalize(cross(p1 - p0, p2 - p0));                     // mesh.h:157-165
        si.uv = V2{ b1, b2 };
        coordinate_system(si.n, si.dp_du, si.dp_dv);
        V3 dp0 = p1 - p0, dp1 = p2 - p0;
        if (m.flags & B2_MESH_HAS_UV) {
            V2 uv0 = tc(i0), uv1 = tc(i1), uv2 = tc(i2);
            si.uv.x = fmaf(uv2.x, b2, fmaf(uv1.x, b1, uv0.x * b0));
            si.uv.y = fmaf(uv2.y, b2, fmaf(uv1.y, b1, uv0.y * b0));
            V2 duv0{ uv1.x - uv0.x, uv1.y - uv0.y }, duv1{ uv2.x - uv0.x, uv2.y - uv0.y };
            float det = fmaf(duv0.x, duv1.y, -(duv0.y * duv1.x)), inv_det = rcp(det);
            if (det != 0.f) {
                si.dp_du = fmadd(V3(duv1.y), dp0, -(dp1 * duv0.y)) * inv_det;   // fmsub(duv1.y, dp0, duv0.y*dp1)
                si.dp_dv = fmadd(V3(-duv1.x), dp0, dp1 * duv0.x) * inv_det;     // fnmadd(duv1.x, dp0, duv0.x*dp1)
            }
        }
        if (m.flags & B2_MESH_HAS_NORMALS) {
            V3 n0 = nrm(i0), n1 = nrm(i1), n2 = nrm(i2);
            V3 n = fmadd(n2, b2, fmadd(n1, b1, n0 * b0));
            float il = rsqrt(squared_norm(n));
            si.sh.n = n * il;
        } else si.sh.n = si.n;
        if (m.flags & B2_MESH_FLIP_NORMALS) { si.n = -si.n; si.sh.n = -si.sh.n; }
        // initialize_sh_frame (interaction.h:278-288)
        si.sh.s = normalize(fmadd(si.sh.n, -dot(si.sh.n, si.dp_du), si.dp_du));
        if (si.dp_du.x == 0.f && si.dp_du.y == 0.f && si.dp_du.z == 0.f) { V3 t; coordinate_system(si.sh.n, si.sh.s, t); }
        si.sh.t = cross(si.sh.n, si.sh.s);
        si.wi = si.sh.to_local(-ray.d);
        return si;
    }

    // Interaction::offset_p / spawn_ray / spawn_ray_to (interaction.h:140-169)
    static V3 offset_p(V3 p, V3 n, V3 dir) {
        float mag = (1.f + hmax(V3(fabsf(p.x), fabsf(p.y), fabsf(p.z)))) * RayEpsilon;
        mag = mulsign(mag, dot(n, dir));
        return fmadd(V3(mag), n, p);
    }
    static Ray spawn_ray(V3 p, V3 n, V3 dir) { return Ray{ offset_p(p, n, dir), dir, Largest }; }
    static Ray spawn_ray_to(V3 p, V3 n, V3 target) {
        V3 o = offset_p(p, n, target - p);
        V3 dd = target - o; float dist = norm(dd); dd = dd / dist;
        return Ray{ o, dd, dist * (1.f - ShadowEpsilon) };
    }

    // ---- dr::Texture::eval_nonaccel (drjit/texture.h:478-551) + wrap (:1360-1389) via
    //      BitmapTextureImpl::interpolate_3 (textures/bitmap.cpp:783-800).  Returns texel indices+weights for PRB.
    static float half_to_float(uint16_t h) {
        uint32_t s = (h >> 15) & 1, e = (h >> 10) & 31, m = h & 1023, bits;
        if (e == 0) { if (m == 0) bits = s << 31; else { int sh = 0; while (!(m & 1024)) { m <<= 1; sh++; } m &= 1023; bits = (s << 31) | ((uint32_t) (127 - 15 - sh + 1) << 23) | (m << 13); } }
        else if (e == 31) bits = (s << 31) | 0x7f800000u | (m << 13);
        else bits = (s << 31) | ((e + 112) << 23) | (m << 13);
        float f; std::memcpy(&f, &bits, 4); return f;
    }
    static int wrap_coord(int pos, int res, int mode) {
        if (mode == B2_WRAP_CLAMP) return std::min(std::max(pos, 0), res - 1);
        int shift = pos < 0 ? pos + 1 : pos, div = shift / res, mod = pos - div * res;
        if (mod < 0) mod += res;
        if (mode == B2_WRAP_MIRROR) mod = (((div & 1) == 0) != (pos < 0)) ? mod : res - 1 - mod;
        return mod;
    }
    V3 texel(const b2_texture &t, uint32_t idx) const {
        if (t.format == B2_TEX_F16) { const uint16_t *p = (const uint16_t *) t.data + 3 * (size_t) idx; return V3(half_to_float(p[0]), half_to_float(p[1]), half_to_float(p[2])); }
        const float *p = (const float *) t.data + 3 * (size_t) idx; return V3(p[0], p[1], p[2]);
    }
    V3 tex_eval(int ti, V2 uv_, uint32_t *idx_out, float *w_out, Counters &c) const {
        const b2_texture &t = d->textures[ti]; c.tex++;
        const float *m = t.to_uv;            // Transform3f::transform_affine(Point2f)
        V2 uv{ fmaf(m[1], uv_.y, fmaf(m[0], uv_.x, m[2])), fmaf(m[4], uv_.y, fmaf(m[3], uv_.x, m[5])) };
        uint32_t idx[4]; float w[4];
        if (t.filter == B2_FILTER_NEAREST) {
            int x = wrap_coord((int) floorf(uv.x * t.width), t.width, t.wrap), y = wrap_coord((int) floorf(uv.y * t.height), t.height, t.wrap);
            idx[0] = idx[1] = idx[2] = idx[3] = (uint32_t) y * t.width + x; w[0] = 1; w[1] = w[2] = w[3] = 0;
            if (idx_out) for (int i = 0; i < 4; ++i) { idx_out[i] = idx[i]; w_out[i] = w[i]; }
            return texel(t, idx[0]);
        }
        float px = fmaf(uv.x, (float) t.width, -.5f), py = fmaf(uv.y, (float) t.height, -.5f);
        int ix = (int) floorf(px), iy = (int) floorf(py);
        float w1x = px - ix, w1y = py - iy, w0x = 1.f - w1x, w0y = 1.f - w1y;
        int x0 = wrap_coord(ix, t.width, t.wrap), x1 = wrap_coord(ix + 1, t.width, t.wrap);
        int y0 = wrap_coord(iy, t.height, t.wrap), y1 = wrap_coord(iy + 1, t.height, t.wrap);
        idx[0] = y0 * t.width + x0; idx[1] = y0 * t.width + x1; idx[2] = y1 * t.width + x0; idx[3] = y1 * t.width + x1;
        w[0] = w0x * w0y; w[1] = w1x * w0y; w[2] = w0x * w1y; w[3] = w1x * w1y;
        V3 out(0.f);
        for (int i = 0; i < 4; ++i) out = fmadd(texel(t, idx[i]), w[i], out);
        if (idx_out) for (int i = 0; i < 4; ++i) { idx_out[i] = idx[i]; w_out[i] = w[i]; }
        return out;
    }
    V3 reflectance(const b2_bsdf &b, const SI &si, Counters &c) const {
        if (b.reflectance_tex >= 0) return tex_eval(b.reflectance_tex, si.uv, nullptr, nullptr, c);
        return V3(b.reflectance[0], b.reflectance[1], b.reflectance[2]);
    }
};

// ---------------------------------------------------------------------------------------------
// Fresnel (include/mitsuba/render/fresnel.h:34-71, 92-117, 276-295)
// ---------------------------------------------------------------------------------------------
static inline void fresnel(float cos_i, float eta, float &r, float &cos_t, float &eta_it, float &eta_ti) {
    bool outside = cos_i >= 0.f;
    float rcp_eta = rcp(eta); eta_it = outside ? eta : rcp_eta; eta_ti = outside ? rcp_eta : eta;
    float cos_t_sqr = fmaf(-fmaf(-cos_i, cos_i, 1.f), eta_ti * eta_ti, 1.f);
    float ci = fabsf(cos_i), ct = safe_sqrt(cos_t_sqr);
    bool index_matched = eta == 1.f, special = index_matched || ci == 0.f;
    float r_sc = index_matched ? 0.f : 1.f;
    float a_s = fmaf(-eta_it, ct, ci) / fmaf(eta_it, ct, ci);
    float a_p = fmaf(-eta_it, ci, ct) / fmaf(eta_it, ci, ct);
    r = 0.5f * (sqr(a_s) + sqr(a_p));
    if (special) r = r_sc;
    cos_t = mulsign_neg(ct, cos_i);
}
static inline float fresnel_conductor(float cos_i, float eta_r, float eta_i) {
    float c2 = cos_i * cos_i, s2 = 1.f - c2, s4 = s2 * s2;
    float temp_1 = eta_r * eta_r - eta_i * eta_i - s2;
    float a_2_pb_2 = safe_sqrt(temp_1 * temp_1 + 4.f * eta_i * eta_i * eta_r * eta_r);
    float a = safe_sqrt(.5f * (a_2_pb_2 + temp_1));
    float term_1 = a_2_pb_2 + c2, term_2 = 2.f * cos_i * a;
    float r_s = (term_1 - term_2) / (term_1 + term_2);
    float term_3 = a_2_pb_2 * c2 + s4, term_4 = term_2 * s2;
    float r_p = r_s * (term_3 - term_4) / (term_3 + term_4);
    return 0.5f * (r_s + r_p);
}

// ---------------------------------------------------------------------------------------------
// MicrofacetDistribution (include/mitsuba/render/microfacet.h:185-421)
// ---------------------------------------------------------------------------------------------
struct Microfacet {
    int type; float au, av; bool visible;
    Microfacet(int t, float a_u, float a_v, bool vis) : type(t), au(fmaxf(a_u, 1e-4f)), av(fmaxf(a_v, 1e-4f)), visible(vis) {}
    void scale_alpha(float v) { au *= v; av *= v; }      // microfacet.h:174-177
    float eval(V3 m) const {
        float alpha_uv = au * av, ct = m.z, ct2 = sqr(ct), result;
        if (type == B2_MF_BECKMANN)
            result = expf(-(sqr(m.x / au) + sqr(m.y / av)) / ct2) / (Pi * alpha_uv * sqr(ct2));
        else
            result = rcp(Pi * alpha_uv * sqr(sqr(m.x / au) + sqr(m.y / av) + sqr(m.z)));
        return result * ct > 1e-20f ? result : 0.f;
    }
    float smith_g1(V3 v, V3 m) const {
        float xy_alpha_2 = sqr(au * v.x) + sqr(av * v.y), tan_theta_alpha_2 = xy_alpha_2 / sqr(v.z), result;
        if (type == B2_MF_BECKMANN) {
            float a = rsqrt(tan_theta_alpha_2), a_sqr = sqr(a);
            result = a >= 1.6f ? 1.f : (3.535f * a + 2.181f * a_sqr) / (1.f + 2.276f * a + 2.577f * a_sqr);
        } else result = 2.f / (1.f + sqrtf(1.f + tan_theta_alpha_2));
        if (xy_alpha_2 == 0.f) result = 1.f;
        if (dot(v, m) * v.z <= 0.f) result = 0.f;
        return result;
    }
    float G(V3 wi, V3 wo, V3 m) const { return smith_g1(wi, m) * smith_g1(wo, m); }
    float pdf(V3 wi, V3 m) const {
        float result = eval(m);
        if (visible) result *= smith_g1(wi, m) * fabsf(dot(wi, m)) / wi.z; else result *= m.z;
        return result;
    }
    V2 sample_visible_11(float cos_theta_i, V2 sample) const {
        if (type == B2_MF_BECKMANN) {
            float tan_theta_i = safe_sqrt(fmaf(-cos_theta_i, cos_theta_i, 1.f)) / cos_theta_i, cot_theta_i = rcp(tan_theta_i);
            float maxval = erff(cot_theta_i);
            sample.x = fmaxf(fminf(sample.x, 1.f - 1e-6f), 1e-6f); sample.y = fmaxf(fminf(sample.y, 1.f - 1e-6f), 1e-6f);
            float x = maxval - (maxval + 1.f) * erff(sqrtf(-logf(sample.x)));
            sample.x *= 1.f + maxval + InvSqrtPi * tan_theta_i * expf(-sqr(cot_theta_i));
            for (int i = 0; i < 3; ++i) {
                float slope = erfinv(x), value = 1.f + x + InvSqrtPi * tan_theta_i * expf(-sqr(slope)) - sample.x,
                      derivative = 1.f - slope * tan_theta_i;
                x -= value / derivative;
            }
            return V2{ erfinv(x), erfinv(fmaf(2.f, sample.y, -1.f)) };
        } else {
            V2 p = square_to_uniform_disk_concentric(sample);
            float s = 0.5f * (1.f + cos_theta_i);
            p.y = lerp(safe_sqrt(1.f - sqr(p.x)), p.y, s);
            float x = p.x, y = p.y, z = safe_sqrt(1.f - fmaf(p.y, p.y, p.x * p.x));
            float sin_theta_i = safe_sqrt(1.f - sqr(cos_theta_i));
            float nrm = rcp(fmaf(sin_theta_i, y, cos_theta_i * z));
            return V2{ fmaf(cos_theta_i, y, -(sin_theta_i * z)) * nrm, x * nrm };
        }
    }
    static void sincos_phi(V3 v, float &sp, float &cp) {   // frame.h:111-122
        float st2 = fmaf(v.x, v.x, sqr(v.y)), inv = rsqrt(st2);
        if (fabsf(st2) <= 4.f * Eps) { cp = 1.f; sp = 0.f; }
        else { cp = clipf(v.x * inv, -1.f, 1.f); sp = clipf(v.y * inv, -1.f, 1.f); }
    }
    V3 sample(V3 wi, V2 sample, float &pdf_out) const {
        if (!visible) {
            float sin_phi, cos_phi, cos_theta, cos_theta_2, alpha_2, pdfv;
            if (au == av) { float a = TwoPi * sample.y; sin_phi = sinf(a); cos_phi = cosf(a); alpha_2 = au * au; }
            else {
                float ratio = av / au, tmp = ratio * tanf(TwoPi * sample.y);
                cos_phi = rsqrt(fmaf(tmp, tmp, 1.f)); cos_phi = mulsign(cos_phi, fabsf(sample.y - .5f) - .25f);
                sin_phi = cos_phi * tmp; alpha_2 = rcp(sqr(cos_phi / au) + sqr(sin_phi / av));
            }
            if (type == B2_MF_BECKMANN) {
                cos_theta = rsqrt(fmaf(-alpha_2, logf(1.f - sample.x), 1.f)); cos_theta_2 = sqr(cos_theta);
                float c3 = fmaxf(cos_theta_2 * cos_theta, 1e-20f);
                pdfv = (1.f - sample.x) / (Pi * au * av * c3);
            } else {
                float tan2 = alpha_2 * sample.x / (1.f - sample.x);
                cos_theta = rsqrt(1.f + tan2); cos_theta_2 = sqr(cos_theta);
                float temp = 1.f + tan2 / alpha_2, c3 = fmaxf(cos_theta_2 * cos_theta, 1e-20f);
                pdfv = rcp(Pi * au * av * c3 * sqr(temp));
            }
            float sin_theta = sqrtf(1.f - cos_theta_2);
            pdf_out = pdfv; return V3(cos_phi * sin_theta, sin_phi * sin_theta, cos_theta);
        }
        V3 wi_p = normalize(V3(au * wi.x, av * wi.y, wi.z));
        float sin_phi, cos_phi; sincos_phi(wi_p, sin_phi, cos_phi);
        float cos_theta = wi_p.z;
        V2 slope = sample_visible_11(cos_theta, sample);
        slope = V2{ fmaf(cos_phi, slope.x, -(sin_phi * slope.y)) * au, fmaf(sin_phi, slope.x, cos_phi * slope.y) * av };
        V3 m = normalize(V3(-slope.x, -slope.y, 1.f));
        pdf_out = eval(m) * smith_g1(wi, m) * fabsf(dot(wi, m)) / wi.z;
        return m;
    }
};

// ---------------------------------------------------------------------------------------------
// Rough-plastic tables (roughplastic.cpp:232-262): external transmittance at 64 cosines and the internal diffuse
// reflectance, by Gauss-Legendre quadrature over the visible-normal sampling domain (microfacet.h:463-570,
// quad.h:27-86, math.h:93-120).
// ---------------------------------------------------------------------------------------------
static inline void legendre_pd(int l, double x, double &lc, double &dc) {
    lc = 0; dc = 0;
    if (l > 1) {
        double l_p_pred = 1, l_pred = x, d_p_pred = 0, d_pred = 1, k0 = 3, k1 = 2, k2 = 1;
        for (int ki = 2; ki <= l; ++ki) {
            lc = (k0 * x * l_pred - k2 * l_p_pred) / k1; dc = d_p_pred + k0 * l_pred;
            l_p_pred = l_pred; l_pred = lc; d_p_pred = d_pred; d_pred = dc; k2 = k1; k0 += 2; k1 += 1;
        }
    } else if (l == 0) { lc = 1; dc = 0; } else { lc = x; dc = 1; }
}
static inline void gauss_legendre(int n, std::vector<float> &nodes, std::vector<float> &weights) {
    nodes.assign(n, 0.f); weights.assign(n, 0.f);
    n--;
    if (n == 0) { nodes[0] = 0.f; weights[0] = 2.f; return; }
    int m = (n + 1) / 2;
    for (int i = 0; i < m; ++i) {
        double x = -cos((double) (2 * i + 1) / (double) (2 * n + 2) * 3.14159265358979323846), lc, dc;
        for (int it = 0; it < 20; ++it) {
            legendre_pd(n + 1, x, lc, dc);
            double step = lc / dc; x -= step;
            if (fabs(step) <= 4 * fabs(x) * 1.1102230246251565e-16) break;
        }
        legendre_pd(n + 1, x, lc, dc);
        weights[i] = weights[n - i] = (float) (2 / ((1 - x * x) * (dc * dc)));
        nodes[i] = (float) x; nodes[n - i] = (float) -x;
    }
    if ((n % 2) == 0) { double lc, dc; legendre_pd(n + 1, 0.0, lc, dc); weights[n / 2] = (float) (2.0 / (dc * dc)); nodes[n / 2] = 0.f; }
}
// eval_reflectance / eval_transmittance (microfacet.h:463-570) for one incident direction
static inline float rough_quadrature(int type, float alpha, V3 wi, float eta, bool transmit) {
    const int res = eta > 1.f ? 32 : 128;
    std::vector<float> nodes, weights; gauss_legendre(res, nodes, weights);
    Microfacet distr(type, alpha, alpha, true);
    float result = 0.f;
    for (int jy = 0; jy < res; ++jy) for (int jx = 0; jx < res; ++jx) {          // dr::meshgrid(nodes, nodes): x runs fastest
        V2 node{ fmaf(nodes[jx], .5f, .5f), fmaf(nodes[jy], .5f, .5f) };
        float pdf; V3 m = distr.sample(wi, node, pdf);
        float f, cos_t, eta_it, eta_ti; fresnel(dot(wi, m), eta, f, cos_t, eta_it, eta_ti);
        float smith;
        if (transmit) {
            V3 wo = fmadd(m, fmaf(dot(wi, m), eta_ti, cos_t), -(wi * eta_ti));   // refract(wi, m, cos_t, eta_ti)
            smith = distr.smith_g1(wo, m) * (1.f - f);
            if (wo.z * wi.z >= 0.f) smith = 0.f;
        } else {
            V3 wo = fmadd(m, 2.f * dot(wi, m), -wi);                              // reflect(wi, m)
            smith = distr.smith_g1(wo, m) * f;
            if (wo.z <= 0.f || wi.z <= 0.f) smith = 0.f;
        }
        result += smith * (weights[jx] * weights[jy]) * 0.25f;
    }
    return result;
}
constexpr int kRoughRes = 64;   // MI_ROUGH_TRANSMITTANCE_RES (roughplastic.cpp:21)
static inline void rough_tables(int type, float alpha, float eta, float *trans, float &internal_reflectance) {
    float acc = 0.f;
    for (int i = 0; i < kRoughRes; ++i) {
        float mu = fmaxf(1e-6f, (float) i * (1.f / (float) (kRoughRes - 1)));
        V3 wi(sqrtf(1.f - mu * mu), 0.f, mu);
        trans[i] = rough_quadrature(type, alpha, wi, eta, true);
        acc += rough_quadrature(type, alpha, wi, 1.f / eta, false) * wi.z;
    }
    internal_reflectance = acc / (float) kRoughRes * 2.f;
}

// ---------------------------------------------------------------------------------------------
// BSDFs.  `unit` = evaluate with the differentiable reflectance set to 1 (linear coefficient, PRB).
// ---------------------------------------------------------------------------------------------
static inline uint32_t bsdf_lobes(const b2_bsdf &b) {
    uint32_t l = b.type == B2_BSDF_DIFFUSE ? LOBE_DIFFUSE : ((b.type == B2_BSDF_ROUGHCONDUCTOR || b.type == B2_BSDF_ROUGHDIELECTRIC) ? LOBE_GLOSSY
               : (b.type == B2_BSDF_ROUGHPLASTIC ? (LOBE_GLOSSY | LOBE_DIFFUSE) : (LOBE_DELTA_R | LOBE_DELTA_T)));
    return (b.flags & B2_BSDF_FLAG_MASK) ? (l | LOBE_NULL) : l;     // mask.cpp:96-110: nested components + a null lobe
}

// RoughPlastic::lerp_gather (roughplastic.cpp:343-351)
static inline float lerp_gather(const float *data, float x, int size) {
    x *= (float) (size - 1);
    uint32_t index = std::min((uint32_t) x, (uint32_t) (size - 2));
    float v0 = data[index], v1 = data[index + 1];
    return lerp(v0, v1, x - (float) index);
}
// RoughPlastic::eval / pdf (roughplastic.cpp:301-341, 353-393); wi.z > 0 and wo.z > 0 are the caller's precondition
static inline void roughplastic_eval_pdf(const Scene &sc, const b2_bsdf &b, const SI &si, V3 wo, V3 &val, float &pdf, bool unit, Counters &c) {
    const RoughTab &tab = sc.rough_of(b);
    const float ci = si.wi.z, co = wo.z;
    Microfacet distr(b.distribution, b.alpha_u, b.alpha_u, b.sample_visible != 0);
    const V3 H = normalize(wo + si.wi);
    const float D = distr.eval(H);
    float F, cos_t, eta_it, eta_ti; fresnel(dot(si.wi, H), b.ior, F, cos_t, eta_it, eta_ti);
    const float G = distr.G(si.wi, wo, H);
    V3 value = V3(F * D * G / (4.f * ci)) * V3(b.transmittance[0], b.transmittance[1], b.transmittance[2]);   // specular_reflectance
    const float t_i = lerp_gather(tab.trans, ci, kRoughRes), t_o = lerp_gather(tab.trans, co, kRoughRes);
    // unit = derivative mode (PRB adjoint): d/d rho of rho / (1 - rho r) = 1 / (1 - rho r)^2 (nonlinear), of rho / (1 - r) = 1 / (1 - r);
    // the specular lobe does not depend on the diffuse reflectance
    V3 diff = sc.reflectance(b, si, c);
    if (unit) {
        V3 den = V3(1.f) - (b.nonlinear ? diff * tab.internal_reflectance : V3(tab.internal_reflectance));
        diff = b.nonlinear ? V3(1.f) / (den * den) : V3(1.f) / den;
        value = V3(0.f);
    } else diff = diff / (V3(1.f) - (b.nonlinear ? diff * tab.internal_reflectance : V3(tab.internal_reflectance)));
    const float inv_eta_2 = 1.f / (b.ior * b.ior);
    val = value + diff * (InvPi * inv_eta_2 * co * t_i * t_o);
    float prob_specular = (1.f - t_i) * b.spec_sampling_weight, prob_diffuse = t_i * (1.f - b.spec_sampling_weight);
    prob_specular = prob_specular / (prob_specular + prob_diffuse); prob_diffuse = 1.f - prob_specular;
    float r = b.sample_visible ? D * distr.smith_g1(si.wi, H) / (4.f * ci) : distr.pdf(si.wi, H) / (4.f * dot(wo, H));
    r *= prob_specular;
    r += prob_diffuse * square_to_cosine_hemisphere_pdf(wo);
    pdf = r;
}

// SmoothDiffuse::eval_pdf (bsdfs/diffuse.cpp:159-179), RoughConductor::eval_pdf (roughconductor.cpp:429-518),
// SmoothDielectric::eval/pdf = 0 (dielectric.cpp:372-380), TwoSidedBRDF::eval_pdf (twosided.cpp:218-260)
static inline void bsdf_eval_pdf_nested(const Scene &sc, const b2_bsdf &b, const SI &si_, V3 wo, V3 &val, float &pdf, bool unit, Counters &c) {
    SI si = si_;
    val = V3(0.f); pdf = 0.f;
    if (b.flags & B2_BSDF_FLAG_TWOSIDED) { wo.z = mulsign(wo.z, si.wi.z); si.wi.z = fabsf(si.wi.z); }
    float ci = si.wi.z, co = wo.z;
    if (b.type == B2_BSDF_DIFFUSE) {
        if (!(ci > 0.f && co > 0.f)) return;
        V3 refl = unit ? V3(1.f) : sc.reflectance(b, si, c);
        val = refl * InvPi * co;
        pdf = square_to_cosine_hemisphere_pdf(wo);
    } else if (b.type == B2_BSDF_ROUGHCONDUCTOR) {
        V3 H = normalize(wo + si.wi);
        if (!(ci > 0.f && co > 0.f && dot(si.wi, H) > 0.f && dot(wo, H) > 0.f)) return;
        Microfacet distr(b.distribution, b.alpha_u, b.alpha_v, b.sample_visible != 0);
        float D = distr.eval(H);
        bool active = D != 0.f;
        float g1_wi = distr.smith_g1(si.wi, H), G = g1_wi * distr.smith_g1(wo, H);
        float value = D * G / (4.f * ci);
        V3 F(fresnel_conductor(dot(si.wi, H), b.eta[0], b.k[0]), fresnel_conductor(dot(si.wi, H), b.eta[1], b.k[1]),
             fresnel_conductor(dot(si.wi, H), b.eta[2], b.k[2]));
        V3 v(value);
        V3 refl = unit ? V3(1.f) : sc.reflectance(b, si, c);
        v = v * refl;
        float p = b.sample_visible ? D * g1_wi / (4.f * ci) : distr.pdf(si.wi, H) / (4.f * dot(wo, H));
        if (active) { val = F * v; pdf = p; }
    } else if (b.type == B2_BSDF_ROUGHPLASTIC) {
        if (!(ci > 0.f && co > 0.f)) return;
        roughplastic_eval_pdf(sc, b, si, wo, val, pdf, unit, c);
    } else if (b.type == B2_BSDF_ROUGHDIELECTRIC) {
        // RoughDielectric::eval_pdf (roughdielectric.cpp:503-605)
        if (ci == 0.f) return;
        const bool reflect = ci * co > 0.f;
        const float eta = ci > 0.f ? b.ior : rcp(b.ior), inv_eta = ci > 0.f ? rcp(b.ior) : b.ior;
        V3 m = normalize(si.wi + wo * (reflect ? 1.f : eta));
        m = V3(mulsign(m.x, m.z), mulsign(m.y, m.z), mulsign(m.z, m.z));
        const float dwi = dot(si.wi, m), dwo = dot(wo, m);
        if (!(dwi * ci > 0.f && dwo * co > 0.f)) return;
        Microfacet distr(b.distribution, b.alpha_u, b.alpha_v, b.sample_visible != 0);
        const float D = distr.eval(m);
        float F, cos_t, eta_it, eta_ti; fresnel(dwi, b.ior, F, cos_t, eta_it, eta_ti);
        const float G = distr.G(si.wi, wo, m);
        if (reflect) {
            val = V3(F * D * G / (4.f * fabsf(ci)));
            val = val * (unit ? V3(1.f) : sc.reflectance(b, si, c));
        } else {
            const float scale = sqr(inv_eta);
            val = V3(fabsf((scale * (1.f - F) * D * G * eta * eta * dwi * dwo) / (ci * sqr(dwi + eta * dwo))));
            val = unit ? V3(0.f) : val * V3(b.transmittance[0], b.transmittance[1], b.transmittance[2]);   // (no dependence on specular_reflectance)
        }
        if (!b.sample_visible) distr.scale_alpha(1.2f - .2f * sqrtf(fabsf(ci)));
        float p = distr.pdf(ci < 0.f ? -si.wi : si.wi, m);
        p *= reflect ? F : 1.f - F;
        const float dwh_dwo = reflect ? rcp(4.f * dwo) : (eta * eta * dwo) / sqr(dwi + eta * dwo);
        pdf = p * fabsf(dwh_dwo);
    }
}

// OpacityMask::eval_pdf (mask.cpp:196-215): nested value and pdf scaled by the opacity
static inline void bsdf_eval_pdf(const Scene &sc, const b2_bsdf &b, const SI &si, V3 wo, V3 &val, float &pdf, bool unit, Counters &c) {
    bsdf_eval_pdf_nested(sc, b, si, wo, val, pdf, unit, c);
    if (b.flags & B2_BSDF_FLAG_MASK) { val *= V3(b.opacity); pdf *= b.opacity; }
}

// SmoothDiffuse::sample (diffuse.cpp:100-124), RoughConductor::sample (roughconductor.cpp:227-313),
// SmoothDielectric::sample (dielectric.cpp:245-370), SmoothConductor::sample (conductor.cpp:249-307),
// TwoSidedBRDF::sample (twosided.cpp:110-147)
static inline void bsdf_sample(const Scene &sc, const b2_bsdf &b, const SI &si_, float s1, V2 s2, BS &bs, V3 &weight, Counters &c) {
    SI si = si_;
    bs = BS(); weight = V3(0.f);
    const bool masked = (b.flags & B2_BSDF_FLAG_MASK) != 0;
    if (masked) {           // OpacityMask::sample (mask.cpp:123-171)
        if (!(s1 < b.opacity)) { bs.wo = -si.wi; bs.eta = 1.f; bs.type = LOBE_NULL; bs.pdf = 1.f - b.opacity; weight = V3(1.f); return; }
        s1 /= b.opacity;
    }
    bool two = (b.flags & B2_BSDF_FLAG_TWOSIDED) != 0;
    if (two) si.wi.z = fabsf(si.wi.z);
    float ci = si.wi.z;
    if (b.type == B2_BSDF_DIFFUSE) {
        if (ci > 0.f) {
            bs.wo = square_to_cosine_hemisphere(s2);
            bs.pdf = square_to_cosine_hemisphere_pdf(bs.wo);
            bs.eta = 1.f; bs.type = LOBE_DIFFUSE;
            if (bs.pdf > 0.f) weight = sc.reflectance(b, si, c);
        }
    } else if (b.type == B2_BSDF_ROUGHCONDUCTOR) {
        if (ci > 0.f) {
            Microfacet distr(b.distribution, b.alpha_u, b.alpha_v, b.sample_visible != 0);
            V3 m = distr.sample(si.wi, s2, bs.pdf);
            bs.wo = fmadd(m, 2.f * dot(si.wi, m), -si.wi);           // reflect(wi, m) = fmsub(m, 2 dot, wi)
            bs.eta = 1.f; bs.type = LOBE_GLOSSY;
            bool active = (bs.pdf != 0.f) && bs.wo.z > 0.f;
            float w = b.sample_visible ? distr.smith_g1(bs.wo, m) : distr.G(si.wi, bs.wo, m) * dot(si.wi, m) / (ci * m.z);
            bs.pdf /= 4.f * dot(bs.wo, m);
            float cm = dot(si.wi, m);
            V3 F(fresnel_conductor(cm, b.eta[0], b.k[0]), fresnel_conductor(cm, b.eta[1], b.k[1]), fresnel_conductor(cm, b.eta[2], b.k[2]));
            V3 wv = V3(w) * sc.reflectance(b, si, c);
            if (active) weight = F * wv;
        }
    } else if (b.type == B2_BSDF_ROUGHPLASTIC) {
        // RoughPlastic::sample (roughplastic.cpp:264-299)
        if (ci > 0.f) {
            const RoughTab &tab = sc.rough_of(b);
            const float t_i = lerp_gather(tab.trans, ci, kRoughRes);
            float prob_specular = (1.f - t_i) * b.spec_sampling_weight, prob_diffuse = t_i * (1.f - b.spec_sampling_weight);
            prob_specular = prob_specular / (prob_specular + prob_diffuse);
            bs.eta = 1.f;
            if (s1 < prob_specular) {
                Microfacet distr(b.distribution, b.alpha_u, b.alpha_u, b.sample_visible != 0);
                float unused; V3 m = distr.sample(si.wi, s2, unused);
                bs.wo = fmadd(m, 2.f * dot(si.wi, m), -si.wi); bs.type = LOBE_GLOSSY;
            } else { bs.wo = square_to_cosine_hemisphere(s2); bs.type = LOBE_DIFFUSE; }
            V3 value(0.f); float p = 0.f;
            if (bs.wo.z > 0.f) roughplastic_eval_pdf(sc, b, si, bs.wo, value, p, false, c);
            bs.pdf = p;
            if (p > 0.f) weight = value / V3(p);
        }
    } else if (b.type == B2_BSDF_ROUGHDIELECTRIC) {
        // RoughDielectric::sample (roughdielectric.cpp:242-357)
        if (ci != 0.f) {
            Microfacet distr(b.distribution, b.alpha_u, b.alpha_v, b.sample_visible != 0), sample_distr = distr;
            if (!b.sample_visible) sample_distr.scale_alpha(1.2f - .2f * sqrtf(fabsf(ci)));
            V3 m = sample_distr.sample(ci < 0.f ? -si.wi : si.wi, s2, bs.pdf);
            const bool active = bs.pdf != 0.f;
            float F, cos_t, eta_it, eta_ti; fresnel(dot(si.wi, m), b.ior, F, cos_t, eta_it, eta_ti);
            const bool sel_r = s1 <= F && active, sel_t = !sel_r && active;
            bs.pdf *= sel_r ? F : 1.f - F;
            bs.eta = sel_r ? 1.f : eta_it; bs.type = LOBE_GLOSSY;
            V3 w(1.f); float dwh_dwo = 0.f;
            if (sel_r) {
                bs.wo = fmadd(m, 2.f * dot(si.wi, m), -si.wi);                       // reflect(wi, m)
                w = w * sc.reflectance(b, si, c);
                dwh_dwo = rcp(4.f * dot(bs.wo, m));
            }
            if (sel_t) {
                bs.wo = fmadd(m, fmaf(dot(si.wi, m), eta_ti, cos_t), -(si.wi * eta_ti));   // refract(wi, m, cos_t, eta_ti)
                w = w * (V3(b.transmittance[0], b.transmittance[1], b.transmittance[2]) * sqr(eta_ti));
                dwh_dwo = (sqr(bs.eta) * dot(bs.wo, m)) / sqr(dot(si.wi, m) + bs.eta * dot(bs.wo, m));
            }
            if (active) {
                w = w * (b.sample_visible ? distr.smith_g1(bs.wo, m) : distr.G(si.wi, bs.wo, m) * dot(si.wi, m) / (ci * m.z));
                weight = w;
            }
            bs.pdf *= fabsf(dwh_dwo);
        }
    } else if (b.type == B2_BSDF_CONDUCTOR) {
        // SmoothConductor::sample (bsdfs/conductor.cpp:249-307): delta reflection, weight = specular_reflectance * F(cos_i)
        if (ci > 0.f) {
            bs.wo = V3(-si.wi.x, -si.wi.y, si.wi.z);                  // reflect(wi)
            bs.pdf = 1.f; bs.eta = 1.f; bs.type = LOBE_DELTA_R;
            V3 F(fresnel_conductor(ci, b.eta[0], b.k[0]), fresnel_conductor(ci, b.eta[1], b.k[1]), fresnel_conductor(ci, b.eta[2], b.k[2]));
            weight = sc.reflectance(b, si, c) * F;
        }
    } else { // dielectric (never two-sided)
        float r_i, cos_t, eta_it, eta_ti; fresnel(ci, b.ior, r_i, cos_t, eta_it, eta_ti);
        float t_i = 1.f - r_i;
        bool sel_r = s1 <= r_i;
        bs.pdf = sel_r ? r_i : t_i;
        bs.type = sel_r ? LOBE_DELTA_R : LOBE_DELTA_T;
        bs.wo = sel_r ? V3(-si.wi.x, -si.wi.y, si.wi.z) : V3(-eta_ti * si.wi.x, -eta_ti * si.wi.y, cos_t);
        bs.eta = sel_r ? 1.f : eta_it;
        weight = V3(1.f);
        if (sel_r) weight *= V3(b.reflectance[0], b.reflectance[1], b.reflectance[2]);
        else { weight *= V3(b.transmittance[0], b.transmittance[1], b.transmittance[2]); weight *= sqr(eta_ti); }
    }
    if (two) bs.wo.z = mulsign(bs.wo.z, si_.wi.z);
    if (masked) bs.pdf *= b.opacity;
}

// ---------------------------------------------------------------------------------------------
// Emitters
// ---------------------------------------------------------------------------------------------
// EnvironmentMapEmitter::eval_spectrum (emitters/envmap.cpp:528-586)
static inline V3 env_eval_uv(const b2_envmap &e, V2 uv) {
    uint32_t rx = e.width, ry = e.height;
    uv.x -= .5f / (rx - 1u);
    uv.x -= floorf(uv.x); uv.y -= floorf(uv.y);
    uv.x *= (float) (rx - 1u); uv.y *= (float) (ry - 1u);
    uint32_t px = std::min((uint32_t) uv.x, rx - 2u), py = std::min((uint32_t) uv.y, ry - 2u);
    float w1x = uv.x - (float) px, w1y = uv.y - (float) py, w0x = 1.f - w1x, w0y = 1.f - w1y;
    uint32_t index = py * rx + px;
    auto px3 = [&](uint32_t i) { const float *p = e.data + 3 * (size_t) i; return V3(p[0], p[1], p[2]); };
    V3 v00 = px3(index), v10 = px3(index + 1), v01 = px3(index + rx), v11 = px3(index + rx + 1);
    V3 v0 = fmadd(V3(w0x), v00, v10 * w1x), v1 = fmadd(V3(w0x), v01, v11 * w1x);
    V3 v = fmadd(V3(w0y), v0, v1 * w1y);
    return v * e.scale;
}
// EnvironmentMapEmitter::eval (envmap.cpp:353-363): direction v = world direction of travel
static inline V3 env_eval_dir(const b2_envmap &e, V3 dir_world) {
    V3 v = xf_vector(e.from_world, dir_world);
    V2 uv{ atan2f(v.x, -v.z) * InvTwoPi, safe_acos(v.y) * InvPi };
    return env_eval_uv(e, uv);
}
// Hierarchical2D<Float,0>::Level::index (core/distr_2d.h:793-797)
static inline uint32_t warp_index(uint32_t x, uint32_t y, uint32_t width) {
    return ((x & 1u) | (((x & ~1u) | (y & 1u)) << 1)) + ((y & ~1u) * width);
}
// Hierarchical2D::sample (distr_2d.h:517-602)
static inline V2 warp_sample(const b2_envmap &e, V2 sample, float &pdf) {
    sample.x = clipf(sample.x, 0.f, 1.f); sample.y = clipf(sample.y, 0.f, 1.f);
    uint32_t ox = 0, oy = 0;
    for (int l = e.n_levels - 2; l > 0; --l) {
        const float *lv = e.warp_data + e.level_offset[l];
        ox <<= 1; oy <<= 1;
        uint32_t i = warp_index(ox, oy, e.level_width[l]);
        float v00 = lv[i], v10 = lv[i + 1], v01 = lv[i + 2], v11 = lv[i + 3];
        sample.x = clipf(sample.x, 0.f, 1.f); sample.y = clipf(sample.y, 0.f, 1.f);
        float r0 = v00 + v10, r1 = v01 + v11;
        sample.y *= r0 + r1;
        bool mask = sample.y > r0;
        if (mask) { oy += 1; sample.y -= r0; }
        sample.y /= mask ? r1 : r0;
        float c0 = mask ? v01 : v00, c1 = mask ? v11 : v10;
        sample.x *= c0 + c1;
        mask = sample.x > c0;
        if (mask) sample.x -= c0;
        sample.x /= mask ? c1 : c0;
        if (mask) ox += 1;
    }
    const float *l0 = e.warp_data + e.level_offset[0]; uint32_t w0 = e.level_width[0];
    uint32_t i = ox + oy * w0;
    float v00 = l0[i], v10 = l0[i + 1], v01 = l0[i + w0], v11 = l0[i + w0 + 1];
    sample = square_to_bilinear(v00, v10, v01, v11, sample, pdf);
    float psx = 1.f / (float) (e.width - 1), psy = 1.f / (float) (e.height - 1);
    return V2{ ((float) (int) ox + sample.x) * psx, ((float) (int) oy + sample.y) * psy };
}
// Alias-table mode (NOT in the reference, which only has Hierarchical2D; north_star asks for alias tables): restates the product's
// alias_sample (b2_shade.cuh) so that the mode has a per-sample checker too.  The patch is drawn from a Vose table over the patch
// masses, the position inside it by the same square_to_bilinear; the density is warp_eval's, which the chi^2 test of
// tests/test_oracle_golden.py checks (style of mitsuba3/src/core/tests/test_distr_2d.py:test04).
static inline V2 alias_sample(const b2_envmap &e, V2 sample, float &pdf) {
    const uint32_t pw = (uint32_t) e.width - 1u, ph = (uint32_t) e.height - 1u, n = pw * ph;
    sample.x = clipf(sample.x, 0.f, 1.f); sample.y = clipf(sample.y, 0.f, 1.f);
    const float scaled = sample.x * (float) n;
    uint32_t bin = std::min((uint32_t) scaled, n - 1u);
    float u = scaled - (float) bin;
    const float q = e.alias_prob[bin];
    if (u < q) u = u / q;
    else { u = (u - q) / (1.f - q); bin = e.alias_idx[bin]; }
    u = clipf(u, 0.f, 1.f);
    const uint32_t oy = bin / pw, ox = bin - oy * pw;
    const float *l0 = e.warp_data + e.level_offset[0]; const uint32_t w0 = e.level_width[0];
    const uint32_t i = ox + oy * w0;
    V2 s = square_to_bilinear(l0[i], l0[i + 1], l0[i + w0], l0[i + w0 + 1], V2{ u, sample.y }, pdf);
    const float psx = 1.f / (float) pw, psy = 1.f / (float) ph;
    return V2{ ((float) (int) ox + s.x) * psx, ((float) (int) oy + s.y) * psy };
}
// Hierarchical2D::eval (distr_2d.h:695-726)
static inline float warp_eval(const b2_envmap &e, V2 pos) {
    pos.x = clipf(pos.x, 0.f, 1.f); pos.y = clipf(pos.y, 0.f, 1.f);
    pos.x *= (float) (e.width - 1); pos.y *= (float) (e.height - 1);
    uint32_t ox = std::min((uint32_t) (int) pos.x, (uint32_t) e.width - 2u), oy = std::min((uint32_t) (int) pos.y, (uint32_t) e.height - 2u);
    pos.x -= (float) (int) ox; pos.y -= (float) (int) oy;
    const float *l0 = e.warp_data + e.level_offset[0]; uint32_t w0 = e.level_width[0];
    uint32_t i = ox + oy * w0;
    return square_to_bilinear_pdf(l0[i], l0[i + 1], l0[i + w0], l0[i + w0 + 1], pos);
}

struct EmitterSys {
    const Scene &sc;
    int env_warp = B2_ENV_WARP_HIERARCHICAL;          // B2_ENV_WARP_ALIAS: the product's fast mode (see alias_sample)
    EmitterSys(const Scene &s) : sc(s) {}
    const b2_scene_desc &d() const { return *sc.d; }

    // SurfaceInteraction::emitter (scene.h:866-878)
    int emitter_of(const SI &si) const { return si.valid ? d().meshes[si.mesh].emitter : sc.env_emitter; }

    // AreaLight::eval (emitters/area.cpp:83-90) | EnvironmentMapEmitter::eval | ConstantBackgroundEmitter::eval
    V3 eval(int em, const SI &si, Counters &c) const {
        if (em < 0) return V3(0.f);
        const b2_emitter &e = d().emitters[em];
        if (e.type == B2_EMITTER_AREA) return si.wi.z > 0.f ? V3(e.radiance[0], e.radiance[1], e.radiance[2]) : V3(0.f);
        c.escapes++;
        if (e.type == B2_EMITTER_CONSTANT) return V3(e.radiance[0], e.radiance[1], e.radiance[2]);
        return env_eval_dir(*d().envmap, -si.wi);
    }

    // Emitter::sample_direction (no visibility): area.cpp:117-168 + shape.cpp:354-371 + rectangle.cpp:182-200
    //   | mesh.cpp:758-812 ; envmap.cpp:414-458
    void sample_direction(int em, const SI &ref, V2 sample, DS &ds, V3 &spec, Counters &c) const {
        const b2_emitter &e = d().emitters[em];
        ds = DS(); ds.emitter = em; spec = V3(0.f);
        if (e.type == B2_EMITTER_AREA) {
            c.nee_area++;
            float pdf_pos;
            if (e.shape_kind == B2_SHAPE_RECTANGLE) {
                ds.p = xf_point_affine(e.to_world, V3(fmaf(sample.x, 2.f, -1.f), fmaf(sample.y, 2.f, -1.f), 0.f));
                ds.n = V3(e.normal[0], e.normal[1], e.normal[2]);
                ds.uv = sample; pdf_pos = e.inv_area;
            } else {
                // Mesh::sample_position: DiscreteDistribution::sample_reuse over the face-area pmf, then uniform triangle
                const b2_mesh &m = d().meshes[e.mesh];
                const float *cdf = d().area_cdf + e.cdf_offset; uint32_t nf = m.n_faces;
                float total = cdf[nf - 1], value = sample.y * total, nrmz = rcp(total);
                uint32_t lo = 0, hi = nf - 1;                 // first index with !(cdf[i] < value) (distr_1d.h:117-183)
                while (lo < hi) { uint32_t mid = (lo + hi) / 2; if (cdf[mid] < value) lo = mid + 1; else hi = mid; }
                uint32_t fi = lo; float prev = fi ? cdf[fi - 1] : 0.f;
                float sy = (sample.y - prev * nrmz) / ((cdf[fi] - prev) * nrmz);
                uint32_t f = m.first_face + fi;
                V3 p0 = sc.pos(d().faces[3 * f]), p1 = sc.pos(d().faces[3 * f + 1]), p2 = sc.pos(d().faces[3 * f + 2]);
                V3 e0 = p1 - p0, e1 = p2 - p0;
                V2 b = square_to_uniform_triangle(V2{ sample.x, sy });
                ds.p = fmadd(e0, b.x, fmadd(e1, b.y, p0));
                ds.n = normalize(cross(e0, e1));
                if (m.flags & B2_MESH_HAS_NORMALS) {
                    V3 n0 = sc.nrm(d().faces[3 * f]), n1 = sc.nrm(d().faces[3 * f + 1]), n2 = sc.nrm(d().faces[3 * f + 2]);
                    ds.n = normalize(fmadd(n0, 1.f - b.x - b.y, fmadd(n1, b.x, n2 * b.y)));
                }
                if (m.flags & B2_MESH_FLIP_NORMALS) ds.n = -ds.n;
                ds.uv = b; pdf_pos = e.inv_area;
            }
            ds.d = ds.p - ref.p;
            float dist2 = squared_norm(ds.d); ds.dist = sqrtf(dist2); ds.d = ds.d / ds.dist;
            float dp = fabsf(dot(ds.d, ds.n)), x = dist2 / dp;
            ds.pdf = pdf_pos * (finite(x) ? x : 0.f);
            bool active = dot(ds.d, ds.n) < 0.f && ds.pdf != 0.f;
            if (active) spec = V3(e.radiance[0], e.radiance[1], e.radiance[2]) / ds.pdf;
        } else if (e.type == B2_EMITTER_ENVMAP) {
            c.nee_env++;
            const b2_envmap &env = *d().envmap;
            float pdf; V2 uv = (env_warp == B2_ENV_WARP_ALIAS && env.alias_prob) ? alias_sample(env, sample, pdf) : warp_sample(env, sample, pdf);
            uv.x += .5f / (float) (env.width - 1);
            bool active = pdf > 0.f;
            float theta = uv.y * Pi, phi = uv.x * TwoPi;
            float st = sinf(theta), ct = cosf(theta), sp = sinf(phi), cp = cosf(phi);
            V3 dsph(cp * st, sp * st, ct);
            V3 dir(dsph.y, dsph.z, -dsph.x);
            V3 bc(env.bsphere_center[0], env.bsphere_center[1], env.bsphere_center[2]);
            float radius = fmaxf(env.bsphere_radius, norm(ref.p - bc)), dist = 2.f * radius;
            float inv_sin_theta = safe_rsqrt(fmaxf(sqr(dir.x) + sqr(dir.z), sqr(Eps)));
            dir = xf_vector(env.to_world, dir);
            ds.p = ref.p + dir * dist; ds.n = -dir; ds.uv = uv;
            ds.pdf = active ? pdf * inv_sin_theta * (1.f / (2.f * sqr(Pi))) : 0.f;
            ds.d = dir; ds.dist = dist;
            if (active) spec = env_eval_uv(env, uv) / ds.pdf;
        } else if (e.type == B2_EMITTER_POINT) {
            // PointLight::sample_direction (emitters/point.cpp:119-148): delta position, pdf 1, intensity / dist^2
            c.nee_area++;
            ds.p = V3(e.to_world[3], e.to_world[7], e.to_world[11]); ds.n = V3(0.f); ds.pdf = 1.f; ds.delta = true;
            ds.d = ds.p - ref.p;
            float dist2 = dot(ds.d, ds.d), inv_dist = 1.f / sqrtf(dist2);
            ds.dist = sqrtf(dist2); ds.d = ds.d * inv_dist;
            spec = V3(e.radiance[0], e.radiance[1], e.radiance[2]) * sqr(inv_dist);
        } else { // constant environment (emitters/constant.cpp): uniform sphere
            c.nee_env++;
            const b2_envmap *env = d().envmap;
            float z = fmaf(-2.f, sample.y, 1.f), r = safe_sqrt(fmaf(-z, z, 1.f));
            float a = TwoPi * sample.x; V3 dir(r * cosf(a), r * sinf(a), z);
            float radius = 1.f; V3 bc(0.f);
            if (env) { bc = V3(env->bsphere_center[0], env->bsphere_center[1], env->bsphere_center[2]); radius = env->bsphere_radius; }
            radius = fmaxf(radius, norm(ref.p - bc)); float dist = 2.f * radius;
            ds.p = ref.p + dir * dist; ds.n = -dir; ds.pdf = 0.25f * InvPi; ds.d = dir; ds.dist = dist;
            spec = V3(e.radiance[0], e.radiance[1], e.radiance[2]) / ds.pdf;
        }
    }

    // Scene::sample_emitter_direction (render/scene.cpp:319-369) incl. sample_emitter (:252-275) and the shadow test
    void sample_emitter_direction(const SI &ref, V2 sample, bool active, DS &ds, V3 &spec, Counters &c) const {
        ds = DS(); spec = V3(0.f);
        uint32_t n = d().n_emitters;
        if (!active || n == 0) return;
        uint32_t index = 0; float weight = 1.f;
        if (n >= 2) {
            float nf = (float) n, scaled = sample.x * nf;
            index = std::min((uint32_t) scaled, n - 1u);
            weight = nf; sample.x = scaled - (float) index;
        }
        sample_direction((int) index, ref, sample, ds, spec, c);
        ds.pdf *= 1.f / (float) n;                                  // pdf_emitter (scene.cpp:277-283)
        spec *= weight;
        if (ds.pdf != 0.f) {
            Ray sr = Scene::spawn_ray_to(ref.p, ref.n, ds.p);
            if (sc.occluded(sr, c)) { spec = V3(0.f); ds.pdf = 0.f; }
        }
    }

    // DirectionSample3f(scene, si, ref) (records.h:173-180)
    DS direction_sample(const SI &si, V3 ref_p) const {
        DS ds; ds.p = si.p; ds.n = si.sh.n; ds.uv = si.uv; ds.pdf = 0.f; ds.delta = false;
        V3 rel = si.p - ref_p; ds.dist = norm(rel);
        ds.d = si.valid ? rel / ds.dist : -si.wi;
        ds.emitter = emitter_of(si);
        return ds;
    }
    // Scene::pdf_emitter_direction (scene.cpp:381-392) -> area.cpp:170-197 + shape.cpp:373-384 | envmap.cpp:460-476
    float pdf_emitter_direction(const DS &ds, bool active) const {
        if (!active || ds.emitter < 0) return 0.f;
        const b2_emitter &e = d().emitters[ds.emitter];
        float pmf = 1.f / (float) d().n_emitters, value;
        if (e.type == B2_EMITTER_AREA) {
            float dp = dot(ds.d, ds.n);
            if (!(dp < 0.f)) return 0.f;
            float adp = fabsf(dp);
            value = e.inv_area * (adp != 0.f ? (ds.dist * ds.dist) / adp : 0.f);
        } else if (e.type == B2_EMITTER_ENVMAP) {
            const b2_envmap &env = *d().envmap;
            V3 dd = xf_vector(env.from_world, ds.d);
            V2 uv{ atan2f(dd.x, -dd.z) * InvTwoPi, safe_acos(dd.y) * InvPi };
            uv.x -= .5f / (float) (env.width - 1u);
            uv.x -= floorf(uv.x); uv.y -= floorf(uv.y);
            float inv_sin_theta = safe_rsqrt(fmaxf(sqr(dd.x) + sqr(dd.z), sqr(Eps)));
            value = warp_eval(env, uv) * inv_sin_theta * (1.f / (2.f * sqr(Pi)));
        } else value = 0.25f * InvPi;
        return value * pmf;
    }
};

} // namespace orc
