// b2_bvh.cuh -- compressed 8-wide BVH: node format + per-ray traversal (closest hit and any hit).
//
// Replaces Embree's rtcIntersect/rtcOccluded (mitsuba3/src/render/scene_embree.inl:245-433) and OptiX's
// built-in triangle traversal (scene_optix.inl:602-743).  The ray/triangle test is the reference's own
// Mesh::moeller_trumbore (mitsuba3/include/mitsuba/render/mesh.h:469-492) with the same operation order, so
// (t, u, v) are bit-identical to what the reference's fallback path computes; the quantised child boxes are
// conservative and only decide which triangles get tested.
//
// Node = 5 x 16 B (80 B), always fetched as five 128-bit loads:
//   w0: origin.xyz (f32 bits) | ex, ey, ez (biased exponents of the per-axis grid step), imask
//   w1: child_base, tri_base, meta[0..3], meta[4..7]
//   w2: qlo_x[8] | qlo_y[8]     w3: qlo_z[8] | qhi_x[8]     w4: qhi_y[8] | qhi_z[8]
// meta: 0 = empty; internal child = 0x20 | (24 + slot); leaf = (unary tri count) << 5 | offset from tri_base.
// Child slots were assigned at build time so that slot ^ ray-octant orders the children front to back.
// Triangle record = 3 x float4: (p0, face index) (e1 = p1 - p0, mesh index) (e2 = p2 - p0, 0).
#pragma once
#include "b2_scene.cuh"

namespace b2 {

struct Hit { float t, u, v; uint32_t face, mesh; B2_DEV bool valid() const { return face != 0xffffffffu; } };

constexpr int kStackSize = 48;

B2_DEV uint32_t byte_of(uint32_t lo, uint32_t hi, int s) { return ((s < 4 ? lo : hi) >> ((s & 3) * 8)) & 0xffu; }

// ANY: stop at the first hit in [0, tmax] (Scene::ray_test); otherwise closest hit (ray_intersect_preliminary).
// Ties at equal t resolve to the lower face index, independent of traversal order.
template <bool ANY, bool COUNT>
B2_DEV bool bvh_trace(const DScene &sc, f3 o, f3 d, float tmax, Hit &hit, Counters<COUNT> &cnt) {
    hit.t = tmax; hit.u = 0.f; hit.v = 0.f; hit.face = 0xffffffffu; hit.mesh = 0;
    if (sc.n_nodes == 0) return false;
    const float ray_maxt = tmax;
    const float tiny = 1e-30f;
    float idx = 1.f / (fabsf(d.x) > tiny ? d.x : copysignf(tiny, d.x));
    float idy = 1.f / (fabsf(d.y) > tiny ? d.y : copysignf(tiny, d.y));
    float idz = 1.f / (fabsf(d.z) > tiny ? d.z : copysignf(tiny, d.z));
    // octinv bit k set <=> the ray travels towards +k: children on the -k side (slot bit clear) are nearer
    const uint32_t octinv = (d.x >= 0.f ? 1u : 0u) | (d.y >= 0.f ? 2u : 0u) | (d.z >= 0.f ? 4u : 0u);

    uint2 stack[kStackSize];
    int sp = 0;
    uint2 ng = make_uint2(0u, (1u << (24 + octinv)) | 1u);   // node group: base | hit bits 24..31, imask 0..7
    uint2 tg = make_uint2(0u, 0u);                           // triangle group: base | hit bits 0..23
    bool found = false;

    while (true) {
        if (ng.y & 0xff000000u) {
            const int bit = 31 - __clz(ng.y);
            ng.y &= ~(1u << bit);
            if (ng.y & 0xff000000u) { if (sp < kStackSize) stack[sp++] = ng; }
            const uint32_t slot = (uint32_t) (bit - 24) ^ octinv;
            const uint32_t rel = __popc(ng.y & 0xffu & ~(0xffffffffu << slot));
            const uint4 *np = sc.nodes + (size_t) (ng.x + rel) * 5;
            const uint4 w0 = __ldg(np), w1 = __ldg(np + 1), w2 = __ldg(np + 2), w3 = __ldg(np + 3), w4 = __ldg(np + 4);
            cnt.add(C_NODES);
            const float sx = __uint_as_float((w0.w & 0xffu) << 23), sy = __uint_as_float(((w0.w >> 8) & 0xffu) << 23),
                        sz = __uint_as_float(((w0.w >> 16) & 0xffu) << 23);
            const uint32_t imask = w0.w >> 24;
            const float ax = sx * idx, ay = sy * idy, az = sz * idz;
            const float bx = (__uint_as_float(w0.x) - o.x) * idx, by = (__uint_as_float(w0.y) - o.y) * idy,
                        bz = (__uint_as_float(w0.z) - o.z) * idz;
            uint32_t nhits = 0, thits = 0;
#pragma unroll
            for (int s = 0; s < 8; ++s) {
                const uint32_t meta = byte_of(w1.z, w1.w, s);
                if (meta == 0) continue;
                const float lx = (float) byte_of(w2.x, w2.y, s), ly = (float) byte_of(w2.z, w2.w, s), lz = (float) byte_of(w3.x, w3.y, s);
                const float hx = (float) byte_of(w3.z, w3.w, s), hy = (float) byte_of(w4.x, w4.y, s), hz = (float) byte_of(w4.z, w4.w, s);
                const float t0x = __fmaf_rn(octinv & 1u ? lx : hx, ax, bx), t1x = __fmaf_rn(octinv & 1u ? hx : lx, ax, bx);
                const float t0y = __fmaf_rn(octinv & 2u ? ly : hy, ay, by), t1y = __fmaf_rn(octinv & 2u ? hy : ly, ay, by);
                const float t0z = __fmaf_rn(octinv & 4u ? lz : hz, az, bz), t1z = __fmaf_rn(octinv & 4u ? hz : lz, az, bz);
                const float tn = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, 0.f));
                const float tf = fminf(fminf(t1x, t1y), t1z) * 1.0000005f;
                if (tn <= fminf(tf, hit.t)) {
                    if ((imask >> s) & 1u) nhits |= 1u << (24 + ((uint32_t) s ^ octinv));
                    else thits |= (meta >> 5) << (meta & 31u);
                }
            }
            ng = make_uint2(w1.x, nhits | imask);
            tg = make_uint2(w1.y, thits);
        } else {
            tg = ng; ng = make_uint2(0u, 0u);
        }

        while (tg.y) {
            const int i = __ffs(tg.y) - 1;
            tg.y &= tg.y - 1;
            const float4 *tp = sc.tris + (size_t) (tg.x + i) * 3;
            const float4 r0 = __ldg(tp), r1 = __ldg(tp + 1), r2 = __ldg(tp + 2);
            cnt.add(C_TRIS);
            // Mesh::moeller_trumbore (mesh.h:469-492)
            const f3 e1(r1.x, r1.y, r1.z), e2(r2.x, r2.y, r2.z);
            const f3 pvec = cross(d, e2);
            const float inv_det = rcp(dot(e1, pvec));
            const f3 tvec = o - f3(r0.x, r0.y, r0.z);
            const float u = dot(tvec, pvec) * inv_det;
            bool ok = u >= 0.f && u <= 1.f;
            const f3 qvec = cross(tvec, e1);
            const float v = dot(d, qvec) * inv_det;
            ok = ok && v >= 0.f && u + v <= 1.f;
            const float t = dot(e2, qvec) * inv_det;
            ok = ok && t >= 0.f && t <= ray_maxt;
            if (ok) {
                if (ANY) { hit.t = t; hit.face = __float_as_uint(r0.w); return true; }
                const uint32_t face = __float_as_uint(r0.w);
                if (!found || t < hit.t || (t == hit.t && face < hit.face)) {
                    found = true; hit.t = t; hit.u = u; hit.v = v; hit.face = face; hit.mesh = __float_as_uint(r1.w);
                }
            }
        }

        if ((ng.y & 0xff000000u) == 0u) {
            if (sp == 0) break;
            ng = stack[--sp];
        }
    }
    return found;
}

}  // namespace b2
