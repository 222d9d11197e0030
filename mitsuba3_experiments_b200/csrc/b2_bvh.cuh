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
// Triangle record = 3 x float4: (p0, face index) (e1 = p1 - p0, mesh index | shading class << 28) (e2 = p2 - p0, 0).
#pragma once
#include "b2_scene.cuh"

namespace b2 {

struct Hit { float t, u, v; uint32_t face, mesh, cls; B2_DEV bool valid() const { return face != 0xffffffffu; } };   // cls: shading class of the hit face (wavefront)


B2_DEV uint32_t byte_of(uint32_t lo, uint32_t hi, int s) { return ((s < 4 ? lo : hi) >> ((s & 3) * 8)) & 0xffu; }

// Quantised plane byte s of the pair (lo, hi) as the float 32768 + q, built by ONE byte permute (exponent byte 0x47,
// q in mantissa bits 8..15) instead of I2F.U8: the conversion runs on the XU pipe at 4 lanes/clk per sub-partition
// and 48 of them per node made the trace kernel XU-bound (profiles/README.md, r01_v4).  The bias is folded into the
// plane offset: t = fma(32768 + q, a, b - 32768 a); that offset is rounded at magnitude 32768 a, i.e. the plane moves
// by <= 2^-9 of a grid cell, which the builder's 1/64-cell quantisation margin covers (bvh_build.cu, k_collapse).
// `magic` (= 0x47) comes from the kernel's constant bank (DScene::qmagic) so that the selector is the immediate of the
// PRMT; with a literal ptxas keeps 0x47 as the immediate and spends IMADs on materialising the selectors.
B2_DEV float qbyte_biased(uint32_t lo, uint32_t hi, int s, uint32_t magic) {
    return __uint_as_float(__byte_perm(s < 4 ? lo : hi, magic, 0x4505u | ((uint32_t) (s & 3) << 4)));
}
constexpr float kQBias = 32768.f;
constexpr float kFarPad = 1.0000006f;      // robust-traversal factor on the exit distance (Ize 2013), folded into the far planes
constexpr float kTinyDir = 1e-20f;         // |d| below this is treated as this (keeps 32768 * scale / d finite)

// The eight slab tests of one wide node.  Returns the raw hit mask: bit 24 + slot for an internal child, the unary
// triangle bits (offset from tri_base) for a leaf child; permute_octant() turns slot bits into traversal-order bits.
B2_DEV uint32_t node_slab_hits(const uint4 &w0, const uint4 &w1, const uint4 &w2, const uint4 &w3, const uint4 &w4,
                               const f3 &o, float idx, float idy, float idz, uint32_t octinv, float tmax, uint32_t magic) {
    const float sx = __uint_as_float((w0.w & 0xffu) << 23), sy = __uint_as_float(((w0.w >> 8) & 0xffu) << 23),
                sz = __uint_as_float(((w0.w >> 16) & 0xffu) << 23);
    const float ax = sx * idx, ay = sy * idy, az = sz * idz;
    const float bx = __fmaf_rn(-kQBias, ax, (__uint_as_float(w0.x) - o.x) * idx), by = __fmaf_rn(-kQBias, ay, (__uint_as_float(w0.y) - o.y) * idy),
                bz = __fmaf_rn(-kQBias, az, (__uint_as_float(w0.z) - o.z) * idz);
    const float fax = ax * kFarPad, fay = ay * kFarPad, faz = az * kFarPad, fbx = bx * kFarPad, fby = by * kFarPad, fbz = bz * kFarPad;
    // entry / exit plane bytes of the eight children, chosen once per node by the ray octant
    const bool px = (octinv & 1u) != 0u, py = (octinv & 2u) != 0u, pz = (octinv & 4u) != 0u;
    const uint32_t nx0 = px ? w2.x : w3.z, nx1 = px ? w2.y : w3.w, fx0 = px ? w3.z : w2.x, fx1 = px ? w3.w : w2.y;
    const uint32_t ny0 = py ? w2.z : w4.x, ny1 = py ? w2.w : w4.y, fy0 = py ? w4.x : w2.z, fy1 = py ? w4.y : w2.w;
    const uint32_t nz0 = pz ? w3.x : w4.z, nz1 = pz ? w3.y : w4.w, fz0 = pz ? w4.z : w3.x, fz1 = pz ? w4.w : w3.y;
    // children are culled against tmax * kFarPad: a box whose entry distance ties with the current hit (coincident
    // triangles in another leaf) is still visited, so that equal-t hits resolve to the lowest face index
    const float tcull = tmax * kFarPad;
    uint32_t hitmask = 0;
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        const uint32_t mw = (s < 4 ? w1.z : w1.w) >> ((s & 3) * 8);      // meta byte of child s in bits 0..7
        const float t0x = __fmaf_rn(qbyte_biased(nx0, nx1, s, magic), ax, bx), t1x = __fmaf_rn(qbyte_biased(fx0, fx1, s, magic), fax, fbx);
        const float t0y = __fmaf_rn(qbyte_biased(ny0, ny1, s, magic), ay, by), t1y = __fmaf_rn(qbyte_biased(fy0, fy1, s, magic), fay, fby);
        const float t0z = __fmaf_rn(qbyte_biased(nz0, nz1, s, magic), az, bz), t1z = __fmaf_rn(qbyte_biased(fz0, fz1, s, magic), faz, fbz);
        const float tn = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, 0.f));
        const float tf = fminf(fminf(t1x, t1y), fminf(t1z, tcull));
        // internal child: meta = 0x20 | (24 + s) -> bit 24 + s; leaf: unary count << offset; empty: 0
        // (funnel shift: the shift amount is taken modulo 32, i.e. meta & 31)
        hitmask |= (tn <= tf) ? __funnelshift_l(0u, (mw >> 5) & 7u, mw) : 0u;
    }
    return hitmask;
}
// child-slot bits (24 + s) -> traversal-order bits (24 + (s ^ octinv)), returned in bits 0..7
B2_DEV uint32_t permute_octant(uint32_t hitmask, uint32_t octinv) {
    uint32_t nh = hitmask >> 24;
    if (octinv & 1u) nh = ((nh & 0x55u) << 1) | ((nh & 0xaau) >> 1);
    if (octinv & 2u) nh = ((nh & 0x33u) << 2) | ((nh & 0xccu) >> 2);
    if (octinv & 4u) nh = ((nh & 0x0fu) << 4) | ((nh & 0xf0u) >> 4);
    return nh;
}

// ANY: stop at the first hit in [0, tmax] (Scene::ray_test); otherwise closest hit (ray_intersect_preliminary).
// Ties at equal t resolve to the lower face index, independent of traversal order.
template <bool ANY, bool COUNT>
B2_DEV bool bvh_trace(const DScene &sc, f3 o, f3 d, float tmax, Hit &hit, Counters<COUNT> &cnt) {
    hit.t = tmax; hit.u = 0.f; hit.v = 0.f; hit.face = 0xffffffffu; hit.mesh = 0; hit.cls = 0;
    if (sc.n_nodes == 0) return false;
    const float ray_maxt = tmax;
    const float tiny = kTinyDir;
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
            const uint32_t imask = w0.w >> 24;
            const uint32_t hm = node_slab_hits(w0, w1, w2, w3, w4, o, idx, idy, idz, octinv, hit.t, sc.qmagic);
            const uint32_t nhits = permute_octant(hm, octinv) << 24, thits = hm & 0x00ffffffu;
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
                    found = true; hit.t = t; hit.u = u; hit.v = v; hit.face = face; hit.mesh = __float_as_uint(r1.w) & 0x0fffffffu; hit.cls = __float_as_uint(r1.w) >> 28;
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
