// b2_scene.cuh -- device-resident scene layout (HBM) shared by all kernels.
//
// Geometry keeps the reference's Mesh buffers as they are (mesh.h:564-568): flat fp32 positions / normals /
// texcoords and u32 faces, because Mesh::compute_surface_interaction (mesh.cpp:1386-1560) re-interpolates
// from the *vertices*.  Traversal uses its own records:
//   * 80-byte compressed 8-wide BVH nodes (5 x 16-byte words, fetched as 128-bit loads)
//   * 48-byte triangle records (3 x float4: p0|face, e1|mesh, e2|-) in leaf order, 16-byte aligned
// Materials / emitters are 64-byte-aligned records.
#pragma once
#include "../../include/b2path.h"
#include "b2_math.cuh"

namespace b2 {

// Capacity of every traversal stack (b2_bvh.cuh, wf_trace.cu): one node group per level of the wide tree.  b2_scene_build_accel
// refuses trees deeper than this, so a push can never be dropped.
constexpr int kStackSize = 48;

struct alignas(16) DBsdf {          // 64 B
    int32_t type, flags, tex, distribution;
    float refl[3]; int32_t sample_visible;
    float eta[3]; float alpha_u;
    float k[3]; float alpha_v;
    float trans[3]; float ior;
    float opacity;                   // mask adapter (B2_BSDF_FLAG_MASK)
    int32_t nonlinear; float spec_weight, int_refl;     // rough plastic: nonlinear flag, specular sampling weight, internal reflectance
    const float *rough_table; uint64_t pad;             // rough plastic: 64 external transmittances (device); 112 B total
};

struct alignas(16) DTexture {
    const void *data;                // [h][w][3] f32 | f16 (device)
    int32_t width, height, format, wrap, filter, pad;
    float to_uv[6];                  // first two rows of the 3x3 affine
};

struct alignas(16) DMesh { int32_t bsdf, emitter; uint32_t flags, first_face, n_faces, cls /* shading class, b2_shade.cuh */, pad[2]; };

struct alignas(16) DEmitter {
    int32_t type, mesh, shape_kind; uint32_t cdf_offset;
    float radiance[3]; float inv_area;
    float normal[3]; float pad;
    float to_world[12];              // rows 0..2 of the rectangle's object->world matrix
};

struct DEnvmap {
    const float *data;               // [h][w][3]
    const float *warp;               // Hierarchical2D levels
    const float *alias_prob; const uint32_t *alias_idx;
    int32_t width, height, n_levels, present;
    float scale, bsphere_radius; float bsphere_center[3];
    float to_world[12], from_world[12];
    uint32_t level_offset[B2_MAX_WARP_LEVELS], level_width[B2_MAX_WARP_LEVELS];
};

struct DSensor {
    float sample_to_camera[16], to_world[16];
    float near_clip, far_clip;
    int32_t crop_offset_x, crop_offset_y, crop_width, crop_height;
    int32_t rfilter; float rfilter_radius, inv_radius;
    float gaussian_coeff[10];
    float scale_x, scale_y, offset_x, offset_y;  // integrator.cpp:432-436
    int32_t grid_w, grid_h, border;              // lane grid = crop window + 2 * border (sample_border, integrator.cpp:134-136)
};

struct DScene {
    const float *pos, *nrm, *uv; const uint32_t *faces;
    const DMesh *meshes; const DBsdf *bsdfs; const DTexture *textures; const DEmitter *emitters;
    const float *area_cdf;
    const uint4 *nodes;              // 5 x uint4 per wide node
    const float4 *tris;              // 3 x float4 per triangle, leaf order
    const float4 *shade;             // 6 x float4 per face (original face order): p0 p1 p2 n0 n1 n2 uv0 uv1 uv2, the inputs of
                                     // Mesh::compute_surface_interaction gathered once at scene creation (one 96-B record per hit
                                     // instead of a face -> 3 vertices x {pos, nrm, uv} pointer chase: 27 scalar loads -> 6 x 128-bit)
    uint32_t n_emitters, n_faces, n_nodes; int32_t env_emitter;
    uint32_t qmagic;                 // = 0x47: exponent byte of 32768 + q, read from the constant bank (b2_bvh.cuh, qbyte_biased)
    DEnvmap env;
    DSensor sensor;
};

// per-thread event counters (reduced per block; see b2_counters in b2path.h)
enum { C_SAMPLES, C_CLOSEST, C_SHADOW, C_VERTICES, C_TEX, C_NEE_AREA, C_NEE_ENV, C_ESCAPES, C_TAPS, C_GRAD, C_NODES, C_TRIS, C_N };

template <bool ON> struct Counters;
template <> struct Counters<true> {
    uint32_t v[C_N];
    B2_DEV Counters() { for (int i = 0; i < C_N; ++i) v[i] = 0; }
    B2_DEV void add(int i, uint32_t n = 1) { v[i] += n; }
};
template <> struct Counters<false> {
    B2_DEV void add(int, uint32_t = 1) {}
};

}  // namespace b2
