// capi.cu -- the extern "C" boundary declared in include/b2path.h (host side: handles, uploads, launches).
// No CPU fallback exists: every entry point either runs the CUDA path or fails with an error string.
#include "b2_build.h"
#include "b2_render.cuh"
#include "b2_wavefront.cuh"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

using namespace b2;

static thread_local std::string g_err;
static int fail(int code, const std::string &msg) { g_err = msg; return code; }
#define CUDA_OK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return fail(-10, std::string(#x) + ": " + cudaGetErrorString(e_)); } while (0)

struct b2_ctx { int device = 0, sm_count = 0; };

struct b2_scene {
    b2_ctx *ctx = nullptr;
    DScene d{};
    Bvh bvh;
    float *host_film = nullptr, *host_image = nullptr;   // device staging of b2_render_host
    cudaStream_t copy_stream = nullptr; cudaEvent_t film_ready = nullptr;   // b2_render_host: film upload overlapped with the bounces
    std::vector<void *> allocs;
    std::vector<b2_bsdf> bsdfs;                 // host mirror (for set_param / grad target validation)
    std::vector<DTexture> textures;             // host mirror of the device texture records
    std::vector<size_t> texture_bytes;
    DBsdf *bsdfs_dev = nullptr;
    uint32_t *face_mesh_dev = nullptr;
    uint32_t *unit_counter = nullptr;
    unsigned long long *counters_dev = nullptr;
    float *wfilm = nullptr;                     // weight film of the PRB sample set
    WfTiming timing; bool timing_on = false;     // kernel history (b2_kernel_times)
    uint32_t class_mask = 0;                    // shading classes present in the scene (bit = CLS_*)
    void *wf_mem = nullptr; uint32_t wf_capacity = 0; WfState wf{};   // wavefront lane state (grown on demand)
    void *wf_hist = nullptr; size_t wf_hist_bytes = 0;                // PRB: closest hits of every bounce of the primal pass (replayed by the adjoint)
    uint64_t launches = 0;

    template <typename T> cudaError_t upload(const T *host, size_t n, T *&dev) {
        void *q = nullptr; cudaError_t e = cudaMalloc(&q, std::max<size_t>(n * sizeof(T), 16)); if (e != cudaSuccess) return e;
        allocs.push_back(q); dev = (T *) q;
        if (n && host) e = cudaMemcpy(q, host, n * sizeof(T), cudaMemcpyHostToDevice);
        return e;
    }
    ~b2_scene() { for (cudaEvent_t e : timing.ev) cudaEventDestroy(e); if (film_ready) cudaEventDestroy(film_ready); if (copy_stream) cudaStreamDestroy(copy_stream); for (void *q : allocs) cudaFree(q); if (wf_mem) cudaFree(wf_mem); if (wf_hist) cudaFree(wf_hist); bvh_free(bvh); }
};

extern "C" {

const char *b2_last_error(void) { return g_err.c_str(); }
int b2_abi_version(void) { return B2_ABI_VERSION; }

int b2_init(int device, b2_ctx **out) {
    if (!out) return fail(-1, "b2_init: null output");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) return fail(-2, std::string("b2_init: no CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e));
    if (device < 0 || device >= n) return fail(-3, "b2_init: device index out of range");
    CUDA_OK(cudaSetDevice(device));
    cudaDeviceProp prop; CUDA_OK(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return fail(-4, "b2_init: this library is built for sm_100a (B200) only");
    b2_ctx *c = new b2_ctx; c->device = device; c->sm_count = prop.multiProcessorCount;
    *out = c; return 0;
}
void b2_shutdown(b2_ctx *ctx) { delete ctx; }

static void m3x4(const float *m16, float *m12) { std::memcpy(m12, m16, 12 * sizeof(float)); }

int b2_scene_create(b2_ctx *ctx, const b2_scene_desc *desc, b2_scene **out) {
    if (!ctx || !desc || !out) return fail(-1, "b2_scene_create: null argument");
    if (desc->abi_version != B2_ABI_VERSION) return fail(-2, "b2_scene_create: ABI version mismatch");
    if (desc->n_emitters && !desc->emitters) return fail(-2, "b2_scene_create: emitters missing");
    CUDA_OK(cudaSetDevice(ctx->device));
    std::unique_ptr<b2_scene> s(new b2_scene); s->ctx = ctx;
    DScene &d = s->d;
    float *pos, *nrm, *uv; uint32_t *faces;
    CUDA_OK(s->upload(desc->positions, (size_t) desc->n_vertices * 3, pos));
    CUDA_OK(s->upload(desc->normals, (size_t) desc->n_vertices * 3, nrm));
    CUDA_OK(s->upload(desc->texcoords, (size_t) desc->n_vertices * 2, uv));
    CUDA_OK(s->upload(desc->faces, (size_t) desc->n_faces * 3, faces));
    d.pos = pos; d.nrm = nrm; d.uv = uv; d.faces = faces; d.n_faces = desc->n_faces;
    for (size_t i = 0; i < (size_t) desc->n_faces * 3; ++i)
        if (desc->faces[i] >= desc->n_vertices) return fail(-2, "b2_scene_create: face index out of bounds");
    {   // per-face shading records (b2_scene.cuh, DScene::shade): the same fp32 values, regrouped
        std::vector<float> rec((size_t) desc->n_faces * 24, 0.f);
        for (size_t f = 0; f < desc->n_faces; ++f) {
            float *r = rec.data() + f * 24;
            for (int k = 0; k < 3; ++k) {
                const size_t v = desc->faces[3 * f + k];
                for (int c = 0; c < 3; ++c) r[3 * k + c] = desc->positions[3 * v + c];
                if (desc->normals) for (int c = 0; c < 3; ++c) r[9 + 3 * k + c] = desc->normals[3 * v + c];
                if (desc->texcoords) for (int c = 0; c < 2; ++c) r[18 + 2 * k + c] = desc->texcoords[2 * v + c];
            }
        }
        float *rec_dev; CUDA_OK(s->upload(rec.data(), rec.size(), rec_dev)); d.shade = (const float4 *) rec_dev;
    }
    // meshes + face -> mesh table
    if (desc->n_meshes >= (1u << 28)) return fail(-2, "b2_scene_create: more than 2^28 meshes");
    std::vector<DMesh> meshes(desc->n_meshes); std::vector<uint32_t> face_mesh(desc->n_faces, 0);
    for (uint32_t i = 0; i < desc->n_meshes; ++i) {
        const b2_mesh &m = desc->meshes[i];
        if (m.bsdf < 0 || (uint32_t) m.bsdf >= desc->n_bsdfs) return fail(-2, "b2_scene_create: mesh references a missing BSDF");
        if ((uint64_t) m.first_face + m.n_faces > desc->n_faces) return fail(-2, "b2_scene_create: mesh face range out of bounds");
        const b2_bsdf &mb = desc->bsdfs[m.bsdf];
        // CLS_* of b2_shade.cuh: 0 diffuse, 1 textured diffuse, 2 rough conductor, 3 delta (dielectric, conductor), 5 rough plastic, 6 rough dielectric
        const uint32_t cls = mb.type == B2_BSDF_DIFFUSE ? (mb.reflectance_tex >= 0 ? 1u : 0u) : mb.type == B2_BSDF_ROUGHCONDUCTOR ? 2u : mb.type == B2_BSDF_ROUGHPLASTIC ? 5u : mb.type == B2_BSDF_ROUGHDIELECTRIC ? 6u : 3u;
        s->class_mask |= 1u << cls;
        meshes[i] = DMesh{ m.bsdf, m.emitter, m.flags, m.first_face, m.n_faces, cls, { 0, 0 } };
        // the triangle records carry mesh | class << 28 (the trace kernel's epilogue needs both and nothing else of the mesh)
        for (uint32_t f = 0; f < m.n_faces; ++f) face_mesh[m.first_face + f] = i | (cls << 28);
    }
    DMesh *meshes_dev; CUDA_OK(s->upload(meshes.data(), meshes.size(), meshes_dev)); d.meshes = meshes_dev;
    CUDA_OK(s->upload(face_mesh.data(), face_mesh.size(), s->face_mesh_dev));
    // textures
    s->textures.resize(desc->n_textures); s->texture_bytes.resize(desc->n_textures);
    for (uint32_t i = 0; i < desc->n_textures; ++i) {
        const b2_texture &t = desc->textures[i];
        size_t bytes = (size_t) t.width * t.height * 3 * (t.format == B2_TEX_F16 ? 2 : 4);
        uint8_t *data; CUDA_OK(s->upload((const uint8_t *) t.data, bytes, data));
        DTexture &o = s->textures[i];
        o.data = data; o.width = t.width; o.height = t.height; o.format = t.format; o.wrap = t.wrap; o.filter = t.filter; o.pad = 0;
        for (int k = 0; k < 6; ++k) o.to_uv[k] = t.to_uv[k];
        s->texture_bytes[i] = bytes;
    }
    DTexture *tex_dev; CUDA_OK(s->upload(s->textures.data(), s->textures.size(), tex_dev)); d.textures = tex_dev;
    // BSDFs
    s->bsdfs.assign(desc->bsdfs, desc->bsdfs + desc->n_bsdfs);
    std::vector<DBsdf> bs(desc->n_bsdfs);
    for (uint32_t i = 0; i < desc->n_bsdfs; ++i) {
        const b2_bsdf &b = desc->bsdfs[i]; DBsdf &o = bs[i];
        if (b.reflectance_tex >= (int) desc->n_textures) return fail(-2, "b2_scene_create: BSDF references a missing texture");
        if (b.type < 0 || b.type > B2_BSDF_ROUGHPLASTIC) return fail(-2, "b2_scene_create: unsupported BSDF type");
        o.type = b.type; o.flags = b.flags; o.tex = b.reflectance_tex; o.distribution = b.distribution; o.sample_visible = b.sample_visible;
        for (int k = 0; k < 3; ++k) { o.refl[k] = b.reflectance[k]; o.eta[k] = b.eta[k]; o.k[k] = b.k[k]; o.trans[k] = b.transmittance[k]; }
        o.alpha_u = b.alpha_u; o.alpha_v = b.alpha_v; o.ior = b.ior;
        o.opacity = (b.flags & B2_BSDF_FLAG_MASK) ? b.opacity : 1.f;
        o.nonlinear = b.nonlinear; o.spec_weight = b.spec_sampling_weight; o.int_refl = 0.f; o.rough_table = nullptr; o.pad = 0;
        if (b.type == B2_BSDF_ROUGHPLASTIC) {
            // tables by (distribution, alpha, eta), shared between BSDFs with equal parameters
            int found = -1;
            for (uint32_t j = 0; j < i; ++j) {
                const b2_bsdf &q = desc->bsdfs[j];
                if (q.type == B2_BSDF_ROUGHPLASTIC && q.distribution == b.distribution && q.alpha_u == b.alpha_u && q.ior == b.ior) { found = (int) j; break; }
            }
            if (found >= 0) { o.rough_table = bs[found].rough_table; o.int_refl = bs[found].int_refl; }
            else {
                float trans[64]; std::string terr; float *tdev = nullptr;
                if (!(b.ior > 0.f) || b.ior == 1.f) return fail(-2, "b2_scene_create: roughplastic needs int_ior != ext_ior");
                if (!rough_tables_build(b.distribution, b.alpha_u, b.ior, trans, &o.int_refl, terr)) return fail(-10, terr);
                CUDA_OK(s->upload(trans, 64, tdev)); o.rough_table = tdev;
            }
        }
        if ((b.flags & B2_BSDF_FLAG_MASK) && !(b.opacity >= 0.f && b.opacity <= 1.f)) return fail(-2, "b2_scene_create: mask opacity outside [0, 1]");
    }
    CUDA_OK(s->upload(bs.data(), bs.size(), s->bsdfs_dev)); d.bsdfs = s->bsdfs_dev;
    // emitters
    std::vector<DEmitter> ems(desc->n_emitters); d.env_emitter = -1;
    for (uint32_t i = 0; i < desc->n_emitters; ++i) {
        const b2_emitter &e = desc->emitters[i]; DEmitter &o = ems[i];
        o.type = e.type; o.mesh = e.mesh; o.shape_kind = e.shape_kind; o.cdf_offset = e.cdf_offset; o.inv_area = e.inv_area; o.pad = 0.f;
        for (int k = 0; k < 3; ++k) { o.radiance[k] = e.radiance[k]; o.normal[k] = e.normal[k]; }
        m3x4(e.to_world, o.to_world);
        if (e.type == B2_EMITTER_ENVMAP || e.type == B2_EMITTER_CONSTANT) d.env_emitter = (int) i;
        else if (e.type == B2_EMITTER_POINT) { o.mesh = -1; }
        else if (e.type != B2_EMITTER_AREA) return fail(-2, "b2_scene_create: unknown emitter type");
        else if (e.mesh < 0 || (uint32_t) e.mesh >= desc->n_meshes) return fail(-2, "b2_scene_create: area emitter without a mesh");
    }
    DEmitter *em_dev; CUDA_OK(s->upload(ems.data(), ems.size(), em_dev)); d.emitters = em_dev; d.n_emitters = desc->n_emitters;
    float *cdf; CUDA_OK(s->upload(desc->area_cdf, desc->n_area_cdf, cdf)); d.area_cdf = cdf;
    // environment
    std::memset(&d.env, 0, sizeof(d.env));
    if (desc->envmap) {
        const b2_envmap &e = *desc->envmap; DEnvmap &o = d.env;
        o.present = 1; o.width = e.width; o.height = e.height; o.n_levels = e.n_levels; o.scale = e.scale; o.bsphere_radius = e.bsphere_radius;
        for (int k = 0; k < 3; ++k) o.bsphere_center[k] = e.bsphere_center[k];
        m3x4(e.to_world, o.to_world); m3x4(e.from_world, o.from_world);
        if (e.n_levels > 0) {
            if (e.n_levels > B2_MAX_WARP_LEVELS) return fail(-2, "b2_scene_create: too many warp levels");
            float *data; CUDA_OK(s->upload(e.data, (size_t) e.width * e.height * 3, data)); o.data = data;
            // repack the Hierarchical2D levels so that every level starts on a 16-byte boundary (2x2 blocks -> one float4 load)
            std::vector<float> packed; uint32_t off = 0;
            for (int l = 0; l < e.n_levels; ++l) {
                o.level_offset[l] = off; o.level_width[l] = e.level_width[l];
                packed.resize(off + ((e.level_size[l] + 3u) & ~3u), 0.f);
                std::memcpy(packed.data() + off, e.warp_data + e.level_offset[l], sizeof(float) * e.level_size[l]);
                off = (uint32_t) packed.size();
            }
            float *warp; CUDA_OK(s->upload(packed.data(), packed.size(), warp)); o.warp = warp;
            if (e.alias_prob && e.alias_idx) {
                size_t n = (size_t) (e.width - 1) * (e.height - 1);
                float *ap; uint32_t *ai; CUDA_OK(s->upload(e.alias_prob, n, ap)); CUDA_OK(s->upload(e.alias_idx, n, ai));
                o.alias_prob = ap; o.alias_idx = ai;
            }
        }
    }
    // sensor
    {
        const b2_sensor &e = desc->sensor; DSensor &o = d.sensor;
        if (e.crop_width <= 0 || e.crop_height <= 0) return fail(-2, "b2_scene_create: empty film");
        std::memcpy(o.sample_to_camera, e.sample_to_camera, sizeof(o.sample_to_camera)); std::memcpy(o.to_world, e.to_world, sizeof(o.to_world));
        o.near_clip = e.near_clip; o.far_clip = e.far_clip;
        o.crop_offset_x = e.crop_offset_x; o.crop_offset_y = e.crop_offset_y; o.crop_width = e.crop_width; o.crop_height = e.crop_height;
        o.rfilter = e.rfilter; o.rfilter_radius = e.rfilter_radius; o.inv_radius = 1.f / e.rfilter_radius;
        std::memcpy(o.gaussian_coeff, e.gaussian_coeff, sizeof(o.gaussian_coeff));
        o.scale_x = 1.f / (float) e.crop_width; o.scale_y = 1.f / (float) e.crop_height;
        o.offset_x = -(float) e.crop_offset_x * o.scale_x; o.offset_y = -(float) e.crop_offset_y * o.scale_y;
        int n = e.rfilter == B2_RFILTER_BOX ? 0 : (int) std::ceil(e.rfilter_radius - .5f);
        if (n > 2) return fail(-2, "b2_scene_create: reconstruction filter footprint larger than 5x5");
        // ReconstructionFilter::border_size (rfilter.cpp:25) = ceil(radius - 1/2 - 2 RayEpsilon)
        o.border = (e.sample_border && e.rfilter != B2_RFILTER_BOX) ? (int) std::ceil(e.rfilter_radius - .5f - 2.f * 1500.f * 5.9604644775390625e-8f) : 0;
        o.grid_w = e.crop_width + 2 * o.border; o.grid_h = e.crop_height + 2 * o.border;
    }
    CUDA_OK(cudaMalloc((void **) &s->unit_counter, 16)); s->allocs.push_back(s->unit_counter);
    CUDA_OK(cudaMalloc((void **) &s->counters_dev, sizeof(unsigned long long) * C_N)); s->allocs.push_back(s->counters_dev);
    d.nodes = nullptr; d.tris = nullptr; d.n_nodes = 0; d.qmagic = 0x47u;
    *out = s.release();
    return 0;
}

void b2_scene_destroy(b2_scene *scene) {
    if (!scene) return;
    cudaSetDevice(scene->ctx->device);
    delete scene;
}

int b2_scene_build_accel(b2_scene *s, int builder, uintptr_t stream, b2_build_stats *stats) {
    if (!s) return fail(-1, "b2_scene_build_accel: null scene");
    if (builder == B2_BUILDER_DEFAULT) { const char *e = getenv("B2_BVH_BUILDER"); builder = e ? atoi(e) : B2_BUILDER_AUTO; }
    if (builder != B2_BUILDER_LBVH && builder != B2_BUILDER_PLOC && builder != B2_BUILDER_AUTO) return fail(-2, "b2_scene_build_accel: unknown builder (0 = LBVH, 1 = PLOC, 2 = both + SAH pick, -1 = default)");
    CUDA_OK(cudaSetDevice(s->ctx->device));
    bvh_free(s->bvh);
    std::string err;
    // The traversal stacks hold kStackSize entries (one node group per level of the wide tree below the root, b2_bvh.cuh /
    // wf_trace.cu); a deeper tree would have to drop subtrees, so it is never handed to the kernels: AUTO falls back to
    // the other builder, an explicitly requested builder fails.
    // B2_MAX_BVH_DEPTH lowers the limit (test hook for this guard)
    const char *e_depth = getenv("B2_MAX_BVH_DEPTH");
    const uint32_t depth_limit = e_depth ? (uint32_t) std::min(std::max(atoi(e_depth), 0), kStackSize) : (uint32_t) kStackSize;
    auto too_deep = [depth_limit](const Bvh &t) { return t.depth > depth_limit; };
    auto deep_msg = [depth_limit](const Bvh &t) { return "b2_scene_build_accel: wide-tree depth " + std::to_string(t.depth) + " exceeds the traversal stack (" + std::to_string(depth_limit) + " levels)"; };
    if (builder == B2_BUILDER_AUTO) {
        // build both trees and keep the one the SAH model prefers (the build is a few ms; the render traces it for seconds)
        Bvh a, b;
        if (!bvh_build(s->d.pos, s->d.faces, s->face_mesh_dev, s->d.n_faces, B2_BUILDER_LBVH, (cudaStream_t) stream, a, err)) return fail(-11, err);
        if (!bvh_build(s->d.pos, s->d.faces, s->face_mesh_dev, s->d.n_faces, B2_BUILDER_PLOC, (cudaStream_t) stream, b, err)) { bvh_free(a); return fail(-11, err); }
        // PLOC's clusters are not octant-shaped, so the slot ^ octant child order is less strictly front to back than on
        // the radix tree: it has to win the SAH estimate by 10 % (bedroom stand-in: 0.78 -> PLOC, +13 % Msamples/s;
        // 10 M random triangles: 0.98 -> LBVH, where PLOC measured 11 % slower; profiles/README.md)
        bool keep_a = !(b.sah_cost < 0.9f * a.sah_cost);
        if (too_deep(a) && too_deep(b)) { const std::string m = deep_msg(a); bvh_free(a); bvh_free(b); return fail(-11, m); }
        if (too_deep(a)) keep_a = false; else if (too_deep(b)) keep_a = true;
        const float ms = a.ms_total + b.ms_total;
        if (keep_a) { bvh_free(b); s->bvh = a; } else { bvh_free(a); s->bvh = b; }
        s->bvh.ms_total = ms;
    } else {
        if (!bvh_build(s->d.pos, s->d.faces, s->face_mesh_dev, s->d.n_faces, builder, (cudaStream_t) stream, s->bvh, err)) return fail(-11, err);
        if (too_deep(s->bvh)) { const std::string m = deep_msg(s->bvh); bvh_free(s->bvh); return fail(-11, m); }
    }
    s->d.nodes = s->bvh.nodes; s->d.tris = s->bvh.tris; s->d.n_nodes = s->bvh.n_nodes;
    if (stats) {
        stats->ms_morton_sort = s->bvh.ms_morton_sort; stats->ms_lbvh = s->bvh.ms_lbvh; stats->ms_collapse = s->bvh.ms_collapse; stats->ms_total = s->bvh.ms_total;
        stats->n_nodes8 = s->bvh.n_nodes; stats->n_tris = s->bvh.n_tris; stats->max_depth = s->bvh.depth; stats->builder = (uint32_t) s->bvh.builder; stats->sah_cost = s->bvh.sah_cost; stats->reserved = 0.f;
        stats->node_bytes = s->bvh.node_bytes; stats->tri_bytes = s->bvh.tri_bytes;
    }
    return 0;
}

int b2_scene_set_param(b2_scene *s, const char *key, const void *dev_ptr, size_t bytes, uintptr_t stream) {
    if (!s || !key || !dev_ptr) return fail(-1, "b2_scene_set_param: null argument");
    CUDA_OK(cudaSetDevice(s->ctx->device));
    int idx = -1; char what[32] = { 0 };
    if (std::sscanf(key, "%d.spec_sampling_weigh%1s", &idx, what) == 2 && idx >= 0 && idx < (int) s->bsdfs.size()) {
        // RoughPlastic::parameters_changed (roughplastic.cpp:216-238) re-derives the lobe sampling weight from the texture means;
        // the caller passes the new value after it changed the diffuse reflectance
        if (bytes != sizeof(float) || s->bsdfs[idx].type != B2_BSDF_ROUGHPLASTIC) return fail(-2, "b2_scene_set_param: spec_sampling_weight is one float of a rough plastic BSDF");
        CUDA_OK(cudaMemcpyAsync(&s->bsdfs_dev[idx].spec_weight, dev_ptr, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t) stream));
        return 0;
    }
    if (std::sscanf(key, "%d.reflectance.%31s", &idx, what) != 2 || idx < 0 || idx >= (int) s->bsdfs.size())
        return fail(-2, std::string("b2_scene_set_param: unknown key '") + key + "'");
    if (!std::strcmp(what, "value")) {
        if (bytes != 3 * sizeof(float)) return fail(-2, "b2_scene_set_param: reflectance.value expects 3 floats");
        if (s->bsdfs[idx].reflectance_tex >= 0) return fail(-2, "b2_scene_set_param: BSDF has a textured reflectance; use .data");
        CUDA_OK(cudaMemcpyAsync(&s->bsdfs_dev[idx].refl[0], dev_ptr, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t) stream));
        return 0;
    }
    if (!std::strcmp(what, "data")) {
        int t = s->bsdfs[idx].reflectance_tex;
        if (t < 0) return fail(-2, "b2_scene_set_param: BSDF has a constant reflectance; use .value");
        if (s->textures[t].format != B2_TEX_F32) return fail(-2, "b2_scene_set_param: only fp32 textures are differentiable parameters (bitmap.cpp:208-218)");
        if (bytes != s->texture_bytes[t]) return fail(-2, "b2_scene_set_param: texture size mismatch");
        CUDA_OK(cudaMemcpyAsync((void *) s->textures[t].data, dev_ptr, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t) stream));
        return 0;
    }
    return fail(-2, std::string("b2_scene_set_param: unknown key '") + key + "'");
}

// wavefront: the lane state of one chunk lives in HBM (208 B per lane); default chunk = 64 Mi lanes (14 GB of the 180 GB)
static int wf_reserve(b2_scene *s, uint64_t shard_lanes, uint32_t chunk_lanes, cudaStream_t st) {
    const uint32_t want = (uint32_t) std::min<uint64_t>(shard_lanes, chunk_lanes ? std::max<uint32_t>(chunk_lanes, 256u) : (64u << 20));
    if (want > s->wf_capacity) {
        CUDA_OK(cudaStreamSynchronize(st));
        if (s->wf_mem) { cudaFree(s->wf_mem); s->wf_mem = nullptr; s->wf_capacity = 0; }
        CUDA_OK(cudaMalloc(&s->wf_mem, wf_state_bytes(want)));
        s->wf_capacity = want;
    }
    wf_bind(s->wf, s->wf_mem, want);
    s->wf.class_mask = s->class_mask;
    return 0;
}

// PRB: room for the closest hit of every bounce of the primal pass (32 B x lanes x max_depth), so that the adjoint replay does not
// trace the same closest rays again.  Optional: beyond B2_HIT_HISTORY_GB (default 32) the adjoint re-traces instead.
static void *wf_reserve_history(b2_scene *s, int max_depth, cudaStream_t st) {
    static const double limit_gb = getenv("B2_HIT_HISTORY_GB") ? atof(getenv("B2_HIT_HISTORY_GB")) : 32.0;
    const size_t need = (size_t) std::max(max_depth, 1) * s->wf.capacity * 32;
    if ((double) need > limit_gb * 1073741824.0) return nullptr;
    if (need > s->wf_hist_bytes) {
        if (cudaStreamSynchronize(st) != cudaSuccess) return nullptr;
        if (s->wf_hist) { cudaFree(s->wf_hist); s->wf_hist = nullptr; s->wf_hist_bytes = 0; }
        if (cudaMalloc(&s->wf_hist, need) != cudaSuccess) { cudaGetLastError(); s->wf_hist = nullptr; return nullptr; }
        s->wf_hist_bytes = need;
    }
    return s->wf_hist;
}

// integrator.cpp:139-152,247-265
static int derive_passes(const b2_scene *s, const b2_render_args *a, uint32_t &spp_per_pass, uint32_t &n_passes, uint64_t &wavefront) {
    if (a->spp == 0) return fail(-2, "render: spp must be > 0");
    const uint64_t npix = (uint64_t) s->d.sensor.grid_w * s->d.sensor.grid_h, limit = 0xffffffffull;
    spp_per_pass = a->spp_per_pass ? std::min(a->spp_per_pass, a->spp) : a->spp;
    wavefront = npix * spp_per_pass;
    if (wavefront > limit) { spp_per_pass /= (uint32_t) ((wavefront + limit - 1) / limit); wavefront = npix * spp_per_pass; }
    if (spp_per_pass == 0 || a->spp % spp_per_pass) return fail(-2, "sample_count must be a multiple of spp_per_pass");
    n_passes = a->spp / spp_per_pass;
    if (a->sample_count) {       // sample shard: the shard's own wavefront has sample_count lanes per pixel
        if ((uint64_t) a->sample_begin + a->sample_count > spp_per_pass) return fail(-2, "sample shard exceeds the samples per pixel of a pass");
        wavefront = npix * a->sample_count;
    }
    return 0;
}

static int setup_params(b2_scene *s, const b2_render_args *a, int mode, bool count, RenderParams &p, int &blocks, size_t &smem, uint64_t lane_begin, uint64_t lane_end) {
    uint32_t spp_per_pass, n_passes; uint64_t wavefront;
    if (int rc = derive_passes(s, a, spp_per_pass, n_passes, wavefront)) return rc;
    if (!s->d.nodes && s->d.n_faces) return fail(-3, "render: b2_scene_build_accel has not been called");
    if (a->max_depth < 0) return fail(-2, "render: max_depth = -1 (unbounded) is not supported; pass a finite depth");
    const DSensor &sn = s->d.sensor;
    p.sc = s->d;
    p.remap_spp = 0; p.remap_off = 0;
    if (a->sample_count) { p.remap_spp = spp_per_pass; p.remap_off = a->sample_begin; spp_per_pass = a->sample_count; }
    p.seed = a->seed; p.spp_per_pass = spp_per_pass; p.n_passes = mode == MODE_RENDER ? n_passes : 1;
    p.spp_shift = 0xffffffffu;
    for (uint32_t k = 0; k < 32; ++k) if ((1u << k) == spp_per_pass) p.spp_shift = k;
    p.max_depth = a->max_depth; p.rr_depth = a->rr_depth; p.env_warp = a->env_warp;
    p.cull_nee = (mode == MODE_RENDER && !(a->reserved & 2u)) ? 1 : 0;   // never in render_backward: a zero albedo still has a gradient
    p.filter_n = sn.rfilter == B2_RFILTER_BOX ? 0 : (int) std::ceil(sn.rfilter_radius - .5f);
    p.lane_begin = lane_begin; p.lane_end = lane_end ? std::min(lane_end, wavefront) : wavefront;
    if (p.lane_begin > p.lane_end) return fail(-2, "render: lane_begin > lane_end");
    const uint32_t W = (uint32_t) sn.grid_w;
    // unit = run of pixels of one row x all samples; aim for >= 8 units per resident CTA, 1k..8k lanes per unit
    const int variant = mode == MODE_RENDER ? a->variant : B2_VARIANT_PRB;
    if (variant < 0 || variant > B2_VARIANT_PRB) return fail(-2, "render: unknown variant");
    uint64_t shard_lanes = p.lane_end - p.lane_begin;
    const bool mega = a->kernel == B2_KERNEL_MEGAKERNEL || mode != MODE_RENDER;
    uint64_t target = (mega && a->chunk_lanes) ? a->chunk_lanes : std::min<uint64_t>(8192, std::max<uint64_t>(1024, shard_lanes / ((uint64_t) s->ctx->sm_count * 2 * 8 + 1)));
    uint32_t P = (uint32_t) std::min<uint64_t>(W, std::max<uint64_t>(1, (target + spp_per_pass - 1) / spp_per_pass));
    p.unit_pixels = P; p.units_per_row = (W + P - 1) / P;
    p.tile_cols = P + 2 * p.filter_n + 1; p.tile_rows = 2 * p.filter_n + 2;
    smem = mode == MODE_BACKWARD ? 0 : sizeof(float4) * (size_t) p.tile_cols * p.tile_rows;
    if (smem > 200 * 1024) return fail(-2, "render: film tile does not fit in shared memory (lower chunk_lanes)");
    if (p.lane_end > p.lane_begin) {
        uint64_t pix0 = p.lane_begin / spp_per_pass, pix1 = (p.lane_end - 1) / spp_per_pass;
        p.unit_begin = (uint32_t) ((pix0 / W) * p.units_per_row + (pix0 % W) / P);
        p.unit_end = (uint32_t) ((pix1 / W) * p.units_per_row + (pix1 % W) / P) + 1;
    } else p.unit_begin = p.unit_end = 0;
    p.unit_counter = s->unit_counter; p.counters = count ? s->counters_dev : nullptr;
    p.grad_in = nullptr; p.wfilm = nullptr; p.n_targets = 0; p.film = nullptr; p.forward = 0;
    int occ = render_occupancy(variant, mode, count, smem);
    if (occ <= 0) return fail(-12, "render: kernel cannot be resident (occupancy 0)");
    uint32_t units = p.unit_end - p.unit_begin;
    blocks = (int) std::max<uint32_t>(1, std::min<uint32_t>(units, (uint32_t) (occ * s->ctx->sm_count)));
    return 0;
}

static void export_counters(const unsigned long long *h, uint64_t launches, b2_counters *o) {
    o->samples = h[C_SAMPLES]; o->closest_rays = h[C_CLOSEST]; o->shadow_rays = h[C_SHADOW]; o->vertices = h[C_VERTICES]; o->tex_evals = h[C_TEX];
    o->nee_area = h[C_NEE_AREA]; o->nee_env = h[C_NEE_ENV]; o->env_escapes = h[C_ESCAPES]; o->splat_taps = h[C_TAPS]; o->grad_atomics = h[C_GRAD];
    o->node_fetches = h[C_NODES]; o->tri_fetches = h[C_TRIS]; o->launches = launches;
}

// The wide nodes (14-21 MB) are the hottest gather target of k_wf_trace; the lane-state streams (GBs per bounce) would
// otherwise keep evicting them from L2.  For the duration of a render the node array is mapped to the persisting part of L2
// (access-policy window on the caller's stream, hit ratio 1), everything else stays normal; the window is removed afterwards.
// Measured on the bedroom stand-in: trace 44.7 -> 43.7 ms.  Also pinning the triangle records (86 MB > the 79 MB persisting
// maximum) was a loss (shade 24 -> 44 ms: the streams lose their share of L2).  B2_L2_PERSIST=0 disables.
struct L2Pin {
    cudaStream_t st; bool on = false;
    L2Pin(const b2_scene *s, cudaStream_t stream) : st(stream) {
        static const int enabled = std::getenv("B2_L2_PERSIST") ? std::atoi(std::getenv("B2_L2_PERSIST")) : 1;
        if (!enabled || !s->d.nodes) return;
        int max_persist = 0, max_window = 0;
        cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, s->ctx->device);
        cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, s->ctx->device);
        const size_t bytes = std::min<size_t>((size_t) s->bvh.n_nodes * 80, (size_t) std::max(max_window, 0));
        if (!bytes || max_persist <= 0) return;
        size_t cur = 0; cudaDeviceGetLimit(&cur, cudaLimitPersistingL2CacheSize);
        if (cur < std::min<size_t>(bytes, (size_t) max_persist)) cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, std::min<size_t>(bytes, (size_t) max_persist));
        cudaStreamAttrValue v{};
        v.accessPolicyWindow.base_ptr = (void *) s->d.nodes; v.accessPolicyWindow.num_bytes = bytes;
        v.accessPolicyWindow.hitRatio = std::min(1.f, (float) max_persist / (float) bytes);
        v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting; v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        on = cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &v) == cudaSuccess;
        cudaGetLastError();
    }
    ~L2Pin() {
        if (!on) return;
        cudaStreamAttrValue v{}; v.accessPolicyWindow.num_bytes = 0;
        cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &v);
        cudaGetLastError();
    }
};

int b2_render(b2_scene *s, const b2_render_args *a, float *film, uintptr_t stream, b2_counters *counters) {
    if (!s || !a || !film) return fail(-1, "b2_render: null argument");
    CUDA_OK(cudaSetDevice(s->ctx->device));
    cudaStream_t st = (cudaStream_t) stream;
    RenderParams p; int blocks; size_t smem; const bool count = counters != nullptr;
    if (int rc = setup_params(s, a, MODE_RENDER, count, p, blocks, smem, a->lane_begin, a->lane_end)) return rc;
    p.film = film;
    uint64_t launches = 0;
    CUDA_OK(cudaMemsetAsync(s->unit_counter, 0, 16, st));
    if (count) CUDA_OK(cudaMemsetAsync(s->counters_dev, 0, sizeof(unsigned long long) * C_N, st));
    if (a->kernel != B2_KERNEL_WAVEFRONT && a->kernel != B2_KERNEL_MEGAKERNEL) return fail(-2, "render: unknown kernel");
    L2Pin pin(s, st);
    if (p.lane_end > p.lane_begin && a->kernel == B2_KERNEL_WAVEFRONT) {
        // wavefront: lane state of one chunk lives in HBM (188 B per lane); default chunk = 16 Mi lanes (3.2 GB)
        if (int rc = wf_reserve(s, p.lane_end - p.lane_begin, a->chunk_lanes, st)) return rc;
        const char *err = nullptr;
        wf_set_timing(s->timing_on ? &s->timing : nullptr);
        const long n = wf_render(a->variant, count, p, s->wf, s->ctx->sm_count, st, &err);
        wf_set_timing(nullptr);
        if (n < 0) return fail(-11, err ? err : "wavefront render failed");
        launches += (uint64_t) n;
    } else if (p.unit_end > p.unit_begin) { launch_render(a->variant, MODE_RENDER, count, p, blocks, smem, st); launches++; }
    CUDA_OK(cudaGetLastError());
    if (count) {
        unsigned long long h[C_N];
        CUDA_OK(cudaMemcpyAsync(h, s->counters_dev, sizeof(h), cudaMemcpyDeviceToHost, st));
        CUDA_OK(cudaStreamSynchronize(st));
        export_counters(h, launches, counters);
    }
    s->launches += launches;
    return 0;
}

int b2_develop(b2_scene *s, const float *film, float *image, uintptr_t stream) {
    if (!s || !film || !image) return fail(-1, "b2_develop: null argument");
    CUDA_OK(cudaSetDevice(s->ctx->device));
    launch_develop(film, image, (uint32_t) s->d.sensor.crop_width * s->d.sensor.crop_height, (cudaStream_t) stream);
    CUDA_OK(cudaGetLastError());
    return 0;
}

// PRB differentiates d value / d reflectance in closed form (b2_shade.cuh, bsdf_eval_pdf in derivative mode): diffuse.reflectance,
// roughconductor.specular_reflectance, roughplastic.diffuse_reflectance (linear and nonlinear, roughplastic.cpp:315-367) and
// roughdielectric.specular_reflectance.  Delta lobes have no eval() and are not targets (detached PRB yields no gradient for them).
static int check_grad_target(const b2_scene *s, const b2_grad_target &g, const char *who) {
    const int t = s->bsdfs[g.bsdf].type;
    if (t == B2_BSDF_DIFFUSE || t == B2_BSDF_ROUGHCONDUCTOR || t == B2_BSDF_ROUGHPLASTIC || t == B2_BSDF_ROUGHDIELECTRIC) return 0;
    return fail(-2, std::string(who) + ": the reflectance of a delta BSDF (conductor, dielectric) is not a differentiable target");
}

int b2_render_backward(b2_scene *s, const b2_render_args *a, const float *grad_in, const b2_grad_target *targets, int n_targets,
                       uintptr_t stream, b2_counters *counters) {
    if (!s || !a || !grad_in || (n_targets && !targets)) return fail(-1, "b2_render_backward: null argument");
    if (n_targets < 0 || n_targets > kMaxGradTargets) return fail(-2, "b2_render_backward: too many gradient targets");
    CUDA_OK(cudaSetDevice(s->ctx->device));
    cudaStream_t st = (cudaStream_t) stream;
    L2Pin pin(s, st);
    uint32_t spp_per_pass, n_passes; uint64_t wavefront;
    if (int rc = derive_passes(s, a, spp_per_pass, n_passes, wavefront)) return rc;
    if (n_passes != 1) return fail(-2, "render_backward: wavefront exceeds 2^32 lanes (common.py:358-363)");
    const size_t film_bytes = sizeof(float) * 4 * (size_t) s->d.sensor.crop_width * s->d.sensor.crop_height;
    if (!s->wfilm) { CUDA_OK(cudaMalloc((void **) &s->wfilm, film_bytes)); s->allocs.push_back(s->wfilm); }
    const bool count = counters != nullptr;
    uint64_t launches = 0;
    // (1) weight film of this sample set: splat weight 1 for every lane of the wavefront (common.py:728-744);
    //     args->reserved bit 0: the weight film of an earlier call with the same (seed, spp) is still valid (sharded bands)
    if (!(a->reserved & 1u)) {
        RenderParams p; int blocks; size_t smem;
        b2_render_args wa = *a; wa.sample_begin = 0; wa.sample_count = 0;      // W of the WHOLE sample set, also for a sample shard
        if (int rc = setup_params(s, &wa, MODE_WEIGHTS, false, p, blocks, smem, 0, 0)) return rc;
        p.film = s->wfilm;
        CUDA_OK(cudaMemsetAsync(s->wfilm, 0, film_bytes, st));
        CUDA_OK(cudaMemsetAsync(s->unit_counter, 0, 16, st));
        if (a->kernel == B2_KERNEL_WAVEFRONT) {
            const char *err = nullptr;
            const long n = wf_weights(p, s->ctx->sm_count, st, &err);
            if (n < 0) return fail(-11, err ? err : "wavefront weight splat failed");
            launches += (uint64_t) n;
        } else { launch_render(B2_VARIANT_PRB, MODE_WEIGHTS, false, p, blocks, smem, st); launches++; }
        CUDA_OK(cudaGetLastError());
    }
    // (2) primal pass + adjoint replay per lane (common.py:750-773)
    {
        RenderParams p; int blocks; size_t smem;
        if (int rc = setup_params(s, a, MODE_BACKWARD, count, p, blocks, smem, a->lane_begin, a->lane_end)) return rc;
        p.grad_in = grad_in; p.wfilm = s->wfilm; p.n_targets = n_targets;
        for (int t = 0; t < n_targets; ++t) {
            const b2_grad_target &g = targets[t];
            if (g.bsdf < 0 || g.bsdf >= (int) s->bsdfs.size() || !g.grad) return fail(-2, "b2_render_backward: bad gradient target");
            const bool textured = s->bsdfs[g.bsdf].reflectance_tex >= 0;
            if ((g.kind == 1) != textured) return fail(-2, "b2_render_backward: target kind does not match the BSDF's reflectance");
            if (int rc = check_grad_target(s, g, "b2_render_backward")) return rc;
            p.targets[t] = GradTargetDev{ g.bsdf, g.kind, g.grad };
        }
        CUDA_OK(cudaMemsetAsync(s->unit_counter, 0, 16, st));
        if (count) CUDA_OK(cudaMemsetAsync(s->counters_dev, 0, sizeof(unsigned long long) * C_N, st));
        if (a->kernel == B2_KERNEL_WAVEFRONT && a->max_depth <= 32) {
            if (p.lane_end > p.lane_begin) {
                if (int rc = wf_reserve(s, p.lane_end - p.lane_begin, a->chunk_lanes, st)) return rc;
                const char *err = nullptr;
                wf_set_timing(s->timing_on ? &s->timing : nullptr);
                const long n = wf_render_backward(count, p, s->wf, wf_reserve_history(s, a->max_depth, st), s->ctx->sm_count, st, &err);
                wf_set_timing(nullptr);
                if (n < 0) return fail(-11, err ? err : "wavefront render_backward failed");
                launches += (uint64_t) n;
            }
        } else if (p.unit_end > p.unit_begin) { launch_render(B2_VARIANT_PRB, MODE_BACKWARD, count, p, blocks, smem, st); launches++; }
        CUDA_OK(cudaGetLastError());
    }
    if (count) {
        unsigned long long h[C_N];
        CUDA_OK(cudaMemcpyAsync(h, s->counters_dev, sizeof(h), cudaMemcpyDeviceToHost, st));
        CUDA_OK(cudaStreamSynchronize(st));
        export_counters(h, launches, counters);
    }
    s->launches += launches;
    return 0;
}

int b2_render_forward(b2_scene *s, const b2_render_args *a, const b2_grad_target *tangents, int n_targets, float *film_rgbw_dev,
                      uintptr_t stream, b2_counters *counters) {
    if (!s || !a || !film_rgbw_dev || (n_targets && !tangents)) return fail(-1, "b2_render_forward: null argument");
    if (n_targets < 0 || n_targets > kMaxGradTargets) return fail(-2, "b2_render_forward: too many tangent targets");
    if (a->kernel != B2_KERNEL_WAVEFRONT || a->max_depth > 32) return fail(-2, "b2_render_forward: wavefront kernel with max_depth <= 32 only");
    CUDA_OK(cudaSetDevice(s->ctx->device));
    cudaStream_t st = (cudaStream_t) stream;
    L2Pin pin(s, st);
    uint32_t spp_per_pass, n_passes; uint64_t wavefront;
    if (int rc = derive_passes(s, a, spp_per_pass, n_passes, wavefront)) return rc;
    if (n_passes != 1) return fail(-2, "render_forward: wavefront exceeds 2^32 lanes (common.py:358-363)");
    const bool count = counters != nullptr;
    uint64_t launches = 0;
    RenderParams p; int blocks; size_t smem;
    if (int rc = setup_params(s, a, MODE_BACKWARD, count, p, blocks, smem, a->lane_begin, a->lane_end)) return rc;
    p.forward = 1; p.film = film_rgbw_dev; p.n_targets = n_targets;
    for (int t = 0; t < n_targets; ++t) {
        const b2_grad_target &g = tangents[t];
        if (g.bsdf < 0 || g.bsdf >= (int) s->bsdfs.size() || !g.grad) return fail(-2, "b2_render_forward: bad tangent target");
        const bool textured = s->bsdfs[g.bsdf].reflectance_tex >= 0;
        if ((g.kind == 1) != textured) return fail(-2, "b2_render_forward: target kind does not match the BSDF's reflectance");
        if (int rc = check_grad_target(s, g, "b2_render_forward")) return rc;
        p.targets[t] = GradTargetDev{ g.bsdf, g.kind, g.grad };
    }
    CUDA_OK(cudaMemsetAsync(s->unit_counter, 0, 16, st));
    if (count) CUDA_OK(cudaMemsetAsync(s->counters_dev, 0, sizeof(unsigned long long) * C_N, st));
    if (p.lane_end > p.lane_begin) {
        if (int rc = wf_reserve(s, p.lane_end - p.lane_begin, a->chunk_lanes, st)) return rc;
        const char *err = nullptr;
        wf_set_timing(s->timing_on ? &s->timing : nullptr);
        const long n = wf_render_backward(count, p, s->wf, wf_reserve_history(s, a->max_depth, st), s->ctx->sm_count, st, &err);
        wf_set_timing(nullptr);
        if (n < 0) return fail(-11, err ? err : "wavefront render_forward failed");
        launches += (uint64_t) n;
    }
    CUDA_OK(cudaGetLastError());
    if (count) {
        unsigned long long h[C_N];
        CUDA_OK(cudaMemcpyAsync(h, s->counters_dev, sizeof(h), cudaMemcpyDeviceToHost, st));
        CUDA_OK(cudaStreamSynchronize(st));
        export_counters(h, launches, counters);
    }
    s->launches += launches;
    return 0;
}

int b2_set_kernel_timing(b2_scene *s, int enabled) {
    if (!s) return fail(-1, "b2_set_kernel_timing: null argument");
    s->timing_on = enabled != 0;
    return 0;
}

int b2_kernel_times(b2_scene *s, double ms[4], uint64_t launches[4]) {
    if (!s || !ms || !launches) return fail(-1, "b2_kernel_times: null argument");
    CUDA_OK(cudaSetDevice(s->ctx->device));
    CUDA_OK(cudaDeviceSynchronize());
    wf_collect_timing(s->timing, ms, launches);
    return 0;
}

int b2_trace(b2_scene *s, const float *rays, uint64_t n, int any_hit, b2_hit *hits, uintptr_t stream) {
    if (!s || (n && (!rays || !hits))) return fail(-1, "b2_trace: null argument");
    if (!s->d.nodes && s->d.n_faces) return fail(-3, "b2_trace: b2_scene_build_accel has not been called");
    CUDA_OK(cudaSetDevice(s->ctx->device));
    if (any_hit & 2) launch_trace(s->d, rays, n, any_hit & 1, hits, (cudaStream_t) stream);      // the megakernel's per-thread traversal (b2_bvh.cuh)
    else if (n) {                                                                                   // the production kernel, k_wf_trace
        if (int rc = wf_reserve(s, n, 0, (cudaStream_t) stream)) return rc;
        RenderParams p{}; p.sc = s->d; p.counters = nullptr;
        const char *err = nullptr;
        const long l = wf_trace_rays(p, s->wf, rays, n, any_hit & 1, hits, s->ctx->sm_count, (cudaStream_t) stream, &err);
        if (l < 0) return fail(-11, err ? err : "wavefront trace failed");
        s->launches += (uint64_t) l;
    }
    CUDA_OK(cudaGetLastError());
    return 0;
}

// No counterpart in the reference: b2_trace without any device allocation, the lane
// state of the rays in the static device array of wf_noalloc.cu.
int b2_trace_noalloc(b2_scene *s, const float *rays, uint64_t n, int any_hit, b2_hit *hits, uintptr_t stream) {
    if (!s || (n && (!rays || !hits))) return fail(-1, "b2_trace_noalloc: null argument");
    if (any_hit & ~1) return fail(-2, "b2_trace_noalloc: any_hit must be 0 (closest hit) or 1 (arrival test)");
    if (n > B2_TRACE_NOALLOC_MAX_RAYS) return fail(-2, "b2_trace_noalloc: n = " + std::to_string(n) + " rays exceeds B2_TRACE_NOALLOC_MAX_RAYS = 32768 (the size of the static scratch array)");
    if (!s->d.nodes && s->d.n_faces) return fail(-3, "b2_trace_noalloc: b2_scene_build_accel has not been called");
    CUDA_OK(cudaSetDevice(s->ctx->device));
    if (n) {
        RenderParams p{}; p.sc = s->d; p.counters = nullptr;
        const char *err = nullptr;
        const long l = wf_trace_rays_noalloc(p, s->class_mask, s->ctx->device, rays, n, any_hit, hits, s->ctx->sm_count, (cudaStream_t) stream, &err);
        if (l < 0) return fail(l == -2 ? -2 : -11, err ? err : "b2_trace_noalloc failed");
        s->launches += (uint64_t) l;
    }
    CUDA_OK(cudaGetLastError());
    return 0;
}

int b2_debug_sampler(b2_scene *s, const b2_render_args *a, uint64_t lane_begin, uint32_t n_lanes, uint32_t n_draws, float *out, float *pos, uintptr_t stream) {
    if (!s || !a || !out) return fail(-1, "b2_debug_sampler: null argument");
    CUDA_OK(cudaSetDevice(s->ctx->device));
    RenderParams p; int blocks; size_t smem;
    if (int rc = setup_params(s, a, MODE_WEIGHTS, false, p, blocks, smem, 0, 0)) return rc;
    launch_debug_sampler(p, lane_begin, n_lanes, n_draws, out, pos, (cudaStream_t) stream);
    CUDA_OK(cudaGetLastError());
    return 0;
}

int b2_debug_primary_hits(b2_scene *s, const b2_render_args *a, uint64_t lane_begin, uint32_t n_lanes, int kernel, b2_hit *hits, uintptr_t stream) {
    if (!s || !a || (n_lanes && !hits)) return fail(-1, "b2_debug_primary_hits: null argument");
    CUDA_OK(cudaSetDevice(s->ctx->device));
    RenderParams p; int blocks; size_t smem;
    if (int rc = setup_params(s, a, MODE_WEIGHTS, false, p, blocks, smem, 0, 0)) return rc;
    if (lane_begin + n_lanes > p.lane_end) return fail(-2, "b2_debug_primary_hits: lanes beyond the wavefront of the pass");
    if (kernel < 0 || kernel > 2) return fail(-2, "b2_debug_primary_hits: kernel must be 0 (per ray), 1 or 2 (packets of 32 / 64 lanes)");
    if (n_lanes) {
        if (int rc = wf_reserve(s, n_lanes, 0, (cudaStream_t) stream)) return rc;
        const char *err = nullptr;
        const long l = wf_trace_primary(p, s->wf, lane_begin, n_lanes, kernel, hits, s->ctx->sm_count, (cudaStream_t) stream, &err);
        if (l < 0) return fail(-11, err ? err : "primary trace failed");
        s->launches += (uint64_t) l;
    }
    CUDA_OK(cudaGetLastError());
    return 0;
}

int b2_render_host(b2_scene *s, const b2_render_args *a, float *film_host, float *image_host, b2_counters *counters) {
    if (!s || !a || !film_host) return fail(-1, "b2_render_host: null argument");
    CUDA_OK(cudaSetDevice(s->ctx->device));
    const size_t npix = (size_t) s->d.sensor.crop_width * s->d.sensor.crop_height;
    // staging buffers live with the scene (allocated on first use): no allocator call per render
    if (!s->host_film) { CUDA_OK(cudaMalloc((void **) &s->host_film, npix * 16)); s->allocs.push_back(s->host_film); }
    if (image_host && !s->host_image) { CUDA_OK(cudaMalloc((void **) &s->host_image, npix * 12)); s->allocs.push_back(s->host_image); }
    cudaStream_t st = 0;
    // the film upload runs on its own stream: the wavefront touches the film only in its splat launches, which wait for the event
    if (!s->copy_stream) { CUDA_OK(cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking)); CUDA_OK(cudaEventCreateWithFlags(&s->film_ready, cudaEventDisableTiming)); }
    CUDA_OK(cudaMemcpyAsync(s->host_film, film_host, npix * 16, cudaMemcpyHostToDevice, s->copy_stream));
    CUDA_OK(cudaEventRecord(s->film_ready, s->copy_stream));
    const bool overlap = a->kernel == B2_KERNEL_WAVEFRONT;
    if (!overlap) CUDA_OK(cudaStreamWaitEvent(st, s->film_ready, 0));
    wf_set_film_ready_event(overlap ? s->film_ready : nullptr);
    int rc = b2_render(s, a, s->host_film, (uintptr_t) st, counters);
    wf_set_film_ready_event(nullptr);
    CUDA_OK(cudaStreamWaitEvent(st, s->film_ready, 0));          // (a render without a splat launch: empty lane range)
    if (rc == 0 && image_host) rc = b2_develop(s, s->host_film, s->host_image, (uintptr_t) st);
    if (rc != 0) return rc;
    CUDA_OK(cudaMemcpyAsync(film_host, s->host_film, npix * 16, cudaMemcpyDeviceToHost, st));
    if (image_host) CUDA_OK(cudaMemcpyAsync(image_host, s->host_image, npix * 12, cudaMemcpyDeviceToHost, st));
    CUDA_OK(cudaStreamSynchronize(st));
    return 0;
}

}  // extern "C"
