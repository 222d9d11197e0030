/*
 * b2path.h -- C-ABI of the B200-native path tracer / PRB adjoint ("b2path").
 *
 * This is the drop-in boundary for ONE hot path of DoeringChristian/mitsuba3-experiments:
 * the unidirectional path tracer of path.py / path-mis.py and the PRB adjoint
 * (mitsuba3/src/python/python/ad/integrators/prb.py).  Everything here is `extern "C"`,
 * plain pointers and sizes; no C++ / torch types cross the boundary.  Every function returns
 * 0 on success and a negative error code otherwise; b2_last_error() returns a thread-local,
 * human readable message (the reference raises std::runtime_error -> Python RuntimeError,
 * mitsuba3/src/render/python/integrator_v.cpp:98-113; the Python host layer re-raises).
 *
 * The same plain-data scene description (b2_scene_desc) is consumed by
 *   - the CUDA product library  (mitsuba3_experiments_b200/csrc  -> libb2path.so)
 *   - the CPU oracle            (oracle/                         -> liboracle.so, test-only)
 *
 * Reference interfaces replaced (file:line relative to /root/reference, mi3 = mitsuba3):
 *   b2_scene_create        <- mi.load_dict / Scene ctor + Mesh buffers
 *                             (mi3/src/render/scene.cpp:22-111, mi3/include/mitsuba/render/mesh.h:564-568)
 *   b2_scene_build_accel   <- rtcCommitScene / optixAccelBuild
 *                             (mi3/src/render/scene_embree.inl:115-167, scene_optix.inl:473-506)
 *   b2_scene_set_param     <- SceneParameters.update() (mi3/src/python/python/util.py:271-337)
 *   b2_render              <- SamplingIntegrator::render JIT branch + render_sample
 *                             (mi3/src/render/integrator.cpp:246-360,419-491) running
 *                             Path.sample (path.py:194-302) / PathIntegrator.sample (path-mis.py:24-155)
 *                             / PRBIntegrator.sample primal (prb.py:63-265)
 *   b2_develop             <- HDRFilm::develop (mi3/src/films/hdrfilm.cpp:304-405)
 *   b2_render_backward     <- RBIntegrator.render_backward (mi3/src/python/python/ad/integrators/common.py:623-781)
 *   b2_render_forward      <- RBIntegrator.render_forward  (mi3/src/python/python/ad/integrators/common.py:495-621)
 *   b2_kernel_times        <- dr.kernel_history() as printed by path.py:332-351
 *   b2_trace               <- Scene::ray_intersect_preliminary / ray_test (mi3/src/render/scene.cpp:191-235)
 *   b2_debug_sampler       <- PCG32Sampler::seed + next_1d (mi3/src/render/sampler.cpp:129-148, samplers/independent.cpp:88-97)
 *   b2_debug_primary_hits  <- SamplingIntegrator::sample_rays + Scene::ray_intersect_preliminary (integrator.cpp:419-436, scene.cpp:191-214)
 *   b2_trace_noalloc       <- no counterpart in the reference
 */
#ifndef B2PATH_H
#define B2PATH_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2_ABI_VERSION 2

/* ------------------------------------------------------------------ enums */
enum { B2_BSDF_DIFFUSE = 0, B2_BSDF_ROUGHCONDUCTOR = 1, B2_BSDF_DIELECTRIC = 2,
       B2_BSDF_CONDUCTOR = 3,      /* smooth conductor (conductor.cpp:249-307), SURVEY 8(f) rank 1 */
       B2_BSDF_ROUGHDIELECTRIC = 4,/* rough dielectric (roughdielectric.cpp:242-605): `ior` = int/ext, alpha, distribution,
                                      reflectance = specular_reflectance, transmittance = specular_transmittance */
       B2_BSDF_ROUGHPLASTIC = 5    /* rough plastic (roughplastic.cpp:216-430): reflectance(+tex) = diffuse_reflectance, transmittance =
                                      specular_reflectance, `ior`, alpha_u (isotropic), distribution, nonlinear, spec_sampling_weight;
                                      the 64-entry transmittance table and the internal reflectance are computed by each implementation */ };
enum { B2_BSDF_FLAG_TWOSIDED = 1,             /* wrapped in a `twosided` adapter (twosided.cpp:110-260) */
       B2_BSDF_FLAG_MASK = 2 };               /* outermost `mask` adapter with constant `opacity` (mask.cpp:123-215), SURVEY 8(f) rank 1 */
enum { B2_MF_BECKMANN = 0, B2_MF_GGX = 1 };   /* MicrofacetType (microfacet.h) */
enum { B2_TEX_F32 = 0, B2_TEX_F16 = 1 };      /* storage (bitmap.cpp:231-246) */
enum { B2_WRAP_REPEAT = 0, B2_WRAP_MIRROR = 1, B2_WRAP_CLAMP = 2 };
enum { B2_FILTER_BILINEAR = 0, B2_FILTER_NEAREST = 1 };
enum { B2_EMITTER_AREA = 0, B2_EMITTER_ENVMAP = 1, B2_EMITTER_CONSTANT = 2,
       B2_EMITTER_POINT = 3 /* emitters/point.cpp: position = translation of to_world, intensity in `radiance` */ };
enum { B2_SHAPE_RECTANGLE = 0, B2_SHAPE_MESH = 1 }; /* how an area light is sampled (rectangle.cpp:182-200 | mesh.cpp:758-812) */
enum { B2_RFILTER_BOX = 0, B2_RFILTER_TENT = 1, B2_RFILTER_GAUSSIAN = 2 };
enum { B2_MESH_HAS_NORMALS = 1, B2_MESH_HAS_UV = 2, B2_MESH_FLIP_NORMALS = 4 };
/* loop formulation: which reference script's arithmetic is reproduced */
enum { B2_VARIANT_PATH = 0 /* path.py */, B2_VARIANT_PATH_MIS = 1 /* path-mis.py == path.cpp */, B2_VARIANT_PRB = 2 /* prb.py primal */ };
enum { B2_KERNEL_WAVEFRONT = 0, B2_KERNEL_MEGAKERNEL = 1 };
enum { B2_ENV_WARP_HIERARCHICAL = 0 /* per-sample parity with envmap.cpp */, B2_ENV_WARP_ALIAS = 1 /* fast mode, parity only in expectation */ };

#define B2_MAX_WARP_LEVELS 24

/* ------------------------------------------------------------ scene input */
typedef struct b2_texture {
    const void *data;       /* [height][width][3], f32 or f16 (host pointer, copied by scene_create) */
    int32_t width, height;
    int32_t format;         /* B2_TEX_* */
    int32_t wrap;           /* B2_WRAP_* */
    int32_t filter;         /* B2_FILTER_* */
    float to_uv[9];         /* row-major 3x3 affine (bitmap.cpp:141) */
} b2_texture;

typedef struct b2_bsdf {
    int32_t type;           /* B2_BSDF_* */
    int32_t flags;          /* B2_BSDF_FLAG_* */
    int32_t reflectance_tex;/* diffuse `reflectance` / conductor `specular_reflectance` texture index, -1 = constant */
    int32_t distribution;   /* B2_MF_* */
    int32_t sample_visible; /* microfacet visible-normal sampling (default 1) */
    float reflectance[3];   /* diffuse.reflectance | roughconductor/dielectric.specular_reflectance */
    float transmittance[3]; /* dielectric.specular_transmittance */
    float eta[3];           /* conductor eta (real part) */
    float k[3];             /* conductor k */
    float alpha_u, alpha_v;
    float ior;              /* dielectric: int_ior / ext_ior */
    float opacity;          /* B2_BSDF_FLAG_MASK: probability of interacting with the nested BSDF; else 1 */
    int32_t nonlinear;      /* roughplastic: account for nonlinear colour shifts due to internal scattering */
    float spec_sampling_weight; /* roughplastic: s_mean / (d_mean + s_mean) from the texture means (roughplastic.cpp:224-230) */
} b2_bsdf;

typedef struct b2_mesh {
    uint32_t first_vertex, n_vertices;
    uint32_t first_face, n_faces;
    int32_t bsdf;           /* index into bsdfs */
    int32_t emitter;        /* index into emitters or -1 */
    uint32_t flags;         /* B2_MESH_* */
} b2_mesh;

typedef struct b2_emitter {
    int32_t type;           /* B2_EMITTER_* */
    int32_t mesh;           /* area: mesh index */
    int32_t shape_kind;     /* B2_SHAPE_* */
    float radiance[3];      /* area/constant radiance */
    float to_world[16];     /* rectangle: object->world, row-major (rectangle.cpp:186-188) */
    float normal[3];        /* rectangle: m_frame.n (rectangle.cpp:104) */
    float inv_area;         /* 1 / surface area (rectangle.cpp:108 | mesh.cpp:443-477) */
    uint32_t cdf_offset;    /* mesh light: offset of its per-face area CDF in `area_cdf` */
} b2_emitter;

typedef struct b2_envmap {
    const float *data;      /* [height][width][3] f32, width = source width + 1 (envmap.cpp:147-149,229-230) */
    int32_t width, height;
    float scale;
    float to_world[16];     /* row-major */
    float from_world[16];
    float bsphere_center[3];
    float bsphere_radius;   /* envmap.cpp:337-344 */
    /* Hierarchical2D<Float,0> (distr_2d.h:403-509): level 0 = normalised input [height][width],
       level l>=1 = 2x2-block-swizzled mip levels */
    int32_t n_levels;
    const float *warp_data;
    uint32_t level_offset[B2_MAX_WARP_LEVELS];
    uint32_t level_width[B2_MAX_WARP_LEVELS];
    uint32_t level_size[B2_MAX_WARP_LEVELS];
    /* optional alias table over the (width-1)x(height-1) bilinear patches (fast mode) */
    const float *alias_prob;
    const uint32_t *alias_idx;
} b2_envmap;

typedef struct b2_sensor {
    float sample_to_camera[16]; /* row-major; perspective.cpp:173-180 */
    float to_world[16];
    float near_clip, far_clip;
    int32_t film_width, film_height;   /* film.size() */
    int32_t crop_offset_x, crop_offset_y, crop_width, crop_height;
    int32_t rfilter;        /* B2_RFILTER_* */
    float rfilter_radius;
    float gaussian_coeff[10]; /* gaussian.cpp:69-89 (llvm Remez/Estrin coefficients) */
    int32_t sample_border;  /* film.sample_border (film.cpp:33): samples are also drawn on a rim of border_size =
                               ceil(radius - 0.5) pixels around the crop window (integrator.cpp:134-136,306-307) */
    int32_t reserved;
} b2_sensor;

typedef struct b2_scene_desc {
    uint32_t abi_version;
    uint32_t n_vertices, n_faces, n_meshes, n_bsdfs, n_textures, n_emitters;
    const float *positions;   /* [n_vertices*3] world space */
    const float *normals;     /* [n_vertices*3] (valid where mesh has B2_MESH_HAS_NORMALS) */
    const float *texcoords;   /* [n_vertices*2] */
    const uint32_t *faces;    /* [n_faces*3] global vertex indices */
    const b2_mesh *meshes;
    const b2_bsdf *bsdfs;
    const b2_texture *textures;
    const b2_emitter *emitters;
    const float *area_cdf;    /* concatenated per-face area CDFs of mesh lights (may be NULL) */
    uint32_t n_area_cdf;
    const b2_envmap *envmap;  /* NULL if none; if present it is also listed in `emitters` */
    b2_sensor sensor;
} b2_scene_desc;

/* ------------------------------------------------------------ render args */
typedef struct b2_render_args {
    uint32_t seed;          /* sampler seed (sampler.cpp:129-148) */
    uint32_t spp;           /* samples per pixel */
    uint32_t spp_per_pass;  /* 0 = derive as integrator.cpp:139-152,247-265 */
    int32_t max_depth, rr_depth;
    int32_t variant;        /* B2_VARIANT_* */
    int32_t kernel;         /* B2_KERNEL_* (CUDA only) */
    int32_t env_warp;       /* B2_ENV_WARP_* */
    uint64_t lane_begin, lane_end; /* shard: global lane range within one pass wavefront; 0,0 = all */
    uint32_t chunk_lanes;   /* wavefront chunk capacity (CUDA only), 0 = default */
    uint32_t reserved;      /* bit 0 (render_backward only): reuse the weight film of the previous call with the same (seed, spp);
                               bit 1 (render): trace every shadow ray, also those of emitter samples whose value is exactly zero
                               (the reference does; the image is identical either way, only b2_counters.shadow_rays differs) */
    uint32_t sample_begin, sample_count; /* sample shard: render only samples [sample_begin, sample_begin + sample_count) of every pixel
                               of a pass (0, 0 = all).  Lane i of the shard is pixel i / sample_count, sample sample_begin + i % sample_count,
                               and draws the stream of the global lane pixel * spp_per_pass + sample, so the union of the shards is
                               the full render; lane_begin / lane_end then index the shard's own lanes.  Perfectly balanced multi-GPU split. */
} b2_render_args;

typedef struct b2_counters {
    uint64_t samples;        /* lanes carried to termination (x passes) */
    uint64_t closest_rays, shadow_rays;
    uint64_t vertices;       /* surface interactions shaded */
    uint64_t tex_evals;
    uint64_t nee_area, nee_env, env_escapes;
    uint64_t splat_taps;
    uint64_t grad_atomics;
    uint64_t node_fetches, tri_fetches; /* traversal work actually done */
    uint64_t launches;       /* CUDA kernels launched */
} b2_counters;

typedef struct b2_build_stats {
    float ms_morton_sort, ms_lbvh, ms_collapse, ms_total;
    uint32_t n_nodes8, n_tris, max_depth;
    uint32_t builder;        /* B2_BUILDER_* that produced the tree */
    uint64_t node_bytes, tri_bytes;
    float sah_cost;          /* SAH cost of the wide tree / root area, c_node = 1, c_prim = 0.3 */
    float reserved;
} b2_build_stats;

typedef struct b2_hit { float t, u, v; uint32_t prim; /* original face index, 0xffffffff = miss */ } b2_hit;

/* PRB: which parameter receives gradients */
typedef struct b2_grad_target {
    int32_t bsdf;           /* bsdf index */
    int32_t kind;           /* 0: constant reflectance (3 floats), 1: reflectance texture data [h][w][3] f32 */
    float *grad;            /* device pointer (CUDA) / host pointer (oracle), accumulated into */
} b2_grad_target;

/* ------------------------------------------------------------ CUDA product */
typedef struct b2_ctx b2_ctx;
typedef struct b2_scene b2_scene;

int  b2_init(int device, b2_ctx **out);
void b2_shutdown(b2_ctx *ctx);
int  b2_scene_create(b2_ctx *ctx, const b2_scene_desc *desc, b2_scene **out);
void b2_scene_destroy(b2_scene *scene);
/* builder: binary tree over the Morton-sorted triangles that is then collapsed (SAH-optimal DP) into the compressed 8-wide BVH.
   LBVH = Karras radix tree; PLOC = parallel locally-ordered clustering (better SAH, ~1.5x the build time); AUTO builds both
   and keeps PLOC when its SAH cost is below 0.9x the radix tree's; DEFAULT = AUTO unless the environment variable
   B2_BVH_BUILDER (0 | 1 | 2) overrides it. */
#define B2_BUILDER_DEFAULT (-1)
#define B2_BUILDER_LBVH    0
#define B2_BUILDER_PLOC    1
#define B2_BUILDER_AUTO    2
int  b2_scene_build_accel(b2_scene *scene, int builder, uintptr_t stream, b2_build_stats *stats);
/* key = "<bsdf index>.reflectance.value" | "<bsdf index>.reflectance.data"; data is a DEVICE pointer */
int  b2_scene_set_param(b2_scene *scene, const char *key, const void *dev_ptr, size_t bytes, uintptr_t stream);
/* film_rgbw_dev: [crop_h][crop_w][4] f32, accumulated into (not cleared) */
int  b2_render(b2_scene *scene, const b2_render_args *args, float *film_rgbw_dev, uintptr_t stream, b2_counters *counters);
int  b2_develop(b2_scene *scene, const float *film_rgbw_dev, float *image_rgb_dev, uintptr_t stream);
/* grad_in_dev: [crop_h][crop_w][3]; runs weight-film splat, primal pass and adjoint replay for (seed, spp) */
int  b2_render_backward(b2_scene *scene, const b2_render_args *args, const float *grad_in_dev,
                        const b2_grad_target *targets, int n_targets, uintptr_t stream, b2_counters *counters);
/* Forward-mode PRB: RBIntegrator.render_forward (common.py:495-621).  tangents[t].grad_dev = DEVICE pointer to the tangent of
   target t (3 floats for a constant reflectance, [h][w][3] for a bitmap; read only).  The lanes' d L are splatted into
   film_rgbw_dev (accumulated, like b2_render); b2_develop of that film is the gradient image. */
int  b2_render_forward(b2_scene *scene, const b2_render_args *args, const b2_grad_target *tangents, int n_targets,
                       float *film_rgbw_dev, uintptr_t stream, b2_counters *counters);
/* unit-test hook: rays_dev [n][8] = o.xyz, d.xyz, tmax, unused.  any_hit bit 0: arrival test (Scene::ray_test) instead of the closest
   hit; bit 1: trace with the megakernel's per-thread traversal instead of the wavefront trace kernel (the product default). */
int  b2_trace(b2_scene *scene, const float *rays_dev, uint64_t n, int any_hit, b2_hit *hits_dev, uintptr_t stream);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc (code under
   CUDA graph capture, tight serving loops, harnesses that forbid it).  b2_trace through the wavefront trace kernel (any_hit bit 0 as
   above; bit 1 is rejected), with the kernel's lane state in a file-scope device array of the library instead of a cudaMalloc'ed
   block: the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).  n > B2_TRACE_NOALLOC_MAX_RAYS is
   rejected with -2.  The array (8 421 632 B, held by every GPU that loads the library until the process exits) is shared by all
   calls and all scenes of a device: calls that overlap on the GPU race -- serialise them (one stream, or events). */
#define B2_TRACE_NOALLOC_MAX_RAYS 32768u
int  b2_trace_noalloc(b2_scene *scene, const float *rays_dev, uint64_t n, int any_hit, b2_hit *hits_dev, uintptr_t stream);
/* unit-test hook for the sampler (integer work, must be bit exact): out_dev [n_lanes][n_draws] = the first n_draws
   next_float32() of lanes lane_begin .. lane_begin + n_lanes - 1, seeded exactly as the render kernels seed them
   (PCG32Sampler::seed, sampler.cpp:129-148: TEA(seed, lane) -> PCG32, drjit random.h:60-79,211-213); the primary-ray
   position of every lane (the first two draws through lane -> pixel, integrator.cpp:293-310) goes to pos_dev [n_lanes][2]
   when it is not NULL.  `args` supplies seed / spp (lane -> pixel); nothing is rendered. */
int  b2_debug_sampler(b2_scene *scene, const b2_render_args *args, uint64_t lane_begin, uint32_t n_lanes, uint32_t n_draws,
                      float *out_dev, float *pos_dev, uintptr_t stream);
/* unit-test hook for the bounce-0 trace: hits_dev [n_lanes] = the closest hit of the primary ray of lanes lane_begin ..
   lane_begin + n_lanes - 1 of pass 0 (sample position from the lane's stream, sensor ray: integrator.cpp:293-310,419-436,
   perspective.cpp:239-280; Scene::ray_intersect_preliminary, scene.cpp:191-214).  kernel 1 / 2 = the packet kernel the renders use
   for bounce 0 from 8 samples per pixel on (one warp-uniform traversal per 32 / 64 consecutive lanes), kernel 0 = the per-ray trace
   kernel: all must agree bit for bit.  `args` supplies seed / spp / sample shard; nothing is rendered. */
int  b2_debug_primary_hits(b2_scene *scene, const b2_render_args *args, uint64_t lane_begin, uint32_t n_lanes, int kernel,
                           b2_hit *hits_dev, uintptr_t stream);
/* host-buffer convenience entry points (what a reference-side binding with host arrays calls):
   upload/download happen inside; film_rgbw_host is accumulated into */
int  b2_render_host(b2_scene *scene, const b2_render_args *args, float *film_rgbw_host, float *image_rgb_host, b2_counters *counters);
/* Kernel history (the reference prints dr.kernel_history() with per-kernel execution times, path.py:332-351):
   when enabled, CUDA events bracket every wavefront launch; b2_kernel_times synchronises the device and returns the
   milliseconds and launch counts accumulated since the previous call, per kernel kind
   [0] trace  [1] shade (+ classify)  [2] splat  [3] other (ray generation). */
int  b2_set_kernel_timing(b2_scene *scene, int enabled);
int  b2_kernel_times(b2_scene *scene, double ms[4], uint64_t launches[4]);
const char *b2_last_error(void);
int  b2_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* B2PATH_H */
