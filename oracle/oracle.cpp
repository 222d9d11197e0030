// ORACLE (test infrastructure, NOT product code).
//
// CPU restatement of the reference hot path: SamplingIntegrator::render / render_sample
// (mitsuba3/src/render/integrator.cpp:246-360,419-491), the three loop formulations
//   Path.sample            path.py:194-302          (variant 0; with SURVEY fact-5 fix: self.max_depth)
//   PathIntegrator.sample  path-mis.py:24-155       (variant 1; == integrators/path.cpp:94-327)
//   PRBIntegrator.sample   ad/integrators/prb.py:63-265 (variant 2; primal + backward)
// the film (imageblock.cpp:174-560, hdrfilm.cpp:304-405, rfilters/*.cpp) and the PRB driver
// RBIntegrator.render_backward (ad/integrators/common.py:623-781).
//
// Parity pinning: PCG32 / TEA / warps / Fresnel / microfacet / Hierarchical2D are checked against the
// reference's own headers compiled in scalar mode (oracle/_ref, oracle/ref_pin.cpp) and against the
// golden constants of the reference's tests (tests/test_oracle_golden.py).  libm supplies
// sin/cos/exp/log/erf/atan2/acos where llvm_ad_rgb uses Dr.Jit's polynomial versions (<= 2 ulp apart).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this library.
#include "oracle_scene.h"
#include <atomic>
#include <cstdio>
#include <memory>
#include <string>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace orc {

// ---------------------------------------------------------------------------------------------
// Reconstruction filters: gaussian.cpp:94-103 (llvm branch: Estrin polynomial in x^2),
// tent.cpp:53-55, box.cpp:41-43
// ---------------------------------------------------------------------------------------------
static inline float rfilter_eval(const b2_sensor &s, float x) {
    if (s.rfilter == B2_RFILTER_GAUSSIAN) return fmaxf(estrin<10>(sqr(x), s.gaussian_coeff), 0.f);
    if (s.rfilter == B2_RFILTER_TENT) return fmaxf(0.f, 1.f - fabsf(x * (1.f / s.rfilter_radius)));
    return (x >= -.5f && x < .5f) ? 1.f : 0.f;
}

struct Film {
    const b2_sensor &s; int w, h; std::vector<double> acc; // [h][w][4]
    Film(const b2_sensor &sensor) : s(sensor), w(sensor.crop_width), h(sensor.crop_height), acc((size_t) w * h * 4, 0.0) {}
    static void atomic_add(double &dst, double v) {
#pragma omp atomic
        dst += v;
    }
    // footprint of ImageBlock::put, coalesced variant (imageblock.cpp:433-508); box fast path (:210-233).
    // Calls f(pixel_index, weight) for every in-bounds tap.
    template <typename F> void taps(V2 pos, F f, Counters &c) const {
        if (s.rfilter == B2_RFILTER_BOX) {
            int px = (int) floorf(pos.x) - s.crop_offset_x, py = (int) floorf(pos.y) - s.crop_offset_y;
            if ((unsigned) px < (unsigned) w && (unsigned) py < (unsigned) h) { c.taps++; f((size_t) py * w + px, 1.f); }
            return;
        }
        int n = (int) ceilf(s.rfilter_radius - .5f), count = 2 * n + 1;
        int pix = (int) floorf(pos.x) - n, piy = (int) floorf(pos.y) - n;
        int lx = pix - s.crop_offset_x, ly = piy - s.crop_offset_y;
        float relx = (float) pix + .5f - pos.x, rely = (float) piy + .5f - pos.y;
        float wx[16], wy[16];
        for (int i = 0; i < count; ++i) { wx[i] = rfilter_eval(s, relx); wy[i] = rfilter_eval(s, rely); relx += 1.f; rely += 1.f; }
        for (int ys = 0; ys < count; ++ys) {
            unsigned y = (unsigned) (ly + ys); if (y >= (unsigned) h) continue;
            for (int xs = 0; xs < count; ++xs) {
                unsigned x = (unsigned) (lx + xs); if (x >= (unsigned) w) continue;
                c.taps++; f((size_t) y * w + x, wy[ys] * wx[xs]);
            }
        }
    }
    void put(V2 pos, V3 rgb, Counters &c) {
        taps(pos, [&](size_t pix, float wgt) {
            double *p = &acc[pix * 4];
            atomic_add(p[0], (double) (rgb.x * wgt)); atomic_add(p[1], (double) (rgb.y * wgt));
            atomic_add(p[2], (double) (rgb.z * wgt)); atomic_add(p[3], (double) (1.f * wgt));
        }, c);
    }
    // the same into a thread-private accumulator (no atomics; merged once per thread at the end of the render)
    void put_private(double *priv, V2 pos, V3 rgb, Counters &c) const {
        taps(pos, [&](size_t pix, float wgt) {
            double *p = priv + pix * 4;
            p[0] += (double) (rgb.x * wgt); p[1] += (double) (rgb.y * wgt); p[2] += (double) (rgb.z * wgt); p[3] += (double) (1.f * wgt);
        }, c);
    }
};

// ---------------------------------------------------------------------------------------------
// Sensor: PerspectiveCamera::sample_ray_differential (sensors/perspective.cpp:239-280)
// ---------------------------------------------------------------------------------------------
static inline Ray camera_ray(const b2_sensor &s, V2 position_sample) {
    V3 near_p = xf_point_persp(s.sample_to_camera, V3(position_sample.x, position_sample.y, 0.f));
    V3 dl = normalize(near_p);
    Ray r;
    r.o = V3(s.to_world[3], s.to_world[7], s.to_world[11]);
    r.d = xf_vector(s.to_world, dl);
    float inv_z = rcp(dl.z), near_t = s.near_clip * inv_z, far_t = s.far_clip * inv_z;
    r.o = r.o + r.d * near_t;
    r.maxt = far_t - near_t;
    return r;
}

struct Sampler {
    PCG32 rng;
    float next_1d() { return rng.next_float32(); }
    V2 next_2d() { float a = next_1d(), b = next_1d(); return V2{ a, b }; }
};

static inline float mis_weight_isfinite(float a, float b) { // path.py:10-18, common.py:1338-1346
    float a2 = sqr(a), b2 = sqr(b), w = a2 / (a2 + b2);
    return finite(w) ? w : 0.f;
}
// path-mis.py:9-15.  DELIBERATE DEVIATION (documented in DESIGN.md): the script's guard is only `pdf_a > 0`;
// for pdf_a > 1.8e19 (grazing light samples) a2 overflows to inf and inf/inf = NaN poisons the pixel.  path.py,
// prb.py and the C++ plugin the script transliterates (path.cpp:340-345) all add `isfinite(w)`; so do we.
static inline float mis_weight_guard(float a, float b) {
    float a2 = sqr(a), w = a2 / fmaf(b, b, a2);
    return (a > 0.f && finite(w)) ? w : 0.f;
}

struct Integrator {
    const Scene &sc; EmitterSys em; int max_depth, rr_depth;
    Integrator(const Scene &s, int md, int rd) : sc(s), em(s), max_depth(md), rr_depth(rd) {}
    const b2_bsdf *bsdf_of(const SI &si) const { return si.valid ? &sc.d->bsdfs[sc.d->meshes[si.mesh].bsdf] : nullptr; }

    // ------------------------------------------------------------------ path.py:194-302
    V3 sample_path_py(Sampler &smp, Ray ray, Counters &c) const {
        V3 L(0.f), f(1.f); float eta = 1.f; uint32_t depth = 1;
        bool active = depth < (uint32_t) max_depth;
        SI si = active ? sc.make_si(ray, sc.intersect(ray, c), c) : sc.make_si(ray, Hit{ 0, 0, 0, 0xffffffffu }, c);
        if (active) L += em.eval(em.emitter_of(si), si, c);      // path.py:238-239: the emitter lookup and eval are masked by `active`
        while (active) {
            const b2_bsdf *bsdf = bsdf_of(si);
            bool active_em = active && bsdf && (bsdf_lobes(*bsdf) & LOBE_SMOOTH);
            DS ds; V3 em_weight;
            em.sample_emitter_direction(si, smp.next_2d(), active_em, ds, em_weight, c);
            active_em = active_em && ds.pdf != 0.f;
            V3 wo = si.valid ? si.sh.to_local(ds.d) : V3(0.f);
            float s1 = smp.next_1d(); V2 s2 = smp.next_2d();
            V3 bsdf_val(0.f), bsdf_weight(0.f); float bsdf_pdf = 0.f; BS bs;
            // (bsdf_val / bsdf_pdf are only consumed under active_em, so they are only evaluated there)
            if (bsdf) { if (active_em) bsdf_eval_pdf(sc, *bsdf, si, wo, bsdf_val, bsdf_pdf, false, c); bsdf_sample(sc, *bsdf, si, s1, s2, bs, bsdf_weight, c); }
            float mis_em = ds.delta ? 1.f : mis_weight_isfinite(ds.pdf, bsdf_pdf);
            if (active_em) L += f * bsdf_val * em_weight * mis_em;
            f *= bsdf_weight; eta *= bs.eta;
            float fmax = hmax(f), rr_prob = fminf(fmax * sqr(eta), 0.95f);
            bool rr_active = depth >= (uint32_t) rr_depth, rr_continue = smp.next_1d() < rr_prob;
            if (rr_active) f *= rcp(rr_prob);
            active = active && fmax != 0.f;
            active = active && (!rr_active || rr_continue);
            SI si2;
            if (active) { ray = Scene::spawn_ray(si.p, si.n, si.sh.to_world(bs.wo)); si2 = sc.make_si(ray, sc.intersect(ray, c), c); }
            else si2.wi = -ray.d;
            bool bsdf_delta = (bs.type & LOBE_DELTA) != 0;
            DS ds2 = em.direction_sample(si2, si.p);
            float em_pdf = em.pdf_emitter_direction(ds2, !bsdf_delta);
            float mis_bsdf = mis_weight_isfinite(bs.pdf, em_pdf);
            if (active) { V3 Le = bs.pdf > 0.f ? em.eval(ds2.emitter, si2, c) : V3(0.f); L += f * Le * mis_bsdf; }
            si = si2;
            if (active) depth += 1;
            active = active && depth < (uint32_t) max_depth;
            active = active && si.valid;
        }
        return L;
    }

    // ------------------------------------------------------------------ path-mis.py:24-155
    V3 sample_path_mis(Sampler &smp, Ray ray, Counters &c) const {
        bool active = true; V3 throughput(1.f), result(0.f); float eta = 1.f; uint32_t depth = 0;
        bool valid_ray = sc.env_emitter >= 0;
        V3 prev_p(0.f); float prev_bsdf_pdf = 1.f; bool prev_bsdf_delta = true;
        for (int it = 0; it < max_depth && active; ++it) {
            SI si = sc.make_si(ray, sc.intersect(ray, c), c);
            DS ds = em.direction_sample(si, prev_p);
            float em_pdf = em.pdf_emitter_direction(ds, !prev_bsdf_delta);
            float mis_bsdf = mis_weight_guard(prev_bsdf_pdf, em_pdf);
            V3 Le = prev_bsdf_pdf > 0.f ? em.eval(ds.emitter, si, c) : V3(0.f);
            result = fmadd(throughput, Le * mis_bsdf, result);
            bool active_next = ((depth + 1) < (uint32_t) max_depth) && si.valid;
            const b2_bsdf *bsdf = bsdf_of(si);
            bool active_em = active_next && bsdf && (bsdf_lobes(*bsdf) & LOBE_SMOOTH);
            V3 em_weight;
            em.sample_emitter_direction(si, smp.next_2d(), active_em, ds, em_weight, c);
            V3 wo = si.valid ? si.sh.to_local(ds.d) : V3(0.f);
            float sample1 = smp.next_1d(); V2 sample2 = smp.next_2d();
            V3 bsdf_val(0.f), bsdf_weight(0.f); float bsdf_pdf = 0.f; BS bs;
            // (eval_pdf is consumed only under active_em and the sample only when the loop continues: evaluated only there)
            if (bsdf) { if (active_em) bsdf_eval_pdf(sc, *bsdf, si, wo, bsdf_val, bsdf_pdf, false, c); if (active_next) bsdf_sample(sc, *bsdf, si, sample1, sample2, bs, bsdf_weight, c); }
            // terminating iteration: the reference still samples the BSDF; only its sampled_type is consumed (valid_ray below),
            // and only an opacity mask can return a null lobe (mask.cpp:123-160: sample1 >= opacity)
            if (bsdf && !active_next && (bsdf->flags & B2_BSDF_FLAG_MASK) && !(sample1 < bsdf->opacity)) bs.type = LOBE_NULL;
            float mi_em = ds.delta ? 1.f : mis_weight_guard(ds.pdf, bsdf_pdf);
            if (active_em) result = fmadd(throughput, bsdf_val * em_weight * mi_em, result);
            if (si.valid) ray = Scene::spawn_ray(si.p, si.n, si.sh.to_world(bs.wo));
            throughput *= bsdf_weight; eta *= bs.eta;
            valid_ray = valid_ray || (active && si.valid && !(bs.type & LOBE_NULL));   // path-mis.py:129-133
            prev_p = si.p; prev_bsdf_pdf = bs.pdf; prev_bsdf_delta = (bs.type & LOBE_DELTA) != 0;
            if (si.valid) depth += 1;
            float tmax = hmax(throughput), rr_prop = fminf(tmax * sqr(eta), 0.95f);
            bool rr_active = depth >= (uint32_t) rr_depth, rr_continue = smp.next_1d() < rr_prop;
            if (rr_active) throughput *= rcp(rr_prop);
            active = active_next && (!rr_active || rr_continue) && (tmax != 0.f);
        }
        return valid_ray ? result : V3(0.f);
    }

    // ------------------------------------------------------------------ prb.py:63-265
    // primal: returns L.  backward: state_in = L of the primal pass, dL = adjoint radiance;
    // scatter-adds into the gradient targets through `grad_cb(bsdf, si, g)`.
    template <typename GradCB>
    V3 sample_prb(bool primal, Sampler &smp, Ray ray, V3 dL, V3 state_in, GradCB grad_cb, Counters &c) const {
        uint32_t depth = 0; V3 L = primal ? V3(0.f) : state_in; V3 beta(1.f); float eta = 1.f; bool active = true;
        Hit pi = sc.intersect(ray, c);
        V3 prev_p(0.f); float prev_bsdf_pdf = 1.f; bool prev_bsdf_delta = true;
        for (int it = 0; it < max_depth && active; ++it) {
            bool active_next = active;
            SI si = sc.make_si(ray, pi, c);
            const b2_bsdf *bsdf = bsdf_of(si);
            DS ds = em.direction_sample(si, prev_p);
            float mis = mis_weight_isfinite(prev_bsdf_pdf, em.pdf_emitter_direction(ds, !prev_bsdf_delta));
            V3 Le = beta * mis * em.eval(ds.emitter, si, c);
            active_next = active_next && (depth + 1 < (uint32_t) max_depth) && si.valid;
            bool active_em = active_next && bsdf && (bsdf_lobes(*bsdf) & LOBE_SMOOTH);
            V3 em_weight;
            em.sample_emitter_direction(si, smp.next_2d(), active_em, ds, em_weight, c);
            active_em = active_em && ds.pdf != 0.f;
            V3 wo = si.valid ? si.sh.to_local(ds.d) : V3(0.f);
            V3 bsdf_value_em(0.f); float bsdf_pdf_em = 0.f;
            if (active_em) bsdf_eval_pdf(sc, *bsdf, si, wo, bsdf_value_em, bsdf_pdf_em, false, c);
            float mis_em = ds.delta ? 1.f : mis_weight_isfinite(ds.pdf, bsdf_pdf_em);
            V3 Lr_dir = beta * mis_em * bsdf_value_em * em_weight;
            float s1 = smp.next_1d(); V2 s2 = smp.next_2d();
            BS bs; V3 bsdf_weight(0.f);
            if (active_next) bsdf_sample(sc, *bsdf, si, s1, s2, bs, bsdf_weight, c);
            V3 beta_before = beta;
            L = primal ? (L + Le + Lr_dir) : (L - Le - Lr_dir);
            if (si.valid) ray = Scene::spawn_ray(si.p, si.n, si.sh.to_world(bs.wo));
            eta *= bs.eta; beta *= bsdf_weight;
            prev_p = si.p; prev_bsdf_pdf = bs.pdf; prev_bsdf_delta = (bs.type & LOBE_DELTA) != 0;
            float beta_max = hmax(beta);
            active_next = active_next && (beta_max != 0.f);
            float rr_prob = fminf(beta_max * sqr(eta), .95f);
            bool rr_active = depth >= (uint32_t) rr_depth;
            if (rr_active) beta *= rcp(rr_prob);
            bool rr_continue = smp.next_1d() < rr_prob;
            active_next = active_next && (!rr_active || rr_continue);
            if (!primal && si.valid) {
                // d/dtheta [ Lr_dir + L * bsdf_val / detach(bsdf_weight * pdf) ]  (prb.py:202-250); all supported
                // BSDFs are linear in their reflectance, so d bsdf_val / d rho_c = bsdf_val(rho := 1)_c
                V3 g(0.f);
                if (active_em) {
                    V3 unit; float dummy; bsdf_eval_pdf(sc, *bsdf, si, wo, unit, dummy, true, c);
                    g += dL * (beta_before * mis_em * unit * em_weight);
                }
                if (active_next) {
                    V3 wo2 = si.sh.to_local(ray.d);
                    V3 unit; float dummy; bsdf_eval_pdf(sc, *bsdf, si, wo2, unit, dummy, true, c);
                    V3 det = bsdf_weight * bs.pdf;
                    V3 inv(det.x != 0.f ? rcp(det.x) : 0.f, det.y != 0.f ? rcp(det.y) : 0.f, det.z != 0.f ? rcp(det.z) : 0.f);
                    g += dL * (L * (inv * unit));
                }
                if (g.x != 0.f || g.y != 0.f || g.z != 0.f) grad_cb(sc.d->meshes[si.mesh].bsdf, si, g);
            }
            if (si.valid) depth += 1;
            active = active_next;
            if (active) pi = sc.intersect(ray, c);
        }
        return L;
    }
};

struct Handle {
    b2_scene_desc desc; Scene scene;
};

static inline int sensor_border(const b2_sensor &s) {
    // ReconstructionFilter::border_size = ceil(radius - 1/2 - 2 RayEpsilon) (rfilter.cpp:25); the box filter (radius 1/2) -> 0
    return (s.sample_border && s.rfilter != B2_RFILTER_BOX) ? (int) std::ceil(s.rfilter_radius - .5f - 2.f * 1500.f * 5.9604644775390625e-8f) : 0;
}
static void derive_passes(const b2_scene_desc &d, const b2_render_args &a, uint32_t &spp_per_pass, uint32_t &n_passes, uint64_t &wavefront) {
    // integrator.cpp:139-152,247-265
    // film_size = crop_size + 2 * border_size when film.sample_border is set (integrator.cpp:134-136)
    const int b = sensor_border(d.sensor);
    uint64_t npix = (uint64_t) (d.sensor.crop_width + 2 * b) * (d.sensor.crop_height + 2 * b);
    spp_per_pass = a.spp_per_pass ? std::min(a.spp_per_pass, a.spp) : a.spp;
    wavefront = npix * spp_per_pass;
    const uint64_t limit = 0xffffffffull;
    if (wavefront > limit) {
        spp_per_pass /= (uint32_t) ((wavefront + limit - 1) / limit);
        wavefront = npix * spp_per_pass;
    }
    n_passes = a.spp / spp_per_pass;
}
static inline void lane_to_pos(const b2_sensor &s, uint64_t lane, uint32_t spp_per_pass, int &px, int &py) {
    // integrator.cpp:293-310
    uint32_t idx = (uint32_t) lane, lg = 0; while ((1u << (lg + 1)) <= spp_per_pass) lg++;
    if ((1u << lg) == spp_per_pass) idx >>= lg; else idx /= spp_per_pass;
    const int b = sensor_border(s); const uint32_t gw = (uint32_t) (s.crop_width + 2 * b);
    uint32_t y = idx / gw, x = idx - gw * y;
    px = (int) x - b + s.crop_offset_x; py = (int) y - b + s.crop_offset_y;      // :306-309 pos -= border_size; pos += crop_offset
}
static inline V2 adjusted_pos(const b2_sensor &s, V2 sample_pos) {
    // integrator.cpp:432-436: scale = 1/crop_size, offset = -crop_offset*scale, fmadd(sample_pos, scale, offset)
    float sx = 1.f / (float) s.crop_width, sy = 1.f / (float) s.crop_height;
    float ox = -(float) s.crop_offset_x * sx, oy = -(float) s.crop_offset_y * sy;
    return V2{ fmaf(sample_pos.x, sx, ox), fmaf(sample_pos.y, sy, oy) };
}
static void export_counters(const Counters &c, b2_counters *o) {
    if (!o) return;
    o->samples = c.samples; o->closest_rays = c.closest; o->shadow_rays = c.shadow; o->vertices = c.vertices;
    o->tex_evals = c.tex; o->nee_area = c.nee_area; o->nee_env = c.nee_env; o->env_escapes = c.escapes;
    o->splat_taps = c.taps; o->grad_atomics = c.grad_atomics; o->node_fetches = c.nodes; o->tri_fetches = c.tris; o->launches = 0;
}

} // namespace orc

using namespace orc;
static thread_local std::string g_err;

extern "C" {

const char *orc_last_error(void) { return g_err.c_str(); }

// The descriptor's arrays are BORROWED (must outlive the handle).
int orc_scene_create(const b2_scene_desc *desc, void **out) {
    if (!desc || desc->abi_version != B2_ABI_VERSION) { g_err = "orc_scene_create: bad descriptor / ABI version"; return -1; }
    Handle *h = new Handle{ *desc, Scene() };
    h->scene.init(&h->desc);
    *out = h; return 0;
}
void orc_scene_destroy(void *h) { delete (Handle *) h; }

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void orc_set_num_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void) n;
#endif
}

int orc_render(void *hp, const b2_render_args *a, float *film_rgbw, b2_counters *counters) {
    Handle *h = (Handle *) hp; const b2_scene_desc &d = h->desc; const b2_sensor &s = d.sensor;
    uint32_t spp_per_pass, n_passes; uint64_t wavefront; derive_passes(d, *a, spp_per_pass, n_passes, wavefront);
    if (spp_per_pass == 0 || a->spp % spp_per_pass) { g_err = "sample_count must be a multiple of spp_per_pass"; return -2; }
    if (a->sample_count) { g_err = "the oracle renders whole pixels only (sample shards are a product-side split)"; return -2; }
    uint64_t lb = a->lane_begin, le = a->lane_end ? a->lane_end : wavefront;
    Film film(s); Integrator integ(h->scene, a->max_depth, a->rr_depth);
    integ.em.env_warp = a->env_warp;
    Counters total;
    // thread-private double films (merged once) instead of one atomic per tap and channel, unless they would not fit in 4 GB
    const bool priv_films = film.acc.size() * sizeof(double) * (size_t) orc_num_threads() <= ((size_t) 4 << 30);
#pragma omp parallel
    {
        Counters c;
        std::vector<double> priv(priv_films ? film.acc.size() : 0, 0.0);
#pragma omp for schedule(dynamic, 1024)
        for (int64_t lane = (int64_t) lb; lane < (int64_t) le; ++lane) {
            Sampler smp; smp.rng = sampler_seed(a->seed, (uint32_t) lane);
            int px, py; lane_to_pos(s, (uint64_t) lane, spp_per_pass, px, py);
            for (uint32_t pass = 0; pass < n_passes; ++pass) {
                V2 j = smp.next_2d();
                V2 sample_pos{ (float) px + j.x, (float) py + j.y };
                Ray ray = camera_ray(s, adjusted_pos(s, sample_pos));
                V3 L = a->variant == B2_VARIANT_PATH ? integ.sample_path_py(smp, ray, c)
                     : a->variant == B2_VARIANT_PATH_MIS ? integ.sample_path_mis(smp, ray, c)
                     : integ.sample_prb(true, smp, ray, V3(0.f), V3(0.f), [](int, const SI &, V3) {}, c);
                V2 splat = s.rfilter == B2_RFILTER_BOX ? V2{ (float) px, (float) py } : sample_pos;
                if (priv_films) film.put_private(priv.data(), splat, L, c); else film.put(splat, L, c);
                c.samples++;
            }
        }
#pragma omp critical
        {
            total.add(c);
            if (priv_films) for (size_t i = 0; i < priv.size(); ++i) film.acc[i] += priv[i];
        }
    }
    for (size_t i = 0; i < film.acc.size(); ++i) film_rgbw[i] += (float) film.acc[i];
    export_counters(total, counters);
    return 0;
}

// debug: run ONE lane and record its closest-hit rays: out [max_records][11] = ray(7), t, u, v, face; returns #records; L -> rgb
int orc_debug_lane(void *hp, const b2_render_args *a, uint32_t lane, float *out, int max_records, float *rgb) {
    Handle *h = (Handle *) hp; const b2_scene_desc &d = h->desc; const b2_sensor &s = d.sensor;
    uint32_t spp_per_pass, n_passes; uint64_t wavefront; derive_passes(d, *a, spp_per_pass, n_passes, wavefront);
    Integrator integ(h->scene, a->max_depth, a->rr_depth); Counters c;
    TraceLog log{ out, max_records, 0 }; g_trace_log = &log;
    Sampler smp; smp.rng = sampler_seed(a->seed, lane);
    int px, py; lane_to_pos(s, lane, spp_per_pass, px, py);
    V2 j = smp.next_2d(); V2 sample_pos{ (float) px + j.x, (float) py + j.y };
    Ray ray = camera_ray(s, adjusted_pos(s, sample_pos));
    V3 L = a->variant == B2_VARIANT_PATH ? integ.sample_path_py(smp, ray, c)
         : a->variant == B2_VARIANT_PATH_MIS ? integ.sample_path_mis(smp, ray, c)
         : integ.sample_prb(true, smp, ray, V3(0.f), V3(0.f), [](int, const SI &, V3) {}, c);
    g_trace_log = nullptr;
    rgb[0] = L.x; rgb[1] = L.y; rgb[2] = L.z;
    return log.n;
}

// HDRFilm::develop (hdrfilm.cpp:304-405): rgb / (W == 0 ? 1 : W)
int orc_develop(void *hp, const float *film_rgbw, float *image_rgb) {
    Handle *h = (Handle *) hp; size_t n = (size_t) h->desc.sensor.crop_width * h->desc.sensor.crop_height;
    for (size_t i = 0; i < n; ++i) {
        float w = film_rgbw[4 * i + 3]; w = w == 0.f ? 1.f : w;
        for (int k = 0; k < 3; ++k) image_rgb[3 * i + k] = film_rgbw[4 * i + k] / w;
    }
    return 0;
}

// RBIntegrator.render_backward (common.py:623-781).  grad_in [h][w][3].  Gradients are ACCUMULATED into targets.
int orc_render_backward(void *hp, const b2_render_args *a, const float *grad_in, const b2_grad_target *targets, int n_targets, b2_counters *counters) {
    Handle *h = (Handle *) hp; const b2_scene_desc &d = h->desc; const b2_sensor &s = d.sensor;
    uint32_t spp_per_pass, n_passes; uint64_t wavefront; derive_passes(d, *a, spp_per_pass, n_passes, wavefront);
    if (n_passes != 1) { g_err = "render_backward: wavefront exceeds 2^32 (common.py:358-363)"; return -2; }
    uint64_t lb = a->lane_begin, le = a->lane_end ? a->lane_end : wavefront;
    Integrator integ(h->scene, a->max_depth, a->rr_depth);
    // (1) weight film of THIS sample set (common.py:728-744: splat L == 1, weight 1)
    Film wfilm(s);
    {
        Counters c;
#pragma omp parallel for schedule(static) firstprivate(c)
        for (int64_t lane = 0; lane < (int64_t) wavefront; ++lane) {
            Sampler smp; smp.rng = sampler_seed(a->seed, (uint32_t) lane);
            int px, py; lane_to_pos(s, (uint64_t) lane, spp_per_pass, px, py);
            V2 j = smp.next_2d();
            V2 pos = s.rfilter == B2_RFILTER_BOX ? V2{ (float) px, (float) py } : V2{ (float) px + j.x, (float) py + j.y };
            wfilm.put(pos, V3(0.f), c);
        }
    }
    std::vector<std::vector<double>> gacc(n_targets);
    for (int t = 0; t < n_targets; ++t) {
        const b2_bsdf &b = d.bsdfs[targets[t].bsdf];
        size_t n = targets[t].kind == 0 ? 3 : (size_t) d.textures[b.reflectance_tex].width * d.textures[b.reflectance_tex].height * 3;
        gacc[t].assign(n, 0.0);
    }
    Counters total;
#pragma omp parallel
    {
        Counters c;
#pragma omp for schedule(dynamic, 1024)
        for (int64_t lane = (int64_t) lb; lane < (int64_t) le; ++lane) {
            Sampler smp; smp.rng = sampler_seed(a->seed, (uint32_t) lane);
            int px, py; lane_to_pos(s, (uint64_t) lane, spp_per_pass, px, py);
            V2 j = smp.next_2d();
            V2 sample_pos{ (float) px + j.x, (float) py + j.y };
            Ray ray = camera_ray(s, adjusted_pos(s, sample_pos));
            V2 splat = s.rfilter == B2_RFILTER_BOX ? V2{ (float) px, (float) py } : sample_pos;
            // (2) dL_k = sum_p grad_in[p] * w_k(p) / W[p]
            V3 dL(0.f); Counters dummy;
            wfilm.taps(splat, [&](size_t pix, float wgt) {
                float W = (float) wfilm.acc[pix * 4 + 3]; W = W == 0.f ? 1.f : W;
                float k = wgt / W;
                dL += V3(grad_in[3 * pix] * k, grad_in[3 * pix + 1] * k, grad_in[3 * pix + 2] * k);
            }, dummy);
            // (3) primal pass on a clone of the sampler, (4) adjoint pass on the sampler itself
            Sampler clone = smp;
            V3 L = integ.sample_prb(true, clone, ray, V3(0.f), V3(0.f), [](int, const SI &, V3) {}, c);
            integ.sample_prb(false, smp, ray, dL, L, [&](int bsdf, const SI &si, V3 g) {
                for (int t = 0; t < n_targets; ++t) {
                    if (targets[t].bsdf != bsdf) continue;
                    if (targets[t].kind == 0) {
                        for (int k = 0; k < 3; ++k) { Film::atomic_add(gacc[t][k], (double) g[k]); c.grad_atomics++; }
                    } else {
                        uint32_t idx[4]; float w[4]; Counters cc;
                        h->scene.tex_eval(d.bsdfs[bsdf].reflectance_tex, si.uv, idx, w, cc);
                        for (int q = 0; q < 4; ++q) for (int k = 0; k < 3; ++k) { Film::atomic_add(gacc[t][(size_t) idx[q] * 3 + k], (double) (g[k] * w[q])); c.grad_atomics++; }
                    }
                }
            }, c);
            c.samples++;
        }
#pragma omp critical
        total.add(c);
    }
    for (int t = 0; t < n_targets; ++t) for (size_t i = 0; i < gacc[t].size(); ++i) targets[t].grad[i] += (float) gacc[t][i];
    export_counters(total, counters);
    return 0;
}

// Forward-mode PRB (RBIntegrator.render_forward, common.py:495-621; prb.py:249-250 dr.forward_to(Lo)): the tangent image of the
// render w.r.t. the reflectance parameters, dI = develop(splat(dL_lane)), dL_lane = sum over vertices of
// (d Lo / d reflectance) . tangent(vertex) -- the same per-vertex coefficient the adjoint scatters, gathered instead.
// tangents[t].grad = the tangent of target t (3 floats, or [h][w][3] for a bitmap).  film_rgbw is accumulated into.
int orc_render_forward(void *hp, const b2_render_args *a, const b2_grad_target *tangents, int n_targets, float *film_rgbw, b2_counters *counters) {
    Handle *h = (Handle *) hp; const b2_scene_desc &d = h->desc; const b2_sensor &s = d.sensor;
    uint32_t spp_per_pass, n_passes; uint64_t wavefront; derive_passes(d, *a, spp_per_pass, n_passes, wavefront);
    if (n_passes != 1) { g_err = "render_forward: wavefront exceeds 2^32 (common.py:358-363)"; return -2; }
    if (a->sample_count) { g_err = "the oracle renders whole pixels only"; return -2; }
    uint64_t lb = a->lane_begin, le = a->lane_end ? a->lane_end : wavefront;
    Film film(s); Integrator integ(h->scene, a->max_depth, a->rr_depth);
    Counters total;
#pragma omp parallel
    {
        Counters c;
#pragma omp for schedule(dynamic, 1024)
        for (int64_t lane = (int64_t) lb; lane < (int64_t) le; ++lane) {
            Sampler smp; smp.rng = sampler_seed(a->seed, (uint32_t) lane);
            int px, py; lane_to_pos(s, (uint64_t) lane, spp_per_pass, px, py);
            V2 j = smp.next_2d();
            V2 sample_pos{ (float) px + j.x, (float) py + j.y };
            Ray ray = camera_ray(s, adjusted_pos(s, sample_pos));
            V2 splat = s.rfilter == B2_RFILTER_BOX ? V2{ (float) px, (float) py } : sample_pos;
            Sampler clone = smp;
            V3 L = integ.sample_prb(true, clone, ray, V3(0.f), V3(0.f), [](int, const SI &, V3) {}, c);
            V3 fwd(0.f);
            integ.sample_prb(false, smp, ray, V3(1.f), L, [&](int bsdf, const SI &si, V3 g) {
                for (int t = 0; t < n_targets; ++t) {
                    if (tangents[t].bsdf != bsdf) continue;
                    const float *tp = tangents[t].grad;
                    if (tangents[t].kind == 0) fwd += g * V3(tp[0], tp[1], tp[2]);
                    else {
                        uint32_t idx[4]; float w[4]; Counters cc;
                        h->scene.tex_eval(d.bsdfs[bsdf].reflectance_tex, si.uv, idx, w, cc);
                        V3 tan(0.f);
                        for (int q = 0; q < 4; ++q) tan += V3(tp[3 * (size_t) idx[q]], tp[3 * (size_t) idx[q] + 1], tp[3 * (size_t) idx[q] + 2]) * w[q];
                        fwd += g * tan;
                    }
                }
            }, c);
            film.put(splat, fwd, c); c.samples++;
        }
#pragma omp critical
        total.add(c);
    }
    for (size_t i = 0; i < film.acc.size(); ++i) film_rgbw[i] += (float) film.acc[i];
    export_counters(total, counters);
    return 0;
}

// rays [n][8] = o, d, maxt, pad
int orc_trace(void *hp, const float *rays, uint64_t n, int any_hit, b2_hit *hits) {
    Handle *h = (Handle *) hp;
#pragma omp parallel for schedule(dynamic, 4096)
    for (int64_t i = 0; i < (int64_t) n; ++i) {
        const float *r = rays + 8 * i; Counters c;
        Ray ray{ V3(r[0], r[1], r[2]), V3(r[3], r[4], r[5]), r[6] };
        if (any_hit) { bool occ = h->scene.occluded(ray, c); hits[i] = b2_hit{ occ ? 0.f : ray.maxt, 0, 0, occ ? 0u : 0xffffffffu }; }
        else { Hit hh = h->scene.intersect(ray, c); hits[i] = b2_hit{ hh.t, hh.u, hh.v, hh.face }; }
    }
    return 0;
}

// ------------------------------------------------------------------ unit-level entry points (tests)
void orc_tea32(uint32_t v0, uint32_t v1, int rounds, uint32_t *out) { sample_tea_32(v0, v1, out[0], out[1], rounds); }
void orc_pcg32_floats(uint64_t initstate, uint64_t initseq, int n, float *out) { PCG32 r; r.seed(initstate, initseq); for (int i = 0; i < n; ++i) out[i] = r.next_float32(); }
void orc_pcg32_uints(uint64_t initstate, uint64_t initseq, int n, uint32_t *out) { PCG32 r; r.seed(initstate, initseq); for (int i = 0; i < n; ++i) out[i] = r.next_uint32(); }
void orc_sampler_floats(uint32_t seed, uint32_t lane, int n, float *out) { PCG32 r = sampler_seed(seed, lane); for (int i = 0; i < n; ++i) out[i] = r.next_float32(); }
float orc_rfilter_eval(const b2_sensor *s, float x) { return rfilter_eval(*s, x); }
void orc_camera_ray(const b2_sensor *s, float x, float y, float *out7) {
    Ray r = camera_ray(*s, adjusted_pos(*s, V2{ x, y }));
    out7[0] = r.o.x; out7[1] = r.o.y; out7[2] = r.o.z; out7[3] = r.d.x; out7[4] = r.d.y; out7[5] = r.d.z; out7[6] = r.maxt;
}
// which: 0 disk_concentric, 1 cosine_hemisphere, 2 uniform_triangle, 3 interval_to_linear(v0,v1,s), 4 erfinv, 5 coordinate_system
void orc_warp(int which, const float *in, float *out) {
    if (which == 0) { V2 r = square_to_uniform_disk_concentric(V2{ in[0], in[1] }); out[0] = r.x; out[1] = r.y; }
    else if (which == 1) { V3 r = square_to_cosine_hemisphere(V2{ in[0], in[1] }); out[0] = r.x; out[1] = r.y; out[2] = r.z; }
    else if (which == 2) { V2 r = square_to_uniform_triangle(V2{ in[0], in[1] }); out[0] = r.x; out[1] = r.y; }
    else if (which == 3) out[0] = interval_to_linear(in[0], in[1], in[2]);
    else if (which == 4) out[0] = erfinv(in[0]);
    else if (which == 5) { V3 s, t; coordinate_system(V3(in[0], in[1], in[2]), s, t); out[0] = s.x; out[1] = s.y; out[2] = s.z; out[3] = t.x; out[4] = t.y; out[5] = t.z; }
}
void orc_fresnel(float cos_i, float eta, float *out4) { fresnel(cos_i, eta, out4[0], out4[1], out4[2], out4[3]); }
float orc_fresnel_conductor(float cos_i, float eta, float k) { return fresnel_conductor(cos_i, eta, k); }
// microfacet: out = m.xyz, pdf | eval D | smith_g1
void orc_microfacet(int type, float au, float av, int visible, int op, const float *in, float *out) {
    Microfacet d(type, au, av, visible != 0);
    if (op == 0) { float pdf; V3 m = d.sample(V3(in[0], in[1], in[2]), V2{ in[3], in[4] }, pdf); out[0] = m.x; out[1] = m.y; out[2] = m.z; out[3] = pdf; }
    else if (op == 1) out[0] = d.eval(V3(in[0], in[1], in[2]));
    else if (op == 2) out[0] = d.smith_g1(V3(in[0], in[1], in[2]), V3(in[3], in[4], in[5]));
    else if (op == 3) out[0] = d.pdf(V3(in[0], in[1], in[2]), V3(in[3], in[4], in[5]));
}
void orc_rough_tables(int type, float alpha, float eta, float *trans64, float *internal_reflectance) { rough_tables(type, alpha, eta, trans64, *internal_reflectance); }
// BSDF in local frame with constant reflectance. sample: out = wo3, pdf, eta, type, weight3 ; eval_pdf: out = val3, pdf
void orc_bsdf_sample(const b2_bsdf *b, const float *wi, float s1, const float *s2, float *out9) {
    static b2_scene_desc d{}; static Scene sc; sc.d = &d;      // static: keeps the rough-plastic table cache between calls (single-threaded test hook)
    SI si; si.valid = true; si.wi = V3(wi[0], wi[1], wi[2]); Counters c;
    BS bs; V3 w; bsdf_sample(sc, *b, si, s1, V2{ s2[0], s2[1] }, bs, w, c);
    out9[0] = bs.wo.x; out9[1] = bs.wo.y; out9[2] = bs.wo.z; out9[3] = bs.pdf; out9[4] = bs.eta; out9[5] = (float) bs.type; out9[6] = w.x; out9[7] = w.y; out9[8] = w.z;
}
void orc_bsdf_eval_pdf(const b2_bsdf *b, const float *wi, const float *wo, float *out4) {
    static b2_scene_desc d{}; static Scene sc; sc.d = &d;
    SI si; si.valid = true; si.wi = V3(wi[0], wi[1], wi[2]); Counters c;
    V3 v; float pdf; bsdf_eval_pdf(sc, *b, si, V3(wo[0], wo[1], wo[2]), v, pdf, false, c);
    out4[0] = v.x; out4[1] = v.y; out4[2] = v.z; out4[3] = pdf;
}
void orc_texture_eval(const b2_texture *t, float u, float v, float *out3, uint32_t *idx4, float *w4) {
    b2_scene_desc d{}; d.textures = t; d.n_textures = 1; Scene sc; sc.d = &d; Counters c;
    V3 r = sc.tex_eval(0, V2{ u, v }, idx4, w4, c); out3[0] = r.x; out3[1] = r.y; out3[2] = r.z;
}
void orc_env_sample(const b2_envmap *e, float u, float v, float *out3) { float pdf; V2 r = warp_sample(*e, V2{ u, v }, pdf); out3[0] = r.x; out3[1] = r.y; out3[2] = pdf; }
float orc_env_warp_eval(const b2_envmap *e, float u, float v) { return warp_eval(*e, V2{ u, v }); }
void orc_env_alias_sample(const b2_envmap *e, float u, float v, float *out3) { float pdf; V2 r = alias_sample(*e, V2{ u, v }, pdf); out3[0] = r.x; out3[1] = r.y; out3[2] = pdf; }
void orc_env_eval_dir(const b2_envmap *e, const float *dir, float *out3) { V3 r = env_eval_dir(*e, V3(dir[0], dir[1], dir[2])); out3[0] = r.x; out3[1] = r.y; out3[2] = r.z; }
// surface interaction of one ray against the scene (a5/a7 pins): out = valid,t,p3,n3,uv2,dp_du3,dp_dv3,sh_n3,wi3, prim_uv2, face
void orc_surface_interaction(void *hp, const float *ray7, float *out) {
    Handle *h = (Handle *) hp; Counters c; Ray r{ V3(ray7[0], ray7[1], ray7[2]), V3(ray7[3], ray7[4], ray7[5]), ray7[6] };
    Hit hit = h->scene.intersect(r, c); SI si = h->scene.make_si(r, hit, c);
    float v[] = { si.valid ? 1.f : 0.f, si.t, si.p.x, si.p.y, si.p.z, si.n.x, si.n.y, si.n.z, si.uv.x, si.uv.y, si.dp_du.x, si.dp_du.y, si.dp_du.z,
                  si.dp_dv.x, si.dp_dv.y, si.dp_dv.z, si.sh.n.x, si.sh.n.y, si.sh.n.z, si.wi.x, si.wi.y, si.wi.z, hit.u, hit.v, (float) hit.face };
    std::memcpy(out, v, sizeof(v));
}
// film splat of explicit samples (imageblock test): pos [n][2], val [n][3] -> film [h][w][4] accumulate
void orc_splat(const b2_sensor *s, const float *pos, const float *val, uint64_t n, float *film_rgbw) {
    Film f(*s); Counters c;
    for (uint64_t i = 0; i < n; ++i) f.put(V2{ pos[2 * i], pos[2 * i + 1] }, V3(val[3 * i], val[3 * i + 1], val[3 * i + 2]), c);
    for (size_t i = 0; i < f.acc.size(); ++i) film_rgbw[i] += (float) f.acc[i];
}

} // extern "C"
