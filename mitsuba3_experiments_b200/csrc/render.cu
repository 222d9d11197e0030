// render.cu -- the path-tracing kernels.
//
// One persistent kernel per (loop formulation, mode):
//   * loop formulations  path.py:194-302 (B2_VARIANT_PATH), path-mis.py:24-155 == integrators/path.cpp:94-327
//     (B2_VARIANT_PATH_MIS), ad/integrators/prb.py:63-265 (B2_VARIANT_PRB, primal and adjoint)
//   * driver             SamplingIntegrator::render / render_sample (src/render/integrator.cpp:246-360,419-491),
//                        RBIntegrator.render_backward (ad/integrators/common.py:623-781)
//   * film               ImageBlock::put (src/render/imageblock.cpp:174-560), rfilters (gaussian.cpp:94-103,
//                        tent.cpp:53-55, box.cpp:41-43), HDRFilm::develop (films/hdrfilm.cpp:304-405)
//
// Execution model (B200): a grid of resident CTAs pulls *units* (a run of pixels of one image row x all of its
// samples) from a global counter.  Inside a unit, threads pull lanes from a shared counter and advance their paths
// one bounce per iteration of a flattened loop; a thread whose path ends immediately regenerates the next lane, so
// warps stay full while path lengths diverge.  Radiance is splatted into a shared-memory film tile
// ((2n+2) rows x (pixels+2n+1) columns of RGBW) and flushed once per unit with 128-bit vector reductions
// (red.global.add.v4.f32), which cuts global atomics from taps/sample to ~taps/pixel/unit.
// Each lane's PCG32 stream is a pure function of (seed, lane): no sampler state ever touches HBM, and the PRB
// adjoint regenerates exactly the same stream by reseeding.
#include "b2_render.cuh"
#include "b2_shade.cuh"

namespace b2 {

// ------------------------------------------------------------------------------------------ filters / film
B2_DEV float rfilter_eval(const DSensor &s, float x) {
    if (s.rfilter == B2_RFILTER_GAUSSIAN) return fmaxf(estrin10(sqr(x), s.gaussian_coeff), 0.f);
    if (s.rfilter == B2_RFILTER_TENT) return fmaxf(0.f, 1.f - fabsf(x * s.inv_radius));
    return (x >= -.5f && x < .5f) ? 1.f : 0.f;
}

// Footprint of ImageBlock::put (coalesced variant imageblock.cpp:433-508; box fast path :210-233).
// Calls f(x, y, weight) with crop-relative pixel coordinates for every in-bounds tap.
template <class F>
B2_DEV uint32_t film_taps(const RenderParams &p, f2 pos, F &&f) {
    const DSensor &s = p.sc.sensor;
    uint32_t taps = 0;
    if (s.rfilter == B2_RFILTER_BOX) {
        const int px = (int) floorf(pos.x) - s.crop_offset_x, py = (int) floorf(pos.y) - s.crop_offset_y;
        if ((unsigned) px < (unsigned) s.crop_width && (unsigned) py < (unsigned) s.crop_height) { f(px, py, 1.f); taps = 1; }
        return taps;
    }
    const int n = p.filter_n, count = 2 * n + 1;
    const int pix = (int) floorf(pos.x) - n, piy = (int) floorf(pos.y) - n;
    const int lx = pix - s.crop_offset_x, ly = piy - s.crop_offset_y;
    float relx = (float) pix + .5f - pos.x, rely = (float) piy + .5f - pos.y;
    float wx[5], wy[5];
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        if (i < count) { wx[i] = rfilter_eval(s, relx); wy[i] = rfilter_eval(s, rely); relx += 1.f; rely += 1.f; }
        else { wx[i] = 0.f; wy[i] = 0.f; }
    }
#pragma unroll
    for (int ys = 0; ys < 5; ++ys) {
        if (ys >= count) break;
        const unsigned y = (unsigned) (ly + ys); if (y >= (unsigned) s.crop_height) continue;
#pragma unroll
        for (int xs = 0; xs < 5; ++xs) {
            if (xs >= count) break;
            const unsigned x = (unsigned) (lx + xs); if (x >= (unsigned) s.crop_width) continue;
            taps++; f((int) x, (int) y, wy[ys] * wx[xs]);
        }
    }
    return taps;
}

// ------------------------------------------------------------------------------------------ sensor
// PerspectiveCamera::sample_ray_differential (sensors/perspective.cpp:239-280)
B2_DEV Ray camera_ray(const DSensor &s, f2 position_sample) {
    const f3 near_p = xf_point_persp(s.sample_to_camera, f3(position_sample.x, position_sample.y, 0.f));
    const f3 dl = normalize(near_p);
    Ray r;
    r.o = f3(s.to_world[3], s.to_world[7], s.to_world[11]);
    r.d = xf_vector(s.to_world, dl);
    const float inv_z = rcp(dl.z), near_t = s.near_clip * inv_z, far_t = s.far_clip * inv_z;
    r.o = r.o + r.d * near_t;
    r.maxt = far_t - near_t;
    return r;
}

B2_DEV float mis_isfinite(float a, float b) {      // path.py:10-18, common.py:1338-1346
    const float a2 = sqr(a), b2 = sqr(b), w = a2 / (a2 + b2);
    return finite_(w) ? w : 0.f;
}
B2_DEV float mis_guard(float a, float b) {         // path-mis.py:9-15 + isfinite (path.cpp:340-345), see DESIGN.md
    const float a2 = sqr(a), w = a2 / fma_(b, b, a2);
    return (a > 0.f && finite_(w)) ? w : 0.f;
}

// ------------------------------------------------------------------------------------------ per-lane path state
struct Path {
    Pcg32 rng;
    Ray ray;
    f3 f, L;                 // throughput (f / throughput / beta) and radiance (L / result)
    float eta;
    uint32_t depth;
    f3 prev_p; float prev_pdf;
    bool prev_delta, first, valid_ray, adjoint;
    f2 pos;                  // splat position (sample_pos, or the pixel for the box filter: integrator.cpp:490)
    f3 dL;                   // adjoint radiance (backward mode)
    uint32_t lane, pass;
};

template <int V>
B2_DEV void start_path(const RenderParams &p, Path &s) {
    const DSensor &sn = p.sc.sensor;
    // lane -> pixel (integrator.cpp:293-310)
    uint32_t idx = s.lane;
    idx = (p.spp_shift != 0xffffffffu) ? (idx >> p.spp_shift) : (idx / p.spp_per_pass);
    const uint32_t y = idx / (uint32_t) sn.crop_width, x = idx - (uint32_t) sn.crop_width * y;
    const int px = (int) x + sn.crop_offset_x, py = (int) y + sn.crop_offset_y;
    const f2 j = s.rng.next_2d();
    const f2 sample_pos{ (float) px + j.x, (float) py + j.y };
    const f2 adjusted{ fma_(sample_pos.x, sn.scale_x, sn.offset_x), fma_(sample_pos.y, sn.scale_y, sn.offset_y) };
    s.ray = camera_ray(sn, adjusted);
    s.pos = sn.rfilter == B2_RFILTER_BOX ? f2{ (float) px, (float) py } : sample_pos;
    s.f = f3(1.f); s.L = f3(0.f); s.eta = 1.f;
    s.depth = (V == B2_VARIANT_PATH) ? 1u : 0u;
    s.prev_p = f3(0.f); s.prev_pdf = 1.f; s.prev_delta = true; s.first = true;
    s.valid_ray = p.sc.env_emitter >= 0;
}

struct GradSink {               // shared-memory accumulators for constant-albedo targets
    float *s_const;             // [kMaxGradTargets][3]
};

// One bounce.  Returns true while the path continues.  RNG discipline: every entered loop iteration of the
// reference consumes exactly 6 draws (2 emitter, 1 + 2 BSDF, 1 roulette), also on its last iteration.
template <int V, bool ADJ, bool COUNT>
B2_DEV bool path_step(const RenderParams &p, Path &s, const GradSink &gs, Counters<COUNT> &cnt) {
    const DScene &sc = p.sc;
    const uint32_t max_depth = (uint32_t) p.max_depth;
    Hit h; h.face = 0xffffffffu; h.t = s.ray.maxt; h.u = h.v = 0.f; h.mesh = 0;
    const bool do_trace = !(V == B2_VARIANT_PATH && s.first && !(s.depth < max_depth));
    if (do_trace) { cnt.add(C_CLOSEST); bvh_trace<false>(sc, s.ray.o, s.ray.d, s.ray.maxt, h, cnt); }
    SI si; make_si(sc, s.ray, h, si, cnt);

    // ---- emission seen by the incoming ray, MIS-weighted against the previous bounce's emitter sampling
    f3 Le(0.f);
    bool active_next;
    if (V == B2_VARIANT_PATH) {
        if (s.first) {
            s.L += emitter_eval(sc, si.emitter, si, cnt);                                   // path.py:238-239
            active_next = s.depth < max_depth;
            if (active_next && !si.valid) { s.rng.skip(6); return false; }                  // body runs once with a null BSDF
        } else {
            DS ds2; direction_sample(si, s.prev_p, ds2);
            const float em_pdf = pdf_emitter_direction(sc, ds2, !s.prev_delta);
            const float mis_bsdf = mis_isfinite(s.prev_pdf, em_pdf);
            const f3 le = s.prev_pdf > 0.f ? emitter_eval(sc, ds2.emitter, si, cnt) : f3(0.f);
            s.L += s.f * le * mis_bsdf;                                                     // path.py:290-292
            s.depth += 1;
            active_next = s.depth < max_depth && si.valid;                                  // path.py:297-300
        }
        s.first = false;
        if (!active_next) return false;
    } else {
        DS ds2; direction_sample(si, s.prev_p, ds2);
        const float em_pdf = pdf_emitter_direction(sc, ds2, !s.prev_delta);
        if (V == B2_VARIANT_PATH_MIS) {
            const float mis_bsdf = mis_guard(s.prev_pdf, em_pdf);
            const f3 le = s.prev_pdf > 0.f ? emitter_eval(sc, ds2.emitter, si, cnt) : f3(0.f);
            s.L = fmadd(s.f, le * mis_bsdf, s.L);                                           // path-mis.py:75-86
        } else {
            const float mis = mis_isfinite(s.prev_pdf, em_pdf);
            Le = s.f * mis * emitter_eval(sc, ds2.emitter, si, cnt);                        // prb.py:113-134
        }
        active_next = (s.depth + 1 < max_depth) && si.valid;
        if (!active_next) {
            if (V == B2_VARIANT_PATH_MIS) { if (si.valid) s.valid_ray = true; }
            else s.L = ADJ ? (s.L - Le) : (s.L + Le);
            s.rng.skip(6);
            return false;
        }
        if (V == B2_VARIANT_PATH_MIS) s.valid_ray = true;
    }

    // ---- emitter sampling (NEE) + BSDF sampling.  si is valid from here on.
    const DBsdf &bsdf = sc.bsdfs[si.bsdf];
    ReflCache rc;
    auto refl = [&]() { return reflectance(sc, bsdf, si, rc, cnt); };
    auto ones = []() { return f3(1.f); };
    bool active_em = (bsdf_lobes(bsdf) & LOBE_SMOOTH) != 0;
    DS ds; f3 em_weight;
    sample_emitter_direction(sc, si, s.rng.next_2d(), active_em, p.env_warp, ds, em_weight, cnt);
    if (V != B2_VARIANT_PATH_MIS) active_em = active_em && ds.pdf != 0.f;
    const f3 wo = si.sh.to_local(ds.d);
    const float s1 = s.rng.next_1d(); const f2 s2 = s.rng.next_2d();
    f3 bsdf_val(0.f), bsdf_weight(0.f); float bsdf_pdf = 0.f; BS bs;
    if (active_em) bsdf_eval_pdf(bsdf, si.wi, wo, refl, bsdf_val, bsdf_pdf);
    bsdf_sample(bsdf, si.wi, s1, s2, refl, bs, bsdf_weight);
    const f3 f_before = s.f;
    float mis_em;
    if (V == B2_VARIANT_PATH) {
        mis_em = ds.delta ? 1.f : mis_isfinite(ds.pdf, bsdf_pdf);
        if (active_em) s.L += s.f * bsdf_val * em_weight * mis_em;                          // path.py:258-259
    } else if (V == B2_VARIANT_PATH_MIS) {
        mis_em = ds.delta ? 1.f : mis_guard(ds.pdf, bsdf_pdf);
        if (active_em) s.L = fmadd(s.f, bsdf_val * em_weight * mis_em, s.L);                // path-mis.py:113-117
    } else {
        mis_em = ds.delta ? 1.f : mis_isfinite(ds.pdf, bsdf_pdf);
        const f3 Lr_dir = s.f * mis_em * bsdf_val * em_weight;                              // prb.py:149-162
        s.L = ADJ ? (s.L - Le - Lr_dir) : (s.L + Le + Lr_dir);                              // prb.py:173
    }
    s.ray = spawn_ray(si.p, si.n, si.sh.to_world(bs.wo));
    s.f *= bsdf_weight; s.eta *= bs.eta;
    s.prev_p = si.p; s.prev_pdf = bs.pdf; s.prev_delta = (bs.type & LOBE_DELTA) != 0;

    // ---- Russian roulette
    bool active;
    if (V == B2_VARIANT_PATH) {
        const float fmax = hmax(s.f), rr_prob = fminf(fmax * sqr(s.eta), 0.95f);
        const bool rr_active = s.depth >= (uint32_t) p.rr_depth, rr_continue = s.rng.next_1d() < rr_prob;
        if (rr_active) s.f *= rcp(rr_prob);
        active = fmax != 0.f && (!rr_active || rr_continue);                                // path.py:268-277
    } else if (V == B2_VARIANT_PATH_MIS) {
        s.depth += 1;
        const float tmax = hmax(s.f), rr_prob = fminf(tmax * sqr(s.eta), 0.95f);
        const bool rr_active = s.depth >= (uint32_t) p.rr_depth, rr_continue = s.rng.next_1d() < rr_prob;
        if (rr_active) s.f *= rcp(rr_prob);
        active = (!rr_active || rr_continue) && tmax != 0.f;                                // path-mis.py:143-153
    } else {
        const float beta_max = hmax(s.f), rr_prob = fminf(beta_max * sqr(s.eta), 0.95f);
        const bool rr_active = s.depth >= (uint32_t) p.rr_depth;
        if (rr_active) s.f *= rcp(rr_prob);
        const bool rr_continue = s.rng.next_1d() < rr_prob;
        active = beta_max != 0.f && (!rr_active || rr_continue);                            // prb.py:188-200
        if (ADJ) {
            // d/d(reflectance) of [ Lr_dir + L * bsdf_val / detach(bsdf_weight * pdf) ] (prb.py:202-250); every
            // supported BSDF is linear in its reflectance, so the derivative is the value with reflectance := 1.
            f3 g(0.f);
            if (active_em) {
                f3 unit; float dummy; bsdf_eval_pdf(bsdf, si.wi, wo, ones, unit, dummy);
                g += s.dL * (f_before * mis_em * unit * em_weight);
            }
            if (active) {
                const f3 wo2 = si.sh.to_local(s.ray.d);
                f3 unit; float dummy; bsdf_eval_pdf(bsdf, si.wi, wo2, ones, unit, dummy);
                const f3 det = bsdf_weight * bs.pdf;
                const f3 inv(det.x != 0.f ? rcp(det.x) : 0.f, det.y != 0.f ? rcp(det.y) : 0.f, det.z != 0.f ? rcp(det.z) : 0.f);
                g += s.dL * (s.L * (inv * unit));
            }
            if (!is_zero(g)) {
                for (int t = 0; t < p.n_targets; ++t) {
                    if (p.targets[t].bsdf != si.bsdf) continue;
                    if (p.targets[t].kind == 0) {
                        atomicAdd(gs.s_const + 3 * t, g.x); atomicAdd(gs.s_const + 3 * t + 1, g.y); atomicAdd(gs.s_const + 3 * t + 2, g.z);
                        cnt.add(C_GRAD, 3);
                    } else {
                        uint32_t idx[4]; float w[4]; Counters<false> none;
                        tex_eval(sc, bsdf.tex, si.uv, idx, w, none);
                        float *gp = p.targets[t].grad;
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            atomicAdd(gp + 3 * (size_t) idx[q], g.x * w[q]); atomicAdd(gp + 3 * (size_t) idx[q] + 1, g.y * w[q]);
                            atomicAdd(gp + 3 * (size_t) idx[q] + 2, g.z * w[q]);
                        }
                        cnt.add(C_GRAD, 12);
                    }
                }
            }
        }
        s.depth += 1;
    }
    return active;
}

// ------------------------------------------------------------------------------------------ persistent kernel
template <int V, int MODE, bool COUNT>
__global__ void __launch_bounds__(kBlockThreads) k_render(const __grid_constant__ RenderParams p) {
    extern __shared__ float4 s_tile[];                 // [tile_rows][tile_cols] RGBW
    __shared__ uint32_t s_unit, s_next;
    __shared__ float s_gconst[kMaxGradTargets * 3];
    const DSensor &sn = p.sc.sensor;
    const int tid = threadIdx.x;
    const uint32_t tile_cells = p.tile_rows * p.tile_cols;
    Counters<COUNT> cnt;
    GradSink gs{ s_gconst };
    if (MODE == MODE_BACKWARD) { if (tid < kMaxGradTargets * 3) s_gconst[tid] = 0.f; }

    for (;;) {
        if (tid == 0) s_unit = p.unit_begin + atomicAdd(p.unit_counter, 1u);
        __syncthreads();
        const uint32_t unit = s_unit;
        if (unit >= p.unit_end) break;
        const uint32_t row = unit / p.units_per_row, ux = unit - row * p.units_per_row;
        const uint32_t x0 = ux * p.unit_pixels, x1 = min(x0 + p.unit_pixels, (uint32_t) sn.crop_width);
        uint64_t lane_lo = ((uint64_t) row * sn.crop_width + x0) * p.spp_per_pass, lane_hi = ((uint64_t) row * sn.crop_width + x1) * p.spp_per_pass;
        lane_lo = max(lane_lo, p.lane_begin); lane_hi = min(lane_hi, p.lane_end);
        const uint32_t n_lanes = lane_hi > lane_lo ? (uint32_t) (lane_hi - lane_lo) : 0u;
        const int tx0 = (int) x0 - p.filter_n, ty0 = (int) row - p.filter_n;
        if (MODE != MODE_BACKWARD) for (uint32_t i = tid; i < tile_cells; i += kBlockThreads) s_tile[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (tid == 0) s_next = 0u;
        __syncthreads();

        Path s; bool have = false, exhausted = false;
        for (;;) {
            if (!have && !exhausted) {
                const uint32_t i = atomicAdd(&s_next, 1u);
                if (i >= n_lanes) exhausted = true;
                else {
                    s.lane = (uint32_t) (lane_lo + i); s.pass = 0; s.adjoint = false;
                    s.rng.seed_lane(p.seed, s.lane);
                    start_path<V>(p, s);
                    have = true;
                }
            }
            if (__all_sync(0xffffffffu, !have)) break;
            if (!have) continue;
            bool alive;
            if (MODE == MODE_WEIGHTS) alive = false;
            else if (MODE == MODE_BACKWARD && s.adjoint) alive = path_step<V, true>(p, s, gs, cnt);
            else alive = path_step<V, false>(p, s, gs, cnt);
            if (alive) continue;
            // ---- path finished
            if (MODE == MODE_BACKWARD) {
                if (!s.adjoint) {
                    // primal pass done (common.py:750-760); replay the identical stream for the adjoint pass (:763-773)
                    const f3 L = s.L;
                    s.rng.seed_lane(p.seed, s.lane);
                    start_path<V>(p, s);
                    s.adjoint = true; s.L = L;
                    // dL = sum_p grad_in[p] * w(p) / W[p]   (adjoint of splat + develop, common.py:728-744)
                    f3 dL(0.f);
                    film_taps(p, s.pos, [&](int x, int y, float wgt) {
                        const size_t pix = (size_t) y * sn.crop_width + x;
                        float W = __ldg(p.wfilm + 4 * pix + 3); W = W == 0.f ? 1.f : W;
                        const float k = wgt / W;
                        dL += f3(__ldg(p.grad_in + 3 * pix) * k, __ldg(p.grad_in + 3 * pix + 1) * k, __ldg(p.grad_in + 3 * pix + 2) * k);
                    });
                    s.dL = dL;
                } else { cnt.add(C_SAMPLES); have = false; }
            } else {
                f3 L = s.L;
                if (MODE == MODE_WEIGHTS) L = f3(0.f);
                else if (V == B2_VARIANT_PATH_MIS && !s.valid_ray) L = f3(0.f);
                const uint32_t taps = film_taps(p, s.pos, [&](int x, int y, float wgt) {
                    const int cx = x - tx0, cy = y - ty0;
                    if ((unsigned) cx < p.tile_cols && (unsigned) cy < p.tile_rows) {
                        float *c = (float *) (s_tile + cy * p.tile_cols + cx);
                        atomicAdd(c, L.x * wgt); atomicAdd(c + 1, L.y * wgt); atomicAdd(c + 2, L.z * wgt); atomicAdd(c + 3, 1.f * wgt);
                    } else {        // cannot happen for lanes of this unit; kept as a safety net
                        float *c = p.film + 4 * ((size_t) y * sn.crop_width + x);
                        atomicAdd(c, L.x * wgt); atomicAdd(c + 1, L.y * wgt); atomicAdd(c + 2, L.z * wgt); atomicAdd(c + 3, 1.f * wgt);
                    }
                });
                cnt.add(C_TAPS, taps); cnt.add(C_SAMPLES);
                s.pass += 1;
                if (s.pass < p.n_passes) start_path<V>(p, s);       // next pass continues the lane's stream (integrator.cpp:319-328)
                else have = false;
            }
        }
        __syncthreads();
        // ---- flush the tile: one 128-bit reduction per touched pixel
        if (MODE != MODE_BACKWARD) {
            for (uint32_t i = tid; i < tile_cells; i += kBlockThreads) {
                const float4 v = s_tile[i];
                if (v.x == 0.f && v.y == 0.f && v.z == 0.f && v.w == 0.f) continue;
                const int cy = i / p.tile_cols, cx = i - cy * p.tile_cols, x = tx0 + cx, y = ty0 + cy;
                if ((unsigned) x < (unsigned) sn.crop_width && (unsigned) y < (unsigned) sn.crop_height)
                    atomicAdd((float4 *) (p.film + 4 * ((size_t) y * sn.crop_width + x)), v);
            }
        }
        __syncthreads();
    }

    if (MODE == MODE_BACKWARD) {
        __syncthreads();
        if (tid < p.n_targets * 3 && p.targets[tid / 3].kind == 0) { const float v = s_gconst[tid]; if (v != 0.f) atomicAdd(p.targets[tid / 3].grad + tid % 3, v); }
    }
    if constexpr (COUNT) {
        if (p.counters) {
#pragma unroll 1
            for (int k = 0; k < C_N; ++k) {
                unsigned long long v = cnt.v[k];
                for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if ((tid & 31) == 0 && v) atomicAdd(p.counters + k, v);
            }
        }
    }
}

template <int V, int MODE, bool COUNT>
static void launch_t(const RenderParams &p, int blocks, size_t smem, cudaStream_t st) {
    cudaFuncSetAttribute(k_render<V, MODE, COUNT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    k_render<V, MODE, COUNT><<<blocks, kBlockThreads, smem, st>>>(p);
}
template <int V, int MODE, bool COUNT>
static int occ_t(size_t smem) {
    int n = 0;
    cudaFuncSetAttribute(k_render<V, MODE, COUNT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_render<V, MODE, COUNT>, kBlockThreads, smem);
    return n;
}

#define B2_DISPATCH(FN, ...)                                                                                   \
    do {                                                                                                       \
        if (mode == MODE_WEIGHTS) return FN<B2_VARIANT_PRB, MODE_WEIGHTS, false>(__VA_ARGS__);                 \
        if (mode == MODE_BACKWARD) return count ? FN<B2_VARIANT_PRB, MODE_BACKWARD, true>(__VA_ARGS__)         \
                                                : FN<B2_VARIANT_PRB, MODE_BACKWARD, false>(__VA_ARGS__);       \
        switch (variant) {                                                                                     \
            case B2_VARIANT_PATH: return count ? FN<B2_VARIANT_PATH, MODE_RENDER, true>(__VA_ARGS__)           \
                                               : FN<B2_VARIANT_PATH, MODE_RENDER, false>(__VA_ARGS__);         \
            case B2_VARIANT_PATH_MIS: return count ? FN<B2_VARIANT_PATH_MIS, MODE_RENDER, true>(__VA_ARGS__)   \
                                                   : FN<B2_VARIANT_PATH_MIS, MODE_RENDER, false>(__VA_ARGS__); \
            default: return count ? FN<B2_VARIANT_PRB, MODE_RENDER, true>(__VA_ARGS__)                         \
                                  : FN<B2_VARIANT_PRB, MODE_RENDER, false>(__VA_ARGS__);                       \
        }                                                                                                      \
    } while (0)

void launch_render(int variant, int mode, bool count, const RenderParams &p, int blocks, size_t smem, cudaStream_t st) {
    B2_DISPATCH(launch_t, p, blocks, smem, st);
}
int render_occupancy(int variant, int mode, bool count, size_t smem) {
    B2_DISPATCH(occ_t, smem);
}

// ------------------------------------------------------------------------------------------ develop / trace
// HDRFilm::develop (hdrfilm.cpp:304-405): rgb / (W == 0 ? 1 : W)
__global__ void k_develop(const float4 *__restrict__ film, float *__restrict__ image, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 v = film[i];
    const float w = v.w == 0.f ? 1.f : v.w;
    image[3 * (size_t) i] = v.x / w; image[3 * (size_t) i + 1] = v.y / w; image[3 * (size_t) i + 2] = v.z / w;
}
void launch_develop(const float *film, float *image, uint32_t n_pixels, cudaStream_t st) {
    if (n_pixels) k_develop<<<(n_pixels + 255) / 256, 256, 0, st>>>((const float4 *) film, image, n_pixels);
}

// Scene::ray_intersect_preliminary / ray_test for explicit rays (unit-test hook)
__global__ void k_trace(const __grid_constant__ DScene sc, const float *__restrict__ rays, uint64_t n, int any_hit, b2_hit *__restrict__ hits) {
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *r = rays + 8 * i;
    const f3 o(r[0], r[1], r[2]), d(r[3], r[4], r[5]); const float maxt = r[6];
    Hit h; Counters<false> cnt;
    if (any_hit) {
        const bool occ = bvh_trace<true>(sc, o, d, maxt, h, cnt);
        hits[i] = b2_hit{ occ ? 0.f : maxt, 0.f, 0.f, occ ? 0u : 0xffffffffu };
    } else {
        bvh_trace<false>(sc, o, d, maxt, h, cnt);
        hits[i] = b2_hit{ h.t, h.u, h.v, h.face };
    }
}
void launch_trace(const DScene &sc, const float *rays, uint64_t n, int any_hit, b2_hit *hits, cudaStream_t st) {
    if (n) k_trace<<<(unsigned) ((n + 127) / 128), 128, 0, st>>>(sc, rays, n, any_hit, hits);
}

}  // namespace b2
