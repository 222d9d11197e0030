// b2_path.cuh -- per-lane path state and ONE bounce of the three loop formulations (device code shared by the
// persistent megakernel in render.cu and the wavefront kernels in wavefront.cu).
//   loops    path.py:194-302 (B2_VARIANT_PATH), path-mis.py:24-155 == integrators/path.cpp:94-327 (B2_VARIANT_PATH_MIS),
//            ad/integrators/prb.py:63-265 (B2_VARIANT_PRB, primal and adjoint)
//   driver   SamplingIntegrator::render_sample (src/render/integrator.cpp:419-491)
//   film     ImageBlock::put (src/render/imageblock.cpp:174-560), rfilters (gaussian.cpp:94-103, tent.cpp:53-55, box.cpp:41-43)
#pragma once
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

// sample shard (b2_render_args.sample_begin / sample_count): the lane of the shard -> the lane of the full wavefront whose
// PCG32 stream it draws (sampler.cpp:129-148 seeds by the global index)
B2_DEV uint32_t global_lane(const RenderParams &p, uint32_t lane) {
    if (p.remap_spp == 0u) return lane;
    const uint32_t pix = (p.spp_shift != 0xffffffffu) ? (lane >> p.spp_shift) : (lane / p.spp_per_pass);
    return pix * p.remap_spp + p.remap_off + (lane - pix * p.spp_per_pass);
}

// lane -> pixel (integrator.cpp:293-310) and the jittered film position (the first two draws of the lane's stream)
B2_DEV f2 lane_sample_pos(const RenderParams &p, uint32_t lane, Pcg32 &rng, int &px, int &py) {
    const DSensor &sn = p.sc.sensor;
    uint32_t idx = lane;
    idx = (p.spp_shift != 0xffffffffu) ? (idx >> p.spp_shift) : (idx / p.spp_per_pass);
    const uint32_t y = idx / (uint32_t) sn.grid_w, x = idx - (uint32_t) sn.grid_w * y;
    px = (int) x - sn.border + sn.crop_offset_x; py = (int) y - sn.border + sn.crop_offset_y;
    const f2 j = rng.next_2d();
    return f2{ (float) px + j.x, (float) py + j.y };
}

// ImageBlock::put for a group of `g` adjacent lanes (g a power of two <= 32) whose samples fall on the same pixel and
// hence share one footprint: every tap value is summed over the group with xor-shuffles and the group's first lane
// issues ONE add per tap and channel -- the warp-aggregated form of the per-sample atomics (the reference aggregates
// the same way inside dr.scatter_reduce, drjit-core cuda_scatter.cpp:63-230).  All 32 lanes must call this together.
// add(x, y, rgbw) receives crop-relative in-bounds coordinates.  Returns this lane's own in-bounds tap count.
template <bool WEIGHTS_ONLY, class ADD>
B2_DEV uint32_t film_taps_group(const RenderParams &p, f2 pos, f3 L, bool valid, uint32_t g, bool leader, ADD &&add) {
    const DSensor &s = p.sc.sensor;
    const bool box = s.rfilter == B2_RFILTER_BOX;
    const int n = box ? 0 : p.filter_n, count = 2 * n + 1;
    const int pix = (int) floorf(pos.x) - n, piy = (int) floorf(pos.y) - n;
    const int lx = pix - s.crop_offset_x, ly = piy - s.crop_offset_y;
    float relx = (float) pix + .5f - pos.x, rely = (float) piy + .5f - pos.y;
    float wx[5], wy[5];
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        if (i < count && valid) { wx[i] = box ? 1.f : rfilter_eval(s, relx); wy[i] = box ? 1.f : rfilter_eval(s, rely); relx += 1.f; rely += 1.f; }
        else { wx[i] = 0.f; wy[i] = 0.f; }
    }
    uint32_t nx = 0, ny = 0;
#pragma unroll
    for (int ys = 0; ys < 5; ++ys) {
        if (ys >= count) break;
        const bool yin = (unsigned) (ly + ys) < (unsigned) s.crop_height; ny += yin;
#pragma unroll
        for (int xs = 0; xs < 5; ++xs) {
            if (xs >= count) break;
            const bool xin = (unsigned) (lx + xs) < (unsigned) s.crop_width; if (ys == 0) nx += xin;
            const float wgt = wy[ys] * wx[xs];
            float4 v = make_float4(L.x * wgt, L.y * wgt, L.z * wgt, 1.f * wgt);
            for (uint32_t o = g >> 1; o; o >>= 1) {
                if (!WEIGHTS_ONLY) { v.x += __shfl_xor_sync(0xffffffffu, v.x, o); v.y += __shfl_xor_sync(0xffffffffu, v.y, o); v.z += __shfl_xor_sync(0xffffffffu, v.z, o); }
                v.w += __shfl_xor_sync(0xffffffffu, v.w, o);
            }
            if (leader && valid && yin && xin) add(lx + xs, ly + ys, v);
        }
    }
    return valid ? nx * ny : 0u;
}

// Butterfly variant of film_taps_group for pixel groups of G = 8 / 16 / 32 adjacent lanes and a C x C footprint (C = 3, 5; RGBW).
// The 4 C^2 tap values of a lane are reduced G at a time by a transposing butterfly -- G - 1 shuffles per G values instead of
// G log2 G -- after which lane j of the group holds the group's sum of value j = (tap j >> 2, channel j & 3) and adds it itself:
// one add per lane and chunk, in parallel, instead of 4 C^2 serial ones by the group's leader (fp32 adds to shared memory are
// compare-and-swap loops).  C^2 is odd and G / 4 even: exactly one tap is left over and goes the xor-reduction way.
// add1(x, y, channel, value) receives crop-relative in-bounds coordinates.  Returns this lane's own in-bounds tap count.
template <int G, int C, class ADD1>
B2_DEV uint32_t film_taps_bfly(const RenderParams &p, f2 pos, f3 L, bool valid, uint32_t lid, ADD1 &&add1) {
    const DSensor &s = p.sc.sensor;
    constexpr int n = (C - 1) / 2, NT = C * C, TPC = G / 4;
    const int pix = (int) floorf(pos.x) - n, piy = (int) floorf(pos.y) - n;
    float relx = (float) pix + .5f - pos.x, rely = (float) piy + .5f - pos.y;
    float wx[C], wy[C];
#pragma unroll
    for (int i = 0; i < C; ++i) {
        wx[i] = valid ? rfilter_eval(s, relx) : 0.f; wy[i] = valid ? rfilter_eval(s, rely) : 0.f; relx += 1.f; rely += 1.f;
    }
    // the footprint of the group is its leader's (lanes past the end of the unit carry zeros and a meaningless position)
    const uint32_t gl = lid & (uint32_t) (G - 1), lead = lid & ~(uint32_t) (G - 1);
    const int lx = __shfl_sync(0xffffffffu, pix - s.crop_offset_x, lead), ly = __shfl_sync(0xffffffffu, piy - s.crop_offset_y, lead);
    const bool gvalid = __shfl_sync(0xffffffffu, valid ? 1 : 0, lead) != 0;
    const float Lc[4] = { L.x, L.y, L.z, 1.f };
#pragma unroll
    for (int t0 = 0; t0 + TPC <= NT; t0 += TPC) {
        float v[G];
#pragma unroll
        for (int j = 0; j < G; ++j) { const int t = t0 + (j >> 2); v[j] = Lc[j & 3] * (wy[t / C] * wx[t % C]); }
#pragma unroll
        for (int o = G / 2; o >= 1; o >>= 1) {
            const bool up = (gl & (uint32_t) o) != 0u;
#pragma unroll
            for (int i = 0; i < o; ++i) {
                const float send = up ? v[i] : v[i + o], keep = up ? v[i + o] : v[i];
                v[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
            }
        }
        const int t = t0 + (int) (gl >> 2), ys = t / C, xs = t - ys * C;
        if (gvalid && (unsigned) (lx + xs) < (unsigned) s.crop_width && (unsigned) (ly + ys) < (unsigned) s.crop_height) add1(lx + xs, ly + ys, (int) (gl & 3u), v[0]);
    }
    {   // the last tap
        const float wgt = wy[C - 1] * wx[C - 1];
        float4 v = make_float4(L.x * wgt, L.y * wgt, L.z * wgt, 1.f * wgt);
#pragma unroll
        for (int o = G / 2; o >= 1; o >>= 1) {
            v.x += __shfl_xor_sync(0xffffffffu, v.x, o); v.y += __shfl_xor_sync(0xffffffffu, v.y, o); v.z += __shfl_xor_sync(0xffffffffu, v.z, o); v.w += __shfl_xor_sync(0xffffffffu, v.w, o);
        }
        const int xs = C - 1, ys = C - 1;
        if (gvalid && gl < 4u && (unsigned) (lx + xs) < (unsigned) s.crop_width && (unsigned) (ly + ys) < (unsigned) s.crop_height)
            add1(lx + xs, ly + ys, (int) gl, gl == 0u ? v.x : gl == 1u ? v.y : gl == 2u ? v.z : v.w);
    }
    uint32_t nx = 0, ny = 0;
#pragma unroll
    for (int i = 0; i < C; ++i) { nx += (unsigned) (lx + i) < (unsigned) s.crop_width; ny += (unsigned) (ly + i) < (unsigned) s.crop_height; }
    return valid ? nx * ny : 0u;
}

// The weight film (W channel only) the same way: the C^2 <= G weights of a lane are one butterfly chunk, lane j < C^2 adds tap j.
template <int G, int C, class ADD1>
B2_DEV uint32_t film_weights_bfly(const RenderParams &p, f2 pos, bool valid, uint32_t lid, ADD1 &&add1) {
    static_assert(C * C <= G, "one chunk");
    const DSensor &s = p.sc.sensor;
    constexpr int n = (C - 1) / 2, NT = C * C;
    const int pix = (int) floorf(pos.x) - n, piy = (int) floorf(pos.y) - n;
    float relx = (float) pix + .5f - pos.x, rely = (float) piy + .5f - pos.y;
    float wx[C], wy[C];
#pragma unroll
    for (int i = 0; i < C; ++i) {
        wx[i] = valid ? rfilter_eval(s, relx) : 0.f; wy[i] = valid ? rfilter_eval(s, rely) : 0.f; relx += 1.f; rely += 1.f;
    }
    const uint32_t gl = lid & (uint32_t) (G - 1), lead = lid & ~(uint32_t) (G - 1);
    const int lx = __shfl_sync(0xffffffffu, pix - s.crop_offset_x, lead), ly = __shfl_sync(0xffffffffu, piy - s.crop_offset_y, lead);
    const bool gvalid = __shfl_sync(0xffffffffu, valid ? 1 : 0, lead) != 0;
    float v[G];
#pragma unroll
    for (int j = 0; j < G; ++j) v[j] = j < NT ? 1.f * (wy[(j < NT ? j : 0) / C] * wx[(j < NT ? j : 0) % C]) : 0.f;
#pragma unroll
    for (int o = G / 2; o >= 1; o >>= 1) {
        const bool up = (gl & (uint32_t) o) != 0u;
#pragma unroll
        for (int i = 0; i < o; ++i) {
            const float send = up ? v[i] : v[i + o], keep = up ? v[i + o] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
    }
    const int ys = (int) gl / C, xs = (int) gl - ys * C;
    if (gvalid && gl < (uint32_t) NT && (unsigned) (lx + xs) < (unsigned) s.crop_width && (unsigned) (ly + ys) < (unsigned) s.crop_height) add1(lx + xs, ly + ys, 3, v[0]);
    uint32_t nx = 0, ny = 0;
#pragma unroll
    for (int i = 0; i < C; ++i) { nx += (unsigned) (lx + i) < (unsigned) s.crop_width; ny += (unsigned) (ly + i) < (unsigned) s.crop_height; }
    return valid ? nx * ny : 0u;
}

// jitter + sensor ray of a lane (SamplingIntegrator::render_sample, integrator.cpp:419-436): consumes the first two draws of
// the lane's stream.  Shared by the trace kernel of bounce 0 and the first shade launch, which must derive the identical ray.
B2_DEV Ray primary_ray(const RenderParams &p, uint32_t lane, Pcg32 &rng, f2 &pos) {
    const DSensor &sn = p.sc.sensor;
    int px, py;
    const f2 sample_pos = lane_sample_pos(p, lane, rng, px, py);
    const f2 adjusted{ fma_(sample_pos.x, sn.scale_x, sn.offset_x), fma_(sample_pos.y, sn.scale_y, sn.offset_y) };
    pos = sn.rfilter == B2_RFILTER_BOX ? f2{ (float) px, (float) py } : sample_pos;
    return camera_ray(sn, adjusted);
}
B2_DEV Ray primary_ray(const RenderParams &p, uint32_t lane, Pcg32 &rng) { f2 pos; return primary_ray(p, lane, rng, pos); }

template <int V>
B2_DEV void start_path(const RenderParams &p, Path &s) {
    s.ray = primary_ray(p, s.lane, s.rng, s.pos);
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
// Deferred next-event-estimation record of the wavefront formulation: the shadow ray and the two factors of the
// contribution, L <- fma(a, b, L) once the ray is known to arrive (b = 1 where the reference writes a plain add).
struct ShadowRec { bool valid, tex, vis; Ray ray; f3 a, b; };   // tex: a texture evaluation is counted iff the ray arrives
                                                                // vis (adjoint replay, input): visibility found by the primal pass

// path.py evaluates ray_intersect with `active &= depth < max_depth` (path.py:235): no first intersection at all when
// max_depth <= 1.
template <int V>
B2_DEV bool first_trace_needed(const RenderParams &p, const Path &s) {
    return !(V == B2_VARIANT_PATH && s.first && !(s.depth < (uint32_t) p.max_depth));
}

// WF = false: the closest hit and the shadow ray are traced here (megakernel).  WF = true: `pre_hit` holds the closest
// hit found by the wavefront trace kernel for s.ray, and the shadow ray is recorded in `so` instead of being traced.
template <int V, bool ADJ, bool COUNT, bool WF = false, int CLS = CLS_GENERIC>
B2_DEV bool path_step(const RenderParams &p, Path &s, const GradSink &gs, Counters<COUNT> &cnt, const Hit *pre_hit = nullptr, ShadowRec *so = nullptr) {
    // WF && ADJ: the adjoint pass replays the lane's sample stream; the visibility of each emitter sample was recorded
    // by the primal pass of the same render_backward call (identical rays), so no shadow ray is traced again.
    const DScene &sc = p.sc;
    const uint32_t max_depth = (uint32_t) p.max_depth;
    Hit h; h.face = 0xffffffffu; h.t = s.ray.maxt; h.u = h.v = 0.f; h.mesh = 0;
    if (WF) { h = *pre_hit; if (CLS == CLS_MISS) h.face = 0xffffffffu; }    // class known at compile time: the ray escaped
    else if (first_trace_needed<V>(p, s)) { cnt.add(C_CLOSEST); bvh_trace<false>(sc, s.ray.o, s.ray.d, s.ray.maxt, h, cnt); }
    SI si; make_si(sc, s.ray, h, si, cnt);

    // ---- emission seen by the incoming ray, MIS-weighted against the previous bounce's emitter sampling
    f3 Le(0.f);
    bool active_next;
    if (V == B2_VARIANT_PATH) {
        if (s.first) {
            active_next = s.depth < max_depth;
            // path.py:235-239: `si.emitter(scene, active).eval(si, active)` with active = depth < max_depth -- with
            // max_depth <= 1 neither the first intersection nor the primary emission is evaluated (L = 0)
            if (active_next) s.L += emitter_eval(sc, si.emitter, si, cnt);
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
            if (V == B2_VARIANT_PATH_MIS) {
                // path-mis.py:129-133 also runs on the terminating iteration: the BSDF is still sampled and
                // valid_ray |= si.is_valid() & ~Null(sampled_type).  Only an opacity mask can sample a null lobe
                // (mask.cpp:123-160: sample1 >= opacity), so only there the third draw of the iteration is looked at.
                if (si.valid) {
                    const DBsdf &bl = sc.bsdfs[si.bsdf];
                    if (bl.flags & B2_BSDF_FLAG_MASK) { s.rng.skip(2); const float s1 = s.rng.next_1d(); if (s1 < bl.opacity) s.valid_ray = true; s.rng.skip(3); return false; }
                    s.valid_ray = true;
                }
            } else s.L = ADJ ? (s.L - Le) : (s.L + Le);
            s.rng.skip(6);
            return false;
        }
    }

    // ---- emitter sampling (NEE) + BSDF sampling.  si is valid from here on.
    const DBsdf &bsdf = sc.bsdfs[si.bsdf];
    ReflCache rc;
    auto refl = [&]() { return reflectance<CLS>(sc, bsdf, si, rc, cnt); };
    bool active_em = (bsdf_lobes<CLS>(bsdf) & LOBE_SMOOTH) != 0;
    DS ds; f3 em_weight;
    sample_emitter_direction(sc, si, s.rng.next_2d(), active_em, p.env_warp, ds, em_weight, cnt, [&](const Ray &sr) {
        if (WF && ADJ) return !so->vis;
        if (WF) { so->valid = true; so->ray = sr; return false; }
        Hit hs; cnt.add(C_SHADOW);
        return bvh_trace<true>(sc, sr.o, sr.d, sr.maxt, hs, cnt);
    });
    if (V != B2_VARIANT_PATH_MIS) active_em = active_em && ds.pdf != 0.f;
    const f3 wo = si.sh.to_local(ds.d);
    const float s1 = s.rng.next_1d(); const f2 s2 = s.rng.next_2d();
    f3 bsdf_val(0.f), bsdf_weight(0.f); float bsdf_pdf = 0.f; BS bs;
    if (WF && !ADJ && V != B2_VARIANT_PATH_MIS) {
        // path.py / prb.py evaluate the BSDF only for unoccluded emitter samples; the wavefront evaluates it before
        // the visibility is known, so the texture-evaluation *event* is counted by the trace kernel on arrival
        Counters<false> quiet; bool tex = false;
        if (active_em) bsdf_eval_pdf<CLS>(bsdf, si.wi, wo, [&]() { tex = cls_textured<CLS>(bsdf.tex); return reflectance<CLS>(sc, bsdf, si, rc, quiet); }, bsdf_val, bsdf_pdf);
        so->tex = tex;
    } else if (active_em) bsdf_eval_pdf<CLS>(bsdf, si.wi, wo, refl, bsdf_val, bsdf_pdf);
    bsdf_sample<CLS>(bsdf, si.wi, s1, s2, refl, bs, bsdf_weight);
    if (V == B2_VARIANT_PATH_MIS && !(bs.type & LOBE_NULL)) s.valid_ray = true;            // path-mis.py:129-133
    const f3 f_before = s.f;
    float mis_em;
    if (V == B2_VARIANT_PATH) {
        mis_em = ds.delta ? 1.f : mis_isfinite(ds.pdf, bsdf_pdf);
        if (active_em) {                                                                    // path.py:258-259
            const f3 c = s.f * bsdf_val * em_weight * mis_em;
            if (WF) { so->a = c; so->b = f3(1.f); } else s.L += c;
        }
    } else if (V == B2_VARIANT_PATH_MIS) {
        mis_em = ds.delta ? 1.f : mis_guard(ds.pdf, bsdf_pdf);
        if (active_em) {                                                                    // path-mis.py:113-117
            const f3 c = bsdf_val * em_weight * mis_em;
            if (WF) { so->a = s.f; so->b = c; } else s.L = fmadd(s.f, c, s.L);
        }
    } else {
        mis_em = ds.delta ? 1.f : mis_isfinite(ds.pdf, bsdf_pdf);
        const f3 Lr_dir = s.f * mis_em * bsdf_val * em_weight;                              // prb.py:149-162
        if (WF && !ADJ) { s.L = s.L + Le; so->a = Lr_dir; so->b = f3(1.f); }
        else s.L = ADJ ? (s.L - Le - Lr_dir) : (s.L + Le + Lr_dir);                         // prb.py:173
    }
    if (WF && !ADJ && p.cull_nee && so->valid) {
        // an emitter sample whose contribution is exactly zero (light behind the surface, back side of the emitter, ...)
        // cannot change L whether or not it arrives: drop the shadow ray (NaN products compare false and are kept)
        const f3 c = so->a * so->b;
        if (c.x == 0.f && c.y == 0.f && c.z == 0.f) so->valid = false;
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
            // d/d(reflectance) of [ Lr_dir + L * bsdf_val / detach(bsdf_weight * pdf) ] (prb.py:202-250): bsdf_eval_pdf in
            // derivative mode (value at reflectance := 1 for the linear BSDFs, the closed form for rough plastic).
            // forward mode (render_forward, common.py:495-621): the same coefficient with dL := 1, contracted with the
            // parameter tangent at this vertex and accumulated in s.dL (the lane's d L, splatted after the replay)
            const f3 dLw = B2_FWD(p) ? f3(1.f) : s.dL;
            f3 g(0.f);
            if (active_em) {
                f3 unit; float dummy; bsdf_eval_pdf<CLS>(bsdf, si.wi, wo, refl, unit, dummy, true);
                g += dLw * (f_before * mis_em * unit * em_weight);
            }
            if (active) {
                const f3 wo2 = si.sh.to_local(s.ray.d);
                f3 unit; float dummy; bsdf_eval_pdf<CLS>(bsdf, si.wi, wo2, refl, unit, dummy, true);
                const f3 det = bsdf_weight * bs.pdf;
                const f3 inv(det.x != 0.f ? rcp(det.x) : 0.f, det.y != 0.f ? rcp(det.y) : 0.f, det.z != 0.f ? rcp(det.z) : 0.f);
                g += dLw * (s.L * (inv * unit));
            }
            if (!is_zero(g)) {
                for (int t = 0; t < p.n_targets; ++t) {
                    if (p.targets[t].bsdf != si.bsdf) continue;
                    float *gp = p.targets[t].grad;
                    if (p.targets[t].kind == 0) {
                        if (B2_FWD(p)) s.dL += g * f3(__ldg(gp), __ldg(gp + 1), __ldg(gp + 2));
                        else {
                            atomicAdd(gs.s_const + 3 * t, g.x); atomicAdd(gs.s_const + 3 * t + 1, g.y); atomicAdd(gs.s_const + 3 * t + 2, g.z);
                            cnt.add(C_GRAD, 3);
                        }
                    } else {
                        uint32_t idx[4]; float w[4]; Counters<false> none;
                        tex_eval(sc, bsdf.tex, si.uv, idx, w, none);
                        if (B2_FWD(p)) {
                            f3 tan(0.f);
#pragma unroll
                            for (int q = 0; q < 4; ++q) tan += f3(__ldg(gp + 3 * (size_t) idx[q]), __ldg(gp + 3 * (size_t) idx[q] + 1), __ldg(gp + 3 * (size_t) idx[q] + 2)) * w[q];
                            s.dL += g * tan;
                        } else {
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
        }
        s.depth += 1;
    }
    return active;
}

}  // namespace b2
