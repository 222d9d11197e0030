// wavefront.cu -- shade / classify / splat kernels of the wavefront render loop and its host driver (default product path
// for b2_render).  The trace kernel lives in wf_trace.cu.
//
//   k_wf_trace      Scene::ray_intersect_preliminary + Scene::ray_test    scene.cpp:191-235 (Embree/OptiX in the reference)
//                   + (bounce 0) lane -> pixel, sampler seeding, sensor ray integrator.cpp:293-310,419-436, sampler.cpp:129-148,
//                                                                          perspective.cpp:239-280
//   k_wf_shade<V>   one bounce of path.py / path-mis.py / prb.py          b2_path.cuh (path_step<WF>)
//   k_wf_splat      ImageBlock::put                                        imageblock.cpp:174-560
#include "b2_path.cuh"
#include "b2_wavefront.cuh"
#include <algorithm>
#include <cstdlib>

namespace b2 {

#ifndef B2_SHADE_MINBLOCKS
#define B2_SHADE_MINBLOCKS 3
#endif

// ------------------------------------------------------------------------------------------ state <-> registers
B2_DEV uint32_t wf_bits(const Path &s, bool pending) {
    return (s.depth & WF_DEPTH_MASK) | (s.prev_delta ? WF_FLAG_PREV_DELTA : 0u) | (s.first ? WF_FLAG_FIRST : 0u) | (s.valid_ray ? WF_FLAG_VALID_RAY : 0u) |
           (pending ? WF_FLAG_PENDING : 0u);
}
B2_DEV Elem wf_rng_elem(const Path &s) {
    Elem e;
    e.a = make_float4(__uint_as_float((uint32_t) s.rng.state), __uint_as_float((uint32_t) (s.rng.state >> 32)), __uint_as_float((uint32_t) s.rng.inc), __uint_as_float((uint32_t) (s.rng.inc >> 32)));
    e.b = make_float4(s.f.x, s.f.y, s.f.z, s.prev_pdf);
    return e;
}
B2_DEV void wf_set_rng(Path &s, const float4 &r) {
    s.rng.state = (uint64_t) __float_as_uint(r.x) | ((uint64_t) __float_as_uint(r.y) << 32);
    s.rng.inc = (uint64_t) __float_as_uint(r.z) | ((uint64_t) __float_as_uint(r.w) << 32);
}
// one 256-bit store per stream; a lane that ended keeps only its radiance element (and its sampler for a later pass)
B2_DEV void wf_store_path(const WfState &st, uint32_t slot, const Path &s, bool alive, bool pending, uint32_t vis, bool keep_rng) {
    Elem e;
    e.a = make_float4(s.L.x, s.L.y, s.L.z, __uint_as_float(wf_bits(s, pending)));
    e.b = make_float4(s.prev_p.x, s.prev_p.y, s.prev_p.z, __uint_as_float(vis));
    st256(&st.rad[slot], e);
    if (alive || keep_rng) st256(&st.rng[slot], wf_rng_elem(s));
    if (alive) {
        e.a = make_float4(s.ray.o.x, s.ray.o.y, s.ray.o.z, s.ray.maxt);
        e.b = make_float4(s.ray.d.x, s.ray.d.y, s.ray.d.z, s.eta);
        st256(&st.ray[slot], e);
    }
}
// returns the lane's flag bits; `vis` = visibility bits recorded so far (PRB)
B2_DEV uint32_t wf_load_path(const WfState &st, uint32_t slot, Path &s, uint32_t &vis) {
    const Elem r = ld256(&st.ray[slot]), g = ld256(&st.rng[slot]), l = ld256(&st.rad[slot]);
    wf_set_rng(s, g.a);
    s.ray.o = f3(r.a.x, r.a.y, r.a.z); s.ray.maxt = r.a.w; s.ray.d = f3(r.b.x, r.b.y, r.b.z); s.eta = r.b.w;
    s.f = f3(g.b.x, g.b.y, g.b.z); s.prev_pdf = g.b.w; s.L = f3(l.a.x, l.a.y, l.a.z);
    const uint32_t bits = __float_as_uint(l.a.w);
    s.depth = bits & WF_DEPTH_MASK; s.prev_delta = (bits & WF_FLAG_PREV_DELTA) != 0; s.first = (bits & WF_FLAG_FIRST) != 0;
    s.valid_ray = (bits & WF_FLAG_VALID_RAY) != 0;
    s.prev_p = f3(l.b.x, l.b.y, l.b.z); vis = __float_as_uint(l.b.w);
    s.adjoint = false; s.dL = f3(0.f); s.pass = 0; s.lane = 0; s.pos = f2{ 0.f, 0.f };
    return bits;
}
// Deferred next-event estimation: the emitter sample of the lane's previous bounce is added here if trace flagged its
// shadow ray as arrived -- L <- fma(a, b, L), the same operation in the same order as the reference's loop.
B2_DEV void wf_apply_pending(const WfState &st, uint32_t slot, f3 &L, uint32_t &vis, bool record_vis) {
    const Elem f = ld256(&st.shf[slot]);
    if (f.a.w != 0.f) {
        L.x = fma_(f.a.x, f.b.x, L.x); L.y = fma_(f.a.y, f.b.y, L.y); L.z = fma_(f.a.z, f.b.z, L.z);
        if (record_vis) vis |= 1u << (__float_as_uint(f.b.w) & 31u);
    }
}

template <bool COUNT>
B2_DEV void wf_flush_counters(const Counters<COUNT> &cnt, unsigned long long *out) {
    if constexpr (COUNT) {
        if (out) {
#pragma unroll 1
            for (int k = 0; k < C_N; ++k) {
                unsigned long long v = cnt.v[k];
                for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if ((threadIdx.x & 31) == 0 && v) atomicAdd(out + k, v);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------ classify
// Sorts the closest hits of iteration `it` by shading class (miss / diffuse / textured diffuse / glossy / delta) into
// per-class queues, so that every warp of a shade kernel runs one BSDF and the kernels can be specialised: block-wide
// counting sort step = five ballots, one atomic per block and class.
constexpr int kClassifyLanes = 4;       // lanes per thread between two block-wide scans: the scan's atomics are a ~1 us round trip
__global__ void __launch_bounds__(256) k_wf_classify(const __grid_constant__ WfLaunch w) {
    __shared__ uint32_t s_cnt[CLS_COUNT][8 * kClassifyLanes], s_base[CLS_COUNT];
    const WfState &st = w.st;
    const uint32_t it = w.it, n = w.first ? w.n : st.ctl->n_closest[it & 3u];
    const uint32_t *qin = st.q[it & 1u];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (blockIdx.x == 0 && tid < 8) {
        const uint32_t z = (it + 2u) & 3u;
        if (tid == 0) { st.ctl->n_closest[z] = 0u; st.ctl->n_shadow[z] = 0u; st.ctl->fetch[z] = 0u; }
        st.ctl->n_class[z][tid] = 0u;
    }
    constexpr uint32_t kGroup = 256u * kClassifyLanes;
    for (uint32_t base = blockIdx.x * kGroup; base < n; base += gridDim.x * kGroup) {
        uint32_t slot[kClassifyLanes], mine[kClassifyLanes]; int cls[kClassifyLanes];
#pragma unroll
        for (int u = 0; u < kClassifyLanes; ++u) {
            const uint32_t i = base + u * 256u + tid;
            slot[u] = 0; cls[u] = -1; mine[u] = 0;
            if (i < n) { slot[u] = w.first ? i : qin[i]; cls[u] = (int) st.cls_q[i]; if (cls[u] >= CLS_COUNT) cls[u] = CLS_MISS; }   // class by queue position, written by trace(it)
        }
#pragma unroll
        for (int u = 0; u < kClassifyLanes; ++u) {
#pragma unroll
            for (int k = 0; k < CLS_COUNT; ++k) {
                const uint32_t m = __ballot_sync(0xffffffffu, cls[u] == k);
                if (lane == 0) s_cnt[k][u * 8 + warp] = __popc(m);
                if (cls[u] == k) mine[u] = m;
            }
        }
        __syncthreads();
        if (tid < CLS_COUNT) {
            uint32_t tot = 0;
            for (int k = 0; k < 8 * kClassifyLanes; ++k) { const uint32_t c = s_cnt[tid][k]; s_cnt[tid][k] = tot; tot += c; }
            s_base[tid] = tot ? atomicAdd(&st.ctl->n_class[it & 3u][tid], tot) : 0u;
        }
        __syncthreads();
#pragma unroll
        for (int u = 0; u < kClassifyLanes; ++u)
            if (cls[u] >= 0) st.qc[cls[u]][s_base[cls[u]] + s_cnt[cls[u]][u * 8 + warp] + __popc(mine[u] & ((1u << lane) - 1u))] = slot[u];
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------ shade
// SORTED: the input is the class queue qc[CLS] filled by k_wf_classify; otherwise the whole live queue (CLS then only
// states what every BSDF of the scene is -- single-class scenes such as the Cornell box skip the sort).
// w.first: bounce 0 of a pass -- the lane state does not exist yet; it is derived from the lane index exactly as the trace
// kernel derived the camera ray (start_path), so nothing but the hit record is read.
#ifndef B2_SHADE_THREADS
#define B2_SHADE_THREADS 256
#endif
constexpr int kShadeT = B2_SHADE_THREADS, kShadeW = kShadeT / 32;
template <int V, bool COUNT, bool ADJ = false, int CLS = CLS_GENERIC, bool SORTED = (CLS != CLS_GENERIC)>
__global__ void __launch_bounds__(kShadeT, ((CLS == CLS_DIFFUSE || CLS == CLS_MISS) ? 4 : B2_SHADE_MINBLOCKS) * (256 / kShadeT)) k_wf_shade(const __grid_constant__ RenderParams p, const __grid_constant__ WfLaunch w) {
    __shared__ uint32_t s_cnt[2][kShadeW], s_base[2];
    __shared__ float s_gconst[kMaxGradTargets * 3];     // adjoint: per-block sums of the constant-albedo gradients
    const WfState &st = w.st;
    const uint32_t it = w.it;
    const uint32_t n = SORTED ? st.ctl->n_class[it & 3u][SORTED ? CLS : 0] : (w.first ? w.n : st.ctl->n_closest[it & 3u]);
    const uint32_t *qin = !SORTED ? st.q[it & 1u] : st.qc[SORTED ? CLS : 0];
    uint32_t *qout = st.q[(it + 1u) & 1u], *qsh = st.q_shadow;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool keep_rng = p.n_passes > 1;
    Counters<COUNT> cnt; GradSink gs{ s_gconst };
    if (ADJ) { if (tid < kMaxGradTargets * 3) s_gconst[tid] = 0.f; __syncthreads(); }
    if (!SORTED && blockIdx.x == 0 && tid == 0) { const uint32_t z = (it + 2u) & 3u; st.ctl->n_closest[z] = 0u; st.ctl->n_shadow[z] = 0u; st.ctl->fetch[z] = 0u; }

    for (uint32_t base = blockIdx.x * (uint32_t) kShadeT; base < n; base += gridDim.x * (uint32_t) kShadeT) {
        const uint32_t i = base + tid;
        bool alive = false, shadow = false; uint32_t slot = 0;
        if (i < n) {
            slot = (!SORTED && w.first) ? i : qin[i];
            Path s; uint32_t vis = 0;
            Hit h;
            if (w.no_hit) { h.t = 0.f; h.u = h.v = 0.f; h.face = 0xffffffffu; h.mesh = 0xffffffffu; }
            else { const Elem hh = ld256(&st.hit[slot]); h.t = hh.a.x; h.u = hh.a.y; h.v = hh.a.z; h.face = __float_as_uint(hh.a.w); h.mesh = __float_as_uint(hh.b.x); }
            if (!ADJ && w.first) {
                s.lane = (uint32_t) (w.chunk_begin + slot); s.pass = w.pass; s.adjoint = false; s.dL = f3(0.f);
                if (w.pass == 0) s.rng.seed_lane(p.seed, global_lane(p, s.lane));
                else wf_set_rng(s, __ldg(&st.rng[slot].a));
                start_path<V>(p, s);
                if (w.store_pos) st.pos[slot] = make_float2(s.pos.x, s.pos.y);
                if (w.no_hit) h.t = s.ray.maxt;
            } else {
                const uint32_t bits = wf_load_path(st, slot, s, vis);
                if (!ADJ && (bits & WF_FLAG_PENDING)) wf_apply_pending(st, slot, s.L, vis, w.record_vis != 0u);
            }
            ShadowRec so; so.valid = false; so.tex = false; so.vis = false; so.a = f3(0.f); so.b = f3(0.f);
            if (ADJ) {
                const float4 g = st.dl[slot]; s.dL = f3(g.x, g.y, g.z); s.adjoint = true;
                so.vis = ((vis >> w.bounce) & 1u) != 0u;
            }
            alive = path_step<V, ADJ, COUNT, true, CLS>(p, s, gs, cnt, &h, &so);
            if (ADJ && !alive) cnt.add(C_SAMPLES);
            if (ADJ && B2_FWD(p)) st.dl[slot] = make_float4(s.dL.x, s.dL.y, s.dL.z, 0.f);
            shadow = so.valid;
            wf_store_path(st, slot, s, alive, shadow, vis, keep_rng);
            if (shadow) {
                Elem e;
                e.a = make_float4(so.ray.o.x, so.ray.o.y, so.ray.o.z, so.ray.maxt);
                e.b = make_float4(so.ray.d.x, so.ray.d.y, so.ray.d.z, so.tex ? 1.f : 0.f);
                st256(&st.shr[slot], e);
                e.a = make_float4(so.a.x, so.a.y, so.a.z, 0.f);                            // .w: set by trace when the ray arrives
                e.b = make_float4(so.b.x, so.b.y, so.b.z, __uint_as_float(w.bounce));     // .w: the bounce of this emitter sample
                st256(&st.shf[slot], e);
            }
        }
        // ---- compaction: warp ballot -> per-warp counts -> one atomic per block and queue (keeps the queue in lane
        // order within a block, which the next trace launch rewards with coherent rays; per-warp atomics measured 10 % slower)
        const uint32_t ma = __ballot_sync(0xffffffffu, alive), ms = __ballot_sync(0xffffffffu, shadow);
        if (lane == 0) { s_cnt[0][warp] = __popc(ma); s_cnt[1][warp] = __popc(ms); }
        __syncthreads();
        if (tid < 2) {
            uint32_t tot = 0;
            for (int k = 0; k < kShadeW; ++k) { const uint32_t c = s_cnt[tid][k]; s_cnt[tid][k] = tot; tot += c; }
            s_base[tid] = tot ? atomicAdd(tid == 0 ? &st.ctl->n_closest[(it + 1u) & 3u] : &st.ctl->n_shadow[it & 3u], tot) : 0u;
        }
        __syncthreads();
        const uint32_t lt = (1u << lane) - 1u;
        if (alive) qout[s_base[0] + s_cnt[0][warp] + __popc(ma & lt)] = slot;
        if (shadow) qsh[s_base[1] + s_cnt[1][warp] + __popc(ms & lt)] = slot;
        __syncthreads();
    }
    if (ADJ) {
        __syncthreads();
        if (tid < p.n_targets * 3 && p.targets[tid / 3].kind == 0) { const float v = s_gconst[tid]; if (v != 0.f) atomicAdd(p.targets[tid / 3].grad + tid % 3, v); }
    }
    wf_flush_counters(cnt, p.counters);
}

// Adjoint pass set-up (RBIntegrator.render_backward, common.py:728-773): the lane keeps the radiance L of its primal
// pass (the last deferred emitter sample is added here), reseeds its sampler to replay the identical stream, and gathers
// dL = sum_p grad_in[p] * w(p) / W[p], the adjoint of splat + develop w.r.t. the splatted value.
__global__ void __launch_bounds__(256) k_wf_generate_adj(const __grid_constant__ RenderParams p, const __grid_constant__ WfLaunch w) {
    const WfState &st = w.st; const DSensor &sn = p.sc.sensor;
    if (blockIdx.x == 0 && threadIdx.x < 4) {
        const uint32_t k = threadIdx.x;
        st.ctl->n_closest[k] = (k == (w.it & 3u)) ? w.n : 0u; st.ctl->n_shadow[k] = 0u; st.ctl->fetch[k] = 0u;
        for (int c = 0; c < 8; ++c) st.ctl->n_class[k][c] = 0u;
    }
    uint32_t *q = st.q[w.it & 1u];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < w.n; i += gridDim.x * blockDim.x) {
        Path s;
        s.lane = (uint32_t) (w.chunk_begin + i); s.pass = 0; s.adjoint = true;
        const Elem l = ld256(&st.rad[i]);
        f3 L(l.a.x, l.a.y, l.a.z); uint32_t vis = __float_as_uint(l.b.w);
        if (__float_as_uint(l.a.w) & WF_FLAG_PENDING) wf_apply_pending(st, i, L, vis, true);
        s.rng.seed_lane(p.seed, global_lane(p, s.lane));
        start_path<B2_VARIANT_PRB>(p, s);
        s.L = L;
        f3 dL(0.f);
        if (!B2_FWD(p)) film_taps(p, s.pos, [&](int x, int y, float wgt) {
            const size_t pix = (size_t) y * sn.crop_width + x;
            float W = __ldg(p.wfilm + 4 * pix + 3); W = W == 0.f ? 1.f : W;
            const float k = wgt / W;
            dL += f3(__ldg(p.grad_in + 3 * pix) * k, __ldg(p.grad_in + 3 * pix + 1) * k, __ldg(p.grad_in + 3 * pix + 2) * k);
        });
        wf_store_path(st, i, s, true, false, vis, false);
        st.dl[i] = make_float4(dL.x, dL.y, dL.z, 0.f);
        q[i] = i;
    }
}

// forward mode: the accumulated d L of every lane becomes the value the splat kernel reads
__global__ void __launch_bounds__(256) k_wf_dl_to_rad(const __grid_constant__ WfLaunch w) {
    const WfState &st = w.st;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < w.n; i += gridDim.x * blockDim.x) {
        const float4 g = st.dl[i]; const float4 l = st.rad[i].a;
        st.rad[i].a = make_float4(g.x, g.y, g.z, __uint_as_float(__float_as_uint(l.w) & ~WF_FLAG_PENDING));
    }
}

// ------------------------------------------------------------------------------------------ splat
// Units (a run of pixels of one image row x all of its samples) are pulled from a global counter; the samples of a
// unit are splatted into a shared-memory RGBW tile which is flushed with one 128-bit reduction per touched pixel.
template <int V, bool COUNT, bool WEIGHTS_ONLY>
__global__ void __launch_bounds__(256) k_wf_splat(const __grid_constant__ RenderParams p, const __grid_constant__ WfLaunch w) {
    extern __shared__ float4 s_tile[];
    __shared__ uint32_t s_unit;
    const DSensor &sn = p.sc.sensor; const WfState &st = w.st;
    const int tid = threadIdx.x;
    const uint32_t tile_cells = p.tile_rows * p.tile_cols;
    const uint64_t c_lo = w.chunk_begin, c_hi = w.chunk_begin + w.n;
    Counters<COUNT> cnt;
    for (;;) {
        if (tid == 0) s_unit = p.unit_begin + atomicAdd(p.unit_counter, 1u);
        __syncthreads();
        const uint32_t unit = s_unit;
        if (unit >= p.unit_end) break;
        const uint32_t row = unit / p.units_per_row, ux = unit - row * p.units_per_row;
        const uint32_t x0 = ux * p.unit_pixels, x1 = min(x0 + p.unit_pixels, (uint32_t) sn.grid_w);
        uint64_t lane_lo = ((uint64_t) row * sn.grid_w + x0) * p.spp_per_pass, lane_hi = ((uint64_t) row * sn.grid_w + x1) * p.spp_per_pass;
        lane_lo = max(lane_lo, c_lo); lane_hi = min(lane_hi, c_hi);
        const uint32_t n_lanes = lane_hi > lane_lo ? (uint32_t) (lane_hi - lane_lo) : 0u;
        const int tx0 = (int) x0 - sn.border - p.filter_n, ty0 = (int) row - sn.border - p.filter_n;   // crop-relative tile origin
        for (uint32_t i = tid; i < tile_cells; i += 256) s_tile[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        __syncthreads();
        const uint32_t g = (p.spp_shift != 0xffffffffu) ? min(p.spp_per_pass, 32u) : 1u;     // lanes per pixel group in a warp
        for (uint32_t base = 0; base < n_lanes; base += 256) {
            const uint32_t i = base + tid; const bool valid = i < n_lanes;
            f2 ps{ 0.f, 0.f }; f3 L(0.f);
            if (valid) {
                if (WEIGHTS_ONLY) {             // weight film of the sample set: positions re-derived from (seed, lane)
                    Pcg32 rng; const uint32_t ln = (uint32_t) (lane_lo + i); rng.seed_lane(p.seed, global_lane(p, ln));
                    int px, py; ps = lane_sample_pos(p, ln, rng, px, py);
                    if (sn.rfilter == B2_RFILTER_BOX) ps = f2{ (float) px, (float) py };
                } else {
                    const uint32_t slot = (uint32_t) (lane_lo + i - c_lo);
                    if (w.store_pos) { const float2 q = st.pos[slot]; ps = f2{ q.x, q.y }; }
                    else {                      // single pass: the position is the first two draws of the lane's stream
                        Pcg32 rng; const uint32_t ln = (uint32_t) (lane_lo + i); rng.seed_lane(p.seed, global_lane(p, ln));
                        int px, py; ps = lane_sample_pos(p, ln, rng, px, py);
                        if (sn.rfilter == B2_RFILTER_BOX) ps = f2{ (float) px, (float) py };
                    }
                    const Elem l = ld256(&st.rad[slot]); L = f3(l.a.x, l.a.y, l.a.z);
                    const uint32_t bits = __float_as_uint(l.a.w);
                    if (bits & WF_FLAG_PENDING) { uint32_t vis = 0; wf_apply_pending(st, slot, L, vis, false); }   // the lane's last emitter sample
                    if (V == B2_VARIANT_PATH_MIS && !(bits & WF_FLAG_VALID_RAY)) L = f3(0.f);
                }
            }
            auto add = [&](int x, int y, float4 v) {
                const int cx = x - tx0, cy = y - ty0;
                float *c = ((unsigned) cx < p.tile_cols && (unsigned) cy < p.tile_rows) ? (float *) (s_tile + cy * p.tile_cols + cx)
                                                                                        : p.film + 4 * ((size_t) y * sn.crop_width + x);
                if (!WEIGHTS_ONLY) { atomicAdd(c, v.x); atomicAdd(c + 1, v.y); atomicAdd(c + 2, v.z); }
                atomicAdd(c + 3, v.w);
            };
            // samples of one pixel sit in adjacent lanes: reduce each tap over the pixel's lanes, one add per group
            bool fast = false; const uint32_t lid = tid & 31u, lead = lid & ~(g - 1u);
            if (g > 1u) {
                const uint32_t key = valid ? (uint32_t) ((lane_lo + i) >> p.spp_shift) : 0xffffffffu;
                const int ax = (int) floorf(ps.x), ay = (int) floorf(ps.y);
                const uint32_t lk = __shfl_sync(0xffffffffu, key, lead);
                const int lax = __shfl_sync(0xffffffffu, ax, lead), lay = __shfl_sync(0xffffffffu, ay, lead);
                fast = __all_sync(0xffffffffu, !valid || (key == lk && ax == lax && ay == lay));
            }
            auto add1 = [&](int x, int y, int ch, float val) {
                const int cx = x - tx0, cy = y - ty0;
                if ((unsigned) cx < p.tile_cols && (unsigned) cy < p.tile_rows) atomicAdd((float *) s_tile + (cy * p.tile_cols + cx) * 4 + ch, val);
                else atomicAdd(p.film + 4 * ((size_t) y * sn.crop_width + x) + ch, val);
            };
            uint32_t taps;
            const uint32_t fn = sn.rfilter == B2_RFILTER_BOX ? 0u : (uint32_t) p.filter_n;
            if (!WEIGHTS_ONLY && fast && g == 32u && fn == 1u) taps = film_taps_bfly<32, 3>(p, ps, L, valid, lid, add1);
            else if (!WEIGHTS_ONLY && fast && g == 32u && fn == 2u) taps = film_taps_bfly<32, 5>(p, ps, L, valid, lid, add1);
            else if (!WEIGHTS_ONLY && fast && g == 16u && fn == 1u) taps = film_taps_bfly<16, 3>(p, ps, L, valid, lid, add1);
            else if (!WEIGHTS_ONLY && fast && g == 16u && fn == 2u) taps = film_taps_bfly<16, 5>(p, ps, L, valid, lid, add1);
            else if (!WEIGHTS_ONLY && fast && g == 8u && fn == 1u) taps = film_taps_bfly<8, 3>(p, ps, L, valid, lid, add1);
            else if (!WEIGHTS_ONLY && fast && g == 8u && fn == 2u) taps = film_taps_bfly<8, 5>(p, ps, L, valid, lid, add1);
            else if (WEIGHTS_ONLY && fast && g == 32u && fn == 1u) taps = film_weights_bfly<32, 3>(p, ps, valid, lid, add1);
            else if (WEIGHTS_ONLY && fast && g == 32u && fn == 2u) taps = film_weights_bfly<32, 5>(p, ps, valid, lid, add1);
            else if (WEIGHTS_ONLY && fast && g == 16u && fn == 1u) taps = film_weights_bfly<16, 3>(p, ps, valid, lid, add1);
            else if (fast) taps = film_taps_group<WEIGHTS_ONLY>(p, ps, L, valid, g, lid == lead, add);
            else taps = film_taps_group<WEIGHTS_ONLY>(p, ps, L, valid, 1u, true, add);
            cnt.add(C_TAPS, taps); if (valid) cnt.add(C_SAMPLES);
        }
        __syncthreads();
        for (uint32_t i = tid; i < tile_cells; i += 256) {
            const float4 v = s_tile[i];
            if (v.x == 0.f && v.y == 0.f && v.z == 0.f && v.w == 0.f) continue;
            const int cy = i / p.tile_cols, cx = i - cy * p.tile_cols, x = tx0 + cx, y = ty0 + cy;
            if ((unsigned) x < (unsigned) sn.crop_width && (unsigned) y < (unsigned) sn.crop_height)
                atomicAdd((float4 *) (p.film + 4 * ((size_t) y * sn.crop_width + x)), v);
        }
        __syncthreads();
    }
    wf_flush_counters(cnt, p.counters);
}

// ------------------------------------------------------------------------------------------ host driver
static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

size_t wf_state_bytes(uint32_t cap) {
    const size_t c = cap;
    return align_up(c * 32, 256) * 6 + align_up(c * 16, 256) + align_up(c * 8, 256) + align_up(c, 256) + align_up(c * 4, 256) * (3 + CLS_COUNT) + 256;
}

void wf_bind(WfState &st, void *base, uint32_t cap) {
    char *q = (char *) base; const size_t c = cap;
    auto take = [&](size_t bytes) { void *r = q; q += align_up(bytes, 256); return r; };
    st.ctl = (WfCtl *) take(256);
    st.ray = (Elem *) take(c * 32); st.hit = (Elem *) take(c * 32); st.rng = (Elem *) take(c * 32); st.rad = (Elem *) take(c * 32);
    st.shr = (Elem *) take(c * 32); st.shf = (Elem *) take(c * 32);
    st.dl = (float4 *) take(c * 16); st.pos = (float2 *) take(c * 8); st.cls_q = (uint8_t *) take(c);
    st.q[0] = (uint32_t *) take(c * 4); st.q[1] = (uint32_t *) take(c * 4); st.q_shadow = (uint32_t *) take(c * 4);
    for (int k = 0; k < CLS_COUNT; ++k) st.qc[k] = (uint32_t *) take(c * 4);
    st.capacity = cap;
}

// optional per-kernel-kind device timing (bench.py's roofline): CUDA events around every launch, summed after the render
static thread_local WfTiming *g_tm = nullptr;
// The film is first touched by the splat launches: a caller whose film arrives on another stream (b2_render_host's upload) hands
// over the event that marks it complete instead of delaying the whole render behind the copy.
static thread_local cudaEvent_t g_film_ready = nullptr;
void wf_set_film_ready_event(cudaEvent_t e) { g_film_ready = e; }
void wf_set_timing(WfTiming *t) { g_tm = t; }
void wf_tm_begin(int kind, cudaStream_t st) {
    WfTiming *t = g_tm; if (!t) return;
    if (t->used + 2 > t->ev.size()) { cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b); t->ev.push_back(a); t->ev.push_back(b); }
    t->kind.push_back(kind);
    cudaEventRecord(t->ev[t->used], st);
}
void wf_tm_end(cudaStream_t st) { WfTiming *t = g_tm; if (!t) return; cudaEventRecord(t->ev[t->used + 1], st); t->used += 2; }
static void tm_begin(int kind, cudaStream_t st) { wf_tm_begin(kind, st); }
static void tm_end(cudaStream_t st) { wf_tm_end(st); }
void wf_collect_timing(WfTiming &t, double ms[WF_K_COUNT], uint64_t n[WF_K_COUNT]) {
    for (int k = 0; k < WF_K_COUNT; ++k) { ms[k] = 0.0; n[k] = 0; }
    for (size_t i = 0; i < t.kind.size(); ++i) {
        float e = 0.f; cudaEventElapsedTime(&e, t.ev[2 * i], t.ev[2 * i + 1]);
        ms[t.kind[i]] += e; n[t.kind[i]]++;
    }
    t.kind.clear(); t.used = 0;
}

// tuning knobs (environment overrides are for sweeps on the GPU box; the defaults are the measured optimum)
static int env_int(const char *name, int def) { const char *v = getenv(name); return v ? std::max(1, atoi(v)) : def; }
static const uint32_t tune_tri = (uint32_t) env_int("B2_TRI_THRESH", 12);
static const uint32_t tune_refill = (uint32_t) env_int("B2_REFILL_THRESH", 6);
static const int tune_sort = getenv("B2_SHADE_SORT") ? atoi(getenv("B2_SHADE_SORT")) : -1;   // 0/1 force, -1 = by scene content

// resident blocks per SM of one kernel instantiation (queried once): every shade specialisation gets the grid its own
// register count allows (the diffuse kernels fit 4 blocks/SM, the glossy ones 3)
template <auto Kern> static int occ_of() {
    static const int occ = [] { int o = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, Kern, kShadeT, 0); return std::max(o, 1); }();
    return occ;
}
static const int tune_shade_mult = getenv("B2_SHADE_GRID_MULT") ? std::max(1, atoi(getenv("B2_SHADE_GRID_MULT"))) : 1;
#define B2_SHADE_LAUNCH(...) do { constexpr auto kern_ = k_wf_shade<__VA_ARGS__>; \
    const int g_ = (int) std::min<uint64_t>((n + kShadeT - 1) / kShadeT, (uint64_t) occ_of<kern_>() * sm_count * tune_shade_mult); \
    tm_begin(WF_K_SHADE, stream); kern_<<<g_, kShadeT, 0, stream>>>(p, w); tm_end(stream); launches++; } while (0)

static WfLaunch wf_launch_defaults(const WfState &st, uint64_t c0, uint32_t n) {
    WfLaunch w{}; w.st = st; w.chunk_begin = c0; w.n = n; w.tri_thresh = tune_tri; w.refill_thresh = std::min(tune_refill, 32u);
    return w;
}

template <int V, bool COUNT>
static long wf_render_t(RenderParams p, WfState &st, int sm_count, cudaStream_t stream, const char **err) {
    long launches = 0;
    const DSensor &sn = p.sc.sensor;
    const uint32_t W = (uint32_t) sn.grid_w, P = p.unit_pixels;
    int occ_shade = 0, occ_splat = 0;
    const size_t smem_splat = sizeof(float4) * (size_t) p.tile_cols * p.tile_rows;
    cudaFuncSetAttribute(k_wf_splat<V, COUNT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem_splat);
    const int g_trace = wf_trace_blocks(COUNT, sm_count);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_shade, k_wf_shade<V, COUNT>, kShadeT, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_splat, k_wf_splat<V, COUNT, false>, 256, smem_splat);
    if (g_trace <= 0 || occ_shade <= 0 || occ_splat <= 0) { *err = "wavefront: a kernel cannot be resident (occupancy 0)"; return -1; }
    const uint64_t lane_lo = p.lane_begin, lane_hi = p.lane_end;
    const bool skip_first_trace = V == B2_VARIANT_PATH && p.max_depth <= 1;      // path.py:235: no first intersection at all
    const bool bounded = p.max_depth <= 64;
    const uint32_t n_iter = bounded ? (uint32_t) std::max(p.max_depth, 1) : 0u;
    uint32_t it = 0;
    for (uint64_t c0 = lane_lo; c0 < lane_hi; c0 += st.capacity) {
        const uint32_t n = (uint32_t) std::min<uint64_t>(st.capacity, lane_hi - c0);
        WfLaunch w = wf_launch_defaults(st, c0, n);
        w.store_pos = p.n_passes > 1 ? 1u : 0u;
        const int g_classify = (int) std::min<uint64_t>((n + 256 * kClassifyLanes - 1) / (256 * kClassifyLanes), (uint64_t) 8 * sm_count);
        for (uint32_t pass = 0; pass < p.n_passes; ++pass) {
            w.pass = pass;
            cudaMemsetAsync(st.ctl, 0, sizeof(WfCtl), stream);          // queue bookkeeping of a pass starts from zero
            for (uint32_t k = 0;; ++k) {
                if (bounded && k >= n_iter) break;
                w.it = it; w.bounce = k; w.first = k == 0 ? 1u : 0u; w.no_hit = (k == 0 && skip_first_trace) ? 1u : 0u;
                if (!w.no_hit) { wf_launch_trace(COUNT, k == 0, p, w, g_trace, stream); launches++; }
                const uint32_t cmask = st.class_mask;
                // (no class sort when the first trace was skipped -- Path with max_depth <= 1: nothing wrote the class bytes
                // and every lane is a miss; the generic kernel takes the whole queue)
                const bool sort = !w.no_hit &&
                                  (tune_sort >= 0 ? tune_sort != 0     // default: sort when the BSDFs differ enough to diverge
                                                  : (__builtin_popcount(cmask) >= 2 && (cmask & ((1u << CLS_CONDUCTOR) | (1u << CLS_DIELECTRIC) | (1u << CLS_PLASTIC) | (1u << CLS_ROUGHDIEL))) != 0));
                if (!sort && cmask == (1u << CLS_DIFFUSE)) B2_SHADE_LAUNCH(V, COUNT, false, CLS_DIFFUSE, false);
                else if (sort) {
                    tm_begin(WF_K_SHADE, stream); k_wf_classify<<<g_classify, 256, 0, stream>>>(w); tm_end(stream); launches++;
                    const uint32_t cm = st.class_mask;
                    if (cm & (1u << CLS_DIFFUSE)) B2_SHADE_LAUNCH(V, COUNT, false, CLS_DIFFUSE);
                    if (cm & (1u << CLS_DIFFUSE_TEX)) B2_SHADE_LAUNCH(V, COUNT, false, CLS_DIFFUSE_TEX);
                    if (cm & (1u << CLS_CONDUCTOR)) B2_SHADE_LAUNCH(V, COUNT, false, CLS_CONDUCTOR);
                    if (cm & (1u << CLS_PLASTIC)) B2_SHADE_LAUNCH(V, COUNT, false, CLS_PLASTIC);
                    if (cm & (1u << CLS_ROUGHDIEL)) B2_SHADE_LAUNCH(V, COUNT, false, CLS_ROUGHDIEL);
                    if (cm & (1u << CLS_DIELECTRIC)) B2_SHADE_LAUNCH(V, COUNT, false, CLS_DIELECTRIC);
                    B2_SHADE_LAUNCH(V, COUNT, false, CLS_MISS);
                } else B2_SHADE_LAUNCH(V, COUNT);
                ++it;
                if (!bounded && (k & 3u) == 3u) {       // unbounded depth: Russian roulette ends the paths; poll the live count
                    uint32_t live = 0;
                    if (cudaMemcpyAsync(&live, &st.ctl->n_closest[it & 3u], 4, cudaMemcpyDeviceToHost, stream) != cudaSuccess ||
                        cudaStreamSynchronize(stream) != cudaSuccess) { *err = "wavefront: polling the live-lane count failed"; return -1; }
                    if (live == 0) break;
                }
            }
            w.it = it; w.first = 0u; w.no_hit = 0u;     // shadow rays of the last bounce
            wf_launch_trace(COUNT, false, p, w, g_trace, stream); launches++;
            ++it;
            // splat this chunk: units covering its pixels
            RenderParams ps = p;
            const uint64_t pix0 = c0 / p.spp_per_pass, pix1 = (c0 + n - 1) / p.spp_per_pass;
            ps.unit_begin = (uint32_t) ((pix0 / W) * p.units_per_row + (pix0 % W) / P);
            ps.unit_end = (uint32_t) ((pix1 / W) * p.units_per_row + (pix1 % W) / P) + 1;
            cudaMemsetAsync(p.unit_counter, 0, 4, stream);
            const int g_splat = (int) std::min<uint32_t>(ps.unit_end - ps.unit_begin, (uint32_t) (occ_splat * sm_count));
            if (g_film_ready) cudaStreamWaitEvent(stream, g_film_ready, 0);      // b2_render_host: the film upload overlaps the bounces
            tm_begin(WF_K_SPLAT, stream); k_wf_splat<V, COUNT, false><<<g_splat, 256, smem_splat, stream>>>(ps, w); tm_end(stream); launches++;
        }
    }
    if (cudaGetLastError() != cudaSuccess) { *err = "wavefront: kernel launch failed"; return -1; }
    return launches;
}

// hist: optional [n_iter][capacity] hit records.  With it, the primal pass keeps the closest hit of every bounce and the adjoint
// replay (identical rays by construction) reads them back instead of tracing every closest ray a second time.
template <bool COUNT>
static long wf_render_backward_t(RenderParams p, WfState &st, Elem *hist, int sm_count, cudaStream_t stream, const char **err) {
    constexpr int V = B2_VARIANT_PRB;
    long launches = 0;
    int occ_shade = 0, occ_adj = 0;
    const int g_trace = wf_trace_blocks(COUNT, sm_count);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_shade, k_wf_shade<V, COUNT, false>, kShadeT, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_adj, k_wf_shade<V, COUNT, true>, kShadeT, 0);
    int occ_splat = 1; const size_t smem_splat = sizeof(float4) * (size_t) p.tile_cols * p.tile_rows;
    if (B2_FWD(p)) {
        cudaFuncSetAttribute(k_wf_splat<V, COUNT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem_splat);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_splat, k_wf_splat<V, COUNT, false>, 256, smem_splat);
        if (occ_splat <= 0) { *err = "wavefront: the splat kernel cannot be resident"; return -1; }
    }
    if (g_trace <= 0 || occ_shade <= 0 || occ_adj <= 0) { *err = "wavefront: a kernel cannot be resident (occupancy 0)"; return -1; }
    if (p.max_depth > 32) { *err = "wavefront render_backward: max_depth > 32"; return -1; }
    const uint32_t n_iter = (uint32_t) std::max(p.max_depth, 1);
    uint32_t it = 0;
    for (uint64_t c0 = p.lane_begin; c0 < p.lane_end; c0 += st.capacity) {
        const uint32_t n = (uint32_t) std::min<uint64_t>(st.capacity, p.lane_end - c0);
        WfLaunch w = wf_launch_defaults(st, c0, n);
        const int g_gen = (int) std::min<uint64_t>((n + 255) / 256, (uint64_t) sm_count * 8);
        const int g_adj = (int) std::min<uint64_t>((n + kShadeT - 1) / kShadeT, (uint64_t) occ_adj * sm_count);
        // ---- primal pass (common.py:750-760): L per lane, and the visibility bit of every emitter sample
        cudaMemsetAsync(st.ctl, 0, sizeof(WfCtl), stream);
        w.record_vis = 1u;
        for (uint32_t k = 0; k < n_iter; ++k) {
            w.it = it; w.bounce = k; w.first = k == 0 ? 1u : 0u;
            if (hist) w.st.hit = hist + (size_t) k * st.capacity;
            wf_launch_trace(COUNT, k == 0, p, w, g_trace, stream); launches++;
            B2_SHADE_LAUNCH(V, COUNT, false);
            ++it;
        }
        w.it = it; w.first = 0u; w.st.hit = st.hit;
        wf_launch_trace(COUNT, false, p, w, g_trace, stream); launches++;
        ++it;
        // ---- adjoint pass (common.py:763-773): replay of the identical stream, closest-hit rays only
        w.it = it; w.record_vis = 0u;
        tm_begin(WF_K_OTHER, stream); k_wf_generate_adj<<<g_gen, 256, 0, stream>>>(p, w); tm_end(stream); launches++;
        for (uint32_t k = 0; k < n_iter; ++k) {
            w.it = it; w.bounce = k;
            if (hist) w.st.hit = hist + (size_t) k * st.capacity;               // the hits the primal pass recorded for this bounce
            else { wf_launch_trace(COUNT, false, p, w, g_trace, stream); launches++; }
            tm_begin(WF_K_SHADE, stream); k_wf_shade<V, COUNT, true><<<g_adj, kShadeT, 0, stream>>>(p, w); tm_end(stream); launches++;
            ++it;
        }
        w.st.hit = st.hit;
        if (B2_FWD(p)) {
            // forward mode: splat the lanes' d L (accumulated in the dL stream) like a render splats L (common.py:585-600)
            const DSensor &sn = p.sc.sensor; const uint32_t W = (uint32_t) sn.grid_w, P = p.unit_pixels;
            tm_begin(WF_K_OTHER, stream); k_wf_dl_to_rad<<<g_gen, 256, 0, stream>>>(w); tm_end(stream); launches++;
            RenderParams ps = p;
            const uint64_t pix0 = c0 / p.spp_per_pass, pix1 = (c0 + n - 1) / p.spp_per_pass;
            ps.unit_begin = (uint32_t) ((pix0 / W) * p.units_per_row + (pix0 % W) / P);
            ps.unit_end = (uint32_t) ((pix1 / W) * p.units_per_row + (pix1 % W) / P) + 1;
            cudaMemsetAsync(p.unit_counter, 0, 4, stream);
            const int g_splat = (int) std::min<uint32_t>(ps.unit_end - ps.unit_begin, (uint32_t) (occ_splat * sm_count));
            if (g_film_ready) cudaStreamWaitEvent(stream, g_film_ready, 0);      // b2_render_host: the film upload overlaps the bounces
            tm_begin(WF_K_SPLAT, stream); k_wf_splat<V, COUNT, false><<<g_splat, 256, smem_splat, stream>>>(ps, w); tm_end(stream); launches++;
        }
    }
    if (cudaGetLastError() != cudaSuccess) { *err = "wavefront: kernel launch failed"; return -1; }
    return launches;
}

// Weight film of a sample set (ADIntegrator: the dummy splat of common.py:728-741): every lane of the pass wavefront
// splats weight 1 at its film position, which depends only on (seed, lane).  p.film = the weight film, units = all.
long wf_weights(RenderParams p, int sm_count, cudaStream_t stream, const char **err) {
    const size_t smem = sizeof(float4) * (size_t) p.tile_cols * p.tile_rows;
    auto k = k_wf_splat<B2_VARIANT_PRB, false, true>;
    int occ = 0;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, 256, smem);
    if (occ <= 0) { *err = "wavefront: the splat kernel cannot be resident"; return -1; }
    if (p.unit_end <= p.unit_begin) return 0;
    WfLaunch w{}; w.chunk_begin = p.lane_begin; w.n = (uint32_t) (p.lane_end - p.lane_begin);
    cudaMemsetAsync(p.unit_counter, 0, 4, stream);
    k<<<(int) std::min<uint32_t>(p.unit_end - p.unit_begin, (uint32_t) (occ * sm_count)), 256, smem, stream>>>(p, w);
    if (cudaGetLastError() != cudaSuccess) { *err = "wavefront: kernel launch failed"; return -1; }
    return 1;
}

// ------------------------------------------------------------------------------------------ b2_trace through the production kernel
// The unit-test hook traces caller-supplied rays with k_wf_trace itself: the rays become the ray (closest) or shadow-ray (any hit)
// stream of a wavefront chunk, the queues are the identity, and the hit records / arrival flags are converted back.
__global__ void __launch_bounds__(256) k_wf_rays_in(const __grid_constant__ WfLaunch w, const float *rays, int any_hit) {
    const WfState &st = w.st;
    if (blockIdx.x == 0 && threadIdx.x == 0) { if (any_hit) st.ctl->n_shadow[(w.it + 3u) & 3u] = w.n; else st.ctl->n_closest[w.it & 3u] = w.n; }
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < w.n; i += gridDim.x * blockDim.x) {
        const float *r = rays + 8 * (size_t) i;
        Elem e; e.a = make_float4(r[0], r[1], r[2], r[6]); e.b = make_float4(r[3], r[4], r[5], 0.f);
        if (any_hit) { st256(&st.shr[i], e); Elem f; f.a = make_float4(0.f, 0.f, 0.f, 0.f); f.b = f.a; st256(&st.shf[i], f); st.q_shadow[i] = i; }
        else { st256(&st.ray[i], e); st.q[w.it & 1u][i] = i; }
    }
}
__global__ void __launch_bounds__(256) k_wf_hits_out(const __grid_constant__ WfLaunch w, const float *rays, int any_hit, b2_hit *hits) {
    const WfState &st = w.st;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < w.n; i += gridDim.x * blockDim.x) {
        if (any_hit) { const bool occ = st.shf[i].a.w == 0.f; hits[i] = b2_hit{ occ ? 0.f : rays[8 * (size_t) i + 6], 0.f, 0.f, occ ? 0u : 0xffffffffu }; }
        else { const float4 h = st.hit[i].a; hits[i] = b2_hit{ h.x, h.y, h.z, __float_as_uint(h.w) }; }
    }
}
long wf_trace_rays(RenderParams p, WfState &st, const float *rays, uint64_t n, int any_hit, b2_hit *hits, int sm_count, cudaStream_t stream, const char **err) {
    const int g_trace = wf_trace_blocks(false, sm_count);
    if (g_trace <= 0) { *err = "wavefront: the trace kernel cannot be resident"; return -1; }
    long launches = 0;
    for (uint64_t c0 = 0; c0 < n; c0 += st.capacity) {
        const uint32_t m = (uint32_t) std::min<uint64_t>(st.capacity, n - c0);
        WfLaunch w = wf_launch_defaults(st, c0, m); w.it = 0;
        const int g = (int) std::min<uint64_t>((m + 255) / 256, (uint64_t) sm_count * 8);
        cudaMemsetAsync(st.ctl, 0, sizeof(WfCtl), stream);
        k_wf_rays_in<<<g, 256, 0, stream>>>(w, rays + 8 * c0, any_hit);
        wf_launch_trace(false, false, p, w, g_trace, stream);
        k_wf_hits_out<<<g, 256, 0, stream>>>(w, rays + 8 * c0, any_hit, hits + c0);
        launches += 3;
    }
    if (cudaGetLastError() != cudaSuccess) { *err = "wavefront: kernel launch failed"; return -1; }
    return launches;
}

long wf_trace_primary(RenderParams p, WfState &st, uint64_t lane_begin, uint32_t n, int packet, b2_hit *hits, int sm_count, cudaStream_t stream, const char **err) {
    const int g_trace = wf_trace_blocks(false, sm_count);
    if (g_trace <= 0) { *err = "wavefront: the trace kernel cannot be resident"; return -1; }
    if (n > st.capacity) { *err = "wavefront: more lanes than the reserved state holds"; return -1; }
    WfLaunch w = wf_launch_defaults(st, lane_begin, n); w.it = 0; w.first = 1u;
    const int g = (int) std::min<uint64_t>((n + 255) / 256, (uint64_t) sm_count * 8);
    cudaMemsetAsync(st.ctl, 0, sizeof(WfCtl), stream);
    wf_launch_trace_primary(packet, p, w, g_trace, stream);
    k_wf_hits_out<<<g, 256, 0, stream>>>(w, nullptr, 0, hits);
    if (cudaGetLastError() != cudaSuccess) { *err = "wavefront: kernel launch failed"; return -1; }
    return 2;
}

long wf_render_backward(bool count, RenderParams p, WfState &st, void *hit_history, int sm_count, cudaStream_t stream, const char **err) {
    return count ? wf_render_backward_t<true>(p, st, (Elem *) hit_history, sm_count, stream, err) : wf_render_backward_t<false>(p, st, (Elem *) hit_history, sm_count, stream, err);
}

long wf_render(int variant, bool count, RenderParams p, WfState &st, int sm_count, cudaStream_t stream, const char **err) {
    switch (variant) {
        case B2_VARIANT_PATH: return count ? wf_render_t<B2_VARIANT_PATH, true>(p, st, sm_count, stream, err) : wf_render_t<B2_VARIANT_PATH, false>(p, st, sm_count, stream, err);
        case B2_VARIANT_PATH_MIS: return count ? wf_render_t<B2_VARIANT_PATH_MIS, true>(p, st, sm_count, stream, err) : wf_render_t<B2_VARIANT_PATH_MIS, false>(p, st, sm_count, stream, err);
        default: return count ? wf_render_t<B2_VARIANT_PRB, true>(p, st, sm_count, stream, err) : wf_render_t<B2_VARIANT_PRB, false>(p, st, sm_count, stream, err);
    }
}

}  // namespace b2
