// wavefront.cu -- wavefront kernels of the render loop (default product path for b2_render).
//
//   k_wf_generate   lane -> pixel, sampler seeding, sensor ray            integrator.cpp:293-310,419-436, sampler.cpp:129-148,
//                                                                          perspective.cpp:239-280
//   k_wf_trace      Scene::ray_intersect_preliminary + Scene::ray_test    scene.cpp:191-235 (Embree/OptiX in the reference)
//   k_wf_shade<V>   one bounce of path.py / path-mis.py / prb.py          b2_path.cuh (path_step<WF>)
//   k_wf_splat      ImageBlock::put                                        imageblock.cpp:174-560
//
// Trace kernel: persistent warps pull batches of rays from a device-side counter into a warp-private pool; a lane whose
// ray terminates is refilled at the next slice boundary instead of idling until the slowest ray of its warp is done
// (the idle lanes of the v1 megakernel were 77 % of all issue slots, profiles/r01_v1_*).  Closest-hit and any-hit rays
// share the launch; the first entries of the traversal stack live in shared memory ([depth][thread], conflict free),
// deeper ones in local memory.  Node = five 128-bit loads, triangle = three.
#include "b2_path.cuh"
#include "b2_wavefront.cuh"
#include <algorithm>
#include <cstdlib>

namespace b2 {

constexpr int kTraceThreads = 256;
constexpr int kSmemStack = 8;          // stack entries per thread held in shared memory
constexpr int kLocalStack = 40;        // overflow entries in local memory
constexpr int kTriQueue = 4;           // parked leaf-triangle groups per thread (shared memory)
constexpr int kFetchBatch = 96;        // rays a warp takes from the global counter at a time
#ifndef B2_TRACE_MINBLOCKS
#define B2_TRACE_MINBLOCKS 4
#endif
#ifndef B2_SHADE_MINBLOCKS
#define B2_SHADE_MINBLOCKS 3
#endif
#ifndef B2_TRACE_SLICE
#define B2_TRACE_SLICE 8               // traversal micro-steps between two refill points (re-tuned on the PLOC tree: 5 -> 8, +3 %)
#endif

// ------------------------------------------------------------------------------------------ state <-> registers
B2_DEV void wf_store_path(const WfState &st, uint32_t slot, const Path &s, bool with_ray) {
    st.rng(slot) = make_uint4((uint32_t) s.rng.state, (uint32_t) (s.rng.state >> 32), (uint32_t) s.rng.inc, (uint32_t) (s.rng.inc >> 32));
    uint32_t bits = (s.depth & WF_DEPTH_MASK) | (s.prev_delta ? WF_FLAG_PREV_DELTA : 0u) | (s.first ? WF_FLAG_FIRST : 0u) | (s.valid_ray ? WF_FLAG_VALID_RAY : 0u);
    st.rad(slot) = make_float4(s.L.x, s.L.y, s.L.z, __uint_as_float(bits));
    if (with_ray) {
        st.ray_o(slot) = make_float4(s.ray.o.x, s.ray.o.y, s.ray.o.z, s.ray.maxt);
        st.ray_d(slot) = make_float4(s.ray.d.x, s.ray.d.y, s.ray.d.z, s.eta);
        st.thr(slot) = make_float4(s.f.x, s.f.y, s.f.z, s.prev_pdf);
        st.prev(slot) = make_float4(s.prev_p.x, s.prev_p.y, s.prev_p.z, 0.f);
    }
}
B2_DEV void wf_load_path(const WfState &st, uint32_t slot, Path &s) {
    const uint4 r = st.rng(slot);
    s.rng.state = (uint64_t) r.x | ((uint64_t) r.y << 32); s.rng.inc = (uint64_t) r.z | ((uint64_t) r.w << 32);
    const float4 o = st.ray_o(slot), d = st.ray_d(slot), t = st.thr(slot), l = st.rad(slot), pv = st.prev(slot);
    s.ray.o = f3(o.x, o.y, o.z); s.ray.maxt = o.w; s.ray.d = f3(d.x, d.y, d.z); s.eta = d.w;
    s.f = f3(t.x, t.y, t.z); s.prev_pdf = t.w; s.L = f3(l.x, l.y, l.z);
    const uint32_t bits = __float_as_uint(l.w);
    s.depth = bits & WF_DEPTH_MASK; s.prev_delta = (bits & WF_FLAG_PREV_DELTA) != 0; s.first = (bits & WF_FLAG_FIRST) != 0;
    s.valid_ray = (bits & WF_FLAG_VALID_RAY) != 0;
    s.prev_p = f3(pv.x, pv.y, pv.z);
    s.adjoint = false; s.dL = f3(0.f); s.pass = 0; s.lane = 0; s.pos = f2{ 0.f, 0.f };
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

// ------------------------------------------------------------------------------------------ generate
template <int V>
__global__ void __launch_bounds__(256) k_wf_generate(const __grid_constant__ RenderParams p, const __grid_constant__ WfLaunch w) {
    const WfState &st = w.st;
    if (blockIdx.x == 0 && threadIdx.x < 4) {       // nothing else is in flight: reset the queue bookkeeping
        const uint32_t k = threadIdx.x;
        st.ctl->n_closest[k] = (k == (w.it & 3u)) ? w.n : 0u; st.ctl->n_shadow[k] = 0u; st.ctl->fetch[k] = 0u;
        for (int c = 0; c < 8; ++c) st.ctl->n_class[k][c] = 0u;
    }
    uint32_t *q = st.q[w.it & 1u];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < w.n; i += gridDim.x * blockDim.x) {
        Path s;
        s.lane = (uint32_t) (w.chunk_begin + i); s.pass = w.pass; s.adjoint = false; s.dL = f3(0.f);
        if (w.pass == 0) s.rng.seed_lane(p.seed, global_lane(p, s.lane));
        else { const uint4 r = st.rng(i); s.rng.state = (uint64_t) r.x | ((uint64_t) r.y << 32); s.rng.inc = (uint64_t) r.z | ((uint64_t) r.w << 32); }
        start_path<V>(p, s);
        wf_store_path(st, i, s, true);
        st.pos(i) = make_float2(s.pos.x, s.pos.y);
        st.hit(i) = make_float4(s.ray.maxt, 0.f, 0.f, __uint_as_float(0xffffffffu));     // a miss, unless trace(it) overwrites it
        st.hit_mesh(i) = 0xffffffffu; st.vis(i) = 0u;
        q[i] = i;
    }
}

// ------------------------------------------------------------------------------------------ trace
B2_DEV void prefetch_l1(const void *p) { asm volatile("prefetch.global.L1 [%0];" :: "l"(p)); }

struct TraceRay {
    f3 o, d; float idx, idy, idz, maxt;
    uint32_t octinv, slot, qi; bool any, tex;
    float t, u, v; uint32_t face, mesh; bool found;
    uint2 ng, tg; int sp, nq;
};

template <bool COUNT>
__global__ void __launch_bounds__(kTraceThreads, B2_TRACE_MINBLOCKS) k_wf_trace(const __grid_constant__ DScene sc, const __grid_constant__ WfLaunch w, unsigned long long *counters) {
    __shared__ uint2 s_stack[kSmemStack][kTraceThreads];
    __shared__ uint2 s_pend[kTriQueue][kTraceThreads];
    const WfState &st = w.st;
    const uint32_t slot_c = w.it & 3u, slot_s = (w.it + 3u) & 3u;
    const uint32_t nc = st.ctl->n_closest[slot_c], ns = st.ctl->n_shadow[slot_s], total = nc + ns;
    const uint32_t *qc = st.q[w.it & 1u], *qs = st.q_shadow[0];
    uint32_t *fetch = &st.ctl->fetch[slot_c];
    const int tid = threadIdx.x, lane = tid & 31;
    const uint32_t lt_mask = (1u << lane) - 1u;
    Counters<COUNT> cnt;
    uint2 l_stack[kLocalStack];

    uint32_t pool_next = 0, pool_end = 0;           // warp-uniform
    bool exhausted = total == 0;
    bool have = false;
    TraceRay r;
    r.sp = 0; r.nq = 0; r.any = false; r.tex = false; r.found = false; r.slot = 0; r.ng = make_uint2(0u, 0u); r.tg = make_uint2(0u, 0u);

    for (;;) {
        // ---- refill idle lanes (warp converged here)
        const uint32_t need = __ballot_sync(0xffffffffu, !have);
        if (need && !exhausted) {
            const uint32_t k = __popc(need), rank = __popc(need & lt_mask);
            uint32_t my = 0xffffffffu;
            const uint32_t avail = pool_end - pool_next;
            if (avail >= k) { if (!have) my = pool_next + rank; pool_next += k; }
            else {
                if (!have && rank < avail) my = pool_next + rank;
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(fetch, (uint32_t) kFetchBatch);
                base = __shfl_sync(0xffffffffu, base, 0);
                const uint32_t nb = base < total ? min((uint32_t) kFetchBatch, total - base) : 0u;
                const uint32_t r2 = rank - avail;
                if (!have && rank >= avail && r2 < nb) my = base + r2;
                pool_next = base + min(k - avail, nb); pool_end = base + nb;
                if (nb == 0u) exhausted = true;
            }
            if (my != 0xffffffffu) {
                r.any = my >= nc; r.qi = my;
                r.slot = r.any ? qs[my - nc] : qc[my];
                const float4 o = r.any ? st.sh_o(r.slot) : st.ray_o(r.slot), d = r.any ? st.sh_d(r.slot) : st.ray_d(r.slot);
                r.o = f3(o.x, o.y, o.z); r.d = f3(d.x, d.y, d.z); r.maxt = o.w; r.tex = r.any && d.w != 0.f;
                const float tiny = kTinyDir;
                r.idx = 1.f / (fabsf(r.d.x) > tiny ? r.d.x : copysignf(tiny, r.d.x));
                r.idy = 1.f / (fabsf(r.d.y) > tiny ? r.d.y : copysignf(tiny, r.d.y));
                r.idz = 1.f / (fabsf(r.d.z) > tiny ? r.d.z : copysignf(tiny, r.d.z));
                r.octinv = (r.d.x >= 0.f ? 1u : 0u) | (r.d.y >= 0.f ? 2u : 0u) | (r.d.z >= 0.f ? 4u : 0u);
                r.t = r.maxt; r.u = 0.f; r.v = 0.f; r.face = 0xffffffffu; r.mesh = 0u; r.found = false;
                r.ng = sc.n_nodes ? make_uint2(0u, (1u << (24 + r.octinv)) | 1u) : make_uint2(0u, 0u);
                r.tg = make_uint2(0u, 0u); r.sp = 0; r.nq = 0;
                cnt.add(r.any ? C_SHADOW : C_CLOSEST);
                have = true;
            }
        }
        const uint32_t smask = __ballot_sync(0xffffffffu, have);
        if (smask == 0u) break;
        if (!have) continue;

        // ---- traverse for up to `slice` micro-steps.  Every micro-step the lanes of this slice vote: a NODE step (pop a
        // child, five 128-bit loads, eight branch-free slab tests) for the lanes that have node work, or a TRIANGLE step
        // (one Moeller-Trumbore test per lane) once enough lanes hold pending leaf triangles.  Leaf hits found by a node
        // step are parked in a per-lane queue (shared memory) instead of being tested on the spot, where only ~4 of
        // 32 lanes had one (profiles/r01_v2_trace_*): the arithmetic per ray is that of bvh_trace (b2_bvh.cuh).
        bool done = false;
#pragma unroll 1
        for (uint32_t step = 0; step < w.slice; ++step) {
            const bool node_work = (r.ng.y & 0xff000000u) != 0u || r.sp > 0;
            const bool can_node = node_work && (r.tg.y == 0u || r.nq < kTriQueue);
            const uint32_t P = __ballot_sync(smask, r.tg.y != 0u), N = __ballot_sync(smask, can_node);
            if ((P | N) == 0u) break;
            if (N == 0u || (uint32_t) __popc(P) >= w.tri_thresh) {
                if (r.tg.y) {
                    const int i = __ffs(r.tg.y) - 1;
                    r.tg.y &= r.tg.y - 1;
                    const float4 *tp = sc.tris + (size_t) (r.tg.x + i) * 3;
                    const float4 r0 = __ldg(tp), r1 = __ldg(tp + 1), r2 = __ldg(tp + 2);
                    cnt.add(C_TRIS);
                    // Mesh::moeller_trumbore (mesh.h:469-492)
                    const f3 e1(r1.x, r1.y, r1.z), e2(r2.x, r2.y, r2.z);
                    const f3 pvec = cross(r.d, e2);
                    const float inv_det = rcp(dot(e1, pvec));
                    const f3 tvec = r.o - f3(r0.x, r0.y, r0.z);
                    const float u = dot(tvec, pvec) * inv_det;
                    bool ok = u >= 0.f && u <= 1.f;
                    const f3 qvec = cross(tvec, e1);
                    const float v = dot(r.d, qvec) * inv_det;
                    ok = ok && v >= 0.f && u + v <= 1.f;
                    const float t = dot(e2, qvec) * inv_det;
                    ok = ok && t >= 0.f && t <= r.maxt;
                    if (ok) {
                        const uint32_t face = __float_as_uint(r0.w);
                        if (r.any) { r.found = true; r.tg.y = 0u; r.ng.y = 0u; r.sp = 0; r.nq = 0; }
                        else if (!r.found || t < r.t || (t == r.t && face < r.face)) {
                            r.found = true; r.t = t; r.u = u; r.v = v; r.face = face; r.mesh = __float_as_uint(r1.w);
                        }
                    }
                    if (r.tg.y == 0u && r.nq > 0) r.tg = s_pend[--r.nq][tid];
                }
            } else if (can_node) {
                if ((r.ng.y & 0xff000000u) == 0u) { --r.sp; r.ng = r.sp < kSmemStack ? s_stack[r.sp][tid] : l_stack[r.sp - kSmemStack]; }
                const int bit = 31 - __clz(r.ng.y);
                r.ng.y &= ~(1u << bit);
                if (r.ng.y & 0xff000000u) {
                    if (r.sp < kSmemStack) s_stack[r.sp][tid] = r.ng;
                    else if (r.sp < kSmemStack + kLocalStack) l_stack[r.sp - kSmemStack] = r.ng;
                    if (r.sp < kSmemStack + kLocalStack) r.sp++;
                }
                const uint32_t cslot = (uint32_t) (bit - 24) ^ r.octinv;
                const uint32_t rel = __popc(r.ng.y & 0xffu & ~(0xffffffffu << cslot));
                const uint4 *np = sc.nodes + (size_t) (r.ng.x + rel) * 5;
                const uint4 w0 = __ldg(np), w1 = __ldg(np + 1), w2 = __ldg(np + 2), w3 = __ldg(np + 3), w4 = __ldg(np + 4);
                cnt.add(C_NODES);
                const uint32_t imask = w0.w >> 24;
                const uint32_t hitmask = node_slab_hits(w0, w1, w2, w3, w4, r.o, r.idx, r.idy, r.idz, r.octinv, r.t, sc.qmagic);
                r.ng = make_uint2(w1.x, (permute_octant(hitmask, r.octinv) << 24) | imask);
                const uint32_t thits = hitmask & 0x00ffffffu;
                if (thits) {
                    // the parked triangles are tested a few micro-steps from now: start their L2 -> L1 transfer here
                    // (ncu r01_v4: the first use of the triangle words was the largest long-scoreboard stall, 14 % of samples)
                    if (w.prefetch & 1u) prefetch_l1(sc.tris + (size_t) (w1.y + (uint32_t) (__ffs(thits) - 1)) * 3);
                    if (r.tg.y == 0u) r.tg = make_uint2(w1.y, thits);
                    else s_pend[r.nq++][tid] = make_uint2(w1.y, thits);
                }
            }
        }
        done = (r.ng.y & 0xff000000u) == 0u && r.sp == 0 && r.tg.y == 0u && r.nq == 0;

        if (done) {
            if (r.any) {
                if (!r.found) {         // the emitter sample arrives: L <- fma(a, b, L)
                    if (r.tex) cnt.add(C_TEX);
                    if (w.vis_bit != 0xffffffffu) st.vis(r.slot) = (uint32_t) st.vis(r.slot) | (1u << w.vis_bit);
                    const float4 a = st.sh_a(r.slot), b = st.sh_b(r.slot); float4 L = st.rad(r.slot);
                    L.x = fma_(a.x, b.x, L.x); L.y = fma_(a.y, b.y, L.y); L.z = fma_(a.z, b.z, L.z);
                    st.rad(r.slot) = L;
                }
            } else {
                st.hit(r.slot) = make_float4(r.t, r.u, r.v, __uint_as_float(r.face));
                st.hit_mesh(r.slot) = r.found ? r.mesh : 0xffffffffu;
                st.cls_q[r.qi] = r.found ? (uint8_t) __ldg(&sc.meshes[r.mesh].cls) : (uint8_t) CLS_MISS;   // for classify(it)
            }
            have = false;
        }
    }
    wf_flush_counters(cnt, counters);
}

// ------------------------------------------------------------------------------------------ shade
// ------------------------------------------------------------------------------------------ classify
// Sorts the closest hits of iteration `it` by shading class (miss / diffuse / textured diffuse / rough conductor /
// dielectric) into per-class queues, so that every warp of a shade kernel runs one BSDF and the kernels can be
// specialised: block-wide counting sort step = five ballots, one atomic per block and class.
__global__ void __launch_bounds__(256) k_wf_classify(const __grid_constant__ WfLaunch w) {
    __shared__ uint32_t s_cnt[CLS_COUNT][8], s_base[CLS_COUNT];
    const WfState &st = w.st;
    const uint32_t it = w.it, n = st.ctl->n_closest[it & 3u];
    const uint32_t *qin = st.q[it & 1u];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (blockIdx.x == 0 && tid < 8) {
        const uint32_t z = (it + 2u) & 3u;
        if (tid == 0) { st.ctl->n_closest[z] = 0u; st.ctl->n_shadow[z] = 0u; st.ctl->fetch[z] = 0u; }
        st.ctl->n_class[z][tid] = 0u;
    }
    for (uint32_t base = blockIdx.x * 256u; base < n; base += gridDim.x * 256u) {
        const uint32_t i = base + tid;
        uint32_t slot = 0; int cls = -1;
        if (i < n) { slot = qin[i]; cls = (int) st.cls_q[i]; }       // class by queue position, written by trace(it)
        uint32_t mine = 0;
#pragma unroll
        for (int k = 0; k < CLS_COUNT; ++k) {
            const uint32_t m = __ballot_sync(0xffffffffu, cls == k);
            if (lane == 0) s_cnt[k][warp] = __popc(m);
            if (cls == k) mine = m;
        }
        __syncthreads();
        if (tid < CLS_COUNT) {
            uint32_t tot = 0;
            for (int k = 0; k < 8; ++k) { const uint32_t c = s_cnt[tid][k]; s_cnt[tid][k] = tot; tot += c; }
            s_base[tid] = tot ? atomicAdd(&st.ctl->n_class[it & 3u][tid], tot) : 0u;
        }
        __syncthreads();
        if (cls >= 0) st.qc[cls][s_base[cls] + s_cnt[cls][warp] + __popc(mine & ((1u << lane) - 1u))] = slot;
        __syncthreads();
    }
}

// SORTED: the input is the class queue qc[CLS] filled by k_wf_classify; otherwise the whole live queue (CLS then only
// states what every BSDF of the scene is -- single-class scenes such as the Cornell box skip the sort).
template <int V, bool COUNT, bool ADJ = false, int CLS = CLS_GENERIC, bool SORTED = (CLS != CLS_GENERIC)>
__global__ void __launch_bounds__(256, (CLS == CLS_DIFFUSE || CLS == CLS_MISS) ? 4 : B2_SHADE_MINBLOCKS) k_wf_shade(const __grid_constant__ RenderParams p, const __grid_constant__ WfLaunch w) {
    __shared__ uint32_t s_cnt[2][8], s_base[2];
    __shared__ float s_gconst[kMaxGradTargets * 3];     // adjoint: per-block sums of the constant-albedo gradients
    const WfState &st = w.st;
    const uint32_t it = w.it, n = !SORTED ? st.ctl->n_closest[it & 3u] : st.ctl->n_class[it & 3u][SORTED ? CLS : 0];
    const uint32_t *qin = !SORTED ? st.q[it & 1u] : st.qc[SORTED ? CLS : 0];
    uint32_t *qout = st.q[(it + 1u) & 1u], *qsh = st.q_shadow[0];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    Counters<COUNT> cnt; GradSink gs{ s_gconst };
    if (ADJ) { if (tid < kMaxGradTargets * 3) s_gconst[tid] = 0.f; __syncthreads(); }
    if (!SORTED && blockIdx.x == 0 && tid == 0) { const uint32_t z = (it + 2u) & 3u; st.ctl->n_closest[z] = 0u; st.ctl->n_shadow[z] = 0u; st.ctl->fetch[z] = 0u; }

    for (uint32_t base = blockIdx.x * 256u; base < n; base += gridDim.x * 256u) {
        const uint32_t i = base + tid;
        bool alive = false, shadow = false; uint32_t slot = 0;
        if (i < n) {
            slot = qin[i];
            Path s; wf_load_path(st, slot, s);
            const float4 hh = st.hit(slot);
            Hit h; h.t = hh.x; h.u = hh.y; h.v = hh.z; h.face = __float_as_uint(hh.w); h.mesh = st.hit_mesh(slot);
            ShadowRec so; so.valid = false; so.tex = false; so.vis = false; so.a = f3(0.f); so.b = f3(0.f);
            if (ADJ) {
                const float4 g = st.dl[slot]; s.dL = f3(g.x, g.y, g.z); s.adjoint = true;
                so.vis = ((st.vis(slot) >> w.vis_bit) & 1u) != 0u;
            }
            alive = path_step<V, ADJ, COUNT, true, CLS>(p, s, gs, cnt, &h, &so);
            if (ADJ && !alive) cnt.add(C_SAMPLES);
            if (ADJ && B2_FWD(p)) st.dl[slot] = make_float4(s.dL.x, s.dL.y, s.dL.z, 0.f);
            wf_store_path(st, slot, s, alive);
            shadow = so.valid;
            if (shadow) {
                st.sh_o(slot) = make_float4(so.ray.o.x, so.ray.o.y, so.ray.o.z, so.ray.maxt);
                st.sh_d(slot) = make_float4(so.ray.d.x, so.ray.d.y, so.ray.d.z, so.tex ? 1.f : 0.f);
                st.sh_a(slot) = make_float4(so.a.x, so.a.y, so.a.z, 0.f);
                st.sh_b(slot) = make_float4(so.b.x, so.b.y, so.b.z, 0.f);
            }
        }
        // ---- compaction: warp ballot -> per-warp counts -> one atomic per block and queue (keeps the queue in lane
        // order within a block, which the next trace launch rewards with coherent rays; per-warp atomics measured 10 % slower)
        const uint32_t ma = __ballot_sync(0xffffffffu, alive), ms = __ballot_sync(0xffffffffu, shadow);
        if (lane == 0) { s_cnt[0][warp] = __popc(ma); s_cnt[1][warp] = __popc(ms); }
        __syncthreads();
        if (tid < 2) {
            uint32_t tot = 0;
            for (int k = 0; k < 8; ++k) { const uint32_t c = s_cnt[tid][k]; s_cnt[tid][k] = tot; tot += c; }
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
// pass, reseeds its sampler to replay the identical stream, and gathers dL = sum_p grad_in[p] * w(p) / W[p], the
// adjoint of splat + develop w.r.t. the splatted value.
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
        const float4 l = st.rad(i);
        s.rng.seed_lane(p.seed, global_lane(p, s.lane));
        start_path<B2_VARIANT_PRB>(p, s);
        s.L = f3(l.x, l.y, l.z);
        f3 dL(0.f);
        if (!B2_FWD(p)) film_taps(p, s.pos, [&](int x, int y, float wgt) {
            const size_t pix = (size_t) y * sn.crop_width + x;
            float W = __ldg(p.wfilm + 4 * pix + 3); W = W == 0.f ? 1.f : W;
            const float k = wgt / W;
            dL += f3(__ldg(p.grad_in + 3 * pix) * k, __ldg(p.grad_in + 3 * pix + 1) * k, __ldg(p.grad_in + 3 * pix + 2) * k);
        });
        wf_store_path(st, i, s, true);
        st.dl[i] = make_float4(dL.x, dL.y, dL.z, 0.f);
        st.hit(i) = make_float4(s.ray.maxt, 0.f, 0.f, __uint_as_float(0xffffffffu));
        st.hit_mesh(i) = 0xffffffffu;
        q[i] = i;
    }
}

// forward mode: the accumulated d L of every lane becomes the value the splat kernel reads
__global__ void __launch_bounds__(256) k_wf_dl_to_rad(const __grid_constant__ WfLaunch w) {
    const WfState &st = w.st;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < w.n; i += gridDim.x * blockDim.x) {
        const float4 g = st.dl[i]; float4 l = st.rad(i);
        st.rad(i) = make_float4(g.x, g.y, g.z, l.w);
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
                    const float2 q = st.pos(slot); ps = f2{ q.x, q.y };
                    const float4 l = st.rad(slot); L = f3(l.x, l.y, l.z);
                    if (V == B2_VARIANT_PATH_MIS && !(__float_as_uint(l.w) & WF_FLAG_VALID_RAY)) L = f3(0.f);
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
            uint32_t taps;
            if (fast) taps = film_taps_group<WEIGHTS_ONLY>(p, ps, L, valid, g, lid == lead, add);
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
    return align_up(c * 32, 256) * 6 + align_up(c * 16, 256) + align_up(c, 256) + align_up(c * 4, 256) * (3 + CLS_COUNT) + 256;
}

void wf_bind(WfState &st, void *base, uint32_t cap) {
    char *q = (char *) base; const size_t c = cap;
    auto take = [&](size_t bytes) { void *r = q; q += align_up(bytes, 256); return r; };
    st.ctl = (WfCtl *) take(256);
    st.p0 = (float4 *) take(c * 32); st.p1 = (float4 *) take(c * 32); st.p2 = (float4 *) take(c * 32); st.p3 = (float4 *) take(c * 32);
    st.s0 = (float4 *) take(c * 32); st.s1 = (float4 *) take(c * 32);
    st.dl = (float4 *) take(c * 16); st.cls_q = (uint8_t *) take(c);
    st.q[0] = (uint32_t *) take(c * 4); st.q[1] = (uint32_t *) take(c * 4); st.q_shadow[0] = (uint32_t *) take(c * 4);
    st.q_shadow[1] = st.q_shadow[0];
    for (int k = 0; k < CLS_COUNT; ++k) st.qc[k] = (uint32_t *) take(c * 4);
    st.capacity = cap;
}

// optional per-kernel-kind device timing (bench.py's roofline): CUDA events around every launch, summed after the render
static thread_local WfTiming *g_tm = nullptr;
void wf_set_timing(WfTiming *t) { g_tm = t; }
static void tm_begin(int kind, cudaStream_t st) {
    WfTiming *t = g_tm; if (!t) return;
    if (t->used + 2 > t->ev.size()) { cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b); t->ev.push_back(a); t->ev.push_back(b); }
    t->kind.push_back(kind);
    cudaEventRecord(t->ev[t->used], st);
}
static void tm_end(cudaStream_t st) { WfTiming *t = g_tm; if (!t) return; cudaEventRecord(t->ev[t->used + 1], st); t->used += 2; }
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
static const uint32_t tune_slice = (uint32_t) env_int("B2_TRACE_SLICE", B2_TRACE_SLICE);
static const uint32_t tune_tri = (uint32_t) env_int("B2_TRI_THRESH", 10);
static const uint32_t tune_prefetch = (uint32_t) env_int("B2_PREFETCH", 1);
static const int tune_trace_blocks = getenv("B2_TRACE_BLOCKS") ? atoi(getenv("B2_TRACE_BLOCKS")) : 0;
static const int tune_sort = getenv("B2_SHADE_SORT") ? atoi(getenv("B2_SHADE_SORT")) : -1;   // 0/1 force, -1 = by scene content

// resident blocks per SM of one kernel instantiation (queried once): every shade specialisation gets the grid its own
// register count allows (the diffuse kernels fit 4 blocks/SM, the glossy ones 3)
template <auto Kern> static int occ_of() {
    static const int occ = [] { int o = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, Kern, 256, 0); return std::max(o, 1); }();
    return occ;
}
static const int tune_shade_mult = getenv("B2_SHADE_GRID_MULT") ? std::max(1, atoi(getenv("B2_SHADE_GRID_MULT"))) : 1;
#define B2_SHADE_LAUNCH(...) do { constexpr auto kern_ = k_wf_shade<__VA_ARGS__>; \
    const int g_ = (int) std::min<uint64_t>((n + 255) / 256, (uint64_t) occ_of<kern_>() * sm_count * tune_shade_mult); \
    tm_begin(WF_K_SHADE, stream); kern_<<<g_, 256, 0, stream>>>(p, w); tm_end(stream); launches++; } while (0)

template <int V, bool COUNT>
static long wf_render_t(RenderParams p, WfState &st, int sm_count, cudaStream_t stream, const char **err) {
    long launches = 0;
    const DSensor &sn = p.sc.sensor;
    const uint32_t W = (uint32_t) sn.grid_w, P = p.unit_pixels;
    int occ_trace = 0, occ_shade = 0, occ_splat = 0;
    const size_t smem_splat = sizeof(float4) * (size_t) p.tile_cols * p.tile_rows;
    cudaFuncSetAttribute(k_wf_splat<V, COUNT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem_splat);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_trace, k_wf_trace<COUNT>, kTraceThreads, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_shade, k_wf_shade<V, COUNT>, 256, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_splat, k_wf_splat<V, COUNT, false>, 256, smem_splat);
    if (occ_trace <= 0 || occ_shade <= 0 || occ_splat <= 0) { *err = "wavefront: a kernel cannot be resident (occupancy 0)"; return -1; }
    const uint64_t lane_lo = p.lane_begin, lane_hi = p.lane_end;
    const bool skip_first_trace = V == B2_VARIANT_PATH && p.max_depth <= 1;
    const bool bounded = p.max_depth <= 64;
    const uint32_t n_iter = bounded ? (uint32_t) std::max(p.max_depth, 1) : 0u;
    uint32_t it = 0;
    for (uint64_t c0 = lane_lo; c0 < lane_hi; c0 += st.capacity) {
        const uint32_t n = (uint32_t) std::min<uint64_t>(st.capacity, lane_hi - c0);
        WfLaunch w; w.st = st; w.chunk_begin = c0; w.n = n; w.slice = tune_slice; w.tri_thresh = tune_tri; w.prefetch = tune_prefetch; w.vis_bit = 0xffffffffu;
        const int g_gen = (int) std::min<uint64_t>((n + 255) / 256, (uint64_t) sm_count * 8);
        const int g_trace = (tune_trace_blocks > 0 ? std::min(tune_trace_blocks, occ_trace) : occ_trace) * sm_count, g_shade = (int) std::min<uint64_t>((n + 255) / 256, (uint64_t) occ_shade * sm_count);
        for (uint32_t pass = 0; pass < p.n_passes; ++pass) {
            w.pass = pass; w.it = it;
            tm_begin(WF_K_OTHER, stream); k_wf_generate<V><<<g_gen, 256, 0, stream>>>(p, w); tm_end(stream); launches++;
            for (uint32_t k = 0;; ++k) {
                if (bounded && k >= n_iter) break;
                w.it = it;
                if (!(k == 0 && skip_first_trace)) { tm_begin(WF_K_TRACE, stream); k_wf_trace<COUNT><<<g_trace, kTraceThreads, 0, stream>>>(p.sc, w, p.counters); tm_end(stream); launches++; }
                const uint32_t cmask = st.class_mask;
                const bool sort = tune_sort >= 0 ? tune_sort != 0     // default: sort when the BSDFs differ enough to diverge
                                                 : (__builtin_popcount(cmask) >= 2 && (cmask & ((1u << CLS_CONDUCTOR) | (1u << CLS_DIELECTRIC))) != 0);
                if (!sort && cmask == (1u << CLS_DIFFUSE)) B2_SHADE_LAUNCH(V, COUNT, false, CLS_DIFFUSE, false);
                else if (sort) {
                    tm_begin(WF_K_SHADE, stream); k_wf_classify<<<g_shade, 256, 0, stream>>>(w); tm_end(stream); launches++;
                    const uint32_t cm = st.class_mask;
                    if (cm & (1u << CLS_DIFFUSE)) B2_SHADE_LAUNCH(V, COUNT, false, CLS_DIFFUSE);
                    if (cm & (1u << CLS_DIFFUSE_TEX)) B2_SHADE_LAUNCH(V, COUNT, false, CLS_DIFFUSE_TEX);
                    if (cm & (1u << CLS_CONDUCTOR)) B2_SHADE_LAUNCH(V, COUNT, false, CLS_CONDUCTOR);
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
            w.it = it;                                  // shadow rays of the last bounce
            tm_begin(WF_K_TRACE, stream); k_wf_trace<COUNT><<<g_trace, kTraceThreads, 0, stream>>>(p.sc, w, p.counters); tm_end(stream); launches++;
            ++it;
            // splat this chunk: units covering its pixels
            RenderParams ps = p;
            const uint64_t pix0 = c0 / p.spp_per_pass, pix1 = (c0 + n - 1) / p.spp_per_pass;
            ps.unit_begin = (uint32_t) ((pix0 / W) * p.units_per_row + (pix0 % W) / P);
            ps.unit_end = (uint32_t) ((pix1 / W) * p.units_per_row + (pix1 % W) / P) + 1;
            cudaMemsetAsync(p.unit_counter, 0, 4, stream);
            const int g_splat = (int) std::min<uint32_t>(ps.unit_end - ps.unit_begin, (uint32_t) (occ_splat * sm_count));
            tm_begin(WF_K_SPLAT, stream); k_wf_splat<V, COUNT, false><<<g_splat, 256, smem_splat, stream>>>(ps, w); tm_end(stream); launches++;
        }
    }
    if (cudaGetLastError() != cudaSuccess) { *err = "wavefront: kernel launch failed"; return -1; }
    return launches;
}

template <bool COUNT>
static long wf_render_backward_t(RenderParams p, WfState &st, int sm_count, cudaStream_t stream, const char **err) {
    constexpr int V = B2_VARIANT_PRB;
    long launches = 0;
    int occ_trace = 0, occ_shade = 0, occ_adj = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_trace, k_wf_trace<COUNT>, kTraceThreads, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_shade, k_wf_shade<V, COUNT, false>, 256, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_adj, k_wf_shade<V, COUNT, true>, 256, 0);
    int occ_splat = 1; const size_t smem_splat = sizeof(float4) * (size_t) p.tile_cols * p.tile_rows;
    if (B2_FWD(p)) {
        cudaFuncSetAttribute(k_wf_splat<V, COUNT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem_splat);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_splat, k_wf_splat<V, COUNT, false>, 256, smem_splat);
        if (occ_splat <= 0) { *err = "wavefront: the splat kernel cannot be resident"; return -1; }
    }
    if (occ_trace <= 0 || occ_shade <= 0 || occ_adj <= 0) { *err = "wavefront: a kernel cannot be resident (occupancy 0)"; return -1; }
    if (p.max_depth > 32) { *err = "wavefront render_backward: max_depth > 32"; return -1; }
    const uint32_t n_iter = (uint32_t) std::max(p.max_depth, 1);
    uint32_t it = 0;
    for (uint64_t c0 = p.lane_begin; c0 < p.lane_end; c0 += st.capacity) {
        const uint32_t n = (uint32_t) std::min<uint64_t>(st.capacity, p.lane_end - c0);
        WfLaunch w; w.st = st; w.chunk_begin = c0; w.n = n; w.slice = tune_slice; w.tri_thresh = tune_tri; w.prefetch = tune_prefetch; w.pass = 0;
        const int g_gen = (int) std::min<uint64_t>((n + 255) / 256, (uint64_t) sm_count * 8);
        const int g_trace = (tune_trace_blocks > 0 ? std::min(tune_trace_blocks, occ_trace) : occ_trace) * sm_count;
        const int g_shade = (int) std::min<uint64_t>((n + 255) / 256, (uint64_t) occ_shade * sm_count);
        const int g_adj = (int) std::min<uint64_t>((n + 255) / 256, (uint64_t) occ_adj * sm_count);
        // ---- primal pass (common.py:750-760): L per lane, and the visibility bit of every emitter sample
        w.it = it; w.vis_bit = 0xffffffffu;
        tm_begin(WF_K_OTHER, stream); k_wf_generate<V><<<g_gen, 256, 0, stream>>>(p, w); tm_end(stream); launches++;
        for (uint32_t k = 0; k < n_iter; ++k) {
            w.it = it; w.vis_bit = k ? k - 1u : 0xffffffffu;
            tm_begin(WF_K_TRACE, stream); k_wf_trace<COUNT><<<g_trace, kTraceThreads, 0, stream>>>(p.sc, w, p.counters); tm_end(stream); launches++;
            B2_SHADE_LAUNCH(V, COUNT, false);
            ++it;
        }
        w.it = it; w.vis_bit = n_iter - 1u;
        tm_begin(WF_K_TRACE, stream); k_wf_trace<COUNT><<<g_trace, kTraceThreads, 0, stream>>>(p.sc, w, p.counters); tm_end(stream); launches++;
        ++it;
        // ---- adjoint pass (common.py:763-773): replay of the identical stream, closest-hit rays only
        w.it = it; w.vis_bit = 0xffffffffu;
        tm_begin(WF_K_OTHER, stream); k_wf_generate_adj<<<g_gen, 256, 0, stream>>>(p, w); tm_end(stream); launches++;
        for (uint32_t k = 0; k < n_iter; ++k) {
            w.it = it; w.vis_bit = 0xffffffffu;
            tm_begin(WF_K_TRACE, stream); k_wf_trace<COUNT><<<g_trace, kTraceThreads, 0, stream>>>(p.sc, w, p.counters); tm_end(stream); launches++;
            w.vis_bit = k;
            tm_begin(WF_K_SHADE, stream); k_wf_shade<V, COUNT, true><<<g_adj, 256, 0, stream>>>(p, w); tm_end(stream); launches++;
            ++it;
        }
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

long wf_render_backward(bool count, RenderParams p, WfState &st, int sm_count, cudaStream_t stream, const char **err) {
    return count ? wf_render_backward_t<true>(p, st, sm_count, stream, err) : wf_render_backward_t<false>(p, st, sm_count, stream, err);
}

long wf_render(int variant, bool count, RenderParams p, WfState &st, int sm_count, cudaStream_t stream, const char **err) {
    switch (variant) {
        case B2_VARIANT_PATH: return count ? wf_render_t<B2_VARIANT_PATH, true>(p, st, sm_count, stream, err) : wf_render_t<B2_VARIANT_PATH, false>(p, st, sm_count, stream, err);
        case B2_VARIANT_PATH_MIS: return count ? wf_render_t<B2_VARIANT_PATH_MIS, true>(p, st, sm_count, stream, err) : wf_render_t<B2_VARIANT_PATH_MIS, false>(p, st, sm_count, stream, err);
        default: return count ? wf_render_t<B2_VARIANT_PRB, true>(p, st, sm_count, stream, err) : wf_render_t<B2_VARIANT_PRB, false>(p, st, sm_count, stream, err);
    }
}

}  // namespace b2
