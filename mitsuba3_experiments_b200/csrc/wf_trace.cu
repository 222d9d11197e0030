// wf_trace.cu -- Scene::ray_intersect_preliminary + Scene::ray_test of the wavefront (scene.cpp:191-235; Embree / OptiX in the
// reference, scene_embree.inl:245-433, scene_optix.inl:602-743).  Three kernels, same queues in and same records out:
//   k_wf_trace          the general one (below): incoherent rays of bounces >= 1, closest and shadow rays in one launch
//   k_wf_trace_packet   bounce 0 from 4 samples per pixel on: one warp-uniform traversal per 32 / 64 consecutive primary rays
//   k_wf_trace_simple   trees of a few nodes (the Cornell box): one thread per ray, the per-thread traversal of b2_bvh.cuh
// All three return bit-identical hit records: the closest hit of a ray is the lexicographic minimum of (t, face) over the
// triangles it intersects (same Moeller-Trumbore arithmetic everywhere), whatever the traversal order.
//
// k_wf_trace:
// Persistent warps.  Each warp owns a three-stage ray pipeline so that a lane whose ray ends can be refilled within a
// few instructions instead of a dependent global-load chain:
//   stage A   lane 0 reserves the next 32 queue entries (atomicAdd on the launch's fetch counter)
//   stage B   the 32 lane slots of that batch are loaded from the queue (closest queue first, then the shadow queue)
//   stage C   the 32 rays are copied global -> shared with cp.async (bounce 0: computed from the lane index, there is
//             no generate kernel: sampler seed, jitter, sensor ray -- integrator.cpp:293-310,419-436, sampler.cpp:129-148,
//             perspective.cpp:239-280)
// Every stage runs one batch ahead of the next, so the latencies of the atomic, of the queue read and of the ray read
// overlap traversal; idle lanes take their next ray from shared memory as soon as `refill_thresh` of them are idle.
//
// Traversal: vote-scheduled micro-steps.  Per micro-step the warp does either a NODE step (pop a child, five 128-bit
// loads, eight branch-free slab tests, b2_bvh.cuh) for the lanes that have node work, or a TRIANGLE step (one
// Moeller-Trumbore test per lane, mesh.h:469-492) once `tri_thresh` lanes hold parked leaf triangles; leaf hits found by a
// node step are parked in a per-lane queue in shared memory.  The traversal stack lives in shared memory ([depth][thread],
// conflict free; entries beyond kSmemStack spill to local memory, which the depth guard of b2_scene_build_accel bounds
// at kStackSize).  Per-ray values that are touched once per ray (lane slot, queue index, u, v, face, mesh) live in
// shared memory, which keeps the kernel at 64 registers without spills.
//
// Epilogue = stores only, done at the next refill by all finished lanes of the warp together: a closest ray writes its 32-byte hit
// record with one 256-bit store (+ the class byte for the counting sort); a shadow ray that arrives sets the `arrived` word of the
// lane's deferred-NEE record.  The reciprocal directions of a staged batch are computed once, by all 32 lanes, when it becomes current.
#include "b2_path.cuh"
#include "b2_wavefront.cuh"

namespace b2 {

constexpr int kTraceThreads = 256;
constexpr int kTraceWarps = kTraceThreads / 32;
constexpr int kSmemStack = 8;                       // stack entries per thread held in shared memory
constexpr int kLocalStack = kStackSize - kSmemStack; // overflow entries in local memory (never reached on the benchmark trees)
constexpr int kTriQueue = 4;                        // parked leaf-triangle groups per thread (shared memory)
#ifndef B2_TRACE_MINBLOCKS
#define B2_TRACE_MINBLOCKS 4
#endif

// lane flag word
enum : uint32_t { TF_OCT = 7u, TF_ANY = 8u, TF_TEX = 16u, TF_FOUND = 32u, TF_NQ_SHIFT = 8, TF_NQ_ONE = 1u << 8, TF_NQ_MASK = 7u << 8, TF_DONE = 1u << 30, TF_HAVE = 1u << 31 };

B2_DEV void cp_async16(void *smem, const void *gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"((uint32_t) __cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
B2_DEV void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
B2_DEV void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

template <bool COUNT>
B2_DEV void trace_flush_counters(const Counters<COUNT> &cnt, unsigned long long *out) {
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

template <bool COUNT, bool PRIMARY>
__global__ void __launch_bounds__(kTraceThreads, B2_TRACE_MINBLOCKS) k_wf_trace(const __grid_constant__ RenderParams p, const __grid_constant__ WfLaunch w) {
    __shared__ uint2 s_stack[kSmemStack][kTraceThreads];
    __shared__ uint2 s_pend[kTriQueue][kTraceThreads];
    __shared__ float4 s_so[kTraceWarps][2][32], s_sd[kTraceWarps][2][32];     // staged rays: o | maxt, d | flag word
    __shared__ uint32_t s_ss[kTraceWarps][2][32];                             // staged lane slot | any << 31
    extern __shared__ float4 s_inv_dyn[];                                     // [kTraceWarps][32] current batch: 1 / d (x, y, z), octant -- computed by all 32 lanes at once
    float4 (*s_inv)[32] = (float4 (*)[32]) s_inv_dyn;                         // (dynamic: the static arrays above fill the 48 KB static limit)
    __shared__ float2 s_uv[kTraceThreads];                                    // closest hit so far: u, v
    __shared__ uint2 s_fm[kTraceThreads], s_id[kTraceThreads];                // face, mesh | class << 28 ; lane slot, queue index
    const DScene &sc = p.sc; const WfState &st = w.st;
    const uint32_t nc = PRIMARY ? w.n : st.ctl->n_closest[w.it & 3u], ns = PRIMARY ? 0u : st.ctl->n_shadow[(w.it + 3u) & 3u], total = nc + ns;
    const uint32_t *qc = st.q[w.it & 1u], *qs = st.q_shadow;
    uint32_t *fetch = &st.ctl->fetch[w.it & 3u];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u, full = 0xffffffffu;
    Counters<COUNT> cnt;
    uint2 l_stack[kLocalStack];

    // ---- ray pipeline (all warp-uniform except pend_a / slot_b)
    uint32_t pend_a = 0;                    // lane 0: result of the stage-A atomic, not yet broadcast
    uint32_t base_b = 0, nb_b = 0, slot_b = 0;   // batch whose lane slots are (being) loaded into slot_b
    uint32_t base_n = 0, cnt_n = 0;         // batch staged in buffer cur ^ 1 (rays in flight)
    uint32_t base_c = 0, cnt_c = 0, cons = 0, cur = 1;   // batch being consumed from buffer cur
    bool exhausted = total == 0;

    auto stage_a = [&]() { if (lane == 0) pend_a = atomicAdd(fetch, 32u); };
    auto stage_b = [&]() {               // consumes pend_a
        base_b = __shfl_sync(full, pend_a, 0);
        nb_b = base_b < total ? min(32u, total - base_b) : 0u;
        if ((uint32_t) lane < nb_b) {
            const uint32_t e = base_b + lane;
            if (PRIMARY) slot_b = e;
            else slot_b = e >= nc ? (qs[e - nc] | 0x80000000u) : qc[e];
        }
    };
    auto stage_c = [&]() {               // consumes slot_b into buffer cur ^ 1
        base_n = base_b; cnt_n = nb_b;
        if ((uint32_t) lane < nb_b) {
            const uint32_t nb = cur ^ 1u;
            if (PRIMARY) {
                Pcg32 rng;
                const uint32_t ln = (uint32_t) (w.chunk_begin + slot_b);
                if (w.pass == 0) rng.seed_lane(p.seed, global_lane(p, ln));
                else { const float4 r = __ldg(&st.rng[slot_b].a); rng.state = (uint64_t) __float_as_uint(r.x) | ((uint64_t) __float_as_uint(r.y) << 32); rng.inc = (uint64_t) __float_as_uint(r.z) | ((uint64_t) __float_as_uint(r.w) << 32); }
                const Ray ray = primary_ray(p, ln, rng);
                s_so[warp][nb][lane] = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.maxt);
                s_sd[warp][nb][lane] = make_float4(ray.d.x, ray.d.y, ray.d.z, 0.f);
            } else {
                const Elem *src = (slot_b >> 31) ? &st.shr[slot_b & 0x7fffffffu] : &st.ray[slot_b];
                cp_async16(&s_so[warp][nb][lane], &src->a);
                cp_async16(&s_sd[warp][nb][lane], &src->b);
            }
            s_ss[warp][nb][lane] = slot_b;
        }
        if (!PRIMARY) cp_async_commit();
    };
    // prime: batch 0 staged, batch 1 slots loading, batch 2 reserved
    if (!exhausted) { stage_a(); stage_b(); stage_a(); stage_c(); stage_b(); stage_a(); }

    f3 o, d; float idx = 0.f, idy = 0.f, idz = 0.f, tbest = 0.f;
    uint2 ng = make_uint2(0u, 0u), tg = make_uint2(0u, 0u);
    int sp = 0; uint32_t fl = 0;

    for (;;) {
        // ---- refill idle lanes from the staged rays (warp converged here)
        const uint32_t idle = __ballot_sync(full, !(fl & TF_HAVE));
        if (idle) {
            const uint32_t k = __popc(idle);
            if (k >= w.refill_thresh || idle == full || exhausted) {
                // ---- epilogue of the finished rays (stores only), for all finished lanes of the warp at once
                if (fl & TF_DONE) {
                    const uint2 id = s_id[tid];
                    if (fl & TF_ANY) {
                        if (!(fl & TF_FOUND)) {         // the emitter sample arrives: the lane's next shade (or the splat) adds fma(a, b, L)
                            if (fl & TF_TEX) cnt.add(C_TEX);
                            st.shf[id.x].a.w = 1.f;
                        }
                    } else {
                        const uint2 fm = s_fm[tid]; const float2 uv = s_uv[tid];
                        const bool found = fm.x != 0xffffffffu;
                        const uint32_t cls = found ? (fm.y >> 28) : (uint32_t) CLS_MISS;
                        Elem h;
                        h.a = make_float4(tbest, uv.x, uv.y, __uint_as_float(fm.x));
                        h.b = make_float4(__uint_as_float(found ? (fm.y & 0x0fffffffu) : 0xffffffffu), __uint_as_float(cls), 0.f, 0.f);
                        st256(&st.hit[id.x], h);
                        st.cls_q[id.y] = (uint8_t) cls;                      // for classify(it), by queue position
                    }
                    fl = 0u;
                }
            }
            if (!exhausted && (k >= w.refill_thresh || idle == full)) {
                if (cons == cnt_c) {            // advance the pipeline: the staged batch becomes current
                    if (!PRIMARY) cp_async_wait_all();
                    __syncwarp();
                    cur ^= 1u; cons = 0; base_c = base_n; cnt_c = cnt_n;
                    {   // reciprocal directions of the whole batch with full lanes (a refill pass runs at ~6 of 32)
                        const float4 dd = s_sd[warp][cur][lane];
                        const float tiny = kTinyDir;
                        const uint32_t oi = (dd.x >= 0.f ? 1u : 0u) | (dd.y >= 0.f ? 2u : 0u) | (dd.z >= 0.f ? 4u : 0u);
                        s_inv[warp][lane] = make_float4(1.f / (fabsf(dd.x) > tiny ? dd.x : copysignf(tiny, dd.x)), 1.f / (fabsf(dd.y) > tiny ? dd.y : copysignf(tiny, dd.y)),
                                                        1.f / (fabsf(dd.z) > tiny ? dd.z : copysignf(tiny, dd.z)), __uint_as_float(oi));
                        __syncwarp();
                    }
                    stage_c(); stage_b(); stage_a();
                    if (cnt_c == 0u) exhausted = true;
                }
                const uint32_t take = min(k, cnt_c - cons), rank = __popc(idle & lt_mask);
                if (!(fl & TF_HAVE) && rank < take) {
                    const uint32_t e = cons + rank;
                    const float4 o4 = s_so[warp][cur][e], d4 = s_sd[warp][cur][e];
                    const uint32_t sl = s_ss[warp][cur][e];
                    o = f3(o4.x, o4.y, o4.z); d = f3(d4.x, d4.y, d4.z); tbest = o4.w;
                    const float4 iv = s_inv[warp][e];
                    idx = iv.x; idy = iv.y; idz = iv.z;
                    const uint32_t octinv = __float_as_uint(iv.w);
                    const bool any = !PRIMARY && (sl >> 31) != 0u;
                    fl = TF_HAVE | octinv | (any ? TF_ANY : 0u) | ((any && d4.w != 0.f) ? TF_TEX : 0u);
                    s_id[tid] = make_uint2(sl & 0x7fffffffu, base_c + e);
                    s_fm[tid] = make_uint2(0xffffffffu, 0u); s_uv[tid] = make_float2(0.f, 0.f);
                    ng = sc.n_nodes ? make_uint2(0u, (1u << (24 + octinv)) | 1u) : make_uint2(0u, 0u);
                    tg = make_uint2(0u, 0u); sp = 0;
                    cnt.add(any ? C_SHADOW : C_CLOSEST);
                }
                cons += take;
            }
        }
        const uint32_t smask = __ballot_sync(full, (fl & TF_HAVE) != 0u);
        if (smask == 0u) { if (exhausted) break; else continue; }

        if (fl & TF_HAVE) {
            // ---- one micro-step: the lanes with a ray vote for a NODE step or a TRIANGLE step
            const bool node_work = (ng.y & 0xff000000u) != 0u || sp > 0;
            const bool can_node = node_work && (tg.y == 0u || (fl & TF_NQ_MASK) < ((uint32_t) kTriQueue << TF_NQ_SHIFT));
            const uint32_t P = __ballot_sync(smask, tg.y != 0u), N = __ballot_sync(smask, can_node);
            if (N == 0u || (uint32_t) __popc(P) >= w.tri_thresh) {
                do if (tg.y) {
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
                    // tbest starts as the ray's maxt and only shrinks: t <= maxt is implied; equal t resolves to the lowest face
                    ok = ok && t >= 0.f && t <= tbest;
                    if (ok) {
                        const uint32_t face = __float_as_uint(r0.w);
                        if (fl & TF_ANY) { fl = (fl | TF_FOUND) & ~TF_NQ_MASK; tg.y = 0u; ng.y = 0u; sp = 0; }
                        else if (t < tbest || face < s_fm[tid].x) {
                            tbest = t; s_uv[tid] = make_float2(u, v); s_fm[tid] = make_uint2(face, __float_as_uint(r1.w));
                        }
                    }
                    if (tg.y == 0u && (fl & TF_NQ_MASK)) { fl -= TF_NQ_ONE; tg = s_pend[(fl & TF_NQ_MASK) >> TF_NQ_SHIFT][tid]; }
                }
                // stay in triangle steps for as long as the vote would choose one again: skips the refill check and the node vote
                // (bedroom trace 38.2 -> 37.1 ms; the 10 M soup 39.5 -> 39.9 ms)
                while ((uint32_t) __popc(__ballot_sync(smask, tg.y != 0u)) >= w.tri_thresh);
            } else if (can_node) {
                if ((ng.y & 0xff000000u) == 0u) { --sp; ng = sp < kSmemStack ? s_stack[sp][tid] : l_stack[sp - kSmemStack]; }
                const int bit = 31 - __clz(ng.y);
                ng.y &= ~(1u << bit);
                if (ng.y & 0xff000000u) {       // siblings left: push the group (depth <= kStackSize, guarded at build time)
                    if (sp < kSmemStack) s_stack[sp][tid] = ng; else l_stack[sp - kSmemStack] = ng;
                    sp++;
                }
                const uint32_t octinv = fl & TF_OCT;
                const uint32_t cslot = (uint32_t) (bit - 24) ^ octinv;
                const uint32_t rel = __popc(ng.y & 0xffu & ~(0xffffffffu << cslot));
                const uint4 *np = sc.nodes + (size_t) (ng.x + rel) * 5;
                const uint4 w0 = __ldg(np), w1 = __ldg(np + 1), w2 = __ldg(np + 2), w3 = __ldg(np + 3), w4 = __ldg(np + 4);
                cnt.add(C_NODES);
                const uint32_t imask = w0.w >> 24;
                const uint32_t hitmask = node_slab_hits(w0, w1, w2, w3, w4, o, idx, idy, idz, octinv, tbest, sc.qmagic);
                ng = make_uint2(w1.x, (permute_octant(hitmask, octinv) << 24) | imask);
                const uint32_t thits = hitmask & 0x00ffffffu;
                if (thits) {
                    if (tg.y == 0u) tg = make_uint2(w1.y, thits);
                    else { s_pend[(fl & TF_NQ_MASK) >> TF_NQ_SHIFT][tid] = make_uint2(w1.y, thits); fl += TF_NQ_ONE; }
                }
            }
            // ---- a finished ray: its stores wait for the next refill, where all finished lanes of the warp do them together
            if ((ng.y & 0xff000000u) == 0u && sp == 0 && tg.y == 0u && (fl & TF_NQ_MASK) == 0u) fl = (fl & (TF_ANY | TF_TEX | TF_FOUND)) | TF_DONE;
        }
    }
    trace_flush_counters(cnt, p.counters);
}

// ------------------------------------------------------------------------------------------ small trees: one thread per ray
// A tree of a few nodes (the Cornell box: 3 nodes, 36 triangles) gives the vote-scheduled kernel nothing to schedule: a ray is
// 1.4 node steps and 4 triangle tests, and the ray pipeline, the votes and the parked-triangle queue cost more than the traversal.
// k_wf_trace_simple walks such trees with the per-thread traversal of b2_bvh.cuh (the megakernel's), one queue entry per thread,
// same queues in, same records out.  Chosen by the host when the tree has at most kSimpleTraceNodes nodes (B2_SIMPLE_TRACE_NODES).
template <bool COUNT>
__global__ void __launch_bounds__(256) k_wf_trace_simple(const __grid_constant__ RenderParams p, const __grid_constant__ WfLaunch w) {
    const DScene &sc = p.sc; const WfState &st = w.st;
    const uint32_t nc = st.ctl->n_closest[w.it & 3u], ns = st.ctl->n_shadow[(w.it + 3u) & 3u], total = nc + ns;
    const uint32_t *qc = st.q[w.it & 1u], *qs = st.q_shadow;
    Counters<COUNT> cnt;
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
        const bool any = e >= nc;
        const uint32_t slot = any ? qs[e - nc] : qc[e];
        const Elem r = ld256(any ? &st.shr[slot] : &st.ray[slot]);
        Hit h;
        cnt.add(any ? C_SHADOW : C_CLOSEST);
        if (any) {
            if (!bvh_trace<true, COUNT>(sc, f3(r.a.x, r.a.y, r.a.z), f3(r.b.x, r.b.y, r.b.z), r.a.w, h, cnt)) {
                if (r.b.w != 0.f) cnt.add(C_TEX);
                st.shf[slot].a.w = 1.f;
            }
        } else {
            const bool found = bvh_trace<false, COUNT>(sc, f3(r.a.x, r.a.y, r.a.z), f3(r.b.x, r.b.y, r.b.z), r.a.w, h, cnt);
            const uint32_t cls = found ? h.cls : (uint32_t) CLS_MISS;
            Elem o;
            o.a = make_float4(h.t, h.u, h.v, __uint_as_float(h.face));
            o.b = make_float4(__uint_as_float(found ? h.mesh : 0xffffffffu), __uint_as_float(cls), 0.f, 0.f);
            st256(&st.hit[slot], o);
            st.cls_q[e] = (uint8_t) cls;
        }
    }
    trace_flush_counters(cnt, p.counters);
}

// ------------------------------------------------------------------------------------------ bounce 0: packet traversal
// The 32 primary rays of a warp are consecutive lanes, i.e. samples of one pixel (or of neighbouring pixels): they visit the same
// nodes.  k_wf_trace_packet walks the tree ONCE per warp: the stack and the node / triangle groups are warp-uniform, lane c < 8
// tests child c of the fetched node against the INTERVAL of the warp's rays (per axis: [min, max] of the origins and of the
// reciprocal directions, outward-rounded, boxes padded), and every lane tests every triangle of the visited leaves against its own
// ray with the same Moeller-Trumbore arithmetic as k_wf_trace.  The closest hit of a ray is the lexicographic minimum of (t, face)
// over the triangles it intersects -- independent of the traversal -- so a traversal that visits a SUPERSET of each ray's nodes
// returns bit-identical hit records.  Children are culled against the largest `tbest` of the warp.
B2_DEV int f_ord(int i) { return i ^ ((i >> 31) & 0x7fffffff); }      // order-preserving float bits <-> int (an involution)
B2_DEV float warp_min_f(float f) { return __int_as_float(f_ord(__reduce_min_sync(0xffffffffu, f_ord(__float_as_int(f))))); }
B2_DEV float warp_max_f(float f) { return __int_as_float(f_ord(__reduce_max_sync(0xffffffffu, f_ord(__float_as_int(f))))); }

constexpr int kPacketThreads = 256;
constexpr float kPkInf = 3.402823466e38f;
#ifndef B2_PACKET_MINBLOCKS
#define B2_PACKET_MINBLOCKS 4
#endif

template <bool COUNT, int R>
__global__ void __launch_bounds__(kPacketThreads, B2_PACKET_MINBLOCKS) k_wf_trace_packet(const __grid_constant__ RenderParams p, const __grid_constant__ WfLaunch w) {
    __shared__ uint2 s_pstack[kPacketThreads / 32][kStackSize];
    const DScene &sc = p.sc; const WfState &st = w.st;
    const uint32_t total = w.n, full = 0xffffffffu;
    uint32_t *fetch = &st.ctl->fetch[w.it & 3u];
    const int lane = threadIdx.x & 31;
    uint2 *stack = s_pstack[threadIdx.x >> 5];      // every lane writes the same value: no ordering needed between the lanes
    const uint32_t c = lane & 7u; const bool chi = c >= 4u;
    const uint32_t psel = 0x4505u | ((c & 3u) << 4);          // byte c & 3 of a plane word -> mantissa bits 8..15 of 32768 + q (qbyte_biased)
    const uint32_t csh = (c & 3u) * 8u;
    Counters<COUNT> cnt;
    uint32_t next = 0;
    if (lane == 0) next = atomicAdd(fetch, 32u * R);
    for (;;) {
        const uint32_t batch = __shfl_sync(full, next, 0);
        if (batch >= total) break;
        if (lane == 0) next = atomicAdd(fetch, 32u * R);
        const uint32_t batch_end = min(batch + 32u * R, total);
        // a batch is cut where an image row ends (the rays of two rows are the left and the right edge of the frustum: their interval
        // would span the whole width); each piece [base, end) is one packet
        for (uint32_t base = batch, end; base < batch_end; base = end) {
        {
            const uint64_t row_lanes = (uint64_t) p.sc.sensor.grid_w * p.spp_per_pass, l0 = w.chunk_begin + base;
            end = (uint32_t) min((uint64_t) batch_end, (l0 / row_lanes + 1u) * row_lanes - w.chunk_begin);
        }
        // ---- the lane's rays (as stage C of k_wf_trace<., true>); lanes past the end of the packet repeat its last ray and store nothing
        f3 o[R], d[R]; float tbest[R], bu[R], bv[R]; uint32_t bface[R], bmesh[R];
        float ix0 = kPkInf, ix1 = -kPkInf, iy0 = kPkInf, iy1 = -kPkInf, iz0 = kPkInf, iz1 = -kPkInf;
        float ox0 = kPkInf, ox1 = -kPkInf, oy0 = kPkInf, oy1 = -kPkInf, oz0 = kPkInf, oz1 = -kPkInf;
        float tbest_w = 0.f;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const uint32_t e = min(base + 32u * r + lane, end - 1u);
            Pcg32 rng;
            const uint32_t ln = (uint32_t) (w.chunk_begin + e);
            if (w.pass == 0) rng.seed_lane(p.seed, global_lane(p, ln));
            else { const float4 q = __ldg(&st.rng[e].a); rng.state = (uint64_t) __float_as_uint(q.x) | ((uint64_t) __float_as_uint(q.y) << 32); rng.inc = (uint64_t) __float_as_uint(q.z) | ((uint64_t) __float_as_uint(q.w) << 32); }
            const Ray ray = primary_ray(p, ln, rng);
            o[r] = ray.o; d[r] = ray.d; tbest[r] = ray.maxt; bu[r] = 0.f; bv[r] = 0.f; bface[r] = 0xffffffffu; bmesh[r] = 0u;
            const float tiny = kTinyDir;
            const float idx = 1.f / (fabsf(ray.d.x) > tiny ? ray.d.x : copysignf(tiny, ray.d.x));
            const float idy = 1.f / (fabsf(ray.d.y) > tiny ? ray.d.y : copysignf(tiny, ray.d.y));
            const float idz = 1.f / (fabsf(ray.d.z) > tiny ? ray.d.z : copysignf(tiny, ray.d.z));
            ix0 = fminf(ix0, idx); ix1 = fmaxf(ix1, idx); iy0 = fminf(iy0, idy); iy1 = fmaxf(iy1, idy); iz0 = fminf(iz0, idz); iz1 = fmaxf(iz1, idz);
            ox0 = fminf(ox0, ray.o.x); ox1 = fmaxf(ox1, ray.o.x); oy0 = fminf(oy0, ray.o.y); oy1 = fmaxf(oy1, ray.o.y); oz0 = fminf(oz0, ray.o.z); oz1 = fmaxf(oz1, ray.o.z);
            tbest_w = fmaxf(tbest_w, ray.maxt);
        }
        cnt.add(C_CLOSEST, end - base > (uint32_t) lane ? (end - base - lane + 31u) / 32u : 0u);
        // ---- intervals of the packet
        ix0 = warp_min_f(ix0); ix1 = warp_max_f(ix1); iy0 = warp_min_f(iy0); iy1 = warp_max_f(iy1); iz0 = warp_min_f(iz0); iz1 = warp_max_f(iz1);
        ox0 = warp_min_f(ox0); ox1 = warp_max_f(ox1); oy0 = warp_min_f(oy0); oy1 = warp_max_f(oy1); oz0 = warp_min_f(oz0); oz1 = warp_max_f(oz1);
        tbest_w = __uint_as_float(__reduce_max_sync(full, __float_as_uint(tbest_w)));       // maxt >= 0: the bit patterns order like the values
        // per axis: all rays travel towards + (pos), all towards - (neg), or mixed (the interval of directions contains 0).  On a mixed
        // axis the rays fan out from the origin: a box above the origins is entered no earlier than (lo - o) / (largest d), a box below
        // them no earlier than (hi - o) / (most negative d), and the axis bounds no exit.  Coded as pos with both reciprocals = 1 / largest d
        // (entry term 1) and m = 1 / most negative d (entry term 2, applied to the exit plane).
        const bool negx = ix1 < 0.f, negy = iy1 < 0.f, negz = iz1 < 0.f, mixx = !negx && !(ix0 > 0.f), mixy = !negy && !(iy0 > 0.f), mixz = !negz && !(iz0 > 0.f);
        float mx = 0.f, my = 0.f, mz = 0.f;
        const bool anymix = mixx || mixy || mixz;
        if (anymix) {
            const float tiny = kTinyDir;
            float px = kPkInf, py = kPkInf, pz = kPkInf, qx = -kPkInf, qy = -kPkInf, qz = -kPkInf;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float ax = 1.f / (fabsf(d[r].x) > tiny ? d[r].x : copysignf(tiny, d[r].x)), ay = 1.f / (fabsf(d[r].y) > tiny ? d[r].y : copysignf(tiny, d[r].y)),
                            az = 1.f / (fabsf(d[r].z) > tiny ? d[r].z : copysignf(tiny, d[r].z));
                if (ax > 0.f) px = fminf(px, ax); else qx = fmaxf(qx, ax);
                if (ay > 0.f) py = fminf(py, ay); else qy = fmaxf(qy, ay);
                if (az > 0.f) pz = fminf(pz, az); else qz = fmaxf(qz, az);
            }
            px = warp_min_f(px); py = warp_min_f(py); pz = warp_min_f(pz); qx = warp_max_f(qx); qy = warp_max_f(qy); qz = warp_max_f(qz);
            if (mixx) { ix0 = px; ix1 = px; mx = qx; }
            if (mixy) { iy0 = py; iy1 = py; my = qy; }
            if (mixz) { iz0 = pz; iz1 = pz; mz = qz; }
        }
        const uint32_t oct0 = __shfl_sync(full, (d[0].x >= 0.f ? 1u : 0u) | (d[0].y >= 0.f ? 2u : 0u) | (d[0].z >= 0.f ? 4u : 0u), 0);   // mixed axes: lane 0 decides the order
        const uint32_t octinv = (mixx ? (oct0 & 1u) : negx ? 0u : 1u) | (mixy ? (oct0 & 2u) : negy ? 0u : 2u) | (mixz ? (oct0 & 4u) : negz ? 0u : 4u);
        // padding of the child boxes: 1/8 grid cell (the builder's own margin is 1/64) + 2e-6 of the largest coordinate involved
        const uint4 root = sc.n_nodes ? __ldg(sc.nodes) : make_uint4(0u, 0u, 0u, 0u);
        const float kx = (fabsf(__uint_as_float(root.x)) + 256.f * __uint_as_float((root.w & 0xffu) << 23) + fmaxf(fabsf(ox0), fabsf(ox1))) * 2e-6f;
        const float ky = (fabsf(__uint_as_float(root.y)) + 256.f * __uint_as_float(((root.w >> 8) & 0xffu) << 23) + fmaxf(fabsf(oy0), fabsf(oy1))) * 2e-6f;
        const float kz = (fabsf(__uint_as_float(root.z)) + 256.f * __uint_as_float(((root.w >> 16) & 0xffu) << 23) + fmaxf(fabsf(oz0), fabsf(oz1))) * 2e-6f;
        // entry plane = lo (pos) / hi (neg) against the origin bound that makes the entry earliest; exit plane likewise latest
        const float onx = negx ? ox0 : ox1, ofx = negx ? ox1 : ox0, ony = negy ? oy0 : oy1, ofy = negy ? oy1 : oy0, onz = negz ? oz0 : oz1, ofz = negz ? oz1 : oz0;
        // (32768 + q) * s + (g -+ k -+ (32768 +- 1/8) s) = g + (q -+ 1/8) s -+ k
        const float cn_x = negx ? -32767.875f : -32768.125f, cf_x = negx ? -32768.125f : -32767.875f, kn_x = negx ? kx : -kx;
        const float cn_y = negy ? -32767.875f : -32768.125f, cf_y = negy ? -32768.125f : -32767.875f, kn_y = negy ? ky : -ky;
        const float cn_z = negz ? -32767.875f : -32768.125f, cf_z = negz ? -32768.125f : -32767.875f, kn_z = negz ? kz : -kz;

        uint2 ng = sc.n_nodes ? make_uint2(0u, (1u << (24 + octinv)) | 1u) : make_uint2(0u, 0u);
        int sp = 0;
        for (;;) {
            if ((ng.y & 0xff000000u) == 0u) { if (sp == 0) break; ng = stack[--sp]; }
            const int bit = 31 - __clz(ng.y);
            ng.y &= ~(1u << bit);
            if (ng.y & 0xff000000u) { stack[sp] = ng; sp++; }          // depth <= kStackSize, guarded at build time
            const uint32_t cslot = (uint32_t) (bit - 24) ^ octinv;
            const uint32_t rel = __popc(ng.y & 0xffu & ~(0xffffffffu << cslot));
            const uint4 *np = sc.nodes + (size_t) (ng.x + rel) * 5;
            const uint4 w0 = __ldg(np), w1 = __ldg(np + 1), w2 = __ldg(np + 2), w3 = __ldg(np + 3), w4 = __ldg(np + 4);
            if (lane == 0) cnt.add(C_NODES);
            // child c of this node against the interval of the rays
            const float sx = __uint_as_float((w0.w & 0xffu) << 23), sy = __uint_as_float(((w0.w >> 8) & 0xffu) << 23), sz = __uint_as_float(((w0.w >> 16) & 0xffu) << 23);
            const float gx = __uint_as_float(w0.x), gy = __uint_as_float(w0.y), gz = __uint_as_float(w0.z);
            const uint32_t lxw = chi ? w2.y : w2.x, lyw = chi ? w2.w : w2.z, lzw = chi ? w3.y : w3.x, hxw = chi ? w3.w : w3.z, hyw = chi ? w4.y : w4.x, hzw = chi ? w4.w : w4.z;
            const float qnx = __uint_as_float(__byte_perm(negx ? hxw : lxw, sc.qmagic, psel)), qfx = __uint_as_float(__byte_perm(negx ? lxw : hxw, sc.qmagic, psel));
            const float qny = __uint_as_float(__byte_perm(negy ? hyw : lyw, sc.qmagic, psel)), qfy = __uint_as_float(__byte_perm(negy ? lyw : hyw, sc.qmagic, psel));
            const float qnz = __uint_as_float(__byte_perm(negz ? hzw : lzw, sc.qmagic, psel)), qfz = __uint_as_float(__byte_perm(negz ? lzw : hzw, sc.qmagic, psel));
            const float nx = fmaf(qnx, sx, fmaf(cn_x, sx, (gx + kn_x) - onx)), fx = fmaf(qfx, sx, fmaf(cf_x, sx, (gx - kn_x) - ofx));
            const float ny = fmaf(qny, sy, fmaf(cn_y, sy, (gy + kn_y) - ony)), fy = fmaf(qfy, sy, fmaf(cf_y, sy, (gy - kn_y) - ofy));
            const float nz = fmaf(qnz, sz, fmaf(cn_z, sz, (gz + kn_z) - onz)), fz = fmaf(qfz, sz, fmaf(cf_z, sz, (gz - kn_z) - ofz));
            float tnx = fminf(nx * ix0, nx * ix1), tny = fminf(ny * iy0, ny * iy1), tnz = fminf(nz * iz0, nz * iz1);
            float tfx = fmaxf(fx * ix0, fx * ix1), tfy = fmaxf(fy * iy0, fy * iy1), tfz = fmaxf(fz * iz0, fz * iz1);
            if (anymix) {
                if (mixx) { tfx = kPkInf; tnx = fmaxf(tnx, fx * mx); }
                if (mixy) { tfy = kPkInf; tny = fmaxf(tny, fy * my); }
                if (mixz) { tfz = kPkInf; tnz = fmaxf(tnz, fz * mz); }
            }
            float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.f));
            float tf = fminf(fminf(tfx, tfy), fminf(tfz, tbest_w * kFarPad));
            tn *= 0.999995f; tf = fmaf(fabsf(tf), 5e-6f, tf);
            const uint32_t mw = ((chi ? w1.w : w1.z) >> csh) & 0xffu;
            const uint32_t mine = (lane < 8 && !(tn > tf)) ? __funnelshift_l(0u, (mw >> 5) & 7u, mw) : 0u;
            const uint32_t hitmask = __reduce_or_sync(full, mine);
            ng = make_uint2(w1.x, (permute_octant(hitmask, octinv) << 24) | (w0.w >> 24));
            uint32_t tm = hitmask & 0x00ffffffu;
            if (tm) {
                const float4 *tb = sc.tris + (size_t) w1.y * 3;
                float tw = 0.f;
                do {
                    const int i = __ffs(tm) - 1;
                    tm &= tm - 1;
                    const float4 *tp = tb + i * 3;
                    const float4 r0 = __ldg(tp), r1 = __ldg(tp + 1), r2 = __ldg(tp + 2);
                    cnt.add(C_TRIS, R);
                    const f3 e1(r1.x, r1.y, r1.z), e2(r2.x, r2.y, r2.z), p0(r0.x, r0.y, r0.z);
                    const uint32_t face = __float_as_uint(r0.w);
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        // Mesh::moeller_trumbore (mesh.h:469-492), operation for operation as in k_wf_trace
                        const f3 pvec = cross(d[r], e2);
                        const float inv_det = rcp(dot(e1, pvec));
                        const f3 tvec = o[r] - p0;
                        const float u = dot(tvec, pvec) * inv_det;
                        bool ok = u >= 0.f && u <= 1.f;
                        const f3 qvec = cross(tvec, e1);
                        const float v = dot(d[r], qvec) * inv_det;
                        ok = ok && v >= 0.f && u + v <= 1.f;
                        const float t = dot(e2, qvec) * inv_det;
                        ok = ok && t >= 0.f && t <= tbest[r];
                        if (ok && (t < tbest[r] || face < bface[r])) { tbest[r] = t; bu[r] = u; bv[r] = v; bface[r] = face; bmesh[r] = __float_as_uint(r1.w); }
                    }
                } while (tm);
#pragma unroll
                for (int r = 0; r < R; ++r) tw = fmaxf(tw, tbest[r]);
                tbest_w = __uint_as_float(__reduce_max_sync(full, __float_as_uint(tw)));
            }
        }
        // ---- epilogue: the hit record and the class byte of each ray (queue position = lane slot on bounce 0)
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const uint32_t e = base + 32u * r + lane;
            if (e < end) {
                const bool found = bface[r] != 0xffffffffu;
                const uint32_t cls = found ? (bmesh[r] >> 28) : (uint32_t) CLS_MISS;
                Elem h;
                h.a = make_float4(tbest[r], bu[r], bv[r], __uint_as_float(bface[r]));
                h.b = make_float4(__uint_as_float(found ? (bmesh[r] & 0x0fffffffu) : 0xffffffffu), __uint_as_float(cls), 0.f, 0.f);
                st256(&st.hit[e], h);
                st.cls_q[e] = (uint8_t) cls;
            }
        }
        }   // pieces of the batch
    }
    trace_flush_counters(cnt, p.counters);
}

constexpr size_t kTraceDynSmem = sizeof(float4) * 32 * kTraceWarps;
static void trace_smem_opt_in() {       // function attributes belong to the current device: set on every call (one per render)
    cudaFuncSetAttribute(k_wf_trace<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kTraceDynSmem);
    cudaFuncSetAttribute(k_wf_trace<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kTraceDynSmem);
    cudaFuncSetAttribute(k_wf_trace<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kTraceDynSmem);
    cudaFuncSetAttribute(k_wf_trace<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kTraceDynSmem);
}
int wf_trace_blocks(bool count, int sm_count) {
    int occ = 0;
    trace_smem_opt_in();
    if (count) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_wf_trace<true, false>, kTraceThreads, kTraceDynSmem);
    else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_wf_trace<false, false>, kTraceThreads, kTraceDynSmem);
    static const int tune_blocks = getenv("B2_TRACE_BLOCKS") ? atoi(getenv("B2_TRACE_BLOCKS")) : 0;
    if (tune_blocks > 0) occ = std::min(occ, tune_blocks);
    return occ * sm_count;
}

// Bounce 0: rays per lane of the packet kernel (0 = the per-ray kernel k_wf_trace<., true>).  A packet of 32 R consecutive lanes spans
// 32 R / spp pixels of a row; the wider it is, the more of the tree it visits that its single rays would not.  Measured on the bedroom
// stand-in, 1280 x 720 (profiles/README.md): 64-lane packets win from 8 samples per pixel on (8 pixels per packet: 10.8 vs 11.2 ms; 64 spp:
// 66.2 vs 73.0 ms), 32-lane packets at 4 .. 7 (6.60 vs 6.74 ms), the per-ray kernel below that (1 spp: 3.5 vs 7.9 ms).
// B2_PRIMARY_PACKET = 0 / 1 / 2 forces the choice (tests, experiments).
static int primary_packet_rays(const RenderParams &p) {
    if (const char *e = getenv("B2_PRIMARY_PACKET")) return std::max(0, std::min(2, atoi(e)));
    return p.spp_per_pass >= 8u ? 2 : p.spp_per_pass >= 4u ? 1 : 0;
}

static uint32_t simple_trace_nodes() { const char *e = getenv("B2_SIMPLE_TRACE_NODES"); return e ? (uint32_t) atoi(e) : 32u; }

static void launch_trace(bool count, bool primary, int packet_rays, const RenderParams &p, const WfLaunch &w, int blocks, cudaStream_t stream) {
    wf_tm_begin(WF_K_TRACE, stream);
    if (!primary && p.sc.n_nodes <= simple_trace_nodes()) {
        const int g = 8 * std::max(blocks / 4, 1);           // blocks = 4 per SM of the persistent kernel -> 8 per SM, grid-stride
        if (count) k_wf_trace_simple<true><<<g, 256, 0, stream>>>(p, w); else k_wf_trace_simple<false><<<g, 256, 0, stream>>>(p, w);
    } else if (primary && packet_rays == 2) {
        if (count) k_wf_trace_packet<true, 2><<<blocks, kPacketThreads, 0, stream>>>(p, w); else k_wf_trace_packet<false, 2><<<blocks, kPacketThreads, 0, stream>>>(p, w);
    } else if (primary && packet_rays == 1) {
        if (count) k_wf_trace_packet<true, 1><<<blocks, kPacketThreads, 0, stream>>>(p, w); else k_wf_trace_packet<false, 1><<<blocks, kPacketThreads, 0, stream>>>(p, w);
    } else if (count) { if (primary) k_wf_trace<true, true><<<blocks, kTraceThreads, kTraceDynSmem, stream>>>(p, w); else k_wf_trace<true, false><<<blocks, kTraceThreads, kTraceDynSmem, stream>>>(p, w); }
    else { if (primary) k_wf_trace<false, true><<<blocks, kTraceThreads, kTraceDynSmem, stream>>>(p, w); else k_wf_trace<false, false><<<blocks, kTraceThreads, kTraceDynSmem, stream>>>(p, w); }
    wf_tm_end(stream);
}
void wf_launch_trace(bool count, bool primary, const RenderParams &p, const WfLaunch &w, int blocks, cudaStream_t stream) {
    launch_trace(count, primary, primary ? primary_packet_rays(p) : 0, p, w, blocks, stream);
}
void wf_launch_trace_primary(int packet_rays, const RenderParams &p, const WfLaunch &w, int blocks, cudaStream_t stream) {
    launch_trace(false, true, packet_rays, p, w, blocks, stream);
}

}  // namespace b2
