// b2_wavefront.cuh -- wavefront formulation of the render loop: state streams, queues and launch interface.
//
// The reference records `sample()` once and lets Dr.Jit compile one megakernel over a wavefront of W*H*spp lanes
// (src/render/integrator.cpp:246-360); lanes that finish early idle inside dr.while_loop.  Here the wavefront is
// explicit: per bounce one TRACE launch (closest rays of the live lanes and the shadow rays of the previous bounce,
// persistent warps with a prefetched ray pipeline, wf_trace.cu) and one SHADE launch (path_step<WF>), with the live lanes
// compacted into the next queue by warp ballot + prefix sums.  Lane state lives in HBM as streams of 32-byte elements
// between launches; every element is moved by ONE 256-bit load / store (LDG.E.256 / STG.E.256, sm_100).
//
// There is no generate kernel: the trace launch of bounce 0 derives the camera ray from the lane index (sampler seed,
// jitter, sensor), and the first shade launch re-derives the same ray and seeds the lane state.
// Deferred next-event estimation: shade(k) writes the shadow ray and the two factors of the contribution (L <- fma(a, b, L)),
// trace(k + 1) only flags the rays that arrive, and the addition happens where the lane's radiance is loaded next
// (shade(k + 1), or the splat for lanes that ended) -- the trace kernel's epilogue is stores only.
#pragma once
#include "b2_render.cuh"
#include "b2_math.cuh"
#include <vector>

namespace b2 {

struct WfCtl {                       // device-side queue bookkeeping, slot = iteration & 3; zeroed at the start of a pass
    uint32_t n_closest[4];           // live lanes entering iteration it (queue q[it & 1])
    uint32_t n_shadow[4];            // shadow rays emitted by shade(it)
    uint32_t fetch[4];               // persistent-warp work counter of trace(it)
    uint32_t n_class[4][8];          // hits of iteration it per shading class (queues qc[cls])
};

// One 32-byte state element = one DRAM / L2 sector, moved by a single 256-bit access.
struct alignas(32) Elem { float4 a, b; };
B2_DEV Elem ld256(const Elem *p) {
    Elem e;
    asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(e.a.x), "=f"(e.a.y), "=f"(e.a.z), "=f"(e.a.w), "=f"(e.b.x), "=f"(e.b.y), "=f"(e.b.z), "=f"(e.b.w) : "l"(p));
    return e;
}
B2_DEV void st256(Elem *p, const Elem &e) {
    asm volatile("st.global.v8.f32 [%8], {%0,%1,%2,%3,%4,%5,%6,%7};"
                 :: "f"(e.a.x), "f"(e.a.y), "f"(e.a.z), "f"(e.a.w), "f"(e.b.x), "f"(e.b.y), "f"(e.b.z), "f"(e.b.w), "l"(p) : "memory");
}

struct WfState {                     // `capacity` lane slots (slot = lane - chunk_begin)
    // A 32-byte element is exactly one sector, so a lane costs whole sectors however sparse or permuted the live lanes are
    // (16-byte SoA streams moved 3x the useful bytes in late, class-sorted bounces), while dense warps still touch only 8
    // cache lines per request.
    Elem *ray;                       // o.xyz | maxt          , d.xyz | eta                        shade -> trace, shade
    Elem *hit;                       // t, u, v | face        , mesh, class, -, -                  trace -> shade
    Elem *rng;                       // PCG32 state, inc      , throughput f.xyz | prev_bsdf_pdf   shade -> shade
    Elem *rad;                       // L.xyz | depth + flags , prev_p.xyz | visibility bits (PRB) shade -> shade, splat
    Elem *shr;                       // deferred NEE: shadow ray o | maxt , d | tex flag           shade -> trace
    Elem *shf;                       //   factor a | arrived (set by trace), factor b | bounce     shade -> shade / splat
    float4 *dl;                      // PRB adjoint: dL (adjoint radiance of the lane's splat), xyz
    float2 *pos;                     // splat position, only for multi-pass renders (a single pass re-derives it from (seed, lane))
    uint8_t *cls_q;                  // shading class of the hit of queue entry i of the current iteration (trace -> classify)
    uint32_t *q[2];                  // closest-ray queues (lane slots), double buffered by iteration parity
    uint32_t *q_shadow;              // shadow-ray queue
    uint32_t *qc[7];                 // closest hits of the current iteration sorted by shading class (CLS_*)
    uint32_t class_mask;             // classes present in the scene
    WfCtl *ctl;
    uint32_t capacity;
};

enum { WF_FLAG_PREV_DELTA = 1u << 24, WF_FLAG_FIRST = 1u << 25, WF_FLAG_VALID_RAY = 1u << 26, WF_FLAG_PENDING = 1u << 27, WF_DEPTH_MASK = 0x00ffffffu };

struct WfLaunch {
    WfState st;
    uint64_t chunk_begin;            // first lane (index within the pass wavefront) of this chunk
    uint32_t n;                      // lanes in this chunk
    uint32_t it;                     // global iteration counter (slot = it & 3)
    uint32_t pass;                   // > 0: continue each lane's PCG32 stream (integrator.cpp:319-328)
    uint32_t first;                  // bounce 0 of a pass: every lane of the chunk is live, rays come from the sensor
    uint32_t no_hit;                 // shade: no trace ran for this bounce (path.py with max_depth <= 1): every lane is a miss
    uint32_t bounce;                 // bounce index of this launch (PRB: the visibility bit a shadow record belongs to)
    uint32_t record_vis;             // PRB primal pass of render_backward: keep the visibility of every emitter sample
    uint32_t tri_thresh;             // trace: lanes with pending leaf triangles that trigger a triangle step
    uint32_t refill_thresh;          // trace: idle lanes that trigger a refill from the staged rays
    uint32_t store_pos;              // multi-pass render: keep the splat position in st.pos
};

// per-kernel-kind timing (host side)
enum { WF_K_TRACE = 0, WF_K_SHADE = 1, WF_K_SPLAT = 2, WF_K_OTHER = 3, WF_K_COUNT = 4 };
struct WfTiming { std::vector<cudaEvent_t> ev; std::vector<int> kind; size_t used = 0; };
void wf_set_timing(WfTiming *t);          // nullptr = off (default)
void wf_collect_timing(WfTiming &t, double ms[WF_K_COUNT], uint64_t n[WF_K_COUNT]);   // after the stream has been synchronised
void wf_set_film_ready_event(cudaEvent_t e);   // nullptr = the film is ready when the render is enqueued (default); thread-local
void wf_tm_begin(int kind, cudaStream_t st);
void wf_tm_end(cudaStream_t st);

// trace launches (wf_trace.cu).  primary: rays of bounce 0 come from the sensor (needs the full RenderParams).
int wf_trace_blocks(bool count, int sm_count);          // resident grid of the persistent trace kernel (0 = cannot be resident)
void wf_launch_trace(bool count, bool primary, const RenderParams &p, const WfLaunch &w, int blocks, cudaStream_t stream);
void wf_launch_trace_primary(int packet_rays, const RenderParams &p, const WfLaunch &w, int blocks, cudaStream_t stream);   // bounce 0 with an explicit kernel: 0 per-ray, 1 / 2 = packets of 32 / 64 lanes

// host entry points (wavefront.cu)
size_t wf_state_bytes(uint32_t capacity);
void wf_bind(WfState &st, void *base, uint32_t capacity);
// Renders lanes [p.lane_begin, p.lane_end) of every pass into p.film.  Returns the number of kernels launched, or -1
// with `err` set.
// PRB adjoint for lanes [p.lane_begin, p.lane_end): primal pass (records emitter-sample visibility), then the replay
// with gradient scatter.  p.wfilm must hold the weight film of the sample set.  Needs max_depth <= 32.
// hit_history: nullptr, or max(max_depth, 1) x capacity 32-byte hit records (the adjoint replay then skips its closest-ray traces)
long wf_render_backward(bool count, RenderParams p, WfState &st, void *hit_history, int sm_count, cudaStream_t stream, const char **err);
// b2_trace through k_wf_trace: rays [n][8] (device) -> hits [n] (device); any_hit: arrival test of shadow rays
long wf_trace_rays(RenderParams p, WfState &st, const float *rays, uint64_t n, int any_hit, b2_hit *hits, int sm_count, cudaStream_t stream, const char **err);
// the same, with the lane state in a file-scope device array instead of st (wf_noalloc.cu): no allocation, n <= kNoAllocMaxRays
constexpr uint64_t kNoAllocMaxRays = 32768;
long wf_trace_rays_noalloc(RenderParams p, uint32_t class_mask, int device, const float *rays, uint64_t n, int any_hit, b2_hit *hits,
                           int sm_count, cudaStream_t stream, const char **err);
// b2_debug_primary_hits: the bounce-0 trace alone for lanes [lane_begin, lane_begin + n) of pass 0; packet = rays per lane of k_wf_trace_packet (1, 2), 0 = k_wf_trace<., true>
long wf_trace_primary(RenderParams p, WfState &st, uint64_t lane_begin, uint32_t n, int packet, b2_hit *hits, int sm_count, cudaStream_t stream, const char **err);
long wf_weights(RenderParams p, int sm_count, cudaStream_t stream, const char **err);
long wf_render(int variant, bool count, RenderParams p, WfState &st, int sm_count, cudaStream_t stream, const char **err);

}  // namespace b2
