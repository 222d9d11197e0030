// b2_wavefront.cuh -- wavefront formulation of the render loop: state streams, queues and launch interface.
//
// The reference records `sample()` once and lets Dr.Jit compile one megakernel over a wavefront of W*H*spp lanes
// (src/render/integrator.cpp:246-360); lanes that finish early idle inside dr.while_loop.  Here the wavefront is
// explicit: per bounce one TRACE launch (closest rays of the live lanes and the shadow rays of the previous bounce,
// persistent warps with dynamic ray fetch) and one SHADE launch (path_step<WF>), with the live lanes compacted into the
// next queue by warp ballot + prefix sums.  Lane state lives in HBM as streams of 32-byte elements between launches.
#pragma once
#include "b2_render.cuh"
#include "b2_math.cuh"
#include <vector>

namespace b2 {

struct WfCtl {                       // device-side queue bookkeeping, slot = iteration & 3
    uint32_t n_closest[4];           // live lanes entering iteration it (queue q[it & 1])
    uint32_t n_shadow[4];            // shadow rays emitted by shade(it)
    uint32_t fetch[4];               // persistent-warp work counter of trace(it)
    uint32_t n_class[4][8];          // hits of iteration it per shading class (queues qc[cls])
};

// Lane-state accessor.  Experiment (B2_STATE_CS=1): every stream element is read once and written once per bounce and a chunk's
// state (GBs) is far larger than L2, so the accesses could carry the evict-first hint (ld.global.cs / st.global.cs) and leave L2
// to the data that is reused (nodes, triangles, per-face records, textures).  Measured on the bedroom stand-in: trace 44.4 ->
// 43.6 ms but shade 24.0 -> 25.8 ms (what a shade kernel writes is what classify / trace read next, and evict-first loses the
// part of it that still sat in L2), net 789 -> 775 Msamples/s: off.  The L2 node pin of capi.cu gets the trace gain alone.
#ifndef B2_STATE_CS
#define B2_STATE_CS 0
#endif
template <class T> struct StateRef {
    T *p;
#if B2_STATE_CS
    B2_DEV operator T() const { return __ldcs(p); }
    B2_DEV const StateRef &operator=(const T &v) const { __stcs(p, v); return *this; }
#else
    B2_DEV operator T() const { return *p; }
    B2_DEV const StateRef &operator=(const T &v) const { *p = v; return *this; }
#endif
    B2_DEV const StateRef &operator=(const StateRef &o) const { return *this = (T) o; }
    B2_DEV T *ptr() const { return p; }
};

struct WfState {                     // `capacity` lane slots (slot = lane - chunk_begin)
    // Lane state = four streams of 32-byte elements (two 16-byte words each, grouped by which kernel touches them
    // together).  A 32-byte element is exactly one DRAM/L2 sector, so a lane costs whole sectors however sparse or
    // permuted the live lanes are (16-byte SoA streams moved 3x the useful bytes in late, class-sorted bounces), while
    // dense warps still touch only 8 cache lines per request (a 128-byte record per lane touched 32: measured +30 % on
    // the coherent Cornell / PRB workloads).
    float4 *p0;                      // ray o.xyz | maxt          , ray d.xyz | eta                     (trace reads)
    float4 *p1;                      // hit t, u, v | face        , hit mesh, vis bits, splat pos xy    (trace writes)
    float4 *p2;                      // PCG32 state, inc          , throughput f.xyz | prev_bsdf_pdf
    float4 *p3;                      // L.xyz | depth + flags     , prev_p.xyz | -
    float4 *s0;                      // deferred NEE: shadow ray o | maxt , d | tex flag                (trace reads)
    float4 *s1;                      //               factor a            , factor b   (L <- fma(a, b, L) on arrival)
    float4 *dl;                      // PRB adjoint: dL (adjoint radiance of the lane's splat), xyz
    uint8_t *cls_q;                  // shading class of the hit of queue entry i of the current iteration (trace -> classify)
    B2_DEV StateRef<float4> ray_o(uint32_t s) const { return StateRef<float4>{ &p0[(size_t) s * 2] }; }
    B2_DEV StateRef<float4> ray_d(uint32_t s) const { return StateRef<float4>{ &p0[(size_t) s * 2 + 1] }; }
    B2_DEV StateRef<float4> hit(uint32_t s) const { return StateRef<float4>{ &p1[(size_t) s * 2] }; }
    B2_DEV StateRef<uint32_t> hit_mesh(uint32_t s) const { return StateRef<uint32_t>{ &((uint32_t *) (p1 + (size_t) s * 2 + 1))[0] }; }
    B2_DEV StateRef<uint32_t> vis(uint32_t s) const { return StateRef<uint32_t>{ &((uint32_t *) (p1 + (size_t) s * 2 + 1))[1] }; }
    B2_DEV StateRef<float2> pos(uint32_t s) const { return StateRef<float2>{ &((float2 *) (p1 + (size_t) s * 2 + 1))[1] }; }
    B2_DEV StateRef<uint4> rng(uint32_t s) const { return StateRef<uint4>{ &((uint4 *) p2)[(size_t) s * 2] }; }
    B2_DEV StateRef<float4> thr(uint32_t s) const { return StateRef<float4>{ &p2[(size_t) s * 2 + 1] }; }
    B2_DEV StateRef<float4> rad(uint32_t s) const { return StateRef<float4>{ &p3[(size_t) s * 2] }; }
    B2_DEV StateRef<float4> prev(uint32_t s) const { return StateRef<float4>{ &p3[(size_t) s * 2 + 1] }; }
    B2_DEV StateRef<float4> sh_o(uint32_t s) const { return StateRef<float4>{ &s0[(size_t) s * 2] }; }
    B2_DEV StateRef<float4> sh_d(uint32_t s) const { return StateRef<float4>{ &s0[(size_t) s * 2 + 1] }; }
    B2_DEV StateRef<float4> sh_a(uint32_t s) const { return StateRef<float4>{ &s1[(size_t) s * 2] }; }
    B2_DEV StateRef<float4> sh_b(uint32_t s) const { return StateRef<float4>{ &s1[(size_t) s * 2 + 1] }; }
    uint32_t *q[2];                  // closest-ray queues (lane slots), double buffered by iteration parity
    uint32_t *q_shadow[2];           // shadow-ray queues, by iteration parity
    uint32_t *qc[5];                 // closest hits of the current iteration sorted by shading class (CLS_*)
    uint32_t class_mask;             // classes present in the scene
    WfCtl *ctl;
    uint32_t capacity;
};

enum { WF_FLAG_PREV_DELTA = 1u << 24, WF_FLAG_FIRST = 1u << 25, WF_FLAG_VALID_RAY = 1u << 26, WF_DEPTH_MASK = 0x00ffffffu };

struct WfLaunch {
    WfState st;
    uint64_t chunk_begin;            // first lane (index within the pass wavefront) of this chunk
    uint32_t n;                      // lanes in this chunk
    uint32_t it;                     // global iteration counter (slot = it & 3)
    uint32_t pass;                   // > 0: continue each lane's PCG32 stream (integrator.cpp:319-328)
    uint32_t slice;                  // trace: traversal micro-steps between two refill points
    uint32_t vis_bit;                // trace: bit to set in vis[] for arriving shadow rays; shade<ADJ>: bit to read; ~0u = off
    uint32_t tri_thresh;             // trace: lanes with pending leaf triangles that trigger a triangle step
    uint32_t prefetch;               // trace: bit 0 = prefetch parked leaf triangles into L1
};

// per-kernel-kind timing (host side)
enum { WF_K_TRACE = 0, WF_K_SHADE = 1, WF_K_SPLAT = 2, WF_K_OTHER = 3, WF_K_COUNT = 4 };
struct WfTiming { std::vector<cudaEvent_t> ev; std::vector<int> kind; size_t used = 0; };
void wf_set_timing(WfTiming *t);          // nullptr = off (default)
void wf_collect_timing(WfTiming &t, double ms[WF_K_COUNT], uint64_t n[WF_K_COUNT]);   // after the stream has been synchronised

// host entry points (wavefront.cu)
size_t wf_state_bytes(uint32_t capacity);
void wf_bind(WfState &st, void *base, uint32_t capacity);
// Renders lanes [p.lane_begin, p.lane_end) of every pass into p.film.  Returns the number of kernels launched, or -1
// with `err` set.
// PRB adjoint for lanes [p.lane_begin, p.lane_end): primal pass (records emitter-sample visibility), then the replay
// with gradient scatter.  p.wfilm must hold the weight film of the sample set.  Needs max_depth <= 32.
long wf_render_backward(bool count, RenderParams p, WfState &st, int sm_count, cudaStream_t stream, const char **err);
long wf_weights(RenderParams p, int sm_count, cudaStream_t stream, const char **err);
long wf_render(int variant, bool count, RenderParams p, WfState &st, int sm_count, cudaStream_t stream, const char **err);

}  // namespace b2
