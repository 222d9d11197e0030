// b2_wavefront.cuh -- wavefront formulation of the render loop: state streams, queues and launch interface.
//
// The reference records `sample()` once and lets Dr.Jit compile one megakernel over a wavefront of W*H*spp lanes
// (src/render/integrator.cpp:246-360); lanes that finish early idle inside dr.while_loop.  Here the wavefront is
// explicit: per bounce one TRACE launch (closest rays of the live lanes and the shadow rays of the previous bounce,
// persistent warps with dynamic ray fetch) and one SHADE launch (path_step<WF>), with the live lanes compacted into the
// next queue by warp ballot + prefix sums.  Lane state lives in HBM as 16-byte SoA streams between launches.
#pragma once
#include "b2_render.cuh"

namespace b2 {

struct WfCtl {                       // device-side queue bookkeeping, slot = iteration & 3
    uint32_t n_closest[4];           // live lanes entering iteration it (queue q[it & 1])
    uint32_t n_shadow[4];            // shadow rays emitted by shade(it)
    uint32_t fetch[4];               // persistent-warp work counter of trace(it)
    uint32_t pad[4];
};

struct WfState {                     // all arrays have `capacity` entries, indexed by lane slot = lane - chunk_begin
    uint4 *rng;                      // PCG32 state, inc
    float4 *ray_o;                   // o.xyz | maxt
    float4 *ray_d;                   // d.xyz | eta
    float4 *thr;                     // f.xyz (throughput) | prev_bsdf_pdf
    float4 *rad;                     // L.xyz | depth (bits 0..23), flags (bits 24..): prev_delta, first, valid_ray
    float4 *prev;                    // prev_p.xyz | -
    float2 *pos;                     // splat position
    float4 *hit;                     // t, u, v | face (bits)
    uint32_t *hit_mesh;
    float4 *dl;                      // PRB adjoint: dL (adjoint radiance of the lane's splat), xyz
    uint32_t *vis;                   // PRB: bit k = the emitter sample of bounce k arrived (primal pass -> adjoint replay)
    float4 *sh_o, *sh_d, *sh_a, *sh_b;   // deferred NEE: shadow ray o|maxt, d|-, factors a, b
    uint32_t *q[2];                  // closest-ray queues (lane slots), double buffered by iteration parity
    uint32_t *q_shadow[2];           // shadow-ray queues, by iteration parity
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
};

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
