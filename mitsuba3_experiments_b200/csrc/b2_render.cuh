// b2_render.cuh -- launch parameters shared by render.cu and capi.cu
#pragma once
#include "b2_scene.cuh"

// experiment switch: -DB2_NO_FORWARD compiles the forward-mode (render_forward) branches out of the adjoint kernels
#ifdef B2_NO_FORWARD
#define B2_FWD(p) false
#else
#define B2_FWD(p) ((p).forward != 0)
#endif

namespace b2 {

constexpr int kMaxGradTargets = 8;
constexpr int kBlockThreads = 256;

enum { MODE_RENDER = 0, MODE_WEIGHTS = 1, MODE_BACKWARD = 2 };

struct GradTargetDev { int32_t bsdf, kind; float *grad; };

struct RenderParams {
    DScene sc;
    uint32_t seed, spp_per_pass, n_passes, spp_shift;   // spp_shift: log2(spp_per_pass) or 0xffffffff (integrator.cpp:293-300)
    uint32_t remap_spp, remap_off;                       // sample shard: global samples per pixel per pass (0 = no shard) and first sample
    int32_t cull_nee;                                    // wavefront: skip shadow rays of zero-valued emitter samples
    int32_t max_depth, rr_depth, env_warp, filter_n;     // filter_n = ceil(radius - 0.5) (imageblock.cpp:438-441); box: 0
    uint64_t lane_begin, lane_end;                       // shard of the per-pass wavefront
    uint32_t unit_pixels, units_per_row, unit_begin, unit_end, tile_cols, tile_rows;
    float *film;                                         // [crop_h][crop_w][4] RGBW, accumulated into
    uint32_t *unit_counter;                              // persistent-thread work counter
    unsigned long long *counters;                        // [C_N] or nullptr
    // PRB adjoint
    const float *grad_in;                                // [crop_h][crop_w][3]
    const float *wfilm;                                  // RGBW film holding the weights of this sample set
    GradTargetDev targets[kMaxGradTargets]; int32_t n_targets;
    int32_t forward;                                     // adjoint replay in forward mode: targets[].grad are TANGENTS (read), dL accumulates d L
};

void launch_render(int variant, int mode, bool count, const RenderParams &p, int blocks, size_t smem, cudaStream_t st);
int render_occupancy(int variant, int mode, bool count, size_t smem);
void launch_develop(const float *film, float *image, uint32_t n_pixels, cudaStream_t st);
void launch_debug_sampler(const RenderParams &p, uint64_t lane_begin, uint32_t n_lanes, uint32_t n_draws, float *out, float *pos, cudaStream_t st);
void launch_trace(const DScene &sc, const float *rays, uint64_t n, int any_hit, b2_hit *hits, cudaStream_t st);

}  // namespace b2
