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
#include "b2_path.cuh"

namespace b2 {


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
        const uint32_t x0 = ux * p.unit_pixels, x1 = min(x0 + p.unit_pixels, (uint32_t) sn.grid_w);
        uint64_t lane_lo = ((uint64_t) row * sn.grid_w + x0) * p.spp_per_pass, lane_hi = ((uint64_t) row * sn.grid_w + x1) * p.spp_per_pass;
        lane_lo = max(lane_lo, p.lane_begin); lane_hi = min(lane_hi, p.lane_end);
        const uint32_t n_lanes = lane_hi > lane_lo ? (uint32_t) (lane_hi - lane_lo) : 0u;
        const int tx0 = (int) x0 - sn.border - p.filter_n, ty0 = (int) row - sn.border - p.filter_n;   // crop-relative tile origin
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
                    s.rng.seed_lane(p.seed, global_lane(p, s.lane));
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
                    s.rng.seed_lane(p.seed, global_lane(p, s.lane));
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
// unit-test hook of the sampler: the draws every render kernel makes for a lane come from Pcg32::seed_lane(seed, global lane)
__global__ void k_debug_sampler(const __grid_constant__ RenderParams p, uint64_t lane_begin, uint32_t n_lanes, uint32_t n_draws, float *out, float *pos) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_lanes) return;
    const uint32_t lane = (uint32_t) (lane_begin + i);
    Pcg32 rng; rng.seed_lane(p.seed, global_lane(p, lane));
    if (pos) { Pcg32 r2 = rng; int px, py; const f2 sp = lane_sample_pos(p, lane, r2, px, py); pos[2 * (size_t) i] = sp.x; pos[2 * (size_t) i + 1] = sp.y; }
    for (uint32_t k = 0; k < n_draws; ++k) out[(size_t) i * n_draws + k] = rng.next_1d();
}
void launch_debug_sampler(const RenderParams &p, uint64_t lane_begin, uint32_t n_lanes, uint32_t n_draws, float *out, float *pos, cudaStream_t st) {
    if (n_lanes) k_debug_sampler<<<(n_lanes + 127) / 128, 128, 0, st>>>(p, lane_begin, n_lanes, n_draws, out, pos);
}

void launch_trace(const DScene &sc, const float *rays, uint64_t n, int any_hit, b2_hit *hits, cudaStream_t st) {
    if (n) k_trace<<<(unsigned) ((n + 127) / 128), 128, 0, st>>>(sc, rays, n, any_hit, hits);
}

}  // namespace b2
