// b2_trace_noalloc: the production traversal kernel (k_wf_trace, wf_trace.cu) over caller-supplied rays with NO device allocation.
// No counterpart in the reference; added for callers that cannot call cudaMalloc (code under
// CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// b2_trace keeps the wavefront lane state of the rays (ray / hit / shadow streams, queues, control block) in a cudaMalloc'ed block
// that grows on demand.  Here the same state lives in ONE file-scope device array of a fixed size, held by every GPU that loads the
// library until the process exits:  kNoAllocMaxRays = 32768 rays  ->  wf_state_bytes(32768) = 8 421 632 B (8.03 MiB).
//
// g_noalloc_scratch below is, in host code, a host-side stand-in for the symbol and NOT a device address: the kernels get the
// address cudaGetSymbolAddress returns, resolved once per device with that device current.
// All calls share the array and it keeps the last call's contents; what must start at zero (the control block: queue counts and
// the work counters) is cleared on the caller's stream at the head of every call (wf_trace_rays), every other byte a call reads is
// written by that call first.  Two calls that overlap on the GPU (different streams, or different host threads without ordering)
// race on the array: callers serialise them.
#include <mutex>

#include "b2_shade.cuh"
#include "b2_wavefront.cuh"

namespace b2 {

static constexpr size_t na_align(size_t x) { return (x + 255) / 256 * 256; }
// the layout of wf_bind (wavefront.cu): control block, six 32-B streams, dl, pos, class bytes, 3 + CLS_COUNT queues
static constexpr size_t na_state_bytes(size_t c) {
    return na_align(c * 32) * 6 + na_align(c * 16) + na_align(c * 8) + na_align(c) + na_align(c * 4) * (3 + CLS_COUNT) + 256;
}
static constexpr size_t kNoAllocBytes = na_state_bytes(kNoAllocMaxRays);

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

static constexpr int kMaxDevices = 64;
static std::mutex g_noalloc_mutex;
static void *g_noalloc_ptr[kMaxDevices] = {};

// the device address of the scratch array on `device` (which must be current)
static void *noalloc_scratch(int device, const char **err) {
    if (device < 0 || device >= kMaxDevices) { *err = "b2_trace_noalloc: device index out of range"; return nullptr; }
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device]) {
        void *ptr = nullptr;
        if (cudaGetSymbolAddress((void **) &ptr, g_noalloc_scratch) != cudaSuccess || !ptr) {
            cudaGetLastError();
            *err = "b2_trace_noalloc: cudaGetSymbolAddress failed for the static scratch array";
            return nullptr;
        }
        g_noalloc_ptr[device] = ptr;
    }
    return g_noalloc_ptr[device];
}

long wf_trace_rays_noalloc(RenderParams p, uint32_t class_mask, int device, const float *rays, uint64_t n, int any_hit, b2_hit *hits,
                           int sm_count, cudaStream_t stream, const char **err) {
    if (n > kNoAllocMaxRays) { *err = "b2_trace_noalloc: more than 32768 rays (the static scratch array holds the state of 32768)"; return -2; }
    if (wf_state_bytes((uint32_t) kNoAllocMaxRays) > kNoAllocBytes) { *err = "b2_trace_noalloc: scratch layout out of date"; return -1; }
    void *base = noalloc_scratch(device, err);
    if (!base) return -1;
    WfState st{};
    wf_bind(st, base, (uint32_t) kNoAllocMaxRays);
    st.class_mask = class_mask;
    return wf_trace_rays(p, st, rays, n, any_hit, hits, sm_count, stream, err);
}

}  // namespace b2
