// bvh_build.cu -- GPU acceleration-structure build: Morton codes -> radix sort -> LBVH (Karras 2012) ->
// bottom-up refit -> level-synchronous collapse into the compressed 8-wide BVH of b2_bvh.cuh.
//
// Replaces rtcCommitScene(RTC_BUILD_QUALITY_HIGH) (mitsuba3/src/render/scene_embree.inl:115-167) and
// optixAccelBuild(PREFER_FAST_TRACE) (scene_optix.inl:473-506); input is the reference's Mesh buffer layout
// (mesh.cpp:1849-1882: fp32 positions, u32 faces).  Everything runs on the device; the host only reads one
// counter per tree level.
#include "b2_build.h"
#include "b2_bvh.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_select.cuh>
#include <cfloat>
#include <cstdlib>
#include <string>

namespace b2 {

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { err = std::string(#x) + ": " + cudaGetErrorString(e_); return false; } } while (0)

constexpr uint32_t kLeafFlag = 0x80000000u;
constexpr uint32_t kLeafMax = 3;       // triangles per leaf child (unary count in 3 meta bits)
constexpr float kQMargin = 1.f / 64.f;  // quantisation margin in grid cells

struct Box { float lo[3], hi[3]; };

B2_DEV uint32_t f2ord(float f) { uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
B2_HD float ord2f(uint32_t u) {
    uint32_t b = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef __CUDA_ARCH__
    return __uint_as_float(b);
#else
    float f; memcpy(&f, &b, 4); return f;
#endif
}

// ---- 1. per-triangle boxes + scene bounds (ordered-uint atomics)
__global__ void k_tri_bounds(const float *__restrict__ pos, const uint32_t *__restrict__ faces, uint32_t n,
                             float4 *__restrict__ blo, float4 *__restrict__ bhi, uint32_t *__restrict__ scene_bounds) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, hi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
    if (i < n) {
        for (int j = 0; j < 3; ++j) {
            uint32_t v = faces[3 * (size_t) i + j];
            for (int k = 0; k < 3; ++k) { float p = pos[3 * (size_t) v + k]; lo[k] = fminf(lo[k], p); hi[k] = fmaxf(hi[k], p); }
        }
        blo[i] = make_float4(lo[0], lo[1], lo[2], 0.f); bhi[i] = make_float4(hi[0], hi[1], hi[2], 0.f);
    }
    for (int k = 0; k < 3; ++k) {      // bounds of the centroids' support = bounds of all boxes
        float l = lo[k], h = hi[k];
        for (int o = 16; o; o >>= 1) { l = fminf(l, __shfl_xor_sync(0xffffffffu, l, o)); h = fmaxf(h, __shfl_xor_sync(0xffffffffu, h, o)); }
        if ((threadIdx.x & 31) == 0 && l <= h) { atomicMin(&scene_bounds[k], f2ord(l)); atomicMax(&scene_bounds[3 + k], f2ord(h)); }
    }
}

B2_DEV uint64_t spread21(uint32_t x) {
    uint64_t v = x & 0x1fffffu;
    v = (v | (v << 32)) & 0x1f00000000ffffULL;
    v = (v | (v << 16)) & 0x1f0000ff0000ffULL;
    v = (v | (v << 8)) & 0x100f00f00f00f00fULL;
    v = (v | (v << 4)) & 0x10c30c30c30c30c3ULL;
    v = (v | (v << 2)) & 0x1249249249249249ULL;
    return v;
}

// ---- 2. 63-bit Morton code of the box centre
__global__ void k_morton(const float4 *__restrict__ blo, const float4 *__restrict__ bhi, uint32_t n,
                         const uint32_t *__restrict__ scene_bounds, uint64_t *__restrict__ keys, uint32_t *__restrict__ vals) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 l = blo[i], h = bhi[i];
    float c[3] = { 0.5f * (l.x + h.x), 0.5f * (l.y + h.y), 0.5f * (l.z + h.z) };
    uint32_t q[3];
    for (int k = 0; k < 3; ++k) {
        float slo = ord2f(scene_bounds[k]), shi = ord2f(scene_bounds[3 + k]);
        float ext = shi - slo, x = ext > 0.f ? (c[k] - slo) / ext : 0.f;
        q[k] = (uint32_t) fminf(fmaxf(x * 2097152.f, 0.f), 2097151.f);
    }
    keys[i] = spread21(q[0]) | (spread21(q[1]) << 1) | (spread21(q[2]) << 2);
    vals[i] = i;
}

// ---- 3. Karras: one thread per internal node of the binary radix tree
B2_DEV int delta(const uint64_t *__restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz((uint32_t) i ^ (uint32_t) j);
    return __clzll((long long) (a ^ b));
}
__global__ void k_karras(const uint64_t *__restrict__ keys, int n, uint32_t *__restrict__ left, uint32_t *__restrict__ right,
                         uint32_t *__restrict__ parent, uint32_t *__restrict__ cnt) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2) if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = delta(keys, n, i, j);
    int s = 0, t = l;
    do { t = (t + 1) >> 1; if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t; } while (t > 1);
    int g = i + s * d + min(d, 0);
    int lo = min(i, j), hi = max(i, j);
    uint32_t L = (lo == g) ? ((uint32_t) g | kLeafFlag) : (uint32_t) g;
    uint32_t R = (hi == g + 1) ? ((uint32_t) (g + 1) | kLeafFlag) : (uint32_t) (g + 1);
    left[i] = L; right[i] = R; cnt[i] = (uint32_t) (hi - lo + 1);
    // parents: internal nodes index [0, n-1), leaves index (n-1) + leaf
    if (L & kLeafFlag) parent[(n - 1) + (L & ~kLeafFlag)] = (uint32_t) i; else parent[L] = (uint32_t) i;
    if (R & kLeafFlag) parent[(n - 1) + (R & ~kLeafFlag)] = (uint32_t) i; else parent[R] = (uint32_t) i;
    if (i == 0) parent[0] = 0xffffffffu;
}

B2_DEV uint32_t ref_index(uint32_t ref, uint32_t n) { return (ref & kLeafFlag) ? (n - 1) + (ref & ~kLeafFlag) : ref; }

// ---- 4. bottom-up refit + collapse DP; the second thread to arrive at a node merges its children.
// DP (Ylitie, Karras, Laine 2017, "Efficient incoherent ray traversal on GPUs through compressed wide BVHs", sec. 4.1):
// cost[x][i-1], i = 1..7, is the least SAH cost of representing the binary subtree x by at most i children of a wide
// node (each a leaf of <= 3 triangles or a wide node of its own):
//   dist(x, j)  = min_{0<k<j} cost[left][k] + cost[right][j-k]                 (j = 2..8)
//   cost[x][1]  = min(A_x P_x c_prim  if P_x <= 3,   A_x c_node + dist(x, 8))
//   cost[x][i]  = min(dist(x, i), cost[x][i-1])
// dec[x][j-1] records the choice: j = 1: 0 = leaf, else internal; j = 2..7: 0 = "as j-1", else the left budget k;
// dec[x][7] = the left budget of dist(x, 8).  A single triangle costs A c_prim for every i.
B2_DEV float half_area(const float4 &lo, const float4 &hi) { float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z; return dx * dy + dy * dz + dz * dx; }

// DP step of binary node p with children L, R (boxes a..b and c..e, node box lo..hi, `count` triangles below p)
B2_DEV void dp_node(uint32_t p, uint32_t L, uint32_t R, const float4 &a, const float4 &b, const float4 &c, const float4 &e,
                    const float4 &lo, const float4 &hi, uint32_t count, float *cost, uint2 *dec, float c_node, float c_prim) {
    float cl[7], cr[7];
    volatile const float *vc = cost;
    if (L & kLeafFlag) { float v = half_area(a, b) * c_prim; for (int k = 0; k < 7; ++k) cl[k] = v; }
    else for (int k = 0; k < 7; ++k) cl[k] = vc[(size_t) L * 7 + k];
    if (R & kLeafFlag) { float v = half_area(c, e) * c_prim; for (int k = 0; k < 7; ++k) cr[k] = v; }
    else for (int k = 0; k < 7; ++k) cr[k] = vc[(size_t) R * 7 + k];
    float dist[9]; uint32_t dk[9];
    for (int j = 2; j <= 8; ++j) {
        float best = FLT_MAX; uint32_t bk = 1;
        for (int k = 1; k < j; ++k) { float v = cl[k - 1] + cr[j - k - 1]; if (v < best) { best = v; bk = (uint32_t) k; } }
        dist[j] = best; dk[j] = bk;
    }
    const float A = half_area(lo, hi);
    const float c_int = A * c_node + dist[8], c_leaf = count <= kLeafMax ? A * (float) count * c_prim : FLT_MAX;
    float out[7]; uint32_t d[8];
    out[0] = fminf(c_leaf, c_int); d[0] = c_leaf <= c_int ? 0u : 1u;
    for (int j = 2; j <= 7; ++j) {
        if (dist[j] < out[j - 2]) { out[j - 1] = dist[j]; d[j - 1] = dk[j]; } else { out[j - 1] = out[j - 2]; d[j - 1] = 0u; }
    }
    d[7] = dk[8];
    for (int k = 0; k < 7; ++k) cost[(size_t) p * 7 + k] = out[k];
    dec[p] = make_uint2(d[0] | (d[1] << 8) | (d[2] << 16) | (d[3] << 24), d[4] | (d[5] << 8) | (d[6] << 16) | (d[7] << 24));
}

__global__ void k_refit(const float4 *__restrict__ blo, const float4 *__restrict__ bhi, const uint32_t *__restrict__ sorted, int n,
                        const uint32_t *__restrict__ left, const uint32_t *__restrict__ right, const uint32_t *__restrict__ parent,
                        const uint32_t *__restrict__ cnt, float4 *__restrict__ nlo, float4 *__restrict__ nhi, uint32_t *__restrict__ flags,
                        float *__restrict__ cost, uint2 *__restrict__ dec, float c_node, float c_prim) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t tri = sorted[i];
    float4 lo = blo[tri], hi = bhi[tri];
    nlo[(n - 1) + i] = lo; nhi[(n - 1) + i] = hi;
    uint32_t p = parent[(n - 1) + i];
    while (p != 0xffffffffu) {
        __threadfence();
        if (atomicAdd(&flags[p], 1u) == 0u) return;
        uint32_t L = left[p], R = right[p];
        uint32_t li = (L & kLeafFlag) ? (n - 1) + (L & ~kLeafFlag) : L, ri = (R & kLeafFlag) ? (n - 1) + (R & ~kLeafFlag) : R;
        volatile const float4 *vlo = nlo, *vhi = nhi;
        float4 a, b, c, e;
        a.x = vlo[li].x; a.y = vlo[li].y; a.z = vlo[li].z; b.x = vhi[li].x; b.y = vhi[li].y; b.z = vhi[li].z;
        c.x = vlo[ri].x; c.y = vlo[ri].y; c.z = vlo[ri].z; e.x = vhi[ri].x; e.y = vhi[ri].y; e.z = vhi[ri].z;
        lo = make_float4(fminf(a.x, c.x), fminf(a.y, c.y), fminf(a.z, c.z), 0.f);
        hi = make_float4(fmaxf(b.x, e.x), fmaxf(b.y, e.y), fmaxf(b.z, e.z), 0.f);
        nlo[p] = lo; nhi[p] = hi;
        dp_node(p, L, R, a, b, c, e, lo, hi, cnt[p], cost, dec, c_node, c_prim);
        p = parent[p];
    }
}

// ---- 4b. PLOC (Meister & Bittner 2018, "Parallel locally-ordered clustering for BVH construction"): the clusters stay
// in Morton order; every iteration each cluster finds the neighbour within +-radius whose union with it has the least
// surface area, mutual pairs merge into a new binary node, and the cluster array is compacted.  Ties are broken by a
// symmetric hash of the pair, so regular meshes (equal areas everywhere) still merge a constant fraction per iteration.
// Internal node indices are handed out downwards, the last merge (the root) gets index 0 as in the Karras tree.
B2_DEV uint32_t pair_hash(uint32_t a, uint32_t b) {
    uint32_t lo = min(a, b), hi = max(a, b), h = lo * 0x9E3779B1u ^ (hi + 0x7F4A7C15u) * 0x85EBCA6Bu;
    h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12; return h;
}
__global__ void k_ploc_init(const float4 *__restrict__ blo, const float4 *__restrict__ bhi, const uint32_t *__restrict__ sorted, uint32_t n,
                            float4 *__restrict__ nlo, float4 *__restrict__ nhi, uint32_t *__restrict__ cid) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t tri = sorted[i];
    nlo[(n - 1) + i] = blo[tri]; nhi[(n - 1) + i] = bhi[tri];
    cid[i] = i | kLeafFlag;
}
__global__ void k_ploc_nn(const uint32_t *__restrict__ cid, uint32_t nc, uint32_t n, const float4 *__restrict__ nlo, const float4 *__restrict__ nhi,
                          int radius, uint32_t *__restrict__ nn) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nc) return;
    const uint32_t ci = cid[i], bi = ref_index(ci, n);
    const float4 l = nlo[bi], h = nhi[bi];
    unsigned long long best = ~0ull; uint32_t bj = i;
    const int j0 = max(0, (int) i - radius), j1 = min((int) nc - 1, (int) i + radius);
    for (int j = j0; j <= j1; ++j) {
        if (j == (int) i) continue;
        const uint32_t cj = cid[j], bjx = ref_index(cj, n);
        const float4 l2 = nlo[bjx], h2 = nhi[bjx];
        const float dx = fmaxf(h.x, h2.x) - fminf(l.x, l2.x), dy = fmaxf(h.y, h2.y) - fminf(l.y, l2.y), dz = fmaxf(h.z, h2.z) - fminf(l.z, l2.z);
        const float area = dx * dy + dy * dz + dz * dx;
        const unsigned long long key = ((unsigned long long) __float_as_uint(area) << 32) | pair_hash(ci, cj);
        if (key < best) { best = key; bj = (uint32_t) j; }
    }
    nn[i] = bj;
}
__global__ void k_ploc_merge(uint32_t *__restrict__ cid, uint32_t nc, uint32_t n, const uint32_t *__restrict__ nn,
                             uint32_t *__restrict__ left, uint32_t *__restrict__ right, uint32_t *__restrict__ cnt,
                             float4 *__restrict__ nlo, float4 *__restrict__ nhi, uint32_t *__restrict__ merge_counter,
                             float *__restrict__ cost, uint2 *__restrict__ dec, float c_node, float c_prim, uint32_t *__restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nc) return;
    const uint32_t j = nn[i];
    const bool mutual = j != i && nn[j] == i;
    if (!mutual) { out[i] = cid[i]; return; }
    if (i > j) { out[i] = 0xffffffffu; return; }
    const uint32_t L = cid[i], R = cid[j], li = ref_index(L, n), ri = ref_index(R, n);
    const uint32_t p = (n - 2u) - atomicAdd(merge_counter, 1u);
    const float4 a = nlo[li], b = nhi[li], c = nlo[ri], e = nhi[ri];
    const float4 lo = make_float4(fminf(a.x, c.x), fminf(a.y, c.y), fminf(a.z, c.z), 0.f), hi = make_float4(fmaxf(b.x, e.x), fmaxf(b.y, e.y), fmaxf(b.z, e.z), 0.f);
    const uint32_t count = ((L & kLeafFlag) ? 1u : cnt[L]) + ((R & kLeafFlag) ? 1u : cnt[R]);
    left[p] = L; right[p] = R; cnt[p] = count; nlo[p] = lo; nhi[p] = hi;
    dp_node(p, L, R, a, b, c, e, lo, hi, count, cost, dec, c_node, c_prim);
    out[i] = p;
}
struct NotRemoved { __device__ bool operator()(uint32_t v) const { return v != 0xffffffffu; } };

// ---- 5. collapse one level: wide node w <- binary subtree wide_to_binary[w]
struct CollapseArgs {
    const float *pos; const uint32_t *faces; const uint32_t *face_mesh; const uint32_t *sorted;
    const uint32_t *left, *right, *cnt; const float4 *nlo, *nhi; const uint2 *dec;   // dec == nullptr: greedy collapse
    uint32_t *wide_to_binary; uint4 *nodes; float4 *tris; uint32_t *counters;   // [0] nodes, [1] tris
    uint32_t n, max_nodes;
};

B2_DEV uint32_t ref_count(const CollapseArgs &a, uint32_t ref) {
    if (ref & kLeafFlag) return 1u;
    return a.cnt[ref];
}
B2_DEV uint32_t exp_byte_for(float extent) {
    // smallest e with 2^(e-127) * 254 >= extent
    float x = extent / 254.f;
    uint32_t b = __float_as_uint(x), e = (b >> 23) & 0xffu;
    if (b & 0x7fffffu) e += 1u;
    return min(max(e, 1u), 254u);
}

__global__ void k_collapse(CollapseArgs a, uint32_t begin, uint32_t end) {
    uint32_t w = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= end) return;
    const uint32_t n = a.n;
    uint32_t root = a.wide_to_binary[w];
    uint32_t ch[8]; int nc = 0; uint32_t leaf_bits = 0;       // leaf_bits: child i is a leaf (<= kLeafMax triangles)
    auto dec_of = [&](uint32_t x, int j) { uint2 d = a.dec[x]; return ((j <= 4 ? d.x : d.y) >> (((j - 1) & 3) * 8)) & 0xffu; };
    if (a.dec && !(root & kLeafFlag)) {
        // follow the DP decisions: (binary node, child budget) pairs; the root is forced to distribute 8
        uint32_t sx[8]; int sj[8]; int sp = 0;
        sx[sp] = root; sj[sp++] = 8; bool forced = true;
        while (sp) {
            uint32_t x = sx[--sp]; int j = sj[sp];
            if (x & kLeafFlag) { leaf_bits |= 1u << nc; ch[nc++] = x; continue; }
            uint32_t k;
            if (forced) { k = dec_of(x, 8); forced = false; }
            else {
                while (j > 1 && dec_of(x, j) == 0u) --j;
                if (j == 1) { if (dec_of(x, 1) == 0u) leaf_bits |= 1u << nc; ch[nc++] = x; continue; }
                k = dec_of(x, j);
            }
            sx[sp] = a.right[x]; sj[sp++] = j - (int) k;
            sx[sp] = a.left[x]; sj[sp++] = (int) k;
        }
    } else {
        nc = 1; ch[0] = root;
        // greedy: open the child with the largest surface area until 8 children
        while (nc < 8) {
            int best = -1; float best_area = -1.f;
            for (int i = 0; i < nc; ++i) {
                if (ref_count(a, ch[i]) <= kLeafMax) continue;
                uint32_t bi = ref_index(ch[i], n);
                float4 l = a.nlo[bi], h = a.nhi[bi];
                float dx = h.x - l.x, dy = h.y - l.y, dz = h.z - l.z, area = dx * dy + dy * dz + dz * dx;
                if (area > best_area) { best_area = area; best = i; }
            }
            if (best < 0) break;
            uint32_t b = ch[best];
            ch[best] = a.left[b]; ch[nc++] = a.right[b];
        }
        for (int i = 0; i < nc; ++i) if (ref_count(a, ch[i]) <= kLeafMax) leaf_bits |= 1u << i;
    }
    // node frame
    uint32_t rbi = ref_index(root, n);
    float4 nl = a.nlo[rbi], nh = a.nhi[rbi];
    float lo[3] = { nl.x, nl.y, nl.z }, hi[3] = { nh.x, nh.y, nh.z };
    float ext_max = fmaxf(fmaxf(hi[0] - lo[0], hi[1] - lo[1]), hi[2] - lo[2]);
    float origin[3], scale[3], padabs[3]; uint32_t eb[3];
    for (int k = 0; k < 3; ++k) {
        // padabs covers fp32 rounding of ray-origin-relative positions; pad >= padabs + the quantisation margin of the children
        padabs[k] = fmaxf(fmaxf(fabsf(lo[k]), fabsf(hi[k])), ext_max) * 2e-6f + 1e-30f;
        float pad = 1.1f * padabs[k] + (hi[k] - lo[k]) * (1.f / 3000.f);
        origin[k] = lo[k] - pad;
        eb[k] = exp_byte_for((hi[k] + pad) - origin[k]);
        scale[k] = __uint_as_float(eb[k] << 23);
    }
    // slot assignment: greedy matching of child centroids to octant directions
    float cx[8], cy[8], cz[8];
    float mx = 0.5f * (lo[0] + hi[0]), my = 0.5f * (lo[1] + hi[1]), mz = 0.5f * (lo[2] + hi[2]);
    for (int i = 0; i < nc; ++i) {
        uint32_t bi = ref_index(ch[i], n); float4 l = a.nlo[bi], h = a.nhi[bi];
        cx[i] = 0.5f * (l.x + h.x) - mx; cy[i] = 0.5f * (l.y + h.y) - my; cz[i] = 0.5f * (l.z + h.z) - mz;
    }
    int slot_child[8]; for (int s = 0; s < 8; ++s) slot_child[s] = -1;
    uint32_t child_done = 0, slot_done = 0;
    for (int it = 0; it < nc; ++it) {
        float bc = -FLT_MAX; int bi = 0, bs = 0;
        for (int i = 0; i < nc; ++i) {
            if (child_done >> i & 1u) continue;
            for (int s = 0; s < 8; ++s) {
                if (slot_done >> s & 1u) continue;
                float c = (s & 1 ? cx[i] : -cx[i]) + (s & 2 ? cy[i] : -cy[i]) + (s & 4 ? cz[i] : -cz[i]);
                if (c > bc) { bc = c; bi = i; bs = s; }
            }
        }
        slot_child[bs] = bi; child_done |= 1u << bi; slot_done |= 1u << bs;
    }
    // allocate children
    uint32_t k_int = 0, k_tri = 0;
    for (int i = 0; i < nc; ++i) { if (leaf_bits >> i & 1u) k_tri += ref_count(a, ch[i]); else k_int++; }
    uint32_t child_base = k_int ? atomicAdd(&a.counters[0], k_int) : 0u;
    uint32_t tri_base = k_tri ? atomicAdd(&a.counters[1], k_tri) : 0u;
    if (k_int && child_base + k_int > a.max_nodes) { a.counters[2] = 1u; return; }   // cannot happen (max_nodes bound), flagged anyway
    uint32_t meta[8], qlo[3][8], qhi[3][8], imask = 0, ni = 0, nt = 0;
    for (int s = 0; s < 8; ++s) {
        meta[s] = 0; for (int k = 0; k < 3; ++k) { qlo[k][s] = 0; qhi[k][s] = 0; }
        int i = slot_child[s]; if (i < 0) continue;
        uint32_t ref = ch[i], bi = ref_index(ref, n), cnt = ref_count(a, ref);
        float4 l = a.nlo[bi], h = a.nhi[bi];
        float cl[3] = { l.x, l.y, l.z }, chh[3] = { h.x, h.y, h.z };
        for (int k = 0; k < 3; ++k) {
            // planes sit padabs + kQMargin cells outside the child box: covers position rounding (as the node frame does)
            // and the <= 2^-9-cell plane shift of the biased-byte slab test (b2_bvh.cuh, qbyte_biased)
            const float m = kQMargin * scale[k] + padabs[k];
            float ql = floorf((cl[k] - m - origin[k]) / scale[k]), qh = ceilf((chh[k] + m - origin[k]) / scale[k]);
            ql = fminf(fmaxf(ql, 0.f), 255.f); qh = fminf(fmaxf(qh, 0.f), 255.f);
            while (ql > 0.f && origin[k] + ql * scale[k] > cl[k] - m) ql -= 1.f;
            while (qh < 255.f && origin[k] + qh * scale[k] < chh[k] + m) qh += 1.f;
            qlo[k][s] = (uint32_t) ql; qhi[k][s] = (uint32_t) qh;
        }
        if (!(leaf_bits >> i & 1u)) {
            imask |= 1u << s; meta[s] = 0x20u | (24u + (uint32_t) s);
            a.wide_to_binary[child_base + ni] = ref; ni++;
        } else {
            meta[s] = (((1u << cnt) - 1u) << 5) | nt;
            // the <= kLeafMax triangles below ref, left to right (= Morton order for the Karras tree)
            uint32_t lpos[kLeafMax], lst[kLeafMax]; int lsp = 0; uint32_t ln = 0; lst[lsp++] = ref;
            while (lsp) { uint32_t x = lst[--lsp]; if (x & kLeafFlag) lpos[ln++] = x & ~kLeafFlag; else { lst[lsp++] = a.right[x]; lst[lsp++] = a.left[x]; } }
            for (uint32_t j = 0; j < cnt; ++j) {
                uint32_t face = a.sorted[lpos[j]];
                uint32_t i0 = a.faces[3 * (size_t) face], i1 = a.faces[3 * (size_t) face + 1], i2 = a.faces[3 * (size_t) face + 2];
                f3 p0(a.pos[3 * (size_t) i0], a.pos[3 * (size_t) i0 + 1], a.pos[3 * (size_t) i0 + 2]);
                f3 p1(a.pos[3 * (size_t) i1], a.pos[3 * (size_t) i1 + 1], a.pos[3 * (size_t) i1 + 2]);
                f3 p2(a.pos[3 * (size_t) i2], a.pos[3 * (size_t) i2 + 1], a.pos[3 * (size_t) i2 + 2]);
                f3 e1 = p1 - p0, e2 = p2 - p0;
                float4 *tp = a.tris + (size_t) (tri_base + nt + j) * 3;
                tp[0] = make_float4(p0.x, p0.y, p0.z, __uint_as_float(face));
                tp[1] = make_float4(e1.x, e1.y, e1.z, __uint_as_float(a.face_mesh[face]));
                tp[2] = make_float4(e2.x, e2.y, e2.z, 0.f);
            }
            nt += cnt;
        }
    }
    auto pack4 = [](const uint32_t *v) { return v[0] | (v[1] << 8) | (v[2] << 16) | (v[3] << 24); };
    uint4 *np = a.nodes + (size_t) w * 5;
    np[0] = make_uint4(__float_as_uint(origin[0]), __float_as_uint(origin[1]), __float_as_uint(origin[2]), eb[0] | (eb[1] << 8) | (eb[2] << 16) | (imask << 24));
    np[1] = make_uint4(child_base, tri_base, pack4(meta), pack4(meta + 4));
    np[2] = make_uint4(pack4(qlo[0]), pack4(qlo[0] + 4), pack4(qlo[1]), pack4(qlo[1] + 4));
    np[3] = make_uint4(pack4(qlo[2]), pack4(qlo[2] + 4), pack4(qhi[0]), pack4(qhi[0] + 4));
    np[4] = make_uint4(pack4(qhi[1]), pack4(qhi[1] + 4), pack4(qhi[2]), pack4(qhi[2] + 4));
}

__global__ void k_init_bounds(uint32_t *b) { if (threadIdx.x < 3) b[threadIdx.x] = 0xffffffffu; else if (threadIdx.x < 6) b[threadIdx.x] = 0u; }

void bvh_free(Bvh &b) {
    if (b.nodes) cudaFree(b.nodes);
    if (b.tris) cudaFree(b.tris);
    b = Bvh();
}

namespace {
// scratch allocations of one build; everything not released() is freed on scope exit
struct Scratch {
    std::vector<void *> ptrs; cudaEvent_t ev[4] = { nullptr, nullptr, nullptr, nullptr };
    template <typename T> cudaError_t alloc(T *&p, size_t bytes) {
        void *q = nullptr; cudaError_t e = cudaMalloc(&q, bytes ? bytes : 16); p = (T *) q; if (e == cudaSuccess) ptrs.push_back(q); return e;
    }
    void release(void *p) { for (auto &q : ptrs) if (q == p) q = nullptr; }
    ~Scratch() { for (void *q : ptrs) if (q) cudaFree(q); for (auto &e : ev) if (e) cudaEventDestroy(e); }
};
}  // namespace

bool bvh_build(const float *pos, const uint32_t *faces, const uint32_t *face_mesh, uint32_t n, int builder, cudaStream_t st, Bvh &out, std::string &err) {
    out = Bvh();
    if (n == 0) return true;
    Scratch S;
    const int T = 256; const uint32_t G = (n + T - 1) / T;
    float4 *blo, *bhi, *nlo, *nhi, *tris;
    uint32_t *scene_bounds, *vals, *vals2, *left, *right, *parent, *flags, *w2b, *counters;
    uint64_t *keys, *keys2; uint32_t *cnt; uint4 *nodes, *nodes_final; void *tmp; size_t tmp_bytes = 0;
    const uint32_t max_nodes = n + 2;     // a wide node per binary internal node at most
    float *cost; uint2 *dec;
    // collapse policy knobs (tuning only; defaults are the measured optimum, profiles/README.md)
    const char *e_mode = getenv("B2_BVH_COLLAPSE"), *e_cn = getenv("B2_BVH_CNODE"), *e_cp = getenv("B2_BVH_CPRIM");
    const bool use_dp = !(e_mode && std::string(e_mode) == "greedy");
    const float c_node = e_cn ? (float) atof(e_cn) : 1.f, c_prim = e_cp ? (float) atof(e_cp) : 0.3f;
    uint32_t level_begin = 0, level_end = 1, depth = 0, host_counters[3] = { 1, 0, 0 };
    for (auto &e : S.ev) CK(cudaEventCreate(&e));
    CK(S.alloc(blo, sizeof(float4) * n)); CK(S.alloc(bhi, sizeof(float4) * n));
    CK(S.alloc(scene_bounds, 6 * sizeof(uint32_t)));
    CK(S.alloc(keys, sizeof(uint64_t) * n)); CK(S.alloc(keys2, sizeof(uint64_t) * n));
    CK(S.alloc(vals, sizeof(uint32_t) * n)); CK(S.alloc(vals2, sizeof(uint32_t) * n));
    CK(cudaEventRecord(S.ev[0], st));
    k_init_bounds<<<1, 32, 0, st>>>(scene_bounds);
    k_tri_bounds<<<G, T, 0, st>>>(pos, faces, n, blo, bhi, scene_bounds);
    k_morton<<<G, T, 0, st>>>(blo, bhi, n, scene_bounds, keys, vals);
    {
        cub::DoubleBuffer<uint64_t> dk(keys, keys2); cub::DoubleBuffer<uint32_t> dv(vals, vals2);
        CK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, dk, dv, (int) n, 0, 63, st));
        CK(S.alloc(tmp, tmp_bytes));
        CK(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, dk, dv, (int) n, 0, 63, st));
        if (dk.Current() != keys) std::swap(keys, keys2);
        if (dv.Current() != vals) std::swap(vals, vals2);
    }
    CK(cudaEventRecord(S.ev[1], st));
    // binary tree (Karras radix tree + refit, or PLOC) with the collapse DP
    CK(S.alloc(left, sizeof(uint32_t) * n)); CK(S.alloc(right, sizeof(uint32_t) * n)); CK(S.alloc(cnt, sizeof(uint32_t) * n));
    CK(S.alloc(nlo, sizeof(float4) * 2 * (size_t) n)); CK(S.alloc(nhi, sizeof(float4) * 2 * (size_t) n));
    CK(S.alloc(cost, sizeof(float) * 7 * (size_t) n)); CK(S.alloc(dec, sizeof(uint2) * (size_t) n));
    if (builder == 0 || n < 2) {
        CK(S.alloc(parent, sizeof(uint32_t) * 2 * (size_t) n));
        CK(S.alloc(flags, sizeof(uint32_t) * n)); CK(cudaMemsetAsync(flags, 0, sizeof(uint32_t) * n, st));
        if (n > 1) k_karras<<<G, T, 0, st>>>(keys, (int) n, left, right, parent, cnt);
        else { CK(cudaMemsetAsync(parent, 0xff, 4, st)); }
        k_refit<<<G, T, 0, st>>>(blo, bhi, vals, (int) n, left, right, parent, cnt, nlo, nhi, flags, cost, dec, c_node, c_prim);
    } else {
        const char *e_r = getenv("B2_PLOC_RADIUS");
        const int radius = std::max(1, e_r ? atoi(e_r) : 10);
        uint32_t *cid, *cid2, *nn, *merged, *d_sel; void *sel_tmp; size_t sel_bytes = 0;
        CK(S.alloc(cid, sizeof(uint32_t) * n)); CK(S.alloc(cid2, sizeof(uint32_t) * n)); CK(S.alloc(nn, sizeof(uint32_t) * n));
        CK(S.alloc(merged, sizeof(uint32_t) * n)); CK(S.alloc(d_sel, 2 * sizeof(uint32_t)));
        CK(cub::DeviceSelect::If(nullptr, sel_bytes, merged, cid2, d_sel, (int) n, NotRemoved(), st));
        CK(S.alloc(sel_tmp, sel_bytes));
        CK(cudaMemsetAsync(d_sel, 0, 2 * sizeof(uint32_t), st));
        k_ploc_init<<<G, T, 0, st>>>(blo, bhi, vals, n, nlo, nhi, cid);
        uint32_t nc = n, iters = 0;
        while (nc > 1) {
            const uint32_t g = (nc + T - 1) / T;
            k_ploc_nn<<<g, T, 0, st>>>(cid, nc, n, nlo, nhi, radius, nn);
            k_ploc_merge<<<g, T, 0, st>>>(cid, nc, n, nn, left, right, cnt, nlo, nhi, d_sel + 1, cost, dec, c_node, c_prim, merged);
            CK(cub::DeviceSelect::If(sel_tmp, sel_bytes, merged, cid2, d_sel, (int) nc, NotRemoved(), st));
            uint32_t nc_new = 0;
            CK(cudaMemcpyAsync(&nc_new, d_sel, 4, cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            if (nc_new >= nc || nc_new == 0) { err = "bvh_build: PLOC made no progress"; return false; }
            nc = nc_new; std::swap(cid, cid2);
            if (++iters > 100000u) { err = "bvh_build: PLOC iteration limit"; return false; }
        }
    }
    CK(cudaEventRecord(S.ev[2], st));
    // collapse, one launch per level of the wide tree
    CK(S.alloc(nodes, sizeof(uint4) * 5 * (size_t) max_nodes)); CK(S.alloc(tris, sizeof(float4) * 3 * (size_t) n));
    CK(S.alloc(w2b, sizeof(uint32_t) * max_nodes)); CK(S.alloc(counters, 3 * sizeof(uint32_t)));
    CK(cudaMemcpyAsync(counters, host_counters, sizeof(host_counters), cudaMemcpyHostToDevice, st));
    { uint32_t root_ref = n > 1 ? 0u : (0u | kLeafFlag); CK(cudaMemcpyAsync(w2b, &root_ref, 4, cudaMemcpyHostToDevice, st)); CK(cudaStreamSynchronize(st)); }
    {
        CollapseArgs ca{ pos, faces, face_mesh, vals, left, right, cnt, nlo, nhi, use_dp ? dec : nullptr, w2b, nodes, tris, counters, n, max_nodes };
        while (level_begin < level_end) {
            uint32_t cnt = level_end - level_begin;
            k_collapse<<<(cnt + 127) / 128, 128, 0, st>>>(ca, level_begin, level_end);
            CK(cudaMemcpyAsync(host_counters, counters, sizeof(host_counters), cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            if (host_counters[2]) { err = "bvh_build: wide node pool exhausted"; return false; }
            level_begin = level_end; level_end = host_counters[0]; depth++;
        }
    }
    float root_cost = 0.f; float4 root_lo = make_float4(0, 0, 0, 0), root_hi = root_lo;
    if (n > 1) {
        CK(cudaMemcpyAsync(&root_cost, cost, 4, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(&root_lo, nlo, 16, cudaMemcpyDeviceToHost, st)); CK(cudaMemcpyAsync(&root_hi, nhi, 16, cudaMemcpyDeviceToHost, st));
    }
    CK(S.alloc(nodes_final, sizeof(uint4) * 5 * (size_t) host_counters[0]));
    CK(cudaMemcpyAsync(nodes_final, nodes, sizeof(uint4) * 5 * (size_t) host_counters[0], cudaMemcpyDeviceToDevice, st));
    CK(cudaEventRecord(S.ev[3], st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    if (host_counters[1] != n) { err = "bvh_build: triangle count mismatch after collapse"; return false; }
    out.nodes = nodes_final; out.tris = tris; out.n_nodes = host_counters[0]; out.n_tris = n; out.depth = depth;
    out.node_bytes = sizeof(uint4) * 5 * (size_t) out.n_nodes; out.tri_bytes = sizeof(float4) * 3 * (size_t) n;
    cudaEventElapsedTime(&out.ms_morton_sort, S.ev[0], S.ev[1]); cudaEventElapsedTime(&out.ms_lbvh, S.ev[1], S.ev[2]);
    cudaEventElapsedTime(&out.ms_collapse, S.ev[2], S.ev[3]); cudaEventElapsedTime(&out.ms_total, S.ev[0], S.ev[3]);
    {   // SAH cost of the collapsed tree (the DP's optimum at the root), in units of c_node per unit of root area
        float dx = root_hi.x - root_lo.x, dy = root_hi.y - root_lo.y, dz = root_hi.z - root_lo.z, A = dx * dy + dy * dz + dz * dx;
        out.sah_cost = A > 0.f ? root_cost / A : 0.f; out.builder = (builder == 0 || n < 2) ? 0 : 1;
    }
    S.release(tris); S.release(nodes_final);
    return true;
}

}  // namespace b2
