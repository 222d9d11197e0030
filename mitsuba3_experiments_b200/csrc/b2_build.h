// b2_build.h -- host interface of the GPU BVH builder (bvh_build.cu)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>

namespace b2 {

struct Bvh {
    uint4 *nodes = nullptr;     // 5 x uint4 per wide node (device)
    float4 *tris = nullptr;     // 3 x float4 per triangle, leaf order (device)
    uint32_t n_nodes = 0, n_tris = 0, depth = 0;
    size_t node_bytes = 0, tri_bytes = 0;
    float ms_morton_sort = 0, ms_lbvh = 0, ms_collapse = 0, ms_total = 0;
    float sah_cost = 0;         // SAH cost of the wide tree / root area (c_node = 1)
    int builder = 0;            // builder that produced this tree
};

// pos / faces / face_mesh are DEVICE pointers (Mesh buffer layout: fp32 xyz, u32 triangle indices, u32 mesh id per face)
// builder: 0 = Karras LBVH, 1 = PLOC (both over the Morton order, both collapsed to the compressed 8-wide BVH by the SAH DP)
bool bvh_build(const float *pos, const uint32_t *faces, const uint32_t *face_mesh, uint32_t n_faces, int builder, cudaStream_t stream, Bvh &out, std::string &err);
void bvh_free(Bvh &b);

// rough-plastic tables (tables.cu): 64 external transmittances + the internal diffuse reflectance, computed on the device
bool rough_tables_build(int distribution, float alpha, float eta, float *trans64_host, float *internal_reflectance, std::string &err);

}  // namespace b2
