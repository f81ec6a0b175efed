// gpu_fast_build.h — performance tree built on the device (gpu_fast_build.cu); same records as fast_bvh.cpp.
#pragma once
#include <string>

#include "host_scene.h"

namespace rb {

struct FastTreeDevice {
    void* nodes = nullptr;  // GpuNode pairs (cudaMalloc); owned by the caller after a successful build
    void* tri = nullptr;    // 4 x float4 per primitive in leaf order
    GpuNode root{};         // link0: pair index 0 or a packed leaf
    uint32_t depth = 0;     // deepest leaf
    uint32_t levels = 0;    // level-synchronous passes over big nodes
    uint32_t small_subtrees = 0;
    float build_ms = 0.0f;  // device time, CUDA events
};

// Inputs: the packed host scene WITHOUT its performance tree (HostScene::{blas, blas_prim_offset,
// blas_node_offset} final) and the device copy of HostScene::tri_pos.  Returns RAYCA_OK or an error code.
// `d_ref_nodes`: the device copy of HostScene::nodes when there is one — the reference leaf of every primitive is then
// looked up on the device (triangle-only scenes); nullptr: on the host (reference_leaves).
int build_fast_tree_device(const HostScene& hs, const float4* d_tri_pos, cudaStream_t stream, FastTreeDevice& out, std::string& err,
                           const void* d_ref_nodes = nullptr);
void reference_leaves(const HostScene& hs, PodVector<uint32_t>& ref_leaf);

// Primitive bounds / centroids already on the device (emit_mesh_device): the build reads them there, leaves its
// primitive order in `order_out` (device, n entries) and in HostBlas::order, and does not touch HostBlas::tris.
struct DeviceBlasInput {
    uint32_t n = 0;
    const float4 *mn = nullptr, *mx = nullptr, *cen = nullptr;
    uint32_t* order_out = nullptr;
};
// Reference BLAS (bvh_build.cpp's build_blas) on the device: same nodes, same primitive order.  RAYCA_E_UNSUPPORTED
// asks the caller to use the host builder (a level broader than the device work lists).
int build_blas_device(HostBlas& blas, uint32_t max_depth, cudaStream_t stream, std::string& err, float* device_ms,
                      const DeviceBlasInput* device_input = nullptr);

// ---- mesh triangles emitted on the device (scene_create's fast path for models made of triangle meshes only) ----------
// One mesh primitive of a model: the triangles [base, base + ntri) of the emit order, the world TRS of its node.
struct DeviceMeshJob {
    uint32_t base, ntri, first_index, first_vertex, vertex_count, material, pad0, pad1;
    float t[4], q[4], s[4]; // translation, rotation quaternion (x, y, z, w), scale
};
struct DeviceBlasMesh {
    uint32_t n = 0, job_count = 0;
    float4* pos = nullptr;      // 3 per triangle, emit order: world-space vertices (Triangle::get_vertex)
    float4 *mn = nullptr, *mx = nullptr, *cen = nullptr; // bounds and the reference's centroid, emit order
    uint32_t* order = nullptr;  // BLAS position -> emit index, after build_blas_device
    DeviceMeshJob* jobs = nullptr;
    uint32_t* bad = nullptr;    // != 0: a vertex index was out of range
};
// `BvhPrimitive::from_triangle_mesh` (primitive.rs:239-268) + `Triangle::new` for every triangle of the jobs, from the
// caller's vertices / indices on the device: the expressions of host_scene.cpp's emit loop, compiled for the device.
int emit_mesh_device(const std::vector<DeviceMeshJob>& jobs, uint32_t n, const void* d_vertices, const void* d_indices, cudaStream_t stream,
                     DeviceBlasMesh& out, std::string& err);
// tri_pos records of the BLAS (packed order): {v0, emit index | v1, material | v2, blas index} at `d_tri_pos` (3 per primitive)
int gather_tri_pos_device(const DeviceBlasMesh& mesh, uint32_t blas_index, float4* d_tri_pos, cudaStream_t stream);
void free_device_blas_mesh(DeviceBlasMesh& mesh, cudaStream_t stream);

// In-place AoS -> plane-major transposition of the sibling-pair records (the layout the walk reads when the library is
// built with RAYCA_SOA_PAIRS, see trace.cuh).  `node_count` 32-byte records, i.e. node_count / 2 pairs.
int transpose_pairs_device(void* d_nodes, size_t node_count, cudaStream_t stream);

// tunables shared with the host builder (fast_bvh.cpp)
void fast_tree_tunables(int& bins, uint32_t& leaf_max, float& traversal_cost);

} // namespace rb
