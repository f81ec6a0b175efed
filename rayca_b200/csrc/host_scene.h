// host_scene.h — host half of `SoftRenderer::draw` before the pixel loop
// (rayca-soft/src/scene.rs:90-115): scene-graph flattening, per-model primitive emission,
// BLAS/TLAS construction, and packing into the HBM layout of gpu_types.h.
#pragma once
#include <cstdint>
#include <functional>
#include <memory>
#include <type_traits>
#include <utility>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/rayca_cuda.h"
#include "gpu_types.h"
#include "rmath.h"

namespace rb {

// std::vector whose resize() leaves trivially-constructible elements uninitialised: the big per-triangle arrays
// (tens of MB) are written once by a parallel loop right after being sized, and value-initialising them first
// costs a sequential pass of page faults (30 of the 42 ms "emit triangles" took for 262k triangles).
// Big blocks (>= 1 MiB) that a PodVector gives back are kept — up to 32 blocks / a quarter of the physical memory, at most 6 GiB — and handed to the next PodVector
// that asks for that many bytes or a little more: `Draw::draw` rebuilds the scene on every call, and without this the
// same 130 MB are mmap'ed, page-faulted in and munmap'ed per draw (measured: 2 ms to free + the faults inside emit / pack).
void* pod_block_acquire(size_t bytes);            // host_scene.cpp; nullptr: nothing suitable cached
bool pod_block_release(void* p, size_t bytes);    // true: kept (the caller must not free it)
void pod_cache_enable_pinning();                  // blocks entering the cache are page-locked from now on (a CUDA context exists)

template <class T>
struct DefaultInitAllocator : std::allocator<T> {
    template <class U> struct rebind { using other = DefaultInitAllocator<U>; };
    using std::allocator<T>::allocator;
    template <class U> void construct(U* p) noexcept(std::is_nothrow_default_constructible<U>::value) { ::new ((void*)p) U; }
    template <class U, class... A> void construct(U* p, A&&... a) { ::new ((void*)p) U(std::forward<A>(a)...); }
    static constexpr size_t kCacheFrom = size_t(1) << 20;
    T* allocate(size_t n) {
        const size_t bytes = n * sizeof(T);
        if (bytes >= kCacheFrom)
            if (void* p = pod_block_acquire(bytes)) return static_cast<T*>(p);
        return static_cast<T*>(::operator new(bytes));
    }
    void deallocate(T* p, size_t n) noexcept {
        const size_t bytes = n * sizeof(T);
        if (bytes >= kCacheFrom && pod_block_release(p, bytes)) return;
        ::operator delete(p);
    }
};
template <class T> using PodVector = std::vector<T, DefaultInitAllocator<T>>;

// One triangle as emitted by BvhPrimitive::from_triangle_mesh_impl (bvh/primitive.rs:194-237),
// already moved to world space with the reference's `trs * vertex` operation order.
struct HostTri {
    V3 v[3];       // world space
    V3 centroid;   // world-space `trs * ((a+b+c)*0.3333)` used by the SAH partition
    V3 bmin, bmax; // world-space bounds of the three vertices
    uint32_t src;  // emit order inside its model (stable id)
    uint32_t sphere; // 0: triangle; k > 0: HostBlas::spheres[k - 1] (v[] unused, bounds / centroid are the sphere's)
};

// BvhGeometry::Sphere (bvh/primitive.rs:10-13,270-282): the sphere stays in MODEL space, the node's
// world TRS is applied to the ray and to the hit point on every query (rayca-geometry/src/sphere.rs:155-163).
struct HostSphere {
    V3 center;      // model space
    float radius2;  // Sphere::new: radius * radius
    TrsF world;     // world TRS of the owning node
    uint32_t material;
};

struct HostAttr {
    float color[3][4];
    V3 normal[3], tangent[3], bitangent[3];
    float uv[3][2];
    uint32_t material; // index into HostScene::materials
};

struct HostNode {
    V3 bmin, bmax;
    uint32_t offset, count; // count != 0: leaf over [offset, offset+count); else children offset, offset+1
};

struct M3 { // rayca-math/src/mat3.rs, row-major m[i][j]
    float m[3][3];
};

// One triangle-mesh primitive of a BLAS: the triangles [base, base + ntri) of the emit order.  Their shading attributes
// are not materialised per triangle: pack_tri_attrs computes them from the caller's vertices (which must still be
// alive) straight into HostScene::tri_attr, in BLAS order.
struct HostMeshJob {
    size_t base = 0;
    uint32_t ntri = 0;
    uint32_t first_index = 0, first_vertex = 0;
    uint32_t material = 0;
    M3 tangent_matrix{}, normal_matrix{};
};

struct HostBlas {
    PodVector<HostTri> tris;     // reordered in place by the build
    PodVector<HostAttr> attrs;   // indexed by HostTri::src; filled for spheres and quad-light triangles only
    std::vector<HostMeshJob> jobs; // ascending base
    // primitives emitted and kept on the DEVICE (scene_create's mesh-only fast path): `tris` stays empty, the build
    // leaves the count and its primitive order (BLAS position -> emit index) here
    size_t device_prims = 0;
    PodVector<uint32_t> order;
    size_t prim_count() const { return tris.empty() ? device_prims : tris.size(); }
    std::vector<HostSphere> spheres;
    std::vector<HostNode> nodes; // node 0 root, node 1 unused (blas.rs:250-256)
    uint32_t model = 0;
    uint32_t depth = 0;
};

struct HostScene {
    std::vector<HostBlas> blas;          // in TLAS-reordered (blas_nodes) order after build
    std::vector<GpuMaterial> materials;  // [0] = Material::DEFAULT -> PbrMaterial::WHITE
    std::vector<GpuImage> images;
    std::vector<unsigned char> texels;   // gathered copy of all images (only when build_host_scene is asked to gather)
    size_t texel_bytes = 0;              // size of the texel buffer; GpuImage::offset addresses it
    std::vector<GpuLight> lights;
    GpuCamera camera{};
    bool has_camera = false;
    bool has_spheres = false; // the reference-tree walks (traversal modes 1 / 2) handle triangles only

    // packed output
    std::vector<GpuNode> nodes;          // all BLAS nodes (offset-adjusted) followed by TLAS top pairs
    PodVector<float4> tri_pos;
    PodVector<float4> tri_attr;
    std::vector<uint32_t> blas_prim_offset, blas_node_offset;
    GpuNode root{};
    uint32_t max_stack = 0;              // deepest push chain any walk of this scene can need (reference tree and performance tree)
    uint32_t ref_stack = 0;              // ... a walk of the REFERENCE tree (traversal modes 1 / 2): TLAS depth + deepest BLAS + 2
    // performance layer
    std::vector<GpuNode> fast_nodes;
    std::vector<float4> fast_tri;
    GpuNode fast_root{};   // link0 = child reference of the root (pair index or packed leaf)
    uint32_t fast_depth = 0;
};

// Returns RAYCA_OK or a negative RAYCA_E_* code; `err` receives a detail string.
// `fast_tree` = false leaves the performance tree to the device builder (gpu_fast_build.cu).
// `blas_builder`, when given, replaces build_blas (the device builder of gpu_fast_build.cu): it returns true when
// it built the BLAS and false to leave it to the host builder.
// `defer_attrs`: HostScene::tri_attr is sized but not filled; the caller runs pack_tri_attrs itself (the device path
// does, while the performance tree is being built on the GPU).
using BlasBuilder = std::function<bool(HostBlas&, uint32_t max_depth)>;
// `mesh_builder`, when given, is offered every model that consists of triangle meshes only (no spheres, no quad lights)
// BEFORE its triangles are emitted on the host: the mesh primitives with the world TRS of their nodes.  It returns 1 when it
// has emitted the triangles and built the BLAS itself (HostBlas::{nodes, depth, device_prims, order} set, `tris` empty),
// 0 to leave the model to the host emit, or a negative RAYCA_E_* code (detail in `err`).
struct HostMeshEmit {
    HostMeshJob job;
    TrsF world;
    uint32_t vertex_count; // of the geometry: indices are checked against it
};
using MeshBlasBuilder = std::function<int(HostBlas&, const std::vector<HostMeshEmit>&, uint32_t max_depth, std::string& err)>;
int build_host_scene(const rayca_scene_desc& desc, bool bvh, HostScene& out, std::string& err, bool fast_tree = true,
                     const BlasBuilder* blas_builder = nullptr, bool gather_texels = true, bool defer_attrs = false,
                     const MeshBlasBuilder* mesh_builder = nullptr);
// Fills HostScene::tri_attr (12 float4 per primitive, BLAS order) from `desc`'s vertices and HostBlas::{tris, jobs, attrs}.
const HostMeshJob* mesh_job_of(const HostBlas& blas, uint32_t src); // the mesh primitive that emitted triangle `src`, or nullptr
void pack_tri_attrs(const rayca_scene_desc& desc, HostScene& scene);
// ... the primitives [g0, g1) of the packed (reference-global) order into `dst` (12 float4 each): the device path packs
// and uploads the array piece by piece through a small page-locked ring instead of materialising it on the host
void pack_tri_attrs_range(const rayca_scene_desc& desc, const HostScene& scene, size_t g0, size_t g1, float4* dst);

// BLAS build: top-down SAH over 63 planes x 3 axes with the reference's tie-breaking
// (bvh/blas.rs:64-127,245-316), evaluated through per-plane bins so that one pass over the
// node's triangles serves all 189 candidates.  `max_depth` = u8::MAX, or 0 when Config::bvh is off.
void build_blas(HostBlas& blas, uint32_t max_depth);

// Performance layer: SAH tree over all triangles of the packed scene (fast_bvh.cpp).  Needs
// HostScene::{blas, tri_pos, blas_prim_offset, blas_node_offset} to be final.
void build_fast_tree(HostScene& scene);

} // namespace rb
