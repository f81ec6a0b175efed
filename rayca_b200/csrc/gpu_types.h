// gpu_types.h — the flattened scene as it lives in HBM (see DESIGN.md "Data layout").
//
// nodes     : 32 B records {min.xyz, link0 | max.xyz, link1}; link1 != 0 => leaf over triangles
//             [link0, link0+link1); link1 == 0 => inner, children are the ADJACENT records
//             link0 and link0+1 (one 64 B aligned fetch), mirroring the reference numbering
//             "left = idx, right = idx+1" (rayca-soft/src/bvh/blas.rs:38-50,302-308).
// tri_pos   : 48 B per triangle, 3 x float4 world-space vertices (w lanes: src id, material, blas).
// tri_attr  : 192 B per triangle, 12 x float4, touched only when a hit is shaded.
#pragma once
#include <cstdint>

#include <cuda_runtime.h>

namespace rb {

struct alignas(16) GpuNode {
    float min_x, min_y, min_z;
    uint32_t link0;
    float max_x, max_y, max_z;
    uint32_t link1;
};
static_assert(sizeof(GpuNode) == 32, "node record is 32 bytes");

struct alignas(16) GpuMaterial {
    float a[4];  // pbr base colour | phong ambient | ggx diffuse
    float b[4];  // phong emission
    float c[4];  // phong diffuse
    float d[4];  // phong / ggx specular
    float p0;    // pbr metallic factor | phong shininess | ggx roughness
    float p1;    // pbr roughness factor
    int32_t tex_albedo, tex_normal, tex_mr; // global image index or -1
    uint32_t kind;                          // RAYCA_MATERIAL_*
    uint32_t emissive;                      // phong: !emission.close(BLACK) (phong.rs:50-52)
    uint32_t pad;
};
static_assert(sizeof(GpuMaterial) == 96, "material record");

struct GpuImage {
    unsigned long long offset; // byte offset into the texel buffer
    uint32_t width, height;
    uint32_t color_type;
    uint32_t pad;
};

struct alignas(16) GpuLight {
    float color[4];
    float intensity;
    float att[3];
    float pos[3];   // Trs::get_translation of the light node's LOCAL trs (SURVEY A-13)
    uint32_t kind;
    float dir[3];   // directional: -rotate(X_AXIS, node rotation) (light/directional.rs:47-51)
    uint32_t pad;
    // quad (rayca-model/src/light/quad.rs): corner a = pos, edges ab / ac, normalize(ab x ac), get_area()
    float ab[3];
    float area;
    float ac[3];
    uint32_t pad2;
    float normal[3];
    uint32_t pad3;
    // the two "wide" triangles QuadLight::intersects tests (quad.rs:135-158), corners a, b, c, d moved to world space
    // with the light node's LOCAL trs: triangles (a, c, b) and (a, d, c); read by the MIS light pdf only
    float wide[4][4];
};

struct GpuCamera {
    float t[3];
    float q[4];
    float s[3];
    float angle; // tan(yfov/2), camera.rs:74-76
};

// Kernel-visible scene (all pointers are device pointers).
struct GpuSceneView {
    const float4* nodes;     // 2 float4 per node
    const float4* tri_pos;   // 3 float4 per triangle
    const float4* tri_attr;  // 12 float4 per triangle
    const GpuMaterial* materials;
    const GpuImage* images;
    const unsigned char* texels;
    const GpuLight* lights;
    const uint32_t* blas_prim_offset; // blas_count + 1 prefix sums (global triangle index -> blas)
    uint32_t blas_count;
    uint32_t light_count;
    uint32_t node_count;
    uint32_t tri_count;
    // root record (copy of the top node): tested first, then the pair walk starts
    float root_min[3];
    uint32_t root_link0;
    float root_max[3];
    uint32_t root_link1;
    GpuCamera camera;
    // performance layer (fast_bvh.cpp): one binary tree over all triangles, sibling pairs adjacent;
    // a child reference (link0) is a pair index or, with bit 31 set, a packed leaf {first, count-1}.
    // fast_tri holds 4 float4 (64 B) per triangle in leaf order:
    //   {n.xyz, d | v0, reference-global triangle index | v1, reference leaf (index into `nodes`)
    //    whose box must also accept the ray | v2, dot(n,n)}
    const float4* fast_nodes;
    const float4* fast_tri;
    float fast_root_min[3];
    uint32_t fast_root_ref;
    float fast_root_max[3];
    uint32_t fast_record_count; // 32-byte records allocated behind fast_nodes (RAYCA_TREELET stages the first ones)
};

} // namespace rb
