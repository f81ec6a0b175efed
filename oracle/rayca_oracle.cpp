// rayca_oracle.cpp — CPU restatement of rayca-soft's per-pixel raytracing hot path.
//
// TEST INFRASTRUCTURE ONLY (see oracle_math.h header).  Built by oracle/Makefile into
// oracle/_build/librayca_oracle.so and loaded through ctypes by tests/, smoke() and
// bench.py's CPU legs.  It consumes the same POD scene description as librayca_cuda
// (include/rayca_cuda.h) so one harness drives both.
//
// Parity status: UNPINNED at image level (the reference holds no golden pixels and cannot
// be compiled here: Rust nightly, no toolchain); unit-level KATs are pinned in tests/.
//
// Restated reference functions (file:line relative to the reference checkout):
//   SceneDrawInfo::new / traverse_*          rayca-soft/src/scene.rs:190-282
//   BvhScene/BvhModel/BvhPrimitive builders  rayca-soft/src/bvh/primitive.rs:194-394
//   Blas build (SAH, 63 planes x 3 axes)     rayca-soft/src/bvh/blas.rs:27-127,245-316
//   Tlas build                               rayca-soft/src/bvh/tlas.rs:74-134,248-269
//   AABB slab test                           rayca-soft/src/bvh/aabb.rs:74-93
//   BvhNode/TlasNode::intersects             blas.rs:129-176, tlas.rs:136-180
//   Triangle::intersects                     rayca-geometry/src/triangle.rs:84-159
//   HitInfo / attribute interpolation        rayca-soft/src/hit.rs, bvh/triangle.rs:33-71
//   Materials / Sampler                      rayca-model/src/material/*.rs, sampler.rs:11-29
//   Lights                                   rayca-model/src/light/{point,directional}.rs
//   Integrators                              rayca-soft/src/integrator/{pathtracer,raytracer,flat}.rs
//   NEE / Cosine samplers                    rayca-soft/src/sampler/{nee,cosine}.rs
//   BRDFs                                    rayca-soft/src/brdf/{ggx,lambertian}.rs
//   Pixel loop                               rayca-soft/src/scene.rs:101-152
//
// Intentional, documented deviations (SURVEY A-19): BLAS order follows ascending model
// handle instead of HashMap iteration order; stochastic samplers draw from a keyed
// Philox4x32-10 counter RNG instead of thread-local unseeded fastrand.
#include "oracle_math.h"

#include "../include/rayca_cuda.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#ifdef _OPENMP
#include <omp.h>
#endif

using namespace orc;

namespace {

// ------------------------------------------------------------------ data model

struct VertexExt {
    Color color;
    Vec3 normal, tangent, bitangent;
    Vec2 uv;
};

// BvhPrimitive (bvh/primitive.rs:52-60) holding a BvhTriangle (bvh/triangle.rs:7-10)
struct Prim {
    Point3 v[3];   // MODEL space (primitive.rs:209-226 keeps positions untransformed)
    Vec3 centroid; // model space, (a+b+c)*0.3333 (rayca-geometry triangle.rs:59-62)
    VertexExt ext[3];
    uint32_t trs;      // index into world_trs (the NodeDrawInfo key)
    uint32_t model;    // owning model
    int32_t material;  // model-local material handle or -1
    uint32_t src;      // stable id: emit order within the model
    // BvhGeometry::Sphere (primitive.rs:10-13, 270-282): model-space sphere, the node TRS is applied per query
    bool sphere = false;
    Point3 center{0, 0, 0};
    float radius = 0, radius2 = 0;
    // cache of the values the reference recomputes on every query (bit-identical, same ops)
    Point3 wv[3];
    Vec3 wcentroid;
    Point3 wmin, wmax;
};

struct AABB {
    Point3 a, b;
};

struct BvhNode {
    AABB bounds;
    uint32_t offset = 0, count = 0;
    bool is_leaf() const { return count != 0; }
};

struct Blas {
    std::vector<BvhNode> nodes;
    std::vector<Prim> prims;
    uint32_t model = 0;
    uint8_t max_depth = 255;
    bool has_spheres = false; // lets triangle-only BLAS keep the branch-free leaf loop (it is the timed CPU baseline)
};

struct TlasNode {
    AABB bounds;
    int32_t left = -1, right = -1;
    uint32_t offset = 0, count = 0;
    bool is_leaf() const { return left < 0 && right < 0; }
};

struct BlasRef {
    uint32_t blas;
};

struct LightDraw {
    uint32_t model;
    uint32_t light; // global light index
    Trs node_trs;   // LOCAL node trs (A-13)
};

struct Hit {
    Ray ray;
    Point3 point;
    uint32_t blas = ~0u, primitive = ~0u;
    float depth = 0;
    Vec2 uv;
};

// "Optimised CPU" arm: what a tuned CPU renderer would run on the same scene —
// ONE binned-SAH tree over the pre-transformed world-space triangles of every BLAS, walked front to back with
// distance culling.  It is NOT the reference's algorithm (that is the exhaustive walk above); it exists so that the
// GPU number can be read against a fair CPU implementation, and as a fast full-frame checker.  It returns the
// reference's result by the same argument the CUDA product walk uses: a candidate counts only
// if the exact slab test of its reference LEAF accepts the ray, ties go to the lowest (blas, primitive) index,
// and the box test is conservative (1e-5 relative slack).
struct OptNode {
    float lo[3], hi[3];
    uint32_t first, count; // count != 0: leaf over opt.prims[first, first+count); else children first, first+1
};
struct OptPrim {
    Point3 v[3];
    uint32_t blas;  // index into blas_nodes (TLAS order) = Hit::blas
    uint32_t prim;  // Hit::primitive
    const AABB* leaf; // bounds of the reference leaf that holds the primitive
};
struct OptTree {
    std::vector<OptNode> nodes;
    std::vector<OptPrim> prims;
    double build_ms = 0;
    uint32_t depth = 0;
};

struct Counters {
    uint64_t rays_primary = 0, rays_shadow = 0, rays_secondary = 0;
    uint64_t ord_nodes = 0, ord_tris = 0; // ordered, distance-culled traversal (algorithmic bytes model)
    uint64_t exh_nodes = 0, exh_tris = 0; // what the reference's exhaustive traversal touches
};

} // namespace

struct rayca_oracle_scene {
    const rayca_scene_desc* desc = nullptr; // borrowed during create only
    // deep copies of what shading needs
    std::vector<rayca_model> models;
    std::vector<rayca_material> materials;
    std::vector<rayca_pbr_material> pbr;
    std::vector<rayca_phong_material> phong;
    std::vector<rayca_ggx_material> ggx;
    std::vector<rayca_texture> textures;
    std::vector<rayca_image> images;
    std::vector<std::vector<uint8_t>> image_data;
    std::vector<rayca_light> lights;
    std::vector<rayca_camera> cameras;

    std::vector<Trs> world_trs;
    std::vector<LightDraw> light_draws;
    bool has_camera = false;
    Trs camera_trs;
    float camera_yfov = 0;

    std::vector<Blas> blass;
    std::vector<BlasRef> blas_nodes; // reordered by the TLAS build
    std::vector<TlasNode> tlas_nodes;
    TlasNode root;
    double build_ms = 0;
    std::string error;
    OptTree* opt = nullptr; // built on the first RAYCA_ORACLE_OPTIMISED render
    ~rayca_oracle_scene() { delete opt; }
};

namespace {

using Scene = rayca_oracle_scene;

Trs to_trs(const rayca_trs& t) {
    Trs r;
    r.translation = {t.translation[0], t.translation[1], t.translation[2]};
    r.rotation = {t.rotation[0], t.rotation[1], t.rotation[2], t.rotation[3]};
    r.scale = {t.scale[0], t.scale[1], t.scale[2]};
    return r;
}
Color to_color(const float c[4]) { return {c[0], c[1], c[2], c[3]}; }

// -------------------------------------------------------------- scene flattening
// rayca-soft/src/scene.rs:190-282

struct ModelDrawInfo {
    std::vector<uint32_t> mesh_nodes;      // world_trs indices, traversal order
    std::vector<uint32_t> mesh_node_local; // model-local node index
    std::vector<uint32_t> quad_nodes;      // world_trs indices of quad-light nodes (scene.rs:265-272)
    std::vector<uint32_t> quad_lights;     // their global light indices
    bool used = false;
};

struct Flattener {
    Scene& s;
    const rayca_scene_desc& d;
    std::vector<ModelDrawInfo> mdi;

    // scene.rs:234-282
    void traverse_model_nodes(const Trs& transform, uint32_t node, uint32_t model_index) {
        const rayca_model& m = d.models[model_index];
        const rayca_node& n = d.nodes[m.node_offset + node];
        Trs current = transform * to_trs(n.trs);
        uint32_t key = ~0u;
        if (n.mesh >= 0 || n.camera >= 0 || n.light >= 0) {
            key = (uint32_t)s.world_trs.size();
            s.world_trs.push_back(current);
        }
        if (n.mesh >= 0) {
            mdi[model_index].mesh_nodes.push_back(key);
            mdi[model_index].mesh_node_local.push_back(node);
            mdi[model_index].used = true;
        }
        if (n.light >= 0) {
            LightDraw ld;
            ld.model = model_index;
            ld.light = m.light_offset + (uint32_t)n.light;
            ld.node_trs = to_trs(n.trs);
            s.light_draws.push_back(ld);
            if (d.lights[ld.light].kind == RAYCA_LIGHT_QUAD) { // the quad becomes two emissive triangles of this model
                mdi[model_index].quad_nodes.push_back(key);
                mdi[model_index].quad_lights.push_back(ld.light);
                mdi[model_index].used = true;
            }
        }
        if (n.camera >= 0 && !s.has_camera) { // camera_draw_infos[0] (scene.rs:110)
            s.has_camera = true;
            s.camera_trs = current;
            s.camera_yfov = d.cameras[m.camera_offset + (uint32_t)n.camera].yfov_radians;
        }
        for (uint32_t i = 0; i < n.child_count; ++i)
            traverse_model_nodes(current, d.children[n.first_child + i], model_index);
    }
    // scene.rs:227-232
    void traverse_model(const Trs& transform, uint32_t model_index) {
        const rayca_model& m = d.models[model_index];
        Trs current = transform * to_trs(m.root_trs);
        for (uint32_t i = 0; i < m.root_child_count; ++i)
            traverse_model_nodes(current, d.children[m.root_first_child + i], model_index);
    }
    // scene.rs:213-225
    void traverse_scene_nodes(const Trs& transform, uint32_t node) {
        const rayca_node& n = d.scene_nodes[node];
        Trs current = transform * to_trs(n.trs);
        if (n.model >= 0) traverse_model(current, (uint32_t)n.model);
        for (uint32_t i = 0; i < n.child_count; ++i) traverse_scene_nodes(current, d.children[n.first_child + i]);
    }
    // scene.rs:206-211
    void run() {
        mdi.resize(d.model_count);
        Trs root = to_trs(d.root_trs);
        for (uint32_t i = 0; i < d.root_child_count; ++i) traverse_scene_nodes(root, d.children[d.root_first_child + i]);
    }
};

// Sphere::{get_center,get_radius,min,max,get_centroid} (rayca-geometry/src/sphere.rs:72-93,165-181)
void sphere_world_bounds(Prim& p, const Trs& trs) {
    Point3 c = transform_point(trs, p.center);
    float r = p.radius * maxf(maxf(maxf(trs.scale.x, trs.scale.y), trs.scale.z), 0.0f); // reduce_max over 4 lanes, w = 0
    p.wcentroid = to_vec3(c);
    p.wmin = c - Vec3{r, r, r};
    p.wmax = c + Vec3{r, r, r};
}

// Triangle::{get_vertex,get_centroid,min,max} (rayca-geometry/src/triangle.rs:59-82,161-181): the cache of what the
// reference recomputes per query
void triangle_world_bounds(Prim& p, const Trs& trs) {
    for (int k = 0; k < 3; ++k) p.wv[k] = transform_point(trs, p.v[k]);
    p.wcentroid = transform_vec(trs, p.centroid);
    Point3 mn{F32_MAX, F32_MAX, F32_MAX}, mx{F32_MIN, F32_MIN, F32_MIN};
    for (int k = 0; k < 3; ++k) {
        mn = {minf(mn.x, p.wv[k].x), minf(mn.y, p.wv[k].y), minf(mn.z, p.wv[k].z)};
        mx = {maxf(mx.x, p.wv[k].x), maxf(mx.y, p.wv[k].y), maxf(mx.z, p.wv[k].z)};
    }
    p.wmin = mn;
    p.wmax = mx;
}


// bvh/primitive.rs:194-237 for one mesh node
void emit_mesh_prims(Scene& s, const rayca_scene_desc& d, uint32_t model_index, uint32_t node_local, uint32_t trs_key,
                     std::vector<Prim>& out) {
    const rayca_model& m = d.models[model_index];
    const rayca_node& n = d.nodes[m.node_offset + node_local];
    const rayca_mesh& mesh = d.meshes[m.mesh_offset + (uint32_t)n.mesh];
    const Trs& trs = s.world_trs[trs_key];
    Mat3 tangent_matrix = mat3_from_trs(trs);
    Mat3 normal_matrix = transpose(mat3_from_inversed_trs(trs));
    for (uint32_t pi = 0; pi < mesh.primitive_count; ++pi) {
        const rayca_primitive& prim = d.primitives[m.primitive_offset + d.mesh_primitives[mesh.first_primitive + pi]];
        const rayca_geometry& g = d.geometries[m.geometry_offset + prim.geometry];
        if (g.kind == RAYCA_GEOMETRY_SPHERE) { // BvhPrimitive::from_sphere (primitive.rs:270-282)
            Prim p{};
            p.sphere = true;
            p.center = {g.center[0], g.center[1], g.center[2]};
            p.radius = g.radius;
            p.radius2 = g.radius * g.radius; // Sphere::new (sphere.rs:51-62)
            p.trs = trs_key;
            p.model = model_index;
            p.material = prim.material;
            p.src = (uint32_t)out.size();
            sphere_world_bounds(p, trs);
            out.push_back(p);
            continue;
        }
        if (g.kind != RAYCA_GEOMETRY_TRIANGLE_MESH) continue;
        for (uint32_t t = 0; t < g.index_count / 3; ++t) {
            Prim p;
            for (int k = 0; k < 3; ++k) {
                const rayca_vertex& v = d.vertices[g.first_vertex + d.indices[g.first_index + t * 3 + k]];
                p.v[k] = {v.pos[0], v.pos[1], v.pos[2]};
                p.ext[k].color = to_color(v.color);
                p.ext[k].normal = normal_matrix * Vec3{v.normal[0], v.normal[1], v.normal[2]};
                p.ext[k].tangent = tangent_matrix * Vec3{v.tangent[0], v.tangent[1], v.tangent[2]};
                p.ext[k].bitangent = tangent_matrix * Vec3{v.bitangent[0], v.bitangent[1], v.bitangent[2]};
                p.ext[k].uv = {v.uv[0], v.uv[1]};
            }
            // Triangle::new (rayca-geometry triangle.rs:59-62)
            p.centroid = ((to_vec3(p.v[0]) + to_vec3(p.v[1])) + to_vec3(p.v[2])) * 0.3333f;
            p.trs = trs_key;
            p.model = model_index;
            p.material = prim.material;
            p.src = (uint32_t)out.size();
            triangle_world_bounds(p, trs);
            out.push_back(p);
        }
    }
}

// BvhPrimitive::from_quad_light (bvh/primitive.rs:310-346): triangles [a, d, b] and [a, c, d] with a = origin,
// b = ab, d = ab + ac, c = ac in the light node's space; vertex normal = the quad normal, untransformed;
// colour WHITE, uv / tangents zero (VertexExtBuilder defaults); material = the light's emissive material
void emit_quad_prims(Scene& s, const rayca_scene_desc& d, uint32_t model_index, uint32_t light_index, uint32_t trs_key,
                     std::vector<Prim>& out) {
    const rayca_light& l = d.lights[light_index];
    const Trs& trs = s.world_trs[trs_key];
    Vec3 ab{l.ab[0], l.ab[1], l.ab[2]}, ac{l.ac[0], l.ac[1], l.ac[2]};
    Vec3 normal = normalized(cross(ab, ac)); // QuadLight::get_normal (quad.rs:36-38)
    Point3 a{0, 0, 0};
    Point3 b = a + ab, dd = (a + ab) + ac, c = a + ac;
    Point3 tri[2][3] = {{a, dd, b}, {a, c, dd}};
    for (int t = 0; t < 2; ++t) {
        Prim p{};
        for (int k = 0; k < 3; ++k) {
            p.v[k] = tri[t][k];
            p.ext[k].color = WHITE;
            p.ext[k].normal = normal;
            p.ext[k].tangent = {0, 0, 0};
            p.ext[k].bitangent = {0, 0, 0};
            p.ext[k].uv = {0, 0};
        }
        p.centroid = ((to_vec3(p.v[0]) + to_vec3(p.v[1])) + to_vec3(p.v[2])) * 0.3333f;
        p.trs = trs_key;
        p.model = model_index;
        p.material = l.material;
        p.src = (uint32_t)out.size();
        for (int k = 0; k < 3; ++k) p.wv[k] = transform_point(trs, p.v[k]);
        p.wcentroid = transform_vec(trs, p.centroid);
        Point3 mn{F32_MAX, F32_MAX, F32_MAX}, mx{F32_MIN, F32_MIN, F32_MIN};
        for (int k = 0; k < 3; ++k) {
            mn = {minf(mn.x, p.wv[k].x), minf(mn.y, p.wv[k].y), minf(mn.z, p.wv[k].z)};
            mx = {maxf(mx.x, p.wv[k].x), maxf(mx.y, p.wv[k].y), maxf(mx.z, p.wv[k].z)};
        }
        p.wmin = mn;
        p.wmax = mx;
        out.push_back(p);
    }
}

// ------------------------------------------------------------------- BLAS build

inline void grow(AABB& box, Point3 p) {
    box.a = {minf(box.a.x, p.x), minf(box.a.y, p.y), minf(box.a.z, p.z)};
    box.b = {maxf(box.b.x, p.x), maxf(box.b.y, p.y), maxf(box.b.z, p.z)};
}
// aabb.rs:20-23
inline float area(const AABB& box) {
    Vec3 e = box.b - box.a;
    return e.x * e.y + e.y * e.z + e.z * e.x;
}
// BvhNode::new, blas.rs:27-36 with aabb.rs:48-59
BvhNode make_node(const Blas& blas, uint32_t offset, uint32_t count) {
    BvhNode n;
    n.bounds.a = {F32_MAX, F32_MAX, F32_MAX};
    n.bounds.b = {F32_MIN, F32_MIN, F32_MIN};
    for (uint32_t i = offset; i < offset + count; ++i) {
        const Prim& p = blas.prims[i];
        n.bounds.a = {minf(n.bounds.a.x, p.wmin.x), minf(n.bounds.a.y, p.wmin.y), minf(n.bounds.a.z, p.wmin.z)};
        n.bounds.b = {maxf(n.bounds.b.x, p.wmax.x), maxf(n.bounds.b.y, p.wmax.y), maxf(n.bounds.b.z, p.wmax.z)};
    }
    n.offset = offset;
    n.count = count;
    return n;
}
// AABB::grow_primitive (aabb.rs:61-72): triangle = its three world vertices, sphere = its min / max points
inline void grow_prim(AABB& box, const Prim& p) {
    if (p.sphere) {
        grow(box, p.wmin);
        grow(box, p.wmax);
        return;
    }
    grow(box, p.wv[0]);
    grow(box, p.wv[1]);
    grow(box, p.wv[2]);
}
// blas.rs:64-89 — candidate boxes start at AABB::default() = two points at the ORIGIN (A-21)
float evaluate_sah(const Blas& blas, const BvhNode& node, int axis, float pos) {
    AABB left_box{{0, 0, 0}, {0, 0, 0}}, right_box{{0, 0, 0}, {0, 0, 0}};
    int left_count = 0, right_count = 0;
    for (uint32_t i = node.offset; i < node.offset + node.count; ++i) {
        const Prim& p = blas.prims[i];
        if (p.wcentroid[axis] < pos) {
            left_count++;
            grow_prim(left_box, p);
        } else {
            right_count++;
            grow_prim(right_box, p);
        }
    }
    float cost = (float)left_count * area(left_box) + (float)right_count * area(right_box);
    return cost > 0.0f ? cost : F32_MAX;
}
// blas.rs:93-123
void find_best_split_plane(const Blas& blas, const BvhNode& node, int& best_axis, float& split_pos, float& best_cost) {
    best_cost = F32_MAX;
    best_axis = 0;
    split_pos = 0.0f;
    struct Cand {
        int axis;
        float pos;
        float cost;
    };
    std::vector<Cand> cands;
    for (int axis = 0; axis < 3; ++axis) {
        float bmin = node.bounds.a[axis], bmax = node.bounds.b[axis];
        if (bmin == bmax) continue;
        const int AREA_COUNT = 64;
        float scale = (bmax - bmin) / (float)AREA_COUNT;
        for (int i = 1; i < AREA_COUNT; ++i) cands.push_back({axis, bmin + (float)i * scale, 0.0f});
    }
    // evaluation order does not affect the values; the selection below is sequential
#pragma omp parallel for schedule(dynamic, 4) if (node.count > 20000)
    for (size_t c = 0; c < cands.size(); ++c) cands[c].cost = evaluate_sah(blas, node, cands[c].axis, cands[c].pos);
    for (const Cand& c : cands) {
        if (c.cost < best_cost) {
            best_cost = c.cost;
            best_axis = c.axis;
            split_pos = c.pos;
        }
    }
}
// blas.rs:261-316
void set_primitives_recursive(Blas& blas, size_t node_index, unsigned level) {
    if (level >= blas.max_depth) return;
    BvhNode node = blas.nodes[node_index];
    int axis;
    float pos, cost;
    find_best_split_plane(blas, node, axis, pos, cost);
    float no_split_cost = (float)node.count * area(node.bounds);
    if (cost > no_split_cost) return;
    size_t i = node.offset, j = node.offset + node.count;
    while (i < j) {
        if (blas.prims[i].wcentroid[axis] < pos) {
            i++;
        } else {
            std::swap(blas.prims[i], blas.prims[j - 1]);
            j--;
        }
    }
    uint32_t left_count = (uint32_t)i - node.offset;
    uint32_t right_count = node.count - left_count;
    if (left_count > 0 && right_count > 0) {
        uint32_t left_index = (uint32_t)blas.nodes.size();
        BvhNode l = make_node(blas, node.offset, left_count);
        BvhNode r = make_node(blas, node.offset + left_count, right_count);
        node.offset = left_index; // set_left_child_index
        node.count = 0;
        blas.nodes.push_back(l);
        blas.nodes.push_back(r);
        set_primitives_recursive(blas, left_index, level + 1);
        set_primitives_recursive(blas, left_index + 1, level + 1);
    }
    blas.nodes[node_index] = node;
}
// blas.rs:245-259
void build_blas(Blas& blas) {
    blas.nodes.clear();
    blas.nodes.push_back(make_node(blas, 0, (uint32_t)blas.prims.size()));
    blas.nodes.push_back(BvhNode{}); // node 1 unused
    set_primitives_recursive(blas, 0, 0);
}

// ------------------------------------------------------------------- TLAS build
// tlas.rs:74-134
void replace_models_recursive(Scene& s, TlasNode& self, uint32_t offset, uint32_t count) {
    self.offset = offset;
    self.count = count;
    self.bounds.a = {F32_MAX, F32_MAX, F32_MAX};
    self.bounds.b = {F32_MIN, F32_MIN, F32_MIN};
    for (uint32_t i = offset; i < offset + count; ++i) {
        const AABB& rb = s.blass[s.blas_nodes[i].blas].nodes[0].bounds;
        self.bounds.a = {minf(self.bounds.a.x, rb.a.x), minf(self.bounds.a.y, rb.a.y), minf(self.bounds.a.z, rb.a.z)};
        self.bounds.b = {maxf(self.bounds.b.x, rb.b.x), maxf(self.bounds.b.y, rb.b.y), maxf(self.bounds.b.z, rb.b.z)};
    }
    Vec3 extent = self.bounds.b - self.bounds.a;
    int axis = 0;
    if (extent.y > extent.x) axis = 1;
    if (extent.z > extent[axis]) axis = 2;
    float split_pos = self.bounds.a[axis] + extent[axis] * 0.5f;
    size_t i = offset, j = offset + count;
    while (i < j) {
        const AABB& rb = s.blass[s.blas_nodes[i].blas].nodes[0].bounds;
        Vec3 centroid = (rb.b - rb.a) / 2.0f; // aabb.rs:95-97 (extent/2 quirk)
        if (centroid[axis] < split_pos) {
            i++;
        } else {
            std::swap(s.blas_nodes[i], s.blas_nodes[j - 1]);
            j--;
        }
    }
    uint32_t left_count = (uint32_t)i - offset;
    uint32_t right_count = count - left_count;
    if (left_count > 0 && right_count > 0) {
        TlasNode l, r;
        replace_models_recursive(s, l, offset, left_count);
        replace_models_recursive(s, r, offset + left_count, right_count);
        s.tlas_nodes.push_back(l);
        self.left = (int32_t)s.tlas_nodes.size() - 1;
        s.tlas_nodes.push_back(r);
        self.right = (int32_t)s.tlas_nodes.size() - 1;
        self.count = 0;
    }
}

// ---------------------------------------------------------------- intersection

// aabb.rs:74-93
inline float aabb_intersects(const AABB& box, const Ray& ray) {
    Vec3 o = to_vec3(ray.origin);
    Point3 t1 = scale_point(box.a - o, ray.rdir);
    Point3 t2 = scale_point(box.b - o, ray.rdir);
    // w lanes: max(1,1)*f32::MAX and min(1,1)*f32::MIN take part in the reductions
    float tmax = minf(minf(minf(maxf(t1.x, t2.x) * 1.0f, maxf(t1.y, t2.y) * 1.0f), maxf(t1.z, t2.z) * 1.0f), 1.0f * F32_MAX);
    float tmin = maxf(maxf(maxf(minf(t1.x, t2.x) * 1.0f, minf(t1.y, t2.y) * 1.0f), minf(t1.z, t2.z) * 1.0f), 1.0f * F32_MIN);
    if (tmax >= tmin && tmax > 0.0f) return tmin;
    return F32_MAX;
}

// rayca-geometry/src/triangle.rs:84-159 on already-transformed vertices
inline bool triangle_intersects_world(Point3 p0, Point3 p1, Point3 p2, const Ray& ray, float& t_out, Point3& p_out, Vec2& uv_out) {
    Vec3 v0 = to_vec3(p0), v1 = to_vec3(p1), v2 = to_vec3(p2);
    Vec3 v0v1 = v1 - v0;
    Vec3 v0v2 = v2 - v0;
    Vec3 n = cross(v0v1, v0v2);
    if (dot(ray.dir, n) > 0.0f) return false;
    float denom = dot(n, n);
    float n_dot_ray_dir = dot(n, ray.dir);
    if (std::fabs(n_dot_ray_dir) < F32_EPSILON) return false;
    float d = -dot(n, v0);
    float t = -(dot(n, to_vec3(ray.origin)) + d) / n_dot_ray_dir;
    if (t < 0.0f) return false;
    Point3 p = ray.origin + ray.dir * t;
    Vec3 edge0 = v1 - v0;
    Vec3 vp0 = to_vec3(p - v0);
    Vec3 c = cross(edge0, vp0);
    if (dot(n, c) < 0.0f) return false;
    Vec3 edge1 = v2 - v1;
    Vec3 vp1 = to_vec3(p - v1);
    c = cross(edge1, vp1);
    float u = dot(n, c);
    if (u < 0.0f) return false;
    Vec3 edge2 = v0 - v2;
    Vec3 vp2 = to_vec3(p - v2);
    c = cross(edge2, vp2);
    float v = dot(n, c);
    if (v < 0.0f) return false;
    t_out = t;
    p_out = p;
    uv_out = {u / denom, v / denom};
    return true;
}

struct Tracer {
    const Scene& s;
    bool faithful; // re-transform the three vertices on every test, as the reference does
    Counters* rays; // always valid: ray tallies
    Counters* cnt;  // non-null only when walk records are being counted
    const OptTree* opt = nullptr; // non-null: the optimised arm answers the query

    // Sphere::intersects + intersects_impl (rayca-geometry/src/sphere.rs:101-163)
    __attribute__((noinline)) bool sphere_intersects(const Prim& p, const Ray& ray, float& t_out, Point3& pt, Vec2& uv) const {
        const Trs& trs = s.world_trs[p.trs];
        Ray r = inverse_transform_ray(trs, ray);
        float a = dot(r.dir, r.dir);
        Vec3 c_to_r = r.origin - p.center;
        float b = dot(c_to_r, r.dir);
        float c = dot(c_to_r, c_to_r) - p.radius2;
        float det = b * b - a * c;
        if (det < 0.0f) return false;
        float det_sqrt = std::sqrt(det);
        float t0 = (-b + det_sqrt) / a;
        float t1 = (-b - det_sqrt) / a;
        if (t0 < 0.0f && t1 < 0.0f) return false;
        float t;
        if (t0 >= 0.0f && t1 >= 0.0f) t = minf(t0, t1);
        else if (t0 >= 0.0f) t = t0;
        else t = t1;
        Point3 point = r.origin + r.dir * t;
        t_out = t;
        pt = transform_point(trs, point);
        uv = {0.0f, 0.0f};
        return true;
    }

    template <bool MAY_BE_SPHERE>
    inline bool prim_intersects(const Prim& p, const Ray& ray, float& t, Point3& pt, Vec2& uv) const {
        if (MAY_BE_SPHERE && p.sphere) { // own temporaries: the triangle path keeps its results in registers
            float st;
            Point3 spt;
            Vec2 suv;
            if (!sphere_intersects(p, ray, st, spt, suv)) return false;
            t = st;
            pt = spt;
            uv = suv;
            return true;
        }
        if (faithful) {
            const Trs& trs = s.world_trs[p.trs]; // get_world_trs (primitive.rs:96)
            return triangle_intersects_world(transform_point(trs, p.v[0]), transform_point(trs, p.v[1]),
                                             transform_point(trs, p.v[2]), ray, t, pt, uv);
        }
        return triangle_intersects_world(p.wv[0], p.wv[1], p.wv[2], ray, t, pt, uv);
    }

    // blas.rs:129-176 (recursive, both children always visited, strict '<')
    // SPH: the BLAS holds spheres; triangle-only BLAS get an instantiation without any sphere code, which is
    // what the timed CPU baseline runs (with the branch compiled in, gcc produced a 1.8x slower triangle loop)
    template <bool SPH>
    bool node_intersects(const Blas& blas, const BvhNode& node, const Ray& ray, Hit& out) const {
        if (cnt) cnt->exh_nodes++;
        float d = aabb_intersects(node.bounds, ray);
        if (d == F32_MAX) return false;
        bool found = false;
        float depth = INFINITY;
        if (node.is_leaf()) {
            if (cnt) cnt->exh_tris += node.count;
            for (uint32_t i = node.offset; i < node.offset + node.count; ++i) {
                float t;
                Point3 pt;
                Vec2 uv;
                if (prim_intersects<SPH>(blas.prims[i], ray, t, pt, uv)) {
                    if (t < depth) {
                        depth = t;
                        out.depth = t;
                        out.point = pt;
                        out.uv = uv;
                        out.primitive = i;
                        found = true;
                    }
                }
            }
        } else {
            Hit h;
            if (node_intersects<SPH>(blas, blas.nodes[node.offset], ray, h)) {
                if (h.depth < depth) {
                    depth = h.depth;
                    out = h;
                    found = true;
                }
            }
            Hit h2;
            if (node_intersects<SPH>(blas, blas.nodes[node.offset + 1], ray, h2)) {
                if (h2.depth < depth) {
                    out = h2;
                    found = true;
                }
            }
        }
        return found;
    }

    // tlas.rs:136-180
    bool tlas_node_intersects(const TlasNode& node, const Ray& ray, Hit& out) const {
        if (aabb_intersects(node.bounds, ray) == F32_MAX) return false;
        bool found = false;
        float depth = INFINITY;
        if (node.is_leaf()) {
            for (uint32_t i = node.offset; i < node.offset + node.count; ++i) {
                const Blas& blas = s.blass[s.blas_nodes[i].blas];
                Hit h;
                if (blas.has_spheres ? node_intersects<true>(blas, blas.nodes[0], ray, h)
                                     : node_intersects<false>(blas, blas.nodes[0], ray, h)) {
                    if (h.depth < depth) {
                        depth = h.depth;
                        h.blas = i;
                        out = h;
                        found = true;
                    }
                }
            }
        } else {
            if (node.left >= 0) {
                Hit h;
                if (tlas_node_intersects(s.tlas_nodes[node.left], ray, h)) {
                    if (h.depth < depth) {
                        depth = h.depth;
                        out = h;
                        found = true;
                    }
                }
            }
            if (node.right >= 0) {
                Hit h;
                if (tlas_node_intersects(s.tlas_nodes[node.right], ray, h)) {
                    if (h.depth < depth) {
                        out = h;
                        found = true;
                    }
                }
            }
        }
        return found;
    }

    // Ordered, distance-culled traversal of ONE BLAS: the algorithm the CUDA kernel runs.
    // Used only to COUNT the node / triangle records such a traversal must read
    // (SURVEY 8(d) algorithmic bytes); the colour path always uses the exhaustive one.
    void count_ordered_blas(const Blas& blas, const Ray& ray, float& best_t, uint64_t& best_key, uint32_t blas_index) const {
        cnt->ord_nodes++;
        if (aabb_intersects(blas.nodes[0].bounds, ray) == F32_MAX) return;
        uint32_t stack[256];
        int sp = 0;
        uint32_t cur = 0;
        for (;;) {
            const BvhNode& node = blas.nodes[cur];
            if (node.is_leaf()) {
                cnt->ord_tris += node.count;
                for (uint32_t i = node.offset; i < node.offset + node.count; ++i) {
                    float t;
                    Point3 pt;
                    Vec2 uv;
                    if (blas.has_spheres ? prim_intersects<true>(blas.prims[i], ray, t, pt, uv)
                                     : prim_intersects<false>(blas.prims[i], ray, t, pt, uv)) {
                        uint64_t key = ((uint64_t)blas_index << 32) | i;
                        if (t < best_t || (t == best_t && key < best_key)) {
                            best_t = t;
                            best_key = key;
                        }
                    }
                }
            } else {
                cnt->ord_nodes += 2;
                uint32_t l = node.offset, r = node.offset + 1;
                float dl = aabb_intersects(blas.nodes[l].bounds, ray);
                float dr = aabb_intersects(blas.nodes[r].bounds, ray);
                bool hl = dl != F32_MAX && !(dl > best_t);
                bool hr = dr != F32_MAX && !(dr > best_t);
                if (hl && hr) {
                    if (dr < dl) std::swap(l, r);
                    stack[sp++] = r;
                    cur = l;
                    continue;
                } else if (hl) {
                    cur = l;
                    continue;
                } else if (hr) {
                    cur = r;
                    continue;
                }
            }
            if (sp == 0) break;
            cur = stack[--sp];
        }
    }

    // conservative slab on an optimised-tree box: [entry, exit] widened by 1e-5 relative, like the CUDA walk
    static inline bool opt_slab(const OptNode& n, const Ray& ray, float& entry) {
        const float ox = ray.origin.x, oy = ray.origin.y, oz = ray.origin.z;
        const float t1x = (n.lo[0] - ox) * ray.rdir.x, t2x = (n.hi[0] - ox) * ray.rdir.x;
        const float t1y = (n.lo[1] - oy) * ray.rdir.y, t2y = (n.hi[1] - oy) * ray.rdir.y;
        const float t1z = (n.lo[2] - oz) * ray.rdir.z, t2z = (n.hi[2] - oz) * ray.rdir.z;
        const float tmax = minf(minf(maxf(t1x, t2x), maxf(t1y, t2y)), maxf(t1z, t2z));
        const float tmin = maxf(maxf(minf(t1x, t2x), minf(t1y, t2y)), minf(t1z, t2z));
        entry = tmin - std::fabs(tmin) * 1.0e-5f;
        const float ext = tmax + std::fabs(tmax) * 1.0e-5f;
        return ext >= entry && ext >= 0.0f;
    }

    // Ordered, distance-culled closest hit over the optimised tree (see OptTree).
    bool opt_intersects(const Ray& ray, Hit& out) const {
        // a zero direction lane has rdir = 0 there: every reference box rejects the ray (SURVEY A-3)
        if (ray.rdir.x == 0.0f || ray.rdir.y == 0.0f || ray.rdir.z == 0.0f) return false;
        const OptTree& t = *opt;
        float best_t = INFINITY;
        uint64_t best_key = ~0ull;
        const AABB* ok_leaf = nullptr;
        struct Entry { uint32_t node; float d; };
        Entry stack[128];
        int sp = 0;
        float e0;
        if (t.nodes.empty() || !opt_slab(t.nodes[0], ray, e0)) return false;
        uint32_t cur = 0;
        for (;;) {
            const OptNode& n = t.nodes[cur];
            if (n.count) {
                for (uint32_t i = n.first; i < n.first + n.count; ++i) {
                    const OptPrim& p = t.prims[i];
                    float tt;
                    Point3 pt;
                    Vec2 uv;
                    if (!triangle_intersects_world(p.v[0], p.v[1], p.v[2], ray, tt, pt, uv)) continue;
                    const uint64_t key = ((uint64_t)p.blas << 32) | p.prim;
                    if (!(tt < best_t || (tt == best_t && key < best_key))) continue;
                    if (p.leaf != ok_leaf) { // the reference tests the triangle iff its leaf box accepts the ray
                        if (aabb_intersects(*p.leaf, ray) == F32_MAX) continue;
                        ok_leaf = p.leaf;
                    }
                    best_t = tt;
                    best_key = key;
                    out.depth = tt;
                    out.point = pt;
                    out.uv = uv;
                    out.blas = p.blas;
                    out.primitive = p.prim;
                }
            } else {
                float dl, dr;
                uint32_t l = n.first, r = n.first + 1;
                bool hl = opt_slab(t.nodes[l], ray, dl) && !(dl > best_t);
                bool hr = opt_slab(t.nodes[r], ray, dr) && !(dr > best_t);
                if (hl && hr) {
                    if (dr < dl) { std::swap(l, r); std::swap(dl, dr); }
                    if (sp < 128) stack[sp++] = {r, dr};
                    cur = l;
                    continue;
                } else if (hl) {
                    cur = l;
                    continue;
                } else if (hr) {
                    cur = r;
                    continue;
                }
            }
            bool popped = false;
            while (sp > 0) {
                --sp;
                if (stack[sp].d > best_t) continue;
                cur = stack[sp].node;
                popped = true;
                break;
            }
            if (!popped) break;
        }
        if (best_key == ~0ull) return false;
        out.ray = ray;
        return true;
    }

    // Tlas::intersects (tlas.rs:271-275)
    bool intersects(const Ray& ray, Hit& out) const {
        if (opt) return opt_intersects(ray, out);
        bool found = tlas_node_intersects(s.root, ray, out);
        if (cnt) {
            float best_t = INFINITY;
            uint64_t best_key = ~0ull;
            for (uint32_t i = 0; i < s.blas_nodes.size(); ++i)
                count_ordered_blas(s.blass[s.blas_nodes[i].blas], ray, best_t, best_key, i);
        }
        if (found) out.ray = ray;
        return found;
    }
};


// ------------------------------------------------------------- optimised-arm tree (see OptTree)
// Binned SAH (16 bins on the centroid bounds, leaves of at most 4 triangles); a split that does not separate
// anything, or a node deeper than 48 levels, falls back to the positional median, which bounds the depth.
struct OptBuilder {
    OptTree& t;
    std::vector<Point3> cen;
    std::vector<uint32_t> order; // permutation of t.prims being partitioned
    std::vector<OptPrim> src;

    static void grow3(float lo[3], float hi[3], const Point3& p) {
        lo[0] = minf(lo[0], p.x); lo[1] = minf(lo[1], p.y); lo[2] = minf(lo[2], p.z);
        hi[0] = maxf(hi[0], p.x); hi[1] = maxf(hi[1], p.y); hi[2] = maxf(hi[2], p.z);
    }
    void bounds(uint32_t first, uint32_t count, float lo[3], float hi[3]) const {
        for (int k = 0; k < 3; ++k) { lo[k] = INFINITY; hi[k] = -INFINITY; }
        for (uint32_t i = first; i < first + count; ++i)
            for (int v = 0; v < 3; ++v) grow3(lo, hi, src[order[i]].v[v]);
    }
    static float half_area(const float lo[3], const float hi[3]) {
        const float ex = hi[0] - lo[0], ey = hi[1] - lo[1], ez = hi[2] - lo[2];
        return ex * ey + ey * ez + ez * ex;
    }
    void build(uint32_t node, uint32_t first, uint32_t count, uint32_t depth) {
        OptNode& n0 = t.nodes[node];
        bounds(first, count, n0.lo, n0.hi);
        t.depth = std::max(t.depth, depth);
        if (count <= 4) {
            n0.first = first;
            n0.count = count;
            return;
        }
        float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (uint32_t i = first; i < first + count; ++i) grow3(clo, chi, cen[order[i]]);
        constexpr int B = 16;
        int best_axis = -1, best_plane = 0;
        float best_cost = INFINITY;
        if (depth < 48) {
            for (int a = 0; a < 3; ++a) {
                const float ext = chi[a] - clo[a];
                if (!(ext > 0.0f)) continue;
                uint32_t cnt[B] = {0};
                float blo[B][3], bhi[B][3];
                for (int b = 0; b < B; ++b)
                    for (int k = 0; k < 3; ++k) { blo[b][k] = INFINITY; bhi[b][k] = -INFINITY; }
                const float scale = (float)B / ext;
                for (uint32_t i = first; i < first + count; ++i) {
                    const Point3& c = cen[order[i]];
                    const float cv = a == 0 ? c.x : (a == 1 ? c.y : c.z);
                    int b = std::min(B - 1, std::max(0, (int)((cv - clo[a]) * scale)));
                    cnt[b]++;
                    for (int v = 0; v < 3; ++v) grow3(blo[b], bhi[b], src[order[i]].v[v]);
                }
                float ra[B];
                uint32_t rc[B];
                float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
                uint32_t c = 0;
                for (int b = B - 1; b > 0; --b) {
                    for (int k = 0; k < 3; ++k) { lo[k] = minf(lo[k], blo[b][k]); hi[k] = maxf(hi[k], bhi[b][k]); }
                    c += cnt[b];
                    ra[b] = c ? half_area(lo, hi) : 0.0f;
                    rc[b] = c;
                }
                for (int k = 0; k < 3; ++k) { lo[k] = INFINITY; hi[k] = -INFINITY; }
                c = 0;
                for (int b = 0; b < B - 1; ++b) {
                    for (int k = 0; k < 3; ++k) { lo[k] = minf(lo[k], blo[b][k]); hi[k] = maxf(hi[k], bhi[b][k]); }
                    c += cnt[b];
                    if (c == 0 || rc[b + 1] == 0) continue;
                    const float cost = half_area(lo, hi) * (float)c + ra[b + 1] * (float)rc[b + 1];
                    if (cost < best_cost) { best_cost = cost; best_axis = a; best_plane = b; }
                }
            }
        }
        uint32_t mid = first + count / 2;
        if (best_axis >= 0) {
            const int a = best_axis;
            const float scale = (float)B / (chi[a] - clo[a]);
            auto left = [&](uint32_t id) {
                const Point3& c = cen[id];
                const float cv = a == 0 ? c.x : (a == 1 ? c.y : c.z);
                return std::min(B - 1, std::max(0, (int)((cv - clo[a]) * scale))) <= best_plane;
            };
            mid = (uint32_t)(std::partition(order.begin() + first, order.begin() + first + count, left) - order.begin());
            if (mid == first || mid == first + count) mid = first + count / 2;
        }
        const uint32_t child = (uint32_t)t.nodes.size();
        t.nodes.emplace_back();
        t.nodes.emplace_back();
        t.nodes[node].first = child;
        t.nodes[node].count = 0;
        build(child, first, mid - first, depth + 1);
        build(child + 1, mid, first + count - mid, depth + 1);
    }
};

// returns false when the scene holds spheres (the optimised arm covers triangle scenes)
bool build_opt_tree(Scene& s) {
    if (s.opt) return true;
    auto t0 = std::chrono::steady_clock::now();
    OptTree* t = new OptTree();
    OptBuilder b{*t, {}, {}, {}};
    for (uint32_t i = 0; i < s.blas_nodes.size(); ++i) {
        const Blas& blas = s.blass[s.blas_nodes[i].blas];
        if (blas.has_spheres) {
            delete t;
            return false;
        }
        for (const BvhNode& n : blas.nodes) {
            if (!n.is_leaf()) continue;
            for (uint32_t k = n.offset; k < n.offset + n.count; ++k) {
                const Prim& p = blas.prims[k];
                b.src.push_back(OptPrim{{p.wv[0], p.wv[1], p.wv[2]}, i, k, &n.bounds});
            }
        }
    }
    const uint32_t n = (uint32_t)b.src.size();
    b.cen.resize(n);
    b.order.resize(n);
    for (uint32_t i = 0; i < n; ++i) {
        const OptPrim& p = b.src[i];
        b.cen[i] = Point3{(p.v[0].x + p.v[1].x + p.v[2].x) * (1.0f / 3.0f), (p.v[0].y + p.v[1].y + p.v[2].y) * (1.0f / 3.0f),
                          (p.v[0].z + p.v[1].z + p.v[2].z) * (1.0f / 3.0f)};
        b.order[i] = i;
    }
    t->nodes.reserve(n ? 2 * (size_t)n : 1);
    t->nodes.emplace_back();
    if (n) b.build(0, 0, n, 0);
    else t->nodes.clear();
    t->prims.resize(n);
    for (uint32_t i = 0; i < n; ++i) t->prims[i] = b.src[b.order[i]];
    t->build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    s.opt = t;
    return true;
}

// ---------------------------------------------------------------------- shading

// rayca-model/src/sampler.rs:11-29
Color sample_image(const Scene& s, uint32_t image_index, Vec2 uv) {
    const rayca_image& img = s.images[image_index];
    float width = (float)img.width, height = (float)img.height;
    float x = (uv.x - std::floor(uv.x) + 1.0f) * width;
    float y = (uv.y - std::floor(uv.y) + 1.0f) * height;
    auto as_u32 = [](float f) -> uint32_t { // Rust `as u32`: saturating, NaN -> 0
        if (!(f == f)) return 0u;
        if (f <= 0.0f) return 0u;
        if (f >= 4294967296.0f) return 0xFFFFFFFFu;
        return (uint32_t)f;
    };
    uint32_t xi = as_u32(x) % img.width;
    uint32_t yi = as_u32(y) % img.height;
    size_t idx = (size_t)yi * img.width + xi;
    const uint8_t* data = s.image_data[image_index].data();
    switch (img.color_type) {
        case RAYCA_COLOR_RGBA32F: {
            const float* f = reinterpret_cast<const float*>(data) + idx * 4;
            return {f[0] / 255.0f, f[1] / 255.0f, f[2] / 255.0f, f[3] / 255.0f}; // color/mod.rs:222-231
        }
        case RAYCA_COLOR_RGBA8: {
            const uint8_t* p = data + idx * 4;
            return {(float)p[0] / 255.0f, (float)p[1] / 255.0f, (float)p[2] / 255.0f, (float)p[3] / 255.0f};
        }
        default: {
            const uint8_t* p = data + idx * 3;
            return {(float)p[0] / 255.0f, (float)p[1] / 255.0f, (float)p[2] / 255.0f, 255.0f / 255.0f};
        }
    }
}

struct Config {
    rayca_config c;
    uint64_t seed;
};

// Philox4x32-10, key = (seed_lo ^ depth-mixed, seed_hi), counter = (pixel, sample, depth, dim)
inline void philox4x32_10(uint32_t ctr[4], uint32_t key[2]) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)M0 * ctr[0];
        uint64_t p1 = (uint64_t)M1 * ctr[2];
        uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
        uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
        uint32_t n0 = hi1 ^ ctr[1] ^ key[0];
        uint32_t n1 = lo1;
        uint32_t n2 = hi0 ^ ctr[3] ^ key[1];
        uint32_t n3 = lo0;
        ctr[0] = n0; ctr[1] = n1; ctr[2] = n2; ctr[3] = n3;
        key[0] += W0;
        key[1] += W1;
    }
}
// uniform in [0,1) with 24 bits, like fastrand::f32
inline float rng_f32(uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t depth, uint32_t dim) {
    uint32_t ctr[4] = {pixel, sample, depth, dim};
    uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    philox4x32_10(ctr, key);
    return (float)(ctr[0] >> 8) * (1.0f / 16777216.0f);
}

// HitInfo (hit.rs:7-246): lazily cached attributes of one hit
struct HitInfo {
    const Scene& s;
    Hit hit;
    const Prim* prim;
    uint32_t pixel = 0, sample = 0; // RNG key
    bool has_color = false, has_normal = false, has_diffuse = false, has_reflection = false, has_next = false;
    Color color, diffuse;
    Vec3 normal, reflection;
    Point3 next_origin;

    HitInfo(const Scene& sc, const Hit& h) : s(sc), hit(h) { prim = &s.blass[s.blas_nodes[h.blas].blas].prims[h.primitive]; }

    const rayca_model& model() const { return s.models[prim->model]; }
    // primitive.rs:103-110 — missing material => Material::DEFAULT = Pbr(NONE)
    rayca_material material() const {
        if (prim->material < 0 || (uint32_t)prim->material >= model().material_count) return {RAYCA_MATERIAL_PBR, RAYCA_NONE};
        return s.materials[model().material_offset + (uint32_t)prim->material];
    }
    rayca_pbr_material pbr(const rayca_material& m) const {
        if (m.index < 0 || (uint32_t)m.index >= model().pbr_count) { // PbrMaterial::WHITE (pbr.rs:73-80)
            rayca_pbr_material w{{1, 1, 1, 1}, RAYCA_NONE, RAYCA_NONE, 0.0f, 1.0f, RAYCA_NONE};
            return w;
        }
        return s.pbr[model().pbr_offset + (uint32_t)m.index];
    }
    rayca_phong_material phong(const rayca_material& m) const {
        if (m.index < 0 || (uint32_t)m.index >= model().phong_count) { // PhongMaterial::DEFAULT
            rayca_phong_material d{{0, 0, 0, 1}, {0, 0, 0, 1}, {0, 0, 0, 1}, {0, 0, 0, 1}, 0.0f};
            return d;
        }
        return s.phong[model().phong_offset + (uint32_t)m.index];
    }
    rayca_ggx_material ggx(const rayca_material& m) const {
        if (m.index < 0 || (uint32_t)m.index >= model().ggx_count) {
            rayca_ggx_material d{{0, 0, 0, 1}, {0, 0, 0, 1}, 0.0f};
            return d;
        }
        return s.ggx[model().ggx_offset + (uint32_t)m.index];
    }
    // texture handle -> global image index or -1 (model.textures.get(handle))
    int tex_image(int32_t tex) const {
        if (tex < 0 || (uint32_t)tex >= model().texture_count) return -1;
        int32_t img = s.textures[model().texture_offset + (uint32_t)tex].image;
        return (int)(model().image_offset + (uint32_t)img);
    }

    // bvh/triangle.rs:34-71 — v2*(1-u-v) + v0*u + v1*v
    Color geometry_color() const {
        if (prim->sphere) return WHITE; // primitive.rs:19
        float u = hit.uv.x, v = hit.uv.y;
        return prim->ext[2].color * (1.0f - u - v) + prim->ext[0].color * u + prim->ext[1].color * v;
    }
    Vec2 get_uv() const {
        if (prim->sphere) return {0.0f, 0.0f}; // primitive.rs:26-27
        float u = hit.uv.x, v = hit.uv.y;
        return prim->ext[2].uv * (1.0f - u - v) + prim->ext[0].uv * u + prim->ext[1].uv * v;
    }
    Vec3 interp(const Vec3 VertexExt::*field) const {
        float u = hit.uv.x, v = hit.uv.y;
        Vec3 n = (prim->ext[2].*field * (1.0f - u - v) + prim->ext[0].*field * u) + prim->ext[1].*field * v;
        return normalized(n);
    }
    // pbr.rs:94-102
    Color pbr_color(const rayca_pbr_material& p, Vec2 uv) const {
        int img = tex_image(p.albedo);
        if (img >= 0) return to_color(p.color) * sample_image(s, (uint32_t)img, uv);
        return to_color(p.color);
    }
    // pbr.rs:125-136 — metallic = B, roughness = R
    void metallic_roughness(float& metallic, float& roughness) const {
        rayca_pbr_material p = pbr(material());
        int img = tex_image(p.metallic_roughness);
        if (img >= 0) {
            Color c = sample_image(s, (uint32_t)img, get_uv());
            metallic = c.b;
            roughness = c.r;
        } else {
            metallic = p.metallic_factor;
            roughness = p.roughness_factor;
        }
    }
    // material/mod.rs:107-114
    Color material_color(Vec2 uv) const {
        rayca_material m = material();
        switch (m.kind) {
            case RAYCA_MATERIAL_PBR: return pbr_color(pbr(m), uv);
            case RAYCA_MATERIAL_PHONG: { // phong.rs:58-60 ambient + emission
                rayca_phong_material p = phong(m);
                return to_color(p.ambient) + to_color(p.emission);
            }
            default: return to_color(ggx(m).diffuse);
        }
    }
    // material/mod.rs:116-123
    Color material_diffuse(Vec2 uv) const {
        rayca_material m = material();
        switch (m.kind) {
            case RAYCA_MATERIAL_PBR: return pbr_color(pbr(m), uv);
            case RAYCA_MATERIAL_PHONG: return to_color(phong(m).diffuse);
            default: return to_color(ggx(m).diffuse);
        }
    }
    // hit.rs:65-71 / primitive.rs:142-148
    Color get_color() {
        if (!has_color) {
            color = geometry_color() * material_color(get_uv());
            has_color = true;
        }
        return color;
    }
    // hit.rs:138-145 / primitive.rs:150-155
    Color get_diffuse() {
        if (!has_diffuse) {
            diffuse = geometry_color() * material_diffuse(get_uv());
            has_diffuse = true;
        }
        return diffuse;
    }
    // material/mod.rs:143-152
    Color get_specular() const {
        rayca_material m = material();
        switch (m.kind) {
            case RAYCA_MATERIAL_PHONG: return to_color(phong(m).specular);
            case RAYCA_MATERIAL_PBR: {
                rayca_pbr_material p = pbr(m);
                Color base = pbr_color(p, get_uv());
                float metallic, roughness;
                metallic_roughness(metallic, roughness);
                return metallic * base;
            }
            default: return to_color(ggx(m).specular);
        }
    }
    // material/mod.rs:171-185
    float get_roughness() const {
        rayca_material m = material();
        switch (m.kind) {
            case RAYCA_MATERIAL_PHONG: {
                float sh = phong(m).shininess;
                return clampf(std::sqrt(2.0f / (sh + 2.0f)), 0.0f, 1.0f);
            }
            case RAYCA_MATERIAL_PBR: {
                float metallic, roughness;
                metallic_roughness(metallic, roughness);
                return roughness;
            }
            default: return ggx(m).roughness;
        }
    }
    float get_shininess() const { return phong(material()).shininess; } // material/mod.rs:154-160 (Phong only)
    // specular weight: phong.rs:64-71, ggx.rs:54-61 (Vec3::reduce_avg = ordered 4-lane sum / 3); todo!() for PBR
    bool t_unsupported = false;
    float get_t() {
        rayca_material m = material();
        float kd[3], ks[3];
        if (m.kind == RAYCA_MATERIAL_PHONG) {
            const rayca_phong_material pm = phong(m);
            for (int k = 0; k < 3; ++k) { kd[k] = pm.diffuse[k]; ks[k] = pm.specular[k]; }
        } else if (m.kind == RAYCA_MATERIAL_GGX) {
            const rayca_ggx_material gm = ggx(m);
            for (int k = 0; k < 3; ++k) { kd[k] = gm.diffuse[k]; ks[k] = gm.specular[k]; }
        } else {
            t_unsupported = true;
            return 0.0f;
        }
        float kd_avg = (((kd[0] + kd[1]) + kd[2]) + 0.0f) / 3.0f, ks_avg = (((ks[0] + ks[1]) + ks[2]) + 0.0f) / 3.0f;
        if (ks_avg == 0.0f && kd_avg == 0.0f) return 1.0f;
        float t = ks_avg / (ks_avg + kd_avg);
        return m.kind == RAYCA_MATERIAL_GGX ? std::max(t, 0.25f) : t;
    }
    // hit.rs:77-83 / primitive.rs:171-181 / pbr.rs:104-123
    Vec3 get_normal() {
        if (!has_normal && prim->sphere) { // primitive.rs:183-190
            const Trs& trs = s.world_trs[prim->trs];
            Point3 hit_point = mat4_from_inversed_trs(trs) * hit.point;
            Vec3 n = normalized(hit_point - prim->center); // Sphere::get_model_normal (sphere.rs:95-97)
            normal = normalized(transpose(mat3_from_inversed_trs(trs)) * n);
            has_normal = true;
        }
        if (!has_normal) {
            Vec3 n = interp(&VertexExt::normal);
            rayca_material m = material();
            if (m.kind == RAYCA_MATERIAL_PBR) {
                rayca_pbr_material p = pbr(m);
                int img = tex_image(p.normal);
                if (img >= 0) {
                    Vec3 t = interp(&VertexExt::tangent);
                    Vec3 b = interp(&VertexExt::bitangent);
                    Vec3 sampled = color_to_vec3(sample_image(s, (uint32_t)img, get_uv()));
                    sampled = sub_scalar(sampled * 2.0f, 1.0f);
                    n = normalized(mat3_tbn(t, b, n) * sampled);
                }
            }
            normal = n;
            has_normal = true;
        }
        return normal;
    }
    Vec3 get_view() const { return -hit.ray.dir; }
    // hit.rs:93-101
    Vec3 get_reflection() {
        if (!has_reflection) {
            reflection = normalized(reflect(hit.ray.dir, get_normal()));
            has_reflection = true;
        }
        return reflection;
    }
    // hit.rs:164-176
    Point3 get_next_ray_origin() {
        if (!has_next) {
            next_origin = hit.point + get_normal() * RAY_BIAS;
            has_next = true;
        }
        return next_origin;
    }
    Ray get_next_ray(Vec3 dir) { return make_ray(get_next_ray_origin(), dir); }
    // primitive.rs:122-130
    // primitive.rs:132-140: Phong emission, BLACK for PBR / GGX
    Color get_emission() const {
        rayca_material m = material();
        if (m.kind == RAYCA_MATERIAL_PHONG) return to_color(phong(m).emission);
        return BLACK;
    }
    bool is_emissive() const {
        rayca_material m = material();
        if (m.kind == RAYCA_MATERIAL_PHONG) return !color_close(to_color(phong(m).emission), BLACK);
        return false;
    }
};

// brdf/ggx.rs:58-127 (path-tracer form)
float ggx_get_d(float a, Vec3 h, Vec3 n) {
    float a2 = a * a;
    float cos_theta = clampf(dot(h, n), 0.0f, 1.0f);
    float theta = std::acos(cos_theta);
    float denominator = std::pow(cos_theta, 4.0f) * std::pow(a2 + std::pow(std::tan(theta), 2.0f), 2.0f);
    if (denominator == 0.0f) return 0.0f;
    return a2 * FRAC_1_PI / denominator;
}
float ggx_get_g1(float a, Vec3 omega, Vec3 n) {
    float cos_theta = dot(omega, n);
    if (cos_theta <= 0.0f) return 0.0f;
    float theta = std::acos(cos_theta);
    float denominator = 1.0f + std::sqrt(1.0f + a * a * std::pow(std::tan(theta), 2.0f));
    return 2.0f / denominator;
}
Color ggx_get_f(Color ks, Vec3 omega_i, Vec3 h) {
    float d = std::fabs(dot(omega_i, h));
    return ks + (WHITE - ks) * std::pow(1.0f - d, 5.0f);
}
Color ggx_get_bsdf(HitInfo& hit, Vec3 omega_i) {
    Vec3 omega_o = hit.get_view();
    Vec3 n = hit.get_normal();
    float i_dot_n = clampf(dot(omega_i, n), 0.0f, 1.0f);
    float o_dot_n = clampf(dot(omega_o, n), 0.0f, 1.0f);
    if (i_dot_n == 0.0f || o_dot_n == 0.0f) return BLACK;
    Color ks = hit.get_specular();
    float a = hit.get_roughness();
    Vec3 h = normalized(omega_i + omega_o);
    Color f = ggx_get_f(ks, omega_i, h);
    float g = ggx_get_g1(a, omega_i, n) * ggx_get_g1(a, omega_o, n);
    float d = ggx_get_d(a, h, n);
    float denominator = 4.0f * i_dot_n * o_dot_n;
    return f * g * d / denominator;
}
Color ggx_get_brdf(HitInfo& hit, Vec3 omega_i) {
    Color kd = hit.get_diffuse();
    return kd * FRAC_1_PI + ggx_get_bsdf(hit, omega_i);
}
// brdf/lambertian.rs:7-15
Color lambertian_get_brdf(HitInfo& hit, Vec3 omega_i) {
    Color lambertian = hit.get_diffuse() * FRAC_1_PI;
    float s = hit.get_shininess();
    Color specular = (hit.get_specular() * (s + 2.0f) * std::pow(dot(hit.get_reflection(), omega_i), s) * FRAC_1_PI) / 2.0f;
    return lambertian + specular;
}
// brdf/lambertian.rs:70-76, brdf/ggx.rs:188-203; hit.rs:238-245
Color get_specular_component(HitInfo& hit, Vec3 omega) {
    if (hit.material().kind == RAYCA_MATERIAL_PHONG) {
        float sh = hit.get_shininess();
        float n_dot_omega = clampf(dot(hit.get_normal(), omega), 0.0f, 1.0f);
        return hit.get_specular() * n_dot_omega * (sh + 2.0f) / (sh + 1.0f);
    }
    Vec3 omega_o = hit.get_view();
    Vec3 n = hit.get_normal();
    float o_dot_n = clampf(dot(omega_o, n), 0.0f, 1.0f);
    Vec3 h = normalized(omega + omega_o);
    float n_dot_h = clampf(dot(n, h), 0.0f, 1.0f);
    if (o_dot_n == 0.0f || n_dot_h == 0.0f) return BLACK;
    Color ks = hit.get_specular();
    float a = hit.get_roughness();
    float h_dot_omega = clampf(dot(h, omega), 0.0f, 1.0f);
    return ggx_get_f(ks, omega, h) * (ggx_get_g1(a, omega, n) * ggx_get_g1(a, omega_o, n)) * h_dot_omega / (o_dot_n * n_dot_h);
}
// lambertian::get_random_dir (lambertian.rs:31-68) / ggx::get_random_dir (ggx.rs:146-186) with e0, e1, e2 supplied
Vec3 material_random_dir(HitInfo& hit, float e0, float e1, float e2) {
    const bool is_phong = hit.material().kind == RAYCA_MATERIAL_PHONG;
    float t = hit.get_t();
    float theta;
    if (e0 <= t) {
        if (is_phong) theta = std::acos(clampf(std::pow(e1, 1.0f / (hit.get_shininess() + 1.0f)), -1.0f, 1.0f));
        else theta = std::atan((hit.get_roughness() * std::sqrt(e1)) / std::sqrt(1.0f - e1));
    } else {
        theta = std::acos(clampf(std::sqrt(e1), -1.0f, 1.0f));
    }
    float omega = 2.0f * PI * e2;
    Vec3 sv{std::cos(omega) * std::sin(theta), std::sin(omega) * std::sin(theta), std::cos(theta)};
    Vec3 w = (is_phong && e0 <= t) ? hit.get_reflection() : hit.get_normal();
    Vec3 a = vec3_close(w, Vec3{0, 1, 0}) ? Vec3{1, 0, 0} : Vec3{0, 1, 0};
    Vec3 u = normalized(cross(a, w));
    Vec3 v = cross(w, u);
    Vec3 dir = (sv.x * u + sv.y * v) + sv.z * w;
    if (!is_phong && e0 <= t) { // the sampled vector is the half vector: reflect -omega_o about it (vec3.rs:219-221)
        Vec3 mo = -hit.get_view();
        return mo - (2.0f * dot(mo, dir)) * dir;
    }
    return dir;
}
// lambertian::get_pdf (lambertian.rs:17-26) / ggx::get_pdf (ggx.rs:129-144); hit.rs:229-236
float material_pdf(HitInfo& hit, Vec3 omega) {
    constexpr float FRAC_2_PI = 0.636619772367581343075535053490057448f;
    if (hit.material().kind == RAYCA_MATERIAL_PHONG) {
        float r_dot_omega = clampf(dot(hit.get_reflection(), omega), 0.0f, 1.0f);
        float sh = hit.get_shininess();
        float spec = (sh + 1.0f) * FRAC_2_PI * std::pow(r_dot_omega, sh);
        float diff = clampf(dot(hit.get_normal(), omega), 0.0f, 1.0f) * FRAC_1_PI;
        float t = hit.get_t();
        return (1.0f - t) * diff + t * spec;
    }
    Vec3 omega_o = hit.get_view();
    Vec3 h = normalized(omega_o + omega);
    float h_dot_omega = clampf(dot(h, omega), 0.0f, 1.0f);
    if (h_dot_omega == 0.0f) return 0.0f;
    Vec3 n = hit.get_normal();
    float n_dot_h = clampf(dot(n, h), 0.0f, 1.0f);
    float spec = ggx_get_d(hit.get_roughness(), h, n) * n_dot_h / (4.0f * h_dot_omega);
    float dif = clampf(dot(n, omega), 0.0f, 1.0f) * FRAC_1_PI;
    float t = hit.get_t();
    return (1.0f - t) * dif + t * spec;
}
// hit.rs:220-227
Color get_brdf(HitInfo& hit, Vec3 omega_i) {
    if (hit.material().kind == RAYCA_MATERIAL_PHONG) return lambertian_get_brdf(hit, omega_i);
    return ggx_get_brdf(hit, omega_i);
}

// hit.rs:251-286
struct Irradiance {
    Color intensity;
    float n_dot_v, n_dot_l;
    Vec3 h;
    float n_dot_h, l_dot_h;
};
Irradiance make_irradiance(Color intensity, HitInfo& hit, Vec3 light_dir) {
    Irradiance ir;
    Vec3 l = light_dir;
    Vec3 n = hit.get_normal();
    Vec3 v = -hit.hit.ray.dir;
    ir.intensity = intensity;
    ir.n_dot_v = clampf(dot(n, v), 0.0f, 1.0f) + 1e-5f;
    ir.n_dot_l = clampf(dot(n, l), 0.0f, 1.0f);
    ir.h = normalized(v + l);
    ir.n_dot_h = clampf(dot(n, ir.h), 0.0f, 1.0f);
    ir.l_dot_h = clampf(dot(l, ir.h), 0.0f, 1.0f);
    return ir;
}
// brdf/ggx.rs:11-52 (raytracer form)
Color ggx_get_radiance(HitInfo& hit, const Irradiance& ir) {
    float metallic, roughness;
    hit.metallic_roughness(metallic, roughness);
    float a = ir.n_dot_h * roughness;
    float k = roughness / (1.0f - ir.n_dot_h * ir.n_dot_h + a * a);
    float d = k * k * FRAC_1_PI;
    Color albedo = hit.get_color();
    Vec3 f0 = Vec3{0.04f, 0.04f, 0.04f} * (1.0f - metallic) + color_to_vec3(albedo) * metallic;
    float fp = std::pow(1.0f - ir.l_dot_h, 5.0f);
    Vec3 f = f0 + (Vec3{1.0f, 1.0f, 1.0f} - f0) * fp;
    Vec3 ks = f;
    Vec3 kd = (Vec3{1.0f, 1.0f, 1.0f} - ks) * (1.0f - metallic);
    float aa = roughness;
    float ggxv = ir.n_dot_l * (ir.n_dot_v * (1.0f - aa) + aa);
    float ggxl = ir.n_dot_v * (ir.n_dot_l * (1.0f - aa) + aa);
    float g = 0.5f / (ggxv + ggxl);
    Color fr = (d * g) * vec3_to_color(f);
    Color fd = vec3_to_color(kd) * albedo * FRAC_1_PI;
    return (fd + fr) * ir.intensity * ir.n_dot_l;
}
// brdf/lambertian.rs:77-81
Color lambertian_get_radiance(HitInfo& hit, const Irradiance& ir) {
    Color diffuse = hit.get_diffuse() * ir.n_dot_l;
    Color specular = hit.get_specular() * std::pow(ir.n_dot_h, hit.get_shininess());
    return (diffuse + specular) * ir.intensity;
}
// hit.rs:201-208
Color get_radiance(HitInfo& hit, const Irradiance& ir) {
    if (hit.material().kind == RAYCA_MATERIAL_PHONG) return lambertian_get_radiance(hit, ir);
    return ggx_get_radiance(hit, ir);
}

// light/point.rs:37-48
float point_fallof(const rayca_light& l, const Trs& trs, Point3 frag) {
    Vec3 dist = to_vec3(frag) - trs_get_translation(trs);
    float r2 = norm(dist);
    float r = std::sqrt(r2);
    float p0 = l.attenuation[0] * 1.0f, p1 = l.attenuation[1] * r, p2 = l.attenuation[2] * r2, p3 = 0.0f * 0.0f;
    return ((p0 + p1) + p2) + p3;
}
Color point_intensity(const rayca_light& l, const Trs& trs, Point3 frag) {
    return (l.intensity * to_color(l.color)) / point_fallof(l, trs, frag);
}

// Russian-roulette paths have no depth limit in the reference (pathtracer.rs:76); with the keyed RNG both this
// oracle and the kernel stop a path that is still alive after this many bounces.
constexpr uint32_t kRouletteDepthCap = 64;

struct Renderer {
    const Scene& s;
    Config cfg;
    Tracer tracer;
    int status = 0;

    // sampler/cosine.rs:65-88 with the keyed RNG in place of fastrand
    Vec3 cosine_random_dir(HitInfo& hit, uint32_t depth, uint32_t light_sample) {
        float e1 = rng_f32(cfg.seed, hit.pixel, hit.sample, depth, 2 * light_sample);
        float e2 = rng_f32(cfg.seed, hit.pixel, hit.sample, depth, 2 * light_sample + 1);
        float theta = std::acos(std::sqrt(e1));
        float omega = 2.0f * PI * e2;
        Vec3 sv{std::cos(omega) * std::sin(theta), std::sin(omega) * std::sin(theta), std::cos(theta)};
        Vec3 w = hit.get_normal();
        Vec3 a = vec3_close(w, Vec3{0, 1, 0}) ? Vec3{1, 0, 0} : Vec3{0, 1, 0};
        Vec3 u = normalized(cross(a, w));
        Vec3 v = cross(w, u);
        return (sv.x * u + sv.y * v) + sv.z * w;
    }

    // sampler/hemisphere.rs:17-40 with the keyed RNG in place of fastrand
    Vec3 hemisphere_random_dir(HitInfo& hit, uint32_t depth, uint32_t light_sample) {
        float e1 = rng_f32(cfg.seed, hit.pixel, hit.sample, depth, 2 * light_sample);
        float e2 = rng_f32(cfg.seed, hit.pixel, hit.sample, depth, 2 * light_sample + 1);
        float theta = std::acos(e1);
        float omega = 2.0f * PI * e2;
        Vec3 sv{std::cos(omega) * std::sin(theta), std::sin(omega) * std::sin(theta), std::cos(theta)};
        Vec3 w = hit.get_normal();
        Vec3 a = vec3_close(w, Vec3{0, 1, 0}) ? Vec3{1, 0, 0} : Vec3{0, 1, 0};
        Vec3 u = normalized(cross(a, w));
        Vec3 v = normalized(cross(w, u));
        return (sv.x * u + sv.y * v) + sv.z * w;
    }

    // Keyed-RNG dimensions at one (pixel, sample, depth): [0, 2 LS) indirect directions, [2 LS, 3 LS) russian
    // roulette, then two per (light, light sample) for the quad-light points.
    uint32_t dim_roulette(uint32_t ls) const { return 2 * cfg.c.light_samples + ls; }
    uint32_t dim_lobe(uint32_t ls) const { return 0x10000u + ls; } // specular / diffuse lobe choice of the Brdf sampler
    uint32_t dim_quad(uint32_t light_slot, uint32_t ls) const { return 3 * cfg.c.light_samples + 2 * (light_slot * cfg.c.light_samples + ls); }

    // NextEventEstimationSampler::get_quad_light_sample (sampler/nee.rs:46-93); QuadLight::get_random_point
    // (quad.rs:112-133).  The sample counts only when the CLOSEST hit along omega is emissive.
    Color nee_quad_sample(HitInfo& hit, const LightDraw& ld, const rayca_light& l, uint32_t light_slot, uint32_t i, uint32_t depth,
                          Vec3* omega_out = nullptr) {
        Color ldc = BLACK;
        const uint32_t ls = cfg.c.light_samples;
        const uint32_t strate_count = cfg.c.light_stratify ? (uint32_t)std::sqrt((float)ls) : 1u; // config.rs:73-79
        Vec3 ab{l.ab[0], l.ab[1], l.ab[2]}, ac{l.ac[0], l.ac[1], l.ac[2]};
        float area = quad_area(l);
        Vec3 step1 = ab / (float)strate_count, step2 = ac / (float)strate_count;
        float u1 = rng_f32(cfg.seed, hit.pixel, hit.sample, depth, dim_quad(light_slot, i)) / (float)strate_count;
        float u2 = rng_f32(cfg.seed, hit.pixel, hit.sample, depth, dim_quad(light_slot, i) + 1) / (float)strate_count;
        Point3 x1 = (quad_corner(l, ld.node_trs, 0) + (u1 * ab)) + (u2 * ac);
        if (cfg.c.light_stratify) {
            float i1 = (float)(i % strate_count), i2 = (float)(i / strate_count);
            x1 = x1 + ((step1 * i1) + (step2 * i2));
        }
        Vec3 x_to_x1 = x1 - hit.hit.point;
        Vec3 omega = normalized(x_to_x1);
        if (omega_out) *omega_out = omega;
        Ray shadow = hit.get_next_ray(omega);
        tracer.rays->rays_shadow++;
        Hit sh;
        if (tracer.intersects(shadow, sh)) {
            HitInfo shi(s, sh);
            if (shi.is_emissive()) {
                Color le = l.intensity * to_color(l.color);
                Color brdf = get_brdf(hit, omega);
                float r_squared = norm(x_to_x1);
                float d_omega = dot(quad_normal(l), omega) / r_squared;
                float n_dot_omega = clampf(dot(hit.get_normal(), omega), 0.0f, 1.0f);
                ldc = le * area * brdf * n_dot_omega * d_omega;
            }
        }
        return ldc;
    }

    // NextEventEstimationSampler::get_point_light_sample (sampler/nee.rs:95-137)
    Color nee_point_sample(HitInfo& hit, const LightDraw& ld, const rayca_light& light, Vec3* omega_out = nullptr) {
        Point3 x1 = to_point3(trs_get_translation(ld.node_trs));
        Point3 x = hit.hit.point;
        Vec3 x_to_x1 = x1 - x;
        float dist = len(x_to_x1);
        Vec3 omega = normalized(x_to_x1);
        if (omega_out) *omega_out = omega;
        Ray shadow = hit.get_next_ray(omega);
        Hit sh;
        tracer.rays->rays_shadow++;
        if (tracer.intersects(shadow, sh) && sh.depth < dist) return BLACK;
        Color le = point_intensity(light, ld.node_trs, x);
        Color brdf = get_brdf(hit, omega);
        float r2 = norm(x_to_x1);
        float d_omega = 1.0f / r2;
        float n_dot_omega = clampf(dot(hit.get_normal(), omega), 0.0f, 1.0f);
        return le * brdf * n_dot_omega * d_omega;
    }

    // sampler/nee.rs:127-216 (directional = todo!())
    Color nee_direct(HitInfo& hit, uint32_t depth = 0) {
        Color ret = BLACK;
        uint32_t light_slot_next = 0;
        for (const LightDraw& ld : s.light_draws) {
            const rayca_light& light = s.lights[ld.light];
            const uint32_t light_slot = light_slot_next++;
            for (uint32_t ls = 0; ls < cfg.c.light_samples; ++ls) {
                if (light.kind == RAYCA_LIGHT_QUAD) {
                    ret = ret + nee_quad_sample(hit, ld, light, light_slot, ls, depth);
                    continue;
                }
                if (light.kind != RAYCA_LIGHT_POINT) {
                    status = RAYCA_E_UNSUPPORTED;
                    continue;
                }
                ret = ret + nee_point_sample(hit, ld, light);
            }
        }
        return ret;
    }

    // QuadLight::intersects over its two "wide" triangles (quad.rs:135-158; the `ab + ab` of corner c is the
    // reference's); first hit in list order, single-sided like every triangle test
    bool quad_light_intersects(const rayca_light& l, const Trs& trs, const Ray& ray, Point3& point) {
        const float SMALL_BIAS = 1e-2f;
        Vec3 ab{l.ab[0], l.ab[1], l.ab[2]}, ac{l.ac[0], l.ac[1], l.ac[2]};
        Point3 a = to_point3(-normalized(ab + ac) * SMALL_BIAS);
        Point3 b = to_point3(ab + normalized(ab - ac) * SMALL_BIAS);
        Point3 c = to_point3((ab + ac) + normalized(ab + ab) * SMALL_BIAS);
        Point3 d = to_point3(ac + normalized(ac - ab) * SMALL_BIAS);
        Point3 tri[2][3] = {{a, c, b}, {a, d, c}};
        for (int k = 0; k < 2; ++k) {
            float t;
            Vec2 uv;
            if (triangle_intersects_world(transform_point(trs, tri[k][0]), transform_point(trs, tri[k][1]), transform_point(trs, tri[k][2]),
                                          ray, t, point, uv))
                return true;
        }
        return false;
    }
    // NextEventEstimationSample::get_pdf (sampler/nee.rs:44-65)
    float nee_pdf(const LightDraw& ld, HitInfo& hit, Vec3 omega) {
        const rayca_light& l = s.lights[ld.light];
        if (l.kind != RAYCA_LIGHT_QUAD) return 0.0f; // Light::intersects is None for point / directional lights
        Ray ray = hit.get_next_ray(omega);
        Point3 light_point;
        if (!quad_light_intersects(l, ld.node_trs, ray, light_point)) return 0.0f;
        float area = quad_area(l);
        if (area == 0.0f) return 0.0f;
        float nl_dot_omega = clampf(dot(quad_normal(l), omega), 0.0f, 1.0f);
        if (nl_dot_omega == 0.0f) return 0.0f;
        float r_squared = norm(light_point - hit.hit.point);
        return r_squared / (area * nl_dot_omega);
    }
    // MultipleImportanceSampling::get_direct_lighting (sampler/mis.rs:41-72) with BrdfSampler::sample_direct
    // (sampler/brdf.rs:48-66); the Brdf sample draws its three numbers at dimensions 0x20000..0x20002
    Color mis_direct(HitInfo& hit, uint32_t depth) {
        struct LightSample { Vec3 omega; Color x; };
        std::vector<LightSample> samples;
        uint32_t light_slot_next = 0;
        for (const LightDraw& ld : s.light_draws) {
            const rayca_light& light = s.lights[ld.light];
            const uint32_t light_slot = light_slot_next++;
            for (uint32_t ls = 0; ls < cfg.c.light_samples; ++ls) {
                LightSample smp;
                if (light.kind == RAYCA_LIGHT_QUAD) {
                    smp.x = nee_quad_sample(hit, ld, light, light_slot, ls, depth, &smp.omega);
                } else if (light.kind == RAYCA_LIGHT_POINT) {
                    smp.x = nee_point_sample(hit, ld, light, &smp.omega);
                } else {
                    status = RAYCA_E_UNSUPPORTED;
                    continue;
                }
                samples.push_back(smp);
            }
        }
        // Brdf sample
        float e0 = rng_f32(cfg.seed, hit.pixel, hit.sample, depth, 0x20000u);
        float e1 = rng_f32(cfg.seed, hit.pixel, hit.sample, depth, 0x20001u);
        float e2 = rng_f32(cfg.seed, hit.pixel, hit.sample, depth, 0x20002u);
        Vec3 omega_b = material_random_dir(hit, e0, e1, e2);
        if (hit.t_unsupported) { status = RAYCA_E_UNSUPPORTED; return BLACK; }
        float pdf_b = material_pdf(hit, omega_b);
        Color x_b = BLACK;
        {
            Ray shadow = hit.get_next_ray(omega_b);
            tracer.rays->rays_shadow++;
            Hit sh;
            if (tracer.intersects(shadow, sh)) {
                HitInfo shi(s, sh);
                if (shi.is_emissive()) {
                    Color li = shi.get_emission();
                    Color cd = hit.get_diffuse();
                    Color cs = get_specular_component(hit, omega_b);
                    x_b = li * (cd + cs);
                }
            }
        }
        auto pdf_nee_of = [&](Vec3 omega) {
            float ret = 0.0f;
            uint32_t k = 0;
            for (const LightDraw& ld : s.light_draws) // one term per light SAMPLE (get_pdf_nee sums over the sample list)
                for (uint32_t ls = 0; ls < cfg.c.light_samples; ++ls, ++k) ret += nee_pdf(ld, hit, omega);
            return ret / (float)samples.size();
        };
        Color ret = BLACK;
        for (const LightSample& smp : samples) {
            float pdf_nee = pdf_nee_of(smp.omega);
            float pdf_brdf = material_pdf(hit, smp.omega);
            float pdf_den = pdf_nee * pdf_nee + pdf_brdf * pdf_brdf;
            float w = pdf_den == 0.0f ? 0.0f : (pdf_nee * pdf_nee) / pdf_den;
            ret = ret + w * smp.x;
        }
        {
            float pdf_nee = pdf_nee_of(omega_b);
            float pdf_den = pdf_nee * pdf_nee + pdf_b * pdf_b;
            float w = pdf_den == 0.0f ? 0.0f : (pdf_b * pdf_b) / pdf_den;
            ret = ret + w * x_b;
        }
        return ret;
    }

    // integrator/pathtracer.rs:23-66
    Color indirect_lighting(HitInfo& hit, uint32_t depth) {
        Color li = BLACK;
        bool collect_emissive = cfg.c.direct_sampler == RAYCA_SAMPLER_NONE;
        const bool hemisphere = cfg.c.indirect_sampler == RAYCA_SAMPLER_HEMISPHERE;
        const bool brdf_sampler = cfg.c.indirect_sampler == RAYCA_SAMPLER_BRDF;
        for (uint32_t ls = 0; ls < cfg.c.light_samples; ++ls) {
            Vec3 omega_i;
            if (brdf_sampler) { // sampler/brdf.rs:70-73: the material's own importance sampling; lobe choice at dim_lobe
                float e0 = rng_f32(cfg.seed, hit.pixel, hit.sample, depth, dim_lobe(ls));
                float e1 = rng_f32(cfg.seed, hit.pixel, hit.sample, depth, 2 * ls);
                float e2 = rng_f32(cfg.seed, hit.pixel, hit.sample, depth, 2 * ls + 1);
                omega_i = material_random_dir(hit, e0, e1, e2);
                if (hit.t_unsupported) { status = RAYCA_E_UNSUPPORTED; return BLACK; } // Material::get_t is todo!() for PBR
            } else {
                omega_i = hemisphere ? hemisphere_random_dir(hit, depth, ls) : cosine_random_dir(hit, depth, ls);
            }
            Color brdf_i = get_brdf(hit, omega_i);
            Ray next = hit.get_next_ray(omega_i);
            float weight = 1.0f;
            if (cfg.c.russian_roulette) { // pathtracer.rs:37-50, ray.rs:95-105
                Color next_throughput = 2.0f * PI * hit.hit.ray.throughput * brdf_i * clampf(dot(hit.get_normal(), omega_i), 0.0f, 1.0f);
                float mx = std::max(std::max(next_throughput.r, next_throughput.g), next_throughput.b);
                float q = 1.0f - std::min(mx, 1.0f);
                float e = clampf(rng_f32(cfg.seed, hit.pixel, hit.sample, depth, dim_roulette(ls)), 0.0f, 1.0f - F32_EPSILON);
                if (!(q < e)) continue; // the path has not survived
                weight = 1.0f / (1.0f - q);
                next.throughput = next_throughput * weight;
                if (depth + 1 >= kRouletteDepthCap) continue; // keyed-RNG paths are cut at the cap on both sides (documented)
            }
            if (cfg.c.russian_roulette || depth + 1 < cfg.c.max_depth) tracer.rays->rays_secondary++;
            Color sample;
            if (pathtracer_trace(next, depth + 1, collect_emissive, hit.pixel, hit.sample, sample)) {
                if (brdf_sampler) { // sampler/brdf.rs:80-90 — indirect_sample * (cd + cs) * weight
                    Color cd = hit.get_diffuse();
                    Color cs = get_specular_component(hit, omega_i);
                    li = li + sample * (cd + cs) * weight;
                } else if (hemisphere) { // sampler/hemisphere.rs:42-52 — 2 PI * brdf * cosine_law * indirect_sample * weight
                    Color brdf = get_brdf(hit, omega_i);
                    float cosine_law = clampf(dot(hit.get_normal(), omega_i), 0.0f, 1.0f);
                    li = li + 2.0f * PI * brdf * cosine_law * sample * weight;
                } else { // sampler/cosine.rs:90-99 — PI * brdf * indirect_sample * weight
                    Color brdf = get_brdf(hit, omega_i);
                    li = li + PI * brdf * sample * weight;
                }
            }
        }
        return li / (float)cfg.c.light_samples;
    }

    // integrator/pathtracer.rs:68-106
    bool pathtracer_trace(const Ray& ray, uint32_t depth, bool collect_emissive, uint32_t pixel, uint32_t sample, Color& out,
                          Hit* primary = nullptr) {
        if (!cfg.c.russian_roulette && depth >= cfg.c.max_depth) return false;
        Hit h;
        if (!tracer.intersects(ray, h)) return false;
        if (primary) *primary = h;
        HitInfo hit(s, h);
        hit.pixel = pixel;
        hit.sample = sample;
        Color ambient_and_emissive = hit.get_color();
        if (collect_emissive && hit.is_emissive()) {
            out = ambient_and_emissive;
            return true;
        }
        Color direct = BLACK;
        if (cfg.c.direct_sampler == RAYCA_SAMPLER_NEE) direct = nee_direct(hit, depth);
        else if (cfg.c.direct_sampler == RAYCA_SAMPLER_MIS) direct = mis_direct(hit, depth);
        uint32_t limit = (cfg.c.direct_sampler != RAYCA_SAMPLER_NONE) ? cfg.c.max_depth - 1 : cfg.c.max_depth;
        Color indirect = BLACK;
        if (cfg.c.russian_roulette || depth < limit) indirect = indirect + indirect_lighting(hit, depth);
        out = direct + indirect;
        return true;
    }

    // integrator/raytracer.rs:16-76
    bool raytracer_trace(const Ray& ray, uint32_t depth, Color& out, Hit* primary = nullptr) {
        if (depth > cfg.c.max_depth) return false;
        Hit h;
        if (!tracer.intersects(ray, h)) return false;
        if (primary) *primary = h;
        HitInfo hit(s, h);
        Color ambient_emissive = hit.get_color();
        Color light_contribution = BLACK;
        for (const LightDraw& ld : s.light_draws) {
            const rayca_light& light = s.lights[ld.light];
            Vec3 light_dir;
            if (light.kind == RAYCA_LIGHT_POINT) { // light/point.rs:50-54
                Vec3 dist = normalized(to_vec3(hit.hit.point) - trs_get_translation(ld.node_trs));
                light_dir = -dist;
            } else if (light.kind == RAYCA_LIGHT_DIRECTIONAL) { // light/directional.rs:47-51
                light_dir = -rotate(Vec3{1.0f, 0.0f, 0.0f}, ld.node_trs.rotation);
            } else {
                status = RAYCA_E_UNSUPPORTED; // Light::get_direction todo!() for quad
                continue;
            }
            Ray shadow = make_ray(hit.get_next_ray_origin(), light_dir);
            tracer.rays->rays_shadow++;
            Hit sh;
            bool is_light;
            if (!tracer.intersects(shadow, sh)) {
                is_light = true;
            } else {
                // light/point.rs:32-35, directional.rs:34-37
                float light_distance;
                if (light.kind == RAYCA_LIGHT_POINT)
                    light_distance = len(to_vec3(hit.hit.point - trs_get_translation(ld.node_trs)));
                else
                    light_distance = len(to_vec3(hit.hit.point) - trs_get_translation(ld.node_trs));
                if (sh.depth > light_distance) {
                    is_light = true;
                } else {
                    HitInfo shi(s, sh);
                    is_light = is_transparent(shi.get_color());
                }
            }
            if (is_light) {
                Color intensity = (light.kind == RAYCA_LIGHT_POINT) ? point_intensity(light, ld.node_trs, hit.hit.point)
                                                                    : light.intensity * to_color(light.color);
                Irradiance ir = make_irradiance(intensity, hit, light_dir);
                light_contribution = light_contribution + get_radiance(hit, ir);
            }
        }
        Ray reflection_ray = make_ray(hit.get_next_ray_origin(), hit.get_reflection());
        Color reflection_color;
        if (depth + 1 <= cfg.c.max_depth) tracer.rays->rays_secondary++;
        if (raytracer_trace(reflection_ray, depth + 1, reflection_color)) {
            float n_dot_r = 1.0f;
            light_contribution = light_contribution + reflection_color * hit.get_specular() * n_dot_r;
        }
        out = ambient_emissive + light_contribution;
        return true;
    }

    // integrator/scratcher.rs:16-89: transmission through transparent surfaces (Color::over), direct light with the
    // raytracer's shadow rule, reflection weighted by the material's radiance function
    bool scratcher_trace(const Ray& ray, uint32_t depth, Color& out, Hit* primary = nullptr) {
        if (depth > cfg.c.max_depth) return false;
        Hit h;
        if (!tracer.intersects(ray, h)) return false;
        if (primary) *primary = h;
        HitInfo hit(s, h);
        Color pixel_color = BLACK;
        if (is_transparent(hit.get_color())) {
            // hit.rs:178-187: origin pushed BELOW the surface, same direction
            Point3 transmit_origin = hit.hit.point + (-hit.get_normal()) * RAY_BIAS;
            Ray transmit_ray = make_ray(transmit_origin, ray.dir);
            if (depth + 1 <= cfg.c.max_depth) tracer.rays->rays_secondary++;
            Color transmit_color;
            if (scratcher_trace(transmit_ray, depth + 1, transmit_color)) {
                Color top = hit.get_color(); // Color::over (color/mod.rs:149-154)
                transmit_color.r = top.r * top.a + transmit_color.r * (1.0f - top.a);
                transmit_color.g = top.g * top.a + transmit_color.g * (1.0f - top.a);
                transmit_color.b = top.b * top.a + transmit_color.b * (1.0f - top.a);
                transmit_color.a = 1.0f;
                pixel_color = pixel_color + transmit_color;
            }
        }
        for (const LightDraw& ld : s.light_draws) {
            const rayca_light& light = s.lights[ld.light];
            Vec3 light_dir;
            if (light.kind == RAYCA_LIGHT_POINT) { // light/point.rs:50-54
                Vec3 dist = normalized(to_vec3(hit.hit.point) - trs_get_translation(ld.node_trs));
                light_dir = -dist;
            } else if (light.kind == RAYCA_LIGHT_DIRECTIONAL) { // light/directional.rs:47-51
                light_dir = -rotate(Vec3{1.0f, 0.0f, 0.0f}, ld.node_trs.rotation);
            } else {
                status = RAYCA_E_UNSUPPORTED; // Light::get_direction todo!() for quad
                continue;
            }
            Ray shadow = make_ray(hit.get_next_ray_origin(), light_dir);
            tracer.rays->rays_shadow++;
            Hit sh;
            bool is_light;
            if (!tracer.intersects(shadow, sh)) {
                is_light = true;
            } else {
                float light_distance;
                if (light.kind == RAYCA_LIGHT_POINT)
                    light_distance = len(to_vec3(hit.hit.point - trs_get_translation(ld.node_trs)));
                else
                    light_distance = len(to_vec3(hit.hit.point) - trs_get_translation(ld.node_trs));
                if (sh.depth > light_distance) {
                    is_light = true;
                } else {
                    HitInfo shi(s, sh);
                    is_light = is_transparent(shi.get_color());
                }
            }
            if (is_light) {
                Color intensity = (light.kind == RAYCA_LIGHT_POINT) ? point_intensity(light, ld.node_trs, hit.hit.point)
                                                                    : light.intensity * to_color(light.color);
                Irradiance ir = make_irradiance(intensity, hit, light_dir);
                pixel_color = pixel_color + get_radiance(hit, ir);
            }
        }
        Vec3 reflection = hit.get_reflection();
        Ray reflection_ray = make_ray(hit.get_next_ray_origin(), reflection);
        if (depth + 1 <= cfg.c.max_depth) tracer.rays->rays_secondary++;
        Color reflection_intensity;
        if (scratcher_trace(reflection_ray, depth + 1, reflection_intensity)) {
            Irradiance ir = make_irradiance(reflection_intensity, hit, reflection);
            pixel_color = pixel_color + get_radiance(hit, ir);
        }
        out = pixel_color;
        return true;
    }

    // ---- QuadLight (rayca-model/src/light/quad.rs); `trs` is the light node's LOCAL trs
    static Point3 quad_corner(const rayca_light& l, const Trs& trs, int k) { // get_a (quad.rs:48-57)
        Point3 a = to_point3(trs_get_translation(trs));
        Vec3 ab{l.ab[0], l.ab[1], l.ab[2]}, ac{l.ac[0], l.ac[1], l.ac[2]};
        switch (k & 3) {
            case 0: return a;
            case 1: return a + ab;
            case 2: return (a + ab) + ac;
            default: return a + ac;
        }
    }
    static Vec3 quad_normal(const rayca_light& l) { // quad.rs:36-38
        return normalized(cross(Vec3{l.ab[0], l.ab[1], l.ab[2]}, Vec3{l.ac[0], l.ac[1], l.ac[2]}));
    }
    static float quad_area(const rayca_light& l) { // quad.rs:40-46 (sin_theta = 1 - cos_theta, as written)
        Vec3 ab{l.ab[0], l.ab[1], l.ab[2]}, ac{l.ac[0], l.ac[1], l.ac[2]};
        float ab_len = len(ab), ac_len = len(ac);
        float cos_theta = dot(ab, ac) / (ab_len * ac_len);
        float sin_theta = 1.0f - cos_theta;
        return sin_theta * ab_len * ac_len;
    }
    // get_radiance / theta / gamma / get_intensity (quad.rs:72-104): edge i runs from corner i to corner i+1
    static Color quad_intensity(const rayca_light& l, const Trs& trs, Point3 frag, Vec3 n) {
        Vec3 ret{0, 0, 0};
        for (int e = 0; e < 4; ++e) {
            Point3 pa = quad_corner(l, trs, e), pb = quad_corner(l, trs, e + 1);
            float theta = std::acos(dot(normalized(pa - frag), normalized(pb - frag)));
            Vec3 gamma = normalized(cross(pa - frag, pb - frag));
            ret = ret + theta * gamma;
        }
        ret = ret / 2.0f;
        float irradiance = dot(ret, n);
        return l.intensity * to_color(l.color) * irradiance;
    }
    // Light::get_intensity (light/mod.rs:54-60)
    Color light_intensity(const rayca_light& l, const Trs& trs, Point3 frag, Vec3 n) {
        if (l.kind == RAYCA_LIGHT_DIRECTIONAL) return l.intensity * to_color(l.color);
        if (l.kind == RAYCA_LIGHT_POINT) return point_intensity(l, trs, frag);
        return quad_intensity(l, trs, frag, n);
    }

    // integrator/analyticdirect.rs:16-49
    bool analytic_trace(const Ray& ray, uint32_t depth, Color& out, Hit* primary) {
        if (depth >= cfg.c.max_depth) return false;
        Hit h;
        if (!tracer.intersects(ray, h)) return false;
        if (primary) *primary = h;
        HitInfo hit(s, h);
        Color ambient_and_emission = hit.get_color();
        if (hit.is_emissive()) {
            out = ambient_and_emission;
            return true;
        }
        Vec3 n = hit.get_normal();
        Color light_contribution = BLACK;
        for (const LightDraw& ld : s.light_draws)
            light_contribution = light_contribution + light_intensity(s.lights[ld.light], ld.node_trs, hit.hit.point, n);
        Color f = hit.get_diffuse() * FRAC_1_PI;
        out = f * light_contribution;
        return true;
    }

    // integrator/direct.rs:16-88; QuadLight::get_random_point (quad.rs:117-137) draws from the keyed RNG:
    // counter (pixel, sample, depth 0, 2 * (light * light_samples + i) + {0, 1})
    bool direct_trace(const Ray& ray, uint32_t depth, uint32_t pixel, uint32_t sample, Color& out, Hit* primary) {
        if (depth >= cfg.c.max_depth) return false;
        Hit h;
        if (!tracer.intersects(ray, h)) return false;
        if (primary) *primary = h;
        HitInfo hit(s, h);
        hit.pixel = pixel;
        hit.sample = sample;
        Color ambient_and_emissive = hit.get_color();
        if (hit.is_emissive()) {
            out = ambient_and_emissive;
            return true;
        }
        Color light_contribution = BLACK;
        const uint32_t ls = cfg.c.light_samples;
        const uint32_t strate_count = cfg.c.light_stratify ? (uint32_t)std::sqrt((float)ls) : 1u; // config.rs:73-79
        uint32_t li = 0;
        for (const LightDraw& ld : s.light_draws) {
            const rayca_light& l = s.lights[ld.light];
            const uint32_t light_slot = li++;
            if (l.kind != RAYCA_LIGHT_QUAD) continue;
            Vec3 ab{l.ab[0], l.ab[1], l.ab[2]}, ac{l.ac[0], l.ac[1], l.ac[2]};
            Color ldc = BLACK;
            for (uint32_t i = 0; i < ls; ++i) {
                Vec3 step1 = ab / (float)strate_count, step2 = ac / (float)strate_count;
                float u1 = rng_f32(cfg.seed, pixel, sample, 0, 2 * (light_slot * ls + i)) / (float)strate_count;
                float u2 = rng_f32(cfg.seed, pixel, sample, 0, 2 * (light_slot * ls + i) + 1) / (float)strate_count;
                Point3 x1 = (quad_corner(l, ld.node_trs, 0) + (u1 * ab)) + (u2 * ac);
                if (cfg.c.light_stratify) {
                    float i1 = (float)(i % strate_count), i2 = (float)(i / strate_count);
                    x1 = x1 + ((step1 * i1) + (step2 * i2));
                }
                Vec3 x1_to_hit = x1 - hit.hit.point;
                Vec3 omega_i = normalized(x1_to_hit);
                Ray shadow = make_ray(hit.get_next_ray_origin(), omega_i);
                tracer.rays->rays_shadow++;
                Hit sh;
                if (tracer.intersects(shadow, sh)) {
                    HitInfo shi(s, sh);
                    if (!shi.is_emissive()) continue;
                }
                Color brdf = get_brdf(hit, omega_i);
                float r = len(x1_to_hit);
                float r_squared = r * r; // len().powf(2.0)
                float d_omega_i = dot(quad_normal(l), omega_i) / r_squared;
                ldc = ldc + (brdf * dot(hit.get_normal(), omega_i)) * d_omega_i;
            }
            Color lic = l.intensity * to_color(l.color);
            float area = quad_area(l);
            ldc = ((lic * area) * ldc) / (float)ls;
            light_contribution = light_contribution + ldc;
        }
        out = light_contribution;
        return true;
    }

    // integrator/flat.rs:16-28
    bool flat_trace(const Ray& ray, Color& out, Hit* primary) {
        Hit h;
        if (!tracer.intersects(ray, h)) return false;
        if (primary) *primary = h;
        HitInfo hit(s, h);
        out = hit.get_color();
        return true;
    }

    // scene.rs:80-86 draw_pixel
    Color draw_pixel(const Ray& ray, uint32_t pixel, uint32_t sample, Hit* primary, bool& hit_any) {
        Color c = BLACK;
        bool ok = false;
        switch (cfg.c.integrator) {
            case RAYCA_INTEGRATOR_PATHTRACER: ok = pathtracer_trace(ray, 0, true, pixel, sample, c, primary); break;
            case RAYCA_INTEGRATOR_RAYTRACER: ok = raytracer_trace(ray, 0, c, primary); break;
            case RAYCA_INTEGRATOR_SCRATCHER: ok = scratcher_trace(ray, 0, c, primary); break;
            case RAYCA_INTEGRATOR_FLAT: ok = flat_trace(ray, c, primary); break;
            case RAYCA_INTEGRATOR_ANALYTIC_DIRECT: ok = analytic_trace(ray, 0, c, primary); break;
            case RAYCA_INTEGRATOR_DIRECT: ok = direct_trace(ray, 0, pixel, sample, c, primary); break;
            default: status = RAYCA_E_UNSUPPORTED; break;
        }
        hit_any = ok && primary && primary->blas != ~0u;
        return ok ? c : BLACK;
    }
};

} // namespace

// ---------------------------------------------------------------------- C API

extern "C" {

typedef struct rayca_oracle_stats {
    double build_ms, trace_ms;
    uint64_t rays_primary, rays_shadow, rays_secondary;
    uint64_t ord_nodes, ord_tris; // ordered + culled traversal record counts (algorithmic bytes model)
    uint64_t exh_nodes, exh_tris; // exhaustive (reference) traversal record counts
    uint32_t threads, reserved; // threads = host threads that actually rendered pixels in this call
    double opt_build_ms;        // RAYCA_ORACLE_OPTIMISED: build time of the optimised-arm tree (once per scene)
} rayca_oracle_stats;

// FAITHFUL: re-transform the vertices on every test like the reference; COUNT: tally walk records;
// OPTIMISED: answer closest-hit queries with the ordered, culled walk over the optimised-arm tree (OptTree)
enum { RAYCA_ORACLE_FAITHFUL = 1, RAYCA_ORACLE_COUNT = 2, RAYCA_ORACLE_OPTIMISED = 4 };

int rayca_oracle_scene_create(const rayca_scene_desc* desc, uint32_t bvh, rayca_oracle_scene** out) {
    if (!desc || !out) return RAYCA_E_INVALID_ARG;
    auto t0 = std::chrono::steady_clock::now();
    Scene* s = new Scene();
    const rayca_scene_desc& d = *desc;
    s->models.assign(d.models, d.models + d.model_count);
    s->materials.assign(d.materials, d.materials + d.material_count);
    s->pbr.assign(d.pbr_materials, d.pbr_materials + d.pbr_count);
    s->phong.assign(d.phong_materials, d.phong_materials + d.phong_count);
    s->ggx.assign(d.ggx_materials, d.ggx_materials + d.ggx_count);
    s->textures.assign(d.textures, d.textures + d.texture_count);
    s->images.assign(d.images, d.images + d.image_count);
    s->lights.assign(d.lights, d.lights + d.light_count);
    s->cameras.assign(d.cameras, d.cameras + d.camera_count);
    s->image_data.resize(d.image_count);
    for (uint32_t i = 0; i < d.image_count; ++i) {
        const rayca_image& im = d.images[i];
        size_t bpp = im.color_type == RAYCA_COLOR_RGB8 ? 3 : (im.color_type == RAYCA_COLOR_RGBA8 ? 4 : 16);
        size_t bytes = (size_t)im.width * im.height * bpp;
        s->image_data[i].assign((const uint8_t*)im.data, (const uint8_t*)im.data + bytes);
        s->images[i].data = nullptr;
    }
    Flattener f{*s, d, {}};
    f.run();
    // BvhScene::from_scene (primitive.rs:385-394) in ascending model order (A-19 deviation)
    for (uint32_t m = 0; m < d.model_count; ++m) {
        if (!f.mdi[m].used) continue;
        Blas blas;
        blas.model = m;
        blas.max_depth = bvh ? 255 : 0; // scene.rs:95-99
        for (size_t k = 0; k < f.mdi[m].mesh_nodes.size(); ++k)
            emit_mesh_prims(*s, d, m, f.mdi[m].mesh_node_local[k], f.mdi[m].mesh_nodes[k], blas.prims);
        for (size_t k = 0; k < f.mdi[m].quad_nodes.size(); ++k) // primitive.rs:363-367: area lights after the meshes
            emit_quad_prims(*s, d, m, f.mdi[m].quad_lights[k], f.mdi[m].quad_nodes[k], blas.prims);
        for (const Prim& p : blas.prims) blas.has_spheres = blas.has_spheres || p.sphere;
        build_blas(blas);
        s->blas_nodes.push_back({(uint32_t)s->blass.size()});
        s->blass.push_back(std::move(blas));
    }
    if (s->blass.empty()) { // replace_models_recursive asserts a non-empty range (tlas.rs:75)
        delete s;
        return RAYCA_E_INVALID_ARG;
    }
    replace_models_recursive(*s, s->root, 0, (uint32_t)s->blas_nodes.size());
    s->build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    *out = s;
    return RAYCA_OK;
}

void rayca_oracle_scene_destroy(rayca_oracle_scene* s) { delete s; }

int rayca_oracle_scene_bvh_info(rayca_oracle_scene* s, rayca_bvh_info* info) {
    if (!s || !info) return RAYCA_E_INVALID_ARG;
    info->blas_count = (uint32_t)s->blas_nodes.size();
    info->node_count = 0;
    info->triangle_count = 0;
    for (const BlasRef& r : s->blas_nodes) {
        info->node_count += (uint32_t)s->blass[r.blas].nodes.size();
        info->triangle_count += (uint32_t)s->blass[r.blas].prims.size();
    }
    info->tlas_node_count = (uint32_t)s->tlas_nodes.size() + 1;
    return RAYCA_OK;
}

// Same layout as rayca_cuda_scene_bvh_read; BLAS in TLAS-reordered (blas_nodes) order.
int rayca_oracle_scene_bvh_read(rayca_oracle_scene* s, float* nodes, uint32_t* blas_node_offset, uint32_t* blas_prim_offset,
                                uint32_t* prim_src) {
    if (!s) return RAYCA_E_INVALID_ARG;
    uint32_t no = 0, po = 0, bi = 0;
    for (const BlasRef& r : s->blas_nodes) {
        const Blas& b = s->blass[r.blas];
        if (blas_node_offset) blas_node_offset[bi] = no;
        if (blas_prim_offset) blas_prim_offset[bi] = po;
        for (const BvhNode& n : b.nodes) {
            if (nodes) {
                float* o = nodes + (size_t)no * 8;
                o[0] = n.bounds.a.x; o[1] = n.bounds.a.y; o[2] = n.bounds.a.z;
                std::memcpy(o + 3, &n.offset, 4);
                o[4] = n.bounds.b.x; o[5] = n.bounds.b.y; o[6] = n.bounds.b.z;
                std::memcpy(o + 7, &n.count, 4);
            }
            no++;
        }
        for (const Prim& p : b.prims) {
            if (prim_src) prim_src[po] = p.src;
            po++;
        }
        bi++;
    }
    if (blas_node_offset) blas_node_offset[bi] = no;
    if (blas_prim_offset) blas_prim_offset[bi] = po;
    return RAYCA_OK;
}

// The pixel loop (scene.rs:101-152).  `rows` (nullable) = [row_begin, row_end) for bounded
// CPU-baseline samples; pixels outside are left untouched.
int rayca_oracle_render(rayca_oracle_scene* s, const rayca_config* config, const rayca_options* options, uint32_t width,
                        uint32_t height, uint8_t* rgba8_out, const rayca_aux_out* aux, uint32_t flags, int threads,
                        const uint32_t* rows, rayca_oracle_stats* stats) {
    if (!s || !config || width == 0 || height == 0) return RAYCA_E_INVALID_ARG;
    if (!s->has_camera) return RAYCA_E_NO_CAMERA;
    if (config->samples_per_pixel == 0) return RAYCA_E_INVALID_ARG;
    Config cfg{*config, options ? options->rng_seed : 0};
    if (cfg.c.integrator == RAYCA_INTEGRATOR_PATHTRACER) {
        if (cfg.c.direct_sampler != RAYCA_SAMPLER_NEE && cfg.c.direct_sampler != RAYCA_SAMPLER_NONE &&
            cfg.c.direct_sampler != RAYCA_SAMPLER_MIS)
            return RAYCA_E_UNSUPPORTED;
        bool needs_indirect = cfg.c.russian_roulette || cfg.c.max_depth > (cfg.c.direct_sampler != RAYCA_SAMPLER_NONE ? 1u : 0u);
        if (needs_indirect && cfg.c.indirect_sampler != RAYCA_SAMPLER_COSINE && cfg.c.indirect_sampler != RAYCA_SAMPLER_HEMISPHERE &&
            cfg.c.indirect_sampler != RAYCA_SAMPLER_BRDF)
            return RAYCA_E_UNSUPPORTED;
    } else if (cfg.c.integrator == RAYCA_INTEGRATOR_DIRECT) {
        if (cfg.c.light_samples == 0) return RAYCA_E_INVALID_ARG;
    } else if (cfg.c.integrator != RAYCA_INTEGRATOR_RAYTRACER && cfg.c.integrator != RAYCA_INTEGRATOR_SCRATCHER &&
               cfg.c.integrator != RAYCA_INTEGRATOR_FLAT &&
               cfg.c.integrator != RAYCA_INTEGRATOR_ANALYTIC_DIRECT) {
        return RAYCA_E_UNSUPPORTED;
    }
    uint32_t row_begin = rows ? rows[0] : 0, row_end = rows ? std::min(rows[1], height) : height;

    const float fw = (float)width, fh = (float)height;
    const float inv_width = 1.0f / fw, inv_height = 1.0f / fh;
    const float aspectratio = fw / fh;
    const float angle = std::tan(s->camera_yfov * 0.5f); // camera.rs:74-76
    const uint32_t spp = cfg.c.samples_per_pixel;
    const float strate_count = std::sqrt((float)spp);
    const float offset = 0.5f / strate_count;
    const float step = 1.0f / strate_count;
    const uint32_t strate_u = (uint32_t)strate_count;

#ifdef _OPENMP
    int nthreads = threads > 0 ? threads : omp_get_max_threads();
#else
    int nthreads = 1;
#endif
    const OptTree* opt = nullptr;
    if (flags & RAYCA_ORACLE_OPTIMISED) {
        if (!build_opt_tree(*s)) return RAYCA_E_UNSUPPORTED;
        opt = s->opt;
    }
    std::vector<Counters> counters((size_t)nthreads);
    std::vector<uint32_t> worked((size_t)nthreads, 0u);
    int status = 0;
    auto t0 = std::chrono::steady_clock::now();
    // The reference fans out over rows AND pixels (`par_iter_mut` on both, rayca-soft/src/scene.rs:117-122): one task
    // per pixel.  Here: one flat pixel index, handed out in chunks of 16 pixels.
    const int64_t px_begin = (int64_t)row_begin * width, px_end = (int64_t)row_end * width;
#pragma omp parallel num_threads(nthreads)
    {
#ifdef _OPENMP
        const size_t tid = (size_t)omp_get_thread_num();
#else
        const size_t tid = 0;
#endif
        Counters* cnt = &counters[tid];
        Renderer r{*s, cfg, Tracer{*s, (flags & RAYCA_ORACLE_FAITHFUL) != 0, cnt, (flags & RAYCA_ORACLE_COUNT) ? cnt : nullptr, opt}};
#pragma omp for schedule(dynamic, 16) nowait
        for (int64_t p_ = px_begin; p_ < px_end; ++p_) {
            worked[tid] = 1;
            const uint32_t y = (uint32_t)(p_ / width), x = (uint32_t)(p_ % width);
            Color color = BLACK;
            Hit primary;
            bool hit_any = false;
            uint32_t pixel = y * width + x;
            for (uint32_t i = 0; i < spp; ++i) {
                float ix = (float)(i % strate_u);
                float iy = (float)(i / strate_u);
                float xx = (2.0f * (((float)x + ix * step + offset) * inv_width) - 1.0f) * angle * aspectratio;
                float yy = (1.0f - 2.0f * (((float)y + iy * step + offset) * inv_height)) * angle;
                Vec3 dir = normalized(Vec3{xx, yy, -1.0f});
                Ray ray = transform_ray(s->camera_trs, make_ray(Point3{0, 0, 0}, dir));
                cnt->rays_primary++;
                primary = Hit{};
                Color c = r.draw_pixel(ray, pixel, i, &primary, hit_any);
                color = color + c;
            }
            color = color / (float)spp;   // DivAssign (color/mod.rs:448-454)
            float g = 1.0f / cfg.c.gamma; // correct_gamma (color/mod.rs:175-180)
            color.r = std::pow(color.r, g);
            color.g = std::pow(color.g, g);
            color.b = std::pow(color.b, g);
            size_t p = (size_t)y * width + x;
            if (rgba8_out) {
                rgba8_out[p * 4 + 0] = quantize(color.r);
                rgba8_out[p * 4 + 1] = quantize(color.g);
                rgba8_out[p * 4 + 2] = quantize(color.b);
                rgba8_out[p * 4 + 3] = quantize(color.a);
            }
            if (aux) {
                if (aux->hit_blas) aux->hit_blas[p] = hit_any ? primary.blas : ~0u;
                if (aux->hit_prim) aux->hit_prim[p] = hit_any ? primary.primitive : ~0u;
                if (aux->hit_src)
                    aux->hit_src[p] = hit_any ? s->blass[s->blas_nodes[primary.blas].blas].prims[primary.primitive].src : ~0u;
                if (aux->depth) aux->depth[p] = hit_any ? primary.depth : INFINITY;
                if (aux->uv) {
                    aux->uv[p * 2] = hit_any ? primary.uv.x : 0.0f;
                    aux->uv[p * 2 + 1] = hit_any ? primary.uv.y : 0.0f;
                }
                if (aux->rgba_f32) {
                    aux->rgba_f32[p * 4 + 0] = color.r;
                    aux->rgba_f32[p * 4 + 1] = color.g;
                    aux->rgba_f32[p * 4 + 2] = color.b;
                    aux->rgba_f32[p * 4 + 3] = color.a;
                }
            }
        }
        if (r.status != 0) {
#pragma omp critical
            status = r.status;
        }
    }
    double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        stats->build_ms = s->build_ms;
        stats->trace_ms = ms;
        stats->threads = 0; // threads that actually rendered pixels (not the size of the pool)
        for (uint32_t w : worked) stats->threads += w;
        stats->opt_build_ms = opt ? opt->build_ms : 0.0;
        for (const Counters& c : counters) {
            stats->rays_primary += c.rays_primary;
            stats->rays_shadow += c.rays_shadow;
            stats->rays_secondary += c.rays_secondary;
            stats->ord_nodes += c.ord_nodes;
            stats->ord_tris += c.ord_tris;
            stats->exh_nodes += c.exh_nodes;
            stats->exh_tris += c.exh_tris;
        }
    }
    return status;
}

// ------------------------------------------------------------- KAT entry points
// Small wrappers so tests can pin the unit-level known answers of the reference's own tests.

void rayca_oracle_kat_reciprocal(const float in[3], float out[3]) {
    Vec3 r = reciprocal(Vec3{in[0], in[1], in[2]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void rayca_oracle_kat_rotate(const float v[3], const float q[4], float out[3]) {
    Vec3 r = rotate(Vec3{v[0], v[1], v[2]}, Quat{q[0], q[1], q[2], q[3]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void rayca_oracle_kat_reflect(const float v[3], const float n[3], float out[3]) {
    Vec3 r = reflect(Vec3{v[0], v[1], v[2]}, Vec3{n[0], n[1], n[2]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void rayca_oracle_kat_trs_mul(const rayca_trs* a, const rayca_trs* b, rayca_trs* out) {
    Trs r = to_trs(*a) * to_trs(*b);
    out->translation[0] = r.translation.x; out->translation[1] = r.translation.y; out->translation[2] = r.translation.z;
    out->rotation[0] = r.rotation.x; out->rotation[1] = r.rotation.y; out->rotation[2] = r.rotation.z; out->rotation[3] = r.rotation.w;
    out->scale[0] = r.scale.x; out->scale[1] = r.scale.y; out->scale[2] = r.scale.z;
}
void rayca_oracle_kat_trs_point(const rayca_trs* t, const float p[3], float out[3]) {
    Point3 r = transform_point(to_trs(*t), Point3{p[0], p[1], p[2]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void rayca_oracle_kat_trs_ray(const rayca_trs* t, const float o[3], const float d[3], float out_o[3], float out_d[3]) {
    Ray r = transform_ray(to_trs(*t), make_ray(Point3{o[0], o[1], o[2]}, Vec3{d[0], d[1], d[2]}));
    out_o[0] = r.origin.x; out_o[1] = r.origin.y; out_o[2] = r.origin.z;
    out_d[0] = r.dir.x; out_d[1] = r.dir.y; out_d[2] = r.dir.z;
}
uint8_t rayca_oracle_kat_quantize(float c) { return quantize(c); }
// returns 1 on hit; out = {t, u, v, px, py, pz}
int rayca_oracle_kat_triangle(const float verts[9], const rayca_trs* trs, const float o[3], const float d[3], float out[6]) {
    Trs t = to_trs(*trs);
    Ray ray = make_ray(Point3{o[0], o[1], o[2]}, Vec3{d[0], d[1], d[2]});
    float tt;
    Point3 p;
    Vec2 uv;
    bool hit = triangle_intersects_world(transform_point(t, Point3{verts[0], verts[1], verts[2]}),
                                         transform_point(t, Point3{verts[3], verts[4], verts[5]}),
                                         transform_point(t, Point3{verts[6], verts[7], verts[8]}), ray, tt, p, uv);
    if (hit) {
        out[0] = tt; out[1] = uv.x; out[2] = uv.y; out[3] = p.x; out[4] = p.y; out[5] = p.z;
    }
    return hit ? 1 : 0;
}
// returns 1 on hit; out = {t, px, py, pz} (Sphere::intersects, rayca-geometry/src/sphere.rs:101-163)
int rayca_oracle_kat_sphere(const float center[3], float radius, const rayca_trs* trs, const float o[3], const float d[3], float out[4]) {
    Scene s;
    s.world_trs.push_back(to_trs(*trs));
    Prim p{};
    p.sphere = true;
    p.center = {center[0], center[1], center[2]};
    p.radius = radius;
    p.radius2 = radius * radius;
    p.trs = 0;
    Counters c;
    Tracer tr{s, false, &c, nullptr};
    float t;
    Point3 pt;
    Vec2 uv;
    bool hit = tr.sphere_intersects(p, make_ray(Point3{o[0], o[1], o[2]}, Vec3{d[0], d[1], d[2]}), t, pt, uv);
    if (hit) {
        out[0] = t; out[1] = pt.x; out[2] = pt.y; out[3] = pt.z;
    }
    return hit ? 1 : 0;
}
// out = {min.xyz, max.xyz, centroid.xyz}
void rayca_oracle_kat_sphere_bounds(const float center[3], float radius, const rayca_trs* trs, float out[9]) {
    Prim p{};
    p.sphere = true;
    p.center = {center[0], center[1], center[2]};
    p.radius = radius;
    sphere_world_bounds(p, to_trs(*trs));
    out[0] = p.wmin.x; out[1] = p.wmin.y; out[2] = p.wmin.z; out[3] = p.wmax.x; out[4] = p.wmax.y; out[5] = p.wmax.z;
    out[6] = p.wcentroid.x; out[7] = p.wcentroid.y; out[8] = p.wcentroid.z;
}
// Sphere::get_model_normal (sphere.rs:95-97) as the shading path evaluates it
void rayca_oracle_kat_sphere_normal(const float center[3], const float point[3], float out[3]) {
    Vec3 n = normalized(Point3{point[0], point[1], point[2]} - Point3{center[0], center[1], center[2]});
    out[0] = n.x; out[1] = n.y; out[2] = n.z;
}
// out = {min.xyz, max.xyz, centroid.xyz} of a triangle under `trs`
void rayca_oracle_kat_triangle_bounds(const float verts[9], const rayca_trs* trs, float out[9]) {
    Prim p{};
    for (int k = 0; k < 3; ++k) p.v[k] = {verts[3 * k], verts[3 * k + 1], verts[3 * k + 2]};
    p.centroid = ((to_vec3(p.v[0]) + to_vec3(p.v[1])) + to_vec3(p.v[2])) * 0.3333f; // Triangle::new (triangle.rs:59-62)
    triangle_world_bounds(p, to_trs(*trs));
    out[0] = p.wmin.x; out[1] = p.wmin.y; out[2] = p.wmin.z; out[3] = p.wmax.x; out[4] = p.wmax.y; out[5] = p.wmax.z;
    out[6] = p.wcentroid.x; out[7] = p.wcentroid.y; out[8] = p.wcentroid.z;
}
// &Inversed<Trs> * Vec3 (rayca-math/src/trs.rs:383-403): translate by -t, rotate by q^-1, scale by 1/s — the operation
// sequence inverse_transform_ray applies to a ray's origin and direction
void rayca_oracle_kat_trs_inverse_vec(const rayca_trs* t, const float v[3], float out[3]) {
    Trs trs = to_trs(*t);
    Vec3 r = Vec3{v[0], v[1], v[2]} + (trs.translation * -1.0f);
    r = rotate(r, conjugate(trs.rotation));
    r = r * reciprocal(trs.scale);
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void rayca_oracle_kat_trs_vec(const rayca_trs* t, const float v[3], float out[3]) {
    Vec3 r = transform_vec(to_trs(*t), Vec3{v[0], v[1], v[2]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
float rayca_oracle_kat_aabb(const float a[3], const float b[3], const float o[3], const float d[3]) {
    AABB box{{a[0], a[1], a[2]}, {b[0], b[1], b[2]}};
    return aabb_intersects(box, make_ray(Point3{o[0], o[1], o[2]}, Vec3{d[0], d[1], d[2]}));
}
// texel index of Sampler::sample
void rayca_oracle_kat_sampler_index(const float uv[2], uint32_t w, uint32_t h, uint32_t out[2]) {
    float x = (uv[0] - std::floor(uv[0]) + 1.0f) * (float)w;
    float y = (uv[1] - std::floor(uv[1]) + 1.0f) * (float)h;
    out[0] = ((uint32_t)x) % w;
    out[1] = ((uint32_t)y) % h;
}
float rayca_oracle_kat_rng(uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t depth, uint32_t dim) {
    return rng_f32(seed, pixel, sample, depth, dim);
}
int rayca_oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

} // extern "C"
