// host_scene.cpp — scene-graph flattening, primitive emission, TLAS and HBM packing.
// Reference behaviour restated (file:line relative to the reference checkout):
//   SceneDrawInfo::new + traverse_*            rayca-soft/src/scene.rs:190-282
//   BvhPrimitive::from_triangle_mesh_impl      rayca-soft/src/bvh/primitive.rs:194-237
//   BvhModel::from_model / BvhScene::from_scene primitive.rs:355-394
//   Tlas::new + replace_models_recursive       rayca-soft/src/bvh/tlas.rs:74-134,248-269
// Layout decisions that differ from the reference are described in DESIGN.md.
#include "host_scene.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>

#include <algorithm>
#include <cstring>
#include <functional>
#include <mutex>
#include <unistd.h>

namespace rb {
namespace {

TrsF to_trs(const rayca_trs& t) {
    TrsF r;
    r.t = V3{t.translation[0], t.translation[1], t.translation[2]};
    r.q = Q4{t.rotation[0], t.rotation[1], t.rotation[2], t.rotation[3]};
    r.s = V3{t.scale[0], t.scale[1], t.scale[2]};
    return r;
}

// ---- 3x3 helpers (rayca-math/src/mat3.rs), row-major m[i][j]; struct M3 lives in host_scene.h
M3 m3_identity() { return M3{{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}}}; }
// mat3.rs:190-216
M3 m3_from_quat(Q4 q) {
    float xx = q.x * q.x, xy = q.x * q.y, xz = q.x * q.z, xw = q.x * q.w;
    float yy = q.y * q.y, yz = q.y * q.z, yw = q.y * q.w;
    float zz = q.z * q.z, zw = q.z * q.w;
    M3 r;
    r.m[0][0] = 1.0f - 2.0f * (yy + zz); r.m[0][1] = 2.0f * (xy - zw);        r.m[0][2] = 2.0f * (xz + yw);
    r.m[1][0] = 2.0f * (xy + zw);        r.m[1][1] = 1.0f - 2.0f * (xx + zz); r.m[1][2] = 2.0f * (yz - xw);
    r.m[2][0] = 2.0f * (xz - yw);        r.m[2][1] = 2.0f * (yz + xw);        r.m[2][2] = 1.0f - 2.0f * (xx + yy);
    return r;
}
// mat3.rs:160-180: e + f + g, left to right
M3 m3_mul(const M3& a, const M3& b) {
    M3 r;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            float e = a.m[i][0] * b.m[0][j], f = a.m[i][1] * b.m[1][j], g = a.m[i][2] * b.m[2][j];
            r.m[i][j] = e + f + g;
        }
    return r;
}
M3 m3_scale(V3 s) {
    M3 r = m3_identity();
    r.m[0][0] *= s.x; r.m[1][1] *= s.y; r.m[2][2] *= s.z;
    return r;
}
M3 m3_transpose(const M3& a) {
    M3 r;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) r.m[i][j] = a.m[j][i];
    return r;
}
// mat3.rs:218-234: ret[i] += m[i][j] * v[j] starting from 0.0
V3 m3_apply(const M3& a, V3 v) {
    float in[3] = {v.x, v.y, v.z}, out[3] = {0.0f, 0.0f, 0.0f};
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) out[i] += a.m[i][j] * in[j];
    return V3{out[0], out[1], out[2]};
}
// Mat3::from(&Trs) (mat3.rs:126-132): from_scale then rotate => Mat3::from(q) * S
M3 tangent_matrix_of(const TrsF& t) { return m3_mul(m3_from_quat(t.q), m3_scale(t.s)); }
// Mat3::from(&Inversed<&Trs>).get_transpose() (mat3.rs:140-144, trs.rs:320-331)
M3 normal_matrix_of(const TrsF& t) {
    M3 rot = m3_mul(m3_from_quat(conj(t.q)), m3_identity()); // Mat3::from_rotation
    return m3_transpose(m3_mul(m3_scale(reciprocal(t.s)), rot));
}

// ---- Mat4 (rayca-math/src/mat4.rs) for the sphere path: Mat4::from(&Inversed<Trs>) =
// from_scale(1/s) * (from_rotation(q^-1) * from_translation(-t)) (mat4.rs:276-283).  The product sums
// row * column over 16 SIMD lanes of which 12 are masked to zero, in lane order (mat4.rs:160-189).
struct M4 {
    float m[4][4];
};
M4 m4_identity() {
    M4 r{};
    for (int i = 0; i < 4; ++i) r.m[i][i] = 1.0f;
    return r;
}
M4 m4_mul(const M4& a, const M4& b) {
    M4 r;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            float acc = ((a.m[i][0] * b.m[0][j] + a.m[i][1] * b.m[1][j]) + a.m[i][2] * b.m[2][j]) + a.m[i][3] * b.m[3][j];
            for (int z = 0; z < 12; ++z) acc += 0.0f;
            r.m[i][j] = acc;
        }
    return r;
}
M4 m4_from_inversed(const TrsF& t) {
    M4 sc = m4_identity();
    V3 rs = reciprocal(t.s);
    sc.m[0][0] *= rs.x; sc.m[1][1] *= rs.y; sc.m[2][2] *= rs.z;
    M3 r3 = m3_from_quat(conj(t.q));
    M4 rq = m4_identity();
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) rq.m[i][j] = r3.m[i][j];
    M4 rot = m4_mul(rq, m4_identity()); // Mat4::rotate: Mat4::from(q) * self (mat4.rs:110-112)
    M4 tr = m4_identity();
    tr.m[0][3] += -t.t.x; tr.m[1][3] += -t.t.y; tr.m[2][3] += -t.t.z;
    return m4_mul(sc, m4_mul(rot, tr));
}

struct MeshDraw {
    uint32_t node_local;
    TrsF world;
};
struct QuadDraw { // a quad-light node: becomes two emissive triangles of its model (scene.rs:265-272)
    uint32_t light; // global light index
    TrsF world;
};

struct Flatten {
    const rayca_scene_desc& d;
    HostScene& out;
    std::vector<std::vector<MeshDraw>> mesh_draws; // per model, traversal order
    std::vector<std::vector<QuadDraw>> quad_draws;
    std::string err;
    int status = RAYCA_OK;

    explicit Flatten(const rayca_scene_desc& desc, HostScene& o) : d(desc), out(o), mesh_draws(desc.model_count), quad_draws(desc.model_count) {}

    void model_nodes(const TrsF& parent, uint32_t node, uint32_t mi, int depth) {
        const rayca_model& m = d.models[mi];
        if (node >= m.node_count || depth > 4096) { status = RAYCA_E_INVALID_ARG; err = "bad node index"; return; }
        const rayca_node& n = d.nodes[m.node_offset + node];
        TrsF cur = trs_mul(parent, to_trs(n.trs));
        if (n.mesh >= 0) mesh_draws[mi].push_back(MeshDraw{node, cur});
        if (n.light >= 0) {
            const rayca_light& l = d.lights[m.light_offset + (uint32_t)n.light];
            GpuLight g{};
            std::memcpy(g.color, l.color, sizeof(g.color));
            g.intensity = l.intensity;
            std::memcpy(g.att, l.attenuation, sizeof(g.att));
            g.kind = l.kind;
            TrsF local = to_trs(n.trs); // lights read the node's LOCAL trs (nee.rs:133-134, raytracer.rs:37-39)
            V3 p = rotate(local.t, local.q); // Trs::get_translation (trs.rs:126-128)
            g.pos[0] = p.x; g.pos[1] = p.y; g.pos[2] = p.z;
            V3 dir = -rotate(V3{1.0f, 0.0f, 0.0f}, local.q);
            g.dir[0] = dir.x; g.dir[1] = dir.y; g.dir[2] = dir.z;
            if (l.kind == RAYCA_LIGHT_QUAD) {
                V3 ab{l.ab[0], l.ab[1], l.ab[2]}, ac{l.ac[0], l.ac[1], l.ac[2]};
                V3 nq = normalize(cross(ab, ac)); // QuadLight::get_normal (quad.rs:36-38)
                float ab_len = length(ab), ac_len = length(ac);
                float cos_theta = dot(ab, ac) / (ab_len * ac_len); // get_area (quad.rs:40-46), sin = 1 - cos as written
                float sin_theta = 1.0f - cos_theta;
                std::memcpy(g.ab, l.ab, sizeof(g.ab));
                std::memcpy(g.ac, l.ac, sizeof(g.ac));
                g.normal[0] = nq.x; g.normal[1] = nq.y; g.normal[2] = nq.z;
                g.area = sin_theta * ab_len * ac_len;
                {   // get_wide_triangles (quad.rs:147-158; `ab + ab` for corner c is the reference's), then `trs * vertex`
                    const float small_bias = 1e-2f;
                    V3 wide[4] = {-normalize(ab + ac) * small_bias, ab + normalize(ab - ac) * small_bias,
                                  (ab + ac) + normalize(ab + ab) * small_bias, ac + normalize(ac - ab) * small_bias};
                    for (int k = 0; k < 4; ++k) {
                        V3 w = trs_point(local, wide[k]);
                        g.wide[k][0] = w.x; g.wide[k][1] = w.y; g.wide[k][2] = w.z; g.wide[k][3] = 0.0f;
                    }
                }
                quad_draws[mi].push_back(QuadDraw{m.light_offset + (uint32_t)n.light, cur});
            }
            out.lights.push_back(g);
        }
        if (n.camera >= 0 && !out.has_camera) { // camera_draw_infos[0] (scene.rs:109-112)
            out.has_camera = true;
            GpuCamera& c = out.camera;
            c.t[0] = cur.t.x; c.t[1] = cur.t.y; c.t[2] = cur.t.z;
            c.q[0] = cur.q.x; c.q[1] = cur.q.y; c.q[2] = cur.q.z; c.q[3] = cur.q.w;
            c.s[0] = cur.s.x; c.s[1] = cur.s.y; c.s[2] = cur.s.z;
            c.angle = tanf(d.cameras[m.camera_offset + (uint32_t)n.camera].yfov_radians * 0.5f);
        }
        for (uint32_t i = 0; i < n.child_count; ++i) model_nodes(cur, d.children[n.first_child + i], mi, depth + 1);
    }
    void scene_nodes(const TrsF& parent, uint32_t node, int depth) {
        if (node >= d.scene_node_count || depth > 4096) { status = RAYCA_E_INVALID_ARG; err = "bad scene node"; return; }
        const rayca_node& n = d.scene_nodes[node];
        TrsF cur = trs_mul(parent, to_trs(n.trs));
        if (n.model >= 0) {
            if ((uint32_t)n.model >= d.model_count) { status = RAYCA_E_INVALID_ARG; err = "bad model index"; return; }
            const rayca_model& m = d.models[n.model];
            TrsF mt = trs_mul(cur, to_trs(m.root_trs));
            for (uint32_t i = 0; i < m.root_child_count; ++i)
                model_nodes(mt, d.children[m.root_first_child + i], (uint32_t)n.model, depth + 1);
        }
        for (uint32_t i = 0; i < n.child_count; ++i) scene_nodes(cur, d.children[n.first_child + i], depth + 1);
    }
    void run() {
        TrsF root = to_trs(d.root_trs);
        for (uint32_t i = 0; i < d.root_child_count; ++i) scene_nodes(root, d.children[d.root_first_child + i], 0);
    }
};

// Resolve every (model, material handle) to one flat GPU record; index 0 is the fallback
// Material::DEFAULT = Pbr(NONE) -> PbrMaterial::WHITE (primitive.rs:103-110, pbr.rs:73-80).
GpuMaterial white_pbr() {
    GpuMaterial g{};
    g.a[0] = g.a[1] = g.a[2] = g.a[3] = 1.0f;
    g.p0 = 0.0f;
    g.p1 = 1.0f;
    g.tex_albedo = g.tex_normal = g.tex_mr = -1;
    g.kind = RAYCA_MATERIAL_PBR;
    return g;
}
bool color_close_black(const float c[4]) { // Color::close (color/mod.rs:160-169) against BLACK
    return fabsf(c[0] - 0.0f) < kF32Eps && fabsf(c[1] - 0.0f) < kF32Eps && fabsf(c[2] - 0.0f) < kF32Eps &&
           fabsf(c[3] - 1.0f) < kF32Eps;
}

} // namespace

// ---- block cache of the PodVector allocator (host_scene.h)
namespace {
struct PodBlock {
    void* p;
    size_t bytes;
    bool pinned;
};
std::mutex g_pod_mutex;
std::vector<PodBlock> g_pod_blocks; // cached (free) blocks
std::vector<PodBlock> g_pod_lent;   // blocks handed out again: their real size and whether they are page-locked
size_t g_pod_bytes = 0;
bool g_pod_pin = false;
constexpr size_t kPodMaxBlocks = 32;
// Cache size: a quarter of the physical memory, at most 6 GiB (a 10 M-triangle scene recycles ~3 GiB of host vectors per
// draw; with the old 1 GiB cap they were mmap'ed, page-faulted and unmapped on every call: 0.6 s of a 1.3 s draw);
// RAYCA_HOST_CACHE_MB overrides.
size_t pod_max_bytes() {
    static const size_t cap = [] {
        if (const char* e = std::getenv("RAYCA_HOST_CACHE_MB")) return (size_t)std::max(0L, std::atol(e)) << 20;
        const long pages = sysconf(_SC_PHYS_PAGES), page = sysconf(_SC_PAGE_SIZE);
        const size_t phys = pages > 0 && page > 0 ? (size_t)pages * (size_t)page : (size_t(4) << 30);
        return std::min(phys / 4, size_t(6) << 30);
    }();
    return cap;
}
} // namespace

// Called once a CUDA context exists (rayca_cuda_init): from then on a block that enters the cache is page-locked
// (cudaHostRegister, once per block), so the scene arrays of the NEXT draw are packed straight into pinned memory and
// their uploads run at pinned speed, asynchronously (pageable uploads were the largest phase of a draw: 5.7 of 21 ms).
void pod_cache_enable_pinning() {
    std::lock_guard<std::mutex> lock(g_pod_mutex);
    g_pod_pin = true;
}

void* pod_block_acquire(size_t bytes) {
    std::lock_guard<std::mutex> lock(g_pod_mutex);
    size_t best = g_pod_blocks.size();
    for (size_t i = 0; i < g_pod_blocks.size(); ++i) // the smallest block that fits without wasting more than a quarter
        if (g_pod_blocks[i].bytes >= bytes && g_pod_blocks[i].bytes - bytes <= bytes / 4 &&
            (best == g_pod_blocks.size() || g_pod_blocks[i].bytes < g_pod_blocks[best].bytes))
            best = i;
    if (best == g_pod_blocks.size()) return nullptr;
    const PodBlock blk = g_pod_blocks[best];
    g_pod_bytes -= blk.bytes;
    g_pod_blocks.erase(g_pod_blocks.begin() + (long)best);
    g_pod_lent.push_back(blk); // its owner will give it back with its own (possibly smaller) size
    return blk.p;
}

bool pod_block_release(void* p, size_t bytes) {
    std::lock_guard<std::mutex> lock(g_pod_mutex);
    PodBlock blk{p, bytes, false};
    for (size_t i = 0; i < g_pod_lent.size(); ++i)
        if (g_pod_lent[i].p == p) {
            blk = g_pod_lent[i];
            g_pod_lent.erase(g_pod_lent.begin() + (long)i);
            break;
        }
    if (g_pod_blocks.size() >= kPodMaxBlocks || g_pod_bytes + blk.bytes > pod_max_bytes()) {
        if (blk.pinned) cudaHostUnregister(blk.p); // the caller frees it
        return false;
    }
    if (g_pod_pin && !blk.pinned) {
        blk.pinned = cudaHostRegister(blk.p, blk.bytes, cudaHostRegisterPortable) == cudaSuccess;
        if (!blk.pinned) cudaGetLastError();
    }
    g_pod_blocks.push_back(blk);
    g_pod_bytes += blk.bytes;
    return true;
}

namespace {
struct PhaseTimer { // RAYCA_TIMING=1 prints the host phases of scene_create to stderr
    bool on = std::getenv("RAYCA_TIMING") != nullptr;
    std::chrono::steady_clock::time_point t = std::chrono::steady_clock::now();
    void lap(const char* what) {
        if (!on) return;
        auto n = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[rayca] %-28s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(n - t).count());
        t = n;
    }
};
} // namespace

// the mesh primitive that emitted triangle `src` of the BLAS (emit order), or nullptr for a sphere / quad-light triangle
const HostMeshJob* mesh_job_of(const HostBlas& b, uint32_t src) {
    size_t lo = 0, hi = b.jobs.size();
    while (lo < hi) { // first job whose base exceeds src
        const size_t mid = (lo + hi) / 2;
        if (b.jobs[mid].base <= src) lo = mid + 1;
        else hi = mid;
    }
    if (lo == 0) return nullptr;
    const HostMeshJob& j = b.jobs[lo - 1];
    return src < j.base + j.ntri ? &j : nullptr;
}
namespace {
inline uint32_t material_of(const HostBlas& b, uint32_t src) {
    if (const HostMeshJob* j = mesh_job_of(b, src)) return j->material;
    return b.attrs[src].material;
}
} // namespace

void pack_tri_attrs(const rayca_scene_desc& d, HostScene& out) {
    const size_t total = out.blas_prim_offset.back();
    out.tri_attr.resize(total * 12);
    pack_tri_attrs_range(d, out, 0, total, out.tri_attr.data());
}

void pack_tri_attrs_range(const rayca_scene_desc& d, const HostScene& out, size_t g0, size_t g1, float4* dst) {
    for (size_t bi = 0; bi < out.blas.size(); ++bi) {
        const HostBlas& b = out.blas[bi];
        const size_t po = out.blas_prim_offset[bi];
        const size_t lo = std::max(g0, po), hi = std::min(g1, po + b.tris.size());
        if (lo >= hi) continue;
#pragma omp parallel for schedule(static) if (hi - lo > 8192)
        for (int64_t g = (int64_t)lo; g < (int64_t)hi; ++g) {
            const HostTri& t = b.tris[(size_t)g - po];
            float4* q = dst + ((size_t)g - g0) * 12;
            if (t.sphere) {
                // sphere prim: attr record = {inverse Mat4 rows 0..2 | normal matrix rows (w: centre) | world t, q, s |
                // material}; what primitive.rs:183-190 derives from the node TRS on every shaded hit
                const HostSphere& sp = b.spheres[t.sphere - 1];
                M4 inv = m4_from_inversed(sp.world);
                M3 nm = normal_matrix_of(sp.world);
                const float cc[3] = {sp.center.x, sp.center.y, sp.center.z};
                for (int r = 0; r < 3; ++r) {
                    q[r] = make_float4(inv.m[r][0], inv.m[r][1], inv.m[r][2], inv.m[r][3]);
                    q[3 + r] = make_float4(nm.m[r][0], nm.m[r][1], nm.m[r][2], cc[r]);
                }
                q[6] = make_float4(sp.world.t.x, sp.world.t.y, sp.world.t.z, 0.0f);
                q[7] = make_float4(sp.world.q.x, sp.world.q.y, sp.world.q.z, sp.world.q.w);
                q[8] = make_float4(sp.world.s.x, sp.world.s.y, sp.world.s.z, 0.0f);
                q[9] = make_float4(0.0f, 0.0f, 0.0f, __builtin_bit_cast(float, sp.material));
                q[10] = q[11] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                continue;
            }
            if (const HostMeshJob* job = mesh_job_of(b, t.src)) {
                // BvhTriangle attributes (bvh/triangle.rs:33-71, primitive.rs:239-268): vertex colour and uv verbatim,
                // normal through the node's normal matrix, tangent and bitangent through its tangent matrix
                const uint32_t local = t.src - (uint32_t)job->base;
                for (int v = 0; v < 3; ++v) {
                    const uint32_t vi = d.indices[job->first_index + (size_t)local * 3 + v]; // range-checked when emitted
                    const rayca_vertex& vx = d.vertices[job->first_vertex + vi];
                    const V3 n = m3_apply(job->normal_matrix, V3{vx.normal[0], vx.normal[1], vx.normal[2]});
                    const V3 tg = m3_apply(job->tangent_matrix, V3{vx.tangent[0], vx.tangent[1], vx.tangent[2]});
                    const V3 bt = m3_apply(job->tangent_matrix, V3{vx.bitangent[0], vx.bitangent[1], vx.bitangent[2]});
                    q[v] = make_float4(vx.color[0], vx.color[1], vx.color[2], vx.color[3]);
                    q[3 + v] = make_float4(n.x, n.y, n.z, vx.uv[0]);
                    q[6 + v] = make_float4(tg.x, tg.y, tg.z, vx.uv[1]);
                    q[9 + v] = make_float4(bt.x, bt.y, bt.z, v == 0 ? __builtin_bit_cast(float, job->material) : 0.0f);
                }
                continue;
            }
            const HostAttr& a = b.attrs[t.src]; // quad-light triangle
            for (int v = 0; v < 3; ++v) {
                q[v] = make_float4(a.color[v][0], a.color[v][1], a.color[v][2], a.color[v][3]);
                q[3 + v] = make_float4(a.normal[v].x, a.normal[v].y, a.normal[v].z, a.uv[v][0]);
                q[6 + v] = make_float4(a.tangent[v].x, a.tangent[v].y, a.tangent[v].z, a.uv[v][1]);
                q[9 + v] = make_float4(a.bitangent[v].x, a.bitangent[v].y, a.bitangent[v].z,
                                       v == 0 ? __builtin_bit_cast(float, a.material) : 0.0f);
            }
        }
    }
}

int build_host_scene(const rayca_scene_desc& d, bool bvh, HostScene& out, std::string& err, bool fast_tree,
                     const BlasBuilder* blas_builder, bool gather_texels, bool defer_attrs, const MeshBlasBuilder* mesh_builder) {
    PhaseTimer phase;
    Flatten f(d, out);
    f.run();
    if (f.status != RAYCA_OK) { err = f.err; return f.status; }

    // ---- images (texels are copied verbatim: nearest, no sRGB decode; sampler.rs:11-29)
    out.images.resize(d.image_count);
    for (uint32_t i = 0; i < d.image_count; ++i) {
        const rayca_image& im = d.images[i];
        size_t bpp = im.color_type == RAYCA_COLOR_RGB8 ? 3 : (im.color_type == RAYCA_COLOR_RGBA8 ? 4 : 16);
        size_t bytes = (size_t)im.width * im.height * bpp;
        size_t off = (out.texel_bytes + 15) & ~size_t(15);
        out.texel_bytes = off + bytes;
        if (gather_texels) { // the device path uploads every image straight from the caller's memory instead (rayca_cuda.cu)
            out.texels.resize(off + bytes);
            if (bytes) std::memcpy(out.texels.data() + off, im.data, bytes);
        }
        out.images[i] = GpuImage{(unsigned long long)off, im.width, im.height, im.color_type, 0};
    }

    phase.lap("images");
    // ---- materials: [0] fallback, then per model in order
    out.materials.push_back(white_pbr());
    std::vector<uint32_t> model_mat_base(d.model_count, 0);
    for (uint32_t mi = 0; mi < d.model_count; ++mi) {
        const rayca_model& m = d.models[mi];
        model_mat_base[mi] = (uint32_t)out.materials.size();
        auto tex = [&](int32_t t) -> int32_t { // model.textures.get(handle) -> image
            if (t < 0 || (uint32_t)t >= m.texture_count) return -1;
            int32_t img = d.textures[m.texture_offset + (uint32_t)t].image;
            if (img < 0 || (uint32_t)img >= m.image_count) return -1;
            return (int32_t)(m.image_offset + (uint32_t)img);
        };
        for (uint32_t k = 0; k < m.material_count; ++k) {
            const rayca_material& mat = d.materials[m.material_offset + k];
            GpuMaterial g{};
            g.tex_albedo = g.tex_normal = g.tex_mr = -1;
            g.kind = mat.kind;
            if (mat.kind == RAYCA_MATERIAL_PBR) {
                if (mat.index < 0 || (uint32_t)mat.index >= m.pbr_count) {
                    g = white_pbr();
                } else {
                    const rayca_pbr_material& p = d.pbr_materials[m.pbr_offset + (uint32_t)mat.index];
                    std::memcpy(g.a, p.color, 16);
                    g.p0 = p.metallic_factor;
                    g.p1 = p.roughness_factor;
                    g.tex_albedo = tex(p.albedo);
                    g.tex_normal = tex(p.normal);
                    g.tex_mr = tex(p.metallic_roughness);
                }
            } else if (mat.kind == RAYCA_MATERIAL_PHONG) {
                rayca_phong_material p{{0, 0, 0, 1}, {0, 0, 0, 1}, {0, 0, 0, 1}, {0, 0, 0, 1}, 0.0f}; // PhongMaterial::DEFAULT
                if (mat.index >= 0 && (uint32_t)mat.index < m.phong_count) p = d.phong_materials[m.phong_offset + (uint32_t)mat.index];
                std::memcpy(g.a, p.ambient, 16);
                std::memcpy(g.b, p.emission, 16);
                std::memcpy(g.c, p.diffuse, 16);
                std::memcpy(g.d, p.specular, 16);
                g.p0 = p.shininess;
                g.emissive = color_close_black(p.emission) ? 0u : 1u;
            } else if (mat.kind == RAYCA_MATERIAL_GGX) {
                rayca_ggx_material p{{0, 0, 0, 1}, {0, 0, 0, 1}, 0.0f};
                if (mat.index >= 0 && (uint32_t)mat.index < m.ggx_count) p = d.ggx_materials[m.ggx_offset + (uint32_t)mat.index];
                std::memcpy(g.a, p.diffuse, 16);
                std::memcpy(g.d, p.specular, 16);
                g.p0 = p.roughness;
            } else {
                err = "unknown material kind";
                return RAYCA_E_INVALID_ARG;
            }
            out.materials.push_back(g);
        }
    }

    phase.lap("materials/lights/camera");
    // ---- BvhScene::from_scene: one BLAS per model that owns mesh nodes, ascending model handle
    // (the reference iterates a HashMap; the order only matters for exact cross-BLAS depth ties)
    for (uint32_t mi = 0; mi < d.model_count; ++mi) {
        if (f.mesh_draws[mi].empty() && f.quad_draws[mi].empty()) continue;
        const rayca_model& m = d.models[mi];
        HostBlas blas;
        blas.model = mi;
        // Emit order is the reference's (primitive.rs:349-367): nodes in draw order, primitives in mesh order.  Pass 1
        // validates, places every primitive (a sphere takes one slot, a mesh its triangle count) and fills the
        // spheres; pass 2 fills the triangles, all meshes of the model in ONE parallel loop over 4096-triangle
        // chunks (a model made of many small meshes would otherwise be emitted by a single thread).
        struct MeshJob {
            const MeshDraw* md;
            const rayca_geometry* g;
            size_t base;
        };
        struct Chunk { uint32_t job, first, count; };
        std::vector<MeshJob> jobs;
        std::vector<Chunk> chunks;
        struct SphereSlot { size_t slot; uint32_t sphere; uint32_t material; float radius; };
        std::vector<SphereSlot> sphere_slots;
        std::vector<HostMeshEmit> emits;
        size_t total_prims = 0;
        for (const MeshDraw& md : f.mesh_draws[mi]) {
            const rayca_node& n = d.nodes[m.node_offset + md.node_local];
            if ((uint32_t)n.mesh >= m.mesh_count) { err = "bad mesh index"; return RAYCA_E_INVALID_ARG; }
            const rayca_mesh& mesh = d.meshes[m.mesh_offset + (uint32_t)n.mesh];
            M3 tangent_matrix = tangent_matrix_of(md.world);
            M3 normal_matrix = normal_matrix_of(md.world);
            for (uint32_t pi = 0; pi < mesh.primitive_count; ++pi) {
                uint32_t ph = d.mesh_primitives[mesh.first_primitive + pi];
                if (ph >= m.primitive_count) { err = "bad primitive handle"; return RAYCA_E_INVALID_ARG; }
                const rayca_primitive& prim = d.primitives[m.primitive_offset + ph];
                if (prim.geometry >= m.geometry_count) { err = "bad geometry handle"; return RAYCA_E_INVALID_ARG; }
                const rayca_geometry& g = d.geometries[m.geometry_offset + prim.geometry];
                uint32_t material = 0;
                if (prim.material >= 0 && (uint32_t)prim.material < m.material_count) material = model_mat_base[mi] + (uint32_t)prim.material;
                if (g.kind == RAYCA_GEOMETRY_SPHERE) { // BvhPrimitive::from_sphere (primitive.rs:270-282)
                    HostSphere sp;
                    sp.center = V3{g.center[0], g.center[1], g.center[2]};
                    sp.radius2 = g.radius * g.radius;
                    sp.world = md.world;
                    sp.material = material;
                    blas.spheres.push_back(sp);
                    sphere_slots.push_back({total_prims, (uint32_t)blas.spheres.size(), material, g.radius});
                    total_prims += 1;
                    out.has_spheres = true;
                    continue;
                }
                if (g.kind != RAYCA_GEOMETRY_TRIANGLE_MESH) { err = "unknown geometry kind"; return RAYCA_E_INVALID_ARG; }
                const uint32_t ntri = g.index_count / 3;
                const uint32_t job = (uint32_t)jobs.size();
                jobs.push_back({&md, &g, total_prims});
                HostMeshJob rec;
                rec.base = total_prims;
                rec.ntri = ntri;
                rec.first_index = g.first_index;
                rec.first_vertex = g.first_vertex;
                rec.material = material;
                rec.tangent_matrix = tangent_matrix;
                rec.normal_matrix = normal_matrix;
                blas.jobs.push_back(rec);
                if (mesh_builder) emits.push_back({rec, md.world, g.vertex_count});
                for (uint32_t first = 0; first < ntri; first += 4096) chunks.push_back({job, first, std::min(4096u, ntri - first)});
                total_prims += ntri;
            }
        }
        const uint32_t max_depth = bvh ? 255u : 0u; // scene.rs:95-99: !config.bvh => max_depth(0)
        if (mesh_builder && sphere_slots.empty() && f.quad_draws[mi].empty() && total_prims >= 16384 && total_prims < 0xFFFFFFF0ull) {
            const int r = (*mesh_builder)(blas, emits, max_depth, err);
            if (r < 0) return r;
            if (r == 1) {
                phase.lap("emit + reference BLAS (device)");
                out.blas.push_back(std::move(blas));
                continue;
            }
        }
        blas.tris.reserve(total_prims + 2 * f.quad_draws[mi].size()); // the quad-light triangles are appended below
        blas.attrs.reserve(total_prims + 2 * f.quad_draws[mi].size());
        blas.tris.resize(total_prims);
        blas.attrs.resize(total_prims); // default-initialised (PodVector): mesh triangles never touch their slots
        for (const SphereSlot& ss : sphere_slots) {
            const HostSphere& sp = blas.spheres[ss.sphere - 1];
            HostTri ht{};
            // Sphere::{get_center,get_radius,min,max} (sphere.rs:72-93,169-181); reduce_max runs over 4 lanes, w = 0
            V3 c = trs_point(sp.world, sp.center);
            float r = ss.radius * fmaxf(fmaxf(fmaxf(sp.world.s.x, sp.world.s.y), sp.world.s.z), 0.0f);
            ht.centroid = c;
            ht.bmin = c - V3{r, r, r};
            ht.bmax = c + V3{r, r, r};
            ht.src = (uint32_t)ss.slot;
            ht.sphere = ss.sphere;
            blas.tris[ss.slot] = ht;
            HostAttr ha{};
            ha.material = ss.material;
            blas.attrs[ss.slot] = ha;
        }
        int bad = 0;
#pragma omp parallel for schedule(dynamic, 1) if (total_prims > 8192)
        for (int64_t ci = 0; ci < (int64_t)chunks.size(); ++ci) {
            const Chunk ch = chunks[(size_t)ci];
            const MeshJob& job = jobs[ch.job];
            const rayca_geometry& g = *job.g;
            const MeshDraw& md = *job.md;
            for (uint32_t t = ch.first; t < ch.first + ch.count; ++t) {
                HostTri& ht = blas.tris[job.base + t];
                V3 mv[3];
                for (int k = 0; k < 3; ++k) { // the shading attributes of the vertices are read by pack_tri_attrs
                    uint32_t vi = d.indices[g.first_index + (size_t)t * 3 + k];
                    if (vi >= g.vertex_count) { bad = 1; vi = 0; }
                    const rayca_vertex& v = d.vertices[g.first_vertex + vi];
                    mv[k] = V3{v.pos[0], v.pos[1], v.pos[2]};
                    ht.v[k] = trs_point(md.world, mv[k]); // Triangle::get_vertex (triangle.rs:67-69)
                }
                // Triangle::new centroid quirk (triangle.rs:59-62) then `trs * Vec3` (trs.rs:253-262)
                V3 c = ((mv[0] + mv[1]) + mv[2]) * 0.3333f;
                ht.centroid = trs_vec(md.world, c);
                ht.bmin = V3{fminf(fminf(fminf(kF32Max, ht.v[0].x), ht.v[1].x), ht.v[2].x),
                             fminf(fminf(fminf(kF32Max, ht.v[0].y), ht.v[1].y), ht.v[2].y),
                             fminf(fminf(fminf(kF32Max, ht.v[0].z), ht.v[1].z), ht.v[2].z)};
                ht.bmax = V3{fmaxf(fmaxf(fmaxf(-kF32Max, ht.v[0].x), ht.v[1].x), ht.v[2].x),
                             fmaxf(fmaxf(fmaxf(-kF32Max, ht.v[0].y), ht.v[1].y), ht.v[2].y),
                             fmaxf(fmaxf(fmaxf(-kF32Max, ht.v[0].z), ht.v[1].z), ht.v[2].z)};
                ht.src = (uint32_t)(job.base + t);
                ht.sphere = 0;
            }
        }
        if (bad) { err = "vertex index out of range"; return RAYCA_E_INVALID_ARG; }
        // BvhPrimitive::from_quad_light (primitive.rs:310-346), after the meshes (primitive.rs:363-367): triangles
        // [a, d, b] and [a, c, d], a = origin of the light node; vertex normal = quad normal (not transformed),
        // colour WHITE, uv / tangents zero; material = the light's emissive Phong material
        for (const QuadDraw& qd : f.quad_draws[mi]) {
            const rayca_light& l = d.lights[qd.light];
            V3 ab{l.ab[0], l.ab[1], l.ab[2]}, ac{l.ac[0], l.ac[1], l.ac[2]};
            V3 nq = normalize(cross(ab, ac));
            V3 a{0.0f, 0.0f, 0.0f};
            V3 b = a + ab, dd = (a + ab) + ac, c = a + ac;
            V3 tri[2][3] = {{a, dd, b}, {a, c, dd}};
            uint32_t material = 0;
            if (l.material >= 0 && (uint32_t)l.material < m.material_count) material = model_mat_base[mi] + (uint32_t)l.material;
            for (int t = 0; t < 2; ++t) {
                HostTri ht{};
                HostAttr ha{};
                for (int k = 0; k < 3; ++k) {
                    ht.v[k] = trs_point(qd.world, tri[t][k]);
                    for (int ch = 0; ch < 4; ++ch) ha.color[k][ch] = 1.0f;
                    ha.normal[k] = nq;
                }
                ha.material = material;
                V3 cen = ((tri[t][0] + tri[t][1]) + tri[t][2]) * 0.3333f;
                ht.centroid = trs_vec(qd.world, cen);
                ht.bmin = V3{fminf(fminf(fminf(kF32Max, ht.v[0].x), ht.v[1].x), ht.v[2].x),
                             fminf(fminf(fminf(kF32Max, ht.v[0].y), ht.v[1].y), ht.v[2].y),
                             fminf(fminf(fminf(kF32Max, ht.v[0].z), ht.v[1].z), ht.v[2].z)};
                ht.bmax = V3{fmaxf(fmaxf(fmaxf(-kF32Max, ht.v[0].x), ht.v[1].x), ht.v[2].x),
                             fmaxf(fmaxf(fmaxf(-kF32Max, ht.v[0].y), ht.v[1].y), ht.v[2].y),
                             fmaxf(fmaxf(fmaxf(-kF32Max, ht.v[0].z), ht.v[1].z), ht.v[2].z)};
                ht.src = (uint32_t)blas.tris.size();
                blas.tris.push_back(ht);
                blas.attrs.push_back(ha);
            }
        }
        phase.lap("emit triangles");
        if (blas_builder && (*blas_builder)(blas, max_depth)) {
            phase.lap("reference BLAS build (device)");
        } else {
            build_blas(blas, max_depth);
            phase.lap("reference BLAS build (host)");
        }
        out.blas.push_back(std::move(blas));
    }
    if (out.blas.empty()) { // Tlas::new would assert on an empty range (tlas.rs:75)
        err = "scene has no triangle geometry";
        return RAYCA_E_INVALID_ARG;
    }

    // ---- TLAS (tlas.rs:74-134): median split of BLAS root boxes on the longest axis, then
    // re-expressed as adjacent node pairs above the BLAS roots.
    struct TNode {
        V3 bmin, bmax;
        int left = -1, right = -1;
        uint32_t offset = 0, count = 0;
    };
    std::vector<uint32_t> order(out.blas.size());
    for (uint32_t i = 0; i < order.size(); ++i) order[i] = i;
    std::vector<TNode> tnodes;
    std::function<int(uint32_t, uint32_t)> tbuild = [&](uint32_t offset, uint32_t count) -> int {
        TNode self;
        self.offset = offset;
        self.count = count;
        self.bmin = V3{kF32Max, kF32Max, kF32Max};
        self.bmax = V3{-kF32Max, -kF32Max, -kF32Max};
        for (uint32_t i = offset; i < offset + count; ++i) {
            const HostNode& r = out.blas[order[i]].nodes[0];
            self.bmin = V3{fminf(self.bmin.x, r.bmin.x), fminf(self.bmin.y, r.bmin.y), fminf(self.bmin.z, r.bmin.z)};
            self.bmax = V3{fmaxf(self.bmax.x, r.bmax.x), fmaxf(self.bmax.y, r.bmax.y), fmaxf(self.bmax.z, r.bmax.z)};
        }
        V3 extent = self.bmax - self.bmin;
        int axis = 0;
        if (extent.y > extent.x) axis = 1;
        float ea = axis == 0 ? extent.x : extent.y;
        if (extent.z > ea) axis = 2;
        auto ax = [&](const V3& v) { return axis == 0 ? v.x : (axis == 1 ? v.y : v.z); };
        float split = ax(self.bmin) + ax(extent) * 0.5f;
        size_t i = offset, j = (size_t)offset + count;
        while (i < j) {
            const HostNode& r = out.blas[order[i]].nodes[0];
            V3 centroid = (r.bmax - r.bmin) / 2.0f; // AABB::get_centroid is the half EXTENT (aabb.rs:95-97)
            if (ax(centroid) < split) {
                ++i;
            } else {
                std::swap(order[i], order[j - 1]);
                --j;
            }
        }
        uint32_t lc = (uint32_t)i - offset, rc = count - lc;
        if (lc > 0 && rc > 0) {
            self.left = tbuild(offset, lc);
            self.right = tbuild(offset + lc, rc);
            self.count = 0;
        }
        tnodes.push_back(self);
        return (int)tnodes.size() - 1;
    };
    int troot = tbuild(0, (uint32_t)order.size());
    {   // apply the TLAS permutation so that blas[i] is blas_nodes[i] (Hit::blas, tlas.rs:152)
        std::vector<HostBlas> sorted;
        sorted.reserve(out.blas.size());
        for (uint32_t i : order) sorted.push_back(std::move(out.blas[i]));
        out.blas.swap(sorted);
    }

    phase.lap("TLAS");
    // ---- pack BLAS nodes / triangles
    size_t total_nodes = 0, total_tris = 0;
    out.blas_node_offset.assign(1, 0);
    out.blas_prim_offset.assign(1, 0);
    for (const HostBlas& b : out.blas) {
        total_nodes += b.nodes.size();
        total_tris += b.prim_count();
        out.blas_node_offset.push_back((uint32_t)total_nodes);
        out.blas_prim_offset.push_back((uint32_t)total_tris);
    }
    if (total_tris >= 0xFFFFFFF0ull || total_nodes >= 0xFFFFFFF0ull) { err = "scene too large"; return RAYCA_E_UNSUPPORTED; }
    out.nodes.resize(total_nodes);
    out.tri_pos.resize(total_tris * 3); // tri_attr: sized by pack_tri_attrs (the device path streams it in pieces instead)
    uint32_t max_blas_depth = 0;
    for (size_t bi = 0; bi < out.blas.size(); ++bi) {
        const HostBlas& b = out.blas[bi];
        uint32_t no = out.blas_node_offset[bi], po = out.blas_prim_offset[bi];
        max_blas_depth = std::max(max_blas_depth, b.depth);
        for (size_t k = 0; k < b.nodes.size(); ++k) {
            const HostNode& n = b.nodes[k];
            GpuNode g;
            g.min_x = n.bmin.x; g.min_y = n.bmin.y; g.min_z = n.bmin.z;
            g.max_x = n.bmax.x; g.max_y = n.bmax.y; g.max_z = n.bmax.z;
            g.link0 = n.count ? po + n.offset : no + n.offset;
            g.link1 = n.count;
            out.nodes[no + k] = g;
        }
#pragma omp parallel for schedule(static) if (b.tris.size() > 8192)
        for (int64_t k = 0; k < (int64_t)b.tris.size(); ++k) {
            const HostTri& t = b.tris[(size_t)k];
            size_t gi = (size_t)po + (size_t)k;
            float4* p = &out.tri_pos[gi * 3];
            if (t.sphere) {
                const HostSphere& sp = b.spheres[t.sphere - 1];
                p[0] = make_float4(sp.center.x, sp.center.y, sp.center.z, __builtin_bit_cast(float, t.src));
                p[1] = make_float4(sp.radius2, 0.0f, 0.0f, __builtin_bit_cast(float, sp.material));
                p[2] = make_float4(0.0f, 0.0f, 0.0f, __builtin_bit_cast(float, (uint32_t)bi));
                continue;
            }
            p[0] = make_float4(t.v[0].x, t.v[0].y, t.v[0].z, __builtin_bit_cast(float, t.src));
            p[1] = make_float4(t.v[1].x, t.v[1].y, t.v[1].z, __builtin_bit_cast(float, material_of(b, t.src)));
            p[2] = make_float4(t.v[2].x, t.v[2].y, t.v[2].z, __builtin_bit_cast(float, (uint32_t)bi));
        }
    }
    if (!defer_attrs) pack_tri_attrs(d, out);

    phase.lap("pack nodes/triangles/attrs");
    // ---- TLAS top pairs: every record is a (box, links) copy of its child, so the walk needs
    // only one node format.  A TLAS leaf holding k > 1 BLAS becomes a right-leaning comb whose
    // boxes are unions of BLAS root boxes (supersets never reject what the reference accepts).
    auto blas_root_record = [&](uint32_t bi) -> GpuNode { return out.nodes[out.blas_node_offset[bi]]; };
    uint32_t top_depth = 0;
    std::function<GpuNode(int, uint32_t)> emit_tnode;
    std::function<GpuNode(uint32_t, uint32_t, uint32_t)> emit_comb = [&](uint32_t first, uint32_t count, uint32_t depth) -> GpuNode {
        top_depth = std::max(top_depth, depth);
        if (count == 1) return blas_root_record(first);
        uint32_t pair = (uint32_t)out.nodes.size();
        out.nodes.resize(out.nodes.size() + 2);
        GpuNode l = blas_root_record(first);
        GpuNode r = emit_comb(first + 1, count - 1, depth + 1);
        out.nodes[pair] = l;
        out.nodes[pair + 1] = r;
        GpuNode self;
        self.min_x = fminf(l.min_x, r.min_x); self.min_y = fminf(l.min_y, r.min_y); self.min_z = fminf(l.min_z, r.min_z);
        self.max_x = fmaxf(l.max_x, r.max_x); self.max_y = fmaxf(l.max_y, r.max_y); self.max_z = fmaxf(l.max_z, r.max_z);
        self.link0 = pair;
        self.link1 = 0;
        return self;
    };
    emit_tnode = [&](int ti, uint32_t depth) -> GpuNode {
        top_depth = std::max(top_depth, depth);
        const TNode tn = tnodes[(size_t)ti];
        if (tn.left < 0 && tn.right < 0) return emit_comb(tn.offset, tn.count, depth);
        uint32_t pair = (uint32_t)out.nodes.size();
        out.nodes.resize(out.nodes.size() + 2);
        GpuNode l = emit_tnode(tn.left, depth + 1);
        GpuNode r = emit_tnode(tn.right, depth + 1);
        out.nodes[pair] = l;
        out.nodes[pair + 1] = r;
        GpuNode self;
        self.min_x = tn.bmin.x; self.min_y = tn.bmin.y; self.min_z = tn.bmin.z;
        self.max_x = tn.bmax.x; self.max_y = tn.bmax.y; self.max_z = tn.bmax.z;
        self.link0 = pair;
        self.link1 = 0;
        return self;
    };
    out.root = emit_tnode(troot, 0);
    out.max_stack = out.ref_stack = top_depth + max_blas_depth + 2;
    phase.lap("TLAS top pairs");
    if (fast_tree) {
        build_fast_tree(out);
        phase.lap("performance tree (host)");
        out.max_stack = std::max(out.max_stack, out.fast_depth + 2);
    }
    return RAYCA_OK;
}

} // namespace rb
