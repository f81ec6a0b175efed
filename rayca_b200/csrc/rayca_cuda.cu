// rayca_cuda.cu — C ABI of librayca_cuda (include/rayca_cuda.h): context, scene upload, render.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/rayca_cuda.h"
#include "gpu_fast_build.h"
#include "host_scene.h"
#include "render.cuh"
#include "wavefront.h"

using namespace rb;

namespace {

thread_local std::string g_last_error;

int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}
#define CUDA_TRY(expr)                                                                                      \
    do {                                                                                                    \
        cudaError_t e__ = (expr);                                                                           \
        if (e__ != cudaSuccess)                                                                             \
            return fail(e__ == cudaErrorMemoryAllocation ? RAYCA_E_OOM : RAYCA_E_CUDA,                     \
                        std::string(#expr) + ": " + cudaGetErrorString(e__));                               \
    } while (0)

double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

} // namespace

struct rayca_ctx {
    int device = 0;
    int sm_count = 1;
    unsigned long long* d_counters = nullptr; // kCounterSlots x u64
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    rayca_stats last{};
    uint64_t last_primary = 0;
    uint32_t last_launches = 1; // kernels of the last frame (row bands and wavefront stages add up)
    // scratch for the host-buffer entry point: device planes, a side stream for the read-back, and
    // one event per row band (kernel done -> copy may start)
    void* d_scratch = nullptr;
    size_t scratch_bytes = 0;
    cudaStream_t render_stream = nullptr, copy_stream = nullptr;
    cudaEvent_t band_done[8] = {};
    // multi-device context (rayca_cuda_init_devices): this context drives extra[0..] as well, each a complete
    // single-device context of its own; scene_create replicates the scene on them and rayca_cuda_render shards the
    // frame's row tiles over all of them from the ONE calling host thread
    std::vector<rayca_ctx*> extra;
    // pinned (portable, mapped) staging frame of a multi-device context: pageable caller buffers are filled through it
    unsigned char* h_stage = nullptr;
    size_t h_stage_bytes = 0;
    rb::WaveScratch wave; // device scratch of the wavefront path (wavefront.cu)
    // scenes created on this context and not destroyed yet: shutdown releases their device arrays and orphans them, so
    // that a scene destroyed AFTER its context (garbage-collected bindings do that) never touches a dead stream
    std::vector<rayca_scene*> scenes;
};

struct rayca_scene {
    rayca_ctx* ctx = nullptr;
    HostScene host; // kept for BVH read-back; texel / attribute vectors are released after upload
    void* d_nodes = nullptr;
    void* d_tri_pos = nullptr;
    void* d_tri_attr = nullptr;
    void* d_materials = nullptr;
    void* d_images = nullptr;
    void* d_texels = nullptr;
    void* d_lights = nullptr;
    void* d_blas_prim_offset = nullptr;
    void* d_fast_nodes = nullptr;
    void* d_fast_tri = nullptr;
    size_t fast_node_count = 0; // 32-byte records in d_fast_nodes
    GpuSceneView view{};
    uint32_t max_stack = 0;
    // multi-device contexts: device copies of the arrays above on ctx->extra[i] (replicas[i]->primary == this); a
    // replica owns its device arrays only — flags, lights and materials the launch code consults stay with the primary
    std::vector<rayca_scene*> replicas;
    rayca_scene* primary = nullptr;
    size_t nbytes[10] = {}; // byte sizes of d_nodes .. d_fast_tri in declaration order (what a replica copies)
};

namespace {
const HostScene& host_of(const rayca_scene& s) { return s.primary ? s.primary->host : s.host; }
} // namespace

namespace {

// index of one of rayca_scene's ten device-array members (d_nodes .. d_fast_tri are consecutive void* members)
size_t slot_of(void* const* member, const rayca_scene* s) { return (size_t)(member - &s->d_nodes); }

// Scene arrays come from the device's stream-ordered pool on the context's render stream (see init_one): a `draw` that
// rebuilds per call gets last call's memory back without cudaMalloc / cudaFree round trips.  scene_create synchronises
// the stream before it returns, so the arrays are safe to read from any stream afterwards.
template <typename V>
int upload(void** dst, const V& src, cudaStream_t st) {
    using T = typename V::value_type;
    size_t bytes = std::max<size_t>(src.size() * sizeof(T), 16);
    CUDA_TRY(cudaMallocAsync(dst, bytes, st));
    if (!src.empty()) CUDA_TRY(cudaMemcpyAsync(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice, st));
    return RAYCA_OK;
}

int validate_config(const rayca_config& c, const rayca_scene& sc) {
    if (c.samples_per_pixel == 0) return fail(RAYCA_E_INVALID_ARG, "samples_per_pixel must be >= 1");
    if (!(c.gamma == c.gamma) || c.gamma == 0.0f) return fail(RAYCA_E_INVALID_ARG, "gamma must be non-zero");
    switch (c.integrator) {
        case RAYCA_INTEGRATOR_FLAT: break;
        case RAYCA_INTEGRATOR_ANALYTIC_DIRECT: break;
        case RAYCA_INTEGRATOR_DIRECT:
            if (c.light_samples == 0) return fail(RAYCA_E_INVALID_ARG, "light_samples must be >= 1");
            break;
        case RAYCA_INTEGRATOR_RAYTRACER:
        case RAYCA_INTEGRATOR_SCRATCHER:
            if (c.max_depth + 1 > (uint32_t)kMaxDepth) return fail(RAYCA_E_UNSUPPORTED, "max_depth > 16");
            for (const GpuLight& l : host_of(sc).lights) // Light::get_direction is todo!() for quad lights (light/mod.rs:79-85)
                if (l.kind == RAYCA_LIGHT_QUAD) return fail(RAYCA_E_UNSUPPORTED, "quad lights under the raytracer / scratcher");
            // Hit::get_radiance -> get_pbr_material panics for Material::Ggx (the Raytracer's ggx::get_radiance reads PBR
            // fields, rayca-soft/src/brdf/ggx.rs:11-52): refused like the reference's other panics
            for (size_t k = 1; k < host_of(sc).materials.size(); ++k)
                if (host_of(sc).materials[k].kind == RAYCA_MATERIAL_GGX)
                    return fail(RAYCA_E_UNSUPPORTED, "GGX materials under the raytracer / scratcher (the reference panics)");
            break;
        case RAYCA_INTEGRATOR_PATHTRACER: {
            if (c.direct_sampler != RAYCA_SAMPLER_NEE && c.direct_sampler != RAYCA_SAMPLER_NONE && c.direct_sampler != RAYCA_SAMPLER_MIS)
                return fail(RAYCA_E_UNSUPPORTED, "direct sampler (None / Nee / Mis)");
            if (c.max_depth > (uint32_t)kMaxDepth) return fail(RAYCA_E_UNSUPPORTED, "max_depth > 17");
            if (c.light_samples == 0) return fail(RAYCA_E_INVALID_ARG, "light_samples must be >= 1");
            bool mis = c.direct_sampler == RAYCA_SAMPLER_MIS;
            bool nee = c.direct_sampler == RAYCA_SAMPLER_NEE || mis;
            bool bounces = c.russian_roulette || c.max_depth > (nee ? 1u : 0u);
            bool material_sampling = mis || (bounces && c.indirect_sampler == RAYCA_SAMPLER_BRDF);
            if (mis && host_of(sc).lights.size() * (size_t)c.light_samples > kMaxMisSamples)
                return fail(RAYCA_E_UNSUPPORTED, "Mis keeps at most 16 light samples per hit");
            if (bounces && c.indirect_sampler != RAYCA_SAMPLER_COSINE && c.indirect_sampler != RAYCA_SAMPLER_HEMISPHERE &&
                c.indirect_sampler != RAYCA_SAMPLER_BRDF)
                return fail(RAYCA_E_UNSUPPORTED, "indirect sampler");
            if (material_sampling) { // Material::get_t is todo!() for PBR (material/mod.rs:163-169)
                const HostScene& hs = host_of(sc);
                bool pbr = hs.materials.size() <= 1;
                for (size_t k = 1; k < hs.materials.size(); ++k) pbr = pbr || hs.materials[k].kind == RAYCA_MATERIAL_PBR;
                if (pbr) return fail(RAYCA_E_UNSUPPORTED, "the Brdf / Mis samplers need Phong / GGX materials");
            }
            if (bounces && c.light_samples != 1) { // a tree of light_samples^depth paths per sample (pathtracer_branching)
                if (c.russian_roulette) return fail(RAYCA_E_UNSUPPORTED, "light_samples > 1 with russian roulette");
                if (c.max_depth > (uint32_t)kMaxBranchDepth) return fail(RAYCA_E_UNSUPPORTED, "light_samples > 1 with max_depth > 8");
            }
            if (nee)
                for (const GpuLight& l : host_of(sc).lights) // NEE: directional is todo!() (nee.rs:177)
                    if (l.kind == RAYCA_LIGHT_DIRECTIONAL) return fail(RAYCA_E_UNSUPPORTED, "NEE with a directional light");
            break;
        }
        default: return fail(RAYCA_E_INVALID_ARG, "unknown integrator");
    }
    return RAYCA_OK;
}

// Launches one instantiation.  Persistent scheduling sizes the grid to what is resident at once
// (SM count x blocks per SM, from the occupancy API, cached per instantiation); otherwise one block
// per 16x8 pixel group.
template <int INTEGRATOR, int MODE, bool COUNT, bool PERSISTENT, bool BOUNCE>
void launch_one(const RenderParams& P, dim3 groups, int sm_count, cudaStream_t st) {
    if (PERSISTENT) {
        // asked per launch: the answer depends on the current device, and a function-local static would be shared by every
        // device and thread of the process; the occupancy query is microseconds and this path is not the default
        int blocks_per_sm = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, render_kernel<INTEGRATOR, MODE, COUNT, true, BOUNCE>, 128, 0) != cudaSuccess ||
            blocks_per_sm < 1)
            blocks_per_sm = 1;
        unsigned long long want = ((unsigned long long)P.tile_count + 3) / 4; // 4 warps per block, one group each
        unsigned long long cap = (unsigned long long)sm_count * blocks_per_sm;
        dim3 grid((unsigned)std::max<unsigned long long>(1, std::min(want, cap)));
        render_kernel<INTEGRATOR, MODE, COUNT, true, BOUNCE><<<grid, 128, 0, st>>>(P);
    } else {
        // A/B: RAYCA_L1_CARVEOUT=<percent of the L1 / shared-memory array asked for shared memory> (the kernels use none: 0 asks
        // for the largest L1); unset: the driver's default.  profiles/r02_ab_walk_variants.txt.
        if (const char* e = std::getenv("RAYCA_L1_CARVEOUT"))
            cudaFuncSetAttribute(render_kernel<INTEGRATOR, MODE, COUNT, false, BOUNCE>, cudaFuncAttributePreferredSharedMemoryCarveout, std::atoi(e));
        render_kernel<INTEGRATOR, MODE, COUNT, false, BOUNCE><<<groups, kBlockThreads, 0, st>>>(P);
    }
}

struct LaunchSel {
    uint32_t mode;
    bool count, persistent, bounce, spheres;
    int sm_count;
};

template <int INTEGRATOR, int MODE>
void launch_mode(const RenderParams& P, dim3 groups, const LaunchSel& k, cudaStream_t st) {
    constexpr bool kPath = INTEGRATOR == RAYCA_INTEGRATOR_PATHTRACER; // BOUNCE only matters to the path tracer
    if (k.count) {
        if (kPath && k.bounce) launch_one<INTEGRATOR, MODE, true, false, kPath>(P, groups, k.sm_count, st);
        else launch_one<INTEGRATOR, MODE, true, false, false>(P, groups, k.sm_count, st);
    } else if (k.persistent) {
        if (kPath && k.bounce) launch_one<INTEGRATOR, MODE, false, true, kPath>(P, groups, k.sm_count, st);
        else launch_one<INTEGRATOR, MODE, false, true, false>(P, groups, k.sm_count, st);
    } else {
        if (kPath && k.bounce) launch_one<INTEGRATOR, MODE, false, false, kPath>(P, groups, k.sm_count, st);
        else launch_one<INTEGRATOR, MODE, false, false, false>(P, groups, k.sm_count, st);
    }
}

template <int INTEGRATOR>
void launch_integrator(const RenderParams& P, dim3 groups, const LaunchSel& k, cudaStream_t st) {
#ifndef RAYCA_FAST_BUILD // tuning variants (tools/variants.py) only carry the product walk
    if (k.mode == RAYCA_TRAVERSAL_EXHAUSTIVE) launch_mode<INTEGRATOR, 1>(P, groups, k, st);
    else if (k.mode == RAYCA_TRAVERSAL_REFERENCE_ORDERED) launch_mode<INTEGRATOR, 2>(P, groups, k, st);
    else if (k.spheres) launch_mode<INTEGRATOR, 3>(P, groups, k, st);
    else
#endif
        launch_mode<INTEGRATOR, 0>(P, groups, k, st);
}

} // namespace

extern "C" {

uint32_t rayca_cuda_abi_version(void) { return 1; }

const char* rayca_cuda_last_error(void) { return g_last_error.c_str(); }

const char* rayca_cuda_strerror(int code) {
    switch (code) {
        case RAYCA_OK: return "ok";
        case RAYCA_E_INVALID_ARG: return "invalid argument";
        case RAYCA_E_UNSUPPORTED: return "unsupported configuration";
        case RAYCA_E_OOM: return "out of memory";
        case RAYCA_E_CUDA: return "CUDA error";
        case RAYCA_E_NO_CAMERA: return "scene has no camera";
        case RAYCA_E_TIMEOUT: return "timed out";
        default: return "unknown error";
    }
}

namespace {
int init_one(rayca_ctx* c, int device) { // every failure leaves `c` in a state rayca_cuda_shutdown can release
    c->device = device;
    CUDA_TRY(cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device));
    {   // keep freed scene memory in the stream-ordered pool instead of returning it to the driver at every sync
        cudaMemPool_t pool = nullptr;
        if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess && pool) {
            unsigned long long keep = ~0ull;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        }
        cudaGetLastError();
    }
    CUDA_TRY(cudaMalloc((void**)&c->d_counters, kCounterSlots * sizeof(unsigned long long)));
    CUDA_TRY(cudaMemset(c->d_counters, 0, kCounterSlots * sizeof(unsigned long long)));
    CUDA_TRY(cudaEventCreate(&c->ev0));
    CUDA_TRY(cudaEventCreate(&c->ev1));
    CUDA_TRY(cudaStreamCreateWithFlags(&c->render_stream, cudaStreamNonBlocking));
    CUDA_TRY(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    for (cudaEvent_t& e : c->band_done) CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    return RAYCA_OK;
}
} // namespace

int rayca_cuda_init(int device, rayca_ctx** out) {
    if (!out) return fail(RAYCA_E_INVALID_ARG, "null out");
    int n = 0;
    CUDA_TRY(cudaGetDeviceCount(&n));
    if (device < 0 || device >= n) return fail(RAYCA_E_INVALID_ARG, "no such CUDA device");
    CUDA_TRY(cudaSetDevice(device));
    rayca_ctx* c = new (std::nothrow) rayca_ctx();
    if (!c) return fail(RAYCA_E_OOM, "host allocation");
    int rc = init_one(c, device);
    if (rc != RAYCA_OK) {
        rayca_cuda_shutdown(c); // null-safe per member: releases whatever was created
        return rc;
    }
    pod_cache_enable_pinning();
    *out = c;
    return RAYCA_OK;
}

int rayca_cuda_init_devices(const int* devices, int device_count, rayca_ctx** out) {
    if (!out || device_count < 0 || (device_count > 0 && !devices)) return fail(RAYCA_E_INVALID_ARG, "null / negative argument");
    int visible = 0;
    CUDA_TRY(cudaGetDeviceCount(&visible));
    std::vector<int> devs;
    if (device_count == 0) {
        for (int d = 0; d < visible; ++d) devs.push_back(d); // all visible devices
    } else {
        devs.assign(devices, devices + device_count);
    }
    if (devs.empty()) return fail(RAYCA_E_CUDA, "no CUDA device");
    if (devs.size() > 32) return fail(RAYCA_E_INVALID_ARG, "at most 32 devices per context");
    rayca_ctx* c = nullptr;
    int rc = rayca_cuda_init(devs[0], &c);
    if (rc != RAYCA_OK) return rc;
    for (size_t k = 1; k < devs.size(); ++k) { // a device may be listed more than once: every entry is a shard of its own
        rayca_ctx* e = nullptr;
        rc = rayca_cuda_init(devs[k], &e);
        if (rc != RAYCA_OK) {
            rayca_cuda_shutdown(c);
            return rc;
        }
        c->extra.push_back(e);
    }
    CUDA_TRY(cudaSetDevice(c->device));
    *out = c;
    return RAYCA_OK;
}

int rayca_cuda_device_count(rayca_ctx* ctx) { return ctx ? 1 + (int)ctx->extra.size() : 0; }

} // extern "C"
namespace {
void release_device_arrays(rayca_scene* s);
}
extern "C" {

void rayca_cuda_shutdown(rayca_ctx* c) {
    if (!c) return;
    for (rayca_ctx* e : c->extra) rayca_cuda_shutdown(e);
    cudaSetDevice(c->device);
    while (!c->scenes.empty()) release_device_arrays(c->scenes.back()); // orphaned: their later destroy frees host memory only
    cudaFree(c->d_counters);
    cudaFree(c->d_scratch);
    cudaFree(c->wave.base);
    if (c->h_stage) cudaFreeHost(c->h_stage);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->render_stream) cudaStreamDestroy(c->render_stream);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    for (cudaEvent_t e : c->band_done)
        if (e) cudaEventDestroy(e);
    delete c;
}

} // extern "C"
namespace {
// gives the scene's device arrays back to the pool and detaches the scene from its context
void release_device_arrays(rayca_scene* s) {
    if (!s->ctx) return;
    cudaSetDevice(s->ctx->device);
    // like cudaFree's implicit synchronisation: no stream may still be reading the arrays; then the memory goes back
    // to the pool in stream order (host-only scenes own no device memory)
    cudaDeviceSynchronize();
    void** arrays = &s->d_nodes;
    for (int k = 0; k < 10; ++k) {
        if (arrays[k]) cudaFreeAsync(arrays[k], s->ctx->render_stream);
        arrays[k] = nullptr;
    }
    auto& live = s->ctx->scenes;
    live.erase(std::remove(live.begin(), live.end(), s), live.end());
    s->ctx = nullptr;
}
} // namespace
extern "C" {

// ---- shading attributes of mesh triangles computed on the device (scene_create) ------------------------------------
// pack_tri_attrs (host_scene.cpp) restated per primitive: the caller's vertices and indices are uploaded as they are
// (13 MB for the 262k-triangle atrium, against 50 MB of packed attributes) and one thread per primitive of the packed
// order looks its mesh job up and writes its 12 quads.  Same operations in the same order as the host pack (the library
// is compiled without FMA contraction on both sides), checked byte for byte under RAYCA_VERIFY_ATTRS=1
// (tests/test_cuda_edge_cases.py).  Spheres and quad-light triangles are left to the host (a handful per scene).
struct DevMeshJob {
    uint32_t base, ntri, first_index, first_vertex, material, pad0, pad1, pad2;
    float tm[9]; // tangent matrix, row-major
    float nm[9]; // normal matrix
    float pad3[2];
};
static_assert(sizeof(DevMeshJob) % 16 == 0, "job records are loaded as whole words");

__device__ __forceinline__ void dev_m3_apply(const float* m, const float* v, float* out) { // m3_apply (host_scene.cpp)
    for (int i = 0; i < 3; ++i) {
        float acc = 0.0f;
        for (int j = 0; j < 3; ++j) acc += m[i * 3 + j] * v[j];
        out[i] = acc;
    }
}

__global__ void mesh_attr_kernel(const float4* __restrict__ tri_pos, uint32_t total, const uint32_t* __restrict__ job_begin,
                                 const DevMeshJob* __restrict__ jobs, const rayca_vertex* __restrict__ vertices,
                                 const uint32_t* __restrict__ indices, float4* __restrict__ tri_attr) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= total) return;
    const uint32_t src = __float_as_uint(tri_pos[3 * (size_t)g].w), bi = __float_as_uint(tri_pos[3 * (size_t)g + 2].w);
    uint32_t lo = job_begin[bi], hi = job_begin[bi + 1];
    const uint32_t first = lo;
    while (lo < hi) { // first job whose base exceeds src
        const uint32_t mid = (lo + hi) >> 1;
        if (jobs[mid].base <= src) lo = mid + 1;
        else hi = mid;
    }
    if (lo == first) return;
    const DevMeshJob& job = jobs[lo - 1];
    if (src >= job.base + job.ntri) return; // a sphere or a quad-light triangle: the host writes its record
    const uint32_t local = src - job.base;
    float4* q = tri_attr + (size_t)g * 12;
    for (int v = 0; v < 3; ++v) {
        const uint32_t vi = indices[job.first_index + (size_t)local * 3 + v];
        const rayca_vertex& vx = vertices[job.first_vertex + vi];
        float n[3], tg[3], bt[3];
        dev_m3_apply(job.nm, vx.normal, n);
        dev_m3_apply(job.tm, vx.tangent, tg);
        dev_m3_apply(job.tm, vx.bitangent, bt);
        q[v] = make_float4(vx.color[0], vx.color[1], vx.color[2], vx.color[3]);
        q[3 + v] = make_float4(n[0], n[1], n[2], vx.uv[0]);
        q[6 + v] = make_float4(tg[0], tg[1], tg[2], vx.uv[1]);
        q[9 + v] = make_float4(bt[0], bt[1], bt[2], v == 0 ? __uint_as_float(job.material) : 0.0f);
    }
}

void rayca_cuda_scene_destroy(rayca_scene* s) {
    if (!s) return;
    const double td0 = now_ms();
    for (rayca_scene* r : s->replicas) rayca_cuda_scene_destroy(r);
    release_device_arrays(s);
    delete s;
    if (std::getenv("RAYCA_TIMING")) std::fprintf(stderr, "[rayca] %-28s %8.2f ms\n", "scene_destroy", now_ms() - td0);
}

int rayca_cuda_scene_create(rayca_ctx* ctx, const rayca_scene_desc* desc, uint32_t bvh, rayca_scene** out, rayca_stats* stats) {
    if (!ctx || !desc || !out) return fail(RAYCA_E_INVALID_ARG, "null argument");
    CUDA_TRY(cudaSetDevice(ctx->device));
    rayca_scene* s = new (std::nothrow) rayca_scene();
    if (!s) return fail(RAYCA_E_OOM, "host allocation");
    s->ctx = ctx;
    ctx->scenes.push_back(s);
    double t0 = now_ms();
    std::string err;
    int rc;
    // the performance tree is built on the device; RAYCA_HOST_FAST_TREE=1 (or a non-default bin count) keeps the
    // host builder for A/B runs
    const bool host_fast_tree = std::getenv("RAYCA_HOST_FAST_TREE") != nullptr || std::getenv("RAYCA_FAST_BINS") != nullptr;
    // so is the reference BLAS of every model with enough primitives to pay for the launches (RAYCA_HOST_REF_TREE=1:
    // host builder everywhere)
    const bool host_ref_tree = std::getenv("RAYCA_HOST_REF_TREE") != nullptr;
    const BlasBuilder device_blas = [&](HostBlas& blas, uint32_t max_depth) -> bool {
        if (host_ref_tree || blas.tris.size() < 16384) return false;
        std::string build_err;
        float ms = 0.0f;
        const int brc = build_blas_device(blas, max_depth, ctx->render_stream, build_err, &ms);
        if (brc == RAYCA_OK && std::getenv("RAYCA_TIMING")) std::fprintf(stderr, "[rayca] %-28s %8.2f ms\n", "  BLAS kernels", ms);
        if (brc != RAYCA_OK) cudaGetLastError();
        return brc == RAYCA_OK;
    };
    // Models made of triangle meshes only are emitted ON THE DEVICE from the caller's vertices and indices, uploaded as
    // they are: emit_mesh_device restates the host emit loop per triangle (same expressions, no FMA on either side),
    // build_blas_device reads the bounds where they are, and the packed positions are a gather by the build's order —
    // the host never holds those triangles (emit 1.4 ms + permutation and packing 0.7 ms of the atrium's draw).  Off for
    // scenes with spheres (their performance tree takes bounds from the host) and under the host-builder / verification
    // switches, which need the host triangles.
    bool any_sphere_geometry = false;
    for (uint32_t i = 0; i < desc->geometry_count; ++i) any_sphere_geometry = any_sphere_geometry || desc->geometries[i].kind == RAYCA_GEOMETRY_SPHERE;
    const bool device_emit = bvh && !host_fast_tree && !host_ref_tree && !any_sphere_geometry && desc->vertex_count > 0 && desc->index_count > 0 &&
                             !std::getenv("RAYCA_HOST_EMIT") && !std::getenv("RAYCA_HOST_ATTRS") && !std::getenv("RAYCA_VERIFY_ATTRS");
    PodVector<unsigned char> raw_stage; // the caller's vertices | indices, page-locked; read by a copy until the final sync
    void* d_raw = nullptr;              // ... on the device (freed in stream order at the end of scene_create)
    const size_t raw_vbytes = (size_t)desc->vertex_count * sizeof(rayca_vertex), raw_ioff = (raw_vbytes + 255) & ~size_t(255);
    const size_t raw_ibytes = (size_t)desc->index_count * 4;
    auto upload_raw = [&]() -> cudaError_t { // gathered by all threads into one page-locked block, one asynchronous copy
        if (d_raw) return cudaSuccess;
        raw_stage.resize(raw_ioff + raw_ibytes);
        struct Piece { const unsigned char* src; size_t dst, bytes; };
        std::vector<Piece> pieces;
        for (size_t o = 0; o < raw_vbytes; o += size_t(1) << 20)
            pieces.push_back({reinterpret_cast<const unsigned char*>(desc->vertices) + o, o, std::min(raw_vbytes - o, size_t(1) << 20)});
        for (size_t o = 0; o < raw_ibytes; o += size_t(1) << 20)
            pieces.push_back({reinterpret_cast<const unsigned char*>(desc->indices) + o, raw_ioff + o, std::min(raw_ibytes - o, size_t(1) << 20)});
#pragma omp parallel for schedule(dynamic, 1) if (pieces.size() > 1)
        for (int64_t k = 0; k < (int64_t)pieces.size(); ++k) std::memcpy(raw_stage.data() + pieces[(size_t)k].dst, pieces[(size_t)k].src, pieces[(size_t)k].bytes);
        cudaError_t ce = cudaMallocAsync(&d_raw, std::max<size_t>(raw_stage.size(), 16), ctx->render_stream);
        if (ce == cudaSuccess && !raw_stage.empty())
            ce = cudaMemcpyAsync(d_raw, raw_stage.data(), raw_stage.size(), cudaMemcpyHostToDevice, ctx->render_stream);
        return ce;
    };
    std::vector<std::pair<uint32_t, DeviceBlasMesh>> device_meshes; // (model, its triangles on the device)
    struct FreeMeshes { // every exit path gives the per-model device arrays back (stream-ordered)
        std::vector<std::pair<uint32_t, DeviceBlasMesh>>& v;
        void*& raw;
        cudaStream_t st;
        ~FreeMeshes() {
            for (auto& m : v) free_device_blas_mesh(m.second, st);
            if (raw) cudaFreeAsync(raw, st);
            raw = nullptr;
        }
    } free_meshes{device_meshes, d_raw, ctx->render_stream};
    const MeshBlasBuilder device_mesh_blas = [&](HostBlas& blas, const std::vector<HostMeshEmit>& emits, uint32_t max_depth, std::string& mesh_err) -> int {
        size_t n = 0;
        std::vector<DeviceMeshJob> jobs;
        for (const HostMeshEmit& e : emits) {
            DeviceMeshJob j{};
            j.base = (uint32_t)e.job.base;
            j.ntri = e.job.ntri;
            j.first_index = e.job.first_index;
            j.first_vertex = e.job.first_vertex;
            j.vertex_count = e.vertex_count;
            j.material = e.job.material;
            j.t[0] = e.world.t.x; j.t[1] = e.world.t.y; j.t[2] = e.world.t.z;
            j.q[0] = e.world.q.x; j.q[1] = e.world.q.y; j.q[2] = e.world.q.z; j.q[3] = e.world.q.w;
            j.s[0] = e.world.s.x; j.s[1] = e.world.s.y; j.s[2] = e.world.s.z;
            if (e.job.base != n) return 0; // the jobs must tile [0, n) in order (they do: emit order)
            n += e.job.ntri;
            jobs.push_back(j);
        }
        if (jobs.empty() || upload_raw() != cudaSuccess) {
            cudaGetLastError();
            return 0;
        }
        DeviceBlasMesh mesh;
        if (emit_mesh_device(jobs, (uint32_t)n, d_raw, static_cast<unsigned char*>(d_raw) + raw_ioff, ctx->render_stream, mesh, mesh_err) != RAYCA_OK) {
            cudaGetLastError();
            return 0;
        }
        DeviceBlasInput in;
        in.n = mesh.n;
        in.mn = mesh.mn;
        in.mx = mesh.mx;
        in.cen = mesh.cen;
        in.order_out = mesh.order;
        float ms = 0.0f;
        std::string build_err;
        const int brc = build_blas_device(blas, max_depth, ctx->render_stream, build_err, &ms, &in); // synchronises the stream
        uint32_t bad = 0;
        if (brc == RAYCA_OK && cudaMemcpy(&bad, mesh.bad, 4, cudaMemcpyDeviceToHost) != cudaSuccess) bad = 0;
        if (brc != RAYCA_OK || bad) {
            cudaGetLastError();
            free_device_blas_mesh(mesh, ctx->render_stream);
            blas.nodes.clear();
            blas.order.clear();
            blas.device_prims = 0;
            if (bad) {
                mesh_err = "vertex index out of range";
                return RAYCA_E_INVALID_ARG;
            }
            return 0; // e.g. a level broader than the device work lists: the host emits and builds
        }
        if (std::getenv("RAYCA_TIMING")) std::fprintf(stderr, "[rayca] %-28s %8.2f ms\n", "  BLAS kernels", ms);
        device_meshes.emplace_back(blas.model, mesh);
        return 1;
    };
    // Texels: the caller's images are pageable as a rule (decoded files), and a pageable cudaMemcpy blocks the calling
    // thread at about 12 GB/s (1.3 of 15 ms of the atrium's draw).  They are gathered into one recycled page-locked block
    // by all host threads instead and leave in ONE asynchronous copy on the copy stream, which runs under everything
    // that follows.  (A helper thread doing the pageable copies was tried first: with the OpenMP workers spinning
    // between the parallel loops of the build, every extra thread cost milliseconds of jitter on a 16-core host.)
    // Offsets as build_host_scene lays the images out (checked below).
    std::vector<size_t> texel_offset(desc->image_count, 0);
    size_t texel_bytes = 0;
    for (uint32_t i = 0; i < desc->image_count; ++i) {
        const rayca_image& im = desc->images[i];
        const size_t bpp = im.color_type == RAYCA_COLOR_RGB8 ? 3 : (im.color_type == RAYCA_COLOR_RGBA8 ? 4 : 16);
        texel_offset[i] = (texel_bytes + 15) & ~size_t(15);
        texel_bytes = texel_offset[i] + (size_t)im.width * im.height * bpp;
    }
    PodVector<unsigned char> texel_stage(texel_bytes);
    {
        struct Piece { const unsigned char* src; size_t dst, bytes; };
        std::vector<Piece> pieces; // 1 MiB pieces so that one large image is spread over the threads too
        for (uint32_t i = 0; i < desc->image_count; ++i) {
            const rayca_image& im = desc->images[i];
            const size_t bpp = im.color_type == RAYCA_COLOR_RGB8 ? 3 : (im.color_type == RAYCA_COLOR_RGBA8 ? 4 : 16);
            const size_t bytes = (size_t)im.width * im.height * bpp;
            for (size_t o = 0; o < bytes; o += size_t(1) << 20)
                pieces.push_back({static_cast<const unsigned char*>(im.data) + o, texel_offset[i] + o, std::min(bytes - o, size_t(1) << 20)});
        }
#pragma omp parallel for schedule(dynamic, 1) if (texel_bytes > (size_t(1) << 20))
        for (int64_t k = 0; k < (int64_t)pieces.size(); ++k) std::memcpy(texel_stage.data() + pieces[(size_t)k].dst, pieces[(size_t)k].src, pieces[(size_t)k].bytes);
    }
    s->nbytes[slot_of(&s->d_texels, s)] = std::max<size_t>(texel_bytes, 16);
    if (int urc = upload(&s->d_texels, texel_stage, ctx->copy_stream); urc != RAYCA_OK) {
        rayca_cuda_scene_destroy(s);
        return urc;
    }
    try {
        rc = build_host_scene(*desc, bvh != 0, s->host, err, host_fast_tree, &device_blas, /*gather_texels=*/false, /*defer_attrs=*/true,
                              device_emit ? &device_mesh_blas : nullptr);
    } catch (const std::bad_alloc&) {
        rayca_cuda_scene_destroy(s);
        return fail(RAYCA_E_OOM, "host allocation while building the scene");
    }
    if (rc != RAYCA_OK) {
        rayca_cuda_scene_destroy(s);
        return fail(rc, err);
    }
    if (s->host.blas_prim_offset.back() >= (1u << 27)) {
        rayca_cuda_scene_destroy(s);
        return fail(RAYCA_E_UNSUPPORTED, "more than 2^27 primitives");
    }
    // The reference allows BLAS depths up to 255 (u8 level, blas.rs:302-316); the product walk (traversal 0) never walks
    // that tree — it only re-validates candidates against reference LEAF boxes — so a deep reference tree is not a
    // reason to refuse the scene: the check moved to render time, for traversal modes 1 / 2 only.
    double t1 = now_ms();
    HostScene& h = s->host;
#define UP(dst, vec)                         \
    do {                                     \
        s->nbytes[slot_of(&s->dst, s)] = std::max<size_t>(h.vec.size() * sizeof(h.vec[0]), 16); \
        int rc__ = upload(&s->dst, h.vec, ctx->render_stream);   \
        if (rc__ != RAYCA_OK) {              \
            rayca_cuda_scene_destroy(s);     \
            return rc__;                     \
        }                                    \
    } while (0)
    const bool timing = std::getenv("RAYCA_TIMING") != nullptr;
    double tl = now_ms();
    auto lap = [&](const char* what) {
        if (!timing) return;
        const double t = now_ms();
        std::fprintf(stderr, "[rayca] %-28s %8.2f ms\n", what, t - tl);
        tl = t;
    };
    if (timing) std::fprintf(stderr, "[rayca] %-28s %8.2f ms\n", "build_host_scene (total)", t1 - t0);
    UP(d_nodes, nodes);
    {   // packed positions: models emitted on the device are gathered there by their build's order, the others come from the host
        bool all_device = !device_meshes.empty();
        for (const HostBlas& b : h.blas) all_device = all_device && b.tris.empty() && b.device_prims > 0;
        if (all_device) {
            s->nbytes[slot_of(&s->d_tri_pos, s)] = std::max<size_t>(h.tri_pos.size() * sizeof(h.tri_pos[0]), 16);
            if (cudaError_t ce = cudaMallocAsync(&s->d_tri_pos, s->nbytes[slot_of(&s->d_tri_pos, s)], ctx->render_stream); ce != cudaSuccess) {
                s->d_tri_pos = nullptr;
                rayca_cuda_scene_destroy(s);
                return fail(ce == cudaErrorMemoryAllocation ? RAYCA_E_OOM : RAYCA_E_CUDA, std::string("triangle positions: ") + cudaGetErrorString(ce));
            }
        } else {
            UP(d_tri_pos, tri_pos);
        }
        for (size_t bi = 0; bi < h.blas.size(); ++bi) {
            if (!h.blas[bi].tris.empty() || h.blas[bi].device_prims == 0) continue;
            for (auto& m : device_meshes)
                if (m.first == h.blas[bi].model && m.second.n == h.blas[bi].device_prims) {
                    if (gather_tri_pos_device(m.second, (uint32_t)bi, static_cast<float4*>(s->d_tri_pos) + (size_t)h.blas_prim_offset[bi] * 3,
                                              ctx->render_stream) != RAYCA_OK) {
                        rayca_cuda_scene_destroy(s);
                        return fail(RAYCA_E_CUDA, "gathering the triangle positions");
                    }
                    break;
                }
        }
    }
    UP(d_materials, materials);
    UP(d_images, images);
    for (uint32_t i = 0; i < desc->image_count; ++i) // the helper thread's layout must be the one the image records carry
        if (h.images[i].offset != texel_offset[i] || h.texel_bytes != texel_bytes) {
            rayca_cuda_scene_destroy(s);
            return fail(RAYCA_E_CUDA, "texel layout mismatch");
        }
    UP(d_lights, lights);
    UP(d_blas_prim_offset, blas_prim_offset);
    lap("uploads (8 arrays)");
    constexpr int kRing = 3;
    PodVector<float4> ring[kRing]; // page-locked staging of the shading attributes (below); read by copies until the final sync
    std::vector<float4> other_attrs;    // ... and the records of the few primitives that are not mesh triangles
    std::vector<DevMeshJob> dev_jobs;   // ... and the job tables of mesh_attr_kernel
    std::vector<uint32_t> dev_job_begin;
    if (host_fast_tree) {
        pack_tri_attrs(*desc, h);
        UP(d_tri_attr, tri_attr);
        UP(d_fast_nodes, fast_nodes);
        UP(d_fast_tri, fast_tri);
        s->fast_node_count = h.fast_nodes.size();
    } else {
        // The shading attributes (the largest array: 192 bytes per primitive) are never materialised on the host: they
        // are packed piece by piece into a ring of three page-locked buffers and leave on the copy stream, the copies
        // running under the packing of the next piece and under the tree build that follows.
        // Scenes made of triangle meshes (all but a handful of primitives) do not even pack them on the host: the caller's
        // vertices and indices go up as they are and mesh_attr_kernel writes the records (above).
        size_t mesh_prims = 0;
        for (const HostBlas& b : h.blas)
            for (const HostMeshJob& j : b.jobs) mesh_prims += j.ntri;
        const size_t total_prims = h.blas_prim_offset.back();
        const bool device_attrs = !device_meshes.empty() || // (those models have no host triangles to pack from)
                                  (mesh_prims > 0 && total_prims - mesh_prims <= 64 && !std::getenv("RAYCA_HOST_ATTRS") &&
                                   desc->vertex_count * sizeof(rayca_vertex) + desc->index_count * 4 <= total_prims * 12 * sizeof(float4));
        if (device_attrs) {
            const size_t attr_bytes = std::max<size_t>(total_prims * 12 * sizeof(float4), 16);
            s->nbytes[slot_of(&s->d_tri_attr, s)] = attr_bytes;
            cudaError_t ce = cudaMallocAsync(&s->d_tri_attr, attr_bytes, ctx->render_stream);
            // the caller's vertices | indices: already on the device when a model was emitted there
            const size_t voff = 0, ioff = raw_ioff;
            if (ce == cudaSuccess) ce = upload_raw();
            dev_jobs.clear();
            std::vector<DevMeshJob>& jobs = dev_jobs;
            std::vector<uint32_t>& job_begin = dev_job_begin;
            job_begin.assign(h.blas.size() + 1, 0);
            for (size_t bi = 0; bi < h.blas.size(); ++bi) {
                for (const HostMeshJob& j : h.blas[bi].jobs) {
                    DevMeshJob dj{};
                    dj.base = (uint32_t)j.base;
                    dj.ntri = j.ntri;
                    dj.first_index = j.first_index;
                    dj.first_vertex = j.first_vertex;
                    dj.material = j.material;
                    for (int r = 0; r < 3; ++r)
                        for (int c = 0; c < 3; ++c) {
                            dj.tm[r * 3 + c] = j.tangent_matrix.m[r][c];
                            dj.nm[r * 3 + c] = j.normal_matrix.m[r][c];
                        }
                    jobs.push_back(dj);
                }
                job_begin[bi + 1] = (uint32_t)jobs.size();
            }
            void *d_jobs = nullptr, *d_job_begin = nullptr;
            if (ce == cudaSuccess) ce = cudaMallocAsync(&d_jobs, jobs.size() * sizeof(DevMeshJob), ctx->render_stream);
            if (ce == cudaSuccess) ce = cudaMallocAsync(&d_job_begin, job_begin.size() * 4, ctx->render_stream);
            if (ce == cudaSuccess) ce = cudaMemcpyAsync(d_jobs, jobs.data(), jobs.size() * sizeof(DevMeshJob), cudaMemcpyHostToDevice, ctx->render_stream);
            if (ce == cudaSuccess) ce = cudaMemcpyAsync(d_job_begin, job_begin.data(), job_begin.size() * 4, cudaMemcpyHostToDevice, ctx->render_stream);
            if (ce == cudaSuccess) {
                mesh_attr_kernel<<<(unsigned)((total_prims + 255) / 256), 256, 0, ctx->render_stream>>>(
                    (const float4*)s->d_tri_pos, (uint32_t)total_prims, (const uint32_t*)d_job_begin, (const DevMeshJob*)d_jobs,
                    reinterpret_cast<const rayca_vertex*>(static_cast<unsigned char*>(d_raw) + voff),
                    reinterpret_cast<const uint32_t*>(static_cast<unsigned char*>(d_raw) + ioff), (float4*)s->d_tri_attr);
                ce = cudaGetLastError();
            }
            // the few primitives that are not mesh triangles: packed here, one small copy each, behind the kernel
            if (ce == cudaSuccess && total_prims != mesh_prims) {
                std::vector<size_t> where;
                for (size_t bi = 0; bi < h.blas.size(); ++bi) {
                    const HostBlas& b = h.blas[bi];
#pragma omp parallel if (b.tris.size() > 8192)
                    {
                        std::vector<size_t> mine;
#pragma omp for schedule(static) nowait
                        for (int64_t k = 0; k < (int64_t)b.tris.size(); ++k) {
                            const HostTri& t = b.tris[(size_t)k];
                            if (t.sphere || !mesh_job_of(b, t.src)) mine.push_back(h.blas_prim_offset[bi] + (size_t)k);
                        }
#pragma omp critical
                        where.insert(where.end(), mine.begin(), mine.end());
                    }
                }
                other_attrs.resize(where.size() * 12);
                for (size_t i = 0; i < where.size() && ce == cudaSuccess; ++i) {
                    pack_tri_attrs_range(*desc, h, where[i], where[i] + 1, other_attrs.data() + i * 12);
                    ce = cudaMemcpyAsync(static_cast<float4*>(s->d_tri_attr) + where[i] * 12, other_attrs.data() + i * 12, 12 * sizeof(float4),
                                         cudaMemcpyHostToDevice, ctx->render_stream);
                }
            }
            // (the staging block and the job tables live until scene_create returns, after the stream has been drained)
            if (d_jobs) cudaFreeAsync(d_jobs, ctx->render_stream);
            if (d_job_begin) cudaFreeAsync(d_job_begin, ctx->render_stream);
            if (ce == cudaSuccess && std::getenv("RAYCA_VERIFY_ATTRS")) { // tests: the device records against the host pack, byte for byte
                pack_tri_attrs(*desc, h);
                std::vector<float4> got(total_prims * 12);
                ce = cudaStreamSynchronize(ctx->render_stream);
                if (ce == cudaSuccess) ce = cudaMemcpy(got.data(), s->d_tri_attr, got.size() * sizeof(float4), cudaMemcpyDeviceToHost);
                if (ce == cudaSuccess && std::memcmp(got.data(), h.tri_attr.data(), got.size() * sizeof(float4)) != 0) {
                    size_t bad = 0;
                    while (std::memcmp(&got[bad * 12], &h.tri_attr[bad * 12], 12 * sizeof(float4)) == 0) ++bad;
                    rayca_cuda_scene_destroy(s);
                    return fail(RAYCA_E_CUDA, "device-computed attributes differ from the host pack at primitive " + std::to_string(bad));
                }
            }
            if (ce != cudaSuccess) {
                rayca_cuda_scene_destroy(s);
                return fail(ce == cudaErrorMemoryAllocation ? RAYCA_E_OOM : RAYCA_E_CUDA, std::string("attribute kernel: ") + cudaGetErrorString(ce));
            }
        } else {
            const size_t total = h.blas_prim_offset.back();
            const size_t attr_bytes = std::max<size_t>(total * 12 * sizeof(float4), 16);
            s->nbytes[slot_of(&s->d_tri_attr, s)] = attr_bytes;
            cudaError_t ce = cudaMallocAsync(&s->d_tri_attr, attr_bytes, ctx->copy_stream);
            constexpr size_t kPiece = 65536; // primitives per piece: 12 MiB
            cudaEvent_t sent[kRing] = {nullptr, nullptr, nullptr};
            for (int k = 0; k < kRing && ce == cudaSuccess; ++k) {
                ring[k].resize(std::min(total, kPiece) * 12);
                ce = cudaEventCreateWithFlags(&sent[k], cudaEventDisableTiming);
            }
            size_t piece = 0;
            for (size_t g0 = 0; g0 < total && ce == cudaSuccess; g0 += kPiece, ++piece) {
                const size_t g1 = std::min(total, g0 + kPiece);
                const int slot = (int)(piece % kRing);
                if (piece >= (size_t)kRing) ce = cudaEventSynchronize(sent[slot]); // the copy that last read this buffer
                if (ce != cudaSuccess) break;
                pack_tri_attrs_range(*desc, h, g0, g1, ring[slot].data());
                ce = cudaMemcpyAsync(static_cast<float4*>(s->d_tri_attr) + g0 * 12, ring[slot].data(), (g1 - g0) * 12 * sizeof(float4),
                                     cudaMemcpyHostToDevice, ctx->copy_stream);
                if (ce == cudaSuccess) ce = cudaEventRecord(sent[slot], ctx->copy_stream);
            }
            // (the ring lives until scene_create returns, after the copy stream has been drained)
            for (int k = 0; k < kRing; ++k)
                if (sent[k]) cudaEventDestroy(sent[k]);
            if (ce != cudaSuccess) {
                cudaStreamSynchronize(ctx->copy_stream);
                rayca_cuda_scene_destroy(s);
                return fail(ce == cudaErrorMemoryAllocation ? RAYCA_E_OOM : RAYCA_E_CUDA, std::string("attribute upload: ") + cudaGetErrorString(ce));
            }
        }
        lap("pack + send attributes");
        FastTreeDevice ft;
        rc = build_fast_tree_device(h, (const float4*)s->d_tri_pos, ctx->render_stream, ft, err, s->d_nodes);
        if (rc != RAYCA_OK) {
            rayca_cuda_scene_destroy(s);
            return fail(rc, err);
        }
        s->d_fast_nodes = ft.nodes;
        s->d_fast_tri = ft.tri;
        s->fast_node_count = (size_t)h.blas_prim_offset.back() * 2 + 2;
        s->nbytes[8] = s->fast_node_count * sizeof(GpuNode);
        s->nbytes[9] = (size_t)h.blas_prim_offset.back() * 4 * sizeof(float4);
        h.fast_root = ft.root;
        h.fast_depth = ft.depth;
        h.max_stack = std::max(h.max_stack, ft.depth + 2);
        if (std::getenv("RAYCA_TIMING"))
            std::fprintf(stderr, "[rayca] %-28s %8.2f ms (%u levels, %u warp subtrees, depth %u)\n", "performance tree (device)",
                         ft.build_ms, ft.levels, ft.small_subtrees, ft.depth);
    }
    lap("performance tree (call)");

    // The walk's stack holds kStackSize entries and saturates beyond (trace.cuh): a performance tree deeper than that
    // would drop subtrees silently, so such a scene is refused here, never mis-rendered.  With 16 SAH bins a level
    // narrows the centroid range 16-fold and the binned builder cannot get near the limit on f32 input; the environment
    // override exists so that tests can exercise the refusal (tests/test_cuda_edge_cases.py).
    {
        uint32_t limit = (uint32_t)kStackSize;
        if (const char* e = std::getenv("RAYCA_MAX_TREE_DEPTH")) limit = std::min<uint32_t>(limit, (uint32_t)std::max(1, std::atoi(e)));
        if (h.fast_depth + 2 > limit) {
            rayca_cuda_scene_destroy(s);
            return fail(RAYCA_E_UNSUPPORTED, "performance tree deeper than the traversal stack (" + std::to_string(h.fast_depth) + " levels, limit " +
                                                 std::to_string(limit - 2) + ")");
        }
    }
#undef UP
#if RAYCA_SOA_PAIRS
    if (s->fast_node_count * sizeof(GpuNode) >= (1ull << 32)) { // the walk addresses pair records with 32-bit byte offsets
        rayca_cuda_scene_destroy(s);
        return fail(RAYCA_E_UNSUPPORTED, "performance tree larger than 4 GiB");
    }
    if (transpose_pairs_device(s->d_fast_nodes, s->fast_node_count & ~size_t(1), ctx->render_stream) != RAYCA_OK ||
        cudaStreamSynchronize(ctx->render_stream) != cudaSuccess) {
        rayca_cuda_scene_destroy(s);
        return fail(RAYCA_E_CUDA, "transposing the pair records");
    }
#endif
    GpuSceneView& v = s->view;
    v.nodes = (const float4*)s->d_nodes;
    v.tri_pos = (const float4*)s->d_tri_pos;
    v.tri_attr = (const float4*)s->d_tri_attr;
    v.materials = (const GpuMaterial*)s->d_materials;
    v.images = (const GpuImage*)s->d_images;
    v.texels = (const unsigned char*)s->d_texels;
    v.lights = (const GpuLight*)s->d_lights;
    v.blas_prim_offset = (const uint32_t*)s->d_blas_prim_offset;
    v.blas_count = (uint32_t)h.blas.size();
    v.light_count = (uint32_t)h.lights.size();
    v.node_count = (uint32_t)h.nodes.size();
    v.tri_count = (uint32_t)(h.tri_pos.size() / 3);
    v.root_min[0] = h.root.min_x; v.root_min[1] = h.root.min_y; v.root_min[2] = h.root.min_z;
    v.root_max[0] = h.root.max_x; v.root_max[1] = h.root.max_y; v.root_max[2] = h.root.max_z;
    v.root_link0 = h.root.link0;
    v.root_link1 = h.root.link1;
    v.camera = h.camera;
    v.fast_nodes = (const float4*)s->d_fast_nodes;
    v.fast_tri = (const float4*)s->d_fast_tri;
    v.fast_root_min[0] = h.fast_root.min_x; v.fast_root_min[1] = h.fast_root.min_y; v.fast_root_min[2] = h.fast_root.min_z;
    v.fast_root_max[0] = h.fast_root.max_x; v.fast_root_max[1] = h.fast_root.max_y; v.fast_root_max[2] = h.fast_root.max_z;
    v.fast_root_ref = h.fast_root.link0;
    v.fast_record_count = (uint32_t)std::min<size_t>(s->fast_node_count, 0xFFFFFFFFu);
    s->max_stack = h.max_stack;
    // host copies that are no longer needed — after the stream has drained: packed into page-locked blocks (host_scene.cpp)
    // the uploads above are truly asynchronous and still read them
    if (cudaError_t ce = cudaStreamSynchronize(ctx->render_stream); ce != cudaSuccess) {
        rayca_cuda_scene_destroy(s);
        return fail(RAYCA_E_CUDA, std::string("scene_create: ") + cudaGetErrorString(ce));
    }
    if (cudaError_t ce = cudaStreamSynchronize(ctx->copy_stream); ce != cudaSuccess) { // texels and shading attributes went this way
        rayca_cuda_scene_destroy(s);
        return fail(RAYCA_E_CUDA, std::string("scene_create (copy stream): ") + cudaGetErrorString(ce));
    }
    std::vector<unsigned char>().swap(h.texels);
    PodVector<float4>().swap(h.tri_attr);
    PodVector<float4>().swap(h.tri_pos);
    std::vector<float4>().swap(h.fast_tri);
    std::vector<GpuNode>().swap(h.fast_nodes);
    for (HostBlas& b : h.blas) PodVector<HostAttr>().swap(b.attrs);
    lap("free host copies");
    // multi-device context: copy the finished device arrays to every other device (device-to-device, no rebuild)
    for (rayca_ctx* e : ctx->extra) {
        rayca_scene* r = new (std::nothrow) rayca_scene();
        if (!r) {
            rayca_cuda_scene_destroy(s);
            return fail(RAYCA_E_OOM, "host allocation");
        }
        r->ctx = e;
        e->scenes.push_back(r);
        r->primary = s;
        s->replicas.push_back(r); // owned by s from here on: every failure below is released by scene_destroy(s)
        void** src = &s->d_nodes;
        void** dst = &r->d_nodes;
        cudaError_t ce = cudaSetDevice(e->device);
        for (int k = 0; k < 10 && ce == cudaSuccess; ++k) {
            r->nbytes[k] = s->nbytes[k];
            if (!src[k] || s->nbytes[k] == 0) continue;
            ce = cudaMallocAsync(&dst[k], s->nbytes[k], e->render_stream);
            if (ce == cudaSuccess) ce = cudaMemcpyPeerAsync(dst[k], e->device, src[k], ctx->device, s->nbytes[k], e->render_stream);
        }
        if (ce == cudaSuccess) ce = cudaStreamSynchronize(e->render_stream);
        if (ce != cudaSuccess) {
            cudaSetDevice(ctx->device);
            rayca_cuda_scene_destroy(s);
            return fail(ce == cudaErrorMemoryAllocation ? RAYCA_E_OOM : RAYCA_E_CUDA, std::string("replicating the scene: ") + cudaGetErrorString(ce));
        }
        r->fast_node_count = s->fast_node_count;
        r->max_stack = s->max_stack;
        r->view = s->view;
        r->view.nodes = (const float4*)r->d_nodes;
        r->view.tri_pos = (const float4*)r->d_tri_pos;
        r->view.tri_attr = (const float4*)r->d_tri_attr;
        r->view.materials = (const GpuMaterial*)r->d_materials;
        r->view.images = (const GpuImage*)r->d_images;
        r->view.texels = (const unsigned char*)r->d_texels;
        r->view.lights = (const GpuLight*)r->d_lights;
        r->view.blas_prim_offset = (const uint32_t*)r->d_blas_prim_offset;
        r->view.fast_nodes = (const float4*)r->d_fast_nodes;
        r->view.fast_tri = (const float4*)r->d_fast_tri;
    }
    if (!ctx->extra.empty()) CUDA_TRY(cudaSetDevice(ctx->device));
    if (cudaError_t ce = cudaStreamSynchronize(ctx->render_stream); ce != cudaSuccess) { // the arrays may now be read from any stream
        rayca_cuda_scene_destroy(s);
        return fail(RAYCA_E_CUDA, std::string("scene_create: ") + cudaGetErrorString(ce));
    }
    double t2 = now_ms();
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        stats->build_ms = t1 - t0;
        stats->upload_ms = t2 - t1;
    }
    *out = s;
    return RAYCA_OK;
}

} // extern "C"

namespace {
// One launch over image rows [row_begin, row_end) (unsharded) or over this shard's tiles.  `first`
// resets the counters and records the start event, `last` records the end event, so that a frame
// rendered as several row bands still reports one kernel time and one set of ray counts.
struct Band {
    uint32_t row_begin, row_end;
    bool first, last;
};
struct FrameTarget {}; // shared-frame launch (rayca_cuda_render_to_frame): pixels at their image row
struct AccumTarget {   // progressive accumulation (rayca_cuda_render_accumulate)
    float4* plane;     // width x height running sums
    uint32_t first, count; // samples [first, first + count) of config->samples_per_pixel
};
int enqueue_render(rayca_ctx* ctx, rayca_scene* scene, const rayca_config* config, const rayca_options* options, uint32_t width,
                   uint32_t height, void* d_rgba8, const rayca_aux_out* d_aux, cudaStream_t st, const Band* band, rayca_stats* stats,
                   const FrameTarget* target = nullptr, const AccumTarget* accum = nullptr);
} // namespace

extern "C" {

int rayca_cuda_render_device(rayca_ctx* ctx, rayca_scene* scene, const rayca_config* config, const rayca_options* options,
                             uint32_t width, uint32_t height, void* d_rgba8, const rayca_aux_out* d_aux, void* cuda_stream,
                             rayca_stats* stats) {
    return enqueue_render(ctx, scene, config, options, width, height, d_rgba8, d_aux, (cudaStream_t)cuda_stream, nullptr, stats);
}

} // extern "C"

namespace {
int enqueue_render(rayca_ctx* ctx, rayca_scene* scene, const rayca_config* config, const rayca_options* options, uint32_t width,
                   uint32_t height, void* d_rgba8, const rayca_aux_out* d_aux, cudaStream_t st, const Band* band, rayca_stats* stats,
                   const FrameTarget* target, const AccumTarget* accum) {
    if (!ctx || !scene || !config || !d_rgba8 || width == 0 || height == 0) return fail(RAYCA_E_INVALID_ARG, "null / zero argument");
    if (scene->ctx != ctx) return fail(RAYCA_E_INVALID_ARG, "scene belongs to another context (or is host-only)");
    if (!host_of(*scene).has_camera) return fail(RAYCA_E_NO_CAMERA, "no camera node in the scene");
    if ((uint64_t)width * height >= 0xFFFFFFFFull) return fail(RAYCA_E_UNSUPPORTED, "image too large");
    int rc = validate_config(*config, *scene);
    if (rc != RAYCA_OK) return rc;
    CUDA_TRY(cudaSetDevice(ctx->device));

    RenderParams P{};
    P.scene = scene->view;
    P.cfg = *config;
    P.seed = options ? options->rng_seed : 0;
    P.width = width;
    P.height = height;
    P.shard_count = (options && options->shard_count > 1) ? options->shard_count : 1;
    P.shard_index = options ? options->shard_index : 0;
    // RAYCA_FLAG_SHARD_GROUPS: shard by 16x8 pixel groups in a diagonal checkerboard instead of row tiles (better balance
    // between shards: every shard samples the whole image); pixels must then go to their final rows
    const bool by_groups = P.shard_count > 1 && options && (options->flags & RAYCA_FLAG_SHARD_GROUPS);
    if (by_groups) {
        if (!(target || (options->flags & RAYCA_FLAG_DIRECT_ROWS)) || band || d_aux || accum || (options->flags & RAYCA_FLAG_PERSISTENT))
            return fail(RAYCA_E_INVALID_ARG, "RAYCA_FLAG_SHARD_GROUPS needs RAYCA_FLAG_DIRECT_ROWS (or a shared frame), whole frames, no aux planes, the static grid");
        if (P.shard_index >= P.shard_count) return fail(RAYCA_E_INVALID_ARG, "shard_index >= shard_count");
        P.gshard_index = P.shard_index;
        P.gshard_count = P.shard_count;
        P.shard_index = 0;
        P.shard_count = 1;
    }
    P.tile_rows = (options && options->tile_rows) ? options->tile_rows : 8;
    if (P.tile_rows % 8 != 0) return fail(RAYCA_E_INVALID_ARG, "tile_rows must be a multiple of 8");
    if (P.shard_index >= P.shard_count) return fail(RAYCA_E_INVALID_ARG, "shard_index >= shard_count");
    uint32_t rows = height;
    if (P.shard_count > 1) {
        uint32_t tiles = (height + P.tile_rows - 1) / P.tile_rows;
        uint32_t per = (tiles + P.shard_count - 1) / P.shard_count; // every shard writes `per` tiles (tail tiles may be empty)
        rows = per * P.tile_rows;
    }
    P.local_rows = rows;
    P.row_begin = 0;
    P.row_end = height;
    if (band) {
        if (P.shard_count > 1 || band->row_begin % 8 != 0 || band->row_begin >= band->row_end || band->row_end > height)
            return fail(RAYCA_E_INVALID_ARG, "bad row band");
        P.row_begin = band->row_begin;
        P.row_end = band->row_end;
        rows = band->row_end - band->row_begin;
    }
    const bool first = !band || band->first, last = !band || band->last;
    float fw = (float)width, fh = (float)height;
    P.inv_w = 1.0f / fw;
    P.inv_h = 1.0f / fh;
    P.aspect = fw / fh;
    P.strate_count = sqrtf((float)config->samples_per_pixel); // scene.rs:125-127
    P.offset = 0.5f / P.strate_count;
    P.step = 1.0f / P.strate_count;
    P.strate_u = (uint32_t)P.strate_count;
    P.inv_gamma = 1.0f / config->gamma;
    P.sample_first = 0;
    P.sample_count = config->samples_per_pixel;
    P.sample_div = (float)config->samples_per_pixel;
    if (accum) {
        if (P.shard_count > 1 || band) return fail(RAYCA_E_INVALID_ARG, "accumulation renders whole frames");
        if (accum->count == 0 || (uint64_t)accum->first + accum->count > config->samples_per_pixel)
            return fail(RAYCA_E_INVALID_ARG, "sample range beyond config->samples_per_pixel");
        P.accum = accum->plane;
        P.accum_load = accum->first > 0 ? 1u : 0u;
        P.sample_first = accum->first;
        P.sample_count = accum->count;
        P.sample_div = (float)(accum->first + accum->count); // the mean of the samples so far; the last call divides by spp
    }
    P.rgba8 = (unsigned char*)d_rgba8;
    if (d_aux) {
        P.hit_blas = d_aux->hit_blas;
        P.hit_prim = d_aux->hit_prim;
        P.hit_src = d_aux->hit_src;
        P.depth = d_aux->depth;
        P.uv = d_aux->uv;
        P.rgba_f32 = d_aux->rgba_f32;
    }
    P.counters = ctx->d_counters;
    if (target || (options && (options->flags & RAYCA_FLAG_DIRECT_ROWS))) P.direct_rows = 1;
    const uint32_t mode = options ? options->traversal : (uint32_t)RAYCA_TRAVERSAL_ORDERED;
    if (mode > RAYCA_TRAVERSAL_REFERENCE_ORDERED) return fail(RAYCA_E_INVALID_ARG, "unknown traversal mode");
    if (mode != RAYCA_TRAVERSAL_ORDERED && host_of(*scene).ref_stack > (uint32_t)kStackSize)
        return fail(RAYCA_E_UNSUPPORTED, "reference tree deeper than the traversal stack: traversal modes 1 / 2 cannot walk it (mode 0 can)");
    if (mode != RAYCA_TRAVERSAL_ORDERED && host_of(*scene).has_spheres)
        return fail(RAYCA_E_UNSUPPORTED, "the reference-tree walks (traversal 1 / 2) handle triangle scenes only");
    const bool count = options && (options->flags & RAYCA_FLAG_COUNT_WALK); // opt-in: walk counters cost registers
    const bool persistent = options && (options->flags & RAYCA_FLAG_PERSISTENT);

    if (first) CUDA_TRY(cudaMemsetAsync(ctx->d_counters, 0, kCounterSlots * sizeof(unsigned long long), st));
    else if (persistent) CUDA_TRY(cudaMemsetAsync(ctx->d_counters + 5, 0, sizeof(unsigned long long), st)); // next-group counter
    dim3 grid((width + kGroupW - 1) / kGroupW, (rows + kGroupH - 1) / kGroupH);
    if (persistent) grid = dim3((width + 15) / 16, (rows + 7) / 8); // the persistent form pulls 16x8 groups
    if (by_groups) {
        P.groups_x_full = grid.x;
        grid.x = (grid.x + P.gshard_count - 1) / P.gshard_count;
    }
    P.groups_x = grid.x;
    P.tile_count = grid.x * grid.y;
    if (first) CUDA_TRY(cudaEventRecord(ctx->ev0, st));
    LaunchSel sel{mode, count, persistent, false, host_of(*scene).has_spheres, ctx->sm_count};
    uint32_t integrator = config->integrator;
    if (integrator == RAYCA_INTEGRATOR_PATHTRACER) { // an indirect ray needs depth < limit (pathtracer.rs:94)
        const bool mis = config->direct_sampler == RAYCA_SAMPLER_MIS;
        const bool nee = config->direct_sampler == RAYCA_SAMPLER_NEE || mis;
        sel.bounce = config->max_depth > (nee ? 1u : 0u);
        // hemisphere bounces, roulette and quad lights under NEE run in the general path tracer (render.cuh)
        bool general = mis || config->russian_roulette != 0 || (sel.bounce && config->indirect_sampler != RAYCA_SAMPLER_COSINE) ||
                       (sel.bounce && config->light_samples > 1);
        if (nee)
            for (const GpuLight& l : host_of(*scene).lights) general = general || l.kind == RAYCA_LIGHT_QUAD;
        if (general) integrator = kInternalPathGeneral;
    }
    uint32_t launches = 1;
    const bool no_wave = options && (options->flags & RAYCA_FLAG_NO_WAVEFRONT);
    if (!no_wave && !by_groups && wave_supported(P, rows, mode, count, persistent, integrator == kInternalPathGeneral, sel.bounce)) {
        // bounce rays: wavefront stages with the live rays sorted between depths (wavefront.cu)
        std::string werr;
        rc = wave_render(P, rows, sel.spheres, ctx->wave, st, &launches, werr);
        if (rc != RAYCA_OK) return fail(rc, werr);
    } else
    switch (integrator) {
        case kInternalPathGeneral: launch_integrator<(int)kInternalPathGeneral>(P, grid, sel, st); break;
        case RAYCA_INTEGRATOR_PATHTRACER: launch_integrator<RAYCA_INTEGRATOR_PATHTRACER>(P, grid, sel, st); break;
        case RAYCA_INTEGRATOR_RAYTRACER: launch_integrator<RAYCA_INTEGRATOR_RAYTRACER>(P, grid, sel, st); break;
        case RAYCA_INTEGRATOR_SCRATCHER: launch_integrator<RAYCA_INTEGRATOR_SCRATCHER>(P, grid, sel, st); break;
        case RAYCA_INTEGRATOR_ANALYTIC_DIRECT: launch_integrator<RAYCA_INTEGRATOR_ANALYTIC_DIRECT>(P, grid, sel, st); break;
        case RAYCA_INTEGRATOR_DIRECT: launch_integrator<RAYCA_INTEGRATOR_DIRECT>(P, grid, sel, st); break;
        default: launch_integrator<RAYCA_INTEGRATOR_FLAT>(P, grid, sel, st); break;
    }
    CUDA_TRY(cudaGetLastError());
    if (last) CUDA_TRY(cudaEventRecord(ctx->ev1, st));
    // primary rays of this shard: pixels it owns x spp
    uint64_t own_rows = 0;
    uint64_t own_pixels = 0;
    if (by_groups) {
        const uint32_t gy_n = (height + kGroupH - 1) / kGroupH;
        for (uint32_t gy = 0; gy < gy_n; ++gy) {
            const uint32_t h = std::min(height, (gy + 1) * kGroupH) - gy * kGroupH;
            for (uint32_t gx = (P.gshard_index + P.gshard_count - gy % P.gshard_count) % P.gshard_count; gx < P.groups_x_full; gx += P.gshard_count)
                own_pixels += (uint64_t)h * (std::min(width, (gx + 1) * kGroupW) - gx * kGroupW);
        }
    } else if (P.shard_count > 1) {
        uint32_t tiles = (height + P.tile_rows - 1) / P.tile_rows;
        for (uint32_t t = P.shard_index; t < tiles; t += P.shard_count) {
            uint32_t y0 = t * P.tile_rows, y1 = std::min(height, y0 + P.tile_rows);
            own_rows += y1 - y0;
        }
    } else {
        own_rows = height;
    }
    if (!by_groups) own_pixels = own_rows * width;
    ctx->last_primary = own_pixels * (uint64_t)P.sample_count;
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        stats->kernel_launches = launches;
        stats->rays_primary = ctx->last_primary;
    }
    ctx->last_launches = first ? launches : ctx->last_launches + launches;
    return RAYCA_OK;
}
} // namespace

extern "C" {

} // extern "C"

namespace {
// Streaming 128-bit reads over a buffer, grid-stride, every load kept alive through a XOR fold.  The loads are
// ld.global.cg (cached in L2 only) and every repetition rotates the block -> slice mapping, so L1 can serve none
// of them: a first version with __ldg re-read a per-block slice that fitted L1 and reported 25 TB/s, above the
// 148 SMs x 64 B/clk L2 -> SM limit.
__global__ void __launch_bounds__(256) probe_read_kernel(const uint4* __restrict__ p, size_t n, uint32_t reps, uint32_t* sink) {
    uint32_t acc = 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (uint32_t r = 0; r < reps; ++r) {
        const size_t blk = (blockIdx.x + (size_t)r * 37u) % gridDim.x;
        for (size_t i = blk * blockDim.x + threadIdx.x; i < n; i += stride) {
            const uint4 v = __ldcg(p + i);
            acc ^= v.x ^ v.y ^ v.z ^ v.w;
        }
    }
    if (acc == 0x9E3779B9u) *sink = acc; // practically never: keeps the loads observable
}
} // namespace

extern "C" {

int rayca_cuda_probe_read_bandwidth(rayca_ctx* ctx, size_t bytes, uint32_t reps, double* gb_per_s) {
    if (!ctx || !gb_per_s || bytes < 4096 || reps == 0) return fail(RAYCA_E_INVALID_ARG, "null / zero argument");
    CUDA_TRY(cudaSetDevice(ctx->device));
    void* buf = nullptr;
    uint32_t* sink = nullptr;
    CUDA_TRY(cudaMalloc(&buf, bytes));
    CUDA_TRY(cudaMalloc((void**)&sink, 4));
    CUDA_TRY(cudaMemset(buf, 1, bytes));
    const size_t n = bytes / sizeof(uint4);
    const int grid = ctx->sm_count * 8;
    probe_read_kernel<<<grid, 256>>>((const uint4*)buf, n, 2, sink); // warm-up: pulls the buffer into L2 if it fits
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaEventCreate(&e0));
    CUDA_TRY(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int k = 0; k < 5; ++k) {
        CUDA_TRY(cudaEventRecord(e0));
        probe_read_kernel<<<grid, 256>>>((const uint4*)buf, n, reps, sink);
        CUDA_TRY(cudaEventRecord(e1));
        CUDA_TRY(cudaEventSynchronize(e1));
        float ms = 0.0f;
        CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
        best = ms < best ? ms : best;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(buf);
    cudaFree(sink);
    *gb_per_s = (double)n * sizeof(uint4) * reps / (best * 1e-3) / 1e9;
    return RAYCA_OK;
}

} // extern "C"

struct rayca_frame {
    rayca_ctx* ctx = nullptr;
    unsigned char* base = nullptr; // flags | pixels
    uint32_t width = 0, height = 0;
    bool imported = false;
    int* timed_out = nullptr; // pinned, device-visible
    unsigned long long timeout_ns = 2000000000ull; // rayca_cuda_frame_set_timeout
};

namespace {
constexpr size_t kFrameFlagBytes = 64 * sizeof(unsigned long long);
constexpr uint32_t kConsumedSlot = 32; // slots 0..31: "rank r has stored frame id"; slot 32: "owner has consumed frame id"
// Publishes `value` in one flag slot (local or peer memory).  Launched on the stream right after the trace
// kernel: a kernel's stores, peer stores included, are performed before the next kernel of the stream starts,
// and the system fence orders this thread's flag store after them for an observer on another GPU.
__global__ void frame_signal_kernel(volatile unsigned long long* slot, unsigned long long value) {
    __threadfence_system();
    *slot = value;
}
// one warp: lane r polls slot first+r; gives up after `timeout_ns` of WALL time (%globaltimer) so that a dead peer
// cannot hang the GPU.  A time-out is sticky in `timed_out` until rayca_cuda_frame_clear_timeout; the frame's
// contents are undefined from then on (a rank may still be storing into it).
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__global__ void frame_wait_kernel(const volatile unsigned long long* flags, uint32_t ranks, unsigned long long frame_id, int* timed_out,
                                  unsigned long long timeout_ns) {
    const uint32_t r = threadIdx.x;
    if (r < ranks) {
        const unsigned long long t0 = global_ns();
        while (flags[r] < frame_id) {
            if (global_ns() - t0 > timeout_ns) {
                *timed_out = 1;
                break;
            }
            __nanosleep(200);
        }
    }
    __threadfence_system();
}
} // namespace

extern "C" {

int rayca_cuda_frame_create(rayca_ctx* ctx, uint32_t width, uint32_t height, rayca_frame** out) {
    if (!ctx || !out || width == 0 || height == 0) return fail(RAYCA_E_INVALID_ARG, "null / zero argument");
    CUDA_TRY(cudaSetDevice(ctx->device));
    rayca_frame* f = new (std::nothrow) rayca_frame();
    if (!f) return fail(RAYCA_E_OOM, "host allocation");
    f->ctx = ctx;
    f->width = width;
    f->height = height;
    const size_t bytes = kFrameFlagBytes + (size_t)width * height * 4;
    cudaError_t ce = cudaMalloc((void**)&f->base, bytes); // a plain cudaMalloc block: IPC handles cover whole allocations
    if (ce == cudaSuccess) ce = cudaMemset(f->base, 0, bytes);
    if (ce == cudaSuccess) ce = cudaHostAlloc((void**)&f->timed_out, sizeof(int), cudaHostAllocMapped);
    if (ce != cudaSuccess) {
        rayca_cuda_frame_destroy(f); // null-safe per member
        return fail(ce == cudaErrorMemoryAllocation ? RAYCA_E_OOM : RAYCA_E_CUDA, std::string("frame_create: ") + cudaGetErrorString(ce));
    }
    *f->timed_out = 0;
    *out = f;
    return RAYCA_OK;
}

int rayca_cuda_frame_export(rayca_frame* f, uint8_t handle[RAYCA_IPC_HANDLE_BYTES]) {
    if (!f || !handle || f->imported) return fail(RAYCA_E_INVALID_ARG, "export needs the owner's frame");
    static_assert(sizeof(cudaIpcMemHandle_t) == RAYCA_IPC_HANDLE_BYTES, "CUDA IPC handle size");
    CUDA_TRY(cudaSetDevice(f->ctx->device));
    cudaIpcMemHandle_t h;
    CUDA_TRY(cudaIpcGetMemHandle(&h, f->base));
    std::memcpy(handle, &h, sizeof(h));
    return RAYCA_OK;
}

int rayca_cuda_frame_import(rayca_ctx* ctx, const uint8_t handle[RAYCA_IPC_HANDLE_BYTES], uint32_t width, uint32_t height,
                            rayca_frame** out) {
    if (!ctx || !handle || !out || width == 0 || height == 0) return fail(RAYCA_E_INVALID_ARG, "null / zero argument");
    CUDA_TRY(cudaSetDevice(ctx->device));
    rayca_frame* f = new (std::nothrow) rayca_frame();
    if (!f) return fail(RAYCA_E_OOM, "host allocation");
    f->ctx = ctx;
    f->width = width;
    f->height = height;
    f->imported = true;
    if (cudaError_t ce = cudaHostAlloc((void**)&f->timed_out, sizeof(int), cudaHostAllocMapped); ce != cudaSuccess) {
        delete f;
        return fail(RAYCA_E_CUDA, std::string("frame_import: ") + cudaGetErrorString(ce));
    }
    *f->timed_out = 0;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof(h));
    cudaError_t e = cudaIpcOpenMemHandle((void**)&f->base, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        cudaFreeHost(f->timed_out);
        delete f;
        return fail(RAYCA_E_CUDA, std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e));
    }
    *out = f;
    return RAYCA_OK;
}

void* rayca_cuda_frame_pixels(rayca_frame* f) { return f ? f->base + kFrameFlagBytes : nullptr; }

void rayca_cuda_frame_destroy(rayca_frame* f) {
    if (!f) return;
    cudaSetDevice(f->ctx->device);
    if (f->imported) {
        if (f->base) cudaIpcCloseMemHandle(f->base);
    } else {
        cudaFree(f->base);
    }
    if (f->timed_out) cudaFreeHost(f->timed_out);
    delete f;
}

int rayca_cuda_render_to_frame(rayca_ctx* ctx, rayca_scene* scene, const rayca_config* config, const rayca_options* options,
                               rayca_frame* frame, uint64_t frame_id, void* cuda_stream) {
    if (!frame || frame->ctx != ctx) return fail(RAYCA_E_INVALID_ARG, "frame belongs to another context");
    const uint32_t rank = (options && options->shard_count > 1) ? options->shard_index : 0;
    if (rank >= kConsumedSlot) return fail(RAYCA_E_INVALID_ARG, "at most 32 ranks share a frame");
    FrameTarget t{};
    int rc = enqueue_render(ctx, scene, config, options, frame->width, frame->height, frame->base + kFrameFlagBytes, nullptr,
                            (cudaStream_t)cuda_stream, nullptr, nullptr, &t);
    if (rc != RAYCA_OK) return rc;
    frame_signal_kernel<<<1, 1, 0, (cudaStream_t)cuda_stream>>>(reinterpret_cast<volatile unsigned long long*>(frame->base) + rank, frame_id);
    CUDA_TRY(cudaGetLastError());
    return RAYCA_OK;
}

int rayca_cuda_frame_release(rayca_ctx* ctx, rayca_frame* frame, uint64_t frame_id, void* cuda_stream) {
    if (!ctx || !frame || frame->imported) return fail(RAYCA_E_INVALID_ARG, "release needs the owner's frame");
    CUDA_TRY(cudaSetDevice(ctx->device));
    frame_signal_kernel<<<1, 1, 0, (cudaStream_t)cuda_stream>>>(reinterpret_cast<volatile unsigned long long*>(frame->base) + kConsumedSlot, frame_id);
    CUDA_TRY(cudaGetLastError());
    return RAYCA_OK;
}

int rayca_cuda_frame_acquire(rayca_ctx* ctx, rayca_frame* frame, uint64_t frame_id, void* cuda_stream) {
    if (!ctx || !frame) return fail(RAYCA_E_INVALID_ARG, "null argument");
    if (frame_id == 0) return RAYCA_OK; // nothing has been stored yet
    CUDA_TRY(cudaSetDevice(ctx->device));
    int* d_flag = nullptr;
    CUDA_TRY(cudaHostGetDevicePointer((void**)&d_flag, frame->timed_out, 0));
    frame_wait_kernel<<<1, 32, 0, (cudaStream_t)cuda_stream>>>(
        reinterpret_cast<const volatile unsigned long long*>(frame->base) + kConsumedSlot, 1, frame_id, d_flag, frame->timeout_ns);
    CUDA_TRY(cudaGetLastError());
    return RAYCA_OK;
}

int rayca_cuda_frame_wait(rayca_ctx* ctx, rayca_frame* frame, uint32_t rank_count, uint64_t frame_id, void* cuda_stream) {
    if (!ctx || !frame || frame->imported || rank_count == 0 || rank_count > 32) return fail(RAYCA_E_INVALID_ARG, "wait needs the owner's frame and 1..32 ranks");
    CUDA_TRY(cudaSetDevice(ctx->device));
    int* d_flag = nullptr;
    CUDA_TRY(cudaHostGetDevicePointer((void**)&d_flag, frame->timed_out, 0));
    frame_wait_kernel<<<1, 32, 0, (cudaStream_t)cuda_stream>>>(reinterpret_cast<const volatile unsigned long long*>(frame->base), rank_count,
                                                              frame_id, d_flag, frame->timeout_ns);
    CUDA_TRY(cudaGetLastError());
    return RAYCA_OK;
}

int rayca_cuda_frame_timed_out(rayca_frame* f) { return (f && f->timed_out) ? *f->timed_out : 0; }

int rayca_cuda_frame_set_timeout(rayca_frame* f, double milliseconds) {
    if (!f || !(milliseconds > 0.0) || milliseconds > 3.6e6) return fail(RAYCA_E_INVALID_ARG, "time-out must be in (0, 1 h]");
    f->timeout_ns = (unsigned long long)(milliseconds * 1.0e6);
    return RAYCA_OK;
}

void rayca_cuda_frame_clear_timeout(rayca_frame* f) {
    if (f && f->timed_out) *f->timed_out = 0;
}

int rayca_cuda_frame_sync(rayca_ctx* ctx, rayca_frame* frame, void* cuda_stream) {
    if (!ctx || !frame) return fail(RAYCA_E_INVALID_ARG, "null argument");
    CUDA_TRY(cudaSetDevice(ctx->device));
    CUDA_TRY(cudaStreamSynchronize((cudaStream_t)cuda_stream));
    if (*frame->timed_out)
        return fail(RAYCA_E_TIMEOUT, "shared frame: a flag wait gave up (a rank is late or dead); the frame's contents are undefined until the "
                                     "ranks are back in step - rayca_cuda_frame_clear_timeout re-arms it");
    return RAYCA_OK;
}

int rayca_cuda_host_alloc(size_t bytes, void** out) {
    if (!out || bytes == 0) return fail(RAYCA_E_INVALID_ARG, "null / zero argument");
    CUDA_TRY(cudaHostAlloc(out, bytes, cudaHostAllocPortable | cudaHostAllocMapped)); // every device may store into it
    return RAYCA_OK;
}

void rayca_cuda_host_free(void* p) {
    if (p) cudaFreeHost(p);
}

int rayca_cuda_host_register(void* p, size_t bytes, void** device_view) {
    if (!p || bytes == 0) return fail(RAYCA_E_INVALID_ARG, "null / zero argument");
    CUDA_TRY(cudaHostRegister(p, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped));
    if (device_view) {
        cudaError_t e = cudaHostGetDevicePointer(device_view, p, 0);
        if (e != cudaSuccess) {
            cudaHostUnregister(p);
            return fail(RAYCA_E_CUDA, cudaGetErrorString(e));
        }
    }
    return RAYCA_OK;
}

int rayca_cuda_host_unregister(void* p) {
    if (!p) return fail(RAYCA_E_INVALID_ARG, "null argument");
    CUDA_TRY(cudaHostUnregister(p));
    return RAYCA_OK;
}

int rayca_cuda_last_counters(rayca_ctx* ctx, rayca_stats* stats) {
    if (!ctx || !stats) return fail(RAYCA_E_INVALID_ARG, "null argument");
    CUDA_TRY(cudaSetDevice(ctx->device));
    unsigned long long c[kCounterSlots];
    CUDA_TRY(cudaEventSynchronize(ctx->ev1));
    CUDA_TRY(cudaMemcpy(c, ctx->d_counters, sizeof(c), cudaMemcpyDeviceToHost));
    float ms = 0.0f;
    CUDA_TRY(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    stats->trace_ms = ms;
    stats->rays_primary = ctx->last_primary;
    stats->rays_shadow = c[0];
    stats->rays_secondary = c[1];
    stats->node_bytes = c[2] * 32ull;
    stats->triangle_bytes = c[3] * 48ull + c[4] * 16ull;
    stats->kernel_launches = ctx->last_launches;
    return RAYCA_OK;
}

} // extern "C"

namespace {
// rayca_cuda_render on a multi-device context: ONE host thread, G = 1 + extra devices.  The frame's 8-row tiles are
// interleaved over the devices (tile t -> device t mod G, like the one-process-per-GPU path: sky and geometry spread
// evenly), every device's launch is enqueued on its own stream before any of them is waited for, and
//   * a pinned / registered caller buffer is filled by the kernels' own stores (RAYCA_FLAG_DIRECT_ROWS): G PCIe links
//     in parallel, no staging plane, no copy;
//   * a pageable caller buffer is filled through a pinned staging frame the context owns: the kernels store into that,
//     the calling thread's OpenMP team copies it out.
// This is the reference's fan-out (`par_iter_mut` over rows and pixels inside one `draw` call,
// rayca-soft/src/scene.rs:117-122) with devices in place of rayon workers.
int render_multi(rayca_ctx* ctx, rayca_scene* scene, const rayca_config* config, const rayca_options* options, uint32_t width,
                 uint32_t height, uint8_t* rgba8_out, rayca_stats* stats) {
    const uint32_t G = 1u + (uint32_t)ctx->extra.size();
    if (scene->replicas.size() + 1 != G) return fail(RAYCA_E_INVALID_ARG, "scene was not created on this multi-device context");
    constexpr uint32_t kTile = 8;
    const size_t frame_bytes = (size_t)width * height * 4;
    void* host_view = nullptr;
    {
        cudaPointerAttributes at{};
        if (cudaPointerGetAttributes(&at, rgba8_out) == cudaSuccess && at.type == cudaMemoryTypeHost) host_view = at.devicePointer;
        cudaGetLastError();
        const char* e = std::getenv("RAYCA_ZERO_COPY");
        if (e && std::atoi(e) == 0) host_view = nullptr; // A/B runs and tests: force the staged path
    }
    // A pageable caller buffer (a Rust Vec<u8>, a numpy array) cannot be stored into by the devices, and strided
    // device-to-pageable copies are slow (measured: 8 GPUs 2.7 Grays/s against 5.2 on ONE device).  So the devices
    // store into a pinned staging frame the context owns (allocated once, reused across calls) and the calling
    // thread's OpenMP team copies it to the caller's buffer.
    const bool staged = host_view == nullptr;
    if (staged) {
        if (frame_bytes > ctx->h_stage_bytes) {
            CUDA_TRY(cudaSetDevice(ctx->device));
            if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
            ctx->h_stage = nullptr;
            ctx->h_stage_bytes = 0;
            CUDA_TRY(cudaHostAlloc((void**)&ctx->h_stage, frame_bytes, cudaHostAllocPortable | cudaHostAllocMapped));
            ctx->h_stage_bytes = frame_bytes;
        }
        CUDA_TRY(cudaHostGetDevicePointer(&host_view, ctx->h_stage, 0));
    }
    double t0 = now_ms();
    int rc = RAYCA_OK;
    for (uint32_t r = 0; r < G && rc == RAYCA_OK; ++r) {
        rayca_ctx* c = r == 0 ? ctx : ctx->extra[r - 1];
        rayca_scene* sc = r == 0 ? scene : scene->replicas[r - 1];
        CUDA_TRY(cudaSetDevice(c->device));
        rayca_options o{};
        if (options) o = *options;
        o.shard_index = r;
        o.shard_count = G;
        o.tile_rows = kTile;
        o.flags |= RAYCA_FLAG_DIRECT_ROWS; // row tiles: measured as fast as the group checkerboard and kinder to the caches
        rc = enqueue_render(c, sc, config, &o, width, height, host_view, nullptr, c->render_stream, nullptr, nullptr);
    }
    // wait for every device even after a failure: nothing may still be writing into the caller's buffer on return
    rayca_stats total{};
    for (uint32_t r = 0; r < G; ++r) {
        rayca_ctx* c = r == 0 ? ctx : ctx->extra[r - 1];
        cudaSetDevice(c->device);
        cudaError_t ce = cudaStreamSynchronize(c->render_stream);
        if (ce != cudaSuccess && rc == RAYCA_OK) rc = fail(RAYCA_E_CUDA, std::string("multi-device render: ") + cudaGetErrorString(ce));
        if (rc == RAYCA_OK && stats) {
            rayca_stats one{};
            int rc2 = rayca_cuda_last_counters(c, &one);
            if (rc2 != RAYCA_OK) rc = rc2;
            total.rays_primary += one.rays_primary;
            total.rays_shadow += one.rays_shadow;
            total.rays_secondary += one.rays_secondary;
            total.node_bytes += one.node_bytes;
            total.triangle_bytes += one.triangle_bytes;
            total.trace_ms = std::max(total.trace_ms, one.trace_ms); // the devices run side by side
            total.kernel_launches += one.kernel_launches;
        }
    }
    cudaSetDevice(ctx->device);
    if (rc != RAYCA_OK) return rc;
    if (staged) { // rows in parallel: one thread moves ~10 GB/s, the frame is 8 MB
        const size_t row = (size_t)width * 4;
        const unsigned char* src = ctx->h_stage;
#pragma omp parallel for schedule(static)
        for (int64_t y = 0; y < (int64_t)height; ++y) std::memcpy(rgba8_out + (size_t)y * row, src + (size_t)y * row, row);
    }
    if (stats) {
        total.readback_ms = std::max(0.0, (now_ms() - t0) - total.trace_ms);
        *stats = total;
    }
    return RAYCA_OK;
}
} // namespace

extern "C" {

int rayca_cuda_render(rayca_ctx* ctx, rayca_scene* scene, const rayca_config* config, const rayca_options* options, uint32_t width,
                      uint32_t height, uint8_t* rgba8_out, const rayca_aux_out* aux, rayca_stats* stats) {
    if (!ctx || !rgba8_out || width == 0 || height == 0) return fail(RAYCA_E_INVALID_ARG, "null / zero argument");
    CUDA_TRY(cudaSetDevice(ctx->device));
    const bool sharded = options && options->shard_count > 1;
    // multi-device context: the whole frame, sharded over the devices from this one call; parity planes (aux) and
    // explicitly sharded calls stay on the context's first device
    if (!ctx->extra.empty() && scene && !aux && !sharded && config) return render_multi(ctx, scene, config, options, width, height, rgba8_out, stats);
    if (sharded) { // one rank's tiles of a frame that several ranks fill: only straight into pinned memory, at the final rows
        cudaPointerAttributes at{};
        const bool pinned = cudaPointerGetAttributes(&at, rgba8_out) == cudaSuccess && at.type == cudaMemoryTypeHost;
        cudaGetLastError();
        if (!pinned || aux || !(options->flags & RAYCA_FLAG_DIRECT_ROWS))
            return fail(RAYCA_E_INVALID_ARG, "a sharded render to host memory needs a pinned frame, RAYCA_FLAG_DIRECT_ROWS and no aux planes");
    }
    const size_t n = (size_t)width * height;
    // scratch layout: rgba8 | hit_blas | hit_prim | hit_src | depth | uv | rgba_f32, each plane 256 B aligned
    auto al = [](size_t b) { return (b + 255) & ~size_t(255); };
    size_t need = al(n * 4) + (aux ? 4 * al(n * 4) + al(n * 8) + al(n * 16) : 0);
    if (need > ctx->scratch_bytes) {
        cudaFree(ctx->d_scratch);
        ctx->d_scratch = nullptr;
        ctx->scratch_bytes = 0;
        CUDA_TRY(cudaMalloc(&ctx->d_scratch, need));
        ctx->scratch_bytes = need;
    }
    unsigned char* base = (unsigned char*)ctx->d_scratch;
    rayca_aux_out d{};
    if (aux) {
        unsigned char* p = base + al(n * 4);
        if (aux->hit_blas) d.hit_blas = (uint32_t*)p;
        p += al(n * 4);
        if (aux->hit_prim) d.hit_prim = (uint32_t*)p;
        p += al(n * 4);
        if (aux->hit_src) d.hit_src = (uint32_t*)p;
        p += al(n * 4);
        if (aux->depth) d.depth = (float*)p;
        p += al(n * 4);
        if (aux->uv) d.uv = (float*)p;
        p += al(n * 8);
        if (aux->rgba_f32) d.rgba_f32 = (float*)p;
    }
    // Row bands: the kernel of band k+1 runs while band k is copied to the host on the side stream, so
    // the PCIe read-back of the frame hides behind the trace instead of following it.  Bands shrink towards the
    // end of the frame: only the last copy is exposed.  A caller buffer
    // from rayca_cuda_host_alloc (pinned) makes the copies truly asynchronous; pageable memory still
    // works (the runtime stages it) but serialises the copies on the calling thread.
    rayca_stats local{};
    const uint32_t nb_max = (uint32_t)(sizeof(ctx->band_done) / sizeof(ctx->band_done[0]));
    // measured on B200, 1080p atrium, pinned output (tools/band_ab.py): 1 band 1.086 ms, 4 equal bands 1.095 (every
    // launch adds its own tail), 3 tapered bands 1.010, 4 tapered 1.028, 8 tapered 1.165
    uint32_t nb = height >= 512 ? 3 : (height >= 128 ? 2 : 1);
    bool taper = true;
    if (const char* e = std::getenv("RAYCA_BANDS")) nb = (uint32_t)std::max(1, std::atoi(e)); // A/B runs
    if (const char* e = std::getenv("RAYCA_BAND_TAPER")) taper = std::atoi(e) != 0;
    if (nb > nb_max) nb = nb_max;
    // band k covers rows [edge[k], edge[k+1]), multiples of 8; tapered: heights proportional to nb - k, so that the
    // last copy (the only one nothing hides) is short
    uint32_t edge[9] = {0};
    {
        const uint32_t total_w = taper ? nb * (nb + 1) / 2 : nb;
        uint32_t acc_w = 0;
        for (uint32_t k = 0; k < nb; ++k) {
            acc_w += taper ? nb - k : 1;
            uint32_t y = (uint32_t)(((uint64_t)height * acc_w + total_w - 1) / total_w);
            y = (y + 7) & ~7u;
            edge[k + 1] = std::min(height, std::max(y, edge[k]));
        }
        edge[nb] = height;
    }
    double t0 = now_ms();
    int rc = RAYCA_OK;
    uint32_t launched = 0, kernels = 0; // launches of enqueue_render (bands); kernels they launched (wavefront stages count)
    uint32_t band_of[8] = {0};
    // Zero copy: when the caller's buffer is pinned host memory (rayca_cuda_host_alloc, or any cudaHostAlloc /
    // cudaHostRegister'ed range) the device can address it, so ONE launch stores the RGBA8 pixels straight into it
    // over PCIe — posted writes that drain while the rest of the frame is traced: no device staging plane, no
    // copy, no band tails.  RAYCA_ZERO_COPY=0 restores the banded copy for A/B runs.
    void* host_view = nullptr;
    {
        const char* e = std::getenv("RAYCA_ZERO_COPY");
        cudaPointerAttributes at{};
        if ((sharded || !(e && std::atoi(e) == 0)) && cudaPointerGetAttributes(&at, rgba8_out) == cudaSuccess && at.type == cudaMemoryTypeHost)
            host_view = at.devicePointer;
        cudaGetLastError(); // a pageable pointer is reported as an error by older runtimes: not ours to keep
    }
    if (host_view) nb = 0;
    // A path-traced frame with bounces runs as wavefront stages whose later stages store their pixels in SORTED path
    // order: straight into host memory those are scattered 4-byte PCIe writes (measured: 3.2 against 3.9 Grays/s).  Such a
    // frame goes to the device plane and leaves in one copy; unsharded frames only (a shard's tiles still go direct).
    const bool bouncing = config && config->integrator == RAYCA_INTEGRATOR_PATHTRACER && config->samples_per_pixel == 1 &&
                          config->max_depth > (config->direct_sampler == RAYCA_SAMPLER_NONE ? 0u : 1u) &&
                          !(options && (options->flags & RAYCA_FLAG_NO_WAVEFRONT)) && !std::getenv("RAYCA_NO_WAVEFRONT") &&
                          (size_t)width * height >= 600000; // wave_supported's size threshold
    if (host_view && bouncing && !sharded) {
        rc = enqueue_render(ctx, scene, config, options, width, height, base, aux ? &d : nullptr, ctx->render_stream, nullptr, &local);
        if (rc != RAYCA_OK) return rc;
        launched = 1;
        kernels += local.kernel_launches;
        CUDA_TRY(cudaMemcpyAsync(rgba8_out, base, n * 4, cudaMemcpyDeviceToHost, ctx->render_stream));
        CUDA_TRY(cudaStreamSynchronize(ctx->render_stream));
    } else if (host_view) {
        rc = enqueue_render(ctx, scene, config, options, width, height, host_view, aux ? &d : nullptr, ctx->render_stream, nullptr, &local);
        if (rc != RAYCA_OK) return rc;
        launched = 1;
        kernels += local.kernel_launches;
        CUDA_TRY(cudaStreamSynchronize(ctx->render_stream));
    }
    for (uint32_t k = 0; k < nb; ++k) {
        if (edge[k] >= edge[k + 1]) continue;
        Band bd{edge[k], edge[k + 1], launched == 0, edge[k + 1] == height};
        rc = enqueue_render(ctx, scene, config, options, width, height, base, aux ? &d : nullptr, ctx->render_stream, &bd, &local);
        if (rc != RAYCA_OK) return rc;
        CUDA_TRY(cudaEventRecord(ctx->band_done[launched], ctx->render_stream));
        band_of[launched++] = k;
        kernels += local.kernel_launches;
    }
    for (uint32_t j = 0; j < launched && !host_view; ++j) {
        const size_t y0 = edge[band_of[j]], y1 = edge[band_of[j] + 1];
        CUDA_TRY(cudaStreamWaitEvent(ctx->copy_stream, ctx->band_done[j], 0));
        CUDA_TRY(cudaMemcpyAsync(rgba8_out + y0 * width * 4, base + y0 * width * 4, (y1 - y0) * width * 4, cudaMemcpyDeviceToHost,
                                 ctx->copy_stream));
    }
    if (!host_view) CUDA_TRY(cudaStreamSynchronize(ctx->copy_stream));
    local.kernel_launches = kernels;
    if (aux) {
        if (aux->hit_blas) CUDA_TRY(cudaMemcpy(aux->hit_blas, d.hit_blas, n * 4, cudaMemcpyDeviceToHost));
        if (aux->hit_prim) CUDA_TRY(cudaMemcpy(aux->hit_prim, d.hit_prim, n * 4, cudaMemcpyDeviceToHost));
        if (aux->hit_src) CUDA_TRY(cudaMemcpy(aux->hit_src, d.hit_src, n * 4, cudaMemcpyDeviceToHost));
        if (aux->depth) CUDA_TRY(cudaMemcpy(aux->depth, d.depth, n * 4, cudaMemcpyDeviceToHost));
        if (aux->uv) CUDA_TRY(cudaMemcpy(aux->uv, d.uv, n * 8, cudaMemcpyDeviceToHost));
        if (aux->rgba_f32) CUDA_TRY(cudaMemcpy(aux->rgba_f32, d.rgba_f32, n * 16, cudaMemcpyDeviceToHost));
    }
    double t1 = now_ms();
    if (stats) {
        const uint32_t launches = local.kernel_launches;
        rc = rayca_cuda_last_counters(ctx, &local);
        if (rc != RAYCA_OK) return rc;
        local.kernel_launches = launches;
        local.readback_ms = (t1 - t0) - local.trace_ms; // what the read-back added beyond the kernels
        if (local.readback_ms < 0.0) local.readback_ms = 0.0;
        *stats = local;
    }
    return RAYCA_OK;
}

int rayca_cuda_draw(const rayca_scene_desc* desc, const rayca_config* config, uint32_t width, uint32_t height, uint8_t* rgba8_out) {
    if (!desc || !config) return fail(RAYCA_E_INVALID_ARG, "null argument");
    rayca_ctx* ctx = nullptr;
    int device = 0;
    if (cudaGetDevice(&device) != cudaSuccess) device = 0;
    int rc = rayca_cuda_init(device, &ctx);
    if (rc != RAYCA_OK) return rc;
    rayca_scene* scene = nullptr;
    rc = rayca_cuda_scene_create(ctx, desc, config->bvh, &scene, nullptr);
    if (rc == RAYCA_OK) {
        rc = rayca_cuda_render(ctx, scene, config, nullptr, width, height, rgba8_out, nullptr, nullptr);
        rayca_cuda_scene_destroy(scene);
    }
    rayca_cuda_shutdown(ctx);
    return rc;
}

int rayca_cuda_scene_create_host(const rayca_scene_desc* desc, uint32_t bvh, rayca_scene** out, rayca_stats* stats) {
    if (!desc || !out) return fail(RAYCA_E_INVALID_ARG, "null argument");
    rayca_scene* s = new (std::nothrow) rayca_scene();
    if (!s) return fail(RAYCA_E_OOM, "host allocation");
    double t0 = now_ms();
    std::string err;
    int rc;
    try {
        rc = build_host_scene(*desc, bvh != 0, s->host, err);
    } catch (const std::bad_alloc&) {
        delete s;
        return fail(RAYCA_E_OOM, "host allocation while building the scene");
    }
    if (rc != RAYCA_OK) {
        delete s;
        return fail(rc, err);
    }
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        stats->build_ms = now_ms() - t0;
    }
    *out = s;
    return RAYCA_OK;
}

int rayca_cuda_scene_bvh_info(rayca_scene* s, rayca_bvh_info* info) {
    if (!s || !info) return fail(RAYCA_E_INVALID_ARG, "null argument");
    info->blas_count = (uint32_t)s->host.blas.size();
    info->node_count = s->host.blas_node_offset.back();
    info->triangle_count = s->host.blas_prim_offset.back();
    info->tlas_node_count = (uint32_t)s->host.nodes.size() - info->node_count;
    return RAYCA_OK;
}

int rayca_cuda_scene_bvh_read(rayca_scene* s, float* nodes, uint32_t* blas_node_offset, uint32_t* blas_prim_offset, uint32_t* prim_src) {
    if (!s) return fail(RAYCA_E_INVALID_ARG, "null argument");
    const HostScene& h = s->host;
    size_t nb = h.blas.size();
    if (blas_node_offset) std::memcpy(blas_node_offset, h.blas_node_offset.data(), (nb + 1) * 4);
    if (blas_prim_offset) std::memcpy(blas_prim_offset, h.blas_prim_offset.data(), (nb + 1) * 4);
    for (size_t bi = 0; bi < nb; ++bi) {
        const HostBlas& b = h.blas[bi];
        if (nodes) {
            float* o = nodes + (size_t)h.blas_node_offset[bi] * 8;
            for (const HostNode& n : b.nodes) { // BLAS-local numbering, like the reference's Vec<BvhNode>
                o[0] = n.bmin.x; o[1] = n.bmin.y; o[2] = n.bmin.z;
                std::memcpy(o + 3, &n.offset, 4);
                o[4] = n.bmax.x; o[5] = n.bmax.y; o[6] = n.bmax.z;
                std::memcpy(o + 7, &n.count, 4);
                o += 8;
            }
        }
        if (prim_src) {
            uint32_t* o = prim_src + h.blas_prim_offset[bi];
            for (const HostTri& t : b.tris) *o++ = t.src;
            if (b.tris.empty()) // a model emitted on the device: the build left its primitive order on the host
                for (uint32_t src : b.order) *o++ = src;
        }
    }
    return RAYCA_OK;
}

} // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// rayca_cuda_render_static — the allocation-free render (include/rayca_cuda.h; no counterpart in the
// reference).
//
// The launch's scratch — the kernel's ray counters, which must start at zero, and the RGBA8 plane the kernel stores
// into before the frame is copied to the caller's host buffer — lives in ONE file-scope device array instead of
// cudaMalloc'ed memory.  In host code `g_static_scratch` is only a host-side stand-in for the symbol, NOT a device
// address: the address the kernel gets comes from cudaGetSymbolAddress, resolved once per device with that device
// current (each device that loads the module has its own copy of the array).
#include <mutex>

namespace {
constexpr size_t kStaticHeaderBytes = 256; // kCounterSlots x u64 ray counters, padded to the plane's alignment
static_assert(kCounterSlots * sizeof(unsigned long long) <= kStaticHeaderBytes, "counters fit the header");
static __device__ __align__(256) unsigned char g_static_scratch[kStaticHeaderBytes + RAYCA_STATIC_FRAME_BYTES];

constexpr int kMaxStaticDevices = 64;
std::mutex g_static_mutex;
unsigned char* g_static_addr[kMaxStaticDevices] = {};

// device address of g_static_scratch on `device` (which must be current)
int static_scratch_address(int device, unsigned char** out) {
    if (device < 0 || device >= kMaxStaticDevices) return fail(RAYCA_E_UNSUPPORTED, "device index beyond the static scratch table");
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_addr[device]) {
        void* p = nullptr;
        CUDA_TRY(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) return fail(RAYCA_E_CUDA, "cudaGetSymbolAddress returned a null address");
        g_static_addr[device] = static_cast<unsigned char*>(p);
    }
    *out = g_static_addr[device];
    return RAYCA_OK;
}

// points ctx->d_counters at the static array for the duration of one call (enqueue_render and
// rayca_cuda_last_counters read the pointer from the context), restoring it on every exit path
struct CounterSwap {
    rayca_ctx* ctx;
    unsigned long long* saved;
    CounterSwap(rayca_ctx* c, unsigned long long* p) : ctx(c), saved(c->d_counters) { c->d_counters = p; }
    ~CounterSwap() { ctx->d_counters = saved; }
};
} // namespace

extern "C" int rayca_cuda_render_static(rayca_ctx* ctx, rayca_scene* scene, const rayca_config* config, const rayca_options* options,
                                        uint32_t width, uint32_t height, uint8_t* rgba8_out, rayca_stats* stats) {
    if (!ctx || !scene || !config || !rgba8_out || width == 0 || height == 0) return fail(RAYCA_E_INVALID_ARG, "null / zero argument");
    const unsigned long long bytes = (unsigned long long)width * height * 4ull;
    if (bytes > RAYCA_STATIC_FRAME_BYTES)
        return fail(RAYCA_E_INVALID_ARG, "rayca_cuda_render_static: a " + std::to_string(width) + " x " + std::to_string(height) +
                                             " frame needs " + std::to_string(bytes) + " bytes, the static scratch holds " +
                                             std::to_string((unsigned long long)RAYCA_STATIC_FRAME_BYTES) +
                                             " (RAYCA_STATIC_FRAME_BYTES); use rayca_cuda_render");
    if (options && options->shard_count > 1) return fail(RAYCA_E_INVALID_ARG, "rayca_cuda_render_static renders whole frames (shard_count <= 1)");
    CUDA_TRY(cudaSetDevice(ctx->device));
    unsigned char* base = nullptr;
    int rc = static_scratch_address(ctx->device, &base);
    if (rc != RAYCA_OK) return rc;
    unsigned char* plane = base + kStaticHeaderBytes;
    double t0 = now_ms();
    rayca_stats local{};
    {
        CounterSwap swap(ctx, reinterpret_cast<unsigned long long*>(base));
        // enqueue_render clears the counters on the stream (cudaMemsetAsync) before the launch: the array keeps the
        // previous call's contents, and the counters are the part that must start at zero; every pixel of the
        // plane inside width x height is overwritten by the kernel
        // the fused kernel only: the wavefront path keeps its ray buffers in cudaMalloc'ed scratch
        rayca_options fused{};
        if (options) fused = *options;
        fused.flags |= RAYCA_FLAG_NO_WAVEFRONT;
        rc = enqueue_render(ctx, scene, config, &fused, width, height, plane, nullptr, ctx->render_stream, nullptr, &local);
        if (rc != RAYCA_OK) return rc;
        CUDA_TRY(cudaMemcpyAsync(rgba8_out, plane, (size_t)bytes, cudaMemcpyDeviceToHost, ctx->render_stream));
        CUDA_TRY(cudaStreamSynchronize(ctx->render_stream));
        if (stats) {
            rc = rayca_cuda_last_counters(ctx, &local); // reads the counters through ctx->d_counters (still the static array)
            if (rc != RAYCA_OK) return rc;
        }
    }
    if (stats) {
        local.kernel_launches = 1;
        local.readback_ms = (now_ms() - t0) - local.trace_ms;
        if (local.readback_ms < 0.0) local.readback_ms = 0.0;
        *stats = local;
    }
    return RAYCA_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Progressive accumulation across calls.
struct rayca_accum {
    rayca_ctx* ctx = nullptr;
    float4* plane = nullptr; // running per-pixel sums, in sample order
    unsigned char* rgba8 = nullptr; // device staging plane for the pixels of the latest call
    uint32_t width = 0, height = 0, done = 0;
};

extern "C" {

int rayca_cuda_accum_create(rayca_ctx* ctx, uint32_t width, uint32_t height, rayca_accum** out) {
    if (!ctx || !out || width == 0 || height == 0 || (uint64_t)width * height >= 0xFFFFFFFFull) return fail(RAYCA_E_INVALID_ARG, "null / zero argument");
    CUDA_TRY(cudaSetDevice(ctx->device));
    rayca_accum* a = new (std::nothrow) rayca_accum();
    if (!a) return fail(RAYCA_E_OOM, "host allocation");
    a->ctx = ctx;
    a->width = width;
    a->height = height;
    const size_t n = (size_t)width * height;
    cudaError_t ce = cudaMalloc((void**)&a->plane, n * sizeof(float4));
    if (ce == cudaSuccess) ce = cudaMalloc((void**)&a->rgba8, n * 4);
    if (ce != cudaSuccess) {
        rayca_cuda_accum_destroy(a);
        return fail(ce == cudaErrorMemoryAllocation ? RAYCA_E_OOM : RAYCA_E_CUDA, std::string("accum_create: ") + cudaGetErrorString(ce));
    }
    *out = a;
    return RAYCA_OK;
}

void rayca_cuda_accum_destroy(rayca_accum* a) {
    if (!a) return;
    if (a->ctx) cudaSetDevice(a->ctx->device);
    cudaFree(a->plane);
    cudaFree(a->rgba8);
    delete a;
}

void rayca_cuda_accum_reset(rayca_accum* a) {
    if (a) a->done = 0; // the next call starts every pixel from BLACK again and overwrites the plane
}

uint32_t rayca_cuda_accum_samples(rayca_accum* a) { return a ? a->done : 0; }

int rayca_cuda_render_accumulate(rayca_ctx* ctx, rayca_scene* scene, const rayca_config* config, const rayca_options* options,
                                 rayca_accum* accum, uint32_t sample_count, uint8_t* rgba8_out, rayca_stats* stats) {
    if (!ctx || !scene || !config || !accum || sample_count == 0) return fail(RAYCA_E_INVALID_ARG, "null / zero argument");
    if (accum->ctx != ctx) return fail(RAYCA_E_INVALID_ARG, "accumulator belongs to another context");
    if ((uint64_t)accum->done + sample_count > config->samples_per_pixel)
        return fail(RAYCA_E_INVALID_ARG, "accumulating " + std::to_string(accum->done) + " + " + std::to_string(sample_count) +
                                             " samples exceeds config->samples_per_pixel = " + std::to_string(config->samples_per_pixel));
    if (options && options->shard_count > 1) return fail(RAYCA_E_INVALID_ARG, "accumulation renders whole frames (shard_count <= 1)");
    CUDA_TRY(cudaSetDevice(ctx->device));
    const AccumTarget t{accum->plane, accum->done, sample_count};
    double t0 = now_ms();
    rayca_stats local{};
    int rc = enqueue_render(ctx, scene, config, options, accum->width, accum->height, accum->rgba8, nullptr, ctx->render_stream, nullptr, &local,
                            nullptr, &t);
    if (rc != RAYCA_OK) return rc;
    if (rgba8_out)
        CUDA_TRY(cudaMemcpyAsync(rgba8_out, accum->rgba8, (size_t)accum->width * accum->height * 4, cudaMemcpyDeviceToHost, ctx->render_stream));
    CUDA_TRY(cudaStreamSynchronize(ctx->render_stream));
    accum->done += sample_count;
    if (stats) {
        rc = rayca_cuda_last_counters(ctx, &local);
        if (rc != RAYCA_OK) return rc;
        local.kernel_launches = 1;
        local.readback_ms = std::max(0.0, (now_ms() - t0) - local.trace_ms);
        *stats = local;
    }
    return RAYCA_OK;
}

} // extern "C"
