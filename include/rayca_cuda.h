/*
 * rayca_cuda.h — C ABI of librayca_cuda, the B200 (sm_100a) backend for the
 * per-pixel raytracing hot path of Fahien/rayca's `rayca-soft`.
 *
 * The boundary is the reference's plug-in trait
 *     pub trait Draw { fn draw(&mut self, scene: &Scene, image: &mut Image); }
 *                                        (rayca-soft/src/draw.rs:7-9)
 * serialised as plain-old-data: the caller (a Rust `rayca-cuda` crate, or the
 * Python/C++ harness in this repository) walks its `rayca_model::Scene` and
 * fills the POD arrays below; the library does everything `SoftRenderer::draw`
 * does (rayca-soft/src/scene.rs:88-154): scene flattening, BVH build, ray
 * generation, traversal, intersection, shading, quantisation.
 *
 * All entry points are `extern "C"`, take plain pointers and sizes, return 0 or
 * a negative RAYCA_E_* code and never unwind across the boundary.  Host pointers
 * in a descriptor are borrowed only for the duration of the call.
 *
 * The same POD structs are consumed by the CPU oracle (oracle/rayca_oracle.cpp),
 * which is test infrastructure and never linked into this library.
 */
#ifndef RAYCA_CUDA_H
#define RAYCA_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ status */
enum {
    RAYCA_OK = 0,
    RAYCA_E_INVALID_ARG = -1,
    RAYCA_E_UNSUPPORTED = -2, /* combos that `todo!()`/panic in the reference (SURVEY A-20) */
    RAYCA_E_OOM = -3,
    RAYCA_E_CUDA = -4,
    RAYCA_E_NO_CAMERA = -5, /* `assert!(!camera_draw_infos.is_empty())` rayca-soft/src/scene.rs:109 */
    RAYCA_E_TIMEOUT = -6    /* rayca_cuda_frame_sync: a shared-frame flag wait gave up */
};

#define RAYCA_NONE (-1) /* Handle::NONE (rayca-util/src/pack.rs:59) */

/* ------------------------------------------------------------- scene graph */

/* rayca-math/src/trs.rs:77-86 — scale, then rotate, then translate. */
typedef struct rayca_trs {
    float translation[3];
    float rotation[4]; /* x, y, z, w */
    float scale[3];
} rayca_trs;

/* rayca-model/src/node.rs:10-31.  Indices are local to the owning graph: for a
 * model node they index that model's meshes / cameras / lights / nodes; for a
 * scene node only `model` and children are meaningful. */
typedef struct rayca_node {
    rayca_trs trs;
    int32_t mesh;
    int32_t camera;
    int32_t light;
    int32_t model;
    uint32_t first_child; /* offset into rayca_scene_desc.children */
    uint32_t child_count;
} rayca_node;

/* rayca-geometry/src/vertex.rs:64-72,139-143 (Vertex + VertexExt). */
typedef struct rayca_vertex {
    float pos[3];
    float color[4];
    float normal[3];
    float tangent[3];
    float bitangent[3];
    float uv[2];
} rayca_vertex;

enum { RAYCA_GEOMETRY_TRIANGLE_MESH = 0, RAYCA_GEOMETRY_SPHERE = 1 };

/* rayca-geometry/src/geometry.rs (Geometry::{TriangleMesh,Sphere}); indices are
 * widened to u32 by the caller and are relative to `first_vertex`. */
typedef struct rayca_geometry {
    uint32_t kind;
    uint32_t first_vertex; /* into rayca_scene_desc.vertices */
    uint32_t vertex_count;
    uint32_t first_index; /* into rayca_scene_desc.indices */
    uint32_t index_count;
    float center[3]; /* sphere */
    float radius;
} rayca_geometry;

/* rayca-model/src/primitive.rs:9-14 */
typedef struct rayca_primitive {
    uint32_t geometry; /* model-local */
    int32_t material;  /* model-local index into materials, or RAYCA_NONE */
} rayca_primitive;

/* rayca-model/src/mesh.rs:31-34: list of primitive handles */
typedef struct rayca_mesh {
    uint32_t first_primitive; /* offset into rayca_scene_desc.mesh_primitives */
    uint32_t primitive_count;
} rayca_mesh;

enum { RAYCA_MATERIAL_PBR = 0, RAYCA_MATERIAL_PHONG = 1, RAYCA_MATERIAL_GGX = 2 };

/* rayca-model/src/material/mod.rs:15-20 */
typedef struct rayca_material {
    uint32_t kind;
    int32_t index; /* model-local index into pbr/phong/ggx arrays, or RAYCA_NONE */
} rayca_material;

/* rayca-model/src/material/pbr.rs:63-72 */
typedef struct rayca_pbr_material {
    float color[4];
    int32_t albedo; /* model-local texture index or RAYCA_NONE */
    int32_t normal;
    float metallic_factor;
    float roughness_factor;
    int32_t metallic_roughness;
} rayca_pbr_material;

/* rayca-model/src/material/phong.rs:10-34 */
typedef struct rayca_phong_material {
    float ambient[4];
    float emission[4];
    float diffuse[4];
    float specular[4];
    float shininess;
} rayca_phong_material;

/* rayca-model/src/material/ggx.rs:10-23 */
typedef struct rayca_ggx_material {
    float diffuse[4];
    float specular[4];
    float roughness;
} rayca_ggx_material;

/* rayca-model/src/texture.rs:36-40 (the sampler handle is unused by the renderer) */
typedef struct rayca_texture {
    int32_t image; /* model-local */
} rayca_texture;

enum { RAYCA_COLOR_RGB8 = 0, RAYCA_COLOR_RGBA8 = 1, RAYCA_COLOR_RGBA32F = 2 }; /* color/mod.rs:19-25 */

/* rayca-model/src/image.rs:26-37: row-major, top-left origin, already decoded */
typedef struct rayca_image {
    uint32_t width;
    uint32_t height;
    uint32_t color_type;
    uint32_t reserved;
    const void* data;
} rayca_image;

enum { RAYCA_LIGHT_DIRECTIONAL = 0, RAYCA_LIGHT_POINT = 1, RAYCA_LIGHT_QUAD = 2 };

/* rayca-model/src/light/{mod,point,directional,quad}.rs */
typedef struct rayca_light {
    uint32_t kind;
    float color[4];
    float intensity;
    float attenuation[3]; /* point: constant, linear, quadratic */
    float ab[3];          /* quad */
    float ac[3];          /* quad */
    int32_t material;     /* quad: model-local material */
} rayca_light;

/* rayca-model/src/camera.rs:11-14 (only yfov is read by the renderer) */
typedef struct rayca_camera {
    float yfov_radians;
} rayca_camera;

/* rayca-model/src/model.rs:31-51: offsets select this model's slice of the
 * global arrays in rayca_scene_desc. */
typedef struct rayca_model {
    rayca_trs root_trs;
    uint32_t root_first_child;
    uint32_t root_child_count;
    uint32_t node_offset, node_count;
    uint32_t mesh_offset, mesh_count;
    uint32_t primitive_offset, primitive_count;
    uint32_t geometry_offset, geometry_count;
    uint32_t material_offset, material_count;
    uint32_t pbr_offset, pbr_count;
    uint32_t phong_offset, phong_count;
    uint32_t ggx_offset, ggx_count;
    uint32_t texture_offset, texture_count;
    uint32_t image_offset, image_count;
    uint32_t light_offset, light_count;
    uint32_t camera_offset, camera_count;
} rayca_model;

/* rayca-model/src/scene.rs:47-53 */
typedef struct rayca_scene_desc {
    rayca_trs root_trs;
    uint32_t root_first_child;
    uint32_t root_child_count;

    const rayca_node* scene_nodes;
    uint32_t scene_node_count;
    const rayca_model* models;
    uint32_t model_count;

    const rayca_node* nodes; /* all models' nodes */
    uint32_t node_count;
    const uint32_t* children;
    uint32_t children_count;
    const rayca_mesh* meshes;
    uint32_t mesh_count;
    const uint32_t* mesh_primitives;
    uint32_t mesh_primitive_count;
    const rayca_primitive* primitives;
    uint32_t primitive_count;
    const rayca_geometry* geometries;
    uint32_t geometry_count;
    const rayca_vertex* vertices;
    uint64_t vertex_count;
    const uint32_t* indices;
    uint64_t index_count;
    const rayca_material* materials;
    uint32_t material_count;
    const rayca_pbr_material* pbr_materials;
    uint32_t pbr_count;
    const rayca_phong_material* phong_materials;
    uint32_t phong_count;
    const rayca_ggx_material* ggx_materials;
    uint32_t ggx_count;
    const rayca_texture* textures;
    uint32_t texture_count;
    const rayca_image* images;
    uint32_t image_count;
    const rayca_light* lights;
    uint32_t light_count;
    const rayca_camera* cameras;
    uint32_t camera_count;
} rayca_scene_desc;

/* ------------------------------------------------------------------ config */

/* `#[repr(u32)] IntegratorStrategy` rayca-soft/src/integrator/mod.rs:32-41 */
enum {
    RAYCA_INTEGRATOR_SCRATCHER = 0,
    RAYCA_INTEGRATOR_RAYTRACER = 1,
    RAYCA_INTEGRATOR_FLAT = 2,
    RAYCA_INTEGRATOR_ANALYTIC_DIRECT = 3,
    RAYCA_INTEGRATOR_DIRECT = 4,
    RAYCA_INTEGRATOR_PATHTRACER = 5
};

/* `#[repr(u32)] SamplerStrategy` rayca-soft/src/sampler/mod.rs:41-50 */
enum {
    RAYCA_SAMPLER_NONE = 0,
    RAYCA_SAMPLER_NEE = 1,
    RAYCA_SAMPLER_HEMISPHERE = 2,
    RAYCA_SAMPLER_COSINE = 3,
    RAYCA_SAMPLER_BRDF = 4,
    RAYCA_SAMPLER_MIS = 5
};

/* rayca-soft/src/config.rs:10-49, field for field (bools widened to u32). */
typedef struct rayca_config {
    uint32_t bvh;               /* default 1 */
    uint32_t light_samples;     /* default 1 */
    uint32_t light_stratify;    /* default 0 */
    uint32_t samples_per_pixel; /* default 1 */
    uint32_t russian_roulette;  /* default 0 */
    uint32_t direct_sampler;    /* default NEE */
    uint32_t indirect_sampler;  /* default COSINE */
    uint32_t integrator;        /* default PATHTRACER */
    uint32_t max_depth;         /* default 5 */
    float gamma;                /* default 1.0 */
} rayca_config;

enum {
    RAYCA_TRAVERSAL_ORDERED = 0,    /* product path: SAH performance tree, front-to-back, distance-culled,
                                       every candidate re-validated against its reference leaf box */
    RAYCA_TRAVERSAL_EXHAUSTIVE = 1, /* the reference tree walked like bvh/blas.rs:129-176: every overlapped node, no cull */
    RAYCA_TRAVERSAL_REFERENCE_ORDERED = 2 /* the flattened reference tree, front-to-back and distance-culled */
};

/* Backend-only knobs, kept out of rayca_config so the reference `Config` stays
 * source compatible. */
typedef struct rayca_options {
    uint64_t rng_seed;   /* keyed counter RNG for stochastic bounces (the reference uses unseeded fastrand) */
    uint32_t traversal;  /* RAYCA_TRAVERSAL_* */
    uint32_t shard_index; /* this process renders image tiles t with t % shard_count == shard_index */
    uint32_t shard_count; /* 0 or 1 = whole image */
    uint32_t tile_rows;   /* rows per shard tile, a multiple of 8 (0 = default 8) */
    uint32_t flags;       /* RAYCA_FLAG_* */
    uint32_t reserved;
} rayca_options;

enum {
    RAYCA_FLAG_COUNT_WALK = 1, /* also tally node / triangle records fetched (roofline accounting; slower) */
    RAYCA_FLAG_PERSISTENT = 2, /* persistent warps that pull 16x8 pixel groups from a global counter instead of
                                  one thread block per group.  NOT the default: measured 15-20 % slower than the
                                  static grid on B200 for the 1080p frames (render.cuh, profiles/r01_ab_registers_*) */
    RAYCA_FLAG_DIRECT_ROWS = 4, /* sharded launches (shard_count > 1): the output pointer addresses the FULL width x height
                                  frame and every owned pixel is stored at its final row-major position, instead of the
                                  shard's tiles packed one after the other — a frame all ranks can address (pinned host
                                  memory registered in every process, rayca_cuda_host_register) is then complete once
                                  every rank's launch is, with no gather and no device-to-host copy */
    RAYCA_FLAG_NO_WAVEFRONT = 16, /* path tracing with bounces normally runs as wavefront stages with the live rays sorted
                                     between depths (wavefront.cu); this flag keeps the one fused kernel (A/B runs, tests:
                                     both produce the same bytes) */
    RAYCA_FLAG_SHARD_GROUPS = 8 /* with RAYCA_FLAG_DIRECT_ROWS (or a shared frame): the shard owns 16x8 pixel groups in a diagonal
                                  checkerboard — group (gx, gy) belongs to shard (gx + gy) mod shard_count — instead of
                                  row tiles, so that every shard samples the whole image.  An option, not the default:
                                  on the 1080p atrium split 8 ways it measured no better than row tiles (slowest / mean
                                  1.09 against 1.12, mean 3 % worse: scattered groups lose cache locality), see
                                  profiles/r02_shard_emulation.txt */
};

/* Optional per-pixel planes for parity checks; any pointer may be NULL.
 * For the sample index spp-1 of each pixel (i.e. the only sample at 1 spp). */
typedef struct rayca_aux_out {
    uint32_t* hit_blas; /* Hit::blas — index into the TLAS-reordered blas_nodes (tlas.rs:152); ~0u on miss */
    uint32_t* hit_prim; /* Hit::primitive — index into the BLAS-reordered primitive vector (blas.rs:153) */
    uint32_t* hit_src;  /* stable id: position of the triangle in BvhModel::from_model emit order, before reordering */
    float* depth;       /* Hit::depth; +inf on miss */
    float* uv;          /* 2 floats per pixel, Hit::uv */
    float* rgba_f32;    /* 4 floats per pixel: the colour before quantisation (after /spp and gamma) */
} rayca_aux_out;

typedef struct rayca_stats {
    double build_ms;   /* host flatten + BVH build (scene_create) */
    double upload_ms;  /* H2D of the scene (scene_create) */
    double trace_ms;   /* CUDA-event time of the render kernel(s) */
    double readback_ms;
    uint64_t rays_primary;
    uint64_t rays_shadow;
    uint64_t rays_secondary;
    uint64_t node_bytes;     /* algorithmic bytes: 64 B per node-pair record fetched */
    uint64_t triangle_bytes; /* algorithmic bytes: product path 16 B per triangle plane fetched + 48 B per triangle
                                whose vertices were fetched; reference-tree modes 48 B per triangle tested */
    uint32_t kernel_launches;
    uint32_t reserved;
} rayca_stats;

/* read-back of the built acceleration structure, for builder parity tests */
typedef struct rayca_bvh_info {
    uint32_t blas_count;
    uint32_t node_count;     /* total 32-byte nodes across all BLAS (incl. the unused node 1 of each) */
    uint32_t triangle_count; /* total triangles */
    uint32_t tlas_node_count;
} rayca_bvh_info;

typedef struct rayca_ctx rayca_ctx;
typedef struct rayca_scene rayca_scene;

/* ------------------------------------------------------------- entry points */

/* Pinned (page-locked) host memory for `rgba8_out`: rayca_cuda_render lets the kernel store the frame straight
 * into it (zero copy, one launch).  The Rust shim allocates Image buffers that it owns through this pair; any
 * other host pointer works too — pageable memory is filled by overlapped copies of row bands instead. */
int rayca_cuda_host_alloc(size_t bytes, void** out);
void rayca_cuda_host_free(void* p);

/* Page-locks and maps host memory the caller already owns (cudaHostRegister, portable + mapped): an Image's Vec<u8>
 * that lives across frames, or a POSIX shared-memory frame that the one-process-per-GPU ranks of a node all mmap —
 * each rank registers its own mapping and its kernel stores its tiles straight into the shared frame over its own
 * PCIe link (RAYCA_FLAG_DIRECT_ROWS).  `device_view` receives the address the device uses for `p` (pass it to
 * rayca_cuda_render_device); rayca_cuda_render accepts `p` itself. */
int rayca_cuda_host_register(void* p, size_t bytes, void** device_view);
int rayca_cuda_host_unregister(void* p);

/* Measurement aid (bench.py): best-of-5 streaming read bandwidth over a `bytes` buffer read `reps` times with
 * 128-bit ld.global.cg loads (L1 bypassed) from all SMs.  A buffer well inside the 126 MB L2 gives the L2 -> SM
 * read peak that bounds the walk on L2-resident scenes; a buffer of several GB gives the HBM read peak. */
int rayca_cuda_probe_read_bandwidth(rayca_ctx* ctx, size_t bytes, uint32_t reps, double* gb_per_s);

/* ---- Shared frame: the multi-GPU gather folded into the trace kernel's own stores ------------------------
 * One process per GPU.  The OWNER (the rank that wants the image) creates a frame in its HBM and exports a
 * 64-byte CUDA IPC handle; every other rank imports it, which maps the owner's memory over NVLink / NVSwitch.
 * rayca_cuda_render_to_frame then makes each rank's trace kernel store its tiles' pixels straight into their
 * final row-major place in the owner's frame, followed on the same stream by a one-thread kernel that
 * publishes `frame_id` in that rank's flag slot.  The owner enqueues rayca_cuda_frame_wait (a one-warp kernel
 * spinning on the flags, with a wall-clock time-out, see rayca_cuda_frame_sync) and may then read the frame on the same stream: no collective,
 * no local staging buffer, no pack / unpack kernel.
 * Back-pressure for a pipeline of frames: the owner calls rayca_cuda_frame_release(frame_id) once it has
 * consumed a frame (e.g. after its device-to-host copy, in stream order); a rank that is about to overwrite the
 * frame first calls rayca_cuda_frame_acquire(previous frame_id), which waits for that release.  Two frames used
 * alternately let the ranks trace frame k+1 while the owner still reads frame k.
 * Layout of the allocation: 64 x u64 flags (512 B: slot r = rank r stored, slot 32 = owner consumed) |
 * width*height*4 bytes RGBA8. */
typedef struct rayca_frame rayca_frame;
#define RAYCA_IPC_HANDLE_BYTES 64
int rayca_cuda_frame_create(rayca_ctx* ctx, uint32_t width, uint32_t height, rayca_frame** out);
int rayca_cuda_frame_export(rayca_frame* frame, uint8_t handle[RAYCA_IPC_HANDLE_BYTES]);
int rayca_cuda_frame_import(rayca_ctx* ctx, const uint8_t handle[RAYCA_IPC_HANDLE_BYTES], uint32_t width, uint32_t height,
                            rayca_frame** out);
void* rayca_cuda_frame_pixels(rayca_frame* frame); /* device pointer (owner: local; importer: peer mapping) */
void rayca_cuda_frame_destroy(rayca_frame* frame);
/* options->shard_index / shard_count / tile_rows select this rank's tiles as in rayca_cuda_render_device
 * (at most 32 ranks per frame). */
int rayca_cuda_render_to_frame(rayca_ctx* ctx, rayca_scene* scene, const rayca_config* config, const rayca_options* options,
                               rayca_frame* frame, uint64_t frame_id, void* cuda_stream);
/* Owner only: returns once (in stream order) ranks 0..rank_count-1 have published a frame id >= frame_id.
 * A time-out sets a flag (device-visible host memory inside the frame object) read by rayca_cuda_frame_timed_out. */
int rayca_cuda_frame_wait(rayca_ctx* ctx, rayca_frame* frame, uint32_t rank_count, uint64_t frame_id, void* cuda_stream);
/* Owner only: in stream order, marks every frame id <= frame_id as consumed. */
int rayca_cuda_frame_release(rayca_ctx* ctx, rayca_frame* frame, uint64_t frame_id, void* cuda_stream);
/* Any rank: in stream order, waits until the owner has released a frame id >= frame_id (0 returns at once). */
int rayca_cuda_frame_acquire(rayca_ctx* ctx, rayca_frame* frame, uint64_t frame_id, void* cuda_stream);
int rayca_cuda_frame_timed_out(rayca_frame* frame);
/* The flag waits (frame_wait / frame_acquire) give up after a WALL-clock time-out (default 2000 ms; a rank can be that
 * late for honest reasons: first-launch module load, a BVH rebuild): set it per frame object.  The enqueueing call has
 * long returned by then, so the condition is reported by the synchronising call: rayca_cuda_frame_sync waits for
 * `cuda_stream` and returns RAYCA_E_TIMEOUT if any wait of this frame object has given up since the last
 * rayca_cuda_frame_clear_timeout.  After a time-out the frame's contents are UNDEFINED (a late rank may still be
 * storing its tiles, or the owner may not have consumed the previous frame): do not present it. */
int rayca_cuda_frame_set_timeout(rayca_frame* frame, double milliseconds);
int rayca_cuda_frame_sync(rayca_ctx* ctx, rayca_frame* frame, void* cuda_stream);
void rayca_cuda_frame_clear_timeout(rayca_frame* frame);

/* Binds a context to one CUDA device (one process per GPU). */
int rayca_cuda_init(int device, rayca_ctx** out);
/* One context over several devices, driven by ONE host thread (SURVEY 8b "Threading"): what the reference's caller
 * sees is a single `renderer.draw(&scene, &mut image)` that fans out over the rayon pool
 * (rayca-soft/src/scene.rs:117-122); here the fan-out is over GPUs.  `devices` = device_count CUDA device indices
 * (NULL / 0 = every visible device; an index may appear more than once, each entry is a shard of its own).
 * rayca_cuda_scene_create builds the scene on devices[0] and copies the finished device arrays to the others;
 * rayca_cuda_render interleaves the frame's 8-row tiles over the devices and every device's kernel stores its
 * tiles at their final rows (pinned caller buffers: zero copy over each device's own PCIe link; pageable buffers:
 * through a pinned staging frame the context owns, copied out by the calling thread's OpenMP team).  Calls with aux planes or an explicit shard run on devices[0].
 * The context's other entry points (render_device, render_to_frame, render_static, frames) address devices[0]. */
int rayca_cuda_init_devices(const int* devices, int device_count, rayca_ctx** out);
int rayca_cuda_device_count(rayca_ctx* ctx); /* shards of the context: 1 for rayca_cuda_init */
void rayca_cuda_shutdown(rayca_ctx* ctx);

/* Replaces scene.rs:90-99: SceneDrawInfo::new, BvhScene::from_scene, Tlas::build;
 * then uploads the flattened arrays.  `bvh` = Config::bvh (0 => max_depth(0)). */
int rayca_cuda_scene_create(rayca_ctx* ctx, const rayca_scene_desc* desc, uint32_t bvh,
                            rayca_scene** out, rayca_stats* stats /*nullable*/);
void rayca_cuda_scene_destroy(rayca_scene* scene);

/* Host half only (flatten + BVH build, no device, nothing uploaded): the result can be
 * inspected with the bvh_info / bvh_read calls below and destroyed; it cannot be rendered. */
int rayca_cuda_scene_create_host(const rayca_scene_desc* desc, uint32_t bvh, rayca_scene** out,
                                 rayca_stats* stats /*nullable*/);

/* Replaces the pixel loop scene.rs:101-152.  `rgba8_out` is HOST memory,
 * width*height*4 bytes, row-major top-left (rayca-model/src/image.rs:29-31).
 * `aux` planes are HOST memory. */
int rayca_cuda_render(rayca_ctx* ctx, rayca_scene* scene, const rayca_config* config,
                      const rayca_options* options /*nullable*/, uint32_t width, uint32_t height,
                      uint8_t* rgba8_out, const rayca_aux_out* aux /*nullable*/,
                      rayca_stats* stats /*nullable*/);

/* Allocation-free render.  NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same result as rayca_cuda_render(ctx, scene, config, options, width, height, rgba8_out, NULL, stats) — the same
 * kernel — but everything the launch needs beyond the scene lives in ONE file-scope `static __device__` array inside
 * the library: the kernel's ray counters (256 bytes, cleared by every call) followed by the RGBA8 staging plane the
 * kernel stores into before it is copied to `rgba8_out` (HOST memory, pageable or pinned).  The call performs no
 * cudaMalloc / cudaFree / cudaHostAlloc and calls no allocating library routine; the array's device address is
 * resolved once per device with cudaGetSymbolAddress, with that device current.
 * Limits: width * height * 4 <= RAYCA_STATIC_FRAME_BYTES (8 MiB: up to 1920 x 1080), else RAYCA_E_INVALID_ARG; no aux
 * planes; no sharding (options->shard_count <= 1).  Every GPU that loads the library holds the 8 MiB + 256 B array
 * until the process exits.  ALL calls on one device share the array and it keeps the previous call's contents: calls
 * that overlap on the GPU (two contexts or two host threads on the same device) race — serialise them. */
#define RAYCA_STATIC_FRAME_BYTES (8u << 20)
int rayca_cuda_render_static(rayca_ctx* ctx, rayca_scene* scene, const rayca_config* config,
                             const rayca_options* options /*nullable*/, uint32_t width, uint32_t height,
                             uint8_t* rgba8_out, rayca_stats* stats /*nullable*/);

/* Same, but every output pointer is DEVICE memory on the context's device and the
 * work is enqueued on `cuda_stream` (a cudaStream_t; NULL = default stream)
 * without synchronising: the multi-GPU caller gathers tiles with NCCL afterwards. */
int rayca_cuda_render_device(rayca_ctx* ctx, rayca_scene* scene, const rayca_config* config,
                             const rayca_options* options /*nullable*/, uint32_t width,
                             uint32_t height, void* d_rgba8, const rayca_aux_out* d_aux /*nullable*/,
                             void* cuda_stream, rayca_stats* stats /*nullable*/);

/* After a render_device + stream sync: fills the ray / byte counters of the last render. */
int rayca_cuda_last_counters(rayca_ctx* ctx, rayca_stats* stats);

/* One-shot `Draw::draw`: scene_create + render + scene_destroy. */
int rayca_cuda_draw(const rayca_scene_desc* desc, const rayca_config* config, uint32_t width,
                    uint32_t height, uint8_t* rgba8_out);

/* BVH read-back (host copies).  nodes: node_count x 8 floats (min.xyz, u32 left_or_first,
 * max.xyz, u32 count), BLAS after BLAS; blas_node_offset/blas_prim_offset: blas_count+1
 * prefix sums; prim_src: triangle_count stable source ids in BLAS-reordered order. */
int rayca_cuda_scene_bvh_info(rayca_scene* scene, rayca_bvh_info* info);
int rayca_cuda_scene_bvh_read(rayca_scene* scene, float* nodes, uint32_t* blas_node_offset,
                              uint32_t* blas_prim_offset, uint32_t* prim_src);

/* Progressive accumulation across calls (the reference has no counterpart — its `draw` always traces
 * all samples_per_pixel at once, rayca-soft/src/scene.rs:123-149).  An accumulator holds one running f32 sum per pixel.
 * Each rayca_cuda_render_accumulate traces the NEXT `sample_count` samples of the config->samples_per_pixel the sequence
 * is heading for (that total fixes the stratification grid, scene.rs:125-131, and must stay the same over the sequence),
 * adds them to the sums in sample order and writes sum / (samples so far) -> gamma -> RGBA8 to `rgba8_out` (HOST,
 * nullable).  Because the additions happen in the order of a single call, the frame written by the call that completes
 * the sequence equals rayca_cuda_render with the same config byte for byte.  Whole frames only, one device. */
typedef struct rayca_accum rayca_accum;
int rayca_cuda_accum_create(rayca_ctx* ctx, uint32_t width, uint32_t height, rayca_accum** out);
void rayca_cuda_accum_destroy(rayca_accum* accum);
void rayca_cuda_accum_reset(rayca_accum* accum);       /* start a new sequence (new camera / scene / config) */
uint32_t rayca_cuda_accum_samples(rayca_accum* accum); /* samples per pixel accumulated so far */
int rayca_cuda_render_accumulate(rayca_ctx* ctx, rayca_scene* scene, const rayca_config* config,
                                 const rayca_options* options /*nullable*/, rayca_accum* accum, uint32_t sample_count,
                                 uint8_t* rgba8_out /*nullable*/, rayca_stats* stats /*nullable*/);

/* `Image::dump_png` (rayca-model/src/image.rs:160-169): the RGBA8 frame (row-major, top-left origin) as an 8-bit RGBA,
 * non-interlaced PNG — to a file, or into a caller buffer.  Host-only (no device needed).  encode_png: `*out_bytes`
 * always receives the encoded size; with `out` NULL or `capacity` too small the call returns RAYCA_E_INVALID_ARG and
 * the caller comes back with a buffer of that size.  Decoding the file yields the input bytes exactly. */
int rayca_cuda_write_png(const char* path, const uint8_t* rgba8, uint32_t width, uint32_t height);
int rayca_cuda_encode_png(const uint8_t* rgba8, uint32_t width, uint32_t height, uint8_t* out, size_t capacity, size_t* out_bytes);

const char* rayca_cuda_strerror(int code);
const char* rayca_cuda_last_error(void); /* thread-local detail string */
uint32_t rayca_cuda_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* RAYCA_CUDA_H */
