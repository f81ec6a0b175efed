// wavefront.cu — wavefront scheduling with ray sorting for incoherent (bounce) rays.  SURVEY 8(f) rank 3; VERDICT r01 item 7.
//
// The fused kernel runs a path lane-per-pixel through all its depths: after the first bounce the 32 rays of a warp point
// all over the scene — 16 of 32 lanes active per instruction, twice the memory stalls, 3.4 x the time per ray
// (profiles/r02_ncu_bounce.json).  Here the path tracer of integrator/pathtracer.rs:23-106 (NEE on point lights + cosine
// bounces, one light sample) is cut at every bounce:
//
//   stage 0   one thread per pixel, 8x4 pixel tiles per warp: primary ray, depth 0 exactly as the fused kernel; a path
//             that goes on stores {next ray, direct light and BRDF factor of this level} and a 24-bit sort key,
//             a path that ends writes its pixel
//   sort      cub::DeviceRadixSort over (key, path index): key = direction octant | Morton code of the ray origin in the
//             scene box; ended paths sort to the end
//   stage k   one thread per SORTED path: neighbouring lanes trace rays that start close together and head the same way
//             (and their shadow rays leave from neighbouring hit points); a path that ends folds its stored levels in the
//             reference's order and writes its pixel
//
// Every value is computed by the same device functions, in the same order, from the same inputs as in the fused kernel
// (pathtracer<> in render.cuh), and the RNG is keyed by (pixel, sample, depth, dimension): the frame is bit-identical to
// the fused kernel's whatever the order the paths are traced in (tests/test_cuda_parity.py::test_wavefront_*).
#include <algorithm>
#include <cstdlib>

#include <cub/device/device_radix_sort.cuh>

#include "../../include/rayca_cuda.h"
#include "render.cuh"
#include "wavefront.h"

namespace rb {

namespace {


struct WaveView {
    float4* ray_o;   // [n] origin
    float4* ray_d;   // [n] direction
    float4* direct;  // [levels][n] direct light of level k (Color)
    float4* brdf;    // [levels][n] BRDF factor of level k (Color)
    uint32_t* keys_in;
    uint32_t* keys_out;
    uint32_t* idx_in;  // identity
    uint32_t* idx_out; // sorted path indices
    uint32_t n;        // paths of this launch = rows x width
    uint32_t key_bits; // Morton bits per axis
    uint32_t dir_bits, key_mode; // see sort_key
    uint32_t dead_key; // one bit above every live key, so that ended paths sort to the end
    float box_min[3], box_scale[3];
    // split stages (see wave_walk_kernel): the hit of the stage's ray, and the key layout of the sort by hit point
    float4* hit_a;     // [n] {t, u, v, primitive}
    float4* hit_b;     // [n] {p, sphere}
    uint32_t hit_bits; // Morton bits per axis of the hit point
    uint32_t miss_key; // rays that hit nothing: after every hit, before the ended paths
    uint32_t* heads;   // [levels] next unfetched sorted path of wave_walk_fetch_kernel, one counter per depth
    uint32_t fetch_min; // lanes that must still be busy for a warp to go on without fetching new rays
};

__device__ __forceinline__ uint32_t spread3(uint32_t v) { // 10 bits -> every third bit
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}

__device__ __forceinline__ uint32_t spread2(uint32_t v) { // 16 bits -> every second bit
    v = (v | (v << 8)) & 0x00FF00FFu;
    v = (v | (v << 4)) & 0x0F0F0F0Fu;
    v = (v | (v << 2)) & 0x33333333u;
    v = (v | (v << 1)) & 0x55555555u;
    return v;
}

__device__ __forceinline__ uint32_t morton_cell(const WaveView& w, V3 o, uint32_t bits) {
    const float cells = (float)(1u << bits);
    const uint32_t qx = min((uint32_t)fmaxf((o.x - w.box_min[0]) * w.box_scale[0] * cells, 0.0f), (1u << bits) - 1u);
    const uint32_t qy = min((uint32_t)fmaxf((o.y - w.box_min[1]) * w.box_scale[1] * cells, 0.0f), (1u << bits) - 1u);
    const uint32_t qz = min((uint32_t)fmaxf((o.z - w.box_min[2]) * w.box_scale[2] * cells, 0.0f), (1u << bits) - 1u);
    return spread3(qx) | (spread3(qy) << 1) | (spread3(qz) << 2);
}

// Sort key of a live ray.  key_mode 0: direction octant | Morton code of the origin (key_bits per axis).
// key_mode 1: Morton code of the origin in a COARSE grid (key_bits per axis) | direction on the octahedral map
// (dir_bits per axis, Morton order): rays that start in the same cell and point the same way — the ones that walk the
// same nodes — become neighbours.  key_mode 2: the same fields, direction first.
__device__ __forceinline__ uint32_t sort_key(const WaveView& w, V3 o, V3 d) {
    const uint32_t morton = morton_cell(w, o, w.key_bits);
    if (w.key_mode == 0) {
        const uint32_t octant = (d.x < 0.0f ? 1u : 0u) | (d.y < 0.0f ? 2u : 0u) | (d.z < 0.0f ? 4u : 0u);
        return (octant << (3u * w.key_bits)) | morton;
    }
    // octahedral map of the direction: the unit octahedron unfolded onto [0,1]^2
    const float inv = 1.0f / (fabsf(d.x) + fabsf(d.y) + fabsf(d.z) + 1e-30f);
    float px = d.x * inv, py = d.y * inv;
    if (d.z < 0.0f) {
        const float tx = (1.0f - fabsf(py)) * (px >= 0.0f ? 1.0f : -1.0f), ty = (1.0f - fabsf(px)) * (py >= 0.0f ? 1.0f : -1.0f);
        px = tx;
        py = ty;
    }
    const float dc = (float)(1u << w.dir_bits);
    const uint32_t du = min((uint32_t)fmaxf((px * 0.5f + 0.5f) * dc, 0.0f), (1u << w.dir_bits) - 1u);
    const uint32_t dv = min((uint32_t)fmaxf((py * 0.5f + 0.5f) * dc, 0.0f), (1u << w.dir_bits) - 1u);
    const uint32_t dir = spread2(du) | (spread2(dv) << 1);
    return w.key_mode == 1 ? (morton << (2u * w.dir_bits)) | dir : (dir << (3u * w.key_bits)) | morton;
}

__device__ __forceinline__ Col ld_c(const float4* p) {
    const float4 v = *p;
    return Col{v.x, v.y, v.z, v.w};
}
__device__ __forceinline__ void st_c(float4* p, Col c) { *p = make_float4(c.r, c.g, c.b, c.a); }

// One depth of Pathtracer::trace_impl for path `i` (pathtracer<MODE, false, true> in render.cuh, one loop iteration),
// from the point where the depth's ray has been traced (`hit`, `h`).  Returns true when the path goes on (its next ray,
// level and key are stored); otherwise `out` / `ok` hold the folded radiance of the whole path (ok = false: `None`, the
// pixel is BLACK).
template <int MODE>
__device__ __forceinline__ bool wave_after_walk(const RenderParams& P, const WaveView& w, uint32_t i, const RayD& ray, uint32_t pixel, uint32_t depth,
                                                const HitD& h, bool hit, PixelCounters& pc, Col& out, bool& ok) {
    const GpuSceneView& s = P.scene;
    const bool nee = P.cfg.direct_sampler == RAYCA_SAMPLER_NEE;
    const uint32_t max_depth = min(P.cfg.max_depth, (uint32_t)kMaxDepth);
    const uint32_t limit = nee ? P.cfg.max_depth - 1 : P.cfg.max_depth;
    const bool collect_emissive = depth == 0 || !nee;
    uint32_t levels = depth; // levels stored so far: every earlier depth stored one
    bool have_tail = false;
    Col tail{0.0f, 0.0f, 0.0f, 1.0f};
    if (hit) {
        Surf sf;
        load_surf<true, MODE == 3>(s, ray, h, sf);
        if (collect_emissive && sf.mat->kind == RAYCA_MATERIAL_PHONG && sf.mat->emissive) {
            tail = surf_color(s, sf);
            have_tail = true;
        } else {
            Col d{0.0f, 0.0f, 0.0f, 1.0f};
            if (nee) d = nee_direct<MODE, false, false>(P, sf, pc);
            if (!(depth < limit)) { // no indirect: Some(direct + BLACK)
                tail = cadd(d, Col{0.0f, 0.0f, 0.0f, 1.0f});
                have_tail = true;
            } else {
                const V3 wi = cosine_dir(P, sf, pixel, 0u, depth, 0);
                st_c(w.direct + (size_t)depth * w.n + i, d);
                st_c(w.brdf + (size_t)depth * w.n + i, surf_brdf(s, sf, wi));
                levels = depth + 1;
                if (depth + 1 < P.cfg.max_depth) pc.secondary++;
                if (depth + 1 < max_depth) { // the loop of trace_impl goes round again
                    const RayD next = make_ray(sf.next_origin, wi);
                    w.ray_o[i] = make_float4(next.o.x, next.o.y, next.o.z, 0.0f);
                    w.ray_d[i] = make_float4(next.d.x, next.d.y, next.d.z, 0.0f);
                    w.keys_in[i] = sort_key(w, next.o, next.d);
                    return true;
                }
            }
        }
    }
    w.keys_in[i] = w.dead_key;
    // fold: value_k = direct_k + (BLACK + (BLACK + PI * brdf_k * value_{k+1} * 1.0) / light_samples)
    ok = levels != 0 || have_tail;
    bool have = have_tail;
    Col val = tail;
    for (int k = (int)levels - 1; k >= 0; --k) {
        Col li{0.0f, 0.0f, 0.0f, 1.0f};
        if (have) li = cadd(li, cscale(cmul(cscale(ld_c(w.brdf + (size_t)k * w.n + i), kPi), val), 1.0f));
        li = cdiv(li, (float)P.cfg.light_samples);
        const Col indirect = cadd(Col{0.0f, 0.0f, 0.0f, 1.0f}, li);
        val = cadd(ld_c(w.direct + (size_t)k * w.n + i), indirect);
        have = true;
    }
    out = val;
    return false;
}

// the whole depth: trace the depth's ray, then the above
template <int MODE>
__device__ __forceinline__ bool wave_step(const RenderParams& P, const WaveView& w, uint32_t i, RayD ray, uint32_t pixel, uint32_t depth,
                                          HitD& primary, PixelCounters& pc, Col& out, bool& ok) {
    HitD h;
    const bool hit = trace_ray<MODE, false, false>(P.scene, ray, h, &pc.walk);
    if (depth == 0) primary = h;
    return wave_after_walk<MODE>(P, w, i, ray, pixel, depth, h, hit, pc, out, ok);
}

__device__ __forceinline__ void flush_counters(const RenderParams& P, const PixelCounters& pc) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t sh = __reduce_add_sync(0xFFFFFFFFu, pc.shadow), se = __reduce_add_sync(0xFFFFFFFFu, pc.secondary);
    if (lane == 0) {
        if (sh) atomicAdd(&P.counters[0], (unsigned long long)sh);
        if (se) atomicAdd(&P.counters[1], (unsigned long long)se);
    }
}

// stage 0: the pixel grid of the fused kernel (16x8 pixel blocks, 8x4 tiles per warp), depth 0
template <int MODE>
__global__ void __launch_bounds__(128, RAYCA_MIN_BLOCKS) wave_stage0_kernel(const __grid_constant__ RenderParams P, const WaveView w) {
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    stage_treelet(P.scene); // nothing unless RAYCA_TREELET
    const uint32_t x = blockIdx.x * 16u + (warp & 1u) * 8u + (lane & 7u), local_row = blockIdx.y * 8u + (warp >> 1) * 4u + (lane >> 3);
    const uint32_t y = image_row(P, local_row);
    const bool mine = x < P.width && y < P.row_end;
    PixelCounters pc;
    if (mine) {
        const uint32_t i = local_row * P.width + x;
        w.idx_in[i] = i;
        HitD primary = no_hit();
        Col c{0.0f, 0.0f, 0.0f, 1.0f};
        bool ok = false;
        const bool alive = wave_step<MODE>(P, w, i, primary_ray(P, x, y, 0u), y * P.width + x, 0u, primary, pc, c, ok);
        if (!alive) {
            if (!ok) c = Col{0.0f, 0.0f, 0.0f, 1.0f}; // unwrap_or(Color::BLACK) (scene.rs:80-86)
            write_pixel_color(P, x, y, local_row, cadd(Col{0.0f, 0.0f, 0.0f, 1.0f}, c));
        }
        write_pixel_aux(P, x, y, local_row, primary);
    }
    flush_counters(P, pc);
}

// pads the sort input beyond the launch's paths (rows x width need not fill the last sort tile) — and marks pixels outside
// the image rows of a ragged last block
__global__ void wave_fill_kernel(uint32_t* keys, uint32_t* idx, uint32_t n, uint32_t dead_key, uint32_t* heads, uint32_t head_count) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < head_count) heads[i] = 0u;
    if (i < n) {
        keys[i] = dead_key;
        idx[i] = i;
    }
}

// stage k >= 1: one thread per SORTED path
template <int MODE>
__global__ void __launch_bounds__(128, RAYCA_MIN_BLOCKS) wave_stage_kernel(const __grid_constant__ RenderParams P, const WaveView w, uint32_t depth) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    stage_treelet(P.scene); // nothing unless RAYCA_TREELET
    PixelCounters pc;
    if (t < w.n && w.keys_out[t] != w.dead_key) {
        const uint32_t i = w.idx_out[t];
        const uint32_t local_row = i / P.width, x = i - local_row * P.width, y = image_row(P, local_row);
        const float4 o = w.ray_o[i], d = w.ray_d[i];
        // make_ray recomputes rd = 1 / d from the stored direction: the same bits as in the fused kernel
        const RayD ray = make_ray(V3{o.x, o.y, o.z}, V3{d.x, d.y, d.z});
        HitD unused = no_hit();
        Col c{0.0f, 0.0f, 0.0f, 1.0f};
        bool ok = false;
        const bool alive = wave_step<MODE>(P, w, i, ray, y * P.width + x, depth, unused, pc, c, ok);
        if (!alive) {
            if (!ok) c = Col{0.0f, 0.0f, 0.0f, 1.0f};
            write_pixel_color(P, x, y, local_row, cadd(Col{0.0f, 0.0f, 0.0f, 1.0f}, c));
        }
    }
    flush_counters(P, pc);
}

// Split stage k >= 1, first half: one thread per path SORTED BY RAY, the depth's ray only.  The hit goes to memory with a
// key made of the hit POINT: the second half (shading, shadow rays towards the point lights, the next bounce) then runs
// on paths sorted by where they are now — their shadow rays leave from neighbouring points for the same lights, like
// the shadow rays of a tile of primary hits — instead of inheriting the order of the rays that got them there.
template <int MODE>
__global__ void __launch_bounds__(128, RAYCA_MIN_BLOCKS) wave_walk_kernel(const __grid_constant__ RenderParams P, const WaveView w) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    stage_treelet(P.scene); // nothing unless RAYCA_TREELET
    PixelCounters pc;
    if (t < w.n && w.keys_out[t] != w.dead_key) {
        const uint32_t i = w.idx_out[t];
        const float4 o = w.ray_o[i], d = w.ray_d[i];
        const RayD ray = make_ray(V3{o.x, o.y, o.z}, V3{d.x, d.y, d.z});
        HitD h;
        const bool hit = trace_ray<MODE, false, false>(P.scene, ray, h, &pc.walk);
        w.hit_a[i] = make_float4(h.t, h.u, h.v, __uint_as_float(hit ? h.prim : 0xFFFFFFFFu));
        w.hit_b[i] = make_float4(h.p.x, h.p.y, h.p.z, __uint_as_float(h.sphere));
        w.keys_in[i] = hit ? morton_cell(w, h.p, w.hit_bits) : w.miss_key;
    }
}

// The same first half as a PERSISTENT kernel with dynamic ray fetch (Aila & Laine, "Understanding the efficiency of ray
// traversal on GPUs", HPG 2009): a lane whose ray is finished does not wait for the longest ray of its warp — when
// fewer than `fetch_min` lanes are still busy, the idle lanes take the next rays of the sorted queue (one atomic per
// warp) and the warp goes on with 32 again.  Bounce rays of one warp differ in length by 2 x and more even when sorted
// (14.8 of 32 lanes active per instruction in the one-ray-per-thread stage, profiles/r02_ncu_bounce_wavefront_stage1.json).
// The walk is walk_fast_body's (trace.cuh), closest hit, triangles only, restated as a resumable loop: the same boxes
// and triangles are tested with the same arithmetic and the same tie rule, so the hit is the same whatever the order.
// SPECULATE: a lane that reaches a leaf while others are still descending puts the leaf aside and goes on with its
// next stack entry (Aila & Laine's "speculative traversal"); the node loop then ends at the lane's SECOND leaf and both
// are tested — the trip counts of the node loop vary less across the warp.  Entries visited before the postponed leaf
// could shorten the ray are extra work, never a different hit (culling only ever drops what cannot win).
template <bool SPECULATE>
__global__ void __launch_bounds__(128, RAYCA_MIN_BLOCKS) wave_walk_fetch_kernel(const __grid_constant__ RenderParams P, const WaveView w, uint32_t* head) {
    const GpuSceneView& s = P.scene;
    const uint32_t lane = threadIdx.x & 31u;
    constexpr uint32_t kFull = 0xFFFFFFFFu, kDone = 0xFFFFFFFFu;
    const float4* __restrict__ fast_nodes = s.fast_nodes;
    const float4* __restrict__ fast_tri = s.fast_tri;
    const float4* __restrict__ ref_nodes = s.nodes;
    uint2 stack[kStackSize + 2];
    bool alive = false, exhausted = false;
    uint32_t i = 0, cur = kDone, best = 0xFFFFFFFFu, ok_leaf = 0xFFFFFFFFu;
    int sp = 1;
    float best_t = 0.0f, bu = 0.0f, bv = 0.0f, bden = 1.0f;
    RayD ray{};
    FastRay fr{};
    auto push = [&](uint32_t ref, float dist) {
        stack[sp] = make_uint2(ref, __float_as_uint(dist));
        sp = min(sp + 1, kStackSize + 1);
    };
    auto pop = [&]() -> uint32_t {
        uint2 e;
        do {
            e = stack[--sp];
        } while (__uint_as_float(e.y) > best_t);
        return e.x;
    };
    auto ref_leaf_accepts = [&](uint32_t leaf) -> bool {
        if (leaf == ok_leaf) return true;
        const float4 lo = __ldg(&ref_nodes[2 * (size_t)leaf]);
        const float4 hi = __ldg(&ref_nodes[2 * (size_t)leaf + 1]);
        if (slab(lo, hi, ray) == kF32Max) return false;
        ok_leaf = leaf;
        return true;
    };
    auto test_leaf = [&](uint32_t ref) { // Triangle::intersects on the leaf's 64-byte records, as in walk_fast_body
        const V3 o = ray.o, d = ray.d;
        const uint32_t first = ref & 0x07FFFFFFu, count = ((ref >> 28) & 7u) + 1u;
        for (uint32_t k = first; k < first + count; ++k) {
            const float4* q = fast_tri + 4 * (size_t)k;
            const float4 q0 = __ldg(q);
            const V3 n{q0.x, q0.y, q0.z};
            const float ndd = dot(d, n);
            if (ndd > 0.0f) continue;
            if (fabsf(ndd) < kF32Eps) continue;
            const float t = -(dot(n, o) + q0.w) / ndd;
            if (t < 0.0f) continue;
            if (t > best_t) continue;
            const float4 q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3);
            const V3 v0{q1.x, q1.y, q1.z}, v1{q2.x, q2.y, q2.z}, v2{q3.x, q3.y, q3.z};
            const V3 p = o + d * t;
            V3 cc = cross(v1 - v0, p - v0);
            if (dot(n, cc) < 0.0f) continue;
            cc = cross(v2 - v1, p - v1);
            const float u = dot(n, cc);
            if (u < 0.0f) continue;
            cc = cross(v0 - v2, p - v2);
            const float v = dot(n, cc);
            if (v < 0.0f) continue;
            const uint32_t id = __float_as_uint(q1.w);
            if (t < best_t || (t == best_t && id < best && best != 0xFFFFFFFFu)) {
                if (ref_leaf_accepts(__float_as_uint(q2.w))) {
                    best_t = t;
                    best = id;
                    bu = u;
                    bv = v;
                    bden = q3.w;
                }
            }
        }
    };
    for (;;) {
        // ---- idle lanes fetch the next rays of the sorted queue
        const uint32_t idle = __ballot_sync(kFull, !alive);
        if (idle && !exhausted) {
            const uint32_t leader = (uint32_t)__ffs((int)idle) - 1u;
            uint32_t base = 0;
            if (lane == leader) base = atomicAdd(head, (uint32_t)__popc(idle));
            base = __shfl_sync(kFull, base, (int)leader);
            bool dry = false;
            if (!alive) {
                const uint32_t t = base + (uint32_t)__popc(idle & ((1u << lane) - 1u));
                if (t < w.n && w.keys_out[t] != w.dead_key) {
                    i = w.idx_out[t];
                    const float4 o = w.ray_o[i], d = w.ray_d[i];
                    ray = make_ray(V3{o.x, o.y, o.z}, V3{d.x, d.y, d.z});
                    fr = make_fast_ray(ray);
                    best_t = __int_as_float(0x7f800000);
                    best = 0xFFFFFFFFu;
                    bu = bv = 0.0f;
                    bden = 1.0f;
                    ok_leaf = 0xFFFFFFFFu;
                    bool go = !(ray.rd.x == 0.0f || ray.rd.y == 0.0f || ray.rd.z == 0.0f);
                    if (go) {
                        const float4 lo = make_float4(s.fast_root_min[0], s.fast_root_min[1], s.fast_root_min[2], 0.0f);
                        const float4 hi = make_float4(s.fast_root_max[0], s.fast_root_max[1], s.fast_root_max[2], 0.0f);
                        float e;
                        go = slab_fast(lo, hi, fr, e);
                    }
                    stack[0] = make_uint2(kDone, 0xff800000u);
                    sp = 1;
                    cur = go ? s.fast_root_ref : kDone;
                    alive = true;
                } else {
                    dry = true; // beyond the live paths (ended paths sort to the end)
                }
            }
            exhausted = __any_sync(kFull, dry);
        }
        if (!__any_sync(kFull, alive)) break;
        const uint32_t keep = exhausted ? 1u : w.fetch_min;
        // ---- walk: one "descend to a leaf, test the leaf" round per iteration, for the lanes that hold a ray
        do {
            if (alive) {
                uint32_t aside = kDone; // SPECULATE: the leaf put aside in this round (kDone: none)
                if (SPECULATE && (cur & kLeafFlag) && cur != kDone) { // the last pop was a leaf already
                    aside = cur;
                    cur = pop();
                }
                while (!(cur & kLeafFlag)) {
                    const float4* np = reinterpret_cast<const float4*>(reinterpret_cast<const char*>(fast_nodes) + (unsigned long long)cur * 32ull);
                    const float4 a0 = __ldg(np + 0), a1 = __ldg(np + 1), b0 = __ldg(np + 2), b1 = __ldg(np + 3);
                    float d0, d1;
                    bool h0 = slab_fast(a0, a1, fr, d0);
                    bool h1 = slab_fast(b0, b1, fr, d1);
                    h0 = h0 && !(d0 > best_t);
                    h1 = h1 && !(d1 > best_t);
                    uint32_t c0 = __float_as_uint(a0.w), c1 = __float_as_uint(b0.w);
                    if (h0 && h1) {
                        if (d1 < d0) {
                            float td = d0; d0 = d1; d1 = td;
                            uint32_t tc = c0; c0 = c1; c1 = tc;
                        }
                        cur = c0;
                        push(c1, d1);
                    } else if (h0) {
                        cur = c0;
                    } else if (h1) {
                        cur = c1;
                    } else {
                        cur = pop();
                    }
                    if (SPECULATE && (cur & kLeafFlag) && cur != kDone && aside == kDone) {
                        aside = cur;
                        cur = pop();
                    }
                }
                if (SPECULATE && aside != kDone) test_leaf(aside);
                if (cur == kDone) { // the ray is finished: its hit goes out, the lane is free
                    const bool hit = best != 0xFFFFFFFFu;
                    const V3 p = ray.o + ray.d * best_t; // Hit::point, the `p` of the winning test (triangle.rs:121)
                    w.hit_a[i] = make_float4(best_t, bu / bden, bv / bden, __uint_as_float(best));
                    w.hit_b[i] = make_float4(p.x, p.y, p.z, __uint_as_float(0u));
                    w.keys_in[i] = hit ? morton_cell(w, p, w.hit_bits) : w.miss_key;
                    alive = false;
                } else {
                    test_leaf(cur);
                    cur = pop();
                }
            }
        } while ((uint32_t)__popc(__ballot_sync(kFull, alive)) >= keep);
    }
}

// Split stage k >= 1, second half: one thread per path SORTED BY HIT POINT
template <int MODE>
__global__ void __launch_bounds__(128, RAYCA_MIN_BLOCKS) wave_shade_kernel(const __grid_constant__ RenderParams P, const WaveView w, uint32_t depth) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    stage_treelet(P.scene); // nothing unless RAYCA_TREELET
    PixelCounters pc;
    if (t < w.n && w.keys_out[t] != w.dead_key) {
        const uint32_t i = w.idx_out[t];
        const uint32_t local_row = i / P.width, x = i - local_row * P.width, y = image_row(P, local_row);
        const float4 o = w.ray_o[i], d = w.ray_d[i], ha = w.hit_a[i], hb = w.hit_b[i];
        const RayD ray = make_ray(V3{o.x, o.y, o.z}, V3{d.x, d.y, d.z});
        HitD h;
        h.t = ha.x; h.u = ha.y; h.v = ha.z; h.prim = __float_as_uint(ha.w);
        h.p = V3{hb.x, hb.y, hb.z};
        h.sphere = __float_as_uint(hb.w);
        Col c{0.0f, 0.0f, 0.0f, 1.0f};
        bool ok = false;
        const bool alive = wave_after_walk<MODE>(P, w, i, ray, y * P.width + x, depth, h, h.prim != 0xFFFFFFFFu, pc, c, ok);
        if (!alive) {
            if (!ok) c = Col{0.0f, 0.0f, 0.0f, 1.0f};
            write_pixel_color(P, x, y, local_row, cadd(Col{0.0f, 0.0f, 0.0f, 1.0f}, c));
        }
    }
    flush_counters(P, pc);
}

// The launch's pixels, device plane -> their place in the target (same index in both): one thread per pixel in
// launch-local order, so a warp moves 128 contiguous bytes of an image row.  Used when the target is HOST memory: the
// sorted stages would otherwise store 4 bytes at a time to scattered addresses over PCIe.
__global__ void wave_copy_kernel(const __grid_constant__ RenderParams P, const uchar4* __restrict__ src, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t local_row = i / P.width, x = i - local_row * P.width, y = image_row(P, local_row);
    if (y >= P.row_end) return;
    const size_t o = (size_t)((P.shard_count > 1 && !P.direct_rows) ? local_row : y) * P.width + x;
    reinterpret_cast<uchar4*>(P.rgba8)[o] = src[o];
}

} // namespace

bool wave_supported(const RenderParams& P, uint32_t rows, uint32_t mode, bool count, bool persistent, bool general_integrator, bool bounce) {
    if (std::getenv("RAYCA_NO_WAVEFRONT")) return false; // A/B runs
    // Eight launches (fill, stage 0, four of the sort, stage 1, copy) each drain on their own: below about half a million
    // paths the fused kernel's single launch is faster (1080p bounce frame over N GPUs, wavefront against fused:
    // 2.07 M paths per GPU +20 %, 1.04 M +9 %, 0.52 M +0.6 %, 0.26 M -16 %; profiles/r02_wavefront_ab.txt)
    size_t min_paths = 600000;
    if (const char* e = std::getenv("RAYCA_WAVE_MIN_PATHS")) min_paths = (size_t)std::max(0, std::atoi(e));
    if ((size_t)rows * P.width < min_paths) return false;
    return P.cfg.integrator == RAYCA_INTEGRATOR_PATHTRACER && !general_integrator && bounce && mode == RAYCA_TRAVERSAL_ORDERED && !count &&
           !persistent && P.cfg.samples_per_pixel == 1 && P.sample_count == 1 && P.sample_first == 0 && P.accum == nullptr &&
           P.cfg.light_samples == 1 && P.gshard_count <= 1;
}

namespace {
template <typename T>
T* carve(unsigned char*& p, size_t count) {
    T* r = reinterpret_cast<T*>(p);
    p += (count * sizeof(T) + 255) & ~size_t(255);
    return r;
}
} // namespace

int wave_render(const RenderParams& P_in, uint32_t rows, bool spheres, WaveScratch& ws, cudaStream_t st, uint32_t* launches, std::string& err) {
    RenderParams P = P_in;
    const uint32_t n = rows * P.width;
    // pixels bound for host memory (zero-copy frames, shared host frames of the multi-GPU paths) are staged on the device
    bool stage_pixels = false;
    {
        cudaPointerAttributes at{};
        stage_pixels = cudaPointerGetAttributes(&at, P.rgba8) == cudaSuccess && at.type == cudaMemoryTypeHost;
        cudaGetLastError();
    }
    // the plane is indexed like the target: whole-frame rows when pixels go to their image rows, launch rows otherwise
    const size_t plane_pixels = (size_t)((P.shard_count > 1 && !P.direct_rows) ? rows : P.height) * P.width;
    const uint32_t max_depth = std::min<uint32_t>(P.cfg.max_depth, (uint32_t)kMaxDepth);
    const uint32_t levels = max_depth; // a level is stored at most once per depth
    // key layout (see sort_key); the defaults are the best of the A/B in profiles/r02_wavefront_ab.txt
    uint32_t key_mode = 1, key_bits = 4, dir_bits = 5;
    if (const char* e = std::getenv("RAYCA_WAVE_KEY_MODE")) key_mode = (uint32_t)std::min(2, std::max(0, std::atoi(e)));
    if (const char* e = std::getenv("RAYCA_WAVE_KEY_BITS")) key_bits = (uint32_t)std::min(9, std::max(1, std::atoi(e)));
    if (const char* e = std::getenv("RAYCA_WAVE_DIR_BITS")) dir_bits = (uint32_t)std::min(8, std::max(1, std::atoi(e)));
    if (key_mode != 0 && 3 * key_bits + 2 * dir_bits > 31) dir_bits = (31 - 3 * key_bits) / 2;
    // split stages (wave_walk_kernel / wave_shade_kernel, a second sort by hit point in between): on unless RAYCA_WAVE_SPLIT=0
    bool split = true;
    uint32_t hit_bits = 7;
    if (const char* e = std::getenv("RAYCA_WAVE_SPLIT")) split = std::atoi(e) != 0;
    if (const char* e = std::getenv("RAYCA_WAVE_HIT_BITS")) hit_bits = (uint32_t)std::min(10, std::max(1, std::atoi(e)));
    // the walk half as a persistent kernel with dynamic ray fetch (triangle-only scenes): on unless RAYCA_WAVE_FETCH=0
    bool fetch = split && !spheres;
    uint32_t fetch_min = 16;
    if (const char* e = std::getenv("RAYCA_WAVE_FETCH")) fetch = fetch && std::atoi(e) != 0;
    if (const char* e = std::getenv("RAYCA_WAVE_FETCH_MIN")) fetch_min = (uint32_t)std::min(32, std::max(1, std::atoi(e)));
    bool speculate = false; // RAYCA_WAVE_SPEC=1: speculative traversal in the walk half (A/B in profiles/r02_wavefront_ab.txt)
    if (const char* e = std::getenv("RAYCA_WAVE_SPEC")) speculate = std::atoi(e) != 0;
    uint32_t live_bits = key_mode == 0 ? 3 * key_bits + 3 : 3 * key_bits + 2 * dir_bits;
    if (split) live_bits = std::max(live_bits, 3 * hit_bits + 1); // hit keys, then the miss key one bit above them
    const int end_bit = (int)live_bits + 1; // ended paths carry the bit above the live keys
    size_t temp_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                    (int)n, 0, end_bit, st);
    auto al = [](size_t b) { return (b + 255) & ~size_t(255); };
    const size_t need = (split ? 4 : 2) * al((size_t)n * 16) + 2 * al((size_t)levels * n * 16) + 4 * al((size_t)n * 4) + al(temp_bytes) +
                        (stage_pixels ? al(plane_pixels * 4) : 0) + al((size_t)levels * 4) + 256;
    if (need > ws.bytes) {
        cudaFree(ws.base);
        ws.base = nullptr;
        ws.bytes = 0;
        cudaError_t e = cudaMalloc(&ws.base, need);
        if (e != cudaSuccess) {
            err = std::string("wavefront scratch: ") + cudaGetErrorString(e);
            return e == cudaErrorMemoryAllocation ? RAYCA_E_OOM : RAYCA_E_CUDA;
        }
        ws.bytes = need;
    }
    unsigned char* p = static_cast<unsigned char*>(ws.base);
    WaveView w{};
    w.ray_o = carve<float4>(p, n);
    w.ray_d = carve<float4>(p, n);
    w.direct = carve<float4>(p, (size_t)levels * n);
    w.brdf = carve<float4>(p, (size_t)levels * n);
    w.keys_in = carve<uint32_t>(p, n);
    w.keys_out = carve<uint32_t>(p, n);
    w.idx_in = carve<uint32_t>(p, n);
    w.idx_out = carve<uint32_t>(p, n);
    if (split) {
        w.hit_a = carve<float4>(p, n);
        w.hit_b = carve<float4>(p, n);
    }
    w.heads = carve<uint32_t>(p, levels);
    w.fetch_min = fetch_min;
    w.hit_bits = hit_bits;
    w.miss_key = 1u << (3 * hit_bits);
    uchar4* plane = stage_pixels ? carve<uchar4>(p, plane_pixels) : nullptr;
    void* temp = p;
    if (stage_pixels) P.rgba8 = reinterpret_cast<unsigned char*>(plane);
    w.n = n;
    w.key_bits = key_bits;
    w.dir_bits = dir_bits;
    w.key_mode = key_mode;
    w.dead_key = 1u << live_bits;
    for (int a = 0; a < 3; ++a) {
        const float lo = P.scene.fast_root_min[a], hi = P.scene.fast_root_max[a];
        w.box_min[a] = lo;
        w.box_scale[a] = hi > lo ? 1.0f / (hi - lo) : 0.0f;
    }
    uint32_t fetch_blocks = 0;
    if (fetch) { // as many blocks as the device keeps resident: every warp loops until the queue is dry
        int dev = 0, sms = 0, per_sm = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (speculate) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, wave_walk_fetch_kernel<true>, 128, 0);
        else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, wave_walk_fetch_kernel<false>, 128, 0);
        fetch_blocks = (uint32_t)std::max(1, sms * std::max(1, per_sm));
    }
    uint32_t k = 0;
    wave_fill_kernel<<<(n + 255) / 256, 256, 0, st>>>(w.keys_in, w.idx_in, n, w.dead_key, w.heads, levels);
    ++k;
    const dim3 grid((P.width + 15) / 16, (rows + 7) / 8);
    if (spheres) wave_stage0_kernel<3><<<grid, 128, 0, st>>>(P, w);
    else wave_stage0_kernel<0><<<grid, 128, 0, st>>>(P, w);
    ++k;
    for (uint32_t depth = 1; depth < max_depth; ++depth) {
        cudaError_t e = cub::DeviceRadixSort::SortPairs(temp, temp_bytes, w.keys_in, w.keys_out, w.idx_in, w.idx_out, (int)n, 0, end_bit, st);
        if (e != cudaSuccess) {
            err = std::string("wavefront sort: ") + cudaGetErrorString(e);
            return RAYCA_E_CUDA;
        }
        if (!split) {
            if (spheres) wave_stage_kernel<3><<<(n + 127) / 128, 128, 0, st>>>(P, w, depth);
            else wave_stage_kernel<0><<<(n + 127) / 128, 128, 0, st>>>(P, w, depth);
            ++k;
            continue;
        }
        if (fetch) {
            const uint32_t blocks = std::min<uint32_t>((n + 127) / 128, fetch_blocks);
            if (speculate) wave_walk_fetch_kernel<true><<<blocks, 128, 0, st>>>(P, w, w.heads + depth);
            else wave_walk_fetch_kernel<false><<<blocks, 128, 0, st>>>(P, w, w.heads + depth);
        } else if (spheres) {
            wave_walk_kernel<3><<<(n + 127) / 128, 128, 0, st>>>(P, w);
        } else {
            wave_walk_kernel<0><<<(n + 127) / 128, 128, 0, st>>>(P, w);
        }
        ++k;
        e = cub::DeviceRadixSort::SortPairs(temp, temp_bytes, w.keys_in, w.keys_out, w.idx_in, w.idx_out, (int)n, 0, end_bit, st);
        if (e != cudaSuccess) {
            err = std::string("wavefront sort: ") + cudaGetErrorString(e);
            return RAYCA_E_CUDA;
        }
        if (spheres) wave_shade_kernel<3><<<(n + 127) / 128, 128, 0, st>>>(P, w, depth);
        else wave_shade_kernel<0><<<(n + 127) / 128, 128, 0, st>>>(P, w, depth);
        ++k;
    }
    if (stage_pixels) {
        wave_copy_kernel<<<(n + 255) / 256, 256, 0, st>>>(P_in, plane, n);
        ++k;
    }
    if (launches) *launches = k;
    return cudaGetLastError() == cudaSuccess ? RAYCA_OK : RAYCA_E_CUDA;
}

} // namespace rb
