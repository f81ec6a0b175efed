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

// Sort key of a live ray.  key_mode 0: direction octant | Morton code of the origin (key_bits per axis).
// key_mode 1: Morton code of the origin in a COARSE grid (key_bits per axis) | direction on the octahedral map
// (dir_bits per axis, Morton order): rays that start in the same cell and point the same way — the ones that walk the
// same nodes — become neighbours.  key_mode 2: the same fields, direction first.
__device__ __forceinline__ uint32_t sort_key(const WaveView& w, V3 o, V3 d) {
    const float cells = (float)(1u << w.key_bits);
    const uint32_t qx = min((uint32_t)fmaxf((o.x - w.box_min[0]) * w.box_scale[0] * cells, 0.0f), (1u << w.key_bits) - 1u);
    const uint32_t qy = min((uint32_t)fmaxf((o.y - w.box_min[1]) * w.box_scale[1] * cells, 0.0f), (1u << w.key_bits) - 1u);
    const uint32_t qz = min((uint32_t)fmaxf((o.z - w.box_min[2]) * w.box_scale[2] * cells, 0.0f), (1u << w.key_bits) - 1u);
    const uint32_t morton = spread3(qx) | (spread3(qy) << 1) | (spread3(qz) << 2);
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

// One depth of Pathtracer::trace_impl for path `i` (pathtracer<MODE, false, true> in render.cuh, one loop iteration).
// Returns true when the path goes on (its next ray, level and key are stored); otherwise `out` / `ok` hold the folded
// radiance of the whole path (ok = false: `None`, the pixel is BLACK).
template <int MODE>
__device__ __forceinline__ bool wave_step(const RenderParams& P, const WaveView& w, uint32_t i, RayD ray, uint32_t pixel, uint32_t depth,
                                          HitD& primary, PixelCounters& pc, Col& out, bool& ok) {
    const GpuSceneView& s = P.scene;
    const bool nee = P.cfg.direct_sampler == RAYCA_SAMPLER_NEE;
    const uint32_t max_depth = min(P.cfg.max_depth, (uint32_t)kMaxDepth);
    const uint32_t limit = nee ? P.cfg.max_depth - 1 : P.cfg.max_depth;
    const bool collect_emissive = depth == 0 || !nee;
    uint32_t levels = depth; // levels stored so far: every earlier depth stored one
    bool have_tail = false;
    Col tail{0.0f, 0.0f, 0.0f, 1.0f};
    HitD h;
    const bool hit = trace_ray<MODE, false, false>(s, ray, h, &pc.walk);
    if (depth == 0) primary = h;
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
__global__ void wave_fill_kernel(uint32_t* keys, uint32_t* idx, uint32_t n, uint32_t dead_key) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
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
    const uint32_t live_bits = key_mode == 0 ? 3 * key_bits + 3 : 3 * key_bits + 2 * dir_bits;
    const int end_bit = (int)live_bits + 1; // ended paths carry the bit above the live keys
    size_t temp_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                    (int)n, 0, end_bit, st);
    auto al = [](size_t b) { return (b + 255) & ~size_t(255); };
    const size_t need = 2 * al((size_t)n * 16) + 2 * al((size_t)levels * n * 16) + 4 * al((size_t)n * 4) + al(temp_bytes) +
                        (stage_pixels ? al(plane_pixels * 4) : 0) + 256;
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
    uint32_t k = 0;
    wave_fill_kernel<<<(n + 255) / 256, 256, 0, st>>>(w.keys_in, w.idx_in, n, w.dead_key);
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
        if (spheres) wave_stage_kernel<3><<<(n + 127) / 128, 128, 0, st>>>(P, w, depth);
        else wave_stage_kernel<0><<<(n + 127) / 128, 128, 0, st>>>(P, w, depth);
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
