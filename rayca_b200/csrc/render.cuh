// render.cuh — the fused per-pixel kernel: primary-ray generation, integrator, accumulation,
// gamma, RGBA8 quantisation.  One thread owns one pixel and loops over its samples in order, so
// the floating-point accumulation order equals the reference's (rayca-soft/src/scene.rs:117-150).
#pragma once
#include <cuda_runtime.h>

#include "../../include/rayca_cuda.h"
#include "gpu_types.h"
#include "rmath.h"
#include "shade.cuh"
#include "trace.cuh"

namespace rb {

constexpr int kMaxDepth = 17; // Config::apply maps maxdepth -1 to 16 (config.rs:59-63); raytracer visits depth 0..=max

struct RenderParams {
    GpuSceneView scene;
    rayca_config cfg;
    unsigned long long seed;
    uint32_t width, height;
    uint32_t shard_index, shard_count, tile_rows, local_rows;
    uint32_t row_begin, row_end; // unsharded: this launch covers image rows [row_begin, row_end)
    float inv_w, inv_h, aspect;
    float strate_count, offset, step, inv_gamma;
    uint32_t strate_u;
    unsigned char* rgba8;
    uint32_t* hit_blas;
    uint32_t* hit_prim;
    uint32_t* hit_src;
    float* depth;
    float* uv;
    float* rgba_f32;
    uint32_t tile_count, groups_x; // persistent scheduling: number of 16x8 pixel groups, groups per row
    // [0] shadow rays [1] secondary rays [2] node records [3] triangle vertex fetches [4] plane fetches
    // [5] next tile (persistent scheduling)
    unsigned long long* counters;
    // shared frame (multi-GPU, rayca_cuda_render_to_frame): rgba8 points into the OWNER GPU's frame (peer memory
    // over NVLink) and pixels go to their final row-major position
    uint32_t direct_rows; // 1: write at the image row even when sharded
    // Progressive accumulation (rayca_cuda_render_accumulate): this launch traces samples [sample_first, sample_first +
    // sample_count) of the cfg.samples_per_pixel the sequence is heading for (which fixes the strata, scene.rs:125-131),
    // starts every pixel's sum from `accum` (when accum_load) and leaves the new sum there; the pixel written is
    // sum / sample_div.  A plain render has sample_first 0, sample_count = sample_div = samples_per_pixel, accum null.
    float4* accum;
    uint32_t sample_first, sample_count, accum_load;
    float sample_div;
    // RAYCA_FLAG_SHARD_GROUPS: the shard owns pixel GROUPS (one thread block each) in a diagonal checkerboard —
    // group (gx, gy) belongs to shard (gx + gy) mod gshard_count — instead of whole row tiles; rows are then image rows
    // (shard_count = 1 for the row arithmetic) and grid.x enumerates the shard's groups of a group row
    uint32_t gshard_index, gshard_count, groups_x_full;
};
constexpr int kCounterSlots = 8;

// Philox4x32-10 keyed by the seed, counter = (pixel, sample, depth, dimension).  The reference draws
// from thread-local unseeded fastrand (sampler/cosine.rs:66-67), which is not reproducible; both
// this kernel and the oracle use this keyed generator instead (documented deviation).
__device__ __forceinline__ float rng_f32(unsigned long long seed, uint32_t pixel, uint32_t sample, uint32_t depth, uint32_t dim) {
    uint32_t c0 = pixel, c1 = sample, c2 = depth, c3 = dim;
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return (float)(c0 >> 8) * (1.0f / 16777216.0f); // 24 random bits, like fastrand::f32
}

// Vec3::close against Y_AXIS with the derived lexicographic PartialOrd (vec3.rs:26,108-111)
__device__ __forceinline__ bool close_to_y(V3 w) {
    float d[3] = {fabsf(w.x - 0.0f), fabsf(w.y - 1.0f), fabsf(w.z - 0.0f)};
    for (int i = 0; i < 3; ++i) {
        if (d[i] < kEps) return true;
        if (d[i] > kEps) return false;
        if (!(d[i] == kEps)) return false;
    }
    return false; // 4th lane: 0 < 0 is false
}

// CosineSampler::get_random_dir (sampler/cosine.rs:65-88)
__device__ __forceinline__ V3 cosine_dir(const RenderParams& P, const Surf& sf, uint32_t pixel, uint32_t sample, uint32_t depth, uint32_t ls) {
    float e1 = rng_f32(P.seed, pixel, sample, depth, 2 * ls);
    float e2 = rng_f32(P.seed, pixel, sample, depth, 2 * ls + 1);
    float theta = acosf(sqrtf(e1));
    float omega = 2.0f * kPi * e2;
    V3 sv{cosf(omega) * sinf(theta), sinf(omega) * sinf(theta), cosf(theta)};
    V3 w = sf.normal;
    V3 a = close_to_y(w) ? V3{1.0f, 0.0f, 0.0f} : V3{0.0f, 1.0f, 0.0f};
    V3 u = normalize(cross(a, w));
    V3 v = cross(w, u);
    return (sv.x * u + sv.y * v) + sv.z * w;
}

struct PixelCounters {
    uint32_t shadow = 0, secondary = 0;
    WalkCounters walk;
};

// NextEventEstimationSampler::get_direct_lighting, point lights (sampler/nee.rs:127-216).
// The shadow rays of up to 32 light samples are walked first (only the hit geometry is live), then the
// material is evaluated once and the unoccluded samples are shaded, summed in the reference's order —
// a reordering of independent computations, so every value is the one the reference computes.
template <int MODE, bool COUNT, bool LAZY_MATERIAL>
__device__ __forceinline__ Col nee_direct(const RenderParams& P, Surf& sf, PixelCounters& pc) {
    const GpuSceneView& s = P.scene;
    Col ret{0.0f, 0.0f, 0.0f, 1.0f};
    const uint32_t ls = P.cfg.light_samples;
    const uint32_t total = s.light_count * ls;
    bool evaluated = !LAZY_MATERIAL;
    for (uint32_t base = 0; base < total; base += 32u) {
        const uint32_t n = min(32u, total - base);
        uint32_t lit = 0;
        for (uint32_t j = 0; j < n; ++j) {
            const GpuLight& L = s.lights[(base + j) / ls];
            V3 d = V3{L.pos[0], L.pos[1], L.pos[2]} - sf.point;
            RayD shadow = make_ray(sf.next_origin, normalize(d));
            HitD sh;
            sh.t = length(d);
            pc.shadow++;
            if (!trace_ray<MODE, true, COUNT>(s, shadow, sh, &pc.walk)) lit |= 1u << j;
        }
        if (lit && !evaluated) {
            eval_material(s, sf);
            evaluated = true;
        }
        for (uint32_t j = 0; j < n; ++j) {
            Col x{0.0f, 0.0f, 0.0f, 1.0f};
            if ((lit >> j) & 1u) {
                const GpuLight& L = s.lights[(base + j) / ls];
                V3 d = V3{L.pos[0], L.pos[1], L.pos[2]} - sf.point;
                V3 omega = normalize(d);
                Col le = point_intensity(L, sf.point);
                Col brdf = surf_brdf(s, sf, omega);
                float r2 = norm2(d);
                float d_omega = 1.0f / r2;
                float ndo = clampf(dot(sf.normal, omega), 0.0f, 1.0f);
                x = cscale(cscale(cmul(le, brdf), ndo), d_omega);
            }
            ret = cadd(ret, x);
        }
    }
    return ret;
}

// Pathtracer::trace_impl (integrator/pathtracer.rs:68-106) for light_samples == 1 chains, unrolled
// into a forward pass (trace, shade, spawn) and a backward pass that folds the radiance with the
// reference's operator order.
// BOUNCE = false compiles the indirect part out (the host picks it when the configuration cannot
// spawn an indirect ray: max_depth <= 1 with NEE, 0 without) — the deterministic primary + shadow slice.
template <int MODE, bool COUNT, bool BOUNCE>
__device__ __forceinline__ bool pathtracer(const RenderParams& P, RayD ray, uint32_t pixel, uint32_t sample, Col& out, HitD& primary,
                                           PixelCounters& pc) {
    const GpuSceneView& s = P.scene;
    const bool nee = P.cfg.direct_sampler == RAYCA_SAMPLER_NEE;
    const uint32_t max_depth = P.cfg.max_depth;
    const uint32_t limit = nee ? max_depth - 1 : max_depth;
    Col direct[BOUNCE ? kMaxDepth : 1];
    Col brdf[BOUNCE ? kMaxDepth : 1];
    int levels = 0;       // number of shaded (non-terminal) levels stored
    bool have_tail = false;
    Col tail{0.0f, 0.0f, 0.0f, 1.0f}; // value of the deepest trace_impl that returned Some without recursing
    bool collect_emissive = true;
    for (uint32_t depth = 0; depth < max_depth && depth < (uint32_t)kMaxDepth; ++depth) {
        HitD h;
        bool hit = trace_ray<MODE, false, COUNT>(s, ray, h, &pc.walk);
        if (depth == 0) primary = h;
        if (!hit) break;
        Surf sf;
        load_surf<BOUNCE, MODE == 3>(s, ray, h, sf); // without bounces the material is evaluated lazily by nee_direct
        if (collect_emissive && sf.mat->kind == RAYCA_MATERIAL_PHONG && sf.mat->emissive) {
            tail = surf_color(s, sf);
            have_tail = true;
            break;
        }
        Col d{0.0f, 0.0f, 0.0f, 1.0f};
        if (nee) d = nee_direct<MODE, COUNT, !BOUNCE>(P, sf, pc);
        if (!BOUNCE || !(depth < limit)) { // no indirect: Some(direct + BLACK)
            tail = cadd(d, Col{0.0f, 0.0f, 0.0f, 1.0f});
            have_tail = true;
            break;
        }
        V3 wi = cosine_dir(P, sf, pixel, sample, depth, 0);
        direct[levels] = d;
        brdf[levels] = surf_brdf(s, sf, wi);
        ++levels;
        ray = make_ray(sf.next_origin, wi);
        if (depth + 1 < max_depth) pc.secondary++;
        collect_emissive = !nee;
    }
    if (levels == 0 && !have_tail) return false;
    // fold: value_k = direct_k + (BLACK + (BLACK + PI*brdf_k*value_{k+1}*1.0) / 1.0)
    bool have = have_tail;
    Col val = tail;
    for (int k = levels - 1; k >= 0; --k) {
        Col li{0.0f, 0.0f, 0.0f, 1.0f};
        if (have) li = cadd(li, cscale(cmul(cscale(brdf[k], kPi), val), 1.0f));
        li = cdiv(li, (float)P.cfg.light_samples);
        Col indirect = cadd(Col{0.0f, 0.0f, 0.0f, 1.0f}, li);
        val = cadd(direct[k], indirect);
        have = true;
    }
    out = val;
    return true;
}

// Raytracer::trace (integrator/raytracer.rs:16-76): mirror chain unrolled the same way.
// The light loop the Raytracer and the Scratcher share (raytracer.rs:34-67, scratcher.rs:45-76): a light
// contributes unless an OPAQUE surface lies between the point and the light (closest hit only).
template <int MODE, bool COUNT>
__device__ __forceinline__ void whitted_lights(const RenderParams& P, const Surf& sf, Col& acc, PixelCounters& pc) {
    const GpuSceneView& s = P.scene;
    for (uint32_t li = 0; li < s.light_count; ++li) {
        const GpuLight& L = s.lights[li];
        V3 lpos{L.pos[0], L.pos[1], L.pos[2]};
        V3 ldir;
        if (L.kind == RAYCA_LIGHT_POINT) ldir = -normalize(sf.point - lpos); // light/point.rs:50-54
        else ldir = V3{L.dir[0], L.dir[1], L.dir[2]};
        RayD shadow = make_ray(sf.next_origin, ldir);
        HitD sh;
        pc.shadow++;
        bool lit = true;
        if (trace_ray<MODE, false, COUNT>(s, shadow, sh, &pc.walk)) {
            float ldist = length(sf.point - lpos); // light/point.rs:32-35, directional.rs:34-37
            if (!(sh.t > ldist)) {
                Surf ss;
                load_surf<true, MODE == 3>(s, shadow, sh, ss);
                lit = surf_color(s, ss).a < 1.0f - kF32Eps; // Color::is_transparent
            }
        }
        if (lit) {
            Col intensity = (L.kind == RAYCA_LIGHT_POINT)
                                ? point_intensity(L, sf.point)
                                : cscale(Col{L.color[0], L.color[1], L.color[2], L.color[3]}, L.intensity);
            acc = cadd(acc, surf_radiance(s, sf, intensity, ldir));
        }
    }
}

// integrator/scratcher.rs:16-89.  The reference recurses twice per hit (transmission through a transparent surface,
// then reflection) and combines the results with functions of the hit (Color::over, Hit::get_radiance), so the
// recursion is run here with an explicit stack of frames: stage 1 waits for the transmitted colour, stage 3 for the
// reflected one; additions happen in the reference's order (transmission, lights, reflection).
template <int MODE, bool COUNT>
__device__ __noinline__ bool scratcher(const RenderParams& P, RayD ray, Col& out, HitD& primary, PixelCounters& pc) {
    const GpuSceneView& s = P.scene;
    struct Frame {
        Surf sf;
        Col acc;
        V3 refl;
        int stage;
    };
    Frame st[kMaxDepth];
    int sp = 0; // frames on the stack = depth of the call being entered
    bool calling = true, ret_ok = false;
    Col ret{0.0f, 0.0f, 0.0f, 1.0f};
    for (;;) {
        if (calling) { // trace(ray, depth = sp)
            HitD h;
            bool ok = (uint32_t)sp <= P.cfg.max_depth && sp < kMaxDepth;
            if (ok) {
                ok = trace_ray<MODE, false, COUNT>(s, ray, h, &pc.walk);
                if (sp == 0) primary = h;
            }
            if (!ok) { // None
                ret_ok = false;
                calling = false;
                if (sp == 0) break;
                continue;
            }
            Frame& f = st[sp++];
            load_surf<true, MODE == 3>(s, ray, h, f.sf);
            f.acc = Col{0.0f, 0.0f, 0.0f, 1.0f};
            f.stage = 2;
            if (surf_color(s, f.sf).a < 1.0f - kF32Eps) { // hit.is_transparent(): continue below the surface (hit.rs:178-187)
                ray = make_ray(f.sf.point + (-f.sf.normal) * kRayBias, f.sf.ray_dir);
                if ((uint32_t)sp <= P.cfg.max_depth) pc.secondary++;
                f.stage = 1;
                continue;
            }
        } else { // a child returned into the top frame
            Frame& f = st[sp - 1];
            if (f.stage == 1) {
                if (ret_ok) { // transmit_color.over(hit.get_color()); pixel_color += transmit_color (color/mod.rs:149-154)
                    const Col top = surf_color(s, f.sf);
                    Col t;
                    t.r = top.r * top.a + ret.r * (1.0f - top.a);
                    t.g = top.g * top.a + ret.g * (1.0f - top.a);
                    t.b = top.b * top.a + ret.b * (1.0f - top.a);
                    t.a = 1.0f;
                    f.acc = cadd(f.acc, t);
                }
                f.stage = 2;
            } else { // stage 3: reflection
                if (ret_ok) f.acc = cadd(f.acc, surf_radiance(s, f.sf, ret, f.refl));
                ret = f.acc;
                ret_ok = true;
                --sp;
                if (sp == 0) break;
                continue;
            }
        }
        // stage 2: direct light, then the reflection call
        Frame& f = st[sp - 1];
        whitted_lights<MODE, COUNT>(P, f.sf, f.acc, pc);
        f.refl = normalize(reflect(f.sf.ray_dir, f.sf.normal)); // hit.rs:93-101
        ray = make_ray(f.sf.next_origin, f.refl);
        if ((uint32_t)sp <= P.cfg.max_depth) pc.secondary++;
        f.stage = 3;
        calling = true;
    }
    out = ret;
    return ret_ok;
}

template <int MODE, bool COUNT>
__device__ __forceinline__ bool raytracer(const RenderParams& P, RayD ray, Col& out, HitD& primary, PixelCounters& pc) {
    const GpuSceneView& s = P.scene;
    Col ambient[kMaxDepth], light[kMaxDepth], spec[kMaxDepth];
    int levels = 0;
    for (uint32_t depth = 0; depth <= P.cfg.max_depth && depth < (uint32_t)kMaxDepth; ++depth) {
        HitD h;
        bool hit = trace_ray<MODE, false, COUNT>(s, ray, h, &pc.walk);
        if (depth == 0) primary = h;
        if (!hit) break;
        Surf sf;
        load_surf<true, MODE == 3>(s, ray, h, sf);
        Col lc{0.0f, 0.0f, 0.0f, 1.0f};
        whitted_lights<MODE, COUNT>(P, sf, lc, pc);
        ambient[levels] = surf_color(s, sf);
        light[levels] = lc;
        spec[levels] = surf_specular(s, sf);
        ++levels;
        V3 refl = normalize(reflect(ray.d, sf.normal)); // hit.rs:93-101
        ray = make_ray(sf.next_origin, refl);
        if (depth + 1 <= P.cfg.max_depth) pc.secondary++;
    }
    if (levels == 0) return false;
    Col val{0.0f, 0.0f, 0.0f, 1.0f};
    bool have = false;
    for (int k = levels - 1; k >= 0; --k) {
        Col lc = light[k];
        if (have) lc = cadd(lc, cscale(cmul(val, spec[k]), 1.0f));
        val = cadd(ambient[k], lc);
        have = true;
    }
    out = val;
    return true;
}

// ---- QuadLight (rayca-model/src/light/quad.rs): corner k of a, a+ab, (a+ab)+ac, a+ac (get_a / get_b :48-70)
__device__ __forceinline__ V3 quad_corner(const GpuLight& L, int k) {
    const V3 a{L.pos[0], L.pos[1], L.pos[2]}, ab{L.ab[0], L.ab[1], L.ab[2]}, ac{L.ac[0], L.ac[1], L.ac[2]};
    switch (k & 3) {
        case 0: return a;
        case 1: return a + ab;
        case 2: return (a + ab) + ac;
        default: return a + ac;
    }
}
// Light::get_intensity (light/mod.rs:54-60); quad: theta / gamma / get_radiance / get_intensity (quad.rs:72-104)
__device__ __forceinline__ Col light_intensity_at(const GpuLight& L, V3 frag, V3 n) {
    const Col lc = cscale(Col{L.color[0], L.color[1], L.color[2], L.color[3]}, L.intensity);
    if (L.kind == RAYCA_LIGHT_DIRECTIONAL) return lc;
    if (L.kind == RAYCA_LIGHT_POINT) return point_intensity(L, frag);
    V3 ret{0.0f, 0.0f, 0.0f};
#pragma unroll 1
    for (int e = 0; e < 4; ++e) {
        const V3 ra = quad_corner(L, e) - frag, rb = quad_corner(L, e + 1) - frag;
        const float theta = acosf(dot(normalize(ra), normalize(rb)));
        ret = ret + theta * normalize(cross(ra, rb));
    }
    ret = ret / 2.0f;
    return cscale(lc, dot(ret, n));
}
__device__ __forceinline__ bool surf_emissive(const Surf& sf) { // primitive.rs:122-130 / phong.rs:50-52
    return sf.mat->kind == RAYCA_MATERIAL_PHONG && sf.mat->emissive;
}

// ---- the general path tracer: everything pathtracer<> above leaves out so that the bench path stays small —
// hemisphere bounces (sampler/hemisphere.rs), quad lights under next event estimation (sampler/nee.rs:46-93) and
// russian roulette (pathtracer.rs:37-50, ray.rs:95-105).  One path per sample (light_samples = 1 whenever there are
// bounces), so radiance = sum over depths of (product of the bounce factors so far) x (direct light at that depth):
// accumulated forwards, which lets roulette paths run without a depth array.  The reference nests the same terms
// from the tail; the two differ by rounding only, inside the tolerance the stochastic cases are compared with.
constexpr uint32_t kInternalPathGeneral = 6u;  // launch-time selector next to the RAYCA_INTEGRATOR_* values
constexpr uint32_t kRouletteDepthCap = 64u;    // a path still alive after this many bounces stops (oracle: same cap)

__device__ __forceinline__ V3 hemisphere_dir(const RenderParams& P, const Surf& sf, uint32_t pixel, uint32_t sample, uint32_t depth,
                                             uint32_t ls = 0) {
    float e1 = rng_f32(P.seed, pixel, sample, depth, 2 * ls);
    float e2 = rng_f32(P.seed, pixel, sample, depth, 2 * ls + 1);
    float theta = acosf(e1);
    float omega = 2.0f * kPi * e2;
    V3 sv{cosf(omega) * sinf(theta), sinf(omega) * sinf(theta), cosf(theta)};
    V3 w = sf.normal;
    V3 a = close_to_y(w) ? V3{1.0f, 0.0f, 0.0f} : V3{0.0f, 1.0f, 0.0f};
    V3 u = normalize(cross(a, w));
    V3 v = normalize(cross(w, u));
    return (sv.x * u + sv.y * v) + sv.z * w;
}

// Material::get_t: the specular weight (phong.rs:64-71, ggx.rs:54-61; todo!() for PBR, refused at launch)
__device__ __forceinline__ float surf_t(const Surf& sf) {
    const GpuMaterial* m = sf.mat;
    const float* kd = m->kind == RAYCA_MATERIAL_PHONG ? m->c : m->a;
    const float* ks = m->d;
    const float kd_avg = (((kd[0] + kd[1]) + kd[2]) + 0.0f) / 3.0f, ks_avg = (((ks[0] + ks[1]) + ks[2]) + 0.0f) / 3.0f;
    if (ks_avg == 0.0f && kd_avg == 0.0f) return 1.0f;
    const float t = ks_avg / (ks_avg + kd_avg);
    return m->kind == RAYCA_MATERIAL_GGX ? fmaxf(t, 0.25f) : t;
}
// lambertian::get_random_dir (lambertian.rs:31-68) / ggx::get_random_dir (ggx.rs:146-186)
__device__ __forceinline__ V3 material_dir(const GpuSceneView& s, const Surf& sf, float e0, float e1, float e2) {
    const bool is_phong = sf.mat->kind == RAYCA_MATERIAL_PHONG;
    const float t = surf_t(sf);
    float theta;
    if (e0 <= t) {
        if (is_phong) theta = acosf(clampf(powf(e1, 1.0f / (sf.mat->p0 + 1.0f)), -1.0f, 1.0f));
        else theta = atanf((surf_roughness(s, sf) * sqrtf(e1)) / sqrtf(1.0f - e1));
    } else {
        theta = acosf(clampf(sqrtf(e1), -1.0f, 1.0f));
    }
    const float omega = 2.0f * kPi * e2;
    const V3 sv{cosf(omega) * sinf(theta), sinf(omega) * sinf(theta), cosf(theta)};
    const V3 w = (is_phong && e0 <= t) ? normalize(reflect(sf.ray_dir, sf.normal)) : sf.normal;
    const V3 a = close_to_y(w) ? V3{1.0f, 0.0f, 0.0f} : V3{0.0f, 1.0f, 0.0f};
    const V3 u = normalize(cross(a, w));
    const V3 v = cross(w, u);
    const V3 dir = (sv.x * u + sv.y * v) + sv.z * w;
    if (!is_phong && e0 <= t) { // the sampled vector is the half vector: reflect -omega_o about it
        const V3 mo = -sf.view_dir;
        return mo - (2.0f * dot(mo, dir)) * dir;
    }
    return dir;
}
// lambertian::get_specular_component (lambertian.rs:70-76) / ggx::get_specular_component (ggx.rs:188-203)
__device__ __forceinline__ Col surf_specular_component(const GpuSceneView& s, const Surf& sf, V3 wi) {
    if (sf.mat->kind == RAYCA_MATERIAL_PHONG) {
        const float sh = sf.mat->p0;
        const float ndw = clampf(dot(sf.normal, wi), 0.0f, 1.0f);
        return cdiv(cscale(cscale(ld_col(sf.mat->d), ndw), sh + 2.0f), sh + 1.0f);
    }
    const V3 wo = sf.view_dir, n = sf.normal;
    const float odn = clampf(dot(wo, n), 0.0f, 1.0f);
    const V3 h = normalize(wi + wo);
    const float ndh = clampf(dot(n, h), 0.0f, 1.0f);
    if (odn == 0.0f || ndh == 0.0f) return Col{0.0f, 0.0f, 0.0f, 1.0f};
    const Col ks = surf_specular(s, sf);
    const float a = surf_roughness(s, sf);
    const float hdw = clampf(dot(h, wi), 0.0f, 1.0f);
    const Col f = cadd(ks, cscale(csub(Col{1.0f, 1.0f, 1.0f, 1.0f}, ks), pow5(1.0f - fabsf(dot(wi, h)))));
    return cdiv(cscale(cscale(f, ggx_g1(a, wi, n) * ggx_g1(a, wo, n)), hdw), odn * ndh);
}

// NextEventEstimationSampler::get_direct_lighting with point AND quad lights; dimensions as in the oracle:
// [0, 2 LS) bounce direction, [2 LS, 3 LS) roulette, then two per (light, light sample)
// One next-event-estimation sample of light `li` (sampler/nee.rs:46-137): its direction and its value
template <int MODE, bool COUNT>
__device__ __forceinline__ Col nee_sample(const RenderParams& P, const Surf& sf, uint32_t li, uint32_t i, uint32_t pixel, uint32_t sample,
                                          uint32_t depth, PixelCounters& pc, V3& omega) {
    const GpuSceneView& s = P.scene;
    const GpuLight& L = s.lights[li];
    const uint32_t ls = P.cfg.light_samples;
    Col xs{0.0f, 0.0f, 0.0f, 1.0f};
    if (L.kind == RAYCA_LIGHT_QUAD) {
        const uint32_t strate = P.cfg.light_stratify ? __float2uint_rz(sqrtf((float)ls)) : 1u; // config.rs:73-79
        const float fstrate = (float)strate;
        const V3 ab{L.ab[0], L.ab[1], L.ab[2]}, ac{L.ac[0], L.ac[1], L.ac[2]}, qn{L.normal[0], L.normal[1], L.normal[2]};
        const uint32_t dim = 3u * ls + 2u * (li * ls + i);
        const float u1 = rng_f32(P.seed, pixel, sample, depth, dim) / fstrate;
        const float u2 = rng_f32(P.seed, pixel, sample, depth, dim + 1u) / fstrate;
        V3 x1 = (quad_corner(L, 0) + u1 * ab) + u2 * ac;
        if (P.cfg.light_stratify) x1 = x1 + ((ab / fstrate) * (float)(i % strate) + (ac / fstrate) * (float)(i / strate));
        const V3 to_x1 = x1 - sf.point;
        omega = normalize(to_x1);
        const RayD shadow = make_ray(sf.next_origin, omega);
        HitD sh;
        pc.shadow++;
        if (trace_ray<MODE, false, COUNT>(s, shadow, sh, &pc.walk)) { // counts only when the closest hit is a light
            const GpuMaterial* sm = s.materials + __float_as_uint(__ldg(s.tri_attr + 12 * (size_t)sh.prim + 9).w);
            if (sm->kind == RAYCA_MATERIAL_PHONG && sm->emissive) {
                const Col le = cscale(Col{L.color[0], L.color[1], L.color[2], L.color[3]}, L.intensity);
                const Col brdf = surf_brdf(s, sf, omega);
                const float d_omega = dot(qn, omega) / norm2(to_x1);
                const float n_dot = fminf(fmaxf(dot(sf.normal, omega), 0.0f), 1.0f);
                xs = cscale(cscale(cmul(cscale(le, L.area), brdf), n_dot), d_omega);
            }
        }
    } else { // point light (nee.rs:95-137)
        const V3 to_x1 = V3{L.pos[0], L.pos[1], L.pos[2]} - sf.point;
        const float dist = length(to_x1);
        omega = normalize(to_x1);
        const RayD shadow = make_ray(sf.next_origin, omega);
        HitD sh;
        pc.shadow++;
        bool occluded = false;
        if (trace_ray<MODE, false, COUNT>(s, shadow, sh, &pc.walk)) occluded = sh.t < dist;
        if (!occluded) {
            const Col le = point_intensity(L, sf.point);
            const Col brdf = surf_brdf(s, sf, omega);
            const float d_omega = 1.0f / norm2(to_x1);
            const float n_dot = fminf(fmaxf(dot(sf.normal, omega), 0.0f), 1.0f);
            xs = cscale(cscale(cmul(le, brdf), n_dot), d_omega);
        }
    }
    return xs;
}
// NextEventEstimationSampler::get_direct_lighting with point AND quad lights; dimensions as in the oracle:
// [0, 2 LS) bounce direction, [2 LS, 3 LS) roulette, then two per (light, light sample)
template <int MODE, bool COUNT>
__device__ __forceinline__ Col nee_general(const RenderParams& P, const Surf& sf, uint32_t pixel, uint32_t sample, uint32_t depth,
                                           PixelCounters& pc) {
    Col ret{0.0f, 0.0f, 0.0f, 1.0f};
    for (uint32_t li = 0; li < P.scene.light_count; ++li)
#pragma unroll 1
        for (uint32_t i = 0; i < P.cfg.light_samples; ++i) {
            V3 omega;
            ret = cadd(ret, nee_sample<MODE, COUNT>(P, sf, li, i, pixel, sample, depth, pc, omega));
        }
    return ret;
}

// lambertian::get_pdf (lambertian.rs:17-26) / ggx::get_pdf (ggx.rs:129-144)
__device__ __forceinline__ float material_pdf(const GpuSceneView& s, const Surf& sf, V3 omega) {
    const float kFrac2Pi = 0.636619772367581343075535053490057448f;
    if (sf.mat->kind == RAYCA_MATERIAL_PHONG) {
        const V3 refl = normalize(reflect(sf.ray_dir, sf.normal));
        const float rdw = clampf(dot(refl, omega), 0.0f, 1.0f);
        const float sh = sf.mat->p0;
        const float spec = (sh + 1.0f) * kFrac2Pi * powf(rdw, sh);
        const float diff = clampf(dot(sf.normal, omega), 0.0f, 1.0f) * kFrac1Pi;
        const float t = surf_t(sf);
        return (1.0f - t) * diff + t * spec;
    }
    const V3 h = normalize(sf.view_dir + omega);
    const float hdw = clampf(dot(h, omega), 0.0f, 1.0f);
    if (hdw == 0.0f) return 0.0f;
    const float ndh = clampf(dot(sf.normal, h), 0.0f, 1.0f);
    const float spec = ggx_d(surf_roughness(s, sf), h, sf.normal) * ndh / (4.0f * hdw);
    const float dif = clampf(dot(sf.normal, omega), 0.0f, 1.0f) * kFrac1Pi;
    const float t = surf_t(sf);
    return (1.0f - t) * dif + t * spec;
}
// NextEventEstimationSample::get_pdf (nee.rs:44-65): QuadLight::intersects over its wide triangles, first hit in order
__device__ __forceinline__ float nee_pdf(const GpuLight& L, const Surf& sf, V3 omega) {
    if (L.kind != RAYCA_LIGHT_QUAD) return 0.0f;
    const RayD ray = make_ray(sf.next_origin, omega);
    const float4 a = make_float4(L.wide[0][0], L.wide[0][1], L.wide[0][2], 0.0f), b = make_float4(L.wide[1][0], L.wide[1][1], L.wide[1][2], 0.0f),
                 c = make_float4(L.wide[2][0], L.wide[2][1], L.wide[2][2], 0.0f), d = make_float4(L.wide[3][0], L.wide[3][1], L.wide[3][2], 0.0f);
    float t, u, v;
    V3 p;
    if (!tri_test(a, c, b, ray, t, u, v, p) && !tri_test(a, d, c, ray, t, u, v, p)) return 0.0f;
    if (L.area == 0.0f) return 0.0f;
    const float nl = clampf(dot(V3{L.normal[0], L.normal[1], L.normal[2]}, omega), 0.0f, 1.0f);
    if (nl == 0.0f) return 0.0f;
    return norm2(p - sf.point) / (L.area * nl);
}
constexpr uint32_t kMaxMisSamples = 16; // lights x light_samples the MIS estimator keeps per hit (refused beyond)
// MultipleImportanceSampling::get_direct_lighting (sampler/mis.rs:41-72) + BrdfSampler::sample_direct (brdf.rs:48-66)
template <int MODE, bool COUNT>
__device__ __noinline__ Col mis_general(const RenderParams& P, const Surf& sf, uint32_t pixel, uint32_t sample, uint32_t depth,
                                        PixelCounters& pc) {
    const GpuSceneView& s = P.scene;
    const uint32_t ls = P.cfg.light_samples;
    V3 omegas[kMaxMisSamples];
    Col xs[kMaxMisSamples];
    uint32_t n = 0;
    for (uint32_t li = 0; li < s.light_count; ++li)
#pragma unroll 1
        for (uint32_t i = 0; i < ls && n < kMaxMisSamples; ++i, ++n) xs[n] = nee_sample<MODE, COUNT>(P, sf, li, i, pixel, sample, depth, pc, omegas[n]);
    const V3 omega_b = material_dir(s, sf, rng_f32(P.seed, pixel, sample, depth, 0x20000u), rng_f32(P.seed, pixel, sample, depth, 0x20001u),
                                    rng_f32(P.seed, pixel, sample, depth, 0x20002u));
    const float pdf_b = material_pdf(s, sf, omega_b);
    Col x_b{0.0f, 0.0f, 0.0f, 1.0f};
    {
        const RayD shadow = make_ray(sf.next_origin, omega_b);
        HitD sh;
        pc.shadow++;
        if (trace_ray<MODE, false, COUNT>(s, shadow, sh, &pc.walk)) {
            const GpuMaterial* sm = s.materials + __float_as_uint(__ldg(s.tri_attr + 12 * (size_t)sh.prim + 9).w);
            if (sm->kind == RAYCA_MATERIAL_PHONG && sm->emissive)
                x_b = cmul(ld_col(sm->b), cadd(surf_diffuse(s, sf), surf_specular_component(s, sf, omega_b)));
        }
    }
    Col ret{0.0f, 0.0f, 0.0f, 1.0f};
#pragma unroll 1
    for (uint32_t k = 0; k <= n; ++k) { // k == n: the Brdf sample
        const V3 omega = k < n ? omegas[k] : omega_b;
        float pdf_nee = 0.0f;
        for (uint32_t li = 0; li < s.light_count; ++li) {
            const float pl = nee_pdf(s.lights[li], sf, omega);
            for (uint32_t i = 0; i < ls; ++i) pdf_nee += pl; // get_pdf_nee sums one term per light SAMPLE
        }
        pdf_nee = pdf_nee / (float)n;
        const float pdf_brdf = k < n ? material_pdf(s, sf, omega) : pdf_b;
        const float den = pdf_nee * pdf_nee + pdf_brdf * pdf_brdf;
        const float num = k < n ? pdf_nee * pdf_nee : pdf_brdf * pdf_brdf;
        const float w = den == 0.0f ? 0.0f : num / den;
        ret = cadd(ret, cscale(k < n ? xs[k] : x_b, w));
    }
    return ret;
}

template <int MODE, bool COUNT>
__device__ __noinline__ bool pathtracer_general(const RenderParams& P, RayD ray, uint32_t pixel, uint32_t sample, Col& out,
                                                HitD& primary, PixelCounters& pc) {
    const GpuSceneView& s = P.scene;
    const bool mis = P.cfg.direct_sampler == RAYCA_SAMPLER_MIS;
    const bool nee = P.cfg.direct_sampler == RAYCA_SAMPLER_NEE || mis; // "a direct sampler is set" (pathtracer.rs:86-90)
    const bool rr = P.cfg.russian_roulette != 0;
    const bool hemisphere = P.cfg.indirect_sampler == RAYCA_SAMPLER_HEMISPHERE;
    const bool brdf_sampler = P.cfg.indirect_sampler == RAYCA_SAMPLER_BRDF;
    const uint32_t max_depth = P.cfg.max_depth;
    const uint32_t limit = nee ? max_depth - 1u : max_depth;
    Col sum{0.0f, 0.0f, 0.0f, 1.0f};
    Col factor{1.0f, 1.0f, 1.0f, 1.0f};      // product of the samplers' get_radiance factors down to this depth
    Col throughput{1.0f, 1.0f, 1.0f, 1.0f};  // Ray::throughput, only used by the roulette
    bool collect_emissive = true, any = false;
    for (uint32_t depth = 0;; ++depth) {
        if (!rr && depth >= max_depth) break;
        if (rr && depth >= kRouletteDepthCap) break;
        HitD h;
        const bool hit = trace_ray<MODE, false, COUNT>(s, ray, h, &pc.walk);
        if (depth == 0) primary = h;
        if (!hit) break;
        any = true;
        Surf sf;
        load_surf<true, MODE == 3>(s, ray, h, sf);
        if (collect_emissive && surf_emissive(sf)) {
            sum = cadd(sum, cmul(factor, surf_color(s, sf)));
            break;
        }
        if (mis) sum = cadd(sum, cmul(factor, mis_general<MODE, COUNT>(P, sf, pixel, sample, depth, pc)));
        else if (nee) sum = cadd(sum, cmul(factor, nee_general<MODE, COUNT>(P, sf, pixel, sample, depth, pc)));
        if (!(rr || depth < limit)) break;
        V3 wi;
        if (brdf_sampler) // sampler/brdf.rs:70-73; lobe choice at dimension 0x10000 (oracle: dim_lobe)
            wi = material_dir(s, sf, rng_f32(P.seed, pixel, sample, depth, 0x10000u), rng_f32(P.seed, pixel, sample, depth, 0),
                              rng_f32(P.seed, pixel, sample, depth, 1));
        else
            wi = hemisphere ? hemisphere_dir(P, sf, pixel, sample, depth) : cosine_dir(P, sf, pixel, sample, depth, 0);
        const Col brdf = surf_brdf(s, sf, wi);
        const float cosine_law = fminf(fmaxf(dot(sf.normal, wi), 0.0f), 1.0f);
        float weight = 1.0f;
        if (rr) {
            const Col next_tp = cscale(cmul(cscale(throughput, 2.0f * kPi), brdf), cosine_law);
            const float q = 1.0f - fminf(fmaxf(fmaxf(next_tp.r, next_tp.g), next_tp.b), 1.0f);
            const float e = fminf(fmaxf(rng_f32(P.seed, pixel, sample, depth, 2u * P.cfg.light_samples), 0.0f), 1.0f - kF32Eps);
            if (!(q < e)) break; // the path has not survived
            weight = 1.0f / (1.0f - q);
            throughput = cscale(next_tp, weight);
        }
        // hemisphere.rs:42-52: 2 PI * brdf * cosine_law * sample * weight; cosine.rs:90-99: PI * brdf * sample * weight
        // brdf.rs:80-90: sample * (cd + cs) * weight
        const Col f = brdf_sampler ? cscale(cadd(surf_diffuse(s, sf), surf_specular_component(s, sf, wi)), weight)
                      : hemisphere ? cscale(cscale(cscale(brdf, 2.0f * kPi), cosine_law), weight)
                                   : cscale(cscale(brdf, kPi), weight);
        factor = cmul(factor, f);
        ray = make_ray(sf.next_origin, wi);
        if (rr || depth + 1u < max_depth) pc.secondary++;
        collect_emissive = !nee;
    }
    out = sum;
    return any;
}

// Pathtracer::get_indirect_lighting with light_samples > 1 (integrator/pathtracer.rs:23-66): EVERY hit below the
// indirect depth limit spawns light_samples indirect rays and each of them recurses, so a pixel sample is a tree of
// up to LS^depth paths.  Walked depth-first on an explicit stack of levels {surface, path factor, next sample};
// radiance is accumulated forwards like pathtracer_general: sum += (product of f_k / LS down the path) * direct, where
// f_k is the indirect sampler's get_radiance factor.  The keyed RNG dimensions of sample `ls` at one (pixel, sample,
// depth) are the oracle's: 2 ls, 2 ls + 1 (+ 0x10000 + ls for the Brdf sampler's lobe choice) — sibling subtrees at
// the same depth draw the same numbers, on both sides.  Without russian roulette only (refused at launch), at most
// kMaxBranchDepth levels.
constexpr int kMaxBranchDepth = 8;
struct BranchLevel {
    Surf sf;
    Col factor;
    uint32_t next_ls;
};

template <int MODE, bool COUNT>
__device__ __noinline__ bool pathtracer_branching(const RenderParams& P, RayD ray, uint32_t pixel, uint32_t sample, Col& out,
                                                  HitD& primary, PixelCounters& pc) {
    const GpuSceneView& s = P.scene;
    const bool mis = P.cfg.direct_sampler == RAYCA_SAMPLER_MIS;
    const bool nee = P.cfg.direct_sampler == RAYCA_SAMPLER_NEE || mis;
    const bool hemisphere = P.cfg.indirect_sampler == RAYCA_SAMPLER_HEMISPHERE;
    const bool brdf_sampler = P.cfg.indirect_sampler == RAYCA_SAMPLER_BRDF;
    const uint32_t max_depth = P.cfg.max_depth;
    const uint32_t limit = nee ? max_depth - 1u : max_depth;
    const uint32_t LS = P.cfg.light_samples;
    Col sum{0.0f, 0.0f, 0.0f, 1.0f};
    out = sum;
    if (max_depth == 0) return false;
    BranchLevel stack[kMaxBranchDepth];
    int top = -1; // stack[top] is the hit at depth == top whose indirect rays are being spawned
    Col factor{1.0f, 1.0f, 1.0f, 1.0f};
    uint32_t depth = 0; // depth of `ray`
    bool collect_emissive = true, any = false;
    for (;;) {
        HitD h;
        const bool hit = trace_ray<MODE, false, COUNT>(s, ray, h, &pc.walk);
        if (depth == 0) {
            primary = h;
            any = hit;
        }
        if (hit) {
            Surf sf;
            load_surf<true, MODE == 3>(s, ray, h, sf);
            if (collect_emissive && surf_emissive(sf)) {
                sum = cadd(sum, cmul(factor, surf_color(s, sf)));
            } else {
                if (mis) sum = cadd(sum, cmul(factor, mis_general<MODE, COUNT>(P, sf, pixel, sample, depth, pc)));
                else if (nee) sum = cadd(sum, cmul(factor, nee_general<MODE, COUNT>(P, sf, pixel, sample, depth, pc)));
                if (depth < limit && top + 1 < kMaxBranchDepth) {
                    ++top; // == depth: levels are pushed in depth order
                    stack[top].sf = sf;
                    stack[top].factor = factor;
                    stack[top].next_ls = 0;
                }
            }
        }
        // the next indirect ray: the deepest level that still has samples to spawn
        bool found = false;
        while (top >= 0 && !found) {
            BranchLevel& L = stack[top];
            if (L.next_ls == LS) {
                --top;
            } else {
                const uint32_t ls = L.next_ls++;
                const uint32_t d = (uint32_t)top;
                // trace_impl(depth + 1) returns None before tracing anything once depth + 1 reaches max_depth
                if (d + 1u < max_depth) {
                    V3 wi;
                    if (brdf_sampler)
                        wi = material_dir(s, L.sf, rng_f32(P.seed, pixel, sample, d, 0x10000u + ls), rng_f32(P.seed, pixel, sample, d, 2u * ls),
                                          rng_f32(P.seed, pixel, sample, d, 2u * ls + 1u));
                    else
                        wi = hemisphere ? hemisphere_dir(P, L.sf, pixel, sample, d, ls) : cosine_dir(P, L.sf, pixel, sample, d, ls);
                    const Col brdf = surf_brdf(s, L.sf, wi);
                    const float cosine_law = fminf(fmaxf(dot(L.sf.normal, wi), 0.0f), 1.0f);
                    const Col f = brdf_sampler ? cscale(cadd(surf_diffuse(s, L.sf), surf_specular_component(s, L.sf, wi)), 1.0f)
                                  : hemisphere ? cscale(cscale(cscale(brdf, 2.0f * kPi), cosine_law), 1.0f)
                                               : cscale(cscale(brdf, kPi), 1.0f);
                    factor = cmul(L.factor, cdiv(f, (float)LS));
                    ray = make_ray(L.sf.next_origin, wi);
                    depth = d + 1u;
                    collect_emissive = !nee;
                    pc.secondary++;
                    found = true;
                }
            }
        }
        if (!found) break;
    }
    out = sum;
    return any;
}

// AnalyticDirect::trace (integrator/analyticdirect.rs:16-49)
template <int MODE, bool COUNT>
__device__ __forceinline__ bool analytic_direct(const RenderParams& P, const RayD& ray, Col& out, HitD& primary, PixelCounters& pc) {
    const GpuSceneView& s = P.scene;
    if (0u >= P.cfg.max_depth) return false;
    if (!trace_ray<MODE, false, COUNT>(s, ray, primary, &pc.walk)) return false;
    Surf sf;
    load_surf<true, MODE == 3>(s, ray, primary, sf);
    if (surf_emissive(sf)) {
        out = surf_color(s, sf);
        return true;
    }
    Col lc{0.0f, 0.0f, 0.0f, 1.0f};
    for (uint32_t li = 0; li < s.light_count; ++li) lc = cadd(lc, light_intensity_at(s.lights[li], sf.point, sf.normal));
    out = cmul(cscale(surf_diffuse(s, sf), kFrac1Pi), lc);
    return true;
}

// Direct::trace (integrator/direct.rs:16-88) with QuadLight::get_random_point (quad.rs:117-137); the two uniform
// numbers of light sample i of light l come from the keyed RNG at (pixel, sample, depth 0, 2*(l*light_samples+i)+{0,1})
template <int MODE, bool COUNT>
__device__ __forceinline__ bool direct_lighting(const RenderParams& P, const RayD& ray, uint32_t pixel, uint32_t sample, Col& out,
                                                HitD& primary, PixelCounters& pc) {
    const GpuSceneView& s = P.scene;
    if (0u >= P.cfg.max_depth) return false;
    if (!trace_ray<MODE, false, COUNT>(s, ray, primary, &pc.walk)) return false;
    Surf sf;
    load_surf<true, MODE == 3>(s, ray, primary, sf);
    if (surf_emissive(sf)) {
        out = surf_color(s, sf);
        return true;
    }
    const uint32_t ls = P.cfg.light_samples;
    const uint32_t strate = P.cfg.light_stratify ? __float2uint_rz(sqrtf((float)ls)) : 1u; // config.rs:73-79
    const float fstrate = (float)strate;
    Col lc{0.0f, 0.0f, 0.0f, 1.0f};
    for (uint32_t li = 0; li < s.light_count; ++li) {
        const GpuLight& L = s.lights[li];
        if (L.kind != RAYCA_LIGHT_QUAD) continue;
        const V3 ab{L.ab[0], L.ab[1], L.ab[2]}, ac{L.ac[0], L.ac[1], L.ac[2]}, qn{L.normal[0], L.normal[1], L.normal[2]};
        const V3 step1 = ab / fstrate, step2 = ac / fstrate;
        Col ld{0.0f, 0.0f, 0.0f, 1.0f};
#pragma unroll 1
        for (uint32_t i = 0; i < ls; ++i) {
            const float u1 = rng_f32(P.seed, pixel, sample, 0, 2u * (li * ls + i)) / fstrate;
            const float u2 = rng_f32(P.seed, pixel, sample, 0, 2u * (li * ls + i) + 1u) / fstrate;
            V3 x1 = (quad_corner(L, 0) + u1 * ab) + u2 * ac;
            if (P.cfg.light_stratify) x1 = x1 + (step1 * (float)(i % strate) + step2 * (float)(i / strate));
            const V3 to_hit = x1 - sf.point;
            const V3 omega = normalize(to_hit);
            const RayD shadow = make_ray(sf.next_origin, omega);
            HitD sh;
            pc.shadow++;
            if (trace_ray<MODE, false, COUNT>(s, shadow, sh, &pc.walk)) { // seen unless the closest hit is not a light
                const GpuMaterial* sm = s.materials + __float_as_uint(__ldg(s.tri_attr + 12 * (size_t)sh.prim + 9).w);
                if (!(sm->kind == RAYCA_MATERIAL_PHONG && sm->emissive)) continue;
            }
            const Col brdf = surf_brdf(s, sf, omega);
            const float r = length(to_hit);
            const float d_omega = dot(qn, omega) / (r * r); // len().powf(2.0)
            ld = cadd(ld, cscale(cscale(brdf, dot(sf.normal, omega)), d_omega));
        }
        const Col li_c = cscale(Col{L.color[0], L.color[1], L.color[2], L.color[3]}, L.intensity);
        ld = cdiv(cmul(cscale(li_c, L.area), ld), (float)ls);
        lc = cadd(lc, ld);
    }
    out = lc;
    return true;
}

// `(c * 255.0).clamp(0.0, 255.0) as u8` (color/rgba8.rs:75-84): truncation, NaN -> 0
__device__ __forceinline__ unsigned char quantize(float c) {
    float v = clampf(c * 255.0f, 0.0f, 255.0f);
    return (unsigned char)__float2uint_rz(v);
}

// One pixel: every sample's primary ray through the integrator, accumulation, gamma, RGBA8
// (rayca-soft/src/scene.rs:123-149).
// correct_gamma (color/mod.rs:175-180): one shared copy of powf
static __device__ __noinline__ float gamma_pow(float c, float inv_gamma) { return powf(c, inv_gamma); }

// image row of a launch-local row (row bands of the host path, interleaved tiles of a shard)
__device__ __forceinline__ uint32_t image_row(const RenderParams& P, uint32_t local_row) {
    if (P.shard_count > 1) {
        uint32_t tile = local_row / P.tile_rows;
        return (tile * P.shard_count + P.shard_index) * P.tile_rows + (local_row % P.tile_rows);
    }
    return local_row + P.row_begin;
}

__device__ __forceinline__ HitD no_hit() {
    HitD h;
    h.prim = 0xFFFFFFFFu;
    h.t = __int_as_float(0x7f800000);
    h.u = h.v = 0.0f;
    h.p = V3{0.0f, 0.0f, 0.0f};
    h.sphere = 0;
    return h;
}

// One sample of one pixel: primary ray (scene.rs:129-141) through the integrator (scene.rs:80-86).
// Primary ray of sample i of pixel (x, y) (scene.rs:129-141)
__device__ __forceinline__ RayD primary_ray(const RenderParams& P, uint32_t x, uint32_t y, uint32_t i) {
    const GpuCamera& cam = P.scene.camera;
    const TrsF ctrs{V3{cam.t[0], cam.t[1], cam.t[2]}, Q4{cam.q[0], cam.q[1], cam.q[2], cam.q[3]}, V3{cam.s[0], cam.s[1], cam.s[2]}};
    float ix = (float)(i % P.strate_u);
    float iy = (float)(i / P.strate_u);
    float xx = (2.0f * (((float)x + ix * P.step + P.offset) * P.inv_w) - 1.0f) * cam.angle * P.aspect;
    float yy = (1.0f - 2.0f * (((float)y + iy * P.step + P.offset) * P.inv_h)) * cam.angle;
    V3 dir = normalize(V3{xx, yy, -1.0f});
    // &camera_trs.trs * Ray::new(origin, dir): scale, rotate, translate (trs.rs:275-284); no re-normalise
    V3 wd = rotate(dir * ctrs.s, ctrs.q);
    V3 wo = rotate(scale_point(V3{0.0f, 0.0f, 0.0f}, ctrs.s), ctrs.q) + ctrs.t;
    return make_ray(wo, wd);
}

template <int INTEGRATOR, int MODE, bool COUNT, bool BOUNCE>
__device__ __forceinline__ void trace_sample(const RenderParams& P, uint32_t x, uint32_t y, uint32_t i, Col& c, HitD& ph, PixelCounters& pc) {
    const uint32_t pixel = y * P.width + x;
    RayD ray = primary_ray(P, x, y, i);
    c = Col{0.0f, 0.0f, 0.0f, 1.0f};
    bool ok;
    ph = no_hit();
    if (INTEGRATOR == (int)kInternalPathGeneral) {
        if (P.cfg.light_samples > 1 && !P.cfg.russian_roulette && P.cfg.max_depth > (P.cfg.direct_sampler == RAYCA_SAMPLER_NONE ? 0u : 1u))
            ok = pathtracer_branching<MODE, COUNT>(P, ray, pixel, i, c, ph, pc); // light_samples indirect rays per hit
        else
            ok = pathtracer_general<MODE, COUNT>(P, ray, pixel, i, c, ph, pc);
    } else if (INTEGRATOR == RAYCA_INTEGRATOR_PATHTRACER) {
        ok = pathtracer<MODE, COUNT, BOUNCE>(P, ray, pixel, i, c, ph, pc);
    } else if (INTEGRATOR == RAYCA_INTEGRATOR_RAYTRACER) {
        ok = raytracer<MODE, COUNT>(P, ray, c, ph, pc);
    } else if (INTEGRATOR == RAYCA_INTEGRATOR_SCRATCHER) {
        ok = scratcher<MODE, COUNT>(P, ray, c, ph, pc);
    } else if (INTEGRATOR == RAYCA_INTEGRATOR_ANALYTIC_DIRECT) {
        ok = analytic_direct<MODE, COUNT>(P, ray, c, ph, pc);
    } else if (INTEGRATOR == RAYCA_INTEGRATOR_DIRECT) {
        ok = direct_lighting<MODE, COUNT>(P, ray, pixel, i, c, ph, pc);
    } else { // Flat (integrator/flat.rs:16-28)
        ok = trace_ray<MODE, false, COUNT>(P.scene, ray, ph, &pc.walk);
        if (ok) {
            Surf sf;
            load_surf<true, MODE == 3>(P.scene, ray, ph, sf);
            c = surf_color(P.scene, sf);
        }
    }
    if (!ok) {
        c = Col{0.0f, 0.0f, 0.0f, 1.0f}; // unwrap_or(Color::BLACK) (scene.rs:80-86)
        if (INTEGRATOR == RAYCA_INTEGRATOR_PATHTRACER && P.cfg.max_depth == 0) ph.prim = 0xFFFFFFFFu;
    }
}

// `color /= spp; correct_gamma; *pixel = color.into()` (scene.rs:145-149) plus the optional parity planes.
// the colour half: `/ spp`, gamma, quantisation, RGBA8 (and the float plane)
__device__ __forceinline__ void write_pixel_color(const RenderParams& P, uint32_t x, uint32_t y, uint32_t local_row, Col color) {
    color = cdiv(color, P.sample_div);
    if (P.inv_gamma != 1.0f) { // powf(c, 1.0) == c in the reference's libm; skipping it keeps the bits
        color.r = gamma_pow(color.r, P.inv_gamma);
        color.g = gamma_pow(color.g, P.inv_gamma);
        color.b = gamma_pow(color.b, P.inv_gamma);
    }
    const size_t o = (size_t)((P.shard_count > 1 && !P.direct_rows) ? local_row : y) * P.width + x;
    uchar4 q = make_uchar4(quantize(color.r), quantize(color.g), quantize(color.b), quantize(color.a));
    reinterpret_cast<uchar4*>(P.rgba8)[o] = q;
    if (P.rgba_f32) reinterpret_cast<float4*>(P.rgba_f32)[o] = make_float4(color.r, color.g, color.b, color.a);
}
// the parity planes of the primary hit
__device__ __forceinline__ void write_pixel_aux(const RenderParams& P, uint32_t x, uint32_t y, uint32_t local_row, const HitD& primary) {
    const size_t o = (size_t)((P.shard_count > 1 && !P.direct_rows) ? local_row : y) * P.width + x;
    const bool hit = primary.prim != 0xFFFFFFFFu;
    if (P.hit_prim || P.hit_blas || P.hit_src) {
        uint32_t blas = 0xFFFFFFFFu, local = 0xFFFFFFFFu, src = 0xFFFFFFFFu;
        if (hit) {
            blas = 0;
            while (blas + 1 < P.scene.blas_count && primary.prim >= P.scene.blas_prim_offset[blas + 1]) ++blas;
            local = primary.prim - P.scene.blas_prim_offset[blas];
            src = __float_as_uint(P.scene.tri_pos[3 * (size_t)primary.prim].w);
        }
        if (P.hit_blas) P.hit_blas[o] = blas;
        if (P.hit_prim) P.hit_prim[o] = local;
        if (P.hit_src) P.hit_src[o] = src;
    }
    if (P.depth) P.depth[o] = hit ? primary.t : __int_as_float(0x7f800000);
    if (P.uv) {
        P.uv[2 * o] = hit ? primary.u : 0.0f;
        P.uv[2 * o + 1] = hit ? primary.v : 0.0f;
    }
}
__device__ __forceinline__ void write_pixel(const RenderParams& P, uint32_t x, uint32_t y, uint32_t local_row, Col color, const HitD& primary) {
    write_pixel_color(P, x, y, local_row, color);
    write_pixel_aux(P, x, y, local_row, primary);
}

// One warp, one 8x4 pixel tile.
//   1 spp: lane = pixel.
//   n spp: the tile's 32*n (pixel, sample) pairs are traced 32 at a time in pixel-major order, so the rays
//     a warp walks together are the samples of ONE or two pixels — far more coherent than 32 different
//     pixels (the 10M-triangle soup at 16 spp ran at 7 of 32 lanes active with lane = pixel).  Lane p owns
//     pixel p of the tile and adds that pixel's sample colours in sample order (shuffles), which is the
//     reference's `color += sample` order (scene.rs:129-144), so the f32 sum has the same bits.
template <int INTEGRATOR, int MODE, bool COUNT, bool BOUNCE>
__device__ __forceinline__ void render_tile(const RenderParams& P, uint32_t x0, uint32_t row0, uint32_t lane, PixelCounters& pc) {
    const uint32_t spp = P.sample_count; // samples traced by THIS launch (all of cfg.samples_per_pixel unless accumulating)
    const uint32_t x = x0 + (lane & 7u), local_row = row0 + (lane >> 3);
    const uint32_t y = image_row(P, local_row);
    const bool mine = x < P.width && y < P.row_end;
    Col color{0.0f, 0.0f, 0.0f, 1.0f};
    if (P.accum && P.accum_load && mine) { // the sum of the samples before sample_first, in sample order
        const float4 a = P.accum[(size_t)y * P.width + x];
        color = Col{a.x, a.y, a.z, a.w};
    }
    HitD primary = no_hit();
    const bool want_aux = P.hit_prim || P.hit_blas || P.hit_src || P.depth || P.uv;
#pragma unroll 1
    for (uint32_t k0 = 0; k0 < 32u * spp; k0 += 32u) {
        const uint32_t k = k0 + lane, pl = k / spp, i = P.sample_first + (k - pl * spp);
        const uint32_t sx = x0 + (pl & 7u), srow = row0 + (pl >> 3), sy = image_row(P, srow);
        Col c{0.0f, 0.0f, 0.0f, 1.0f};
        HitD ph = no_hit();
        if (sx < P.width && sy < P.row_end) trace_sample<INTEGRATOR, MODE, COUNT, BOUNCE>(P, sx, sy, i, c, ph, pc);
        if (spp == 1) { // lane = pixel: nothing to exchange
            color = cadd(color, c);
            primary = ph;
            break;
        }
#pragma unroll 1
        for (uint32_t src = 0; src < 32u; ++src) {
            const uint32_t ks = k0 + src, ps = ks / spp; // warp-uniform
            Col cs{__shfl_sync(0xFFFFFFFFu, c.r, src), __shfl_sync(0xFFFFFFFFu, c.g, src), __shfl_sync(0xFFFFFFFFu, c.b, src),
                   __shfl_sync(0xFFFFFFFFu, c.a, src)};
            if (ps == lane) color = cadd(color, cs);
            if (want_aux && ks - ps * spp == spp - 1u) { // the parity planes describe the last sample traced
                const uint32_t hp = __shfl_sync(0xFFFFFFFFu, ph.prim, src);
                const float ht = __shfl_sync(0xFFFFFFFFu, ph.t, src), hu = __shfl_sync(0xFFFFFFFFu, ph.u, src),
                            hv = __shfl_sync(0xFFFFFFFFu, ph.v, src);
                if (ps == lane) {
                    primary.prim = hp;
                    primary.t = ht;
                    primary.u = hu;
                    primary.v = hv;
                }
            }
        }
    }
    if (P.accum && mine) P.accum[(size_t)y * P.width + x] = make_float4(color.r, color.g, color.b, color.a);
    if (mine) write_pixel(P, x, y, local_row, color, primary);
}

// 8 blocks x 4 warps per SM = 64 registers per thread: the walk is latency- and issue-bound, and half
// occupancy with a few spilled shading temporaries beats 25 % occupancy without (profiles/: 1.28 ms at
// 146 registers, 0.98 at 80, 0.92 at 64 for the 1080p atrium frame).
#ifndef RAYCA_MIN_BLOCKS
#define RAYCA_MIN_BLOCKS 8
#endif
// Static grid: warps per block along x and y (each warp an 8x4 pixel tile).  Default 2 x 2 = 128 threads = 16x8 pixels.
// A/B knobs (tools/variants.py): 2 x 1 = 64 threads, 1 x 1 = one warp per block, 4 x 2 = 256 threads.
#ifndef RAYCA_WARPS_X
#define RAYCA_WARPS_X 2
#endif
#ifndef RAYCA_WARPS_Y
#define RAYCA_WARPS_Y 2
#endif
constexpr uint32_t kWarpsX = RAYCA_WARPS_X, kWarpsY = RAYCA_WARPS_Y, kBlockThreads = 32u * kWarpsX * kWarpsY;
constexpr uint32_t kGroupW = 8u * kWarpsX, kGroupH = 4u * kWarpsY; // pixels per block

// Block = 128 threads = 4 warps; a warp owns an 8x4 pixel tile (coherent rays share node fetches
// through L1), four consecutive tiles form a 16x8 pixel group.
//   default: one block per pixel group, scheduled by the hardware's block dispatcher.
//   PERSISTENT (RAYCA_FLAG_PERSISTENT): the grid is sized to the resident capacity of the GPU and every
//     warp pulls its next group (4 tiles, walked one after the other) from a global counter — one
//     atomicAdd per warp, broadcast by shuffle.  Measured 15-20 % SLOWER than the static grid on B200 for
//     the 1080p frames (profiles/r01_ab_*): the block dispatcher already is a dynamic queue with 16k
//     entries per frame, and the persistent form adds the fetch latency to every group.
template <int INTEGRATOR, int MODE, bool COUNT, bool PERSISTENT, bool BOUNCE>
__global__ void __launch_bounds__(kBlockThreads, RAYCA_MIN_BLOCKS) render_kernel(const __grid_constant__ RenderParams P) {
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    if (MODE == 0 || MODE == 3) stage_treelet(P.scene); // nothing unless RAYCA_TREELET
    PixelCounters pc;
    if (PERSISTENT) {
        // one atomic per warp per GROUP (4 tiles): same-address atomics retire at about one per L2 clock,
        // so per-tile fetches (260k per 1080p frame) would cost more than the imbalance they remove
        for (;;) {
            unsigned long long t = 0;
            if (lane == 0) t = atomicAdd(&P.counters[5], 1ull);
            const uint32_t group = (uint32_t)__shfl_sync(0xFFFFFFFFu, t, 0);
            if (group >= P.tile_count) break;
            const uint32_t gx = group % P.groups_x, gy = group / P.groups_x;
#pragma unroll 1
            for (uint32_t sub = 0; sub < 4u; ++sub) {
                render_tile<INTEGRATOR, MODE, COUNT, BOUNCE>(P, gx * 16u + (sub & 1u) * 8u, gy * 8u + (sub >> 1) * 4u, lane, pc);
            }
        }
    } else {
        uint32_t gx = blockIdx.x;
        if (P.gshard_count > 1) { // this shard's blockIdx.x-th group of group row blockIdx.y
            gx = (P.gshard_index + P.gshard_count - blockIdx.y % P.gshard_count) % P.gshard_count + blockIdx.x * P.gshard_count;
            if (gx >= P.groups_x_full) return; // nothing was traced: the counters below would add zeros
        }
        render_tile<INTEGRATOR, MODE, COUNT, BOUNCE>(P, gx * kGroupW + (warp % kWarpsX) * 8u, blockIdx.y * kGroupH + (warp / kWarpsX) * 4u, lane, pc);
    }
    // warp-aggregated counters: one atomic per warp per counter
    uint32_t sh = __reduce_add_sync(0xFFFFFFFFu, pc.shadow);
    uint32_t se = __reduce_add_sync(0xFFFFFFFFu, pc.secondary);
    if (lane == 0) {
        if (sh) atomicAdd(&P.counters[0], (unsigned long long)sh);
        if (se) atomicAdd(&P.counters[1], (unsigned long long)se);
    }
    if (COUNT) {
        unsigned long long n = pc.walk.nodes, t = pc.walk.tris, pl = pc.walk.planes;
        for (int ofs = 16; ofs > 0; ofs >>= 1) {
            n += __shfl_down_sync(0xFFFFFFFFu, n, ofs);
            t += __shfl_down_sync(0xFFFFFFFFu, t, ofs);
            pl += __shfl_down_sync(0xFFFFFFFFu, pl, ofs);
        }
        if (lane == 0) {
            atomicAdd(&P.counters[2], n);
            atomicAdd(&P.counters[3], t);
            atomicAdd(&P.counters[4], pl);
        }
    }
}

} // namespace rb
