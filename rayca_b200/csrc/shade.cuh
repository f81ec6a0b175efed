// shade.cuh — hit attribute fetch, texturing, BRDFs and light evaluation (device).
//
// Restates HitInfo (rayca-soft/src/hit.rs:7-246), BvhTriangle::interpolate_* (bvh/triangle.rs:33-71),
// BvhPrimitive::{get_color,get_diffuse,get_normal} (bvh/primitive.rs:142-181), Sampler::sample
// (rayca-model/src/sampler.rs:11-29), Material::* (rayca-model/src/material/mod.rs:107-185),
// PbrMaterial::* (material/pbr.rs:94-136), ggx / lambertian BRDFs (rayca-soft/src/brdf/*.rs),
// PointLight (rayca-model/src/light/point.rs:37-54) and Irradiance::new (hit.rs:267-286).
// Colour arithmetic keeps the reference's alpha rules (see rmath.h).
#pragma once
#include <cuda_runtime.h>

#include "../../include/rayca_cuda.h"
#include "gpu_types.h"
#include "rmath.h"
#include "trace.cuh"

namespace rb {

__device__ __forceinline__ Col ld_col(const float* p) { return Col{p[0], p[1], p[2], p[3]}; }

// Rust `as u32`: saturating, NaN -> 0
__device__ __forceinline__ uint32_t f32_as_u32(float f) { return __float2uint_rz(f); }

#if RAYCA_SHADE_OPT
// (float)t / 255.0f for a byte t, in three instructions: q = t * r with r = RN(1/255), one residual step
// q' = fma(fma(-q, 255, t), r, q).  Equal to the IEEE quotient for ALL 256 inputs (checked exhaustively with exact
// rational arithmetic in tests/test_oracle_kat.py::test_byte_over_255_identity); the general division is ~10
// instructions plus a slow path, and a shaded pixel converts twelve texel channels.
__device__ __forceinline__ float byte_to_unit(unsigned char t) {
    const float ft = (float)t, r = 0x1.010102p-8f;
    const float q = ft * r;
    return __fmaf_rn(__fmaf_rn(-q, 255.0f, ft), r, q);
}
// x % n: a mask when n is a power of two (every glTF texture of the fixtures is), the general remainder otherwise
__device__ __forceinline__ uint32_t wrap_index(uint32_t x, uint32_t n) { return (n & (n - 1u)) == 0u ? (x & (n - 1u)) : (x % n); }
#else
__device__ __forceinline__ float byte_to_unit(unsigned char t) { return (float)t / 255.0f; }
__device__ __forceinline__ uint32_t wrap_index(uint32_t x, uint32_t n) { return x % n; }
#endif

// Sampler::sample — nearest, repeat, no filtering, no sRGB decode
__device__ __forceinline__ Col sample_image(const GpuSceneView& s, int image, float uvx, float uvy) {
    const GpuImage im = s.images[image];
    float x = (uvx - floorf(uvx) + 1.0f) * (float)im.width;
    float y = (uvy - floorf(uvy) + 1.0f) * (float)im.height;
    uint32_t xi = wrap_index(f32_as_u32(x), im.width);
    uint32_t yi = wrap_index(f32_as_u32(y), im.height);
    size_t idx = (size_t)yi * im.width + xi;
    const unsigned char* base = s.texels + im.offset;
    if (im.color_type == RAYCA_COLOR_RGBA8) {
        uchar4 t = __ldg(reinterpret_cast<const uchar4*>(base) + idx);
        return Col{byte_to_unit(t.x), byte_to_unit(t.y), byte_to_unit(t.z), byte_to_unit(t.w)};
    } else if (im.color_type == RAYCA_COLOR_RGB8) {
        const unsigned char* p = base + idx * 3;
        return Col{byte_to_unit(__ldg(p)), byte_to_unit(__ldg(p + 1)), byte_to_unit(__ldg(p + 2)), 255.0f / 255.0f};
    } else {
        float4 t = __ldg(reinterpret_cast<const float4*>(base) + idx);
        return Col{t.x / 255.0f, t.y / 255.0f, t.z / 255.0f, t.w / 255.0f}; // color/mod.rs:222-231
    }
}

// Everything the integrators ask a HitInfo for, evaluated once per shaded hit.
struct Surf {
    V3 point;      // Hit::point
    V3 view_dir;   // -ray.dir
    V3 ray_dir;
    V3 normal;     // shading normal (normal-mapped for PBR)
    V3 next_origin;
    float uvx, uvy;
    Col geom_color;
    const GpuMaterial* mat;
    // textured PBR lookups, evaluated once per hit by eval_material (the reference re-samples the same
    // texels on every Material::get_* call; the values are identical)
    Col pbr;                    // PbrMaterial::get_color
    float metallic, roughness;  // PbrMaterial::get_metallic_roughness
};

__device__ __forceinline__ V3 interp3(float4 a0, float4 a1, float4 a2, float w, float u, float v) {
    V3 n = (V3{a2.x, a2.y, a2.z} * w + V3{a0.x, a0.y, a0.z} * u) + V3{a1.x, a1.y, a1.z} * v;
    return normalize(n);
}

__device__ __forceinline__ Col pbr_color(const GpuSceneView& s, const GpuMaterial* m, float uvx, float uvy) {
    Col c = ld_col(m->a);
    if (m->tex_albedo >= 0) return cmul(c, sample_image(s, m->tex_albedo, uvx, uvy));
    return c;
}

// PbrMaterial::get_metallic_roughness (pbr.rs:125-136): metallic = B, roughness = R
__device__ __forceinline__ void eval_material(const GpuSceneView& s, Surf& sf) {
    const GpuMaterial* m = sf.mat;
    if (m->kind != RAYCA_MATERIAL_PBR) return;
    sf.pbr = pbr_color(s, m, sf.uvx, sf.uvy);
    if (m->tex_mr >= 0) {
        Col c = sample_image(s, m->tex_mr, sf.uvx, sf.uvy);
        sf.metallic = c.b;
        sf.roughness = c.r;
    } else {
        sf.metallic = m->p0;
        sf.roughness = m->p1;
    }
}

template <bool MATERIAL = true, bool SPHERES = true>
__device__ __forceinline__ void load_surf(const GpuSceneView& s, const RayD& ray, const HitD& h, Surf& sf) {
    const float4* q = s.tri_attr + 12 * (size_t)h.prim;
    if (SPHERES && h.sphere) {
        // BvhGeometry::Sphere: colour WHITE, uv (0,0) (primitive.rs:16-29); normal = normalize(N * normalize(
        // inv * hit.point - centre)) with inv = Mat4::from(&Inversed(trs)), N = Mat3::from(&inv)^T (primitive.rs:183-190);
        // the attr record carries inv rows 0..2 and N's rows (host_scene.cpp), the 4th row of inv is (0,0,0,1)
        const float4 r0 = __ldg(q + 0), r1 = __ldg(q + 1), r2 = __ldg(q + 2);
        const float4 n0 = __ldg(q + 3), n1 = __ldg(q + 4), n2 = __ldg(q + 5);
        const V3 hp{((r0.x * h.p.x + r0.y * h.p.y) + r0.z * h.p.z) + r0.w * 1.0f, ((r1.x * h.p.x + r1.y * h.p.y) + r1.z * h.p.z) + r1.w * 1.0f,
                    ((r2.x * h.p.x + r2.y * h.p.y) + r2.z * h.p.z) + r2.w * 1.0f};
        const V3 mn = normalize(hp - V3{n0.w, n1.w, n2.w});
        V3 wn; // Mat3 * Vec3 accumulates from 0.0 (mat3.rs:218-234)
        wn.x = ((0.0f + n0.x * mn.x) + n0.y * mn.y) + n0.z * mn.z;
        wn.y = ((0.0f + n1.x * mn.x) + n1.y * mn.y) + n1.z * mn.z;
        wn.z = ((0.0f + n2.x * mn.x) + n2.y * mn.y) + n2.z * mn.z;
        sf.geom_color = Col{1.0f, 1.0f, 1.0f, 1.0f};
        sf.uvx = sf.uvy = 0.0f;
        sf.mat = s.materials + __float_as_uint(__ldg(q + 9).w);
        sf.point = h.p;
        sf.ray_dir = ray.d;
        sf.view_dir = -ray.d;
        sf.normal = normalize(wn);
        sf.next_origin = sf.point + sf.normal * kRayBias; // hit.rs:164-171
        if (MATERIAL) eval_material(s, sf);
        return;
    }
    float u = h.u, v = h.v;
    float w = 1.0f - u - v;
    // bvh/triangle.rs:34-48 — v2*(1-u-v) + v0*u + v1*v with Color / Vec2 operators
    float4 c0 = __ldg(q + 0), c1 = __ldg(q + 1), c2 = __ldg(q + 2);
    Col gc = cscale(Col{c2.x, c2.y, c2.z, c2.w}, w);
    gc = cadd(gc, cscale(Col{c0.x, c0.y, c0.z, c0.w}, u));
    gc = cadd(gc, cscale(Col{c1.x, c1.y, c1.z, c1.w}, v));
    sf.geom_color = gc;
    float4 n0 = __ldg(q + 3), n1 = __ldg(q + 4), n2 = __ldg(q + 5);
    float4 t0 = __ldg(q + 6), t1 = __ldg(q + 7), t2 = __ldg(q + 8);
    float4 b0 = __ldg(q + 9);
    sf.uvx = (n2.w * w + n0.w * u) + n1.w * v;
    sf.uvy = (t2.w * w + t0.w * u) + t1.w * v;
    sf.mat = s.materials + __float_as_uint(b0.w);
    sf.point = h.p;
    sf.ray_dir = ray.d;
    sf.view_dir = -ray.d;
    V3 n = interp3(n0, n1, n2, w, u, v);
    if (sf.mat->kind == RAYCA_MATERIAL_PBR && sf.mat->tex_normal >= 0) { // pbr.rs:104-123
        float4 b1 = __ldg(q + 10), b2 = __ldg(q + 11);
        V3 tg = interp3(t0, t1, t2, w, u, v);
        V3 bt = interp3(b0, b1, b2, w, u, v);
        V3 sn = col_to_v3(sample_image(s, sf.mat->tex_normal, sf.uvx, sf.uvy));
        sn = V3{sn.x * 2.0f - 1.0f, sn.y * 2.0f - 1.0f, sn.z * 2.0f - 1.0f};
        // Mat3::tbn * v (mat3.rs:57-65,218-234): rows (t.i, b.i, n.i), accumulated from 0.0
        V3 r;
        r.x = ((0.0f + tg.x * sn.x) + bt.x * sn.y) + n.x * sn.z;
        r.y = ((0.0f + tg.y * sn.x) + bt.y * sn.y) + n.y * sn.z;
        r.z = ((0.0f + tg.z * sn.x) + bt.z * sn.y) + n.z * sn.z;
        n = normalize(r);
    }
    sf.normal = n;
    sf.next_origin = sf.point + n * kRayBias; // hit.rs:164-171
    if (MATERIAL) eval_material(s, sf);
}

// Material::get_color (material/mod.rs:107-114) x geometry colour (primitive.rs:142-148)
__device__ __forceinline__ Col surf_color(const GpuSceneView& s, const Surf& sf) {
    const GpuMaterial* m = sf.mat;
    Col mc;
    if (m->kind == RAYCA_MATERIAL_PBR) mc = sf.pbr;
    else if (m->kind == RAYCA_MATERIAL_PHONG) mc = cadd(ld_col(m->a), ld_col(m->b)); // ambient + emission
    else mc = ld_col(m->a);
    return cmul(sf.geom_color, mc);
}
// Material::get_diffuse (material/mod.rs:116-123) x geometry colour (primitive.rs:150-155)
__device__ __forceinline__ Col surf_diffuse(const GpuSceneView& s, const Surf& sf) {
    const GpuMaterial* m = sf.mat;
    Col mc;
    if (m->kind == RAYCA_MATERIAL_PBR) mc = sf.pbr;
    else if (m->kind == RAYCA_MATERIAL_PHONG) mc = ld_col(m->c);
    else mc = ld_col(m->a);
    return cmul(sf.geom_color, mc);
}
__device__ __forceinline__ void surf_metal_rough(const GpuSceneView& s, const Surf& sf, float& metallic, float& roughness) {
    metallic = sf.metallic;
    roughness = sf.roughness;
}
// Material::get_specular (material/mod.rs:143-152)
__device__ __forceinline__ Col surf_specular(const GpuSceneView& s, const Surf& sf) {
    const GpuMaterial* m = sf.mat;
    if (m->kind == RAYCA_MATERIAL_PBR) {
        float metallic, roughness;
        surf_metal_rough(s, sf, metallic, roughness);
        return cscale(sf.pbr, metallic);
    }
    return ld_col(m->d);
}
// Material::get_roughness (material/mod.rs:171-185)
__device__ __forceinline__ float surf_roughness(const GpuSceneView& s, const Surf& sf) {
    const GpuMaterial* m = sf.mat;
    if (m->kind == RAYCA_MATERIAL_PHONG) return clampf(sqrtf(2.0f / (m->p0 + 2.0f)), 0.0f, 1.0f);
    if (m->kind == RAYCA_MATERIAL_PBR) {
        float metallic, roughness;
        surf_metal_rough(s, sf, metallic, roughness);
        return roughness;
    }
    return m->p0;
}

// powf(x, 2.0 / 4.0 / 5.0) of the reference (f32::powf -> libm) written as products: correctly rounded
// squares, at most an ulp or two away from libm for the 4th / 5th power, i.e. inside the spread
// libdevice's own powf has against glibc, and far inside the 1e-3 RGB tolerance.  Keeps ~1k instructions
// of inlined powf out of the kernel (the i-cache is what this kernel starves on, see profiles/).
__device__ __forceinline__ float pow2(float x) { return x * x; }
__device__ __forceinline__ float pow4(float x) { float x2 = x * x; return x2 * x2; }
__device__ __forceinline__ float pow5(float x) { float x2 = x * x; return (x2 * x2) * x; }

// ---- brdf/ggx.rs:58-127 (path-tracer form)
// tan(acos(c))^2 == (1 - c^2) / c^2.  The reference spells it `theta = acos(c); ... tan(theta)` (brdf/ggx.rs:58-84,
// the oracle keeps that form); libm and CUDA already differ by ulps there, and near grazing angles the float
// angle itself limits tan to ~1e-3 relative, so the identity sits inside the same 1e-3 RGB contract — and is
// twelve acosf / tanf expansions shorter per shaded pixel (1080p atrium frame: identical RGBA8 bytes, 0.878 ->
// 0.859 ms).  c > 1 (a dot product of unit vectors one ulp above 1) stays a NaN like acos(c).
__device__ __forceinline__ float tan2_of_acos(float c) {
    if (c > 1.0f) return __int_as_float(0x7fc00000);
    const float c2 = c * c;
    return (1.0f - c2) / c2;
}
__device__ __forceinline__ float ggx_d(float a, V3 h, V3 n) {
    float a2 = a * a;
    float ct = clampf(dot(h, n), 0.0f, 1.0f);
    float p4 = pow4(ct);
    if (p4 == 0.0f) return 0.0f; // the reference's den is pow4(ct) * (finite)^2 == 0 here
    float den = p4 * pow2(a2 + tan2_of_acos(ct));
    if (den == 0.0f) return 0.0f;
    return a2 * kFrac1Pi / den;
}
__device__ __forceinline__ float ggx_g1(float a, V3 omega, V3 n) {
    float ct = dot(omega, n);
    if (ct <= 0.0f) return 0.0f;
    float den = 1.0f + sqrtf(1.0f + a * a * tan2_of_acos(ct));
    return 2.0f / den;
}
__device__ __forceinline__ Col ggx_bsdf(const GpuSceneView& s, const Surf& sf, V3 wi) {
    const Col black{0.0f, 0.0f, 0.0f, 1.0f};
    V3 wo = sf.view_dir;
    V3 n = sf.normal;
    float idn = clampf(dot(wi, n), 0.0f, 1.0f);
    float odn = clampf(dot(wo, n), 0.0f, 1.0f);
    if (idn == 0.0f || odn == 0.0f) return black;
    Col ks = surf_specular(s, sf);
    float a = surf_roughness(s, sf);
    V3 h = normalize(wi + wo);
    // get_f: ks + (WHITE - ks) * (1 - |wi.h|)^5
    float idh = fabsf(dot(wi, h));
    Col f = cadd(ks, cscale(csub(Col{1.0f, 1.0f, 1.0f, 1.0f}, ks), pow5(1.0f - idh)));
    float g = ggx_g1(a, wi, n) * ggx_g1(a, wo, n);
    float d = ggx_d(a, h, n);
    float den = 4.0f * idn * odn;
    return cdiv(cscale(cscale(f, g), d), den);
}
// hit.rs:220-227 -> ggx.rs:124-127 / lambertian.rs:7-15
__device__ __forceinline__ Col surf_brdf(const GpuSceneView& s, const Surf& sf, V3 wi) {
    if (sf.mat->kind == RAYCA_MATERIAL_PHONG) {
        Col lambertian = cscale(surf_diffuse(s, sf), kFrac1Pi);
        float sh = sf.mat->p0;
        V3 refl = normalize(reflect(sf.ray_dir, sf.normal)); // hit.rs:93-101
        Col spec = cdiv(cscale(cscale(cscale(ld_col(sf.mat->d), sh + 2.0f), powf(dot(refl, wi), sh)), kFrac1Pi), 2.0f);
        return cadd(lambertian, spec);
    }
    Col kd = surf_diffuse(s, sf);
    return cadd(cscale(kd, kFrac1Pi), ggx_bsdf(s, sf, wi));
}

// ---- Raytracer radiance: Irradiance::new (hit.rs:267-286) + ggx::get_radiance (ggx.rs:11-52)
// + lambertian::get_radiance (lambertian.rs:77-81)
__device__ __forceinline__ Col surf_radiance(const GpuSceneView& s, const Surf& sf, Col intensity, V3 l) {
    V3 n = sf.normal;
    V3 v = sf.view_dir;
    float ndv = clampf(dot(n, v), 0.0f, 1.0f) + 1e-5f;
    float ndl = clampf(dot(n, l), 0.0f, 1.0f);
    V3 h = normalize(v + l);
    float ndh = clampf(dot(n, h), 0.0f, 1.0f);
    float ldh = clampf(dot(l, h), 0.0f, 1.0f);
    if (sf.mat->kind == RAYCA_MATERIAL_PHONG) {
        Col diffuse = cscale(surf_diffuse(s, sf), ndl);
        Col specular = cscale(ld_col(sf.mat->d), powf(ndh, sf.mat->p0));
        return cmul(cadd(diffuse, specular), intensity);
    }
    float metallic, roughness;
    if (sf.mat->kind == RAYCA_MATERIAL_PBR) {
        surf_metal_rough(s, sf, metallic, roughness);
    } else { // get_pbr_material panics for Material::Ggx in the reference; treated as (0, roughness)
        metallic = 0.0f;
        roughness = sf.mat->p0;
    }
    float a = ndh * roughness;
    float k = roughness / (1.0f - ndh * ndh + a * a);
    float d = k * k * kFrac1Pi;
    Col albedo = surf_color(s, sf);
    V3 f0 = V3{0.04f, 0.04f, 0.04f} * (1.0f - metallic) + col_to_v3(albedo) * metallic;
    float fp = pow5(1.0f - ldh);
    V3 f = f0 + (V3{1.0f, 1.0f, 1.0f} - f0) * fp;
    V3 kd = (V3{1.0f, 1.0f, 1.0f} - f) * (1.0f - metallic);
    float ggxv = ndl * (ndv * (1.0f - roughness) + roughness);
    float ggxl = ndv * (ndl * (1.0f - roughness) + roughness);
    float g = 0.5f / (ggxv + ggxl);
    Col fr = cscale(v3_to_col(f), d * g);
    Col fd = cscale(cmul(v3_to_col(kd), albedo), kFrac1Pi);
    return cscale(cmul(cadd(fd, fr), intensity), ndl);
}

// PointLight::get_intensity / get_fallof (light/point.rs:37-48)
__device__ __forceinline__ Col point_intensity(const GpuLight& l, V3 frag) {
    V3 dist = frag - V3{l.pos[0], l.pos[1], l.pos[2]};
    float r2 = norm2(dist);
    float r = sqrtf(r2);
    float fall = ((l.att[0] * 1.0f + l.att[1] * r) + l.att[2] * r2) + 0.0f;
    Col c = cscale(Col{l.color[0], l.color[1], l.color[2], l.color[3]}, l.intensity);
    return cdiv(c, fall);
}

} // namespace rb
