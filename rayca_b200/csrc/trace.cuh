// trace.cuh — closest-hit / any-hit BVH walk and the ray-triangle test (device).
//
// Replaces Tlas::intersects -> TlasNode::intersects -> BvhNode::intersects ->
// Triangle::intersects (rayca-soft/src/bvh/tlas.rs:136-180,271-275, bvh/blas.rs:129-176,
// rayca-geometry/src/triangle.rs:84-159).
//
// Equivalence with the reference's exhaustive recursive walk:
//   * a triangle is tested iff the slab test of its LEAF box accepts the ray; every ancestor box
//     is an exact f32 superset, and the slab arithmetic is monotone under rounding, so ancestors
//     never reject what the leaf accepts — the walk order is therefore free;
//   * the reference keeps the first hit in depth-first order on equal depth (strict '<'), and
//     depth-first order equals ascending (blas, primitive) index, so ties resolve to the lowest
//     global triangle index here;
//   * ordered mode additionally skips a subtree when its slab entry distance exceeds the best
//     depth so far (never on equality).  Exhaustive mode visits everything the reference visits.
#pragma once
#include <cuda_runtime.h>

#include "gpu_types.h"
#include "rmath.h"

namespace rb {

constexpr int kStackSize = 96;

struct RayD {
    V3 o, d, rd;
};
__device__ __forceinline__ RayD make_ray(V3 o, V3 d) { return RayD{o, d, reciprocal(d)}; } // ray.rs:63-72

struct HitD {
    float t;
    uint32_t prim; // global triangle index, ~0u = miss
    float u, v;
    V3 p;
};

struct WalkCounters {
    unsigned long long nodes = 0; // 32 B node records fetched
    unsigned long long tris = 0;  // 48 B triangle records fetched
};

// AABB::intersects (bvh/aabb.rs:74-93); returns tmin or f32::MAX
__device__ __forceinline__ float slab(const float4 lo, const float4 hi, const RayD& r) {
    float t1x = (lo.x - r.o.x) * r.rd.x, t2x = (hi.x - r.o.x) * r.rd.x;
    float t1y = (lo.y - r.o.y) * r.rd.y, t2y = (hi.y - r.o.y) * r.rd.y;
    float t1z = (lo.z - r.o.z) * r.rd.z, t2z = (hi.z - r.o.z) * r.rd.z;
    float tmax = fminf(fminf(fminf(fmaxf(t1x, t2x), fmaxf(t1y, t2y)), fmaxf(t1z, t2z)), kF32Max);
    float tmin = fmaxf(fmaxf(fmaxf(fminf(t1x, t2x), fminf(t1y, t2y)), fminf(t1z, t2z)), -kF32Max);
    return (tmax >= tmin && tmax > 0.0f) ? tmin : kF32Max;
}

// Triangle::intersects (rayca-geometry/src/triangle.rs:84-159) on world-space vertices.
__device__ __forceinline__ bool tri_test(const float4 a, const float4 b, const float4 c, const RayD& r, float& t_out,
                                         float& u_out, float& v_out, V3& p_out) {
    V3 v0{a.x, a.y, a.z}, v1{b.x, b.y, b.z}, v2{c.x, c.y, c.z};
    V3 e01 = v1 - v0;
    V3 e02 = v2 - v0;
    V3 n = cross(e01, e02);
    float ndd = dot(r.d, n); // == n.dot(ray.dir): same products, same order
    if (ndd > 0.0f) return false;               // back face (:96)
    if (fabsf(ndd) < kF32Eps) return false;     // parallel (:106)
    float denom = dot(n, n);
    float d = -dot(n, v0);
    float t = -(dot(n, r.o) + d) / ndd;
    if (t < 0.0f) return false;                 // behind (:117)
    V3 p = r.o + r.d * t;
    V3 cc = cross(e01, p - v0);
    if (dot(n, cc) < 0.0f) return false;
    cc = cross(v2 - v1, p - v1);
    float u = dot(n, cc);
    if (u < 0.0f) return false;
    cc = cross(v0 - v2, p - v2);
    float v = dot(n, cc);
    if (v < 0.0f) return false;
    t_out = t;
    u_out = u / denom;
    v_out = v / denom;
    p_out = p;
    return true;
}

// CULL: ordered + distance-culled.  ANY: stop at the first hit with t < hit.t (shadow query for
// NEE point lights: "shadowed iff closest.depth < distance" <=> "some hit has depth < distance",
// sampler/nee.rs:152-156).  COUNT: tally node / triangle records for the roofline model.
template <bool CULL, bool ANY, bool COUNT>
__device__ __forceinline__ bool walk(const GpuSceneView& s, const RayD& ray, HitD& hit, WalkCounters* wc) {
    float best_t = ANY ? hit.t : __int_as_float(0x7f800000); // +inf (blas.rs:143)
    uint32_t best = 0xFFFFFFFFu;
    float bu = 0.0f, bv = 0.0f;
    V3 bp{0.0f, 0.0f, 0.0f};

    auto leaf = [&](uint32_t first, uint32_t count) -> bool {
        if (COUNT) wc->tris += count;
        for (uint32_t i = first; i < first + count; ++i) {
            const float4 a = __ldg(&s.tri_pos[3 * (size_t)i + 0]);
            const float4 b = __ldg(&s.tri_pos[3 * (size_t)i + 1]);
            const float4 c = __ldg(&s.tri_pos[3 * (size_t)i + 2]);
            float t, u, v;
            V3 p;
            if (tri_test(a, b, c, ray, t, u, v, p)) {
                if (ANY) {
                    if (t < best_t) return true;
                } else if (t < best_t || (t == best_t && i < best)) {
                    best_t = t;
                    best = i;
                    bu = u;
                    bv = v;
                    bp = p;
                }
            }
        }
        return false;
    };
    auto finish = [&]() -> bool {
        if (ANY) return false;
        hit.t = best_t;
        hit.prim = best;
        hit.u = bu;
        hit.v = bv;
        hit.p = bp;
        return best != 0xFFFFFFFFu;
    };

    // root record
    {
        if (COUNT) wc->nodes += 1;
        float4 lo = make_float4(s.root_min[0], s.root_min[1], s.root_min[2], 0.0f);
        float4 hi = make_float4(s.root_max[0], s.root_max[1], s.root_max[2], 0.0f);
        if (slab(lo, hi, ray) == kF32Max) return finish();
        if (s.root_link1 != 0) {
            if (leaf(s.root_link0, s.root_link1)) return true;
            return finish();
        }
    }

    uint32_t stack_i[kStackSize];
    float stack_d[kStackSize];
    int sp = 0;
    uint32_t cur = s.root_link0;
    for (;;) {
        const float4* np = s.nodes + 2 * (size_t)cur;
        float4 a0 = __ldg(np + 0), a1 = __ldg(np + 1), b0 = __ldg(np + 2), b1 = __ldg(np + 3);
        if (COUNT) wc->nodes += 2;
        float d0 = slab(a0, a1, ray), d1 = slab(b0, b1, ray);
        uint32_t l0 = __float_as_uint(a0.w), c0 = __float_as_uint(a1.w);
        uint32_t l1 = __float_as_uint(b0.w), c1 = __float_as_uint(b1.w);
        bool h0 = d0 != kF32Max && !(CULL && d0 > best_t);
        bool h1 = d1 != kF32Max && !(CULL && d1 > best_t);
        if (CULL && h0 && h1 && d1 < d0) { // nearer child first
            float td = d0; d0 = d1; d1 = td;
            uint32_t tl = l0; l0 = l1; l1 = tl;
            uint32_t tc = c0; c0 = c1; c1 = tc;
        }
        if (h0 && c0 != 0) {
            if (leaf(l0, c0)) return true;
            h0 = false;
            if (CULL && h1 && d1 > best_t) h1 = false;
        }
        if (h1 && c1 != 0) {
            if (leaf(l1, c1)) return true;
            h1 = false;
        }
        if (h0) {
            cur = l0;
            if (h1) {
                if (sp < kStackSize) {
                    stack_i[sp] = l1;
                    stack_d[sp] = d1;
                    ++sp;
                }
            }
            continue;
        }
        if (h1) {
            cur = l1;
            continue;
        }
        bool popped = false;
        while (sp > 0) {
            --sp;
            if (CULL && stack_d[sp] > best_t) continue;
            cur = stack_i[sp];
            popped = true;
            break;
        }
        if (!popped) break;
    }
    return finish();
}

// ---------------------------------------------------------------------------------------------
// Performance-layer walk (fast_bvh.cpp).  Same result as walk<> by construction:
//   * the slab test here is CONSERVATIVE: it accepts whenever the exact test would, plus a relative
//     tolerance in t, so rounding can never hide a triangle the reference would reach;
//   * a candidate hit only counts if the exact slab test of its REFERENCE leaf box (nodes[w1]) accepts
//     the ray — that is precisely the condition under which the reference tests the triangle;
//   * ties on depth resolve to the lowest reference-global index (w0), like depth-first order does.
constexpr float kSlabTol = 1.0e-5f;

__device__ __forceinline__ bool slab_conservative(const float4 lo, const float4 hi, const RayD& r, float& entry) {
    float t1x = (lo.x - r.o.x) * r.rd.x, t2x = (hi.x - r.o.x) * r.rd.x;
    float t1y = (lo.y - r.o.y) * r.rd.y, t2y = (hi.y - r.o.y) * r.rd.y;
    float t1z = (lo.z - r.o.z) * r.rd.z, t2z = (hi.z - r.o.z) * r.rd.z;
    float tmax = fminf(fminf(fmaxf(t1x, t2x), fmaxf(t1y, t2y)), fmaxf(t1z, t2z));
    float tmin = fmaxf(fmaxf(fminf(t1x, t2x), fminf(t1y, t2y)), fminf(t1z, t2z));
    float tol = kSlabTol * fmaxf(fabsf(tmin), fabsf(tmax));
    entry = tmin - tol;
    return (tmax + tol >= tmin) && (tmax + tol >= 0.0f);
}

template <bool ANY, bool COUNT>
__device__ __forceinline__ bool walk_fast(const GpuSceneView& s, const RayD& ray, HitD& hit, WalkCounters* wc) {
    float best_t = ANY ? hit.t : __int_as_float(0x7f800000);
    uint32_t best = 0xFFFFFFFFu;
    float bu = 0.0f, bv = 0.0f;
    V3 bp{0.0f, 0.0f, 0.0f};
    uint32_t ok_leaf = 0xFFFFFFFFu; // last reference leaf whose exact slab test accepted this ray

    auto ref_leaf_accepts = [&](uint32_t leaf) -> bool {
        if (leaf == ok_leaf) return true;
        const float4 lo = __ldg(&s.nodes[2 * (size_t)leaf]);
        const float4 hi = __ldg(&s.nodes[2 * (size_t)leaf + 1]);
        if (COUNT) wc->nodes += 1;
        if (slab(lo, hi, ray) == kF32Max) return false;
        ok_leaf = leaf;
        return true;
    };
    auto leaf = [&](uint32_t first, uint32_t count) -> bool {
        if (COUNT) wc->tris += count;
        for (uint32_t i = first; i < first + count; ++i) {
            const float4 a = __ldg(&s.fast_tri[3 * (size_t)i + 0]);
            const float4 b = __ldg(&s.fast_tri[3 * (size_t)i + 1]);
            const float4 c = __ldg(&s.fast_tri[3 * (size_t)i + 2]);
            float t, u, v;
            V3 p;
            if (tri_test(a, b, c, ray, t, u, v, p)) {
                const uint32_t ref = __float_as_uint(a.w);
                if (ANY) {
                    if (t < best_t && ref_leaf_accepts(__float_as_uint(b.w))) return true;
                } else if (t < best_t || (t == best_t && ref < best)) {
                    if (ref_leaf_accepts(__float_as_uint(b.w))) {
                        best_t = t;
                        best = ref;
                        bu = u;
                        bv = v;
                        bp = p;
                    }
                }
            }
        }
        return false;
    };
    auto finish = [&]() -> bool {
        if (ANY) return false;
        hit.t = best_t;
        hit.prim = best;
        hit.u = bu;
        hit.v = bv;
        hit.p = bp;
        return best != 0xFFFFFFFFu;
    };

    {
        if (COUNT) wc->nodes += 1;
        float4 lo = make_float4(s.fast_root_min[0], s.fast_root_min[1], s.fast_root_min[2], 0.0f);
        float4 hi = make_float4(s.fast_root_max[0], s.fast_root_max[1], s.fast_root_max[2], 0.0f);
        float e;
        if (!slab_conservative(lo, hi, ray, e)) return finish();
        if (s.fast_root_link1 != 0) {
            if (leaf(s.fast_root_link0, s.fast_root_link1)) return true;
            return finish();
        }
    }
    uint32_t stack_i[kStackSize];
    float stack_d[kStackSize];
    int sp = 0;
    uint32_t cur = s.fast_root_link0;
    for (;;) {
        const float4* np = s.fast_nodes + 2 * (size_t)cur;
        float4 a0 = __ldg(np + 0), a1 = __ldg(np + 1), b0 = __ldg(np + 2), b1 = __ldg(np + 3);
        if (COUNT) wc->nodes += 2;
        float d0, d1;
        bool h0 = slab_conservative(a0, a1, ray, d0);
        bool h1 = slab_conservative(b0, b1, ray, d1);
        h0 = h0 && !(d0 > best_t);
        h1 = h1 && !(d1 > best_t);
        uint32_t l0 = __float_as_uint(a0.w), c0 = __float_as_uint(a1.w);
        uint32_t l1 = __float_as_uint(b0.w), c1 = __float_as_uint(b1.w);
        if (h0 && h1 && d1 < d0) {
            float td = d0; d0 = d1; d1 = td;
            uint32_t tl = l0; l0 = l1; l1 = tl;
            uint32_t tc = c0; c0 = c1; c1 = tc;
        }
        if (h0 && c0 != 0) {
            if (leaf(l0, c0)) return true;
            h0 = false;
            if (h1 && d1 > best_t) h1 = false;
        }
        if (h1 && c1 != 0) {
            if (leaf(l1, c1)) return true;
            h1 = false;
        }
        if (h0) {
            cur = l0;
            if (h1 && sp < kStackSize) {
                stack_i[sp] = l1;
                stack_d[sp] = d1;
                ++sp;
            }
            continue;
        }
        if (h1) {
            cur = l1;
            continue;
        }
        bool popped = false;
        while (sp > 0) {
            --sp;
            if (stack_d[sp] > best_t) continue;
            cur = stack_i[sp];
            popped = true;
            break;
        }
        if (!popped) break;
    }
    return finish();
}

// MODE 0: performance layer (product path); 1: reference tree, exhaustive, reference visiting rule;
// 2: reference tree, ordered + distance-culled ("flattened reference BVH").
template <int MODE, bool ANY, bool COUNT>
__device__ __forceinline__ bool trace_ray(const GpuSceneView& s, const RayD& ray, HitD& hit, WalkCounters* wc) {
    if (MODE == 0) return walk_fast<ANY, COUNT>(s, ray, hit, wc);
    if (MODE == 1) return walk<false, ANY, COUNT>(s, ray, hit, wc);
    return walk<true, ANY, COUNT>(s, ray, hit, wc);
}

} // namespace rb
