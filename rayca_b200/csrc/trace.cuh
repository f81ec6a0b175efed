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
#ifndef RAYCA_SMEM_STACK
#define RAYCA_SMEM_STACK 0
#endif

// RAYCA_TREELET = K > 0: the first K sibling pairs of the performance tree (64 B each; the device builder lays the
// levels above the 32-primitive subtrees out breadth-first, so these are its top levels) are copied into shared
// memory once per block with one bulk asynchronous copy (cp.async.bulk, 1-D TMA) and the walk reads them from there.
// A/B only (tools/variants.py "tl*"), measured no better: profiles/r02_ab_walk_variants.txt.
#ifndef RAYCA_ANY_NODIST
#define RAYCA_ANY_NODIST 1 // any-hit walks keep 4-byte stack entries (below); 0 restores 8-byte entries everywhere for A/B runs
#endif
#ifndef RAYCA_PREFETCH_FAR
#define RAYCA_PREFETCH_FAR 0 // prefetch the pushed (far) child's record into L1 (tools/variants.py "pf"): measured in profiles/r02_ab_walk_variants.txt
#endif
#ifndef RAYCA_TREELET
#define RAYCA_TREELET 0
#endif
#if RAYCA_TREELET > 0
__shared__ __align__(128) float4 g_treelet[RAYCA_TREELET * 4];
__shared__ __align__(8) unsigned long long g_treelet_bar;
__device__ __forceinline__ uint32_t treelet_records(const GpuSceneView& s) {
    return min((uint32_t)RAYCA_TREELET * 2u, s.fast_record_count & ~1u);
}
// every thread of the block calls this once, before its first walk
__device__ __forceinline__ void stage_treelet(const GpuSceneView& s) {
    const uint32_t bytes = treelet_records(s) * 32u;
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&g_treelet_bar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        if (bytes) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                             (uint32_t)__cvta_generic_to_shared(g_treelet)),
                         "l"(s.fast_nodes), "r"(bytes), "r"(bar)
                         : "memory");
        } else {
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
        }
    }
    __syncthreads(); // the barrier is initialised for everyone
    uint32_t done = 0;
    while (!done) {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0, 1, 0, p;\n}"
                     : "=r"(done)
                     : "r"(bar)
                     : "memory");
    }
}
#else
__device__ __forceinline__ void stage_treelet(const GpuSceneView&) {}
#endif

struct RayD {
    V3 o, d, rd;
};
__device__ __forceinline__ RayD make_ray(V3 o, V3 d) { return RayD{o, d, reciprocal(d)}; } // ray.rs:63-72

struct HitD {
    float t;
    uint32_t prim; // global triangle index, ~0u = miss
    float u, v;
    V3 p;
    uint32_t sphere; // the hit primitive is a sphere (BvhGeometry::Sphere): uv = 0, see load_surf
};

struct WalkCounters {
    unsigned long long nodes = 0;  // 32 B node records fetched
    unsigned long long tris = 0;   // 48 B: three vertex quads of a triangle record fetched
    unsigned long long planes = 0; // 16 B: plane quad of a performance-layer triangle record fetched
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
                } else if (t < best_t || (t == best_t && i < best && best != 0xFFFFFFFFu)) {
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
        hit.sphere = 0; // the reference-tree walks handle triangles only (scene_create refuses otherwise)
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
//   * the slab test here is CONSERVATIVE (see FastRay), so rounding cannot hide a triangle the
//     reference would reach;
//   * a candidate hit only counts if the exact slab test of its REFERENCE leaf box (nodes[leaf])
//     accepts the ray — that is precisely the condition under which the reference tests the triangle;
//   * ties on depth resolve to the lowest reference-global index, like depth-first order does;
//   * the triangle test performs the reference's operations on the reference's operands: the plane
//     (n, d) and dot(n,n) of the 64-byte record are precomputed on the host with the very same f32
//     operation order (rmath.h), so every accepted (t, u, v) is bit-identical.  The test is only
//     REORDERED: plane distance first (16 bytes), then `t > best` rejects before the edge functions
//     — a hit with t > best would be discarded anyway.
//
// Instruction budget (this loop is issue-bound, profiles/): per box 6 LOP3 + 6 FFMA + 2 FMNMX3 + 2 FFMA.
//   exact      t  = RN(RN(p - o) * rd)                         (aabb.rs:76-80)
//   here       t' = RN(fma(p, rd, a)),  a = RN(-o*rd) -/+ e,   e = 2^-22 |o*rd| per axis
//   |t - t'| <= u|o*rd| + 3u|t| with u = 2^-24; e = 4u|o*rd| covers the first term (and the rounding of
//   a itself), folded into the addend of the near (-e) and far (+e) plane; the second term and the
//   triangle-test-vs-box rounding for grazing rays are covered by widening the resulting interval by
//   1e-5 relative (fma(|t|, -/+1e-5, t)).  Near/far planes are picked per axis by the sign of rd, so no
//   min/max per axis is needed.  NaNs (0 * inf for a denormal direction lane) are dropped by
//   fminf/fmaxf, which only ever widens the interval.
constexpr float kSlabTol = 1.0e-5f;
// child reference: bit31 leaf, bits 28..30 count-1, bit 27 the leaf holds spheres, bits 0..26 first record
constexpr uint32_t kLeafFlag = 0x80000000u, kSphereFlag = 0x08000000u;

#ifndef RAYCA_FOLD_TOL
#define RAYCA_FOLD_TOL 0
#endif
// RAYCA_SOA_PAIRS: sibling pairs are stored plane-major — {lo0.x lo1.x hi0.x hi1.x | y | z | ref0 ref1 - -}, 64 bytes —
// so the near / far planes of BOTH children are one 64-bit load each at a per-ray offset (0 or 8 bytes into the
// axis' quad, chosen once from the sign of rd): the 12 LOP3 selects per visited pair (the hottest source line of the
// round-1 kernel, 13 % of its samples) become address arithmetic.  The records are transposed on the device after the
// build (gpu_fast_build.cu::transpose_pairs_device).
#ifndef RAYCA_SOA_PAIRS
#define RAYCA_SOA_PAIRS 0
#endif
#if RAYCA_SOA_PAIRS && !RAYCA_FOLD_TOL
#error "RAYCA_SOA_PAIRS needs RAYCA_FOLD_TOL (the pair test reads the folded per-ray constants)"
#endif
#if RAYCA_FOLD_TOL
// The relative widening of the interval is FOLDED into the per-ray constants: the near planes use rd (1 - tol') and
// a_near (1 - tol'), the far planes rd (1 + tol') and a_far (1 + tol'), so a box costs 6 LOP3 + 6 FFMA + 2 FMNMX3 and
// nothing else.  Why this accepts everything the explicit form `fma(|t|, -/+tol, t)` accepted:
//   * tn = max_i v_i.  If tn > 0 the maximum is attained by a positive lane and scaling every lane by (1 - tol')
//     scales the maximum; negative lanes move towards zero but stay negative, and a negative entry distance is never
//     compared with anything but `exit >= entry` (true whenever exit >= 0) and `entry > best` (false).
//   * tf = min_i f_i.  If every far lane is >= 0 the minimum scales by (1 + tol'); if one is negative the box is
//     behind the origin and both forms reject.
//   * the products rd (1 -/+ tol') and a (1 -/+ tol') are rounded (2^-24 relative, and p*rd, a can each be as large as
//     |o*rd| while their sum is small), so the per-axis bound e doubles to 2^-21 |o*rd| and tol' = 1.1e-5 > 1e-5.
struct FastRay {
    float rnx, rny, rnz; // rd (1 - tol')
    float rfx, rfy, rfz; // rd (1 + tol')
    float anx, any, anz; // near-plane addend  (RN(-o*rd) - e)(1 - tol')
    float afx, afy, afz; // far-plane addend   (RN(-o*rd) + e)(1 + tol')
    uint32_t sx, sy, sz; // rd < 0 ? ~0 : 0
};
__device__ __forceinline__ float pick(uint32_t m, float a, float b) { // m ? a : b on the bit patterns
    return __uint_as_float((__float_as_uint(a) & m) | (__float_as_uint(b) & ~m));
}
__device__ __forceinline__ FastRay make_fast_ray(const RayD& r) {
    FastRay f;
    constexpr float kLo = 1.0f - 1.1e-5f, kHi = 1.0f + 1.1e-5f;
    f.rnx = r.rd.x * kLo; f.rny = r.rd.y * kLo; f.rnz = r.rd.z * kLo;
    f.rfx = r.rd.x * kHi; f.rfy = r.rd.y * kHi; f.rfz = r.rd.z * kHi;
    const float nx = -(r.o.x * r.rd.x), ny = -(r.o.y * r.rd.y), nz = -(r.o.z * r.rd.z);
    const float ex = fabsf(nx) * 4.76837158203125e-07f, ey = fabsf(ny) * 4.76837158203125e-07f,
                ez = fabsf(nz) * 4.76837158203125e-07f; // 2^-21
    f.anx = (nx - ex) * kLo; f.any = (ny - ey) * kLo; f.anz = (nz - ez) * kLo;
    f.afx = (nx + ex) * kHi; f.afy = (ny + ey) * kHi; f.afz = (nz + ez) * kHi;
    asm("shr.s32 %0, %1, 31;" : "=r"(f.sx) : "r"(__float_as_int(r.rd.x)));
    asm("shr.s32 %0, %1, 31;" : "=r"(f.sy) : "r"(__float_as_int(r.rd.y)));
    asm("shr.s32 %0, %1, 31;" : "=r"(f.sz) : "r"(__float_as_int(r.rd.z)));
#if RAYCA_SOA_PAIRS
    f.sx = f.sx & 8u;          // near planes of x: lo at +0 (rd > 0) or hi at +8 (rd < 0); the far planes sit at offset ^ 8
    f.sy = 16u + (f.sy & 8u);
    f.sz = 32u + (f.sz & 8u);
#endif
    return f;
}
#if RAYCA_SOA_PAIRS
__device__ __forceinline__ float2 ldg_f2(const char* base, uint32_t off) {
    return __ldg(reinterpret_cast<const float2*>(base + off));
}
// both children of the pair at byte offset `pair` of `nodes`: entry distances d0 / d1, "may intersect" h0 / h1, references
__device__ __forceinline__ void slab_pair(const char* nodes, uint32_t pair, const FastRay& f, float& d0, float& d1, bool& h0, bool& h1,
                                          uint32_t& c0, uint32_t& c1) {
    const float2 nx = ldg_f2(nodes, pair + f.sx), fx = ldg_f2(nodes, pair + (f.sx ^ 8u));
    const float2 ny = ldg_f2(nodes, pair + f.sy), fy = ldg_f2(nodes, pair + (f.sy ^ 8u));
    const float2 nz = ldg_f2(nodes, pair + f.sz), fz = ldg_f2(nodes, pair + (f.sz ^ 8u));
    const uint2 refs = __ldg(reinterpret_cast<const uint2*>(nodes + (pair + 48u)));
    d0 = fmaxf(fmaxf(__fmaf_rn(nx.x, f.rnx, f.anx), __fmaf_rn(ny.x, f.rny, f.any)), __fmaf_rn(nz.x, f.rnz, f.anz));
    d1 = fmaxf(fmaxf(__fmaf_rn(nx.y, f.rnx, f.anx), __fmaf_rn(ny.y, f.rny, f.any)), __fmaf_rn(nz.y, f.rnz, f.anz));
    const float e0 = fminf(fminf(__fmaf_rn(fx.x, f.rfx, f.afx), __fmaf_rn(fy.x, f.rfy, f.afy)), __fmaf_rn(fz.x, f.rfz, f.afz));
    const float e1 = fminf(fminf(__fmaf_rn(fx.y, f.rfx, f.afx), __fmaf_rn(fy.y, f.rfy, f.afy)), __fmaf_rn(fz.y, f.rfz, f.afz));
    h0 = e0 >= d0 && e0 >= 0.0f;
    h1 = e1 >= d1 && e1 >= 0.0f;
    c0 = refs.x;
    c1 = refs.y;
}
#endif
// returns "may intersect"; entry <= the exact slab entry distance of the same box whenever that is positive
__device__ __forceinline__ bool slab_fast(const float4 lo, const float4 hi, const FastRay& f, float& entry) {
#if RAYCA_SOA_PAIRS
    const uint32_t mx = 0u - ((f.sx >> 3) & 1u), my = 0u - ((f.sy >> 3) & 1u), mz = 0u - ((f.sz >> 3) & 1u); // root box only
#else
    const uint32_t mx = f.sx, my = f.sy, mz = f.sz;
#endif
    const float tn = fmaxf(fmaxf(__fmaf_rn(pick(mx, hi.x, lo.x), f.rnx, f.anx), __fmaf_rn(pick(my, hi.y, lo.y), f.rny, f.any)),
                           __fmaf_rn(pick(mz, hi.z, lo.z), f.rnz, f.anz));
    const float tf = fminf(fminf(__fmaf_rn(pick(mx, lo.x, hi.x), f.rfx, f.afx), __fmaf_rn(pick(my, lo.y, hi.y), f.rfy, f.afy)),
                           __fmaf_rn(pick(mz, lo.z, hi.z), f.rfz, f.afz));
    entry = tn;
    return tf >= tn && tf >= 0.0f;
}
#else
struct FastRay {
    float rdx, rdy, rdz;
    float anx, any, anz; // near-plane addend  RN(-o*rd) - e
    float afx, afy, afz; // far-plane addend   RN(-o*rd) + e
    uint32_t sx, sy, sz; // rd < 0 ? ~0 : 0 — near/far planes are picked with one LOP3 each (FSEL would need the
                         // three sign predicates, which the compiler re-derives with 3 FSETP on every visit)
};
__device__ __forceinline__ float pick(uint32_t m, float a, float b) { // m ? a : b on the bit patterns
    return __uint_as_float((__float_as_uint(a) & m) | (__float_as_uint(b) & ~m));
}

__device__ __forceinline__ FastRay make_fast_ray(const RayD& r) {
    FastRay f;
    f.rdx = r.rd.x; f.rdy = r.rd.y; f.rdz = r.rd.z;
    const float nx = -(r.o.x * r.rd.x), ny = -(r.o.y * r.rd.y), nz = -(r.o.z * r.rd.z);
    const float ex = fabsf(nx) * 2.384185791015625e-07f, ey = fabsf(ny) * 2.384185791015625e-07f,
                ez = fabsf(nz) * 2.384185791015625e-07f; // 2^-22
    f.anx = nx - ex; f.any = ny - ey; f.anz = nz - ez;
    f.afx = nx + ex; f.afy = ny + ey; f.afz = nz + ez;
    // sign bit smeared over the word; asm keeps the compiler from turning the masks back into predicates + FSEL
    // (rd is never +-0 in the walk: zero-direction rays skip it, see walk_fast_body)
    asm("shr.s32 %0, %1, 31;" : "=r"(f.sx) : "r"(__float_as_int(r.rd.x)));
    asm("shr.s32 %0, %1, 31;" : "=r"(f.sy) : "r"(__float_as_int(r.rd.y)));
    asm("shr.s32 %0, %1, 31;" : "=r"(f.sz) : "r"(__float_as_int(r.rd.z)));
    return f;
}

// returns "may intersect"; entry <= the exact slab entry distance of the same box
__device__ __forceinline__ bool slab_fast(const float4 lo, const float4 hi, const FastRay& f, float& entry) {
    const float tn = fmaxf(fmaxf(__fmaf_rn(pick(f.sx, hi.x, lo.x), f.rdx, f.anx), __fmaf_rn(pick(f.sy, hi.y, lo.y), f.rdy, f.any)),
                           __fmaf_rn(pick(f.sz, hi.z, lo.z), f.rdz, f.anz));
    const float tf = fminf(fminf(__fmaf_rn(pick(f.sx, lo.x, hi.x), f.rdx, f.afx), __fmaf_rn(pick(f.sy, lo.y, hi.y), f.rdy, f.afy)),
                           __fmaf_rn(pick(f.sz, lo.z, hi.z), f.rdz, f.afz));
    entry = __fmaf_rn(fabsf(tn), -kSlabTol, tn);
    const float ext = __fmaf_rn(fabsf(tf), kSlabTol, tf);
    return ext >= entry && ext >= 0.0f;
}

#endif // RAYCA_FOLD_TOL

// SPHERES: the scene holds sphere primitives (kernel instantiation MODE 3); triangle-only scenes run the
// instantiation without the sphere leaf code (it costs 6 % on the atrium frame when compiled in).
template <bool COUNT, bool SPHERES>
__device__ __forceinline__ bool walk_fast_body(const GpuSceneView& s, const RayD& ray, HitD& hit, WalkCounters* wc, const bool ANY) {
    const V3 o = ray.o, d = ray.d;
    float best_t = ANY ? hit.t : __int_as_float(0x7f800000);
    uint32_t best = 0xFFFFFFFFu;
    float bu = 0.0f, bv = 0.0f, bden = 1.0f;
    uint32_t bsphere = 0;
    uint32_t ok_leaf = 0xFFFFFFFFu; // last reference leaf whose exact slab test accepted this ray

    auto finish = [&]() -> bool {
        if (ANY) return false;
        hit.t = best_t;
        hit.prim = best;
        hit.u = bu / bden; // uv = (u / |n|^2, v / |n|^2) (triangle.rs:152-153)
        hit.v = bv / bden;
        hit.sphere = SPHERES ? bsphere : 0u;
        if (SPHERES && bsphere) { // hit.point = trs * (inverse ray origin + dir * t) (sphere.rs:141,160-161)
            const float4* a = s.tri_attr + 12 * (size_t)best;
            const float4 wt = __ldg(a + 6), wq = __ldg(a + 7), ws = __ldg(a + 8);
            const TrsF w{V3{wt.x, wt.y, wt.z}, Q4{wq.x, wq.y, wq.z, wq.w}, V3{ws.x, ws.y, ws.z}};
            const Q4 qi = conj(w.q);
            const V3 si = reciprocal(w.s);
            const V3 lo = scale_point(rotate(o + (-w.t), qi), si);
            const V3 ld = rotate(d, qi) * si;
            hit.p = trs_point(w, lo + ld * best_t);
        } else {
            hit.p = o + d * best_t; // Hit::point, the `p` of the winning test (triangle.rs:121)
        }
        return best != 0xFFFFFFFFu;
    };
    // A ray with an exactly-zero direction lane has rdir = 0 there (vec3.rs:195-216): both slab
    // distances of that axis are 0, tmax <= 0, and `tmax > 0.0` (aabb.rs:88) rejects EVERY box, the
    // TLAS root included — the reference returns a miss for such rays (SURVEY A-3).
    //
    // Control flow is kept STRUCTURED — no return from inside the loops, one break target per loop, an
    // empty stack shows up as the sentinel reference kDone popped from slot 0 — so that, inlined into
    // a caller's loop, the compiler still brackets the inner node loop and the leaf loop with convergence
    // barriers (a first inlined version with early returns lost them: lanes that left the node loop for
    // a leaf never rejoined the others, 0.90 -> 1.43 ms on the 1080p atrium frame).
    bool found = false; // ANY: an accepted hit below hit.t exists
    bool go = !(ray.rd.x == 0.0f || ray.rd.y == 0.0f || ray.rd.z == 0.0f);

    const float4* __restrict__ fast_nodes = s.fast_nodes;
    const float4* __restrict__ fast_tri = s.fast_tri;
    const float4* __restrict__ ref_nodes = s.nodes;
    const FastRay fr = make_fast_ray(ray);
#if RAYCA_TREELET > 0
    const uint32_t treelet_top = treelet_records(s);
#endif

    auto ref_leaf_accepts = [&](uint32_t leaf) -> bool {
        if (leaf == ok_leaf) return true;
        const float4 lo = __ldg(&ref_nodes[2 * (size_t)leaf]);
        const float4 hi = __ldg(&ref_nodes[2 * (size_t)leaf + 1]);
        if (COUNT) wc->nodes += 1;
        if (slab(lo, hi, ray) == kF32Max) return false;
        ok_leaf = leaf;
        return true;
    };
    if (go) {
        if (COUNT) wc->nodes += 1;
        float4 lo = make_float4(s.fast_root_min[0], s.fast_root_min[1], s.fast_root_min[2], 0.0f);
        float4 hi = make_float4(s.fast_root_max[0], s.fast_root_max[1], s.fast_root_max[2], 0.0f);
        float e;
        go = slab_fast(lo, hi, fr, e);
    }
    constexpr uint32_t kDone = 0xFFFFFFFFu; // leaf flag set: leaves the node loop like any leaf reference
    // {child reference, entry distance}; slot 0 holds the sentinel (distance -inf is never culled), the last
    // slot absorbs pushes beyond kStackSize entries (the count saturates; scene_create refuses trees that deep)
#if RAYCA_SMEM_STACK > 0
    // The first RAYCA_SMEM_STACK slots of every thread live in SHARED memory, [slot][thread] so that a warp's
    // accesses to one slot are 256 contiguous bytes (no bank conflicts); deeper slots spill to the local array,
    // which a well-built tree touches rarely.  With the whole stack in local memory every resident thread owned
    // 784 B of it, 119 MB per GPU — more than L2 keeps next to the scene — and the evictions showed up as 334 MB
    // of DRAM writes per 1080p frame (profiles/r01_ncu_render_kernel_v4.json).
    constexpr int kSm = RAYCA_SMEM_STACK;
    __shared__ uint2 s_stack[kSm][256];
    uint2 stack[kStackSize + 2 - kSm];
    const uint32_t tid = threadIdx.x;
    s_stack[0][tid] = make_uint2(kDone, 0xff800000u);
    int sp = 1;
    uint32_t cur = go ? s.fast_root_ref : kDone;
    auto push = [&](uint32_t ref, float dist) {
        const uint2 e = make_uint2(ref, __float_as_uint(dist));
        if (sp < kSm) s_stack[sp][tid] = e;
        else stack[sp - kSm] = e;
        sp = min(sp + 1, kStackSize + 1);
    };
    auto pop = [&]() -> uint32_t { // next entry whose entry distance does not exceed the best depth
        uint2 e;
        do {
            --sp;
            e = sp < kSm ? s_stack[sp][tid] : stack[sp - kSm];
        } while (__uint_as_float(e.y) > best_t);
        return e.x;
    };
#else
    uint2 stack[kStackSize + 2];
    stack[0] = make_uint2(kDone, 0xff800000u);
    int sp = 1;
    uint32_t cur = go ? s.fast_root_ref : kDone;
#if RAYCA_ANY_NODIST
    // An any-hit query never shortens its ray, so an entry that was pushed is never culled later: its distance word is
    // neither stored nor read — the two words of an entry live in different 128-byte lines of local memory, so the stack
    // of a shadow ray takes half the L1 lines (+0.2 % on the atrium frame, +0.8 % on the bounce frame: profiles/r02_ab_walk_variants.txt).
    auto push = [&](uint32_t ref, float dist) {
        if (ANY) stack[sp].x = ref;
        else stack[sp] = make_uint2(ref, __float_as_uint(dist));
        sp = min(sp + 1, kStackSize + 1);
    };
    auto pop = [&]() -> uint32_t {
        if (ANY) return stack[--sp].x;
        uint2 e;
        do {
            e = stack[--sp];
        } while (__uint_as_float(e.y) > best_t);
        return e.x;
    };
#else
    auto push = [&](uint32_t ref, float dist) {
        stack[sp] = make_uint2(ref, __float_as_uint(dist)); // unguarded store, saturating count
        sp = min(sp + 1, kStackSize + 1);
    };
    auto pop = [&]() -> uint32_t { // next entry whose entry distance does not exceed the best depth
        uint2 e;
        do {
            e = stack[--sp];
        } while (__uint_as_float(e.y) > best_t);
        return e.x;
    };
#endif
#endif
    for (;;) {
        // inner nodes: descend to the nearer accepted child, push the farther one
        while (!(cur & kLeafFlag)) {
#if RAYCA_SOA_PAIRS
            if (COUNT) wc->nodes += 2;
            float d0, d1;
            bool h0, h1;
            uint32_t c0, c1;
            slab_pair(reinterpret_cast<const char*>(fast_nodes), cur * 32u, fr, d0, d1, h0, h1, c0, c1);
            h0 = h0 && !(d0 > best_t);
            h1 = h1 && !(d1 > best_t);
#else
#if RAYCA_TREELET > 0
            float4 a0, a1, b0, b1;
            if (cur < treelet_top) {
                const float4* np = g_treelet + cur * 2u;
                a0 = np[0]; a1 = np[1]; b0 = np[2]; b1 = np[3];
            } else {
                const float4* np = reinterpret_cast<const float4*>(reinterpret_cast<const char*>(fast_nodes) + (unsigned long long)cur * 32ull);
                a0 = __ldg(np + 0); a1 = __ldg(np + 1); b0 = __ldg(np + 2); b1 = __ldg(np + 3);
            }
#else
            const float4* np = reinterpret_cast<const float4*>(reinterpret_cast<const char*>(fast_nodes) + (unsigned long long)cur * 32ull);
            const float4 a0 = __ldg(np + 0), a1 = __ldg(np + 1), b0 = __ldg(np + 2), b1 = __ldg(np + 3);
#endif
            if (COUNT) wc->nodes += 2;
            float d0, d1;
            bool h0 = slab_fast(a0, a1, fr, d0);
            bool h1 = slab_fast(b0, b1, fr, d1);
            h0 = h0 && !(d0 > best_t);
            h1 = h1 && !(d1 > best_t);
            uint32_t c0 = __float_as_uint(a0.w), c1 = __float_as_uint(b0.w);
#endif
            if (h0 && h1) {
                if (d1 < d0) {
                    float td = d0; d0 = d1; d1 = td;
                    uint32_t tc = c0; c0 = c1; c1 = tc;
                }
                cur = c0;
                push(c1, d1);
#if RAYCA_PREFETCH_FAR
                // the far child will be popped later unless a closer hit culls it: ask for its line now (A/B only)
                {
                    const char* fa = (c1 & kLeafFlag) ? reinterpret_cast<const char*>(fast_tri) + (unsigned long long)(c1 & 0x07FFFFFFu) * 64ull
                                                      : reinterpret_cast<const char*>(fast_nodes) + (unsigned long long)c1 * 32ull;
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(fa));
                }
#endif
            } else if (h0) {
                cur = c0;
            } else if (h1) {
                cur = c1;
            } else {
                cur = pop();
            }
        }
        if (cur == kDone) break;
        // leaf: Triangle::intersects (rayca-geometry/src/triangle.rs:84-159) on the 64-byte records
        // {n.xyz, d | v0, ref index | v1, ref leaf | v2, dot(n,n)}
        const uint32_t first = cur & 0x07FFFFFFu, count = ((cur >> 28) & 7u) + 1u;
        if (SPHERES && (cur & kSphereFlag)) {
            // Sphere::intersects (rayca-geometry/src/sphere.rs:101-163) on the 64-byte records
            // {-t, r^2 | q^-1 | 1/s, ref index | centre, ref leaf}: the ray goes to model space by translate,
            // rotate, scale (trs.rs:405-414 with ray.rs:74-91), then the geometric quadratic
            for (uint32_t i = first; i < first + count && !found; ++i) {
                const float4* q = fast_tri + 4 * (size_t)i;
                const float4 q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3);
                if (COUNT) { wc->planes += 1; wc->tris += 1; }
                const Q4 qi{q1.x, q1.y, q1.z, q1.w};
                const V3 si{q2.x, q2.y, q2.z};
                const V3 lo = scale_point(rotate(o + V3{q0.x, q0.y, q0.z}, qi), si);
                const V3 ld = rotate(d, qi) * si;
                const float a = dot(ld, ld);
                const V3 c_to_r = lo - V3{q3.x, q3.y, q3.z};
                const float b = dot(c_to_r, ld);
                const float c = dot(c_to_r, c_to_r) - q0.w;
                const float det = b * b - a * c;
                if (det < 0.0f) continue;
                const float det_sqrt = sqrtf(det);
                const float t0 = (-b + det_sqrt) / a;
                const float t1 = (-b - det_sqrt) / a;
                if (t0 < 0.0f && t1 < 0.0f) continue;
                float t;
                if (t0 >= 0.0f && t1 >= 0.0f) t = fminf(t0, t1);
                else if (t0 >= 0.0f) t = t0;
                else t = t1;
                const uint32_t id = __float_as_uint(q2.w);
                if (ANY) {
                    if (t < best_t && ref_leaf_accepts(__float_as_uint(q3.w))) found = true;
                } else if (t < best_t || (t == best_t && id < best && best != 0xFFFFFFFFu)) {
                    if (ref_leaf_accepts(__float_as_uint(q3.w))) {
                        best_t = t;
                        best = id;
                        bu = 0.0f;
                        bv = 0.0f;
                        bden = 1.0f;
                        bsphere = 1;
                    }
                }
            }
        } else {
            for (uint32_t i = first; i < first + count && !found; ++i) {
                const float4* q = fast_tri + 4 * (size_t)i;
                const float4 q0 = __ldg(q);
                if (COUNT) wc->planes += 1;
                const V3 n{q0.x, q0.y, q0.z};
                const float ndd = dot(d, n);
                if (ndd > 0.0f) continue;             // back face (:96)
                if (fabsf(ndd) < kF32Eps) continue;   // parallel (:106)
                const float t = -(dot(n, o) + q0.w) / ndd;
                if (t < 0.0f) continue;               // behind (:117)
                if (ANY ? !(t < best_t) : (t > best_t)) continue; // cannot win: skip the edge functions
                const float4 q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3);
                if (COUNT) wc->tris += 1;
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
                if (ANY) {
                    if (ref_leaf_accepts(__float_as_uint(q2.w))) found = true;
                } else if (t < best_t || (t == best_t && id < best && best != 0xFFFFFFFFu)) { // a NaN t never replaces a hit (strict `<`, blas.rs:150)
                    if (ref_leaf_accepts(__float_as_uint(q2.w))) {
                        best_t = t;
                        best = id;
                        bu = u;
                        bv = v;
                        bden = q3.w;
                        bsphere = 0;
                    }
                }
            }
        }
        if (found) break;
        cur = pop();
    }
    if (found) return true;
    return finish();
}

// One shared copy of the walk serves closest-hit and any-hit queries of every call site: the kernel
// starves on instruction fetch when the walk is inlined two or three times (measured 1.9 ms -> 1.25 ms
// for the 1080p atrium frame, profiles/).  RAYCA_INLINE_WALK restores the inlined form for A/B runs.
#ifndef RAYCA_INLINE_WALK
template <bool COUNT, bool SPHERES>
__device__ __noinline__ bool walk_fast_shared(const GpuSceneView& s, const RayD& ray, HitD& hit, WalkCounters* wc, bool any) {
    return walk_fast_body<COUNT, SPHERES>(s, ray, hit, wc, any);
}
template <bool ANY, bool COUNT, bool SPHERES>
__device__ __forceinline__ bool walk_fast(const GpuSceneView& s, const RayD& ray, HitD& hit, WalkCounters* wc) {
    return walk_fast_shared<COUNT, SPHERES>(s, ray, hit, wc, ANY);
}
#else
template <bool ANY, bool COUNT, bool SPHERES>
__device__ __forceinline__ bool walk_fast(const GpuSceneView& s, const RayD& ray, HitD& hit, WalkCounters* wc) {
    return walk_fast_body<COUNT, SPHERES>(s, ray, hit, wc, ANY);
}
#endif

// MODE 0: performance layer (product path), 3: the same for scenes with spheres; 1: reference tree, exhaustive, reference visiting rule;
// 2: reference tree, ordered + distance-culled ("flattened reference BVH").
template <int MODE, bool ANY, bool COUNT>
__device__ __forceinline__ bool trace_ray(const GpuSceneView& s, const RayD& ray, HitD& hit, WalkCounters* wc) {
    if (MODE == 0) return walk_fast<ANY, COUNT, false>(s, ray, hit, wc);
    if (MODE == 3) return walk_fast<ANY, COUNT, true>(s, ray, hit, wc);
    if (MODE == 1) return walk<false, ANY, COUNT>(s, ray, hit, wc);
    return walk<true, ANY, COUNT>(s, ray, hit, wc);
}

} // namespace rb
