// rmath.h — float32 vector maths shared by the host flattening code and the device kernels.
//
// Bit-parity contract with rayca-math (the reference's `std::simd` f32x4 types): every helper
// performs the same IEEE-754 single-precision operations in the same order as the Rust code it
// cites, and nothing here may be contracted into FMAs.  Device code is therefore compiled with
// `--fmad=false`, host code with `-ffp-contract=off` (see rayca_b200/build.py).
//   dot   : ((x*x' + y*y') + z*z') + 0        rayca-math/src/vec3.rs:240-243 (ordered reduce_sum)
//   cross : a.yzx*b.zxy - (a.yzx*b).yzx       rayca-math/src/vec3.rs:134-142
//   rotate: 2(u.v)u + (s*s - u.u)v + 2s(u x v) rayca-math/src/vec3.rs:148-159, point3.rs:65-76
#pragma once
#include <cmath>
#include <cstdint>

#if defined(__CUDACC__)
#define RHD __host__ __device__ __forceinline__
#else
#define RHD inline
#endif

#ifndef RAYCA_SHADE_OPT
#define RAYCA_SHADE_OPT 0
#endif

namespace rb {

constexpr float kF32Eps = 1.1920929e-07f;          // f32::EPSILON
constexpr float kEps = kF32Eps * 8192.0f;          // rayca-math/src/lib.rs:33
constexpr float kF32Max = 3.402823466e+38f;        // f32::MAX
constexpr float kPi = 3.14159265358979323846f;
constexpr float kFrac1Pi = 0.318309886183790671538f;
constexpr float kRayBias = 1e-4f;                  // rayca-math/src/ray.rs:57

struct V3 {
    float x, y, z;
};
struct Q4 {
    float x, y, z, w;
};
struct Col {
    float r, g, b, a;
};
struct TrsF {
    V3 t;
    Q4 q;
    V3 s;
};

RHD V3 v3(float x, float y, float z) { return V3{x, y, z}; }
RHD V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
RHD V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
RHD V3 operator-(V3 a) { return V3{-a.x, -a.y, -a.z}; }
RHD V3 operator*(V3 a, V3 b) { return V3{a.x * b.x, a.y * b.y, a.z * b.z}; }
RHD V3 operator*(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
RHD V3 operator*(float s, V3 a) { return V3{a.x * s, a.y * s, a.z * s}; }
RHD V3 operator/(V3 a, float s) { return V3{a.x / s, a.y / s, a.z / s}; }

RHD float dot(V3 a, V3 b) {
    float p0 = a.x * b.x, p1 = a.y * b.y, p2 = a.z * b.z;
    return ((p0 + p1) + p2) + 0.0f;
}
RHD V3 cross(V3 a, V3 b) {
    float t2x = a.y * b.x, t2y = a.z * b.y, t2z = a.x * b.z;
    float t3x = a.y * b.z, t3y = a.z * b.x, t3z = a.x * b.y;
    return V3{t3x - t2y, t3y - t2z, t3z - t2x};
}
RHD float norm2(V3 a) { return dot(a, a); }
RHD float length(V3 a) { return sqrtf(dot(a, a)); }
// vec3.rs:183-188 — no-op when len <= EPS
#if defined(__CUDACC__)
// one out-of-line copy: three IEEE divisions and a square root are ~50 SASS instructions, and the kernel
// inlines some twenty normalisations into straight-line shading code that already starves on instruction fetch
__host__ __device__ __noinline__ inline V3 normalize(V3 a) {
#else
RHD V3 normalize(V3 a) {
#endif
    float l = length(a);
    if (l > kEps) return V3{a.x / l, a.y / l, a.z / l};
    return a;
}
// vec3.rs:195-216 — zero lanes give a ZERO reciprocal, never an infinity
#if defined(__CUDA_ARCH__) && RAYCA_SHADE_OPT
// rcp.rn.f32 is the correctly rounded reciprocal, i.e. the very value of the IEEE division 1.0f / d, without the
// general division's quotient refinement (make_ray runs three of these per ray)
RHD float recip_lane(float d) { return d == 0.0f ? 0.0f : __frcp_rn(d); }
#else
RHD float recip_lane(float d) { return d == 0.0f ? 0.0f : 1.0f / d; }
#endif
RHD V3 reciprocal(V3 a) { return V3{recip_lane(a.x), recip_lane(a.y), recip_lane(a.z)}; }
// vec3.rs:219-221
RHD V3 reflect(V3 v, V3 n) {
    float k = 2.0f * dot(v, n);
    return v - n * k;
}
RHD V3 rotate(V3 v, Q4 q) {
    V3 u{q.x, q.y, q.z};
    float s = q.w;
    V3 a = (2.0f * dot(u, v)) * u;
    V3 b = (s * s - dot(u, u)) * v;
    V3 c = (2.0f * s) * cross(u, v);
    return (a + b) + c;
}
// point3.rs:59-63 — the one explicit mul_add of the hot path; fma(x, s, +0) == RN(x*s)
RHD V3 scale_point(V3 p, V3 s) { return V3{fmaf(p.x, s.x, 0.0f), fmaf(p.y, s.y, 0.0f), fmaf(p.z, s.z, 0.0f)}; }
// trs.rs:264-273 — scale, rotate, translate
RHD V3 trs_point(const TrsF& t, V3 p) { return rotate(scale_point(p, t.s), t.q) + t.t; }
// trs.rs:253-262
RHD V3 trs_vec(const TrsF& t, V3 v) { return rotate(v * t.s, t.q) + t.t; }
RHD Q4 conj(Q4 q) { return Q4{-q.x, -q.y, -q.z, q.w}; }
// quat.rs:236-258
RHD Q4 qmul(Q4 a, Q4 b) {
    Q4 r;
    r.x = a.x * b.w + a.y * b.z - a.z * b.y + a.w * b.x;
    r.y = -a.x * b.z + a.y * b.w + a.z * b.x + a.w * b.y;
    r.z = a.x * b.y - a.y * b.x + a.z * b.w + a.w * b.z;
    r.w = -a.x * b.x - a.y * b.y - a.z * b.z + a.w * b.w;
    return r;
}
// trs.rs:211-221
RHD TrsF trs_mul(const TrsF& a, const TrsF& b) {
    TrsF r;
    r.t = a.t + rotate(a.s * b.t, a.q);
    r.q = qmul(a.q, b.q);
    r.s = rotate(a.s * rotate(b.s, b.q), conj(b.q));
    return r;
}

// Rust f32::clamp
RHD float clampf(float x, float lo, float hi) { return x < lo ? lo : (x > hi ? hi : x); }

// ---- Color (rayca-math/src/color/mod.rs:239-454): Add folds rhs alpha into rhs rgb and keeps
// the lhs alpha; Mul<Color> multiplies alpha; Mul<f32>/Div<f32> keep alpha.
RHD Col col(float r, float g, float b, float a) { return Col{r, g, b, a}; }
RHD Col cadd(Col l, Col r) { return Col{l.r + r.r * r.a, l.g + r.g * r.a, l.b + r.b * r.a, l.a}; }
RHD Col csub(Col l, Col r) { return Col{l.r - r.r * r.a, l.g - r.g * r.a, l.b - r.b * r.a, l.a}; }
RHD Col cmul(Col l, Col r) { return Col{l.r * r.r, l.g * r.g, l.b * r.b, l.a * r.a}; }
RHD Col cscale(Col c, float s) { return Col{c.r * s, c.g * s, c.b * s, c.a}; }
RHD Col cdiv(Col c, float s) { return Col{c.r / s, c.g / s, c.b / s, c.a}; }
RHD V3 col_to_v3(Col c) { return V3{c.r * c.a, c.g * c.a, c.b * c.a}; } // vec3.rs:407-411
RHD Col v3_to_col(V3 v) { return Col{v.x, v.y, v.z, 1.0f}; }

} // namespace rb
