// oracle_math.h — CPU restatement of rayca-math for the parity ORACLE.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is linked, imported or executed
// by the product path (rayca_b200/, librayca_cuda).  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may use it, as the checker.
//
// Parity status: the reference (Rust nightly, std::simd) cannot be built in this
// environment and its tests hold no pixel-level golden data, so image-level parity is
// UNPINNED; the unit-level known answers the reference's own tests assert are pinned in
// tests/test_oracle_kat.py.
//
// Every function cites the reference file:line it restates (paths relative to the
// reference checkout).  Lane semantics follow std::simd f32x4:
//   * Vec3 has w = 0, Point3 has w = 1 (vec3.rs:68-81, point3.rs:25-29)
//   * dot = lane-wise multiply then ORDERED reduce_sum ((p0+p1)+p2)+p3 (vec3.rs:240-243)
//   * Rust never contracts a*b+c; this file must be compiled -ffp-contract=off
//   * the only fused op is Point3::scale = mul_add(self, scale, [0,0,0,1]) (point3.rs:59-63)
#pragma once
#include <cmath>
#include <cstdint>
#include <limits>

namespace orc {

static constexpr float F32_EPSILON = 1.1920929e-07f;          // f32::EPSILON
static constexpr float EPS = F32_EPSILON * 8192.0f;            // rayca-math/src/lib.rs:33
static constexpr float F32_MAX = std::numeric_limits<float>::max();
static constexpr float F32_MIN = -std::numeric_limits<float>::max(); // f32::MIN
static constexpr float PI = 3.14159265358979323846264338327950288f;
static constexpr float FRAC_1_PI = 0.318309886183790671537767526745028724f;
static constexpr float RAY_BIAS = 1e-4f;                       // ray.rs:57

// Rust f32::clamp: NaN stays NaN, -0.0 stays -0.0.
static inline float clampf(float x, float lo, float hi) {
    if (x < lo) return lo;
    if (x > hi) return hi;
    return x;
}
// simd_min / simd_max and f32::min/max return the non-NaN operand == fminf/fmaxf.
static inline float minf(float a, float b) { return std::fmin(a, b); }
static inline float maxf(float a, float b) { return std::fmax(a, b); }

struct Vec2 {
    float x = 0, y = 0;
};
// vec2.rs Add / Mul<f32>
static inline Vec2 operator+(Vec2 a, Vec2 b) { return {a.x + b.x, a.y + b.y}; }
static inline Vec2 operator*(Vec2 a, float s) { return {a.x * s, a.y * s}; }

// Vec3: xyz lanes; the w lane is identically 0 and is only materialised where its
// contribution (+0.0f) takes part in an ordered sum.
struct Vec3 {
    float x = 0, y = 0, z = 0;
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
struct Point3 {
    float x = 0, y = 0, z = 0; // w == 1
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
struct Quat {
    float x = 0, y = 0, z = 0, w = 1;
};

static inline Vec3 operator+(Vec3 a, Vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline Vec3 operator-(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline Vec3 operator-(Vec3 a) { return {-a.x, -a.y, -a.z}; }
static inline Vec3 operator*(Vec3 a, Vec3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
static inline Vec3 operator*(Vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
static inline Vec3 operator*(float s, Vec3 a) { return {a.x * s, a.y * s, a.z * s}; }
static inline Vec3 operator/(Vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
// Vec3 - f32 / + f32 (vec3.rs:281-288,332-336) — splat over xyz (w lane discarded by users)
static inline Vec3 sub_scalar(Vec3 a, float s) { return {a.x - s, a.y - s, a.z - s}; }

// vec3.rs:240-243 — (self.simd * rhs.simd).reduce_sum(), 4 lanes, w product = 0*0
static inline float dot(Vec3 a, Vec3 b) {
    float p0 = a.x * b.x, p1 = a.y * b.y, p2 = a.z * b.z, p3 = 0.0f * 0.0f;
    return ((p0 + p1) + p2) + p3;
}
// vec3.rs:134-142
static inline Vec3 cross(Vec3 a, Vec3 b) {
    // tmp0 = a.yzx, tmp1 = b.zxy, tmp2 = tmp0*b, tmp3 = tmp0*tmp1, tmp4 = tmp2.yzx
    float t2x = a.y * b.x, t2y = a.z * b.y, t2z = a.x * b.z;
    float t3x = a.y * b.z, t3y = a.z * b.x, t3z = a.x * b.y;
    return {t3x - t2y, t3y - t2z, t3z - t2x};
}
static inline float norm(Vec3 a) { return dot(a, a); }          // vec3.rs:171-173
static inline float len(Vec3 a) { return std::sqrt(norm(a)); }  // vec3.rs:175-177
// vec3.rs:183-188
static inline Vec3 normalized(Vec3 a) {
    float l = len(a);
    if (l > EPS) return {a.x / l, a.y / l, a.z / l};
    return a;
}
// vec3.rs:195-216 — lanes with dir == 0 give 0 (num 0 / den 1)
static inline Vec3 reciprocal(Vec3 a) {
    Vec3 r;
    r.x = (a.x == 0.0f) ? (0.0f / (a.x + 1.0f)) : (1.0f / a.x);
    r.y = (a.y == 0.0f) ? (0.0f / (a.y + 1.0f)) : (1.0f / a.y);
    r.z = (a.z == 0.0f) ? (0.0f / (a.z + 1.0f)) : (1.0f / a.z);
    return r;
}
// vec3.rs:219-221 — self - 2.0 * self.dot(normal) * normal
static inline Vec3 reflect(Vec3 v, Vec3 n) {
    float k = 2.0f * dot(v, n);
    return v - (n * k);
}
// vec3.rs:108-111 — derived PartialOrd on f32x4 => LEXICOGRAPHIC `<` over 4 lanes
static inline bool vec3_close(Vec3 a, Vec3 b) {
    float d[4] = {std::fabs(a.x - b.x), std::fabs(a.y - b.y), std::fabs(a.z - b.z), std::fabs(0.0f - 0.0f)};
    float e[4] = {EPS, EPS, EPS, 0.0f};
    for (int i = 0; i < 4; ++i) {
        if (d[i] < e[i]) return true;
        if (d[i] > e[i]) return false;
        if (!(d[i] == e[i])) return false; // NaN: partial_cmp is None => `<` false
    }
    return false;
}

static inline Vec3 to_vec3(Point3 p) { return {p.x, p.y, p.z}; }   // vec3.rs:394-399
static inline Point3 to_point3(Vec3 v) { return {v.x, v.y, v.z}; } // point3.rs:117-122
static inline Point3 operator+(Point3 p, Vec3 v) { return {p.x + v.x, p.y + v.y, p.z + v.z}; }
static inline Point3 operator-(Point3 p, Vec3 v) { return {p.x - v.x, p.y - v.y, p.z - v.z}; }
static inline Vec3 operator-(Point3 a, Point3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }

// vec3.rs:148-159 / point3.rs:65-76
static inline Vec3 rotate(Vec3 v, Quat q) {
    Vec3 u{q.x * 1.0f, q.y * 1.0f, q.z * 1.0f};
    float s = q.w;
    Vec3 a = (2.0f * dot(u, v)) * u;
    Vec3 b = (s * s - dot(u, u)) * v;
    Vec3 c = (2.0f * s) * cross(u, v);
    return (a + b) + c;
}
// point3.rs:59-63 — fused multiply-add with +0 on xyz
static inline Point3 scale_point(Point3 p, Vec3 s) {
    return {std::fmaf(p.x, s.x, 0.0f), std::fmaf(p.y, s.y, 0.0f), std::fmaf(p.z, s.z, 0.0f)};
}

// quat.rs:236-258
static inline Quat operator*(Quat a, Quat b) {
    Quat r;
    r.x = a.x * b.w + a.y * b.z - a.z * b.y + a.w * b.x;
    r.y = -a.x * b.z + a.y * b.w + a.z * b.x + a.w * b.y;
    r.z = a.x * b.y - a.y * b.x + a.z * b.w + a.w * b.z;
    r.w = -a.x * b.x - a.y * b.y - a.z * b.z + a.w * b.w;
    return r;
}
// quat.rs:95-103 (the is_normalized assert is reported by the caller)
static inline Quat conjugate(Quat q) { return {q.x * -1.0f, q.y * -1.0f, q.z * -1.0f, q.w * 1.0f}; }
static inline float quat_len(Quat q) {
    float p0 = q.x * q.x, p1 = q.y * q.y, p2 = q.z * q.z, p3 = q.w * q.w;
    return std::sqrt(((p0 + p1) + p2) + p3);
}

struct Trs {
    Vec3 translation{0, 0, 0};
    Quat rotation{0, 0, 0, 1};
    Vec3 scale{1, 1, 1};
};
// trs.rs:211-221
static inline Trs operator*(const Trs& a, const Trs& b) {
    Trs r;
    r.translation = a.translation + rotate(a.scale * b.translation, a.rotation);
    r.rotation = a.rotation * b.rotation;
    r.scale = rotate(a.scale * rotate(b.scale, b.rotation), conjugate(b.rotation));
    return r;
}
// trs.rs:264-273
static inline Point3 transform_point(const Trs& t, Point3 p) {
    p = scale_point(p, t.scale);
    p = to_point3(rotate(to_vec3(p), t.rotation));
    return p + t.translation;
}
// trs.rs:253-262
static inline Vec3 transform_vec(const Trs& t, Vec3 v) {
    v = v * t.scale;
    v = rotate(v, t.rotation);
    return v + t.translation;
}
// trs.rs:126-128
static inline Vec3 trs_get_translation(const Trs& t) { return rotate(t.translation, t.rotation); }

struct Mat3 {
    float m[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
};
static inline Mat3 mat3_identity() {
    Mat3 r;
    r.m[0][0] = r.m[1][1] = r.m[2][2] = 1.0f;
    return r;
}
// mat3.rs:190-216
static inline Mat3 mat3_from_quat(Quat q) {
    float xq[4] = {q.x * q.x, q.x * q.y, q.x * q.z, q.x * q.w};
    float yq[4] = {q.y * q.x, q.y * q.y, q.y * q.z, q.y * q.w};
    float zq[4] = {q.z * q.x, q.z * q.y, q.z * q.z, q.z * q.w};
    Mat3 r;
    r.m[0][0] = 1.0f - 2.0f * (yq[1] + zq[2]);
    r.m[0][1] = 2.0f * (xq[1] - zq[3]);
    r.m[0][2] = 2.0f * (xq[2] + yq[3]);
    r.m[1][0] = 2.0f * (xq[1] + zq[3]);
    r.m[1][1] = 1.0f - 2.0f * (xq[0] + zq[2]);
    r.m[1][2] = 2.0f * (yq[2] - xq[3]);
    r.m[2][0] = 2.0f * (xq[2] - yq[3]);
    r.m[2][1] = 2.0f * (yq[2] + xq[3]);
    r.m[2][2] = 1.0f - 2.0f * (xq[0] + yq[1]);
    return r;
}
// mat3.rs:160-180
static inline Mat3 operator*(const Mat3& a, const Mat3& b) {
    Mat3 r;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            float e = a.m[i][0] * b.m[0][j];
            float f = a.m[i][1] * b.m[1][j];
            float g = a.m[i][2] * b.m[2][j];
            r.m[i][j] = e + f + g;
        }
    return r;
}
// mat3.rs:67-71
static inline Mat3 mat3_from_scale(Vec3 s) {
    Mat3 r = mat3_identity();
    r.m[0][0] *= s.x;
    r.m[1][1] *= s.y;
    r.m[2][2] *= s.z;
    return r;
}
// mat3.rs:37-41,73-75
static inline Mat3 mat3_from_rotation(Quat q) { return mat3_from_quat(q) * mat3_identity(); }
// mat3.rs:126-132 — From<&Trs>
static inline Mat3 mat3_from_trs(const Trs& t) { return mat3_from_quat(t.rotation) * mat3_from_scale(t.scale); }
// mat3.rs:140-144 — From<&Inversed<&Trs>>; scale = reciprocal (trs.rs:329-331), rotation = inverse
static inline Mat3 mat3_from_inversed_trs(const Trs& t) {
    return mat3_from_scale(reciprocal(t.scale)) * mat3_from_rotation(conjugate(t.rotation));
}
static inline Mat3 transpose(const Mat3& a) {
    Mat3 r;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) r.m[i][j] = a.m[j][i];
    return r;
}
// mat3.rs:218-234
static inline Vec3 operator*(const Mat3& a, Vec3 v) {
    float ret[3] = {0.0f, 0.0f, 0.0f};
    float vv[3] = {v.x, v.y, v.z};
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) ret[i] += a.m[i][j] * vv[j];
    return {ret[0], ret[1], ret[2]};
}
// mat3.rs:57-65
static inline Mat3 mat3_tbn(Vec3 t, Vec3 b, Vec3 n) {
    Mat3 r;
    r.m[0][0] = t.x; r.m[0][1] = b.x; r.m[0][2] = n.x;
    r.m[1][0] = t.y; r.m[1][1] = b.y; r.m[1][2] = n.y;
    r.m[2][0] = t.z; r.m[2][1] = b.z; r.m[2][2] = n.z;
    return r;
}

// color/mod.rs:59-66 and the operator zoo at :239-454
struct Color {
    float r = 1, g = 1, b = 1, a = 1;
};
static constexpr Color BLACK{0, 0, 0, 1};
static constexpr Color WHITE{1, 1, 1, 1};
// Add: lhs.rgb + rhs.rgb * rhs.a, keeps lhs alpha (color/mod.rs:239-286)
static inline Color operator+(Color l, Color r) { return {l.r + r.r * r.a, l.g + r.g * r.a, l.b + r.b * r.a, l.a}; }
// Sub (color/mod.rs:288-298)
static inline Color operator-(Color l, Color r) { return {l.r - r.r * r.a, l.g - r.g * r.a, l.b - r.b * r.a, l.a}; }
// Mul<f32> keeps alpha (color/mod.rs:300-360)
static inline Color operator*(Color c, float s) { return {c.r * s, c.g * s, c.b * s, c.a}; }
static inline Color operator*(float s, Color c) { return {c.r * s, c.g * s, c.b * s, c.a}; }
// Mul<Color> multiplies alpha too (color/mod.rs:362-408)
static inline Color operator*(Color l, Color r) { return {l.r * r.r, l.g * r.g, l.b * r.b, l.a * r.a}; }
// Div<f32> (color/mod.rs:410-424)
static inline Color operator/(Color c, float s) { return {c.r / s, c.g / s, c.b / s, c.a}; }
// vec3.rs:407-411 — premultiplies alpha
static inline Vec3 color_to_vec3(Color c) { return {c.r * c.a, c.g * c.a, c.b * c.a}; }
// color/mod.rs:199-203
static inline Color vec3_to_color(Vec3 v) { return {v.x, v.y, v.z, 1.0f}; }
// color/mod.rs:156-158
static inline bool is_transparent(Color c) { return c.a < 1.0f - F32_EPSILON; }
// color/mod.rs:160-169
static inline bool color_close(Color a, Color b) {
    return std::fabs(a.r - b.r) < F32_EPSILON && std::fabs(a.g - b.g) < F32_EPSILON &&
           std::fabs(a.b - b.b) < F32_EPSILON && std::fabs(a.a - b.a) < F32_EPSILON;
}
// color/rgba8.rs:75-84 — `(c * 255.0).clamp(0.0, 255.0) as u8`: truncation, NaN -> 0
static inline uint8_t quantize(float c) {
    float v = clampf(c * 255.0f, 0.0f, 255.0f);
    if (v != v) return 0;
    return (uint8_t)v;
}

// ray.rs:41-72
struct Ray {
    Point3 origin;
    Vec3 dir{0, 0, -1};
    Vec3 rdir;
    Color throughput = WHITE;
};
static inline Ray make_ray(Point3 origin, Vec3 dir) {
    Ray r;
    r.origin = origin;
    r.dir = dir;
    r.rdir = reciprocal(dir);
    return r;
}
// trs.rs:275-284 with ray.rs:74-91
static inline Ray transform_ray(const Trs& t, Ray r) {
    r.dir = r.dir * t.scale;
    r.rdir = reciprocal(r.dir);
    r.origin = scale_point(r.origin, t.scale);
    r.dir = rotate(r.dir, t.rotation);
    r.rdir = reciprocal(r.dir);
    r.origin = to_point3(rotate(to_vec3(r.origin), t.rotation));
    r.origin = r.origin + t.translation;
    return r;
}

// ---- Mat4 (rayca-math/src/mat4.rs), only what the sphere path needs
struct Mat4 {
    float m[4][4];
};
static inline Mat4 mat4_identity() {
    Mat4 r{};
    for (int i = 0; i < 4; ++i) r.m[i][i] = 1.0f;
    return r;
}
// mat4.rs:160-189 — row * column over 16 lanes of which 12 are masked to zero; ordered reduce_sum
static inline Mat4 operator*(const Mat4& a, const Mat4& b) {
    Mat4 r;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            float acc = ((a.m[i][0] * b.m[0][j] + a.m[i][1] * b.m[1][j]) + a.m[i][2] * b.m[2][j]) + a.m[i][3] * b.m[3][j];
            for (int z = 0; z < 12; ++z) acc += 0.0f; // the masked lanes
            r.m[i][j] = acc;
        }
    return r;
}
// mat4.rs:207-235
static inline Mat4 mat4_from_quat(Quat q) {
    Mat3 r3 = mat3_from_quat(q);
    Mat4 r = mat4_identity();
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) r.m[i][j] = r3.m[i][j];
    return r;
}
// mat4.rs:63-79,104-118: from_translation / from_rotation (= Mat4::from(q) * identity) / from_scale
static inline Mat4 mat4_from_translation(Vec3 t) {
    Mat4 r = mat4_identity();
    r.m[0][3] += t.x;
    r.m[1][3] += t.y;
    r.m[2][3] += t.z;
    return r;
}
static inline Mat4 mat4_from_rotation(Quat q) { return mat4_from_quat(q) * mat4_identity(); }
static inline Mat4 mat4_from_scale(Vec3 s) {
    Mat4 r = mat4_identity();
    r.m[0][0] *= s.x;
    r.m[1][1] *= s.y;
    r.m[2][2] *= s.z;
    return r;
}
// mat4.rs:276-283 — From<&Inversed<Trs>>: S^-1 * (R^-1 * T^-1)
static inline Mat4 mat4_from_inversed_trs(const Trs& t) {
    return mat4_from_scale(reciprocal(t.scale)) * (mat4_from_rotation(conjugate(t.rotation)) * mat4_from_translation(t.translation * -1.0f));
}
// mat4.rs:290-313 (impl_mul3!): rows dot [x, y, z, 1], divide by the 4th unless it is zero
static inline Point3 operator*(const Mat4& a, Point3 p) {
    float ret[4];
    for (int i = 0; i < 4; ++i) ret[i] = ((a.m[i][0] * p.x + a.m[i][1] * p.y) + a.m[i][2] * p.z) + a.m[i][3] * 1.0f;
    float den = ret[3] != 0.0f ? ret[3] : 1.0f;
    return {ret[0] / den, ret[1] / den, ret[2] / den};
}
// trs.rs:405-414 with ray.rs:74-91: translate, rotate, scale by the inverse parts
static inline Ray inverse_transform_ray(const Trs& t, Ray r) {
    r.origin = r.origin + (t.translation * -1.0f);
    Quat qi = conjugate(t.rotation);
    r.dir = rotate(r.dir, qi);
    r.rdir = reciprocal(r.dir);
    r.origin = to_point3(rotate(to_vec3(r.origin), qi));
    Vec3 si = reciprocal(t.scale);
    r.dir = r.dir * si;
    r.rdir = reciprocal(r.dir);
    r.origin = scale_point(r.origin, si);
    return r;
}

} // namespace orc
