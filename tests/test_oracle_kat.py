"""Known-answer tests that PIN the oracle: every case restates an assertion of the reference's own
unit tests (file:line cited) and checks the oracle's restatement of the same function against it."""
import ctypes as C

import numpy as np

from rayca_b200 import model as M

EPS = np.float32(np.finfo(np.float32).eps) * 8192  # rayca-math/src/lib.rs:33


def f3(*v):
    return (C.c_float * 3)(*v)


def f4(*v):
    return (C.c_float * 4)(*v)


def trs(t=(0, 0, 0), q=(0, 0, 0, 1), s=(1, 1, 1)):
    return M.rayca_trs(f3(*t), f4(*q), f3(*s))


def rotate(lib, v, q):
    out = f3()
    lib.rayca_oracle_kat_rotate(f3(*v), f4(*q), out)
    return np.array(out[:], dtype=np.float32)


def test_vec3_rotate(oracle):
    """rayca-math/src/vec3.rs:526-553"""
    lib = oracle.lib()
    cases = [((1, 0, 0), (0, 1, 0, 0), (-1, 0, 0)), ((1, 0, 0), (0, 0.707, 0, 0.707), (0, 0, -1)),
             ((1, 0, 0), (0, 0, 1, 0), (-1, 0, 0)), ((1, 0, 0), (0, 0, 0.707, 0.707), (0, 1, 0)),
             ((0, 0, 1), (-0.383, 0, 0, 0.924), (0, 0.707, 0.707))]
    for v, q, want in cases:
        assert np.all(np.abs(rotate(lib, v, q) - np.array(want, dtype=np.float32)) < EPS), (v, q)


def test_ray_rotate(oracle):
    """rayca-math/src/ray.rs:158-165: default ray dir (0,0,-1) rotated by x:-45deg"""
    got = rotate(oracle.lib(), (0, 0, -1), (-0.383, 0, 0, 0.924))
    assert np.all(np.abs(got - np.array([0, -0.707, -0.707], dtype=np.float32)) < EPS)


def test_reciprocal(oracle):
    """rayca-math/src/vec3.rs:574-581: zero lanes give 0, not inf"""
    out = f3()
    oracle.lib().rayca_oracle_kat_reciprocal(f3(2.0, 4.0, 0.0), out)
    assert abs(out[0] - 0.5) < 1e-6 and abs(out[1] - 0.25) < 1e-6 and out[2] == 0.0
    oracle.lib().rayca_oracle_kat_reciprocal(f3(-0.0, 0.0, -8.0), out)
    assert out[0] == 0.0 and out[1] == 0.0 and out[2] == -0.125


def test_reflect(oracle):
    """rayca-math/src/vec3.rs:583-589"""
    out = f3()
    oracle.lib().rayca_oracle_kat_reflect(f3(1, -1, 0), f3(0, 1, 0), out)
    assert out[:] == [1.0, 1.0, 0.0]


def test_quantize(oracle):
    """rayca-math/src/color/rgba8.rs:110-132: 1.0->255, 0.0->0, 0.5->127, 2.0->255, -1.0->0, 1.5->255; NaN->0"""
    q = oracle.lib().rayca_oracle_kat_quantize
    assert [q(1.0), q(0.0), q(0.5), q(2.0), q(-1.0), q(1.5)] == [255, 0, 127, 255, 0, 255]
    assert q(float("nan")) == 0


def test_triangle_front_and_back_face(oracle):
    """rayca-geometry/src/triangle.rs:570-576 and rayca-soft/src/bvh/triangle.rs:110-113"""
    lib = oracle.lib()
    verts = (C.c_float * 9)(-1, 0, 0, 1, 0, 0, 0, 1, 0)
    out = (C.c_float * 6)()
    ident = trs()
    assert lib.rayca_oracle_kat_triangle(verts, C.byref(ident), f3(0, 0, 1), f3(0, 0, -1), out) == 1
    assert out[0] == 1.0  # depth
    # barycentrics: u weights vertex 0, v weights vertex 1 (triangle.rs:136-153); hit at (0,0,0)
    assert abs(out[1] - 0.5) < 1e-6 and abs(out[2] - 0.5) < 1e-6
    assert out[3:6] == [0.0, 0.0, 0.0]
    assert lib.rayca_oracle_kat_triangle(verts, C.byref(ident), f3(0, 0, 1), f3(0, 0, 1), out) == 0
    # single sided: the same ray from behind is culled (triangle.rs:96)
    assert lib.rayca_oracle_kat_triangle(verts, C.byref(ident), f3(0, 0, -1), f3(0, 0, 1), out) == 0


def test_trs_compose(oracle):
    """rayca-math/src/trs.rs:461-500"""
    lib = oracle.lib()
    out = M.rayca_trs()
    a, b = trs(), trs()
    lib.rayca_oracle_kat_trs_mul(C.byref(a), C.byref(b), C.byref(out))
    assert out.translation[:] == [0, 0, 0] and out.rotation[:] == [0, 0, 0, 1] and out.scale[:] == [1, 1, 1]
    a, b = trs(t=(1, 2, 3)), trs(t=(4, 5, 6))
    lib.rayca_oracle_kat_trs_mul(C.byref(a), C.byref(b), C.byref(out))
    assert out.translation[:] == [5, 7, 9]
    a, b = trs(s=(2, 2, 2)), trs(s=(0.5, 0.5, 0.5))
    lib.rayca_oracle_kat_trs_mul(C.byref(a), C.byref(b), C.byref(out))
    assert out.scale[:] == [1, 1, 1]
    # two quarter turns about Y compose to a half turn (trs.rs:491-500)
    h = np.float32(np.sqrt(0.5))
    a = b = trs(q=(0, h, 0, h))
    lib.rayca_oracle_kat_trs_mul(C.byref(a), C.byref(b), C.byref(out))
    assert np.all(np.abs(np.array(out.rotation[:]) - np.array([0, 1, 0, 0])) < 1e-5)


def test_trs_point_order(oracle):
    """trs.rs:264-273: scale, then rotate, then translate"""
    out = f3()
    h = float(np.float32(np.sqrt(0.5)))
    t = trs(t=(1, 2, 3), q=(0, 0, h, h), s=(2, 2, 2))
    oracle.lib().rayca_oracle_kat_trs_point(C.byref(t), f3(1, 0, 0), out)
    assert np.all(np.abs(np.array(out[:]) - np.array([1, 4, 3])) < 1e-5)


def test_ray_transform_keeps_direction_unnormalised(oracle):
    """ray.rs:74-91,206-213 and scene.rs:141: scaling the camera scales the direction"""
    oo, od = f3(), f3()
    t = trs(t=(0, 0, 4), s=(2, 2, 2))
    oracle.lib().rayca_oracle_kat_trs_ray(C.byref(t), f3(0, 0, 0), f3(0, 0, -1), oo, od)
    assert oo[:] == [0, 0, 4] and od[:] == [0, 0, -2]


def test_sampler_wrap(oracle):
    """rayca-model/src/sampler.rs:36-50: uv (0,0) and (1,1) address a valid texel of a 1x1 image"""
    out = (C.c_uint32 * 2)()
    for uv in ((0.0, 0.0), (1.0, 1.0)):
        oracle.lib().rayca_oracle_kat_sampler_index((C.c_float * 2)(*uv), 1, 1, out)
        assert out[:] == [0, 0]
    oracle.lib().rayca_oracle_kat_sampler_index((C.c_float * 2)(0.5, 1.25), 4, 8, out)
    assert out[:] == [2, 2]  # ((0.5+1)*4) % 4 = 2 ; ((0.25+1)*8) % 8 = 2


def test_slab_zero_direction_lane_misses(oracle):
    """SURVEY A-3 (aabb.rs:74-93 with vec3.rs:195-216): rdir = 0 on a lane makes every box miss"""
    lib = oracle.lib()
    lib.rayca_oracle_kat_aabb.argtypes = [C.c_float * 3] * 4
    fmax = np.finfo(np.float32).max
    assert lib.rayca_oracle_kat_aabb(f3(-1, -1, -1), f3(1, 1, 1), f3(0, 0, 5), f3(0, 0, -1)) == fmax
    got = lib.rayca_oracle_kat_aabb(f3(-1, -1, -1), f3(1, 1, 1), f3(0.1, 0.2, 5), f3(0.01, 0.01, -1))
    assert abs(got - 4.0) < 1e-5
    # origin inside: accepted, returns the (negative) entry distance (aabb.rs:88-92)
    got = lib.rayca_oracle_kat_aabb(f3(-1, -1, -1), f3(1, 1, 1), f3(0, 0, 0), f3(0.5, 0.5, -1))
    assert got < 0.0


def _single_blas(oracle, positions):
    s = M.Scene()
    s.push_model(M.triangle_soup_model(positions))
    s.push_model(M.create_default_model())
    osc = oracle.OracleScene(s.to_desc())
    return osc.bvh()


def test_bvh_one_triangle_is_a_leaf(oracle):
    """rayca-soft/src/bvh/blas.rs:403-435: a single triangle leaves the root a leaf"""
    info, nodes, no, po, src = _single_blas(oracle, [[[-1, 0, 0], [1, 0, 0], [0, 1, 0]]])
    assert info.blas_count == 1 and info.node_count == 2 and info.triangle_count == 1
    assert nodes[0, 7].view(np.uint32) == 1  # root holds the primitive


def test_bvh_two_separated_triangles_split(oracle):
    """rayca-soft/src/bvh/blas.rs:437-490: two triangles far apart split the root into two leaves"""
    tris = [[[-1, 0, 0], [1, 0, 0], [0, 1, 0]], [[9, 0, 0], [11, 0, 0], [10, 1, 0]]]
    info, nodes, no, po, src = _single_blas(oracle, tris)
    assert info.node_count == 4
    assert nodes[0, 7].view(np.uint32) == 0 and nodes[0, 3].view(np.uint32) == 2  # inner, children 2 and 3
    assert nodes[2, 7].view(np.uint32) == 1 and nodes[3, 7].view(np.uint32) == 1


def test_keyed_rng_is_deterministic_and_uniform(oracle):
    lib = oracle.lib()
    a = [lib.rayca_oracle_kat_rng(0x5EED, p, 0, 0, 0) for p in range(2000)]
    b = [lib.rayca_oracle_kat_rng(0x5EED, p, 0, 0, 0) for p in range(2000)]
    assert a == b and all(0.0 <= x < 1.0 for x in a)
    assert abs(np.mean(a) - 0.5) < 0.03
    assert lib.rayca_oracle_kat_rng(0x5EED, 1, 0, 0, 0) != lib.rayca_oracle_kat_rng(0x5EED, 1, 0, 0, 1)


def f9(*v):
    return (C.c_float * 9)(*v)


def test_sphere_intersects(oracle):
    """rayca-geometry/src/sphere.rs:186-203 `intersect`: unit sphere at the origin is hit from inside along +x, missed
    from (2,0,0) along +x; a sphere at (4,0,0) is missed by a ray along -z."""
    lib = oracle.lib()
    out = f4()
    ident = trs()
    assert lib.rayca_oracle_kat_sphere(f3(0, 0, 0), C.c_float(1.0), C.byref(ident), f3(0, 0, 0), f3(1, 0, 0), out) == 1
    assert abs(out[0] - 1.0) < 1e-6 and abs(out[1] - 1.0) < 1e-6  # leaves the sphere at x = 1
    assert lib.rayca_oracle_kat_sphere(f3(0, 0, 0), C.c_float(1.0), C.byref(ident), f3(2, 0, 0), f3(1, 0, 0), out) == 0
    assert lib.rayca_oracle_kat_sphere(f3(4, 0, 0), C.c_float(1.0), C.byref(ident), f3(0, 0, 0), f3(0, 0, -1), out) == 0


def test_sphere_min_max_centroid(oracle):
    """rayca-geometry/src/sphere.rs:205-215 `min_max_and_centroid` (exact equality in the reference)."""
    out = f9()
    ident = trs()
    oracle.lib().rayca_oracle_kat_sphere_bounds(f3(1, 2, 3), C.c_float(2.0), C.byref(ident), out)
    assert out[:] == [-1.0, 0.0, 1.0, 3.0, 4.0, 5.0, 1.0, 2.0, 3.0]


def test_sphere_normal(oracle):
    """rayca-geometry/src/sphere.rs:217-223 `get_normal`."""
    out = f3()
    oracle.lib().rayca_oracle_kat_sphere_normal(f3(0, 0, 0), f3(0, 1, 0), out)
    assert np.linalg.norm(np.array(out[:]) - np.array([0.0, 1.0, 0.0])) < 1e-6


def test_triangle_centroid_min_max(oracle):
    """rayca-geometry/src/triangle.rs:578-593 `centroid_and_min_max`: Triangle::default() = (-1,0,0), (1,0,0), (0,1,0);
    centroid within 1e-3 of (a+b+c)*0.3333, min / max exact."""
    out = f9()
    ident = trs()
    oracle.lib().rayca_oracle_kat_triangle_bounds(f9(-1, 0, 0, 1, 0, 0, 0, 1, 0), C.byref(ident), out)
    assert out[0:3] == [-1.0, 0.0, 0.0] and out[3:6] == [1.0, 1.0, 0.0]
    want = np.array([0.0, 1.0, 0.0], dtype=np.float32) * np.float32(0.3333)
    assert np.linalg.norm(np.array(out[6:9], dtype=np.float32) - want) < 1e-3


def _roundtrip(oracle, t, v):
    fwd, back = f3(), f3()
    oracle.lib().rayca_oracle_kat_trs_vec(C.byref(t), f3(*v), fwd)
    oracle.lib().rayca_oracle_kat_trs_inverse_vec(C.byref(t), fwd, back)
    return np.array(back[:], dtype=np.float32)


def test_trs_inversion(oracle):
    """rayca-math/src/trs.rs:505-553 `inverse_trs_{translation,scale,rotation}_only`, `inverse_trs_full`:
    &inv * (&trs * v) is close to v (Vec3::close: every lane within EPS)."""
    s2 = float(np.sqrt(np.float32(0.5)))
    rot_z90 = (0.0, 0.0, s2, s2)  # Quat::axis_angle(Z, FRAC_PI_2)
    cases = [(trs(t=(1, 2, 3)), (4, 5, 6)), (trs(s=(2, 3, 4)), (8, 9, 12)), (trs(q=rot_z90), (1, 0, 0)),
             (trs(t=(1, 2, 3), q=rot_z90, s=(2, 2, 2)), (1, 1, 1))]
    for t, v in cases:
        got = _roundtrip(oracle, t, v)
        assert np.all(np.abs(got - np.array(v, dtype=np.float32)) < EPS), (v, got)


def test_byte_over_255_identity():
    """shade.cuh::byte_to_unit: q = t * r, q' = fma(fma(-q, 255, t), r, q) with r = RN(1/255) equals the correctly
    rounded t / 255 for every byte t (exact rational arithmetic; the CUDA path relies on it for bit parity of texels)."""
    from fractions import Fraction

    def rn(fr):
        x = np.float32(float(fr))
        best = None
        for c in (np.nextafter(x, np.float32(-np.inf)), x, np.nextafter(x, np.float32(np.inf))):
            d = abs(Fraction(float(c)) - fr)
            even = (int(np.float32(c).view(np.uint32)) & 1) == 0
            if best is None or d < best[0] or (d == best[0] and even):
                best = (d, np.float32(c))
        return best[1]

    r = np.float32(float.fromhex("0x1.010102p-8"))
    assert r == rn(Fraction(1, 255))
    for t in range(256):
        q = rn(Fraction(t) * Fraction(float(r)))
        rem = rn(Fraction(t) - Fraction(float(q)) * 255)
        q2 = rn(Fraction(float(rem)) * Fraction(float(r)) + Fraction(float(q)))
        assert q2 == rn(Fraction(t, 255)) == np.float32(t) / np.float32(255.0), t
