"""Edge cases of the hot path, CUDA against the oracle run live on the same input (through the C ABI on both sides):
empty and degenerate geometry, exact ties, rays the reference's slab arithmetic drops, image sizes around the
kernel's 16x8 groups and 8x4 warp tiles, coordinates far from the origin.  Bit-exact on hit ids / depth /
barycentrics, RGB within 1e-3 (tests/test_cuda_parity.py::check_against)."""
import numpy as np
import pytest

from rayca_b200 import model as M
from rayca_b200 import renderer as R

from conftest import load_desc
from test_cuda_parity import check_against

pytestmark = pytest.mark.gpu
f32 = np.float32


@pytest.fixture(scope="module")
def ctx():
    c = R.Context(0)
    yield c
    c.close()


def camera_and_lights(z=4.0, yfov=np.pi / 4, lights=((-1.0, 4.0, 3.0), (1.0, 4.0, 3.0)), intensity=1024.0):
    """The default model of rayca-soft/src/scene.rs:18-55: camera on +z looking down -z, two point lights."""
    dm = M.Model()
    dm.cameras.append(M.Camera(float(f32(yfov))))
    dm.nodes.append(M.Node(camera=0, trs=M.Trs(translation=np.array([0.0, 0.0, z], dtype=f32))))
    dm.root.children.append(0)
    dm.lights.append(M.Light(kind=M.LIGHT_POINT, intensity=intensity))
    for k, p in enumerate(lights):
        dm.nodes.append(M.Node(light=0, trs=M.Trs(translation=np.array(p, dtype=f32))))
        dm.root.children.append(1 + k)
    return dm


def soup_scene(tris, **cam):
    s = M.Scene()
    tris = np.asarray(tris, dtype=f32).reshape(-1, 3, 3)
    if len(tris):
        s.push_model(M.triangle_soup_model(tris, M.PbrMaterial((0.8, 0.7, 0.6, 1), M.NONE, M.NONE, 0.0, 1.0, M.NONE)))
    s.push_model(camera_and_lights(**cam))
    return s


def compare(ctx, oracle, scene, w, h, cfg=None, bvh=True):
    cfg = cfg or R.Config(max_depth=1)
    cfg.bvh = bvh
    desc = scene.to_desc()
    osc = oracle.OracleScene(desc, bvh=bvh)
    o8, oaux, ost = osc.render(cfg.to_c(), w, h)
    dev = R.DeviceScene(ctx, desc, bvh=bvh)
    try:
        for traversal in (R.TRAVERSAL_ORDERED, R.TRAVERSAL_EXHAUSTIVE, R.TRAVERSAL_REFERENCE_ORDERED):
            rgba8, aux, st = dev.render(cfg, w, h, options=M.rayca_options(0x5EED, traversal, 0, 0, 0, 0, 0), want_aux=True)
            want = dict(hit_src=oaux.hit_src.reshape(h, w), hit_prim=oaux.hit_prim.reshape(h, w),
                        hit_blas=oaux.hit_blas.reshape(h, w), depth=oaux.depth.reshape(h, w), uv=oaux.uv.reshape(h, w, 2),
                        rgba_f32=oaux.rgba_f32.reshape(h, w, 4), rgba8=o8)
            check_against(want, rgba8, aux, h, w, False)
            assert [st.rays_primary, st.rays_shadow, st.rays_secondary] == [ost.rays_primary, ost.rays_shadow, ost.rays_secondary]
    finally:
        dev.close()
        osc.close()
    return oaux


def test_scene_without_geometry_is_refused(ctx, oracle):
    """Camera and lights only: the reference asserts a non-empty model range when it builds the TLAS
    (tlas.rs:75), i.e. `draw` panics; both sides report an invalid argument instead of rendering."""
    desc = soup_scene([]).to_desc()
    with pytest.raises(RuntimeError):
        oracle.OracleScene(desc, bvh=True)
    with pytest.raises(R.RaycaError) as e:
        R.DeviceScene(ctx, desc, bvh=True)
    assert e.value.code == -1


def test_every_ray_misses(ctx, oracle):
    """Geometry entirely behind the camera: every ray misses, the frame is opaque black
    (scene.rs:80-86 unwrap_or(BLACK))."""
    aux = compare(ctx, oracle, soup_scene([[[-1, -1, 9], [1, -1, 9], [0, 1, 9]]]), 37, 21)
    assert (aux.hit_src == 0xFFFFFFFF).all()


@pytest.mark.parametrize("w,h", [(1, 1), (2, 3), (7, 5), (8, 4), (9, 5), (16, 8), (17, 9), (31, 33), (129, 1), (1, 67)])
def test_image_sizes_around_the_tile_shape(ctx, oracle, w, h):
    """Sizes below, at and just above the 8x4 warp tile and the 16x8 block group, single rows and columns."""
    tris = [[[-1, -1, 0], [1, -1, 0], [0, 1, 0]], [[-2, -1, -1], [2, -1, -1], [0, 2, -1]]]
    compare(ctx, oracle, soup_scene(tris), w, h)


def test_rays_with_zero_direction_lanes_miss(ctx, oracle):
    """Odd image sizes put pixel centres on the camera's axes: the whole middle row has dir.y == 0 and the whole
    middle column dir.x == 0.  The reference defines the reciprocal of a zero lane as 0 (vec3.rs:195-216), so the slab
    test rejects every box, the root included (aabb.rs:88): those pixels stay black although a triangle covers them
    (SURVEY A-3).  Config::bvh(false) only caps the tree depth, the root box is still tested: same cross."""
    tris = [[[-1, -1, 0], [1, -1, 0], [0, 1, 0]]]
    w = h = 33
    for bvh in (True, False):
        hit = compare(ctx, oracle, soup_scene(tris), w, h, bvh=bvh).hit_src.reshape(h, w) != 0xFFFFFFFF
        assert not hit[16, :].any() and not hit[:, 16].any()
        assert hit[15, 15] and hit[17, 17] and hit[15, 17] and hit[17, 15]


def test_coincident_triangles_resolve_to_the_first(ctx, oracle):
    """Exact depth ties: the reference keeps the first hit in depth-first order (strict '<', blas.rs:141-172), i.e.
    the lowest primitive index after its own partitioning; 64 copies of one triangle plus a shifted fan."""
    t = np.array([[-1, -1, 0], [1, -1, 0], [0, 1, 0]], dtype=f32)
    tris = [t] * 64 + [t + np.array([0.25 * k, 0, -0.5], dtype=f32) for k in range(-3, 4)]
    compare(ctx, oracle, soup_scene(tris), 96, 64)


def test_degenerate_triangles_are_ignored(ctx, oracle):
    """Zero-area triangles (repeated vertices, collinear vertices, a point) have n = 0: |n.dir| < eps rejects them
    (triangle.rs:106); their boxes still enter the SAH builder.  NaN normals from get_normalized on a zero vector
    do not reach a shaded pixel."""
    good = [[[-1, -1, 0], [1, -1, 0], [0, 1, 0]], [[-1.5, 0, -1], [1.5, 0, -1], [0, 1.5, -1]]]
    bad = [[[0, 0, 1], [0, 0, 1], [0, 0, 1]], [[-1, 0, 0.5], [0, 0, 0.5], [1, 0, 0.5]], [[0.2, 0.2, 1], [0.2, 0.2, 1], [0.5, 0.5, 1]]]
    n = np.zeros((len(good) + len(bad), 3, 3), dtype=f32)
    n[..., 2] = 1.0  # explicit normals: the face-normal helper would divide by zero for the degenerate ones
    s = M.Scene()
    s.push_model(M.triangle_soup_model(np.asarray(good + bad, dtype=f32), M.PbrMaterial((1, 1, 1, 1), M.NONE, M.NONE, 0.0, 1.0, M.NONE),
                                       normals=n.reshape(-1, 3)))
    s.push_model(camera_and_lights())
    aux = compare(ctx, oracle, s, 80, 60)
    hit = aux.hit_src[aux.hit_src != 0xFFFFFFFF]
    assert len(hit) and hit.max() < len(good), "only the two proper triangles can be hit"


def test_shared_edges_and_vertices(ctx, oracle):
    """A regular grid seen head-on: many pixel centres fall exactly on shared edges and vertices, where the three
    edge functions are exactly zero (accepted: the tests reject on '< 0' only, triangle.rs:123-150) and the two
    neighbours tie in depth."""
    k = 8
    xs = np.linspace(-1.0, 1.0, k + 1, dtype=f32)
    tris = []
    for i in range(k):
        for j in range(k):
            a, b, c, d = (xs[i], xs[j], 0), (xs[i + 1], xs[j], 0), (xs[i], xs[j + 1], 0), (xs[i + 1], xs[j + 1], 0)
            tris += [[a, b, d], [a, d, c]]
    # yfov chosen so that the grid spans the image exactly: tan(yfov/2) * 4 = 1
    compare(ctx, oracle, soup_scene(tris, yfov=2 * np.arctan(0.25)), 64, 64)
    compare(ctx, oracle, soup_scene(tris, yfov=2 * np.arctan(0.25)), 128, 128, cfg=R.Config(integrator=R.IntegratorStrategy.Raytracer, max_depth=0))


def test_geometry_far_from_the_origin(ctx, oracle):
    """Coordinates around 1e5 (f32 spacing 1/128): the conservative slab test of the performance tree widens with
    |o * rd| per axis and must still never hide a triangle the exact test reaches."""
    rng = np.random.default_rng(7)
    c = rng.uniform(-3.0, 3.0, (400, 1, 3))
    t = (c + rng.normal(scale=0.4, size=(400, 3, 3))).astype(f32)
    off = np.array([1.0e5, -2.0e5, 3.0e5], dtype=f32)
    s = M.Scene()
    s.push_model(M.triangle_soup_model(t + off, M.PbrMaterial((1, 1, 1, 1), M.NONE, M.NONE, 0.0, 1.0, M.NONE)))
    dm = camera_and_lights()
    for n in dm.nodes:
        n.trs.translation = (n.trs.translation + off + np.array([0, 0, 8.0], dtype=f32) * (n.camera is not None)).astype(f32)
    s.push_model(dm)
    compare(ctx, oracle, s, 120, 90)


def test_single_huge_and_single_tiny_triangle(ctx, oracle):
    """One triangle far larger than the view and one far smaller than a pixel, 12 orders of magnitude apart in area."""
    tris = [[[-1e4, -1e4, -5], [1e4, -1e4, -5], [0, 1e4, -5]], [[0.1, 0.1, 0], [0.1001, 0.1, 0], [0.1, 0.1001, 0]]]
    compare(ctx, oracle, soup_scene(tris), 101, 77)


def test_many_samples_per_pixel_on_a_ragged_image(ctx, oracle):
    """Stratified 3x3 and 4x4 samples (scene.rs:125-131) on sizes that leave partly filled warp tiles: the
    sample-parallel tiles must add a pixel's samples in the reference's order."""
    tris = [[[-1, -1, 0], [1, -1, 0], [0, 1, 0]], [[-2, -1, -1], [2, -1, -1], [0, 2, -1]]]
    for spp, (w, h) in ((9, (21, 13)), (16, (10, 7)), (5, (19, 3))):
        compare(ctx, oracle, soup_scene(tris), w, h, cfg=R.Config(max_depth=1, samples_per_pixel=spp))


def test_shared_frame_late_rank_is_reported_not_blended(ctx):
    """a flag wait that gives up must surface as an error from a synchronising call, not as a silently torn
    frame.  The owner waits for two ranks, only rank 0 ever stores: with a 50 ms limit rayca_cuda_frame_sync returns
    RAYCA_E_TIMEOUT (sticky until cleared); once the missing rank has stored, the same wait succeeds."""
    import time

    sc = R.DeviceScene(ctx, load_desc("box"))
    cfg = R.Config(max_depth=1)
    frame = R.SharedFrame(ctx, 64, 64)
    frame.set_timeout(50.0)
    sc.render_to_frame(cfg, frame, 1, options=M.rayca_options(0, 0, 0, 2, 8, 0, 0))  # rank 0 of 2
    t0 = time.perf_counter()
    frame.wait(2, 1)
    with pytest.raises(R.RaycaError) as e:
        frame.sync()
    waited = time.perf_counter() - t0
    assert e.value.code == -6 and "undefined" in str(e.value)
    assert 0.04 <= waited < 1.5, f"the wait gave up after {waited:.3f} s, limit 50 ms"
    assert frame.timed_out()
    with pytest.raises(R.RaycaError):  # sticky: still reported
        frame.sync()
    frame.clear_timeout()
    sc.render_to_frame(cfg, frame, 1, options=M.rayca_options(0, 0, 1, 2, 8, 0, 0))  # the late rank arrives
    frame.wait(2, 1)
    frame.sync()
    assert not frame.timed_out()
    with pytest.raises(R.RaycaError):
        frame.set_timeout(0.0)
    frame.close()
    sc.close()


@pytest.mark.parametrize("scene,cfg,w,h,chunks", [
    ("suzanne", R.Config(max_depth=1, samples_per_pixel=16), 200, 120, [1, 3, 4, 8]),
    ("cubes_over_plane", R.Config(max_depth=3, samples_per_pixel=9), 160, 96, [2, 2, 5]),  # bounces: keyed by sample index
    ("box", R.Config(max_depth=1, samples_per_pixel=1), 64, 64, [1]),
    ("glass", R.Config(integrator=R.IntegratorStrategy.Scratcher, max_depth=3, samples_per_pixel=4, gamma=2.2), 96, 96, [3, 1]),
])
def test_progressive_accumulation_equals_one_call(ctx, scene, cfg, w, h, chunks):
    """samples accumulated over several calls, in sample order, give the bytes of ONE call with the
    summed samples_per_pixel (same strata, same f32 additions); intermediate frames are the mean so far."""
    dev = R.DeviceScene(ctx, load_desc(scene))
    opts = M.rayca_options(0x5EED, 0, 0, 0, 0, 0, 0)
    want, _, wst = dev.render(cfg, w, h, options=opts)
    acc = R.Accumulator(ctx, w, h)
    for rnd in range(2):  # a reset accumulator starts over
        rays = 0
        for n in chunks:
            got, st = dev.render_accumulate(cfg, acc, n, options=opts)
            rays += st.rays_primary
        assert acc.samples == cfg.samples_per_pixel == sum(chunks)
        assert np.array_equal(got, want), f"round {rnd}: {(got != want).sum()} bytes differ from the single call"
        assert rays == wst.rays_primary
        with pytest.raises(R.RaycaError):  # the sequence is complete
            dev.render_accumulate(cfg, acc, 1, options=opts)
        acc.reset()
    if len(chunks) > 1:  # an intermediate frame is the mean of the samples so far: not black, not the final frame
        mid, _ = dev.render_accumulate(cfg, acc, chunks[0], options=opts)
        assert mid.any() and mid.shape == want.shape
    acc.close()
    dev.close()


def test_over_deep_performance_tree_is_refused_not_misrendered(ctx, monkeypatch):
    """VERDICT r01 hygiene: the walk's stack saturates beyond kStackSize entries, so a tree that deep must be refused at
    scene_create.  The binned builder cannot produce such a tree from f32 input, hence the limit override."""
    monkeypatch.setenv("RAYCA_MAX_TREE_DEPTH", "6")
    with pytest.raises(R.RaycaError) as e:
        R.DeviceScene(ctx, load_desc("suzanne"))
    assert e.value.code == -2 and "deeper than the traversal stack" in str(e.value)
    monkeypatch.delenv("RAYCA_MAX_TREE_DEPTH")
    sc = R.DeviceScene(ctx, load_desc("suzanne"))  # the same scene is accepted at the real limit
    sc.close()


def test_ggx_material_under_the_raytracer_is_refused(ctx):
    """The reference panics (Hit::get_radiance -> get_pbr_material on Material::Ggx): RAYCA_E_UNSUPPORTED, like its other panics."""
    sc = R.DeviceScene(ctx, load_desc("ggx"))
    for integ in (R.IntegratorStrategy.Raytracer, R.IntegratorStrategy.Scratcher):
        with pytest.raises(R.RaycaError) as e:
            sc.render(R.Config(integrator=integ, max_depth=1), 32, 32)
        assert e.value.code == -2
    sc.close()


def test_scene_outliving_its_context_is_destroyed_safely(cuda_lib):
    """Garbage-collected bindings may destroy a scene after its context: shutdown releases the scene's device arrays and
    orphans it, the later destroy only frees host memory (a crash here once surfaced as a segfault at interpreter exit)."""
    c = R.Context(0)
    sc = R.DeviceScene(c, load_desc("box"))
    sc.render(R.Config(max_depth=1), 32, 32)
    multi = R.Context(devices=[0, 0])
    ms = R.DeviceScene(multi, load_desc("box"))
    c.close()
    multi.close()
    sc.close()
    ms.close()
    c2 = R.Context(0)  # the library is still usable
    sc2 = R.DeviceScene(c2, load_desc("box"))
    rgba8, _, _ = sc2.render(R.Config(max_depth=1), 32, 32)
    assert rgba8.any()
    sc2.close()
    c2.close()


def test_out_of_range_vertex_index_is_refused_on_both_emit_paths(ctx, monkeypatch):
    """An index beyond its geometry's vertices: the host emit loop refuses the scene with RAYCA_E_INVALID_ARG; so must the
    device emit (models of >= 16384 mesh triangles are emitted by a kernel, which raises a flag the host reads back)."""
    import ctypes as C

    from rayca_b200 import procedural as P

    desc = P.soup(20_000).to_desc()
    idx = np.ctypeslib.as_array((C.c_uint32 * int(desc.desc.index_count)).from_address(int(desc.desc.indices)))
    good = int(idx[12345])
    idx[12345] = 0x7FFFFFFF
    try:
        for host in (False, True):
            if host:
                monkeypatch.setenv("RAYCA_HOST_EMIT", "1")
            with pytest.raises(R.RaycaError) as e:
                R.DeviceScene(ctx, desc)
            assert e.value.code == -1 and "vertex index out of range" in str(e.value)
    finally:
        idx[12345] = good
    monkeypatch.delenv("RAYCA_HOST_EMIT", raising=False)
    R.DeviceScene(ctx, desc).close()  # the repaired scene builds
