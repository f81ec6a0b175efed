"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle and the golden vectors.

Bars (BASELINE.json north_star): per-pixel hit ids / depth / barycentrics bit-exact; RGB within
1e-3 absolute (PSNR >= 60 dB) — transcendental functions (powf/acosf/tanf/cosf/sinf) differ by ulps
between glibc and CUDA's libdevice and feed only the colour."""
import numpy as np
import pytest

from conftest import case_config, golden_cases, load_case, load_desc
from rayca_b200 import model as M
from rayca_b200 import renderer as R

pytestmark = pytest.mark.gpu

RGB_TOL = 1e-3


@pytest.fixture(scope="module")
def ctx(cuda_lib):
    c = R.Context(0)
    yield c
    c.close()


_dev = {}


def dev_scene(ctx, scene, bvh=True):
    key = (scene, bvh)
    if key not in _dev:
        _dev[key] = R.DeviceScene(ctx, load_desc(scene), bvh=bvh)
    return _dev[key]


def psnr(a, b):
    a = np.clip(np.nan_to_num(a[..., :3]), 0, 1)
    b = np.clip(np.nan_to_num(b[..., :3]), 0, 1)
    mse = np.mean((a - b) ** 2)
    return 99.0 if mse == 0 else 10 * np.log10(1.0 / mse)


def check_against(z, rgba8, aux, h, w, stochastic):
    hit_src = aux.hit_src.reshape(h, w)
    assert np.array_equal(hit_src, z["hit_src"]), f"{(hit_src != z['hit_src']).sum()} primary hit ids differ"
    assert np.array_equal(aux.hit_prim.reshape(h, w), z["hit_prim"])
    assert np.array_equal(aux.hit_blas.reshape(h, w), z["hit_blas"])
    assert np.array_equal(aux.depth.reshape(h, w), z["depth"]), "primary depth must be bit-exact"
    assert np.array_equal(aux.uv.reshape(h, w, 2), z["uv"]), "barycentrics must be bit-exact"
    got, want = aux.rgba_f32.reshape(h, w, 4), z["rgba_f32"]
    assert np.array_equal(np.isnan(got), np.isnan(want)), "NaN pattern differs"
    diff = np.abs(np.nan_to_num(got) - np.nan_to_num(want))[..., :3]
    # relative for HDR values > 1, absolute below (the 8-bit image saturates at 1.0)
    scaled = diff / np.maximum(1.0, np.abs(np.nan_to_num(want))[..., :3])
    if stochastic:
        # a bounce direction is built from acosf/cosf/sinf: ulp-level differences can flip which
        # triangle a secondary ray near an edge hits, so a few pixels may differ visibly
        bad = (scaled.max(axis=-1) > RGB_TOL).mean()
        assert bad <= 0.01, f"{bad:.4%} of pixels beyond tolerance"
    else:
        assert scaled.max() <= RGB_TOL, f"max scaled RGB error {scaled.max()}"
        d8 = np.abs(rgba8.astype(np.int32) - z["rgba8"].astype(np.int32))
        assert d8.max() <= 1 and (d8 > 0).mean() <= 0.01, "RGBA8 may differ by one step on <=1% of channels"
    assert psnr(got, want) >= 60.0


@pytest.mark.parametrize("traversal", [R.TRAVERSAL_ORDERED, R.TRAVERSAL_EXHAUSTIVE, R.TRAVERSAL_REFERENCE_ORDERED])
@pytest.mark.parametrize("scene,case", golden_cases())
def test_cuda_matches_golden(ctx, scene, case, traversal):
    z = load_case(scene, case)
    cfg = case_config(z)
    h, w = z["hit_src"].shape
    sc = dev_scene(ctx, scene, cfg.bvh)
    opts = M.rayca_options(int(z["seed"][0]), traversal, 0, 0, 0, 0, 0)
    try:
        rgba8, aux, st = sc.render(cfg, w, h, options=opts, want_aux=True)
    except R.RaycaError as e:
        if traversal != R.TRAVERSAL_ORDERED and e.code == -2 and "triangle scenes only" in str(e):
            pytest.skip("the reference-tree walks are triangle-only; sphere scenes run on the product walk")
        raise
    # directions built with acosf / cosf / sinf (bounces, the Brdf sample of Mis) differ from glibc's by ulps
    stochastic = cfg.integrator == R.IntegratorStrategy.Pathtracer and (
        cfg.max_depth > 1 or cfg.russian_roulette or cfg.direct_sampler == R.SamplerStrategy.Mis)
    check_against(z, rgba8, aux, h, w, stochastic)
    rays = [int(x) for x in z["rays"]]
    assert st.rays_primary == rays[0]
    if not stochastic:
        assert [st.rays_shadow, st.rays_secondary] == rays[1:]


@pytest.mark.parametrize("scene,case,w,h", [("box", "pt1", 333, 217), ("suzanne", "pt1", 200, 120),
                                            ("cubes_over_plane", "rt3", 150, 100), ("orientation", "rt0", 257, 129)])
def test_cuda_matches_live_oracle(ctx, oracle, scene, case, w, h):
    """Ragged sizes (not multiples of the 16x8 block) against the oracle run live on the same input."""
    z = load_case(scene, case)
    cfg = case_config(z)
    osc = oracle.OracleScene(load_desc(scene), bvh=cfg.bvh)
    o8, oaux, ost = osc.render(cfg.to_c(), w, h)
    rgba8, aux, st = dev_scene(ctx, scene, cfg.bvh).render(cfg, w, h, want_aux=True)
    want = dict(hit_src=oaux.hit_src.reshape(h, w), hit_prim=oaux.hit_prim.reshape(h, w),
                hit_blas=oaux.hit_blas.reshape(h, w), depth=oaux.depth.reshape(h, w), uv=oaux.uv.reshape(h, w, 2),
                rgba_f32=oaux.rgba_f32.reshape(h, w, 4), rgba8=o8)
    check_against(want, rgba8, aux, h, w, False)
    assert [st.rays_primary, st.rays_shadow, st.rays_secondary] == [ost.rays_primary, ost.rays_shadow, ost.rays_secondary]


def test_ordered_equals_exhaustive_at_full_size(ctx):
    """Size-independent property at 1080p: the distance-culled walk returns exactly what the
    reference-order exhaustive walk returns (ids, depth, barycentrics, colour bits)."""
    sc = dev_scene(ctx, "suzanne")
    cfg = R.Config(max_depth=1)
    a8, a, _ = sc.render(cfg, 1920, 1080, options=M.rayca_options(0, R.TRAVERSAL_ORDERED, 0, 0, 0, 0, 0), want_aux=True)
    b8, b, _ = sc.render(cfg, 1920, 1080, options=M.rayca_options(0, R.TRAVERSAL_EXHAUSTIVE, 0, 0, 0, 0, 0), want_aux=True)
    assert np.array_equal(a.hit_src, b.hit_src) and np.array_equal(a.depth, b.depth) and np.array_equal(a.uv, b.uv)
    assert np.array_equal(a.rgba_f32.view(np.uint32), b.rgba_f32.view(np.uint32)) and np.array_equal(a8, b8)


def test_render_is_deterministic(ctx):
    sc = dev_scene(ctx, "cubes_over_plane")
    cfg = R.Config(max_depth=3)
    opts = M.rayca_options(1234, 0, 0, 0, 0, 0, 0)
    a8, a, _ = sc.render(cfg, 128, 128, options=opts, want_aux=True)
    b8, b, _ = sc.render(cfg, 128, 128, options=opts, want_aux=True)
    assert np.array_equal(a8, b8) and np.array_equal(a.rgba_f32.view(np.uint32), b.rgba_f32.view(np.uint32))
    _, c, _ = sc.render(cfg, 128, 128, options=M.rayca_options(99, 0, 0, 0, 0, 0, 0), want_aux=True)
    assert not np.array_equal(a.rgba_f32, c.rgba_f32), "the seed must matter once a bounce is sampled"


def test_draw_trait_mirror(ctx, cuda_lib):
    """`Draw::draw(&Scene, &mut Image)` through CudaRenderer and the one-shot C entry point."""
    import ctypes as C

    z = load_case("triangle", "pt1")
    s = M.Scene()
    tri = M.Model()
    tri.geometries.append(M.TriangleMesh(M.make_vertices([[0, 0, 0], [1, 0, 0], [0, 1, 0]]), np.array([0, 1, 2], dtype=np.uint32)))
    tri.primitives.append(M.Primitive(0))
    tri.meshes.append(M.Mesh([0]))
    tri.nodes.append(M.Node(mesh=0))
    tri.root.children.append(0)
    s.push_model(tri)
    s.push_model(R.CudaRenderer.create_default_model())
    img = R.Image(64, 64)
    r = R.CudaRenderer(R.Config(max_depth=1))
    r.draw(s, img)
    assert img.buffer[..., 3].min() == 255 and img.buffer[..., :3].max() > 0
    out = np.zeros((64, 64, 4), dtype=np.uint8)
    cfg = R.Config(max_depth=1).to_c()
    assert cuda_lib.rayca_cuda_draw(C.cast(s.to_desc().ptr(), C.c_void_p), C.byref(cfg), 64, 64, out.ctypes.data) == 0
    assert np.array_equal(out, img.buffer)
    del z


def test_unsupported_and_error_paths(ctx):
    sc = dev_scene(ctx, "box")
    for cfg in (R.Config(direct_sampler=R.SamplerStrategy.Mis), R.Config(indirect_sampler=R.SamplerStrategy.Brdf),
                R.Config(light_samples=4, russian_roulette=True), R.Config(light_samples=2, max_depth=9)):
        # Mis / Brdf need Phong or GGX materials (box.gltf is PBR); branching bounces run without roulette, 8 levels deep
        with pytest.raises(R.RaycaError) as e:
            sc.render(cfg, 16, 16)
        assert e.value.code == -2
    with pytest.raises(R.RaycaError) as e:
        sc.render(R.Config(samples_per_pixel=0), 16, 16)
    assert e.value.code == -1
    # no camera in the scene (scene.rs:109 asserts) -> RAYCA_E_NO_CAMERA
    s = M.Scene()
    s.push_model(M.triangle_soup_model([[[0, 0, 0], [1, 0, 0], [0, 1, 0]]]))
    nocam = R.DeviceScene(ctx, s.to_desc())
    with pytest.raises(R.RaycaError) as e:
        nocam.render(R.Config(max_depth=1), 16, 16)
    assert e.value.code == -5


def test_sharded_rows_reassemble(ctx):
    """The multi-GPU split (tile t -> shard t % G) rendered shard by shard on one device
    reassembles into exactly the unsharded frame."""
    import torch

    sc = dev_scene(ctx, "cubes_over_plane")
    cfg = R.Config(max_depth=1)
    w, h, G, tile = 200, 100, 3, 8
    full, _, _ = sc.render(cfg, w, h)
    tiles = (h + tile - 1) // tile
    per = (tiles + G - 1) // G
    out = np.zeros((per * G * tile, w, 4), dtype=np.uint8)
    for g in range(G):
        buf = torch.zeros((per * tile, w, 4), dtype=torch.uint8, device="cuda:0")
        sc.render_device(cfg, w, h, buf.data_ptr(), options=M.rayca_options(0, 0, g, G, tile, 0, 0))
        torch.cuda.synchronize()
        local = buf.cpu().numpy().reshape(per, tile, w, 4)
        for k in range(per):
            t = k * G + g
            out[t * tile:(t + 1) * tile] = local[k]
    assert np.array_equal(out[:h], full)


@pytest.fixture(scope="module")
def atrium(ctx):
    """BASELINE M3 stand-in at full size: the procedural Sponza-class atrium bench.py times."""
    from rayca_b200 import procedural as P

    sc = R.DeviceScene(ctx, P.atrium(262_267).to_desc())
    yield sc
    sc.close()


def _bits(aux):
    return aux.hit_src, aux.hit_prim, aux.hit_blas, aux.depth.view(np.uint32), aux.uv.view(np.uint32), aux.rgba_f32.view(np.uint32)


@pytest.mark.parametrize("other", [R.TRAVERSAL_EXHAUSTIVE, R.TRAVERSAL_REFERENCE_ORDERED])
def test_product_walk_equals_reference_walks_on_the_bench_frame(ctx, atrium, other):
    """1920x1080, 262k triangles, primary + shadow rays: the performance-layer walk returns bit for bit what
    the reference tree walked with the reference's visiting rule (exhaustive, depth-first, first hit wins
    ties) returns — ids, depth, barycentrics, float colour and RGBA8."""
    cfg = R.Config(max_depth=1)
    a8, a, sa = atrium.render(cfg, 1920, 1080, options=M.rayca_options(0, R.TRAVERSAL_ORDERED, 0, 0, 0, 0, 0), want_aux=True)
    b8, b, sb = atrium.render(cfg, 1920, 1080, options=M.rayca_options(0, other, 0, 0, 0, 0, 0), want_aux=True)
    for x, y in zip(_bits(a), _bits(b)):
        assert np.array_equal(x, y)
    assert np.array_equal(a8, b8)
    assert (sa.rays_primary, sa.rays_shadow) == (sb.rays_primary, sb.rays_shadow)
    assert (a.hit_src != 0xFFFFFFFF).mean() > 0.9, "the camera sits inside the atrium: nearly every pixel hits"


def test_bounce_frame_equals_reference_walk(ctx, atrium):
    """BASELINE M5 (one cosine bounce, incoherent secondary + shadow rays) at 960x540: same keyed RNG, so the
    product walk and the reference-tree ordered walk must agree bit for bit."""
    cfg = R.Config(max_depth=2)
    a8, a, _ = atrium.render(cfg, 960, 540, options=M.rayca_options(0x5EED, R.TRAVERSAL_ORDERED, 0, 0, 0, 0, 0), want_aux=True)
    b8, b, _ = atrium.render(cfg, 960, 540, options=M.rayca_options(0x5EED, R.TRAVERSAL_REFERENCE_ORDERED, 0, 0, 0, 0, 0), want_aux=True)
    for x, y in zip(_bits(a), _bits(b)):
        assert np.array_equal(x, y)
    assert np.array_equal(a8, b8)


def test_persistent_scheduling_equals_static_grid(ctx, atrium):
    cfg = R.Config(max_depth=1, samples_per_pixel=4)
    a8, a, sa = atrium.render(cfg, 1001, 563, options=M.rayca_options(0, 0, 0, 0, 0, 0, 0), want_aux=True)
    b8, b, sb = atrium.render(cfg, 1001, 563, options=M.rayca_options(0, 0, 0, 0, 0, R.FLAG_PERSISTENT, 0), want_aux=True)
    for x, y in zip(_bits(a), _bits(b)):
        assert np.array_equal(x, y)
    assert np.array_equal(a8, b8)
    assert (sa.rays_primary, sa.rays_shadow) == (sb.rays_primary, sb.rays_shadow)


def test_host_render_bands_equal_single_launch(ctx, atrium):
    """rayca_cuda_render stores the frame straight into a pinned caller buffer from one launch (zero copy), and for
    pageable memory traces it in row bands, each read back on a side stream; both must equal one render_device
    launch over the whole frame."""
    import torch

    cfg = R.Config(max_depth=1)
    w, h = 1283, 721
    img = R.Image(w, h, pinned=True)
    img.buffer[...] = 7
    _, _, zc = atrium.render(cfg, w, h, out=img.buffer)
    assert zc.rays_primary == w * h and zc.kernel_launches == 1
    pageable, _, st = atrium.render(cfg, w, h)
    buf = torch.zeros((h, w, 4), dtype=torch.uint8, device="cuda:0")
    atrium.render_device(cfg, w, h, buf.data_ptr())
    torch.cuda.synchronize()
    one = buf.cpu().numpy()
    assert np.array_equal(img.buffer, one) and np.array_equal(pageable, one)
    assert st.rays_primary == w * h and st.kernel_launches == 3
    c = atrium.last_counters()
    assert c.rays_primary == w * h and c.rays_shadow == st.rays_shadow


def test_shared_frame_of_three_shards_equals_single_launch(ctx, atrium):
    """rayca_cuda_render_to_frame: three shards (here three launches of one process; on N GPUs one per rank,
    through the CUDA IPC mapping) store their interleaved 8-row tiles straight into ONE frame and publish their
    completion flags; frame_wait sees all three and the frame equals a single full launch, ragged last tile
    included."""
    import torch

    cfg = R.Config(max_depth=1)
    w, h = 1283, 717
    frame = R.SharedFrame(ctx, w, h)
    assert len(frame.export()) == 64
    for frame_id in (1, 2):
        for r in (2, 0, 1):
            atrium.render_to_frame(cfg, frame, frame_id, options=M.rayca_options(0x5EED, 0, r, 3, 8, 0, 0))
        frame.wait(3, frame_id)
        torch.cuda.synchronize()
        assert not frame.timed_out()
    buf = torch.zeros((h, w, 4), dtype=torch.uint8, device="cuda:0")
    atrium.render_device(cfg, w, h, buf.data_ptr())
    torch.cuda.synchronize()
    assert torch.equal(torch.as_tensor(frame, device="cuda:0"), buf)
    frame.close()


def test_host_frame_of_three_shards_equals_single_launch(ctx, atrium):
    """Multi-GPU host path on one device: three shards stored with RAYCA_FLAG_DIRECT_ROWS straight into one
    page-locked shared-memory frame (rayca_cuda_host_register) — through rayca_cuda_render and through
    render_device with the frame's device view — give the frame of a single launch; a sharded render into pageable
    memory or without the flag is refused."""
    import torch

    cfg = R.Config(max_depth=1)
    w, h, G = 1283, 721, 3
    full, _, _ = atrium.render(cfg, w, h)
    f = R.HostFrame(w, h)
    try:
        for g in range(G):
            _, _, st = atrium.render(cfg, w, h, out=f.pixels, options=M.rayca_options(0, 0, g, G, 8, R.FLAG_DIRECT_ROWS, 0))
            assert st.kernel_launches == 1
        assert np.array_equal(f.pixels, full)
        f.pixels[...] = 0
        for g in range(G):
            atrium.render_device(cfg, w, h, f.device_ptr, options=M.rayca_options(0, 0, g, G, 8, R.FLAG_DIRECT_ROWS, 0))
        torch.cuda.synchronize()
        assert np.array_equal(f.pixels, full)
        for out, flags in ((f.pixels, 0), (np.zeros((h, w, 4), dtype=np.uint8), R.FLAG_DIRECT_ROWS)):
            with pytest.raises(R.RaycaError) as e:
                atrium.render(cfg, w, h, out=out, options=M.rayca_options(0, 0, 0, G, 8, flags, 0))
            assert e.value.code == -1
    finally:
        f.close()


@pytest.mark.parametrize("G", [2, 3, 8])
def test_group_checkerboard_shards_equal_single_launch(ctx, atrium, G):
    """RAYCA_FLAG_SHARD_GROUPS: shard (gx + gy) mod G owns 16x8 pixel group (gx, gy); every pixel is stored exactly once,
    into a host frame (DIRECT_ROWS) and into a shared device frame; the ray counts of the shards add up to the frame's."""
    import torch

    cfg = R.Config(max_depth=1)
    w, h = 1283, 721  # ragged in x and y
    full, _, fst = atrium.render(cfg, w, h)
    flags = R.FLAG_DIRECT_ROWS | R.FLAG_SHARD_GROUPS
    f = R.HostFrame(w, h)
    try:
        f.pixels[...] = 0
        rays = [0, 0]
        for g in range(G):
            _, _, st = atrium.render(cfg, w, h, out=f.pixels, options=M.rayca_options(0, 0, g, G, 8, flags, 0))
            rays[0] += st.rays_primary
            rays[1] += st.rays_shadow
        assert np.array_equal(f.pixels, full), f"{(f.pixels != full).sum()} bytes differ"
        assert rays == [fst.rays_primary, fst.rays_shadow]
    finally:
        f.close()
    frame = R.SharedFrame(ctx, w, h)
    for g in range(G):
        atrium.render_to_frame(cfg, frame, 1, options=M.rayca_options(0, 0, g, G, 8, R.FLAG_SHARD_GROUPS, 0))
    torch.cuda.synchronize()
    assert np.array_equal(torch.as_tensor(frame, device="cuda:0").cpu().numpy(), full)
    frame.close()
    with pytest.raises(R.RaycaError) as e:  # compact (tile-packed) output cannot be sharded by groups
        buf = torch.zeros((h, w, 4), dtype=torch.uint8, device="cuda:0")
        atrium.render_device(cfg, w, h, buf.data_ptr(), options=M.rayca_options(0, 0, 0, G, 8, R.FLAG_SHARD_GROUPS, 0))
    assert e.value.code == -1


@pytest.mark.parametrize("scene,cfg,w,h", [
    ("suzanne", R.Config(max_depth=2), 333, 217),
    ("cubes_over_plane", R.Config(max_depth=3), 256, 144),
    ("cubes_over_plane", R.Config(max_depth=5, direct_sampler=R.SamplerStrategy.NONE), 160, 90),  # no NEE: emissive-only paths
    ("normalquad", R.Config(max_depth=4), 200, 120),
    ("box", R.Config(max_depth=17), 96, 96),  # the deepest path the fused kernel keeps
    ("scratcher_sphere", R.Config(max_depth=2), 128, 128),  # spheres: kernel instantiation MODE 3
])
# the stage of a depth: "fetch" (default) = persistent ray walk with dynamic fetch, then shading re-sorted by hit point;
# "spec" = the same with speculative traversal (a leaf is put aside while the lane goes on descending);
# "split" = the same two halves, one ray per thread; "stage" = one kernel per depth
@pytest.mark.parametrize("stage", ["fetch", "spec", "split", "stage"])
def test_wavefront_stages_equal_the_fused_kernel(ctx, scene, cfg, w, h, stage, monkeypatch):
    """bounce rays run as wavefront stages with the live rays radix-sorted between depths
    (wavefront.cu).  The RNG is keyed by (pixel, sample, depth, dimension) and every value is computed by the same
    device functions in the same order, so the frame, the parity planes and the ray counts equal the fused kernel's
    (RAYCA_FLAG_NO_WAVEFRONT) bit for bit — through rayca_cuda_render (row bands), and sharded."""
    import torch

    monkeypatch.setenv("RAYCA_WAVE_MIN_PATHS", "0")  # small frames normally keep the fused kernel
    monkeypatch.setenv("RAYCA_WAVE_SPLIT", "0" if stage == "stage" else "1")
    monkeypatch.setenv("RAYCA_WAVE_FETCH", "1" if stage in ("fetch", "spec") else "0")
    monkeypatch.setenv("RAYCA_WAVE_SPEC", "1" if stage == "spec" else "0")  # speculative traversal in the walk half
    sc = dev_scene(ctx, scene)
    wave = M.rayca_options(0x5EED, 0, 0, 0, 0, 0, 0)
    fused = M.rayca_options(0x5EED, 0, 0, 0, 0, R.FLAG_NO_WAVEFRONT, 0)
    a8, a, ast = sc.render(cfg, w, h, options=wave, want_aux=True)
    b8, b, bst = sc.render(cfg, w, h, options=fused, want_aux=True)
    assert np.array_equal(a8, b8), f"{(a8 != b8).sum()} bytes differ"
    assert np.array_equal(a.rgba_f32.view(np.uint32), b.rgba_f32.view(np.uint32))
    assert np.array_equal(a.hit_src, b.hit_src) and np.array_equal(a.depth, b.depth) and np.array_equal(a.uv, b.uv)
    assert (ast.rays_primary, ast.rays_shadow, ast.rays_secondary) == (bst.rays_primary, bst.rays_shadow, bst.rays_secondary)
    assert ast.kernel_launches > bst.kernel_launches  # the stages really ran
    # one launch over the whole frame on device memory, and three compact shards reassembled
    buf = torch.zeros((h, w, 4), dtype=torch.uint8, device="cuda:0")
    sc.render_device(cfg, w, h, buf.data_ptr(), options=wave)
    torch.cuda.synchronize()
    assert np.array_equal(buf.cpu().numpy(), b8)
    G, tile = 3, 8
    tiles = (h + tile - 1) // tile
    per = (tiles + G - 1) // G
    out = np.zeros((per * G * tile, w, 4), dtype=np.uint8)
    for g in range(G):
        part = torch.zeros((per * tile, w, 4), dtype=torch.uint8, device="cuda:0")
        sc.render_device(cfg, w, h, part.data_ptr(), options=M.rayca_options(0x5EED, 0, g, G, tile, 0, 0))
        torch.cuda.synchronize()
        out.reshape(per, G, tile, w, 4)[:, g] = part.cpu().numpy().reshape(per, tile, w, 4)
    assert np.array_equal(out[:h], b8), "sharded wavefront frame differs"
    # three shards stored at their final rows of a page-locked HOST frame (the multi-GPU default): the sorted stages'
    # pixels are staged on the device and leave through one coalesced copy kernel; and the zero-copy single launch
    f = R.HostFrame(w, h)
    try:
        for g in range(G):
            sc.render(cfg, w, h, out=f.pixels, options=M.rayca_options(0x5EED, 0, g, G, tile, R.FLAG_DIRECT_ROWS, 0))
        assert np.array_equal(f.pixels, b8), "host-frame wavefront shards differ"
        f.pixels[...] = 0
        sc.render_device(cfg, w, h, f.device_ptr, options=wave)
        torch.cuda.synchronize()
        assert np.array_equal(f.pixels, b8), "zero-copy wavefront frame differs"
    finally:
        f.close()


def test_shared_frame_wait_times_out_instead_of_hanging(ctx, atrium):
    """A rank that never publishes its flag must not hang the owner: the wait kernel gives up after ~2 s and
    frame_timed_out reports it."""
    import torch

    frame = R.SharedFrame(ctx, 64, 64)
    atrium.render_to_frame(R.Config(max_depth=1), frame, 7, options=M.rayca_options(0x5EED, 0, 0, 2, 8, 0, 0))
    frame.wait(2, 7)
    torch.cuda.synchronize()
    assert frame.timed_out()
    frame.close()


def test_dragon_direct_lighting_at_full_size(ctx):
    """The reference's sdtf/2/dragon.sdtf at its own 640x480 with its own settings (Direct integrator, two quad
    lights, 25 stratified light samples: 0.3 M primary + 15 M closest-hit shadow rays over 100,002 triangles): the
    product walk and the reference-tree ordered walk agree bit for bit."""
    sc = dev_scene(ctx, "dragon")
    cfg = R.Config(integrator=R.IntegratorStrategy.Direct, light_samples=25, light_stratify=True, max_depth=5)
    a8, a, sa = sc.render(cfg, 640, 480, options=M.rayca_options(0x5EED, R.TRAVERSAL_ORDERED, 0, 0, 0, 0, 0), want_aux=True)
    b8, b, sb = sc.render(cfg, 640, 480, options=M.rayca_options(0x5EED, R.TRAVERSAL_REFERENCE_ORDERED, 0, 0, 0, 0, 0), want_aux=True)
    for x, y in zip(_bits(a), _bits(b)):
        assert np.array_equal(x, y)
    assert np.array_equal(a8, b8)
    assert sa.rays_shadow == sb.rays_shadow and sa.rays_shadow > 10_000_000


def test_thousand_spheres_at_full_size(ctx):
    """sdtf/1/scene5.sdtf (1000 spheres, Raytracer depth 5) at 1280x960: deterministic, every hit is a sphere hit,
    and the static and persistent schedules agree."""
    sc = dev_scene(ctx, "scene5")
    cfg = R.Config(integrator=R.IntegratorStrategy.Raytracer, max_depth=5)
    a8, a, sa = sc.render(cfg, 1280, 960, want_aux=True)
    b8, b, sb = sc.render(cfg, 1280, 960, options=M.rayca_options(0, 0, 0, 0, 0, R.FLAG_PERSISTENT, 0), want_aux=True)
    for x, y in zip(_bits(a), _bits(b)):
        assert np.array_equal(x, y)
    assert np.array_equal(a8, b8)
    hit = a.hit_src != 0xFFFFFFFF
    assert 0.2 < hit.mean() < 0.9 and np.all(a.uv.reshape(-1, 2)[hit] == 0.0)
    assert sa.rays_secondary > 0 and sa.rays_shadow == sb.rays_shadow


def _host_tree_scene(ctx, desc):
    import os

    os.environ["RAYCA_HOST_FAST_TREE"] = "1"
    try:
        return R.DeviceScene(ctx, desc)
    finally:
        del os.environ["RAYCA_HOST_FAST_TREE"]


@pytest.mark.parametrize("scene", ["box", "suzanne", "scene5", "quadsphere", "dragon", "cornell"])
def test_device_built_tree_renders_what_the_host_built_tree_renders(ctx, scene):
    """The performance tree is built on the GPU (gpu_fast_build.cu); RAYCA_HOST_FAST_TREE=1 keeps the host builder.
    Both trees must give the same frame bit for bit (ids, depth, barycentrics, colour): small scenes (one warp
    builds the whole tree), spheres (pure leaves), 100k triangles (level passes + warp subtrees)."""
    try:
        desc = load_desc(scene)
    except Exception as e:  # scene not among the fixtures of this checkout
        pytest.skip(str(e))
    a = R.DeviceScene(ctx, desc)
    b = _host_tree_scene(ctx, desc)
    quad_lit = scene in ("quadsphere", "dragon", "cornell")  # quad lights: Direct integrator (sdtf/2 settings)
    cfg = R.Config(integrator=R.IntegratorStrategy.Direct, light_samples=4, light_stratify=True, max_depth=5) if quad_lit \
        else R.Config(integrator=R.IntegratorStrategy.Raytracer, max_depth=3)
    a8, aa, sa = a.render(cfg, 256, 192, want_aux=True)
    b8, bb, sb = b.render(cfg, 256, 192, want_aux=True)
    for x, y in zip(_bits(aa), _bits(bb)):
        assert np.array_equal(x, y)
    assert np.array_equal(a8, b8)
    assert (sa.rays_primary, sa.rays_shadow, sa.rays_secondary) == (sb.rays_primary, sb.rays_shadow, sb.rays_secondary)
    a.close()
    b.close()


def test_device_built_tree_on_the_bench_scene(ctx, atrium):
    """262k triangles at 1080p: device-built against host-built tree, and the device build is repeatable (the
    layout is decided by scans, not by atomics)."""
    from rayca_b200 import procedural as P

    desc = P.atrium(262_267).to_desc()
    host = _host_tree_scene(ctx, desc)
    again = R.DeviceScene(ctx, desc)
    cfg = R.Config(max_depth=1)
    a8, aa, _ = atrium.render(cfg, 1920, 1080, want_aux=True)
    b8, bb, _ = host.render(cfg, 1920, 1080, want_aux=True)
    c8, cc, _ = again.render(cfg, 1920, 1080, want_aux=True)
    for x, y, z in zip(_bits(aa), _bits(bb), _bits(cc)):
        assert np.array_equal(x, y) and np.array_equal(x, z)
    assert np.array_equal(a8, b8) and np.array_equal(a8, c8)
    host.close()
    again.close()


def _host_ref_scene(ctx, desc):
    import os

    os.environ["RAYCA_HOST_REF_TREE"] = "1"
    try:
        return R.DeviceScene(ctx, desc)
    finally:
        del os.environ["RAYCA_HOST_REF_TREE"]


def _big_scene_descs():
    from rayca_b200 import procedural as P

    yield "atrium", P.atrium(262_267).to_desc()
    yield "soup", P.soup(200_000).to_desc()
    for name in ("dragon", "scene7"):
        yield name, load_desc(name)


def test_device_built_reference_blas_equals_host_built(ctx):
    """The reference BLAS (blas.rs:64-127,245-316) of models with >= 16384 primitives is built on the GPU: node
    boxes, child links, leaf ranges and the primitive order left by the reference's swap partition must equal the
    host builder's (which the CPU tests pin against the oracle) bit for bit."""
    for name, desc in _big_scene_descs():
        a = R.DeviceScene(ctx, desc)
        b = _host_ref_scene(ctx, desc)
        ia, na, noa, poa, sa = a.bvh()
        ib, nb, nob, pob, sb = b.bvh()
        assert (ia.blas_count, ia.node_count, ia.triangle_count, ia.tlas_node_count) == \
            (ib.blas_count, ib.node_count, ib.triangle_count, ib.tlas_node_count), name
        assert np.array_equal(na.view(np.uint32), nb.view(np.uint32)), name
        assert np.array_equal(noa, nob) and np.array_equal(poa, pob) and np.array_equal(sa, sb), name
        a.close()
        b.close()


def test_branching_bounces_estimate_the_same_image(ctx):
    """light_samples > 1 with bounces (pathtracer.rs:23-66: every hit spawns light_samples indirect rays): the tree of
    paths is an estimator of the same depth-truncated image as the single chain, so the means agree (within 4 %),
    and the secondary-ray count is bounded by the full tree.  Without a direct sampler: the reference's NEE sums its
    light samples without dividing by their number (the oracle and the golden cases keep that), so only the
    indirect estimator, which does divide (pathtracer.rs:65), is independent of light_samples."""
    sc = dev_scene(ctx, "cornell")
    P, S = R.IntegratorStrategy.Pathtracer, R.SamplerStrategy
    base = dict(integrator=P, max_depth=4, light_stratify=False, direct_sampler=S.NONE)
    means = {}
    for name, kw in {"chain": dict(indirect_sampler=S.Cosine, light_samples=1, samples_per_pixel=64),
                     "tree2 cosine": dict(indirect_sampler=S.Cosine, light_samples=2, samples_per_pixel=32),
                     "tree3 hemisphere": dict(indirect_sampler=S.Hemisphere, light_samples=3, samples_per_pixel=16)}.items():
        _, aux, st = sc.render(R.Config(**base, **kw), 240, 240, want_aux=True)
        rgb = aux.rgba_f32.reshape(240, 240, 4)[..., :3]
        assert np.isfinite(rgb).all()
        means[name] = float(rgb.mean())
        ls, spp = kw["light_samples"], kw["samples_per_pixel"]
        assert st.rays_primary == 240 * 240 * spp
        assert 0 < st.rays_secondary <= st.rays_primary * (ls + ls ** 2 + ls ** 3)  # max_depth 4: traced indirect rays leave depths 0, 1, 2
    for name, m in means.items():
        assert abs(m - means["chain"]) <= 0.04 * means["chain"], means


def test_cornell_estimators_agree_in_the_mean(ctx):
    """sdtf/3 + sdtf/4 settings at the files' own 480x480: hemisphere and cosine bounces, with and without next event
    estimation on the quad light, with and without russian roulette, the Brdf sampler and multiple importance
    sampling are estimators of the same image, so their mean radiance must agree (64 spp: within 3 %)."""
    sc = dev_scene(ctx, "cornell")
    P, S = R.IntegratorStrategy.Pathtracer, R.SamplerStrategy
    base = dict(integrator=P, max_depth=16, samples_per_pixel=64, light_samples=1, light_stratify=False)
    means = {}
    for name, kw in {"hemisphere": dict(direct_sampler=S.NONE, indirect_sampler=S.Hemisphere),
                     "cosine": dict(direct_sampler=S.NONE, indirect_sampler=S.Cosine),
                     "nee": dict(direct_sampler=S.Nee, indirect_sampler=S.Hemisphere),
                     "nee+rr": dict(direct_sampler=S.Nee, indirect_sampler=S.Cosine, russian_roulette=True),
                     "nee+brdf+rr": dict(direct_sampler=S.Nee, indirect_sampler=S.Brdf, russian_roulette=True),
                     "mis+brdf+rr": dict(direct_sampler=S.Mis, indirect_sampler=S.Brdf, russian_roulette=True)}.items():
        _, aux, st = sc.render(R.Config(**base, **kw), 480, 480, want_aux=True)
        rgb = aux.rgba_f32.reshape(480, 480, 4)[..., :3]
        assert np.isfinite(rgb).all()
        means[name] = float(rgb.mean())
        assert st.rays_primary == 480 * 480 * 64
    ref = means["nee"]
    for name, m in means.items():
        assert abs(m - ref) <= 0.03 * ref, means


@pytest.mark.parametrize("scene", ["suzanne", "cubes_over_plane", "normalquad", "orientation", "cornell", "quadsphere", "dragon",
                                   "flighthelmet", "atrium"])
def test_device_computed_attributes_equal_the_host_pack(ctx, scene, monkeypatch):
    """scene_create computes the shading attributes of mesh triangles on the device from the caller's vertices
    (mesh_attr_kernel) instead of packing them on the host; under RAYCA_VERIFY_ATTRS it also runs the host pack and
    refuses the scene unless the two arrays are equal byte for byte.  Scenes with scaled / rotated nodes, several BLAS,
    vertex colours, tangents, quad-light triangles (cornell: written by the host, next to the kernel's) and spheres."""
    monkeypatch.setenv("RAYCA_VERIFY_ATTRS", "1")
    if scene == "atrium":
        from rayca_b200 import procedural as P

        desc = P.atrium(262_267).to_desc()
    else:
        desc = load_desc(scene)
    sc = R.DeviceScene(ctx, desc)  # raises RaycaError on a mismatch
    sc.close()
    monkeypatch.setenv("RAYCA_HOST_ATTRS", "1")  # the host ring path still works
    R.DeviceScene(ctx, desc).close()


def test_device_emitted_models_equal_host_emitted(ctx, monkeypatch):
    """Models made of triangle meshes only (>= 16384 triangles) are emitted on the device from the caller's vertices and
    indices (emit_mesh_device), their reference BLAS is built from bounds that never left the device and the packed
    positions are a gather by the build's order.  Against the host emit (RAYCA_HOST_EMIT=1): the same BLAS and primitive
    order, and the same frame bit for bit in every plane — on the atrium (one mesh, rotated root), the dragon, a 200k soup
    and FlightHelmet (several meshes under their own node transforms)."""
    from rayca_b200 import procedural as P

    descs = [("atrium", P.atrium(262_267).to_desc()), ("soup", P.soup(200_000).to_desc()), ("dragon", load_desc("dragon")),
             ("flighthelmet", load_desc("flighthelmet"))]
    cfg = R.Config(max_depth=1)
    for name, desc in descs:
        a = R.DeviceScene(ctx, desc)
        monkeypatch.setenv("RAYCA_HOST_EMIT", "1")
        b = R.DeviceScene(ctx, desc)
        monkeypatch.delenv("RAYCA_HOST_EMIT")
        ia, na, noa, poa, sa = a.bvh()
        ib, nb, nob, pob, sb = b.bvh()
        assert (ia.blas_count, ia.node_count, ia.triangle_count) == (ib.blas_count, ib.node_count, ib.triangle_count), name
        assert np.array_equal(na.view(np.uint32), nb.view(np.uint32)), name
        assert np.array_equal(noa, nob) and np.array_equal(poa, pob) and np.array_equal(sa, sb), name
        a8, aa, ast = a.render(cfg, 480, 270, want_aux=True)
        b8, ba, bst = b.render(cfg, 480, 270, want_aux=True)
        assert np.array_equal(a8, b8), f"{name}: {(a8 != b8).sum()} bytes differ"
        for plane in ("hit_src", "hit_prim", "hit_blas"):
            assert np.array_equal(getattr(aa, plane), getattr(ba, plane)), (name, plane)
        for plane in ("depth", "uv", "rgba_f32"):
            assert np.array_equal(getattr(aa, plane).view(np.uint32), getattr(ba, plane).view(np.uint32)), (name, plane)
        assert (ast.rays_primary, ast.rays_shadow) == (bst.rays_primary, bst.rays_shadow), name
        a.close()
        b.close()
