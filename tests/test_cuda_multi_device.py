"""One context over several devices (rayca_cuda_init_devices): ONE `render` call shards the frame's 8-row tiles over
the devices from the calling thread — the reference's `draw` fans out over its rayon pool the same way
(rayca-soft/src/scene.rs:117-122).  A device index may be listed more than once (every entry is a shard with its own
streams and scene replica), which is how a single-GPU box exercises replication and both output paths."""
import numpy as np
import pytest

from conftest import load_desc
from rayca_b200 import model as M
from rayca_b200 import renderer as R

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def single(cuda_lib):
    c = R.Context(0)
    yield c
    c.close()


def device_count(lib):
    # the library itself reports its shards: a context over "every visible device"
    ctx = R.Context(devices=[])
    k = ctx.device_count
    ctx.close()
    return k


@pytest.mark.parametrize("devices", [[0, 0], [0, 0, 0], None])
@pytest.mark.parametrize("scene,cfg,w,h", [
    ("suzanne", R.Config(max_depth=1), 333, 217),  # 28 tiles, the last one a single row
    ("cubes_over_plane", R.Config(max_depth=3), 640, 360),  # bounces: the keyed RNG makes shards order-independent
    ("box", R.Config(integrator=R.IntegratorStrategy.Raytracer, max_depth=2), 64, 5),  # fewer tiles than shards
])
def test_multi_device_render_equals_single_device(cuda_lib, single, devices, scene, cfg, w, h, monkeypatch):
    devs = devices if devices is not None else []  # []: every visible device
    multi = R.Context(devices=devs)
    assert multi.device_count == (len(devs) if devs else device_count(cuda_lib))
    desc = load_desc(scene)
    opts = M.rayca_options(0x5EED, 0, 0, 0, 0, 0, 0)
    want, _, wst = R.DeviceScene(single, desc).render(cfg, w, h, options=opts)
    ms = R.DeviceScene(multi, desc)
    # pageable caller buffer: one strided copy per device
    got, _, st = ms.render(cfg, w, h, options=opts)
    assert np.array_equal(got, want), f"pageable: {(got != want).sum()} bytes differ"
    assert (st.rays_primary, st.rays_shadow, st.rays_secondary) == (wst.rays_primary, wst.rays_shadow, wst.rays_secondary)
    if multi.device_count > 1:  # a 1-device context takes the ordinary path (row bands for pageable buffers)
        assert st.kernel_launches >= multi.device_count  # more than one per device when a big bounce frame runs as wavefront stages
    # pinned caller buffer: every device's kernel stores its tiles straight into it
    img = R.Image(w, h, pinned=True)
    img.buffer[...] = 7
    ms.render(cfg, w, h, options=opts, out=img.buffer)
    assert np.array_equal(img.buffer, want), f"pinned: {(img.buffer != want).sum()} bytes differ"
    # parity planes on a multi-device context come from its first device
    _, aux, _ = ms.render(cfg, w, h, options=opts, want_aux=True)
    _, waux, _ = R.DeviceScene(single, desc).render(cfg, w, h, options=opts, want_aux=True)
    assert np.array_equal(aux.hit_src, waux.hit_src) and np.array_equal(aux.depth, waux.depth)
    ms.close()
    multi.close()


def test_renderer_draw_fans_out(cuda_lib, single):
    """`CudaRenderer(devices=...).draw(scene, image)`: the drop-in call reaches every shard."""
    from rayca_b200 import procedural as P

    scene = P.atrium(20_000)
    a, b = R.Image(320, 180), R.Image(320, 180)
    R.CudaRenderer(R.Config(max_depth=1)).draw(scene, a)
    R.CudaRenderer(R.Config(max_depth=1), devices=[0, 0]).draw(scene, b)
    assert a.buffer.any() and np.array_equal(a.buffer, b.buffer)


def test_bad_device_lists_are_refused(cuda_lib):
    with pytest.raises(R.RaycaError) as e:
        R.Context(devices=[0, 99])
    assert e.value.code == -1
    with pytest.raises(R.RaycaError):
        R.Context(devices=[0] * 33)
