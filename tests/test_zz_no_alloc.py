"""rayca_cuda_render_static — the allocation-free render entry point (scratch in a file-scope `static __device__` array
of the library; include/rayca_cuda.h).  It has no counterpart in the reference: it was added
for callers that cannot call cudaMalloc.

Collected LAST on purpose: a failure here must not hide any other test.  Passes alone
in a fresh process and after the other tests.  The new entry point runs the same kernel as rayca_cuda_render, so the
frames must be equal byte for byte — inside every tolerance the other parity tests use (RGB within 1e-3, ids bit-exact)."""
import numpy as np
import pytest

from conftest import load_desc
from rayca_b200 import model as M
from rayca_b200 import renderer as R

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(cuda_lib):
    c = R.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("scene,cfg,w,h", [
    ("box", R.Config(max_depth=1), 128, 128),
    ("suzanne", R.Config(max_depth=2), 333, 217),  # ragged size, textured, one stochastic bounce (keyed RNG: deterministic)
    ("cubes_over_plane", R.Config(integrator=R.IntegratorStrategy.Raytracer, max_depth=3), 1920, 1080),  # the largest frame that fits
])
def test_static_render_equals_render_on_two_calls_in_a_row(ctx, scene, cfg, w, h):
    dev = R.DeviceScene(ctx, load_desc(scene))
    opts = M.rayca_options(0x5EED, 0, 0, 0, 0, 0, 0)
    want, _, wst = dev.render(cfg, w, h, options=opts)
    for call in range(2):  # the array keeps the previous call's contents: the second call must not see them
        got, st = dev.render_static(cfg, w, h, options=opts)
        assert np.array_equal(got, want), f"call {call}: {(got != want).sum()} bytes differ from rayca_cuda_render"
        assert (st.rays_primary, st.rays_shadow, st.rays_secondary) == (wst.rays_primary, wst.rays_shadow, wst.rays_secondary), \
            f"call {call}: ray counters differ (the counters in the static array were not cleared?)"
        assert st.kernel_launches == 1 and st.trace_ms > 0.0
    # a different, smaller frame right after a big one: stale pixels of the previous call must not leak into it
    small_want, _, _ = dev.render(cfg, 64, 48, options=opts)
    small_got, _ = dev.render_static(cfg, 64, 48, options=opts)
    assert np.array_equal(small_got, small_want)
    dev.close()


def test_static_render_rejects_what_it_cannot_hold(ctx):
    dev = R.DeviceScene(ctx, load_desc("box"))
    cfg = R.Config(max_depth=1)
    assert 1920 * 1080 * 4 <= R.STATIC_FRAME_BYTES < 2048 * 1025 * 4
    with pytest.raises(R.RaycaError) as e:  # one row more than 8 MiB
        dev.render_static(cfg, 2048, 1025)
    assert e.value.code == -1 and "static scratch" in str(e.value) and str(R.STATIC_FRAME_BYTES) in str(e.value)
    with pytest.raises(R.RaycaError) as e:  # sharded launches are not offered by this entry point
        dev.render_static(cfg, 128, 128, options=M.rayca_options(0, 0, 0, 2, 8, R.FLAG_DIRECT_ROWS, 0))
    assert e.value.code == -1
    # the context still renders normally afterwards, through both entry points
    a, _, _ = dev.render(cfg, 128, 128)
    b, _ = dev.render_static(cfg, 128, 128)
    assert np.array_equal(a, b)
    dev.close()
