"""The frames bench.py times, compared with the LIVE oracle (VERDICT r01 "next" item 1): the CUDA product path through
the C ABI against oracle rows rendered on the same input in the same process.

  * 1080p atrium (BASELINE config 3 stand-in, config.workload of the bench line), max_depth 1 and 2: eight spread rows
    against the oracle's exhaustive reference walk, and the WHOLE frame against the oracle's optimised arm, which is
    itself checked against the exhaustive walk on those rows.
  * soup(200_000) at 16 spp (BASELINE config 4 at a size the oracle finishes in seconds): four rows.

Bars as everywhere (test_cuda_parity.py): primary hit ids, depth and barycentrics bit-exact; RGB within 1e-3 (stochastic
frames: on >= 99 % of the pixels, relative above 1.0)."""
import numpy as np
import pytest

from rayca_b200 import model as M
from rayca_b200 import procedural as P
from rayca_b200 import renderer as R

pytestmark = pytest.mark.gpu

W, H = 1920, 1080
ROWS = [int((k + 0.5) / 8 * H) for k in range(8)]
RGB_TOL = 1e-3


@pytest.fixture(scope="module")
def ctx(cuda_lib):
    c = R.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def atrium(ctx, oracle):
    desc = P.atrium(262_267).to_desc()
    dev = R.DeviceScene(ctx, desc)
    osc = oracle.OracleScene(desc)
    yield desc, dev, osc
    dev.close()
    osc.close()


def row_planes(aux, rgba8, y, w):
    sl = slice(y * w, (y + 1) * w)
    return dict(hit_src=aux.hit_src[sl], hit_prim=aux.hit_prim[sl], hit_blas=aux.hit_blas[sl], depth=aux.depth[sl],
                uv=aux.uv[2 * y * w:2 * (y + 1) * w], rgba_f32=aux.rgba_f32[4 * y * w:4 * (y + 1) * w].reshape(w, 4),
                rgba8=rgba8[y])


def compare(got, want, stochastic, what):
    for k in ("hit_src", "hit_prim", "hit_blas"):
        assert np.array_equal(got[k], want[k]), f"{what}: {(got[k] != want[k]).sum()} {k} values differ"
    assert np.array_equal(got["depth"].view(np.uint32), want["depth"].view(np.uint32)), f"{what}: depth bits differ"
    assert np.array_equal(got["uv"].view(np.uint32), want["uv"].view(np.uint32)), f"{what}: barycentric bits differ"
    a, b = got["rgba_f32"].reshape(-1, 4), want["rgba_f32"].reshape(-1, 4)
    assert np.array_equal(np.isnan(a), np.isnan(b)), f"{what}: NaN pattern differs"
    err = np.abs(np.nan_to_num(a) - np.nan_to_num(b))[:, :3]
    if stochastic:
        err = err / np.maximum(1.0, np.abs(np.nan_to_num(b))[:, :3])
        beyond = (err.max(axis=1) > RGB_TOL).mean()
        assert beyond <= 0.01, f"{what}: {beyond:.3%} of the pixels beyond {RGB_TOL}"
    else:
        assert err.max() <= RGB_TOL, f"{what}: max RGB error {err.max()}"
        d8 = np.abs(got["rgba8"].astype(np.int16) - want["rgba8"].astype(np.int16))
        assert d8.max() <= 1 and (d8 > 0).mean() <= 0.01, f"{what}: RGBA8 differs by more than one step / on > 1 % of the channels"


@pytest.mark.parametrize("max_depth", [1, 2])
def test_atrium_bench_frame_against_the_oracle(atrium, oracle, max_depth):
    desc, dev, osc = atrium
    cfg = R.Config(max_depth=max_depth)
    opts = M.rayca_options(0x5EED, 0, 0, 0, 0, 0, 0)  # bench.py's seed
    g8, gaux, gst = dev.render(cfg, W, H, options=opts, want_aux=True)
    stochastic = max_depth > 1
    # (1) eight spread rows against the exhaustive reference walk
    exh = {}
    for y in ROWS:
        o8, oaux, _ = osc.render(cfg.to_c(), W, H, options=opts, rows=(y, y + 1))
        exh[y] = row_planes(oaux, o8, y, W)
        exh[y] = {k: np.array(v) for k, v in exh[y].items()}
        compare(row_planes(gaux, g8, y, W), exh[y], stochastic, f"atrium max_depth={max_depth} row {y} vs exhaustive oracle")
    # (2) the whole frame against the optimised arm; that arm must itself equal the exhaustive walk on the rows
    f8, faux, fst = osc.render(cfg.to_c(), W, H, options=opts, flags=oracle.OPTIMISED)
    for y in ROWS:
        got = row_planes(faux, f8, y, W)
        for k in ("hit_src", "hit_prim", "hit_blas"):
            assert np.array_equal(got[k], exh[y][k])
        for k in ("depth", "uv", "rgba_f32"):
            assert np.array_equal(got[k].view(np.uint32), exh[y][k].view(np.uint32)), f"optimised arm != exhaustive walk ({k}, row {y})"
    whole = dict(hit_src=faux.hit_src, hit_prim=faux.hit_prim, hit_blas=faux.hit_blas, depth=faux.depth, uv=faux.uv,
                 rgba_f32=faux.rgba_f32, rgba8=f8)
    mine = dict(hit_src=gaux.hit_src, hit_prim=gaux.hit_prim, hit_blas=gaux.hit_blas, depth=gaux.depth, uv=gaux.uv,
                rgba_f32=gaux.rgba_f32, rgba8=g8)
    compare(mine, whole, stochastic, f"atrium max_depth={max_depth} whole frame vs optimised oracle arm")
    assert gst.rays_primary == fst.rays_primary == W * H
    if not stochastic:
        assert (gst.rays_shadow, gst.rays_secondary) == (fst.rays_shadow, fst.rays_secondary)


def test_soup_16spp_rows_against_the_oracle(ctx, oracle):
    """BASELINE config 4 (procedural soup, 16 spp = 4 x 4 strata, warp-cooperative sample loop) at 200,000 triangles and
    960 x 540: four rows against the exhaustive reference walk."""
    desc = P.soup(200_000).to_desc()
    dev = R.DeviceScene(ctx, desc)
    osc = oracle.OracleScene(desc)
    w, h = 960, 540
    cfg = R.Config(max_depth=1, samples_per_pixel=16)
    opts = M.rayca_options(0x5EED, 0, 0, 0, 0, 0, 0)
    g8, gaux, gst = dev.render(cfg, w, h, options=opts, want_aux=True)
    shadow = 0
    for y in [int((k + 0.5) / 4 * h) for k in range(4)]:
        o8, oaux, ost = osc.render(cfg.to_c(), w, h, options=opts, rows=(y, y + 1))
        compare(row_planes(gaux, g8, y, w), row_planes(oaux, o8, y, w), False, f"soup 16 spp row {y}")
        shadow += ost.rays_shadow
    assert gst.rays_primary == w * h * 16 and shadow > 0
    dev.close()
    osc.close()


def test_atrium_bounce_frame_wavefront_equals_fused(atrium, monkeypatch):
    """The 1080p bounce frame of the bench (BASELINE config 5) at full size: the wavefront stages — persistent ray-walk
    kernel with dynamic fetch over two million sorted paths, shading re-sorted by hit point — against the fused kernel,
    byte for byte and ray count for ray count; and the one-kernel-per-depth stage likewise."""
    desc, dev, osc = atrium
    cfg = R.Config(max_depth=2)
    fused8, _, fst = dev.render(cfg, W, H, options=M.rayca_options(0x5EED, 0, 0, 0, 0, R.FLAG_NO_WAVEFRONT, 0))
    wave8, _, wst = dev.render(cfg, W, H, options=M.rayca_options(0x5EED, 0, 0, 0, 0, 0, 0))
    assert wst.kernel_launches > fst.kernel_launches, "the frame is large enough for the stages: they must have run"
    assert np.array_equal(wave8, fused8), f"{(wave8 != fused8).sum()} bytes differ"
    assert (wst.rays_primary, wst.rays_shadow, wst.rays_secondary) == (fst.rays_primary, fst.rays_shadow, fst.rays_secondary)
    monkeypatch.setenv("RAYCA_WAVE_SPLIT", "0")
    one8, _, ost = dev.render(cfg, W, H, options=M.rayca_options(0x5EED, 0, 0, 0, 0, 0, 0))
    assert np.array_equal(one8, fused8)
    assert (ost.rays_primary, ost.rays_shadow, ost.rays_secondary) == (fst.rays_primary, fst.rays_shadow, fst.rays_secondary)
