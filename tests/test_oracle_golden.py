"""The oracle must keep reproducing the committed golden vectors (tests/golden/make_golden.py)."""
import numpy as np
import pytest

from conftest import case_config, golden_cases, load_case, load_desc
from rayca_b200 import model as M

_scenes = {}


def _oracle_scene(oracle, scene, bvh):
    key = (scene, bvh)
    if key not in _scenes:
        _scenes[key] = oracle.OracleScene(load_desc(scene), bvh=bvh)
    return _scenes[key]


@pytest.mark.parametrize("scene,case", golden_cases())
def test_oracle_matches_golden(oracle, scene, case):
    z = load_case(scene, case)
    cfg = case_config(z)
    h, w = z["hit_src"].shape
    if scene == "scene7":
        h, w = 24, 32  # bounded: exhaustive traversal of 100k triangles is slow on CPU
    osc = _oracle_scene(oracle, scene, cfg.bvh)
    opts = M.rayca_options(int(z["seed"][0]), 0, 0, 0, 0, 0, 0)
    if scene == "scene7":
        rows = None
        rgba8, aux, st = osc.render(cfg.to_c(), 80, 60, options=opts, rows=(20, 26))
        sl = slice(20, 26)
        assert np.array_equal(aux.hit_src.reshape(60, 80)[sl], z["hit_src"][sl])
        assert np.array_equal(aux.depth.reshape(60, 80)[sl], z["depth"][sl])
        assert np.array_equal(rgba8[sl], z["rgba8"][sl])
        return
    rgba8, aux, st = osc.render(cfg.to_c(), w, h, options=opts)
    assert np.array_equal(aux.hit_src.reshape(h, w), z["hit_src"])
    assert np.array_equal(aux.hit_prim.reshape(h, w), z["hit_prim"])
    assert np.array_equal(aux.hit_blas.reshape(h, w), z["hit_blas"])
    assert np.array_equal(aux.depth.reshape(h, w), z["depth"])
    assert np.array_equal(aux.uv.reshape(h, w, 2), z["uv"])
    np.testing.assert_allclose(aux.rgba_f32.reshape(h, w, 4), z["rgba_f32"], rtol=1e-6, atol=1e-7, equal_nan=True)
    assert np.array_equal(rgba8, z["rgba8"])
    assert [st.rays_primary, st.rays_shadow, st.rays_secondary] == [int(x) for x in z["rays"]]


def test_faithful_mode_is_bit_identical(oracle):
    """Re-transforming vertices on every test (what the reference does) gives the same bits as the cache."""
    z = load_case("box", "pt1")
    cfg = case_config(z)
    osc = _oracle_scene(oracle, "box", True)
    a8, a, _ = osc.render(cfg.to_c(), 64, 64)
    b8, b, _ = osc.render(cfg.to_c(), 64, 64, flags=oracle.FAITHFUL)
    assert np.array_equal(a8, b8) and np.array_equal(a.depth, b.depth) and np.array_equal(a.rgba_f32, b.rgba_f32)


def test_image_conventions(oracle):
    """Row-major, top-left origin (image.rs:29-31); alpha 255 everywhere (Color::BLACK alpha survives +=)."""
    z = load_case("unit_triangle", "pt1")
    assert np.all(z["rgba8"][..., 3] == 255)
    hit_rows = np.where((z["hit_src"] != 0xFFFFFFFF).any(axis=1))[0]
    # the triangle spans y in [-1, 1] world with its apex up: wider at the bottom rows of the image
    top, bottom = hit_rows[0], hit_rows[-1]
    assert (z["hit_src"][top] != 0xFFFFFFFF).sum() < (z["hit_src"][bottom] != 0xFFFFFFFF).sum()
