"""CPU-side checks of the product's host code: the C-ABI library loads and exports every symbol
include/rayca_cuda.h declares, POD struct sizes agree with the ctypes mirror, and the binned BLAS
builder reproduces the reference builder (as restated verbatim by the oracle) node for node."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, golden_scenes, load_desc
from rayca_b200 import model as M
from rayca_b200 import renderer as R


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "rayca_cuda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rayca_cuda_[a-z_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(cuda_lib):
    names = declared_symbols()
    assert len(names) >= 14
    for n in names:
        assert hasattr(cuda_lib, n), f"{n} is declared in include/rayca_cuda.h but not exported"
    assert cuda_lib.rayca_cuda_abi_version() == 1
    assert cuda_lib.rayca_cuda_strerror(-2) == b"unsupported configuration"


def test_pod_struct_sizes():
    assert C.sizeof(M.rayca_trs) == 40
    assert C.sizeof(M.rayca_node) == 64
    assert C.sizeof(M.rayca_geometry) == 36
    assert C.sizeof(M.rayca_config) == 40
    assert C.sizeof(M.rayca_options) == 32
    assert C.sizeof(M.rayca_light) == 64
    assert C.sizeof(M.rayca_model) == 40 + 8 + 12 * 8
    assert M.VERTEX_FLOATS * 4 == 72


def test_invalid_arguments_return_codes(cuda_lib):
    h = C.c_void_p()
    assert cuda_lib.rayca_cuda_scene_create_host(None, 1, C.byref(h), None) == -1
    # a scene with no geometry: Tlas::new would assert (tlas.rs:75) -> INVALID_ARG, never a crash
    s = M.Scene()
    s.push_model(M.create_default_model())
    with pytest.raises(R.RaycaError) as e:
        R.DeviceScene(None, s.to_desc())
    assert e.value.code == -1


@pytest.mark.parametrize("scene", golden_scenes())
@pytest.mark.parametrize("bvh", [True, False])
def test_builder_matches_reference_builder(oracle, cuda_lib, scene, bvh):
    if scene == "scene7" and not bvh:
        pytest.skip("covered by the bvh=True case")
    desc = load_desc(scene)
    osc = oracle.OracleScene(desc, bvh=bvh)
    psc = R.DeviceScene(None, desc, bvh=bvh)
    oi, on, ono, opo, osrc = osc.bvh()
    pi, pn, pno, ppo, psrc = psc.bvh()
    assert (oi.blas_count, oi.node_count, oi.triangle_count) == (pi.blas_count, pi.node_count, pi.triangle_count)
    assert np.array_equal(ono, pno) and np.array_equal(opo, ppo)
    assert np.array_equal(on.view(np.uint32), pn.view(np.uint32)), "node boxes / links differ"
    assert np.array_equal(osrc, psrc), "primitive order differs"
    if not bvh:
        assert oi.node_count == 2 * oi.blas_count  # max_depth(0): every BLAS root stays a leaf


def test_builder_on_random_soup(oracle, cuda_lib):
    rng = np.random.default_rng(7)
    n = 3000
    c = rng.uniform(-5, 5, (n, 1, 3))
    tris = (c + rng.normal(0, 0.3, (n, 3, 3))).astype(np.float32)
    tris[:50] = tris[0]  # duplicates: identical centroids must not break the partition
    s = M.Scene()
    s.push_model(M.triangle_soup_model(tris))
    s.push_model(M.create_default_model())
    d = s.to_desc()
    oi, on, _, _, osrc = oracle.OracleScene(d).bvh()
    pi, pn, _, _, psrc = R.DeviceScene(None, d).bvh()
    assert oi.node_count == pi.node_count
    assert np.array_equal(on.view(np.uint32), pn.view(np.uint32)) and np.array_equal(osrc, psrc)


def test_product_does_not_import_the_oracle():
    """The oracle is test infrastructure: nothing under rayca_b200/ may import, link or execute it."""
    for root, _, files in os.walk(os.path.join(ROOT, "rayca_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(root, f)).read()
                for needle in ("import oracle", "from oracle", "librayca_oracle", "rayca_oracle_", "oracle_math"):
                    assert needle not in text, f"{f} references the oracle ({needle})"
