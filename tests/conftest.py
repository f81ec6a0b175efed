import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def golden_cases():
    """[(scene, case)] for every committed tests/golden/<scene>.<case>.npz."""
    out = []
    for f in sorted(os.listdir(GOLDEN)):
        if f.endswith(".npz") and not f.endswith(".scene.npz"):
            scene, case = f[:-4].split(".", 1)
            out.append((scene, case))
    return out


def golden_scenes():
    return sorted({s for s, _ in golden_cases()})


_desc_cache = {}


def load_desc(scene):
    from rayca_b200 import model as M

    if scene not in _desc_cache:
        _desc_cache[scene] = M.SceneDesc.load(os.path.join(GOLDEN, f"{scene}.scene.npz"))
    return _desc_cache[scene]


def load_case(scene, case):
    return np.load(os.path.join(GOLDEN, f"{scene}.{case}.npz"))


def case_config(z):
    """Rebuilds the renderer Config a golden case was rendered with."""
    from rayca_b200.renderer import Config, IntegratorStrategy, SamplerStrategy

    c = [int(x) for x in z["config"]]
    return Config(bvh=bool(c[0]), light_samples=c[1], light_stratify=bool(c[2]), samples_per_pixel=c[3],
                  russian_roulette=bool(c[4]), direct_sampler=SamplerStrategy(c[5]),
                  indirect_sampler=SamplerStrategy(c[6]), integrator=IntegratorStrategy(c[7]), max_depth=c[8],
                  gamma=float(z["gamma"][0]))


@pytest.fixture(scope="session")
def oracle():
    from oracle import binding

    binding.lib()
    return binding


@pytest.fixture(scope="session")
def cuda_lib():
    from rayca_b200 import build, renderer

    build.build()
    return renderer.load_library()
