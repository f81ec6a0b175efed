#!/usr/bin/env python
"""Prints the host phases of rayca_cuda_scene_create (RAYCA_TIMING) for a bench workload, twice (cold, warm)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["RAYCA_TIMING"] = "1"
import bench
from rayca_b200 import renderer as R

args = bench.parse_args()
scene, what = bench.make_scene(args)
ctx = None
try:
    import torch

    if torch.cuda.is_available():
        ctx = R.Context(0)
except Exception:
    pass
for rep in range(2):
    t = time.perf_counter()
    desc = scene.to_desc()
    t1 = time.perf_counter()
    dev = R.DeviceScene(ctx, desc, bvh=True)
    t2 = time.perf_counter()
    print(f"[{what}] to_desc {1e3 * (t1 - t):.1f} ms, scene_create {1e3 * (t2 - t1):.1f} ms ({'device' if ctx else 'host only'})", flush=True)
    dev.close()
