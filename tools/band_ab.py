#!/usr/bin/env python
"""A/B of the row-band split of rayca_cuda_render (RAYCA_BANDS / RAYCA_BAND_TAPER): ms per frame with the frame
landing in pinned host memory, bench workload."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from rayca_b200 import model as M, renderer as R

args = bench.parse_args()
scene, _ = bench.make_scene(args)
cfg = bench.render_config(args)
ctx = R.Context(0)
dev = R.DeviceScene(ctx, scene.to_desc())
img = R.Image(args.width, args.height, pinned=True)
for bands, taper in [(1, 0), (2, 0), (4, 0), (8, 0), (3, 1), (4, 1), (6, 1), (8, 1), (4, 0)]:
    os.environ["RAYCA_BANDS"], os.environ["RAYCA_BAND_TAPER"] = str(bands), str(taper)
    for _ in range(5):
        dev.render(cfg, args.width, args.height, out=img.buffer)
    best = []
    for rep in range(3):
        t = time.perf_counter()
        for _ in range(40):
            dev.render(cfg, args.width, args.height, out=img.buffer)
        best.append((time.perf_counter() - t) / 40 * 1e3)
    print(f"bands {bands} taper {taper}: {min(best):.4f} ms (median {sorted(best)[1]:.4f})", flush=True)
