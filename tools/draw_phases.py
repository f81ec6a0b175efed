#!/usr/bin/env python
"""Times CudaRenderer.draw (the reference's Draw::draw: rebuild + render + read-back) a few times with the host
phases printed (RAYCA_TIMING)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["RAYCA_TIMING"] = "1"
import bench
from rayca_b200 import renderer as R

args = bench.parse_args()
scene, what = bench.make_scene(args)
cfg = bench.render_config(args)
rnd = R.CudaRenderer(cfg)
img = R.Image(args.width, args.height)
for rep in range(4):
    t = time.perf_counter()
    rnd.draw(scene, img)
    print(f"draw {rep}: {1e3 * (time.perf_counter() - t):.1f} ms", flush=True)
