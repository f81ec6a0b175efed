#!/usr/bin/env python
"""One `render` call on a multi-device context (rayca_cuda_init_devices) against the same call on one device: what a
`CudaRenderer::draw` caller gets from a box with several GPUs without torchrun.  End to end, frame in HOST memory.

    python tools/multi_device_bench.py [--reps 30] [--max-depth 1]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=30)
    ap.add_argument("--max-depth", type=int, default=1)
    a = ap.parse_args()
    from rayca_b200 import build, model as M, procedural as P, renderer as R

    build.build()
    W, H = 1920, 1080
    desc = P.atrium(262_267).to_desc()
    cfg = R.Config(max_depth=a.max_depth)
    opts = M.rayca_options(0x5EED, 0, 0, 0, 0, 0, 0)
    out = {}
    ref = None
    for name, ctx in (("one_device", R.Context(0)), ("all_devices", R.Context(devices=[]))):
        t0 = time.perf_counter()
        dev = R.DeviceScene(ctx, desc)
        create_ms = (time.perf_counter() - t0) * 1e3
        res = {"devices": ctx.device_count, "scene_create_ms": create_ms}
        for kind in ("pinned", "pageable"):
            img = R.Image(W, H, pinned=(kind == "pinned"))
            for _ in range(5):
                _, _, st = dev.render(cfg, W, H, options=opts, out=img.buffer)
            t1 = time.perf_counter()
            for _ in range(a.reps):
                _, _, st = dev.render(cfg, W, H, options=opts, out=img.buffer)
            ms = (time.perf_counter() - t1) / a.reps * 1e3
            rays = st.rays_primary + st.rays_shadow + st.rays_secondary
            res[kind] = {"ms_per_frame": ms, "Mrays_per_s": rays / ms / 1e3, "kernel_ms_max_over_devices": st.trace_ms}
            if ref is None:
                ref = img.buffer.copy()
            res[kind]["equals_one_device_frame"] = bool(np.array_equal(img.buffer, ref))
        out[name] = res
        dev.close()
        ctx.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
