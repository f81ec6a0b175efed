#!/usr/bin/env python
"""Per-rank kernel time of an N-GPU run, measured on ONE GPU: renders shard r of K (the tiles rank r of a K-GPU run owns,
rayca_options.shard_index / shard_count) for every r and prints the kernel time of each — the spread is the load imbalance
between ranks, the ratio to (whole frame / K) is what the tail of a short launch costs.

    [RAYCA_CUDA_LIB=rayca_b200/variants/x.so] python tools/shard_ab.py [--shards 8] [--reps 30] [--tile 8] [--max-depth 1]
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shards", type=int, default=8)
    ap.add_argument("--reps", type=int, default=30)
    ap.add_argument("--tile", type=int, default=8)
    ap.add_argument("--max-depth", type=int, default=1)
    ap.add_argument("--flags", type=int, default=0)
    ap.add_argument("--no-flush", action="store_true", help="leave the L2 warm between timed launches")
    a = ap.parse_args()
    import torch

    from rayca_b200 import build, model as M, procedural as P, renderer as R

    build.build()
    W, H = 1920, 1080
    ctx = R.Context(0)
    dev = R.DeviceScene(ctx, P.atrium(262_267).to_desc())
    cfg = R.Config(max_depth=a.max_depth)
    buf = torch.zeros((H + 64, W, 4), dtype=torch.uint8, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream

    def time_shard(r, K):
        fl = a.flags | (R.FLAG_DIRECT_ROWS if (a.flags & R.FLAG_SHARD_GROUPS) and K > 1 else 0)
        opts = M.rayca_options(0x5EED, 0, r, K, a.tile, fl if K > 1 else (a.flags & ~R.FLAG_SHARD_GROUPS), 0)
        for _ in range(3):
            dev.render_device(cfg, W, H, buf.data_ptr(), stream=stream, options=opts)
        ms = []
        for k in range(a.reps):
            if not a.no_flush:
                flush.fill_(k & 0xFF)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            dev.render_device(cfg, W, H, buf.data_ptr(), stream=stream, options=opts)
            e1.record()
            torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        return float(np.median(ms))

    whole = time_shard(0, 1)
    per = [time_shard(r, a.shards) for r in range(a.shards)]
    ideal = whole / a.shards
    print(f"{'warm L2 ' if a.no_flush else ''}lib {os.environ.get('RAYCA_CUDA_LIB', 'product')} tile {a.tile} depth {a.max_depth}: whole frame {whole:.4f} ms, ideal per rank {ideal:.4f} ms; "
          f"shards of {a.shards}: min {min(per):.4f} mean {np.mean(per):.4f} max {max(per):.4f} ms -> efficiency {ideal / max(per):.3f} "
          f"(imbalance max/mean {max(per) / np.mean(per):.3f}, tail mean/ideal {np.mean(per) / ideal:.3f})")


if __name__ == "__main__":
    main()
