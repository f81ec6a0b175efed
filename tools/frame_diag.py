#!/usr/bin/env python
"""Per-phase timing of the shared-frame (peer store) gather under torchrun, to see where an e2e step goes.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/frame_diag.py
Prints one line per variant on rank 0: wall ms per step, max over ranks."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

import bench
from rayca_b200 import model as M, renderer as R


def main():
    rank, G, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    sys.argv = [sys.argv[0]]
    args = bench.parse_args()
    W, H = args.width, args.height
    scene, _ = bench.make_scene(args)
    cfg = bench.render_config(args)
    ctx = R.Context(lr)
    dev = R.DeviceScene(ctx, scene.to_desc(), bvh=True)
    opts = M.rayca_options(0x5EED, 0, rank, G, 8, 0, 0)
    stream = torch.cuda.current_stream().cuda_stream
    frames = []
    for _ in range(2):
        h = [None]
        if rank == 0:
            frames.append(R.SharedFrame(ctx, W, H))
            h[0] = frames[-1].export()
        dist.broadcast_object_list(h, src=0)
        if rank != 0:
            frames.append(R.SharedFrame(ctx, W, H, handle=h[0]))
    srcs = [torch.as_tensor(f, device="cuda") for f in frames] if rank == 0 else None
    host = torch.empty((H, W, 4), dtype=torch.uint8).pin_memory()
    ids = [0, 0]

    def run(name, d2h, flow, two, host_barrier, n=60):
        k = [0]

        def one():
            b = (k[0] & 1) if two else 0
            k[0] += 1
            f = frames[b]
            if flow:
                f.acquire(ids[b], stream=stream)
            ids[b] += 1
            dev.render_to_frame(cfg, f, ids[b], stream=stream, options=opts)
            if rank == 0:
                f.wait(G, ids[b], stream=stream)
                if d2h:
                    host.copy_(srcs[b], non_blocking=True)
                f.release(ids[b], stream=stream)
            torch.cuda.synchronize()
            if host_barrier:
                dist.barrier()

        for _ in range(4):
            one()
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n):
            one()
        mine = (time.perf_counter() - t0) / n * 1e3
        dist.barrier()
        torch.cuda.synchronize()
        t = torch.tensor([mine], dtype=torch.float64, device="cuda")
        allt = [torch.zeros_like(t) for _ in range(G)]
        dist.all_gather(allt, t)
        if rank == 0:
            print(f"{name:46s} " + " ".join(f"{float(x):.3f}" for x in allt), flush=True)

    run("trace+wait", False, False, False, False)
    run("trace+wait+d2h", True, False, False, False)
    run("trace+wait+d2h, host barrier", True, False, False, True)
    run("acquire/release, one frame, no d2h", False, True, False, False)
    run("acquire/release, one frame, d2h", True, True, False, False)
    run("acquire/release, two frames, no d2h", False, True, True, False)
    run("acquire/release, two frames, d2h", True, True, True, False)
    if any(f.timed_out() for f in frames):
        print("TIMED OUT", rank, flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
