"""N>1 host logic on CPU: two gloo ranks fill their shard buffers, all_gather, reassemble."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from rayca_b200 import sharding


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, height, width, tile, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    _, per, local_rows = sharding.shard_layout(height, tile, world)
    local = torch.zeros((local_rows, width, 4), dtype=torch.uint8)
    for i, y in enumerate(sharding.owned_rows(height, tile, world, rank)):
        if y >= 0:  # stand-in for the render: encode (row, column) into the pixel
            local[i, :, 0] = y % 256
            local[i, :, 1] = y // 256
            local[i, :, 2] = torch.arange(width) % 256
            local[i, :, 3] = 255
    gathered = [torch.zeros_like(local) for _ in range(world)]
    dist.all_gather(gathered, local)
    frame = sharding.reassemble(torch.stack(gathered), height, tile)
    if rank == 0:
        np.save(out_path, frame.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("height,tile", [(100, 8), (1080, 8), (37, 16)])
def test_two_rank_gather_reassembles_the_frame(tmp_path, height, tile):
    world, width = 2, 24
    out = str(tmp_path / "frame.npy")
    mp.spawn(_worker, args=(world, _free_port(), height, width, tile, out), nprocs=world, join=True)
    frame = np.load(out)
    assert frame.shape == (height, width, 4)
    ys = np.arange(height)
    assert np.array_equal(frame[:, 0, 0], ys % 256) and np.array_equal(frame[:, 0, 1], ys // 256)
    assert np.all(frame[..., 3] == 255)


def test_layout_covers_every_row_once():
    for height, tile, world in [(1080, 8, 8), (1080, 8, 3), (2160, 8, 8), (5, 8, 4), (64, 16, 2)]:
        seen = []
        for r in range(world):
            seen += [y for y in sharding.owned_rows(height, tile, world, r) if y >= 0]
        assert sorted(seen) == list(range(height))
