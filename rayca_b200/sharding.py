"""Image-tile sharding across GPUs: row tiles of `tile_rows` rows, tile t -> rank t % G.

Each rank renders its tiles into a compact local buffer (rayca_cuda_render_device with
rayca_options.shard_*), the buffers are gathered rank-major with one NCCL all_gather, and
`reassemble` turns the gathered (G, per*tile, W, C) tensor back into the row-major frame — a pure
view permutation, no extra kernel.  The scene is replicated; there is no other exchange.
"""
from __future__ import annotations


def shard_layout(height: int, tile_rows: int, world: int):
    """Returns (tiles, tiles_per_rank, local_rows): every rank owns `tiles_per_rank` tile slots
    (tail slots of the last ranks may lie below the image and stay unwritten)."""
    tiles = (height + tile_rows - 1) // tile_rows
    per = (tiles + world - 1) // world
    return tiles, per, per * tile_rows


def owned_rows(height: int, tile_rows: int, world: int, rank: int):
    """Global row indices rank `rank` renders, in its local buffer order."""
    tiles, per, _ = shard_layout(height, tile_rows, world)
    rows = []
    for k in range(per):
        t = k * world + rank
        for r in range(tile_rows):
            y = t * tile_rows + r
            rows.append(y if (t < tiles and y < height) else -1)
    return rows


def reassemble(gathered, height: int, tile_rows: int):
    """gathered: tensor (G, per*tile_rows, W, C) from all_gather -> (height, W, C) frame."""
    G, local_rows, W, Cc = gathered.shape
    per = local_rows // tile_rows
    x = gathered.reshape(G, per, tile_rows, W, Cc).permute(1, 0, 2, 3, 4).reshape(per * G * tile_rows, W, Cc)
    return x[:height]
