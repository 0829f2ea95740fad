"""Screen-tile sharding of one frame over the GPUs of a box.

Partition: RT_TILE_W x RT_TILE_H tiles, tile k -> rank k % world, where tile k sits in tile row k // tilesX at column
(k % tilesX + row) % tilesX (every row rotated by its number: a rank's tiles form diagonals, not the vertical stripes a plain
row-major numbering gives whenever tilesX is a multiple of world); the scene is replicated.  Each rank renders a compact tile-major shard (rt_set_shard); the one
exchange step is a gather of the shards to rank 0 (NCCL over NVLink via torch.distributed, gloo in
the CPU tests), after which K3 (rt_untile) restores row-major order.  No other collective.
"""
import numpy as np

TILE_W, TILE_H = 32, 8
TILE_PX = TILE_W * TILE_H


def tile_grid(W, H):
    return (W + TILE_W - 1) // TILE_W, (H + TILE_H - 1) // TILE_H


def shard_tiles(W, H, rank, world):
    """Tile indices owned by `rank`."""
    tx, ty = tile_grid(W, H)
    return list(range(rank, tx * ty, world))


def shard_pixels(W, H, world):
    """Pixels in every rank's compact buffer (padded to rank 0's tile count; == rt_shard_pixels)."""
    if world == 1:
        return W * H
    tx, ty = tile_grid(W, H)
    return ((tx * ty + world - 1) // world) * TILE_PX


def tile_origin(W, H, tile, world):
    """Pixel origin (x0, y0) of tile `tile` (include/rt_api.h: row-major for one GPU, every row rotated by its number when sharded)."""
    tx, _ = tile_grid(W, H)
    row, c0 = divmod(tile, tx)
    col = c0 if world == 1 else (c0 + row) % tx
    return col * TILE_W, row * TILE_H


def tile_pixels(W, H, tile, world):
    """(ys, xs) int arrays of the TILE_PX pixel coordinates of a tile in compact order (may exceed the image)."""
    x0, y0 = tile_origin(W, H, tile, world)
    ys, xs = np.divmod(np.arange(TILE_PX), TILE_W)
    return ys + y0, xs + x0


def untile_numpy(shards, W, H, world):
    """Host restatement of K3: shards [world, shard_pixels, C] -> [H, W, C]."""
    shards = np.asarray(shards)
    out = np.zeros((H, W) + shards.shape[2:], dtype=shards.dtype)
    for rank in range(world):
        for j, tile in enumerate(shard_tiles(W, H, rank, world)):
            ys, xs = tile_pixels(W, H, tile, world)
            ok = (ys < H) & (xs < W)
            out[ys[ok], xs[ok]] = shards[rank, j * TILE_PX:(j + 1) * TILE_PX][ok]
    return out


def tile_numpy(image, W, H, rank, world):
    """Inverse for one rank: [H, W, C] -> compact shard [shard_pixels, C] (padding pixels zero)."""
    image = np.asarray(image)
    out = np.zeros((shard_pixels(W, H, world),) + image.shape[2:], dtype=image.dtype)
    for j, tile in enumerate(shard_tiles(W, H, rank, world)):
        ys, xs = tile_pixels(W, H, tile, world)
        ok = (ys < H) & (xs < W)
        out[j * TILE_PX:(j + 1) * TILE_PX][ok] = image[ys[ok], xs[ok]]
    return out


def gather_shards(shard, rank, world, dst=0):
    """The exchange step: gather equal-sized shard tensors to `dst` with torch.distributed
    (backend nccl on the GPUs, gloo in the CPU tests).  Returns [world, ...] on dst, None elsewhere."""
    import torch
    import torch.distributed as dist
    if world == 1:
        return shard.unsqueeze(0)
    if rank == dst:
        out = torch.empty((world,) + tuple(shard.shape), dtype=shard.dtype, device=shard.device)
        dist.gather(shard, list(out.unbind(0)), dst=dst)
        return out
    dist.gather(shard, None, dst=dst)
    return None
