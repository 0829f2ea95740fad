"""CPU, world_size 2 over gloo: the N>1 host path — interleaved-tile partition, the gather (the path's
one exchange step) and the untile — with each rank's shard produced by the oracle (checker only)."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H = 100, 60            # ragged: 4x8 tiles, edge tiles padded


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as O
    import rtload
    rt = rtload.load()
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sc = rt.Scene("toy", W, H)
    p = rt.make_params(W, H)
    full, *_ = O.render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, p, W, H, threads=1, want_counters=False)
    mine = torch.from_numpy(rt.shard.tile_numpy(full, W, H, rank, world))      # what this rank's GPU would render
    assert mine.shape[0] == rt.shard.shard_pixels(W, H, world)
    got = rt.shard.gather_shards(mine, rank, world)
    if rank == 0:
        img = rt.shard.untile_numpy(got.numpy(), W, H, world)
        q.put(bool(np.array_equal(img, full)))
    else:
        assert got is None
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gather_reassembles_the_frame():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok = q.get(timeout=120)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert ok


def test_partition_covers_every_pixel_once(rt=None):
    sys.path.insert(0, ROOT)
    import rtload
    sh = rtload.load().shard
    for world in (1, 2, 3, 4, 8):
        seen = np.zeros((H, W), int)
        total = 0
        for r in range(world):
            tiles = sh.shard_tiles(W, H, r, world)
            total += len(tiles)
            assert len(tiles) * sh.TILE_PX <= sh.shard_pixels(W, H, world) or world == 1
            for t in tiles:
                ys, xs = sh.tile_pixels(W, H, t, world)
                ok = (ys < H) & (xs < W)
                seen[ys[ok], xs[ok]] += 1
        assert (seen == 1).all() and total == 4 * 8
