"""Properties of the screen-tile partition (shard.py is the host restatement of K1's tile arithmetic and of K3) at
random frame sizes and rank counts: every pixel belongs to exactly one rank, the compact shards untile to the frame, tiling is
the inverse, and a rank's buffer is as large as rt_shard_pixels says."""
import numpy as np
from hypothesis import given, settings, strategies as st


@settings(max_examples=60, deadline=None)
@given(W=st.integers(1, 700), H=st.integers(1, 300), world=st.integers(1, 9), seed=st.integers(0, 10 ** 6))
def test_tile_untile_roundtrip(rt, W, H, world, seed):
    sh = rt.shard
    img = np.random.default_rng(seed).integers(0, 2 ** 31, size=(H, W, 4), dtype=np.int64).astype(np.float32)
    px = sh.shard_pixels(W, H, world)
    tx, ty = sh.tile_grid(W, H)
    if world > 1:
        assert px == ((tx * ty + world - 1) // world) * sh.TILE_PX
    shards = np.stack([sh.tile_numpy(img, W, H, r, world).reshape(-1, 4)[:px] if world > 1 else img.reshape(-1, 4) for r in range(world)])
    if world == 1:
        assert px == W * H
        return
    assert shards.shape == (world, px, 4)
    back = sh.untile_numpy(shards, W, H, world)
    assert np.array_equal(back, img)
    seen = np.zeros((H, W), np.int32)
    for r in range(world):
        tiles = sh.shard_tiles(W, H, r, world)
        assert len(tiles) * sh.TILE_PX <= px
        for t in tiles:
            ys, xs = sh.tile_pixels(W, H, t, world)
            ok = (ys < H) & (xs < W)
            seen[ys[ok], xs[ok]] += 1
    assert (seen == 1).all()


@settings(max_examples=40, deadline=None)
@given(W=st.integers(33, 4000), H=st.integers(9, 2200), world=st.integers(2, 8))
def test_a_ranks_tiles_do_not_form_columns(rt, W, H, world):
    """Every tile row is rotated by its number (include/rt_api.h): when the row length is a multiple of the rank count, a rank's
    tiles in two consecutive rows sit in different columns (diagonals, not vertical stripes)."""
    sh = rt.shard
    tx, ty = sh.tile_grid(W, H)
    if ty < 2 or tx < 2 or tx % world != 0:
        return
    cols = [sorted(sh.tile_origin(W, H, t, world)[0] for t in sh.shard_tiles(W, H, 0, world) if t // tx == row) for row in (0, 1)]
    assert cols[0] != cols[1]
