#!/usr/bin/env python3
"""Per-pixel work AOV of the CUDA path (nodes visited + primitive tests): where the heavy 8x4 blocks are."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, rtload
rt = rtload.load()
name = sys.argv[1] if len(sys.argv) > 1 else "car"
W = int(sys.argv[2]) if len(sys.argv) > 2 else 3840
H = int(sys.argv[3]) if len(sys.argv) > 3 else 2160
b = int(sys.argv[4]) if len(sys.argv) > 4 else 3
sc = rt.Scene(name, W, H)
tr = rt.RayTracer(W, H, flags=0)
tr.upload(sc)
cost = np.zeros(W * H, np.int32)
p = rt.make_params(W, H, maxBounces=b)
st = tr.render_stats(p, cost=cost)
cost = cost.reshape(H, W)
blk = cost[:H // 4 * 4, :W // 8 * 8].reshape(H // 4, 4, W // 8, 8)
bsum = blk.sum(axis=(1, 3)); bmax = blk.max(axis=(1, 3))
print("pixel cost: mean %.1f max %d ; block(8x4) sum: mean %.0f max %d ; block max-lane: mean %.1f max %d" % (cost.mean(), cost.max(), bsum.mean(), bsum.max(), bmax.mean(), bmax.max()))
ys, xs = np.unravel_index(np.argsort(bmax.ravel())[-5:], bmax.shape)
for y, x in zip(ys, xs):
    print("heavy block at px (%d,%d): lane costs" % (x * 8, y * 4), blk[y, :, x, :].ravel().tolist())
print(tr.item_stats())
tr.close()
