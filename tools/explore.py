#!/usr/bin/env python3
"""Diagnostics on a GPU box: work distribution per pixel / per warp block and kernel time."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rtload
rt = rtload.load()

def warp_blocks(cost, H, W):
    h8, w8 = H // 4 * 4, W // 8 * 8
    c = cost[:h8, :w8].reshape(h8 // 4, 4, w8 // 8, 8)
    return c.max(axis=(1, 3)), c.sum(axis=(1, 3))

for name, W, H in [("car", 800, 600), ("car", 3840, 2160), ("monkey", 800, 600)]:
    sc = rt.Scene(name, W, H)
    p = rt.make_params(W, H)
    tr = rt.RayTracer(W, H, flags=0)
    tr.upload(sc)
    cost = np.zeros(W * H, np.int32)
    st = tr.render_stats(p, cost)
    cost = cost.reshape(H, W)
    mx, sm = warp_blocks(cost, H, W)
    for _ in range(3): tr.render_device(p)
    tr.synchronize()
    ts = []
    for _ in range(10):
        tr.render_device(p); ts.append(tr.timings()["render_ms"])
    print(name, W, H, "render ms min/med %.3f %.3f" % (min(ts), sorted(ts)[5]), "rays", st["rays"],
          "steps/pixel mean %.0f max %d p99 %.0f" % (cost.mean(), cost.max(), np.percentile(cost, 99)),
          "| warp-block max-lane steps: mean %.0f max %d p99 %.0f; SIMT efficiency sum/(32*max) %.2f" %
          (mx.mean(), mx.max(), np.percentile(mx, 99), sm.sum() / (32.0 * mx.sum())))
    tr.close()
