#!/usr/bin/env python3
"""C4 probe on one GPU: python tools/soup_probe.py [n W H]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtload
rt = rtload.load()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
W = int(sys.argv[2]) if len(sys.argv) > 2 else 3840
H = int(sys.argv[3]) if len(sys.argv) > 3 else 2160
t = time.time(); sc = rt.Scene("soup", W, H, n_override=n); print("host scene %.2f s" % (time.time() - t), sc.bvh_stats())
tr = rt.RayTracer(W, H, flags=0)
t = time.time(); tr.upload(sc); tr.synchronize(); print("first upload %.3f s" % (time.time() - t), tr.timings())
p = rt.make_params(W, H)
for i in range(3):
    t = time.time(); tr.render_device(p); tr.synchronize(); print("render %.3f ms wall" % ((time.time() - t) * 1e3), tr.timings()["render_ms"])
st = tr.render_stats(p); print(st); print(tr.walk_stats())
print("Mrays/s %.1f" % (st["rays"] / tr.timings()["render_ms"] / 1e3))
t = time.time(); tr.upload(sc); tr.synchronize(); print("re-upload (same topology) %.3f s" % (time.time() - t), tr.timings())
tr.close()
