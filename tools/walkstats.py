#!/usr/bin/env python3
"""Walk-efficiency diagnostics on the GPU: python tools/walkstats.py [scene W H]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtload
rt = rtload.load()
name = sys.argv[1] if len(sys.argv) > 1 else "car"
W = int(sys.argv[2]) if len(sys.argv) > 2 else 3840
H = int(sys.argv[3]) if len(sys.argv) > 3 else 2160
n = int(sys.argv[4]) if len(sys.argv) > 4 else 0
sc = rt.Scene(name, W, H, n_override=n)
tr = rt.RayTracer(W, H, flags=0)
tr.upload(sc)
for b in (1, 3):
    p = rt.make_params(W, H, maxBounces=b)
    st = tr.render_stats(p)
    print("bounces", b, st)
    print("   ", tr.walk_stats())
    print("   ", tr.item_stats())
tr.close()
