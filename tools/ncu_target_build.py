#!/usr/bin/env python3
"""A short launch sequence of the device builder (K8 / K9) for ncu: python tools/ncu_target_build.py [scene] [depth] [n] [build flags]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtload
rt = rtload.load()
name = sys.argv[1] if len(sys.argv) > 1 else "monkey"
depth = int(sys.argv[2]) if len(sys.argv) > 2 else 15
n = int(sys.argv[3]) if len(sys.argv) > 3 else 0
flags = int(sys.argv[4]) if len(sys.argv) > 4 else 0     # RT_BUILD_* (2: no single block, 8: no cluster)
sc = rt.Scene(name, 64, 48, n_override=n)
tr = rt.RayTracer(64, 48, flags=0)
tr.upload_scene(sc.light, sc.camera, sc.shapes)
for _ in range(4):
    tr.build_bvh(depth, flags, fetch=False)
print("built", name, tr.last_build_ms(), "ms")
tr.close()
