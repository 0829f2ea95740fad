#!/usr/bin/env python3
"""A few frames of the animated monkey (BASELINE configs[4], device animation: K6 K7 K4 K2, the queue order, K1) for a launch list:
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv python tools/ncu_target_anim.py [frames]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, rtload
rt = rtload.load()
W, H = 800, 600
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 6
sc = rt.Scene("monkey", W, H); p = rt.make_params(W, H)
tr = rt.RayTracer(W, H, flags=0); tr.upload(sc); tr.register_animations(sc)
bufs = [torch.empty((H * W, 4), dtype=torch.float32, pin_memory=True).numpy() for _ in range(2)]
for i in range(frames):
    tr.wait_frames(1); tr.animate(i / 60.0, 1 / 60.0); tr.render_async(p, bufs[i & 1])
tr.synchronize()
print("done")
tr.close()
