#!/usr/bin/env python3
"""BASELINE configs[4] loop (animated monkey, pipelined rt_render_async) with and without the deferred host copy, beside the
device-only loop: python tools/c5_probe.py [frames]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import json, torch, rtload
rt = rtload.load()
W, H = 800, 600
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 300
sc = rt.Scene("monkey", W, H); p = rt.make_params(W, H)
bufs = [torch.empty((H * W, 4), dtype=torch.float32, pin_memory=True).numpy() for _ in range(2)]
dev = torch.empty((H * W, 4), dtype=torch.float32, device="cuda")
out = {}
for label, flags in (("deferred_copy", 0), ("copy_at_render_end", rt.RT_FLAG_NO_DEFERRED_COPY)):
    tr = rt.RayTracer(W, H, flags=flags); tr.upload(sc); tr.register_animations(sc)
    def step(kind, i):
        if kind == "animate+async":
            tr.wait_frames(1); tr.animate(i / 60.0, 1 / 60.0); tr.render_async(p, bufs[i & 1])
        elif kind == "async":
            tr.wait_frames(1); tr.render_async(p, bufs[i & 1])
        elif kind == "animate+device":
            tr.animate(i / 60.0, 1 / 60.0); tr.render_device(p, dev.data_ptr())
        else:
            tr.render_device(p, dev.data_ptr())
    for kind in ("animate+async", "async", "animate+device", "device"):
        best = 1e9
        for rep in range(3):
            for i in range(10): step(kind, i)
            tr.synchronize(); t0 = time.perf_counter()
            for i in range(10, 10 + frames): step(kind, i)
            tr.synchronize(); best = min(best, (time.perf_counter() - t0) / frames * 1e3)
        out["%s/%s" % (label, kind)] = round(best, 4)
    tr.close()
print(json.dumps(out, indent=1))
