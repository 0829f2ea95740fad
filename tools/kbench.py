#!/usr/bin/env python3
"""Kernel-variant timing (diagnostic, not a bench line): render-kernel ms of a few workloads for the library named by
RT_B200_LIB (default: the in-tree build), L2 flushed between frames, CUDA events on the launching stream, plus a hash of
every frame so that variants can be checked to be bit-identical.

    RT_B200_LIB=build_var/x.so python tools/kbench.py [--flags N] [--steps 20] [--cases car4k,car800,monkey800,soup4k]"""
import argparse
import hashlib
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import rtload  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--flags", type=int, default=0)
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--warmup", type=int, default=5)
ap.add_argument("--cases", default="car4k,car800,monkey800")
ap.add_argument("--world", type=int, default=1)
ap.add_argument("--rank", type=int, default=0)
ap.add_argument("--noflush", action="store_true", help="leave L2 as the previous frame left it (diagnostic: how much of a frame is cold misses)")
a = ap.parse_args()
rt = rtload.load()
CASES = {"car4k": ("car", 3840, 2160, 0), "car800": ("car", 800, 600, 0), "monkey800": ("monkey", 800, 600, 0),
         "soup4k": ("soup", 3840, 2160, 1000000), "soup800": ("soup", 800, 600, 1000000), "soup300k4k": ("soup", 3840, 2160, 300000), "soup60k4k": ("soup", 3840, 2160, 60000), "soup15k4k": ("soup", 3840, 2160, 15000), "car4k_mtfr": ("car", 3840, 2160, 0)}
out = {"lib": os.environ.get("RT_B200_LIB", "in-tree"), "flags": a.flags}
stream = torch.cuda.Stream()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
with torch.cuda.stream(stream):
    for case in a.cases.split(","):
        name, W, H, n = CASES[case]
        sc = rt.Scene(name, W, H, n_override=n)
        p = rt.make_params(W, H, useMollerTrumbore=case.endswith("mtfr"), useFresnel=case.endswith("mtfr"))
        tr = rt.RayTracer(W, H, device=0, flags=a.flags)
        tr.set_stream(stream.cuda_stream)
        tr.set_shard(a.rank, a.world)
        tr.upload(sc)
        px = tr.shard_pixels()
        img = torch.empty((px, 4), dtype=torch.float32, device="cuda")
        evs = []
        for i in range(a.warmup + a.steps):
            if not a.noflush:
                flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); tr.render_device(p, img.data_ptr()); e1.record()
            if i >= a.warmup:
                evs.append((e0, e1))
        torch.cuda.synchronize()
        ms = [x.elapsed_time(y) for x, y in evs]
        st = tr.render_stats(p)
        out[case] = {"ms": round(sum(ms) / len(ms), 4), "min": round(min(ms), 4), "hash": hashlib.md5(img.cpu().numpy().tobytes()).hexdigest()[:10],
                     "node_visits": st["node_visits"], "prim_tests": st["prim_tests"], "rays": st["rays"]}
        tr.close()
print(json.dumps(out))
