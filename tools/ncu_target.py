#!/usr/bin/env python3
"""A short, deterministic launch sequence of the render kernel for ncu (profiling recipe: keep the command short, 1 GPU).

    python tools/ncu_target.py [--scene car] [--width 3840] [--height 2160] [--world 1] [--rank 0] [--frames 6] [--n 0]

Uploads the scene and renders `--frames` frames into device memory through the C ABI (rt_render_device), as the rank
`--rank` of `--world` tile-interleaved ranks (scene replicated, BASELINE configs[2]).  From the second frame on the
heavy-first order of the previous frame's costs is in effect, as in the timed loop of bench.py — profile a late launch
(ncu -k regex:render_kernel -s 3 -c 1)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtload  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--scene", default="car")
ap.add_argument("--width", type=int, default=3840)
ap.add_argument("--height", type=int, default=2160)
ap.add_argument("--world", type=int, default=1)
ap.add_argument("--rank", type=int, default=0)
ap.add_argument("--frames", type=int, default=6)
ap.add_argument("--n", type=int, default=0, help="primitive count override (soup)")
ap.add_argument("--flags", type=int, default=0)
a = ap.parse_args()
rt = rtload.load()
sc = rt.Scene(a.scene, a.width, a.height, n_override=a.n)
p = rt.make_params(a.width, a.height)
tr = rt.RayTracer(a.width, a.height, device=0, flags=a.flags)
tr.set_shard(a.rank, a.world)
tr.upload(sc)
for _ in range(a.frames):
    tr.render_device(p)
    tr.synchronize()
print("rendered %d frames of %s %dx%d as rank %d/%d, %d launches" % (a.frames, a.scene, a.width, a.height, a.rank, a.world, tr.launch_count()))
tr.close()
