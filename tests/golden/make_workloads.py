#!/usr/bin/env python3
"""Workload constants for bench.py: the REFERENCE ALGORITHM's deterministic per-frame work
 for the named configurations, counted by the oracle on the exact scene bytes.

    python tests/golden/make_workloads.py        # writes tests/golden/workloads.json

rays = intersectScene2-equivalent queries (primary + shadow + reflection); alg_bytes =
48*node_visits + 4*index_reads + per-type primitive bytes + 48*ray_hits.  bench.py divides these by
the measured render time; it cross-checks `rays` against the CUDA path's own counter at run time.
"""
import json
import os
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_py as O  # noqa: E402
import rtload  # noqa: E402

rt = rtload.load()
CASES = [("car", 800, 600, 0), ("car", 3840, 2160, 0), ("car", 800, 600, 1), ("car", 3840, 2160, 1),
         ("monkey", 800, 600, 0), ("monkey", 3840, 2160, 0)]


def crc(sc):
    c = 0
    for a in (sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices):
        c = zlib.crc32(np.ascontiguousarray(a).tobytes(), c)
    return c


def key(name, W, H, mt, fresnel=0, bounces=3):
    return "%s_%dx%d_mt%d_fr%d_b%d" % (name, W, H, mt, fresnel, bounces)


if __name__ == "__main__":
    out = {}
    for name, W, H, mt in CASES:
        sc = rt.Scene(name, W, H)
        p = rt.make_params(W, H, useMollerTrumbore=bool(mt))
        _, ids, _, c = O.render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, p, W, H)
        c.update(scene_crc=crc(sc), bvh=sc.bvh_stats(), pixels=W * H, hit_pixels=int((ids >= 0).sum()),
                 shapes=len(sc.shapes), scene_bytes=int(sc.shapes.nbytes + sc.nodes.nbytes + sc.indices.nbytes))
        out[key(name, W, H, mt)] = c
        print(key(name, W, H, mt), "rays/px %.3f  B/ray %.0f" % (c["rays"] / (W * H), c["alg_bytes"] / c["rays"]))
    json.dump(out, open(os.path.join(HERE, "workloads.json"), "w"), indent=1, sort_keys=True)
