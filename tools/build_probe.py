"""Times K8 (rt_build_bvh) against the host builder on the same shapes (no torch needed): device time by CUDA
events, wall time of the call, host builder wall time, byte equality.  Usage: python tools/build_probe.py [soup_n [scene,scene...]]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtload  # noqa: E402

rt = rtload.load()
soup_n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
only = sys.argv[2].split(",") if len(sys.argv) > 2 else None
out = {}
for name, n, depth in (("monkey", 0, 15), ("car", 0, 25), ("soup", soup_n, 40)):
    if only and name not in only:
        continue
    sc = rt.Scene(name, 800, 600, n_override=n)
    t0 = time.perf_counter()
    sc.animate(0.0, 0.0, 1)                       # buildBVH + serializeBVH on the host
    t_host = time.perf_counter() - t0
    tr = rt.RayTracer(64, 48)
    tr.upload_scene(sc.light, sc.camera, sc.shapes)
    res = {}
    for flags, label in ((0, "default"), (2, "per_phase_launches"), (3, "per_phase_launches_plain_atomics")):
        dev, wall = [], []
        for _ in range(3):
            t0 = time.perf_counter()
            nn, ni = tr.build_bvh(depth, flags, fetch=False)
            wall.append((time.perf_counter() - t0) * 1e3)
            dev.append(tr.last_build_ms())
        nodes, idx = tr.build_bvh(depth, flags)
        res[label] = {"device_ms": min(dev), "call_wall_ms": min(wall),
                      "bytes_equal_host_builder": nodes.tobytes() == sc.nodes.tobytes() and idx.tobytes() == sc.indices.tobytes()}
    tr.close()
    out["%s_%d" % (name, len(sc.shapes))] = dict(shapes=len(sc.shapes), nodes=int(nn), maxDepth=depth, host_builder_ms=t_host * 1e3, **res)
print(json.dumps(out, indent=1))
