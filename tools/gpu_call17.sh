#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_build.py tests/test_gpu_upstream.py -m gpu -x -q > gpurun_out/gpu_tests17.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/gpu_tests17.log
timeout 300 python - <<'PY'
import sys, json, torch
sys.path.insert(0, '.')
import rtload
rt = rtload.load()
for name, depth in (("toy", 8), ("monkey", 15), ("car", 25)):
    sc = rt.Scene(name, 64, 48)
    tr = rt.RayTracer(64, 48, flags=0)
    tr.upload_scene(sc.light, sc.camera, sc.shapes)
    ms = []
    for _ in range(5):
        tr.build_bvh(depth, fetch=False); ms.append(tr.last_build_ms())
    print(name, "K8 device ms: %.4f (best of 5: %s)" % (min(ms), ["%.3f" % m for m in ms]))
    tr.close()
PY
