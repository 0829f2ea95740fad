#!/usr/bin/env python3
"""Host-side layout build (rt_layout.h) timed per phase on this machine's CPU: python tools/layout_probe.py [scene ...]
Compiles tools/layout_probe.cpp with g++ into gpurun_out/ (scratch)."""
import ctypes as C, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, rtload
rt = rtload.load()
out_dir = os.path.join(ROOT, "gpurun_out"); os.makedirs(out_dir, exist_ok=True)
so = os.path.join(out_dir, "liblayout_probe.so")
subprocess.check_call(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-pthread", "-I", os.path.join(ROOT, "include"),
                       "-I", os.path.join(ROOT, "opengl-ray-tracer_b200", "csrc"), os.path.join(ROOT, "tools", "layout_probe.cpp"), "-o", so])
L = C.CDLL(so)
for spec in (sys.argv[1:] or ["monkey", "car", "car:12"]):
    name, _, depth = spec.partition(":")
    sc = rt.Scene(name, 800, 600, **({"bvh_depth": int(depth)} if depth else {}))
    nd = np.ascontiguousarray(sc.nodes); ix = np.ascontiguousarray(sc.indices); sh = np.ascontiguousarray(sc.shapes)
    out = (C.c_double * 8)()
    L.time_layout(nd.ctypes.data_as(C.c_void_p), len(nd), ix.ctypes.data_as(C.c_void_p), len(ix), sh.ctypes.data_as(C.c_void_p), len(sh), C.c_double(1e4), out)
    print("%s: build_layout %.3f ms, build_top %.3f ms, refit graph %.3f ms; list A %d, list B %d, sub-tree nodes %d, SAH estimate A %.1f B %.1f" %
          (spec, out[0], out[1], out[2], out[4], out[5], out[6], out[7], out[3]))
