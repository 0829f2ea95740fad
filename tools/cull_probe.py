#!/usr/bin/env python3
"""EXPERIMENT (CPU, estimate): node visits per frame with distance culling in list order / near-first order.
   python tools/cull_probe.py [scene W H [n_override]]"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, rtload
rt = rtload.load()
name = sys.argv[1] if len(sys.argv) > 1 else "car"
W = int(sys.argv[2]) if len(sys.argv) > 2 else 480
H = int(sys.argv[3]) if len(sys.argv) > 3 else 270
nov = int(sys.argv[4]) if len(sys.argv) > 4 else 0
sc = rt.Scene(name, W, H, n_override=nov)
sh, nd, ix, cam, li = sc.shapes, sc.nodes, sc.indices, sc.camera, sc.light
L = C.CDLL(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "build_var", "libcull_probe.so"))
vp, ip = C.c_void_p, C.c_int
L.emu_create.argtypes = [vp, ip, vp, ip, vp, ip, ip, C.c_double]; L.emu_create.restype = vp
L.probe_frame.argtypes = [vp, vp, ip, vp, vp, vp, ip, ip, vp]
p = lambda a: a.ctypes.data_as(vp)
root = max(1.0, float(np.abs(nd["boundsMin"][-1]).max()), float(np.abs(nd["boundsMax"][-1]).max()))
campos = np.array(cam["Position"], np.float64).ravel(); lpos = np.array(li["position"], np.float32).ravel()
omax = 4.0 * max(root, np.abs(campos).max(), np.abs(lpos).max())
h = L.emu_create(p(sh), len(sh), p(nd), len(nd), p(ix), len(ix), 3, float(omax))
fov = float(np.array(cam["fov"]).ravel()[0]); asp = float(np.array(cam["aspectRatio"]).ravel()[0])
ih = 2 * np.tan(np.radians(fov / 2)); iw = ih * asp
x, y = np.meshgrid(np.arange(W), np.arange(H))
nx = 2.0 * x / W - 1; ny = 1 - 2.0 * y / H
F, U, R = (np.array(cam[k], np.float64).ravel() for k in ("Front", "Up", "Right"))
d = F[None, None] + (nx * iw / 2)[..., None] * R + (ny * ih / 2)[..., None] * U
d /= np.linalg.norm(d, axis=2, keepdims=True)
D = np.ascontiguousarray(d.reshape(-1, 3), np.float32); O = np.ascontiguousarray(np.repeat(campos[None].astype(np.float32), W * H, 0))
for mode, nm in ((0, "list order"), (1, "list order + cull"), (2, "near first + cull")):
    out = (C.c_uint64 * 8)()
    L.probe_frame(h, p(sh), W * H, p(O), p(D), p(lpos), 3, mode, out)
    print("%-18s primary: visits %d prims %d rays %d | secondary: visits %d prims %d rays %d | total visits %d" % (nm, out[0], out[1], out[2], out[4], out[5], out[6], out[0] + out[4]))
