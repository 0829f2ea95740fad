"""ctypes binding of the CPU oracle (TEST INFRASTRUCTURE).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs import this module.  The product path never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


class Counters(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("primary_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("reflect_rays", C.c_uint64), ("node_visits", C.c_uint64), ("index_reads", C.c_uint64),
                ("tests", C.c_uint64 * 4), ("tri_stage2", C.c_uint64), ("ray_hits", C.c_uint64),
                ("alg_bytes", C.c_uint64)]

    def as_dict(self):
        d = {k: int(getattr(self, k)) for k, _ in self._fields_ if k != "tests"}
        d["tests"] = [int(x) for x in self.tests]
        return d


def build(force=False):
    out = os.path.join(HERE, "liboracle_rt.so")
    srcs = [os.path.join(HERE, "oracle_rt.cpp"), os.path.join(HERE, "oracle_rt.h")]
    if force or not os.path.exists(out) or any(os.path.getmtime(s) > os.path.getmtime(out) for s in srcs):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-pthread",
                               "-o", out, srcs[0]])
    return out


def build_ref():
    """(Re)build oracle/_ref from the reference tree when it is present; returns the .so path or None."""
    subprocess.call(["python3", os.path.join(HERE, "build_ref.py")])
    p = os.path.join(HERE, "_ref", "libref_shader.so")
    return p if os.path.exists(p) else None


_lib = None
_ref = None


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        vp, ip = C.c_void_p, C.c_int
        L.oracle_render.argtypes = [vp, vp, vp, ip, vp, ip, vp, ip, vp, ip, ip, ip, ip, vp, vp, vp, vp, ip]
        L.oracle_render.restype = ip
        L.oracle_intersect_shape.argtypes = [vp, vp, vp, ip, vp]; L.oracle_intersect_shape.restype = ip
        L.oracle_intersect_scene.argtypes = [vp, ip, vp, ip, vp, ip, vp, vp, ip, vp, vp, vp]; L.oracle_intersect_scene.restype = ip
        L.oracle_intersect_scene_batch.argtypes = [vp, ip, vp, ip, vp, ip, ip, vp, vp, ip, vp, vp, vp]; L.oracle_intersect_scene_batch.restype = None
        L.oracle_primary_ties.argtypes = [vp, vp, ip, vp, ip, vp, ip, vp, ip, ip, ip, ip, vp, ip]; L.oracle_primary_ties.restype = C.c_int64
        L.oracle_present.argtypes = [vp, vp, C.c_int64]; L.oracle_present.restype = None
        _lib = L
    return _lib


def ref_lib():
    """The reference shader compiled as C++ (oracle/_ref/libref_shader.so), or None if never built."""
    global _ref
    if _ref is None:
        p = os.path.join(HERE, "_ref", "libref_shader.so")
        if not os.path.exists(p):
            return None
        R = C.CDLL(p)
        vp, ip = C.c_void_p, C.c_int
        R.ref_render.argtypes = [vp, vp, vp, ip, vp, ip, vp, ip, vp, ip, ip, ip, ip, vp, ip]
        R.ref_render.restype = ip
        _ref = R
    return _ref


_ref_glsl = None


def ref_lib_glsl():
    """The reference shader compiled as C++ with min / max / clamp / mix as the GLSL specification words them
    (oracle/_ref/libref_shader_glsl.so, -DREF_GLSL_BUILTINS in ref_shader_glue.cpp), or None if never built."""
    global _ref_glsl
    if _ref_glsl is None:
        p = os.path.join(HERE, "_ref", "libref_shader_glsl.so")
        if not os.path.exists(p):
            return None
        R = C.CDLL(p)
        vp, ip = C.c_void_p, C.c_int
        R.ref_render.argtypes = [vp, vp, vp, ip, vp, ip, vp, ip, vp, ip, ip, ip, ip, vp, ip]
        R.ref_render.restype = ip
        _ref_glsl = R
    return _ref_glsl


_bvh = None


def refbvh_lib():
    """The reference's own builder / serialiser / animation step cut out of src/main.cpp and compiled in place
    (oracle/_ref/libref_bvh.so, see oracle/ref_bvh_glue.cpp), or None if never built."""
    global _bvh
    if _bvh is None:
        p = os.path.join(HERE, "_ref", "libref_bvh.so")
        if not os.path.exists(p):
            return None
        B = C.CDLL(p)
        vp, ip, fl = C.c_void_p, C.c_int, C.c_float
        B.refbvh_load.argtypes = [vp, ip, vp]; B.refbvh_load.restype = ip
        B.refbvh_build.argtypes = [ip, C.POINTER(ip), C.POINTER(ip)]; B.refbvh_build.restype = None
        B.refbvh_get.argtypes = [vp, vp]; B.refbvh_get.restype = None
        B.refbvh_get_shapes.argtypes = [vp]; B.refbvh_get_shapes.restype = None
        B.refbvh_wheel.argtypes = [ip, vp, ip, vp]; B.refbvh_wheel.restype = ip
        B.refbvh_bounce.argtypes = [ip, fl, fl, fl]; B.refbvh_bounce.restype = ip
        B.refbvh_wheels_step.argtypes = [fl, fl]; B.refbvh_wheels_step.restype = None
        B.refbvh_frame.argtypes = [ip, ip, C.POINTER(ip), C.POINTER(ip)]; B.refbvh_frame.restype = None
        _bvh = B
    return _bvh


class RefScene:
    """The reference's `scene` (src/main.cpp:93-101) filled from flat shapes; ONE instance at a time (upstream globals).
    build() = buildBVH + serializeBVH; frame() = updateScene + updateBVH | buildBVH + serializeBVH (main.cpp:328-344)."""

    def __init__(self, shapes, animated=(), node_dtype=None):
        self.B = refbvh_lib()
        if self.B is None:
            raise RuntimeError("oracle/_ref/libref_bvh.so not built")
        self.shapes0 = np.ascontiguousarray(shapes)
        self.n = len(self.shapes0)
        flag = np.zeros(max(self.n, 1), np.uint8)
        flag[np.asarray(list(animated), dtype=np.int64)] = 1
        if self.B.refbvh_load(_p(self.shapes0), self.n, _p(flag)) != 0:
            raise ValueError("unknown shape type")
        self.node_dtype = node_dtype

    def _fetch(self, nn, ni):
        nodes = np.zeros(nn.value, dtype=self.node_dtype if self.node_dtype is not None else np.dtype(("u1", 48)))
        idx = np.zeros(ni.value, np.int32)
        self.B.refbvh_get(_p(nodes), _p(idx))
        return nodes, idx

    def build(self, max_depth):
        nn, ni = C.c_int(0), C.c_int(0)
        self.B.refbvh_build(int(max_depth), C.byref(nn), C.byref(ni))
        return self._fetch(nn, ni)

    def wheel(self, k, indices, center):
        idx = np.ascontiguousarray(indices, np.int32); c = np.ascontiguousarray(center, np.float32)
        assert self.B.refbvh_wheel(int(k), _p(idx), len(idx), _p(c)) == 0

    def bounce(self, shape, t, amplitude, frequency):
        assert self.B.refbvh_bounce(int(shape), t, amplitude, frequency) == 0

    def wheels_step(self, t, dt):
        self.B.refbvh_wheels_step(t, dt)

    def frame(self, rebuild=False, max_depth=15):
        nn, ni = C.c_int(0), C.c_int(0)
        self.B.refbvh_frame(int(bool(rebuild)), int(max_depth), C.byref(nn), C.byref(ni))
        return self._fetch(nn, ni)

    def shapes(self):
        out = self.shapes0.copy()
        self.B.refbvh_get_shapes(_p(out))
        return out


def render(light, camera, shapes, nodes, indices, params, W, H, threads=None, row_step=1, row_phase=0,
           want_counters=True):
    """Returns (rgba[H,W,4] f32, ids[H,W] i32, dists[H,W] f32, counters dict|None)."""
    threads = threads or os.cpu_count() or 1
    rgba = np.zeros((H, W, 4), np.float32)
    ids = np.full((H, W), -1, np.int32)
    dists = np.full((H, W), 1e20, np.float32)
    cnt = Counters() if want_counters else None
    arrs = [np.ascontiguousarray(a) for a in (light, camera, shapes, nodes, indices)]
    rc = lib().oracle_render(_p(arrs[0]), _p(arrs[1]), _p(arrs[2]), len(arrs[2]), _p(arrs[3]), len(arrs[3]),
                             _p(arrs[4]), len(arrs[4]), C.addressof(params), W, H, row_step, row_phase,
                             _p(rgba), _p(ids), _p(dists), C.addressof(cnt) if cnt is not None else None, threads)
    if rc != 0:
        raise RuntimeError("oracle_render failed")
    return rgba, ids, dists, (cnt.as_dict() if cnt is not None else None)


def primary_ties(camera, shapes, nodes, indices, params, W, H, threads=None, row_step=1, row_phase=0, want_map=False):
    """Number of tie pixels among the sampled rows (a candidate of another shape within 1 ulp of the winner's
    closestDist, gpu_shader.comp:410-411); with want_map also the 0/1 map."""
    threads = threads or os.cpu_count() or 1
    arrs = [np.ascontiguousarray(a) for a in (camera, shapes, nodes, indices)]
    ties = np.zeros((H, W), np.uint8) if want_map else None
    n = lib().oracle_primary_ties(_p(arrs[0]), _p(arrs[1]), len(arrs[1]), _p(arrs[2]), len(arrs[2]), _p(arrs[3]), len(arrs[3]),
                                  C.addressof(params), W, H, row_step, row_phase, _p(ties), threads)
    if n < 0:
        raise RuntimeError("oracle_primary_ties failed")
    return (int(n), ties) if want_map else int(n)


def ref_render(light, camera, shapes, nodes, indices, params, W, H, threads=None, row_step=1, row_phase=0, glsl_builtins=False):
    R = ref_lib_glsl() if glsl_builtins else ref_lib()
    if R is None:
        raise RuntimeError("oracle/_ref not built")
    threads = threads or os.cpu_count() or 1
    rgba = np.zeros((H, W, 4), np.float32)
    arrs = [np.ascontiguousarray(a) for a in (light, camera, shapes, nodes, indices)]
    rc = R.ref_render(_p(arrs[0]), _p(arrs[1]), _p(arrs[2]), len(arrs[2]), _p(arrs[3]), len(arrs[3]),
                      _p(arrs[4]), len(arrs[4]), C.addressof(params), W, H, row_step, row_phase, _p(rgba), threads)
    if rc != 0:
        raise RuntimeError("ref_render failed")
    return rgba


def intersect_shape(shape, start, direction, useMT=False):
    hp = np.zeros(3, np.float32)
    s = np.ascontiguousarray(shape)
    t = lib().oracle_intersect_shape(_p(s), _p(np.asarray(start, np.float32)), _p(np.asarray(direction, np.float32)),
                                     int(useMT), _p(hp))
    return t, hp


def intersect_scene_batch(shapes, nodes, indices, starts, dirs, useMT=False):
    """Closest-hit BVH query (intersectScene2) for n rays: returns (shape ids, closestDist, hit points)."""
    arrs = [np.ascontiguousarray(a) for a in (shapes, nodes, indices)]
    starts = np.ascontiguousarray(starts, np.float32); dirs = np.ascontiguousarray(dirs, np.float32)
    n = len(starts)
    ids = np.empty(n, np.int32); dist = np.empty(n, np.float32); pts = np.empty((n, 3), np.float32)
    lib().oracle_intersect_scene_batch(_p(arrs[0]), len(arrs[0]), _p(arrs[1]), len(arrs[1]), _p(arrs[2]), len(arrs[2]),
                                       n, _p(starts), _p(dirs), int(useMT), _p(ids), _p(dist), _p(pts))
    return ids, dist, pts


def present(rgba):
    rgba = np.ascontiguousarray(rgba, np.float32)
    out = np.zeros(rgba.shape[:-1] + (4,), np.uint8)
    lib().oracle_present(_p(rgba), _p(out), rgba.size // 4)
    return out
