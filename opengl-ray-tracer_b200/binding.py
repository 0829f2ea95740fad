"""ctypes bindings of the C ABI (include/rt_api.h) and of the headless host library.

This is harness plumbing for tests/ and bench.py; the product is librt_b200.so.
There is no fallback: if the CUDA library is missing or cannot be loaded, import
of `lib()` raises.
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

# ---- std430 wire format (include/rt_flat.h; reference src/flatStructures.hpp:7-106) ----
LIGHT_DTYPE = np.dtype({"names": ["position", "color"], "formats": [("f4", 3), ("f4", 3)],
                        "offsets": [0, 16], "itemsize": 32})
CAMERA_DTYPE = np.dtype({"names": ["Position", "aspectRatio", "Front", "Up", "Right", "fov"],
                         "formats": [("f4", 3), "f4", ("f4", 3), ("f4", 3), ("f4", 3), "f4"],
                         "offsets": [0, 12, 16, 32, 48, 64], "itemsize": 80})
SHAPE_DTYPE = np.dtype({
    "names": ["type", "color", "fresnelStrength", "ambientStrength", "diffuseStrength", "specularStrength",
              "shininess", "sphereCenter", "sphereRadius", "planeNormal", "planeD", "wallStart", "wallWidth",
              "wallHeight", "triP1", "triP2", "triP3"],
    "formats": ["i4", ("f4", 3), "f4", "f4", "f4", "f4", "i4", ("f4", 3), "f4", ("f4", 3), "f4", ("f4", 3), "f4",
                "f4", ("f4", 3), ("f4", 3), ("f4", 3)],
    "offsets": [0, 32, 44, 48, 52, 56, 60, 64, 76, 80, 92, 96, 108, 112, 144, 160, 176],
    "itemsize": 192})
NODE_DTYPE = np.dtype({"names": ["boundsMin", "boundsMax", "leftChild", "rightChild", "startShapeIdx", "numShapes"],
                       "formats": [("f4", 3), ("f4", 3), "i4", "i4", "i4", "i4"],
                       "offsets": [0, 16, 32, 36, 40, 44], "itemsize": 48})

RT_FLAG_REFERENCE_WALK = 1
RT_FLAG_HIT_IDS = 2
RT_FLAG_NO_TOP_TREE = 4
RT_FLAG_NO_HEAVY_FIRST = 8
RT_FLAG_NO_DEFERRED_COPY = 16
RT_FLAG_NO_ONE_BLOCK_UPDATE = 32
RT_BUILD_DEVICE_LAYOUT = 4
TILE_W, TILE_H = 32, 8


class Params(C.Structure):
    """rt_params == the five uniforms of gpu_shader.comp:126-130."""
    _fields_ = [("screenRes", C.c_float * 2), ("maxBounces", C.c_int), ("useBVH", C.c_int),
                ("useFresnel", C.c_int), ("useMollerTrumbore", C.c_int)]


def make_params(width, height, maxBounces=3, useBVH=True, useFresnel=False, useMollerTrumbore=False):
    """Defaults as in src/main.cpp:118,135-137."""
    p = Params()
    p.screenRes[0], p.screenRes[1] = float(width), float(height)
    p.maxBounces, p.useBVH, p.useFresnel, p.useMollerTrumbore = int(maxBounces), int(useBVH), int(useFresnel), int(useMollerTrumbore)
    return p


_lib = None
_hostlib = None

# every symbol include/rt_api.h declares
API_SYMBOLS = ["rt_create", "rt_destroy", "rt_last_error", "rt_upload_scene", "rt_update_camera", "rt_update_light",
               "rt_update_shapes", "rt_update_nodes", "rt_set_shard", "rt_shard_pixels", "rt_render",
               "rt_render_async", "rt_render_rgba8", "rt_render_rgba8_async", "rt_wait_frames", "rt_render_device", "rt_render_stats", "rt_walk_stats", "rt_item_stats", "rt_set_frame_target", "rt_frame_export", "rt_frame_open", "rt_frame_download", "rt_sync_local", "rt_sync_export", "rt_sync_peer", "rt_sync_enable", "rt_frame_acquire", "rt_frame_release", "rt_host_frame_register", "rt_host_frame_unregister", "rt_anim_clear", "rt_anim_bounce", "rt_anim_rotate_z", "rt_animate", "rt_animate_shapes", "rt_download_scene", "rt_build_bvh", "rt_build_bvh_static", "rt_get_bvh", "rt_last_build_ms", "rt_layout_info", "rt_untile", "rt_present", "rt_stream", "rt_set_stream", "rt_synchronize", "rt_device_image",
               "rt_last_timings", "rt_launch_count", "rt_build_info"]


RT_STATIC_BUILD_MAX_SHAPES = 2048


def build_bvh_static(dev_shapes, n_shapes, max_depth, dev_nodes_out, node_capacity, dev_indices_out, index_capacity,
                     dev_counts_out, stream=None, library=None):
    """rt_build_bvh_static (include/rt_api.h): K8's single-block build with its tables in the library's static device
    array; every argument is a raw device address (int) or None.  Allocates nothing, does not synchronise.  Raises
    RuntimeError with the library's message on failure."""
    L = library or lib()
    err = C.create_string_buffer(512)
    rc = L.rt_build_bvh_static(dev_shapes, int(n_shapes), int(max_depth), dev_nodes_out, int(node_capacity),
                               dev_indices_out, int(index_capacity), dev_counts_out, stream, err, len(err))
    if rc != 0:
        raise RuntimeError("rt_build_bvh_static failed (%d): %s" % (rc, err.value.decode()))


def _records(a, dtype, what):
    """The C side strides over std430 records (192 / 48 bytes): numpy silently PACKS a padded structured dtype in some
    operations (np.concatenate does), and such an array would be read past its end."""
    if a is None:
        return None
    a = np.ascontiguousarray(a)
    if a.dtype.itemsize != dtype.itemsize:
        raise ValueError("%s: records of %d bytes, the wire format has %d (a packed copy of the structured dtype?)" % (what, a.dtype.itemsize, dtype.itemsize))
    return a


def obj_triangles(path, origin=(0.0, 0.0, 0.0), material_from=None, apply_rule=False):
    """rh_obj_triangles: a Wavefront OBJ as flat triangles (SHAPE_DTYPE) + the first triangle of every mesh (last entry: the total)."""
    H = hostlib()
    org = (C.c_float * 3)(*[float(x) for x in origin])
    mat = None
    if material_from is not None:
        mat = np.ascontiguousarray(material_from).reshape(-1)[:1]
        if mat.dtype.itemsize != SHAPE_DTYPE.itemsize:
            raise ValueError("material_from: one SHAPE_DTYPE record")
    nm = C.c_int(0)
    n = H.rh_obj_triangles(str(path).encode(), org, _ptr(mat), int(bool(apply_rule)), None, 0, None, 0, C.byref(nm))
    if n < 0:
        raise RuntimeError("rh_obj_triangles(%s): %s" % (path, {-1: "bad argument", -2: "cannot open the file", -3: "malformed v / f record"}.get(n, n)))
    out = np.zeros(n, dtype=SHAPE_DTYPE)
    first = np.zeros(nm.value + 1, dtype=np.int32)
    H.rh_obj_triangles(str(path).encode(), org, _ptr(mat), int(bool(apply_rule)), _ptr(out) if n else None, n, _ptr(first), len(first), C.byref(nm))
    return out, first


def lib_path():
    return os.path.join(_build.CSRC, "librt_b200.so")


_lib_fast = None


def lib_fast():
    """librt_b200_fast.so: the same source with FMA contraction allowed (bench.py's fast_build block only)."""
    global _lib_fast
    if _lib_fast is None:
        _lib_fast = _declare(C.CDLL(_build.build_cuda(fast=True)))
    return _lib_fast


def lib():
    """Load librt_b200.so (building it if nvcc is here).  Raises if unavailable."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("RT_B200_LIB") or _build.build_cuda()      # RT_B200_LIB: kernel-variant experiments (tools/)
    _lib = _declare(C.CDLL(path))
    return _lib


def _declare(L):
    vp, ip, fp = C.c_void_p, C.c_int, C.POINTER(C.c_float)
    L.rt_create.argtypes = [C.POINTER(vp), ip, ip, ip, C.c_uint]; L.rt_create.restype = ip
    L.rt_destroy.argtypes = [vp]; L.rt_destroy.restype = None
    L.rt_last_error.argtypes = [vp]; L.rt_last_error.restype = C.c_char_p
    L.rt_upload_scene.argtypes = [vp, vp, vp, vp, ip, vp, ip, vp, ip]; L.rt_upload_scene.restype = ip
    L.rt_update_camera.argtypes = [vp, vp]; L.rt_update_camera.restype = ip
    L.rt_update_light.argtypes = [vp, vp]; L.rt_update_light.restype = ip
    L.rt_update_shapes.argtypes = [vp, ip, ip, vp]; L.rt_update_shapes.restype = ip
    L.rt_update_nodes.argtypes = [vp, vp, ip]; L.rt_update_nodes.restype = ip
    L.rt_set_shard.argtypes = [vp, ip, ip]; L.rt_set_shard.restype = ip
    L.rt_shard_pixels.argtypes = [vp]; L.rt_shard_pixels.restype = C.c_int64
    L.rt_render.argtypes = [vp, C.POINTER(Params), vp, vp]; L.rt_render.restype = ip
    L.rt_render_async.argtypes = [vp, C.POINTER(Params), vp, vp]; L.rt_render_async.restype = ip
    L.rt_render_rgba8.argtypes = [vp, C.POINTER(Params), vp, vp]; L.rt_render_rgba8.restype = ip
    L.rt_render_rgba8_async.argtypes = [vp, C.POINTER(Params), vp, vp]; L.rt_render_rgba8_async.restype = ip
    L.rt_render_device.argtypes = [vp, C.POINTER(Params), vp, vp]; L.rt_render_device.restype = ip
    L.rt_render_stats.argtypes = [vp, C.POINTER(Params), C.POINTER(C.c_uint64), vp]; L.rt_render_stats.restype = ip
    L.rt_walk_stats.argtypes = [vp, C.POINTER(C.c_uint64)]; L.rt_walk_stats.restype = ip
    L.rt_set_frame_target.argtypes = [vp, vp]; L.rt_set_frame_target.restype = ip
    L.rt_frame_export.argtypes = [vp, C.c_char_p, C.POINTER(vp)]; L.rt_frame_export.restype = ip
    L.rt_frame_open.argtypes = [vp, C.c_char_p]; L.rt_frame_open.restype = ip
    L.rt_frame_download.argtypes = [vp, vp]; L.rt_frame_download.restype = ip
    L.rt_sync_local.argtypes = [vp, C.POINTER(vp)]; L.rt_sync_local.restype = ip
    L.rt_sync_export.argtypes = [vp, C.c_char_p]; L.rt_sync_export.restype = ip
    L.rt_sync_peer.argtypes = [vp, ip, vp, C.c_char_p]; L.rt_sync_peer.restype = ip
    L.rt_sync_enable.argtypes = [vp, ip]; L.rt_sync_enable.restype = ip
    L.rt_frame_acquire.argtypes = [vp, C.POINTER(vp)]; L.rt_frame_acquire.restype = ip
    L.rt_frame_release.argtypes = [vp]; L.rt_frame_release.restype = ip
    L.rt_host_frame_register.argtypes = [vp, vp, C.c_size_t, C.POINTER(vp)]; L.rt_host_frame_register.restype = ip
    L.rt_host_frame_unregister.argtypes = [vp, vp]; L.rt_host_frame_unregister.restype = ip
    L.rt_wait_frames.argtypes = [vp, ip]; L.rt_wait_frames.restype = ip
    L.rt_anim_clear.argtypes = [vp]; L.rt_anim_clear.restype = ip
    L.rt_anim_bounce.argtypes = [vp, ip, C.c_float, C.c_float, C.c_float]; L.rt_anim_bounce.restype = ip
    L.rt_anim_rotate_z.argtypes = [vp, vp, ip, vp, C.c_float]; L.rt_anim_rotate_z.restype = ip
    L.rt_animate.argtypes = [vp, C.c_float, C.c_float]; L.rt_animate.restype = ip
    L.rt_animate_shapes.argtypes = [vp, C.c_float, C.c_float]; L.rt_animate_shapes.restype = ip
    L.rt_download_scene.argtypes = [vp, vp, vp]; L.rt_download_scene.restype = ip
    L.rt_build_bvh.argtypes = [vp, ip, C.c_uint, C.POINTER(C.c_int), C.POINTER(C.c_int)]; L.rt_build_bvh.restype = ip
    L.rt_get_bvh.argtypes = [vp, vp, ip, vp, ip]; L.rt_get_bvh.restype = ip
    L.rt_last_build_ms.argtypes = [vp]; L.rt_last_build_ms.restype = C.c_float
    L.rt_layout_info.argtypes = [vp, C.POINTER(C.c_int * 4)]; L.rt_layout_info.restype = ip
    L.rt_build_bvh_static.argtypes = [vp, ip, ip, vp, ip, vp, ip, vp, vp, C.c_char_p, C.c_size_t]; L.rt_build_bvh_static.restype = ip
    L.rt_item_stats.argtypes = [vp, C.POINTER(C.c_uint64)]; L.rt_item_stats.restype = ip
    L.rt_untile.argtypes = [vp, vp, ip, vp]; L.rt_untile.restype = ip
    L.rt_present.argtypes = [vp, vp, vp, C.c_int64]; L.rt_present.restype = ip
    L.rt_stream.argtypes = [vp]; L.rt_stream.restype = vp
    L.rt_set_stream.argtypes = [vp, vp]; L.rt_set_stream.restype = ip
    L.rt_synchronize.argtypes = [vp]; L.rt_synchronize.restype = ip
    L.rt_device_image.argtypes = [vp]; L.rt_device_image.restype = vp
    L.rt_last_timings.argtypes = [vp, fp]; L.rt_last_timings.restype = ip
    L.rt_launch_count.argtypes = [vp]; L.rt_launch_count.restype = C.c_int64
    L.rt_build_info.argtypes = []; L.rt_build_info.restype = C.c_char_p
    return L


def hostlib():
    global _hostlib
    if _hostlib is not None:
        return _hostlib
    H = C.CDLL(_build.build_host())
    vp, ip = C.c_void_p, C.c_int
    H.rh_scene_create.argtypes = [C.c_char_p, ip, ip, ip, ip]; H.rh_scene_create.restype = vp
    H.rh_scene_from_shapes.argtypes = [vp, ip, vp, vp, ip]; H.rh_scene_from_shapes.restype = vp
    H.rh_scene_destroy.argtypes = [vp]; H.rh_scene_destroy.restype = None
    H.rh_obj_triangles.argtypes = [C.c_char_p, vp, vp, ip, vp, ip, vp, ip, C.POINTER(ip)]; H.rh_obj_triangles.restype = ip
    for name in ("rh_light", "rh_camera"):
        getattr(H, name).argtypes = [vp]; getattr(H, name).restype = vp
    for name in ("rh_shapes", "rh_nodes", "rh_indices", "rh_animated"):
        getattr(H, name).argtypes = [vp, C.POINTER(ip)]; getattr(H, name).restype = vp
    H.rh_bvh_stats.argtypes = [vp, C.POINTER(ip * 4)]; H.rh_bvh_stats.restype = None
    H.rh_scene_animate.argtypes = [vp, C.c_float, C.c_float, ip]; H.rh_scene_animate.restype = ip
    H.rh_anim_count.argtypes = [vp]; H.rh_anim_count.restype = ip
    H.rh_anim_get.argtypes = [vp, ip, C.POINTER(ip), C.POINTER(ip), C.POINTER(C.c_float * 4)]; H.rh_anim_get.restype = vp
    H.rh_set_camera.argtypes = [vp, C.POINTER(C.c_float * 3), C.POINTER(C.c_float * 3), C.c_float, C.c_float]
    H.rh_set_camera.restype = None
    f3 = C.c_float * 3
    H.rh_cam_create.argtypes = [f3, C.c_float]; H.rh_cam_create.restype = vp
    H.rh_cam_destroy.argtypes = [vp]; H.rh_cam_destroy.restype = None
    H.rh_cam_look_at.argtypes = [vp, f3]; H.rh_cam_look_at.restype = None
    H.rh_cam_keyboard.argtypes = [vp, ip, C.c_float]; H.rh_cam_keyboard.restype = None
    H.rh_cam_mouse.argtypes = [vp, C.c_float, C.c_float, ip]; H.rh_cam_mouse.restype = None
    H.rh_cam_scroll.argtypes = [vp, C.c_float]; H.rh_cam_scroll.restype = None
    H.rh_cam_shift.argtypes = [vp, ip]; H.rh_cam_shift.restype = None
    H.rh_cam_state.argtypes = [vp, C.c_float * 4]; H.rh_cam_state.restype = None
    H.rh_cam_serialize.argtypes = [vp, vp]; H.rh_cam_serialize.restype = None
    _hostlib = H
    return H


class RtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("rt error %d: %s" % (code, msg))
        self.code = code


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return C.c_void_p(a.__array_interface__["data"][0])      # (a.ctypes.data_as costs 4 us a pointer; the caller keeps `a` alive)
    if isinstance(a, int):
        return C.c_void_p(a)
    return C.cast(a, C.c_void_p)


class Scene:
    """A host-side scene: the reference's FlatScene + flatNodes + bvhIndices (numpy views)."""

    def __init__(self, name, width=800, height=600, n_override=0, bvh_depth=0, _handle=None):
        self._H = hostlib()
        self._h = _handle if _handle is not None else self._H.rh_scene_create(name.encode(), width, height, n_override, bvh_depth)
        if not self._h:
            raise ValueError("unknown scene %r" % name)
        self.name, self.width, self.height = name, width, height

    @classmethod
    def from_shapes(cls, shapes, light, camera, bvh_depth=15, width=800, height=600):
        H = hostlib()
        shapes = _records(shapes, SHAPE_DTYPE, "shapes")
        h = H.rh_scene_from_shapes(_ptr(shapes), len(shapes), _ptr(light), _ptr(camera), bvh_depth)
        return cls("custom", width, height, _handle=h)

    def _arr(self, fn, dtype):
        n = C.c_int(0)
        p = fn(self._h, C.byref(n))
        if n.value == 0:
            return np.zeros(0, dtype=dtype)
        buf = (C.c_char * (n.value * dtype.itemsize)).from_address(p)
        return np.frombuffer(buf, dtype=np.uint8).copy().view(dtype)      # byte copy: keeps the std430 padding

    @property
    def light(self):
        return np.frombuffer((C.c_char * 32).from_address(self._H.rh_light(self._h)), dtype=np.uint8).copy().view(LIGHT_DTYPE)

    @property
    def camera(self):
        return np.frombuffer((C.c_char * 80).from_address(self._H.rh_camera(self._h)), dtype=np.uint8).copy().view(CAMERA_DTYPE)

    @property
    def shapes(self):
        return self._arr(self._H.rh_shapes, SHAPE_DTYPE)

    @property
    def nodes(self):
        return self._arr(self._H.rh_nodes, NODE_DTYPE)

    @property
    def indices(self):
        return self._arr(self._H.rh_indices, np.dtype("i4"))

    @property
    def animated(self):
        return self._arr(self._H.rh_animated, np.dtype("i4"))

    def bvh_stats(self):
        out = (C.c_int * 4)()
        self._H.rh_bvh_stats(self._h, C.byref(out))
        return dict(nodes=out[0], leaves=out[1], max_leaf=out[2], depth=out[3])

    def animate(self, t, dt, mode=0):
        return self._H.rh_scene_animate(self._h, float(t), float(dt), int(mode))

    def animations(self):
        """[(kind, shape indices, params)]: kind 0 bounce (origin y, amplitude, frequency), 1 rotation about z (centre xyz, speed)."""
        out = []
        for k in range(self._H.rh_anim_count(self._h)):
            kind, count, params = C.c_int(), C.c_int(), (C.c_float * 4)()
            p = self._H.rh_anim_get(self._h, k, C.byref(kind), C.byref(count), C.byref(params))
            idx = np.frombuffer((C.c_int32 * count.value).from_address(p), dtype=np.int32).copy()
            out.append((kind.value, idx, [float(x) for x in params]))
        return out

    def set_camera(self, pos, target, fov=60.0, aspect=None):
        a = aspect if aspect is not None else self.width / self.height
        self._H.rh_set_camera(self._h, C.byref((C.c_float * 3)(*pos)), C.byref((C.c_float * 3)(*target)), fov, a)

    def close(self):
        if self._h:
            self._H.rh_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Camera:
    """The reference's interactive camera (camera.hpp:30-165) as its render loop drives it (processInput main.cpp:509-533,
    mouse_callback :536-551); `flat` is serializeCamera's 80 bytes (:806-816), the argument of RayTracer.update_camera."""
    FORWARD, BACKWARD, LEFT, RIGHT, UP, DOWN = range(6)

    def __init__(self, position=(0.0, 0.0, 0.0), aspect=1.0):
        self._H = hostlib()
        self._h = self._H.rh_cam_create((C.c_float * 3)(*[float(v) for v in position]), float(aspect))

    def look_at(self, target):
        self._H.rh_cam_look_at(self._h, (C.c_float * 3)(*[float(v) for v in target]))

    def keyboard(self, direction, dt):
        self._H.rh_cam_keyboard(self._h, int(direction), float(dt))

    def mouse(self, dx, dy, constrain_pitch=True):
        self._H.rh_cam_mouse(self._h, float(dx), float(dy), int(bool(constrain_pitch)))

    def scroll(self, dy):
        self._H.rh_cam_scroll(self._h, float(dy))

    def shift(self, pressed):
        self._H.rh_cam_shift(self._h, int(bool(pressed)))

    @property
    def state(self):
        out = (C.c_float * 4)()
        self._H.rh_cam_state(self._h, out)
        return dict(yaw=out[0], pitch=out[1], zoom=out[2], speed=out[3])

    @property
    def flat(self):
        out = np.zeros(1, dtype=CAMERA_DTYPE)
        assert out.nbytes == 80
        self._H.rh_cam_serialize(self._h, _ptr(out))
        return out

    def close(self):
        if self._h:
            self._H.rh_cam_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class RayTracer:
    """Thin object wrapper over rt_create/rt_upload_scene/rt_render/rt_destroy."""

    def __init__(self, width, height, device=0, flags=RT_FLAG_HIT_IDS, library=None):
        self._L = library if library is not None else lib()
        self._c = C.c_void_p()
        rc = self._L.rt_create(C.byref(self._c), width, height, device, flags)
        if rc != 0:
            raise RtError(rc, "rt_create failed (no CUDA device? there is no CPU fallback)")
        self.width, self.height, self.flags = width, height, flags

    def _chk(self, rc):
        if rc != 0:
            raise RtError(rc, (self._L.rt_last_error(self._c) or b"").decode())

    def upload_scene(self, light=None, camera=None, shapes=None, nodes=None, indices=None):
        keep = [np.ascontiguousarray(a) if a is not None else None for a in (light, camera, shapes, nodes, indices)]
        li, ca, sh, no, ix = keep
        sh, no = _records(sh, SHAPE_DTYPE, "shapes"), _records(no, NODE_DTYPE, "nodes")
        self._chk(self._L.rt_upload_scene(self._c, _ptr(li), _ptr(ca), _ptr(sh), 0 if sh is None else len(sh),
                                          _ptr(no), 0 if no is None else len(no), _ptr(ix), 0 if ix is None else len(ix)))

    def upload(self, scene):
        self.upload_scene(scene.light, scene.camera, scene.shapes, scene.nodes, scene.indices)

    def update_camera(self, camera):
        self._chk(self._L.rt_update_camera(self._c, _ptr(np.ascontiguousarray(camera))))

    def update_light(self, light):
        self._chk(self._L.rt_update_light(self._c, _ptr(np.ascontiguousarray(light))))

    def update_shapes(self, first, shapes):
        shapes = _records(shapes, SHAPE_DTYPE, "shapes")
        self._chk(self._L.rt_update_shapes(self._c, int(first), len(shapes), _ptr(shapes)))

    def update_nodes(self, nodes):
        nodes = _records(nodes, NODE_DTYPE, "nodes")
        self._chk(self._L.rt_update_nodes(self._c, _ptr(nodes), len(nodes)))

    def set_shard(self, rank, world):
        self._chk(self._L.rt_set_shard(self._c, rank, world))

    def shard_pixels(self):
        return int(self._L.rt_shard_pixels(self._c))

    def render(self, params, want_ids=True, out=None, out_ids=None):
        """Synchronous render into host numpy arrays (the rt_render entry point)."""
        px = self.shard_pixels()
        full = px == self.width * self.height
        rgba = out if out is not None else np.empty((px, 4), dtype=np.float32)
        ids = None
        if want_ids and (self.flags & RT_FLAG_HIT_IDS):
            ids = out_ids if out_ids is not None else np.empty(px, dtype=np.int32)
        self._chk(self._L.rt_render(self._c, C.byref(params), _ptr(rgba), _ptr(ids)))
        if full and out is None:
            rgba = rgba.reshape(self.height, self.width, 4)
            ids = ids.reshape(self.height, self.width) if ids is not None else None
        return rgba, ids

    def render_async(self, params, out, out_ids=None):
        """rt_render_async: returns at once; `out` (pinned) is complete after synchronize()."""
        self._chk(self._L.rt_render_async(self._c, C.byref(params), _ptr(out), _ptr(out_ids)))

    def render_rgba8(self, params, out=None, out_ids=None):
        """rt_render_rgba8: the frame as the reference shows it (compute dispatch + display quad), W*H RGBA8 pixels on the host."""
        if out is None:
            out = np.empty((self.shard_pixels(), 4), dtype=np.uint8)
        self._chk(self._L.rt_render_rgba8(self._c, C.byref(params), _ptr(out), _ptr(out_ids)))
        return out

    def render_rgba8_async(self, params, out, out_ids=None):
        """rt_render_rgba8_async: returns at once; `out` (pinned uint8) is complete after wait_frames / synchronize()."""
        self._chk(self._L.rt_render_rgba8_async(self._c, C.byref(params), _ptr(out), _ptr(out_ids)))

    def wait_frames(self, max_in_flight=1):
        self._chk(self._L.rt_wait_frames(self._c, int(max_in_flight)))

    def render_device(self, params, dev_rgba_ptr=None, dev_ids_ptr=None):
        """Asynchronous render into device memory (pointers as ints, e.g. tensor.data_ptr())."""
        self._chk(self._L.rt_render_device(self._c, C.byref(params), _ptr(dev_rgba_ptr), _ptr(dev_ids_ptr)))

    def render_stats(self, params, cost=None):
        """Untimed counting render: dict of work counters of the CUDA path."""
        out = (C.c_uint64 * 8)()
        self._chk(self._L.rt_render_stats(self._c, C.byref(params), out, _ptr(cost)))
        keys = ["rays", "primary_rays", "shadow_rays", "reflect_rays", "node_visits", "prim_tests", "ray_hits", "shadowed"]
        return dict(zip(keys, [int(x) for x in out]))

    def walk_stats(self):
        out = (C.c_uint64 * 8)()
        self._chk(self._L.rt_walk_stats(self._c, out))
        keys = ["packet_iters", "lane_iters", "packet_served", "lane_served"]
        v = [int(x) for x in out]
        return {"primary": dict(zip(keys, v[:4])), "bounce": dict(zip(keys, v[4:]))}

    def set_frame_target(self, dev_ptr):
        self._chk(self._L.rt_set_frame_target(self._c, _ptr(dev_ptr) if dev_ptr else None))

    def frame_export(self):
        """Rank 0: returns (64-byte IPC handle, device pointer of the full frame)."""
        h = C.create_string_buffer(64)
        p = C.c_void_p()
        self._chk(self._L.rt_frame_export(self._c, h, C.byref(p)))
        return h.raw, p.value

    def frame_open(self, handle):
        self._chk(self._L.rt_frame_open(self._c, C.create_string_buffer(handle, 64)))

    def frame_download(self, host):
        self._chk(self._L.rt_frame_download(self._c, _ptr(host)))

    def sync_local(self):
        """This rank's release word (device pointer) for the signalled fused gather."""
        p = C.c_void_p()
        self._chk(self._L.rt_sync_local(self._c, C.byref(p)))
        return p.value

    def sync_export(self):
        h = C.create_string_buffer(64)
        self._chk(self._L.rt_sync_export(self._c, h))
        return h.raw

    def sync_peer(self, rank, dev_word=None, handle=None):
        self._chk(self._L.rt_sync_peer(self._c, int(rank), C.c_void_p(dev_word) if dev_word else None,
                                       C.create_string_buffer(handle, 64) if handle else None))

    def sync_enable(self, world):
        self._chk(self._L.rt_sync_enable(self._c, int(world)))

    def frame_acquire(self):
        p = C.c_void_p()
        self._chk(self._L.rt_frame_acquire(self._c, C.byref(p)))
        return p.value

    def frame_release(self):
        self._chk(self._L.rt_frame_release(self._c))

    def host_frame_register(self, host):
        """host: a C-contiguous numpy float32 array of W*H*4 (may live in shared memory); returns the device alias."""
        d = C.c_void_p()
        self._chk(self._L.rt_host_frame_register(self._c, _ptr(host), host.nbytes, C.byref(d)))
        return d.value

    def register_animations(self, scene):
        """Registers the scene's animation list (bouncing spheres / rotating wheels) for rt_animate."""
        self._chk(self._L.rt_anim_clear(self._c))
        for kind, shapes, params in scene.animations():
            if kind == 0:
                self._chk(self._L.rt_anim_bounce(self._c, int(shapes[0]), params[0], params[1], params[2]))
            else:
                ctr = np.array(params[:3], np.float32)
                self._chk(self._L.rt_anim_rotate_z(self._c, _ptr(np.ascontiguousarray(shapes, np.int32)), len(shapes), _ptr(ctr), params[3]))

    def animate(self, t, dt):
        self._chk(self._L.rt_animate(self._c, t, dt))

    def animate_shapes(self, t, dt):
        """K6 only (no K7): for loops that rebuild the tree afterwards."""
        self._chk(self._L.rt_animate_shapes(self._c, t, dt))

    def download_scene(self, n_shapes, n_nodes):
        shapes = np.zeros(n_shapes, dtype=SHAPE_DTYPE); nodes = np.zeros(n_nodes, dtype=NODE_DTYPE)
        self._chk(self._L.rt_download_scene(self._c, _ptr(shapes), _ptr(nodes)))
        return shapes, nodes

    def build_bvh(self, max_depth, build_flags=0, fetch=True):
        """K8: build the reference's BVH on the device from the context's shapes; returns (nodes, indices) when `fetch`."""
        nn, ni = C.c_int(0), C.c_int(0)
        self._chk(self._L.rt_build_bvh(self._c, int(max_depth), int(build_flags), C.byref(nn), C.byref(ni)))
        if not fetch:
            return nn.value, ni.value
        nodes = np.zeros(nn.value, dtype=NODE_DTYPE); idx = np.zeros(ni.value, dtype=np.int32)
        self._chk(self._L.rt_get_bvh(self._c, _ptr(nodes), nn.value, _ptr(idx), ni.value))
        return nodes, idx

    def layout_info(self):
        out = (C.c_int * 4)()
        self._chk(self._L.rt_layout_info(self._c, C.byref(out)))
        return dict(nodes=out[0], nodes_b=out[1], sub=out[2], on_device=bool(out[3]))

    def last_build_ms(self):
        return float(self._L.rt_last_build_ms(self._c))

    def host_frame_unregister(self, host):
        self._chk(self._L.rt_host_frame_unregister(self._c, _ptr(host)))

    def item_stats(self):
        out = (C.c_uint64 * 3)()
        self._chk(self._L.rt_item_stats(self._c, out))
        return dict(max_item_cycles=int(out[0]), mean_item_cycles=int(out[1]) / max(1, int(out[2])), items=int(out[2]))

    def untile(self, dev_shards_ptr, world, dev_out_ptr):
        self._chk(self._L.rt_untile(self._c, _ptr(dev_shards_ptr), world, _ptr(dev_out_ptr)))

    def present(self, dev_rgba_ptr, dev_rgba8_ptr, pixels):
        self._chk(self._L.rt_present(self._c, _ptr(dev_rgba_ptr), _ptr(dev_rgba8_ptr), pixels))

    def stream(self):
        return self._L.rt_stream(self._c)

    def set_stream(self, cuda_stream):
        """Share a caller-owned stream (int handle, e.g. torch.cuda.current_stream().cuda_stream)."""
        self._chk(self._L.rt_set_stream(self._c, C.c_void_p(cuda_stream) if cuda_stream else None))

    def synchronize(self):
        self._chk(self._L.rt_synchronize(self._c))

    def device_image(self):
        return self._L.rt_device_image(self._c)

    def timings(self):
        out = (C.c_float * 4)()
        self._chk(self._L.rt_last_timings(self._c, out))
        return dict(h2d_ms=out[0], repack_ms=out[1], render_ms=out[2], d2h_ms=out[3])

    def launch_count(self):
        return int(self._L.rt_launch_count(self._c))

    def close(self):
        if self._c:
            self._L.rt_destroy(self._c)
            self._c = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
