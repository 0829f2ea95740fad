"""The per-frame camera step of the reference's render loop (processInput main.cpp:509-533, mouse_callback :536-551,
Camera camera.hpp:30-165, serializeCamera main.cpp:806-816) in the headless host (rh_cam_*): the 80 bytes handed to
rt_update_camera equal, bit for bit, those of the reference's own Camera class compiled in place
(oracle/_ref/libref_camera.so) — live where the reference tree is present, and against committed traces generated from it
(tests/golden/camera_traces.json, tests/golden/make_camera_golden.py) everywhere."""
import ctypes as C
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(os.path.dirname(HERE), "oracle", "_ref", "libref_camera.so")


class RefCamera:
    """The reference class behind oracle/ref_camera_glue.cpp."""

    def __init__(self, pos, aspect):
        L = C.CDLL(REF_SO)
        f3, vp = C.c_float * 3, C.c_void_p
        L.refcam_create.argtypes = [f3, C.c_float]; L.refcam_create.restype = vp
        L.refcam_destroy.argtypes = [vp]
        L.refcam_look_at.argtypes = [vp, f3]
        L.refcam_keyboard.argtypes = [vp, C.c_int, C.c_float]
        L.refcam_mouse.argtypes = [vp, C.c_float, C.c_float, C.c_int]
        L.refcam_scroll.argtypes = [vp, C.c_float]
        L.refcam_shift.argtypes = [vp, C.c_int]
        L.refcam_state.argtypes = [vp, C.c_float * 4]
        L.refcam_serialize.argtypes = [vp, vp]
        self._L, self._f3 = L, f3
        self._h = L.refcam_create(f3(*pos), aspect)

    def look_at(self, t): self._L.refcam_look_at(self._h, self._f3(*t))
    def keyboard(self, d, dt): self._L.refcam_keyboard(self._h, d, dt)
    def mouse(self, dx, dy, constrain=True): self._L.refcam_mouse(self._h, dx, dy, int(constrain))
    def scroll(self, dy): self._L.refcam_scroll(self._h, dy)
    def shift(self, p): self._L.refcam_shift(self._h, int(p))

    @property
    def state(self):
        out = (C.c_float * 4)()
        self._L.refcam_state(self._h, out)
        return dict(yaw=out[0], pitch=out[1], zoom=out[2], speed=out[3])

    @property
    def flat(self):
        out = np.zeros(80, np.uint8)
        self._L.refcam_serialize(self._h, out.ctypes.data)
        return out


def make_trace(seed, n):
    """A seeded input trace: (start position, aspect, [(event, args...)]).  Frame times as a 60 - 240 Hz loop produces them,
    mouse offsets up to a fast flick, pitch driven into the +-89 degree clamp now and then."""
    rng = np.random.default_rng(1000 + seed)
    f = lambda v: float(np.float32(v))
    start = [f(v) for v in rng.uniform(-40, 40, 3)]
    aspect = f(rng.choice([800 / 600, 3840 / 2160, 1.0]))
    ev = [("look_at", [f(v) for v in rng.uniform(-30, 30, 3)])]
    for _ in range(n - 1):
        k = rng.integers(0, 12)
        if k < 5:
            ev.append(("keyboard", int(rng.integers(0, 6)), f(rng.uniform(1 / 240, 1 / 60))))
        elif k < 9:
            big = rng.integers(0, 6) == 0
            ev.append(("mouse", f(rng.uniform(-60, 60)), f(rng.uniform(-900, 900) if big else rng.uniform(-40, 40)), bool(rng.integers(0, 8) > 0)))
        elif k == 9:
            ev.append(("scroll", f(rng.uniform(-20, 20))))
        elif k == 10:
            ev.append(("shift", bool(rng.integers(0, 2))))
        else:
            ev.append(("look_at", [f(v) for v in rng.uniform(-30, 30, 3)]))
    return start, aspect, ev


def play(cam, events):
    frames = []
    for e in events:
        getattr(cam, e[0])(*e[1:])
        frames.append(np.asarray(cam.flat).view(np.uint8).ravel().copy())
    return frames


@pytest.fixture(scope="module")
def host_camera(rt):
    import sys
    return sys.modules[rt.__name__ + ".binding"].Camera


def test_default_camera_and_serialised_layout(rt, host_camera):
    """Camera() (camera.hpp:49-56): yaw -90, pitch 0, fov 60; FlatCamera field offsets (flatStructures.hpp:55-70)."""
    cam = host_camera((1.0, 2.0, 3.0), 1.5)
    fl = cam.flat
    assert fl.nbytes == 80 and cam.state == dict(yaw=-90.0, pitch=0.0, zoom=45.0, speed=15.0)
    assert np.allclose(fl["Position"][0], (1, 2, 3)) and fl["aspectRatio"][0] == 1.5 and fl["fov"][0] == 60.0
    assert np.allclose(fl["Front"][0], (0, 0, -1), atol=1e-7) and np.allclose(fl["Up"][0], (0, 1, 0)) and np.allclose(fl["Right"][0], (1, 0, 0), atol=1e-7)
    cam.shift(True); assert cam.state["speed"] == 30.0                       # SPEED * SPEEDAMPLIFIER, main.cpp:528
    cam.scroll(100.0); assert cam.state["zoom"] == 1.0                        # camera.hpp:118-121
    cam.mouse(0.0, 1e4); assert cam.state["pitch"] == 89.0                    # :103-108
    cam.mouse(0.0, 1e4, constrain_pitch=False); assert cam.state["pitch"] > 89.0
    cam.close()


def test_up_key_moves_against_the_up_vector(rt, host_camera):
    """ProcessKeyboard(UP) SUBTRACTS Up * velocity (camera.hpp:86-89): kept as upstream has it."""
    cam = host_camera((0.0, 0.0, 0.0), 1.0)
    cam.keyboard(4, 0.1)
    assert np.allclose(cam.flat["Position"][0], (0, -1.5, 0))
    cam.keyboard(0, 0.1)
    assert np.allclose(cam.flat["Position"][0], (0, -1.5, -1.5), atol=1e-6)
    cam.close()


@pytest.mark.parametrize("seed", range(6))
def test_traces_equal_the_committed_reference_bytes(rt, host_camera, seed):
    with open(os.path.join(HERE, "golden", "camera_traces.json")) as f:
        gold = json.load(f)["traces"][seed]
    start, aspect, events = make_trace(seed, gold["n"])
    frames = play(host_camera(start, aspect), events)
    for i, (fr, hx) in enumerate(zip(frames, gold["flat_hex"])):
        assert fr.tobytes().hex() == hx, "seed %d: camera bytes after event %d %r" % (seed, i, events[i])


@pytest.mark.parametrize("seed", range(100, 140))
def test_traces_equal_the_reference_class_live(rt, host_camera, seed):
    if not os.path.exists(REF_SO):
        pytest.skip("oracle/_ref/libref_camera.so not built (no reference tree here)")
    start, aspect, events = make_trace(seed, 200)
    ref = RefCamera(start, aspect)
    mine = host_camera(start, aspect)
    for i, e in enumerate(events):
        getattr(ref, e[0])(*e[1:]); getattr(mine, e[0])(*e[1:])
        a, b = np.asarray(ref.flat).view(np.uint8).ravel(), np.asarray(mine.flat).view(np.uint8).ravel()
        assert np.array_equal(a, b), "seed %d: camera bytes after event %d %r: %s vs %s" % (seed, i, e, a.view(np.float32), b.view(np.float32))
        assert ref.state == mine.state


@pytest.mark.parametrize("scene,pos,intensity", [("monkey", (0.0, -14.0, 0.0), 50.0), ("car", (14.8, -17.0, 17.0), 26.0), ("triangle", (14.8, -17.0, 17.0), 26.0)])
def test_scene_lights_equal_the_reference_light_class(rt, scene, pos, intensity):
    """Light(position, vec3(1), intensity) (light.hpp:22-35; main.cpp:591, :725, :1203) through serializeLight (:818-823):
    the 32 bytes of the headless scenes' lights equal those of the reference class compiled in place."""
    if not os.path.exists(REF_SO):
        pytest.skip("oracle/_ref/libref_camera.so not built (no reference tree here)")
    L = C.CDLL(REF_SO)
    f3 = C.c_float * 3
    L.reflight_serialize.argtypes = [f3, f3, C.c_float, C.c_void_p]
    out = np.zeros(32, np.uint8)
    L.reflight_serialize(f3(*pos), f3(1.0, 1.0, 1.0), intensity, out.ctypes.data)
    sc = rt.Scene(scene, 800, 600)
    assert np.array_equal(np.asarray(sc.light).view(np.uint8).ravel(), out), (out.view(np.float32), sc.light)
