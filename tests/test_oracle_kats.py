"""CPU: primitive-level known answers.  kat_shapes.npz holds outputs of the reference's own C++
shape classes (Sphere/Wall/Triangle::get_intersection); the rest are hand-computed."""
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
NONE, INNER, OUTER = 0, 1, 2


def _shape(rt, **kw):
    s = np.zeros(1, dtype=rt.SHAPE_DTYPE)
    for k, v in kw.items():
        s[k] = v
    return s


def test_reference_shape_classes(rt, oracle):
    rec = np.load(os.path.join(GOLD, "kat_shapes.npz"))["rec"]
    n_inner = 0
    for r in rec:
        t = int(r["type"])
        if t == 0:
            s = _shape(rt, type=0, sphereCenter=r["a"], sphereRadius=r["s0"])
        elif t == 2:
            s = _shape(rt, type=2, planeNormal=r["nd"][:3], planeD=r["nd"][3], wallStart=r["a"], wallWidth=r["s0"], wallHeight=r["s1"])
        else:
            s = _shape(rt, type=3, planeNormal=r["nd"][:3], planeD=r["nd"][3], triP1=r["a"], triP2=r["b"], triP3=r["c"])
        res, hp = oracle.intersect_shape(s, r["start"], r["dir"], useMT=False)
        if int(r["res"]) == INNER:
            assert res == INNER
            assert np.array_equal(hp.view(np.uint32), r["hit"].view(np.uint32)), "hit point differs from the reference class"
            n_inner += 1
        else:
            assert res != INNER            # the shader returns early on OUTER; both mean "no hit"
    assert n_inner > 200


def test_unit_sphere(rt, oracle):
    s = _shape(rt, type=0, sphereCenter=(0, 0, 0), sphereRadius=1.0)
    res, hp = oracle.intersect_shape(s, (0, 0, 5), (0, 0, -1))
    assert res == INNER and np.allclose(hp, (0, 0, 1))
    res, _ = oracle.intersect_shape(s, (0, 0, 0), (0, 0, -1))      # from inside: only the far root -> OUTER
    assert res == OUTER
    res, _ = oracle.intersect_shape(s, (0, 0, 5), (0, 0, 1))       # behind
    assert res == NONE
    res, _ = oracle.intersect_shape(s, (0, 1, 5), (0, 0, -1))      # tangent: D == 0 is a miss (D > 0 strict)
    assert res == NONE


def test_scene3_triangle_is_back_face_only(rt, oracle):
    """Triangle (0,0,0),(5,0,0),(2.5,-5,0) has n=(0,0,-1): INNER needs dot(n,dir) > 0."""
    sc = rt.Scene("triangle", 64, 48)
    s = sc.shapes[0:1]
    assert np.allclose(s["planeNormal"][0], (0, 0, -1))
    res, hp = oracle.intersect_shape(s, (2.5, -2, 5), (0, 0, -1))
    assert res == INNER and np.allclose(hp, (2.5, -2, 0))
    res, _ = oracle.intersect_shape(s, (2.5, -2, -5), (0, 0, 1))
    assert res != INNER
    res, _ = oracle.intersect_shape(s, (2.5, -2, -5), (0, 0, 1), useMT=True)    # Moller-Trumbore is two-sided
    assert res == INNER
    res, _ = oracle.intersect_shape(s, (9, -2, 5), (0, 0, -1))                   # outside the triangle
    assert res == NONE


def test_nan_floor_wall_is_unbounded(rt, oracle):
    """normal (0,1,0): u = normalize(0) = NaN, the 1e-5 guard is false for NaN, bounds never reject."""
    s = _shape(rt, type=2, planeNormal=(0, 1, 0), planeD=-25.0, wallStart=(-100, 25, -100), wallWidth=210, wallHeight=210)
    res, hp = oracle.intersect_shape(s, (5000, 0, 7000), (0, 1, 0))              # far outside the 210x210 rectangle
    assert res == INNER and np.allclose(hp, (5000, 25, 7000))
    s2 = _shape(rt, type=2, planeNormal=(1, 0, 0), planeD=0.0, wallStart=(0, 0, 0), wallWidth=2, wallHeight=2)
    res, _ = oracle.intersect_shape(s2, (-5, 50, 50), (1, 0, 0))                 # a regular wall does clip
    assert res == NONE


def test_plane(rt, oracle):
    s = _shape(rt, type=1, planeNormal=(0, 0, 1), planeD=-2.0)                   # z = 2
    res, hp = oracle.intersect_shape(s, (0, 0, 0), (0, 0, 1))
    assert res == INNER and np.allclose(hp, (0, 0, 2))
    res, _ = oracle.intersect_shape(s, (0, 0, 5), (0, 0, -1))
    assert res == OUTER
    res, _ = oracle.intersect_shape(s, (0, 0, 0), (1, 0, 0))                     # parallel
    assert res == NONE


def test_present_quantisation(oracle):
    img = np.array([[[0.0, 1.0, 0.5, 0.3], [-1.0, 7.5, np.nan, 0.0]]], np.float32)
    out = oracle.present(img)
    assert out.tolist() == [[[0, 255, 128, 255], [0, 255, 0, 255]]]
