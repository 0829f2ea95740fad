"""CPU: the headless host library — scene contents and the reference BVH builder's invariants
(post-order, root last, siblings adjacent, junk leaf fields on interior nodes)."""
import numpy as np
import pytest


def test_scene_primitive_counts(rt):
    m = rt.Scene("monkey").shapes
    assert len(m) == 1240                                         # README.md:4
    assert (m["type"] == 3).sum() == 1 + 968 + 240 and (m["type"] == 0).sum() == 29 and (m["type"] == 2).sum() == 2
    c = rt.Scene("car").shapes
    assert (c["type"] == 3).sum() == 4022 and (c["type"] == 0).sum() == 100
    s = rt.Scene("soup", n_override=2000)
    assert len(s.shapes) == 2020
    assert len(rt.Scene("triangle").shapes) == 1


def test_scene_bytes_are_deterministic(rt):
    a, b = rt.Scene("car", 800, 600), rt.Scene("car", 800, 600)
    for k in ("light", "camera", "shapes", "nodes", "indices"):
        assert getattr(a, k).tobytes() == getattr(b, k).tobytes()


@pytest.mark.parametrize("name", ["toy", "monkey", "car"])
def test_bvh_invariants(rt, name):
    sc = rt.Scene(name)
    nodes, idx, shapes = sc.nodes, sc.indices, sc.shapes
    n = len(nodes)
    leaf = nodes["leftChild"] == -1
    inner = ~leaf
    # root is last; children precede parents; siblings adjacent
    assert (nodes["leftChild"][inner] < np.nonzero(inner)[0]).all()
    assert (nodes["rightChild"][inner] == nodes["leftChild"][inner] + 1).all()
    reached = np.zeros(n, bool)
    stack = [n - 1]
    while stack:
        k = stack.pop()
        assert not reached[k]
        reached[k] = True
        if not leaf[k]:
            stack += [int(nodes["leftChild"][k]), int(nodes["rightChild"][k])]
    assert reached.all()
    # leaves partition the bounded shapes
    got = np.concatenate([idx[s:s + c] for s, c in zip(nodes["startShapeIdx"][leaf], nodes["numShapes"][leaf])])
    assert sorted(got.tolist()) == list(range(len(shapes)))
    # interior nodes carry subtree sizes in numShapes (must not be used as the leaf test)
    assert nodes["numShapes"][-1] == len(shapes)
    # child boxes are inside parent boxes
    for k in np.nonzero(inner)[0]:
        for ch in (nodes["leftChild"][k], nodes["rightChild"][k]):
            assert (nodes["boundsMin"][ch] >= nodes["boundsMin"][k]).all() and (nodes["boundsMax"][ch] <= nodes["boundsMax"][k]).all()
    st = sc.bvh_stats()
    assert st["nodes"] == n and st["leaves"] == leaf.sum() and st["depth"] <= {"toy": 8, "monkey": 15, "car": 25}[name]


def test_refit_only_grows_and_keeps_topology(rt):
    sc = rt.Scene("monkey")
    n0 = sc.nodes.copy()
    sc.animate(0.7, 1 / 60, mode=0)
    n1 = sc.nodes
    for k in ("leftChild", "rightChild", "startShapeIdx", "numShapes"):
        assert np.array_equal(n0[k], n1[k])
    assert (n1["boundsMin"] <= n0["boundsMin"]).all() and (n1["boundsMax"] >= n0["boundsMax"]).all()
    assert (n1["boundsMax"] > n0["boundsMax"]).any()
    assert list(sc.animated) == [0, 1, 2]
    assert sc.shapes["sphereCenter"][0][1] != 10.0


def test_wheel_animation_keeps_stale_planes(rt):
    """updateWheelAnimations rotates a,b,c but not m_normal/d (src/main.cpp:1101-1106)."""
    sc = rt.Scene("car")
    s0 = sc.shapes.copy()
    assert sc.animate(0.1, 0.1, mode=0) == 960
    s1 = sc.shapes
    a = sc.animated
    assert not np.array_equal(s0["triP1"][a], s1["triP1"][a])
    assert np.array_equal(s0["planeNormal"][a], s1["planeNormal"][a]) and np.array_equal(s0["planeD"][a], s1["planeD"][a])


def test_camera_basis_is_orthonormal(rt):
    cam = rt.Scene("car").camera[0]
    F, U, R = cam["Front"], cam["Up"], cam["Right"]
    for v in (F, U, R):
        assert abs(np.linalg.norm(v) - 1) < 1e-6
    assert abs(F @ U) < 1e-6 and abs(F @ R) < 1e-6 and abs(U @ R) < 1e-6
    assert cam["fov"] == 60 and abs(cam["aspectRatio"] - 800 / 600) < 1e-6


def test_animation_list_reproduces_the_host_animation(rt):
    """Scene.animations() (what rt_anim_bounce / rt_anim_rotate_z are fed) carries the parameters of the host's own
    animation step: applying the reference formulas (main.cpp:1079-1109; the wheels through GLM's translate * rotate *
    translate and mat4 * vec4) to them with libm's sinf / cosf and fp32 arithmetic gives the animated shapes of Scene.animate
    bit for bit — the arithmetic rt_animate / K6 perform."""
    import ctypes as C
    m = C.CDLL("libm.so.6")
    m.sinf.restype = C.c_float; m.sinf.argtypes = [C.c_float]
    m.cosf.restype = C.c_float; m.cosf.argtypes = [C.c_float]
    f32 = np.float32
    # bouncing spheres
    sc = rt.Scene("monkey", 200, 150)
    an = sc.animations()
    assert sorted(int(i[0]) for _, i, _ in an) == sorted(int(x) for x in sc.animated)
    t, dt = f32(0.7), f32(1.0 / 60.0)
    sc.animate(float(t), float(dt), mode=0)
    for kind, idx, p in an:
        assert kind == 0
        want = f32(p[0]) + f32(p[1]) * f32(m.sinf(f32(p[2]) * t))
        got = sc.shapes["sphereCenter"][int(idx[0])][1]
        assert f32(want).tobytes() == f32(got).tobytes()
    # rotating wheels (vertices only: the planes stay stale, as upstream)
    sc = rt.Scene("car", 200, 150)
    before = sc.shapes.copy()
    an = sc.animations()
    assert sorted(int(x) for _, i, _ in an for x in i) == sorted(int(x) for x in sc.animated)
    sc.animate(0.3, 0.3, mode=0)
    after = sc.shapes
    def wheel_matrix(c, cs, sn):
        """translate(I, c) * rotate(I, angle, (0,0,1)) * translate(I, -c) in GLM 0.9.8.5's operation order, fp32 (the
        restatement host/rt_anim_math.h is pinned to GLM by test_wheel_transform_restatement_equals_glm)."""
        I = np.eye(4, dtype=f32)                                  # M[row, col]
        def translate(m, v):
            out = m.copy()
            out[:, 3] = ((m[:, 0] * v[0] + m[:, 1] * v[1]) + m[:, 2] * v[2]) + m[:, 3]
            return out
        ax = np.array([0, 0, 1], f32) * (f32(1.0) / np.sqrt(f32(1.0)))
        omc = f32(1.0) - cs
        tt = omc * ax
        R = np.zeros((3, 3), f32)                                 # R[col][k] as in gtc/matrix_transform.inl
        R[0, 0] = cs + tt[0] * ax[0]; R[0, 1] = tt[0] * ax[1] + sn * ax[2]; R[0, 2] = tt[0] * ax[2] - sn * ax[1]
        R[1, 0] = tt[1] * ax[0] - sn * ax[2]; R[1, 1] = cs + tt[1] * ax[1]; R[1, 2] = tt[1] * ax[2] + sn * ax[0]
        R[2, 0] = tt[2] * ax[0] + sn * ax[1]; R[2, 1] = tt[2] * ax[1] - sn * ax[0]; R[2, 2] = cs + tt[2] * ax[2]
        rot = I.copy()
        for col in range(3):
            rot[:, col] = (I[:, 0] * R[col, 0] + I[:, 1] * R[col, 1]) + I[:, 2] * R[col, 2]
        def mul(a, b):
            out = np.zeros((4, 4), f32)
            for col in range(4):
                out[:, col] = ((a[:, 0] * b[0, col] + a[:, 1] * b[1, col]) + a[:, 2] * b[2, col]) + a[:, 3] * b[3, col]
            return out
        return mul(mul(translate(I, c), rot), translate(I, -c))

    for kind, idx, p in an:
        assert kind == 1
        ang = f32(p[3]) * f32(0.3)
        cs, sn = f32(m.cosf(ang)), f32(m.sinf(ang))
        M = wheel_matrix(np.array(p[:3], f32), cs, sn)
        for key in ("triP1", "triP2", "triP3"):
            v = before[key][idx].astype(f32)
            for r in range(3):                                    # vec3(M * vec4(p, 1)): (m0 x + m1 y) + (m2 z + m3 1)
                want = (M[r, 0] * v[:, 0] + M[r, 1] * v[:, 1]) + (M[r, 2] * v[:, 2] + M[r, 3] * f32(1.0))
                assert want.astype(f32).tobytes() == after[key][idx][:, r].astype(f32).tobytes(), (key, r)
        assert np.array_equal(before["planeNormal"][idx], after["planeNormal"][idx])


@pytest.mark.parametrize("name", ["monkey", "car", "toy"])
def test_leaf_growth_plus_child_union_equals_updateBVH(rt, name):
    """K7's rule (grow every leaf by its animated shapes, then make every inner node contain its children) gives the
    node array of the reference's updateBVH (main.cpp:1068-1077), which grows EVERY node by the animated shapes it
    holds — true for trees whose parents contain their children, which the reference builder's do."""
    f32 = np.float32
    sc = rt.Scene(name, 200, 150)
    nodes0 = sc.nodes.copy()
    idx = sc.indices
    animated = set(int(x) for x in sc.animated)
    for frame in range(1, 4):
        sc.animate(frame * 0.4, 0.4, mode=0)
        shapes, want = sc.shapes, sc.nodes
        mn = nodes0["boundsMin"].astype(f32).copy(); mx = nodes0["boundsMax"].astype(f32).copy()
        for n in range(len(nodes0)):                               # post-order array: children come before parents
            if nodes0["leftChild"][n] == -1:
                s0, c = int(nodes0["startShapeIdx"][n]), int(nodes0["numShapes"][n])
                for s in idx[s0:s0 + c]:
                    if int(s) not in animated:
                        continue
                    sh = shapes[int(s)]
                    if sh["type"] == 0:
                        r = f32(sh["sphereRadius"]); ctr = sh["sphereCenter"].astype(f32)
                        pts = [ctr + r, ctr - r]
                    elif sh["type"] == 3:
                        pts = [sh["triP1"].astype(f32), sh["triP2"].astype(f32), sh["triP3"].astype(f32)]
                    else:
                        pts = []
                    for p in pts:
                        mn[n] = np.minimum(mn[n], p); mx[n] = np.maximum(mx[n], p)
            else:
                for ch in (int(nodes0["leftChild"][n]), int(nodes0["rightChild"][n])):
                    mn[n] = np.minimum(mn[n], mn[ch]); mx[n] = np.maximum(mx[n], mx[ch])
        assert mn.tobytes() == want["boundsMin"].astype(f32).tobytes(), "frame %d" % frame
        assert mx.tobytes() == want["boundsMax"].astype(f32).tobytes(), "frame %d" % frame
        nodes0 = want.copy()


def test_wheel_transform_restatement_equals_glm():
    """host/rt_anim_math.h (prepared for K6: the reference's translate * rotate * translate matrix and the matrix-vector
    product, operation by operation in GLM's order) against the reference's vendored GLM evaluating the very expression of
    updateWheelAnimations (src/main.cpp:1091-1105): matrix and transformed points bit for bit, any axis."""
    import ctypes as C
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ref = os.path.join(root, "oracle", "_ref", "libref_anim.so")
    if not os.path.exists(ref):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    src = os.path.join(root, "tests", "emu", "anim_math_emu.cpp")
    so = os.path.join(root, "tests", "emu", "libanim_math_emu.so")
    hdr = os.path.join(root, "opengl-ray-tracer_b200", "host", "rt_anim_math.h")
    if not os.path.exists(so) or max(os.path.getmtime(src), os.path.getmtime(hdr)) > os.path.getmtime(so):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-o", so, src])
    R, E = C.CDLL(ref), C.CDLL(so)
    sig = [C.c_void_p, C.c_float, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    R.refanim_wheel.argtypes = sig; E.rta_wheel.argtypes = sig
    rng = np.random.default_rng(77)
    for trial in range(200):
        center = rng.uniform(-15, 15, 3).astype(np.float32)
        axis = np.array([0, 0, 1], np.float32) if trial % 2 == 0 else rng.normal(size=3).astype(np.float32)
        angle = np.float32(rng.uniform(-0.2, 0.2) if trial % 3 else (trial + 1) / 60.0)
        pts = rng.uniform(-20, 20, (64, 3)).astype(np.float32)
        a, b = np.zeros_like(pts), np.zeros_like(pts)
        ma, mb = np.zeros(16, np.float32), np.zeros(16, np.float32)
        R.refanim_wheel(center.ctypes.data, angle, axis.ctypes.data, len(pts), pts.ctypes.data, a.ctypes.data, ma.ctypes.data)
        E.rta_wheel(center.ctypes.data, angle, axis.ctypes.data, len(pts), pts.ctypes.data, b.ctypes.data, mb.ctypes.data)
        assert np.array_equal(ma.view(np.uint32), mb.view(np.uint32)), "transform matrix, trial %d" % trial
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), "transformed points, trial %d" % trial
