"""The builder, the serialiser and the animated frame step pinned to UPSTREAM CODE: oracle/_ref/libref_bvh.so is the reference's
own class Node / split / buildBVH / serializeBVH / serializeShape / updateScene / updateBVH / bounceSphere /
updateWheelAnimations, cut out of src/main.cpp mechanically by oracle/build_ref.py and compiled in place.  The repo's host
side (host/rt_host.cpp: what feeds the kernel, the oracle and both bench arms) must produce the same bytes.

Compared per FIELD: upstream leaves the std430 padding words and the fields of foreign shape types indeterminate.
"""
import numpy as np
import pytest

from util import random_shapes

NODE_FIELDS = ("boundsMin", "boundsMax", "leftChild", "rightChild", "startShapeIdx", "numShapes")
SHAPE_FIELDS = {0: ("sphereCenter", "sphereRadius"), 1: ("planeNormal", "planeD"),
                2: ("planeNormal", "planeD", "wallStart", "wallWidth", "wallHeight"),
                3: ("planeNormal", "planeD", "triP1", "triP2", "triP3")}
MATERIAL = ("type", "color", "fresnelStrength", "ambientStrength", "diffuseStrength", "specularStrength", "shininess")


@pytest.fixture(scope="module")
def ref(oracle):
    if oracle.refbvh_lib() is None:
        pytest.skip("oracle/_ref/libref_bvh.so not built (no reference tree here)")
    return oracle


def same_nodes(a, b):
    return len(a) == len(b) and all(np.array_equal(a[f].view(np.uint32), b[f].view(np.uint32)) for f in NODE_FIELDS)


def same_shapes(a, b):
    if len(a) != len(b) or not np.array_equal(a["type"], b["type"]):
        return False
    for f in MATERIAL:
        if not np.array_equal(a[f].view(np.uint32), b[f].view(np.uint32)):
            return False
    for t, fields in SHAPE_FIELDS.items():
        m = a["type"] == t
        for f in fields:
            if not np.array_equal(a[f][m].view(np.uint32), b[f][m].view(np.uint32)):
                return False
    return True


@pytest.mark.parametrize("name,n", [("toy", 0), ("triangle", 0), ("monkey", 0), ("car", 0), ("soup", 30000)])
def test_host_builder_equals_upstream_on_the_named_scenes(rt, ref, name, n):
    sc = rt.Scene(name, 800, 600, n_override=n)
    depth = {"toy": 8, "car": 25, "soup": 40}.get(name, 15)
    R = ref.RefScene(sc.shapes, node_dtype=rt.NODE_DTYPE)
    nodes, idx = R.build(depth)
    assert same_nodes(nodes, sc.nodes), name
    assert np.array_equal(idx, sc.indices)
    assert same_shapes(R.shapes(), sc.shapes)


@pytest.mark.parametrize("seed", range(6))
def test_host_builder_equals_upstream_on_random_scenes(rt, ref, seed):
    """Spheres, walls and triangles (an infinite plane's split centre is indeterminate upstream: main.cpp:1123 leaves `center`
    unset for it, so planes have no defined place in the tree), coincident centres, signed zeros, depth budgets that cut
    the recursion short."""
    rng = np.random.default_rng(100 + seed)
    base = rt.Scene("toy", 64, 48)
    n = int(rng.integers(1, 400))
    shapes = random_shapes(rt, rng, n, types=(0, 2, 3), extent=float(rng.choice([0.5, 10.0, 1000.0])))
    if seed % 2:                                     # duplicates: equal centres end up on the same side, a split may come out empty
        k = rng.integers(0, n, n // 3)
        shapes[k] = shapes[rng.integers(0, n, n // 3)]
    if seed == 3:
        shapes["sphereCenter"][shapes["type"] == 0] *= np.float32(-0.0)
    depth = int(rng.choice([1, 3, 15, 40]))
    sc = rt.Scene.from_shapes(shapes, base.light, base.camera, bvh_depth=depth)
    nodes, idx = ref.RefScene(shapes, node_dtype=rt.NODE_DTYPE).build(depth)
    assert same_nodes(nodes, sc.nodes)
    assert np.array_equal(idx, sc.indices)


def test_animated_monkey_equals_upstream(rt, ref):
    """bounceSphere + updateScene + updateBVH | buildBVH + serializeBVH, 40 frames (src/main.cpp:328-344, 1068-1082)."""
    for rebuild in (False, True):
        sc = rt.Scene("monkey", 800, 600)
        R = ref.RefScene(sc.shapes, animated=sc.animated, node_dtype=rt.NODE_DTYPE)
        R.build(15)
        anims = sc.animations()
        for i in range(40):
            t, dt = i / 60.0, 1.0 / 60.0
            sc.animate(t, dt, 1 if rebuild else 0)
            for kind, idx, par in anims:
                assert kind == 0
                R.bounce(int(idx[0]), t, par[1], par[2])
            nodes, ind = R.frame(rebuild, 15)
            assert same_shapes(R.shapes(), sc.shapes), "frame %d" % i
            assert same_nodes(nodes, sc.nodes), "frame %d rebuild=%s" % (i, rebuild)
            assert np.array_equal(ind, sc.indices)


def test_animated_car_wheels_equal_upstream(rt, ref):
    """updateWheelAnimations (GLM translate * rotate * translate, mat4 * vec4; src/main.cpp:1084-1109) over 60 frames: the
    incremental rotation accumulates, so any ulp of difference in one frame shows in all later ones.  Stale planes included."""
    for rebuild in (False, True):
        sc = rt.Scene("car", 800, 600)
        R = ref.RefScene(sc.shapes, animated=sc.animated, node_dtype=rt.NODE_DTYPE)
        R.build(25)
        for k, (kind, idx, par) in enumerate(sc.animations()):
            assert kind == 1 and par[3] == 1.0                      # rotationSpeed = 1 (main.cpp:1085)
            R.wheel(k, idx, par[:3])
        for i in range(60):
            t, dt = i * 0.013, 0.013 + 0.001 * (i % 5)
            sc.animate(t, dt, 1 if rebuild else 0)
            R.wheels_step(t, dt)
            nodes, ind = R.frame(rebuild, 25)
            assert same_shapes(R.shapes(), sc.shapes), "frame %d" % i
            assert same_nodes(nodes, sc.nodes), "frame %d rebuild=%s" % (i, rebuild)
            assert np.array_equal(ind, sc.indices)
