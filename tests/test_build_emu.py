"""CPU: the device BVH builder K8 (opengl-ray-tracer_b200/csrc/rt_build.cuh).  tests/emu/build_emu.cpp compiles the
per-thread bodies of its kernels with g++ and runs them one thread at a time, kernel after kernel; the emitted
FlatNode[] and bvhIndices[] must equal the host restatement of the reference builder (buildBVH / split /
serializeBVH, src/main.cpp:955-979, 1111-1193; opengl-ray-tracer_b200/host/rt_host.cpp) BYTE FOR BYTE, whatever the
order the threads of a kernel run in."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from util import random_shapes

HERE = os.path.dirname(os.path.abspath(__file__))
EMU_SRC = os.path.join(HERE, "emu", "build_emu.cpp")
EMU_SO = os.path.join(HERE, "emu", "libbuild_emu.so")
CSRC = os.path.join(os.path.dirname(HERE), "opengl-ray-tracer_b200", "csrc")


@pytest.fixture(scope="module")
def emu():
    deps = [EMU_SRC, os.path.join(CSRC, "rt_build.cuh"), os.path.join(CSRC, "rt_build_list.cuh"), os.path.join(CSRC, "rt_device.cuh"),
            os.path.join(CSRC, "rt_layout.h")]
    if not os.path.exists(EMU_SO) or any(os.path.getmtime(d) > os.path.getmtime(EMU_SO) for d in deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared",
                               "-I/usr/local/cuda/include", "-o", EMU_SO, EMU_SRC])
    L = C.CDLL(EMU_SO)
    L.emu_build_bvh.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_uint]
    L.emu_build_bvh.restype = C.c_int
    vp = C.c_void_p
    L.emu_build_bvh_list.argtypes = [vp, C.c_int, C.c_int, vp, vp, C.c_uint, vp, vp, vp, vp]; L.emu_build_bvh_list.restype = C.c_int
    L.emu_shape_centres.argtypes = [vp, C.c_int, vp]; L.emu_shape_centres.restype = None
    L.emu_host_list.argtypes = [vp, C.c_int, vp, C.c_int, vp, C.c_int, vp, vp, vp, vp]; L.emu_host_list.restype = C.c_int
    return L


def device_build(emu, rt, shapes, depth, order=0):
    shapes = np.ascontiguousarray(shapes)
    n = len(shapes)
    nodes = np.zeros(2 * n, dtype=rt.NODE_DTYPE)
    idx = np.full(n, -7, dtype=np.int32)
    nn = emu.emu_build_bvh(shapes.ctypes.data, n, depth, nodes.ctypes.data, idx.ctypes.data, order)
    assert 1 <= nn <= 2 * n - 1
    return nodes[:nn], idx


def check(emu, rt, shapes, depth, orders=(0, 1, 7)):
    ref = rt.Scene.from_shapes(shapes, np.zeros(1, rt.LIGHT_DTYPE), np.zeros(1, rt.CAMERA_DTYPE), bvh_depth=depth)
    want_nodes, want_idx = ref.nodes, ref.indices
    for order in orders:
        nodes, idx = device_build(emu, rt, shapes, depth, order)
        assert len(nodes) == len(want_nodes), (order, len(nodes), len(want_nodes))
        assert nodes.tobytes() == want_nodes.tobytes(), "node bytes differ (thread order %d)" % order
        assert idx.tobytes() == want_idx.tobytes(), "bvhIndices differ (thread order %d)" % order
    return ref


@pytest.mark.parametrize("name,n,depth", [("toy", 0, 0), ("triangle", 0, 0), ("monkey", 0, 0), ("car", 0, 0),
                                          ("soup", 20000, 0), ("monkey", 0, 3), ("car", 0, 40), ("monkey", 0, 1)])
def test_scene_trees_are_byte_identical(rt, emu, name, n, depth):
    sc = rt.Scene(name, 160, 120, n_override=n, bvh_depth=depth)
    d = depth if depth > 0 else {"toy": None, "triangle": None, "monkey": 15, "car": 25, "soup": 40}[name]
    if d is None:                      # the scene's own default: recover it from a rebuild with a known depth below
        d = 15
        ref = check(emu, rt, sc.shapes, d)
        return
    ref = check(emu, rt, sc.shapes, d)
    assert ref.nodes.tobytes() == sc.nodes.tobytes() and ref.indices.tobytes() == sc.indices.tobytes()


def test_rebuild_follows_the_animation(rt, emu):
    """C5's rebuild mode: after every animated frame the tree built from the moved shapes equals the host's rebuild."""
    for name, depth in (("monkey", 15), ("car", 25)):
        sc = rt.Scene(name, 160, 120)
        for f in range(1, 6):
            sc.animate(f / 60.0, 1 / 60.0, 1)
            nodes, idx = device_build(emu, rt, sc.shapes, depth, order=f)
            assert nodes.tobytes() == sc.nodes.tobytes() and idx.tobytes() == sc.indices.tobytes(), (name, f)


@pytest.mark.parametrize("seed", range(6))
def test_random_shapes_all_types(rt, emu, seed):
    rng = np.random.default_rng(100 + seed)
    n = int(rng.integers(1, 400))
    sh = random_shapes(rt, rng, n, types=(0, 1, 2, 3), extent=float(rng.choice([0.5, 10.0, 1e3])))
    check(emu, rt, sh, int(rng.integers(1, 20)))


def test_edge_cases(rt, emu):
    rng = np.random.default_rng(5)
    # one and two shapes, depth budgets 0 / 1
    for n in (1, 2, 3):
        sh = random_shapes(rt, rng, n, types=(0, 3))
        for depth in (1, 2, 64):
            check(emu, rt, sh, depth)
        nodes, idx = device_build(emu, rt, sh, 0)          # no depth budget: the root is the only node (split, :1113-1118)
        assert len(nodes) == 1 and nodes["leftChild"][0] == -1 and nodes["numShapes"][0] == n and idx.tolist() == list(range(n))
    # coincident centres: one side of the split stays empty, the node is a (fat) leaf
    sh = random_shapes(rt, rng, 40, types=(0,))
    sh["sphereCenter"][:] = sh["sphereCenter"][0]
    check(emu, rt, sh, 10)
    # signed zeros: bounds that are +0 in one shape and -0 in another keep the sign of the first in list order
    sh = random_shapes(rt, rng, 64, types=(3,))
    for k, f in enumerate(("triP1", "triP2", "triP3")):
        v = sh[f].copy()
        v[rng.uniform(size=v.shape) < 0.3] = 0.0
        v[rng.uniform(size=v.shape) < 0.15] = -0.0
        v = np.minimum(v, 0.0) if k == 0 else v
        sh[f] = v
    check(emu, rt, sh, 12)
    sh2 = sh.copy()
    sh2["triP1"] = np.abs(sh2["triP1"]) * 0.0 - np.where(rng.uniform(size=sh2["triP1"].shape) < 0.5, 0.0, -0.0)
    check(emu, rt, sh2, 12)
    # non-finite coordinates: a triangle with a non-finite x is skipped as a whole, NaN in y / z is ignored per
    # component, infinities are kept; a NaN wall basis (normal parallel to no axis test) gives NaN corners
    sh = random_shapes(rt, rng, 80, types=(0, 2, 3))
    tri = np.nonzero(sh["type"] == 3)[0]
    sh["triP1"][tri[0], 0] = np.nan
    sh["triP2"][tri[1], 0] = np.inf
    sh["triP3"][tri[2], 1] = np.nan
    sh["triP1"][tri[3], 2] = np.inf
    sh["triP2"][tri[4], 1] = -np.inf
    wall = np.nonzero(sh["type"] == 2)[0]
    sh["planeNormal"][wall[0]] = (0.0, 0.0, 0.0)
    sph = np.nonzero(sh["type"] == 0)[0]
    sh["sphereRadius"][sph[0]] = np.nan
    sh["sphereCenter"][sph[1], 1] = np.inf
    check(emu, rt, sh, 12)
    # infinite planes only: nothing bounds the root, nothing splits
    check(emu, rt, random_shapes(rt, rng, 5, types=(1,)), 8)


@pytest.mark.parametrize("name,n,depth,order", [("toy", 0, 15, 0), ("monkey", 0, 15, 3), ("car", 0, 25, 0), ("soup", 20000, 40, 5), ("monkey", 0, 2, 1)])
def test_list_a_from_the_build_tables(rt, emu, name, n, depth, order):
    """K9 (rt_build_list.cuh, prepared for the next round): the threaded node list and the primitive slots that
    rtl::build_layout derives on the host from FlatNode[] / bvhIndices[] (plain reference walk, no sub-trees) follow
    word for word from K8's own tables, node by node, in any thread order."""
    sc = rt.Scene(name, 160, 120, n_override=n, bvh_depth=depth)
    shapes = np.ascontiguousarray(sc.shapes)
    ns = len(shapes)
    nodes = np.zeros(2 * ns, dtype=rt.NODE_DTYPE); idx = np.zeros(ns, np.int32)
    src = np.full(2 * ns, -9, np.int32); lc = np.full((2 * ns, 2), -9, np.int32)
    psh = np.full(ns, -9, np.int32); prk = np.full(ns, -9, np.int32)
    nn = emu.emu_build_bvh_list(shapes.ctypes.data, ns, depth, nodes.ctypes.data, idx.ctypes.data, order,
                                src.ctypes.data, lc.ctypes.data, psh.ctypes.data, prk.ctypes.data)
    assert nodes[:nn].tobytes() == np.ascontiguousarray(sc.nodes).tobytes()
    hsrc = np.zeros(nn, np.int32); hlc = np.zeros((nn, 2), np.int32); hpsh = np.zeros(ns, np.int32); hprk = np.zeros(ns, np.int32)
    hn = np.ascontiguousarray(sc.nodes); hi = np.ascontiguousarray(sc.indices)
    m = emu.emu_host_list(shapes.ctypes.data, ns, hn.ctypes.data, len(hn), hi.ctypes.data, len(hi),
                          hsrc.ctypes.data, hlc.ctypes.data, hpsh.ctypes.data, hprk.ctypes.data)
    assert m == nn
    assert np.array_equal(src[:nn], hsrc) and np.array_equal(lc[:nn], hlc)
    assert np.array_equal(psh, hpsh) and np.array_equal(prk, hprk)


def test_property_random_bit_patterns(rt, emu):
    """Property test (hypothesis): shapes whose coordinates are arbitrary fp32 BIT PATTERNS — denormals, huge values,
    infinities, NaNs with payloads, both zeros — in any mix of the four types (and an unknown one): the emulated device
    build still equals the host builder byte for byte."""
    hyp = pytest.importorskip("hypothesis")
    from hypothesis import given, settings, strategies as st, HealthCheck

    special = [0x00000000, 0x80000000, 0x7f800000, 0xff800000, 0x7fc00000, 0xffc00001, 0x00000001, 0x80000001,
               0x7f7fffff, 0xff7fffff, 0x3f800000, 0xbf800000]
    word = st.one_of(st.sampled_from(special), st.integers(0, 2 ** 32 - 1),
                     st.floats(-50, 50, width=32).map(lambda f: int(np.float32(f).view(np.uint32))))

    @settings(max_examples=60, deadline=None, suppress_health_check=[HealthCheck.too_slow])
    @given(n=st.integers(1, 40), depth=st.integers(1, 12), data=st.data())
    def run(n, depth, data):
        sh = np.zeros(n, dtype=rt.SHAPE_DTYPE)
        raw = sh.view(np.uint32).reshape(n, 48)
        for i in range(n):
            sh["type"][i] = data.draw(st.sampled_from([0, 0, 1, 2, 3, 3, 3, 7]))
            for w in (16, 17, 18, 19, 20, 21, 22, 23, 24, 25, 26, 27, 28, 36, 37, 38, 40, 41, 42, 44, 45, 46):
                raw[i, w] = data.draw(word)      # centre, radius, normal, D, wall start / width / height, the three vertices
        check(emu, rt, sh, depth, orders=(0, 5))

    with np.errstate(all="ignore"):
        run()


@pytest.mark.parametrize("seed", range(6))
def test_boxes_and_centres_are_the_reference_classes(rt, emu, seed):
    """Pins the arithmetic the builder is made of to the REFERENCE'S OWN CODE (oracle/_ref/libref_shapes.so compiles
    src/BoundingBox.hpp, shapes/wall.hpp, shapes/triangle.hpp, shapes/sphere.hpp in place): one reference BoundingBox grown
    by random spheres / walls / triangles in list order through growToInclude(unique_ptr<Shape>&) — the call buildBVH
    makes — must equal, bit for bit, the root box of the host builder and of the emulated device build; split's centres
    (sphere centre, (start + Wall::end()) / 2, Triangle::center()) must equal K8's per-shape records."""
    ref = os.path.join(os.path.dirname(HERE), "oracle", "_ref", "libref_shapes.so")
    if not os.path.exists(ref):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    R = C.CDLL(ref)
    R.refbbox_grow.argtypes = [C.c_int] + [C.c_void_p] * 5; R.refbbox_grow.restype = None
    rng = np.random.default_rng(4000 + seed)
    n = int(rng.integers(1, 200))
    types = rng.choice([0, 2, 3], size=n).astype(np.int32)
    par = np.zeros((n, 12), np.float32)
    scale = float(rng.choice([1.0, 30.0, 1e-3]))
    par[:, 0:9] = rng.uniform(-scale, scale, (n, 9))
    par[:, 9:11] = rng.uniform(0.1, 5.0, (n, 2)) * scale
    if seed % 2:                                   # signed zeros and axis-aligned walls
        par[rng.uniform(size=par.shape) < 0.15] = 0.0
        par[rng.uniform(size=par.shape) < 0.08] = -0.0
        par[:, 9:11] = np.abs(par[:, 9:11]) + 0.5
        w = types == 2
        par[w, 3:6] = np.where(np.abs(par[w, 3:6]).sum(axis=1, keepdims=True) == 0, 1.0, par[w, 3:6])   # a wall needs a normal
    box = np.zeros(6, np.float32); nd = np.zeros((n, 4), np.float32); cen = np.zeros((n, 3), np.float32)
    R.refbbox_grow(n, types.ctypes.data, par.ctypes.data, box.ctypes.data, nd.ctypes.data, cen.ctypes.data)
    sh = np.zeros(n, dtype=rt.SHAPE_DTYPE)
    sh["type"] = types
    for i in range(n):                              # what serializeShape writes (src/main.cpp:994-1066)
        if types[i] == 0:
            sh["sphereCenter"][i] = par[i, 0:3]; sh["sphereRadius"][i] = par[i, 9]
        elif types[i] == 2:
            sh["wallStart"][i] = par[i, 0:3]; sh["wallWidth"][i] = par[i, 9]; sh["wallHeight"][i] = par[i, 10]
            sh["planeNormal"][i] = nd[i, 0:3]; sh["planeD"][i] = nd[i, 3]
        else:
            sh["triP1"][i], sh["triP2"][i], sh["triP3"][i] = par[i, 0:3], par[i, 3:6], par[i, 6:9]
            sh["planeNormal"][i] = nd[i, 0:3]; sh["planeD"][i] = nd[i, 3]
    host = rt.Scene.from_shapes(sh, np.zeros(1, rt.LIGHT_DTYPE), np.zeros(1, rt.CAMERA_DTYPE), bvh_depth=6)
    root = host.nodes[-1]
    got = np.concatenate([root["boundsMin"], root["boundsMax"]]).astype(np.float32)
    assert got.view(np.uint32).tolist() == box.view(np.uint32).tolist(), "host builder's root box vs the reference BoundingBox"
    nodes, _ = device_build(emu, rt, sh, 6, order=seed)
    got = np.concatenate([nodes[-1]["boundsMin"], nodes[-1]["boundsMax"]]).astype(np.float32)
    assert got.view(np.uint32).tolist() == box.view(np.uint32).tolist(), "emulated device build's root box vs the reference BoundingBox"
    k8 = np.zeros((n, 3), np.float32)
    emu.emu_shape_centres(np.ascontiguousarray(sh).ctypes.data, n, k8.ctypes.data)
    assert np.array_equal(k8.view(np.uint32), cen.view(np.uint32)), "K8's split centres vs sphere centre / Wall::end / Triangle::center"
