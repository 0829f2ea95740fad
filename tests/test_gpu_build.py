"""GPU: K8, the reference's BVH builder on the device (rt_build_bvh, opengl-ray-tracer_b200/csrc/rt_build.cuh), through
the C ABI.  The emitted FlatNode[] / bvhIndices[] must equal the host builder's (buildBVH / split / serializeBVH,
src/main.cpp:955-979, 1111-1193, restated in host/rt_host.cpp) BYTE FOR BYTE, and a frame rendered over the
device-built tree must be bit-identical to one rendered over the uploaded host tree."""
import numpy as np
import pytest

from util import bit_fuzzed_shapes, random_shapes

pytestmark = pytest.mark.gpu

NO_AGG = 1      # RT_BUILD_NO_WARP_AGGREGATION
NO_SINGLE = 2   # RT_BUILD_NO_SINGLE_BLOCK: scenes of <= 2048 shapes do not take the single-block path
NO_CLUSTER = 8  # RT_BUILD_NO_CLUSTER: scenes of <= 28000 shapes do not take the single-cluster path
NO_COOP = 16    # RT_BUILD_NO_COOPERATIVE: scenes of <= 120000 shapes do not take the single cooperative launch
# default (single block / cluster / cooperative grid / per-phase by size), cluster where it fits, cooperative grid (2049 .. 120000 shapes), per-phase
# launches, per-phase launches without aggregation
PATHS = [0, NO_SINGLE, NO_SINGLE | NO_CLUSTER, NO_SINGLE | NO_CLUSTER | NO_COOP, NO_SINGLE | NO_CLUSTER | NO_COOP | NO_AGG]


def _same_frame(a, b):
    return np.array_equal(a[1], b[1]) and np.array_equal(a[0].view(np.uint32), b[0].view(np.uint32))


@pytest.mark.parametrize("name,n,depth", [("toy", 0, 15), ("monkey", 0, 15), ("car", 0, 25), ("soup", 60000, 40), ("monkey", 0, 2),
                                          ("soup", 9000, 40), ("soup", 20000, 40), ("soup", 27000, 40), ("soup", 110000, 40)])   # clusters of 16 blocks, 1 - 2 positions per thread
@pytest.mark.parametrize("flags", PATHS)
def test_device_tree_is_the_host_tree(rt, name, n, depth, flags):
    W, H = 200, 150
    sc = rt.Scene(name, W, H, n_override=n, bvh_depth=depth)
    p = rt.make_params(W, H)
    a = rt.RayTracer(W, H); a.upload(sc)
    want = a.render(p)
    b = rt.RayTracer(W, H)
    b.upload_scene(sc.light, sc.camera, sc.shapes)                # shapes only: no host tree ever reaches this context
    nodes, idx = b.build_bvh(depth, flags)
    assert len(nodes) == len(sc.nodes) and nodes.tobytes() == np.ascontiguousarray(sc.nodes).tobytes(), "FlatNode[] differs"
    assert idx.tobytes() == np.ascontiguousarray(sc.indices).tobytes(), "bvhIndices[] differs"
    assert _same_frame(b.render(p), want)
    # rebuilding over an uploaded tree of the same topology keeps the layout and the frame
    nodes2, idx2 = a.build_bvh(depth, flags)
    assert nodes2.tobytes() == nodes.tobytes() and idx2.tobytes() == idx.tobytes()
    assert _same_frame(a.render(p), want)
    a.close(); b.close()


@pytest.mark.parametrize("name,depth", [("monkey", 15), ("car", 25)])
def test_rebuild_per_frame_equals_host_rebuild(rt, name, depth):
    """BASELINE configs[4] in rebuild mode with nothing but a few scalars crossing PCIe: rt_animate moves the shapes on
    the device (K6), rt_build_bvh rebuilds the tree there (K8); the host animates and rebuilds (buildBVH) beside it."""
    W, H = 200, 150
    sc = rt.Scene(name, W, H)
    p = rt.make_params(W, H)
    a = rt.RayTracer(W, H); b = rt.RayTracer(W, H)
    a.upload(sc); b.upload(sc)
    b.register_animations(sc)
    for frame in range(1, 5):
        t, dt = frame * 0.21, 0.21
        sc.animate(t, dt, mode=1)
        a.upload_scene(None, None, sc.shapes, sc.nodes, sc.indices)
        b.animate(t, dt)
        nodes, idx = b.build_bvh(depth, PATHS[frame % len(PATHS)])
        assert nodes.tobytes() == np.ascontiguousarray(sc.nodes).tobytes(), "nodes, frame %d" % frame
        assert idx.tobytes() == np.ascontiguousarray(sc.indices).tobytes(), "indices, frame %d" % frame
        assert _same_frame(a.render(p), b.render(p)), "frame %d" % frame
    a.close(); b.close()


def test_device_build_edge_cases(rt):
    """Every shape type incl. infinite planes, coincident centres, signed zeros, non-finite coordinates, depth budgets."""
    rng = np.random.default_rng(11)
    light = np.zeros(1, rt.LIGHT_DTYPE); cam = np.zeros(1, rt.CAMERA_DTYPE)
    cases = []
    for n in (1, 2, 3, 33, 700):
        cases.append((random_shapes(rt, rng, n, types=(0, 1, 2, 3)), int(rng.integers(1, 20))))
    sh = random_shapes(rt, rng, 40, types=(0,)); sh["sphereCenter"][:] = sh["sphereCenter"][0]
    cases.append((sh, 10))
    sh = random_shapes(rt, rng, 300, types=(3,))
    for f in ("triP1", "triP2", "triP3"):
        v = sh[f].copy()
        v[rng.uniform(size=v.shape) < 0.3] = 0.0
        v[rng.uniform(size=v.shape) < 0.15] = -0.0
        sh[f] = v
    cases.append((sh, 12))
    sh = random_shapes(rt, rng, 200, types=(0, 2, 3))
    tri = np.nonzero(sh["type"] == 3)[0]
    sh["triP1"][tri[0], 0] = np.nan; sh["triP2"][tri[1], 0] = np.inf; sh["triP3"][tri[2], 1] = np.nan
    sh["triP1"][tri[3], 2] = np.inf; sh["triP2"][tri[4], 1] = -np.inf
    sh["planeNormal"][np.nonzero(sh["type"] == 2)[0][0]] = (0.0, 0.0, 0.0)
    sph = np.nonzero(sh["type"] == 0)[0]
    sh["sphereRadius"][sph[0]] = np.nan; sh["sphereCenter"][sph[1], 1] = np.inf
    cases.append((sh, 12))
    tr = rt.RayTracer(64, 48)
    for shapes, depth in cases:
        ref = rt.Scene.from_shapes(shapes, light, cam, bvh_depth=depth)
        for flags in PATHS:
            tr.upload_scene(light, cam, shapes)
            nodes, idx = tr.build_bvh(depth, flags)
            assert nodes.tobytes() == np.ascontiguousarray(ref.nodes).tobytes(), (len(shapes), depth, flags)
            assert idx.tobytes() == np.ascontiguousarray(ref.indices).tobytes(), (len(shapes), depth, flags)
    # no depth budget: the root is the only node
    nodes, idx = tr.build_bvh(0)
    assert len(nodes) == 1 and nodes["leftChild"][0] == -1 and nodes["numShapes"][0] == len(cases[-1][0])
    tr.close()


@pytest.mark.parametrize("seed", list(range(24)))
def test_device_build_bit_fuzz(rt, seed):
    """Arbitrary bit patterns in the geometry fields (NaN, inf, denormals, signed zeros): K8's tree is still the host
    builder's byte for byte — the ordered keys and their sequence numbers reproduce what `x < mn ? x : mn` keeps."""
    light = np.zeros(1, rt.LIGHT_DTYPE); cam = np.zeros(1, rt.CAMERA_DTYPE)
    sh, depth = bit_fuzzed_shapes(rt, seed)
    ref = rt.Scene.from_shapes(sh, light, cam, bvh_depth=depth)
    tr = rt.RayTracer(64, 48)
    for flags in PATHS:
        tr.upload_scene(light, cam, sh)
        nodes, idx = tr.build_bvh(depth, flags)
        assert nodes.tobytes() == np.ascontiguousarray(ref.nodes).tobytes(), (seed, len(sh), depth, flags)
        assert idx.tobytes() == np.ascontiguousarray(ref.indices).tobytes(), (seed, len(sh), depth, flags)
    tr.close()


DEVICE_LAYOUT = 4   # RT_BUILD_DEVICE_LAYOUT


@pytest.mark.parametrize("name,n,depth,on_device", [("toy", 0, 8, True), ("soup", 60000, 40, True), ("soup", 300000, 40, True),
                                                     ("monkey", 0, 15, True), ("car", 0, 25, False)])
def test_device_layout_k9(rt, oracle, name, n, depth, on_device):
    """K9 (rt_build_bvh with RT_BUILD_DEVICE_LAYOUT): list A, the primitive slots and the ranks laid out on the device
    from K8's tables, no host layout call, no copy of the tree to the host.  The frame is bit-identical to the one a
    context renders over the host builder's tree and the host layout builder's lists; the tree comes back on demand
    (rt_get_bvh) byte for byte; leaves of up to 16 shapes are walked as they are (monkey: 13), a tree with fatter leaves (car: 277)
    falls back to the host layout; later host-side
    updates (shape moves, a type change, a fresh upload) work on from the device-laid-out state."""
    W, H = 200, 150
    sc = rt.Scene(name, W, H, n_override=n, bvh_depth=depth)
    p = rt.make_params(W, H, useFresnel=True)
    a = rt.RayTracer(W, H, flags=rt.RT_FLAG_HIT_IDS | rt.RT_FLAG_NO_TOP_TREE); a.upload(sc)
    want = a.render(p)
    b = rt.RayTracer(W, H, flags=rt.RT_FLAG_HIT_IDS | rt.RT_FLAG_NO_TOP_TREE)
    b.upload_scene(sc.light, sc.camera, sc.shapes)
    nn, ni = b.build_bvh(depth, DEVICE_LAYOUT, fetch=False)
    info = b.layout_info()
    assert info["on_device"] == on_device, info
    assert nn == len(sc.nodes) and ni == len(sc.indices)
    if a.layout_info()["sub"] == 0 or not on_device:
        assert info["nodes"] == a.layout_info()["nodes"]
    else:       # leaves of 9 .. 16 shapes: the host layout expands them into sub-trees, the device layout walks them as they are
        assert info["sub"] == 0 and info["nodes"] == len(sc.nodes)
    assert _same_frame(b.render(p), want)
    o_rgba, o_ids, _, _ = oracle.render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, p, W, H, want_counters=False)
    from util import assert_parity, compare
    got = b.render(p)
    assert_parity(compare(got[0], got[1], o_rgba, o_ids), "%s over the device-laid-out tree" % name)
    nodes = np.zeros(nn, dtype=rt.NODE_DTYPE); idx = np.zeros(ni, dtype=np.int32)
    b._chk(b._L.rt_get_bvh(b._c, nodes.ctypes.data, nn, idx.ctypes.data, ni))
    assert nodes.tobytes() == np.ascontiguousarray(sc.nodes).tobytes() and idx.tobytes() == np.ascontiguousarray(sc.indices).tobytes()
    # a second device build of the same scene, then a moved shape, a type change, and a fresh upload
    b.build_bvh(depth, DEVICE_LAYOUT, fetch=False)
    assert _same_frame(b.render(p), want)
    shapes = sc.shapes
    k = int(np.flatnonzero(shapes["type"] == 0)[0])
    shapes["sphereCenter"][k] += np.float32(0.25)
    for t in (a, b):
        t.update_shapes(k, shapes[k:k + 1])
    assert _same_frame(b.render(p), a.render(p))
    shapes["type"][k] = 3                                          # topology change: the layout goes back to the host builder
    for t in (a, b):
        t.update_shapes(k, shapes[k:k + 1])
    assert _same_frame(b.render(p), a.render(p))
    assert not b.layout_info()["on_device"]
    b.upload(sc)
    assert _same_frame(b.render(p), want)
    a.close(); b.close()


def test_device_layout_then_device_animation(rt):
    """rt_animate (K6 + K7 need the tree's parent / leaf tables, built from the host copy) after a device-side layout."""
    W, H = 200, 150
    sc = rt.Scene("toy", W, H)
    p = rt.make_params(W, H)
    a = rt.RayTracer(W, H); a.upload(sc); a.register_animations(sc)
    b = rt.RayTracer(W, H); b.upload_scene(sc.light, sc.camera, sc.shapes)
    b.build_bvh(8, DEVICE_LAYOUT, fetch=False)
    b.register_animations(sc)
    for frame in range(1, 4):
        a.animate(frame * 0.3, 0.3); b.animate(frame * 0.3, 0.3)
        assert _same_frame(a.render(p), b.render(p)), "frame %d" % frame
    a.close(); b.close()
