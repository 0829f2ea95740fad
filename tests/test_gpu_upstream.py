"""GPU: the device builder K8 (rt_build_bvh) and the device animation K6 / K7 (rt_animate) against UPSTREAM CODE — the
reference's own split / buildBVH / serializeBVH / updateWheelAnimations / bounceSphere / updateScene / updateBVH cut out of
src/main.cpp and compiled in place (oracle/_ref/libref_bvh.so, see oracle/ref_bvh_glue.cpp) — through the C ABI.
Skipped when oracle/_ref did not travel to the box."""
import numpy as np
import pytest

from test_ref_builder import same_nodes, same_shapes
from util import random_shapes

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ref(oracle):
    if oracle.refbvh_lib() is None:
        pytest.skip("oracle/_ref/libref_bvh.so did not travel")
    return oracle


@pytest.mark.parametrize("name,n,depth", [("toy", 0, 8), ("monkey", 0, 15), ("car", 0, 25), ("soup", 200000, 40)])
def test_k8_tree_is_upstreams_tree(rt, ref, name, n, depth):
    sc = rt.Scene(name, 64, 48, n_override=n)
    want_nodes, want_idx = ref.RefScene(sc.shapes, node_dtype=rt.NODE_DTYPE).build(depth)
    tr = rt.RayTracer(64, 48)
    tr.upload_scene(sc.light, sc.camera, sc.shapes)
    for flags in (0, 2, 3):                                        # single block | per-phase launches | without warp aggregation
        nodes, idx = tr.build_bvh(depth, flags)
        assert same_nodes(nodes, want_nodes), (name, flags)
        assert np.array_equal(idx, want_idx), (name, flags)
    tr.close()


@pytest.mark.parametrize("seed", range(4))
def test_k8_tree_is_upstreams_tree_random(rt, ref, seed):
    rng = np.random.default_rng(500 + seed)
    base = rt.Scene("toy", 64, 48)
    n = int(rng.integers(2, 3000))
    shapes = random_shapes(rt, rng, n, types=(0, 2, 3), extent=float(rng.choice([1.0, 50.0])))
    k = rng.integers(0, n, n // 4)
    shapes[k] = shapes[rng.integers(0, n, n // 4)]                 # coincident centres
    depth = int(rng.choice([2, 15, 40]))
    want_nodes, want_idx = ref.RefScene(shapes, node_dtype=rt.NODE_DTYPE).build(depth)
    tr = rt.RayTracer(64, 48)
    tr.upload_scene(base.light, base.camera, shapes)
    for flags in (0, 2):
        nodes, idx = tr.build_bvh(depth, flags)
        assert same_nodes(nodes, want_nodes) and np.array_equal(idx, want_idx), (seed, flags)
    tr.close()


@pytest.mark.parametrize("name,depth,frames", [("monkey", 15, 40), ("car", 25, 60)])
@pytest.mark.parametrize("rebuild", [False, True])
def test_device_animation_is_upstreams_frame_step(rt, ref, name, depth, frames, rebuild):
    """rt_animate (K6 + K7), and rt_animate + rt_build_bvh (K8), frame after frame against the reference's own frame step
    (src/main.cpp:328-344): the shapes on the device (wheel vertices through GLM's translate * rotate * translate and
    mat4 * vec4 — the incremental rotation accumulates any ulp of difference) and the grown / rebuilt node arrays are
    bit-identical to upstream's."""
    sc = rt.Scene(name, 160, 120)
    R = ref.RefScene(sc.shapes, animated=sc.animated, node_dtype=rt.NODE_DTYPE)
    R.build(depth)
    anims = sc.animations()
    for k, (kind, idx, par) in enumerate(anims):
        if kind == 1:
            R.wheel(k, idx, par[:3])
    tr = rt.RayTracer(160, 120)
    tr.upload(sc)
    tr.register_animations(sc)
    for i in range(frames):
        t, dt = i * 0.017, 0.017 + 0.001 * (i % 3)
        for kind, idx, par in anims:
            if kind == 0:
                R.bounce(int(idx[0]), t, par[1], par[2])
        if name == "car":
            R.wheels_step(t, dt)
        want_nodes, want_idx = R.frame(rebuild, depth)
        want_shapes = R.shapes()
        tr.animate(t, dt)
        if rebuild:
            got_nodes, got_idx = tr.build_bvh(depth)
            assert np.array_equal(got_idx, want_idx), "%s rebuilt indices, frame %d" % (name, i)
            got_shapes, _ = tr.download_scene(len(sc.shapes), len(got_nodes))
        else:
            got_shapes, got_nodes = tr.download_scene(len(sc.shapes), len(sc.nodes))
        assert same_shapes(got_shapes, want_shapes), "%s shapes, frame %d" % (name, i)
        assert same_nodes(got_nodes, want_nodes), "%s %s nodes, frame %d" % (name, "rebuilt" if rebuild else "grown", i)
    tr.close()
