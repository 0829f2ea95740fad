"""GPU parity: the CUDA path through the C ABI vs the CPU oracle on identical scene bytes.

Bar (north_star): primary-hit shape ids bit-exact; colour within max-abs 1e-3 per channel on
>= 99.9 % of pixels.  Everything here calls librt_b200.so through ctypes (the C ABI).
"""
import numpy as np
import pytest

from util import assert_parity, bit_fuzzed_shapes, compare, random_shapes

pytestmark = pytest.mark.gpu


def _render_both(rt, oracle, sc, W, H, **kw):
    p = rt.make_params(W, H, **kw)
    o_rgba, o_ids, o_d, _ = oracle.render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, p, W, H, want_counters=False)
    tr = rt.RayTracer(W, H)
    tr.upload(sc)
    rgba, ids = tr.render(p)
    tr.close()
    return compare(rgba, ids, o_rgba, o_ids, o_d), rgba, ids, o_rgba, o_ids


@pytest.mark.parametrize("name,W,H", [("toy", 320, 240), ("triangle", 200, 150), ("monkey", 400, 300), ("car", 400, 300)])
@pytest.mark.parametrize("mt", [False, True])
@pytest.mark.parametrize("fresnel", [False, True])
def test_scene_parity(rt, oracle, name, W, H, mt, fresnel):
    sc = rt.Scene(name, W, H)
    stats, *_ = _render_both(rt, oracle, sc, W, H, useMollerTrumbore=mt, useFresnel=fresnel)
    assert_parity(stats, "%s mt=%s fresnel=%s" % (name, mt, fresnel))


@pytest.mark.parametrize("name", ["toy", "monkey"])
def test_bruteforce_twin(rt, oracle, name):
    """useBVH=false path (gpu_shader.comp:523-620)."""
    W, H = 160, 120
    sc = rt.Scene(name, W, H)
    stats, *_ = _render_both(rt, oracle, sc, W, H, useBVH=False)
    assert_parity(stats, name + " brute force")


@pytest.mark.parametrize("bounces", [0, 1, 2, 10])
def test_bounce_counts(rt, oracle, bounces):
    W, H = 200, 150
    sc = rt.Scene("monkey", W, H)
    stats, *_ = _render_both(rt, oracle, sc, W, H, maxBounces=bounces, useFresnel=True)
    assert_parity(stats, "bounces=%d" % bounces)


def test_full_size_car(rt, oracle):
    """BASELINE configs[1]: car scene at 800x600."""
    W, H = 800, 600
    sc = rt.Scene("car", W, H)
    stats, *_ = _render_both(rt, oracle, sc, W, H)
    assert_parity(stats, "car 800x600")


def test_ragged_resolution(rt, oracle):
    """Image size not a multiple of the 32x8 tile."""
    W, H = 123, 77
    sc = rt.Scene("toy", W, H)
    stats, *_ = _render_both(rt, oracle, sc, W, H)
    assert_parity(stats, "ragged")


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_random_scenes(rt, oracle, seed):
    rng = np.random.default_rng(seed)
    W, H = 160, 120
    base = rt.Scene("toy", W, H)
    shapes = random_shapes(rt, rng, 200, types=(0, 2, 3))
    sc = rt.Scene.from_shapes(shapes, base.light, base.camera, bvh_depth=12, width=W, height=H)
    for mt in (False, True):
        stats, *_ = _render_both(rt, oracle, sc, W, H, useMollerTrumbore=mt, useFresnel=True)
        assert_parity(stats, "random seed=%d mt=%s" % (seed, mt))


def _look_at(rt, pos, target, W, H, fov):
    """Camera bytes as Camera::LookAt leaves them (camera.hpp:99-110): Front, Right = Front x worldUp, Up = Right x Front, fp32."""
    cam = np.zeros(1, dtype=rt.CAMERA_DTYPE)
    f32 = np.float32
    pos, target = np.asarray(pos, f32), np.asarray(target, f32)
    nrm = lambda v: (v * f32(1.0 / np.sqrt(f32((v * v).sum(dtype=f32))))).astype(f32)   # noqa: E731
    front = nrm(target - pos)
    right = nrm(np.cross(front, np.array([0, 1, 0], f32)).astype(f32))
    up = nrm(np.cross(right, front).astype(f32))
    cam["Position"], cam["Front"], cam["Up"], cam["Right"] = pos, front, up, right
    cam["aspectRatio"], cam["fov"] = f32(W) / f32(H), f32(fov)
    return cam


@pytest.mark.parametrize("seed", list(range(100, 140)))
def test_fuzz_scenes_cameras_and_settings(rt, oracle, seed):
    """Random everything at once: 1 .. 500 shapes of the three boundable types with a tenth of them DUPLICATED under another
    material (equal distances: the first in the reference's visit order must win, gpu_shader.comp:410-411), a camera somewhere
    inside or outside the scene looking anywhere, a light anywhere, BVH depth 1 .. 20 (fat leaves of hundreds of shapes at the
    low end), 0 .. 5 bounces, both triangle tests, Fresnel on / off, a ragged frame size."""
    rng = np.random.default_rng(seed)
    W, H = int(rng.integers(33, 131)), int(rng.integers(17, 97))
    n = int(rng.choice([1, 2, 7, 40, 150, 500]))
    shapes = random_shapes(rt, rng, n, types=(0, 2, 3), extent=float(rng.choice([3.0, 10.0, 40.0])))
    ndup = max(1, n // 10)
    dup = shapes[rng.integers(0, n, ndup)].copy()
    dup["color"] = rng.uniform(0, 1, (ndup, 3))
    dup["specularStrength"] = rng.choice([0.0, 0.5, 1.0], ndup)
    both = np.zeros(n + ndup, dtype=rt.SHAPE_DTYPE)          # (np.concatenate would PACK the padded dtype: 124-byte records)
    both[:n], both[n:] = shapes, dup
    shapes = both[rng.permutation(n + ndup)]
    assert shapes.dtype.itemsize == 192
    base = rt.Scene("toy", W, H)
    light = base.light.copy()
    light["position"] = rng.uniform(-25, 25, 3)
    light["color"] = rng.uniform(5, 40) * np.ones(3)
    cam = _look_at(rt, rng.uniform(-20, 20, 3), rng.uniform(-5, 5, 3), W, H, float(rng.choice([30.0, 60.0, 100.0])))
    depth = int(rng.integers(1, 21))
    sc = rt.Scene.from_shapes(shapes, light, cam, bvh_depth=depth, width=W, height=H)
    kw = dict(maxBounces=int(rng.integers(0, 6)), useMollerTrumbore=bool(rng.integers(0, 2)), useFresnel=bool(rng.integers(0, 2)))
    stats, *_ = _render_both(rt, oracle, sc, W, H, **kw)
    assert_parity(stats, "fuzz seed=%d %dx%d n=%d %r" % (seed, W, H, len(shapes), kw))
    # the same scene through the device builder + device layout, and in the reference's own walk order
    p = rt.make_params(W, H, **kw)
    a = rt.RayTracer(W, H); a.upload(sc); want = a.render(p); a.close()
    b = rt.RayTracer(W, H); b.upload_scene(sc.light, sc.camera, sc.shapes)
    nodes, idx = b.build_bvh(depth, rt.RT_BUILD_DEVICE_LAYOUT)
    assert nodes.tobytes() == np.ascontiguousarray(sc.nodes).tobytes() and idx.tobytes() == np.ascontiguousarray(sc.indices).tobytes()
    got = b.render(p); b.close()
    assert np.array_equal(got[1], want[1]) and np.array_equal(got[0].view(np.uint32), want[0].view(np.uint32)), "device build + layout differs, seed %d" % seed
    c = rt.RayTracer(W, H, flags=rt.RT_FLAG_REFERENCE_WALK | rt.RT_FLAG_HIT_IDS); c.upload(sc); got = c.render(p); c.close()
    assert np.array_equal(got[1], want[1]) and np.array_equal(got[0].view(np.uint32), want[0].view(np.uint32)), "reference walk differs, seed %d" % seed


@pytest.mark.parametrize("seed", list(range(24)))
def test_bit_fuzzed_scenes(rt, oracle, seed):
    """Geometry fields with arbitrary bit patterns (NaN, inf, denormals, signed zeros; util.bit_fuzzed_shapes): the walk's
    fast paths must hand such rays and boxes to the exact GLSL-select code (slab / Ray::fast / Ray::sub), the device layout must
    stay conservative or fall back — ids and colours still equal the oracle's, with the BVH, both triangle tests, and brute force."""
    sh, depth = bit_fuzzed_shapes(rt, seed)
    if seed % 3:
        sh = sh[sh["type"] != 1]                       # infinite planes only in every third case
    rng = np.random.default_rng(seed)
    W, H = 96, 64
    base = rt.Scene("toy", W, H)
    light = base.light.copy()
    light["position"] = rng.uniform(-25, 25, 3)
    cam = _look_at(rt, rng.uniform(-20, 20, 3), rng.uniform(-5, 5, 3), W, H, 60.0)
    sc = rt.Scene.from_shapes(sh, light, cam, bvh_depth=depth, width=W, height=H)
    for kw in (dict(), dict(useMollerTrumbore=True, useFresnel=True), dict(useBVH=False)):
        stats, *_ = _render_both(rt, oracle, sc, W, H, **kw)
        assert_parity(stats, "bit fuzz seed=%d n=%d depth=%d %r" % (seed, len(sh), depth, kw))


def test_planes_bruteforce(rt, oracle):
    """Infinite planes (type 1) only work in the brute-force path upstream."""
    rng = np.random.default_rng(7)
    W, H = 160, 120
    base = rt.Scene("toy", W, H)
    shapes = random_shapes(rt, rng, 40, types=(0, 1, 2, 3))
    sc = rt.Scene.from_shapes(shapes, base.light, base.camera, bvh_depth=8, width=W, height=H)
    stats, *_ = _render_both(rt, oracle, sc, W, H, useBVH=False)
    assert_parity(stats, "planes brute force")
    stats, *_ = _render_both(rt, oracle, sc, W, H, useBVH=True)
    assert_parity(stats, "planes in bvh")


def test_empty_scene(rt, oracle):
    """No shapes: a root leaf with zero shapes -> background everywhere."""
    W, H = 64, 48
    base = rt.Scene("toy", W, H)
    sc = rt.Scene.from_shapes(np.zeros(0, dtype=rt.SHAPE_DTYPE), base.light, base.camera, width=W, height=H)
    stats, rgba, ids, *_ = _render_both(rt, oracle, sc, W, H)
    assert_parity(stats, "empty")
    assert (ids == -1).all()


def test_partial_updates_match_full_upload(rt, oracle):
    """Animated path: rt_update_shapes + rt_update_nodes (refit) == a fresh full upload == oracle."""
    W, H = 320, 240
    sc = rt.Scene("monkey", W, H)
    p = rt.make_params(W, H)
    tr = rt.RayTracer(W, H)
    tr.upload(sc)
    tr.render(p)
    for frame in range(1, 4):
        sc.animate(frame / 10.0, 0.1, mode=0)
        shapes = sc.shapes
        for i in sc.animated:                       # one 192-byte update per animated shape, as upstream
            tr.update_shapes(int(i), shapes[int(i):int(i) + 1])
        tr.update_nodes(sc.nodes)
        rgba, ids = tr.render(p)
        o_rgba, o_ids, _, _ = oracle.render(sc.light, sc.camera, shapes, sc.nodes, sc.indices, p, W, H, want_counters=False)
        assert_parity(compare(rgba, ids, o_rgba, o_ids), "animated frame %d" % frame)
    sc.animate(0.5, 0.1, mode=1)                    # full rebuild: topology may change
    tr.upload_scene(shapes=sc.shapes, nodes=sc.nodes, indices=sc.indices)
    rgba, ids = tr.render(p)
    o_rgba, o_ids, _, _ = oracle.render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, p, W, H, want_counters=False)
    assert_parity(compare(rgba, ids, o_rgba, o_ids), "rebuild")
    tr.close()


def test_sharded_tiles_reassemble(rt):
    """Multi-GPU partition on one device: every rank's compact shard, untiled, equals the 1-GPU image bit for bit."""
    import torch
    W, H = 333, 201
    sc = rt.Scene("monkey", W, H)
    p = rt.make_params(W, H)
    tr = rt.RayTracer(W, H)
    tr.upload(sc)
    full, _ = tr.render(p)
    for world in (2, 3, 8):
        tr.set_shard(0, world)
        px = tr.shard_pixels()
        shards = torch.zeros((world, px, 4), dtype=torch.float32, device="cuda")
        for r in range(world):
            tr.set_shard(r, world)
            tr.render_device(p, shards[r].data_ptr())
        out = torch.empty((H, W, 4), dtype=torch.float32, device="cuda")
        tr.untile(shards.data_ptr(), world, out.data_ptr())
        tr.synchronize()
        assert np.array_equal(out.cpu().numpy().view(np.uint32), full.view(np.uint32)), "world=%d" % world
    tr.close()


def test_sharded_scheduled_frames(rt):
    """The compact tile-major shard path with the heavy-first order, split blocks and paired walks in effect (one
    context per rank, several frames each so that the previous frame's costs exist): every frame untiles to the
    1-GPU image, ids included."""
    import torch
    W, H = 640, 360
    sc = rt.Scene("car", W, H)
    p = rt.make_params(W, H)
    one = rt.RayTracer(W, H)
    one.upload(sc)
    full, full_ids = one.render(p)
    world = 3
    ranks = []
    for r in range(world):
        t = rt.RayTracer(W, H)
        t.upload(sc); t.set_shard(r, world)
        ranks.append(t)
    px = ranks[0].shard_pixels()
    shards = torch.zeros((world, px, 4), dtype=torch.float32, device="cuda")
    sid = torch.zeros((world, px), dtype=torch.int32, device="cuda")
    out = torch.empty((H, W, 4), dtype=torch.float32, device="cuda")
    for frame in range(4):
        shards.fill_(-2.0)
        for r, t in enumerate(ranks):
            t.render_device(p, shards[r].data_ptr(), sid[r].data_ptr())
        for t in ranks:
            t.synchronize()
        one.untile(shards.data_ptr(), world, out.data_ptr())
        one.synchronize()
        assert np.array_equal(out.cpu().numpy().view(np.uint32), full.view(np.uint32)), "frame %d" % frame
    # ids of rank 0's shard against the full id map (tile k of the shard = tile k*world of the frame)
    ids0 = sid[0].cpu().numpy()
    tx = (W + 31) // 32
    for k in range(0, px // 256, 7):
        tile = k * world
        x0, y0 = rt.shard.tile_origin(W, H, tile, world)
        if y0 + 8 > H or x0 + 32 > W:
            continue
        assert np.array_equal(ids0[k * 256:(k + 1) * 256].reshape(8, 32), full_ids[y0:y0 + 8, x0:x0 + 32]), "tile %d" % tile
    for t in ranks:
        t.close()
    one.close()


@pytest.mark.parametrize("n", [1, 255, 1023, 1025, 4097, 262147])
def test_present_edge_values_and_ragged_sizes(rt, oracle, n):
    """K5 on synthetic pixels: below 0, above 1, NaN, +-inf, exact quantisation boundaries; sizes around the four-pixels-per-
    thread passes of the kernel (a pass = grid x 256 x 4 pixels)."""
    import torch
    rng = np.random.default_rng(n)
    rgba = rng.uniform(-0.5, 1.5, (n, 4)).astype(np.float32)
    special = np.array([np.nan, np.inf, -np.inf, 0.0, -0.0, 1.0, 0.5, 127.5 / 255.0, 0.5 / 255.0, 254.5 / 255.0], dtype=np.float32)
    rgba.reshape(-1)[:: 7] = special[np.arange(len(rgba.reshape(-1)[:: 7])) % len(special)]
    tr = rt.RayTracer(64, 48)
    src = torch.from_numpy(rgba).cuda()
    img8 = torch.full((n + 8, 4), 77, dtype=torch.uint8, device="cuda")
    tr.present(src.data_ptr(), img8.data_ptr(), n)
    tr.synchronize()
    got = img8.cpu().numpy()
    assert np.array_equal(got[:n], oracle.present(rgba.reshape(1, n, 4)).reshape(n, 4))
    assert (got[n:] == 77).all(), "wrote past the end"
    tr.close()


def test_present_matches_oracle(rt, oracle):
    import torch
    W, H = 200, 150
    sc = rt.Scene("monkey", W, H)
    p = rt.make_params(W, H)
    tr = rt.RayTracer(W, H)
    tr.upload(sc)
    rgba, _ = tr.render(p)
    img8 = torch.empty((H, W, 4), dtype=torch.uint8, device="cuda")
    tr.present(tr.device_image(), img8.data_ptr(), W * H)
    tr.synchronize()
    assert np.array_equal(img8.cpu().numpy(), oracle.present(rgba))
    tr.close()


def test_errors(rt):
    tr = rt.RayTracer(64, 48)
    p = rt.make_params(64, 48)
    with pytest.raises(rt.RtError) as e:
        tr.render(p)
    assert e.value.code == -4                       # RT_ERR_STATE: render before upload
    sc = rt.Scene("toy", 64, 48)
    nodes = sc.nodes.copy()
    nodes["leftChild"][-1] = 10 ** 6
    with pytest.raises(rt.RtError) as e:
        tr.upload_scene(sc.light, sc.camera, sc.shapes, nodes, sc.indices)
    assert e.value.code == -3                       # RT_ERR_SCENE
    idx = sc.indices.copy()
    idx[0] = 999
    with pytest.raises(rt.RtError):
        tr.upload_scene(sc.light, sc.camera, sc.shapes, sc.nodes, idx)
    tr.upload(sc)
    tr.render(p)
    tr.close()


def test_reference_walk_flag_matches_default(rt, oracle):
    """RT_FLAG_REFERENCE_WALK (no sub-trees: the reference's exact node/primitive sequence) and the default
    (fat leaves expanded into conservative sub-trees) give the same ids and bit-identical colours."""
    W, H = 400, 300
    sc = rt.Scene("car", W, H)
    for kw in (dict(), dict(useMollerTrumbore=True, useFresnel=True)):
        p = rt.make_params(W, H, **kw)
        outs = []
        for flags in (rt.RT_FLAG_HIT_IDS, rt.RT_FLAG_HIT_IDS | rt.RT_FLAG_REFERENCE_WALK):
            tr = rt.RayTracer(W, H, flags=flags)
            tr.upload(sc)
            rgba, ids = tr.render(p)
            stats = tr.render_stats(p)
            tr.close()
            outs.append((rgba, ids, stats))
        assert np.array_equal(outs[0][1], outs[1][1])
        assert np.array_equal(outs[0][0].view(np.uint32), outs[1][0].view(np.uint32))
        assert outs[0][2]["rays"] == outs[1][2]["rays"]
        assert outs[0][2]["prim_tests"] < outs[1][2]["prim_tests"] / 3


def test_animated_car_refit(rt, oracle):
    """Rotating wheels (stale planes upstream, src/main.cpp:1101-1106) + grow-only refit: the sub-tree boxes are
    refitted on rt_update_shapes and every frame still matches the oracle."""
    W, H = 400, 300
    sc = rt.Scene("car", W, H)
    p = rt.make_params(W, H)
    tr = rt.RayTracer(W, H)
    tr.upload(sc)
    tr.render(p)
    for frame in range(1, 4):
        sc.animate(frame * 0.21, 0.21, mode=0)
        shapes = sc.shapes
        an = sc.animated
        lo, hi = int(an.min()), int(an.max()) + 1
        tr.update_shapes(lo, shapes[lo:hi])
        tr.update_nodes(sc.nodes)
        rgba, ids = tr.render(p)
        o_rgba, o_ids, _, _ = oracle.render(sc.light, sc.camera, shapes, sc.nodes, sc.indices, p, W, H, want_counters=False)
        assert_parity(compare(rgba, ids, o_rgba, o_ids), "animated car frame %d" % frame)
    tr.close()


def test_device_refit_keeps_the_boxes_tight(rt):
    """K4 (refit_kernel): after rt_update_shapes + rt_update_nodes the boxes refitted ON THE DEVICE give the same image
    as a context that got the moved scene as a fresh upload (boxes built on the host), with the same amount of
    work: no box degenerated to the all-covering fallback; a shape that stops being boundable (a triangle squashed
    to a sliver) still renders exactly and the layout is rebuilt at the next upload."""
    W, H = 400, 300
    sc = rt.Scene("car", W, H)
    p = rt.make_params(W, H)
    a = rt.RayTracer(W, H)
    a.upload(sc)
    a.render(p)
    for frame in range(1, 3):
        sc.animate(frame * 0.33, 0.33, mode=0)
        an = sc.animated
        lo, hi = int(an.min()), int(an.max()) + 1
        a.update_shapes(lo, sc.shapes[lo:hi]); a.update_nodes(sc.nodes)
        ra, ia = a.render(p)
        b = rt.RayTracer(W, H)
        b.upload(sc)
        rb, ib = b.render(p)
        assert np.array_equal(ia, ib) and np.array_equal(ra.view(np.uint32), rb.view(np.uint32))
        sa, sb = a.render_stats(p), b.render_stats(p)
        assert sa["rays"] == sb["rays"]
        assert sa["prim_tests"] <= 1.25 * sb["prim_tests"] and sa["node_visits"] <= 1.25 * sb["node_visits"], (sa, sb)
        b.close()
    # squash every 7th triangle into a sliver: prim_bound gives up on those (some of them live in fat leaves)
    shapes = sc.shapes.copy()
    tri = np.flatnonzero(shapes["type"] == 3)[::7]
    shapes["triP3"][tri] = shapes["triP1"][tri] + (shapes["triP2"][tri] - shapes["triP1"][tri]) * np.float32(0.5)
    k = 0
    a.update_shapes(0, shapes)
    ra, ia = a.render(p)
    b = rt.RayTracer(W, H)
    b.upload_scene(sc.light, sc.camera, shapes, sc.nodes, sc.indices)
    rb, ib = b.render(p)
    assert np.array_equal(ia, ib) and np.array_equal(ra.view(np.uint32), rb.view(np.uint32))
    a.update_shapes(k, shapes[k:k + 1])            # next upload: the degraded flag has arrived, the layout is rebuilt
    ra2, ia2 = a.render(p)
    assert np.array_equal(ia2, ib) and np.array_equal(ra2.view(np.uint32), rb.view(np.uint32))
    sa, sb = a.render_stats(p), b.render_stats(p)
    assert sa["prim_tests"] == sb["prim_tests"], (sa, sb)
    a.close(); b.close()


@pytest.mark.parametrize("general", [False, True])
@pytest.mark.parametrize("name", ["monkey", "car"])
def test_device_animation_equals_host_path(rt, name, general):
    """rt_animate (K6 + K7 on the device copies: one block that moves the shapes and grows the animated leaves and their
    ancestors level by level, or — general, RT_FLAG_NO_ONE_BLOCK_UPDATE — anim_kernel + grow_kernel) against the
    reference's frame step done on the host (animate, per-shape update, grow-only updateBVH, node re-upload): the
    shapes, the node boxes and the rendered frames are bit-identical, frame after frame."""
    W, H = 320, 240
    sc = rt.Scene(name, W, H)
    p = rt.make_params(W, H)
    fl = rt.RT_FLAG_HIT_IDS | (rt.RT_FLAG_NO_ONE_BLOCK_UPDATE if general else 0)
    a = rt.RayTracer(W, H); b = rt.RayTracer(W, H, flags=fl)
    a.upload(sc); b.upload(sc)
    b.register_animations(sc)
    an = sc.animated
    lo, hi = int(an.min()), int(an.max()) + 1
    for frame in range(1, 5):
        t, dt = frame * 0.21, 0.21
        sc.animate(t, dt, mode=0)
        a.update_shapes(lo, sc.shapes[lo:hi]); a.update_nodes(sc.nodes)
        b.animate(t, dt)
        ra, ia = a.render(p); rb, ib = b.render(p)
        assert np.array_equal(ia, ib) and np.array_equal(ra.view(np.uint32), rb.view(np.uint32)), "frame %d" % frame
        shapes_d, nodes_d = b.download_scene(len(sc.shapes), len(sc.nodes))
        assert shapes_d.tobytes() == np.ascontiguousarray(sc.shapes).tobytes(), "shapes, frame %d" % frame
        assert nodes_d.tobytes() == np.ascontiguousarray(sc.nodes).tobytes(), "nodes, frame %d" % frame
    # a host-side change after device animation: the layout is rebuilt from the downloaded state
    shapes = shapes_d.copy()
    k = int(np.flatnonzero(shapes["type"] == 0)[-1])
    shapes["type"][k] = 3                                        # a sphere becomes a (degenerate) triangle: topology change
    b.update_shapes(k, shapes[k:k + 1])
    rb, ib = b.render(p)
    c = rt.RayTracer(W, H)
    c.upload_scene(sc.light, sc.camera, shapes, sc.nodes, sc.indices)
    rc, ic = c.render(p)
    assert np.array_equal(ib, ic) and np.array_equal(rb.view(np.uint32), rc.view(np.uint32))
    a.close(); b.close(); c.close()


def test_camera_far_away(rt, oracle):
    """Camera beyond the region the sub-tree margins were computed for (omax): margins are recomputed / ignored."""
    W, H = 200, 150
    sc = rt.Scene("car", W, H)
    p = rt.make_params(W, H)
    tr = rt.RayTracer(W, H)
    tr.upload(sc)
    tr.render(p)
    for dist in (500.0, 20000.0):
        sc.set_camera((0.0, -10.0, dist), (0.0, 0.0, 0.0), fov=2.0 if dist > 1000 else 10.0)
        tr.update_camera(sc.camera)
        rgba, ids = tr.render(p)
        o_rgba, o_ids, _, _ = oracle.render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, p, W, H, want_counters=False)
        assert_parity(compare(rgba, ids, o_rgba, o_ids), "camera at %g" % dist)
        assert (o_ids >= 0).any()
    tr.close()


@pytest.mark.parametrize("n,W,H", [(20000, 320, 240), (1000000, 256, 144)])
def test_soup_parity(rt, oracle, n, W, H):
    """BASELINE configs[3]: triangle soup + spheres, deep BVH (traversal-bound); the full 1M-triangle scene at a
    resolution the oracle finishes in seconds."""
    sc = rt.Scene("soup", W, H, n_override=n)
    stats, *_ = _render_both(rt, oracle, sc, W, H)
    assert_parity(stats, "soup %d" % n)


def test_render_async_and_banded_copy(rt, oracle):
    """rt_render_async (banded launches, D2H overlapped) delivers the same bytes as rt_render / rt_render_device."""
    import torch
    W, H = 640, 520                                  # 65 tile rows -> 4 bands, ragged last tile row
    sc = rt.Scene("monkey", W, H)
    p = rt.make_params(W, H, useFresnel=True)
    tr = rt.RayTracer(W, H)
    tr.upload(sc)
    ref, ref_ids = tr.render(p)
    pin = torch.empty((H * W, 4), dtype=torch.float32, pin_memory=True).numpy()
    pin_ids = torch.empty(H * W, dtype=torch.int32, pin_memory=True).numpy()
    pin[:] = -7.0
    tr.render_async(p, pin, pin_ids)
    tr.synchronize()
    assert np.array_equal(pin.reshape(H, W, 4).view(np.uint32), ref.view(np.uint32))
    assert np.array_equal(pin_ids.reshape(H, W), ref_ids)
    dev = torch.empty((H, W, 4), dtype=torch.float32, device="cuda")
    tr.render_device(p, dev.data_ptr())
    tr.synchronize()
    assert np.array_equal(dev.cpu().numpy().view(np.uint32), ref.view(np.uint32))
    o_rgba, o_ids, _, _ = oracle.render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, p, W, H, want_counters=False)
    assert_parity(compare(ref, ref_ids, o_rgba, o_ids), "banded")
    tr.close()


@pytest.mark.parametrize("W,H", [(400, 300), (2048, 1100)])
def test_pipelined_async_frames(rt, W, H):
    """rt_render_async pipelines frames over two device images (small frames: one launch with the heavy-first order;
    large ones banded): with two host buffers and rt_wait_frames(1) every frame of an animated sequence arrives
    complete and equal to a synchronous render of the same scene state."""
    import torch
    sc = rt.Scene("monkey", W, H)
    p = rt.make_params(W, H)
    a = rt.RayTracer(W, H, flags=0); b = rt.RayTracer(W, H, flags=0)
    a.upload(sc); b.upload(sc)
    an = sc.animated
    lo, hi = int(an.min()), int(an.max()) + 1
    bufs = [torch.empty((H * W, 4), dtype=torch.float32, pin_memory=True).numpy() for _ in range(2)]
    expected = []
    n = 7
    for i in range(n):
        sc.animate(i * 0.17, 0.17, mode=0)
        for t in (a, b):
            t.update_shapes(lo, sc.shapes[lo:hi]); t.update_nodes(sc.nodes)
        expected.append(b.render(p, want_ids=False)[0].copy())
        a.wait_frames(1)
        if i >= 2:
            assert np.array_equal(bufs[i & 1].view(np.uint32).reshape(H, W, 4), expected[i - 2].view(np.uint32)), "frame %d" % (i - 2)
        bufs[i & 1][:] = -3.0
        a.render_async(p, bufs[i & 1])
    a.wait_frames(0)
    for i in (n - 2, n - 1):
        assert np.array_equal(bufs[i & 1].view(np.uint32).reshape(H, W, 4), expected[i].view(np.uint32)), "frame %d" % i
    # a synchronous render after pipelined ones still works on the same context
    r, _ = a.render(p, want_ids=False)
    assert np.array_equal(r.view(np.uint32), expected[-1].view(np.uint32))
    a.close(); b.close()


@pytest.mark.parametrize("name,W,H", [("monkey", 333, 217), ("car", 1920, 1200)])
def test_render_rgba8_is_the_presented_frame(rt, oracle, name, W, H):
    """rt_render_rgba8 (K1 with K5 behind it band by band, RGBA8 to the host) delivers exactly the display quad's arithmetic
    (oracle.present: shader.frag:8-12 + framebuffer conversion) on rt_render's own rgba32f frame — one launch (small frame),
    banded (2.3 M pixels), pipelined over two images, and as a compact shard."""
    import torch
    sc = rt.Scene(name, W, H)
    p = rt.make_params(W, H)
    tr = rt.RayTracer(W, H, flags=rt.RT_FLAG_HIT_IDS); tr.upload(sc)
    rgba, ids = tr.render(p, want_ids=True)
    want = oracle.present(rgba.reshape(-1, 4)).reshape(-1, 4)
    ids8 = np.full(W * H, -5, np.int32)
    got = tr.render_rgba8(p, out_ids=ids8)
    assert got.dtype == np.uint8 and got.shape == (W * H, 4)
    assert np.array_equal(got, want) and np.array_equal(ids8, ids.ravel())
    img8 = torch.empty((H * W, 4), dtype=torch.uint8, device="cuda")        # the float frame stays on the device
    tr.present(tr.device_image(), img8.data_ptr(), W * H); tr.synchronize()
    assert np.array_equal(img8.cpu().numpy(), want)
    bufs = [torch.empty((H * W, 4), dtype=torch.uint8, pin_memory=True).numpy() for _ in range(2)]
    for i in range(5):
        tr.wait_frames(1)
        bufs[i & 1][:] = 7
        tr.render_rgba8_async(p, bufs[i & 1])
    tr.wait_frames(0)
    assert np.array_equal(bufs[0], want) and np.array_equal(bufs[1], want)
    # a shard: the compact tile-major pixels of rank 1 of 3, presented
    tr.set_shard(1, 3)
    sh, _ = tr.render(p, want_ids=False)
    got = tr.render_rgba8(p)
    assert got.shape[0] == tr.shard_pixels() and np.array_equal(got, oracle.present(sh.reshape(-1, 4)).reshape(-1, 4))
    tr.close()


@pytest.mark.parametrize("flags", [0, "no_deferred"])
def test_deferred_copy_flush_points(rt, flags):
    """The host copy of a pipelined single-launch frame is issued with the next frame's render kernel or by whoever waits
    first (rt_api.h: rt_render_async): every way of ending the pipeline must deliver the frame, ids included."""
    import torch
    W, H = 320, 200
    sc = rt.Scene("monkey", W, H)
    p = rt.make_params(W, H)
    fl = rt.RT_FLAG_HIT_IDS | (rt.RT_FLAG_NO_DEFERRED_COPY if flags else 0)
    a = rt.RayTracer(W, H, flags=fl); a.upload(sc)
    want, want_ids = a.render(p, want_ids=True)
    want = want.copy().view(np.uint32).reshape(-1, 4); want_ids = want_ids.copy().ravel()
    buf = [torch.empty((H * W, 4), dtype=torch.float32, pin_memory=True).numpy() for _ in range(2)]
    ids = [torch.empty(H * W, dtype=torch.int32, pin_memory=True).numpy() for _ in range(2)]
    dev = torch.empty((H * W, 4), dtype=torch.float32, device="cuda")

    def clear():
        for k in range(2):
            buf[k][:] = -7.0; ids[k][:] = -9

    def check(k):
        assert np.array_equal(buf[k].view(np.uint32), want), "colours of buffer %d" % k
        assert np.array_equal(ids[k], want_ids), "ids of buffer %d" % k

    clear(); a.render_async(p, buf[0], ids[0]); a.wait_frames(0); check(0)                       # one frame, waited for
    clear(); a.render_async(p, buf[0], ids[0]); a.synchronize(); check(0)                        # ended by rt_synchronize
    clear(); a.render_async(p, buf[0], ids[0]); a.timings(); check(0)                            # ended by rt_last_timings
    clear(); a.render_async(p, buf[0], ids[0]); a.render_device(p, dev.data_ptr()); a.synchronize(); check(0)   # a device render follows
    assert np.array_equal(dev.cpu().numpy().view(np.uint32), want)
    clear(); a.render_async(p, buf[0], ids[0]); r, ri = a.render(p, want_ids=True); check(0)     # a synchronous render follows
    assert np.array_equal(r.view(np.uint32).reshape(-1, 4), want) and np.array_equal(ri.ravel(), want_ids)
    clear(); a.render_async(p, buf[0], ids[0]); a.render_async(p, buf[1], ids[1]); a.wait_frames(1); check(0)   # two in flight
    a.wait_frames(0); check(1)
    clear()
    for i in range(5):                                                                           # no wait at all between frames
        a.render_async(p, buf[i & 1], ids[i & 1])
    a.wait_frames(0); check(0); check(1)
    clear(); a.render_async(p, buf[0], ids[0]); a.set_shard(0, 2); a.set_shard(0, 1); a.synchronize(); check(0)  # the shard changes
    a.close()


@pytest.mark.parametrize("mt,fresnel,bounces", [(False, False, 3), (True, True, 3), (False, True, 6)])
def test_scheduled_frames_identical(rt, mt, fresnel, bounces):
    """From the second frame on a context orders its blocks by last frame's costs, splits the heaviest into lane groups
    and lets the idle partner lanes of a group walk the reflected ray beside the shadow ray (paired mode): every frame
    must stay bit-identical to the first one and to a context with RT_FLAG_NO_HEAVY_FIRST."""
    import torch
    W, H = 800, 600
    sc = rt.Scene("car", W, H)
    p = rt.make_params(W, H, maxBounces=bounces, useFresnel=fresnel, useMollerTrumbore=mt)
    a = rt.RayTracer(W, H, flags=rt.RT_FLAG_HIT_IDS)
    b = rt.RayTracer(W, H, flags=rt.RT_FLAG_HIT_IDS | rt.RT_FLAG_NO_HEAVY_FIRST)
    a.upload(sc); b.upload(sc)
    dev = torch.empty((H, W, 4), dtype=torch.float32, device="cuda")
    ids = torch.empty((H, W), dtype=torch.int32, device="cuda")
    b.render_device(p, dev.data_ptr(), ids.data_ptr()); b.synchronize()
    want, want_ids = dev.cpu().numpy().copy(), ids.cpu().numpy().copy()
    for frame in range(5):
        dev.fill_(-1.0); ids.fill_(-5)
        a.render_device(p, dev.data_ptr(), ids.data_ptr()); a.synchronize()
        assert np.array_equal(ids.cpu().numpy(), want_ids), "ids, frame %d" % frame
        assert np.array_equal(dev.cpu().numpy().view(np.uint32), want.view(np.uint32)), "frame %d" % frame
    a.close(); b.close()


def test_fused_gather_frame_target(rt):
    """Fused gather on one device: the contexts of `world` ranks store their tiles straight into one row-major
    frame (rt_set_frame_target; across processes the same pointer is a CUDA-IPC peer mapping) == the 1-GPU image."""
    import torch
    W, H = 333, 201
    sc = rt.Scene("monkey", W, H)
    p = rt.make_params(W, H)
    tr = rt.RayTracer(W, H)
    tr.upload(sc)
    full, _ = tr.render(p)
    tr.close()
    for world in (2, 5):
        frame = torch.full((H, W, 4), -1.0, dtype=torch.float32, device="cuda")
        ranks = []
        for r in range(world):
            t = rt.RayTracer(W, H, flags=0)
            t.upload(sc)
            t.set_shard(r, world)
            t.set_frame_target(frame.data_ptr())
            t.render_device(p)
            ranks.append(t)
        for t in ranks:
            t.synchronize()
            t.close()
        assert np.array_equal(frame.cpu().numpy().view(np.uint32), full.view(np.uint32)), "world=%d" % world


def test_host_frame_target(rt):
    """rt_host_frame_register: the kernels store straight into a pinned, mapped host frame (here an mmap'd
    /dev/shm file, as several ranks would share it) == rt_render's frame, for world 1 and for 3 ranks."""
    import os
    W, H = 333, 201
    sc = rt.Scene("monkey", W, H)
    p = rt.make_params(W, H)
    tr = rt.RayTracer(W, H, flags=0)
    tr.upload(sc)
    full, _ = tr.render(p, want_ids=False)
    path = "/dev/shm/rt_b200_test_%d" % os.getpid()
    host = np.memmap(path, dtype=np.float32, mode="w+", shape=(H * W * 4,))
    os.unlink(path)
    try:
        for world in (1, 3):
            host[:] = -1.0
            ranks = []
            for r in range(world):
                t = rt.RayTracer(W, H, flags=0)
                t.upload(sc)
                t.set_shard(r, world)
                alias = t.host_frame_register(host) if r == 0 else alias
                t.set_frame_target(alias)
                t.render_device(p)
                ranks.append(t)
            for t in ranks:
                t.synchronize()
            assert np.array_equal(np.asarray(host).view(np.uint32), full.view(np.uint32).reshape(-1)), "world=%d" % world
            ranks[0].host_frame_unregister(host)
            for t in ranks:
                t.close()
    finally:
        tr.close()


def test_top_tree_flag_matches(rt, oracle):
    """List B (own SAH tree over the reference leaves) against RT_FLAG_NO_TOP_TREE: same ids, bit-identical colours,
    fewer box tests; also after a grow-only refit of the reference boxes (rt_update_nodes)."""
    W, H = 400, 300
    sc = rt.Scene("car", W, H)
    p = rt.make_params(W, H, useFresnel=True)
    a = rt.RayTracer(W, H, flags=rt.RT_FLAG_HIT_IDS)
    b = rt.RayTracer(W, H, flags=rt.RT_FLAG_HIT_IDS | rt.RT_FLAG_NO_TOP_TREE)
    a.upload(sc); b.upload(sc)
    for frame in range(3):
        if frame:
            sc.animate(frame * 0.3, 0.3, mode=0)
            an = sc.animated
            lo, hi = int(an.min()), int(an.max()) + 1
            for t in (a, b):
                t.update_shapes(lo, sc.shapes[lo:hi]); t.update_nodes(sc.nodes)
        ra, ia = a.render(p); rb, ib = b.render(p)
        assert np.array_equal(ia, ib) and np.array_equal(ra.view(np.uint32), rb.view(np.uint32)), "frame %d" % frame
        sa, sb = a.render_stats(p), b.render_stats(p)
        assert sa["rays"] == sb["rays"] and sa["node_visits"] < sb["node_visits"]
    o_rgba, o_ids, _, _ = oracle.render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, p, W, H, want_counters=False)
    assert_parity(compare(ra, ia, o_rgba, o_ids), "top tree, animated")
    a.close(); b.close()


def test_disjoint_host_updates_after_device_animation(rt, oracle):
    """rt_animate moves shapes on the device (the host shadow goes stale); two DISJOINT rt_update_shapes calls afterwards
    must not drag the pre-animation values of the shapes between them back onto the device (one copy per interval),
    also when the update changes the topology (the layout is rebuilt from the read-back device state + the host's writes)."""
    W, H = 320, 240
    sc = rt.Scene("car", W, H)
    p = rt.make_params(W, H)
    tr = rt.RayTracer(W, H); tr.upload(sc); tr.register_animations(sc)
    for frame in range(1, 4):
        sc.animate(frame * 0.25, 0.25, mode=0)
        tr.animate(frame * 0.25, 0.25)
    shapes = sc.shapes                                              # a fresh byte copy (ndarray.copy() would not keep the std430 padding)
                                                                    # == the device state (test_device_animation_equals_host_path)
    an = sc.animated
    a, b = int(an.min()) - 5, int(an.max()) + 5                     # two static shapes either side of the animated wheels
    for k, topo in ((0, False), (1, True)):
        for i in (a, b):
            shapes["color"][i] = (0.9, 0.1 * k, 0.2)
            if topo:
                shapes["type"][i] = 0; shapes["sphereCenter"][i] = shapes["triP1"][i]; shapes["sphereRadius"][i] = 0.7
            tr.update_shapes(i, shapes[i:i + 1])
        rgba, ids = tr.render(p)
        got, _ = tr.download_scene(len(shapes), len(sc.nodes))
        assert got.tobytes() == shapes.tobytes(), "device shapes after disjoint updates (topology change: %s)" % topo
        o_rgba, o_ids, _, _ = oracle.render(sc.light, sc.camera, shapes, sc.nodes, sc.indices, p, W, H, want_counters=False)
        assert_parity(compare(rgba, ids, o_rgba, o_ids), "disjoint updates after rt_animate, topo=%s" % topo)
    tr.close()


def test_zero_bounces_after_a_warmed_up_heavy_list(rt, oracle):
    """maxBounces = 0: the shader's bounce loop never runs and the pixel is (0,0,0,1) (gpu_shader.comp:450); also when the
    previous frames left a heavy list whose blocks are split into paired row groups."""
    W, H = 800, 600
    sc = rt.Scene("car", W, H)
    tr = rt.RayTracer(W, H); tr.upload(sc)
    p3 = rt.make_params(W, H, maxBounces=3)
    for _ in range(3):
        tr.render(p3)
    p0 = rt.make_params(W, H, maxBounces=0)
    rgba, ids = tr.render(p0)
    assert np.array_equal(rgba[..., :3], np.zeros_like(rgba[..., :3])) and (rgba[..., 3] == 1).all()
    tr.close()


@pytest.mark.parametrize("world", [2, 4])
def test_signalled_fused_gather_animated(rt, world):
    """Compute + signal, no collective (rt_api.h "signalled fused gather"), emulated on one device with one context per
    rank: the other ranks run one frame ahead, rank 0 renders, acquires, downloads and releases frame after frame.  The camera moves every frame, so a frame torn by a rank that ran ahead (overwriting a
    buffer still being read), or one acquired before a rank's tiles landed, differs from the single-GPU render."""
    import torch
    W, H, frames = 333, 201, 7
    sc = rt.Scene("monkey", W, H)
    p = rt.make_params(W, H)
    cams = []
    for f in range(frames):
        sc.set_camera((30.0 - 3.0 * f, -5.0, 40.0), (0.0, 10.0, -8.0))
        cams.append(sc.camera)
    solo = rt.RayTracer(W, H, flags=0); solo.upload(sc)
    want = []
    for f in range(frames):
        solo.update_camera(cams[f])
        want.append(solo.render(p, want_ids=False)[0].copy())
    solo.close()
    ranks = [rt.RayTracer(W, H, flags=0) for _ in range(world)]
    for r, t in enumerate(ranks):
        t.upload(sc); t.set_shard(r, world)
    _, base = ranks[0].frame_export()
    for r in range(1, world):
        ranks[r].set_frame_target(base)
        ranks[0].sync_peer(r, dev_word=ranks[r].sync_local())
    for t in ranks:
        t.sync_enable(world)
    def others(f):                                # the other ranks enqueue frame f
        for r in range(1, world):
            ranks[r].update_camera(cams[f]); ranks[r].render_device(p)
    host = torch.empty((H, W, 4), dtype=torch.float32, pin_memory=True)
    others(0)
    for f in range(frames):
        # the other ranks run ONE frame ahead (into the other buffer) while rank 0 renders and reads frame f.  (On ONE
        # device nothing may be enqueued whose wait is not already satisfied: streams of one process can share a hardware
        # queue, and a pending wait would block the very work it waits for.  Across GPUs the waits gate real run-ahead.)
        if f + 1 < frames:
            others(f + 1)
        ranks[0].update_camera(cams[f]); ranks[0].render_device(p)
        host.fill_(-1.0)
        ranks[0].frame_download(host.data_ptr())
        ranks[0].synchronize()
        assert np.array_equal(host.numpy().view(np.uint32), want[f].view(np.uint32)), "frame %d, world %d" % (f, world)
    for t in ranks:
        t.synchronize(); t.close()
