"""Stateful fuzz of the C ABI's frame loop (gpu): random SEQUENCES of the calls a host makes between two frames — camera and
light updates, per-shape sub-updates (colour only, and type changes = topology changes), the reference's animation step done on
the host (grow-only refit or full rebuild, main.cpp:336-345, 1068-1077, 1175-1193) or on the device (rt_animate), full
re-uploads, a device rebuild (rt_build_bvh), shard changes — interleaved with every render entry point (rt_render,
rt_render_async, rt_render_rgba8, rt_render_device).  After every render the long-lived context must deliver exactly the bytes
of a FRESH context that was handed the same scene state in one rt_upload_scene, and (full frames) the oracle's image: whatever
the dirty flags, coalesced ranges, stale host shadows, kept layouts, heavy lists and deferred copies of the earlier calls were."""
import numpy as np
import pytest

from util import compare, assert_parity

pytestmark = pytest.mark.gpu

W, H = 208, 152          # ragged against the 32x8 tiles


class Truth:
    """The scene state the host believes the device has: the headless host scene + the test's own shape edits."""

    def __init__(self, rt, name):
        self.sc = rt.Scene(name, W, H)
        self.edits = {}                      # shape index -> record (applied over the host scene's shapes)
        self.light = self.sc.light
        self.nodes, self.indices = self.sc.nodes, self.sc.indices
        self.tree_from_device = False        # after rt_build_bvh the host scene's own tree is no longer the device's

    @property
    def camera(self):
        return self.sc.camera

    @property
    def shapes(self):
        s = self.sc.shapes
        for i, rec in self.edits.items():
            s[i:i + 1] = rec
        return s

    def sync_tree(self):
        self.nodes, self.indices = self.sc.nodes, self.sc.indices


def valid_pixels(rt, shard):
    """Mask of the compact shard buffer's pixels that belong to the image (the padding of the last tile slot and the pixels of
    ragged tiles beyond the frame are never written)."""
    rank, world = shard
    if world == 1:
        return np.ones(W * H, bool)
    m = np.zeros(rt.shard.shard_pixels(W, H, world), bool)
    for j, tile in enumerate(rt.shard.shard_tiles(W, H, rank, world)):
        ys, xs = rt.shard.tile_pixels(W, H, tile, world)
        m[j * rt.shard.TILE_PX:(j + 1) * rt.shard.TILE_PX] = (ys < H) & (xs < W)
    return m


def fresh_render(rt, T, p, shard):
    c = rt.RayTracer(W, H)
    c.upload_scene(T.light, T.camera, T.shapes, T.nodes, T.indices)
    if shard != (0, 1):
        c.set_shard(*shard)
    rgba, ids = c.render(p)
    c.close()
    return rgba.reshape(-1, 4), ids.ravel()


def render_some_way(rt, tr, p, how, oracle):
    """-> (rgba float32 (px, 4) or None, rgba8 (px, 4) or None, ids)"""
    import torch
    px = tr.shard_pixels()
    if how == "sync":
        rgba, ids = tr.render(p)
        return rgba.reshape(-1, 4), None, ids.ravel()
    if how == "async":
        pin = torch.empty((px, 4), dtype=torch.float32, pin_memory=True).numpy(); pin[:] = -9.0
        pin_ids = torch.empty(px, dtype=torch.int32, pin_memory=True).numpy(); pin_ids[:] = -9
        tr.render_async(p, pin, pin_ids)
        tr.wait_frames(0)
        return pin.copy(), None, pin_ids.copy()
    if how == "rgba8":
        ids = np.full(px, -9, np.int32)
        return None, tr.render_rgba8(p, out_ids=ids), ids
    dev = torch.empty((px, 4), dtype=torch.float32, device="cuda")
    dev_ids = torch.empty(px, dtype=torch.int32, device="cuda")
    tr.render_device(p, dev.data_ptr(), dev_ids.data_ptr())
    tr.synchronize()
    return dev.cpu().numpy(), None, dev_ids.cpu().numpy()


@pytest.mark.parametrize("name,seed", [(("monkey", "car", "toy", "car")[s % 4], s) for s in range(1, 25)])
def test_random_call_sequences(rt, oracle, name, seed):
    rng = np.random.default_rng(seed)
    T = Truth(rt, name)
    flags = rt.RT_FLAG_HIT_IDS | int(rng.choice([0, 0, rt.RT_FLAG_NO_TOP_TREE, rt.RT_FLAG_NO_HEAVY_FIRST, rt.RT_FLAG_NO_DEFERRED_COPY,
                                                 rt.RT_FLAG_NO_ONE_BLOCK_UPDATE, rt.RT_FLAG_REFERENCE_WALK]))
    tr = rt.RayTracer(W, H, flags=flags)
    tr.upload(T.sc)
    tr.register_animations(T.sc)
    an = T.sc.animated
    static = np.setdiff1d(np.arange(len(T.sc.shapes)), an)
    lo, hi = (int(an.min()), int(an.max()) + 1) if len(an) else (0, 0)
    cam0 = T.camera["Position"][0].copy()
    t, shard, log, checks, oracle_checks = 0.0, (0, 1), [], 0, 0
    ops = ["camera", "light", "colour", "type", "host_refit", "host_rebuild", "device_anim", "upload", "device_build", "shard", "params", "move"]
    wts = np.array([3, 2, 4, 2, 3, 2, 4, 1, 1, 1, 2, 2], float)
    p = rt.make_params(W, H)
    NSTEPS = 48
    for step in range(NSTEPS):
        op = str(rng.choice(ops, p=wts / wts.sum()))
        if op in ("host_refit", "host_rebuild", "device_anim") and (len(an) == 0 or T.tree_from_device):
            op = "colour"
        log.append(op)
        if op == "camera":
            pos = cam0 + rng.uniform(-6, 6, 3).astype(np.float32)
            T.sc.set_camera([float(v) for v in pos], [float(v) for v in rng.uniform(-4, 4, 3)])
            tr.update_camera(T.camera)
        elif op == "light":
            l = T.light.copy()
            l["position"][0][:3] += rng.uniform(-3, 3, 3).astype(np.float32)
            T.light = l
            tr.update_light(l)
        elif op in ("colour", "type"):
            for i in rng.choice(static, size=int(rng.integers(1, 4)), replace=False):
                i = int(i)
                rec = T.shapes[i:i + 1]
                rec["color"][0] = rng.uniform(0, 1, 3).astype(np.float32)
                if op == "type" and rec["type"][0] == 3:
                    rec["type"][0] = 0; rec["sphereCenter"][0] = rec["triP1"][0]; rec["sphereRadius"][0] = 0.6
                T.edits[i] = rec
                tr.update_shapes(i, rec)
        elif op == "move":                   # geometry changes under an unchanged tree (the shader trusts the boxes it is given)
            cur = T.shapes
            spheres = static[cur["type"][static] == 0]
            if len(spheres):
                i = int(rng.choice(spheres))
                rec = cur[i:i + 1]
                rec["sphereCenter"][0] += rng.uniform(-0.4, 0.4, 3).astype(np.float32)
                T.edits[i] = rec
                tr.update_shapes(i, rec)
        elif op == "host_refit":
            t += 0.19
            T.sc.animate(t, 0.19, mode=0); T.sync_tree()
            tr.update_shapes(lo, T.shapes[lo:hi]); tr.update_nodes(T.nodes)
        elif op == "host_rebuild":
            t += 0.23
            T.sc.animate(t, 0.23, mode=1); T.sync_tree()
            tr.update_shapes(lo, T.shapes[lo:hi])
            tr.upload_scene(nodes=T.nodes, indices=T.indices)
        elif op == "device_anim":
            t += 0.17
            T.sc.animate(t, 0.17, mode=0); T.sync_tree()
            tr.animate(t, 0.17)
        elif op == "upload":
            tr.upload_scene(T.light, T.camera, T.shapes, T.nodes, T.indices)
        elif op == "device_build":
            depth = int(rng.integers(4, 20))
            T.nodes, T.indices = tr.build_bvh(depth, build_flags=int(rng.integers(0, 2)) * rt.RT_BUILD_DEVICE_LAYOUT)
            T.tree_from_device = True
        elif op == "shard":
            world = int(rng.integers(1, 4))
            shard = (int(rng.integers(0, world)), world)
            tr.set_shard(*shard)
        elif op == "params":
            p = rt.make_params(W, H, maxBounces=int(rng.integers(0, 5)), useFresnel=bool(rng.integers(0, 2)),
                               useMollerTrumbore=bool(rng.integers(0, 2)), useBVH=bool(rng.integers(0, 8) > 0))
        if rng.uniform() < 0.6 or step == NSTEPS - 1:
            how = str(rng.choice(["sync", "async", "rgba8", "device"]))
            log.append("render:" + how)
            rgba, rgba8, ids = render_some_way(rt, tr, p, how, oracle)
            f_rgba, f_ids = fresh_render(rt, T, p, shard)
            what = "%s seed %d flags %d step %d shard %s after %s" % (name, seed, flags, step, shard, " ".join(log[-12:]))
            m = valid_pixels(rt, shard)
            assert np.array_equal(ids[m], f_ids[m]), "hit ids differ from a fresh context: " + what
            if rgba is not None:
                assert np.array_equal(rgba.view(np.uint32)[m], f_rgba.view(np.uint32)[m]), "frame differs from a fresh context: " + what
            else:
                assert np.array_equal(rgba8[m], oracle.present(f_rgba).reshape(-1, 4)[m]), "shown frame differs from a fresh context's: " + what
            checks += 1
            if shard == (0, 1) and rgba is not None and checks % 3 == 0:
                o_rgba, o_ids, _, _ = oracle.render(T.light, T.camera, T.shapes, T.nodes, T.indices, p, W, H, want_counters=False)
                assert_parity(compare(rgba, ids, o_rgba, o_ids), "oracle: " + what)
                oracle_checks += 1
    assert checks >= 10
    tr.close()


@pytest.mark.parametrize("name,seed,shown", [("monkey", 31, False), ("car", 32, False), ("monkey", 33, True), ("car", 34, True), ("toy", 35, False)])
def test_random_pipelined_sequences(rt, oracle, name, seed, shown):
    """The same call mix between PIPELINED frames (rt_render_async / rt_render_rgba8_async over two host buffers with
    rt_wait_frames(1): a frame's host copy is deferred to the next frame's render kernel): every delivered frame equals a
    fresh context's synchronous render of the scene state at the time the frame was issued, although the host keeps
    rewriting the pinned shadows while the previous frame is still in flight."""
    import torch
    rng = np.random.default_rng(seed)
    T = Truth(rt, name)
    tr = rt.RayTracer(W, H, flags=0)
    tr.upload(T.sc)
    tr.register_animations(T.sc)
    an = T.sc.animated
    static = np.setdiff1d(np.arange(len(T.sc.shapes)), an)
    lo, hi = (int(an.min()), int(an.max()) + 1) if len(an) else (0, 0)
    cam0 = T.camera["Position"][0].copy()
    p = rt.make_params(W, H)
    dt_, n_ = (torch.uint8, np.uint8) if shown else (torch.float32, np.float32)
    bufs = [torch.empty((H * W, 4), dtype=dt_, pin_memory=True).numpy() for _ in range(2)]
    expected, t, N = [], 0.0, 40
    for i in range(N):
        for _ in range(int(rng.integers(1, 4))):
            op = str(rng.choice(["camera", "light", "colour", "type", "host_refit", "host_rebuild", "device_anim", "upload"],
                                p=np.array([3, 1, 3, 1, 3, 1, 4, 1]) / 17.0))
            if op == "camera":
                T.sc.set_camera([float(v) for v in cam0 + rng.uniform(-5, 5, 3).astype(np.float32)], [float(v) for v in rng.uniform(-4, 4, 3)])
                tr.update_camera(T.camera)
            elif op == "light":
                l = T.light.copy(); l["position"][0][:3] += rng.uniform(-2, 2, 3).astype(np.float32)
                T.light = l; tr.update_light(l)
            elif op in ("colour", "type"):
                k = int(rng.choice(static))
                rec = T.shapes[k:k + 1]
                rec["color"][0] = rng.uniform(0, 1, 3).astype(np.float32)
                if op == "type" and rec["type"][0] == 3:
                    rec["type"][0] = 0; rec["sphereCenter"][0] = rec["triP1"][0]; rec["sphereRadius"][0] = 0.5
                T.edits[k] = rec; tr.update_shapes(k, rec)
            elif op == "host_refit":
                t += 0.19; T.sc.animate(t, 0.19, mode=0); T.sync_tree()
                tr.update_shapes(lo, T.shapes[lo:hi]); tr.update_nodes(T.nodes)
            elif op == "host_rebuild":
                t += 0.23; T.sc.animate(t, 0.23, mode=1); T.sync_tree()
                tr.update_shapes(lo, T.shapes[lo:hi]); tr.upload_scene(nodes=T.nodes, indices=T.indices)
            elif op == "device_anim":
                t += 0.17; T.sc.animate(t, 0.17, mode=0); T.sync_tree(); tr.animate(t, 0.17)
            else:
                tr.upload_scene(T.light, T.camera, T.shapes, T.nodes, T.indices)
        f_rgba, _ = fresh_render(rt, T, p, (0, 1))
        expected.append(oracle.present(f_rgba).reshape(-1, 4) if shown else f_rgba.copy())
        tr.wait_frames(1)
        if i >= 2:
            assert np.array_equal(bufs[i & 1], expected[i - 2]), "%s seed %d: delivered frame %d" % (name, seed, i - 2)
        bufs[i & 1][:] = 3
        (tr.render_rgba8_async if shown else tr.render_async)(p, bufs[i & 1])
    tr.wait_frames(0)
    for i in (N - 2, N - 1):
        assert np.array_equal(bufs[i & 1], expected[i]), "%s seed %d: delivered frame %d" % (name, seed, i)
    tr.close()


@pytest.mark.parametrize("Wx,Hx", [(1, 1), (1, 300), (4097, 1), (33, 9), (31, 7), (8, 4)])
def test_extreme_frame_sizes(rt, oracle, Wx, Hx):
    """Frames of one pixel, one column, one row (more than 128 tiles wide), a tile plus one pixel: every entry point against the
    oracle, and as three compact shards untiled."""
    sc = rt.Scene("toy", Wx, Hx)
    p = rt.make_params(Wx, Hx, useFresnel=True)
    tr = rt.RayTracer(Wx, Hx); tr.upload(sc)
    rgba, ids = tr.render(p)
    o_rgba, o_ids, _, _ = oracle.render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, p, Wx, Hx, want_counters=False)
    assert_parity(compare(rgba, ids, o_rgba, o_ids), "%dx%d" % (Wx, Hx))
    assert np.array_equal(tr.render_rgba8(p), oracle.present(rgba.reshape(-1, 4)).reshape(-1, 4))
    shards = []
    for r in range(3):
        tr.set_shard(r, 3)
        s, _ = tr.render(p)
        shards.append(np.where(np.repeat(_valid(rt, Wx, Hx, (r, 3))[:, None], 4, 1), s.reshape(-1, 4), 0.0))
    whole = rt.shard.untile_numpy(np.stack(shards), Wx, Hx, 3)
    assert np.array_equal(whole.view(np.uint32), rgba.reshape(Hx, Wx, 4).view(np.uint32))
    tr.close()


def _valid(rt, Wx, Hx, shard):
    rank, world = shard
    m = np.zeros(rt.shard.shard_pixels(Wx, Hx, world), bool)
    for j, tile in enumerate(rt.shard.shard_tiles(Wx, Hx, rank, world)):
        ys, xs = rt.shard.tile_pixels(Wx, Hx, tile, world)
        m[j * rt.shard.TILE_PX:(j + 1) * rt.shard.TILE_PX] = (ys < Hx) & (xs < Wx)
    return m


def test_two_contexts_on_two_host_threads(rt, oracle):
    """A context is single-threaded like a GL context, but two contexts may be driven from two host threads at once (ctypes
    drops the GIL during the calls): an animated car on one, an animated monkey with device animation on the other — every frame
    equals the frame the same calls give when the threads run one after the other."""
    import threading

    def run(name, device_anim, out):
        sc = rt.Scene(name, W, H)
        p = rt.make_params(W, H, useFresnel=(name == "car"))
        tr = rt.RayTracer(W, H); tr.upload(sc)
        if device_anim:
            tr.register_animations(sc)
        an = sc.animated
        lo, hi = int(an.min()), int(an.max()) + 1
        for f in range(1, 25):
            if device_anim:
                tr.animate(f * 0.1, 0.1)
            else:
                sc.animate(f * 0.1, 0.1, mode=f % 2)
                tr.update_shapes(lo, sc.shapes[lo:hi])
                if f % 2:
                    tr.upload_scene(nodes=sc.nodes, indices=sc.indices)
                else:
                    tr.update_nodes(sc.nodes)
            rgba, ids = tr.render(p)
            out.append((rgba.copy(), ids.copy()))
        tr.close()

    serial = ([], [])
    run("car", False, serial[0]); run("monkey", True, serial[1])
    both = ([], [])
    errs = []

    def guarded(*a):
        try:
            run(*a)
        except Exception as e:                                  # surfaces in the main thread below
            errs.append(e)
    ths = [threading.Thread(target=guarded, args=("car", False, both[0])), threading.Thread(target=guarded, args=("monkey", True, both[1]))]
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    assert not errs, errs
    for k in range(2):
        assert len(both[k]) == len(serial[k]) == 24
        for f, ((ra, ia), (rb, ib)) in enumerate(zip(both[k], serial[k])):
            assert np.array_equal(ia, ib) and np.array_equal(ra.view(np.uint32), rb.view(np.uint32)), "context %d frame %d" % (k, f)


def test_argument_errors_leave_the_context_usable(rt):
    """Bad arguments come back as negative codes with a message (never abort, never print: computeShader.hpp:48-79 prints and
    carries on); the context renders the same frame afterwards."""
    import ctypes as C
    L = rt.lib()
    h = C.c_void_p()
    assert L.rt_create(C.byref(h), 0, 10, 0, 0) < 0 and L.rt_create(C.byref(h), 10, -1, 0, 0) < 0
    assert L.rt_create(C.byref(h), 16, 16, 4096, 0) < 0                  # no such device
    sc = rt.Scene("toy", 64, 48)
    p = rt.make_params(64, 48)
    tr = rt.RayTracer(64, 48); tr.upload(sc)
    ref, ref_ids = tr.render(p)
    n = len(sc.shapes)
    for bad in (lambda: tr.update_shapes(n, sc.shapes[:1]), lambda: tr.update_shapes(-1, sc.shapes[:1]),
                lambda: tr.update_shapes(n - 1, sc.shapes[:2]), lambda: tr.set_shard(2, 2), lambda: tr.set_shard(0, 0),
                lambda: tr._chk(L.rt_render(tr._c, None, None, None)),
                lambda: tr.build_bvh(65), lambda: tr.build_bvh(-1)):
        with pytest.raises(rt.RtError) as e:
            bad()
        assert e.value.code < 0 and str(e.value)
    again, again_ids = tr.render(p)
    assert np.array_equal(again.view(np.uint32), ref.view(np.uint32)) and np.array_equal(again_ids, ref_ids)
    tr.close()
