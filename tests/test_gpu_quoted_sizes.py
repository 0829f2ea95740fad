"""GPU parity ON THE CONFIGURATIONS THE NUMBERS ARE QUOTED ON (BASELINE.json configs[2], [3], [4]) at the north_star bar:
primary-hit ids bit-exact, colour max-abs 1e-3 per channel on >= 99.9 % of pixels, tie pixels counted and printed.

  * car 3840x2160 (the headline frame), default and Moller-Trumbore + Fresnel, against the oracle port AND — when
    oracle/_ref travelled to the box — against the reference shader file itself compiled as C++ (gpu_shader.comp:433-624);
  * soup-1M at 7680x4320, a row sample (the oracle needs ~1 ms per soup ray), single GPU layout and the 8-rank tile shards;
  * the 300-frame animated monkey sequence, every 25th frame, in the four modes bench.time_animated measures.

Everything goes through the C ABI (librt_b200.so via ctypes).
"""
import ctypes as C

import numpy as np
import pytest

from util import assert_parity, compare

pytestmark = pytest.mark.gpu


def _ties(oracle, sc, p, W, H, **kw):
    return oracle.primary_ties(sc.camera, sc.shapes, sc.nodes, sc.indices, p, W, H, **kw)


@pytest.mark.parametrize("mt,fresnel", [(False, False), (True, True)])
def test_car_4k_against_oracle_and_reference_build(rt, oracle, mt, fresnel):
    W, H = 3840, 2160
    sc = rt.Scene("car", W, H)
    p = rt.make_params(W, H, useMollerTrumbore=mt, useFresnel=fresnel)
    tr = rt.RayTracer(W, H)
    tr.upload(sc)
    tr.render(p)                                           # frame 2 runs with the heavy-first order of frame 1's costs
    rgba, ids = tr.render(p)
    tr.close()
    o_rgba, o_ids, o_d, _ = oracle.render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, p, W, H, want_counters=False)
    st = compare(rgba, ids, o_rgba, o_ids, o_d)
    st["tie_pixels"] = _ties(oracle, sc, p, W, H)
    assert_parity(st, "car 4K mt=%s fresnel=%s vs oracle" % (mt, fresnel))
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref did not travel: the CUDA <-> reference-build comparison needs it")
    r_rgba = oracle.ref_render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, p, W, H)
    st = compare(rgba, None, r_rgba, o_ids)
    assert_parity(st, "car 4K mt=%s fresnel=%s vs the reference shader compiled as C++" % (mt, fresnel))


@pytest.mark.parametrize("name,W,H", [("car", 800, 600), ("monkey", 800, 600), ("toy", 320, 240)])
def test_cuda_against_the_reference_build_directly(rt, oracle, name, W, H):
    """The chain CUDA <-> reference closed on the hardware: librt_b200.so against oracle/_ref/libref_shader.so."""
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref did not travel")
    sc = rt.Scene(name, W, H)
    for kw in (dict(), dict(useFresnel=True, useMollerTrumbore=True), dict(useBVH=False)):
        if name == "car" and kw.get("useBVH") is False:
            continue                                       # 4122 shapes x 480 k pixels by brute force on the CPU: too slow
        p = rt.make_params(W, H, **kw)
        tr = rt.RayTracer(W, H)
        tr.upload(sc)
        rgba, _ = tr.render(p)
        tr.close()
        r_rgba = oracle.ref_render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, p, W, H)
        assert_parity(compare(rgba, None, r_rgba, None), "%s %s vs reference build" % (name, kw))


def test_soup_1m_8k_row_sample_and_shards(rt, oracle):
    """BASELINE configs[3]: 1M-triangle soup + 10k spheres at 7680x4320.  The whole frame is rendered on the device (one
    GPU's layout, then as the 8 interleaved tile shards of the multi-GPU run, untiled); the oracle renders every 60th row."""
    import torch
    W, H = 7680, 4320
    step, phase = 60, 17
    sc = rt.Scene("soup", W, H, n_override=1000000)
    p = rt.make_params(W, H)
    tr = rt.RayTracer(W, H)
    tr.upload(sc)
    dev = torch.empty((H, W, 4), dtype=torch.float32, device="cuda")
    dids = torch.empty((H, W), dtype=torch.int32, device="cuda")
    tr.render_device(p, dev.data_ptr(), dids.data_ptr()); tr.synchronize()
    rows = np.arange(phase, H, step)
    rgba = dev[rows].cpu().numpy(); ids = dids[rows].cpu().numpy()
    o_rgba, o_ids, o_d, _ = oracle.render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, p, W, H,
                                          row_step=step, row_phase=phase, want_counters=False)
    st = compare(rgba, ids, o_rgba[rows], o_ids[rows], o_d[rows])
    st["tie_pixels"] = _ties(oracle, sc, p, W, H, row_step=step, row_phase=phase)
    assert_parity(st, "soup-1M 8K rows y%%%d==%d (%d rows)" % (step, phase, len(rows)))
    # the sharded layout of the 8-GPU run: every rank's compact shard, untiled, equals the single-GPU frame bit for bit
    world = 8
    tr.set_shard(0, world)
    px = tr.shard_pixels()
    shards = torch.zeros((world, px, 4), dtype=torch.float32, device="cuda")
    for r in range(world):
        tr.set_shard(r, world)
        tr.render_device(p, shards[r].data_ptr())
    out = torch.empty((H, W, 4), dtype=torch.float32, device="cuda")
    tr.untile(shards.data_ptr(), world, out.data_ptr()); tr.synchronize()
    assert torch.equal(out.view(torch.int32), dev.view(torch.int32)), "8 shards untiled != single-GPU frame"
    tr.close()


@pytest.mark.parametrize("mode,label", [(0, "refit"), (1, "rebuild"), (2, "device_animation"), (3, "device_rebuild"), (4, "device_rebuild_k9")])
def test_animated_monkey_300_frames(rt, oracle, mode, label):
    """BASELINE configs[4] exactly as bench.time_animated drives it (pipelined rt_render_async, two host buffers), every
    25th frame compared with the oracle on the host-side scene state of that frame (the reference's own frame step:
    bounceSphere / updateBVH or buildBVH, src/main.cpp:328-348, 1068-1082)."""
    import torch
    W, H, frames = 800, 600, 300
    L, Hl = rt.lib(), rt.hostlib()
    sc = rt.Scene("monkey", W, H)
    p = rt.make_params(W, H)
    tr = rt.RayTracer(W, H)
    tr.upload(sc)
    if mode >= 2:
        tr.register_animations(sc)
    bufs = [torch.empty((H * W, 4), dtype=torch.float32, pin_memory=True).numpy() for _ in range(2)]
    idb = [torch.empty(H * W, dtype=torch.int32, pin_memory=True).numpy() for _ in range(2)]
    anim = sc.animated
    lo, hi = int(anim.min()), int(anim.max()) + 1
    n = C.c_int(0)
    checked = 0
    worst = 1.0
    for i in range(frames):
        t, dt = i / 60.0, 1.0 / 60.0
        sc.animate(t, dt, 1 if mode in (1, 3, 4) else 0)   # the host's state of this frame = the oracle's input
        if mode == 2:
            tr.wait_frames(1); tr.animate(t, dt)
        elif mode == 3:
            tr.wait_frames(1); tr.animate(t, dt); tr.build_bvh(15, fetch=False)
        elif mode == 4:
            tr.wait_frames(1); tr.animate_shapes(t, dt); tr.build_bvh(15, rt.RT_BUILD_DEVICE_LAYOUT, fetch=False)
        else:
            shp = Hl.rh_shapes(sc._h, C.byref(n)); ns = n.value
            nod = Hl.rh_nodes(sc._h, C.byref(n)); nn = n.value
            if mode == 0:
                tr._chk(L.rt_update_shapes(tr._c, lo, hi - lo, C.c_void_p(shp + 192 * lo)))
                tr._chk(L.rt_update_nodes(tr._c, C.c_void_p(nod), nn))
            else:
                idx = Hl.rh_indices(sc._h, C.byref(n)); ni = n.value
                tr._chk(L.rt_upload_scene(tr._c, None, None, C.c_void_p(shp), ns, C.c_void_p(nod), nn, C.c_void_p(idx), ni))
            tr.wait_frames(1)
        tr.render_async(p, bufs[i & 1], idb[i & 1])
        if i % 25 == 0 or i == frames - 1:
            tr.synchronize()
            o_rgba, o_ids, o_d, _ = oracle.render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, p, W, H, want_counters=False)
            st = compare(bufs[i & 1].reshape(H, W, 4), idb[i & 1].reshape(H, W), o_rgba, o_ids, o_d)
            assert st["id_mismatch"] == 0 and st["frac_ok"] >= 0.999, "%s frame %d: %r" % (label, i, st)
            worst = min(worst, st["frac_ok"])
            checked += 1
    tr.close()
    print("C5 %s: %d of %d frames checked, ids bit-exact, worst fraction within 1e-3: %.6f" % (label, checked, frames, worst))
