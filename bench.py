#!/usr/bin/env python3
"""bench.py — the reference's headline metric (Mrays/s, frames/s on the car scene) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one frame of the hot path (primary ray -> BVH traversal -> closest hit -> shadow ray ->
Phong/Fresnel -> reflection bounces) over the whole image.  Workload: BASELINE.json configs[2]/[1],
the car scene (4022 triangles + 100 spheres; synthetic stand-in mesh, the .obj is absent upstream)
at 3840x2160, defaults maxBounces=3 useBVH=1 useFresnel=0 useMollerTrumbore=0, tile-sharded over
the N ranks with an NCCL gather to rank 0 (strong scaling: the frame is fixed).  At N=1 the 800x600
frame of configs[1] is measured too and reported under "extra".

  value     Mrays/s with the scene resident in HBM: rays of the REFERENCE algorithm for this frame
            (oracle-counted constants in tests/golden/workloads.json, cross-checked against the CUDA
            path's own counter) / device time of render (+ gather + untile when N > 1), CUDA events,
            max over ranks, L2 flushed between steps.
  e2e       same metric through the C ABI with HOST buffers: rt_upload_scene (pinned staging, H2D,
            repack) + render + D2H of the rgba32f frame inside the timed region.
  roofline  the render kernel is bound by warp-instruction ISSUE (the scene is L1/L2 resident): achieved = warp
            instructions per launch (ncu capture of this kernel, profiles/traffic.json) / live kernel time, peak = 4
            schedulers x 148 SMs x the SM clock sampled during the run; beside it the MEASURED traffic of every memory level
            over the live kernel time against MEASURED peaks (HBM: MEASURED_PEAKS.json, L1 / L2: tools/l2_probe.cu ->
            profiles/l2_peak.json) and, under alg_normalised, SURVEY 8d's algorithmic bytes of the reference traversal.
  parity    the timed frame against the oracle: ids differing, fraction of pixels within max-abs 1e-3, tie pixels.
  cpu_baseline  the reference shader compiled as C++ (oracle/_ref) or the oracle port, on the host
            cores, on a bounded row sample of the same frame (N=1, rank 0 only).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--scene", default="car")
    ap.add_argument("--width", type=int, default=3840)
    ap.add_argument("--height", type=int, default=2160)
    ap.add_argument("--mt", type=int, default=0)
    ap.add_argument("--fresnel", type=int, default=0)
    ap.add_argument("--bounces", type=int, default=3)
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="CPU-baseline budget (core-seconds are bounded by a row sample)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle comparison of the timed frame")
    ap.add_argument("--gather", default="fused", choices=["fused", "barrier", "nccl"],
                    help="N>1: fused = every rank's render kernel stores its tiles into rank 0's frame over NVLink peer memory and its "
                         "last block signals completion (no collective per frame); barrier = the same stores with one tiny all-reduce "
                         "as completion barrier; nccl = compact shards + dist.gather + untile kernel")
    ap.add_argument("--no-flush", action="store_true", help="diagnostic only: keep L2 warm between steps (not a valid bench line)")
    return ap.parse_args()


def workload_constants(scene, W, H, mt, fresnel, bounces):
    path = os.path.join(ROOT, "tests", "golden", "workloads.json")
    key = "%s_%dx%d_mt%d_fr%d_b%d" % (scene, W, H, mt, fresnel, bounces)
    try:
        return json.load(open(path)).get(key)
    except OSError:
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------
# CPU legs (the ONLY place bench.py touches oracle/): cpu_baseline and --impl reference
# ---------------------------------------------------------------------------------------------
def cpu_sample_plan(rt, O, sc, params, W, H, budget_s, threads):
    """Choose a row stride so that one sample frame costs about budget_s seconds of wall time."""
    probe_step = max(1, H // 16)
    t0 = time.perf_counter()
    O.render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, params, W, H, threads=threads,
             row_step=probe_step, row_phase=probe_step // 2, want_counters=False)
    dt = max(time.perf_counter() - t0, 1e-4)
    rows_probe = len(range(probe_step // 2, H, probe_step))
    per_row = dt / rows_probe
    rows = int(min(H, max(8, budget_s / per_row)))
    step = max(1, H // rows)
    return step, step // 2


_RAYS_CACHE = {}


SCENE_BVH_DEPTH = {"monkey": 15, "car": 25, "soup": 40, "triangle": 15, "toy": 8}    # buildBVH's argument per scene (host/rt_host.cpp, main.cpp:712, :800)


class _SceneView:
    """The scene bytes of the reference arm / the CPU baseline: shapes, light and camera of the synthetic scene, the BVH as
    UPSTREAM's own buildBVH + serializeBVH (cut out of src/main.cpp, oracle/_ref/libref_bvh.so) builds it when that library
    travelled with the snapshot — checked to be the byte-identical tree the library under test is given."""

    def __init__(self, rt, O, sc, name):
        self.light, self.camera, self.shapes = sc.light, sc.camera, sc.shapes
        self.nodes, self.indices = sc.nodes, sc.indices
        self.bvh = {"builder": "host/rt_host.cpp (pinned to upstream by tests/test_ref_builder.py)"}
        try:
            if O.refbvh_lib() is not None and name in SCENE_BVH_DEPTH and len(sc.shapes) <= 200000:
                nodes, idx = O.RefScene(sc.shapes, node_dtype=rt.NODE_DTYPE).build(SCENE_BVH_DEPTH[name])
                mine = np.asarray(sc.nodes)
                same = (len(nodes) == len(mine) and np.array_equal(idx, sc.indices) and
                        all(np.array_equal(nodes[f].view(np.uint32), mine[f].view(np.uint32)) for f in rt.NODE_DTYPE.names))
                self.nodes, self.indices = nodes, idx
                self.bvh = {"builder": "upstream buildBVH + serializeBVH compiled in place (oracle/_ref/libref_bvh.so)",
                            "every_field_identical_to_the_tree_given_to_the_library": bool(same)}   # (upstream leaves FlatNode's padding unset)
        except Exception as e:      # the baseline must not depend on it
            self.bvh["upstream_builder_error"] = repr(e)


def cpu_render_timed(O, sc, params, W, H, threads, step, phase, use_ref):
    """Time one sample frame on the host cores; returns (seconds, rays in the sampled rows)."""
    args = (sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, params, W, H)
    key = (W, H, step, phase)
    if key not in _RAYS_CACHE:      # rays in exactly these rows, from the oracle's counters (untimed, once)
        _RAYS_CACHE[key] = O.render(*args, threads=threads, row_step=step, row_phase=phase)[3]["rays"]
    t0 = time.perf_counter()
    if use_ref:
        O.ref_render(*args, threads=threads, row_step=step, row_phase=phase)
    else:
        O.render(*args, threads=threads, row_step=step, row_phase=phase, want_counters=False)
    return time.perf_counter() - t0, _RAYS_CACHE[key]


def cpu_baseline(rt, sc, params, W, H, budget_s, repeats=1, scene_name=None):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as O
    O.build()
    sc = _SceneView(rt, O, sc, scene_name)
    threads = os.cpu_count() or 1
    use_ref = O.ref_lib() is not None
    step, phase = cpu_sample_plan(rt, O, sc, params, W, H, budget_s, threads)
    best, rays = None, 0
    for _ in range(repeats):
        dt, rays = cpu_render_timed(O, sc, params, W, H, threads, step, phase, use_ref)
        best = dt if best is None else min(best, dt)
    rows = len(range(phase, H, step))
    return {"value": rays / best / 1e6, "unit": "Mrays/s", "cores": threads,
            "kind": "reference" if use_ref else "port",
            "sample": "rows y%%%d==%d (%d of %d rows, %d rays) of the %dx%d frame, %.2f s, %s" %
                      (step, phase, rows, H, rays, W, H, best,
                       "reference gpu_shader.comp compiled as C++ (oracle/_ref)" if use_ref else "oracle port (oracle_rt.cpp)"),
            "ms_per_frame_extrapolated": best * 1e3 * H / rows, "bvh": sc.bvh}, (step, phase, use_ref, threads)


def run_reference_arm(args):
    """--impl reference: the reference's own CPU implementation of the path (its shader compiled as
    C++ when oracle/_ref travelled with the snapshot, else the oracle port), all host threads,
    each step a bounded row sample of the same frame."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import rtload
    rt = rtload.load()
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as O
    O.build()
    W, H = args.width, args.height
    sc = _SceneView(rt, O, rt.Scene(args.scene, W, H), args.scene)
    params = rt.make_params(W, H, maxBounces=args.bounces, useFresnel=bool(args.fresnel), useMollerTrumbore=bool(args.mt))
    threads = os.cpu_count() or 1
    use_ref = O.ref_lib() is not None
    total = max(1, args.steps + args.warmup)
    step, phase = cpu_sample_plan(rt, O, sc, params, W, H, min(args.cpu_seconds, 120.0 / total), threads)
    for _ in range(args.warmup):
        cpu_render_timed(O, sc, params, W, H, threads, step, phase, use_ref)
    times, rays = [], 0
    for _ in range(args.steps):
        dt, rays = cpu_render_timed(O, sc, params, W, H, threads, step, phase, use_ref)
        times.append(dt)
    ms = 1e3 * sum(times) / len(times)
    value = rays / (ms / 1e3) / 1e6
    rows = len(range(phase, H, step))
    kind = "reference" if use_ref else "port"
    sample = "rows y%%%d==%d (%d of %d rows, %d rays/step) of the %s %dx%d frame" % (step, phase, rows, H, rays, args.scene, W, H)
    print(json.dumps({
        "impl": "reference", "metric": "Mrays/s (%s scene)" % args.scene, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, sample=sample),
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": threads, "kind": kind, "sample": sample, "bvh": sc.bvh},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "fps_extrapolated": 1e3 / (ms * H / rows),
        "note": "reference GLSL through llvmpipe/GL: unavailable on this box; this is the shader's arithmetic on the host cores",
    }))


def memory_levels(prof, kernel_ms, hbm_peak):
    """The render kernel's MEASURED traffic per launch (one ncu --set full capture of this kernel on this workload,
    profiles/traffic.json) over the LIVE kernel time, level by level, against measured peaks: HBM from
    MEASURED_PEAKS.json, L1 / L2 read bandwidth from tools/l2_probe.cu (profiles/l2_peak.json)."""
    try:
        l1 = prof["l1_global_load_bytes"] + prof["l1_local_load_bytes"]
        out = {"l1_load_GBps": l1 / kernel_ms / 1e6, "l1_global_load_GBps": prof["l1_global_load_bytes"] / kernel_ms / 1e6,
               "l1_local_load_GBps (register spills)": prof["l1_local_load_bytes"] / kernel_ms / 1e6,
               "l2_read_from_sm_GBps": prof["l2_read_bytes_from_sm"] / kernel_ms / 1e6,
               "dram_GBps": prof["dram_bytes"] / kernel_ms / 1e6, "dram_read_GBps": prof["dram_read_bytes"] / kernel_ms / 1e6,
               "hbm_peak_GBps": hbm_peak, "hbm_frac": prof["dram_bytes"] / kernel_ms / 1e6 / hbm_peak,
               "l1_hit_pct": prof.get("l1_hit_pct"), "source": prof.get("source"), "capture_commit": prof.get("commit")}
    except (KeyError, TypeError, ZeroDivisionError):
        return None
    try:
        pk = json.load(open(os.path.join(ROOT, "profiles", "l2_peak.json")))
        out["l1_peak_GBps"], out["l2_peak_GBps"] = pk.get("l1_gbs"), pk.get("l2_gbs")
        out["peaks_source"] = "measured: tools/l2_probe.cu on this pool's B200 (profiles/l2_peak.json); HBM: MEASURED_PEAKS.json"
        if pk.get("l2_gbs"):
            out["l2_frac"] = out["l2_read_from_sm_GBps"] / pk["l2_gbs"]
        if pk.get("l1_gbs"):
            out["l1_frac"] = out["l1_load_GBps"] / pk["l1_gbs"]
    except (OSError, ValueError):
        out["l1_peak_GBps"] = out["l2_peak_GBps"] = None
    return out


def parity_block(rt, sc, params, W, H, rgba, ids):
    """The timed frame against the oracle (the checker, never the thing measured): north_star's bar."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as O
    O.build()
    o_rgba, o_ids, _, _ = O.render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, params, W, H, want_counters=False)
    diff = np.abs(np.asarray(rgba, np.float64).reshape(o_rgba.shape) - o_rgba.astype(np.float64))
    diff[np.isnan(diff)] = np.inf
    out = {"ids_mismatch": int((np.asarray(ids).reshape(o_ids.shape) != o_ids).sum()) if ids is not None else None,
           "frac_within_1e-3": float((diff <= 1e-3).all(axis=-1).mean()), "frac_within_1e-4": float((diff <= 1e-4).all(axis=-1).mean()),
           "max_abs": float(diff[np.isfinite(diff)].max()) if np.isfinite(diff).any() else None,
           "tie_pixels": O.primary_ties(sc.camera, sc.shapes, sc.nodes, sc.indices, params, W, H),
           "pixels": int(W * H), "against": "oracle/oracle_rt.cpp (pinned to the reference shader compiled as C++, tests/test_oracle_golden.py)"}
    if O.ref_lib() is not None:                     # the reference shader file itself, when oracle/_ref travelled
        r = O.ref_render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, params, W, H)
        d2 = np.abs(np.asarray(rgba, np.float64).reshape(r.shape) - r.astype(np.float64))
        d2[np.isnan(d2)] = np.inf
        out["frac_within_1e-3_vs_reference_build"] = float((d2 <= 1e-3).all(axis=-1).mean())
    return out


def workload_config(args, **more):
    which = {("car", 3840, 2160): "BASELINE.json configs[2]", ("car", 800, 600): "BASELINE.json configs[1]",
             ("monkey", 800, 600): "BASELINE.json configs[0]", ("soup", 7680, 4320): "BASELINE.json configs[3]"}.get(
                 (args.scene, args.width, args.height), "not a BASELINE.json configuration")
    cfg = {"workload": "%s scene %dx%d, maxBounces=%d useBVH=1 useFresnel=%d useMollerTrumbore=%d "
                       "(%s; synthetic stand-in mesh, same primitive counts)" %
                       (args.scene, args.width, args.height, args.bounces, args.fresnel, args.mt, which),
           "scene": args.scene, "width": args.width, "height": args.height,
           "parallelism": "tiles%d" % args.gpus, "tile": "32x8 interleaved",
           "l2": "NOT flushed (diagnostic run)" if args.no_flush else "flushed between steps (256 MiB write)"}
    cfg.update(more)
    return cfg


# ---------------------------------------------------------------------------------------------
# The B200 arm
# ---------------------------------------------------------------------------------------------
def setup_fused(torch, dist, tr, rank, world, signalled=True):
    """Rank 0 exports its (double-buffered) frame over CUDA IPC, the others map it (peer access over NVLink).  Signalled
    mode on top: every other rank exports its release word to rank 0 — from then on a frame needs no collective at all
    (completion words written by the render kernels' last blocks, stream memory operations on the waiting side)."""
    box = [tr.frame_export() if rank == 0 else None]
    handle = [box[0][0] if rank == 0 else None]
    dist.broadcast_object_list(handle, src=0)
    if rank != 0:
        tr.frame_open(handle[0])
    mode = "nccl_barrier"
    if signalled:
        try:
            mine = tr.sync_export() if rank != 0 else None
            words = [None] * world
            dist.all_gather_object(words, mine)
            if rank == 0:
                for r in range(1, world):
                    tr.sync_peer(r, handle=words[r])
            tr.sync_enable(world)
            ok = torch.ones(1, dtype=torch.int32, device="cuda")
        except Exception:
            ok = torch.zeros(1, dtype=torch.int32, device="cuda")
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)          # all ranks or none (also the barrier after the words were zeroed)
        if int(ok[0]) == 1:
            mode = "signalled"
        else:
            tr.sync_enable(0)
    return (box[0][1] if rank == 0 else None), mode


def setup_host_frame(dist, tr, rank, world, W, H):
    """One W*H rgba32f frame in POSIX shared memory, mapped by every rank and pinned + mapped into each
    rank's GPU (rt_host_frame_register): the render kernels store their tiles straight into it, every
    GPU over its own PCIe link.  Returns (numpy view, device alias)."""
    name = ["/dev/shm/rt_b200_frame_%d" % os.getpid() if rank == 0 else None]
    if rank == 0:
        arr = np.memmap(name[0], dtype=np.float32, mode="w+", shape=(H * W * 4,))
        arr[:] = 0
    if world > 1:
        dist.broadcast_object_list(name, src=0)
        if rank != 0:
            arr = np.memmap(name[0], dtype=np.float32, mode="r+", shape=(H * W * 4,))
        dist.barrier()
    if rank == 0:
        os.unlink(name[0])                      # the mappings keep the segment alive
    return arr, tr.host_frame_register(arr)


def time_frames(torch, dist, rt, tr, params, W, H, rank, world, steps, warmup, flush, fused=False):
    """Device-timed frames with the scene resident.  Returns per-rank mean ms (total, render, gather).
    fused: False = NCCL gather + untile; "nccl_barrier" = peer stores + one all-reduce; "signalled" = peer stores + completion
    words (rank 0 acquires and releases the frame in its stream; nothing collective)."""
    px = tr.shard_pixels()
    signalled = fused == "signalled"
    token = torch.zeros(1, dtype=torch.int32, device="cuda") if (fused and world > 1 and not signalled) else None
    shard = torch.empty((px, 4), dtype=torch.float32, device="cuda")
    full = torch.empty((H, W, 4), dtype=torch.float32, device="cuda") if (rank == 0 and world > 1) else None
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731

    def frame(timed):
        e0, e1, e2 = ev(), ev(), ev()
        if flush is not None:
            flush.zero_()                       # evict the scene and the image from L2 (untimed)
        e0.record()
        if signalled:
            tr.render_device(params)            # tiles go straight into rank 0's frame; the last block signals
            e1.record()
            if rank == 0:
                tr.frame_acquire()              # the stream waits for every rank's completion word
                tr.frame_release()              # ... and tells them the buffer may be reused two frames on
        elif token is not None:
            tr.render_device(params)            # tiles go straight into rank 0's frame (peer stores)
            e1.record()
            dist.all_reduce(token)              # completion barrier: the only collective left
        else:
            tr.render_device(params, shard.data_ptr())
            e1.record()
            if world > 1:
                got = rt.shard.gather_shards(shard, rank, world)
                if rank == 0:
                    tr.untile(got.data_ptr(), world, full.data_ptr())
        e2.record()
        return (e0, e1, e2) if timed else None

    for _ in range(warmup):
        frame(False)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    evs = [frame(True) for _ in range(steps)]
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    tot = sum(a.elapsed_time(c) for a, _, c in evs) / steps
    ren = sum(a.elapsed_time(b) for a, b, _ in evs) / steps
    return tot, ren, tot - ren, (full if full is not None else shard)


def time_e2e(torch, dist, rt, tr, sc, params, W, H, rank, world, steps, warmup, frame_ptr=None, fused=False, host_alias=None, rgba8=False):
    """End to end through the C ABI from HOST buffers: upload (pinned staging + H2D + repack), render,
    (gather + untile), D2H of the rgba32f frame into pinned host memory.  Wall clock, sync both sides.
    host_alias: the frame target is a mapped host frame (setup_host_frame), the kernels' own stores are the D2H."""
    light, cam, shapes, nodes, idx = sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices
    px = tr.shard_pixels()
    if host_alias is not None:
        token = torch.zeros(1, dtype=torch.int32, device="cuda")
    elif world == 1 and rgba8:                  # the frame as the reference shows it: rt_render_rgba8 (K1 + K5, 4 bytes per pixel to the host)
        host = torch.empty((H * W, 4), dtype=torch.uint8, pin_memory=True).numpy()
    elif world == 1:
        host = torch.empty((H * W, 4), dtype=torch.float32, pin_memory=True).numpy()
    else:
        shard = torch.empty((px, 4), dtype=torch.float32, device="cuda")
        full = torch.empty((H, W, 4), dtype=torch.float32, device="cuda") if rank == 0 else None
        host = torch.empty((H, W, 4), dtype=torch.float32, pin_memory=True) if rank == 0 else None
        token = torch.zeros(1, dtype=torch.int32, device="cuda")

    def frame():
        tr.upload_scene(light, cam, shapes, nodes, idx)
        if host_alias is not None:
            tr.render_device(params)
            if world > 1:
                dist.all_reduce(token)              # completion: every rank's tiles are in the host frame
            torch.cuda.synchronize()
            if world > 1:
                dist.all_reduce(token)              # release: the consumer is done with the frame, the next step may overwrite it
                torch.cuda.synchronize()
        elif world == 1 and rgba8:
            tr.render_rgba8(params, out=host)
        elif world == 1:
            tr.render(params, want_ids=False, out=host)
        elif fused == "signalled":
            tr.render_device(params)
            if rank == 0:
                tr.frame_download(host.data_ptr())  # acquire (completion words) + copy + release: no collective, no host barrier
            torch.cuda.synchronize()
        elif fused:
            tr.render_device(params)
            dist.all_reduce(token)                  # completion barrier
            if rank == 0:
                tr.frame_download(host.data_ptr())
            torch.cuda.synchronize()
            dist.all_reduce(token)                  # release barrier: rank 0 has read its frame (rt_api.h: rt_frame_export)
            torch.cuda.synchronize()
        else:
            tr.render_device(params, shard.data_ptr())
            got = rt.shard.gather_shards(shard, rank, world)
            if rank == 0:
                tr.untile(got.data_ptr(), world, full.data_ptr())
                host.copy_(full, non_blocking=True)
            torch.cuda.synchronize()

    for _ in range(warmup):
        frame()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        frame()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = (time.perf_counter() - t0) * 1e3 / steps
    # bytes that cross PCIe per step on every rank / on rank 0
    h2d = int(shapes.nbytes + nodes.nbytes + len(nodes) * 12 + len(idx) * 4)     # raw scene + visit-order maps
    d2h = int(W * H * (4 if rgba8 else 16))
    return ms, h2d, d2h


def time_animated(torch, rt, stream, frames=300, W=800, H=600):
    """BASELINE.json configs[4]: the animated monkey scene, `frames` frames, per-frame CPU animation +
    BVH refit (reference behaviour: grow-only updateBVH, main.cpp:1068-1077) or full rebuild, per-shape
    rt_update_shapes + rt_update_nodes / full upload through pinned staging, render, D2H of the frame.
    End-to-end wall-clock fps, host buffers double-buffered (rt_render_async / rt_wait_frames) so that the
    host's animation step and the render of frame i+1 overlap the copy of frame i."""
    import ctypes as C
    out = {}
    L, Hl = rt.lib(), rt.hostlib()
    for mode, label in ((0, "refit"), (1, "rebuild"), (2, "device_animation"), (3, "device_rebuild"), (4, "device_rebuild_k9")):
        sc = rt.Scene("monkey", W, H)
        params = rt.make_params(W, H)
        tr = rt.RayTracer(W, H, device=torch.cuda.current_device(), flags=0)
        tr.set_stream(stream.cuda_stream)
        tr.upload(sc)
        if mode >= 2:
            tr.register_animations(sc)
        bufs = [torch.empty((H * W, 4), dtype=torch.float32, pin_memory=True).numpy() for _ in range(2)]
        n = C.c_int(0)
        anim = sc.animated
        lo, hi = int(anim.min()), int(anim.max()) + 1

        def frame(i):
            if mode == 2:                         # K6 + K7 on the device copies: nothing to upload
                tr.wait_frames(1)
                tr.animate(i / 60.0, 1.0 / 60.0)
                tr.render_async(params, bufs[i & 1])
                return
            if mode == 3:                         # K6 moves the shapes, K8 rebuilds the reference tree on the device
                tr.wait_frames(1)
                tr.animate(i / 60.0, 1.0 / 60.0)
                tr.build_bvh(15, fetch=False)
                tr.render_async(params, bufs[i & 1])
                return
            if mode == 4:                         # K6 only, K8 + K9: tree AND layout on the device, nothing comes back but 296 bytes
                tr.wait_frames(1)
                tr.animate_shapes(i / 60.0, 1.0 / 60.0)
                tr.build_bvh(15, rt.RT_BUILD_DEVICE_LAYOUT, fetch=False)
                tr.render_async(params, bufs[i & 1])
                return
            sc.animate(i / 60.0, 1.0 / 60.0, mode)
            shp = Hl.rh_shapes(sc._h, C.byref(n)); ns = n.value
            nod = Hl.rh_nodes(sc._h, C.byref(n)); nn = n.value
            if mode == 0:
                tr._chk(L.rt_update_shapes(tr._c, lo, hi - lo, C.c_void_p(shp + 192 * lo)))
                tr._chk(L.rt_update_nodes(tr._c, C.c_void_p(nod), nn))
            else:
                idx = Hl.rh_indices(sc._h, C.byref(n)); ni = n.value
                tr._chk(L.rt_upload_scene(tr._c, None, None, C.c_void_p(shp), ns, C.c_void_p(nod), nn, C.c_void_p(idx), ni))
            tr.wait_frames(1)                     # frame i-2 is complete: its buffer (this one) may be consumed / reused
            tr.render_async(params, bufs[i & 1])

        try:
            for i in range(10):
                frame(i)
            tr.synchronize()
            t0 = time.perf_counter()
            for i in range(10, 10 + frames):
                frame(i)
            tr.synchronize()
            dt = time.perf_counter() - t0
            out[label] = {"fps": frames / dt, "ms_per_frame": dt * 1e3 / frames}
        except Exception as e:                    # a diagnostic mode must not cost the bench line
            out[label] = {"error": repr(e)}
        tr.close()
    out["workload"] = "BASELINE.json configs[4]: monkey-1240 %dx%d, %d frames, CPU animation + BVH %s, pinned async upload, D2H of every frame" % (
        W, H, frames, "refit | rebuild; device_animation = rt_animate (shapes and node boxes move on the device, no upload); "
                      "device_rebuild = rt_animate + rt_build_bvh (K8: the reference tree rebuilt on the device every frame); "
                      "device_rebuild_k9 = rt_animate_shapes + rt_build_bvh(RT_BUILD_DEVICE_LAYOUT) (K8 + K9: tree and layout on the device)")
    return out


def time_device_build(torch, rt, stream):
    """K8 (rt_build_bvh): the reference's buildBVH + serializeBVH on the device against the host builder
    (host/rt_host.cpp restating src/main.cpp:955-979, 1111-1193), same shapes: device time by CUDA events (best of 3),
    the call's wall time incl. the D2H of the tree, the host builder's wall time, and whether the bytes agree."""
    out = {}
    for name, n, depth in (("monkey", 0, 15), ("car", 0, 25), ("soup", 1000000, 40)):
        try:
            t0 = time.perf_counter()
            sc = rt.Scene(name, 800, 600, n_override=n)
            t_scene = time.perf_counter() - t0
            t0 = time.perf_counter()
            sc.animate(0.0, 0.0, 1)                                   # mode 1 = buildBVH + serializeBVH on the host
            t_host = time.perf_counter() - t0
            tr = rt.RayTracer(64, 48, device=torch.cuda.current_device(), flags=0)
            tr.set_stream(stream.cuda_stream)
            tr.upload_scene(sc.light, sc.camera, sc.shapes)
            dev, wall = [], []
            for _ in range(3):
                t0 = time.perf_counter()
                nn, ni = tr.build_bvh(depth, fetch=False)
                wall.append(time.perf_counter() - t0)
                dev.append(tr.last_build_ms())
            nodes, idx = tr.build_bvh(depth)
            same = nodes.tobytes() == sc.nodes.tobytes() and idx.tobytes() == sc.indices.tobytes()
            # K9: the layout on the device too (no host layout, no copy of the tree): build + first frame, wall clock
            k9 = None
            try:
                p64 = rt.make_params(64, 48)
                tr.upload_scene(sc.light, sc.camera, sc.shapes)          # a fresh scene state: nothing cached
                tr.synchronize()
                t0 = time.perf_counter()
                tr.build_bvh(depth, rt.RT_BUILD_DEVICE_LAYOUT, fetch=False)
                tr.render_device(p64); tr.synchronize()
                k9 = {"build_plus_first_frame_wall_ms": (time.perf_counter() - t0) * 1e3, "device_ms": tr.last_build_ms(),
                      "layout_on_device": tr.layout_info()["on_device"]}
                tr.upload_scene(sc.light, sc.camera, sc.shapes)
                tr.synchronize()
                t0 = time.perf_counter()
                tr.build_bvh(depth, fetch=False)
                tr.render_device(p64); tr.synchronize()
                k9["host_layout_build_plus_first_frame_wall_ms"] = (time.perf_counter() - t0) * 1e3
            except Exception as e:
                k9 = {"error": repr(e)}
            tr.close()
            out["%s_%d" % (name, len(sc.shapes))] = {"shapes": len(sc.shapes), "nodes": int(nn), "maxDepth": depth,
                "device_ms": min(dev), "call_wall_ms": min(wall) * 1e3, "host_builder_ms": t_host * 1e3,
                "bytes_equal_host_builder": bool(same), "algorithmic_bytes": 192 * len(sc.shapes) + 48 * int(nn) + 4 * int(ni),
                "device_layout_k9": k9}
        except Exception as e:
            out[name] = {"error": repr(e)}
    return out


def time_k3_k5(torch, rt, stream, local, W, H, flush, steps=20, warmup=5):
    """K3 (rt_untile: 8 tile-major shards -> one row-major frame) and K5 (rt_present: float4 -> RGBA8) are pure streaming
    kernels: device time with L2 flushed, algorithmic GB/s (32 and 20 bytes per pixel) against the measured HBM copy peak."""
    peak = None
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs")
    except (OSError, ValueError):
        pass
    tr = rt.RayTracer(W, H, device=local)
    tr.set_stream(stream.cuda_stream)
    world = 8
    tr.set_shard(0, world)
    px = tr.shard_pixels()
    shards = torch.rand((world * px, 4), dtype=torch.float32, device="cuda")
    full = torch.empty((H, W, 4), dtype=torch.float32, device="cuda")
    img8 = torch.empty((H, W, 4), dtype=torch.uint8, device="cuda")
    out = {"workload": "%dx%d frame, 8 shards" % (W, H), "hbm_peak_GBps": peak or 6650.0,
           "hbm_peak_source": "MEASURED_PEAKS.json hbm_gbs" if peak else "fallback (B200_PROFILING.md)"}
    for name, fn, nbytes in (("k3_untile", lambda: tr.untile(shards.data_ptr(), world, full.data_ptr()), 32 * W * H),
                             ("k5_present", lambda: tr.present(full.data_ptr(), img8.data_ptr(), W * H), 20 * W * H)):
        evs = []
        for i in range(warmup + steps):
            if flush is not None:
                flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            if i >= warmup:
                evs.append((e0, e1))
        torch.cuda.synchronize()
        ms = sum(x.elapsed_time(y) for x, y in evs) / len(evs)
        out[name] = {"ms": ms, "algorithmic_bytes": nbytes, "GBps": nbytes / ms / 1e6, "frac_of_hbm_peak": nbytes / ms / 1e6 / (peak or 6650.0)}
    tr.close()
    return out


def soup_extra(torch, dist, rt, stream, local, rank, world, W, H, flush, steps, warmup):
    """BASELINE.json configs[3]: the 1M-triangle soup + 10k spheres (deep BVH, traversal-bound), tile-sharded over the ranks
    like the headline frame.  The scene is uploaded ONCE; an end-to-end step updates the camera, renders and brings the frame
    to the host.  Rays are counted by the kernel's own counter (no oracle constant exists at this size); parity is
    checked on a row sample against the oracle."""
    out = {"workload": "BASELINE.json configs[3]: soup of 1,000,000 triangles + 10,000 spheres, %dx%d, %d GPU(s), maxBounces=3" % (W, H, world)}
    t0 = time.perf_counter()
    sc = rt.Scene("soup", W, H, n_override=1000000)
    params = rt.make_params(W, H)
    out["scene_build_s"] = time.perf_counter() - t0
    tr = rt.RayTracer(W, H, device=local, flags=0)
    tr.set_stream(stream.cuda_stream)
    tr.set_shard(rank, world)
    t0 = time.perf_counter()
    tr.upload(sc); tr.synchronize()
    out["first_upload_s (290 MB H2D + host layout build)"] = time.perf_counter() - t0
    st = tr.render_stats(params)
    cnt = torch.tensor([st["rays"], st["node_visits"], st["prim_tests"]], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(cnt)
    rays = int(cnt[0])
    fused = False
    if world > 1:
        _, fused = setup_fused(torch, dist, tr, rank, world)
    tot, ren, gat, frame = time_frames(torch, dist, rt, tr, params, W, H, rank, world, steps, warmup, flush, fused=fused)
    t = torch.tensor([tot, ren, gat], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    tot, ren, gat = t.tolist()
    # end to end: camera update + render (+ gather) + the frame on the host
    host = torch.empty((H, W, 4), dtype=torch.float32, pin_memory=True) if rank == 0 else None
    token = torch.zeros(1, dtype=torch.int32, device="cuda")

    def step():
        tr.update_camera(sc.camera)
        if world == 1:
            tr.render(params, want_ids=False, out=host.numpy().reshape(-1, 4))
        elif fused == "signalled":
            tr.render_device(params)
            if rank == 0:
                tr.frame_download(host.data_ptr())      # acquire (completion words) + copy + release
            torch.cuda.synchronize()
        else:
            tr.render_device(params)
            dist.all_reduce(token)                      # every rank's tiles are in rank 0's frame
            if rank == 0:
                tr.frame_download(host.data_ptr())
            torch.cuda.synchronize()
            dist.all_reduce(token)                      # rank 0 has read the frame: the next step may overwrite it
    for _ in range(2):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = max(3, steps // 2)
    for _ in range(n):
        step()
    torch.cuda.synchronize()
    e2e = torch.tensor([(time.perf_counter() - t0) * 1e3 / n], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e, op=dist.ReduceOp.MAX)
    e2e = float(e2e[0])
    out.update({"ms_per_frame": tot, "render_ms": ren, "gather_untile_ms": gat, "fps": 1e3 / tot, "rays_counted_by_kernel": rays,
                "Mrays_per_s": rays / tot / 1e3, "node_visits": int(cnt[1]), "prim_tests": int(cnt[2]),
                "e2e_ms": e2e, "e2e_fps": 1e3 / e2e, "e2e_Mrays_per_s": rays / e2e / 1e3,
                "e2e_h2d_bytes_per_step": 80, "e2e_d2h_bytes_per_step": int(W * H * 16)})
    if rank == 0:       # parity on a row sample (the oracle needs ~1 us per node visit: the whole frame would take minutes)
        try:
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import oracle_py as O
            O.build()
            rs, ph = max(1, H // 24), max(1, H // 48)
            o_rgba, _, _, _ = O.render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, params, W, H, row_step=rs, row_phase=ph, want_counters=False)
            rows = np.arange(ph, H, rs)
            got = host.numpy()[rows]
            d = np.abs(got.astype(np.float64) - o_rgba[rows].astype(np.float64))
            out["parity"] = {"rows_checked": int(len(rows)), "pixels": int(len(rows) * W), "frac_within_1e-3": float((d <= 1e-3).all(axis=-1).mean()),
                             "max_abs": float(d.max()), "tie_pixels": O.primary_ties(sc.camera, sc.shapes, sc.nodes, sc.indices, params, W, H, row_step=rs, row_phase=ph)}
        except Exception as e:
            out["parity"] = {"error": repr(e)}
    tr.close()
    return out


def fast_build_block(torch, dist, rt, stream, local, sc, params, W, H, flush, steps, warmup, parity_ms):
    """SURVEY par. 7's "fast build": librt_b200_fast.so = the same source compiled with FMA contraction allowed.  Its speed next
    to the parity build's, and how far its frame is from the parity build's (ids and colours) — reported, never the default."""
    out = {"library": "librt_b200_fast.so (-fmad=true; everything else as the parity build)"}
    Lf = rt.lib_fast()
    out["build_info"] = Lf.rt_build_info().decode()
    a = rt.RayTracer(W, H, device=local)                      # parity build, with ids
    b = rt.RayTracer(W, H, device=local, library=Lf)
    for t in (a, b):
        t.set_stream(stream.cuda_stream); t.upload(sc)
    ra, ia = a.render(params); rb, ib = b.render(params)
    d = np.abs(ra.astype(np.float64) - rb.astype(np.float64))
    d[np.isnan(d)] = np.inf
    out.update({"ids_mismatch_vs_parity_build": int((ia != ib).sum()), "pixels": int(W * H),
                "frac_bit_identical_vs_parity_build": float((ra.view(np.uint32) == rb.view(np.uint32)).all(axis=-1).mean()),
                "frac_within_1e-3_vs_parity_build": float((d <= 1e-3).all(axis=-1).mean()),
                "max_abs_vs_parity_build": float(d[np.isfinite(d)].max())})
    a.close()
    f = rt.RayTracer(W, H, device=local, flags=0, library=Lf)
    f.set_stream(stream.cuda_stream); f.upload(sc)
    tot, ren, _, _ = time_frames(torch, dist, rt, f, params, W, H, 0, 1, steps, warmup, flush)
    f.close(); b.close()
    out.update({"ms_per_frame": tot, "render_ms": ren, "parity_build_ms_per_frame": parity_ms, "speedup_vs_parity_build": parity_ms / tot})
    return out


def main():
    args = parse()
    if args.impl == "reference":
        run_reference_arm(args)
        return
    import torch
    import torch.distributed as dist
    import rtload
    rt = rtload.load()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback for the hot path")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W, H = args.width, args.height
    sc = rt.Scene(args.scene, W, H)
    params = rt.make_params(W, H, maxBounces=args.bounces, useFresnel=bool(args.fresnel), useMollerTrumbore=bool(args.mt))
    consts = workload_constants(args.scene, W, H, args.mt, args.fresnel, args.bounces)

    stream = torch.cuda.Stream()
    out = {}
    with torch.cuda.stream(stream):
        tr = rt.RayTracer(W, H, device=local, flags=0)
        tr.set_stream(stream.cuda_stream)
        tr.set_shard(rank, world)
        tr.upload(sc)
        # the CUDA path's own ray count for this rank's tiles (untimed counting launch)
        mine = tr.render_stats(params)
        rays_t = torch.tensor([mine["rays"], mine["node_visits"], mine["prim_tests"]], dtype=torch.int64, device="cuda")
        if world > 1:
            dist.all_reduce(rays_t)
        rays_gpu, nodes_gpu, tests_gpu = (int(x) for x in rays_t.tolist())
        rays = consts["rays"] if consts else rays_gpu
        rays_match = (consts is None) or (rays_gpu == consts["rays"])

        flush = None if args.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        clocks = ClockSampler(local)
        launches0 = tr.launch_count()
        if rank == 0:
            clocks.start()
        fused = False
        frame_ptr = None
        if world > 1 and args.gather != "nccl":
            frame_ptr, fused = setup_fused(torch, dist, tr, rank, world, signalled=args.gather == "fused")
        tot, ren, gat, last_frame = time_frames(torch, dist, rt, tr, params, W, H, rank, world, args.steps, args.warmup, flush, fused=fused)
        launches = (tr.launch_count() - launches0) * args.steps // (args.steps + args.warmup)
        nccl_ms, barrier_ms = None, None
        if fused:                               # the other exchange paths beside it (north_star: gather timed separately)
            if fused == "signalled":            # the same peer stores with an all-reduce as completion barrier
                tr.sync_enable(0)
                dist.barrier()
                t2 = torch.tensor(list(time_frames(torch, dist, rt, tr, params, W, H, rank, world, args.steps, args.warmup, flush, fused="nccl_barrier")[:3]),
                                  dtype=torch.float64, device="cuda")
                dist.all_reduce(t2, op=dist.ReduceOp.MAX)
                barrier_ms = dict(zip(["total_ms", "render_ms", "gather_untile_ms"], t2.tolist()))
                torch.cuda.synchronize(); dist.barrier()
                tr.sync_enable(world)
                torch.cuda.synchronize(); dist.barrier()
            t2 = torch.tensor(list(time_frames(torch, dist, rt, tr, params, W, H, rank, world, args.steps, args.warmup, flush)[:3]),
                              dtype=torch.float64, device="cuda")
            dist.all_reduce(t2, op=dist.ReduceOp.MAX)
            nccl_ms = dict(zip(["total_ms", "render_ms", "gather_untile_ms"], t2.tolist()))
        clk = clocks.stop() if rank == 0 else None
        frame_ok = None
        if fused:                               # the gathered frame must equal a single-GPU render bit for bit
            tr.render_device(params)
            if fused != "signalled":
                dist.all_reduce(torch.zeros(1, dtype=torch.int32, device="cuda"))
            torch.cuda.synchronize()
            got = None
            if rank == 0:
                got = torch.empty((H, W, 4), dtype=torch.float32, pin_memory=True)
                tr.frame_download(got.data_ptr()); tr.synchronize()      # signalled: acquire + copy + release
                solo = rt.RayTracer(W, H, device=local, flags=0)
                solo.upload(sc)
                want, _ = solo.render(params, want_ids=False)
                solo.close()
                frame_ok = bool(np.array_equal(got.numpy().view(np.uint32), want.view(np.uint32)))
            dist.barrier()
        t = torch.tensor([tot, ren, gat], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        tot, ren, gat = t.tolist()
        e2e_copy, h2d, d2h = time_e2e(torch, dist, rt, tr, sc, params, W, H, rank, world, max(3, args.steps // 3), 3,
                                      frame_ptr=frame_ptr, fused=fused)
        e2e_rgba8 = None
        if world == 1:                          # the displayed frame (compute dispatch + display quad), RGBA8 to the host
            e2e_rgba8, _, d2h8 = time_e2e(torch, dist, rt, tr, sc, params, W, H, rank, world, max(3, args.steps // 3), 3, rgba8=True)
        # the same step with the frame target in mapped host memory: the kernels' own stores are the D2H
        hostf, alias = setup_host_frame(dist, tr, rank, world, W, H)
        tr.set_frame_target(alias)                               # (leaves signalled mode: this target is one host frame)
        e2e_host, _, _ = time_e2e(torch, dist, rt, tr, sc, params, W, H, rank, world, max(3, args.steps // 3), 3, host_alias=alias)
        host_ok = None
        if rank == 0:                           # the host frame must equal a single-GPU rt_render bit for bit
            solo = rt.RayTracer(W, H, device=local, flags=0)
            solo.upload(sc)
            want, _ = solo.render(params, want_ids=False)
            solo.close()
            host_ok = bool(np.array_equal(np.asarray(hostf).view(np.uint32).reshape(-1), want.view(np.uint32).reshape(-1)))
        tr.set_frame_target(None)
        tr.host_frame_unregister(hostf)
        t = torch.tensor([e2e_copy, e2e_host], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_copy, e2e_host = t.tolist()
        e2e_ms = min(e2e_copy, e2e_host)
        e2e_path = "host_frame" if e2e_host <= e2e_copy else "device_frame_copy"

        # a step whose upload CHANGES the topology (two trees of the same shapes, alternating): the host-side layout build is in
        cold_ms = None
        if world == 1 and args.scene != "soup":
            try:
                alt = [sc, rt.Scene(args.scene, W, H, bvh_depth=12)]
                host = torch.empty((H * W, 4), dtype=torch.float32, pin_memory=True).numpy()
                for i in range(2):
                    tr.upload(alt[i & 1]); tr.render(params, want_ids=False, out=host)
                t0 = time.perf_counter()
                for i in range(6):
                    tr.upload(alt[i & 1]); tr.render(params, want_ids=False, out=host)
                cold_ms = (time.perf_counter() - t0) * 1e3 / 6
                tr.upload(sc)
            except Exception as e:
                cold_ms = repr(e)

        # parity of the timed frame (rank 0): the frame the timed loop produces against the oracle
        parity = None
        if rank == 0 and not args.no_parity and args.scene != "soup":
            try:
                if world == 1:
                    chk = rt.RayTracer(W, H, device=local)          # RT_FLAG_HIT_IDS
                    chk.upload(sc)
                    chk.render(params)                               # the second frame runs in the heavy-first order, as the timed ones
                    f_rgba, f_ids = chk.render(params)
                    chk.close()
                else:
                    f_rgba, f_ids = (got.numpy() if fused else last_frame.cpu().numpy()), None
                parity = parity_block(rt, sc, params, W, H, f_rgba, f_ids)
            except Exception as e:
                parity = {"error": repr(e)}

        extra = {}
        if world == 1 and not args.no_extra and (W, H) != (800, 600):
            # BASELINE.json configs[1]: the same scene at the reference's own 800x600
            sc2 = rt.Scene(args.scene, 800, 600)
            p2 = rt.make_params(800, 600, maxBounces=args.bounces, useFresnel=bool(args.fresnel), useMollerTrumbore=bool(args.mt))
            tr2 = rt.RayTracer(800, 600, device=local, flags=0)
            tr2.set_stream(stream.cuda_stream)
            tr2.upload(sc2)
            c2 = workload_constants(args.scene, 800, 600, args.mt, args.fresnel, args.bounces)
            r2 = tr2.render_stats(p2)["rays"]
            tot2, ren2, _, _ = time_frames(torch, dist, rt, tr2, p2, 800, 600, 0, 1, args.steps, args.warmup, flush)
            e2, _, _ = time_e2e(torch, dist, rt, tr2, sc2, p2, 800, 600, 0, 1, max(3, args.steps // 3), 3)
            e28, _, _ = time_e2e(torch, dist, rt, tr2, sc2, p2, 800, 600, 0, 1, max(3, args.steps // 3), 3, rgba8=True)
            extra["car_800x600"] = {"workload": "BASELINE.json configs[1]", "Mrays_per_s": (c2["rays"] if c2 else r2) / ren2 / 1e3,
                                    "ms_per_frame": tot2, "fps": 1e3 / tot2, "e2e_ms": e2, "e2e_fps": 1e3 / e2, "e2e_rgba8_ms": e28,
                                    "rays": r2, "rays_match_oracle": (c2 is None) or r2 == c2["rays"]}
            if c2:
                extra["car_800x600"]["alg_GBps"] = c2["alg_bytes"] / ren2 / 1e6
            tr2.close()
            extra["monkey_animated_300f"] = time_animated(torch, rt, stream)
            extra["device_bvh_build"] = time_device_build(torch, rt, stream)
            try:
                extra["fast_build"] = fast_build_block(torch, dist, rt, stream, local, sc, params, W, H, flush, args.steps, args.warmup, tot)
            except Exception as e:
                extra["fast_build"] = {"error": repr(e)}
            try:
                extra["k3_k5_hbm"] = {"3840x2160": time_k3_k5(torch, rt, stream, local, 3840, 2160, flush),
                                      "7680x4320": time_k3_k5(torch, rt, stream, local, 7680, 4320, flush)}
            except Exception as e:
                extra["k3_k5_hbm"] = {"error": repr(e)}
            try:
                extra["soup_1M_3840x2160"] = soup_extra(torch, dist, rt, stream, local, rank, world, 3840, 2160, flush, 6, 3)
            except Exception as e:
                extra["soup_1M_3840x2160"] = {"error": repr(e)}
        if world == 8 and not args.no_extra and args.scene == "car":
            try:        # every rank takes part; rank 0 keeps the result
                extra["soup_1M_7680x4320"] = soup_extra(torch, dist, rt, stream, local, rank, world, 7680, 4320, flush, 6, 3)
            except Exception as e:
                extra["soup_1M_7680x4320"] = {"error": repr(e)}

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        peak, peak_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)") if "hbm_gbs" in peaks else (6650.0, "fallback (B200_PROFILING.md)")
        alg = (consts["alg_bytes"] / world) if consts else None        # per launch: one rank's share of the frame
        prof = {}
        try:        # the ncu capture of THIS kernel on THIS workload (one rank's launch when world > 1)
            prof = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(
                "%s_%dx%d_w%d" % (args.scene, W, H, world) if not (args.mt or args.fresnel or args.bounces != 3) else "-", {})
        except (OSError, ValueError):
            pass
        sm_clk = (clk or {}).get("sm_mhz") or 1965.0
        issue_peak = 148 * 4 * sm_clk * 1e6                      # warp instructions / s: one per SM sub-partition per cycle
        winst = prof.get("warp_instructions")
        line = {
            "metric": "Mrays/s (%s scene)" % args.scene, "value": rays / tot / 1e3, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": tot, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(args),
            "fps": 1e3 / tot, "render_ms": ren, "gather_untile_ms": gat,
            "gather": ({"signalled": "fused + signalled: peer stores from the render kernel, its last block publishes a completion word on rank 0, "
                                     "rank 0's stream waits on the words (stream memory operation) and releases the buffer: no collective per frame",
                        "nccl_barrier": "fused: peer stores from the render kernel + 1 all-reduce barrier"}.get(fused, "nccl gather + untile")) if world > 1 else None,
            "barrier_path": barrier_ms, "nccl_gather_path": nccl_ms, "gathered_frame_equals_single_gpu": frame_ok,
            "rays_per_frame": rays, "rays_counted_by_kernel": rays_gpu, "rays_match_oracle": rays_match,
            "kernel_work": {"node_visits": nodes_gpu, "prim_tests": tests_gpu,
                            "reference_node_visits": consts["node_visits"] if consts else None,
                            "reference_prim_tests": sum(consts["tests"]) if consts else None},
            "parity": parity,
            "e2e": {"value": rays / e2e_ms / 1e3, "unit": "Mrays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms, "fps": 1e3 / e2e_ms, "path": e2e_path,
                    "paths_ms": {"host_frame: kernels store into mapped (shared) host memory, every GPU over its own PCIe link": e2e_host,
                                 "device_frame_copy: render (+ gather) on the device, D2H copy of the frame from rank 0": e2e_copy},
                    "host_frame_equals_single_gpu": host_ok,
                    "displayed_frame_rgba8": None if e2e_rgba8 is None else {
                        "what": "rt_render_rgba8: the same step, but the frame leaves as the reference SHOWS it (display quad: alpha 1, clamp, 8 bits; "
                                "shader.frag:8-12) — K5 behind K1 band by band, 4 bytes per pixel over PCIe.  Not the headline: value / ms_per_step above are the rgba32f frame",
                        "ms_per_step": e2e_rgba8, "fps": 1e3 / e2e_rgba8, "d2h_bytes_per_step": d2h8},
                    "note": "every step re-uploads the whole scene from host buffers; the topology is unchanged from step to step, so the "
                            "library keeps its threaded layout and only refits (K4) + repacks (K2).  A step whose upload CHANGES the "
                            "topology also pays the host-side layout build: cold_topology_ms",
                    "cold_topology_ms": cold_ms},
            "gpu_launches": int(launches),
            "clocks": clk,
            "roofline": {"bound": "issue", "achieved": (winst / ren / 1e6) if winst else None, "peak": issue_peak / 1e9,
                         "unit": "Gwarp-inst/s", "frac": (winst / (ren * 1e-3) / issue_peak) if winst else None,
                         "traffic": prof.get("dram_bytes"), "kernel": "rt::render_kernel", "kernel_ms": ren,
                         "peak_source": "4 warp schedulers x 148 SMs x %.0f MHz (SM clock sampled during the timed region)" % sm_clk,
                         "warp_instructions_per_launch": winst, "ncu_issue_active_pct": prof.get("issue_active_pct"),
                         "ncu_duration_ms": prof.get("ncu_duration_ms"), "active_lanes_per_instruction": prof.get("active_lanes_per_instruction"),
                         "memory": memory_levels(prof, ren, peak) if prof else None,
                         "hbm_frac": (prof["dram_bytes"] / ren / 1e6 / peak) if prof else None,
                         "alg_normalised": {"what": "SURVEY 8d: algorithmic bytes of the REFERENCE's unculled traversal (oracle-counted) per launch / "
                                                    "kernel time / HBM peak.  Not a fraction of anything physical: the scene is cache-resident and "
                                                    "the walk skips most of that traffic (DRAM sees the image write: traffic)",
                                            "alg_bytes_per_launch": alg, "GBps": (alg / ren / 1e6) if alg else None,
                                            "over_hbm_peak": (alg / ren / 1e6 / peak) if alg else None, "hbm_peak_GBps": peak, "hbm_peak_source": peak_src},
                         "note": "instruction-issue bound: the kernel issues on %.0f %% of the scheduler cycles in the ncu capture; frac = the same "
                                 "ratio from the LIVE kernel time" % (prof.get("issue_active_pct") or 0)},
            "extra": extra,
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"], _ = cpu_baseline(rt, sc, params, W, H, args.cpu_seconds, scene_name=args.scene)
        print(json.dumps(line))
    tr.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
