#!/usr/bin/env python3
"""Generate the committed golden fixtures.  Run in the build container, where
/root/reference is mounted (the GPU box has no reference tree):

    python tests/golden/make_golden.py

Sources of truth:
  * ref_*  — the REFERENCE SHADER FILE compiled as C++ (oracle/_ref/libref_shader.so,
             built by oracle/build_ref.py from /root/reference/src/shaders/gpu_shader.comp);
  * kat_*  — the REFERENCE'S OWN C++ shape classes (oracle/_ref/libref_shapes.so);
  * ora_*  — the oracle restatement (ids / dists / counters the reference cannot emit).
Scene bytes come from the deterministic headless host library.
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import oracle_py as O  # noqa: E402
import rtload  # noqa: E402

rt = rtload.load()
W, H = 96, 72
CASES = [(s, mt, fr) for s in ("toy", "triangle", "monkey", "car") for mt in (0, 1) for fr in (0, 1)]


def scene_images():
    assert O.build_ref() is not None, "needs /root/reference"
    for name, mt, fr in CASES:
        sc = rt.Scene(name, W, H)
        p = rt.make_params(W, H, useMollerTrumbore=bool(mt), useFresnel=bool(fr))
        ref = O.ref_render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, p, W, H, threads=1)
        rgba, ids, dists, cnt = O.render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, p, W, H, threads=1)
        keys = ["rays", "node_visits", "index_reads", "tri_stage2", "ray_hits", "alg_bytes"]
        np.savez_compressed(os.path.join(HERE, "img_%s_mt%d_fr%d.npz" % (name, mt, fr)),
                            ref_rgb=ref[..., :3].astype(np.float32), ora_rgb=rgba[..., :3], ora_ids=ids.astype(np.int16),
                            ora_dists=dists, counters=np.array([cnt[k] for k in keys] + cnt["tests"], dtype=np.uint64),
                            scene_crc=np.array([scene_crc(sc)], dtype=np.uint64))
        print(name, mt, fr, "max|ref-oracle| = %.3g" % np.abs(ref[..., :3] - rgba[..., :3]).max())
    # brute-force twin
    for name in ("toy", "monkey"):
        sc = rt.Scene(name, W, H)
        p = rt.make_params(W, H, useBVH=False, useFresnel=True)
        ref = O.ref_render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, p, W, H, threads=1)
        np.savez_compressed(os.path.join(HERE, "img_%s_nobvh.npz" % name), ref_rgb=ref[..., :3].astype(np.float32))


def scene_crc(sc):
    import zlib
    crc = 0
    for a in (sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices):
        crc = zlib.crc32(np.ascontiguousarray(a).tobytes(), crc)
    return crc


def shape_kats():
    """Random rays against the reference's Sphere / Wall / Triangle classes."""
    R = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_shapes.so"))
    fp = C.POINTER(C.c_float)
    R.refshape_intersect.argtypes = [C.c_int, fp, fp, fp, C.c_float, C.c_float, fp, fp, fp, fp]
    R.refshape_intersect.restype = C.c_int
    rng = np.random.default_rng(2024)
    n = 3000
    rec = np.zeros(n, dtype=[("type", "i4"), ("a", "f4", 3), ("b", "f4", 3), ("c", "f4", 3), ("s0", "f4"), ("s1", "f4"),
                             ("start", "f4", 3), ("dir", "f4", 3), ("res", "i4"), ("hit", "f4", 3), ("nd", "f4", 4)])
    for i in range(n):
        t = int(rng.choice([0, 2, 3]))
        r = rec[i]
        r["type"] = t
        r["a"] = rng.uniform(-5, 5, 3)
        r["b"] = rng.uniform(-5, 5, 3) if t == 3 else rng.normal(size=3)
        r["c"] = rng.uniform(-5, 5, 3)
        if t == 2 and i % 7 == 0:
            r["b"] = (0, 1, 0)                       # the NaN tangent-basis wall (scene 1 floor)
        r["s0"] = rng.uniform(0.5, 4) if t == 0 else (rng.uniform(2, 9) if t == 2 else float(i % 2))
        r["s1"] = rng.uniform(2, 9)
        r["start"] = rng.uniform(-12, 12, 3)
        tgt = rng.uniform(-4, 4, 3)
        d = (tgt - r["start"]).astype(np.float32)
        r["dir"] = d / np.float32(np.sqrt((d * d).sum()))
        hit = np.zeros(3, np.float32)
        nd = np.zeros(4, np.float32)
        arr = lambda x: np.ascontiguousarray(x, np.float32).ctypes.data_as(fp)  # noqa: E731
        a, b, c_, st, di = (np.array(r[k], np.float32) for k in ("a", "b", "c", "start", "dir"))
        r["res"] = R.refshape_intersect(t, arr(a), arr(b), arr(c_), float(r["s0"]), float(r["s1"]), arr(st), arr(di),
                                        hit.ctypes.data_as(fp), nd.ctypes.data_as(fp))
        r["hit"], r["nd"] = hit, nd
    np.savez_compressed(os.path.join(HERE, "kat_shapes.npz"), rec=rec)
    print("kat_shapes: INNER %d OUTER %d NONE %d" % ((rec["res"] == 1).sum(), (rec["res"] == 2).sum(), (rec["res"] == 0).sum()))


if __name__ == "__main__":
    scene_images()
    shape_kats()
