"""CPU: the host-side layout builder (rt_layout.h: threaded BVH, conservative sub-trees in fat leaves,
rank tie-break) walked by a scalar emulation of the kernel's query (tests/emu/walk_emu.cpp, same
primitive/box test source as the kernel, compiled by g++ -ffp-contract=off) against the oracle's
intersectScene2, ray by ray: shape ids and closestDist must be bit-identical."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from util import random_shapes

HERE = os.path.dirname(os.path.abspath(__file__))
EMU_SRC = os.path.join(HERE, "emu", "walk_emu.cpp")
EMU_SO = os.path.join(HERE, "emu", "libwalk_emu.so")
CSRC = os.path.join(os.path.dirname(HERE), "opengl-ray-tracer_b200", "csrc")


@pytest.fixture(scope="module")
def emu():
    deps = [EMU_SRC, os.path.join(CSRC, "rt_layout.h"), os.path.join(CSRC, "rt_device.cuh")]
    if not os.path.exists(EMU_SO) or any(os.path.getmtime(d) > os.path.getmtime(EMU_SO) for d in deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared",
                               "-I/usr/local/cuda/include", "-o", EMU_SO, EMU_SRC])
    L = C.CDLL(EMU_SO)
    vp, ip = C.c_void_p, C.c_int
    L.emu_create.argtypes = [vp, ip, vp, ip, vp, ip, ip, C.c_double]; L.emu_create.restype = vp
    L.emu_destroy.argtypes = [vp]; L.emu_destroy.restype = None
    L.emu_info.argtypes = [vp, vp]; L.emu_info.restype = None
    L.emu_refit.argtypes = [vp, vp, vp]; L.emu_refit.restype = ip
    L.emu_refit_graph_check.argtypes = [vp, vp, vp]; L.emu_refit_graph_check.restype = ip
    L.emu_query.argtypes = [vp, ip, vp, vp, ip, ip, vp, vp, vp, vp, vp]; L.emu_query.restype = None
    L.emu_layout_hash.argtypes = [vp]; L.emu_layout_hash.restype = C.c_uint64
    L.emu_subboxes.argtypes = [vp, vp, ip]; L.emu_subboxes.restype = ip
    L.emu_scan_nodes.argtypes = [vp, vp, ip, vp]; L.emu_scan_nodes.restype = ip
    return L


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Emu:
    def __init__(self, L, shapes, nodes, indices, subtrees=True, omax=None, top=True):
        self.L = L
        self.keep = [np.ascontiguousarray(a) for a in (shapes, nodes, indices)]
        sh, nd, ix = self.keep
        if omax is None:
            omax = 4.0 * max(1.0, float(np.abs(nd["boundsMin"][-1]).max()), float(np.abs(nd["boundsMax"][-1]).max())) if len(nd) else 4.0
        self.omax = omax
        self.h = L.emu_create(_p(sh), len(sh), _p(nd), len(nd), _p(ix), len(ix), int(subtrees) | (2 if (top and subtrees) else 0), float(omax))
        assert self.h

    def info(self):
        out = (C.c_int * 6)()
        self.L.emu_info(self.h, out)
        return dict(nodes=out[0], sub=out[1], prims=out[2], fat=out[3], max_leaf=out[4], nodes_b=out[5])

    def query(self, o, d, mt=False, bounds=None):
        o = np.ascontiguousarray(o, np.float32); d = np.ascontiguousarray(d, np.float32)
        n = len(o)
        ids = np.empty(n, np.int32); dist = np.empty(n, np.float32); pts = np.empty((n, 3), np.float32)
        cnt = (C.c_uint64 * 2)()
        b = None if bounds is None else np.ascontiguousarray(bounds, np.float32)
        self.L.emu_query(self.h, n, _p(o), _p(d), int(mt), int(bounds is not None), _p(b), _p(ids), _p(dist), _p(pts), cnt)
        return ids, dist, pts, (int(cnt[0]), int(cnt[1]))

    def refit(self, shapes, nodes):
        self.keep[0] = np.ascontiguousarray(shapes); self.keep[1] = np.ascontiguousarray(nodes)
        return self.L.emu_refit(self.h, _p(self.keep[0]), _p(self.keep[1]))

    def refit_graph_check(self):
        return self.L.emu_refit_graph_check(self.h, _p(self.keep[0]), _p(self.keep[1]))

    def close(self):
        self.L.emu_destroy(self.h)


def normalize(v):
    v = v.astype(np.float32)
    return (v * (np.float32(1.0) / np.sqrt((v * v).sum(axis=1, keepdims=True, dtype=np.float32)))).astype(np.float32)


def make_rays(rng, sc_shapes, nodes, cam_pos, n):
    """Camera rays, rays between random points of the scene box, and adversarial rays aimed at triangle
    edges/vertices and sphere silhouettes (where a wrong conservative bound would show)."""
    lo, hi = nodes["boundsMin"][-1].astype(np.float64), nodes["boundsMax"][-1].astype(np.float64)
    lo, hi = np.where(np.isfinite(lo), lo, -50), np.where(np.isfinite(hi), hi, 50)
    O, D = [], []
    # (1) from the camera towards random points of the box
    tgt = rng.uniform(lo, hi, (n, 3))
    O.append(np.repeat(cam_pos[None], n, 0)); D.append(tgt - cam_pos)
    # (2) between random points (secondary-ray like)
    a = rng.uniform(lo, hi, (n, 3)); b = rng.uniform(lo, hi, (n, 3))
    O.append(a); D.append(b - a)
    # (3) adversarial: aim at points on triangle edges / vertices / just outside, and sphere limbs
    tri = sc_shapes[sc_shapes["type"] == 3]
    if len(tri):
        k = rng.integers(0, len(tri), n)
        P1, P2, P3 = tri["triP1"][k].astype(np.float64), tri["triP2"][k].astype(np.float64), tri["triP3"][k].astype(np.float64)
        u = rng.uniform(0, 1, (n, 1)); mode = rng.integers(0, 4, (n, 1))
        edge = np.where(mode == 0, P1 + u * (P2 - P1), np.where(mode == 1, P2 + u * (P3 - P2), np.where(mode == 2, P3 + u * (P1 - P3), P1)))
        edge = edge + rng.normal(0, 1, (n, 3)) * rng.choice([0.0, 1e-6, 1e-4, 1e-2], (n, 1))
        src = np.where(rng.uniform(0, 1, (n, 1)) < 0.5, np.repeat(cam_pos[None], n, 0), rng.uniform(lo, hi, (n, 3)))
        O.append(src); D.append(edge - src)
    sph = sc_shapes[sc_shapes["type"] == 0]
    if len(sph):
        k = rng.integers(0, len(sph), n)
        c, r = sph["sphereCenter"][k].astype(np.float64), sph["sphereRadius"][k].astype(np.float64)[:, None]
        src = np.where(rng.uniform(0, 1, (n, 1)) < 0.5, np.repeat(cam_pos[None], n, 0), rng.uniform(lo, hi, (n, 3)))
        to = c - src
        perp = np.cross(to, rng.normal(0, 1, (n, 3)))
        perp /= np.linalg.norm(perp, axis=1, keepdims=True) + 1e-30
        limb = c + perp * r * (1 + rng.choice([0.0, 1e-6, -1e-6, 1e-4, -1e-4, 1e-2], (n, 1)))
        O.append(src); D.append(limb - src)
    O = np.concatenate(O).astype(np.float32); D = np.concatenate(D)
    ok = np.linalg.norm(D, axis=1) > 1e-9
    return O[ok], normalize(D[ok])


def check_scene(emu, oracle, sc, rng, n=20000, label=""):
    shapes, nodes, idx = sc.shapes, sc.nodes, sc.indices
    cam = sc.camera["Position"][0].astype(np.float64)
    e = Emu(emu, shapes, nodes, idx, subtrees=True)
    o, d = make_rays(rng, shapes, nodes, cam, n)
    out = {}
    for mt in (False, True):
        rid, rdist, rpt = oracle.intersect_scene_batch(shapes, nodes, idx, o, d, mt)
        gid, gdist, gpt, cnt = e.query(o, d, mt)
        bad = np.nonzero((gid != rid) | (gdist.view(np.uint32) != rdist.view(np.uint32)))[0]
        assert len(bad) == 0, "%s mt=%s: %d of %d rays differ, first %d: emu (%d, %r) oracle (%d, %r)" % (
            label, mt, len(bad), len(o), bad[0], gid[bad[0]], gdist[bad[0]], rid[bad[0]], rdist[bad[0]])
        hitm = rid >= 0
        assert np.array_equal(gpt[hitm].view(np.uint32), rpt[hitm].view(np.uint32))
        # any-hit (shadow) queries: the boolean "closest INNER hit is nearer than the bound"
        bounds = (rdist * rng.choice([0.5, 0.999999, 1.0, 1.000001, 2.0], len(o))).astype(np.float32)
        bounds = np.where(hitm, bounds, rng.uniform(1, 100, len(o)).astype(np.float32))
        aid, _, _, _ = e.query(o, d, mt, bounds=np.minimum(bounds, np.float32(1e20)))
        assert np.array_equal(aid >= 0, rdist < bounds), "%s mt=%s any-hit differs" % (label, mt)
        out[mt] = cnt
    info = e.info()
    e.close()
    return info, out, len(o)


@pytest.mark.parametrize("name", ["toy", "triangle", "monkey", "car"])
def test_named_scenes(rt, oracle, emu, name):
    rng = np.random.default_rng(11)
    sc = rt.Scene(name, 800, 600)
    info, cnt, n = check_scene(emu, oracle, sc, rng, label=name)
    if name == "car":
        assert info["fat"] > 0 and info["max_leaf"] <= 8


@pytest.mark.parametrize("seed", range(6))
def test_random_scenes_fat_leaves(rt, oracle, emu, seed):
    """Shallow reference BVHs (depth 1-3) put hundreds of mixed shapes into one leaf."""
    rng = np.random.default_rng(100 + seed)
    base = rt.Scene("toy", 160, 120)
    shapes = random_shapes(rt, rng, 300, types=(0, 2, 3) if seed % 2 else (0, 1, 2, 3), extent=10.0 * (1 + 10 * (seed % 3)))
    sc = rt.Scene.from_shapes(shapes, base.light, base.camera, bvh_depth=1 + seed % 3, width=160, height=120)
    info, _, _ = check_scene(emu, oracle, sc, rng, n=8000, label="random %d" % seed)
    assert info["sub"] > 0


def test_degenerate_and_stale_planes(rt, oracle, emu):
    """Slivers, zero-area triangles, NaN vertices and triangles whose stored plane no longer matches their
    vertices (upstream's animated wheels, src/main.cpp:1101-1106) must not be culled wrongly."""
    rng = np.random.default_rng(5)
    base = rt.Scene("toy", 160, 120)
    shapes = random_shapes(rt, rng, 200, types=(3,), extent=8.0)
    shapes["triP3"][0:10] = shapes["triP2"][0:10]                                     # zero area
    shapes["triP3"][10:20] = shapes["triP1"][10:20] + (shapes["triP2"][10:20] - shapes["triP1"][10:20]) * np.float32(1.5)   # collinear
    shapes["triP3"][20:30] += np.float32(1e-6)
    shapes["triP2"][30:33] = np.nan
    # stale planes: rotate the vertices of 60 triangles about z, keep planeNormal/planeD
    for i in range(40, 100):
        a = rng.uniform(0, 2 * np.pi)
        R = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]], np.float32)
        for f in ("triP1", "triP2", "triP3"):
            shapes[f][i] = R @ shapes[f][i]
    sc = rt.Scene.from_shapes(shapes, base.light, base.camera, bvh_depth=2, width=160, height=120)
    check_scene(emu, oracle, sc, rng, n=20000, label="degenerate")


def test_far_origins_ignore_subtrees(rt, oracle, emu):
    """Origins beyond omax: the margins do not hold, the walk must treat sub-tree boxes as hit."""
    rng = np.random.default_rng(9)
    sc = rt.Scene("car", 800, 600)
    shapes, nodes, idx = sc.shapes, sc.nodes, sc.indices
    e = Emu(emu, shapes, nodes, idx)
    n = 4000
    tgt = rng.uniform(-10, 10, (n, 3))
    src = normalize(rng.normal(0, 1, (n, 3))).astype(np.float64) * rng.choice([1e3, 1e4, 1e5, 1e6], (n, 1))
    d = normalize(tgt - src)
    rid, rdist, _ = oracle.intersect_scene_batch(shapes, nodes, idx, src.astype(np.float32), d)
    gid, gdist, _, _ = e.query(src.astype(np.float32), d)
    assert np.array_equal(gid, rid) and np.array_equal(gdist.view(np.uint32), rdist.view(np.uint32))
    e.close()


def test_refit_after_animation(rt, oracle, emu):
    """Wheels rotate (stale planes), boxes grow: refit of the sub-tree boxes keeps every decision."""
    rng = np.random.default_rng(3)
    sc = rt.Scene("car", 800, 600)
    e = Emu(emu, sc.shapes, sc.nodes, sc.indices)
    cam = sc.camera["Position"][0].astype(np.float64)
    for frame in range(1, 4):
        sc.animate(frame * 0.37, 0.37, mode=0)
        shapes, nodes = sc.shapes, sc.nodes
        assert e.refit(shapes, nodes) == 1
        o, d = make_rays(rng, shapes, nodes, cam, 6000)
        for mt in (False, True):
            rid, rdist, _ = oracle.intersect_scene_batch(shapes, nodes, sc.indices, o, d, mt)
            gid, gdist, _, _ = e.query(o, d, mt)
            assert np.array_equal(gid, rid) and np.array_equal(gdist.view(np.uint32), rdist.view(np.uint32)), "frame %d mt=%s" % (frame, mt)
    e.close()


@pytest.mark.parametrize("name", ["car", "monkey", "toy"])
def test_refit_graph_reproduces_host_refit(rt, emu, name):
    """The dependency graph the device refit kernel executes (rtl::build_refit_graph), replayed on the host with
    the kernel's arrival rule, gives bit for bit the boxes of the host refit, before and after the scene moved."""
    sc = rt.Scene(name, 800, 600)
    e = Emu(emu, sc.shapes, sc.nodes, sc.indices)
    if e.info()["sub"] == 0 and e.info()["nodes_b"] == 0:
        e.close(); pytest.skip("no boxes of ours in this scene")
    assert e.refit(sc.shapes, sc.nodes) == 1
    assert e.refit_graph_check() == 0
    for frame in range(1, 3):
        sc.animate(frame * 0.41, 0.41, mode=0)
        assert e.refit(sc.shapes, sc.nodes) == 1
        assert e.refit_graph_check() == 0
    e.close()


def test_subtrees_cut_the_work(rt, oracle, emu):
    """The point of the sub-trees: far fewer primitive tests on the car scene, same answers."""
    rng = np.random.default_rng(1)
    sc = rt.Scene("car", 800, 600)
    cam = sc.camera["Position"][0].astype(np.float64)
    o, d = make_rays(rng, sc.shapes, sc.nodes, cam, 5000)
    a = Emu(emu, sc.shapes, sc.nodes, sc.indices, subtrees=False)
    b = Emu(emu, sc.shapes, sc.nodes, sc.indices, subtrees=True)
    ia, da, _, ca = a.query(o, d)
    ib, db, _, cb = b.query(o, d)
    assert np.array_equal(ia, ib) and np.array_equal(da.view(np.uint32), db.view(np.uint32))
    assert cb[1] < ca[1] / 3, (ca, cb)
    a.close(); b.close()


def test_list_b_used_and_equivalent(rt, oracle, emu):
    """List B (SAH tree over the reference leaves) is chosen on the car scene and changes no answer."""
    rng = np.random.default_rng(21)
    sc = rt.Scene("car", 800, 600)
    cam = sc.camera["Position"][0].astype(np.float64)
    o, d = make_rays(rng, sc.shapes, sc.nodes, cam, 8000)
    a = Emu(emu, sc.shapes, sc.nodes, sc.indices, top=False)
    b = Emu(emu, sc.shapes, sc.nodes, sc.indices, top=True)
    assert a.info()["nodes_b"] == 0 and b.info()["nodes_b"] > 0
    for mt in (False, True):
        ia, da, _, ca = a.query(o, d, mt)
        ib, db, _, cb = b.query(o, d, mt)
        assert np.array_equal(ia, ib) and np.array_equal(da.view(np.uint32), db.view(np.uint32))
        assert cb[0] < ca[0]
    a.close(); b.close()


def test_axis_parallel_and_degenerate_rays(rt, oracle, emu):
    """Directions with exact zero components (1/d = inf: NaN slab arithmetic, GLSL select semantics) and
    non-finite origins must take the reference tree (list A) and still match the oracle bit for bit."""
    rng = np.random.default_rng(4)
    for name in ("car", "monkey", "toy"):
        sc = rt.Scene(name, 800, 600)
        shapes, nodes, idx = sc.shapes, sc.nodes, sc.indices
        e = Emu(emu, shapes, nodes, idx)
        n = 6000
        lo, hi = nodes["boundsMin"][-1].astype(np.float64), nodes["boundsMax"][-1].astype(np.float64)
        lo, hi = np.where(np.isfinite(lo), lo, -50), np.where(np.isfinite(hi), hi, 50)
        o = rng.uniform(lo - 5, hi + 5, (n, 3)).astype(np.float32)
        # snap some origins onto box planes of random nodes: (bound - o) == 0 against 1/d == inf gives NaN
        pick = rng.integers(0, len(nodes), n)
        ax = rng.integers(0, 3, n)
        snap = rng.uniform(0, 1, n) < 0.5
        o[snap, ax[snap]] = nodes["boundsMin"][pick[snap], ax[snap]]
        d = rng.normal(0, 1, (n, 3))
        zero = rng.integers(0, 3, n)
        d[np.arange(n), zero] = 0.0
        two = rng.uniform(0, 1, n) < 0.3
        d[two, (zero[two] + 1) % 3] = 0.0
        d = normalize(d)
        o[:5] = np.nan; o[5:10, 0] = np.inf
        for mt in (False, True):
            rid, rdist, _ = oracle.intersect_scene_batch(shapes, nodes, idx, o, d, mt)
            gid, gdist, _, _ = e.query(o, d, mt)
            assert np.array_equal(gid, rid), "%s mt=%s: %d ids differ" % (name, mt, int((gid != rid).sum()))
            assert np.array_equal(gdist.view(np.uint32), rdist.view(np.uint32))
        e.close()


@pytest.mark.parametrize("name,mt", [("car", False), ("car", True), ("monkey", False)])
def test_grazing_rays_at_the_inflated_box_faces(rt, oracle, emu, name, mt):
    """Adversarial check of the conservative margins (rtl::prim_bound: 64 E S, 128 E ... / sqrt(sin2) / |cos|): rays aimed AT
    the faces, edges and corners of every sub-tree box — tangent to a face, or crossing it at a grazing angle a few ulp
    inside / outside — from nearby, from far away and from the edge of the origin range the margins were computed for
    (|o| = omax).  A ray that misses such a box must not be accepted by anything below it: ids and distances must equal the
    walk WITHOUT sub-trees (the reference's own leaves) and the oracle."""
    rng = np.random.default_rng(77)
    sc = rt.Scene(name, 800, 600)
    shapes, nodes, idx = sc.shapes, sc.nodes, sc.indices
    a = Emu(emu, shapes, nodes, idx, subtrees=False)
    b = Emu(emu, shapes, nodes, idx, subtrees=True)
    nsub = b.info()["sub"]
    if nsub == 0:
        a.close(); b.close(); pytest.skip("no sub-trees in this scene")
    boxes = np.zeros((nsub, 6), np.float32)
    assert emu.emu_subboxes(b.h, _p(boxes), nsub) == nsub
    pick = boxes[rng.permutation(nsub)[:600]].astype(np.float64)
    O, D = [], []
    omax = float(b.omax)
    for bx in pick:
        lo, hi = bx[:3], bx[3:]
        for _ in range(6):
            ax = int(rng.integers(0, 3)); side = int(rng.integers(0, 2))
            # a point of the face: interior, on an edge or at a corner
            u = rng.choice([0.0, 1.0, rng.uniform()], 3)
            p = lo + u * (hi - lo)
            p[ax] = hi[ax] if side else lo[ax]
            # direction: in the face plane, plus a grazing component through the face of 0, +-1e-7, +-1e-5, +-1e-3
            t = rng.normal(0, 1, 3); t[ax] = 0.0
            t /= np.linalg.norm(t) + 1e-30
            t[ax] = rng.choice([0.0, 1e-7, -1e-7, 1e-5, -1e-5, 1e-3, -1e-3])
            # nudge the target a few fp32 ulp in / out of the box along the face normal
            q = p.copy()
            q[ax] = np.nextafter(np.float32(q[ax]), np.float32(np.inf if rng.uniform() < 0.5 else -np.inf), dtype=np.float32) if rng.uniform() < 0.7 else q[ax]
            for dist in (0.5, 30.0, None):
                if dist is None:                                   # origin on the boundary of the margins' origin range
                    s = np.abs(t).max()
                    far = (0.999 * omax - np.abs(q).max()) / max(s, 1e-9)
                    o = q - t * far
                    o = np.clip(o, -0.999 * omax, 0.999 * omax)
                else:
                    o = q - t * dist
                O.append(o); D.append(q - o)
    o = np.asarray(O, np.float32); d = normalize(np.asarray(D, np.float64))
    ok = np.isfinite(d).all(axis=1)
    o, d = o[ok], d[ok]
    ia, da, _, _ = a.query(o, d, mt)
    ib, db, _, _ = b.query(o, d, mt)
    assert np.array_equal(ia, ib), "%d of %d grazing rays change their hit under the sub-tree boxes" % (int((ia != ib).sum()), len(o))
    assert np.array_equal(da.view(np.uint32), db.view(np.uint32))
    rid, rdist, _ = oracle.intersect_scene_batch(shapes, nodes, idx, o, d, mt)
    assert np.array_equal(ib, rid) and np.array_equal(db.view(np.uint32), rdist.view(np.uint32))
    assert (ib >= 0).mean() > 0.05, "the rays do hit the geometry under the boxes"
    a.close(); b.close()


# FNV-1a of every layout array (lists A / B, slots, boxes, refit graph), omax = 200, sub-trees and list B on.  These are
# the layouts the render kernel was validated with on the B200; the values were produced by the builder as it was then
# (commit 79331da) and by the record-based / multi-threaded SAH builder that replaced it — any change to rt_layout.h
# that moves a single word shows up here.  soup 80 000 (> 65 536 reference leaves) takes the multi-threaded path.
LAYOUT_HASHES = [("toy", 0, 0x0a6055fbc85e3700), ("triangle", 0, 0xcd9ed5b4917cac12), ("monkey", 0, 0x49a569ec8332fdfe),
                 ("car", 0, 0xac2e7573db02538d), ("soup", 80000, 0xaee99b7a581aefff)]


@pytest.mark.parametrize("name,n,want", LAYOUT_HASHES)
def test_layout_is_the_validated_one(rt, emu, name, n, want):
    sc = rt.Scene(name, 800, 600, n_override=n)
    e = Emu(emu, sc.shapes, sc.nodes, sc.indices, omax=200.0)
    assert emu.emu_layout_hash(e.h) == want, "%s: layout hash %016x" % (name, emu.emu_layout_hash(e.h))


@pytest.mark.parametrize("name,n,depth,want", [("soup", 140000, 1, 0x11f969cb14bdb73c), ("soup", 300000, 2, 0x5fafe17f02e94c90)])
def test_threaded_fat_leaf_subtrees_are_the_validated_ones(rt, emu, name, n, depth, want):
    """Fat leaves of 65 536 shapes and more (a soup under a tree of depth 1 / 2: leaves of 70 k - 100 k shapes) take
    SubBuilder::build_par, the forked build with leaves of several shapes; the soup of 80 000 above takes build_fixed (list B's
    tree, one item per leaf).  Both give the layout the sequential builder gave (hashes recorded before the builders were
    threaded this way; node ids may differ, the tree may not)."""
    sc = rt.Scene(name, 800, 600, n_override=n, bvh_depth=depth)
    e = Emu(emu, sc.shapes, sc.nodes, sc.indices, omax=200.0)
    assert emu.emu_layout_hash(e.h) == want, "%s %d depth %d: layout hash %016x" % (name, n, depth, emu.emu_layout_hash(e.h))


def test_scan_nodes_equals_the_three_passes(emu, rt):
    """rtl::scan_nodes (one pass over an uploaded node array: topology compare with the shadow, copy, safe coordinates,
    nestedness) against the three separate passes it replaced, on real trees and on every kind of damage."""
    rng = np.random.default_rng(5)
    cases = 0
    for name in ("toy", "monkey", "car"):
        sc = rt.Scene(name, 64, 48)
        base = np.ascontiguousarray(sc.nodes).copy()
        n = len(base)
        variants = [("same", base.copy(), base)]
        for kind in ("grown", "shrunk", "nan", "inf", "huge", "child", "range", "leafidx", "inneridx", "empty_leaf", "no_old"):
            for _ in range(4):
                v = base.copy()
                i = int(rng.integers(0, n))
                inner = np.flatnonzero(base["leftChild"] != -1); leaves = np.flatnonzero(base["leftChild"] == -1)
                if kind == "grown": v["boundsMax"][inner[rng.integers(0, len(inner))]] += 1.0
                elif kind == "shrunk": v["boundsMax"][inner[rng.integers(0, len(inner))]] -= 0.5
                elif kind == "nan": v["boundsMin"][i, rng.integers(0, 3)] = np.nan
                elif kind == "inf": v["boundsMax"][i, rng.integers(0, 3)] = np.inf
                elif kind == "huge": v["boundsMax"][n - 1, rng.integers(0, 3)] = 2e37
                elif kind == "child" and len(inner) > 1:
                    a, b = inner[rng.integers(0, len(inner), 2)]
                    v["leftChild"][a] = base["leftChild"][b]
                elif kind == "range": v["rightChild"][inner[rng.integers(0, len(inner))]] = n + 3
                elif kind == "leafidx": v["numShapes"][leaves[rng.integers(0, len(leaves))]] += 1
                elif kind == "inneridx": v["startShapeIdx"][inner[rng.integers(0, len(inner))]] += 7      # ignored for inner nodes
                elif kind == "empty_leaf":
                    lf = leaves[rng.integers(0, len(leaves))]
                    v["boundsMin"][lf] = np.inf; v["boundsMax"][lf] = -np.inf
                variants.append((kind, v, None if kind == "no_old" else base))
        for kind, v, old in variants:
            out = np.zeros(6, np.int32)
            copied = emu.emu_scan_nodes(_p(v), _p(old), n, _p(out))
            assert copied == 1, (name, kind)
            assert out[:3].tolist() == out[3:].tolist(), (name, kind, out.tolist())
            cases += 1
    out = np.zeros(6, np.int32)
    assert emu.emu_scan_nodes(None, None, 0, _p(out)) == 1 and out[1] == 1 and out[2] == 1     # no nodes: safe, nested, nothing to compare
    assert cases > 100
