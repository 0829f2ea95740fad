"""Shared helpers for the parity tests."""
import numpy as np

# north_star tolerance: colour max-abs 1e-3 per channel on >= 99.9 % of pixels; ids bit-exact.
COLOR_TOL = 1e-3
COLOR_FRAC = 0.999


def compare(rgba, ids, o_rgba, o_ids, o_dists=None):
    """Returns a dict of parity statistics between a render and the oracle's."""
    rgba = np.asarray(rgba).reshape(o_rgba.shape)
    with np.errstate(invalid="ignore"):              # inf - inf: handled below
        diff = np.abs(rgba.astype(np.float64) - o_rgba.astype(np.float64))
    both_nan = np.isnan(rgba) & np.isnan(o_rgba)
    diff[both_nan] = 0.0
    diff[np.isinf(o_rgba) & (rgba == o_rgba)] = 0.0       # the same infinity on both sides (inf - inf is NaN)
    diff[np.isnan(diff)] = np.inf
    # THE BAR is north_star's: max-abs 1e-3 per channel on >= 99.9 % of pixels (frac_ok, asserted).  Two more counts are
    # reported beside it: a 10x stricter one (1e-4) and the round-1 looser one (1e-3 + 1e-4 relative; radiance is unclamped).
    ok = (diff <= COLOR_TOL).all(axis=-1)
    out = dict(frac_ok=float(ok.mean()), max_abs=float(diff[np.isfinite(diff)].max() if np.isfinite(diff).any() else 0.0),
               n_bad=int((~ok).sum()), frac_within_1e_4=float((diff <= 1e-4).all(axis=-1).mean()),
               frac_within_1e_3_plus_rel=float((diff <= COLOR_TOL + 1e-4 * np.abs(o_rgba)).all(axis=-1).mean()),
               pixels=int(ok.size))
    if ids is not None:
        ids = np.asarray(ids).reshape(o_ids.shape)
        out["id_mismatch"] = int((ids != o_ids).sum())
    return out


def assert_parity(stats, what=""):
    assert stats.get("id_mismatch", 0) == 0, "%s: %d primary-hit ids differ from the oracle" % (what, stats["id_mismatch"])
    assert stats["frac_ok"] >= COLOR_FRAC, "%s: only %.5f of pixels within max-abs 1e-3 (max abs %.3g)" % (what, stats["frac_ok"], stats["max_abs"])
    print("parity %s: ids differ %s, within 1e-3 %.6f, within 1e-4 %.6f, within 1e-3+1e-4|v| %.6f, max abs %.3g, tie pixels %s" %
          (what, stats.get("id_mismatch"), stats["frac_ok"], stats["frac_within_1e_4"], stats["frac_within_1e_3_plus_rel"],
           stats["max_abs"], stats.get("tie_pixels", "n/a")))


def random_shapes(rt, rng, n, types=(0, 2, 3), extent=10.0):
    """n random flat shapes of the given types around the origin (plane normals/D filled like the reference ctors)."""
    sh = np.zeros(n, dtype=rt.SHAPE_DTYPE)
    for i in range(n):
        t = int(rng.choice(types))
        s = sh[i]
        s["type"] = t
        s["color"] = rng.uniform(0, 1, 3)
        s["fresnelStrength"] = rng.uniform(0, 1)
        s["ambientStrength"] = rng.uniform(0, 0.5)
        s["diffuseStrength"] = rng.uniform(0, 1)
        s["specularStrength"] = rng.choice([0.0, 0.0, 0.3, 0.5, 1.0])
        s["shininess"] = int(rng.choice([1, 8, 32, 64]))
        c = rng.uniform(-extent, extent, 3).astype(np.float32)
        if t == 0:
            s["sphereCenter"] = c
            s["sphereRadius"] = rng.uniform(0.3, 2.5)
        else:
            if t == 3:
                p1 = c
                p2 = (c + rng.uniform(-3, 3, 3)).astype(np.float32)
                p3 = (c + rng.uniform(-3, 3, 3)).astype(np.float32)
                nrm = np.cross((p2 - p1).astype(np.float32), (p3 - p1).astype(np.float32)).astype(np.float32)
                s["triP1"], s["triP2"], s["triP3"] = p1, p2, p3
            else:
                nrm = rng.normal(size=3).astype(np.float32)
            nrm = (nrm * np.float32(1.0 / np.sqrt(np.float32((nrm * nrm).sum())))).astype(np.float32)
            if rng.uniform() < 0.5:
                nrm = -nrm
            s["planeNormal"] = nrm
            s["planeD"] = -np.float32((nrm * c).sum())
            if t == 2:
                s["wallStart"] = c
                s["wallWidth"] = rng.uniform(2, 12)
                s["wallHeight"] = rng.uniform(2, 12)
    return sh


FLOAT_FIELDS = ["sphereCenter", "sphereRadius", "planeNormal", "planeD", "wallStart", "wallWidth", "wallHeight", "triP1", "triP2", "triP3"]


def bit_fuzzed_shapes(rt, seed):
    """Shapes of all four types whose geometry fields carry arbitrary bit patterns in places (NaNs of both signs, infinities,
    denormals, signed zeros, FLT_MAX): 2 % or 30 % of the words, any pattern or a list of special ones.  Returns (shapes, depth)."""
    rng = np.random.default_rng(1000 + seed)
    n = int(rng.choice([3, 20, 200, 900]))
    sh = random_shapes(rt, rng, n, types=(0, 1, 2, 3))
    mode = seed % 4
    special = np.array([0x7fc00000, 0xffc00000, 0x7f800000, 0xff800000, 0x00000001, 0x80000001, 0, 0x80000000, 0x7f7fffff], dtype=np.uint32)
    for f in FLOAT_FIELDS:
        v = sh[f].copy().view(np.uint32)
        m = rng.uniform(size=v.shape) < (0.02 if mode < 2 else 0.3)
        if mode % 2 == 0:
            v[m] = rng.integers(0, 2 ** 32, size=int(m.sum()), dtype=np.uint64).astype(np.uint32)
        else:
            v[m] = rng.choice(special, size=int(m.sum()))
        sh[f] = v.view(np.float32)
    return sh, int(rng.integers(1, 20))
