"""CPU: the Wavefront OBJ reader of the host library (rh_obj_triangles) — the reference's model path (model.hpp:49-168
through assimp, mesh.hpp:163-189 mesh2triangles, main.cpp:645-675) restated without assimp — against a numpy restatement
of the same rules, and through the oracle as a scene."""
import numpy as np
import pytest

OBJ = """# a unit cube of quads (v/vt/vn and v//vn forms), then a second object with a triangle and a pentagon by negative indices
o cube
v 0 0 0
v 1 0 0
v 1 1 0
v 0 1 0
v 0 0 1
v 1 0 1
v 1 1 1
v 0 1 1
vn 0 0 -1
vt 0.5 0.5
f 1/1/1 2/1/1 3/1/1 4/1/1
f 5//1 8//1 7//1 6//1
f 1 5 6 2
usemtl other
f 2 6 7 3
f 3 7 8 4
f 4 8 5 1
o fan
v 3 0 0
v 4 0 0
v 4.5 1 0
v 3.5 2 0
v 2.5 1 0
f -5 -4 -3
f -5 -4 -3 -2 -1
"""


def _restate(origin, apply_rule):
    """The same rules in numpy: fans, one mesh per o / g / usemtl run with faces, Mesh::center over the face corners."""
    v = []
    meshes, cur = [], None
    for line in OBJ.splitlines():
        t = line.split()
        if not t:
            continue
        if t[0] == "v":
            v.append([float(x) for x in t[1:4]])
        elif t[0] == "f":
            idx = []
            for w in t[1:]:
                k = int(w.split("/")[0])
                idx.append(k - 1 if k > 0 else len(v) + k)
            if cur is None:
                cur = []
                meshes.append(cur)
            for i in range(1, len(idx) - 1):
                cur.append((v[idx[0]], v[idx[i]], v[idx[i + 1]]))
        elif t[0] in ("o", "g", "usemtl"):
            cur = None
    org = np.asarray(origin, np.float32)
    tris, normals, first = [], [], [0]
    for m in meshes:
        raw = np.asarray(m, np.float32)                                  # (T, 3, 3)
        pts = raw + org
        n = np.cross(pts[:, 1] - pts[:, 0], pts[:, 2] - pts[:, 0]).astype(np.float32)
        n = n / np.linalg.norm(n, axis=1, keepdims=True)
        if apply_rule:
            c = org.copy()
            for p in raw.reshape(-1, 3):
                c = (c + (org + p)).astype(np.float32)
            c = c / np.float32(len(raw) * 3)
            flip = (n * c).sum(axis=1) > 0
            n[flip] = -n[flip]
        tris.append(pts); normals.append(n)
        first.append(first[-1] + len(raw))
    return np.concatenate(tris), np.concatenate(normals), np.asarray(first, np.int32)


@pytest.mark.parametrize("apply_rule", [False, True])
def test_obj_reader(rt, tmp_path, apply_rule):
    path = tmp_path / "two.obj"
    path.write_text(OBJ)
    origin = (0.5, -2.0, -30.0)
    mat = np.zeros(1, rt.SHAPE_DTYPE)
    mat["color"] = (0.7, 0.65, 0.24); mat["fresnelStrength"] = 1.0; mat["ambientStrength"] = 0.2
    mat["diffuseStrength"] = 0.8; mat["specularStrength"] = 0.1; mat["shininess"] = 16
    tris, first = rt.obj_triangles(path, origin, mat, apply_rule)
    want, wn, wfirst = _restate(origin, apply_rule)
    assert len(tris) == 6 * 2 + 1 + 3 and first.tolist() == wfirst.tolist() == [0, 6, 12, 16]     # cube split by usemtl, the fan object
    assert (tris["type"] == 3).all()
    got = np.stack([tris["triP1"], tris["triP2"], tris["triP3"]], axis=1)
    assert np.array_equal(got, want)
    assert np.allclose(tris["planeNormal"], wn, atol=1e-6)
    assert np.allclose(tris["planeD"], -(wn * want[:, 0]).sum(axis=1), atol=1e-4)
    if apply_rule:                                                         # the rule flipped something, and not everything
        nat, _ = rt.obj_triangles(path, origin, mat, False)
        flipped = (np.sign(tris["planeNormal"]) != np.sign(nat["planeNormal"])).any(axis=1)
        assert 0 < flipped.sum() < len(tris)
    for k in ("color", "fresnelStrength", "ambientStrength", "diffuseStrength", "specularStrength", "shininess"):
        assert (tris[k] == mat[k]).all()
    # default material (material.hpp:23) without a template
    d, _ = rt.obj_triangles(path)
    assert np.allclose(d["color"], 1.0) and (d["shininess"] == 32).all() and np.allclose(d["specularStrength"], 0.5)


def test_obj_errors(rt, tmp_path):
    with pytest.raises(RuntimeError, match="cannot open"):
        rt.obj_triangles(tmp_path / "missing.obj")
    bad = tmp_path / "bad.obj"
    bad.write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 9\n")
    with pytest.raises(RuntimeError, match="malformed"):
        rt.obj_triangles(bad)
    empty = tmp_path / "empty.obj"
    empty.write_text("# nothing\n")
    t, first = rt.obj_triangles(empty)
    assert len(t) == 0 and first.tolist() == [0]


def test_obj_scene_through_the_oracle(rt, oracle, tmp_path):
    """The loaded mesh as a scene: reference builder over it (rh_scene_from_shapes), rendered by the oracle."""
    path = tmp_path / "two.obj"
    path.write_text(OBJ)
    tris, _ = rt.obj_triangles(path, (0.0, 0.0, -6.0))
    base = rt.Scene("toy", 64, 48)
    sc = rt.Scene.from_shapes(tris, base.light, base.camera, bvh_depth=8, width=64, height=48)
    assert len(sc.shapes) == len(tris) and len(sc.nodes) >= 1
    p = rt.make_params(64, 48)
    rgba, ids, _, _ = oracle.render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, p, 64, 48, want_counters=False)
    assert np.isfinite(rgba).all() and rgba.shape == (48, 64, 4)
