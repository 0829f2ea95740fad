"""The C++ host program of examples/render_frame.cpp (what a reference maintainer's main.cpp does through
the C ABI): builds against include/rt_api.h, fails loudly without a GPU, matches the oracle with one."""
import os
import subprocess

import numpy as np
import pytest

from util import assert_parity, compare

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "examples", "render_frame.cpp")
EXE = os.path.join(ROOT, "examples", "render_frame")


def build(rt):
    rt.build.build_all()
    csrc, host = rt.build.CSRC, rt.build.HOST
    deps = [SRC, os.path.join(ROOT, "include", "rt_api.h")]
    if not os.path.exists(EXE) or any(os.path.getmtime(d) > os.path.getmtime(EXE) for d in deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-I" + os.path.join(ROOT, "include"), SRC, "-o", EXE,
                               os.path.join(csrc, "librt_b200.so"), os.path.join(host, "librt_host.so"),
                               "-Wl,-rpath," + csrc, "-Wl,-rpath," + host])
    return EXE


def test_example_builds_and_has_no_cpu_fallback(rt, tmp_path):
    import torch
    exe = build(rt)
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run([exe, "toy", "64", "48", str(tmp_path / "x.pfm")], capture_output=True, text=True)
    assert r.returncode != 0 and "rt_create" in r.stderr and not (tmp_path / "x.pfm").exists()


@pytest.mark.gpu
def test_example_matches_oracle(rt, oracle, tmp_path):
    exe = build(rt)
    W, H = 320, 240
    out = tmp_path / "car.pfm"
    r = subprocess.run([exe, "car", str(W), str(H), str(out), "3"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    raw = open(out, "rb").read()
    hdr_end = 0
    for _ in range(3):
        hdr_end = raw.index(b"\n", hdr_end) + 1
    img = np.frombuffer(raw[hdr_end:], dtype="<f4").reshape(H, W, 3)[::-1]
    sc = rt.Scene("car", W, H)
    for f in (1, 2):
        sc.animate(f / 60.0, 1.0 / 60.0, 0)
    p = rt.make_params(W, H)
    o_rgba, o_ids, _, _ = oracle.render(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices, p, W, H, want_counters=False)
    rgba = np.concatenate([img, np.ones((H, W, 1), np.float32)], axis=-1)
    assert_parity(compare(rgba, None, o_rgba, o_ids), "C++ example, 3 animated frames")
    ppm = open(str(out) + ".ppm", "rb").read()              # the displayed frame: the display quad's arithmetic on that image
    assert ppm.startswith(b"P6\n%d %d\n255\n" % (W, H))
    shown = np.frombuffer(ppm[len(b"P6\n%d %d\n255\n" % (W, H)):], dtype=np.uint8).reshape(H, W, 3)
    assert np.array_equal(shown, oracle.present(np.ascontiguousarray(rgba))[..., :3])
