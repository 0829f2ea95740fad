"""CPU: memory- and thread-safety of the code that cannot be put under compute-sanitizer on the GPU pool (closed there).
The per-thread bodies of the device BVH builder (K8 / K9: the kernels' own index arithmetic, compiled by g++ through
tests/emu/build_emu.cpp) run under AddressSanitizer + UndefinedBehaviorSanitizer; the host layout builder (rt_layout.h, with
its multi-threaded SAH builder) runs under ThreadSanitizer and ASan and must still produce the pinned layouts."""
import os
import subprocess

import pytest

from test_layout_emu import LAYOUT_HASHES

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
HOST = os.path.join(ROOT, "opengl-ray-tracer_b200", "host")
COMMON = ["g++", "-std=c++17", "-O1", "-g", "-ffp-contract=off", "-pthread", "-I/usr/local/cuda/include"]
LINK = [os.path.join(HOST, "librt_host.so"), "-Wl,-rpath," + HOST]


def _build(rt, src, exe, flags):
    rt.build.build_host()
    exe = os.path.join(HERE, "emu", exe)
    r = subprocess.run(COMMON + flags + [os.path.join(HERE, "emu", src), "-o", exe] + LINK, capture_output=True, text=True)
    if r.returncode != 0:
        pytest.skip("this g++ cannot build with %s: %s" % (" ".join(flags), r.stderr[-300:]))
    return exe


def _run(cmd):
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=1:abort_on_error=0", TSAN_OPTIONS="halt_on_error=1")
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=900)
    if r.returncode != 0 and "unexpected memory mapping" in r.stderr:      # TSan vs. address-space randomisation of some kernels
        r = subprocess.run(["setarch", "-R"] + cmd, capture_output=True, text=True, env=env, timeout=900)
    return r


def test_device_builder_bodies_under_asan_ubsan(rt):
    exe = _build(rt, "sanitize_build_main.cpp", "sanitize_build_asan", ["-fsanitize=address,undefined", "-fno-sanitize-recover=undefined"])
    for args in (["toy", "0", "15"], ["monkey", "0", "15"], ["car", "0", "25"], ["soup", "8000", "40"], ["monkey", "0", "1"]):
        r = _run([exe] + args)
        assert r.returncode == 0, r.stderr[-3000:]
        assert r.stdout.count("identical 1") == 2 and "identical 0" not in r.stdout, r.stdout


@pytest.mark.parametrize("flags,tag", [(["-fsanitize=thread"], "tsan"), (["-fsanitize=address,undefined", "-fno-sanitize-recover=undefined"], "asan")])
def test_layout_builder_under_sanitizers(rt, flags, tag):
    exe = _build(rt, "sanitize_layout_main.cpp", "sanitize_layout_" + tag, flags)
    want = {(name, n): h for name, n, h in LAYOUT_HASHES}
    for name, n in (("car", 0), ("monkey", 0), ("soup", 80000)):           # soup 80 000: the multi-threaded path
        r = _run([exe, name, str(n)])
        assert r.returncode == 0, r.stderr[-3000:]
        assert "layout hash %016x" % want[(name, n)] in r.stdout, r.stdout
    # fat leaves of 70 000 shapes (a soup under a tree of depth 1): SubBuilder::build_par, the forked build with leaves of
    # several shapes (hash: tests/test_layout_emu.py::test_threaded_fat_leaf_subtrees_are_the_validated_ones)
    r = _run([exe, "soup", "140000", "1"])
    assert r.returncode == 0, r.stderr[-3000:]
    assert "layout hash 11f969cb14bdb73c" in r.stdout, r.stdout


def test_walk_emulation_under_asan_ubsan(rt):
    """The kernel's primitive / box tests (rt_device.cuh) walking the layouts of rt_layout.h (reference walk, sub-trees,
    list B): every node, link and slot access checked; the number of hits must not depend on the layout."""
    exe = _build(rt, "sanitize_walk_main.cpp", "sanitize_walk_asan", ["-fsanitize=address,undefined", "-fno-sanitize-recover=undefined"])
    for args in (["car", "0"], ["monkey", "0"], ["toy", "0"]):
        r = _run([exe] + args)
        assert r.returncode == 0, r.stderr[-3000:]
        hits = {}
        for line in r.stdout.splitlines():
            w = line.replace(":", "").replace(",", "").split()
            hits.setdefault((w[4], w[6]), set()).add(int(w[9]))            # (mt, any) -> hit counts over the three layouts
        assert len(hits) == 4 and all(len(v) == 1 for v in hits.values()), r.stdout
