"""GPU: rt_build_bvh_static (include/rt_api.h) — the K8 single-block build with its working tables in the library's
file-scope __device__ array, for callers that may not call cudaMalloc.  It has no counterpart in the reference;
what it must equal is the existing entry point: rt_build_bvh's FlatNode[] and
bvhIndices[] (themselves byte-identical to buildBVH + serializeBVH, src/main.cpp:955-979, :1111-1193, see
test_gpu_build.py), BYTE FOR BYTE — the tolerance of every builder test here is zero.

Called through the C ABI (ctypes) with torch tensors as the caller's device buffers, several calls in a row on the shared
scratch (different scenes and sizes, so that a call sees the previous call's leftovers).  The file sorts last on purpose:
it must pass alone in a fresh process and after every other test.
"""
import ctypes as C

import numpy as np
import pytest

from util import random_shapes

pytestmark = pytest.mark.gpu


def _existing(rt, shapes, depth):
    """The existing entry point: shapes only into a context, rt_build_bvh, rt_get_bvh."""
    tr = rt.RayTracer(64, 48)
    sc = rt.Scene("toy", 64, 48)
    tr.upload_scene(sc.light, sc.camera, shapes)
    nodes, idx = tr.build_bvh(depth)
    tr.close()
    return nodes, idx


def _static(rt, torch, shapes, depth, stream=None, check_free=False):
    n = len(shapes)
    raw = np.ascontiguousarray(shapes).view(np.uint8).reshape(-1)
    d_shapes = torch.from_numpy(raw.copy()).cuda() if n else None
    cap_nodes = max(1, 2 * n - 1)
    d_nodes = torch.full((cap_nodes * rt.NODE_DTYPE.itemsize,), 0xEE, dtype=torch.uint8, device="cuda")
    d_idx = torch.full((max(1, n),), -7, dtype=torch.int32, device="cuda")
    d_cnt = torch.full((2,), -1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    rt.build_bvh_static(d_shapes.data_ptr() if n else None, n, depth, d_nodes.data_ptr(), cap_nodes,
                        d_idx.data_ptr() if n else None, n, d_cnt.data_ptr(), stream)
    torch.cuda.synchronize()
    if check_free:                                             # cudaMemGetInfo: the call took no device memory (the first call of a
        assert torch.cuda.mem_get_info()[0] == free_before     # process may load the module lazily, so it is not checked)
    nn, ni = (int(v) for v in d_cnt.cpu().numpy())
    assert 1 <= nn <= cap_nodes and ni == n
    nodes = d_nodes.cpu().numpy()[: nn * rt.NODE_DTYPE.itemsize].view(rt.NODE_DTYPE)
    idx = d_idx.cpu().numpy()[:ni]
    return nodes, idx


def _scene_shapes(rt, name, **kw):
    return np.ascontiguousarray(rt.Scene(name, 64, 48, **kw).shapes)


def test_static_build_equals_rt_build_bvh_on_calls_in_a_row(rt):
    import torch
    rt.lib()                                                   # fails loudly without the CUDA library
    cases = [(_scene_shapes(rt, "monkey"), 15), (_scene_shapes(rt, "toy"), 15), (_scene_shapes(rt, "monkey"), 15),
             (_scene_shapes(rt, "monkey"), 2), (random_shapes(rt, np.random.default_rng(5), 2048), 40),
             (random_shapes(rt, np.random.default_rng(6), 1), 15), (random_shapes(rt, np.random.default_rng(7), 777), 0),
             (_scene_shapes(rt, "monkey"), 15)]
    assert any(len(s) > 1024 for s, _ in cases) and all(len(s) <= rt.RT_STATIC_BUILD_MAX_SHAPES for s, _ in cases)
    want = [_existing(rt, s, d) for s, d in cases]
    side = torch.cuda.Stream()
    for i, ((shapes, depth), (w_nodes, w_idx)) in enumerate(zip(cases, want)):
        stream = None if i % 2 == 0 else side.cuda_stream      # the caller's stream; the synchronize in _static orders the calls
        nodes, idx = _static(rt, torch, shapes, depth, stream, check_free=i > 0)
        assert len(nodes) == len(w_nodes), "call %d: %d nodes, rt_build_bvh built %d" % (i, len(nodes), len(w_nodes))
        assert nodes.tobytes() == w_nodes.tobytes(), "call %d: FlatNode[] differs from rt_build_bvh's" % i
        assert idx.tobytes() == w_idx.tobytes(), "call %d: bvhIndices[] differs from rt_build_bvh's" % i


def test_static_build_back_to_back_on_one_stream_without_a_sync_between(rt):
    """Two launches queued behind each other (stream order is the only ordering the entry point asks for)."""
    import torch
    a, b = _scene_shapes(rt, "monkey"), _scene_shapes(rt, "toy")
    wa, wb = _existing(rt, a, 15), _existing(rt, b, 15)
    bufs = []
    for s in (a, b, a):
        n = len(s)
        bufs.append((torch.from_numpy(np.ascontiguousarray(s).view(np.uint8).reshape(-1).copy()).cuda(),
                     torch.zeros(((2 * n - 1) * rt.NODE_DTYPE.itemsize,), dtype=torch.uint8, device="cuda"),
                     torch.zeros((n,), dtype=torch.int32, device="cuda"), torch.zeros((2,), dtype=torch.int32, device="cuda"), n))
    torch.cuda.synchronize()
    for sh, nd, ix, ct, n in bufs:
        rt.build_bvh_static(sh.data_ptr(), n, 15, nd.data_ptr(), 2 * n - 1, ix.data_ptr(), n, ct.data_ptr(), None)
    torch.cuda.synchronize()
    for (sh, nd, ix, ct, n), (w_nodes, w_idx) in zip(bufs, (wa, wb, wa)):
        nn = int(ct[0])
        assert nn == len(w_nodes) and int(ct[1]) == n
        assert nd.cpu().numpy()[: nn * rt.NODE_DTYPE.itemsize].tobytes() == w_nodes.tobytes()
        assert ix.cpu().numpy().tobytes() == w_idx.tobytes()


def test_static_build_empty_scene_is_one_empty_leaf(rt):
    """buildBVH over nothing: one empty leaf (main.cpp:1175-1193), as the host builder (host/rt_host.cpp) emits it."""
    import torch
    base = rt.Scene("toy", 64, 48)
    sc = rt.Scene.from_shapes(np.zeros(0, dtype=rt.SHAPE_DTYPE), base.light, base.camera, width=64, height=48)
    nodes, idx = _static(rt, torch, np.zeros(0, dtype=rt.SHAPE_DTYPE), 15)
    assert len(nodes) == 1 and nodes.tobytes() == np.ascontiguousarray(sc.nodes).tobytes() and len(idx) == 0


def test_static_build_rejects_what_does_not_fit(rt):
    import torch
    L = rt.lib()
    err = C.create_string_buffer(512)
    dummy = torch.zeros(64, dtype=torch.int32, device="cuda")
    p = dummy.data_ptr()
    big = rt.RT_STATIC_BUILD_MAX_SHAPES + 1
    assert L.rt_build_bvh_static(p, big, 15, p, 2 * big, p, big, p, None, err, len(err)) == -1
    assert b"2048" in err.value and b"RT_STATIC_BUILD_MAX_SHAPES" in err.value, err.value
    assert L.rt_build_bvh_static(p, 10, 15, p, 18, p, 10, p, None, err, len(err)) == -1 and b"19 nodes" in err.value, err.value
    assert L.rt_build_bvh_static(p, 10, 15, p, 19, p, 9, p, None, err, len(err)) == -1
    assert L.rt_build_bvh_static(p, 10, 65, p, 19, p, 10, p, None, err, len(err)) == -1 and b"maxDepth" in err.value
    assert L.rt_build_bvh_static(p, -1, 15, p, 19, p, 10, p, None, None, 0) == -1
    assert L.rt_build_bvh_static(None, 10, 15, p, 19, p, 10, p, None, err, len(err)) == -1 and b"NULL" in err.value
    with pytest.raises(RuntimeError, match="RT_STATIC_BUILD_MAX_SHAPES"):
        rt.build_bvh_static(p, big, 15, p, 2 * big, p, big, p)
    assert L.rt_build_bvh_static(None, 0, 0, None, 0, None, 0, None, None, err, len(err)) == 0 and err.value == b""   # set-up only
    torch.cuda.synchronize()
