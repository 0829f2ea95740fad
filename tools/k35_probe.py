#!/usr/bin/env python3
"""K3 (rt_untile) and K5 (rt_present) against the HBM roofline: device time (CUDA events on the context's stream, L2 flushed
between launches, mean of 20) and algorithmic GB/s = (bytes read + bytes written) / time.

    python tools/k35_probe.py [W H]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, rtload
rt = rtload.load()
sizes = [(int(sys.argv[1]), int(sys.argv[2]))] if len(sys.argv) > 2 else [(3840, 2160), (7680, 4320)]
peak = None
try:
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
except Exception:
    pass
out = {"hbm_peak": peak}
stream = torch.cuda.Stream()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
with torch.cuda.stream(stream):
    for W, H in sizes:
        tr = rt.RayTracer(W, H)
        tr.set_stream(stream.cuda_stream)
        for world in (1, 8):
            tr.set_shard(0, world)
            px = tr.shard_pixels()
            shards = torch.rand((world * px, 4), dtype=torch.float32, device="cuda")
            full = torch.empty((H, W, 4), dtype=torch.float32, device="cuda")
            img8 = torch.empty((H, W, 4), dtype=torch.uint8, device="cuda")
            res = {}
            for name, fn, nbytes in (("untile", lambda: tr.untile(shards.data_ptr(), world, full.data_ptr()), 32 * W * H),
                                     ("present", lambda: tr.present(full.data_ptr(), img8.data_ptr(), W * H), 20 * W * H)):
                evs = []
                for i in range(25):
                    flush.zero_()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(); fn(); e1.record()
                    if i >= 5:
                        evs.append((e0, e1))
                torch.cuda.synchronize()
                ms = sum(a.elapsed_time(b) for a, b in evs) / len(evs)
                res[name] = {"ms": round(ms, 4), "GBps": round(nbytes / ms / 1e6, 1)}
            out["%dx%d_w%d" % (W, H, world)] = res
        tr.close()
print(json.dumps(out))
