#!/usr/bin/env python3
"""Timeline of the PRODUCT schedule (heavy-first order, split blocks, paired walks) of one rank's shard: needs a library built
with -DRT_TIMELINE=1 (diagnostic: the hit-id plane then carries, per part of an 8x4 item, the end time, the duration in SM
cycles, the SM, the warp slot and the split).

    RT_B200_LIB=build_var/tl.so python tools/timeline2.py [scene W H world rank]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, rtload
rt = rtload.load()
name = sys.argv[1] if len(sys.argv) > 1 else "car"
W = int(sys.argv[2]) if len(sys.argv) > 2 else 3840
H = int(sys.argv[3]) if len(sys.argv) > 3 else 2160
world = int(sys.argv[4]) if len(sys.argv) > 4 else 8
rank = int(sys.argv[5]) if len(sys.argv) > 5 else 0
MHZ = 1965.0
sc = rt.Scene(name, W, H)
tr = rt.RayTracer(W, H, flags=rt.RT_FLAG_HIT_IDS)
tr.set_shard(rank, world)
tr.upload(sc)
p = rt.make_params(W, H)
px = tr.shard_pixels()
img = torch.empty((px, 4), dtype=torch.float32, device="cuda")
ids = torch.empty((px,), dtype=torch.int32, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(6):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); tr.render_device(p, img.data_ptr(), ids.data_ptr()); e1.record()
    torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
a = ids.cpu().numpy()
if world == 1:
    t = a.reshape(H // 8, 8, W // 32, 32).transpose(0, 2, 1, 3).reshape(-1, 8, 32)
else:
    t = a.reshape(-1, 8, 32)
# item (tile, b): rows (b >> 2) * 4 .. + 4, columns (b & 3) * 8 .. + 8, lane = row * 8 + column
blk = t.reshape(-1, 2, 4, 4, 8).transpose(0, 1, 3, 2, 4).reshape(-1, 32)        # [item][lane]
rec = []
for it in range(len(blk)):
    lanes = blk[it]
    parts = (int(lanes[2]) >> 16) & 0xff
    if parts not in (1, 2, 4, 8):
        continue
    for q in range(parts):
        f = q * 32 // parts
        meta = int(lanes[f + 2])
        rec.append((it, q, parts, int(lanes[f]) & 0x7fffffff, int(lanes[f + 1]) * 4, meta & 0xff, (meta >> 8) & 0xff, int(lanes[f + 3])))
r = np.array(rec, dtype=np.int64)
end = r[:, 3]; dur = r[:, 4] * 1000.0 / MHZ                # ns
t0 = (end - dur).min(); end = end - t0; start = end - dur
print("%s %dx%d rank %d of %d: event time %.4f ms; %d items, %d parts; span of the parts %.4f ms" % (name, W, H, rank, world, ms, len(blk), len(r), (end.max() - start.min()) / 1e6))
print("part duration us: mean %.1f p50 %.1f p99 %.1f max %.1f ; sum of durations / (span x warps that ran) = %.2f" % (
    dur.mean() / 1e3, np.median(dur) / 1e3, np.percentile(dur, 99) / 1e3, dur.max() / 1e3,
    dur.sum() / ((end.max() - start.min()) * len(np.unique(r[:, 5] * 64 + r[:, 6])))))
print("split: " + ", ".join("%d parts: %d items" % (k, (r[r[:, 1] == 0][:, 2] == k).sum()) for k in (1, 2, 4, 8)))
print("the parts that end last:")
for i in np.argsort(-end)[:12]:
    print("  item %6d part %d/%d sm %3d slot %2d start %.1f us end %.1f us dur %.1f us, %d walk iterations, %.0f cycles each" % (r[i, 0], r[i, 1], r[i, 2], r[i, 5], r[i, 6], start[i] / 1e3, end[i] / 1e3, dur[i] / 1e3, r[i, 7], r[i, 4] / max(1, r[i, 7])))
print("the longest parts:")
for i in np.argsort(-dur)[:12]:
    print("  item %6d part %d/%d sm %3d slot %2d start %.1f us end %.1f us dur %.1f us, %d walk iterations, %.0f cycles each" % (r[i, 0], r[i, 1], r[i, 2], r[i, 5], r[i, 6], start[i] / 1e3, end[i] / 1e3, dur[i] / 1e3, r[i, 7], r[i, 4] / max(1, r[i, 7])))
print("cycles per walk iteration over all parts: %.0f (parts that started in the first 10 us: %.0f; in the last third: %.0f)" % (
    r[:, 4].sum() / r[:, 7].sum(), r[start < 1e4][:, 4].sum() / r[start < 1e4][:, 7].sum(), r[start > end.max() * 2 / 3][:, 4].sum() / max(1, r[start > end.max() * 2 / 3][:, 7].sum())))
se = np.sort(end)
print("parts ended: 50%% by %.1f us, 90%% by %.1f, 99%% by %.1f, 99.9%% by %.1f, all by %.1f" % tuple(se[min(len(se) - 1, int(len(se) * f))] / 1e3 for f in (.5, .9, .99, .999, 1.0)))
# running warps over time
edges = np.linspace(0, end.max(), 21)
act = [int(((start < b) & (end > a_)).sum()) for a_, b in zip(edges[:-1], edges[1:])]
print("parts in flight per 5%% slice of the span:", act)
tr.close()
