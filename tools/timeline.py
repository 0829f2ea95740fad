#!/usr/bin/env python3
"""Timeline of the warp items of one counting launch: end time (globaltimer ns), duration (SM cycles), SM, warp slot."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, rtload
rt = rtload.load()
name = sys.argv[1] if len(sys.argv) > 1 else "car"
W = int(sys.argv[2]) if len(sys.argv) > 2 else 3840
H = int(sys.argv[3]) if len(sys.argv) > 3 else 2160
b = int(sys.argv[4]) if len(sys.argv) > 4 else 3
sc = rt.Scene(name, W, H)
tr = rt.RayTracer(W, H, flags=0)
tr.upload(sc)
aov = np.zeros(W * H, np.int32)
p = rt.make_params(W, H, maxBounces=b)
for _ in range(2):
    st = tr.render_stats(p, cost=aov)
a = aov.reshape(H, W)[:H // 4 * 4, :W // 8 * 8].reshape(H // 4, 4, W // 8, 8)
end = a[:, 0, :, 0].astype(np.int64).ravel(); dur = a[:, 0, :, 1].astype(np.int64).ravel() * 4
sm = a[:, 0, :, 2].ravel(); wid = a[:, 0, :, 3].ravel()
its = [a[:, 0, :, 4 + j].ravel() for j in range(4)]
t0 = end.min(); end = (end - t0) % (1 << 31)
dur_ns = dur / 1.965
start = end - dur_ns
print("items %d, kernel span %.3f ms, item duration ns: mean %.0f p50 %.0f p99 %.0f max %.0f" % (
    len(end), (end.max() - start.min()) / 1e6, dur_ns.mean(), np.median(dur_ns), np.percentile(dur_ns, 99), dur_ns.max()))
order = np.argsort(-dur_ns)[:8]
for i in order:
    print("  item blk(%d,%d) sm %d warp %d start %.3f ms end %.3f ms dur %.3f ms | iters primary packet/lane %d/%d bounce packet/lane %d/%d" % (i % (W // 8), i // (W // 8), sm[i], wid[i], start[i] / 1e6, end[i] / 1e6, dur_ns[i] / 1e6, its[0][i], its[1][i], its[2][i], its[3][i]))
tot_it = sum(x.astype(np.int64) for x in its)
print("iterations per item: mean %.0f p99 %.0f max %d ; ns per iteration overall %.0f" % (tot_it.mean(), np.percentile(tot_it, 99), tot_it.max(), dur_ns.sum() / max(1, tot_it.sum())))
# how much of the tail: time after which fewer than 10% of items are still running
last = np.sort(end)
print("last item ends %.3f ms; 99%% of items ended by %.3f ms; 99.9%% by %.3f ms" % (last[-1] / 1e6, last[int(len(last) * .99)] / 1e6, last[int(len(last) * .999)] / 1e6))
# per warp slot: items processed
key = sm.astype(np.int64) * 64 + wid
u, cnt = np.unique(key, return_counts=True)
print("warp slots used %d, items per slot: min %d mean %.1f max %d" % (len(u), cnt.min(), cnt.mean(), cnt.max()))
wc = np.bincount(wid, minlength=64)
print("items by warp id:", wc[wc > 0].tolist())
tr.close()
