import os, sys, time, json
sys.path.insert(0, '.')
import numpy as np, torch, rtload
rt = rtload.load()
out = {"lib": os.environ.get("RT_B200_LIB", "in-tree")}
for name, W, H in (("car", 3840, 2160), ("car", 800, 600), ("monkey", 800, 600)):
    sc = rt.Scene(name, W, H); p = rt.make_params(W, H)
    tr = rt.RayTracer(W, H, flags=0); tr.upload(sc)
    host = torch.empty((H * W, 4), dtype=torch.float32, pin_memory=True).numpy()
    light, cam, shapes, nodes, idx = sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices
    def frame():
        tr.upload_scene(light, cam, shapes, nodes, idx); tr.render(p, want_ids=False, out=host)
    for _ in range(5): frame()
    t0 = time.perf_counter()
    for _ in range(20): frame()
    out["%s_%dx%d" % (name, W, H)] = round((time.perf_counter() - t0) * 1e3 / 20, 4)
    tr.close()
print(json.dumps(out))
