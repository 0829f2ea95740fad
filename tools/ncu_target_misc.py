#!/usr/bin/env python3
"""A short launch sequence of the streaming kernels (K3 rt_untile, K5 rt_present at 7680x4320) and of the cluster build (K8 on the
car) for ncu:  ncu --set full -k regex:'untile_kernel|present_kernel|bld_cluster_kernel' -s 3 -c 3 python tools/ncu_target_misc.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, rtload
rt = rtload.load()
W, H, world = 7680, 4320, 8
tr = rt.RayTracer(W, H)
tr.set_shard(0, world)
px = tr.shard_pixels()
shards = torch.rand((world * px, 4), dtype=torch.float32, device="cuda")
full = torch.empty((H, W, 4), dtype=torch.float32, device="cuda")
img8 = torch.empty((H, W, 4), dtype=torch.uint8, device="cuda")
sc = rt.Scene("car", 64, 48)
b = rt.RayTracer(64, 48, flags=0)
b.upload_scene(sc.light, sc.camera, sc.shapes)
for _ in range(2):                      # launches 0-2 warm up, 3-5 are the ones to capture
    tr.untile(shards.data_ptr(), world, full.data_ptr())
    tr.present(full.data_ptr(), img8.data_ptr(), W * H)
    b.build_bvh(25, fetch=False)
    tr.synchronize(); b.synchronize()
print("done")
tr.close(); b.close()
