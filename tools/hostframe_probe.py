#!/usr/bin/env python3
"""Where does the host-frame path saturate?  torchrun --nproc-per-node N tools/hostframe_probe.py
Concurrent D2H bandwidth of N GPUs (copy engine, private pinned vs one shared segment) and the render kernel
storing into the shared host frame (no upload), all ranks at once."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist, rtload
rt = rtload.load()
rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
W, H = 3840, 2160
total = W * H * 16
mine = total // world
if rank == 0:
    os.system("nvidia-smi topo -m | head -30; numactl -H 2>/dev/null | head -8")

def timed(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize(); dist.barrier()
    t = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); dist.barrier()
    ms = (time.perf_counter() - t) * 1e3 / n
    tt = torch.tensor([ms], device="cuda", dtype=torch.float64); dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    return tt.item()

src = torch.empty(mine, dtype=torch.uint8, device="cuda")
priv = torch.empty(mine, dtype=torch.uint8, pin_memory=True)
ms = timed(lambda: priv.copy_(src, non_blocking=True))
if rank == 0: print("N=%d  CE D2H %d MB per rank to private pinned: %.3f ms (%.1f GB/s per GPU, %.1f aggregate)" % (world, mine >> 20, ms, mine / ms / 1e6, total / ms / 1e6))

sc = rt.Scene("car", W, H)
tr = rt.RayTracer(W, H, device=local, flags=0)
tr.set_stream(torch.cuda.current_stream().cuda_stream)
tr.set_shard(rank, world); tr.upload(sc)
p = rt.make_params(W, H)
name = ["/dev/shm/rt_probe_%d" % os.getpid() if rank == 0 else None]
if rank == 0:
    arr = np.memmap(name[0], dtype=np.float32, mode="w+", shape=(H * W * 4,)); arr[:] = 0
dist.broadcast_object_list(name, src=0)
if rank != 0: arr = np.memmap(name[0], dtype=np.float32, mode="r+", shape=(H * W * 4,))
dist.barrier()
if rank == 0: os.unlink(name[0])
alias = tr.host_frame_register(arr)
# CE copy into this rank's contiguous slice of the shared segment
import ctypes
cudart = torch.cuda.cudart()
dst_ptr = alias + rank * mine
def ce_shared():
    ctypes.memmove  # no-op, keep ctypes imported
    rt.lib().rt_frame_download  # no-op
    torch.cuda.current_stream()
    err = cudart.cudaMemcpyAsync(dst_ptr, src.data_ptr(), mine, 2, torch.cuda.current_stream().cuda_stream) if hasattr(cudart, "cudaMemcpyAsync") else None
try:
    if hasattr(cudart, "cudaMemcpyAsync"):
        ms = timed(ce_shared)
        if rank == 0: print("N=%d  CE D2H to a contiguous slice of the shared segment: %.3f ms (%.1f GB/s aggregate)" % (world, ms, total / ms / 1e6))
except Exception as e:
    if rank == 0: print("CE shared skipped:", e)
tr.set_frame_target(alias)
ms = timed(lambda: tr.render_device(p))
if rank == 0: print("N=%d  render kernel storing into the shared host frame: %.3f ms (%.1f GB/s aggregate)" % (world, ms, total / ms / 1e6))
# private pinned frame per rank (no sharing): isolates the shared-segment effects
privf = torch.empty(H * W * 4, dtype=torch.float32, pin_memory=True)
tr.set_frame_target(privf.data_ptr())
ms = timed(lambda: tr.render_device(p))
if rank == 0: print("N=%d  render kernel storing into a PRIVATE pinned frame per rank: %.3f ms" % (world, ms))
tr.set_frame_target(None)
ms = timed(lambda: tr.upload_scene(sc.light, sc.camera, sc.shapes, sc.nodes, sc.indices))
if rank == 0: print("N=%d  upload_scene alone: %.3f ms" % (world, ms))
ms = timed(lambda: tr.render_device(p))
if rank == 0: print("N=%d  render to the device shard: %.3f ms" % (world, ms))
tr.host_frame_unregister(arr); tr.close()
dist.barrier(); dist.destroy_process_group()
