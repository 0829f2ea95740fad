"""In-tree builds: librt_b200.so (CUDA, sm_100a), librt_host.so (headless host side).

The .so files are git-ignored but travel to the GPU box with the gpurun snapshot.
"""
import os
import shutil
import subprocess

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
HOST = os.path.join(PKG, "host")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-Xcompiler", "-fPIC", "-shared"]
CXX_FLAGS = ["-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-pthread"]


def _newer(target, sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources)


def _run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("build failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))


def cuda_sources():
    return [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cu", ".cuh", ".h"))] + \
           [os.path.join(ROOT, "include", "rt_api.h"), os.path.join(ROOT, "include", "rt_flat.h"), os.path.join(HOST, "rt_anim_math.h")]


def build_cuda(force=False, fast=False):
    """librt_b200.so: the PARITY build (-fmad=false: every hit / miss decision in the shader's own IEEE operations).
    fast=True: librt_b200_fast.so, the same source with FMA contraction allowed (SURVEY par. 7's "fast build"); bench.py
    reports its speed and how many ids / colours differ from the parity build.  Never the default."""
    out = os.path.join(CSRC, "librt_b200_fast.so" if fast else "librt_b200.so")
    if not force and _newer(out, cuda_sources()):
        return out
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        if os.path.exists(out):
            return out
        raise RuntimeError("nvcc not found and no prebuilt " + os.path.basename(out))
    flags = [f for f in NVCC_FLAGS if f != "-fmad=false"] + ["-fmad=true", "-DRT_FAST_BUILD=1"] if fast else NVCC_FLAGS
    _run([nvcc] + flags + ["-o", out, os.path.join(CSRC, "rt_api.cu")])
    return out


def build_host(force=False):
    out = os.path.join(HOST, "librt_host.so")
    srcs = [os.path.join(HOST, "rt_host.cpp"), os.path.join(HOST, "rt_host.h"), os.path.join(HOST, "rt_anim_math.h")]
    if not force and _newer(out, srcs):
        return out
    _run(["g++"] + CXX_FLAGS + ["-o", out, srcs[0]])
    return out


def build_all(force=False):
    return build_cuda(force), build_host(force), build_cuda(force, fast=True)


if __name__ == "__main__":
    print(build_all(force=True))
