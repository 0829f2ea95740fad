"""B200-native drop-in for the opengl-ray-tracer compute-shader dispatch.

The product is csrc/librt_b200.so (C ABI in include/rt_api.h).  This package only
holds the build recipe, the ctypes harness bindings and the headless host library.
The directory name contains '-', so load it with `rtload.load()` at the repo root.
"""
from . import build, shard  # noqa: F401
from .binding import (CAMERA_DTYPE, LIGHT_DTYPE, NODE_DTYPE, SHAPE_DTYPE, RT_FLAG_REFERENCE_WALK, RT_FLAG_HIT_IDS, RT_FLAG_NO_TOP_TREE, RT_FLAG_NO_HEAVY_FIRST, RT_FLAG_NO_DEFERRED_COPY, RT_FLAG_NO_ONE_BLOCK_UPDATE, RT_BUILD_DEVICE_LAYOUT,  # noqa: F401
                      API_SYMBOLS, RT_STATIC_BUILD_MAX_SHAPES, build_bvh_static, obj_triangles, Params, RayTracer, RtError, Scene, hostlib, lib, lib_fast, lib_path, make_params)
