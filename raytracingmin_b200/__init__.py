"""raytracingmin_b200 — B200-native renderer for RaytracingMin's per-pixel radiance loop.

The product is the CUDA library `librtm_b200.so` (csrc/, C ABI in
include/rtm_b200.h) plus the C++ host front end (host/).  This Python package is
the thin ctypes front end used by the tests and bench.py.  There is no CPU
fallback: importing works anywhere, rendering needs the built library and a GPU.
"""
from .scene import Scene, SceneError, load_scene, scene_from_dict, scene_to_dict, random_scene  # noqa: F401
from .api import (  # noqa: F401
    Renderer, RtmError, RenderResult, Stats, lib_path, load_library,
    COMPAT_HEAD, COMPAT_ABC, PRECISION_F32, PRECISION_F64, VARIANT_MEGA, VARIANT_WAVEFRONT,
    ACCEL_AUTO, ACCEL_BRUTE, ACCEL_BVH,
)

from .partition import sample_slice, plan_partition, reduce_partials  # noqa: F401

__all__ = [
    "sample_slice", "plan_partition", "reduce_partials",
    "Scene", "SceneError", "load_scene", "scene_from_dict", "scene_to_dict", "random_scene",
    "Renderer", "RtmError", "RenderResult", "Stats", "lib_path", "load_library",
]
