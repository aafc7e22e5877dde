"""ctypes binding of the C ABI in include/rtm_b200.h (librtm_b200.so).

The class below plays the role png::Renderer plays in the reference
(src/Renderer.h:9-17): construct from the parsed scene, call render().  There
is no CPU fallback — a missing library or GPU raises RtmError.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Optional

import numpy as np

from .scene import Scene

COMPAT_HEAD, COMPAT_ABC = 0, 1
PRECISION_F32, PRECISION_F64 = 0, 1
VARIANT_MEGA, VARIANT_WAVEFRONT = 0, 1
ACCEL_AUTO, ACCEL_BRUTE, ACCEL_BVH = 0, 1, 2

_COMPAT = {"head": COMPAT_HEAD, "abc": COMPAT_ABC}
_PRECISION = {"f32": PRECISION_F32, "f64": PRECISION_F64}
_VARIANT = {"mega": VARIANT_MEGA, "wavefront": VARIANT_WAVEFRONT}
_ACCEL = {"auto": ACCEL_AUTO, "brute": ACCEL_BRUTE, "bvh": ACCEL_BVH}
ESTIMATOR_REFERENCE, ESTIMATOR_NEE = 0, 1
_ESTIMATOR = {"reference": ESTIMATOR_REFERENCE, "nee": ESTIMATOR_NEE}


class RtmError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"rtm error {code}: {msg}")
        self.code = code


class _SceneDesc(C.Structure):
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32), ("samples", C.c_int32), ("superSamples", C.c_int32),
        ("cam_origin", C.c_double * 3), ("cam_target", C.c_double * 3), ("cam_up", C.c_double * 3),
        ("fov", C.c_float), ("n_spheres", C.c_int32),
        ("position", C.POINTER(C.c_double)), ("radius", C.POINTER(C.c_float)),
        ("color", C.POINTER(C.c_double)), ("emission", C.POINTER(C.c_double)),
        # ABI 3: PlaneObject extension
        ("n_planes", C.c_int32),
        ("plane_position", C.POINTER(C.c_double)), ("plane_up", C.POINTER(C.c_double)), ("plane_target", C.POINTER(C.c_double)),
        ("plane_width", C.POINTER(C.c_double)), ("plane_color", C.POINTER(C.c_double)), ("plane_emission", C.POINTER(C.c_double)),
    ]


class _Options(C.Structure):
    _fields_ = [
        ("device", C.c_int32), ("n_gpus", C.c_int32), ("compat", C.c_int32), ("precision", C.c_int32),
        ("variant", C.c_int32), ("accel", C.c_int32), ("max_depth", C.c_int32),
        ("sample_begin", C.c_int32), ("sample_end", C.c_int32), ("seed", C.c_uint64),
        ("sub_begin", C.c_int32), ("sub_end", C.c_int32), ("estimator", C.c_int32),
        # ABI 4: test / A-B knobs (0 = product defaults); they replace round 1's environment variables
        ("debug_retry", C.c_int32), ("debug_bvh_walk", C.c_int32), ("debug_bvh_leaf_max", C.c_int32),
        ("debug_bvh_ctrav", C.c_float), ("debug_clamp", C.c_int32),
    ]


class _Stats(C.Structure):
    _fields_ = [
        ("samples", C.c_uint64), ("rays", C.c_uint64), ("rays_traced", C.c_uint64), ("sphere_tests", C.c_uint64),
        ("device_ms", C.c_double), ("e2e_ms", C.c_double), ("kernel_launches", C.c_int32), ("n_gpus", C.c_int32),
        ("bvh_nodes", C.c_uint64),
    ]


@dataclass
class Stats:
    samples: int
    rays: int
    rays_traced: int
    sphere_tests: int
    device_ms: float
    e2e_ms: float
    kernel_launches: int
    n_gpus: int
    bvh_nodes: int = 0


@dataclass
class RenderResult:
    image: Optional[np.ndarray]  # float32 [H][W][3] linear sums (src/Renderer.cpp:246-248)
    hits: Optional[np.ndarray]   # int32 [H][W][ss][ss], -1 = miss
    stats: Stats


EXPORTS = [
    "renderer_abi_version", "renderer_default_options", "renderer_create", "renderer_render",
    "renderer_render_async", "renderer_render_noalloc", "renderer_read_stats", "renderer_read_kernel_ms", "renderer_device_accum", "renderer_device_hits",
    "renderer_set_accum_buffer", "renderer_peer_export", "renderer_peer_open", "renderer_sum_partials", "renderer_read_rgb", "renderer_read_image_f64", "renderer_read_u8", "renderer_destroy",
    "renderer_last_error", "rtm_debug_philox", "rtm_debug_closest_hit", "rtm_debug_fp32_peak", "rtm_debug_bvh", "rtm_debug_bvh16", "rtm_debug_bvh16x4", "rtm_debug_bvh4", "rtm_debug_acne_eps",
]


def lib_path() -> str:
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "librtm_b200.so")


_lib = None


def load_library():
    """Loads librtm_b200.so (built in-tree by `make -C raytracingmin_b200/csrc`).  Fails loudly."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise RtmError(-2, f"{path} is not built (run `make -C raytracingmin_b200/csrc`); there is no CPU fallback")
    lib = C.CDLL(path)
    P = C.POINTER
    vp = C.c_void_p
    lib.renderer_abi_version.restype = C.c_int
    lib.renderer_default_options.argtypes = [P(_Options)]
    lib.renderer_default_options.restype = None
    lib.renderer_create.argtypes = [P(_SceneDesc), P(_Options), P(vp)]
    lib.renderer_render.argtypes = [vp, vp, vp, P(_Stats)]
    lib.renderer_render_async.argtypes = [vp, vp]
    lib.renderer_render_noalloc.argtypes = [vp, vp, vp]
    lib.renderer_render_noalloc.restype = C.c_int
    lib.renderer_read_stats.argtypes = [vp, P(_Stats)]
    lib.renderer_read_kernel_ms.argtypes = [vp, vp]
    lib.renderer_device_accum.argtypes = [vp]
    lib.renderer_device_accum.restype = vp
    lib.renderer_device_hits.argtypes = [vp]
    lib.renderer_device_hits.restype = vp
    lib.renderer_set_accum_buffer.argtypes = [vp, vp]
    lib.renderer_peer_export.argtypes = [vp, C.c_int32, vp, P(vp)]
    lib.renderer_peer_open.argtypes = [vp, vp, P(vp)]
    lib.renderer_sum_partials.argtypes = [vp, vp, C.c_int32, vp, vp]
    lib.renderer_read_rgb.argtypes = [vp, vp, vp]
    lib.renderer_read_image_f64.argtypes = [vp, vp]
    lib.renderer_read_u8.argtypes = [vp, vp]
    lib.renderer_destroy.argtypes = [vp]
    lib.renderer_destroy.restype = None
    lib.renderer_last_error.restype = C.c_char_p
    lib.rtm_debug_philox.argtypes = [C.c_int32] + [C.c_uint32] * 6 + [vp]
    lib.rtm_debug_closest_hit.argtypes = [vp, C.c_int32, C.c_uint32, vp, vp, vp, vp, vp]
    lib.rtm_debug_fp32_peak.argtypes = [C.c_int32, P(C.c_double), P(C.c_double)]
    lib.rtm_debug_bvh.argtypes = [P(_SceneDesc), vp, vp, vp, vp]
    lib.rtm_debug_bvh16.argtypes = [P(_SceneDesc), vp, vp, vp]
    lib.rtm_debug_bvh16x4.argtypes = [P(_SceneDesc), vp, vp]
    lib.rtm_debug_bvh4.argtypes = [P(_SceneDesc), vp, vp]
    lib.rtm_debug_acne_eps.argtypes = [C.c_int32, C.c_uint32, vp, vp]
    _lib = lib
    return lib


def _check(lib, rc: int):
    if rc != 0:
        raise RtmError(rc, lib.renderer_last_error().decode("utf-8", "replace"))


def _stats(s: _Stats) -> Stats:
    return Stats(s.samples, s.rays, s.rays_traced, s.sphere_tests, s.device_ms, s.e2e_ms, s.kernel_launches, s.n_gpus, s.bvh_nodes)


class Renderer:
    """One handle = one scene on one device (or n_gpus devices of this process)."""

    def __init__(self, scene: Scene, *, device: int = 0, n_gpus: int = 1, compat: str = "abc", precision: str = "f32",
                 variant: str = "mega", accel: str = "auto", max_depth: int = 0, seed: int = 0x5EED,
                 sample_begin: int = 0, sample_end: int = 0, sub_begin: int = 0, sub_end: int = 0,
                 estimator: str = "reference", debug_retry: int = 0, debug_bvh_walk: int = 0, debug_bvh_leaf_max: int = 0,
                 debug_bvh_ctrav: float = 0.0, debug_clamp: int = 0):
        self._lib = load_library()
        self._h = C.c_void_p()
        sc = scene.normalized()
        self.scene = sc
        d = self._desc(sc)
        o = _Options()
        self._lib.renderer_default_options(C.byref(o))
        o.device, o.n_gpus = device, n_gpus
        o.compat, o.precision = _COMPAT[compat], _PRECISION[precision]
        o.variant, o.accel = _VARIANT[variant], _ACCEL[accel]
        o.max_depth, o.seed = max_depth, seed
        o.sample_begin, o.sample_end = sample_begin, sample_end
        o.sub_begin, o.sub_end = sub_begin, sub_end
        o.estimator = _ESTIMATOR[estimator]
        o.debug_retry, o.debug_bvh_walk, o.debug_bvh_leaf_max = debug_retry, debug_bvh_walk, debug_bvh_leaf_max
        o.debug_bvh_ctrav, o.debug_clamp = debug_bvh_ctrav, debug_clamp
        self.compat = compat
        _check(self._lib, self._lib.renderer_create(C.byref(d), C.byref(o), C.byref(self._h)))

    @staticmethod
    def _desc(sc: Scene) -> "_SceneDesc":
        d = _SceneDesc()
        d.width, d.height, d.samples, d.superSamples = sc.width, sc.height, sc.samples, sc.superSamples
        d.cam_origin[:] = list(sc.cam_origin)
        d.cam_target[:] = list(sc.cam_target)
        d.cam_up[:] = list(sc.cam_up)
        d.fov = sc.fov
        d.n_spheres = sc.n
        d.position = sc.position.ctypes.data_as(C.POINTER(C.c_double))
        d.radius = sc.radius.ctypes.data_as(C.POINTER(C.c_float))
        d.color = sc.color.ctypes.data_as(C.POINTER(C.c_double))
        d.emission = sc.emission.ctypes.data_as(C.POINTER(C.c_double))
        d.n_planes = sc.n_planes
        if sc.n_planes:
            dp = C.POINTER(C.c_double)
            d.plane_position, d.plane_up = sc.pl_position.ctypes.data_as(dp), sc.pl_up.ctypes.data_as(dp)
            d.plane_target, d.plane_width = sc.pl_target.ctypes.data_as(dp), sc.pl_width.ctypes.data_as(dp)
            d.plane_color, d.plane_emission = sc.pl_color.ctypes.data_as(dp), sc.pl_emission.ctypes.data_as(dp)
        return d

    # -- the drop-in call: host buffers in, host buffers out
    def render(self, want_image: bool = True, want_hits: bool = False, image_out: Optional[np.ndarray] = None) -> RenderResult:
        """renderer_render: host buffers out.  `image_out` (float32 [H][W][3], C-contiguous) is reused when given."""
        sc = self.scene
        if image_out is not None:
            assert image_out.dtype == np.float32 and image_out.shape == (sc.height, sc.width, 3) and image_out.flags.c_contiguous
        img = (image_out if image_out is not None else np.empty((sc.height, sc.width, 3), np.float32)) if want_image else None
        hits = np.empty((sc.height, sc.width, sc.superSamples, sc.superSamples), np.int32) if want_hits else None
        st = _Stats()
        _check(self._lib, self._lib.renderer_render(
            self._h, img.ctypes.data if img is not None else None, hits.ctypes.data if hits is not None else None, C.byref(st)))
        return RenderResult(img, hits, _stats(st))

    # -- device-resident pass on a caller-supplied CUDA stream (0 = default stream)
    def render_async(self, stream: int = 0) -> None:
        _check(self._lib, self._lib.renderer_render_async(self._h, C.c_void_p(stream)))

    def render_noalloc(self, accum_device_ptr: int, stream: int = 0) -> None:
        """renderer_render_noalloc: one FP32 pass into the caller's float4[H][W] device buffer with the pass's scratch in
        the library's fixed device array (no cudaMalloc in the call; include/rtm_b200.h for the limit and the sharing rule)."""
        _check(self._lib, self._lib.renderer_render_noalloc(self._h, C.c_void_p(accum_device_ptr), C.c_void_p(stream)))

    def read_stats(self) -> Stats:
        st = _Stats()
        _check(self._lib, self._lib.renderer_read_stats(self._h, C.byref(st)))
        return _stats(st)

    def read_kernel_ms(self):
        """(primary, radiance, accumulate) CUDA-event ms of the last FP32 pass."""
        ms = (C.c_double * 3)()
        _check(self._lib, self._lib.renderer_read_kernel_ms(self._h, ms))
        return tuple(ms)

    def device_accum_ptr(self) -> int:
        return int(self._lib.renderer_device_accum(self._h) or 0)

    def set_accum_buffer(self, device_ptr: int) -> None:
        _check(self._lib, self._lib.renderer_set_accum_buffer(self._h, C.c_void_p(device_ptr)))

    # -- peer-memory gather (one process per GPU), see include/rtm_b200.h
    def peer_export(self, n_slots: int):
        h = (C.c_char * 64)()
        p = C.c_void_p()
        _check(self._lib, self._lib.renderer_peer_export(self._h, n_slots, h, C.byref(p)))
        return bytes(h.raw), int(p.value)

    def peer_open(self, handle: bytes) -> int:
        h = (C.c_char * 64).from_buffer_copy(handle)
        p = C.c_void_p()
        _check(self._lib, self._lib.renderer_peer_open(self._h, h, C.byref(p)))
        return int(p.value)

    def sum_partials(self, partials_ptr: int, n_slots: int, out_ptr: int, stream: int = 0) -> None:
        _check(self._lib, self._lib.renderer_sum_partials(self._h, C.c_void_p(partials_ptr), n_slots, C.c_void_p(out_ptr),
                                                          C.c_void_p(stream)))

    def read_rgb(self, out: Optional[np.ndarray] = None, stream: int = 0) -> np.ndarray:
        """The accumulation buffer as float32 [H][W][3] in host memory (renderer_read_rgb)."""
        sc = self.scene
        if out is None:
            out = np.empty((sc.height, sc.width, 3), np.float32)
        assert out.dtype == np.float32 and out.shape == (sc.height, sc.width, 3) and out.flags.c_contiguous
        _check(self._lib, self._lib.renderer_read_rgb(self._h, out.ctypes.data, C.c_void_p(stream)))
        return out

    def read_image_f64(self) -> np.ndarray:
        sc = self.scene
        out = np.empty((sc.height, sc.width, 3), np.float64)
        _check(self._lib, self._lib.renderer_read_image_f64(self._h, out.ctypes.data))
        return out

    def read_u8(self) -> np.ndarray:
        sc = self.scene
        out = np.empty((sc.height, sc.width, 3), np.uint8)
        _check(self._lib, self._lib.renderer_read_u8(self._h, out.ctypes.data))
        return out

    def closest_hit(self, org: np.ndarray, dir: np.ndarray, mode: str = "f64", origin_id: Optional[np.ndarray] = None):
        org = np.ascontiguousarray(org, np.float64).reshape(-1, 3)
        dir = np.ascontiguousarray(dir, np.float64).reshape(-1, 3)
        n = org.shape[0]
        hit = np.empty(n, np.int32)
        t = np.empty(n, np.float64)
        oid = None if origin_id is None else np.ascontiguousarray(origin_id, np.int32)
        _check(self._lib, self._lib.rtm_debug_closest_hit(
            self._h, 0 if mode == "f64" else 1, n, org.ctypes.data, dir.ctypes.data,
            oid.ctypes.data if oid is not None else None, hit.ctypes.data, t.ctypes.data))
        return hit, t

    def close(self) -> None:
        if getattr(self, "_h", None) and self._h.value:
            self._lib.renderer_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def philox_device(n: int, c1: int, c2: int, c3: int, k0: int, k1: int, device: int = 0) -> np.ndarray:
    lib = load_library()
    out = np.empty((n, 4), np.uint32)
    _check(lib, lib.rtm_debug_philox(device, n, c1, c2, c3, k0, k1, out.ctypes.data))
    return out


def build_bvh(scene: Scene) -> dict:
    """The sphere hierarchy renderer_create builds for `scene` (host code only; no GPU needed)."""
    lib = load_library()
    sc = scene.normalized()
    d = Renderer._desc(sc)
    info = np.zeros(13, np.int32)
    rc = lib.rtm_debug_bvh(C.byref(d), info.ctypes.data, None, None, None)
    nodes = np.zeros((int(info[0]), 4, 4), np.float32)
    leaf_geo = np.zeros((int(info[1]), 4), np.float32)
    leaf_id = np.zeros(int(info[1]), np.int32)
    rc = lib.rtm_debug_bvh(C.byref(d), info.ctypes.data, nodes.ctypes.data, leaf_geo.ctypes.data, leaf_id.ctypes.data)
    _check(lib, rc)
    return dict(nodes=nodes, leaf_geo=leaf_geo, leaf_id=leaf_id, root=int(info[2]), depth=int(info[3]),
                big_id=[int(v) for v in info[5:5 + int(info[4])]])


def build_bvh16(scene: Scene) -> dict:
    """The compact half-precision form of build_bvh's tree (host code only): nodes[k] = 8 uint32."""
    lib = load_library()
    d = Renderer._desc(scene.normalized())
    n, meta = C.c_int32(), np.zeros(6, np.float32)
    _check(lib, lib.rtm_debug_bvh16(C.byref(d), C.byref(n), meta.ctypes.data, None))
    nodes = np.zeros((n.value, 8), np.uint32)
    _check(lib, lib.rtm_debug_bvh16(C.byref(d), C.byref(n), meta.ctypes.data, nodes.ctypes.data))
    return dict(nodes=nodes, ref=meta[:3].astype(np.float64), inv_scale=float(meta[3]), inflation=float(meta[4]), ok=bool(meta[5]))


def build_bvh16x4(scene: Scene) -> dict:
    """The 4-wide collapse of build_bvh16's tree, which the FP32 walk uses (host code only): nodes[k] = 16 uint32."""
    lib = load_library()
    d = Renderer._desc(scene.normalized())
    info = np.zeros(2, np.int32)
    _check(lib, lib.rtm_debug_bvh16x4(C.byref(d), info.ctypes.data, None))
    nodes = np.zeros((int(info[0]), 16), np.uint32)
    _check(lib, lib.rtm_debug_bvh16x4(C.byref(d), info.ctypes.data, nodes.ctypes.data))
    return dict(nodes=nodes, depth=int(info[1]), ok=bool(info[0] > 0))


def build_bvh4(scene: Scene) -> dict:
    """The 4-wide collapse of build_bvh's tree (host code only): nodes[k] = 8 rows of 4 floats."""
    lib = load_library()
    d = Renderer._desc(scene.normalized())
    info = np.zeros(3, np.int32)
    _check(lib, lib.rtm_debug_bvh4(C.byref(d), info.ctypes.data, None))
    nodes = np.zeros((int(info[0]), 8, 4), np.float32)
    _check(lib, lib.rtm_debug_bvh4(C.byref(d), info.ctypes.data, nodes.ctypes.data))
    return dict(nodes=nodes, root=int(info[1]), depth=int(info[2]))


def acne_eps_device(bits: np.ndarray, device: int = 0) -> np.ndarray:
    lib = load_library()
    bits = np.ascontiguousarray(bits, np.uint32)
    out = np.empty(bits.shape, np.float32)
    _check(lib, lib.rtm_debug_acne_eps(device, bits.size, bits.ctypes.data, out.ctypes.data))
    return out


def fp32_peak_tflops(device: int = 0):
    lib = load_library()
    tf, mhz = C.c_double(), C.c_double()
    _check(lib, lib.rtm_debug_fp32_peak(device, C.byref(tf), C.byref(mhz)))
    return tf.value, mhz.value
