"""ctypes access to the CPU oracle (oracle/librtm_oracle.so) and, when built, to
the reference's own object code (oracle/_ref/libref_{head,abc}.so).

TEST INFRASTRUCTURE: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs only — never by raytracingmin_b200/.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LEVEL_HEAD, LEVEL_ABC = 0, 1


class _OScene(C.Structure):
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32), ("samples", C.c_int32), ("superSamples", C.c_int32),
        ("cam_origin", C.c_double * 3), ("cam_target", C.c_double * 3), ("cam_up", C.c_double * 3),
        ("fov", C.c_float), ("n", C.c_int32),
        ("pos", C.POINTER(C.c_double)), ("radius", C.POINTER(C.c_float)),
        ("color", C.POINTER(C.c_double)), ("emission", C.POINTER(C.c_double)),
        ("n_planes", C.c_int32),
        ("pl_position", C.POINTER(C.c_double)), ("pl_up", C.POINTER(C.c_double)), ("pl_target", C.POINTER(C.c_double)),
        ("pl_width", C.POINTER(C.c_double)), ("pl_color", C.POINTER(C.c_double)), ("pl_emission", C.POINTER(C.c_double)),
    ]


def build_oracle(force: bool = False) -> None:
    so = os.path.join(ORACLE_DIR, "librtm_oracle.so")
    if force or not os.path.exists(so):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "librtm_oracle.so"], stdout=subprocess.DEVNULL)


_oracle = None


def oracle():
    global _oracle
    if _oracle is None:
        build_oracle()
        lib = C.CDLL(os.path.join(ORACLE_DIR, "librtm_oracle.so"))
        vp, P = C.c_void_p, C.POINTER
        lib.rtmo_intersect_sphere.argtypes = [vp, C.c_float, vp, vp, P(C.c_double), vp]
        lib.rtmo_camera_frame.argtypes = [P(_OScene), vp]
        lib.rtmo_camera_dir.argtypes = [P(_OScene), vp, C.c_int, C.c_int, C.c_int, C.c_int, vp]
        lib.rtmo_closest_hit.argtypes = [P(_OScene), vp, vp, P(C.c_double), vp]
        lib.rtmo_primary_hits.argtypes = [P(_OScene), vp]
        lib.rtmo_render.argtypes = [P(_OScene), C.c_int, C.c_uint32, vp, P(C.c_uint64), P(C.c_uint64), C.c_int]
        lib.rtmo_to_u8.argtypes = [vp, vp, C.c_size_t]
        lib.rtmo_render_head_exact.argtypes = [P(_OScene), vp]
        lib.rtmo_render_nee.argtypes = [P(_OScene), C.c_uint32, vp, P(C.c_uint64), P(C.c_uint64), P(C.c_uint64)]
        lib.rtmo_plane_frame.argtypes = [vp, vp, vp, C.c_double, vp]
        lib.rtmo_plane_frame.restype = None
        lib.rtmo_intersect_plane.argtypes = [vp, vp, vp, C.c_double, vp, vp, P(C.c_double), vp]
        _oracle = lib
    return _oracle


def _oscene(sc):
    sc = sc.normalized()
    s = _OScene()
    s.width, s.height, s.samples, s.superSamples = sc.width, sc.height, sc.samples, sc.superSamples
    s.cam_origin[:] = list(sc.cam_origin)
    s.cam_target[:] = list(sc.cam_target)
    s.cam_up[:] = list(sc.cam_up)
    s.fov = sc.fov
    s.n = sc.n
    s.pos = sc.position.ctypes.data_as(C.POINTER(C.c_double))
    s.radius = sc.radius.ctypes.data_as(C.POINTER(C.c_float))
    s.color = sc.color.ctypes.data_as(C.POINTER(C.c_double))
    s.emission = sc.emission.ctypes.data_as(C.POINTER(C.c_double))
    s.n_planes = sc.n_planes
    for name in ("pl_position", "pl_up", "pl_target", "pl_width", "pl_color", "pl_emission"):
        setattr(s, name, getattr(sc, name).ctypes.data_as(C.POINTER(C.c_double)))
    return s, sc  # keep sc alive


def oracle_primary_hits(sc) -> np.ndarray:
    s, keep = _oscene(sc)
    out = np.empty((sc.height, sc.width, sc.superSamples, sc.superSamples), np.int32)
    oracle().rtmo_primary_hits(C.byref(s), out.ctypes.data)
    return out


def oracle_render(sc, level: int = LEVEL_ABC, seed: int = 1234, nthreads: int = 0):
    s, keep = _oscene(sc)
    img = np.empty((sc.height, sc.width, 3), np.float64)
    rays, samples = C.c_uint64(), C.c_uint64()
    oracle().rtmo_render(C.byref(s), level, seed, img.ctypes.data, C.byref(rays), C.byref(samples), nthreads)
    return img, rays.value, samples.value


def oracle_render_nee(sc, seed: int = 1234):
    """The next-event-estimation extension: (image, path rays, samples, shadow rays)."""
    s, keep = _oscene(sc)
    img = np.empty((sc.height, sc.width, 3), np.float64)
    rays, samples, shadow = C.c_uint64(), C.c_uint64(), C.c_uint64()
    oracle().rtmo_render_nee(C.byref(s), seed, img.ctypes.data, C.byref(rays), C.byref(samples), C.byref(shadow))
    return img, rays.value, samples.value, shadow.value


def oracle_render_head_exact(sc) -> np.ndarray:
    s, keep = _oscene(sc)
    img = np.empty((sc.height, sc.width, 3), np.float64)
    oracle().rtmo_render_head_exact(C.byref(s), img.ctypes.data)
    return img


def oracle_to_u8(img: np.ndarray) -> np.ndarray:
    img = np.ascontiguousarray(img, np.float64)
    out = np.empty(img.shape, np.uint8)
    oracle().rtmo_to_u8(img.ctypes.data, out.ctypes.data, img.size)
    return out


def oracle_closest_hit(sc, org, dir):
    s, keep = _oscene(sc)
    org = np.ascontiguousarray(org, np.float64).reshape(-1, 3)
    dir = np.ascontiguousarray(dir, np.float64).reshape(-1, 3)
    n = org.shape[0]
    hit = np.empty(n, np.int32)
    t = np.empty(n, np.float64)
    nrm = np.zeros((n, 3), np.float64)
    lib = oracle()
    tt = C.c_double()
    for i in range(n):
        hit[i] = lib.rtmo_closest_hit(C.byref(s), org[i].ctypes.data, dir[i].ctypes.data, C.byref(tt), nrm[i].ctypes.data)
        t[i] = tt.value
    return hit, t, nrm


def oracle_camera_dirs(sc, xs, ys, sxs, sys_):
    s, keep = _oscene(sc)
    fr = np.empty(11, np.float64)
    lib = oracle()
    lib.rtmo_camera_frame(C.byref(s), fr.ctypes.data)
    out = np.empty((len(xs), 3), np.float64)
    for i in range(len(xs)):
        lib.rtmo_camera_dir(C.byref(s), fr.ctypes.data, int(xs[i]), int(ys[i]), int(sxs[i]), int(sys_[i]), out[i].ctypes.data)
    return out


# ---- the reference's own object code (oracle/_ref) -----------------------------

def ref_available(level: str) -> bool:
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", f"libref_{level}.so"))


_refs = {}


def ref_lib(level: str):
    if level not in _refs:
        lib = C.CDLL(os.path.join(ORACLE_DIR, "_ref", f"libref_{level}.so"))
        vp = C.c_void_p
        lib.refh_scene_from_json.argtypes = [C.c_char_p]
        lib.refh_scene_from_json.restype = vp
        lib.refh_scene_from_arrays.argtypes = [C.c_int] * 4 + [vp, C.c_float, C.c_int, vp, vp, vp, vp]
        lib.refh_scene_from_arrays.restype = vp
        lib.refh_scene_destroy.argtypes = [vp]
        lib.refh_scene_destroy.restype = None
        lib.refh_scene_info.argtypes = [vp] + [C.POINTER(C.c_int)] * 5
        lib.refh_scene_set_dims.argtypes = [vp] + [C.c_int] * 4
        lib.refh_scene_get.argtypes = [vp, vp, C.POINTER(C.c_float), vp, vp, vp, vp]
        lib.refh_intersect.argtypes = [vp, C.c_int, vp, vp, C.POINTER(C.c_double), vp]
        lib.refh_primary_hits.argtypes = [vp, vp, C.c_int]
        lib.refh_render.argtypes = [vp, C.c_uint32, vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.c_int]
        lib.refh_to_u8.argtypes = [vp, vp, C.c_size_t]
        if hasattr(lib, "refh_stb_write"):
            lib.refh_stb_write.argtypes = [C.c_char_p, vp, C.c_int, C.c_int, C.c_int]
        if hasattr(lib, "refh_render_window"):
            lib.refh_render_window.argtypes = [vp, C.c_uint32] + [C.c_int] * 4 + [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.c_int]
        _refs[level] = lib
    return _refs[level]


class RefScene:
    """A scene living inside the reference's own SettingData / SphereObject objects."""

    def __init__(self, level: str, sc=None, json_path: str | None = None):
        self.lib = ref_lib(level)
        if json_path is not None:
            self.h = self.lib.refh_scene_from_json(json_path.encode())
        else:
            sc = sc.normalized()
            cam9 = np.concatenate([sc.cam_origin, sc.cam_target, sc.cam_up]).astype(np.float64)
            self.h = self.lib.refh_scene_from_arrays(
                sc.width, sc.height, sc.samples, sc.superSamples, cam9.ctypes.data, C.c_float(sc.fov), sc.n,
                sc.position.ctypes.data, sc.radius.ctypes.data, sc.color.ctypes.data, sc.emission.ctypes.data)
        w, h, s, ss, n = (C.c_int() for _ in range(5))
        self.lib.refh_scene_info(self.h, w, h, s, ss, n)
        self.width, self.height, self.samples, self.superSamples, self.n = w.value, h.value, s.value, ss.value, n.value

    def set_dims(self, width=0, height=0, samples=0, superSamples=0):
        self.lib.refh_scene_set_dims(self.h, width, height, samples, superSamples)
        w, h, s, ss, n = (C.c_int() for _ in range(5))
        self.lib.refh_scene_info(self.h, w, h, s, ss, n)
        self.width, self.height, self.samples, self.superSamples, self.n = w.value, h.value, s.value, ss.value, n.value

    def get(self):
        cam9 = np.empty(9)
        fov = C.c_float()
        pos, col, emi = np.empty((self.n, 3)), np.empty((self.n, 3)), np.empty((self.n, 3))
        rad = np.empty(self.n, np.float32)
        self.lib.refh_scene_get(self.h, cam9.ctypes.data, C.byref(fov), pos.ctypes.data, rad.ctypes.data, col.ctypes.data, emi.ctypes.data)
        return cam9, fov.value, pos, rad, col, emi

    def primary_hits(self, mode: int = 0) -> np.ndarray:
        out = np.empty((self.height, self.width, self.superSamples, self.superSamples), np.int32)
        self.lib.refh_primary_hits(self.h, out.ctypes.data, mode)
        return out

    def render(self, seed: int = 1234, nthreads: int = 0):
        img = np.empty((self.height, self.width, 3), np.float64)
        rays, samples = C.c_uint64(), C.c_uint64()
        self.lib.refh_render(self.h, seed, img.ctypes.data, C.byref(rays), C.byref(samples), nthreads)
        return img, rays.value, samples.value

    def render_window(self, seed: int, x0: int, y0: int, w: int, h: int, nthreads: int = 0):
        """Only the pixels [x0,x0+w) x [y0,y0+h) of the full frame (same camera rays, row generators keyed by the
        full-frame row)."""
        img = np.empty((h, w, 3), np.float64)
        rays, samples = C.c_uint64(), C.c_uint64()
        rc = self.lib.refh_render_window(self.h, seed, x0, y0, w, h, img.ctypes.data, C.byref(rays), C.byref(samples), nthreads)
        if rc:
            raise ValueError(f"refh_render_window: {rc}")
        return img, rays.value, samples.value

    def intersect(self, obj: int, org, dir):
        org = np.ascontiguousarray(org, np.float64)
        dir = np.ascontiguousarray(dir, np.float64)
        t = C.c_double()
        nrm = np.zeros(3)
        hit = self.lib.refh_intersect(self.h, obj, org.ctypes.data, dir.ctypes.data, C.byref(t), nrm.ctypes.data)
        return bool(hit), t.value, nrm

    def close(self):
        if self.h:
            self.lib.refh_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
