"""PlaneObject extension, CPU side: the oracle's definition against analytic
known answers, and both loaders (Python, C++) on objectType 2.

PARITY UNPINNED for this row by construction: the reference declares PlaneObject
(src/SettingData.h:33-42) but never parses it and its Intersect only rejects parallel rays and
falls off the end (src/SettingData.cpp:243-246).  What the reference DOES define — the ctor's
frame, src/SettingData.cpp:235-242, and the parallel-ray rejection — is checked here literally.
"""
import ctypes as C
import json
import os
import subprocess

import numpy as np
import pytest

import oracle_lib as O
import raytracingmin_b200 as rtm
from raytracingmin_b200.scene import SceneError, scene_from_dict, scene_to_dict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PLANES = os.path.join(ROOT, "scenes", "cornellBoxPlanes.json")
FLT_EPSILON = float(np.finfo(np.float32).eps)


def _frame(pos, up, tgt, w):
    out = np.empty(11)
    a = [np.ascontiguousarray(v, np.float64) for v in (pos, up, tgt)]
    O.oracle().rtmo_plane_frame(a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data, float(w), out.ctypes.data)
    return out[0:3], out[3:6], out[6:9], out[9], out[10]


def _hit(pos, up, tgt, w, org, d):
    a = [np.ascontiguousarray(v, np.float64) for v in (pos, up, tgt, org, d)]
    t = C.c_double()
    nrm = np.zeros(3)
    h = O.oracle().rtmo_intersect_plane(a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data, float(w), a[3].ctypes.data,
                                        a[4].ctypes.data, C.byref(t), nrm.ctypes.data)
    return bool(h), t.value, nrm


def test_plane_frame_follows_the_reference_ctor():
    # src/SettingData.cpp:241-242 with the float-sqrt Normalize of src/Ray.h:67-72, restated in numpy
    pos, up, tgt, w = np.array([1.0, 2.0, 3.0]), np.array([0.0, 1.0, 0.2]), np.array([4.0, -1.0, 0.5]), 3.0

    def normalize(v):
        return v / np.float64(np.sqrt(np.float32(v[0] * v[0] + v[1] * v[1] + v[2] * v[2])))

    def cross(a, b):  # src/Ray.h:64-66
        return np.array([a[1] * b[2] - a[2] * b[1], -a[0] * b[2] + a[2] * b[0], a[0] * b[1] - a[1] * b[0]])

    n = normalize(tgt - pos)
    right = normalize(cross(n, up)) * 0.5 * w
    upp = cross(right, n)
    gn, gr, gu, rr, uu = _frame(pos, up, tgt, w)
    assert np.array_equal(gn, n) and np.array_equal(gr, right) and np.array_equal(gu, upp)
    assert rr == right[0] * right[0] + right[1] * right[1] + right[2] * right[2]
    assert abs(np.sqrt(rr) - 1.5) < 1e-6 and abs(np.sqrt(uu) - 1.5) < 1e-6     # half width
    assert abs(gr @ gn) < 1e-12 and abs(gu @ gn) < 1e-12 and abs(gr @ gu) < 1e-12


def test_plane_intersect_known_answers():
    pos, up, tgt, w = [0, 0, 5], [0, 1, 0], [0, 0, 0], 4.0       # square |x|,|y| <= 2 in the plane z = 5, normal -z
    h, t, n = _hit(pos, up, tgt, w, [0, 0, 0], [0, 0, 1])
    assert h and t == 5.0 and n.tolist() == [0, 0, -1]           # the normal is NOT oriented by Intersect
    h, t, n = _hit(pos, up, tgt, w, [0, 0, 9], [0, 0, -1])       # from behind: same normal, PathTracing orients it
    assert h and t == 4.0 and n.tolist() == [0, 0, -1]
    assert _hit(pos, up, tgt, w, [1.5, -1.5, 0], [0, 0, 1])[0]
    assert _hit(pos, up, tgt, w, [2.0, 2.0, 0], [0, 0, 1])[0]                 # the edge belongs to the square
    assert not _hit(pos, up, tgt, w, [2.0 + 1e-9, 0, 0], [0, 0, 1])[0]
    assert not _hit(pos, up, tgt, w, [0, -2.001, 0], [0, 0, 1])[0]
    assert not _hit(pos, up, tgt, w, [0, 0, 0], [0, 0, -1])[0]                # plane behind the ray
    # src/SettingData.cpp:244-245: |Dot(normal, dir)| < FLT_EPSILON is rejected
    assert not _hit(pos, up, tgt, w, [0, 0, 0], [1, 0, 0])[0]
    assert not _hit(pos, up, tgt, w, [0, 0, 5 - 1e-9], [1, 0, 0.5 * FLT_EPSILON])[0]
    assert _hit(pos, up, tgt, w, [0, 0, 5 - 1e-7], [1e-3, 0, 2 * FLT_EPSILON])[1] > 0.4    # just above the threshold: a hit at t ~ 0.42
    # the near threshold of the sphere test (src/SettingData.cpp:209): t must exceed 1e-3
    assert not _hit(pos, up, tgt, w, [0, 0, 5 - 0.0009], [0, 0, 1])[0]
    assert _hit(pos, up, tgt, w, [0, 0, 5 - 0.0011], [0, 0, 1])[0]
    assert not _hit(pos, up, tgt, w, [0, 0, 5], [0.6, 0, 0.8])[0]             # leaving the surface: no self-hit


def test_closest_hit_orders_spheres_before_planes_on_ties():
    # a sphere touching the plane from the front: the ray along the axis reaches both at t = 3 exactly
    sc = rtm.Scene(width=4, height=4, samples=1, superSamples=1, cam_origin=np.array([0.0, 0, 0]), cam_target=np.array([0.0, 0, 1]),
                   cam_up=np.array([0.0, 1, 0]), fov=1.0, position=np.array([[0.0, 0, 4]]), radius=np.array([1.0], np.float32),
                   color=np.array([[0.5, 0.5, 0.5]]), emission=np.zeros((1, 3)),
                   pl_position=np.array([[0.0, 0, 3]]), pl_up=np.array([[0.0, 1, 0]]), pl_target=np.array([[0.0, 0, 0]]),
                   pl_width=np.array([8.0]), pl_color=np.array([[0.5, 0.5, 0.5]]), pl_emission=np.zeros((1, 3))).normalized()
    hit, t, nrm = O.oracle_closest_hit(sc, [[0, 0, 0], [2.5, 0, 0]], [[0, 0, 1], [0, 0, 1]])
    assert hit.tolist() == [0, 1] and t.tolist() == [3.0, 3.0]
    assert nrm[1].tolist() == [0, 0, -1]


def test_python_loader_planes_and_order(tmp_path):
    sc = rtm.load_scene(PLANES)
    assert (sc.n, sc.n_planes, sc.n_objects) == (3, 6, 9)
    assert sc.object_order.tolist() == [0, 7, 8, 1, 2, 3, 4, 5, 6]       # spheres keep JSON order, planes follow
    assert sc.pl_width.tolist() == [20.0] * 6 and sc.pl_position[2].tolist() == [0, 10, 0]
    d = scene_to_dict(sc)
    sc2 = scene_from_dict(d)
    for name in ("position", "radius", "color", "emission", "pl_position", "pl_up", "pl_target", "pl_width", "pl_color", "pl_emission"):
        assert np.array_equal(getattr(sc, name), getattr(sc2, name)), name
    assert sc2.object_order.tolist() == list(range(9))
    d["02 scene"]["00 object"][3]["00 objectType"] = 3
    with pytest.raises(SceneError, match="neither a sphere"):
        scene_from_dict(d)
    d["02 scene"]["00 object"][3]["00 objectType"] = 2
    d["02 scene"]["00 object"][3]["01 width"] = 0
    with pytest.raises(SceneError, match="width must be positive"):
        scene_from_dict(d)
    with pytest.raises(SceneError, match="L-ABC"):
        rtm.load_scene(PLANES, "head")


def test_cpp_loader_planes_match_python(tmp_path):
    host = os.path.join(ROOT, "raytracingmin_b200", "host")
    subprocess.check_call(["make", "-C", host], stdout=subprocess.DEVNULL)
    lib = C.CDLL(os.path.join(host, "librtm_host.so"))
    lib.rtmh_last_error.restype = C.c_char_p
    vp = C.c_void_p
    lib.rtmh_load_planes.argtypes = [C.c_char_p, C.c_int] + [vp] * 6 + [C.c_int, vp]
    lib.rtmh_roundtrip_json.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
    sc = rtm.load_scene(PLANES)

    def load(path):
        m = lib.rtmh_load_planes(path.encode(), 0, None, None, None, None, None, None, 0, None)
        assert m >= 0, lib.rtmh_last_error()
        a = [np.empty((m, 3)) for _ in range(3)] + [np.empty(m)] + [np.empty((m, 3)) for _ in range(2)]
        order = np.empty(64, np.int32)
        lib.rtmh_load_planes(path.encode(), m, *[x.ctypes.data for x in a], 64, order.ctypes.data)
        return a, order

    a, order = load(PLANES)
    for got, name in zip(a, ("pl_position", "pl_up", "pl_target", "pl_width", "pl_color", "pl_emission")):
        assert np.array_equal(got, getattr(sc, name)), name
    assert order[:9].tolist() == sc.object_order.tolist()
    # to_json round trip keeps the planes
    buf = C.create_string_buffer(1 << 16)
    assert lib.rtmh_roundtrip_json(PLANES.encode(), buf, len(buf)) > 0
    p = tmp_path / "rt.json"
    p.write_bytes(buf.value)
    b, _ = load(str(p))
    for x, y in zip(a, b):
        assert np.array_equal(x, y)
    assert rtm.load_scene(str(p)).n == 3
    bad = json.load(open(PLANES))
    del bad["02 scene"]["00 object"][1]["01 up"]
    p.write_text(json.dumps(bad))
    assert lib.rtmh_load_planes(str(p).encode(), 0, None, None, None, None, None, None, 0, None) == -1
    assert b"01 up" in lib.rtmh_last_error()


def test_planar_cornell_oracle_known_answers():
    """The planar Cornell box is closed: every primary ray hits something, the wall behind the
    camera (object id 8) is never seen, and the histogram is symmetric left/right up to the two
    extra spheres."""
    sc = rtm.load_scene(PLANES).with_dims(120, 63, 2, 2)
    h = O.oracle_primary_hits(sc)
    hist = np.bincount(h.ravel() + 1, minlength=10)
    assert hist.sum() == 120 * 63 * 4 and hist[1 + 8] == 0
    assert hist[0] <= 4                       # only rays through the exact box corners may escape
    assert hist[1 + 3] > 0 and hist[1 + 4] > 0 and abs(int(hist[1 + 3]) - int(hist[1 + 4])) < 0.35 * hist[1 + 3]
    img, rays, samples = O.oracle_render(sc.with_dims(60, 32, 8, 1), O.LEVEL_ABC, 3)
    assert np.isfinite(img).all() and img.min() >= 0 and 4.0 < rays / samples < 7.0


def test_extension_known_answers_are_stable():
    """Regression pins of the two extension definitions (generated by this oracle, committed here: there is no
    reference output for them): the planar Cornell hit histogram at 480x252 ss=2, and one small image each."""
    sc = rtm.load_scene(PLANES).with_dims(480, 252, 1, 2)
    hist = np.bincount(O.oracle_primary_hits(sc).ravel() + 1, minlength=10).tolist()
    assert hist == [4, 27554, 12178, 15889, 163419, 163203, 22171, 26245, 53177, 0]
    # the light sphere's 27554 is the sphere Cornell box's count at the same size: same sphere, same camera
    small = sc.with_dims(32, 18, 4, 1)
    img, rays, samples = O.oracle_render(small, O.LEVEL_ABC, 1234)
    assert (rays, samples) == (11996, 2304) and abs(img.sum() - 931.4044779993226) < 1e-9
    img, rays, samples, shadow = O.oracle_render_nee(small, 1234)
    assert (rays, samples, shadow) == (11896, 2304, 8718) and abs(img.sum() - 1010.8715379767282) < 1e-9
