"""Scene/config loading for the Python front end (tests, bench).

Mirrors what the reference's parser produces (png::LoadData::from_json,
reference src/SettingData.cpp:129-186) as flat arrays ready for the C ABI.
The shipping host loader is the C++ one under raytracingmin_b200/host/; this is
its Python twin with the same policy:

  * unknown keys are ignored (src/SettingData.cpp:130-186 only looks at known ones),
  * "00 sample" is accepted as an alias of "00 samples"; missing "00 superSamples" = 1,
  * an object without "00 objectType" is a sphere (type 1); `{}` is skipped,
  * objectType 2 is the PlaneObject EXTENSION (the reference declares the
    type, src/SettingData.h:33-42, but never parses or intersects it): keys "00 position",
    "01 up", "01 target", "01 width", "02 material" = the ctor arguments of
    src/SettingData.cpp:235.  Planes are numbered AFTER the spheres (object ids n..n+m-1);
    `object_order` maps those ids back to positions in the JSON list,
  * any other objectType is rejected (the reference pushes a nullptr and segfaults),
  * fov is ALWAYS the tangent-space half width (src/Renderer.cpp:207,229),
  * compat="head" reproduces F2: the reference assigns posi.x three times
    (src/SettingData.cpp:165-167), so every sphere lands at (position[2], 0, 0).
"""
from __future__ import annotations

import json
from dataclasses import dataclass, field, replace

import numpy as np


@dataclass
class Scene:
    width: int
    height: int
    samples: int
    superSamples: int
    cam_origin: np.ndarray
    cam_target: np.ndarray
    cam_up: np.ndarray
    fov: float
    position: np.ndarray = field(default_factory=lambda: np.zeros((0, 3)))
    radius: np.ndarray = field(default_factory=lambda: np.zeros((0,), np.float32))
    color: np.ndarray = field(default_factory=lambda: np.zeros((0, 3)))
    emission: np.ndarray = field(default_factory=lambda: np.zeros((0, 3)))
    # PlaneObject extension (objectType 2): finite squares, object ids n .. n+n_planes-1
    pl_position: np.ndarray = field(default_factory=lambda: np.zeros((0, 3)))
    pl_up: np.ndarray = field(default_factory=lambda: np.zeros((0, 3)))
    pl_target: np.ndarray = field(default_factory=lambda: np.zeros((0, 3)))
    pl_width: np.ndarray = field(default_factory=lambda: np.zeros((0,)))
    pl_color: np.ndarray = field(default_factory=lambda: np.zeros((0, 3)))
    pl_emission: np.ndarray = field(default_factory=lambda: np.zeros((0, 3)))
    object_order: np.ndarray = field(default_factory=lambda: np.zeros((0,), np.int32))  # id -> JSON position

    @property
    def n(self) -> int:
        return int(self.position.shape[0])

    @property
    def n_planes(self) -> int:
        return int(np.asarray(self.pl_position).reshape(-1, 3).shape[0])

    @property
    def n_objects(self) -> int:
        return self.n + self.n_planes

    @property
    def spp(self) -> int:
        return self.samples * self.superSamples ** 2

    @property
    def n_sub(self) -> int:
        return self.width * self.height * self.superSamples ** 2

    def with_dims(self, width=None, height=None, samples=None, superSamples=None) -> "Scene":
        return replace(
            self,
            width=self.width if width is None else int(width),
            height=self.height if height is None else int(height),
            samples=self.samples if samples is None else int(samples),
            superSamples=self.superSamples if superSamples is None else int(superSamples),
        )

    def normalized(self) -> "Scene":
        """Arrays in the exact dtypes/layout the C ABI expects."""
        return replace(
            self,
            cam_origin=np.ascontiguousarray(self.cam_origin, np.float64).reshape(3),
            cam_target=np.ascontiguousarray(self.cam_target, np.float64).reshape(3),
            cam_up=np.ascontiguousarray(self.cam_up, np.float64).reshape(3),
            position=np.ascontiguousarray(self.position, np.float64).reshape(-1, 3),
            radius=np.ascontiguousarray(self.radius, np.float32).reshape(-1),
            color=np.ascontiguousarray(self.color, np.float64).reshape(-1, 3),
            emission=np.ascontiguousarray(self.emission, np.float64).reshape(-1, 3),
            pl_position=np.ascontiguousarray(self.pl_position, np.float64).reshape(-1, 3),
            pl_up=np.ascontiguousarray(self.pl_up, np.float64).reshape(-1, 3),
            pl_target=np.ascontiguousarray(self.pl_target, np.float64).reshape(-1, 3),
            pl_width=np.ascontiguousarray(self.pl_width, np.float64).reshape(-1),
            pl_color=np.ascontiguousarray(self.pl_color, np.float64).reshape(-1, 3),
            pl_emission=np.ascontiguousarray(self.pl_emission, np.float64).reshape(-1, 3),
            object_order=np.ascontiguousarray(self.object_order, np.int32).reshape(-1),
        )


class SceneError(ValueError):
    pass


def scene_from_dict(d: dict, compat: str = "abc") -> Scene:
    if compat not in ("abc", "head"):
        raise SceneError(f"compat must be 'abc' or 'head', got {compat!r}")

    def need(obj, key, what):
        if key not in obj:
            raise SceneError(f"{what}: missing key {key!r}")
        return obj[key]

    samples = d.get("00 samples", d.get("00 sample"))
    if samples is None:
        raise SceneError("missing '00 samples'")
    cam = need(d, "01 camera", "scene")
    pos, rad, col, emi = [], [], [], []
    ppos, pup, ptgt, pwid, pcol, pemi = [], [], [], [], [], []
    order_s, order_p = [], []
    k = 0
    for i, o in enumerate(d.get("02 scene", {}).get("00 object", [])):
        if not o:
            continue
        typ = o.get("00 objectType", 1)
        if typ == 2:
            if compat == "head":
                raise SceneError(f"object {i}: planes are an extension of the L-ABC level")
            mat = need(o, "02 material", f"object {i}")
            ppos.append([float(v) for v in need(o, "00 position", f"object {i}")])
            pup.append([float(v) for v in need(o, "01 up", f"object {i}")])
            ptgt.append([float(v) for v in need(o, "01 target", f"object {i}")])
            pwid.append(float(need(o, "01 width", f"object {i}")))
            pcol.append([float(v) for v in need(mat, "color", f"object {i} material")])
            pemi.append([float(v) for v in need(mat, "emission", f"object {i} material")])
            if not (pwid[-1] > 0):
                raise SceneError(f"object {i}: plane width must be positive")
            order_p.append(k)
            k += 1
            continue
        if typ != 1:
            raise SceneError(f"object {i}: objectType {typ} is neither a sphere (1) nor a plane (2) (the reference would crash)")
        order_s.append(k)
        k += 1
        p = need(o, "00 position", f"object {i}")
        mat = need(o, "02 material", f"object {i}")
        if compat == "head":
            p = [p[2], 0.0, 0.0]  # F2
        pos.append([float(v) for v in p])
        rad.append(float(need(o, "01 size", f"object {i}")))
        col.append([float(v) for v in need(mat, "color", f"object {i} material")])
        emi.append([float(v) for v in need(mat, "emission", f"object {i} material")])
    sc = Scene(
        width=int(need(d, "00 width", "scene")),
        height=int(need(d, "00 height", "scene")),
        samples=int(samples),
        superSamples=int(d.get("00 superSamples", 1)),
        cam_origin=np.array(need(cam, "origin", "camera"), np.float64),
        cam_target=np.array(need(cam, "target", "camera"), np.float64),
        cam_up=np.array(need(cam, "upVec", "camera"), np.float64),
        fov=float(np.float32(need(cam, "fov", "camera"))),
        position=np.array(pos, np.float64).reshape(-1, 3),
        radius=np.array(rad, np.float64).astype(np.float32),  # double size -> float m_size
        color=np.array(col, np.float64).reshape(-1, 3),
        emission=np.array(emi, np.float64).reshape(-1, 3),
        pl_position=np.array(ppos, np.float64).reshape(-1, 3), pl_up=np.array(pup, np.float64).reshape(-1, 3),
        pl_target=np.array(ptgt, np.float64).reshape(-1, 3), pl_width=np.array(pwid, np.float64).reshape(-1),
        pl_color=np.array(pcol, np.float64).reshape(-1, 3), pl_emission=np.array(pemi, np.float64).reshape(-1, 3),
        object_order=np.array(order_s + order_p, np.int32),
    )
    return sc.normalized()


def load_scene(path: str, compat: str = "abc") -> Scene:
    with open(path) as f:
        return scene_from_dict(json.load(f), compat)


def scene_to_dict(sc: Scene) -> dict:
    """HEAD-schema JSON (what LoadData::to_json would write if its scene part were
    not commented out, src/SettingData.cpp:106-128)."""
    objs = []
    for i in range(sc.n):
        objs.append({
            "00 objectType": 1,
            "00 position": [float(v) for v in sc.position[i]],
            "01 size": float(sc.radius[i]),
            "02 material": {"color": [float(v) for v in sc.color[i]], "emission": [float(v) for v in sc.emission[i]]},
        })
    for j in range(sc.n_planes):
        objs.append({
            "00 objectType": 2,
            "00 position": [float(v) for v in sc.pl_position[j]],
            "01 up": [float(v) for v in sc.pl_up[j]],
            "01 target": [float(v) for v in sc.pl_target[j]],
            "01 width": float(sc.pl_width[j]),
            "02 material": {"color": [float(v) for v in sc.pl_color[j]], "emission": [float(v) for v in sc.pl_emission[j]]},
        })
    return {
        "00 width": sc.width, "00 height": sc.height, "00 samples": sc.samples, "00 superSamples": sc.superSamples,
        "01 camera": {"origin": [float(v) for v in sc.cam_origin], "target": [float(v) for v in sc.cam_target],
                      "upVec": [float(v) for v in sc.cam_up], "fov": float(sc.fov)},
        "02 scene": {"00 object": objs},
    }


def random_scene(seed: int, n: int, width=64, height=48, samples=4, superSamples=2, closed=True) -> Scene:
    """Small random scene for parity tests: optional closed room of six big
    spheres plus n random small ones; radii exact in float; kd <= 0.9."""
    rng = np.random.default_rng(seed)
    pos, rad, col, emi = [], [], [], []
    if closed:
        for axis in range(3):
            for sgn in (1, -1):
                c = [0.0, 0.0, 0.0]
                c[axis] = sgn * 1010.0
                pos.append(c)
                rad.append(1000.0)
                col.append(list(rng.uniform(0.2, 0.9, 3)))
                emi.append([0.0, 0.0, 0.0])
    for i in range(n):
        pos.append(list(rng.uniform(-7, 7, 3)))
        rad.append(float(np.round(rng.uniform(0.3, 2.0) * 256) / 256))
        col.append(list(rng.uniform(0.1, 0.9, 3)))
        e = float(rng.uniform(2, 12)) if (i % 4 == 0) else 0.0
        emi.append([e, e, e])
    return Scene(
        width=width, height=height, samples=samples, superSamples=superSamples,
        cam_origin=np.array([0.5, 0.25, -9.0]), cam_target=np.array([0.0, 0.0, 0.0]), cam_up=np.array([0.0, 1.0, 0.0]),
        fov=float(np.float32(0.8)),
        position=np.array(pos).reshape(-1, 3), radius=np.array(rad, np.float32), color=np.array(col).reshape(-1, 3),
        emission=np.array(emi).reshape(-1, 3),
    ).normalized()
