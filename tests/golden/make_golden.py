#!/usr/bin/env python3
"""Generates tests/golden/*.npz from the reference's OWN object code
(oracle/_ref/libref_{head,abc}.so = /root/reference/src compiled unmodified,
resp. with the 8-line L-ABC patch; see oracle/Makefile).  Run in the build
container after `make -C oracle`; the fixtures are committed because the GPU box
has no /root/reference.

Per scene (the four shipped scene files as HEAD-schema copies, scenes/*.head.json,
parsed by the reference's own LoadData) and per level (head / abc):
  hits_small   int8  [34][64][2][2]   primary-hit map at 64x34, ss=2, through png::PathTracing (mode 1)
  hist_480     int64 [n+1]            histogram (first entry = misses) at 480x252, ss=2
  image_small  f64   [34][64][3]      linear image at 64x34, ss=2, samples=3, mt19937(1234+row) per row
  rays_small   int64                  PathTracing invocations of that render
  parsed_*                            what the reference's parser produced (exposes F2 at HEAD)
plus known-answer ray/sphere cases through SphereObject::Intersect.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)

import oracle_lib as O  # noqa: E402

SCENES = ["settingData", "cornellBox", "simple1", "simple2"]


def main():
    out = {}
    for name in SCENES:
        path = os.path.join(ROOT, "scenes", name + ".head.json")
        for level in ("head", "abc"):
            rs = O.RefScene(level, json_path=path)
            cam9, fov, pos, rad, col, emi = rs.get()
            k = f"{name}.{level}."
            out[k + "parsed_cam9"], out[k + "parsed_fov"] = cam9, np.float32(fov)
            out[k + "parsed_pos"], out[k + "parsed_rad"], out[k + "parsed_col"], out[k + "parsed_emi"] = pos, rad, col, emi
            out[k + "parsed_dims"] = np.array([rs.width, rs.height, rs.samples, rs.superSamples])
            rs.set_dims(480, 252, 1, 2)
            h = rs.primary_hits(1)
            assert np.array_equal(h, rs.primary_hits(0))
            out[k + "hist_480"] = np.bincount(h.ravel() + 1, minlength=rs.n + 1).astype(np.int64)
            rs.set_dims(64, 34, 3, 2)
            hs = rs.primary_hits(1)
            out[k + "hits_small"] = hs.astype(np.int8)
            img, rays, samples = rs.render(1234)
            out[k + "image_small"], out[k + "rays_small"] = img, np.int64(rays)
            rs.close()
    # known-answer ray/sphere tests on the reference's SphereObject::Intersect (L-ABC build returns the normal)
    sc_path = os.path.join(ROOT, "scenes", "simple2.head.json")
    rs = O.RefScene("abc", json_path=sc_path)
    cases = []
    def case(obj, org, d):
        d = np.array(d, float)
        hit, t, n = rs.intersect(obj, np.array(org, float), d)
        cases.append([obj, *org, *d, float(hit), t if hit else 0.0, *n])
    case(0, [0, 0, -10], [0, 0, 1])            # through the centre of the unit sphere
    case(0, [1, 0, -10], [0, 0, 1])            # tangent
    case(0, [1.0000001, 0, -10], [0, 0, 1])    # just missing
    case(0, [0, 0, 0], [0, 0, 1])              # origin at the centre (inside)
    case(4, [0, 0, -10], [0, 0, 1])            # camera ON the r=10 sphere: t1~0 -> far root
    case(4, [0, 0, -10], [0, 0, -1])           # on the surface, pointing away
    case(0, [0, 0, -1.0005], [0, 0, 1])        # t1 = 5e-4 in (1e-5, 1e-3]: skipped in favour of t2
    case(0, [0, 0, -1.002], [0, 0, 1])         # t1 = 2e-3 > 1e-3: taken
    case(0, [0, 0, 5], [0, 0, 1])              # sphere behind the ray
    case(3, [0.3, -2, 1], [0.1, 0.7, 0.7])     # non-unit direction (no 'a' term in the formula)
    case(1, [3, 0, -0.999995], [0, 0, -1])     # inside, 5e-6 from the surface: both roots < 1e-5? no: t2 tiny
    out["intersect_cases"] = np.array(cases)
    rs.close()
    np.savez_compressed(os.path.join(HERE, "reference_golden.npz"), **out)
    print("wrote reference_golden.npz with", len(out), "arrays")


if __name__ == "__main__":
    main()
