#!/usr/bin/env python3
"""Generates tests/golden/synth10k_golden.npz — known answers for config C5:
the 10 000-sphere scene scenes/synth10k.json, whose closest hit the CUDA path finds through
a hierarchy while the reference walks all N spheres (src/Renderer.cpp:58-73).

  hits_480     int16 [270][480]   primary-hit map at 480x270, ss=1, from the REFERENCE'S OWN
                                  objects (oracle/_ref/libref_abc.so: png::PathTracing over
                                  SphereObject::Intersect), cross-checked against the C oracle
  hist_1080    int64 [n+1]        histogram of the full 1920x1080 ss=1 map (first entry = misses)
  sha_1080     bytes              sha256 of that map as little-endian int32 [1080][1920][1][1]
  hist_1080 / sha_1080 come from the C oracle (oracle/rtm_oracle.c), which the 480x270 run
  pins against the reference objects on this very scene.
  image_48     f64 [27][48][3]    L-ABC linear image at 48x27, 512 spp, reference objects,
                                  per-row mt19937(1234+row)
  rays_48, samples_48             PathTracing invocations / samples of that render

Run in the build container after `make -C oracle` (needs /root/reference for oracle/_ref).
About 5 minutes on 8 cores.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)

import oracle_lib as O  # noqa: E402
import raytracingmin_b200 as rtm  # noqa: E402  (scene loader only)


def main():
    sc = rtm.load_scene(os.path.join(ROOT, "scenes", "synth10k.json"), "abc")
    assert sc.n == 10000
    out = {}
    rs = O.RefScene("abc", sc.with_dims(480, 270, 1, 1))
    h_ref = rs.primary_hits(1)
    rs.close()
    h_orc = O.oracle_primary_hits(sc.with_dims(480, 270, 1, 1))
    assert np.array_equal(h_ref, h_orc), "C oracle differs from the reference objects on synth10k"
    out["hits_480"] = h_ref[:, :, 0, 0].astype(np.int16)

    full = O.oracle_primary_hits(sc.with_dims(1920, 1080, 1, 1))
    out["hist_1080"] = np.bincount(full.ravel() + 1, minlength=sc.n + 1).astype(np.int64)
    out["sha_1080"] = np.frombuffer(hashlib.sha256(np.ascontiguousarray(full, "<i4").tobytes()).digest(), np.uint8)

    rs = O.RefScene("abc", sc.with_dims(48, 27, 512, 1))
    img, rays, samples = rs.render(1234)
    rs.close()
    out["image_48"], out["rays_48"], out["samples_48"] = img, np.int64(rays), np.int64(samples)
    np.savez_compressed(os.path.join(HERE, "synth10k_golden.npz"), **out)
    print("wrote synth10k_golden.npz:", {k: getattr(v, "shape", v) for k, v in out.items()})


if __name__ == "__main__":
    sys.exit(main())
