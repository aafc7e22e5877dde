#!/usr/bin/env python3
"""Generates tests/golden/c2_quarter_golden.npz — the radiance bar of BASELINE.json's config 2 (Cornell at its
default sampling: 25 samples x 4x4 sub-pixels = 400 spp) from the REFERENCE'S OWN objects
(oracle/_ref/libref_abc.so: png::PathTracing, per-row mt19937), at a quarter of the default resolution so that the
fixture stays small (240x126; the full 960x504 reference render at 6400 spp would take 35 minutes here):

  hi          f32 [126][240][3]  linear image at 16 x the test spp (400 samples x 4x4 = 6400 spp), seed 424242
  band_rmse   f64 [8][3]         per-channel RMSE against `hi` of K = 8 independent renders at the test spp
                                 (seeds 10007*k: disjoint per-row generator sets) = the reference's seed-to-seed band
  band_mean   f64 [8][3]         per-channel mean signed error of those renders
  rays_per_sample                of the 6400-spp render
Run in the build container after `make -C oracle`.  About 5 minutes on 8 cores.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)

import oracle_lib as O  # noqa: E402
import raytracingmin_b200 as rtm  # noqa: E402


def rmse(a, b):
    return np.sqrt(((a - b) ** 2).reshape(-1, 3).mean(axis=0))


def main():
    sc = rtm.load_scene(os.path.join(ROOT, "scenes", "cornellBox.head.json"), "abc").with_dims(240, 126, 25, 4)
    rs = O.RefScene("abc", sc.with_dims(samples=400))
    hi, rays, samples = rs.render(424242)
    rs.close()
    band, bias = [], []
    for k in range(1, 9):
        rs = O.RefScene("abc", sc)
        img, _, _ = rs.render(10007 * k)
        rs.close()
        band.append(rmse(img, hi))
        bias.append((img - hi).reshape(-1, 3).mean(axis=0))
    np.savez_compressed(os.path.join(HERE, "c2_quarter_golden.npz"), hi=hi.astype(np.float32), band_rmse=np.array(band),
                        band_mean=np.array(bias), rays_per_sample=np.float64(rays / samples))
    print("band rmse", np.array(band).mean(axis=0), "+-", np.array(band).std(axis=0), "rps", rays / samples)


if __name__ == "__main__":
    sys.exit(main())
