#!/usr/bin/env python3
"""Generates tests/golden/band_<cfg>.npz — the north-star radiance bar ("per-channel RMSE of the GPU image against a
high-spp reference render stays within the reference's own seed-to-seed noise band at equal spp") for every BASELINE
config, from the REFERENCE'S OWN objects (oracle/_ref/libref_abc.so: png::PathTracing over SphereObject::Intersect,
src/Renderer.cpp:57-117 accumulated as :223-248, a private mt19937(seed + row) per image row):

  hi            f16 [h][w][3]   linear image at `hi_mult` x the config's samples (seed 424242).  float16 keeps the
                                fixture small; its rounding (2^-11 relative, unbiased) is 100x below the noise compared
  band_rmse     f64 [K][3]      per-channel RMSE against `hi` (as stored, i.e. after the f16 rounding) of K = 8
                                independent renders at the config's spp (seeds 10007*k) = the seed-to-seed band
  band_mean     f64 [K][3]      per-channel mean signed error of those renders
  rays_per_sample               of the high-spp render
  window        i32 [4]         x0, y0, w, h of the evaluated window inside the full frame
  dims          i32 [4]         full-frame width, height, samples, superSamples of the config
  hi_mult, seconds

What is evaluated (and what is reduced, and why):
  c1  simpleSetting1 960x504 x 100 spp              full frame, hi = 16x
  c2  cornellBox     960x504 x 25 x 4x4 = 400 spp   full frame, hi = 16x (6400 spp)
  c3  simpleSetting2 1920x1080 x 256 spp            the centred 640x360 window of the full frame (same camera rays,
                                                    1/9 of the pixels: the fixture stays at 1.4 MB), hi = 16x
  c4  cornellBox     3840x2160 x 64 x 4x4 = 1024    the 960x540 window at (540, 180) — 1/16 of the area: left wall,
                                                    ceiling, the rim of the light and a corner of the back wall — at
                                                    the config's REAL sampling; hi = 8x (8192 spp; 16x would be 25 more
                                                    minutes on 8 cores)
  c5  synth10k       1920x1080 x 256 spp            the frame at 64x36 (every 30th pixel's ray, whole scene in view),
                                                    256 spp; hi = 8x.  The reference walks 10^4 spheres per ray: one
                                                    64x36x256 render is ~2 min on 8 cores, the full frame 45 hours
Run in the build container after `make -C oracle` (needs /root/reference for oracle/_ref):
    python tests/golden/make_golden_bands.py c1 c2 c3 c4 c5      (~75 minutes on 8 cores; c4 and c5 dominate)
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)

import oracle_lib as O  # noqa: E402
import raytracingmin_b200 as rtm  # noqa: E402  (scene loader only)

CONFIGS = {
    #      scene file              W     H     S   ss  window (x0,y0,w,h) or None   hi_mult
    "c1": ("simple1.head.json",    960,  504,  100, 1, None,                         16),
    "c2": ("cornellBox.head.json", 960,  504,  25,  4, None,                         16),
    "c3": ("simple2.head.json",    1920, 1080, 256, 1, (640, 360, 640, 360),         16),
    "c4": ("cornellBox.head.json", 3840, 2160, 64,  4, (540, 180, 960, 540),         8),
    "c5": ("synth10k.json",        64,   36,   256, 1, None,                         8),
}
K = 8


def rmse(a, b):
    return np.sqrt(((a - b) ** 2).reshape(-1, 3).mean(axis=0))


def make(cfg, threads=0):
    fname, W, H, S, ss, win, mult = CONFIGS[cfg]
    sc = rtm.load_scene(os.path.join(ROOT, "scenes", fname), "abc").with_dims(W, H, S, ss)
    x0, y0, w, h = win if win else (0, 0, W, H)
    t0 = time.time()
    rs = O.RefScene("abc", sc.with_dims(samples=S * mult))
    hi, rays, samples = rs.render_window(424242, x0, y0, w, h, threads)
    rs.close()
    hi16 = hi.astype(np.float16)
    assert np.isfinite(hi16).all()
    hi_q = hi16.astype(np.float64)
    print(f"{cfg}: hi done in {time.time() - t0:.0f} s, rays/sample {rays / samples:.4f}", flush=True)
    band, bias = [], []
    for k in range(1, K + 1):
        rs = O.RefScene("abc", sc)
        img, _, _ = rs.render_window(10007 * k, x0, y0, w, h, threads)
        rs.close()
        band.append(rmse(img, hi_q))
        bias.append((img - hi_q).reshape(-1, 3).mean(axis=0))
        print(f"{cfg}: seed {k}/{K} rmse {band[-1]}", flush=True)
    out = os.path.join(HERE, f"band_{cfg}.npz")
    np.savez_compressed(out, hi=hi16, band_rmse=np.array(band), band_mean=np.array(bias),
                        rays_per_sample=np.float64(rays / samples), window=np.array([x0, y0, w, h], np.int32),
                        dims=np.array([W, H, S, ss], np.int32), hi_mult=np.int32(mult), seconds=np.float64(time.time() - t0))
    print(f"{cfg}: wrote {out} ({os.path.getsize(out) / 1e6:.2f} MB) in {time.time() - t0:.0f} s; band",
          np.array(band).mean(axis=0), "+-", np.array(band).std(axis=0), flush=True)


if __name__ == "__main__":
    names = [a for a in sys.argv[1:] if a in CONFIGS] or list(CONFIGS)
    thr = int(os.environ.get("BAND_THREADS", "0"))
    for c in names:
        make(c, thr)
