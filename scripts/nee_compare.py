#!/usr/bin/env python3
"""Equal-quality comparison of the two estimators on the Cornell box (C2 size): device time and RMSE against a
16k-spp reference-estimator image, at several sample counts.  Prints one JSON line per (estimator, spp)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import raytracingmin_b200 as rtm  # noqa: E402

for scene, tag in (("cornellBox.head.json", "cornellBox"), ("cornellBoxPlanes.json", "cornellBoxPlanes")):
    sc = rtm.load_scene(os.path.join(ROOT, "scenes", scene), "abc").with_dims(960, 504, 25, 4)
    with rtm.Renderer(sc.with_dims(samples=1024), seed=99) as r:
        ref = r.render().image.astype(np.float64)
    for est in ("reference", "nee"):
        for S in (1, 4, 25, 100):
            with rtm.Renderer(sc.with_dims(samples=S), estimator=est, seed=7) as r:
                r.render()                       # warm-up
                res = r.render()
            img = res.image.astype(np.float64)
            rmse = float(np.sqrt(((img - ref) ** 2).mean()))
            print(json.dumps({"scene": tag, "estimator": est, "spp": S * 16, "device_ms": round(res.stats.device_ms, 3),
                              "rmse": round(rmse, 5), "mean": round(float(img.mean()), 5),
                              "rays_per_sample": round(res.stats.rays / res.stats.samples, 3),
                              "traced_per_sample": round(res.stats.rays_traced / res.stats.samples, 3)}))
