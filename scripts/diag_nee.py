#!/usr/bin/env python3
"""NEE diagnostics: reference vs nee means on variations of a random scene."""
import os, sys
from dataclasses import replace
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import raytracingmin_b200 as rtm

def means(sc, tag):
    out = []
    for est, seed in (("reference", 1), ("nee", 2), ("nee", 3)):
        with rtm.Renderer(sc, estimator=est, seed=seed) as r:
            res = r.render()
        out.append(res.image.astype(np.float64).mean())
    print(f"{tag:50s} ref {out[0]:.5f} nee {out[1]:.5f} {out[2]:.5f}  ratio {out[1]/out[0]:.4f}")

base = rtm.random_scene(17, 12, 64, 36, 2048, 1)
means(base, "base n=18, 3 diffuse lights")
emi = base.emission.copy(); lights = np.nonzero(emi[:, 0] > 0)[0]
for keep in lights:
    e = np.zeros_like(emi); e[keep] = emi[keep]
    means(replace(base, emission=e).normalized(), f"only light {keep} r={base.radius[keep]:.2f}")
col = base.color.copy(); col[lights] = 0.0
means(replace(base, color=col).normalized(), "3 lights, kd=0")
# fewer spheres: packed-pair path (n <= 8)
idx = [0, 1, 2, 3, 4, 5] + [int(lights[0]), int(lights[1])]
small = replace(base, position=base.position[idx], radius=base.radius[idx], color=base.color[idx], emission=base.emission[idx]).normalized()
means(small, "n=8 (walls + 2 lights), packed pairs")
idx9 = idx + [7]
s9 = replace(base, position=base.position[idx9], radius=base.radius[idx9], color=base.color[idx9], emission=base.emission[idx9]).normalized()
means(s9, "n=9 (walls + 2 lights + 1), loop")
