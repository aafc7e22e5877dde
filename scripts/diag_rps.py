import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import raytracingmin_b200 as rtm
import oracle_lib as O
g = np.load(os.path.join(ROOT, "tests/golden/synth10k_golden.npz"))
print("ref rps", float(g["rays_48"]) / float(g["samples_48"]), "mean", g["image_48"].mean(axis=(0, 1)))
sc = rtm.load_scene(os.path.join(ROOT, "scenes/synth10k.json")).with_dims(48, 27, 1024, 1)
for prec in ("f32", "f64"):
    for md in (0, 2, 3):
        with rtm.Renderer(sc, precision=prec, seed=11, max_depth=md) as r:
            res = r.render()
        print("synth10k", prec, "max_depth", md, "rps", res.stats.rays / res.stats.samples, "mean", res.image.reshape(-1, 3).mean(axis=0),
              "tests/ray", res.stats.sphere_tests / res.stats.rays_traced)
for n in (40, 200):
    s2 = rtm.random_scene(21, n, width=48, height=27, samples=256, superSamples=1)
    ref, rays, samples = O.oracle_render(s2, O.LEVEL_ABC, 5)
    print("random", n, "oracle rps", rays / samples, "mean", ref.reshape(-1, 3).mean(axis=0))
    for prec, accel in (("f32", "brute"), ("f32", "bvh"), ("f64", "brute"), ("f64", "bvh")):
        with rtm.Renderer(s2.with_dims(samples=1024), precision=prec, accel=accel, seed=3) as r:
            res = r.render()
        print("   ", prec, accel, "rps", res.stats.rays / res.stats.samples, "mean", res.image.reshape(-1, 3).mean(axis=0))
