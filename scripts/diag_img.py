import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import raytracingmin_b200 as rtm
synth = rtm.load_scene(os.path.join(ROOT, "scenes/synth10k.json"))
sc = synth.with_dims(96, 54, 2048, 1)
out = {}
for prec in ("f32", "f64"):
    with rtm.Renderer(sc, precision=prec, seed=11, max_depth=2) as r:
        res = r.render(want_hits=True)
    out[prec] = res.image.astype(np.float64)
    hits = res.hits[:, :, 0, 0]
a, b = out["f32"], out["f64"]
em = synth.emission[np.maximum(hits, 0)].sum(axis=-1) > 0
kd = synth.color[np.maximum(hits, 0)].max(axis=-1)
for name, m in (("primary on wall", (hits < 7) & (hits > 0)), ("primary on light", hits == 0), ("primary on small", hits >= 7),
                ("small emitter", (hits >= 7) & em), ("small non-emitter", (hits >= 7) & ~em)):
    if m.sum():
        print(name, m.sum(), "f32", a[m].mean(), "f64", b[m].mean(), "ratio", a[m].mean() / b[m].mean())
m = (hits >= 7) & ~em
rad = synth.radius[np.maximum(hits, 0)]
for lo, hi in ((0, 0.08), (0.08, 0.15), (0.15, 0.31)):
    mm = m & (rad >= lo) & (rad < hi)
    print("small non-emitter r in", lo, hi, mm.sum(), "ratio", a[mm].mean() / b[mm].mean())
z = synth.position[np.maximum(hits, 0)][..., 2]
for lo, hi in ((-6, -2), (-2, 3), (3, 9.1)):
    mm = m & (z >= lo) & (z < hi)
    print("small non-emitter z in", lo, hi, mm.sum(), "ratio", a[mm].mean() / b[mm].mean())
