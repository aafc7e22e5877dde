import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import raytracingmin_b200 as rtm
synth = rtm.load_scene(os.path.join(ROOT, "scenes/synth10k.json"))
rng = np.random.default_rng(2)
n = 200000
sc = synth.with_dims(16, 9, 1, 1)
org0 = np.tile(synth.cam_origin, (n, 1))
d0 = rng.normal(size=(n, 3)); d0[:, 2] = np.abs(d0[:, 2]) + 0.5
d0 /= np.linalg.norm(d0, axis=1, keepdims=True)
with rtm.Renderer(sc) as r:
    hit0, t0 = r.closest_hit(org0, d0, "f64")
    p = org0 + d0 * t0[:, None]
    nrm = p - synth.position[hit0]
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    nrm[np.einsum("ij,ij->i", nrm, d0) > 0] *= -1
    s = rng.normal(size=(n, 3)); s /= np.linalg.norm(s, axis=1, keepdims=True)
    d1 = nrm + s; d1 /= np.linalg.norm(d1, axis=1, keepdims=True)      # cosine weighted
    h64, t64 = r.closest_hit(p, d1, "f64")
    h32, t32 = r.closest_hit(p, d1, "f32", origin_id=hit0)
dis = h64 != h32
print("disagree", dis.mean(), "n", dis.sum())
print("origin small:", (hit0 >= 7).mean(), " disagree | origin small", dis[hit0 >= 7].mean(), " | origin big", dis[hit0 < 7].mean())
k = np.where(dis)[0]
print("f64 hits self:", (h64[k] == hit0[k]).mean(), "f32 hits self:", (h32[k] == hit0[k]).mean())
print("f64 small:", (h64[k] >= 7).mean(), "f32 small:", (h32[k] >= 7).mean(), "f32 miss", (h32[k] < 0).mean(), "f64 miss", (h64[k] < 0).mean())
print("t64 median", np.median(t64[k]), "t32 median", np.median(t32[k]), " t32>t64:", (t32[k] > t64[k]).mean())
cosn = np.einsum("ij,ij->i", nrm, d1)
print("cos at disagreements: median", np.median(cosn[k]), " overall", np.median(cosn))
for i in k[:25]:
    print(i, "org", hit0[i], "f64", h64[i], t64[i], "f32", h32[i], t32[i], "cos", cosn[i])
