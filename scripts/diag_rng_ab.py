import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import raytracingmin_b200 as rtm
tag = sys.argv[1]
for name, dims in (("simple1", (64, 34, 512, 1)), ("simple2", (64, 34, 512, 1)), ("cornellBox", (64, 34, 32, 2))):
    sc = rtm.load_scene(os.path.join(ROOT, "scenes", name + ".head.json")).with_dims(*dims)
    ms, imgs = [], []
    for seed in range(1, 49):
        with rtm.Renderer(sc, seed=seed * 7919) as r:
            res = r.render()
        ms.append(res.image.mean()); imgs.append(res.image.astype(np.float64))
    imgs = np.array(imgs)
    pix_var = imgs.var(axis=0).mean()          # per-pixel variance across seeds
    mean_var = np.var(ms)                      # variance of the image mean: inflated by inter-pixel correlation
    npx = imgs.shape[1] * imgs.shape[2] * 3
    print(tag, name, "mean %.5f  std(mean) %.5f  per-pixel std %.5f  var(mean)*Npix/pixvar %.2f" %
          (np.mean(ms), np.std(ms), np.sqrt(pix_var), mean_var * npx / pix_var))
