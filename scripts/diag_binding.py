import json, os, subprocess, sys, tempfile
import numpy as np
from PIL import Image
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")
name = sys.argv[1] if len(sys.argv) > 1 else "simple1"
d = json.load(open(os.path.join(ROOT, "scenes", name + ".head.json")))
d["00 width"], d["00 height"], d["00 samples"], d["00 superSamples"] = 160, 84, 3, 2
tmp = tempfile.mkdtemp()
p = os.path.join(tmp, "scene.json"); open(p, "w").write(json.dumps(d))
out = {}
for tag, exe, env in (("cpu", "ref_head_cli", {}), ("gpu", "ref_head_b200_cli", {"RTM_COMPAT": "head"})):
    cwd = os.path.join(tmp, tag); os.makedirs(cwd)
    e = dict(os.environ); e.update(env)
    r = subprocess.run([os.path.join(REF, exe), "-json", p], cwd=cwd, capture_output=True, env=e)
    print(tag, r.returncode, r.stderr[-200:])
    out[tag] = np.asarray(Image.open(os.path.join(cwd, "result.bmp")).convert("RGB")).astype(int)
a, b = out["cpu"], out["gpu"]
diff = np.argwhere(a != b)
print("differing entries", len(diff), "of", a.size, "unique cpu", np.unique(a)[:10], "unique gpu", np.unique(b)[:10])
for y, x, c in diff[:10]:
    print(y, x, c, a[y, x, c], b[y, x, c])
