#!/bin/bash
# compute-sanitizer on the smallest configuration for the three kernel organisations.
# memcheck: every access in bounds / aligned; racecheck: shared-memory hazards (traversal stacks, per-lane primary records);
# synccheck: barrier misuse.  Logs -> gpurun_out/sanitize_*.log
# NOTE: the GPU pool this round ran on refuses compute-sanitizer (rc 86, profiles/r02/compute_sanitizer_closed.log); the
# evidence committed instead comes from scripts/gpu_checks.sh (device-side invariants).  Kept for pools where the tool runs.
mkdir -p gpurun_out
S=/usr/local/cuda/bin/compute-sanitizer
cat > /tmp/san_run.py <<'PY'
import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np
import raytracingmin_b200 as rtm
which = sys.argv[1]
if which == "mega":
    sc, kw = rtm.load_scene("scenes/cornellBox.head.json", "abc").with_dims(96, 54, 4, 2), {}
elif which == "mega_rt":
    sc, kw = rtm.load_scene("scenes/simple2.head.json", "abc").with_dims(96, 54, 8, 1), {}
elif which == "bvh":
    sc, kw = rtm.load_scene("scenes/synth10k.json", "abc").with_dims(64, 36, 4, 1), {}
elif which == "bvh_float_local":
    sc, kw = rtm.load_scene("scenes/synth10k.json", "abc").with_dims(64, 36, 4, 1), dict(debug_bvh_walk=3)
elif which == "wavefront":
    sc, kw = rtm.load_scene("scenes/cornellBox.head.json", "abc").with_dims(64, 36, 2, 2), dict(variant="wavefront")
elif which == "planes_nee":
    sc, kw = rtm.load_scene("scenes/cornellBoxPlanes.json").with_dims(64, 36, 4, 2), dict(estimator="nee")
elif which == "head":
    sc, kw = rtm.load_scene("scenes/cornellBox.head.json", "head").with_dims(64, 36, 2, 2), dict(compat="head")
with rtm.Renderer(sc, **kw) as r:
    res = r.render(want_hits=True)
    u8 = r.read_u8()
print(which, "ok", float(res.image.mean()), res.stats.rays)
PY
for tool in memcheck racecheck synccheck; do
  for w in mega mega_rt bvh bvh_float_local wavefront planes_nee head; do
    timeout 600 $S --tool $tool --error-exitcode 9 python /tmp/san_run.py $w > gpurun_out/sanitize_${tool}_$w.log 2>&1
    echo "$tool $w rc=$? $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' gpurun_out/sanitize_${tool}_$w.log | tail -1)"
  done
done
