#!/usr/bin/env python3
"""One-shot GPU probe of the experimental 4-wide walk (RTM_BVH_WIDE=1), sized for a few seconds:
(1) a small synth10k render must be bit-identical between the default two-child walk and the wide walk,
(2) one C5-size frame each, device-timed.  Child processes because the switch is read at renderer_create from the
environment.  Appends JSON lines to gpurun_out/wide_probe.jsonl as it goes."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out", "wide_probe.jsonl")

CHILD = r"""
import os, sys, json, hashlib, numpy as np
sys.path.insert(0, %r)
import raytracingmin_b200 as rtm
synth = rtm.load_scene(os.path.join(%r, 'scenes', 'synth10k.json'), 'abc')
out = {'wide': os.environ.get('RTM_BVH_WIDE', '0')}
with rtm.Renderer(synth.with_dims(64, 36, 8, 2), seed=1) as r:
    res = r.render()
out['small'] = {'sha': hashlib.sha256(np.ascontiguousarray(res.image).tobytes()).hexdigest(), 'rays': int(res.stats.rays),
                'tests': int(res.stats.sphere_tests), 'nodes': int(res.stats.bvh_nodes), 'mean': float(res.image.mean())}
print(json.dumps(out), flush=True)
with rtm.Renderer(synth.with_dims(1920, 1080, 256, 1), seed=0x5EED) as r:
    r.render(want_image=False)
    res = r.render(want_image=False)
out['c5'] = {'device_ms': float(res.stats.device_ms), 'rays': int(res.stats.rays), 'grays_per_s': res.stats.rays / res.stats.device_ms / 1e6,
             'nodes_per_ray': res.stats.bvh_nodes / max(1, res.stats.rays_traced), 'tests_per_ray': res.stats.sphere_tests / max(1, res.stats.rays_traced)}
print(json.dumps(out), flush=True)
""" % (ROOT, ROOT)


def main():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    procs = []
    for wide in ("1", "0"):
        env = dict(os.environ, RTM_BVH_WIDE=wide)
        p = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True, timeout=60)
        with open(OUT, "a") as f:
            for ln in p.stdout.splitlines():
                if ln.startswith("{"):
                    f.write(ln + "\n")
            if p.returncode != 0:
                f.write(json.dumps({"wide": wide, "rc": p.returncode, "stderr": p.stderr[-600:]}) + "\n")
        print(wide, p.returncode, p.stdout.strip().splitlines()[-1][:400] if p.stdout.strip() else p.stderr[-300:], flush=True)


if __name__ == "__main__":
    main()
