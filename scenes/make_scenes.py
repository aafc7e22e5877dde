#!/usr/bin/env python3
"""Generates the scene fixtures under scenes/ (run in the build container, where
/root/reference exists; the outputs are committed because the GPU box has no
/root/reference).

  *.head.json   HEAD-schema copies of the reference's four shipped scene files
                (SURVEY.md §5.1): the reference's parser at HEAD aborts on every
                shipped file (src/SettingData.cpp:162 reads "00 objectType",
                which none of them has), so the oracle needs shimmed copies:
                  - "00 objectType": 1 added to every object,
                  - empty objects `{}` dropped (cornellBoxSetting.json:47),
                  - legacy "00 sample" renamed to "00 samples",
                  - missing "00 superSamples" set to 1.
                Geometry, materials and camera are untouched.
  synth10k.json 10 000-sphere stress scene in the same schema.
"""
import json
import os
import sys

import numpy as np

REF = os.environ.get("RTM_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))

SHIPPED = {
    "settingData": "settingData.json",
    "cornellBox": "ExampleScene/cornellBoxSetting.json",
    "simple1": "ExampleScene/simpleSetting1.json",
    "simple2": "ExampleScene/simpleSetting2.json",
}


def shim(d):
    out = {}
    for k, v in d.items():
        if k == "00 sample":
            out["00 samples"] = v
        elif k == "02 scene":
            objs = []
            for o in v.get("00 object", []):
                if not o:
                    continue
                o = {kk: vv for kk, vv in o.items() if kk != "comment"}
                o.setdefault("00 objectType", 1)
                objs.append(o)
            out[k] = {"00 object": objs}
        else:
            out[k] = v
    out.setdefault("00 superSamples", 1)
    return out


def synth10k(cornell, seed=12345, n_total=10000):
    rng = np.random.default_rng(seed)
    objs = list(cornell["02 scene"]["00 object"])
    n = n_total - len(objs)
    centre = np.stack([rng.uniform(-9, 9, n), rng.uniform(-9, 9, n), rng.uniform(-6, 9, n)], axis=1)
    radius = np.exp(rng.uniform(np.log(0.05), np.log(0.30), n))
    radius = np.round(radius * 1024.0) / 1024.0  # exact in float (m_size is float)
    colour = rng.uniform(0.1, 0.9, (n, 3))
    emitter = rng.uniform(0, 1, n) < 0.01
    emis = rng.uniform(5, 20, n)
    for i in range(n):
        e = float(round(emis[i], 3)) if emitter[i] else 0.0
        objs.append({
            "00 objectType": 1,
            "00 position": [float(round(c, 4)) for c in centre[i]],
            "01 size": float(radius[i]),
            "02 material": {"color": [float(round(c, 4)) for c in colour[i]], "emission": [e, e, e]},
        })
    out = dict(cornell)
    out["00 width"], out["00 height"], out["00 samples"], out["00 superSamples"] = 1920, 1080, 256, 1
    out["02 scene"] = {"00 object": objs}
    return out


def main():
    made = {}
    for name, rel in SHIPPED.items():
        with open(os.path.join(REF, rel)) as f:
            d = shim(json.load(f))
        made[name] = d
        with open(os.path.join(HERE, name + ".head.json"), "w") as f:
            json.dump(d, f, indent=1)
            f.write("\n")
    with open(os.path.join(HERE, "synth10k.json"), "w") as f:
        json.dump(synth10k(made["cornellBox"]), f, separators=(",", ":"))
        f.write("\n")
    print("wrote", sorted(os.listdir(HERE)))


if __name__ == "__main__":
    sys.exit(main())
