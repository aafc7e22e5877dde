#!/usr/bin/env python3
"""Design study for DESIGN.md §12 item 1 (no GPU, no product code): how many DEPENDENT node fetches per ray would a
4-wide hierarchy with 8-bit quantised child boxes need on the 10 000-sphere scene (config C5), against the two-child
tree the kernels walk today?

Takes the tree bvh_build.cpp builds (api.build_bvh), collapses it to 4 children per node (the standard greedy
collapse: open the child with the largest box area until the node holds 4), optionally snaps each child box outward
to a 256-step grid of its parent's box, and walks both trees for the same bounce-like rays (origins on sphere
surfaces seen from the camera, cosine-ish random directions), counting per ray:
    fetches  = inner nodes visited = dependent memory round trips (what the latency-bound kernel pays for),
    boxes    = slab tests,
    spheres  = leaf sphere tests,
with the closest hit checked equal between all variants.  Output: one JSON line (profiles/r01/study_bvh4.json).
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as O          # noqa: E402  (study tooling, not product)
import raytracingmin_b200 as rtm  # noqa: E402
from raytracingmin_b200 import api  # noqa: E402

NONE = 0x7FFFFFFF


def children2(nd):
    c = nd[3].view(np.int32)
    b0 = np.array([nd[0][0], nd[0][2], nd[2][0], nd[0][1], nd[0][3], nd[2][1]], np.float64)   # lo xyz, hi xyz
    b1 = np.array([nd[1][0], nd[1][2], nd[2][2], nd[1][1], nd[1][3], nd[2][3]], np.float64)
    return [(int(c[0]), b0), (int(c[1]), b1)]


def leaf(code):
    v = (~code) & 0xFFFFFFFF
    return v >> 3, (v & 7) + 1


def area(b):
    e = b[3:] - b[:3]
    return 2.0 * (e[0] * e[1] + e[1] * e[2] + e[2] * e[0])


def collapse4(b):
    """wide[k] = list of (child_code, box); child_code >= 0 indexes `wide`, negative = the binary tree's leaf code."""
    wide, index = [], {}

    def make(code2):
        if code2 in index:
            return index[code2]
        k = len(wide)
        index[code2] = k
        wide.append(None)
        kids = children2(b["nodes"][code2])
        while len(kids) < 4:
            inner = [(area(bx), i) for i, (c, bx) in enumerate(kids) if c >= 0]
            if not inner:
                break
            _, i = max(inner)
            c, _ = kids.pop(i)
            kids.extend(children2(b["nodes"][c]))
        wide[k] = [((make(c) if c >= 0 else c), bx) for c, bx in kids]
        return k

    root = b["root"]
    if root < 0 or root == NONE:
        return wide, root
    sys.setrecursionlimit(10000)
    return wide, make(root)


def quantise(wide):
    """child boxes snapped OUTWARD to a 256-step grid of the node's own bounds (union of its children)."""
    out = []
    for kids in wide:
        lo = np.min([bx[:3] for _, bx in kids], axis=0)
        hi = np.max([bx[3:] for _, bx in kids], axis=0)
        step = np.maximum(hi - lo, 1e-30) / 255.0
        q = []
        for c, bx in kids:
            ql = lo + np.floor((bx[:3] - lo) / step) * step
            qh = lo + np.ceil((bx[3:] - lo) / step) * step
            q.append((c, np.concatenate([np.minimum(ql, bx[:3]), np.maximum(qh, bx[3:])])))
        out.append(q)
    return out


def slab(bx, orel, inv, best):
    with np.errstate(invalid="ignore"):
        t1, t2 = (bx[:3] - orel) * inv, (bx[3:] - orel) * inv
    tmin = max(0.0, np.nanmax(np.fmin(t1, t2)))
    tmax = min(best, np.nanmin(np.fmax(t1, t2)))
    return tmin <= tmax, tmin


def sphere_t(c, r, o, d):
    p = c - o
    bq = float(p @ d)
    D = bq * bq - float(p @ p) + r * r
    if D < 0:
        return None
    s = np.sqrt(D)
    for t in (bq - s, bq + s):
        if t > 1e-3:
            return t
    return None


def walk(sc, b, get_kids, root, o, d, pad):
    """generic n-ary walk, near child first; returns (hit, fetches, boxes, spheres)"""
    best, hit = np.inf, -1
    fetches = boxes = spheres = 0
    for i in b["big_id"]:
        t = sphere_t(sc.position[i], float(sc.radius[i]), o, d)
        if t is not None and t < best:
            best, hit = t, int(i)
    orel = o - sc.cam_origin
    with np.errstate(divide="ignore"):
        inv = 1.0 / d
    stack = [root]
    while stack:
        node = stack.pop()
        if node == NONE:
            continue
        if node < 0:
            first, cnt = leaf(node)
            for k in range(first, first + cnt):
                i = int(b["leaf_id"][k])
                spheres += 1
                t = sphere_t(sc.position[i], float(sc.radius[i]), o, d)
                if t is not None and (t < best or (t == best and i < hit)):
                    best, hit = t, i
            continue
        fetches += 1
        cand = []
        for c, bx in get_kids(node):
            boxes += 1
            bxp = bx + np.array([-pad, -pad, -pad, pad, pad, pad])
            ok, tn = slab(bxp, orel, inv, best)
            if ok:
                cand.append((tn, c))
        for tn, c in sorted(cand, key=lambda x: -x[0]):   # farthest pushed first: nearest popped first
            stack.append(c)
    return hit, fetches, boxes, spheres


def main():
    n_rays = int(sys.argv[1]) if len(sys.argv) > 1 else 1500
    sc = rtm.load_scene(os.path.join(ROOT, "scenes", "synth10k.json"))
    b = api.build_bvh(sc)
    wide, wroot = collapse4(b)
    wq = quantise(wide)

    small = sc.with_dims(96, 54, 1, 1)
    rng = np.random.default_rng(7)
    xs, ys = rng.integers(0, small.width, n_rays), rng.integers(0, small.height, n_rays)
    d0 = O.oracle_camera_dirs(small, xs, ys, np.ones(n_rays, np.int64), np.ones(n_rays, np.int64))
    org = np.tile(sc.cam_origin, (n_rays, 1))
    hit0, t0, _ = O.oracle_closest_hit(sc, org, d0)
    ok = hit0 >= 0
    p = (org + d0 * t0[:, None])[ok]
    hid = hit0[ok]
    nrm = (p - sc.position[hid]) / sc.radius[hid][:, None].astype(np.float64)
    nrm *= np.where((nrm * d0[ok]).sum(axis=1) < 0, 1.0, -1.0)[:, None]
    u = rng.normal(size=p.shape)
    u /= np.linalg.norm(u, axis=1)[:, None]
    d1 = nrm + u
    d1 /= np.linalg.norm(d1, axis=1)[:, None]        # cosine-weighted about the oriented normal (kernel's construction)
    p = p + nrm * 2e-3                                 # the study's sphere test has no self-hit logic: step off the surface

    kids2 = lambda k: children2(b["nodes"][k])
    variants = {"bvh2": (kids2, b["root"], 0.0), "bvh4": (lambda k: wide[k], wroot, 0.0),
                "bvh4_q8": (lambda k: wq[k], wroot, 0.0)}
    tot = {k: np.zeros(3) for k in variants}
    mism = 0
    for i in range(len(p)):
        ref = None
        for name, (gk, root, pad) in variants.items():
            h, f, bx, s = walk(sc, b, gk, root, p[i], d1[i], pad)
            tot[name] += (f, bx, s)
            if ref is None:
                ref = h
            elif h != ref:
                mism += 1
    n = len(p)
    depth4 = 0
    stack = [(wroot, 1)]
    while stack:
        k, dpt = stack.pop()
        depth4 = max(depth4, dpt)
        stack.extend((c, dpt + 1) for c, _ in wide[k] if c >= 0)
    out = {"study": "bvh4 collapse + 8-bit child boxes vs the two-child tree, synth10k, bounce rays from the primary hits",
           "rays": n, "hit_mismatches_between_variants": mism,
           "nodes": {"bvh2": int(len(b["nodes"])), "bvh4": len(wide)},
           "bytes": {"bvh2": int(len(b["nodes"])) * 64, "bvh4_f32_boxes": len(wide) * (4 * 24 + 16),
                     "bvh4_q8": len(wide) * (24 + 4 * 6 + 16)},
           "depth": {"bvh2": int(b["depth"]), "bvh4": depth4},
           "mean_children_per_bvh4_node": float(np.mean([len(k) for k in wide])),
           "per_ray": {k: {"fetches": v[0] / n, "box_tests": v[1] / n, "sphere_tests": v[2] / n} for k, v in tot.items()}}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
