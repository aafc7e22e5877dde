"""Host-side checks of the sphere hierarchy (raytracingmin_b200/csrc/bvh_build.cpp) — no GPU.

The hierarchy only selects candidates; the winner must be the one the reference's O(N) loop
(src/Renderer.cpp:58-73) picks.  Here: structural invariants of the built tree, and a Python
emulation of the device traversal of kernels_exact.cu (same per-ray box inflation) checked
against the oracle's brute-force closest hit on the 10 000-sphere scene.
"""
import os

import numpy as np
import pytest

import oracle_lib as O
import raytracingmin_b200 as rtm
from raytracingmin_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NONE = 0x7FFFFFFF


def _children(node):
    c = node[3].view(np.int32)
    return int(c[0]), int(c[1])


def _boxes(node):
    b0 = (node[0][0], node[0][1], node[0][2], node[0][3], node[2][0], node[2][1])
    b1 = (node[1][0], node[1][1], node[1][2], node[1][3], node[2][2], node[2][3])
    return b0, b1


def _leaf(code):
    v = (~code) & 0xFFFFFFFF
    return v >> 3, (v & 7) + 1


def _collect(b, code, out):
    if code < 0:
        first, cnt = _leaf(code)
        out.extend(range(first, first + cnt))
        return
    c0, c1 = _children(b["nodes"][code])
    _collect(b, c0, out)
    _collect(b, c1, out)


@pytest.fixture(scope="module")
def synth():
    sc = rtm.load_scene(os.path.join(ROOT, "scenes", "synth10k.json"))
    return sc, api.build_bvh(sc)


def _real(leaf_id):
    """leaf slots without the pad entries (id -1) that keep leaves of two or more spheres on even slots"""
    return [int(v) for v in leaf_id if v >= 0]


def test_leaves_of_two_or_more_start_on_even_slots(synth):
    sc, b = synth
    seen = 0
    for nd in b["nodes"]:
        for code in _children(nd):
            if code < 0:
                first, cnt = _leaf(int(code))
                assert cnt == 1 or first % 2 == 0, (first, cnt)
                assert (b["leaf_id"][first:first + cnt] >= 0).all()
                seen += cnt
    assert seen == sc.n - len(b["big_id"])


def test_every_sphere_exactly_once(synth):
    sc, b = synth
    assert sorted(_real(b["leaf_id"]) + b["big_id"]) == list(range(sc.n))
    assert b["big_id"] == [0, 1, 2, 3, 4, 5, 6]           # the Cornell room of scenes/make_scenes.py
    assert b["depth"] + 2 < 64


def test_boxes_contain_their_spheres_and_rmin_is_a_lower_bound(synth):
    sc, b = synth
    ref = sc.cam_origin
    rel = sc.position - ref
    rad = sc.radius.astype(np.float64)
    for ni in range(0, len(b["nodes"]), 37):              # every 37th node, full subtree each
        node = b["nodes"][ni]
        for code, box, rmin in zip(_children(node), _boxes(node), (node[3][2], node[3][3])):
            slots = []
            _collect(b, code, slots)
            ids = b["leaf_id"][slots]
            lo = (rel[ids] - rad[ids, None]).min(axis=0)
            hi = (rel[ids] + rad[ids, None]).max(axis=0)
            assert box[0] < lo[0] and box[2] < lo[1] and box[4] < lo[2]
            assert box[1] > hi[0] and box[3] > hi[1] and box[5] > hi[2]
            assert rmin <= rad[ids].min()
    # leaf slots carry the float centre the FP32 kernels use and r^2
    real = b["leaf_id"] >= 0                               # pad slots (id -1) keep leaves of 2+ spheres on even slots
    k = b["leaf_id"][real]
    assert np.array_equal(b["leaf_geo"][real][:, :3], rel[k].astype(np.float32))
    assert np.array_equal(b["leaf_geo"][real][:, 3], (rad[k] * rad[k]).astype(np.float32))
    assert len(b["leaf_id"]) % 2 == 0 and (b["leaf_geo"][~real][:, 3] == 0).all()


def _sphere_test(c, r, o, d):
    """src/SettingData.cpp:197-223 in numpy doubles (r*r in float, minValue 1e-5f)."""
    p = c - o
    bq = p[0] * d[0] + p[1] * d[1] + p[2] * d[2]
    D4 = bq * bq - (p[0] * p[0] + p[1] * p[1] + p[2] * p[2]) + float(np.float32(r) * np.float32(r))
    if D4 < 0.0:
        return None
    s = np.sqrt(D4)
    t1, t2 = bq - s, bq + s
    mv = float(np.float32(1e-5))
    if t1 < mv and t2 < mv:
        return None
    return t1 if t1 > 0.001 else t2


def _traverse(sc, b, o, d):
    """kernels_exact.cu: closest_hit_bvh, statement by statement."""
    dd = d[0] * d[0] + d[1] * d[1] + d[2] * d[2]
    assert abs(dd - 1.0) <= 4e-7
    best, bid, tested = np.finfo(np.float64).max, -1, 0

    def consider(i):
        nonlocal best, bid, tested
        tested += 1
        t = _sphere_test(sc.position[i], sc.radius[i], o, d)
        if t is not None and t > 0.0 and (t < best or (t == best and i < bid)):
            best, bid = t, i

    for i in b["big_id"]:
        consider(i)
    ke = 4.5 * (abs(dd - 1.0) + 1e-14)
    orel = o - sc.cam_origin
    with np.errstate(divide="ignore"):
        inv = 1.0 / d

    def box(bx, rmin):
        lo = np.array([bx[0], bx[2], bx[4]], np.float64)
        hi = np.array([bx[1], bx[3], bx[5]], np.float64)
        far = np.maximum(np.abs(lo - orel), np.abs(hi - orel))
        infl = ke * float(far @ far) / float(rmin)
        with np.errstate(invalid="ignore"):
            t1, t2 = (lo - infl - orel) * inv, (hi + infl - orel) * inv
        tmin = max(0.0, np.nanmax(np.fmin(t1, t2)))
        tmax = min(best, np.nanmin(np.fmax(t1, t2)))
        return tmin <= tmax, tmin

    stack, node = [NONE], b["root"]
    while node != NONE:
        # the bound the FP32 kernel's shared-memory stack instantiation relies on (kernels_radiance.cu, SST):
        # one entry per level of the current path + the sentinel
        assert len(stack) <= b["depth"] + 1
        if node >= 0:
            nd = b["nodes"][node]
            c0, c1 = _children(nd)
            b0, b1 = _boxes(nd)
            h0, n0 = box(b0, nd[3][2])
            h1, n1 = box(b1, nd[3][3])
            if h0 and h1:
                if n1 < n0:
                    stack.append(c0)
                    node = c1
                else:
                    stack.append(c1)
                    node = c0
            elif h0 or h1:
                node = c0 if h0 else c1
            else:
                node = stack.pop()
        else:
            first, cnt = _leaf(node)
            for k in range(first, first + cnt):
                consider(int(b["leaf_id"][k]))
            node = stack.pop()
    return bid, best, tested


def test_emulated_traversal_equals_brute_force(synth):
    sc, b = synth
    small = sc.with_dims(96, 54, 1, 1)
    rng = np.random.default_rng(5)
    n = 700
    xs, ys = rng.integers(0, small.width, n), rng.integers(0, small.height, n)
    d = O.oracle_camera_dirs(small, xs, ys, np.ones(n, np.int64), np.ones(n, np.int64))
    org = np.tile(sc.cam_origin, (n, 1))
    hit_o, t_o, _ = O.oracle_closest_hit(sc, org, d)
    tested = 0
    for i in range(n):
        bid, best, k = _traverse(sc, b, org[i], d[i])
        tested += k
        assert bid == hit_o[i], (i, bid, hit_o[i])
        if bid >= 0:
            assert best == t_o[i]
    assert tested / n < 200          # the point of the hierarchy: far fewer than 10 000 tests per ray
    # secondary rays: start on the primary hit, random unit directions normalised the reference's way
    p = org + d * t_o[:, None]
    d2 = rng.normal(size=(n, 3))
    m = np.sqrt((d2 * d2).sum(axis=1).astype(np.float32)).astype(np.float64)     # Ray.h:67-72
    d2 = d2 / m[:, None]
    hit2, t2, _ = O.oracle_closest_hit(sc, p, d2)
    for i in range(0, n, 2):
        bid, best, _ = _traverse(sc, b, p[i], d2[i])
        assert bid == hit2[i], (i, bid, hit2[i])
        if bid >= 0:
            assert best == t2[i]


def test_small_and_degenerate_inputs():
    for seed, n, closed in [(1, 0, False), (2, 1, False), (3, 2, True), (4, 5, True), (5, 70, False)]:
        sc = rtm.random_scene(seed, n, closed=closed)
        b = api.build_bvh(sc)
        assert sorted(_real(b["leaf_id"]) + b["big_id"]) == list(range(sc.n))
    # many identical spheres (what COMPAT_HEAD's position collapse produces): median splits must terminate
    sc = rtm.random_scene(7, 300, closed=False)
    sc.position[:] = sc.position[0]
    sc.radius[:] = sc.radius[0]
    b = api.build_bvh(sc.normalized())
    assert sorted(_real(b["leaf_id"]) + b["big_id"]) == list(range(sc.n))
    assert b["depth"] < 40


# ---------------- 4-wide collapse (bvh.h: RtmBvh4Host; experimental walk, RTM_BVH_WIDE=1) ----------------
def _wide_children(nd):
    codes = nd[6].view(np.int32)
    out = []
    for j in range(4):
        if int(codes[j]) != NONE:
            out.append((int(codes[j]), np.array([nd[0][j], nd[2][j], nd[4][j], nd[1][j], nd[3][j], nd[5][j]], np.float64)))
    return out


def test_wide_collapse_structure(synth):
    sc, b = synth
    w = api.build_bvh4(sc)
    assert w["root"] == 0 and len(w["nodes"]) < 0.6 * len(b["nodes"])
    assert w["depth"] <= (b["depth"] + 1) // 2 + 1
    # the binary tree's child boxes, keyed by child code: the wide tree must reuse them unchanged
    boxes2 = {}
    for nd in b["nodes"]:
        (c0, c1), (b0, b1) = _children(nd), _boxes(nd)
        boxes2[c0] = np.array([b0[0], b0[2], b0[4], b0[1], b0[3], b0[5]], np.float64)
        boxes2[c1] = np.array([b1[0], b1[2], b1[4], b1[1], b1[3], b1[5]], np.float64)
    seen_leaf_slots, seen_nodes, depth = [], set(), 0
    stack = [(0, 0)]
    while stack:
        k, lvl = stack.pop()
        assert k not in seen_nodes
        seen_nodes.add(k)
        kids = _wide_children(w["nodes"][k])
        assert 2 <= len(kids) <= 4
        for code, bx in kids:
            if code >= 0:
                stack.append((code, lvl + 1))
            else:
                assert np.array_equal(bx, boxes2[code])          # a leaf keeps the box it had under its binary parent
                first, cnt = _leaf(code)
                seen_leaf_slots.extend(range(first, first + cnt))
                depth = max(depth, lvl + 1)
    assert len(seen_nodes) == len(w["nodes"])
    assert sorted(seen_leaf_slots) == [k for k in range(len(b["leaf_id"])) if b["leaf_id"][k] >= 0]   # every real leaf slot exactly once
    assert depth == w["depth"]


def test_wide_walk_equals_brute_force(synth):
    """Python statement of bvh4_inner_step's order (hit children near to far, far ones pushed) on the C++-built wide
    tree: same winner as testing every sphere, fewer node visits than the binary walk, stack within 3*depth + 1."""
    sc, b = synth
    w = api.build_bvh4(sc)
    rng = np.random.default_rng(21)
    n = 300
    org = sc.cam_origin + rng.uniform(-4, 4, size=(n, 3)) * np.array([1.0, 1.0, 0.5])
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1)[:, None]
    pos, rad = np.asarray(sc.position, np.float64), np.asarray(sc.radius, np.float64)

    def sphere_t(i, o, dd):
        p = pos[i] - o
        bq = float(p @ dd)
        D = bq * bq - float(p @ p) + rad[i] * rad[i]
        if D < 0:
            return None
        s = np.sqrt(D)
        for t in (bq - s, bq + s):
            if t > 1e-3:
                return t
        return None

    visits_w = 0
    for i in range(n):
        o, dd = org[i], d[i]
        # brute force, vectorised
        p = pos - o
        bq = p @ dd
        D = bq * bq - (p * p).sum(axis=1) + rad * rad
        s = np.sqrt(np.where(D >= 0, D, np.nan))
        t1, t2 = bq - s, bq + s
        t = np.where(t1 > 1e-3, t1, np.where(t2 > 1e-3, t2, np.inf))
        t = np.where(np.isnan(t), np.inf, t)
        want = int(np.argmin(t)) if np.isfinite(t.min()) else -1
        best, hit = np.inf, -1
        for j in b["big_id"]:
            tt = sphere_t(j, o, dd)
            if tt is not None and (tt < best or (tt == best and j < hit)):
                best, hit = tt, j
        orel = o - sc.cam_origin
        with np.errstate(divide="ignore"):
            inv = 1.0 / dd
        stack, node = [NONE], w["root"]
        while node != NONE:
            assert len(stack) <= 3 * w["depth"] + 1
            if node >= 0:
                visits_w += 1
                cand = []
                for code, bx in _wide_children(w["nodes"][node]):
                    with np.errstate(invalid="ignore"):
                        a1, a2 = (bx[:3] - orel) * inv, (bx[3:] - orel) * inv
                    tmin = max(0.0, np.nanmax(np.fmin(a1, a2)))
                    tmax = min(best, np.nanmin(np.fmax(a1, a2)))
                    if tmin <= tmax:
                        cand.append((tmin, code))
                cand.sort(key=lambda x: x[0])
                for _, code in reversed(cand[1:]):
                    stack.append(code)
                node = cand[0][1] if cand else stack.pop()
            else:
                first, cnt = _leaf(node)
                for k in range(first, first + cnt):
                    j = int(b["leaf_id"][k])
                    tt = sphere_t(j, o, dd)
                    if tt is not None and (tt < best or (tt == best and j < hit)):
                        best, hit = tt, j
                node = stack.pop()
        assert hit == want, (i, hit, want)
    assert visits_w / n < 40


# ---------------- compact half-precision nodes (bvh.h: RtmBvh16Host): what the FP32 hierarchy kernel walks ----------------
def _half_box(words, ref, inv_scale):
    """u32[3] -> (lo[3], hi[3]) in the float nodes' frame."""
    h = np.array(words, np.uint32).view(np.float16).astype(np.float64)     # cx cy cz hx hy hz
    c, e = h[:3], h[3:]
    return (c - e) * inv_scale + ref, (c + e) * inv_scale + ref


def test_compact16_boxes_contain_the_float_boxes(synth):
    """Conservativeness: every child box of the compact tree, decoded as the kernel decodes it, contains the padded float
    box of the same child — the compact walk can only visit MORE candidates than the float walk, never miss one.  Codes and
    node numbering are shared with the float tree."""
    sc, b = synth
    h = api.build_bvh16(sc)
    assert h["ok"] and len(h["nodes"]) == len(b["nodes"])
    assert 0.0 <= h["inflation"] < 0.1, h["inflation"]                     # synth10k: boxes of r >= 0.05 spheres in a 20-unit room
    worst = 0.0
    for nd, w in zip(b["nodes"], h["nodes"]):
        codes, boxes = _children(nd), _boxes(nd)
        assert (int(np.int32(w[6])), int(np.int32(w[7]))) == tuple(int(c) for c in codes)
        for j in range(2):
            lo, hi = _half_box(w[3 * j:3 * j + 3], h["ref"], h["inv_scale"])
            bx = boxes[j]
            flo, fhi = np.array([bx[0], bx[2], bx[4]], np.float64), np.array([bx[1], bx[3], bx[5]], np.float64)
            assert (lo <= flo).all() and (hi >= fhi).all(), (lo, flo, hi, fhi)
            worst = max(worst, float(((hi - lo) / np.maximum(fhi - flo, 1e-30)).max()))
    assert worst < 1.5, worst                                              # no box more than half again as wide


def test_compact16_reports_what_half_precision_cannot_hold():
    """A scene whose features are tiny against its extent (two clusters of small spheres a million units apart): the
    median leaf box would be inflated far beyond the host's threshold (0.25), so renderer_create keeps the float nodes."""
    sc = rtm.random_scene(3, 200, width=8, height=8, samples=1, superSamples=1, closed=False)
    sc.position[:100, 0] += 1e6
    h = api.build_bvh16(sc.normalized())
    assert (not h["ok"]) or h["inflation"] > 0.25, h


# ---------------- 4-wide compact nodes (bvh.h: RtmBvh16x4Host): what the FP32 walk uses by default ----------------
def test_compact16x4_is_a_collapse_of_the_compact_tree(synth):
    """Every child of a 4-wide node is a child of SOME two-child compact node with the very same box words and, below it,
    the very same leaves; every leaf slot is reached exactly once; empty slots carry the NONE code behind a box that no
    ray can hit; the depth that bounds the walk's stack (3 entries per level + the sentinel) is the real one."""
    sc, b = synth
    h2, h4 = api.build_bvh16(sc), api.build_bvh16x4(sc)
    assert h4["ok"] and 0 < len(h4["nodes"]) < len(h2["nodes"])
    by_words = {}
    for k, w in enumerate(h2["nodes"]):
        for j in range(2):
            by_words.setdefault((int(w[3 * j]), int(w[3 * j + 1]), int(w[3 * j + 2]), ), []).append(int(np.int32(w[6 + j])))

    def slots_below2(code):
        out = []
        _collect(b, code, out)
        return sorted(out)

    seen, depth, filled = [], 0, 0
    stack = [(0, 0)]
    visited = set()
    while stack:
        k, lvl = stack.pop()
        assert k not in visited
        visited.add(k)
        w = h4["nodes"][k]
        for j in range(4):
            box = tuple(int(v) for v in w[8 * (j >> 1) + 3 * (j & 1):8 * (j >> 1) + 3 * (j & 1) + 3])
            code = int(np.int32(w[8 * (j >> 1) + 6 + (j & 1)]))
            if code == NONE:
                assert box == (0, 0xFC000000, 0xFC00FC00)          # centre 0, half-extents -inf
                continue
            filled += 1
            assert box in by_words, (k, j)
            if code < 0:
                assert code in by_words[box]                        # the same leaf behind the same box
                first, cnt = _leaf(code)
                seen.extend(range(first, first + cnt))
                depth = max(depth, lvl + 1)
            else:
                stack.append((code, lvl + 1))
    assert len(visited) == len(h4["nodes"])
    real = [s for s in seen if b["leaf_id"][s] >= 0]
    assert sorted(int(b["leaf_id"][s]) for s in real) == sorted(_real(b["leaf_id"]))
    assert len(set(seen)) == len(seen)
    assert depth == h4["depth"] and 3 * depth + 1 <= 64
    assert filled / (4 * len(h4["nodes"])) > 0.7                   # synth10k: 2.98 of 4 slots in use


def test_compact16x4_walk_finds_the_brute_force_hit(synth):
    """The walk as the kernel does it (nearest hit child first, the others pushed in slot order, half-precision boxes)
    against brute force over all spheres, for bounce-like rays."""
    sc, b = synth
    h2, h4 = api.build_bvh16(sc), api.build_bvh16x4(sc)
    rng = np.random.default_rng(5)
    n = 300
    pos, rad = np.asarray(sc.position, np.float64), np.asarray(sc.radius, np.float64)
    src = rng.integers(0, sc.n, n)
    nrm = rng.normal(size=(n, 3))
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    org = pos[src] + nrm * rad[src][:, None] * (1 + 1e-3)
    d = nrm + rng.normal(size=(n, 3)) * 0.7
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    ref = np.asarray(sc.cam_origin, np.float64)
    fetches = 0
    for i in range(n):
        o, dd = org[i], d[i]
        p = pos - o
        bq = p @ dd
        D = bq * bq - (p * p).sum(1) + rad * rad
        t = np.where(D >= 0, bq - np.sqrt(np.maximum(D, 0)), np.inf)
        t = np.where(t > 1e-6, t, np.where(D >= 0, bq + np.sqrt(np.maximum(D, 0)), np.inf))
        t = np.where(t > 1e-6, t, np.inf)
        want = int(np.argmin(t)) if np.isfinite(t.min()) else -1
        best, hit = np.inf, -1
        for j in b["big_id"]:
            if t[j] < best:
                best, hit = t[j], int(j)
        with np.errstate(divide="ignore", invalid="ignore"):
            inv = 1.0 / dd
            st = [0]
            while st:
                code = st.pop()
                if code < 0:
                    first, cnt = _leaf(code)
                    for s in range(first, first + cnt):
                        k = int(b["leaf_id"][s])
                        if k >= 0 and (t[k] < best or (t[k] == best and k < hit)):
                            best, hit = t[k], k
                    continue
                fetches += 1
                w = h4["nodes"][code]
                cand = []
                for j in range(4):
                    c = int(np.int32(w[8 * (j >> 1) + 6 + (j & 1)]))
                    # the kernel's form (box_test_h16): near / far = (c - o) / d -+ h |1 / d|; an empty slot has h = -inf
                    hw = np.array(w[8 * (j >> 1) + 3 * (j & 1):8 * (j >> 1) + 3 * (j & 1) + 3], np.uint32).view(np.float16).astype(np.float64)
                    tc = (hw[:3] * h2["inv_scale"] + h2["ref"] + ref - o) * inv
                    he = hw[3:] * h2["inv_scale"] * np.abs(inv)
                    tn, tf = max(0.0, np.nanmax(tc - he)), min(best, np.nanmin(tc + he))
                    if tn <= tf:
                        assert c != NONE
                        cand.append((tn, j, c))
                if cand:
                    near = min(cand)[1]
                    st.extend(c for _, j, c in cand if j != near)      # slot order
                    st.append([c for _, j, c in cand if j == near][0])
        assert hit == want, (i, hit, want)
    assert fetches / n < 30
