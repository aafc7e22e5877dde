#!/usr/bin/env python3
"""Host-side study of the inner-node NUMBERING of the sphere hierarchy (bvh_build.cpp: layout 0 = breadth-first
throughout, layout 1 = breadth-first top + treelets of 4): L1 hit rate of the node fetches of incoherent bounce rays under a
sectored LRU model of one SM's L1 (128-byte lines, 32-byte sectors), with as many rays in flight as the hierarchy kernel
keeps (7 blocks x 128 lanes, about half of them walking).  No GPU needed.

    python scripts/study_bvh_layout.py [n_rays] [l1_kb]
"""
import sys
from collections import OrderedDict

import numpy as np

sys.path.insert(0, ".")
import raytracingmin_b200 as rtm  # noqa: E402
from raytracingmin_b200 import api  # noqa: E402

NONE = 0x7FFFFFFF


def bfs_numbering(ch):
    n = ch.shape[0]
    order, new_id = [0], np.full(n, -1, np.int64)
    k = 0
    while k < len(order):
        for c in ch[order[k]]:
            if c >= 0:
                order.append(int(c))
        k += 1
    new_id[np.array(order)] = np.arange(len(order))
    return new_id


def treelet_numbering(ch, top=1024):
    """bvh_build.cpp layout 1: `top` nodes breadth-first, then clusters of 4 filled breadth-first from a depth-first stack."""
    n = ch.shape[0]
    order, fifo, head = [], [0], 0
    while head < len(fifo) and len(order) < top:
        k = fifo[head]
        head += 1
        order.append(k)
        fifo.extend(int(c) for c in ch[k] if c >= 0)
    stack = fifo[head:][::-1]
    while stack:
        lq, lh, in_line = [], 0, len(order) & 3
        while True:
            if lh == len(lq):
                if not stack:
                    break
                lq.append(stack.pop())
            k = lq[lh]
            lh += 1
            order.append(k)
            lq.extend(int(c) for c in ch[k] if c >= 0)
            in_line += 1
            if in_line >= 4:
                break
        stack.extend(lq[lh:][::-1])
    new_id = np.full(n, -1, np.int64)
    new_id[np.array(order)] = np.arange(len(order))
    return new_id


def make_rays(rng, geo, n):
    """Bounce rays: a point on a random sphere, cosine-distributed direction about the outward normal."""
    real = geo[:, 3] > 0
    idx = rng.choice(np.nonzero(real)[0], n)
    c, r = geo[idx, :3].astype(np.float64), np.sqrt(geo[idx, 3].astype(np.float64))
    nrm = rng.normal(size=(n, 3))
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    o = c + nrm * r[:, None] * (1 + 1e-6)
    u1, u2 = rng.random(n), rng.random(n)
    a = np.where(np.abs(nrm[:, :1]) > 0.9, [[0.0, 1.0, 0.0]], [[1.0, 0.0, 0.0]])
    t = np.cross(a, nrm)
    t /= np.linalg.norm(t, axis=1, keepdims=True)
    b = np.cross(nrm, t)
    rr, ph = np.sqrt(u1), 2 * np.pi * u2
    d = t * (rr * np.cos(ph))[:, None] + b * (rr * np.sin(ph))[:, None] + nrm * np.sqrt(1 - u1)[:, None]
    return o, d, idx


class L1:
    def __init__(self, kb):
        self.cap = kb * 1024 // 128
        self.lines = OrderedDict()
        self.hit = self.miss = 0

    def access(self, byte_addr):
        line, sec = byte_addr >> 7, (byte_addr >> 5) & 3
        m = self.lines.get(line)
        if m is None:
            if len(self.lines) >= self.cap:
                self.lines.popitem(last=False)
            self.lines[line] = 1 << sec
            self.miss += 1
        else:
            self.lines.move_to_end(line)
            if m & (1 << sec):
                self.hit += 1
            else:
                self.lines[line] = m | (1 << sec)
                self.miss += 1


def walk(tree, o, d, self_slot, numberings, l1_kb, pool=448, node_bytes=32, pollute=0):
    nodes, geo = tree["nodes"].astype(np.float64), tree["leaf_geo"].astype(np.float64)
    ch = nodes[:, 3, :2].astype(np.float32).view(np.int32).reshape(-1, 2) if False else tree["nodes"][:, 3, :2].copy().view(np.int32)
    lo = np.stack([nodes[:, 0, 0], nodes[:, 0, 2], nodes[:, 2, 0]], 1), np.stack([nodes[:, 1, 0], nodes[:, 1, 2], nodes[:, 2, 2]], 1)
    hi = np.stack([nodes[:, 0, 1], nodes[:, 0, 3], nodes[:, 2, 1]], 1), np.stack([nodes[:, 1, 1], nodes[:, 1, 3], nodes[:, 2, 3]], 1)
    caches = {k: L1(l1_kb) for k in numberings}
    prng = np.random.default_rng(7)
    leaf_base = 1 << 30
    n = o.shape[0]
    nxt = min(pool, n)
    ray = np.arange(nxt)                       # ray index per pool slot
    node = np.zeros(nxt, np.int64)
    stack = np.full((nxt, 64), NONE, np.int64)
    sp = np.ones(nxt, np.int64)                # stack[0] = sentinel
    best = np.full(nxt, np.inf)
    visits = leafs = 0
    inv = 1.0 / d
    while ray.size:
        inner = node >= 0
        k = np.nonzero(inner & (node != NONE))[0]
        if k.size:
            visits += k.size
            nd = node[k]
            for name, num in numberings.items():
                c = caches[name]
                for a in (num[nd] * node_bytes).tolist():
                    c.access(a)
                    if node_bytes == 64:
                        c.access(a + 32)
            r = ray[k]
            tn, tf = [], []
            for j in range(2):
                t0, t1 = (lo[j][nd] - o[r]) * inv[r], (hi[j][nd] - o[r]) * inv[r]
                near, far = np.minimum(t0, t1).max(1), np.maximum(t0, t1).min(1)
                tn.append(np.maximum(near, 0.0))
                tf.append(np.minimum(far, best[k]))
            h0, h1 = tn[0] <= tf[0], tn[1] <= tf[1]
            c0, c1 = ch[nd, 0].astype(np.int64), ch[nd, 1].astype(np.int64)
            first0 = tn[0] <= tn[1]
            both = h0 & h1
            near_c, far_c = np.where(first0, c0, c1), np.where(first0, c1, c0)
            new = np.where(both, near_c, np.where(h0, c0, np.where(h1, c1, NONE)))
            kb = k[both]
            stack[kb, sp[kb]] = far_c[both]
            sp[kb] += 1
            pop = k[~(h0 | h1)]
            sp[pop] -= 1
            new[~(h0 | h1)] = stack[pop, sp[pop]]
            node[k] = new
        k = np.nonzero(node < 0)[0]
        if k.size:
            leafs += k.size
            code = ~node[k]
            first, cnt = code >> 3, (code & 7) + 1
            r = ray[k]
            for j in range(int(cnt.max())):
                m = cnt > j
                s = first[m] + j
                if j % 2 == 0:   # pair fetch: 32 bytes per two slots
                    for name in numberings:
                        c = caches[name]
                        for a in (leaf_base + s * 16).tolist():
                            c.access(a)
                rr = r[m]
                pc = geo[s, :3] - o[rr]
                b = (pc * d[rr]).sum(1)
                disc = b * b - (pc * pc).sum(1) + geo[s, 3]
                t = b - np.sqrt(np.maximum(disc, 0))
                t = np.where(t > 1e-4, t, b + np.sqrt(np.maximum(disc, 0)))
                ok = (disc >= 0) & (t > 1e-4) & (s != self_slot[rr]) & (t < best[k[m]])
                bb = best[k[m]]
                bb[ok] = t[ok]
                best[k[m]] = bb
            sp[k] -= 1
            node[k] = stack[k, sp[k]]
        fin = np.nonzero(node == NONE)[0]
        if fin.size:
            if pollute:   # the shading block's material records (geo, emi, alb of 10^4 spheres: 480 KB, random) allocate too
                for a in prng.integers(0, 480 * 1024 // 128, fin.size * pollute).tolist():
                    for c in caches.values():
                        c.access((2 << 30) + a * 128)
            take = min(fin.size, n - nxt)
            rep = fin[:take]
            ray[rep] = np.arange(nxt, nxt + take)
            nxt += take
            node[rep], sp[rep], best[rep] = 0, 1, np.inf
            stack[rep, 0] = NONE
            drop = fin[take:]
            if drop.size:
                keep = np.ones(ray.size, bool)
                keep[drop] = False
                ray, node, stack, sp, best = ray[keep], node[keep], stack[keep], sp[keep], best[keep]
    return visits, leafs, caches


def main():
    n_rays = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
    l1_kb = int(sys.argv[2]) if len(sys.argv) > 2 else 150
    sc = rtm.load_scene("scenes/synth10k.json", "abc")
    tree = api.build_bvh(sc)
    n = tree["nodes"].shape[0]
    ch = tree["nodes"][:, 3, :2].copy().view(np.int32)
    rng = np.random.default_rng(1)
    geo = tree["leaf_geo"]
    o, d, slot = make_rays(rng, geo, n_rays)
    numberings = {"breadth-first (as built)": np.arange(n), "top + treelets of 4": treelet_numbering(ch)}
    for nb, pol in ((32, 0), (32, 4), (64, 0)):
        v, lf, caches = walk(tree, o, d, slot, numberings, l1_kb, node_bytes=nb, pollute=pol)
        print(f"{n} inner nodes x {nb} B, {n_rays} rays, {v / n_rays:.1f} node visits and {lf / n_rays:.2f} leaves per ray, L1 model {l1_kb} KB, "
              f"{pol} material lines allocated per ray")
        for name, c in caches.items():
            print(f"  {name:24s} sector hit rate {100 * c.hit / (c.hit + c.miss):5.1f} %   ({c.miss / n_rays:.2f} L2 sectors per ray)")


if __name__ == "__main__":
    main()
