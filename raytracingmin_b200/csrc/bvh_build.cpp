// bvh_build.cpp — host-side binned-SAH build of the sphere hierarchy (see bvh.h).
// Runs once per renderer_create; 10 000 spheres take a few milliseconds.
#include "bvh.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>

namespace {

struct Prim {
    double lo[3], hi[3], c[3];
    double r;
    int32_t id;
};

struct Box {
    double lo[3], hi[3];
    void reset() {
        for (int k = 0; k < 3; ++k) { lo[k] = DBL_MAX; hi[k] = -DBL_MAX; }
    }
    void grow(const double* l, const double* h) {
        for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], l[k]); hi[k] = std::max(hi[k], h[k]); }
    }
    double area() const {
        const double dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        return (dx < 0) ? 0.0 : 2.0 * (dx * dy + dy * dz + dz * dx);
    }
};

float round_down(double v) {
    float f = (float)v;
    if ((double)f > v) f = std::nextafterf(f, -INFINITY);
    return f;
}
float round_up(double v) {
    float f = (float)v;
    if ((double)f < v) f = std::nextafterf(f, INFINITY);
    return f;
}

struct Builder {
    std::vector<Prim> prims;
    RtmBvhHost* out;
    int max_depth = 0;

    struct Sub {
        int32_t code;
        Box box;
        double rmin;
    };

    Sub make_leaf(int b, int e) {
        Sub s;
        s.box.reset();
        s.rmin = DBL_MAX;
        // a leaf of two or more spheres starts on an EVEN slot: the FP32 walk fetches its spheres two per 32-byte request.
        // The pad slot holds a sphere nothing can hit (id -1); no leaf code covers it.
        if (e - b >= 2 && (out->leaf_id.size() & 1u)) {
            out->leaf_geo.push_back(make_float4(3e18f, 3e18f, 3e18f, 0.f));
            out->leaf_id.push_back(-1);
        }
        const uint32_t first = (uint32_t)out->leaf_id.size();
        for (int i = b; i < e; ++i) {
            const Prim& p = prims[i];
            s.box.grow(p.lo, p.hi);
            s.rmin = std::min(s.rmin, p.r);
            out->leaf_geo.push_back(make_float4((float)p.c[0], (float)p.c[1], (float)p.c[2], (float)(p.r * p.r)));
            out->leaf_id.push_back(p.id);
        }
        s.code = rtm_bvh_leaf_code(first, (uint32_t)(e - b));
        return s;
    }

    // SAH constants: cost of visiting an inner node (two box tests, a stack operation, a dependent 64-byte load)
    // relative to one sphere test, and the largest leaf (the leaf code holds up to 8)
    double c_trav = RTM_BVH_CTRAV;
    int leaf_max = RTM_BVH_LEAF_MAX;

    Sub build(int b, int e, int depth) {
        max_depth = std::max(max_depth, depth);
        const int n = e - b;
        if (n <= 2) return make_leaf(b, e);   // a 3..4 range is still split below when SAH says it pays
        Box bounds, cb;
        bounds.reset();
        cb.reset();
        for (int i = b; i < e; ++i) {
            bounds.grow(prims[i].lo, prims[i].hi);
            cb.grow(prims[i].c, prims[i].c);
        }
        int axis = -1, split_at = -1;
        if (depth < 32) {
            // binned SAH over the centroid bounds, 16 bins per axis
            const int NB = 16;
            double best_cost = (n <= leaf_max) ? (double)n * bounds.area() : DBL_MAX;
            int best_axis = -1, best_bin = -1;
            for (int ax = 0; ax < 3; ++ax) {
                const double ext = cb.hi[ax] - cb.lo[ax];
                if (!(ext > 0)) continue;
                Box bb[NB];
                int cnt[NB] = {0};
                for (auto& x : bb) x.reset();
                const double scale = NB / ext;
                for (int i = b; i < e; ++i) {
                    int k = (int)((prims[i].c[ax] - cb.lo[ax]) * scale);
                    k = std::min(std::max(k, 0), NB - 1);
                    bb[k].grow(prims[i].lo, prims[i].hi);
                    ++cnt[k];
                }
                double right_area[NB];
                int right_cnt[NB];
                Box acc;
                acc.reset();
                int c = 0;
                for (int k = NB - 1; k >= 1; --k) {
                    if (cnt[k]) acc.grow(bb[k].lo, bb[k].hi);
                    c += cnt[k];
                    right_area[k] = acc.area();
                    right_cnt[k] = c;
                }
                acc.reset();
                c = 0;
                for (int k = 0; k < NB - 1; ++k) {
                    if (cnt[k]) acc.grow(bb[k].lo, bb[k].hi);
                    c += cnt[k];
                    if (c == 0 || right_cnt[k + 1] == 0) continue;
                    const double cost = c_trav * bounds.area() + acc.area() * c + right_area[k + 1] * right_cnt[k + 1];
                    if (cost < best_cost) { best_cost = cost; best_axis = ax; best_bin = k; }
                }
            }
            if (best_axis >= 0) {
                const double ext = cb.hi[best_axis] - cb.lo[best_axis];
                const double scale = NB / ext;
                const double lo = cb.lo[best_axis];
                auto mid = std::partition(prims.begin() + b, prims.begin() + e, [&](const Prim& p) {
                    int k = (int)((p.c[best_axis] - lo) * scale);
                    k = std::min(std::max(k, 0), NB - 1);
                    return k <= best_bin;
                });
                split_at = (int)(mid - prims.begin());
                axis = best_axis;
            }
        }
        if (axis < 0 || split_at <= b || split_at >= e) {
            if (n <= leaf_max) return make_leaf(b, e);
            // median split by count along the widest centroid axis (also the depth >= 32 rule:
            // halving bounds the remaining depth by log2(n))
            axis = 0;
            for (int k = 1; k < 3; ++k)
                if (cb.hi[k] - cb.lo[k] > cb.hi[axis] - cb.lo[axis]) axis = k;
            split_at = b + n / 2;
            std::nth_element(prims.begin() + b, prims.begin() + split_at, prims.begin() + e,
                             [axis](const Prim& x, const Prim& y) { return x.c[axis] < y.c[axis] || (x.c[axis] == y.c[axis] && x.id < y.id); });
        }
        const size_t me = out->nodes.size();
        out->nodes.resize(me + 4);
        const Sub l = build(b, split_at, depth + 1);
        const Sub r = build(split_at, e, depth + 1);
        auto pad = [](const Box& bx, const double rmin, float* lo, float* hi) {
            (void)rmin;
            for (int k = 0; k < 3; ++k) {
                const double m = std::max(std::fabs(bx.lo[k]), std::fabs(bx.hi[k]));
                const double p = 1e-5 + 2e-6 * m;
                lo[k] = round_down(bx.lo[k] - p);
                hi[k] = round_up(bx.hi[k] + p);
            }
        };
        float l0[3], h0[3], l1[3], h1[3];
        pad(l.box, l.rmin, l0, h0);
        pad(r.box, r.rmin, l1, h1);
        float4* q = out->nodes.data() + me;
        q[0] = make_float4(l0[0], h0[0], l0[1], h0[1]);
        q[1] = make_float4(l1[0], h1[0], l1[1], h1[1]);
        q[2] = make_float4(l0[2], h0[2], l1[2], h1[2]);
        float4 c;
        memcpy(&c.x, &l.code, 4);
        memcpy(&c.y, &r.code, 4);
        c.z = round_down(l.rmin);
        c.w = round_down(r.rmin);
        q[3] = c;
        Sub s;
        s.code = (int32_t)(me / 4);
        s.box = l.box;
        s.box.grow(r.box.lo, r.box.hi);
        s.rmin = std::min(l.rmin, r.rmin);
        return s;
    }
};

}  // namespace

bool rtm_bvh_build(int32_t n, const double* position, const float* radius, const double ref[3], RtmBvhHost& out, double c_trav,
                   int leaf_max, int layout) {
    out = RtmBvhHost();
    if (n <= 0) return true;
    // Spheres far larger than the typical one (the r = 1e4 "walls" of the Cornell scenes) would
    // overlap every box: up to RTM_BVH_MAX_BIG of them are kept out and tested for every ray.
    std::vector<float> rs(radius, radius + n);
    std::nth_element(rs.begin(), rs.begin() + n / 2, rs.end());
    const double median = rs[n / 2];
    std::vector<int32_t> cand;
    for (int i = 0; i < n; ++i)
        if ((double)radius[i] > 8.0 * median || (double)radius[i] >= RTM_BVH_BIG_RADIUS) cand.push_back(i);
    std::sort(cand.begin(), cand.end(), [&](int a, int b) { return radius[a] > radius[b] || (radius[a] == radius[b] && a < b); });
    if ((int)cand.size() > RTM_BVH_MAX_BIG) {
        for (size_t k = RTM_BVH_MAX_BIG; k < cand.size(); ++k)
            if ((double)radius[cand[k]] >= RTM_BVH_BIG_RADIUS) ++out.n_big_overflow;
        cand.resize(RTM_BVH_MAX_BIG);
    }
    std::sort(cand.begin(), cand.end());  // ascending original index: the tie rule of Renderer.cpp:67
    std::vector<char> is_big(n, 0);
    for (int32_t i : cand) {
        is_big[i] = 1;
        out.big_id[out.n_big++] = i;
    }
    Builder bl;
    bl.out = &out;
    if (c_trav > 0.0) bl.c_trav = c_trav;                                   // rtm_options.debug_bvh_ctrav / _leaf_max
    if (leaf_max > 0) bl.leaf_max = std::min(8, std::max(1, leaf_max));
    bl.prims.reserve(n - out.n_big);
    for (int i = 0; i < n; ++i) {
        if (is_big[i]) continue;
        Prim p;
        p.r = (double)radius[i];
        p.id = i;
        for (int k = 0; k < 3; ++k) {
            p.c[k] = position[3 * i + k] - ref[k];
            p.lo[k] = p.c[k] - p.r;
            p.hi[k] = p.c[k] + p.r;
        }
        bl.prims.push_back(p);
    }
    if (bl.prims.empty()) return true;
    const Builder::Sub root = bl.build(0, (int)bl.prims.size(), 0);
    out.root = root.code;
    out.depth = bl.max_depth;
    if (out.leaf_id.size() & 1u) {   // even length: the last pair fetch stays inside the array
        out.leaf_geo.push_back(make_float4(3e18f, 3e18f, 3e18f, 0.f));
        out.leaf_id.push_back(-1);
    }
    if (root.code >= 0) {
        // Renumber the inner nodes.  layout 0 (default): breadth-first — node 0 is the root and the first 2^k - 1 indices
        // are the top k levels (neighbouring cache lines for the part of the tree every ray walks; a shared-memory copy of
        // these levels was measured and was slower than L1: profiles/r01).  layout 1 (A/B knob): the first
        // RTM_BVH_LAYOUT_TOP nodes breadth-first, the rest in TREELETS of 4 (a node, its inner children, one grandchild:
        // four compact 32-byte nodes in one 128-byte line).  Measured equal (6.44 Grays/s either way on C5, and an LRU
        // model of the L1 agrees: scripts/study_bvh_layout.py) — a warp waits for its slowest lane, and with 17 lanes on
        // 17 different nodes some lane misses L1 under any numbering.  Only addresses change: same walk, same image.
        const size_t n_nodes = out.nodes.size() / 4;
        auto kids = [&](int32_t node, int32_t ch[2]) {
            const float4 c = out.nodes[4 * (size_t)node + 3];
            memcpy(&ch[0], &c.x, 4);
            memcpy(&ch[1], &c.y, 4);
        };
        std::vector<int32_t> order, new_id(n_nodes, -1);
        order.reserve(n_nodes);
        std::vector<int32_t> fifo{root.code};
        size_t head = 0;
        const size_t top = (layout == 1) ? (size_t)RTM_BVH_LAYOUT_TOP : n_nodes;
        while (head < fifo.size() && order.size() < top) {
            const int32_t node = fifo[head++];
            order.push_back(node);
            int32_t ch[2];
            kids(node, ch);
            for (int j = 0; j < 2; ++j)
                if (ch[j] >= 0) fifo.push_back(ch[j]);
        }
        // the subtrees hanging below the top, depth-first so that neighbouring subtrees get neighbouring addresses
        std::vector<int32_t> stack(fifo.rbegin(), fifo.rend() - (ptrdiff_t)head);
        std::vector<int32_t> lq;
        while (!stack.empty()) {
            lq.clear();
            size_t lh = 0, in_line = order.size() & 3u;   // keep filling the line the top part left open
            do {
                if (lh == lq.size()) {
                    if (stack.empty()) break;
                    lq.push_back(stack.back());
                    stack.pop_back();
                }
                const int32_t node = lq[lh++];
                order.push_back(node);
                int32_t ch[2];
                kids(node, ch);
                for (int j = 0; j < 2; ++j)
                    if (ch[j] >= 0) lq.push_back(ch[j]);
            } while (++in_line < 4);
            for (size_t k = lq.size(); k > lh; --k) stack.push_back(lq[k - 1]);
        }
        for (size_t k = 0; k < order.size(); ++k) new_id[order[k]] = (int32_t)k;
        std::vector<float4> nn(out.nodes.size());
        for (size_t k = 0; k < order.size(); ++k) {
            for (int j = 0; j < 4; ++j) nn[4 * k + j] = out.nodes[4 * (size_t)order[k] + j];
            float4& c = nn[4 * k + 3];
            int32_t ch[2];
            memcpy(&ch[0], &c.x, 4);
            memcpy(&ch[1], &c.y, 4);
            for (int j = 0; j < 2; ++j)
                if (ch[j] >= 0) ch[j] = new_id[ch[j]];
            memcpy(&c.x, &ch[0], 4);
            memcpy(&c.y, &ch[1], 4);
        }
        out.nodes.swap(nn);
        out.root = 0;
    }
    return bl.max_depth + 2 < RTM_BVH_STACK;
}

namespace {
struct Kid4 {
    int32_t code;
    float lo[3], hi[3], rmin;
    double area() const {
        const double ex = (double)hi[0] - lo[0], ey = (double)hi[1] - lo[1], ez = (double)hi[2] - lo[2];
        return 2.0 * (ex * ey + ey * ez + ez * ex);
    }
};
void kids_of(const RtmBvhHost& in, int32_t node, Kid4 out[2]) {
    const float4 q0 = in.nodes[4 * (size_t)node], q1 = in.nodes[4 * (size_t)node + 1], q2 = in.nodes[4 * (size_t)node + 2],
                 q3 = in.nodes[4 * (size_t)node + 3];
    memcpy(&out[0].code, &q3.x, 4);
    memcpy(&out[1].code, &q3.y, 4);
    out[0].lo[0] = q0.x; out[0].hi[0] = q0.y; out[0].lo[1] = q0.z; out[0].hi[1] = q0.w; out[0].lo[2] = q2.x; out[0].hi[2] = q2.y;
    out[1].lo[0] = q1.x; out[1].hi[0] = q1.y; out[1].lo[1] = q1.z; out[1].hi[1] = q1.w; out[1].lo[2] = q2.z; out[1].hi[2] = q2.w;
    out[0].rmin = q3.z;
    out[1].rmin = q3.w;
}
}  // namespace

// ---- compact half-precision form (bvh.h: RtmBvh16Host) ---------------------------------------------------------
namespace {
// IEEE binary16 from a finite double, round to nearest even / toward +infinity (up = true).  Returns 0x7C00 (inf) on overflow.
uint16_t to_half(double v, bool up) {
    const uint16_t sign = std::signbit(v) ? 0x8000 : 0;
    const double a = std::fabs(v);
    if (a == 0.0) return sign;
    int e;
    const double m = std::frexp(a, &e);          // a = m 2^e, m in [0.5, 1)
    const int E = e - 1;                         // a = (2m) 2^E, 2m in [1, 2)
    double q;                                    // fraction of a in units of the half's ulp at that magnitude
    int bias_e;
    if (E < -14) { q = std::ldexp(a, 24); bias_e = 0; }          // subnormal: ulp = 2^-24
    else { q = std::ldexp(2.0 * m - 1.0, 10); bias_e = E + 15; } // 10 fraction bits
    double f = std::floor(q);
    const double rem = q - f;
    // toward +infinity: a positive value moves to the next half, a negative one is truncated
    const bool round_up = up ? (rem > 0.0 && !sign) : (rem > 0.5 || (rem == 0.5 && std::fmod(f, 2.0) == 1.0));
    if (round_up) f += 1.0;
    uint32_t bits;
    if (bias_e == 0) bits = (uint32_t)f;                          // may carry into the smallest normal: still correct
    else {
        uint32_t frac = (uint32_t)f;
        if (frac == 1024u) { frac = 0; ++bias_e; }
        if (bias_e >= 31) return (uint16_t)(sign | 0x7C00);
        bits = ((uint32_t)bias_e << 10) | frac;
    }
    return (uint16_t)(sign | bits);
}
double from_half(uint16_t h) {
    const int e = (h >> 10) & 31, f = h & 1023;
    const double v = (e == 0) ? std::ldexp((double)f, -24) : (e == 31 ? INFINITY : std::ldexp(1.0 + f / 1024.0, e - 15));
    return (h & 0x8000) ? -v : v;
}
}  // namespace

void rtm_bvh_compact16(const RtmBvhHost& in, RtmBvh16Host& out) {
    out = RtmBvh16Host();
    if (in.root < 0 || in.root == RTM_BVH_NONE) return;           // empty tree or a single leaf: nothing to walk
    const size_t n_nodes = in.nodes.size() / 4;
    // frame: centre of the root box, power-of-two scale that maps the largest |coordinate - centre| into [2^13, 2^14)
    Kid4 rk[2];
    kids_of(in, in.root, rk);
    double lo[3], hi[3];
    for (int a = 0; a < 3; ++a) {
        lo[a] = std::min((double)rk[0].lo[a], (double)rk[1].lo[a]);
        hi[a] = std::max((double)rk[0].hi[a], (double)rk[1].hi[a]);
        out.ref[a] = (float)(0.5 * (lo[a] + hi[a]));
    }
    double ext = 0.0;
    for (int a = 0; a < 3; ++a) ext = std::max(ext, std::max(std::fabs(lo[a] - out.ref[a]), std::fabs(hi[a] - out.ref[a])));
    if (!(ext > 0.0) || !std::isfinite(ext)) return;
    int e;
    std::frexp(ext, &e);                                          // ext < 2^e
    const double scale = std::ldexp(1.0, 14 - e);                 // ext * scale < 2^14
    out.inv_scale = (float)(1.0 / scale);
    out.nodes.assign(n_nodes * 8, 0u);
    std::vector<double> infl;
    for (size_t i = 0; i < n_nodes; ++i) {
        Kid4 k[2];
        kids_of(in, (int32_t)i, k);
        uint32_t* w = out.nodes.data() + 8 * i;
        for (int j = 0; j < 2; ++j) {
            uint16_t c16[3], h16[3];
            double worst = 0.0;
            for (int a = 0; a < 3; ++a) {
                const double l = ((double)k[j].lo[a] - out.ref[a]) * scale, h = ((double)k[j].hi[a] - out.ref[a]) * scale;
                c16[a] = to_half(0.5 * (l + h), false);
                const double c = from_half(c16[a]);
                const double need = std::max(h - c, c - l);       // half-extent that covers [l, h] around the ROUNDED centre
                h16[a] = to_half(need * (1.0 + 1e-7), true);
                if (h16[a] == 0x7C00 || (c16[a] & 0x7FFF) == 0x7C00) { out.nodes.clear(); return; }
                const double true_h = 0.5 * (h - l);
                if (true_h > 0.0) worst = std::max(worst, (from_half(h16[a]) - true_h) / true_h);
            }
            if (k[j].code < 0) infl.push_back(worst);
            w[3 * j + 0] = (uint32_t)c16[0] | ((uint32_t)c16[1] << 16);
            w[3 * j + 1] = (uint32_t)c16[2] | ((uint32_t)h16[0] << 16);
            w[3 * j + 2] = (uint32_t)h16[1] | ((uint32_t)h16[2] << 16);
            w[6 + j] = (uint32_t)k[j].code;
        }
    }
    if (!infl.empty()) {
        std::nth_element(infl.begin(), infl.begin() + infl.size() / 2, infl.end());
        out.inflation = infl[infl.size() / 2];
    }
    out.ok = true;
}

// ---- 4-wide compact form (bvh.h: RtmBvh16x4Host) -----------------------------------------------------------------
void rtm_bvh_compact16x4(const RtmBvhHost& in, const RtmBvh16Host& c16, RtmBvh16x4Host& out) {
    out = RtmBvh16x4Host();
    if (!c16.ok || in.root < 0 || in.root == RTM_BVH_NONE) return;
    struct Wide {
        int32_t code;      // the two-child tree's child code
        double area;
        uint32_t w[3];     // its half-precision box: the words of the two-child compact node that holds it
    };
    auto kids2 = [&](int32_t node, Wide k[2]) {
        Kid4 f[2];
        kids_of(in, node, f);
        for (int j = 0; j < 2; ++j) {
            k[j].code = f[j].code;
            k[j].area = f[j].area();
            for (int a = 0; a < 3; ++a) k[j].w[a] = c16.nodes[8 * (size_t)node + 3 * j + a];
        }
    };
    std::vector<int32_t> src{in.root}, level{0};
    for (size_t k = 0; k < src.size(); ++k) {
        std::vector<Wide> kids(2);
        kids2(src[k], kids.data());
        while (kids.size() < 4) {
            int pick = -1;
            double best = -1.0;
            for (size_t i = 0; i < kids.size(); ++i)
                if (kids[i].code >= 0 && kids[i].area > best) {
                    best = kids[i].area;
                    pick = (int)i;
                }
            if (pick < 0) break;
            Wide two[2];
            kids2(kids[pick].code, two);
            kids[pick] = two[0];
            kids.push_back(two[1]);
        }
        uint32_t w[16];
        for (int j = 0; j < 4; ++j) {
            uint32_t* box = w + 8 * (j >> 1) + 3 * (j & 1);
            uint32_t& code = w[8 * (j >> 1) + 6 + (j & 1)];
            if (j < (int)kids.size()) {
                for (int a = 0; a < 3; ++a) box[a] = kids[j].w[a];
                if (kids[j].code >= 0) {
                    code = (uint32_t)src.size();                   // index the child will get
                    src.push_back(kids[j].code);
                    level.push_back(level[k] + 1);
                } else {
                    code = (uint32_t)kids[j].code;                 // leaf: the two-child tree's leaf code
                    out.depth = std::max(out.depth, level[k] + 1);
                }
            } else {
                box[0] = 0u;                                       // centre 0, half-extents -inf: near = +inf, far = -inf
                box[1] = 0xFC000000u;
                box[2] = 0xFC00FC00u;
                code = (uint32_t)RTM_BVH_NONE;
            }
        }
        out.nodes.insert(out.nodes.end(), w, w + 16);
    }
    out.ok = true;
}

// ---- 4-wide collapse (bvh.h) -----------------------------------------------------------------------------

void rtm_bvh_collapse4(const RtmBvhHost& in, RtmBvh4Host& out) {
    out.nodes.clear();
    out.root = in.root;
    out.depth = 0;
    if (in.root < 0 || in.root == RTM_BVH_NONE) return;   // empty tree or a single leaf: nothing to widen
    // breadth-first over the wide nodes; each queue entry is the two-child node the wide node is grown from
    std::vector<int32_t> src{in.root}, level{0};
    for (size_t k = 0; k < src.size(); ++k) {
        std::vector<Kid4> kids(2);
        kids_of(in, src[k], kids.data());
        while (kids.size() < 4) {
            int pick = -1;
            double best = -1.0;
            for (size_t i = 0; i < kids.size(); ++i)
                if (kids[i].code >= 0 && kids[i].area() > best) {
                    best = kids[i].area();
                    pick = (int)i;
                }
            if (pick < 0) break;
            Kid4 two[2];
            kids_of(in, kids[pick].code, two);
            kids[pick] = two[0];
            kids.push_back(two[1]);
        }
        float v[8][4];
        int32_t code[4];
        for (int j = 0; j < 4; ++j) {
            const bool real = j < (int)kids.size();
            for (int a = 0; a < 3; ++a) {
                v[2 * a][j] = real ? kids[j].lo[a] : 1e30f;       // an empty slot: a far, empty box and the NONE code
                v[2 * a + 1][j] = real ? kids[j].hi[a] : 1e30f;
            }
            v[7][j] = real ? kids[j].rmin : 1.f;
            code[j] = RTM_BVH_NONE;
            if (real) {
                if (kids[j].code >= 0) {
                    code[j] = (int32_t)src.size();                 // index the child will get
                    src.push_back(kids[j].code);
                    level.push_back(level[k] + 1);
                } else {
                    code[j] = kids[j].code;                        // leaf: the two-child tree's leaf code
                    out.depth = std::max(out.depth, level[k] + 1);
                }
            }
        }
        memcpy(v[6], code, 16);
        for (int j = 0; j < 8; ++j) out.nodes.push_back(make_float4(v[j][0], v[j][1], v[j][2], v[j][3]));
    }
    out.root = 0;
}
