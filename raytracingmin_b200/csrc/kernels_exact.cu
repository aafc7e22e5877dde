// kernels_exact.cu — the FP64 kernels whose results are compared BIT-FOR-BIT with
// the reference: camera-ray generation, ray/sphere test, closest-hit search,
// the L-HEAD image, and the reference-faithful FP64 bounce loop.
//
// This translation unit is compiled with -fmad=false: the reference is built
// without FMA contraction (x86-64 SSE2), and one fused multiply-add would change
// hit indices.  IEEE sqrt/div for float and double are nvcc's defaults
// (-prec-sqrt=true -prec-div=true).  Citations are to /root/reference/src.
#include "rtm_internal.h"
#include "philox.cuh"

#include <float.h>

namespace {

struct d3 {
    double x, y, z;
};
__device__ __forceinline__ d3 mk(double x, double y, double z) { return d3{x, y, z}; }
__device__ __forceinline__ d3 add(d3 a, d3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ d3 sub(d3 a, d3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ d3 mul(d3 a, d3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ d3 scale(d3 a, double s) { return mk(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ d3 divs(d3 a, double s) { return mk(a.x / s, a.y / s, a.z / s); }
// Ray.h:61-63: (ax*bx + ay*by) + az*bz
__device__ __forceinline__ double dot(d3 a, d3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// Ray.h:64-66: the y component is written -a.x*b.z + a.z*b.x
__device__ __forceinline__ d3 cross(d3 a, d3 b) {
    return mk(a.y * b.z - a.z * b.y, -a.x * b.z + a.z * b.x, a.x * b.y - a.y * b.x);
}
// Ray.h:67-72: Magnitude is sqrtf of the double sum (float argument, float
// sqrt); Normalize is three divides by that float-rounded length.
__device__ __forceinline__ d3 normalize(d3 a) {
    const double m = (double)__fsqrt_rn(__double2float_rn(a.x * a.x + a.y * a.y + a.z * a.z));
    return divs(a, m);
}

// SettingData.cpp:197-223 (distance only; the normal is derived for the winner).
__device__ __forceinline__ bool sphere_test(d3 c, double r2, d3 o, d3 d, double& t) {
    const d3 p_o = sub(c, o);
    const double b = dot(p_o, d);
    const double D4 = b * b - dot(p_o, p_o) + r2;
    if (D4 < 0.0) return false;  // a NaN discriminant passes, as in the reference
    const double s = sqrt(D4);
    const double t1 = b - s, t2 = b + s;
    const double minValue = (double)1e-5f;  // `const float minValue = 1e-5`
    if (t1 < minValue && t2 < minValue) return false;
    t = (t1 > 0.001) ? t1 : t2;
    return true;
}

// PlaneObject EXTENSION (no reference semantics: SettingData.cpp:243-246 only rejects parallel
// rays and falls off the end).  Finite square, operation for operation the definition in DESIGN.md §10:
// intersect_plane: t = n.(position - o) / n.d, accepted for t > 0.001 inside |q.right| <= rr,
// |q.upp| <= uu.  The frame (normal, right, upp, rr, uu) is derived on the host.
__device__ __forceinline__ bool plane_test(const double* __restrict__ pl, d3 o, d3 d, double& t) {
    const d3 n = mk(__ldg(pl + 0), __ldg(pl + 1), __ldg(pl + 2));
    const double dn = dot(n, d);
    if (fabs(dn) < (double)FLT_EPSILON) return false;   // SettingData.cpp:244-245
    const d3 pos = mk(__ldg(pl + 3), __ldg(pl + 4), __ldg(pl + 5));
    const double tt = dot(n, sub(pos, o)) / dn;
    if (!(tt > 0.001)) return false;
    const d3 q = sub(add(o, scale(d, tt)), pos);
    if (fabs(dot(q, mk(__ldg(pl + 6), __ldg(pl + 7), __ldg(pl + 8)))) > __ldg(pl + 12)) return false;
    if (fabs(dot(q, mk(__ldg(pl + 9), __ldg(pl + 10), __ldg(pl + 11)))) > __ldg(pl + 13)) return false;
    t = tt;
    return true;
}

// planes follow the spheres in the object list: strict `<` keeps a sphere on an exact tie
__device__ __forceinline__ void consider_planes(const RtmSceneF64& sc, d3 o, d3 d, double& best, int& id) {
    for (int j = 0; j < sc.n_planes; ++j) {
        double t;
        if (plane_test(sc.planes + 16 * j, o, d, t) && t < best && t > 0.0) {
            best = t;
            id = sc.n + j;
        }
    }
}

// Renderer.cpp:58-73: strict `<` keeps the lowest index on exact ties; `t > 0`.
__device__ __forceinline__ int closest_hit_brute(const RtmSceneF64& sc, d3 o, d3 d, double& best) {
    int id = -1;
    best = DBL_MAX;
    for (int i = 0; i < sc.n; ++i) {
        double t;
        if (sphere_test(mk(__ldg(sc.cx + i), __ldg(sc.cy + i), __ldg(sc.cz + i)), __ldg(sc.r2 + i), o, d, t) &&
            t < best && t > 0.0) {
            best = t;
            id = i;
        }
    }
    consider_planes(sc, o, d, best, id);
    return id;
}

// ---- the same search behind the hierarchy ------------------------------------------
// Candidates are visited in tree order, so the sequential rule "t < best, lowest index
// on ties" becomes the order-independent "(t, i) lexicographically smallest": identical
// winner.  What must hold for bit-exactness is that no node holding the winner is culled.
// A candidate reported by sphere_test at parameter t sits at x = o + t d with
//     |x - c|^2 = r^2 + t^2 (|d|^2 - 1)
// (t solves the UNIT-direction quadratic, but Normalize() divides by a float-rounded
// length, Ray.h:67-72, so |d|^2 = 1 +- 1.8e-7): x can lie OUTSIDE the sphere by
// t^2 eps / 2r.  With t <= 3 D (D = distance from o to the farthest corner of the node's
// box, which holds c) the box is inflated per ray by 4.5 D^2 eps / rmin before the slab
// test; eps is this ray's own | |d|^2 - 1 |.  Rays that are not unit to 4e-7 (only the
// debug hook can supply them) take the brute-force loop.
struct BvhWork {
    unsigned tests, nodes;
};
__device__ __forceinline__ void consider(const RtmSceneF64& sc, int i, d3 o, d3 d, double& best, int& id) {
    double t;
    if (sphere_test(mk(__ldg(sc.cx + i), __ldg(sc.cy + i), __ldg(sc.cz + i)), __ldg(sc.r2 + i), o, d, t) && t > 0.0 &&
        (t < best || (t == best && i < id))) {
        best = t;
        id = i;
    }
}

__device__ __forceinline__ bool box_test_f64(double lx, double hx, double ly, double hy, double lz, double hz, double rmin,
                                             d3 o, d3 inv, double ke, double best, double& tnear) {
    const double fx = fmax(fabs(lx - o.x), fabs(hx - o.x)), fy = fmax(fabs(ly - o.y), fabs(hy - o.y)),
                 fz = fmax(fabs(lz - o.z), fabs(hz - o.z));
    const double infl = ke * (fx * fx + fy * fy + fz * fz) / rmin;
    // fmin/fmax return the non-NaN operand: a 0*inf slab (ray inside a face plane of the
    // padded box, which no sphere touches) degrades to the other bound.
    double t1 = (lx - infl - o.x) * inv.x, t2 = (hx + infl - o.x) * inv.x;
    double tmin = fmax(0.0, fmin(t1, t2)), tmax = fmin(best, fmax(t1, t2));
    t1 = (ly - infl - o.y) * inv.y; t2 = (hy + infl - o.y) * inv.y;
    tmin = fmax(tmin, fmin(t1, t2)); tmax = fmin(tmax, fmax(t1, t2));
    t1 = (lz - infl - o.z) * inv.z; t2 = (hz + infl - o.z) * inv.z;
    tmin = fmax(tmin, fmin(t1, t2)); tmax = fmin(tmax, fmax(t1, t2));
    tnear = tmin;
    return tmin <= tmax;
}

__device__ int closest_hit_bvh(const RtmSceneF64& sc, d3 o, d3 d, double& best, BvhWork* work = nullptr) {
    const double dd = d.x * d.x + d.y * d.y + d.z * d.z;
    if (!(fabs(dd - 1.0) <= 4e-7)) {
        if (work) work->tests += sc.n;
        return closest_hit_brute(sc, o, d, best);
    }
    if (work) work->tests += sc.bvh.n_big;
    int id = -1;
    best = DBL_MAX;
    for (int j = 0; j < sc.bvh.n_big; ++j) consider(sc, sc.bvh.big_id[j], o, d, best, id);
    int node = sc.bvh.root;
    if (node == 0x7FFFFFFF) {
        consider_planes(sc, o, d, best, id);
        return id;
    }
    const double ke = 4.5 * (fabs(dd - 1.0) + 1e-14);
    const d3 orel = mk(o.x - sc.ref[0], o.y - sc.ref[1], o.z - sc.ref[2]);
    const d3 inv = mk(1.0 / d.x, 1.0 / d.y, 1.0 / d.z);
    int stack[64];
    int sp = 0;
    stack[sp++] = 0x7FFFFFFF;
    while (node != 0x7FFFFFFF) {
        if (node >= 0) {
            const float4 q0 = __ldg(sc.bvh.nodes + 4 * node), q1 = __ldg(sc.bvh.nodes + 4 * node + 1),
                         q2 = __ldg(sc.bvh.nodes + 4 * node + 2), q3 = __ldg(sc.bvh.nodes + 4 * node + 3);
            double n0, n1;
            const bool h0 = box_test_f64(q0.x, q0.y, q0.z, q0.w, q2.x, q2.y, q3.z, orel, inv, ke, best, n0);
            const bool h1 = box_test_f64(q1.x, q1.y, q1.z, q1.w, q2.z, q2.w, q3.w, orel, inv, ke, best, n1);
            const int c0 = __float_as_int(q3.x), c1 = __float_as_int(q3.y);
            if (work) ++work->nodes;
            if (h0 && h1) {
                const bool swap = n1 < n0;
                stack[sp++] = swap ? c0 : c1;
                node = swap ? c1 : c0;
            } else if (h0 || h1) {
                node = h0 ? c0 : c1;
            } else {
                node = stack[--sp];
            }
        } else {
            const unsigned code = ~(unsigned)node;
            const unsigned first = code >> 3, cnt = (code & 7u) + 1u;
            for (unsigned k = 0; k < cnt; ++k) consider(sc, __ldg(sc.bvh.leaf_id + first + k), o, d, best, id);
            if (work) work->tests += cnt;
            node = stack[--sp];
        }
    }
    consider_planes(sc, o, d, best, id);
    return id;
}

__device__ __forceinline__ int closest_hit(const RtmSceneF64& sc, d3 o, d3 d, double& best, BvhWork* work = nullptr) {
    return sc.use_bvh ? closest_hit_bvh(sc, o, d, best, work) : closest_hit_brute(sc, o, d, best);
}

__device__ __forceinline__ d3 sphere_center(const RtmSceneF64& sc, int i) {
    return mk(__ldg(sc.cx + i), __ldg(sc.cy + i), __ldg(sc.cz + i));
}

// the normal Intersect hands back: SettingData.cpp:214-215 for a sphere, the plane's own normal (extension)
__device__ __forceinline__ d3 hit_normal(const RtmSceneF64& sc, int id, d3 o, d3 d, double t) {
    if (id >= sc.n) {
        const double* pl = sc.planes + 16 * (id - sc.n);
        return mk(__ldg(pl + 0), __ldg(pl + 1), __ldg(pl + 2));
    }
    return normalize(sub(add(o, scale(d, t)), sphere_center(sc, id)));
}

// Renderer.cpp:227-232.  `rate` is float, rate*sx is a float product, the
// 2.0f / 1.0f literals widen to double; evaluation is left to right.
__device__ __forceinline__ d3 camera_dir(const RtmFrame& f, int x, int y, int sx, int sy) {
    const float fx = f.rate * (float)sx, fy = f.rate * (float)sy;
    const double px = 2.0 * ((double)x + (double)fx) / (double)f.W - 1.0;
    const double py = 2.0 * ((double)y + (double)fy) / (double)f.H - 1.0;
    const d3 v = mk(f.ax[0] * px + f.ay[0] * py + f.az[0], f.ax[1] * px + f.ay[1] * py + f.az[1],
                    f.ax[2] * px + f.ay[2] * py + f.az[2]);
    return normalize(v);
}

// local record index ql (pixel-major over this handle's slice of the sub-pixel grid) ->
// pixel, sub-pixel coordinates (1-based as src/Renderer.cpp:224-225) and the GLOBAL
// sub-pixel index qg = pixel*ss*ss + (sx-1)*ss + (sy-1) that keys the RNG and the hit map.
__device__ __forceinline__ void decode_sub(const RtmPassParams& p, uint64_t ql, int& x, int& y, int& sx, int& sy, uint64_t& qg) {
    const uint32_t ss = (uint32_t)p.frame.ss, sc = (uint32_t)p.sub_count;
    const uint32_t pix = (uint32_t)(ql / sc);
    const uint32_t j = (uint32_t)p.sub_begin + (uint32_t)(ql % sc);
    sy = (int)(j % ss) + 1;
    sx = (int)(j / ss) + 1;
    x = (int)(pix % (uint32_t)p.frame.W);
    y = (int)(pix / (uint32_t)p.frame.W);
    qg = (uint64_t)pix * (ss * ss) + j;
}

// Renderer.cpp:43-49 with std::max/std::min semantics
__device__ __forceinline__ double clamp01(double v) {
    const double lo = (v < 0.0) ? 0.0 : v;
    return (1.0 < lo) ? 1.0 : lo;
}

// ---- K1: primary pass, one thread per sub-pixel ----------------------------
__device__ __forceinline__ void primary_one(const RtmSceneF64& sc, const RtmPassParams& p, uint64_t q, int32_t* __restrict__ hits,
                                            float4* __restrict__ records, float* __restrict__ cps0, BvhWork* work) {
    int x, y, sx, sy;
    uint64_t qg;
    decode_sub(p, q, x, y, sx, sy, qg);
    const d3 o = mk(p.frame.o[0], p.frame.o[1], p.frame.o[2]);
    const d3 d = camera_dir(p.frame, x, y, sx, sy);
    double t;
    const int id = closest_hit(sc, o, d, t, work);
    if (hits) hits[qg] = id;
    if (records) {
        float4 rec;
        if (id < 0) {
            rec = make_float4(0.f, 0.f, 0.f, __uint_as_float(RTM_REC_MISS));
        } else {
            const d3 hp = add(scale(d, t), o);                          // Renderer.cpp:79
            const d3 n = hit_normal(sc, id, o, d, t);
            const uint32_t flip = (dot(n, d) < 0.0) ? 0u : RTM_REC_FLIP;   // Renderer.cpp:82-83
            rec = make_float4((float)(hp.x - p.refd[0]), (float)(hp.y - p.refd[1]), (float)(hp.z - p.refd[2]),
                              __uint_as_float((uint32_t)id | flip));
            if (cps0 && id >= sc.n) cps0[q] = 0.f;
            else if (cps0) {
                // what SettingData.cpp:198-200 will see for THIS sphere from the bounce origin hp:
                // c' = |p_o|^2 - r^2, nonzero because d is unit only to float precision (Ray.h:67-72)
                const d3 p_o = sub(sphere_center(sc, id), hp);
                cps0[q] = (float)(dot(p_o, p_o) - __ldg(sc.r2 + id));
            }
        }
        records[q] = rec;
    }
}

__global__ void __launch_bounds__(256) primary_f64_kernel(RtmSceneF64 sc, RtmPassParams p, int32_t* __restrict__ hits,
                                                          float4* __restrict__ records, RtmCounters* __restrict__ counters,
                                                          float* __restrict__ cps0) {
    const uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    BvhWork work{0u, 0u};
    if (q < p.n_sub) primary_one(sc, p, q, hits, records, cps0, sc.use_bvh ? &work : nullptr);
    if (sc.use_bvh && counters) {   // hierarchy: tests / nodes actually executed, one atomic pair per warp
        unsigned long long t = work.tests, n = work.nodes;
        for (int off = 16; off > 0; off >>= 1) {
            t += __shfl_down_sync(0xffffffffu, t, off);
            n += __shfl_down_sync(0xffffffffu, n, off);
        }
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(&counters->tests, t);
            atomicAdd(&counters->nodes, n);
        }
    }
}

// ---- L-HEAD image: one thread per pixel -------------------------------------
// At HEAD the hit normal never reaches PathTracing (SettingData.h:21 takes it by
// value), the bounce ray is NaN and misses, so every sample is the emission of
// the primary hit (Renderer.cpp:109,112); accumulation as Renderer.cpp:223-248.
__global__ void __launch_bounds__(128) head_image_kernel(RtmSceneF64 sc, RtmPassParams p, const int32_t* __restrict__ hits,
                                                         double* __restrict__ image64, float4* __restrict__ accum) {
    const uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t npix = (uint32_t)p.frame.W * (uint32_t)p.frame.H;
    if (pix >= npix) return;
    const int ss = p.frame.ss;
    d3 acc = mk(0, 0, 0);
    for (int j = 0; j < ss * ss; ++j) {
        const int id = hits[(uint64_t)pix * (ss * ss) + j];
        d3 cal = (id < 0) ? mk(0, 0, 0)
                          : mk(__ldg(sc.emission + 3 * id), __ldg(sc.emission + 3 * id + 1), __ldg(sc.emission + 3 * id + 2));
        cal = divs(divs(divs(cal, (double)ss), (double)ss), (double)p.frame.S);  // Renderer.cpp:240
        cal = mk(clamp01(cal.x), clamp01(cal.y), clamp01(cal.z));               // :241
        for (int s = p.s_begin; s < p.s_end; ++s) acc = add(acc, cal);          // :242
    }
    if (image64) {
        image64[3 * (uint64_t)pix + 0] = acc.x;
        image64[3 * (uint64_t)pix + 1] = acc.y;
        image64[3 * (uint64_t)pix + 2] = acc.z;
    }
    accum[pix] = make_float4((float)acc.x, (float)acc.y, (float)acc.z, 0.f);
}

// ---- reference-faithful FP64 bounce loop: one thread per sub-pixel -----------
// Renderer.cpp:57-117 line by line (L-ABC), the recursion unrolled forward:
// L = e0 + k0*(e1 + k1*(...)) is accumulated as L += T*e, T *= k.
__global__ void __launch_bounds__(128) radiance_f64_kernel(RtmSceneF64 sc, RtmPassParams p, float4* __restrict__ records,
                                                           RtmCounters* __restrict__ counters) {
    const uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long rays = 0;
    BvhWork work{0u, 0u};
    if (q < p.n_sub) {
        int x, y, sx, sy;
        uint64_t qg;
        decode_sub(p, q, x, y, sx, sy, qg);
        const d3 o0 = mk(p.frame.o[0], p.frame.o[1], p.frame.o[2]);
        const d3 d0 = camera_dir(p.frame, x, y, sx, sy);
        const double PI = 3.14159265358979323846;
        d3 acc = mk(0, 0, 0);
        for (int s = p.s_begin; s < p.s_end; ++s) {
            d3 o = o0, d = d0, L = mk(0, 0, 0), T = mk(1, 1, 1);
            for (int depth = 0; depth < p.max_depth; ++depth) {
                ++rays;
                double t;
                const int id = closest_hit(sc, o, d, t, &work);
                if (id < 0) break;
                const d3 e = mk(__ldg(sc.emission + 3 * id), __ldg(sc.emission + 3 * id + 1), __ldg(sc.emission + 3 * id + 2));
                L = add(L, mul(T, e));
                const uint4 r = rtm_philox4x32_10(make_uint4((uint32_t)qg, (uint32_t)(qg >> 32), (uint32_t)s, (uint32_t)depth),
                                                  p.seed_lo, p.seed_hi);
                const double kd = __ldg(sc.kd + id);
                if (!(rtm_u01_f64(r.x, r.y) <= kd)) break;  // :78 / :112
                const d3 col = mk(__ldg(sc.color + 3 * id), __ldg(sc.color + 3 * id + 1), __ldg(sc.color + 3 * id + 2));
                T = mul(T, divs(col, kd));                   // colorKD(), SettingData.h:11-13
                const d3 hp = add(scale(d, t), o);           // :79
                const d3 n = hit_normal(sc, id, o, d, t);
                const d3 w = (dot(n, d) < 0.0) ? n : scale(n, -1.0);  // :82-83,87
                const double r1 = 2 * PI * rtm_u01_f64_32(r.z);       // :88
                const double r2 = rtm_u01_f64_32(r.w);                // :89
                const double r2s = sqrt(r2);                          // :90
                d3 u;
                if (fabs(w.x) > (double)FLT_MIN) u = normalize(cross(mk(0, 1, 0), w));  // :96-101
                else u = normalize(cross(mk(1, 0, 0), w));
                const d3 v = cross(w, u);                                               // :102
                d = normalize(add(add(scale(scale(u, cos(r1)), r2s), scale(scale(v, sin(r1)), r2s)),
                                  scale(w, sqrt(1.0 - r2))));                          // :103-107
                o = hp;
            }
            d3 cal = divs(divs(divs(L, (double)p.frame.ss), (double)p.frame.ss), (double)p.frame.S);
            acc = add(acc, mk(clamp01(cal.x), clamp01(cal.y), clamp01(cal.z)));
        }
        records[q] = make_float4((float)acc.x, (float)acc.y, (float)acc.z, 0.f);
    }
    // one atomic per warp
    unsigned long long tests = work.tests, nodes = work.nodes;
    for (int off = 16; off > 0; off >>= 1) {
        rays += __shfl_down_sync(0xffffffffu, rays, off);
        tests += __shfl_down_sync(0xffffffffu, tests, off);
        nodes += __shfl_down_sync(0xffffffffu, nodes, off);
    }
    if ((threadIdx.x & 31) == 0 && rays) {
        atomicAdd(&counters->rays_ref, rays);
        atomicAdd(&counters->rays_traced, rays);
        atomicAdd(&counters->tests, sc.use_bvh ? tests : rays * (unsigned long long)sc.n);
        if (sc.use_bvh) atomicAdd(&counters->nodes, nodes);
    }
}

// ---- test hook: arbitrary rays through the exact closest-hit loop -----------
__global__ void closest_hit_f64_kernel(RtmSceneF64 sc, uint32_t n, const double* __restrict__ org,
                                       const double* __restrict__ dir, int32_t* __restrict__ hit, double* __restrict__ tout) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double t;
    const int id = closest_hit(sc, mk(org[3 * i], org[3 * i + 1], org[3 * i + 2]), mk(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]), t);
    hit[i] = id;
    tout[i] = t;
}

}  // namespace

void rtm_launch_primary_f64(const RtmSceneF64& sc, const RtmPassParams& p, int32_t* hits, float4* records,
                            RtmCounters* counters, float* cps0, cudaStream_t st) {
    const uint32_t blocks = (uint32_t)((p.n_sub + 255) / 256);
    primary_f64_kernel<<<blocks, 256, 0, st>>>(sc, p, hits, records, counters, cps0);
}

void rtm_launch_head_image(const RtmSceneF64& sc, const RtmPassParams& p, const int32_t* hits, double* image64,
                           float4* accum, cudaStream_t st) {
    const uint32_t npix = (uint32_t)p.frame.W * (uint32_t)p.frame.H;
    head_image_kernel<<<(npix + 127) / 128, 128, 0, st>>>(sc, p, hits, image64, accum);
}

void rtm_launch_radiance_f64(const RtmSceneF64& sc, const RtmPassParams& p, float4* records, RtmCounters* counters,
                             cudaStream_t st) {
    const uint32_t blocks = (uint32_t)((p.n_sub + 127) / 128);
    radiance_f64_kernel<<<blocks, 128, 0, st>>>(sc, p, records, counters);
}

void rtm_launch_closest_hit_f64(const RtmSceneF64& sc, uint32_t n, const double* org, const double* dir, int32_t* hit,
                                double* t, cudaStream_t st) {
    closest_hit_f64_kernel<<<(n + 127) / 128, 128, 0, st>>>(sc, n, org, dir, hit, t);
}
