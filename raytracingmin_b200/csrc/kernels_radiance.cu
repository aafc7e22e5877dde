// kernels_radiance.cu — the FP32 bounce loop (product path), the sub-pixel ->
// pixel accumulation pass and the 8-bit conversion.
//
// What the reference does per sample (src/Renderer.cpp:57-117, L-ABC): closest
// hit over all spheres, Russian roulette on kd, cosine-weighted bounce, recurse.
// Here: a persistent megakernel.  Each lane owns one sub-pixel at a time (its
// exact primary hit comes from the FP64 pass), walks that sub-pixel's samples
// one hit per loop iteration, and pulls the next sub-pixel from a warp-level
// work cursor as soon as it finishes, so a warp never waits for its longest path.
//
// The sphere test is NOT the reference's formula evaluated in float (that flips
// 7 % of Cornell hits).  It is rearranged so that FP32 carries it:
//   * positions are relative to the camera origin; c' = |o-c|^2 - r^2 is expanded
//     as |o|^2 - 2 o.c + (|c|^2 - r^2), the last term precomputed in double, so the
//     1e8-magnitude squares of the r=1e4 wall spheres never meet in float;
//   * roots come from q = b + sign(b) sqrt(b^2 - c'), t = {q, c'/q}: no cancellation;
//   * a ray leaving sphere k's surface gets c'_k = 0 exactly, which is what the
//     reference's 1e-5 / 1e-3 epsilons (SettingData.cpp:208-212) achieve in double.
#include "rtm_internal.h"
#include "philox.cuh"

#include <float.h>
#include <stdlib.h>

__constant__ RtmSceneF32Host c_scene;

cudaError_t rtm_upload_scene_f32(const RtmSceneF32Host& h, cudaStream_t st) {
    return cudaMemcpyToSymbolAsync(c_scene, &h, sizeof(RtmSceneF32Host), 0, cudaMemcpyHostToDevice, st);
}

namespace {

struct f3 {
    float x, y, z;
};
__device__ __forceinline__ f3 mk(float x, float y, float z) { return f3{x, y, z}; }
__device__ __forceinline__ float dot(f3 a, f3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }

__device__ __forceinline__ float fast_sqrt(float x) {
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float fast_rcp(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

__device__ __forceinline__ float fast_rsqrt(float x) {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// ---- ray/sphere test of the FP32 path ------------------------------------------
// Semantics of SettingData.cpp:197-223 + the accept rule of Renderer.cpp:67, stated as
// "the smallest root greater than RTM_T_EPS; lowest index wins exact ties".  The
// reference takes the near root t1 when t1 > 1e-3, else the far root t2 when t2 >= 1e-5;
// the two rules differ only for a far root inside [1e-5, 1e-3], i.e. for chords shorter
// than 1e-3 (documented in DESIGN.md; invisible at the radiance tolerance).
//
// Roots without cancellation: tq = b + sign(b) sqrt(b^2 - c'), the other root is c'/tq.
// A negative discriminant makes sqrt.approx return NaN, which propagates into both
// candidates; the unsigned-integer minimum below ignores NaNs and negative values by
// construction (their bit patterns are above +inf), so no explicit D >= 0 / t > 0 tests
// are issued.
#define RTM_T_EPS 1e-3f
#define RTM_KEY_INF 0x7F800000u

__device__ __forceinline__ void sphere_roots_f32(bool self, f3 o, f3 d, float oo, float dO, float cx, float cy, float cz,
                                                 float m2cx, float m2cy, float m2cz, float K, float& ta, float& tb) {
    const float b = fmaf(d.x, cx, fmaf(d.y, cy, fmaf(d.z, cz, -dO)));           // (c-o).d
    float cp = fmaf(o.x, m2cx, fmaf(o.y, m2cy, fmaf(o.z, m2cz, oo + K)));       // |c-o|^2 - r^2
    cp = self ? 0.f : cp;                                                       // leaving this sphere's surface
    const float s = fast_sqrt(fmaf(b, b, -cp));
    const float tq = b + copysignf(s, b);
    ta = tq - RTM_T_EPS;
    tb = fmaf(cp, fast_rcp(tq), -RTM_T_EPS);
}

// ---- packed FP32 (sm_100 f32x2): two spheres per instruction ----------------------
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float lo, float hi) {
    u64 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ u64 pk(float2 v) { return pk(v.x, v.y); }
__device__ __forceinline__ void unpk(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {
    u64 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ u64 add2(u64 a, u64 b) {
    u64 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ u64 mul2(u64 a, u64 b) {
    u64 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
// Row oid+1 of this table is 1 everywhere except column oid: multiplying -c' by it zeroes
// c' of the sphere the ray leaves (2 LDS.128 + one FMUL2 per pair instead of a compare and
// a select per sphere).  Row 0 (oid = -1: no origin sphere) is all ones.
struct SelfMask {
    float4 row[9][2];
};
__device__ __forceinline__ void init_selfmask(SelfMask& m) {
    for (int i = threadIdx.x; i < 9 * 8; i += blockDim.x) {
        const int r = i >> 3, c = i & 7;
        reinterpret_cast<float*>(&m.row[0][0])[i] = (c == r - 1) ? 0.f : 1.f;
    }
}
// (x & mask) | idx in ONE LOP3: the mask sits in a register (an immediate mask plus an
// immediate index would take two instructions).
__device__ __forceinline__ unsigned key_pack(unsigned x, unsigned mask_reg, unsigned idx) {
    unsigned r;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(x), "r"(mask_reg), "r"(idx));
    return r;
}

// Closest hit for N <= 8: spheres are tested in PAIRS with FFMA2/FADD2 (one instruction
// does the same step for two spheres; the ray operand is broadcast, the sphere constants
// come as 64-bit uniform-register pairs from constant memory).  The candidate (t - eps)
// and the sphere index share one 32-bit key, t in the upper 29 bits (8 ulp of t are given
// up: below the FP32 error of t itself), the index in the low 3; positive floats order
// like unsigned integers, so an unsigned min does the accept test, the tie rule and the
// bookkeeping.  Returns the key; RTM_KEY_INF or above = miss.
// NS = number of slots that hold (or may hold) a sphere: with NS odd the upper half of the last pair is padding, and
// its scalar tail (sqrt, copysign, rcp, min, key) is not issued — 5 of ~280 instructions of a Cornell-box bounce.
template <int NS>
__device__ __forceinline__ unsigned trace_pairs_f32(const SelfMask& sm, int oid, f3 o, f3 d, unsigned key_mask) {
    constexpr int NP = (NS + 1) / 2;
    const float noo = -dot(o, o), ndO = -dot(d, o);
    const u64 dx = pk(d.x, d.x), dy = pk(d.y, d.y), dz = pk(d.z, d.z);
    // o.(2c) is formed as (2o).c — bit-identical (power-of-two scaling) and the pair needs 4 constants, not 7
    const float o2x = o.x + o.x, o2y = o.y + o.y, o2z = o.z + o.z;
    const u64 ox = pk(o2x, o2x), oy = pk(o2y, o2y), oz = pk(o2z, o2z);
    const u64 noo2 = pk(noo, noo), ndO2 = pk(ndO, ndO), neps2 = pk(-RTM_T_EPS, -RTM_T_EPS);
    unsigned best = RTM_KEY_INF;
    const float4 m0 = sm.row[oid + 1][0], m1 = sm.row[oid + 1][1];
    const u64 mask[4] = {pk(m0.x, m0.y), pk(m0.z, m0.w), pk(m1.x, m1.y), pk(m1.z, m1.w)};
#pragma unroll
    for (int j = 0; j < NP; ++j) {
        const RtmPairF32& c = c_scene.pairs[j];
        const u64 b2 = fma2(dx, pk(c.cx), fma2(dy, pk(c.cy), fma2(dz, pk(c.cz), ndO2)));                   // (c-o).d
        const u64 ncp2 = fma2(ox, pk(c.cx), fma2(oy, pk(c.cy), fma2(oz, pk(c.cz), add2(noo2, pk(c.nK)))));     // r^2 - |c-o|^2
        const u64 ncp = mul2(ncp2, mask[j]);   // leaving this sphere's surface: c' = 0 exactly
        float b0, b1;
        unpk(b2, b0, b1);
        float D0, D1;
        unpk(fma2(b2, b2, ncp), D0, D1);
        const bool pad = (NS & 1) && j == NP - 1;                    // compile-time: slot 2j+1 is padding
        const float s0 = fast_sqrt(D0), s1 = pad ? 0.f : fast_sqrt(D1);   // NaN when the ray misses the sphere
        const u64 tq2 = add2(b2, pk(copysignf(s0, b0), pad ? 0.f : copysignf(s1, b1)));
        float q0, q1;
        unpk(tq2, q0, q1);
        const u64 rn = pk(fast_rcp(-q0), pad ? 0.f : fast_rcp(-q1));
        float ta0, ta1, tb0, tb1;
        unpk(add2(tq2, neps2), ta0, ta1);                            // tq - eps
        unpk(fma2(ncp, rn, neps2), tb0, tb1);                        // c'/tq - eps
        const unsigned k0 = key_pack(min(__float_as_uint(ta0), __float_as_uint(tb0)), key_mask, 2 * j);
        if (pad) {
            best = min(best, k0);
        } else {
            const unsigned k1 = key_pack(min(__float_as_uint(ta1), __float_as_uint(tb1)), key_mask, 2 * j + 1);
            best = min(best, min(k0, k1));
        }
    }
    return best;
}

// Runtime-N version (any N that fits constant memory): explicit (t, id) pair.
__device__ __forceinline__ int trace_loop_f32(int n, int oid, f3 o, f3 d, float& t) {
    const float oo = dot(o, o), dO = dot(d, o);
    float best = FLT_MAX;
    int id = -1;
#pragma unroll 4
    for (int i = 0; i < n; ++i) {
        float ta, tb;
        sphere_roots_f32(i == oid, o, d, oo, dO, c_scene.cx[i], c_scene.cy[i], c_scene.cz[i], c_scene.m2cx[i],
                         c_scene.m2cy[i], c_scene.m2cz[i], c_scene.K[i], ta, tb);
        // candidates are (root - eps); valid iff > 0; NaN fails every comparison
        const float c1 = (ta > 0.f) ? ta : FLT_MAX, c2 = (tb > 0.f) ? tb : FLT_MAX;
        const float c = fminf(c1, c2);
        const bool ok = c < best;
        best = ok ? c : best;
        id = ok ? i : id;
    }
    t = best + RTM_T_EPS;
    return id;
}

template <int N>
__device__ __forceinline__ int trace_f32(const SelfMask& sm, int n_rt, int oid, f3 o, f3 d, float& t, unsigned key_mask) {
    if (N > 0) {
        const unsigned key = trace_pairs_f32<N>(sm, oid, o, d, key_mask);
        t = __uint_as_float(key & ~7u) + RTM_T_EPS;
        return (key < RTM_KEY_INF) ? (int)(key & 7u) : -1;
    }
    return trace_loop_f32(n_rt, oid, o, d, t);
}

// ---- PlaneObject extension ---------------------------------------
// Finite squares from constant memory, two per instruction (f32x2) like the spheres, branch-free: for
// each pair t = (d0 - n.o) * rcp(n.d), the point q = o + t d - position, the two half-extent dots, and
// one predicate per plane (inside, not parallel, t > eps, not the plane the ray leaves, closer than
// the best so far) that selects (t, id).  Tested after the spheres with a strict `<`, so a sphere keeps
// an exact tie.  The plane a ray leaves is skipped by id (the FP64 definition rejects it through
// t > 1e-3: t is 0 there).
__device__ __forceinline__ void trace_planes_f32(int n_sph, int n_pl, int oid, f3 o, f3 d, float& t, int& id) {
    const u64 dx = pk(d.x, d.x), dy = pk(d.y, d.y), dz = pk(d.z, d.z);
    const u64 ox = pk(o.x, o.x), oy = pk(o.y, o.y), oz = pk(o.z, o.z);
    const u64 nox = pk(-o.x, -o.x), noy = pk(-o.y, -o.y), noz = pk(-o.z, -o.z);
    const int self = oid - n_sph;   // slot of the plane the ray leaves (negative: none)
    const int np = (n_pl + 1) >> 1;
    // unrolled with a uniform guard: constant indices let the pair's constants come straight from the
    // constant bank (LDCU / c[][] operands) instead of one indexed LDC per value
#pragma unroll
    for (int j = 0; j < RTM_MAX_PLANES / 2; ++j) {
        if (j >= np) break;
        const RtmPlanePairF32& c = c_scene.plane_pairs[j];
        const u64 nx = pk(c.nx), ny = pk(c.ny), nz = pk(c.nz);
        float dn0, dn1;
        unpk(fma2(nx, dx, fma2(ny, dy, mul2(nz, dz))), dn0, dn1);                       // n.d
        const u64 num2 = fma2(nx, nox, fma2(ny, noy, fma2(nz, noz, pk(c.d0))));         // n.(position - o)
        const u64 tt2 = mul2(num2, pk(fast_rcp(dn0), fast_rcp(dn1)));
        const u64 qx = fma2(dx, tt2, add2(ox, pk(c.npx))), qy = fma2(dy, tt2, add2(oy, pk(c.npy))),
                  qz = fma2(dz, tt2, add2(oz, pk(c.npz)));
        float a0, a1, b0, b1, t0, t1;
        unpk(fma2(qx, pk(c.rx), fma2(qy, pk(c.ry), mul2(qz, pk(c.rz)))), a0, a1);
        unpk(fma2(qx, pk(c.ux), fma2(qy, pk(c.uy), mul2(qz, pk(c.uz)))), b0, b1);
        unpk(tt2, t0, t1);
        const bool v0 = fmaxf(fabsf(a0), fabsf(b0)) <= 1.f && fabsf(dn0) >= FLT_EPSILON && t0 > RTM_T_EPS && 2 * j != self && t0 < t;
        t = v0 ? t0 : t;
        id = v0 ? n_sph + 2 * j : id;
        const bool v1 = fmaxf(fabsf(a1), fabsf(b1)) <= 1.f && fabsf(dn1) >= FLT_EPSILON && t1 > RTM_T_EPS && 2 * j + 1 != self && t1 < t;
        t = v1 ? t1 : t;
        id = v1 ? n_sph + 2 * j + 1 : id;
    }
}

// ---- closest hit through the hierarchy (bvh.h), FP32 ------------------------------
// Up to 8 "big" spheres (the r = 1e4 walls) go through the packed pair test above with the
// cancellation-free K form; everything else sits in a two-child hierarchy walked per lane
// (while-while, near child first, explicit stack in local memory).  Small spheres use the
// direct form p = c - o, c' = |p|^2 - r^2: for r <= 1 the squares are small and the direct
// form is the more accurate one next to the surface the ray leaves.
__device__ __forceinline__ bool box_test_f32(float lx, float hx, float ly, float hy, float lz, float hz, f3 oi, f3 inv,
                                             float tmax_in, float& tnear) {
    // (plane - o) / d as one FMA per plane; NaN (0 * inf) drops out of fminf/fmaxf: that slab is ignored
    const float ax = fmaf(lx, inv.x, -oi.x), bx = fmaf(hx, inv.x, -oi.x);
    const float ay = fmaf(ly, inv.y, -oi.y), by = fmaf(hy, inv.y, -oi.y);
    const float az = fmaf(lz, inv.z, -oi.z), bz = fmaf(hz, inv.z, -oi.z);
    // nested so that the compiler can form 3-input FMNMX3
    const float tmin = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.f));
    const float tmax = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), tmax_in));
    tnear = tmin;
    return tmin <= tmax;
}

struct SharedBig {
    int id[8];
};

#define RTM_BVH_NONE_DEV 0x7FFFFFFF
__device__ __forceinline__ bool bvh_is_inner(int node) { return (unsigned)node < 0x7FFFFFFFu; }   // leaves are negative

// Ray state of one lane inside the hierarchy.  best = candidate (root - eps), as in the packed test.
struct BvhRay {
    f3 inv, oi;      // 1/d and o/d
    float best;
    int hit;         // original sphere index of the current winner, -1 = none
    int node;        // current node (child code), RTM_BVH_NONE_DEV when the traversal is over
    int sp;
};

// the packed test over the big list, then position the lane on the root
// ST = distance in ints between two consecutive entries of a lane's stack: 1 for a private (local-memory) array,
// the block size for the interleaved shared-memory layout stack[level][thread]
template <int NP, int ST = 1>
__device__ __forceinline__ void bvh_ray_begin(const SelfMask& sm, const SharedBig& big, const RtmBvhDev& bvh, int oslot, f3 o, f3 d,
                                              unsigned key_mask, BvhRay& r, int* stack, unsigned& n_tests) {
    r.best = FLT_MAX;
    r.hit = -1;
    if (NP > 0) {
        const unsigned key = trace_pairs_f32<2 * NP>(sm, oslot, o, d, key_mask);   // the big list is padded to even at run time
        if (key < RTM_KEY_INF) {
            r.best = __uint_as_float(key & ~7u);
            r.hit = big.id[key & 7u];
        }
        n_tests += 2 * NP;
    }
    r.inv = mk(fast_rcp(d.x), fast_rcp(d.y), fast_rcp(d.z));
    r.oi = mk(o.x * r.inv.x, o.y * r.inv.y, o.z * r.inv.z);
    r.node = bvh.root;
    r.sp = 0;
    stack[ST * r.sp++] = RTM_BVH_NONE_DEV;
}

// one inner node: test both children, descend into the nearer one, push the other
template <int ST = 1>
__device__ __forceinline__ void bvh_inner_step(const RtmBvhDev& bvh, BvhRay& r, int* stack) {
    const float4* q = bvh.nodes + 4 * r.node;
    const float4 q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3);
    float n0, n1;
    const float tm = r.best + 2.f * RTM_T_EPS;
    const bool h0 = box_test_f32(q0.x, q0.y, q0.z, q0.w, q2.x, q2.y, r.oi, r.inv, tm, n0);
    const bool h1 = box_test_f32(q1.x, q1.y, q1.z, q1.w, q2.z, q2.w, r.oi, r.inv, tm, n1);
    const int c0 = __float_as_int(q3.x), c1 = __float_as_int(q3.y);
    if (h0 && h1) {
        const bool swap = n1 < n0;
        stack[ST * r.sp++] = swap ? c0 : c1;
        r.node = swap ? c1 : c0;
    } else if (h0 || h1) {
        r.node = h0 ? c0 : c1;
    } else {
        r.node = stack[ST * --r.sp];
    }
}

// one inner node of the 4-wide tree (bvh.h: RtmBvh4Host; experimental, RTM_BVH_WIDE=1): one 128-byte fetch, four
// slab tests, the hit children ordered near to far.  The entry distance (>= 0, so its bits order like an unsigned
// integer) and the child's slot share one key — two low mantissa bits are given up to the slot — and a 5-exchange
// min/max network sorts the four keys; a missed or empty child has the all-ones key.
template <int ST = 1>
__device__ __forceinline__ void bvh4_inner_step(const RtmBvhDev& bvh, BvhRay& r, int* stack) {
    const float4* q = bvh.nodes4 + 8 * r.node;
    const float4 lx = __ldg(q), hx = __ldg(q + 1), ly = __ldg(q + 2), hy = __ldg(q + 3), lz = __ldg(q + 4), hz = __ldg(q + 5);
    const float4 cf = __ldg(q + 6);
    const int c0 = __float_as_int(cf.x), c1 = __float_as_int(cf.y), c2 = __float_as_int(cf.z), c3 = __float_as_int(cf.w);
    const float tm = r.best + 2.f * RTM_T_EPS;
    float n0, n1, n2, n3;
    const bool h0 = box_test_f32(lx.x, hx.x, ly.x, hy.x, lz.x, hz.x, r.oi, r.inv, tm, n0) && c0 != RTM_BVH_NONE_DEV;
    const bool h1 = box_test_f32(lx.y, hx.y, ly.y, hy.y, lz.y, hz.y, r.oi, r.inv, tm, n1) && c1 != RTM_BVH_NONE_DEV;
    const bool h2 = box_test_f32(lx.z, hx.z, ly.z, hy.z, lz.z, hz.z, r.oi, r.inv, tm, n2) && c2 != RTM_BVH_NONE_DEV;
    const bool h3 = box_test_f32(lx.w, hx.w, ly.w, hy.w, lz.w, hz.w, r.oi, r.inv, tm, n3) && c3 != RTM_BVH_NONE_DEV;
    unsigned k0 = h0 ? ((__float_as_uint(n0) & ~3u) | 0u) : 0xFFFFFFFFu;
    unsigned k1 = h1 ? ((__float_as_uint(n1) & ~3u) | 1u) : 0xFFFFFFFFu;
    unsigned k2 = h2 ? ((__float_as_uint(n2) & ~3u) | 2u) : 0xFFFFFFFFu;
    unsigned k3 = h3 ? ((__float_as_uint(n3) & ~3u) | 3u) : 0xFFFFFFFFu;
    unsigned a, b;
    a = min(k0, k1); b = max(k0, k1); k0 = a; k1 = b;
    a = min(k2, k3); b = max(k2, k3); k2 = a; k3 = b;
    a = min(k0, k2); b = max(k0, k2); k0 = a; k2 = b;
    a = min(k1, k3); b = max(k1, k3); k1 = a; k3 = b;
    a = min(k1, k2); b = max(k1, k2); k1 = a; k2 = b;
    auto child = [&](unsigned k) { const unsigned i = k & 3u; return (i == 0u) ? c0 : (i == 1u) ? c1 : (i == 2u) ? c2 : c3; };
    // far to near onto the stack, the nearest is walked next
    if (k3 != 0xFFFFFFFFu) stack[ST * r.sp++] = child(k3);
    if (k2 != 0xFFFFFFFFu) stack[ST * r.sp++] = child(k2);
    if (k1 != 0xFFFFFFFFu) stack[ST * r.sp++] = child(k1);
    if (k0 != 0xFFFFFFFFu) r.node = child(k0);
    else r.node = stack[ST * --r.sp];
}

// one leaf: 1..4 spheres in the direct form
__device__ __forceinline__ void bvh_leaf_test(const RtmBvhDev& bvh, int leaf, int oid, f3 o, f3 d, BvhRay& r, unsigned& n_tests) {
    const unsigned code = ~(unsigned)leaf;
    const unsigned first = code >> 3, cnt = (code & 7u) + 1u;
    for (unsigned k = 0; k < cnt; ++k) {
        const float4 g = __ldg(bvh.leaf_geo + first + k);
        const int sid = __ldg(bvh.leaf_id + first + k);
        const f3 p = mk(g.x - o.x, g.y - o.y, g.z - o.z);
        const float b = dot(p, d);
        // discriminant as r^2 - |p - b d|^2: b^2 - (|p|^2 - r^2) cancels at |p|^2 ~ 400 and
        // leaves an error of 3e-4 against r^2 = 2.5e-3 for the smallest spheres
        const f3 l = mk(fmaf(-b, d.x, p.x), fmaf(-b, d.y, p.y), fmaf(-b, d.z, p.z));
        float D = g.w - dot(l, l);
        float cp = dot(p, p) - g.w;
        const bool self = (sid == oid);                    // leaving this sphere's surface
        cp = self ? 0.f : cp;
        D = self ? b * b : D;
        const float s = fast_sqrt(D);                      // NaN: the ray misses
        const float tq = b + copysignf(s, b);
        const float ta = tq - RTM_T_EPS, tb = fmaf(cp, fast_rcp(tq), -RTM_T_EPS);
        const float c1 = (ta > 0.f) ? ta : FLT_MAX, c2 = (tb > 0.f) ? tb : FLT_MAX;
        const float c = fminf(c1, c2);
        const bool ok = (c < r.best) || (c == r.best && sid < r.hit);
        r.best = ok ? c : r.best;
        r.hit = ok ? sid : r.hit;
    }
    n_tests += cnt;
}
// ... then pop
template <int ST = 1>
__device__ __forceinline__ void bvh_leaf_step(const RtmBvhDev& bvh, int oid, f3 o, f3 d, BvhRay& r, int* stack, unsigned& n_tests) {
    bvh_leaf_test(bvh, r.node, oid, o, d, r, n_tests);
    r.node = stack[ST * --r.sp];
}

// PlaneObject extension: planes are tested like the big spheres, before the descent, so that the hierarchy is
// walked with their distance as the pruning bound
__device__ __forceinline__ void bvh_ray_planes(int n_sph, int n_pl, int oid, f3 o, f3 d, BvhRay& r) {
    float t = r.best + RTM_T_EPS;
    int id = r.hit;
    trace_planes_f32(n_sph, n_pl, oid, o, d, t, id);
    if (id != r.hit) {
        r.best = t - RTM_T_EPS;
        r.hit = id;
    }
}

// whole query (test hook; the render kernel interleaves the same steps across its lanes)
template <int NP>
__device__ __forceinline__ int trace_bvh_f32(const SelfMask& sm, const SharedBig& big, const RtmBvhDev& bvh, int oid, int oslot,
                                             f3 o, f3 d, float& t, unsigned key_mask, unsigned& n_tests, unsigned& n_nodes,
                                             int n_sph = 0, int n_pl = 0) {
    BvhRay r;
    int stack[RTM_BVH_STACK_DEV];
    bvh_ray_begin<NP>(sm, big, bvh, oslot, o, d, key_mask, r, stack, n_tests);
    if (n_pl > 0) bvh_ray_planes(n_sph, n_pl, oid, o, d, r);
    while (r.node != RTM_BVH_NONE_DEV) {
        while (bvh_is_inner(r.node)) {
            bvh_inner_step(bvh, r, stack);
            ++n_nodes;
        }
        if (r.node < 0) bvh_leaf_step(bvh, oid, o, d, r, stack, n_tests);
    }
    t = r.best + RTM_T_EPS;
    return r.hit;
}

// Per-sphere data addressed with a per-lane index (the hit id): shared memory, three
// 16-byte records per sphere so that a lane needs three LDS.128 per hit.
struct SharedScene {
    float4 geo[RTM_MAX_CONST_SPHERES];  // cx, cy, cz, 1/r
    float4 emi[RTM_MAX_CONST_SPHERES];  // emission rgb, roulette threshold (uint bits)
    float4 alb[RTM_MAX_CONST_SPHERES];  // color/kd rgb (colorKD, SettingData.h:11-13), "listed light" flag (NEE)
};

// Hierarchy path: the per-sphere records stay in global memory (L1/L2 resident); shared
// memory only holds the original indices of the "big" spheres the packed test covers.
template <bool BVH>
struct SceneRecs {
    const SharedScene* sh;
    RtmMatDev m;
    __device__ __forceinline__ float4 geo(int i) const { return BVH ? __ldg(m.geo + i) : sh->geo[i]; }
    __device__ __forceinline__ float4 emi(int i) const { return BVH ? __ldg(m.emi + i) : sh->emi[i]; }
    __device__ __forceinline__ float4 alb(int i) const { return BVH ? __ldg(m.alb + i) : sh->alb[i]; }
};

__device__ __forceinline__ void load_shared_scene(SharedScene& s, int n) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        s.geo[i] = make_float4(c_scene.cx[i], c_scene.cy[i], c_scene.cz[i], c_scene.inv_r[i]);
        s.emi[i] = make_float4(c_scene.ex[i], c_scene.ey[i], c_scene.ez[i], __uint_as_float(c_scene.kd_thr[i]));
        s.alb[i] = make_float4(c_scene.kx[i], c_scene.ky[i], c_scene.kz[i], c_scene.lit[i]);
    }
    __syncthreads();
}

// Cosine-weighted direction about the unit normal w (Renderer.cpp:88-107): a uniform
// point s on the unit sphere added to w and normalised is cosine-distributed about w,
// which needs no tangent frame.  The reference builds u,v from Cross(axis,w); the
// azimuth is uniform, so the distribution of the sampled direction is the same.
//   phi = 2*pi*xi1 (from 32 bits), z = 1 - 2*xi2 (22 bits, centred: never exactly +-1)
__device__ __forceinline__ f3 cosine_dir(f3 w, uint32_t u_phi, uint32_t u_z22) {
    const float z = fmaf((float)u_z22, -4.76837158203125e-7f, 1.0f - 2.384185791015625e-7f);  // 1 - (2k+1)/2^22
    const float rxy = fast_sqrt(fmaf(-z, z, 1.0f));
    float sn, cs;
    __sincosf((float)u_phi * 1.4629180792671596e-9f, &sn, &cs);  // 2*pi / 2^32
    const f3 v = mk(fmaf(rxy, cs, w.x), fmaf(rxy, sn, w.y), w.z + z);
    const float inv = fast_rsqrt(fmaxf(dot(v, v), 1e-12f));
    return mk(v.x * inv, v.y * inv, v.z * inv);
}

__device__ __forceinline__ float clamp01f(float v) { return __saturatef(v); }

// ---- the reference's self-intersection on small spheres ("ACNE") ---------------------
// Ray.h:67-72 normalises with a FLOAT sqrt, so every direction has |d|^2 = 1 + eps,
// eps ~ 1e-7.  SettingData.cpp:197-223 assumes a unit direction, so a hit reported at
// distance t sits at |x - c|^2 = r^2 + t^2 eps: OFF the surface by t^2 eps / 2r, which for
// r = 0.05, t = 15 is 2e-4.  When that point is inside the sphere (eps < 0), the next,
// outward ray finds the far root t2 = |c'| / (|b| + sqrt(b^2 + |c'|)) of its own sphere
// with c' = t^2 eps, b = -r cos(theta); it is accepted when t2 >= 1e-5 (:208-209), i.e. when
//     -c' >= 2 tau r cos(theta) + tau^2,   tau = 1e-5f.
// The path then continues from the same point with the normal flipped INWARD and stays
// inside the sphere.  The r = 1e4 walls never qualify (cos(theta) < 1e-3); 0.05..0.3 spheres
// do for about a fifth of the bounces and half of the primary hits, which darkens them
// visibly in the reference's images — so the FP32 path carries c' as explicit lane state:
// exact for primary hits (from the FP64 pass), and for bounce rays drawn from the
// distribution the reference's arithmetic produces: eta = |v|^2 - 1 before normalisation
// (std 5.3e-8, four summed bytes), x = fl32(1 + eta), m = sqrtf(x), eps = eta - 2(m - 1).
// State: cpsn = c' / (2 tau r) (0 when the normal is flipped: an inward ray takes the
// ordinary far root), test: cpsn + cos(theta) <= 0.
#define RTM_ACNE_INV2TAU 5.0000001e4f   // 1 / (2 * (double)1e-5f)
#define RTM_ACNE_2TAU 1.99999995e-5f
__device__ __forceinline__ float ref_normalize_eps(uint32_t bits) {
    const int S = (int)__dp4a(bits, 0x01010101u, 0u);
    const float eta = (float)(S - 510) * 3.586e-10f;
    const float m = __fsqrt_rn(1.0f + eta);
    return eta - 2.0f * (m - 1.0f);
}

#define RTM_MEGA_THREADS 128
#define RTM_MEGA_CHUNK 64u  // sub-pixels a warp reserves per global atomic
#ifndef RTM_MEGA_RETRY
#define RTM_MEGA_RETRY 12   // lanes killed by the roulette in one round that trigger another round before the bounce block
#endif

// ---- the radiance megakernel ---------------------------------------------------
// Persistent: grid = SMs x resident blocks.  A lane owns one sub-pixel (unit) at a time
// and walks its samples; one loop iteration = one roulette decision + one bounce ray.
// The lane state is "standing on a hit": origin o on sphere id with oriented normal w,
// throughput T, and radiance L that already includes the emission of this hit
// (Renderer.cpp:109 unrolled forward: L += T*e, T *= colorKD).  A path ends when the
// roulette kills it (Renderer.cpp:112), the bounce ray misses (:116), the depth cap is
// reached, or it lands on a surface with kd == 0 (the roulette there can never continue,
// so the sample is closed in the same iteration instead of spending another one).
// RNG: one rtm_bounce_rng call per hit (pcg3d hash, philox.cuh), counter = (sub-pixel, sample<<9 | depth), key
// from the seed: x -> roulette (32-bit compare against kd*2^32), y -> azimuth,
// the low bits of x and y -> height.
// PL: the scene also holds n_pl planes (extension); their records sit behind the spheres', ids n .. n+n_pl-1.
// NEE: next-event estimation (extension).  At every vertex that survives the roulette one light
// of the list (c_scene.lgeo / lemi) is chosen uniformly, a direction is drawn uniformly inside the cone the
// light sphere subtends (pdf 1 / (2 pi (1 - cos_max))), a shadow ray is traced by EVERY lane of the branch
// (lanes without a valid direction shoot along w and ignore the answer: no divergence around the trace) and
//     L += T * e_light * max(0, w.l) * 2 (1 - cos_max) * n_lights
// if it reaches the light.  A bounce ray that lands on a listed light from outside adds no emission (it was
// counted by the sampling at the previous vertex); from inside it does (no cone there).  Same expectation
// as the reference's estimator, ~8x less variance on the Cornell box.
// RT: roulette rounds (see the loop body) — chosen by the host for scenes whose visible surfaces kill most paths.
template <int N, bool BVH, bool ACNE, bool PL, bool NEE, bool RT>
__global__ void __launch_bounds__(RTM_MEGA_THREADS, BVH ? 4 : 5) radiance_mega_kernel(RtmPassParams p, float4* __restrict__ records,
                                                                                      RtmCounters* __restrict__ counters, int n_rt,
                                                                                      RtmBvhDev bvh, RtmMatDev mat,
                                                                                      const float* __restrict__ cps0, int n_pl) {
    __shared__ float4 sh_lgeo[RTM_MAX_LIGHTS], sh_lemi[RTM_MAX_LIGHTS];
    if (NEE && threadIdx.x < RTM_MAX_LIGHTS) {
        sh_lgeo[threadIdx.x] = c_scene.lgeo[threadIdx.x];
        sh_lemi[threadIdx.x] = c_scene.lemi[threadIdx.x];
    }
    unsigned long long shadow = 0;
    // brute force: N spheres (0 = runtime n_rt) with their records in shared memory;
    // hierarchy: N = slots of the packed big-sphere test (even, 0..8), records in global memory
    __shared__ SharedScene sh;
    __shared__ SelfMask sh_mask;
    __shared__ SharedBig sh_big;
    const int n = (N > 0) ? N : n_rt;
    init_selfmask(sh_mask);
    if (BVH) {
        if (threadIdx.x < 8) sh_big.id[threadIdx.x] = (threadIdx.x < (unsigned)bvh.n_big) ? bvh.big_id[threadIdx.x] : 0;
        __syncthreads();
    } else {
        load_shared_scene(sh, PL ? n + n_pl : n);
    }
    SceneRecs<BVH> recs;
    recs.sh = &sh;
    recs.m = mat;
    unsigned n_tests = 0, n_nodes = 0;

    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    const unsigned n_sub = (unsigned)p.n_sub;   // renderer_create guarantees n_sub < 2^32
    const int s_begin = p.s_begin, s_end = p.s_end;
    const float inv_norm = p.inv_norm;

    unsigned w_next = 0, w_end = 0;  // warp-uniform slice of the work cursor
    bool have = false, done = false;

    const unsigned sub_begin = (unsigned)p.sub_begin, sub_count = (unsigned)p.sub_count;
    const unsigned grid = (unsigned)(p.frame.ss * p.frame.ss);
    const bool full_grid = (sub_count == grid);
    // unit (sub-pixel) state: ql = record index in this handle's slice, q = global sub-pixel index
    unsigned q = 0, ql = 0;
    f3 P = mk(0, 0, 0), Pw = mk(0, 0, 1);  // primary hit point, oriented normal there
    int Pid = 0;
    f3 acc = mk(0, 0, 0);
    // path state
    f3 o = mk(0, 0, 0), w = mk(0, 0, 1), T = mk(1, 1, 1), L = mk(0, 0, 0);
    int id = 0, s = 0, depth = 0;
    float cpsn = 0.f, Pcpsn = 0.f;   // ACNE: c' / (2 tau r) at the current origin / at the primary hit

    unsigned long long rays_ref = 0, traced = 0;
    const unsigned key_mask = p.key_mask;  // 0xFFFFFFF8 from the host: kept in a register for key_pack

    for (;;) {
        // ---- refill: lanes without a unit take the next sub-pixels of the warp's chunk
        unsigned need = __ballot_sync(0xffffffffu, !have && !done);
        while (need) {
            if (w_next == w_end) {
                unsigned long long start = 0;
                if (lane == 0) start = atomicAdd(&counters->work_next, (unsigned long long)RTM_MEGA_CHUNK);
                start = __shfl_sync(0xffffffffu, start, 0);
                if (start >= (unsigned long long)n_sub) {
                    if (!have) done = true;
                    break;
                }
                w_next = (unsigned)start;
                w_end = (start + RTM_MEGA_CHUNK < (unsigned long long)n_sub) ? (unsigned)start + RTM_MEGA_CHUNK : n_sub;
            }
            const unsigned avail = w_end - w_next;
            const unsigned rank = __popc(need & lt_mask);
            const bool mine = ((need >> lane) & 1u) && rank < avail;
            if (mine) {
                ql = w_next + rank;
                // RNG counter = GLOBAL sub-pixel index, so the image does not depend on the partition
                q = full_grid ? ql : (ql / sub_count) * grid + sub_begin + (ql % sub_count);
                const float4 rec = records[ql];
                const unsigned wbits = __float_as_uint(rec.w);
                const unsigned ns = (unsigned)(s_end - s_begin);
                const bool miss = (wbits == RTM_REC_MISS);
                Pid = miss ? 0 : (int)(wbits & RTM_REC_IDMASK);
                const float4 e0 = recs.emi(Pid);
                if (miss || __float_as_uint(e0.w) == 0u) {
                    // every sample of this sub-pixel is the same: nothing (miss) or the
                    // emission of a surface that never continues (kd == 0, Renderer.cpp:112)
                    float4 out = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (!miss) {
                        out.x = clamp01f(e0.x * inv_norm) * (float)ns;
                        out.y = clamp01f(e0.y * inv_norm) * (float)ns;
                        out.z = clamp01f(e0.z * inv_norm) * (float)ns;
                    }
                    records[ql] = out;
                    rays_ref += ns;
                } else {
                    have = true;
                    P = mk(rec.x, rec.y, rec.z);
                    const float4 g = recs.geo(Pid);
                    const float ir = (wbits & RTM_REC_FLIP) ? -g.w : g.w;
                    Pw = mk((P.x - g.x) * ir, (P.y - g.y) * ir, (P.z - g.z) * ir);
                    if (PL && Pid >= n) Pw = (wbits & RTM_REC_FLIP) ? mk(-g.x, -g.y, -g.z) : mk(g.x, g.y, g.z);
                    acc = mk(0, 0, 0);
                    s = s_begin;
                    o = P; id = Pid; w = Pw;
                    T = mk(1, 1, 1); L = mk(e0.x, e0.y, e0.z); depth = 0;
                    if (ACNE) {
                        Pcpsn = (wbits & RTM_REC_FLIP) ? 0.f : cps0[ql] * g.w * RTM_ACNE_INV2TAU;
                        cpsn = Pcpsn;
                    }
                }
            }
            const unsigned took = (__popc(need) < avail) ? __popc(need) : avail;
            w_next += took;
            need = __ballot_sync(0xffffffffu, !have && !done && !mine);
        }
        if (!__any_sync(0xffffffffu, have)) {
            if (__all_sync(0xffffffffu, done)) break;
            continue;
        }

        // closes the sample (Renderer.cpp:240-242), then finishes the sub-pixel or restarts at the primary hit
        auto close_sample = [&]() {
            acc.x += clamp01f(L.x * inv_norm);
            acc.y += clamp01f(L.y * inv_norm);
            acc.z += clamp01f(L.z * inv_norm);
            ++s;
            if (s == s_end) {
                records[ql] = make_float4(acc.x, acc.y, acc.z, 0.f);
                rays_ref += (unsigned)(s_end - s_begin);
                have = false;
            } else {
                const float4 e0 = recs.emi(Pid);
                o = P; id = Pid; w = Pw;
                T = mk(1, 1, 1); L = mk(e0.x, e0.y, e0.z); depth = 0;
                if (ACNE) cpsn = Pcpsn;
            }
        };
        // Russian roulette, Renderer.cpp:78.  A killed path closes its sample and the lane stands on the primary hit
        // of the next one.  RT instantiation: when RTM_MEGA_RETRY or more lanes of the warp are in that position they
        // draw again before the (expensive) bounce block, so that scenes where most paths die at once
        // (simpleSetting1/2: kd = 0.25 over most of the image, 1.1-1.3 rays per sample) do not run it at a quarter of
        // the lanes: C3 108.7 -> 145 Grays/s.  On the Cornell box (19 % kills per round, ~5 lanes) the rounds never
        // trigger, but the loop's mere presence costs 8 % (111 vs 120.5 Grays/s), hence a template flag set by the
        // host from the scene (rtm_api.cu).  Draws are keyed by (sub-pixel, sample, depth): the image does not depend
        // on the number of rounds, nor on the flag.
        uint2 rnd = make_uint2(0u, 0u);
        bool go = false, killed = false;
        auto roulette = [&]() {
            rnd = rtm_bounce_rng(make_uint2(q, ((unsigned)s << 9) | (unsigned)depth), p.rk);
            const float4 em = recs.emi(id);
            go = (rnd.x < __float_as_uint(em.w)) && (depth + 1 < p.max_depth);
            killed = !go;
        };
        if (have) roulette();
        while (RT && __popc(__ballot_sync(0xffffffffu, killed)) >= RTM_MEGA_RETRY) {   // warp-uniform
            if (killed) {
                killed = false;
                close_sample();
                if (have) roulette();
            }
        }
        // the common case closes a killed path's sample together with the paths the bounce block ends (one pass)
        bool end = killed;
        if (have && go) {
            {
                const float4 al = recs.alb(id);
                T.x *= al.x; T.y *= al.y; T.z *= al.z;   // colorKD, SettingData.h:11-13
                if (NEE) {
                    const uint2 r2 = rtm_bounce_rng(make_uint2(q, ((unsigned)s << 9) | (unsigned)depth), p.rk + 1);
                    const unsigned nl = (unsigned)c_scene.n_lights;
                    const unsigned k = __umulhi(r2.x, nl);
                    const float4 lg = sh_lgeo[k], le = sh_lemi[k];
                    const int lid = __float_as_int(le.w);
                    const f3 v = mk(lg.x - o.x, lg.y - o.y, lg.z - o.z);
                    const float d2 = dot(v, v), rr2 = lg.w * lg.w;
                    const float q2 = rr2 * fast_rcp(d2);                                   // sin^2 of the cone's half angle
                    const float omc = q2 * fast_rcp(1.f + fast_sqrt(fmaxf(1.f - q2, 0.f)));   // 1 - cos_max, no cancellation
                    const float x = fmaf((float)(r2.x & 0x7FFFFFu), 1.1920929e-7f, 5.9604645e-8f) * omc;   // 1 - cos(theta)
                    const float ct = 1.f - x, st = fast_sqrt(x * (2.f - x));
                    float sn, cs;
                    __sincosf((float)r2.y * 1.4629180792671596e-9f, &sn, &cs);
                    const float inv_d = fast_rsqrt(d2);
                    const f3 a = mk(v.x * inv_d, v.y * inv_d, v.z * inv_d);
                    // branch-free orthonormal basis about a (Duff et al., "Building an Orthonormal Basis, Revisited", JCGT 2017)
                    const float sg = copysignf(1.f, a.z), aa = -fast_rcp(sg + a.z), bb = a.x * a.y * aa;
                    const f3 t1 = mk(fmaf(sg * a.x, a.x * aa, 1.f), sg * bb, -sg * a.x), t2 = mk(bb, fmaf(a.y, a.y * aa, sg), -a.y);
                    const float c1 = st * cs, c2 = st * sn;
                    const f3 l = mk(fmaf(a.x, ct, fmaf(t1.x, c1, t2.x * c2)), fmaf(a.y, ct, fmaf(t1.y, c1, t2.y * c2)),
                                    fmaf(a.z, ct, fmaf(t1.z, c1, t2.z * c2)));
                    const float cw = dot(w, l);
                    const bool ok = (lid != id) && (d2 > rr2) && (cw > 0.f);
                    const f3 ls = ok ? l : w;
                    float ts;
                    int ids = trace_f32<N>(sh_mask, n, (PL && id >= n) ? -1 : id, o, ls, ts, key_mask);
                    if (PL) trace_planes_f32(n, n_pl, id, o, ls, ts, ids);
                    shadow += ok ? 1u : 0u;
                    if (ok && ids == lid) {
                        const float g = cw * 2.f * omc * (float)nl;
                        L.x = fmaf(T.x * le.x, g, L.x); L.y = fmaf(T.y * le.y, g, L.y); L.z = fmaf(T.z * le.z, g, L.z);
                    }
                }
                const f3 d = cosine_dir(w, rnd.y, ((rnd.x & 0xFFFu) << 10) | (rnd.y & 0x3FFu));
                float t;
                int id2;
                bool selfhit = false;
                float cosw = 0.f;
                if (ACNE) {
                    cosw = dot(w, d);
                    selfhit = (cpsn + cosw <= 0.f) && !(PL && id >= n);
                }
                if (ACNE && selfhit) {
                    // SettingData.cpp:208-212: t1 <= 0.001 -> t2 of the sphere the ray stands (just) inside
                    const float rr = fast_rcp(recs.geo(id).w);
                    const float q = -cpsn * rr * RTM_ACNE_2TAU, bq = rr * cosw;
                    t = q * fast_rcp(bq + fast_sqrt(fmaf(bq, bq, q)));
                    id2 = id;
                } else if (BVH) id2 = trace_bvh_f32<N / 2>(sh_mask, sh_big, bvh, id, __float_as_int(al.w), o, d, t, key_mask, n_tests, n_nodes);
                else {
                    // a ray leaving a plane has no origin sphere (the self mask only has rows for spheres)
                    id2 = trace_f32<N>(sh_mask, n, (PL && id >= n) ? -1 : id, o, d, t, key_mask);
                    if (PL) trace_planes_f32(n, n_pl, id, o, d, t, id2);
                }
                ++traced;
                if (id2 >= 0) {
                    o = mk(fmaf(d.x, t, o.x), fmaf(d.y, t, o.y), fmaf(d.z, t, o.z));  // Renderer.cpp:79
                    const float4 g = recs.geo(id2);
                    f3 nn = mk((o.x - g.x) * g.w, (o.y - g.y) * g.w, (o.z - g.z) * g.w);
                    if (PL && id2 >= n) nn = mk(g.x, g.y, g.z);
                    if (BVH && g.w > 1e-2f) {
                        // small spheres: t carries an absolute error of ~1e-5 that is not small against
                        // r = 0.05; put the hit back on the surface and make the normal unit
                        const float inv = fast_rsqrt(dot(nn, nn));
                        nn = mk(nn.x * inv, nn.y * inv, nn.z * inv);
                        const float rr = fast_rcp(g.w);
                        o = mk(fmaf(rr, nn.x, g.x), fmaf(rr, nn.y, g.y), fmaf(rr, nn.z, g.z));
                    }
                    const bool flip = !(dot(nn, d) < 0.f);                            // :82-83
                    w = flip ? mk(-nn.x, -nn.y, -nn.z) : nn;
                    id = id2;
                    ++depth;
                    if (ACNE) cpsn = flip ? 0.f : t * t * ref_normalize_eps(rnd.x ^ (rnd.y * 0x9E3779B9u)) * g.w * RTM_ACNE_INV2TAU;
                    const float4 e2 = recs.emi(id2);
                    // NEE: a listed light reached from outside was already counted by the sampling above
                    if (!NEE || flip || __float_as_int(recs.alb(id2).w) == 0) {
                        L.x = fmaf(T.x, e2.x, L.x); L.y = fmaf(T.y, e2.y, L.y); L.z = fmaf(T.z, e2.z, L.z);
                    }
                    end = (__float_as_uint(e2.w) == 0u);   // kd == 0: the next roulette cannot continue
                } else {
                    end = true;  // miss: the recursion returns 0 (Renderer.cpp:116)
                }
            }
        }
        if (end) close_sample();
    }

    rays_ref += traced;
    traced += shadow;   // shadow rays are closest-hit queries too, but not rays of the reference's estimator
    unsigned long long tests = BVH ? (unsigned long long)n_tests : traced * (unsigned long long)n;
    unsigned long long nodes = n_nodes;
    for (int off = 16; off > 0; off >>= 1) {
        rays_ref += __shfl_down_sync(0xffffffffu, rays_ref, off);
        traced += __shfl_down_sync(0xffffffffu, traced, off);
        tests += __shfl_down_sync(0xffffffffu, tests, off);
        nodes += __shfl_down_sync(0xffffffffu, nodes, off);
    }
    if (lane == 0) {
        atomicAdd(&counters->rays_ref, rays_ref);
        atomicAdd(&counters->rays_traced, traced);
        atomicAdd(&counters->tests, tests);
        if (BVH) atomicAdd(&counters->nodes, nodes);
    }
}

// ---- the radiance kernel of the hierarchy path -------------------------------------------
// Same estimator, same lane-owns-a-sub-pixel work distribution as radiance_mega_kernel, but a
// closest-hit query now costs ~25 node visits + ~11 sphere tests with a large spread, so "one
// loop iteration = one bounce" would leave a warp waiting for its longest traversal with the
// shading code interleaved at 1/32 occupancy (measured: 6.3 of 32 lanes active on average).
// Here a lane is in one of two phases and the warp runs them as separate, coherent blocks:
//   S (shade)  standing on a hit: apply a finished ray, roulette, sample the bounce, start the
//              next ray (or close the sample / the sub-pixel);
//   T (trace)  walk the hierarchy, while-while: all lanes step through inner nodes until each
//              holds a leaf, then all test their leaf's spheres.
// The T block keeps stepping until at least RTM_BVH_S_THRESH lanes have finished their rays;
// only then is the S block run (for those lanes together), so S code runs at >= 25 % and T code
// at >= 75 % lane occupancy (Aila & Laine's persistent threads with a refill threshold).
#ifndef RTM_BVH_S_THRESH
#define RTM_BVH_S_THRESH 16
#endif
#ifndef RTM_BVH_BURST
#define RTM_BVH_BURST 8
#endif
#ifndef RTM_BVH_MINBLOCKS
#define RTM_BVH_MINBLOCKS 7
#endif
#ifndef RTM_BVH_HOLD
#define RTM_BVH_HOLD 0
#endif
#ifndef RTM_BVH_S_ROUND
#define RTM_BVH_S_ROUND 1   // lanes that must need another roulette decision for the S block to run another round (swept: 1 4.50, 2 4.46, 4 4.45, 8 4.34 Grays/s on C5)
#endif

// PL: n_pl planes (extension) behind the n_sph spheres in the material records, tested at the start of every ray
// SST: the per-lane traversal stack lives in shared memory, stack[level][thread], RTM_BVH_SSTACK levels (the host
// picks it when the tree is shallow enough: depth + 2 entries are the most a walk can hold); otherwise a private
// 64-entry array in local memory.
#ifndef RTM_BVH_SSTACK
#define RTM_BVH_SSTACK 24
#endif
#ifndef RTM_BVH_SSTACK_W4
#define RTM_BVH_SSTACK_W4 28   // the wide walk pushes up to 3 entries a level: 3 * depth4 + 1 (synth10k: 25)
#endif
// W4: walk the 4-wide form of the tree (bvh4_inner_step) — experimental, opt-in, same leaves and same candidates.
template <int NP, bool ACNE, bool PL, bool SST, bool W4 = false>
__global__ void __launch_bounds__(RTM_MEGA_THREADS, W4 ? 6 : RTM_BVH_MINBLOCKS) radiance_bvh_kernel(RtmPassParams p, float4* __restrict__ records,
                                                                           RtmCounters* __restrict__ counters, RtmBvhDev bvh,
                                                                           RtmMatDev mat, const float* __restrict__ cps0, int n_sph,
                                                                           int n_pl) {
    __shared__ SelfMask sh_mask;
    __shared__ SharedBig sh_big;
    init_selfmask(sh_mask);
    if (threadIdx.x < 8) sh_big.id[threadIdx.x] = (threadIdx.x < (unsigned)bvh.n_big) ? bvh.big_id[threadIdx.x] : 0;
    __syncthreads();

    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    const unsigned n_sub = (unsigned)p.n_sub;
    const int s_begin = p.s_begin, s_end = p.s_end;
    const float inv_norm = p.inv_norm;
    const unsigned sub_begin = (unsigned)p.sub_begin, sub_count = (unsigned)p.sub_count;
    const unsigned grid = (unsigned)(p.frame.ss * p.frame.ss);
    const bool full_grid = (sub_count == grid);
    const unsigned key_mask = p.key_mask;

    unsigned w_next = 0, w_end = 0;
    bool have = false, done = false, tracing = false, pending = false;
    unsigned q = 0, ql = 0;
    f3 acc = mk(0, 0, 0);
    f3 o = mk(0, 0, 0), w = mk(0, 0, 1), T = mk(1, 1, 1), L = mk(0, 0, 0), d = mk(0, 0, 1);
    int id = 0, s = 0, depth = 0;
    float cpsn = 0.f;
    uint32_t eps_bits = 0;
    // the primary hit of the lane's sub-pixel is NOT lane state here: it is re-read from records[ql] (which holds it
    // until the sub-pixel's sum is written) at every sample restart — 8 registers less, which is what lets
    // RTM_BVH_MINBLOCKS blocks stay resident
    auto stand_on_primary = [&](const float4 rec) {
        const unsigned wbits = __float_as_uint(rec.w);
        const int pid = (int)(wbits & RTM_REC_IDMASK);
        const float4 g = __ldg(mat.geo + pid), e0 = __ldg(mat.emi + pid);
        const float ir = (wbits & RTM_REC_FLIP) ? -g.w : g.w;
        o = mk(rec.x, rec.y, rec.z);
        w = mk((o.x - g.x) * ir, (o.y - g.y) * ir, (o.z - g.z) * ir);
        if (PL && pid >= n_sph) w = (wbits & RTM_REC_FLIP) ? mk(-g.x, -g.y, -g.z) : mk(g.x, g.y, g.z);
        id = pid;
        T = mk(1, 1, 1); L = mk(e0.x, e0.y, e0.z); depth = 0;
        if (ACNE) cpsn = (wbits & RTM_REC_FLIP) ? 0.f : cps0[ql] * g.w * RTM_ACNE_INV2TAU;
    };
    BvhRay ray;
    ray.node = RTM_BVH_NONE_DEV;
    constexpr int ST = SST ? RTM_MEGA_THREADS : 1;
    __shared__ int sh_stack[SST ? (W4 ? RTM_BVH_SSTACK_W4 : RTM_BVH_SSTACK) * RTM_MEGA_THREADS : 1];
    int lstack[SST ? 1 : RTM_BVH_STACK_DEV];
    int* const stack = SST ? (sh_stack + threadIdx.x) : lstack;
    unsigned long long rays_ref = 0, traced = 0;
    unsigned n_tests = 0, n_nodes = 0;

    for (;;) {
        // ---- refill (warp-uniform): lanes without a sub-pixel take the next ones of the warp's chunk
        unsigned need = __ballot_sync(0xffffffffu, !have && !done);
        while (need) {
            if (w_next == w_end) {
                unsigned long long start = 0;
                if (lane == 0) start = atomicAdd(&counters->work_next, (unsigned long long)RTM_MEGA_CHUNK);
                start = __shfl_sync(0xffffffffu, start, 0);
                if (start >= (unsigned long long)n_sub) {
                    if (!have) done = true;
                    break;
                }
                w_next = (unsigned)start;
                w_end = (start + RTM_MEGA_CHUNK < (unsigned long long)n_sub) ? (unsigned)start + RTM_MEGA_CHUNK : n_sub;
            }
            const unsigned avail = w_end - w_next;
            const unsigned rank = __popc(need & lt_mask);
            const bool mine = ((need >> lane) & 1u) && rank < avail;
            if (mine) {
                ql = w_next + rank;
                q = full_grid ? ql : (ql / sub_count) * grid + sub_begin + (ql % sub_count);
                const float4 rec = records[ql];
                const unsigned wbits = __float_as_uint(rec.w);
                const unsigned ns = (unsigned)(s_end - s_begin);
                const bool miss = (wbits == RTM_REC_MISS);
                const int Pid = miss ? 0 : (int)(wbits & RTM_REC_IDMASK);
                const float4 e0 = __ldg(mat.emi + Pid);
                if (miss || __float_as_uint(e0.w) == 0u) {
                    float4 out = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (!miss) {
                        out.x = clamp01f(e0.x * inv_norm) * (float)ns;
                        out.y = clamp01f(e0.y * inv_norm) * (float)ns;
                        out.z = clamp01f(e0.z * inv_norm) * (float)ns;
                    }
                    records[ql] = out;
                    rays_ref += ns;
                } else {
                    have = true;
                    acc = mk(0, 0, 0);
                    s = s_begin;
                    stand_on_primary(rec);
                }
            }
            const unsigned took = (__popc(need) < avail) ? __popc(need) : avail;
            w_next += took;
            need = __ballot_sync(0xffffffffu, !have && !done && !mine);
        }
        const unsigned tr = __ballot_sync(0xffffffffu, tracing);
        const unsigned sready = __ballot_sync(0xffffffffu, have && !tracing);
        if ((tr | sready) == 0u) {
            if (__all_sync(0xffffffffu, done)) break;
            continue;
        }

        // ---- S block.  Entered by the whole warp (warp-uniform control flow, lanes outside it are predicated off):
        // the roulette runs in ROUNDS shared by all S lanes, and the packed big-sphere test + ray start run once,
        // after the rounds, for every lane that has a direction.  (With per-lane loops the compiler let each lane
        // leave the roulette loop on its own: the ~140-instruction ray start then ran once per group of lanes that
        // left together, 3.3 times per entry at 6 of 32 lanes — ncu_bvh_v3.)  A lane whose path is killed in the last
        // round closes its sample and stays S-ready; another round is run only while RTM_BVH_S_ROUND or more lanes
        // need one.
        const bool in_s = (tr == 0u || __popc(sready) >= RTM_BVH_S_THRESH) && have && !tracing;
        if (__any_sync(0xffffffffu, in_s)) {
            // a finished ray is applied to the path; returns true when the path ends there
            auto apply_hit = [&](int id2, float t) -> bool {
                if (id2 < 0) return true;                                            // miss: Renderer.cpp:116
                o = mk(fmaf(d.x, t, o.x), fmaf(d.y, t, o.y), fmaf(d.z, t, o.z));     // :79
                const float4 g = __ldg(mat.geo + id2);
                f3 nn = mk((o.x - g.x) * g.w, (o.y - g.y) * g.w, (o.z - g.z) * g.w);
                if (PL && id2 >= n_sph) nn = mk(g.x, g.y, g.z);   // a plane's record holds its normal, 1/r = 0
                if (g.w > 1e-2f) {
                    // small spheres: t carries an absolute error of ~1e-5 that is not small against
                    // r = 0.05; put the hit back on the surface and make the normal unit
                    const float inv = fast_rsqrt(dot(nn, nn));
                    nn = mk(nn.x * inv, nn.y * inv, nn.z * inv);
                    const float rr = fast_rcp(g.w);
                    o = mk(fmaf(rr, nn.x, g.x), fmaf(rr, nn.y, g.y), fmaf(rr, nn.z, g.z));
                }
                const bool flip = !(dot(nn, d) < 0.f);                               // :82-83
                w = flip ? mk(-nn.x, -nn.y, -nn.z) : nn;
                id = id2;
                ++depth;
                if (ACNE) cpsn = flip ? 0.f : t * t * ref_normalize_eps(eps_bits) * g.w * RTM_ACNE_INV2TAU;
                const float4 e2 = __ldg(mat.emi + id2);
                L.x = fmaf(T.x, e2.x, L.x); L.y = fmaf(T.y, e2.y, L.y); L.z = fmaf(T.z, e2.z, L.z);
                return __float_as_uint(e2.w) == 0u;                                  // kd == 0: the next roulette cannot continue
            };
            bool end = false;
            if (in_s && pending) {
                pending = false;
                end = apply_hit(ray.hit, ray.best + RTM_T_EPS);
            }
            bool deciding = in_s, start = false;
            int oslot = -1;
            for (;;) {
                if (deciding) {
                    if (!end) {
                        const uint2 rnd = rtm_bounce_rng(make_uint2(q, ((unsigned)s << 9) | (unsigned)depth), p.rk);
                        const float4 em = __ldg(mat.emi + id);
                        end = !((rnd.x < __float_as_uint(em.w)) && (depth + 1 < p.max_depth));   // Renderer.cpp:78
                        if (!end) {
                            const float4 al = __ldg(mat.alb + id);
                            T.x *= al.x; T.y *= al.y; T.z *= al.z;                               // colorKD
                            d = cosine_dir(w, rnd.y, ((rnd.x & 0xFFFu) << 10) | (rnd.y & 0x3FFu));
                            eps_bits = rnd.x ^ (rnd.y * 0x9E3779B9u);
                            ++traced;
                            bool self = false;
                            if (ACNE) {
                                const float cosw = dot(w, d);
                                if (cpsn + cosw <= 0.f && !(PL && id >= n_sph)) {
                                    // SettingData.cpp:208-212: the far root of the sphere the ray stands (just) inside
                                    const float rr = fast_rcp(__ldg(mat.geo + id).w);
                                    const float qq = -cpsn * rr * RTM_ACNE_2TAU, bq = rr * cosw;
                                    end = apply_hit(id, qq * fast_rcp(bq + fast_sqrt(fmaf(bq, bq, qq))));
                                    self = true;   // stands on the next vertex: decide again
                                }
                            }
                            if (!self) {
                                oslot = __float_as_int(al.w);
                                start = true;
                                deciding = false;
                            }
                        }
                    }
                    if (deciding && end) {
                        // the sample is complete: Renderer.cpp:240-242
                        acc.x += clamp01f(L.x * inv_norm);
                        acc.y += clamp01f(L.y * inv_norm);
                        acc.z += clamp01f(L.z * inv_norm);
                        ++s;
                        if (s == s_end) {
                            records[ql] = make_float4(acc.x, acc.y, acc.z, 0.f);
                            rays_ref += (unsigned)(s_end - s_begin);
                            have = false;
                            deciding = false;
                        } else {
                            stand_on_primary(records[ql]);
                            end = false;
                        }
                    }
                }
                if (__popc(__ballot_sync(0xffffffffu, deciding)) < RTM_BVH_S_ROUND) break;
            }
            if (start) {
                // the packed big-sphere test (~140 instructions) and the planes: once, for all lanes that start a ray
                bvh_ray_begin<NP, ST>(sh_mask, sh_big, bvh, oslot, o, d, key_mask, ray, stack, n_tests);
                if (W4) ray.node = bvh.root4;
                if (PL) bvh_ray_planes(n_sph, n_pl, id, o, d, ray);
                if (ray.node != RTM_BVH_NONE_DEV) tracing = true;
                else pending = true;   // empty hierarchy: the hit (big spheres / planes) is applied at the next entry
            }
        }

        // ---- T block: while-while until enough lanes are ready for the S block
        unsigned still = __ballot_sync(0xffffffffu, tracing);
        while (still) {
            // at most RTM_BVH_BURST inner nodes per round: a lane on its long first descent must not
            // keep the lanes that hop between neighbouring leaves (2-3 nodes apart) waiting
#if RTM_BVH_HOLD
            // a lane that reaches a leaf inside the burst parks it and goes on with the next node of its stack, so it
            // keeps the inner-node block company instead of idling until the leaf block (Aila & Laine's speculative
            // traversal); a second leaf makes it wait
            int held = 0;   // leaf codes are negative
#pragma unroll 1
            for (int k = 0; k < RTM_BVH_BURST; ++k) {
                if (tracing && ray.node < 0 && held == 0) {
                    held = ray.node;
                    ray.node = stack[ST * --ray.sp];
                }
                if (!(tracing && bvh_is_inner(ray.node))) break;
                bvh_inner_step<ST>(bvh, ray, stack);
                ++n_nodes;
            }
            if (tracing && held != 0) bvh_leaf_test(bvh, held, id, o, d, ray, n_tests);
            else if (tracing && ray.node < 0) bvh_leaf_step<ST>(bvh, id, o, d, ray, stack, n_tests);
#else
#pragma unroll 1
            for (int k = 0; k < (W4 ? (RTM_BVH_BURST + 1) / 2 : RTM_BVH_BURST) && tracing && bvh_is_inner(ray.node); ++k) {
                if (W4) bvh4_inner_step<ST>(bvh, ray, stack);
                else bvh_inner_step<ST>(bvh, ray, stack);
                ++n_nodes;
            }
            if (tracing && ray.node < 0) bvh_leaf_step<ST>(bvh, id, o, d, ray, stack, n_tests);
#endif
            if (tracing && ray.node == RTM_BVH_NONE_DEV) {
                tracing = false;
                pending = true;
            }
            still = __ballot_sync(0xffffffffu, tracing);
            const unsigned ready = __ballot_sync(0xffffffffu, !tracing && !done);
            if (__popc(ready) >= RTM_BVH_S_THRESH) break;
        }
    }

    rays_ref += traced;
    unsigned long long tests = n_tests, nodes = n_nodes;
    for (int off = 16; off > 0; off >>= 1) {
        rays_ref += __shfl_down_sync(0xffffffffu, rays_ref, off);
        traced += __shfl_down_sync(0xffffffffu, traced, off);
        tests += __shfl_down_sync(0xffffffffu, tests, off);
        nodes += __shfl_down_sync(0xffffffffu, nodes, off);
    }
    if (lane == 0) {
        atomicAdd(&counters->rays_ref, rays_ref);
        atomicAdd(&counters->rays_traced, traced);
        atomicAdd(&counters->tests, tests);
        atomicAdd(&counters->nodes, nodes);
    }
}

// ---- accumulate: sub-pixel sums -> pixel, coalesced float4 in and out ---------
// records float4[n_pix][g] (g = ss*ss) -> accum float4[n_pix].  For power-of-two
// g <= 32 one lane per record and a butterfly over each group of g lanes.
__global__ void __launch_bounds__(256) accumulate_pow2_kernel(const float4* __restrict__ records, float4* __restrict__ accum,
                                                              unsigned long long n_sub, int g) {
    const unsigned long long q = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    float4 v = (q < n_sub) ? records[q] : make_float4(0, 0, 0, 0);
    for (int off = 1; off < g; off <<= 1) {
        v.x += __shfl_xor_sync(0xffffffffu, v.x, off);
        v.y += __shfl_xor_sync(0xffffffffu, v.y, off);
        v.z += __shfl_xor_sync(0xffffffffu, v.z, off);
    }
    if (q < n_sub && (q & (unsigned long long)(g - 1)) == 0) accum[q / g] = make_float4(v.x, v.y, v.z, 0.f);
}

__global__ void __launch_bounds__(256) accumulate_generic_kernel(const float4* __restrict__ records, float4* __restrict__ accum,
                                                                 unsigned long long n_pix, int g) {
    const unsigned long long pix = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= n_pix) return;
    float x = 0, y = 0, z = 0;
    for (int j = 0; j < g; ++j) {
        const float4 v = records[pix * g + j];
        x += v.x; y += v.y; z += v.z;
    }
    accum[pix] = make_float4(x, y, z, 0.f);
}

// ---- Renderer.cpp:251-254: (unsigned char)255 * std::min(image[i], 1.0) --------
__global__ void __launch_bounds__(256) to_u8_kernel(const float4* __restrict__ accum, unsigned char* __restrict__ out,
                                                    unsigned long long n_pix) {
    const unsigned long long pix = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= n_pix) return;
    const float4 v = accum[pix];
    const double c[3] = {(double)v.x, (double)v.y, (double)v.z};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double m = (1.0 < c[k]) ? 1.0 : c[k];
        out[3 * pix + k] = (unsigned char)(int)(255.0 * m);
    }
}

// ---- sum of G partial framebuffers that the ranks' accumulate kernels stored into THIS GPU's memory over
// NVLink (peer stores): out[pix] = sum_g partials[g][pix], fixed order g = 0..G-1 (deterministic) --------------
__global__ void __launch_bounds__(256) sum_partials_kernel(const float4* __restrict__ partials, float4* __restrict__ out,
                                                           unsigned long long n_pix, int G) {
    const unsigned long long pix = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= n_pix) return;
    float4 a = partials[pix];
    for (int g = 1; g < G; ++g) {
        const float4 v = partials[(unsigned long long)g * n_pix + pix];
        a.x += v.x; a.y += v.y; a.z += v.z;
    }
    a.w = 0.f;
    out[pix] = a;
}

// ---- float4[H][W] {r,g,b,0} -> float[H][W][3], the layout of renderer_render's rgb_out ----
__global__ void __launch_bounds__(256) pack_rgb_kernel(const float4* __restrict__ accum, float* __restrict__ out,
                                                       unsigned long long n_pix) {
    const unsigned long long pix = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= n_pix) return;
    const float4 v = accum[pix];
    out[3 * pix] = v.x; out[3 * pix + 1] = v.y; out[3 * pix + 2] = v.z;
}

// ---- test hooks ---------------------------------------------------------------
__global__ void closest_hit_f32_kernel(int n_sph, int n_pl, unsigned key_mask, double rx, double ry, double rz, uint32_t n, const double* __restrict__ org,
                                       const double* __restrict__ dir, const int32_t* __restrict__ origin_id,
                                       int32_t* __restrict__ hit, double* __restrict__ tout) {
    __shared__ SelfMask sh_mask;
    init_selfmask(sh_mask);
    __syncthreads();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const f3 o = mk((float)(org[3 * i] - rx), (float)(org[3 * i + 1] - ry), (float)(org[3 * i + 2] - rz));
    const f3 d = mk((float)dir[3 * i], (float)dir[3 * i + 1], (float)dir[3 * i + 2]);
    float t;
    int oid_all = origin_id ? origin_id[i] : -1;
    if (oid_all < -1 || oid_all >= n_sph + n_pl) oid_all = -1;
    const int oid = (oid_all >= n_sph) ? -1 : oid_all;
    int id;
    // same code paths as the megakernel: packed pairs up to 8 spheres, the loop beyond
    if (n_sph >= 1 && n_sph <= 2) id = trace_f32<2>(sh_mask, n_sph, oid, o, d, t, key_mask);
    else if (n_sph <= 4 && n_sph >= 3) id = trace_f32<4>(sh_mask, n_sph, oid, o, d, t, key_mask);
    else if (n_sph <= 6 && n_sph >= 5) id = trace_f32<6>(sh_mask, n_sph, oid, o, d, t, key_mask);
    else if (n_sph <= 8 && n_sph >= 7) id = trace_f32<8>(sh_mask, n_sph, oid, o, d, t, key_mask);
    else id = trace_f32<0>(sh_mask, n_sph, oid, o, d, t, key_mask);
    trace_planes_f32(n_sph, n_pl, oid_all, o, d, t, id);
    hit[i] = id;
    tout[i] = (double)t;
}

// the hierarchy variant of the hook: same trace function as radiance_mega_kernel<*, true>
template <int NP>
__global__ void closest_hit_bvh_f32_kernel(RtmBvhDev bvh, RtmMatDev mat, unsigned key_mask, double rx, double ry, double rz, uint32_t n,
                                           const double* __restrict__ org, const double* __restrict__ dir,
                                           const int32_t* __restrict__ origin_id, int n_sph, int n_pl, int32_t* __restrict__ hit,
                                           double* __restrict__ tout) {
    __shared__ SelfMask sh_mask;
    __shared__ SharedBig sh_big;
    init_selfmask(sh_mask);
    if (threadIdx.x < 8) sh_big.id[threadIdx.x] = (threadIdx.x < (unsigned)bvh.n_big) ? bvh.big_id[threadIdx.x] : 0;
    __syncthreads();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const f3 o = mk((float)(org[3 * i] - rx), (float)(org[3 * i + 1] - ry), (float)(org[3 * i + 2] - rz));
    const f3 d = mk((float)dir[3 * i], (float)dir[3 * i + 1], (float)dir[3 * i + 2]);
    int oid = origin_id ? origin_id[i] : -1;
    if (oid < -1 || oid >= n_sph + n_pl) oid = -1;
    const int oslot = (oid >= 0) ? __float_as_int(__ldg(mat.alb + oid).w) : -1;
    float t;
    unsigned a = 0, b = 0;
    hit[i] = trace_bvh_f32<NP>(sh_mask, sh_big, bvh, oid, oslot, o, d, t, key_mask, a, b, n_sph, n_pl);
    tout[i] = (double)t;
}

__global__ void acne_eps_kernel(uint32_t n, const uint32_t* __restrict__ bits, float* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = ref_normalize_eps(bits[i]);
}

__global__ void philox_kernel(uint32_t n, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 r = rtm_philox4x32_10(make_uint4(i, c1, c2, c3), k0, k1);
    out[4 * i] = r.x; out[4 * i + 1] = r.y; out[4 * i + 2] = r.z; out[4 * i + 3] = r.w;
}

// FP32 peak: 8 independent FMA chains per thread, register resident.
__global__ void __launch_bounds__(256) fma_peak_kernel(float* __restrict__ sink, int iters) {
    float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const float m = 0.999f, c = 1e-3f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            a0 = fmaf(a0, m, c); a1 = fmaf(a1, m, c); a2 = fmaf(a2, m, c); a3 = fmaf(a3, m, c);
            a4 = fmaf(a4, m, c); a5 = fmaf(a5, m, c); a6 = fmaf(a6, m, c); a7 = fmaf(a7, m, c);
        }
    }
    const float r = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
    if (r == 12345.678f) sink[0] = r;
}

template <int N, bool BVH, bool ACNE, bool PL, bool NEE = false, bool RT = false>
void launch_mega2(const RtmPassParams& p, float4* records, RtmCounters* counters, int n_rt, int n_pl, int sm_count, cudaStream_t st,
                  const RtmBvhDev& bvh, const RtmMatDev& mat, const float* cps0) {
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, radiance_mega_kernel<N, BVH, ACNE, PL, NEE, RT>, RTM_MEGA_THREADS, 0);
    if (per_sm < 1) per_sm = 1;
    radiance_mega_kernel<N, BVH, ACNE, PL, NEE, RT><<<sm_count * per_sm, RTM_MEGA_THREADS, 0, st>>>(p, records, counters, n_rt, bvh, mat, cps0, n_pl);
}
// cps0 != null selects the ACNE instantiation (see ref_normalize_eps), n_pl > 0 the one with the plane loop
template <int N, bool BVH>
void launch_mega(const RtmPassParams& p, float4* records, RtmCounters* counters, int n_rt, int n_pl, bool nee, bool retry, int sm_count,
                 cudaStream_t st, const float* cps0, const RtmBvhDev& bvh = RtmBvhDev{}, const RtmMatDev& mat = RtmMatDev{}) {
    if (retry && !nee && n_pl == 0) {   // roulette rounds: instantiated for the reference's own scene model
        if (cps0) launch_mega2<N, BVH, true, false, false, true>(p, records, counters, n_rt, 0, sm_count, st, bvh, mat, cps0);
        else launch_mega2<N, BVH, false, false, false, true>(p, records, counters, n_rt, 0, sm_count, st, bvh, mat, nullptr);
    } else if (nee) {
        if (n_pl > 0) {
            if (cps0) launch_mega2<N, BVH, true, true, true>(p, records, counters, n_rt, n_pl, sm_count, st, bvh, mat, cps0);
            else launch_mega2<N, BVH, false, true, true>(p, records, counters, n_rt, n_pl, sm_count, st, bvh, mat, nullptr);
        } else {
            if (cps0) launch_mega2<N, BVH, true, false, true>(p, records, counters, n_rt, 0, sm_count, st, bvh, mat, cps0);
            else launch_mega2<N, BVH, false, false, true>(p, records, counters, n_rt, 0, sm_count, st, bvh, mat, nullptr);
        }
    } else if (n_pl > 0) {
        if (cps0) launch_mega2<N, BVH, true, true>(p, records, counters, n_rt, n_pl, sm_count, st, bvh, mat, cps0);
        else launch_mega2<N, BVH, false, true>(p, records, counters, n_rt, n_pl, sm_count, st, bvh, mat, nullptr);
    } else {
        if (cps0) launch_mega2<N, BVH, true, false>(p, records, counters, n_rt, 0, sm_count, st, bvh, mat, cps0);
        else launch_mega2<N, BVH, false, false>(p, records, counters, n_rt, 0, sm_count, st, bvh, mat, nullptr);
    }
}

}  // namespace

void rtm_launch_radiance_mega(int32_t n, int32_t n_pl, bool nee, bool retry, const RtmPassParams& p, float4* records,
                              RtmCounters* counters, int32_t sm_count, const float* cps0, cudaStream_t st) {
    switch (n) {
        case 1: launch_mega<1, false>(p, records, counters, n, n_pl, nee, retry, sm_count, st, cps0); break;
        case 2: launch_mega<2, false>(p, records, counters, n, n_pl, nee, retry, sm_count, st, cps0); break;
        case 3: launch_mega<3, false>(p, records, counters, n, n_pl, nee, retry, sm_count, st, cps0); break;
        case 4: launch_mega<4, false>(p, records, counters, n, n_pl, nee, retry, sm_count, st, cps0); break;
        case 5: launch_mega<5, false>(p, records, counters, n, n_pl, nee, retry, sm_count, st, cps0); break;
        case 6: launch_mega<6, false>(p, records, counters, n, n_pl, nee, retry, sm_count, st, cps0); break;
        case 7: launch_mega<7, false>(p, records, counters, n, n_pl, nee, retry, sm_count, st, cps0); break;
        case 8: launch_mega<8, false>(p, records, counters, n, n_pl, nee, retry, sm_count, st, cps0); break;
        default: launch_mega<0, false>(p, records, counters, n, n_pl, nee, retry, sm_count, st, cps0); break;
    }
}

// ---- wavefront variant -------------------------------------------------------------------
namespace {
#include "kernels_wavefront.cuh"

template <int N, bool BVH>
void launch_wf_shade(const RtmBvhDev& bvh, const RtmMatDev& mat, const RtmPassParams& p, float4* records, RtmCounters* counters,
                     int n_rt, const float* cps0, const RtmWaveDev& wv, int wave, cudaStream_t st) {
    const unsigned g = (unsigned)((wv.n_slots + 127) / 128);
    if (cps0) wf_shade_kernel<N, BVH, true><<<g, 128, 0, st>>>(p, records, counters, n_rt, bvh, mat, cps0, wv, wave);
    else wf_shade_kernel<N, BVH, false><<<g, 128, 0, st>>>(p, records, counters, n_rt, bvh, mat, nullptr, wv, wave);
}
}  // namespace

void rtm_launch_wf_shade(int32_t n, bool use_bvh, const RtmBvhDev& bvh, const RtmMatDev& mat, const RtmPassParams& p,
                         float4* records, RtmCounters* counters, const float* cps0, const RtmWaveDev& wv, int32_t wave,
                         cudaStream_t st) {
    if (use_bvh) {
        switch ((bvh.n_big + 1) / 2) {
            case 0: launch_wf_shade<0, true>(bvh, mat, p, records, counters, 0, cps0, wv, wave, st); break;
            case 1: launch_wf_shade<2, true>(bvh, mat, p, records, counters, 0, cps0, wv, wave, st); break;
            case 2: launch_wf_shade<4, true>(bvh, mat, p, records, counters, 0, cps0, wv, wave, st); break;
            case 3: launch_wf_shade<6, true>(bvh, mat, p, records, counters, 0, cps0, wv, wave, st); break;
            default: launch_wf_shade<8, true>(bvh, mat, p, records, counters, 0, cps0, wv, wave, st); break;
        }
        return;
    }
    // brute force: the shade kernel only needs the per-sphere records (shared memory, any n <= 256)
    launch_wf_shade<0, false>(bvh, mat, p, records, counters, n, cps0, wv, wave, st);
}

void rtm_launch_wf_extend(int32_t n, bool use_bvh, const RtmBvhDev& bvh, const RtmWaveDev& wv, int32_t wave, RtmCounters* counters,
                          int32_t sm_count, cudaStream_t st) {
    if (use_bvh) {
        int per_sm = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, wf_extend_bvh_kernel, 128, 0);
        if (per_sm < 1) per_sm = 1;
        wf_extend_bvh_kernel<<<sm_count * per_sm, 128, 0, st>>>(bvh, wv, wave, counters);
        return;
    }
    const unsigned g = (unsigned)((wv.n_slots + 127) / 128);
    switch (n) {
#define RTM_CASE(K) case K: wf_extend_brute_kernel<K><<<g, 128, 0, st>>>(wv, wave, n, 0xFFFFFFF8u, counters); break
        RTM_CASE(1); RTM_CASE(2); RTM_CASE(3); RTM_CASE(4); RTM_CASE(5); RTM_CASE(6); RTM_CASE(7); RTM_CASE(8);
#undef RTM_CASE
        default: wf_extend_brute_kernel<0><<<g, 128, 0, st>>>(wv, wave, n, 0xFFFFFFF8u, counters); break;
    }
}

// Hierarchy path: the packed test covers the (padded-to-even) list of big spheres.
namespace {
template <int NP, bool ACNE, bool PL, bool SST>
void launch_bvh3(const RtmBvhDev& bvh, const RtmMatDev& mat, const RtmPassParams& p, float4* records, RtmCounters* counters,
                 int sm_count, const float* cps0, int n_sph, int n_pl, cudaStream_t st) {
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, radiance_bvh_kernel<NP, ACNE, PL, SST>, RTM_MEGA_THREADS, 0);
    if (per_sm < 1) per_sm = 1;
    radiance_bvh_kernel<NP, ACNE, PL, SST><<<sm_count * per_sm, RTM_MEGA_THREADS, 0, st>>>(p, records, counters, bvh, mat, cps0, n_sph, n_pl);
}
template <int NP, bool ACNE, bool PL>
void launch_bvh2(const RtmBvhDev& bvh, const RtmMatDev& mat, const RtmPassParams& p, float4* records, RtmCounters* counters,
                 int sm_count, const float* cps0, int n_sph, int n_pl, cudaStream_t st) {
    // a walk holds at most depth + 2 stack entries (one per level + the sentinel)
    static const bool want = [] { const char* e = getenv("RTM_BVH_SSTACK_OFF"); return !(e && *e == '1'); }();
    if (!PL && bvh.nodes4 && bvh.root4 >= 0 && 3 * bvh.depth4 + 1 <= RTM_BVH_STACK_DEV) {
        // experimental wide walk (RTM_BVH_WIDE=1 at renderer_create: local-memory stack; =2: shared-memory stack), spheres only
        int per_sm = 0;
        if (bvh.wide_mode == 2 && 3 * bvh.depth4 + 1 <= RTM_BVH_SSTACK_W4) {
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, radiance_bvh_kernel<NP, ACNE, false, true, true>, RTM_MEGA_THREADS, 0);
            if (per_sm < 1) per_sm = 1;
            radiance_bvh_kernel<NP, ACNE, false, true, true><<<sm_count * per_sm, RTM_MEGA_THREADS, 0, st>>>(p, records, counters, bvh, mat, cps0, n_sph, n_pl);
            return;
        }
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, radiance_bvh_kernel<NP, ACNE, false, false, true>, RTM_MEGA_THREADS, 0);
        if (per_sm < 1) per_sm = 1;
        radiance_bvh_kernel<NP, ACNE, false, false, true><<<sm_count * per_sm, RTM_MEGA_THREADS, 0, st>>>(p, records, counters, bvh, mat, cps0, n_sph, n_pl);
        return;
    }
    if (want && bvh.depth + 2 <= RTM_BVH_SSTACK) launch_bvh3<NP, ACNE, PL, true>(bvh, mat, p, records, counters, sm_count, cps0, n_sph, n_pl, st);
    else launch_bvh3<NP, ACNE, PL, false>(bvh, mat, p, records, counters, sm_count, cps0, n_sph, n_pl, st);
}
template <int NP>
void launch_bvh(const RtmBvhDev& bvh, const RtmMatDev& mat, const RtmPassParams& p, float4* records, RtmCounters* counters,
                int sm_count, const float* cps0, int n_sph, int n_pl, cudaStream_t st) {
    if (n_pl > 0) {
        if (cps0) launch_bvh2<NP, true, true>(bvh, mat, p, records, counters, sm_count, cps0, n_sph, n_pl, st);
        else launch_bvh2<NP, false, true>(bvh, mat, p, records, counters, sm_count, nullptr, n_sph, n_pl, st);
    } else {
        if (cps0) launch_bvh2<NP, true, false>(bvh, mat, p, records, counters, sm_count, cps0, n_sph, 0, st);
        else launch_bvh2<NP, false, false>(bvh, mat, p, records, counters, sm_count, nullptr, n_sph, 0, st);
    }
}
}  // namespace

void rtm_launch_radiance_bvh(const RtmBvhDev& bvh, const RtmMatDev& mat, const RtmPassParams& p, float4* records,
                             RtmCounters* counters, int32_t sm_count, const float* cps0, int32_t n_sph, int32_t n_pl, cudaStream_t st) {
    switch ((bvh.n_big + 1) / 2) {
        case 0: launch_bvh<0>(bvh, mat, p, records, counters, sm_count, cps0, n_sph, n_pl, st); break;
        case 1: launch_bvh<1>(bvh, mat, p, records, counters, sm_count, cps0, n_sph, n_pl, st); break;
        case 2: launch_bvh<2>(bvh, mat, p, records, counters, sm_count, cps0, n_sph, n_pl, st); break;
        case 3: launch_bvh<3>(bvh, mat, p, records, counters, sm_count, cps0, n_sph, n_pl, st); break;
        default: launch_bvh<4>(bvh, mat, p, records, counters, sm_count, cps0, n_sph, n_pl, st); break;
    }
}

void rtm_launch_accumulate(const RtmPassParams& p, const float4* records, float4* accum, cudaStream_t st) {
    const int g = p.sub_count;
    const unsigned long long n_pix = (unsigned long long)p.frame.W * p.frame.H;
    if (g <= 32 && (g & (g - 1)) == 0) {
        accumulate_pow2_kernel<<<(unsigned)((p.n_sub + 255) / 256), 256, 0, st>>>(records, accum, p.n_sub, g);
    } else {
        accumulate_generic_kernel<<<(unsigned)((n_pix + 255) / 256), 256, 0, st>>>(records, accum, n_pix, g);
    }
}

void rtm_launch_sum_partials(const float4* partials, float4* out, uint64_t n_pix, int32_t G, cudaStream_t st) {
    sum_partials_kernel<<<(unsigned)((n_pix + 255) / 256), 256, 0, st>>>(partials, out, n_pix, G);
}

void rtm_launch_pack_rgb(const float4* accum, float* out, uint64_t n_pix, cudaStream_t st) {
    pack_rgb_kernel<<<(unsigned)((n_pix + 255) / 256), 256, 0, st>>>(accum, out, n_pix);
}

void rtm_launch_to_u8(const float4* accum, unsigned char* out, uint64_t n_pix, cudaStream_t st) {
    to_u8_kernel<<<(unsigned)((n_pix + 255) / 256), 256, 0, st>>>(accum, out, n_pix);
}

void rtm_launch_closest_hit_f32(int32_t n_spheres, int32_t n_planes, const double* refd, uint32_t n, const double* org, const double* dir,
                                const int32_t* origin_id, int32_t* hit, double* t, cudaStream_t st) {
    closest_hit_f32_kernel<<<(n + 127) / 128, 128, 0, st>>>(n_spheres, n_planes, 0xFFFFFFF8u, refd[0], refd[1], refd[2], n, org, dir, origin_id, hit, t);
}

void rtm_launch_closest_hit_bvh_f32(const RtmBvhDev& bvh, const RtmMatDev& mat, int32_t n_spheres, int32_t n_planes, const double* refd, uint32_t n,
                                    const double* org, const double* dir, const int32_t* origin_id, int32_t* hit, double* t,
                                    cudaStream_t st) {
    const unsigned g = (n + 127) / 128;
    switch ((bvh.n_big + 1) / 2) {
#define RTM_CASE(NP) closest_hit_bvh_f32_kernel<NP><<<g, 128, 0, st>>>(bvh, mat, 0xFFFFFFF8u, refd[0], refd[1], refd[2], n, org, dir, origin_id, n_spheres, n_planes, hit, t)
        case 0: RTM_CASE(0); break;
        case 1: RTM_CASE(1); break;
        case 2: RTM_CASE(2); break;
        case 3: RTM_CASE(3); break;
        default: RTM_CASE(4); break;
#undef RTM_CASE
    }
}

void rtm_launch_acne_eps(uint32_t n, const uint32_t* bits, float* out, cudaStream_t st) {
    acne_eps_kernel<<<(n + 127) / 128, 128, 0, st>>>(n, bits, out);
}

void rtm_launch_philox(uint32_t n, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t* out, cudaStream_t st) {
    philox_kernel<<<(n + 127) / 128, 128, 0, st>>>(n, c1, c2, c3, k0, k1, out);
}

void rtm_launch_fma_peak(float* sink, int32_t blocks, int32_t iters, cudaStream_t st) {
    fma_peak_kernel<<<blocks, 256, 0, st>>>(sink, iters);
}
