// bvh.h — bounding-volume hierarchy over the scene's spheres (host build, device
// traversal).  Used when the brute-force closest-hit loop of the reference
// (src/Renderer.cpp:62-72, O(N) per ray) stops being affordable: SURVEY.md §8(d) C5,
// 10 000 primitives.  The hierarchy only decides WHICH spheres a ray is tested
// against; the per-candidate arithmetic and the accept/tie rule stay those of
// src/SettingData.cpp:197-223 / src/Renderer.cpp:67, so the FP64 hit map remains
// bit-exact against brute force (see closest_hit_bvh in kernels_exact.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

#define RTM_BVH_NONE 0x7FFFFFFF   // "no node": empty hierarchy / bottom of the traversal stack
#define RTM_BVH_MAX_BIG 8         // spheres too large for the hierarchy, tested for every ray
#define RTM_BVH_LEAF_MAX 4
#define RTM_BVH_CTRAV 1.0         // SAH: cost of an inner-node visit in sphere tests
#define RTM_BVH_STACK 64
#define RTM_BVH_BIG_RADIUS 256.0  // absolute "too large for the FP32 leaf test" radius
#ifndef RTM_BVH_LAYOUT_TOP
#define RTM_BVH_LAYOUT_TOP 1024   // layout 1: inner nodes numbered breadth-first (32 KB of compact nodes) before the treelets start
#endif

// child code: >= 0  inner node index
//             <  0  leaf: ~code = (first_slot << 3) | (count - 1)
static inline int32_t rtm_bvh_leaf_code(uint32_t first, uint32_t count) { return ~(int32_t)((first << 3) | (count - 1u)); }

// Inner node = 4 x float4 (64 B), boxes of BOTH children, coordinates relative to the
// scene reference point (the camera origin), rounded OUTWARD to float after padding:
//   q[0] = c0.lo.x c0.hi.x c0.lo.y c0.hi.y
//   q[1] = c1.lo.x c1.hi.x c1.lo.y c1.hi.y
//   q[2] = c0.lo.z c0.hi.z c1.lo.z c1.hi.z
//   q[3] = child0 child1 (int bits)  rmin0 rmin1 (smallest radius below each child)
struct RtmBvhHost {
    std::vector<float4> nodes;       // 4 per inner node
    std::vector<float4> leaf_geo;    // per leaf slot: centre (relative, float), r^2 (float)
    std::vector<int32_t> leaf_id;    // per leaf slot: original sphere index
    int32_t root = RTM_BVH_NONE;     // child code of the root
    int32_t n_big = 0;
    int32_t big_id[RTM_BVH_MAX_BIG] = {0};
    int32_t depth = 0;
    int32_t n_big_overflow = 0;      // spheres that qualify as big but did not fit the big list
};

// position[n][3] / radius[n] as in rtm_scene_desc; ref = scene reference point.
// Throws nothing; returns false when the tree would exceed RTM_BVH_STACK levels.
// c_trav / leaf_max: SAH constants (0 = the defaults above).  A sphere is kept out of the tree ("big") when its radius
// exceeds 8 x the median radius or RTM_BVH_BIG_RADIUS: the FP32 leaf test evaluates |c - o|^2 - r^2 directly, which is
// only accurate for small r; out.n_big_overflow counts the big spheres beyond RTM_BVH_MAX_BIG that had to stay in leaves
// (fine for the exact FP64 walk; the FP32 path refuses such a scene).
// layout: numbering of the inner nodes — 0 = breadth-first (default), 1 = breadth-first top + treelets of 4 below (A/B knob).
bool rtm_bvh_build(int32_t n, const double* position, const float* radius, const double ref[3], RtmBvhHost& out,
                   double c_trav = 0.0, int leaf_max = 0, int layout = 0);

// ---- compact form of the same tree for the FP32 walk: 32 bytes per inner node ---------------------------------
// The FP32 hierarchy kernel is bound by L1 wavefronts (one per lane per load instruction, every lane on a different
// node: ncu, profiles/r02).  A node that fits ONE 32-byte request halves them again after the 2 x 32-byte fetch of the
// float node.  Child boxes are stored as (centre, half-extent) in IEEE half precision, in a frame centred on the root
// box and scaled by a power of two so that the scene spans [-2^14, 2^14):
//   u32[0..2]  child 0: cx|cy, cz|hx, hy|hz   (low half first)      u32[3..5]  child 1      u32[6], u32[7]  child codes
// The (centre, half-extent) form needs no per-axis min/max in the slab test (near = c.inv - oi - |h.inv|, far = ... + |h.inv|).
// Conservative by construction: the half box {c16 -+ h16} contains the padded float box of RtmBvhHost (h16 is rounded
// up after the centre has been rounded), so the compact walk visits a superset of the candidates and finds the same
// winner.  `inflation` = median over the leaf children of (h16 - h) / h on the widest-error axis: the host uses the form
// only when that stays small (a scene whose features are tiny against its extent keeps the float nodes).
struct RtmBvh16Host {
    std::vector<uint32_t> nodes;     // 8 per inner node, same node numbering as RtmBvhHost
    float ref[3] = {0.f, 0.f, 0.f};  // centre of the frame, relative to the scene reference point
    float inv_scale = 1.f;           // 1 / scale (a power of two)
    double inflation = 0.0;
    bool ok = false;                 // false: empty tree, a single leaf, or a coordinate out of half range
};
void rtm_bvh_compact16(const RtmBvhHost& in, RtmBvh16Host& out);

// ---- 4-wide compact form (A/B alternative of the FP32 walk: rtm_options.debug_bvh_walk bit 3) ---------------------
// A node that holds up to four children — greedy collapse of the two-child tree: the inner child with the largest box is
// replaced by its own two children until there are four — halves the DEPENDENT fetches of a walk (synth10k: 29.4 -> 15.6
// per ray for the same 59 box tests: scripts/study_bvh4.py).  Built to test whether the walk is bound by the latency of
// those fetches; it is not: 6.42 Grays/s on C5 against 6.44 for the two-child walk (equal work per burst), because a step
// costs 125 instructions against 66 and the instruction stream, not the fetch latency, is what the kernel waits for.
// 64 bytes = two halves laid out exactly like a two-child compact node:
//   u32[0..2] child 0   u32[3..5] child 1   u32[6], u32[7] codes 0, 1   |   u32[8..10] child 2   u32[11..13] child 3   u32[14], u32[15] codes 2, 3
// The child boxes ARE RtmBvh16Host's (same words, same frame), the leaves are shared, child codes >= 0 index THIS array;
// an empty slot holds the RTM_BVH_NONE code behind a box that no ray hits (half-extents = -infinity).
struct RtmBvh16x4Host {
    std::vector<uint32_t> nodes;     // 16 per node, breadth-first, node 0 = root
    int32_t depth = 0;               // deepest leaf, root = 0: a walk holds at most 3 * depth + 1 stack entries
    bool ok = false;
};
void rtm_bvh_compact16x4(const RtmBvhHost& in, const RtmBvh16Host& c16, RtmBvh16x4Host& out);

// ---- 4-wide form of the same tree with float boxes (host-side design study) ----------------
// Greedy collapse of the two-child tree: a node adopts its grandchildren (largest box first) until it holds 4
// children.  The child boxes ARE the two-child tree's boxes, the leaves are shared (same leaf codes, same
// leaf_geo / leaf_id), so the set of candidates a ray reaches can only shrink in visits, never change in winners.
// Node = 8 x float4 (128 B, one cache line), children side by side:
//   q[0] lo.x[4]  q[1] hi.x[4]  q[2] lo.y[4]  q[3] hi.y[4]  q[4] lo.z[4]  q[5] hi.z[4]
//   q[6] child codes[4] (int bits; RTM_BVH_NONE = empty slot)   q[7] rmin[4]
struct RtmBvh4Host {
    std::vector<float4> nodes;       // 8 per node, breadth-first
    int32_t root = RTM_BVH_NONE;     // child code of the root (a leaf code when the tree is a single leaf)
    int32_t depth = 0;               // deepest leaf, root = 0: a walk holds at most 3 * depth + 1 stack entries
};
void rtm_bvh_collapse4(const RtmBvhHost& in, RtmBvh4Host& out);
