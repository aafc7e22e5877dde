// rtm_internal.h — types shared by the C-ABI implementation (rtm_api.cu) and the
// kernel translation units.  Not part of the public interface.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define RTM_MAX_CONST_SPHERES 256  // FP32 brute-force scene lives in __constant__ memory up to this size
#define RTM_WARP 32
// Device-side invariants (array bounds, traversal-stack depth, object ids).  compute-sanitizer is closed on the GPU pool this
// was developed on, so the memory-safety evidence is a build with these asserts compiled in (make checks -> librtm_b200_checks.so)
// run under the GPU test-suite: a violated invariant fails the launch with cudaErrorAssert.  The product build compiles them out.
#ifdef RTM_DEVICE_CHECKS
#include <cassert>
#define RTM_ASSERT(x) assert(x)
#else
#define RTM_ASSERT(x) ((void)0)
#endif
#define RTM_BVH_STACK_DEV 64       // traversal stack entries per lane (bvh.h: the build caps the depth below this)

// Camera frame of reference src/Renderer.cpp:202-208, precomputed on the host
// in double without FMA: ax = camX*fovx, ay = camY*fovy, az = camZ.
struct RtmFrame {
    double o[3];
    double ax[3], ay[3], az[3];
    float rate;          // (float)(1.0/(1+ss))  src/Renderer.cpp:227
    int32_t W, H, ss, S; // S = global "samples"
};

// Device view of the sphere hierarchy (bvh.h): nodes / leaf slots in global memory.
struct RtmBvhDev {
    const float4* nodes;      // 4 float4 per inner node
    const float4* leaf_geo;   // per leaf slot: centre relative to the reference point, r^2 (float)
    const int32_t* leaf_id;   // per leaf slot: original sphere index
    int32_t root;             // child code of the root, RTM_BVH_NONE when empty
    int32_t n_big;            // spheres kept out of the hierarchy, tested for every ray
    int32_t big_id[8];        //   ascending original index
    int32_t depth;            // levels of the tree (a walk holds at most depth + 2 stack entries)
    int32_t n_nodes, n_slots; // array sizes (device-side checks: -DRTM_DEVICE_CHECKS)
    // compact form for the FP32 walk (bvh.h: RtmBvh16Host): 8 x u32 per inner node, NULL = not used
    const uint32_t* nodes16;
    float ref16[3];           // centre of the half-precision frame, relative to the scene reference point
    float inv_scale16;        // 1 / (power-of-two scale of that frame)
    // 4-wide compact form (bvh.h: RtmBvh16x4Host): 16 x u32 per node in the same frame, NULL = not used.  When set, the
    // FP32 walk starts at its node 0 and child codes >= 0 index THIS array; the leaves are shared with the other forms
    const uint32_t* nodes16x4;
    int32_t n_nodes4, depth4;
};

// Per-sphere records of the FP32 path in GLOBAL memory (hierarchy path, any N): the
// same three float4 the brute-force kernel keeps in shared memory, plus the slot of
// the sphere in the "big" list (or -1) in alb.w.
struct RtmMatDev {
    const float4* geo;  // cx cy cz 1/r
    const float4* emi;  // emission rgb, roulette threshold (uint bits)
    const float4* alb;  // color/kd rgb, big-list slot (int bits)
};

// Scene for the exact FP64 kernels: SoA in global memory (any N).
struct RtmSceneF64 {
    int32_t n;
    int32_t use_bvh;         // closest hit through `bvh` (same candidates arithmetic, same winner)
    double ref[3];           // reference point the hierarchy's boxes are relative to
    RtmBvhDev bvh;
    const double* cx;
    const double* cy;
    const double* cz;
    const double* r2;        // (double)(float)(m_size*m_size)   src/SettingData.cpp:200
    const double* emission;  // [n][3]
    const double* color;     // [n][3]
    const double* kd;        // (double)(float)max(r,g,b)        src/SettingData.h:14-16
    // PlaneObject extension: ids n .. n+n_planes-1; kd/emission/color hold n+n_planes entries.
    // planes[16*j + ..] = normal[0..2], position[3..5], right[6..8], upp[9..11], rr[12], uu[13]
    int32_t n_planes;
    const double* planes;
};

// Two spheres side by side for the packed (f32x2) sphere test of the FP32 megakernel:
// lane .x = sphere 2j, lane .y = sphere 2j+1 (a never-hit dummy pads an odd count).
// p2 = +2c and nK = r^2 - |c|^2 so that the kernel forms -c' = o.(2c) + (nK - |o|^2).
struct __align__(16) RtmPairF32 {
    float2 cx, cy, cz, p2x, p2y, p2z, nK, pad;
};
#define RTM_MAX_PAIRS 4
#define RTM_MAX_PLANES 16
#define RTM_MAX_LIGHTS 16

// Two planes of the FP32 path side by side (lane .x = plane 2j, .y = plane 2j+1), relative to the scene
// reference point: unit normal, d0 = n.position, -position, and the two half-extent vectors divided by their
// squared lengths (inside the square iff both |dots| <= 1).  An odd count is padded with a never-hit dummy
// (n = 0, d0 = -1).
struct __align__(16) RtmPlanePairF32 {
    float2 nx, ny, nz, d0, npx, npy, npz, rx, ry, rz, ux, uy, uz, pad0, pad1, pad2;
};

// Per-sphere data of the FP32 path, relative to the scene reference point
// (the camera origin), SoA.  Filled on the host in double, rounded once.
struct RtmSceneF32Host {
    int32_t n;
    float cx[RTM_MAX_CONST_SPHERES], cy[RTM_MAX_CONST_SPHERES], cz[RTM_MAX_CONST_SPHERES];
    float m2cx[RTM_MAX_CONST_SPHERES], m2cy[RTM_MAX_CONST_SPHERES], m2cz[RTM_MAX_CONST_SPHERES];  // -2c
    float K[RTM_MAX_CONST_SPHERES];      // |c|^2 - r^2
    float inv_r[RTM_MAX_CONST_SPHERES];
    float ex[RTM_MAX_CONST_SPHERES], ey[RTM_MAX_CONST_SPHERES], ez[RTM_MAX_CONST_SPHERES];        // emission
    float kx[RTM_MAX_CONST_SPHERES], ky[RTM_MAX_CONST_SPHERES], kz[RTM_MAX_CONST_SPHERES];        // color/kd
    float kd[RTM_MAX_CONST_SPHERES];
    RtmPairF32 pairs[RTM_MAX_PAIRS];     // the first 8 spheres again, paired
    uint32_t kd_thr[RTM_MAX_CONST_SPHERES];  // min(kd * 2^32, 2^32-1): roulette continues iff a 32-bit draw < kd_thr
    // PlaneObject extension: materials of plane j live at index n + j of the arrays above (cx,cy,cz = its normal,
    // inv_r = 0); geometry here
    int32_t n_planes;
    RtmPlanePairF32 plane_pairs[RTM_MAX_PLANES / 2];
    // next-event estimation (extension): the light list and, per sphere, "is in the list"
    int32_t n_lights;
    float4 lgeo[RTM_MAX_LIGHTS];   // centre (relative to the reference point), radius
    float4 lemi[RTM_MAX_LIGHTS];   // emission rgb, sphere index (int bits)
    float lit[RTM_MAX_CONST_SPHERES];   // 1 for listed lights (int bits), else 0
};

// What the FP32 bounce kernels read through constant-bank operands.  Handed to every launch BY VALUE as a
// __grid_constant__ kernel parameter, so it is private to the launch (and to the handle that made it): there is
// no process-global scene, two handles may render concurrently on one device.  `full` is the handle's device
// copy of the whole per-sphere block; the kernels stage what they need of it into shared memory.
struct RtmHotScene {
    RtmPairF32 pairs[RTM_MAX_PAIRS];                  // the first 8 spheres (brute force) or the big list (hierarchy), paired
    RtmPlanePairF32 plane_pairs[RTM_MAX_PLANES / 2];
    int32_t n, n_planes, n_lights, pad;
    const RtmSceneF32Host* full;
};

// Sub-pixel record written by the exact primary pass and consumed by the FP32
// bounce kernels: primary hit point (relative to the reference point, rounded
// to float from the double hit point) and the hit id.  w = id | flip<<30, or
// 0xFFFFFFFF for a miss.  The bounce kernel overwrites the record with the
// sub-pixel's radiance sum {r,g,b,0}.
#define RTM_REC_MISS 0xFFFFFFFFu
#define RTM_REC_FLIP 0x40000000u
#define RTM_REC_IDMASK 0x3FFFFFFFu

struct RtmCounters {
    unsigned long long rays_ref;     // reference-equivalent PathTracing invocations
    unsigned long long rays_traced;  // closest-hit queries executed
    unsigned long long tests;        // ray-sphere tests executed
    unsigned long long work_next;    // dynamic work cursor of the persistent megakernel
    unsigned long long nodes;        // hierarchy nodes visited (hierarchy path only)
};

struct RtmPassParams {
    RtmFrame frame;
    int32_t s_begin, s_end;   // sample slice
    int32_t max_depth;
    uint32_t seed_lo, seed_hi;
    uint32_t key_mask;        // 0xFFFFFFF8, passed in so that it lives in a register (see key_pack)
    uint32_t rk[11];          // rk[0] = fold(seed): key of the bounce hash; rk[r] = rk[0] + r*0x9E3779B9 (Philox2x32 round keys, -DRTM_RNG_PHILOX)
    float inv_norm;           // 1/(ss*ss*S) for the FP32 path
    float ref[3];             // reference point (as float, informational)
    double refd[3];           // reference point
    uint64_t n_sub;           // W*H*sub_count: sub-pixels THIS handle renders (records are indexed by it)
    int32_t sub_begin, sub_count;  // slice of the ss*ss sub-pixel grid
};

// Wavefront variant (RTM_VARIANT_WAVEFRONT): path state of `n_slots` concurrent sub-pixels in
// global memory, one wave = shade kernel (apply hit, roulette, sample, emit ray) + extend kernel
// (closest hit for every queued ray).  All arrays are per slot.
struct RtmWaveDev {
    float4* a0;        // o.xyz, id
    float4* a1;        // w.xyz (oriented normal), cpsn
    float4* a2;        // T.xyz, s << 9 | depth
    float4* a3;        // L.xyz, flags (0 empty, 1 sub-pixel in flight, 2 + ray pending)
    float4* a4;        // acc.xyz, ql
    float4* a5;        // P.xyz, Pid             (constant while the slot holds a sub-pixel)
    float4* a6;        // Pw.xyz, Pcpsn
    float4* ray;       // d.xyz, -
    float2* hit;       // t, hit id (shade seeds it with the big-sphere result on the hierarchy path)
    int32_t* queue[2]; // slots with a pending ray, per wave parity
    unsigned* ctrl;    // [0],[1] queue lengths per parity; [2] extend's fetch cursor; [3] 1 + last wave that queued a ray
    int32_t n_slots;
};

// ---- launchers implemented by the kernel TUs (all asynchronous on `st`) ----
void rtm_launch_primary_f64(const RtmSceneF64& sc, const RtmPassParams& p, int32_t* hits /*or null*/,
                            float4* records /*or null*/, RtmCounters* counters /*hierarchy work, or null*/,
                            float* cps0 /*or null*/, cudaStream_t st);
void rtm_launch_head_image(const RtmSceneF64& sc, const RtmPassParams& p, const int32_t* hits,
                           double* image64 /*[H][W][3]*/, float4* accum, cudaStream_t st);
void rtm_launch_radiance_f64(const RtmSceneF64& sc, const RtmPassParams& p, float4* records,
                             RtmCounters* counters, cudaStream_t st);
void rtm_launch_closest_hit_f64(const RtmSceneF64& sc, uint32_t n, const double* org, const double* dir,
                                int32_t* hit, double* t, cudaStream_t st);

// cps0: per-sub-pixel c' = |c - P|^2 - r^2 at the primary hit from the FP64 pass, or null when the scene
// has no sphere small enough for the reference's self-intersection to matter (kernels_common.cuh: ACNE)
// nee: the next-event-estimation instantiation (needs hot.n_lights > 0)
// retry: the roulette-rounds instantiation (scenes whose visible surfaces kill most paths at once)
// clamp: evaluate the per-sample clamp of Renderer.cpp:240-241 (false only when the host proved it can never engage)
void rtm_launch_radiance_mega(const RtmHotScene& hot, int32_t n_spheres, int32_t n_planes, bool nee, bool retry, bool clamp,
                              const RtmPassParams& p, float4* records, RtmCounters* counters, int32_t sm_count, const float* cps0,
                              cudaStream_t st);
// walk: RTM_BVH_WALK_* (rtm_options.debug_bvh_walk), 0 = the product default
void rtm_launch_radiance_bvh(const RtmHotScene& hot, const RtmBvhDev& bvh, const RtmMatDev& mat, const RtmPassParams& p, float4* records,
                             RtmCounters* counters, int32_t sm_count, const float* cps0, int32_t n_spheres, int32_t n_planes,
                             int32_t walk, cudaStream_t st);
// the same pass with 64 rays per warp parked in shared memory (kernels_bvh_pool.cu); rtm_launch_radiance_bvh picks it when
// rtm_pool_kernel_takes() the scene (compact nodes, no planes, a tree its per-slot stacks can hold)
bool rtm_pool_kernel_takes(const RtmBvhDev& bvh, int32_t n_planes);
void rtm_launch_radiance_bvh_pool(const RtmHotScene& hot, const RtmBvhDev& bvh, const RtmMatDev& mat, const RtmPassParams& p, float4* records,
                                  RtmCounters* counters, int32_t sm_count, const float* cps0, int32_t n_spheres, cudaStream_t st);
// one wave of the wavefront variant; `wave` counts from 0 (its parity selects the queue)
void rtm_launch_wf_shade(const RtmHotScene& hot, int32_t n_spheres, bool use_bvh, const RtmBvhDev& bvh, const RtmMatDev& mat,
                         const RtmPassParams& p, float4* records, RtmCounters* counters, const float* cps0, const RtmWaveDev& wv,
                         int32_t wave, cudaStream_t st);
void rtm_launch_wf_extend(const RtmHotScene& hot, int32_t n_spheres, bool use_bvh, const RtmBvhDev& bvh, const RtmWaveDev& wv, int32_t wave,
                          RtmCounters* counters, int32_t sm_count, cudaStream_t st);
void rtm_launch_accumulate(const RtmPassParams& p, const float4* records, float4* accum, cudaStream_t st);
void rtm_launch_sum_partials(const float4* partials /*[G][n_pix]*/, float4* out, uint64_t n_pix, int32_t G, cudaStream_t st);
void rtm_launch_pack_rgb(const float4* accum, float* out /*[n_pix][3]*/, uint64_t n_pix, cudaStream_t st);
void rtm_launch_to_u8(const float4* accum, unsigned char* out, uint64_t n_pix, cudaStream_t st);
void rtm_launch_to_u8_f64(const double* image64 /*[n_pix][3]*/, unsigned char* out, uint64_t n_pix, cudaStream_t st);
void rtm_launch_closest_hit_f32(const RtmHotScene& hot, int32_t n_spheres, int32_t n_planes, const double* refd, uint32_t n, const double* org,
                                const double* dir, const int32_t* origin_id, int32_t* hit, double* t,
                                cudaStream_t st);
void rtm_launch_closest_hit_bvh_f32(const RtmHotScene& hot, const RtmBvhDev& bvh, const RtmMatDev& mat, int32_t n_spheres, int32_t n_planes,
                                    const double* refd, uint32_t n, const double* org, const double* dir, const int32_t* origin_id,
                                    int32_t* hit, double* t, cudaStream_t st);
void rtm_launch_acne_eps(uint32_t n, const uint32_t* bits, float* out, cudaStream_t st);
void rtm_launch_philox(uint32_t n, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t* out,
                       cudaStream_t st);
void rtm_launch_fma_peak(float* sink, int32_t blocks, int32_t iters, cudaStream_t st);
