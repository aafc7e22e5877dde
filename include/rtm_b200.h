/* include/rtm_b200.h — C ABI of the B200 renderer for RaytracingMin's radiance loop.
 *
 * The reference has no FFI of its own: the hot path sits behind the C++ class
 * png::Renderer (reference src/Renderer.h:9-17: ctor taking SettingData&, and
 * Render(std::string)), fed by the struct png::SettingData
 * (src/SettingData.h:43-51) and driven by src/main.cpp:41-44.  This header is
 * what an FFI for that path binds:
 *
 *   renderer_create   <-  png::Renderer::Renderer(SettingData&)   src/Renderer.cpp:20-23
 *   renderer_render   <-  png::Renderer::Render(std::string), the loop nest only
 *                         (src/Renderer.cpp:202-248); the 8-bit conversion of
 *                         :251-254 is renderer_read_u8, the stb writes of
 *                         :256-257 stay with the host caller
 *   renderer_destroy  <-  ~Renderer (implicit)
 *
 * Plain pointers and sizes only.  The library copies everything it needs in
 * renderer_create (the caller may free the scene arrays immediately), owns all
 * device memory, and never throws across the boundary: every entry point
 * returns 0 or a negative RTM_E_* code; renderer_last_error() holds the text
 * (thread-local).  There is no CPU fallback: without a CUDA device
 * renderer_create fails with RTM_E_CUDA.
 */
#ifndef RTM_B200_H
#define RTM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTM_ABI_VERSION 4

enum {
    RTM_OK = 0,
    RTM_E_INVALID = -1,   /* bad argument / scene the reference would crash on */
    RTM_E_CUDA = -2,      /* CUDA runtime error (text in renderer_last_error) */
    RTM_E_NCCL = -3,      /* NCCL error or libnccl not loadable (multi-GPU only) */
    RTM_E_UNSUPPORTED = -4
};

/* Semantics level. */
enum {
    RTM_COMPAT_HEAD = 0,  /* the reference exactly as written: normals never reach the
                             caller (src/SettingData.h:21), every sample evaluates to the
                             emission of the primary hit.  Deterministic; the double image
                             is bit-exact against the reference. */
    RTM_COMPAT_ABC = 1    /* the intended renderer (8 patched lines): diffuse GI with
                             Russian roulette.  Default. */
};

/* Arithmetic of the bounce loop (the primary-hit pass is always the exact FP64 one). */
enum {
    RTM_PRECISION_F32 = 0,  /* product path: FP32 bounces, cancellation-free sphere test */
    RTM_PRECISION_F64 = 1   /* reference-faithful: the reference's own double/float-mixed
                               formulas line by line (incl. sqrtf Normalize); slow, used as
                               the on-device statistical anchor */
};

/* Kernel organisation of the FP32 bounce loop. */
enum {
    RTM_VARIANT_MEGA = 0,      /* persistent megakernel with per-lane path refill */
    RTM_VARIANT_WAVEFRONT = 1  /* extend / shade kernels over compacted SoA ray queues */
};

/* Closest-hit search. */
enum {
    RTM_ACCEL_AUTO = 0,   /* brute force when n_spheres <= 64, else BVH */
    RTM_ACCEL_BRUTE = 1,  /* the reference's O(N) loop (src/Renderer.cpp:62-72) */
    RTM_ACCEL_BVH = 2     /* same per-candidate arithmetic and tie rule behind a BVH */
};

/* Estimator of the bounce loop. */
enum {
    RTM_ESTIMATOR_REFERENCE = 0,  /* src/Renderer.cpp:57-117: emission is collected where paths land */
    RTM_ESTIMATOR_NEE = 1         /* EXTENSION (the reference's SurfaeSample, src/Renderer.cpp:119-198,
                                     is dead, unfinished code): next-event estimation, one uniformly chosen sphere
                                     light per vertex, cone sampling + shadow ray.  Same expectation as the reference's
                                     estimator, less variance.  FP32 brute-force megakernel only. */
};

/* What png::SettingData holds after LoadData::from_json (src/SettingData.cpp:129-186),
 * flattened.  Doubles in, so the FP64 primary pass sees exactly what the reference parsed. */
typedef struct rtm_scene_desc {
    int32_t width, height;          /* "00 width", "00 height" */
    int32_t samples, superSamples;  /* "00 samples", "00 superSamples"; spp = samples*superSamples^2 */
    double cam_origin[3];           /* "01 camera".origin */
    double cam_target[3];           /* .target */
    double cam_up[3];               /* .upVec */
    float fov;                      /* .fov: FLOAT (src/SettingData.h:45), tangent-space half width */
    int32_t n_spheres;
    const double* position;         /* [n][3] SphereObject::m_position */
    const float* radius;            /* [n]    SphereObject::m_size (FLOAT, src/SettingData.h:31) */
    const double* color;            /* [n][3] Material::color */
    const double* emission;         /* [n][3] Material::emission */
    /* EXTENSION (ABI 3): PlaneObject, "00 objectType" 2 — declared by the reference
     * (src/SettingData.h:33-42) but never parsed and never intersected (src/SettingData.cpp:243-246
     * falls off the end), so there is no reference behaviour to match; DESIGN.md §10 defines it as a
     * finite square.  Arrays are the ctor arguments of src/SettingData.cpp:235.  Planes take the
     * object ids n_spheres .. n_spheres+n_planes-1 (hit map, tie rule).  COMPAT_ABC only; the FP32
     * path takes up to 16 of them, in both megakernels (brute force and hierarchy), not in the wavefront variant.
     * n_planes = 0 (and NULL arrays) = the reference's scenes. */
    int32_t n_planes;
    const double* plane_position;   /* [m][3] */
    const double* plane_up;         /* [m][3] */
    const double* plane_target;     /* [m][3] */
    const double* plane_width;      /* [m]    */
    const double* plane_color;      /* [m][3] */
    const double* plane_emission;   /* [m][3] */
} rtm_scene_desc;

typedef struct rtm_options {
    int32_t device;        /* CUDA ordinal of the (first) GPU */
    int32_t n_gpus;        /* 1: this device only.  >1: devices device..device+n_gpus-1 driven by
                              this process, sample budget sliced across them, partial buffers
                              summed with one ncclReduce to `device`.  COMPAT_HEAD is deterministic and
                              one cheap kernel: it always runs on `device` alone (rtm_stats.n_gpus = 1), so
                              its double image stays bit-exact */
    int32_t compat;        /* RTM_COMPAT_* */
    int32_t precision;     /* RTM_PRECISION_* */
    int32_t variant;       /* RTM_VARIANT_* */
    int32_t accel;         /* RTM_ACCEL_* */
    int32_t max_depth;     /* hits per path before forced termination; 0 = default (256).  The
                              reference recurses without bound (src/Renderer.cpp:108). */
    int32_t sample_begin;  /* slice [sample_begin, sample_end) of the "samples" loop */
    int32_t sample_end;    /*   (src/Renderer.cpp:226) this handle renders; 0,0 = all.  The divisor
                              of :240 stays the GLOBAL samples count, so slices add up exactly. */
    uint64_t seed;         /* RNG key; the image is a pure function of (scene, seed) */
    int32_t sub_begin;     /* slice [sub_begin, sub_end) of the superSamples^2 sub-pixel grid */
    int32_t sub_end;       /*   (src/Renderer.cpp:224-225, index (sx-1)*ss + (sy-1)) this handle renders;
                              0,0 = all.  The preferred multi-GPU partition when ss^2 divides by the GPU
                              count: every rank keeps full-length sample loops and a 1/G primary pass.
                              The primary-hit map needs the full grid. */
    int32_t estimator;     /* RTM_ESTIMATOR_* (ABI 3) */
    /* ABI 4: test / A-B knobs.  They replace the environment variables round 1's library read (a product .so takes
     * its switches through its interface, not through getenv); 0 everywhere = the product defaults. */
    int32_t debug_retry;        /* roulette-rounds megakernel: 0 = chosen from the scene, 1 = forced off, 2 = forced on
                                   (both render the same image bit for bit) */
    int32_t debug_bvh_walk;     /* hierarchy kernel, bit mask: 0 = default (traversal stack in shared memory when the tree
                                   is shallow enough; compact half-precision nodes when they fit the scene),
                                   bit 0 = stack in local memory, bit 1 = full-precision 64-byte nodes,
                                   bit 2 = inner nodes numbered breadth-first at the top and in 4-node treelets (one
                                   128-byte line each) below (default: breadth-first throughout; same walk, same image),
                                   bit 3 = the 4-wide collapse of the compact nodes (default: two children per node),
                                   bit 4 = the ray-pool kernel (64 rays per warp parked in shared memory, phases pick 32 in
                                   the same phase; default: one ray per lane).  Bits 3 and 4 are measured design
                                   alternatives that render the same image, not faster */
    int32_t debug_bvh_leaf_max; /* SAH build: largest leaf, 1..8; 0 = default (4) */
    float   debug_bvh_ctrav;    /* SAH build: cost of an inner-node visit in sphere tests; 0 = default (1.0) */
    int32_t debug_clamp;        /* 0 = the per-sample clamp (src/Renderer.cpp:240-241) is compiled out when the host can
                                   prove it never engages, 1 = always evaluated */
} rtm_options;

typedef struct rtm_stats {
    uint64_t samples;        /* innermost-loop iterations (src/Renderer.cpp:226) evaluated */
    uint64_t rays;           /* PathTracing invocations the reference would have made for them
                                (src/Renderer.cpp:57): sum of path lengths, primary included */
    uint64_t rays_traced;    /* closest-hit queries actually executed on the device (the primary
                                query is shared by the `samples` samples of a sub-pixel; with
                                RTM_ESTIMATOR_NEE the shadow rays are counted here, not in `rays`) */
    uint64_t sphere_tests;   /* ray-sphere tests actually executed */
    double device_ms;        /* CUDA-event time of the kernels (+ reduce), max over GPUs */
    double e2e_ms;           /* host wall time of the renderer_render call */
    int32_t kernel_launches; /* kernels launched by the call */
    int32_t n_gpus;
    uint64_t bvh_nodes;      /* hierarchy nodes visited (RTM_ACCEL_BVH), 0 for brute force */
} rtm_stats;

typedef struct rtm_renderer rtm_renderer;

int renderer_abi_version(void);

/* Fills *opt with the defaults (device 0, 1 GPU, ABC, F32, MEGA, AUTO, seed 0x5EED). */
void renderer_default_options(rtm_options* opt);

int renderer_create(const rtm_scene_desc* scene, const rtm_options* opt, rtm_renderer** out);

/* Host-buffer entry point (the drop-in call).  Renders the whole image and copies back.
 *   rgb_out          float[H][W][3] linear radiance sums as src/Renderer.cpp:246-248, or NULL
 *   primary_hit_out  int32[H][W][ss][ss] (axes y, x, sx-1, sy-1), -1 = miss, or NULL
 *   stats            or NULL */
int renderer_render(rtm_renderer* r, float* rgb_out, int32_t* primary_hit_out, rtm_stats* stats);

/* Device-resident entry point: enqueues the kernels of ONE pass on `stream`
 * (a cudaStream_t; NULL = the legacy default stream) and returns without
 * synchronising.  Results stay in the buffers below. */
int renderer_render_async(rtm_renderer* r, void* stream);

/* After the stream has been synchronised: counters of the last pass. */
int renderer_read_stats(rtm_renderer* r, rtm_stats* stats);

/* After the stream has been synchronised: CUDA-event durations (ms) of the three
 * kernels of the last FP32 ABC pass, measured on the stream they ran on:
 * ms3 = {primary FP64 pass, radiance megakernel, sub-pixel accumulate}. */
int renderer_read_kernel_ms(rtm_renderer* r, double* ms3);

/* Device pointers (valid until destroy): the float4[H][W] accumulation buffer
 * {r,g,b,0} and the int32 primary-hit map. */
void* renderer_device_accum(rtm_renderer* r);
void* renderer_device_hits(rtm_renderer* r);

/* Let the caller own the accumulation buffer (e.g. a registered NCCL send
 * buffer): float4[H][W] device memory on the handle's device. */
int renderer_set_accum_buffer(rtm_renderer* r, void* device_float4);

/* Peer-memory gather for one-process-per-GPU runs (replaces the NCCL reduce of the partial framebuffers):
 *   root:  renderer_peer_export(r, G, handle, &base)   allocates float4[G][H][W] on the root GPU, CUDA IPC handle (64 B)
 *   rank:  renderer_peer_open(r, handle, &base)        maps it; renderer_set_accum_buffer(r, base + g*H*W*16) makes the
 *          rank's LAST kernel store its partial framebuffer straight into the root's HBM over NVLink
 *   root:  after a cross-rank barrier on the stream, renderer_sum_partials(r, base, G, out, stream): out = sum of slots */
int renderer_peer_export(rtm_renderer* r, int32_t n_slots, void* handle64, void** device_ptr);
int renderer_peer_open(rtm_renderer* r, const void* handle64, void** device_ptr);
int renderer_sum_partials(rtm_renderer* r, const void* partials, int32_t n_slots, void* out_float4, void* stream);

/* The accumulation buffer as renderer_render returns it — float[H][W][3] in HOST memory — for callers that drive the
 * device-resident entry points themselves (one process per GPU: render_async on every rank, a sum-reduce of the partial
 * framebuffers into rank 0's buffer, then this call on rank 0).  The repack kernel and the copy are enqueued on `stream`
 * (a cudaStream_t, the one the pass / the reduce ran on; NULL = the legacy default stream, as in renderer_render_async),
 * which is synchronised before returning.  After a renderer_render call any stream will do (that call has synchronised). */
int renderer_read_rgb(rtm_renderer* r, float* rgb_out, void* stream);

/* COMPAT_HEAD only: the double image double[H][W][3], bit-exact against the reference. */
int renderer_read_image_f64(rtm_renderer* r, double* image_out);

/* src/Renderer.cpp:251-254 on the device: u8[H][W][3] = (unsigned char)(255*min(v,1)).  COMPAT_HEAD converts the
 * DOUBLE image (bytes identical to the reference's); COMPAT_ABC converts the float accumulation buffer. */
int renderer_read_u8(rtm_renderer* r, unsigned char* rgb8_out);

/* One FP32 pass that allocates nothing, for callers that may not call cudaMalloc (CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  No counterpart in the reference.
 * Same kernels and same result as renderer_render_async (primary pass, bounce kernel, sub-pixel accumulate), but the
 * pass's scratch — 16 B of record per sub-pixel (+ 4 B on scenes with very small spheres) and the work cursor — lives in
 * ONE fixed array of RTM_NOALLOC_SCRATCH_BYTES inside the library (every GPU that loads the library holds it until the
 * process exits) instead of the handle's own buffers.
 *   r                  a COMPAT_ABC / PRECISION_F32 / VARIANT_MEGA handle on one GPU (inputs: its scene and options)
 *   accum_out_float4   output, caller-owned device memory on the handle's device: float4[H][W] = {r,g,b,0}, overwritten
 *   stream             a cudaStream_t (NULL = the legacy default stream); the call enqueues and returns
 * Limit: W*H*(sub_end-sub_begin) sub-pixels must fit the array — at most 1,048,560 (838,848 on scenes with very small
 * spheres); larger frames are refused with RTM_E_UNSUPPORTED and a message naming both sizes.  The call runs no
 * cudaMalloc / cudaFree and no allocating library routine.  All calls on a device share the array: it keeps the last
 * call's contents (the call clears what must start at zero, the counters and the work cursor), and two calls that
 * overlap on the GPU — different streams, or two handles — race; order them on one stream or with events.
 * renderer_read_stats / renderer_read_kernel_ms keep describing the handle's last ordinary pass. */
#define RTM_NOALLOC_SCRATCH_BYTES (16u << 20)
int renderer_render_noalloc(rtm_renderer* r, void* accum_out_float4, void* stream);

void renderer_destroy(rtm_renderer* r);

const char* renderer_last_error(void);

/* ---- test hooks (deterministic building blocks, used by tests/ only) ---- */

/* Philox4x32-10 on the device: out[i] = philox(ctr = {i, c1, c2, c3}, key). */
int rtm_debug_philox(int32_t device, uint32_t n, uint32_t c1, uint32_t c2, uint32_t c3,
                     uint32_t key0, uint32_t key1, uint32_t* out4n);

/* One closest-hit query per ray through the handle's scene.
 *   mode 0: the exact FP64 test (bit-comparable with src/SettingData.cpp:197-226)
 *   mode 1: the FP32 product test; origin_id[i] >= 0 marks a ray leaving that sphere's surface
 * org/dir double[n][3]; hit_out int32[n]; t_out double[n]. */
int rtm_debug_closest_hit(rtm_renderer* r, int32_t mode, uint32_t n, const double* org, const double* dir,
                          const int32_t* origin_id, int32_t* hit_out, double* t_out);

/* The FP32 path's model of the reference's Normalize() error (ACNE): out[i] = eps drawn
 * from the 32 random bits bits[i]. */
int rtm_debug_acne_eps(int32_t device, uint32_t n, const uint32_t* bits, float* eps_out);

/* The sphere hierarchy renderer_create would build for `scene` with default options (host only, no GPU needed).
 * info[13] = {inner nodes, leaf slots, root code, depth, n_big, big_id[8]}.  Call once with the
 * three arrays NULL to size them: nodes_out float[16*nodes] (4 float4 per node: child boxes,
 * child codes, rmin), leaf_geo_out float[4*slots], leaf_id_out int32[slots]. */
int rtm_debug_bvh(const rtm_scene_desc* scene, int32_t* info, float* nodes_out, float* leaf_geo_out,
                  int32_t* leaf_id_out);

/* The compact form of that hierarchy which the FP32 walk uses (half-precision (centre, half-extent) child boxes,
 * 32 bytes per node; host only).  meta6 = {frame centre x y z, 1/scale, median leaf-box inflation, usable (0/1)};
 * nodes16_out uint32[8*nodes] or NULL: cx|cy, cz|hx, hy|hz of child 0, the same of child 1, the two child codes. */
int rtm_debug_bvh16(const rtm_scene_desc* scene, int32_t* n_nodes_out, float* meta6, uint32_t* nodes16_out);

/* The 4-wide collapse of the compact form (the walk behind rtm_options.debug_bvh_walk bit 3; host only).
 * info2 = {nodes (0: not usable), depth}; nodes_out uint32[16*nodes] or NULL: two halves laid out like a node of
 * rtm_debug_bvh16 — children 0,1 + their codes, children 2,3 + their codes; codes >= 0 index this array, negative codes
 * are the leaf codes of rtm_debug_bvh, 0x7FFFFFFF marks an empty slot (behind a box no ray hits). */
int rtm_debug_bvh16x4(const rtm_scene_desc* scene, int32_t* info2, uint32_t* nodes_out);

/* The 4-wide collapse of that hierarchy with float boxes (design study; host only).
 * info4[3] = {wide nodes, root code, depth}.  nodes4_out float[32*nodes] or NULL: 8 float4 per node —
 * lo.x[4] hi.x[4] lo.y[4] hi.y[4] lo.z[4] hi.z[4], child codes[4] (int bits, 0x7FFFFFFF = empty), rmin[4].
 * Leaf codes index the leaf arrays of rtm_debug_bvh. */
int rtm_debug_bvh4(const rtm_scene_desc* scene, int32_t* info4, float* nodes4_out);

/* FP32 FMA-chain microbenchmark: measured FP32 peak of the device in TFLOP/s. */
int rtm_debug_fp32_peak(int32_t device, double* tflops_out, double* sm_clock_mhz_out);

#ifdef __cplusplus
}
#endif
#endif /* RTM_B200_H */
