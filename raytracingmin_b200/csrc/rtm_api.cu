// rtm_api.cu — implementation of the C ABI in include/rtm_b200.h: scene
// flattening, device memory, launch orchestration, the single-process multi-GPU
// path (sample-budget slices + one ncclReduce) and the test hooks.
//
// Host arithmetic that feeds the bit-exact kernels (camera frame, r*r, kd) is
// done here in the reference's own types and order (src/Renderer.cpp:202-208,
// src/SettingData.cpp:200, src/SettingData.h:14-16); this file is compiled with
// -ffp-contract=off.
#include "../../include/rtm_b200.h"
#include "rtm_internal.h"
#include "bvh.h"

#include <dlfcn.h>
#include <nccl.h>  // types only; the library is dlopen'ed when n_gpus > 1

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

namespace {

thread_local std::string g_err;

int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return fail(RTM_E_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

// ---- NCCL through dlopen (only the six calls the reduce needs) ----------------
struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi g_nccl;
std::mutex g_nccl_mu;

int load_nccl() {
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    if (g_nccl.lib) return RTM_OK;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    void* lib = nullptr;
    for (const char* nm : names) {
        lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (lib) break;
    }
    if (!lib) return fail(RTM_E_NCCL, "cannot dlopen libnccl.so.2: %s", dlerror());
    NcclApi a;
    a.lib = lib;
    a.CommInitAll = (decltype(a.CommInitAll))dlsym(lib, "ncclCommInitAll");
    a.CommDestroy = (decltype(a.CommDestroy))dlsym(lib, "ncclCommDestroy");
    a.Reduce = (decltype(a.Reduce))dlsym(lib, "ncclReduce");
    a.GroupStart = (decltype(a.GroupStart))dlsym(lib, "ncclGroupStart");
    a.GroupEnd = (decltype(a.GroupEnd))dlsym(lib, "ncclGroupEnd");
    a.GetErrorString = (decltype(a.GetErrorString))dlsym(lib, "ncclGetErrorString");
    if (!a.CommInitAll || !a.CommDestroy || !a.Reduce || !a.GroupStart || !a.GroupEnd || !a.GetErrorString)
        return fail(RTM_E_NCCL, "libnccl is missing a required symbol");
    g_nccl = a;
    return RTM_OK;
}

#define NC(call)                                                                                            \
    do {                                                                                                    \
        ncclResult_t r_ = (call);                                                                           \
        if (r_ != ncclSuccess) return fail(RTM_E_NCCL, "%s: %s", #call, g_nccl.GetErrorString(r_));        \
    } while (0)

// ---- host-side reference arithmetic ------------------------------------------
struct v3 {
    double x, y, z;
};
v3 vsub(v3 a, v3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
v3 vneg(v3 a) { return {-a.x, -a.y, -a.z}; }
v3 vscale(v3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
double vdot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }  // Ray.h:61-63
v3 vcross(v3 a, v3 b) { return {a.y * b.z - a.z * b.y, -a.x * b.z + a.z * b.x, a.x * b.y - a.y * b.x}; }  // Ray.h:64-66
v3 vnormalize(v3 a) {                                                                                        // Ray.h:67-72
    const double m = (double)sqrtf((float)(a.x * a.x + a.y * a.y + a.z * a.z));
    return {a.x / m, a.y / m, a.z / m};
}

struct GpuCtx {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t ek[4] = {nullptr, nullptr, nullptr, nullptr};  // around primary / radiance / accumulate
    bool ek_valid = false;
    double* d_scene64 = nullptr;  // one allocation: cx cy cz r2 kd [n] + emission color [3n]
    RtmSceneF64 s64{};
    float4* records = nullptr;
    float4* accum = nullptr;
    bool accum_owned = true;
    int32_t* hits = nullptr;
    double* image64 = nullptr;
    unsigned char* u8 = nullptr;
    RtmCounters* counters = nullptr;
    float* cps0 = nullptr;      // per sub-pixel c' at the primary hit (ACNE scenes only)
    float4* d_wave = nullptr;   // wavefront variant: one allocation behind `wv`
    RtmWaveDev wv{};
    unsigned* h_live = nullptr; // pinned: wv.ctrl[3] read back between batches of waves
    int32_t waves = 0;          // waves of the last wavefront pass
    float4* d_bvh = nullptr;    // one allocation: nodes | leaf_geo | geo | emi | alb | leaf_id
    RtmBvhDev bvh{};
    RtmMatDev mat{};
    float* d_rgb = nullptr;     // float[H][W][3] staging for renderer_render's rgb_out
    RtmSceneF32Host* d_f32 = nullptr;   // this handle's device copy of the FP32 per-sphere block (no process-global scene)
    RtmHotScene hot{};                  // kernel-parameter part of it (hot.full = d_f32)
    float4* peer_stage = nullptr;   // exported gather buffer: G partial framebuffers written by peer GPUs
    void* peer_opened = nullptr;    // a peer's gather buffer opened through CUDA IPC
    int32_t s_begin = 0, s_end = 0;
    ncclComm_t comm = nullptr;
};

}  // namespace

struct rtm_renderer {
    rtm_options opt{};
    int32_t W = 0, H = 0, S = 0, ss = 0, n = 0, n_planes = 0;
    uint64_t n_pix = 0, n_sub = 0;
    RtmFrame frame{};
    RtmSceneF32Host f32{};
    double refd[3]{};
    std::vector<GpuCtx> gpus;
    int32_t last_launches = 0;
    bool hits_valid = false;
    bool retry = false;         // roulette-rounds megakernel: the surfaces the camera sees kill most paths at once
    bool clamp = true;          // the per-sample clamp of Renderer.cpp:240-241 can engage (false: proven impossible for this scene)
    bool nee = false;           // next-event-estimation kernels (estimator NEE and at least one listed light)
    bool acne = false;          // the FP32 path models the reference's self-intersection on small spheres
    bool use_bvh = false;       // closest hit through the hierarchy (bvh.h) instead of the O(N) loop
    RtmBvhHost bvh_host;
    RtmBvh16Host bvh16;         // compact form of bvh_host for the FP32 walk (ok = false: the float nodes are walked)
    RtmBvh16x4Host bvh16x4;     // its 4-wide collapse (ok = false: the two-child compact nodes are walked)
    std::vector<float4> mat_geo, mat_emi, mat_alb;
};

namespace {

void free_ctx(GpuCtx& c) {
    cudaSetDevice(c.device);
    if (c.comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c.comm);
    if (c.d_scene64) cudaFree(c.d_scene64);
    if (c.records) cudaFree(c.records);
    if (c.accum && c.accum_owned) cudaFree(c.accum);
    if (c.hits) cudaFree(c.hits);
    if (c.image64) cudaFree(c.image64);
    if (c.u8) cudaFree(c.u8);
    if (c.counters) cudaFree(c.counters);
    if (c.d_bvh) cudaFree(c.d_bvh);
    if (c.cps0) cudaFree(c.cps0);
    if (c.d_wave) cudaFree(c.d_wave);
    if (c.h_live) cudaFreeHost(c.h_live);
    if (c.d_rgb) cudaFree(c.d_rgb);
    if (c.d_f32) cudaFree(c.d_f32);
    if (c.peer_opened) cudaIpcCloseMemHandle(c.peer_opened);
    if (c.peer_stage) cudaFree(c.peer_stage);
    if (c.ev0) cudaEventDestroy(c.ev0);
    if (c.ev1) cudaEventDestroy(c.ev1);
    for (auto& e : c.ek) if (e) cudaEventDestroy(e);
    if (c.stream) cudaStreamDestroy(c.stream);
    c = GpuCtx{};
}

RtmPassParams make_params(const rtm_renderer* r, const GpuCtx& c) {
    RtmPassParams p{};
    p.frame = r->frame;
    p.s_begin = c.s_begin;
    p.s_end = c.s_end;
    p.max_depth = r->opt.max_depth;
    p.seed_lo = (uint32_t)r->opt.seed;
    p.seed_hi = (uint32_t)(r->opt.seed >> 32);
    p.inv_norm = (float)(1.0 / ((double)r->ss * r->ss * r->S));
    p.key_mask = 0xFFFFFFF8u;
    {
        // 32-bit Philox2x32 key from the 64-bit seed
        uint32_t k = p.seed_lo ^ (p.seed_hi * 0x85EBCA6Bu) ^ (p.seed_hi >> 15);
        for (int i = 0; i < 11; ++i) p.rk[i] = k + (uint32_t)i * 0x9E3779B9u;   // rk[1]: key of the NEE draws
    }
    for (int k = 0; k < 3; ++k) {
        p.refd[k] = r->refd[k];
        p.ref[k] = (float)r->refd[k];
    }
    p.n_sub = r->n_sub;
    p.sub_begin = r->opt.sub_begin;
    p.sub_count = r->opt.sub_end - r->opt.sub_begin;
    return p;
}

int ensure_hits(rtm_renderer* r, GpuCtx& c) {
    if (!c.hits) CU(cudaMalloc(&c.hits, r->n_sub * sizeof(int32_t)));
    return RTM_OK;
}

// Enqueue one full pass on ctx c / stream st.  want_hits: also write the hit map.
int enqueue_pass(rtm_renderer* r, GpuCtx& c, cudaStream_t st, bool want_hits, int* launches) {
    CU(cudaSetDevice(c.device));
    const RtmPassParams p = make_params(r, c);
    CU(cudaMemsetAsync(c.counters, 0, sizeof(RtmCounters), st));
    if (r->opt.compat == RTM_COMPAT_HEAD) {
        int rc = ensure_hits(r, c);
        if (rc) return rc;
        if (!c.image64) CU(cudaMalloc(&c.image64, r->n_pix * 3 * sizeof(double)));
        rtm_launch_primary_f64(c.s64, p, c.hits, nullptr, nullptr, nullptr, st);
        rtm_launch_head_image(c.s64, p, c.hits, c.image64, c.accum, st);
        *launches += 2;
    } else if (r->opt.precision == RTM_PRECISION_F64) {
        if (want_hits) {
            int rc = ensure_hits(r, c);
            if (rc) return rc;
            rtm_launch_primary_f64(c.s64, p, c.hits, nullptr, nullptr, nullptr, st);
            *launches += 1;
        }
        rtm_launch_radiance_f64(c.s64, p, c.records, c.counters, st);
        rtm_launch_accumulate(p, c.records, c.accum, st);
        *launches += 2;
    } else {
        if (want_hits) {
            int rc = ensure_hits(r, c);
            if (rc) return rc;
        }
        CU(cudaEventRecord(c.ek[0], st));
        rtm_launch_primary_f64(c.s64, p, want_hits ? c.hits : nullptr, c.records, c.counters, c.cps0, st);
        CU(cudaEventRecord(c.ek[1], st));
        if (r->opt.variant == RTM_VARIANT_WAVEFRONT) {
            // waves until a shade kernel queues no ray; the host looks at the device flag every 32 waves
            // (the only host synchronisations of the pass: this variant is not fully asynchronous)
            CU(cudaMemsetAsync(c.wv.ctrl, 0, 4 * sizeof(unsigned), st));
            int wave = 0;
            for (;;) {
                for (int k = 0; k < 32; ++k, ++wave) {
                    rtm_launch_wf_shade(c.hot, r->n, r->use_bvh, c.bvh, c.mat, p, c.records, c.counters, c.cps0, c.wv, wave, st);
                    rtm_launch_wf_extend(c.hot, r->n, r->use_bvh, c.bvh, c.wv, wave, c.counters, c.sm_count, st);
                }
                *launches += 64;
                CU(cudaMemcpyAsync(c.h_live, c.wv.ctrl + 3, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
                CU(cudaStreamSynchronize(st));
                if (*c.h_live != (unsigned)wave) break;   // the last wave of the batch queued nothing
            }
            c.waves = wave;
        } else if (r->use_bvh) rtm_launch_radiance_bvh(c.hot, c.bvh, c.mat, p, c.records, c.counters, c.sm_count, c.cps0, r->n, r->n_planes, r->opt.debug_bvh_walk, st);
        else rtm_launch_radiance_mega(c.hot, r->n, r->n_planes, r->nee, r->retry, r->clamp, p, c.records, c.counters, c.sm_count, c.cps0, st);
        CU(cudaEventRecord(c.ek[2], st));
        rtm_launch_accumulate(p, c.records, c.accum, st);
        CU(cudaEventRecord(c.ek[3], st));
        c.ek_valid = true;
        *launches += 3;
    }
    CU(cudaGetLastError());
    return RTM_OK;
}

int fill_stats(rtm_renderer* r, rtm_stats* st) {
    memset(st, 0, sizeof *st);
    for (auto& c : r->gpus) {
        CU(cudaSetDevice(c.device));
        RtmCounters h{};
        CU(cudaMemcpy(&h, c.counters, sizeof h, cudaMemcpyDeviceToHost));
        const uint64_t ns = (uint64_t)(c.s_end - c.s_begin);
        st->samples += r->n_sub * ns;
        if (r->opt.compat == RTM_COMPAT_HEAD) {
            // one closest-hit query per sub-pixel on the device; the reference makes
            // 2 PathTracing calls per sample that hits a kd>0 surface, else 1 — reported
            // conservatively as 1 per sample.
            st->rays += r->n_sub * ns;
            st->rays_traced += r->n_sub;
            st->sphere_tests += r->n_sub * (uint64_t)r->n;
        } else if (r->opt.precision == RTM_PRECISION_F64) {
            st->rays += h.rays_ref;
            st->rays_traced += h.rays_traced;
            st->sphere_tests += h.tests;
        } else {
            st->rays += h.rays_ref;
            st->rays_traced += h.rays_traced + r->n_sub;
            // hierarchy path: the primary pass counts its own tests into the same counter
            st->sphere_tests += h.tests + (r->use_bvh ? 0 : r->n_sub * (uint64_t)r->n);
        }
        st->bvh_nodes += h.nodes;
    }
    st->n_gpus = (int32_t)r->gpus.size();
    st->kernel_launches = r->last_launches;
    return RTM_OK;
}

}  // namespace

// ---- renderer_render_noalloc: the pass with its scratch in a fixed device array -------------------------------
// Scratch of one FP32 pass = the sub-pixel records the primary pass hands to the bounce kernel (16 B each), the
// per-sub-pixel c' of ACNE scenes (4 B each) and the counter block with the work cursor.  The handle keeps them in
// cudaMalloc'd buffers; this entry point keeps them HERE, in an array that every device which loads the library holds
// until the process exits, so the call itself allocates nothing.  All calls on a device share the array.
namespace {

__device__ __align__(256) static unsigned char g_noalloc_scratch[RTM_NOALLOC_SCRATCH_BYTES];

constexpr int kNoallocMaxDevices = 64;
std::mutex g_noalloc_mu;
unsigned char* g_noalloc_dev[kNoallocMaxDevices];   // device address of g_noalloc_scratch per device, resolved once

// In host code `g_noalloc_scratch` is only the symbol's host-side stand-in; the address a kernel may use comes from
// cudaGetSymbolAddress with the owning device current.
int noalloc_scratch(int device, unsigned char** out) {
    if (device < 0 || device >= kNoallocMaxDevices) return fail(RTM_E_UNSUPPORTED, "device %d beyond the no-alloc table (%d)", device, kNoallocMaxDevices);
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_dev[device]) {
        CU(cudaSetDevice(device));
        void* ptr = nullptr;
        CU(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
        if (!ptr || ((uintptr_t)ptr & 255u)) return fail(RTM_E_CUDA, "scratch symbol resolved to %p", ptr);
        g_noalloc_dev[device] = (unsigned char*)ptr;
    }
    *out = g_noalloc_dev[device];
    return RTM_OK;
}

}  // namespace

extern "C" {

int renderer_render_noalloc(rtm_renderer* r, void* accum_out_float4, void* stream) {
    if (!r || !accum_out_float4) return fail(RTM_E_INVALID, "null argument");
    if (r->gpus.size() != 1) return fail(RTM_E_UNSUPPORTED, "renderer_render_noalloc drives one GPU per handle");
    if (r->opt.compat != RTM_COMPAT_ABC || r->opt.precision != RTM_PRECISION_F32 || r->opt.variant == RTM_VARIANT_WAVEFRONT)
        return fail(RTM_E_UNSUPPORTED, "renderer_render_noalloc runs the FP32 ABC megakernel pass only");
    GpuCtx& c = r->gpus[0];
    // layout: [counters, 256 B] [records, 16 B x n_sub] [c', 4 B x n_sub (ACNE scenes)]
    const uint64_t per_sub = sizeof(float4) + (c.cps0 ? sizeof(float) : 0);
    const uint64_t max_sub = ((uint64_t)RTM_NOALLOC_SCRATCH_BYTES - 256u) / per_sub;
    if (r->n_sub > max_sub)
        return fail(RTM_E_UNSUPPORTED, "renderer_render_noalloc: %llu sub-pixels need %llu bytes of scratch, the fixed array holds %llu (at most %llu sub-pixels of this scene)",
                    (unsigned long long)r->n_sub, (unsigned long long)(256u + r->n_sub * per_sub),
                    (unsigned long long)RTM_NOALLOC_SCRATCH_BYTES, (unsigned long long)max_sub);
    unsigned char* base = nullptr;
    int rc = noalloc_scratch(c.device, &base);
    if (rc) return rc;
    CU(cudaSetDevice(c.device));
    cudaStream_t st = (cudaStream_t)stream;
    RtmCounters* counters = (RtmCounters*)base;
    float4* records = (float4*)(base + 256);
    float* cps0 = c.cps0 ? (float*)(base + 256 + r->n_sub * sizeof(float4)) : nullptr;
    const RtmPassParams p = make_params(r, c);
    // the array keeps the previous call's contents: the work cursor and the counters must start at zero; every record
    // (and c') is written by the primary pass before the bounce kernel reads it
    CU(cudaMemsetAsync(counters, 0, sizeof(RtmCounters), st));
    rtm_launch_primary_f64(c.s64, p, nullptr, records, counters, cps0, st);
    if (r->use_bvh) rtm_launch_radiance_bvh(c.hot, c.bvh, c.mat, p, records, counters, c.sm_count, cps0, r->n, r->n_planes, r->opt.debug_bvh_walk, st);
    else rtm_launch_radiance_mega(c.hot, r->n, r->n_planes, r->nee, r->retry, r->clamp, p, records, counters, c.sm_count, cps0, st);
    rtm_launch_accumulate(p, records, (float4*)accum_out_float4, st);
    CU(cudaGetLastError());
    return RTM_OK;
}

int renderer_abi_version(void) { return RTM_ABI_VERSION; }

const char* renderer_last_error(void) { return g_err.c_str(); }

void renderer_default_options(rtm_options* o) {
    if (!o) return;
    memset(o, 0, sizeof *o);
    o->device = 0;
    o->n_gpus = 1;
    o->compat = RTM_COMPAT_ABC;
    o->precision = RTM_PRECISION_F32;
    o->variant = RTM_VARIANT_MEGA;
    o->accel = RTM_ACCEL_AUTO;
    o->max_depth = 0;
    o->seed = 0x5EEDull;
    o->estimator = RTM_ESTIMATOR_REFERENCE;
    // debug_* stay 0: the product defaults
}

void renderer_destroy(rtm_renderer* r) {
    if (!r) return;
    for (auto& c : r->gpus) free_ctx(c);
    delete r;
}

int renderer_create(const rtm_scene_desc* sc, const rtm_options* opt_in, rtm_renderer** out) {
    if (!sc || !out) return fail(RTM_E_INVALID, "renderer_create: null argument");
    *out = nullptr;
    rtm_options opt;
    if (opt_in) opt = *opt_in;
    else renderer_default_options(&opt);
    if (sc->width <= 0 || sc->height <= 0 || sc->samples <= 0 || sc->superSamples <= 0)
        return fail(RTM_E_INVALID, "width/height/samples/superSamples must be positive (got %d %d %d %d)", sc->width,
                    sc->height, sc->samples, sc->superSamples);
    if (sc->n_spheres < 0 || (sc->n_spheres > 0 && (!sc->position || !sc->radius || !sc->color || !sc->emission)))
        return fail(RTM_E_INVALID, "scene arrays missing for %d spheres", sc->n_spheres);
    if (sc->n_spheres > (int32_t)RTM_REC_IDMASK) return fail(RTM_E_INVALID, "too many spheres");
    const int m = sc->n_planes;
    if (m < 0 || (m > 0 && (!sc->plane_position || !sc->plane_up || !sc->plane_target || !sc->plane_width || !sc->plane_color ||
                            !sc->plane_emission)))
        return fail(RTM_E_INVALID, "scene arrays missing for %d planes", m);
    if (opt.n_gpus < 1) opt.n_gpus = 1;
    // COMPAT_HEAD is deterministic and one cheap kernel: all samples run on `device`, so that image64 (and the
    // bytes derived from it) stay the reference's, whatever n_gpus says
    if (opt.compat == RTM_COMPAT_HEAD) opt.n_gpus = 1;
    if (opt.max_depth <= 0) opt.max_depth = 256;
    if (opt.compat != RTM_COMPAT_HEAD && opt.compat != RTM_COMPAT_ABC) return fail(RTM_E_INVALID, "bad compat %d", opt.compat);
    if (opt.precision != RTM_PRECISION_F32 && opt.precision != RTM_PRECISION_F64)
        return fail(RTM_E_INVALID, "bad precision %d", opt.precision);
    if (opt.sample_begin == 0 && opt.sample_end == 0) opt.sample_end = sc->samples;
    if (opt.sample_begin < 0 || opt.sample_end > sc->samples || opt.sample_begin >= opt.sample_end)
        return fail(RTM_E_INVALID, "bad sample slice [%d,%d) of %d", opt.sample_begin, opt.sample_end, sc->samples);
    if (opt.max_depth > 511) return fail(RTM_E_INVALID, "max_depth %d exceeds 511 (the RNG counter holds 9 depth bits)", opt.max_depth);
    if (sc->samples >= (1 << 23)) return fail(RTM_E_INVALID, "samples %d exceeds 2^23-1", sc->samples);
    if ((uint64_t)sc->width * sc->height * sc->superSamples * sc->superSamples >= (1ull << 32))
        return fail(RTM_E_INVALID, "width*height*superSamples^2 must stay below 2^32 sub-pixels");
    if (opt.sub_begin == 0 && opt.sub_end == 0) opt.sub_end = sc->superSamples * sc->superSamples;
    if (opt.sub_begin < 0 || opt.sub_end > sc->superSamples * sc->superSamples || opt.sub_begin >= opt.sub_end)
        return fail(RTM_E_INVALID, "bad sub-pixel slice [%d,%d) of %d", opt.sub_begin, opt.sub_end,
                    sc->superSamples * sc->superSamples);
    const bool full_grid = (opt.sub_end - opt.sub_begin) == sc->superSamples * sc->superSamples;
    if (!full_grid && (opt.compat == RTM_COMPAT_HEAD || opt.n_gpus > 1))
        return fail(RTM_E_UNSUPPORTED, "sub-pixel slices need COMPAT_ABC and one GPU per handle");
    if (opt.variant != RTM_VARIANT_MEGA && opt.variant != RTM_VARIANT_WAVEFRONT) return fail(RTM_E_INVALID, "bad variant %d", opt.variant);
    const bool f32_path = (opt.compat == RTM_COMPAT_ABC && opt.precision == RTM_PRECISION_F32);
    if (opt.variant == RTM_VARIANT_WAVEFRONT && (!f32_path || opt.n_gpus > 1))
        return fail(RTM_E_UNSUPPORTED, "the wavefront variant exists for COMPAT_ABC + PRECISION_F32 with one GPU per handle");
    if (opt.accel != RTM_ACCEL_AUTO && opt.accel != RTM_ACCEL_BRUTE && opt.accel != RTM_ACCEL_BVH)
        return fail(RTM_E_INVALID, "bad accel %d", opt.accel);
    const bool use_bvh = opt.accel == RTM_ACCEL_BVH || (opt.accel == RTM_ACCEL_AUTO && sc->n_spheres > 64);
    if (f32_path && !use_bvh && sc->n_spheres + m > RTM_MAX_CONST_SPHERES)
        return fail(RTM_E_UNSUPPORTED, "FP32 brute-force path holds at most %d objects (use RTM_ACCEL_BVH / AUTO)",
                    RTM_MAX_CONST_SPHERES);
    if (opt.debug_retry < 0 || opt.debug_retry > 2 || opt.debug_bvh_walk < 0 || opt.debug_bvh_walk > 31 || opt.debug_bvh_leaf_max < 0 ||
        opt.debug_bvh_leaf_max > 8 || !(opt.debug_bvh_ctrav >= 0.f) || opt.debug_clamp < 0 || opt.debug_clamp > 1)
        return fail(RTM_E_INVALID, "bad debug_* option");
    if (opt.variant == RTM_VARIANT_WAVEFRONT && sc->samples >= (1 << 22))
        return fail(RTM_E_INVALID, "samples %d exceeds 2^22-1 (wavefront variant)", sc->samples);
    if (opt.estimator != RTM_ESTIMATOR_REFERENCE && opt.estimator != RTM_ESTIMATOR_NEE)
        return fail(RTM_E_INVALID, "bad estimator %d", opt.estimator);
    if (opt.estimator == RTM_ESTIMATOR_NEE && (!f32_path || use_bvh || opt.variant != RTM_VARIANT_MEGA))
        return fail(RTM_E_UNSUPPORTED, "next-event estimation exists in the FP32 brute-force megakernel (COMPAT_ABC, PRECISION_F32, "
                    "VARIANT_MEGA, <= 64 spheres or RTM_ACCEL_BRUTE)");
    if (m > 0 && opt.compat != RTM_COMPAT_ABC) return fail(RTM_E_UNSUPPORTED, "planes are an extension of COMPAT_ABC");
    if (m > 0 && f32_path && (opt.variant != RTM_VARIANT_MEGA || m > RTM_MAX_PLANES))
        return fail(RTM_E_UNSUPPORTED, "FP32 path: planes go through the megakernels (RTM_VARIANT_MEGA, <= %d planes); "
                    "use RTM_PRECISION_F64 otherwise", RTM_MAX_PLANES);

    int ndev = 0;
    cudaError_t ce = cudaGetDeviceCount(&ndev);
    if (ce != cudaSuccess || ndev == 0)
        return fail(RTM_E_CUDA, "no CUDA device (%s); this library has no CPU fallback", cudaGetErrorString(ce));
    if (opt.device < 0 || opt.device + opt.n_gpus > ndev)
        return fail(RTM_E_INVALID, "devices %d..%d requested, %d present", opt.device, opt.device + opt.n_gpus - 1, ndev);
    if (opt.n_gpus > 1 && (opt.sample_end - opt.sample_begin) < opt.n_gpus)
        return fail(RTM_E_INVALID, "cannot slice %d samples over %d GPUs", opt.sample_end - opt.sample_begin, opt.n_gpus);

    rtm_renderer* r = new rtm_renderer();
    r->opt = opt;
    r->W = sc->width; r->H = sc->height; r->S = sc->samples; r->ss = sc->superSamples; r->n = sc->n_spheres;
    r->n_planes = m;
    r->use_bvh = use_bvh;
    r->n_pix = (uint64_t)r->W * r->H;
    r->n_sub = r->n_pix * (uint64_t)(opt.sub_end - opt.sub_begin);   // sub-pixels this handle renders

    // camera frame, src/Renderer.cpp:202-208
    {
        const v3 origin{sc->cam_origin[0], sc->cam_origin[1], sc->cam_origin[2]};
        const v3 target{sc->cam_target[0], sc->cam_target[1], sc->cam_target[2]};
        const v3 up{sc->cam_up[0], sc->cam_up[1], sc->cam_up[2]};
        const v3 direction = vnormalize(vsub(target, origin));
        const v3 camX = vneg(vnormalize(vcross(direction, up)));
        const v3 camY = vcross(camX, direction);
        const double fovx = (double)sc->fov;
        const double fovy = fovx * r->H / r->W;
        const v3 ax = vscale(camX, fovx), ay = vscale(camY, fovy);
        RtmFrame& f = r->frame;
        f.o[0] = origin.x; f.o[1] = origin.y; f.o[2] = origin.z;
        f.ax[0] = ax.x; f.ax[1] = ax.y; f.ax[2] = ax.z;
        f.ay[0] = ay.x; f.ay[1] = ay.y; f.ay[2] = ay.z;
        f.az[0] = direction.x; f.az[1] = direction.y; f.az[2] = direction.z;
        f.rate = (float)(1.0 / (1 + r->ss));  // Renderer.cpp:227
        f.W = r->W; f.H = r->H; f.ss = r->ss; f.S = r->S;
        r->refd[0] = origin.x; r->refd[1] = origin.y; r->refd[2] = origin.z;
    }

    // scene arrays
    const int n = r->n;
    // cx cy cz r2 [n] | kd [n+m] | emission [3(n+m)] | color [3(n+m)] | planes [16 m]
    const size_t nm = (size_t)n + (size_t)m;
    std::vector<double> h64((size_t)n * 4 + nm * 7 + (size_t)m * 16 + 1);
    double* cx = h64.data();
    double* cy = cx + n;
    double* cz = cy + n;
    double* r2 = cz + n;
    double* kd = r2 + n;
    double* em = kd + nm;
    double* col = em + 3 * nm;
    double* planes = col + 3 * nm;
    r->f32.n = n;
    r->f32.n_planes = m;
    for (int i = 0; i < n; ++i) {
        cx[i] = sc->position[3 * i]; cy[i] = sc->position[3 * i + 1]; cz[i] = sc->position[3 * i + 2];
        const float rad = sc->radius[i];
        const float rr = rad * rad;  // float product, SettingData.cpp:200
        r2[i] = (double)rr;
        const double c0 = sc->color[3 * i], c1 = sc->color[3 * i + 1], c2 = sc->color[3 * i + 2];
        const double m01 = (c0 < c1) ? c1 : c0;
        const float kdf = (float)((m01 < c2) ? c2 : m01);  // SettingData.h:14-16
        kd[i] = (double)kdf;
        for (int k = 0; k < 3; ++k) {
            em[3 * i + k] = sc->emission[3 * i + k];
            col[3 * i + k] = sc->color[3 * i + k];
        }
        if (f32_path) {
            const double px = cx[i] - r->refd[0], py = cy[i] - r->refd[1], pz = cz[i] - r->refd[2];
            const float fcx = (float)px, fcy = (float)py, fcz = (float)pz;
            const float inv_r = (float)(1.0 / (double)rad);
            const double th = (double)kdf * 4294967296.0;
            const uint32_t thr = (th >= 4294967295.0) ? 0xFFFFFFFFu : (th <= 0.0 ? 0u : (uint32_t)th);
            float kx = 0.f, ky = 0.f, kz = 0.f;
            if (kdf > 0.f) { kx = (float)(c0 / (double)kdf); ky = (float)(c1 / (double)kdf); kz = (float)(c2 / (double)kdf); }
            if (use_bvh) {
                float4 e = make_float4((float)em[3 * i], (float)em[3 * i + 1], (float)em[3 * i + 2], 0.f);
                memcpy(&e.w, &thr, 4);
                float4 a = make_float4(kx, ky, kz, 0.f);
                const int32_t none = -1;
                memcpy(&a.w, &none, 4);
                r->mat_geo.push_back(make_float4(fcx, fcy, fcz, inv_r));
                r->mat_emi.push_back(e);
                r->mat_alb.push_back(a);
            } else {
                RtmSceneF32Host& s = r->f32;
                s.cx[i] = fcx; s.cy[i] = fcy; s.cz[i] = fcz;
                s.m2cx[i] = -2.0f * fcx; s.m2cy[i] = -2.0f * fcy; s.m2cz[i] = -2.0f * fcz;
                s.inv_r[i] = inv_r;
                s.ex[i] = (float)em[3 * i]; s.ey[i] = (float)em[3 * i + 1]; s.ez[i] = (float)em[3 * i + 2];
                s.kd[i] = kdf;
                s.kd_thr[i] = thr;
                s.kx[i] = kx; s.ky[i] = ky; s.kz[i] = kz;
            }
        }
    }

    // PlaneObject extension: frame as the ctor of src/SettingData.cpp:235-242 derives it (float-sqrt Normalize),
    // operation order fixed by DESIGN.md §10 (the tests compare the hit maps bit for bit)
    for (int j = 0; j < m; ++j) {
        const v3 position{sc->plane_position[3 * j], sc->plane_position[3 * j + 1], sc->plane_position[3 * j + 2]};
        const v3 up{sc->plane_up[3 * j], sc->plane_up[3 * j + 1], sc->plane_up[3 * j + 2]};
        const v3 target{sc->plane_target[3 * j], sc->plane_target[3 * j + 1], sc->plane_target[3 * j + 2]};
        const v3 normal = vnormalize(vsub(target, position));
        const v3 right = vscale(vscale(vnormalize(vcross(normal, up)), 0.5), sc->plane_width[j]);
        const v3 upp = vcross(right, normal);
        const double rr = vdot(right, right), uu = vdot(upp, upp);
        if (!(rr > 0.0) || !(uu > 0.0) || !std::isfinite(rr) || !std::isfinite(uu) || !std::isfinite(vdot(normal, normal))) {
            delete r;
            return fail(RTM_E_INVALID, "plane %d is degenerate (target == position, up parallel to the normal, or width <= 0)", j);
        }
        double* pl = planes + 16 * (size_t)j;
        pl[0] = normal.x; pl[1] = normal.y; pl[2] = normal.z;
        pl[3] = position.x; pl[4] = position.y; pl[5] = position.z;
        pl[6] = right.x; pl[7] = right.y; pl[8] = right.z;
        pl[9] = upp.x; pl[10] = upp.y; pl[11] = upp.z;
        pl[12] = rr; pl[13] = uu; pl[14] = pl[15] = 0.0;
        const size_t i = (size_t)n + j;
        const double c0 = sc->plane_color[3 * j], c1 = sc->plane_color[3 * j + 1], c2 = sc->plane_color[3 * j + 2];
        const double m01 = (c0 < c1) ? c1 : c0;
        const float kdf = (float)((m01 < c2) ? c2 : m01);  // SettingData.h:14-16
        kd[i] = (double)kdf;
        for (int k = 0; k < 3; ++k) {
            em[3 * i + k] = sc->plane_emission[3 * j + k];
            col[3 * i + k] = sc->plane_color[3 * j + k];
        }
        if (f32_path && use_bvh) {
            // hierarchy path: the plane's records follow the spheres' in the global-memory arrays
            const double th = (double)kdf * 4294967296.0;
            const uint32_t thr = (th >= 4294967295.0) ? 0xFFFFFFFFu : (th <= 0.0 ? 0u : (uint32_t)th);
            float4 e = make_float4((float)em[3 * i], (float)em[3 * i + 1], (float)em[3 * i + 2], 0.f);
            memcpy(&e.w, &thr, 4);
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
            if (kdf > 0.f) a = make_float4((float)(c0 / (double)kdf), (float)(c1 / (double)kdf), (float)(c2 / (double)kdf), 0.f);
            const int32_t none = -1;
            memcpy(&a.w, &none, 4);
            r->mat_geo.push_back(make_float4((float)normal.x, (float)normal.y, (float)normal.z, 0.f));
            r->mat_emi.push_back(e);
            r->mat_alb.push_back(a);
        }
        if (f32_path) {
            RtmSceneF32Host& s = r->f32;
            if (!use_bvh) {   // brute force: records in constant -> shared memory, behind the spheres'
                const double th = (double)kdf * 4294967296.0;
                s.kd_thr[i] = (th >= 4294967295.0) ? 0xFFFFFFFFu : (th <= 0.0 ? 0u : (uint32_t)th);
                s.kd[i] = kdf;
                s.cx[i] = (float)normal.x; s.cy[i] = (float)normal.y; s.cz[i] = (float)normal.z;   // geo of a plane: its normal, 1/r = 0
                s.inv_r[i] = 0.f;
                s.ex[i] = (float)em[3 * i]; s.ey[i] = (float)em[3 * i + 1]; s.ez[i] = (float)em[3 * i + 2];
                s.kx[i] = s.ky[i] = s.kz[i] = 0.f;
                if (kdf > 0.f) { s.kx[i] = (float)(c0 / (double)kdf); s.ky[i] = (float)(c1 / (double)kdf); s.kz[i] = (float)(c2 / (double)kdf); }
            }
            float* q = reinterpret_cast<float*>(&s.plane_pairs[j >> 1]) + (j & 1);   // lane of the pair, stride 2 floats
            const v3 prel{position.x - r->refd[0], position.y - r->refd[1], position.z - r->refd[2]};
            const double vals[13] = {normal.x, normal.y, normal.z, vdot(normal, prel), -prel.x, -prel.y, -prel.z,
                                     right.x / rr, right.y / rr, right.z / rr, upp.x / uu, upp.y / uu, upp.z / uu};
            for (int k = 0; k < 13; ++k) q[2 * k] = (float)vals[k];
        }
    }

    if (opt.estimator == RTM_ESTIMATOR_NEE) {
        // light list: the first RTM_MAX_LIGHTS spheres in index order with a positive emission component and
        // radius <= 1000.  Without a listed light the estimator IS the reference's.
        RtmSceneF32Host& s = r->f32;
        for (int i = 0; i < n && s.n_lights < RTM_MAX_LIGHTS; ++i) {
            if (!(em[3 * i] > 0.0 || em[3 * i + 1] > 0.0 || em[3 * i + 2] > 0.0) || !(sc->radius[i] <= 1000.f)) continue;
            const int k = s.n_lights++;
            s.lgeo[k] = make_float4(s.cx[i], s.cy[i], s.cz[i], sc->radius[i]);
            s.lemi[k] = make_float4(s.ex[i], s.ey[i], s.ez[i], 0.f);
            memcpy(&s.lemi[k].w, &i, 4);
            const int one = 1;
            memcpy(&s.lit[i], &one, 4);
        }
        r->nee = s.n_lights > 0;
    }

    if (f32_path && !use_bvh) {
        // Which megakernel: estimate the roulette's survival probability at the primary vertices as the mean kd of the
        // reflecting spheres weighted by their apparent size from the camera (r^2 / d^2, 1 when the camera is inside;
        // occlusion ignored).  Below 0.6 (simpleSetting1/2: 0.25 surfaces fill the image) the bounce block would run
        // at a fraction of the lanes -> roulette rounds.  Either kernel renders the same image, bit for bit.
        double wsum = 0.0, ksum = 0.0;
        for (int i = 0; i < n; ++i) {
            if (!(kd[i] > 0.0)) continue;
            const double dx = cx[i] - r->refd[0], dy = cy[i] - r->refd[1], dz = cz[i] - r->refd[2];
            const double d2 = dx * dx + dy * dy + dz * dz, rr = (double)sc->radius[i] * (double)sc->radius[i];
            const double wgt = (d2 > rr) ? rr / d2 : 1.0;
            wsum += wgt;
            ksum += wgt * kd[i];
        }
        r->retry = wsum > 0.0 && ksum / wsum < 0.6;
        if (opt.debug_retry) r->retry = (opt.debug_retry == 2);   // test hook: both kernels, same scene
    }

    if (f32_path && (m & 1)) reinterpret_cast<float*>(&r->f32.plane_pairs[m >> 1])[2 * 3 + 1] = -1.0f;   // dummy lane: n = 0, d0 = -1

    if (f32_path) {
        // ACNE (kernels_radiance.cu): the reference re-hits the sphere a ray leaves when
        // t^2 |eps| >= 2 tau r cos(theta).  With t <= D (extent of the finite part of the scene seen from
        // the camera), |eps| <= 1.8e-7, tau = 1e-5 this concerns directions with cos(theta) <= 9e-3 D^2 / r;
        // cosine-weighted, that is a fraction (9e-3 D^2 / r)^2 of the bounces.  Modelled when any
        // diffuse sphere reaches 1e-4 (the r = 1e4 walls of the Cornell scenes: 6e-7).
        double D = 0.0;
        for (int i = 0; i < n; ++i) {
            if (sc->radius[i] >= 1000.f) continue;
            const double dx = cx[i] - r->refd[0], dy = cy[i] - r->refd[1], dz = cz[i] - r->refd[2];
            D = std::max(D, std::sqrt(dx * dx + dy * dy + dz * dz) + (double)sc->radius[i]);
        }
        if (D == 0.0) D = 30.0;
        for (int i = 0; i < n; ++i)
            if (kd[i] > 0.0 && 9e-3 * D * D / (double)sc->radius[i] > 1e-2) r->acne = true;
    }

    if (f32_path) {
        // Per-sample clamp (Renderer.cpp:240-241): clamp(L / spp, 0, 1).  L >= 0 when every emission is >= 0.  colorKD has
        // components <= 1 (+1 ulp), so a path's throughput never exceeds ~1 and L <= (number of emitters a path can
        // collect) x (largest emission): 1 when every emitter has kd == 0 (the path ends on it), else max_depth.
        // When that bound / spp stays below 1 the clamp is the identity and the megakernel drops it (and with it the
        // per-sample accumulator: one running sum per sub-pixel).
        double e_max = 0.0;
        bool nonneg = true, lights_end = true;
        for (size_t i = 0; i < nm; ++i) {
            bool emits = false;
            for (int k = 0; k < 3; ++k) {
                const double e = em[3 * i + k];
                if (!(e >= 0.0) || !std::isfinite(e)) nonneg = false;
                if (e > 0.0) emits = true;
                e_max = std::max(e_max, e);
            }
            if (emits && kd[i] > 0.0) lights_end = false;
        }
        const double bound = e_max * (lights_end ? 1.0 : (double)opt.max_depth) * 1.001;
        r->clamp = opt.debug_clamp == 1 || !nonneg || !(bound / ((double)r->ss * r->ss * r->S) <= 1.0);
    }

    if (use_bvh) {
        if (!rtm_bvh_build(n, sc->position, sc->radius, r->refd, r->bvh_host, (double)opt.debug_bvh_ctrav, opt.debug_bvh_leaf_max, (opt.debug_bvh_walk & 4) ? 1 : 0 /* treelet numbering */)) {
            const int d = r->bvh_host.depth;
            delete r;
            return fail(RTM_E_UNSUPPORTED, "sphere hierarchy too deep (%d levels)", d);
        }
        if (f32_path && r->bvh_host.n_big_overflow > 0) {
            const int k = r->bvh_host.n_big_overflow;
            delete r;
            return fail(RTM_E_UNSUPPORTED, "%d sphere(s) of radius >= %g beyond the %d the FP32 hierarchy path tests for every ray: "
                        "use RTM_ACCEL_BRUTE or RTM_PRECISION_F64", k, RTM_BVH_BIG_RADIUS, RTM_BVH_MAX_BIG);
        }
        if (f32_path && !(opt.debug_bvh_walk & 2)) {
            // compact half-precision nodes for the FP32 walk, unless rounding the boxes to half precision would inflate
            // the typical leaf box by more than a quarter (features tiny against the scene's extent)
            rtm_bvh_compact16(r->bvh_host, r->bvh16);
            if (r->bvh16.ok && r->bvh16.inflation > 0.25) r->bvh16.ok = false;
            // ... and, on request (A/B knob), their 4-wide collapse: half the dependent fetches per ray for the same box tests,
            // measured no faster than the two-child walk (profiles/r02); needs a walk's worst-case stack (3 entries per
            // level + the sentinel) to fit the kernels' private array
            if (r->bvh16.ok && (opt.debug_bvh_walk & 8)) {
                rtm_bvh_compact16x4(r->bvh_host, r->bvh16, r->bvh16x4);
                if (r->bvh16x4.ok && 3 * r->bvh16x4.depth + 1 > RTM_BVH_STACK_DEV) r->bvh16x4.ok = false;
            }
        }
        if (f32_path) {
            // alb.w of a sphere: its slot in the big list (0..7) or 8 + its leaf slot (kernels_common.cuh: bvh_big_slot / bvh_leaf_slot)
            for (int j = 0; j < r->bvh_host.n_big; ++j) memcpy(&r->mat_alb[r->bvh_host.big_id[j]].w, &j, 4);
            for (size_t sl = 0; sl < r->bvh_host.leaf_id.size(); ++sl) {
                if (r->bvh_host.leaf_id[sl] < 0) continue;   // pad slot (bvh_build.cpp: even-aligned leaves)
                const int32_t code = 8 + (int32_t)sl;
                memcpy(&r->mat_alb[r->bvh_host.leaf_id[sl]].w, &code, 4);
            }
        }
    }

    if (f32_path) {
        // the packed pair test: the first 8 spheres (brute force) or the big list (hierarchy)
        RtmSceneF32Host& s = r->f32;
        const int n_packed = use_bvh ? r->bvh_host.n_big : (n < 2 * RTM_MAX_PAIRS ? n : 2 * RTM_MAX_PAIRS);
        for (int j = 0; j < RTM_MAX_PAIRS; ++j) {
            float* f = reinterpret_cast<float*>(&s.pairs[j]);
            for (int h = 0; h < 2; ++h) {
                const int slot = 2 * j + h;
                const bool real = slot < n_packed;
                const int i = real ? (use_bvh ? r->bvh_host.big_id[slot] : slot) : 0;
                float fcx = 0.f, fcy = 0.f, fcz = 0.f, K = 1e30f;
                if (real) {
                    fcx = (float)(cx[i] - r->refd[0]); fcy = (float)(cy[i] - r->refd[1]); fcz = (float)(cz[i] - r->refd[2]);
                    // use the rounded centre consistently: K belongs to the sphere the FP32 kernel sees
                    const double fx = fcx, fy = fcy, fz = fcz, rad = (double)sc->radius[i];
                    K = (float)(fx * fx + fy * fy + fz * fz - rad * rad);
                    if (!use_bvh) s.K[i] = K;
                }
                // dummy: centre at the reference point, -c' = -1e30 -> negative discriminant -> NaN -> never a hit
                f[0 + h] = fcx;  f[2 + h] = fcy;  f[4 + h] = fcz;
                f[6 + h] = 2.0f * fcx; f[8 + h] = 2.0f * fcy; f[10 + h] = 2.0f * fcz;
                f[12 + h] = -K;
                f[14 + h] = 0.f;
            }
        }
        if (!use_bvh)
            for (int i = 2 * RTM_MAX_PAIRS; i < n; ++i) {
                const double fx = s.cx[i], fy = s.cy[i], fz = s.cz[i], rad = (double)sc->radius[i];
                s.K[i] = (float)(fx * fx + fy * fy + fz * fz - rad * rad);
            }
    }

    // devices
    const int S0 = opt.sample_begin, SN = opt.sample_end - opt.sample_begin;
    r->gpus.resize(opt.n_gpus);
    int rc = RTM_OK;
    auto body = [&]() -> int {
        for (int g = 0; g < opt.n_gpus; ++g) {
            GpuCtx& c = r->gpus[g];
            c.device = opt.device + g;
            c.s_begin = S0 + (int)((int64_t)SN * g / opt.n_gpus);
            c.s_end = S0 + (int)((int64_t)SN * (g + 1) / opt.n_gpus);
            CU(cudaSetDevice(c.device));
            cudaDeviceProp prop;
            CU(cudaGetDeviceProperties(&prop, c.device));
            c.sm_count = prop.multiProcessorCount;
            CU(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
            CU(cudaEventCreate(&c.ev0));
            CU(cudaEventCreate(&c.ev1));
            for (auto& e : c.ek) CU(cudaEventCreate(&e));
            CU(cudaMalloc(&c.d_scene64, h64.size() * sizeof(double)));
            CU(cudaMemcpy(c.d_scene64, h64.data(), h64.size() * sizeof(double), cudaMemcpyHostToDevice));
            c.s64.n = n;
            c.s64.cx = c.d_scene64; c.s64.cy = c.s64.cx + n; c.s64.cz = c.s64.cy + n; c.s64.r2 = c.s64.cz + n;
            c.s64.kd = c.s64.r2 + n; c.s64.emission = c.s64.kd + nm; c.s64.color = c.s64.emission + 3 * nm;
            c.s64.n_planes = m;
            c.s64.planes = c.s64.color + 3 * nm;
            if (use_bvh) {
                // one allocation, float4 units: nodes | leaf_geo | geo | emi | alb | leaf_id (int32, padded)
                const RtmBvhHost& B = r->bvh_host;
                const size_t n_nodes4 = B.nodes.size(), n_slots = B.leaf_geo.size(), n_mat = r->mat_geo.size();
                const size_t n_16 = r->bvh16.ok ? r->bvh16.nodes.size() / 4 : 0;     // float4 units: 2 per node
                const size_t n_16x4 = r->bvh16x4.ok ? r->bvh16x4.nodes.size() / 4 : 0; // float4 units: 4 per node
                const size_t total4 = n_nodes4 + n_16x4 + n_16 + n_slots + 3 * n_mat + (n_slots + 3) / 4 + 1;
                CU(cudaMalloc(&c.d_bvh, total4 * sizeof(float4)));
                float4* d = c.d_bvh;
                cudaError_t put_err = cudaSuccess;
                auto put = [&](const void* src, size_t count4, size_t bytes) -> float4* {
                    float4* at = d;
                    if (bytes && put_err == cudaSuccess) put_err = cudaMemcpy(at, src, bytes, cudaMemcpyHostToDevice);
                    d += count4;
                    return at;
                };
                c.bvh.nodes = put(B.nodes.data(), n_nodes4, n_nodes4 * sizeof(float4));
                // right behind the 64-byte float nodes: 64-byte aligned (both halves of a 4-wide node in one line), then
                // 32-byte aligned, as the 256-bit loads need
                c.bvh.nodes16x4 = n_16x4 ? (const uint32_t*)put(r->bvh16x4.nodes.data(), n_16x4, n_16x4 * sizeof(float4)) : nullptr;
                c.bvh.n_nodes4 = (int32_t)(n_16x4 / 4);
                c.bvh.depth4 = r->bvh16x4.depth;
                c.bvh.nodes16 = n_16 ? (const uint32_t*)put(r->bvh16.nodes.data(), n_16, n_16 * sizeof(float4)) : nullptr;
                for (int k = 0; k < 3; ++k) c.bvh.ref16[k] = r->bvh16.ref[k];
                c.bvh.inv_scale16 = r->bvh16.inv_scale;
                c.bvh.leaf_geo = put(B.leaf_geo.data(), n_slots, n_slots * sizeof(float4));
                c.mat.geo = put(r->mat_geo.data(), n_mat, n_mat * sizeof(float4));
                c.mat.emi = put(r->mat_emi.data(), n_mat, n_mat * sizeof(float4));
                c.mat.alb = put(r->mat_alb.data(), n_mat, n_mat * sizeof(float4));
                c.bvh.leaf_id = (const int32_t*)put(B.leaf_id.data(), (n_slots + 3) / 4, n_slots * sizeof(int32_t));
                CU(put_err);
                c.bvh.root = B.root;
                c.bvh.n_big = B.n_big;
                for (int j = 0; j < 8; ++j) c.bvh.big_id[j] = B.big_id[j];
                c.bvh.depth = B.depth;
                c.bvh.n_nodes = (int32_t)(n_nodes4 / 4);
                c.bvh.n_slots = (int32_t)n_slots;
                c.s64.use_bvh = 1;
                c.s64.bvh = c.bvh;
                for (int k = 0; k < 3; ++k) c.s64.ref[k] = r->refd[k];
            }
            if (f32_path) {
                CU(cudaMalloc(&c.d_f32, sizeof(RtmSceneF32Host)));
                CU(cudaMemcpy(c.d_f32, &r->f32, sizeof(RtmSceneF32Host), cudaMemcpyHostToDevice));
                memcpy(c.hot.pairs, r->f32.pairs, sizeof c.hot.pairs);
                memcpy(c.hot.plane_pairs, r->f32.plane_pairs, sizeof c.hot.plane_pairs);
                c.hot.n = r->f32.n;
                c.hot.n_planes = r->f32.n_planes;
                c.hot.n_lights = r->f32.n_lights;
                c.hot.full = c.d_f32;
            }
            if (r->acne) CU(cudaMalloc(&c.cps0, r->n_sub * sizeof(float)));
            if (opt.variant == RTM_VARIANT_WAVEFRONT) {
                // path slots: 8 resident warps' worth per SM partition x 4, at most one per sub-pixel
                uint64_t slots = (uint64_t)c.sm_count * 2048ull * 4ull;
                if (slots > r->n_sub) slots = r->n_sub;
                slots = (slots + 127) / 128 * 128;
                const size_t per_slot4 = 8;   // a0..a6, ray
                const size_t total4 = slots * per_slot4 + slots / 2 /*hit: float2*/ + 2 * (slots / 4) /*queues*/ + 4;
                CU(cudaMalloc(&c.d_wave, total4 * sizeof(float4)));
                float4* d = c.d_wave;
                RtmWaveDev& w = c.wv;
                w.a0 = d; d += slots; w.a1 = d; d += slots; w.a2 = d; d += slots; w.a3 = d; d += slots;
                w.a4 = d; d += slots; w.a5 = d; d += slots; w.a6 = d; d += slots; w.ray = d; d += slots;
                w.hit = (float2*)d; d += slots / 2;
                w.queue[0] = (int32_t*)d; d += slots / 4;
                w.queue[1] = (int32_t*)d; d += slots / 4;
                w.ctrl = (unsigned*)d;
                w.n_slots = (int32_t)slots;
                CU(cudaMallocHost(&c.h_live, sizeof(unsigned)));
            }
            if (opt.compat == RTM_COMPAT_ABC) CU(cudaMalloc(&c.records, r->n_sub * sizeof(float4)));
            CU(cudaMalloc(&c.accum, r->n_pix * sizeof(float4)));
            CU(cudaMemset(c.accum, 0, r->n_pix * sizeof(float4)));
            CU(cudaMalloc(&c.counters, sizeof(RtmCounters)));
            CU(cudaMemset(c.counters, 0, sizeof(RtmCounters)));
        }
        if (opt.n_gpus > 1) {
            int lr = load_nccl();
            if (lr) return lr;
            std::vector<int> devs(opt.n_gpus);
            std::vector<ncclComm_t> comms(opt.n_gpus);
            for (int g = 0; g < opt.n_gpus; ++g) devs[g] = opt.device + g;
            NC(g_nccl.CommInitAll(comms.data(), opt.n_gpus, devs.data()));
            for (int g = 0; g < opt.n_gpus; ++g) r->gpus[g].comm = comms[g];
        }
        return RTM_OK;
    };
    rc = body();
    if (rc != RTM_OK) {
        const std::string keep = g_err;
        renderer_destroy(r);
        g_err = keep;
        return rc;
    }
    *out = r;
    return RTM_OK;
}

int renderer_render_async(rtm_renderer* r, void* stream) {
    if (!r) return fail(RTM_E_INVALID, "null renderer");
    if (r->gpus.size() != 1) return fail(RTM_E_UNSUPPORTED, "renderer_render_async drives one GPU per handle");
    int launches = 0;
    int rc = enqueue_pass(r, r->gpus[0], (cudaStream_t)stream, false, &launches);
    r->last_launches = launches;
    r->hits_valid = (r->opt.compat == RTM_COMPAT_HEAD);
    return rc;
}

int renderer_read_stats(rtm_renderer* r, rtm_stats* st) {
    if (!r || !st) return fail(RTM_E_INVALID, "null argument");
    return fill_stats(r, st);
}

int renderer_read_kernel_ms(rtm_renderer* r, double* ms3) {
    if (!r || !ms3) return fail(RTM_E_INVALID, "null argument");
    GpuCtx& c = r->gpus[0];
    if (!c.ek_valid) return fail(RTM_E_UNSUPPORTED, "per-kernel timing exists for the FP32 ABC path after a pass");
    CU(cudaSetDevice(c.device));
    CU(cudaEventSynchronize(c.ek[3]));
    for (int k = 0; k < 3; ++k) {
        float ms = 0;
        CU(cudaEventElapsedTime(&ms, c.ek[k], c.ek[k + 1]));
        ms3[k] = ms;
    }
    return RTM_OK;
}

void* renderer_device_accum(rtm_renderer* r) { return r ? (void*)r->gpus[0].accum : nullptr; }
void* renderer_device_hits(rtm_renderer* r) { return r ? (void*)r->gpus[0].hits : nullptr; }

int renderer_set_accum_buffer(rtm_renderer* r, void* p) {
    if (!r || !p) return fail(RTM_E_INVALID, "null argument");
    if (r->gpus.size() != 1) return fail(RTM_E_UNSUPPORTED, "one GPU per handle");
    GpuCtx& c = r->gpus[0];
    CU(cudaSetDevice(c.device));
    if (c.accum && c.accum_owned) CU(cudaFree(c.accum));
    c.accum = (float4*)p;
    c.accum_owned = false;
    return RTM_OK;
}

// ---- peer-memory gather (one process per GPU): the root exports G framebuffer slots, every rank points its
// accumulate kernel at its slot (coalesced 16 B stores over NVLink), the root sums the slots.
int renderer_peer_export(rtm_renderer* r, int32_t n_slots, void* handle64, void** device_ptr) {
    if (!r || !handle64 || !device_ptr || n_slots < 1) return fail(RTM_E_INVALID, "bad argument");
    if (r->gpus.size() != 1) return fail(RTM_E_UNSUPPORTED, "one GPU per handle");
    GpuCtx& c = r->gpus[0];
    CU(cudaSetDevice(c.device));
    if (c.peer_stage) { CU(cudaFree(c.peer_stage)); c.peer_stage = nullptr; }
    CU(cudaMalloc(&c.peer_stage, (size_t)n_slots * r->n_pix * sizeof(float4)));
    CU(cudaMemset(c.peer_stage, 0, (size_t)n_slots * r->n_pix * sizeof(float4)));
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, c.peer_stage));
    memcpy(handle64, &h, 64);
    *device_ptr = c.peer_stage;
    return RTM_OK;
}

int renderer_peer_open(rtm_renderer* r, const void* handle64, void** device_ptr) {
    if (!r || !handle64 || !device_ptr) return fail(RTM_E_INVALID, "null argument");
    if (r->gpus.size() != 1) return fail(RTM_E_UNSUPPORTED, "one GPU per handle");
    GpuCtx& c = r->gpus[0];
    CU(cudaSetDevice(c.device));
    if (c.peer_opened) { CU(cudaIpcCloseMemHandle(c.peer_opened)); c.peer_opened = nullptr; }
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    CU(cudaIpcOpenMemHandle(&c.peer_opened, h, cudaIpcMemLazyEnablePeerAccess));
    *device_ptr = c.peer_opened;
    return RTM_OK;
}

int renderer_sum_partials(rtm_renderer* r, const void* partials, int32_t n_slots, void* out_float4, void* stream) {
    if (!r || !partials || !out_float4 || n_slots < 1) return fail(RTM_E_INVALID, "bad argument");
    GpuCtx& c = r->gpus[0];
    CU(cudaSetDevice(c.device));
    rtm_launch_sum_partials((const float4*)partials, (float4*)out_float4, r->n_pix, n_slots, (cudaStream_t)stream);
    CU(cudaGetLastError());
    return RTM_OK;
}

int renderer_render(rtm_renderer* r, float* rgb_out, int32_t* hit_out, rtm_stats* stats) {
    if (!r) return fail(RTM_E_INVALID, "null renderer");
    if (hit_out && (r->opt.sub_end - r->opt.sub_begin) != r->ss * r->ss)
        return fail(RTM_E_UNSUPPORTED, "the primary-hit map needs a handle that covers the full sub-pixel grid");
    const auto t0 = std::chrono::steady_clock::now();
    const int G = (int)r->gpus.size();
    int launches = 0;
    for (int g = 0; g < G; ++g) {
        GpuCtx& c = r->gpus[g];
        CU(cudaSetDevice(c.device));
        CU(cudaEventRecord(c.ev0, c.stream));
        // the host-buffer entry point takes its inputs from host memory on every call: the scene block travels again
        if (c.d_f32) CU(cudaMemcpyAsync(c.d_f32, &r->f32, sizeof(RtmSceneF32Host), cudaMemcpyHostToDevice, c.stream));
        int rc = enqueue_pass(r, c, c.stream, hit_out != nullptr && g == 0, &launches);
        if (rc) return rc;
    }
    if (G > 1) {
        // one sum-reduce of the partial accumulation buffers to the first GPU, enqueued
        // behind each GPU's last kernel on the same stream (no host sync in between)
        NC(g_nccl.GroupStart());
        for (int g = 0; g < G; ++g) {
            GpuCtx& c = r->gpus[g];
            NC(g_nccl.Reduce(c.accum, c.accum, r->n_pix * 4, ncclFloat32, ncclSum, 0, c.comm, c.stream));
        }
        NC(g_nccl.GroupEnd());
    }
    for (int g = 0; g < G; ++g) {
        GpuCtx& c = r->gpus[g];
        CU(cudaSetDevice(c.device));
        CU(cudaEventRecord(c.ev1, c.stream));
    }
    GpuCtx& c0 = r->gpus[0];
    CU(cudaSetDevice(c0.device));
    if (rgb_out) {
        // repacked on the device, then ONE copy of 12 B/pixel straight into the caller's buffer
        if (!c0.d_rgb) CU(cudaMalloc(&c0.d_rgb, r->n_pix * 3 * sizeof(float)));
        rtm_launch_pack_rgb(c0.accum, c0.d_rgb, r->n_pix, c0.stream);
        CU(cudaGetLastError());
        CU(cudaMemcpyAsync(rgb_out, c0.d_rgb, r->n_pix * 3 * sizeof(float), cudaMemcpyDeviceToHost, c0.stream));
    }
    if (hit_out) CU(cudaMemcpyAsync(hit_out, c0.hits, r->n_sub * sizeof(int32_t), cudaMemcpyDeviceToHost, c0.stream));
    double dev_ms = 0;
    for (int g = 0; g < G; ++g) {
        GpuCtx& c = r->gpus[g];
        CU(cudaSetDevice(c.device));
        CU(cudaStreamSynchronize(c.stream));
        float ms = 0;
        CU(cudaEventElapsedTime(&ms, c.ev0, c.ev1));
        if (ms > dev_ms) dev_ms = ms;
    }
    r->last_launches = launches;
    r->hits_valid = (hit_out != nullptr) || r->opt.compat == RTM_COMPAT_HEAD;
    if (stats) {
        int rc = fill_stats(r, stats);
        if (rc) return rc;
        stats->device_ms = dev_ms;
        stats->e2e_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    return RTM_OK;
}

int renderer_read_rgb(rtm_renderer* r, float* rgb_out, void* stream) {
    if (!r || !rgb_out) return fail(RTM_E_INVALID, "null argument");
    GpuCtx& c = r->gpus[0];
    CU(cudaSetDevice(c.device));
    // `stream` as in renderer_render_async: the caller's stream, NULL = the legacy default stream — the copy is ordered behind
    // whatever the caller enqueued there (its render_async pass, its reduce).  (The handle's own stream is non-blocking: using
    // it for NULL would read the buffer before a reduce on the default stream has finished.)
    cudaStream_t st = (cudaStream_t)stream;
    if (!c.d_rgb) CU(cudaMalloc(&c.d_rgb, r->n_pix * 3 * sizeof(float)));
    rtm_launch_pack_rgb(c.accum, c.d_rgb, r->n_pix, st);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(rgb_out, c.d_rgb, r->n_pix * 3 * sizeof(float), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return RTM_OK;
}

int renderer_read_image_f64(rtm_renderer* r, double* out) {
    if (!r || !out) return fail(RTM_E_INVALID, "null argument");
    if (r->opt.compat != RTM_COMPAT_HEAD) return fail(RTM_E_UNSUPPORTED, "the double image exists in COMPAT_HEAD only");
    GpuCtx& c = r->gpus[0];
    if (!c.image64) return fail(RTM_E_INVALID, "render first");
    CU(cudaSetDevice(c.device));
    CU(cudaMemcpy(out, c.image64, r->n_pix * 3 * sizeof(double), cudaMemcpyDeviceToHost));
    return RTM_OK;
}

int renderer_read_u8(rtm_renderer* r, unsigned char* out) {
    if (!r || !out) return fail(RTM_E_INVALID, "null argument");
    GpuCtx& c = r->gpus[0];
    CU(cudaSetDevice(c.device));
    if (!c.u8) CU(cudaMalloc(&c.u8, r->n_pix * 3));
    if (r->opt.compat == RTM_COMPAT_HEAD) {
        if (!c.image64) return fail(RTM_E_INVALID, "render first");
        rtm_launch_to_u8_f64(c.image64, c.u8, r->n_pix, c.stream);   // the reference converts its DOUBLE image
    } else {
        rtm_launch_to_u8(c.accum, c.u8, r->n_pix, c.stream);
    }
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(out, c.u8, r->n_pix * 3, cudaMemcpyDeviceToHost, c.stream));
    CU(cudaStreamSynchronize(c.stream));
    return RTM_OK;
}

// ---- test hooks ---------------------------------------------------------------

int rtm_debug_philox(int32_t device, uint32_t n, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t* out) {
    if (!out || n == 0) return fail(RTM_E_INVALID, "bad argument");
    CU(cudaSetDevice(device));
    uint32_t* d = nullptr;
    CU(cudaMalloc(&d, (size_t)n * 16));
    rtm_launch_philox(n, c1, c2, c3, k0, k1, d, nullptr);
    cudaError_t e = cudaMemcpy(out, d, (size_t)n * 16, cudaMemcpyDeviceToHost);
    cudaFree(d);
    CU(e);
    return RTM_OK;
}

int rtm_debug_closest_hit(rtm_renderer* r, int32_t mode, uint32_t n, const double* org, const double* dir,
                          const int32_t* origin_id, int32_t* hit_out, double* t_out) {
    if (!r || !org || !dir || !hit_out || !t_out || n == 0) return fail(RTM_E_INVALID, "bad argument");
    GpuCtx& c = r->gpus[0];
    CU(cudaSetDevice(c.device));
    if (mode == 1 && !r->use_bvh && r->n > RTM_MAX_CONST_SPHERES) return fail(RTM_E_UNSUPPORTED, "too many spheres for the FP32 test");
    double *d_org = nullptr, *d_dir = nullptr, *d_t = nullptr;
    int32_t *d_oid = nullptr, *d_hit = nullptr;
    int rc = RTM_OK;
    auto body = [&]() -> int {
        CU(cudaMalloc(&d_org, (size_t)n * 24));
        CU(cudaMalloc(&d_dir, (size_t)n * 24));
        CU(cudaMalloc(&d_t, (size_t)n * 8));
        CU(cudaMalloc(&d_hit, (size_t)n * 4));
        CU(cudaMemcpy(d_org, org, (size_t)n * 24, cudaMemcpyHostToDevice));
        CU(cudaMemcpy(d_dir, dir, (size_t)n * 24, cudaMemcpyHostToDevice));
        if (origin_id) {
            CU(cudaMalloc(&d_oid, (size_t)n * 4));
            CU(cudaMemcpy(d_oid, origin_id, (size_t)n * 4, cudaMemcpyHostToDevice));
        }
        if (mode == 0) {
            rtm_launch_closest_hit_f64(c.s64, n, d_org, d_dir, d_hit, d_t, c.stream);
        } else {
            if (r->f32.n != r->n) return fail(RTM_E_UNSUPPORTED, "handle was not created for the FP32 path");
            if (!c.d_f32) return fail(RTM_E_UNSUPPORTED, "handle was not created for the FP32 path");
            if (r->use_bvh) rtm_launch_closest_hit_bvh_f32(c.hot, c.bvh, c.mat, r->n, r->n_planes, r->refd, n, d_org, d_dir, d_oid, d_hit, d_t, c.stream);
            else rtm_launch_closest_hit_f32(c.hot, r->n, r->n_planes, r->refd, n, d_org, d_dir, d_oid, d_hit, d_t, c.stream);
        }
        CU(cudaGetLastError());
        CU(cudaStreamSynchronize(c.stream));
        CU(cudaMemcpy(hit_out, d_hit, (size_t)n * 4, cudaMemcpyDeviceToHost));
        CU(cudaMemcpy(t_out, d_t, (size_t)n * 8, cudaMemcpyDeviceToHost));
        return RTM_OK;
    };
    rc = body();
    cudaFree(d_org); cudaFree(d_dir); cudaFree(d_t); cudaFree(d_hit); cudaFree(d_oid);
    return rc;
}

int rtm_debug_acne_eps(int32_t device, uint32_t n, const uint32_t* bits, float* out) {
    if (!bits || !out || n == 0) return fail(RTM_E_INVALID, "bad argument");
    CU(cudaSetDevice(device));
    uint32_t* d_bits = nullptr;
    float* d_out = nullptr;
    CU(cudaMalloc(&d_bits, (size_t)n * 4));
    cudaError_t e = cudaMalloc(&d_out, (size_t)n * 4);
    if (e == cudaSuccess) e = cudaMemcpy(d_bits, bits, (size_t)n * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        rtm_launch_acne_eps(n, d_bits, d_out, nullptr);
        e = cudaMemcpy(out, d_out, (size_t)n * 4, cudaMemcpyDeviceToHost);
    }
    cudaFree(d_bits); cudaFree(d_out);
    CU(e);
    return RTM_OK;
}

int rtm_debug_bvh(const rtm_scene_desc* sc, int32_t* info, float* nodes_out, float* leaf_geo_out, int32_t* leaf_id_out) {
    if (!sc || !info) return fail(RTM_E_INVALID, "null argument");
    RtmBvhHost B;
    const bool ok = rtm_bvh_build(sc->n_spheres, sc->position, sc->radius, sc->cam_origin, B);
    info[0] = (int32_t)(B.nodes.size() / 4);
    info[1] = (int32_t)B.leaf_id.size();
    info[2] = B.root;
    info[3] = B.depth;
    info[4] = B.n_big;
    for (int j = 0; j < 8; ++j) info[5 + j] = B.big_id[j];
    if (nodes_out) memcpy(nodes_out, B.nodes.data(), B.nodes.size() * sizeof(float4));
    if (leaf_geo_out) memcpy(leaf_geo_out, B.leaf_geo.data(), B.leaf_geo.size() * sizeof(float4));
    if (leaf_id_out) memcpy(leaf_id_out, B.leaf_id.data(), B.leaf_id.size() * sizeof(int32_t));
    return ok ? RTM_OK : fail(RTM_E_UNSUPPORTED, "sphere hierarchy too deep (%d levels)", B.depth);
}

int rtm_debug_bvh16(const rtm_scene_desc* sc, int32_t* n_nodes_out, float* meta6, uint32_t* nodes16_out) {
    if (!sc || !n_nodes_out || !meta6) return fail(RTM_E_INVALID, "null argument");
    RtmBvhHost B;
    if (!rtm_bvh_build(sc->n_spheres, sc->position, sc->radius, sc->cam_origin, B))
        return fail(RTM_E_UNSUPPORTED, "sphere hierarchy too deep (%d levels)", B.depth);
    RtmBvh16Host H;
    rtm_bvh_compact16(B, H);
    *n_nodes_out = (int32_t)(H.nodes.size() / 8);
    meta6[0] = H.ref[0]; meta6[1] = H.ref[1]; meta6[2] = H.ref[2];
    meta6[3] = H.inv_scale; meta6[4] = (float)H.inflation; meta6[5] = H.ok ? 1.f : 0.f;
    if (nodes16_out && H.ok) memcpy(nodes16_out, H.nodes.data(), H.nodes.size() * sizeof(uint32_t));
    return RTM_OK;
}

int rtm_debug_bvh16x4(const rtm_scene_desc* sc, int32_t* info2, uint32_t* nodes_out) {
    if (!sc || !info2) return fail(RTM_E_INVALID, "null argument");
    RtmBvhHost B;
    if (!rtm_bvh_build(sc->n_spheres, sc->position, sc->radius, sc->cam_origin, B))
        return fail(RTM_E_UNSUPPORTED, "hierarchy deeper than %d levels", RTM_BVH_STACK);
    RtmBvh16Host H;
    rtm_bvh_compact16(B, H);
    RtmBvh16x4Host Q;
    rtm_bvh_compact16x4(B, H, Q);
    info2[0] = Q.ok ? (int32_t)(Q.nodes.size() / 16) : 0;
    info2[1] = Q.depth;
    if (nodes_out && Q.ok) memcpy(nodes_out, Q.nodes.data(), Q.nodes.size() * sizeof(uint32_t));
    return RTM_OK;
}

int rtm_debug_bvh4(const rtm_scene_desc* sc, int32_t* info4, float* nodes4_out) {
    if (!sc || !info4) return fail(RTM_E_INVALID, "null argument");
    RtmBvhHost B;
    if (!rtm_bvh_build(sc->n_spheres, sc->position, sc->radius, sc->cam_origin, B))
        return fail(RTM_E_UNSUPPORTED, "sphere hierarchy too deep (%d levels)", B.depth);
    RtmBvh4Host W;
    rtm_bvh_collapse4(B, W);
    info4[0] = (int32_t)(W.nodes.size() / 8);
    info4[1] = W.root;
    info4[2] = W.depth;
    if (nodes4_out) memcpy(nodes4_out, W.nodes.data(), W.nodes.size() * sizeof(float4));
    return RTM_OK;
}

int rtm_debug_fp32_peak(int32_t device, double* tflops_out, double* mhz_out) {
    if (!tflops_out) return fail(RTM_E_INVALID, "null argument");
    CU(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    float* sink = nullptr;
    CU(cudaMalloc(&sink, 4));
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    const int blocks = prop.multiProcessorCount * 8, iters = 4096;
    rtm_launch_fma_peak(sink, blocks, 64, nullptr);  // warm-up
    double best = 0;
    for (int rep = 0; rep < 5; ++rep) {
        CU(cudaEventRecord(e0, nullptr));
        rtm_launch_fma_peak(sink, blocks, iters, nullptr);
        CU(cudaEventRecord(e1, nullptr));
        CU(cudaEventSynchronize(e1));
        float ms = 0;
        CU(cudaEventElapsedTime(&ms, e0, e1));
        const double flops = (double)blocks * 256.0 * iters * 16.0 * 8.0 * 2.0;
        const double tf = flops / (ms * 1e-3) / 1e12;
        if (tf > best) best = tf;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(sink);
    *tflops_out = best;
    if (mhz_out) *mhz_out = best * 1e12 / ((double)prop.multiProcessorCount * 128.0 * 2.0) / 1e6;  // implied SM clock
    return RTM_OK;
}

}  // extern "C"
