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

// One ray/sphere test, FP32.  Semantics of SettingData.cpp:197-223 + the accept
// rule of Renderer.cpp:67 (`t < best`, lowest index wins ties).
__device__ __forceinline__ void sphere_test_f32(int i, int oid, f3 o, f3 d, float oo, float dO, float cx, float cy, float cz,
                                                float m2cx, float m2cy, float m2cz, float K, float& best, int& best_id) {
    const float b = fmaf(d.x, cx, fmaf(d.y, cy, fmaf(d.z, cz, -dO)));           // (c-o).d
    float cp = fmaf(o.x, m2cx, fmaf(o.y, m2cy, fmaf(o.z, m2cz, oo + K)));       // |c-o|^2 - r^2
    cp = (i == oid) ? 0.f : cp;
    const float D = fmaf(b, b, -cp);
    const float sq = fast_sqrt(fmaxf(D, 0.f));
    const float q = b + copysignf(sq, b);
    const float r = cp * fast_rcp(q);
    const bool bpos = b > 0.f;
    const float t2 = bpos ? q : r;                          // far root
    const float t1 = bpos ? r : q;                          // near root
    const float t = (t1 > 0.001f) ? t1 : t2;                // SettingData.cpp:212-223
    const bool ok = (D >= 0.f) && (t2 >= 1e-5f) && (t < best);  // :202, :208-210, Renderer.cpp:67
    best = ok ? t : best;
    best_id = ok ? i : best_id;
}

template <int N>
__device__ __forceinline__ int trace_f32(int n_rt, int oid, f3 o, f3 d, float& t) {
    const float oo = dot(o, o), dO = dot(d, o);
    float best = FLT_MAX;
    int id = -1;
    if (N > 0) {
#pragma unroll
        for (int i = 0; i < N; ++i)
            sphere_test_f32(i, oid, o, d, oo, dO, c_scene.cx[i], c_scene.cy[i], c_scene.cz[i], c_scene.m2cx[i],
                            c_scene.m2cy[i], c_scene.m2cz[i], c_scene.K[i], best, id);
    } else {
#pragma unroll 4
        for (int i = 0; i < n_rt; ++i)
            sphere_test_f32(i, oid, o, d, oo, dO, c_scene.cx[i], c_scene.cy[i], c_scene.cz[i], c_scene.m2cx[i],
                            c_scene.m2cy[i], c_scene.m2cz[i], c_scene.K[i], best, id);
    }
    t = best;
    return id;
}

// Per-sphere data addressed with a per-lane index (hit id) lives in shared memory.
struct SharedScene {
    float cx[RTM_MAX_CONST_SPHERES], cy[RTM_MAX_CONST_SPHERES], cz[RTM_MAX_CONST_SPHERES], inv_r[RTM_MAX_CONST_SPHERES];
    float ex[RTM_MAX_CONST_SPHERES], ey[RTM_MAX_CONST_SPHERES], ez[RTM_MAX_CONST_SPHERES];
    float kx[RTM_MAX_CONST_SPHERES], ky[RTM_MAX_CONST_SPHERES], kz[RTM_MAX_CONST_SPHERES], kd[RTM_MAX_CONST_SPHERES];
};

__device__ __forceinline__ void load_shared_scene(SharedScene& s, int n) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        s.cx[i] = c_scene.cx[i]; s.cy[i] = c_scene.cy[i]; s.cz[i] = c_scene.cz[i]; s.inv_r[i] = c_scene.inv_r[i];
        s.ex[i] = c_scene.ex[i]; s.ey[i] = c_scene.ey[i]; s.ez[i] = c_scene.ez[i];
        s.kx[i] = c_scene.kx[i]; s.ky[i] = c_scene.ky[i]; s.kz[i] = c_scene.kz[i]; s.kd[i] = c_scene.kd[i];
    }
    __syncthreads();
}

// Outward unit normal at hp on sphere id, oriented against the incoming ray.
__device__ __forceinline__ f3 oriented_normal(const SharedScene& s, int id, f3 hp, bool flip) {
    const float ir = flip ? -s.inv_r[id] : s.inv_r[id];
    return mk((hp.x - s.cx[id]) * ir, (hp.y - s.cy[id]) * ir, (hp.z - s.cz[id]) * ir);
}

// Cosine-weighted direction about w (Renderer.cpp:88-107).  The tangent frame is
// the branchless one of Duff et al. 2017 instead of the reference's
// Normalize(Cross(axis, w)): the azimuth is uniform, so the distribution of the
// sampled direction is the same for any orthonormal frame.
__device__ __forceinline__ f3 cosine_dir(f3 w, float u1, float u2) {
    const float sgn = copysignf(1.0f, w.z);
    const float a = -fast_rcp(sgn + w.z);
    const float b = w.x * w.y * a;
    const f3 u = mk(fmaf(sgn * w.x, w.x * a, 1.0f), sgn * b, -sgn * w.x);
    const f3 v = mk(b, fmaf(w.y, w.y * a, sgn), -w.y);
    float sn, cs;
    __sincosf(6.283185307179586f * u1, &sn, &cs);  // r1 = 2*PI*xi1
    const float r2s = fast_sqrt(u2);                // sqrt(r2)
    const float cu = cs * r2s, cv = sn * r2s, cw = fast_sqrt(1.0f - u2);
    return mk(fmaf(u.x, cu, fmaf(v.x, cv, w.x * cw)), fmaf(u.y, cu, fmaf(v.y, cv, w.y * cw)),
              fmaf(u.z, cu, fmaf(v.z, cv, w.z * cw)));
}

__device__ __forceinline__ float clamp01f(float v) { return fminf(fmaxf(v, 0.f), 1.f); }

#define RTM_MEGA_THREADS 128
#define RTM_MEGA_CHUNK 64u  // sub-pixels a warp reserves per global atomic

template <int N>
__global__ void __launch_bounds__(RTM_MEGA_THREADS, 4) radiance_mega_kernel(RtmPassParams p, float4* __restrict__ records,
                                                                            RtmCounters* __restrict__ counters, int n_rt) {
    __shared__ SharedScene sh;
    const int n = (N > 0) ? N : n_rt;
    load_shared_scene(sh, n);

    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    const unsigned long long n_sub = p.n_sub;
    const int s_begin = p.s_begin, s_end = p.s_end;
    const float inv_norm = p.inv_norm;

    unsigned long long w_next = 0, w_end = 0;  // warp-uniform slice of the work cursor
    bool have = false, done = false;

    // unit (sub-pixel) state
    unsigned long long q = 0;
    f3 P = mk(0, 0, 0);      // primary hit point
    int Pid = 0;
    bool Pflip = false;
    double ax = 0, ay = 0, az = 0;
    // path state: sitting at a hit (o, id) with oriented normal w
    f3 o = mk(0, 0, 0), w = mk(0, 0, 1), T = mk(1, 1, 1), L = mk(0, 0, 0);
    int id = 0, s = 0, depth = 0;

    unsigned long long rays_ref = 0, traced = 0;

    for (;;) {
        // ---- refill: lanes without a unit take the next sub-pixels of the warp's chunk
        unsigned need = __ballot_sync(0xffffffffu, !have && !done);
        while (need) {
            if (w_next == w_end) {
                unsigned long long start = 0;
                if (lane == 0) start = atomicAdd(&counters->work_next, (unsigned long long)RTM_MEGA_CHUNK);
                start = __shfl_sync(0xffffffffu, start, 0);
                if (start >= n_sub) {
                    if (!have) done = true;
                    break;
                }
                w_next = start;
                w_end = (start + RTM_MEGA_CHUNK < n_sub) ? start + RTM_MEGA_CHUNK : n_sub;
            }
            const unsigned avail = (unsigned)(w_end - w_next);
            const unsigned rank = __popc(need & lt_mask);
            const bool mine = ((need >> lane) & 1u) && rank < avail;
            if (mine) {
                q = w_next + rank;
                const float4 rec = records[q];
                const unsigned wbits = __float_as_uint(rec.w);
                const unsigned ns = (unsigned)(s_end - s_begin);
                const bool miss = (wbits == RTM_REC_MISS);
                Pid = miss ? 0 : (int)(wbits & RTM_REC_IDMASK);
                if (miss || !(sh.kd[Pid] > 0.f)) {
                    // every sample of this sub-pixel is the same: nothing (miss) or the
                    // emission of a surface that never continues (kd == 0, Renderer.cpp:112)
                    float4 out = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (!miss) {
                        out.x = (float)((double)clamp01f(sh.ex[Pid] * inv_norm) * ns);
                        out.y = (float)((double)clamp01f(sh.ey[Pid] * inv_norm) * ns);
                        out.z = (float)((double)clamp01f(sh.ez[Pid] * inv_norm) * ns);
                    }
                    records[q] = out;
                    rays_ref += ns;
                } else {
                    have = true;
                    P = mk(rec.x, rec.y, rec.z);
                    Pflip = (wbits & RTM_REC_FLIP) != 0u;
                    ax = ay = az = 0.0;
                    s = s_begin;
                    o = P; id = Pid; w = oriented_normal(sh, Pid, P, Pflip);
                    T = mk(1, 1, 1); L = mk(0, 0, 0); depth = 0;
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

        // ---- one hit per iteration
        if (have) {
            L.x = fmaf(T.x, sh.ex[id], L.x); L.y = fmaf(T.y, sh.ey[id], L.y); L.z = fmaf(T.z, sh.ez[id], L.z);
            const uint4 rnd = rtm_philox4x32_10(make_uint4((uint32_t)q, (uint32_t)(q >> 32), (uint32_t)s, (uint32_t)depth),
                                                p.seed_lo, p.seed_hi);
            // Russian roulette, Renderer.cpp:78.  `<` on a 24-bit uniform: kd == 0 never continues.
            bool cont = (rtm_u01_f32(rnd.x) < sh.kd[id]) && (depth + 1 < p.max_depth);
            if (cont) {
                T.x *= sh.kx[id]; T.y *= sh.ky[id]; T.z *= sh.kz[id];   // colorKD, SettingData.h:11-13
                const f3 d = cosine_dir(w, rtm_u01_f32(rnd.y), rtm_u01_f32(rnd.z));
                float t;
                const int id2 = trace_f32<N>(n, id, o, d, t);
                ++traced;
                if (id2 >= 0) {
                    o = mk(fmaf(d.x, t, o.x), fmaf(d.y, t, o.y), fmaf(d.z, t, o.z));  // Renderer.cpp:79
                    const f3 nn = oriented_normal(sh, id2, o, false);
                    const bool flip = !(dot(nn, d) < 0.f);                            // :82-83
                    w = flip ? mk(-nn.x, -nn.y, -nn.z) : nn;
                    id = id2;
                    ++depth;
                } else {
                    cont = false;  // miss: the recursion returns 0 (Renderer.cpp:116)
                }
            }
            if (!cont) {
                // Renderer.cpp:240-242
                ax += (double)clamp01f(L.x * inv_norm);
                ay += (double)clamp01f(L.y * inv_norm);
                az += (double)clamp01f(L.z * inv_norm);
                ++s;
                if (s == s_end) {
                    records[q] = make_float4((float)ax, (float)ay, (float)az, 0.f);
                    rays_ref += (unsigned)(s_end - s_begin);
                    have = false;
                } else {
                    o = P; id = Pid; w = oriented_normal(sh, Pid, P, Pflip);
                    T = mk(1, 1, 1); L = mk(0, 0, 0); depth = 0;
                }
            }
        }
    }

    rays_ref += traced;
    for (int off = 16; off > 0; off >>= 1) {
        rays_ref += __shfl_down_sync(0xffffffffu, rays_ref, off);
        traced += __shfl_down_sync(0xffffffffu, traced, off);
    }
    if (lane == 0) {
        atomicAdd(&counters->rays_ref, rays_ref);
        atomicAdd(&counters->rays_traced, traced);
        atomicAdd(&counters->tests, traced * (unsigned long long)n);
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

// ---- test hooks ---------------------------------------------------------------
__global__ void closest_hit_f32_kernel(int n_sph, double rx, double ry, double rz, uint32_t n, const double* __restrict__ org,
                                       const double* __restrict__ dir, const int32_t* __restrict__ origin_id,
                                       int32_t* __restrict__ hit, double* __restrict__ tout) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const f3 o = mk((float)(org[3 * i] - rx), (float)(org[3 * i + 1] - ry), (float)(org[3 * i + 2] - rz));
    const f3 d = mk((float)dir[3 * i], (float)dir[3 * i + 1], (float)dir[3 * i + 2]);
    float t;
    const int id = trace_f32<0>(n_sph, origin_id ? origin_id[i] : -1, o, d, t);
    hit[i] = id;
    tout[i] = (double)t;
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

template <int N>
void launch_mega(const RtmPassParams& p, float4* records, RtmCounters* counters, int n_rt, int sm_count, cudaStream_t st) {
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, radiance_mega_kernel<N>, RTM_MEGA_THREADS, 0);
    if (per_sm < 1) per_sm = 1;
    radiance_mega_kernel<N><<<sm_count * per_sm, RTM_MEGA_THREADS, 0, st>>>(p, records, counters, n_rt);
}

}  // namespace

void rtm_launch_radiance_mega(int32_t n, const RtmPassParams& p, float4* records, RtmCounters* counters, int32_t sm_count,
                              cudaStream_t st) {
    switch (n) {
        case 1: launch_mega<1>(p, records, counters, n, sm_count, st); break;
        case 2: launch_mega<2>(p, records, counters, n, sm_count, st); break;
        case 3: launch_mega<3>(p, records, counters, n, sm_count, st); break;
        case 4: launch_mega<4>(p, records, counters, n, sm_count, st); break;
        case 5: launch_mega<5>(p, records, counters, n, sm_count, st); break;
        case 6: launch_mega<6>(p, records, counters, n, sm_count, st); break;
        case 7: launch_mega<7>(p, records, counters, n, sm_count, st); break;
        case 8: launch_mega<8>(p, records, counters, n, sm_count, st); break;
        default: launch_mega<0>(p, records, counters, n, sm_count, st); break;
    }
}

void rtm_launch_accumulate(const RtmPassParams& p, const float4* records, float4* accum, cudaStream_t st) {
    const int g = p.frame.ss * p.frame.ss;
    const unsigned long long n_pix = (unsigned long long)p.frame.W * p.frame.H;
    if (g <= 32 && (g & (g - 1)) == 0) {
        accumulate_pow2_kernel<<<(unsigned)((p.n_sub + 255) / 256), 256, 0, st>>>(records, accum, p.n_sub, g);
    } else {
        accumulate_generic_kernel<<<(unsigned)((n_pix + 255) / 256), 256, 0, st>>>(records, accum, n_pix, g);
    }
}

void rtm_launch_to_u8(const float4* accum, unsigned char* out, uint64_t n_pix, cudaStream_t st) {
    to_u8_kernel<<<(unsigned)((n_pix + 255) / 256), 256, 0, st>>>(accum, out, n_pix);
}

void rtm_launch_closest_hit_f32(int32_t n_spheres, const double* refd, uint32_t n, const double* org, const double* dir,
                                const int32_t* origin_id, int32_t* hit, double* t, cudaStream_t st) {
    closest_hit_f32_kernel<<<(n + 127) / 128, 128, 0, st>>>(n_spheres, refd[0], refd[1], refd[2], n, org, dir, origin_id, hit, t);
}

void rtm_launch_philox(uint32_t n, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t* out, cudaStream_t st) {
    philox_kernel<<<(n + 127) / 128, 128, 0, st>>>(n, c1, c2, c3, k0, k1, out);
}

void rtm_launch_fma_peak(float* sink, int32_t blocks, int32_t iters, cudaStream_t st) {
    fma_peak_kernel<<<blocks, 256, 0, st>>>(sink, iters);
}
