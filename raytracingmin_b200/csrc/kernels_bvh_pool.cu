// kernels_bvh_pool.cu — the hierarchy path's radiance kernel with MORE RAYS THAN LANES (config C5).
//
// radiance_bvh_kernel (kernels_bvh.cu) gives every lane one ray and runs the warp in phases (walk inner nodes / test a
// leaf / shade); ncu puts it at 15 of 32 lanes per instruction — 17 in the inner-node block, 10 in the leaf block, 6 in the
// shading rounds — and every sweep of its phase thresholds is flat: with 32 rays per warp there are never 32 in the same
// phase.  Here a warp owns RTM_POOL_SLOTS = 64 rays.  Their state lives in shared memory (24 words + a traversal stack
// per ray); a phase picks 32 rays that are IN that phase, loads what the phase needs into registers, runs the very same
// device functions as radiance_bvh_kernel, and writes back what changed:
//   S  shade: radiance_bvh_kernel's S block as it was when this kernel was measured (refill, apply the finished ray,
//      warp-wide roulette rounds, start the next ray; the lane-owned kernel has since moved restarts into a private loop),
//      run when half of the warp's live rays wait for it — 32 lanes enter instead of 12;
//   T  trace: up to RTM_POOL_BURST inner nodes, then the leaf for the lanes that reached one, RTM_POOL_ROUNDS times;
//      all 32 lanes start on an inner node.
// Same per-ray arithmetic, same RNG counters, same tie rules: the image is bit-identical to radiance_bvh_kernel's.
// Scenes with planes keep the lane-owned kernel (the host chooses, kernels_bvh.cu).
#include "kernels_common.cuh"

namespace {

#ifndef RTM_POOL_BURST
#define RTM_POOL_BURST 8
#endif
#ifndef RTM_POOL_ROUNDS
#define RTM_POOL_ROUNDS 2
#endif
#ifndef RTM_POOL_MINBLOCKS
#define RTM_POOL_MINBLOCKS 5
#endif

constexpr int kSlots = RTM_POOL_BLOCK_SLOTS;   // ray slots per block: RTM_POOL_SLOTS per warp

// SoA over the block's slots.  Two lanes of a warp can meet on a bank (slots s and s + 32): 2-way conflicts at worst.
struct PoolState {
    float ox[kSlots], oy[kSlots], oz[kSlots];   // vertex the path stands on = origin of its ray
    float dx[kSlots], dy[kSlots], dz[kSlots];   // ray direction
    float best[kSlots];                         // BvhRay::best
    int hit[kSlots];                            // BvhRay::hit
    int node[kSlots];                           // BvhRay::node; RTM_BVH_NONE_DEV = not walking (waits for the S phase)
    int sp[kSlots];                             // BvhRay::sp (shared-memory address of the next free stack entry)
    int sslot[kSlots];                          // leaf slot of the sphere the ray leaves
    float Tx[kSlots], Ty[kSlots], Tz[kSlots];   // throughput
    float Lx[kSlots], Ly[kSlots], Lz[kSlots];   // radiance of the current sample
    float ax[kSlots], ay[kSlots], az[kSlots];   // sum over the sub-pixel's finished samples
    int id[kSlots];                             // object the path stands on
    unsigned sd[kSlots];                        // sample << 9 | depth
    unsigned ql[kSlots];                        // record index of the sub-pixel
    unsigned eps[kSlots];                       // ACNE: the 32 random bits behind the model of the reference's Normalize() error
    unsigned flags[kSlots];                     // kHave | kFresh | kPending | kDone
};
constexpr unsigned kHave = 1u, kFresh = 2u, kPending = 4u, kDone = 8u;

template <int NP, bool ACNE>
__global__ void __launch_bounds__(RTM_MEGA_THREADS, RTM_POOL_MINBLOCKS)
radiance_bvh_pool_kernel(const __grid_constant__ RtmHotScene hot, const __grid_constant__ RtmPassParams p, float4* __restrict__ records,
                         RtmCounters* __restrict__ counters, RtmBvhDev bvh, RtmMatDev mat, const float* __restrict__ cps0, int n_sph) {
    __shared__ SelfMask sh_mask;
    __shared__ SharedBig sh_big;
    __shared__ PoolState ps;
    __shared__ int sh_stack[RTM_POOL_LEVELS * kSlots];
    __shared__ unsigned char sh_list[RTM_MEGA_THREADS / 32][RTM_POOL_SLOTS];
    init_selfmask(sh_mask);
    if (threadIdx.x < 8) sh_big.id[threadIdx.x] = (threadIdx.x < (unsigned)bvh.n_big) ? bvh.big_id[threadIdx.x] : 0;
    for (int k = threadIdx.x; k < kSlots; k += RTM_MEGA_THREADS) {
        ps.flags[k] = 0u;
        ps.node[k] = RTM_BVH_NONE_DEV;
    }
    __syncthreads();

    constexpr int ST = kSlots;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    const unsigned sbase = warp * RTM_POOL_SLOTS;
    const unsigned n_sub = (unsigned)p.n_sub;
    const int s_begin = p.s_begin;
    const float inv_norm = p.inv_norm;
    const unsigned sub_begin = (unsigned)p.sub_begin, sub_count = (unsigned)p.sub_count;
    const unsigned grid = (unsigned)(p.frame.ss * p.frame.ss);
    const bool full_grid = (sub_count == grid);
    const unsigned key_mask = p.key_mask;
    const unsigned sd_end = (unsigned)p.s_end << 9;
    const unsigned ns_all = (unsigned)(p.s_end - p.s_begin);

    unsigned w_next = 0, w_end = 0;
    bool wdone = false;              // warp-uniform: the work cursor is exhausted
    unsigned units = 0, traced = 0, n_tests = 0, n_nodes = 0;
    auto flush_if_big = [&]() {
        if ((traced | n_tests | n_nodes) & 0x80000000u) {
            atomicAdd(&counters->rays_ref, (unsigned long long)traced);
            atomicAdd(&counters->rays_traced, (unsigned long long)traced);
            atomicAdd(&counters->tests, (unsigned long long)n_tests);
            atomicAdd(&counters->nodes, (unsigned long long)n_nodes);
            traced = n_tests = n_nodes = 0;
        }
    };

    for (;;) {
        // ---- which of the warp's 64 rays are walking, which wait for the S phase
        const int nd0 = ps.node[sbase + lane], nd1 = ps.node[sbase + lane + 32];
        const unsigned fl0 = ps.flags[sbase + lane], fl1 = ps.flags[sbase + lane + 32];
        const unsigned t0 = __ballot_sync(0xffffffffu, nd0 != RTM_BVH_NONE_DEV), t1 = __ballot_sync(0xffffffffu, nd1 != RTM_BVH_NONE_DEV);
        const unsigned s0 = __ballot_sync(0xffffffffu, nd0 == RTM_BVH_NONE_DEV && !(fl0 & kDone));
        const unsigned s1 = __ballot_sync(0xffffffffu, nd1 == RTM_BVH_NONE_DEV && !(fl1 & kDone));
        const int nT = __popc(t0) + __popc(t1), nS = __popc(s0) + __popc(s1);
        if ((nT | nS) == 0) break;
        // shade when half of the live rays wait for it (32 of 64 until the work runs out)
        const int half = (nT + nS + 1) >> 1;
        const bool doS = nT == 0 || nS >= (half < 32 ? half : 32);
        const unsigned m0 = doS ? s0 : t0, m1 = doS ? s1 : t1;
        if ((m0 >> lane) & 1u) sh_list[warp][__popc(m0 & lt_mask)] = (unsigned char)lane;
        if ((m1 >> lane) & 1u) sh_list[warp][__popc(m0) + __popc(m1 & lt_mask)] = (unsigned char)(lane + 32u);
        __syncwarp();
        const int n_el = __popc(m0) + __popc(m1);
        const bool act = (int)lane < n_el;
        const unsigned slot = sbase + (act ? (unsigned)sh_list[warp][lane] : lane);
        int* const stack = sh_stack + slot;
        f3 o = mk(0, 0, 0), d = mk(0, 0, 1);
        BvhRay ray;
        ray.best = 0.f; ray.hit = -1; ray.node = RTM_BVH_NONE_DEV; ray.sp = 0;
        int sslot = -1;

        if (!doS) {
            // ---- T phase
            bool tracing = act;
            if (act) {
                o = mk(ps.ox[slot], ps.oy[slot], ps.oz[slot]);
                d = mk(ps.dx[slot], ps.dy[slot], ps.dz[slot]);
                ray.best = ps.best[slot]; ray.hit = ps.hit[slot]; ray.node = ps.node[slot]; ray.sp = ps.sp[slot];
                sslot = ps.sslot[slot];
                bvh_ray_frame<true>(bvh, o, d, ray);
            }
#pragma unroll 1
            for (int round = 0; round < RTM_POOL_ROUNDS; ++round) {
                int k = 0;
#pragma unroll 1
                for (; k < RTM_POOL_BURST && tracing && bvh_is_inner(ray.node); ++k) bvh16_inner_step<ST>(bvh, ray, stack);
                n_nodes += (unsigned)k;
                if (tracing && ray.node < 0) bvh_leaf_step<ST>(bvh, sslot, o, d, ray, stack, n_tests);
                if (tracing && ray.node == RTM_BVH_NONE_DEV) tracing = false;
                if (!__any_sync(0xffffffffu, tracing)) break;
            }
            if (act) {
                ps.best[slot] = ray.best; ps.hit[slot] = ray.hit; ps.node[slot] = ray.node; ps.sp[slot] = ray.sp;
                if (ray.node == RTM_BVH_NONE_DEV) ps.flags[slot] = kHave | kPending;
            }
            __syncwarp();
            continue;
        }

        // ---- S phase: radiance_bvh_kernel's S block on the lane's slot
        unsigned fl = act ? ps.flags[slot] : kDone;
        bool have = (fl & kHave) != 0u, fresh = (fl & kFresh) != 0u, pending = (fl & kPending) != 0u, done = (fl & kDone) != 0u;
        unsigned ql = 0, sd = 0;
        uint32_t eps_bits = 0;
        f3 acc = mk(0, 0, 0), T = mk(1, 1, 1), L = mk(0, 0, 0);
        int id = 0;
        if (act && have) {
            ql = ps.ql[slot]; sd = ps.sd[slot]; id = ps.id[slot];
            acc = mk(ps.ax[slot], ps.ay[slot], ps.az[slot]);
            T = mk(ps.Tx[slot], ps.Ty[slot], ps.Tz[slot]);
            L = mk(ps.Lx[slot], ps.Ly[slot], ps.Lz[slot]);
            o = mk(ps.ox[slot], ps.oy[slot], ps.oz[slot]);
            d = mk(ps.dx[slot], ps.dy[slot], ps.dz[slot]);
            ray.best = ps.best[slot]; ray.hit = ps.hit[slot];
            if (ACNE) eps_bits = ps.eps[slot];
        }
        // refill (warp-uniform): slots without a sub-pixel take the next ones of the warp's chunk
        unsigned need = wdone ? 0u : __ballot_sync(0xffffffffu, act && !have);
        while (need) {
            if (w_next == w_end) {
                unsigned long long start = 0;
                unsigned chunk = RTM_MEGA_CHUNK;
                if (lane == 0) {
                    const unsigned long long seen = *(volatile unsigned long long*)&counters->work_next;
                    const unsigned long long fair = (seen < n_sub ? n_sub - seen : 0ull) / (2ull * gridDim.x * (RTM_MEGA_THREADS / 32));
                    chunk = fair >= RTM_MEGA_CHUNK ? RTM_MEGA_CHUNK : (fair < 4ull ? 4u : (unsigned)fair);
                    start = atomicAdd(&counters->work_next, (unsigned long long)chunk);
                }
                start = __shfl_sync(0xffffffffu, start, 0);
                chunk = __shfl_sync(0xffffffffu, chunk, 0);
                if (start >= (unsigned long long)n_sub) {
                    wdone = true;
                    break;
                }
                w_next = (unsigned)start;
                w_end = (start + chunk < (unsigned long long)n_sub) ? (unsigned)start + chunk : n_sub;
            }
            const unsigned avail = w_end - w_next;
            const unsigned rank = __popc(need & lt_mask);
            const bool mine = ((need >> lane) & 1u) && rank < avail;
            if (mine) {
                ql = w_next + rank;
                const float4 rec = records[ql];
                const unsigned wbits = __float_as_uint(rec.w);
                const bool miss = (wbits == RTM_REC_MISS);
                const int Pid = miss ? 0 : (int)(wbits & RTM_REC_IDMASK);
                RTM_ASSERT(ql < n_sub && (miss || Pid < n_sph));
                const float4 e0 = miss ? make_float4(0.f, 0.f, 0.f, 0.f) : __ldg(mat.emi + Pid);
                if (miss || __float_as_uint(e0.w) == 0u) {
                    float4 out = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (!miss) {
                        out.x = clamp01f(e0.x * inv_norm) * (float)ns_all;
                        out.y = clamp01f(e0.y * inv_norm) * (float)ns_all;
                        out.z = clamp01f(e0.z * inv_norm) * (float)ns_all;
                    }
                    records[ql] = out;
                    ++units;
                } else {
                    have = true;
                    acc = mk(0, 0, 0);
                    sd = (unsigned)s_begin << 9;
                    fresh = true;
                }
            }
            const unsigned took = (__popc(need) < avail) ? __popc(need) : avail;
            w_next += took;
            need = __ballot_sync(0xffffffffu, act && !have && !mine);
        }
        if (wdone && act && !have) done = true;   // nothing left to fetch: the slot retires

        const bool in_s = act && have;
        {
            f3 w = mk(0, 0, 1);
            float cpsn = 0.f;   // ACNE: c' / (2 tau r) at the vertex the path stands on
            unsigned thr = 0;   // roulette threshold kd * 2^32 of the surface the path stands on
            auto stand_on_primary = [&](const float4 rec) {
                const unsigned wbits = __float_as_uint(rec.w);
                const int pid = (int)(wbits & RTM_REC_IDMASK);
                const float4 g = __ldg(mat.geo + pid), e0 = __ldg(mat.emi + pid);
                const float ir = (wbits & RTM_REC_FLIP) ? -g.w : g.w;
                o = mk(rec.x, rec.y, rec.z);
                w = mk((o.x - g.x) * ir, (o.y - g.y) * ir, (o.z - g.z) * ir);
                id = pid;
                T = mk(1, 1, 1); L = mk(e0.x, e0.y, e0.z); sd &= ~511u;
                thr = __float_as_uint(e0.w);
                if (ACNE) cpsn = (wbits & RTM_REC_FLIP) ? 0.f : cps0[ql] * g.w * RTM_ACNE_INV2TAU;
            };
            auto apply_hit = [&](int id2, float t) -> bool {
                if (id2 < 0) return true;                                            // miss: Renderer.cpp:116
                RTM_ASSERT(id2 < n_sph);
                o = mk(fmaf(d.x, t, o.x), fmaf(d.y, t, o.y), fmaf(d.z, t, o.z));     // :79
                const float4 g = __ldg(mat.geo + id2);
                f3 nn = mk((o.x - g.x) * g.w, (o.y - g.y) * g.w, (o.z - g.z) * g.w);
                if (g.w > 1e-2f) {   // small spheres: back onto the surface, unit normal (as radiance_bvh_kernel)
                    const float inv = fast_rsqrt(dot(nn, nn));
                    nn = mk(nn.x * inv, nn.y * inv, nn.z * inv);
                    const float rr = fast_rcp(g.w);
                    o = mk(fmaf(rr, nn.x, g.x), fmaf(rr, nn.y, g.y), fmaf(rr, nn.z, g.z));
                }
                const bool flip = !(dot(nn, d) < 0.f);                               // :82-83
                w = flip ? mk(-nn.x, -nn.y, -nn.z) : nn;
                id = id2;
                ++sd;
                if (ACNE) cpsn = flip ? 0.f : t * t * ref_normalize_eps(eps_bits) * g.w * RTM_ACNE_INV2TAU;
                const float4 e2 = __ldg(mat.emi + id2);
                L.x = fmaf(T.x, e2.x, L.x); L.y = fmaf(T.y, e2.y, L.y); L.z = fmaf(T.z, e2.z, L.z);
                thr = __float_as_uint(e2.w);
                return thr == 0u;                                                    // kd == 0: the next roulette cannot continue
            };
            bool end = false;
            if (in_s && fresh) {
                fresh = false;
                stand_on_primary(records[ql]);
            }
            if (in_s && pending) {
                pending = false;
                end = apply_hit(bvh_resolve_hit(bvh, ray.hit), ray.best + RTM_T_EPS);
            }
            bool deciding = in_s, start = false;
            int ocode = -1;
            for (;;) {
                if (deciding) {
                    if (!end) {
                        const unsigned q = full_grid ? ql : (ql / sub_count) * grid + sub_begin + (ql % sub_count);
                        uint2 rnd = rtm_bounce_rng(make_uint2(q, sd), p.rk);
                        end = !((rnd.x < thr) && ((sd & 511u) + 1u < (unsigned)p.max_depth));   // Renderer.cpp:78
#if RTM_BVH_FASTFWD
                        // A sample killed at its PRIMARY vertex (depth 0: L = the primary hit's emission, T = 1) leaves the
                        // lane exactly where stand_on_primary put it, so the lane closes it here and draws for the next
                        // sample at once instead of asking the whole warp for another round (a quarter of all restarts on
                        // synth10k).  The sub-pixel's last sample is left to the block below, which also finishes the unit.
                        while (end && (sd & 511u) == 0u && sd + 512u != sd_end) {
                            acc.x += clamp01f(L.x * inv_norm);
                            acc.y += clamp01f(L.y * inv_norm);
                            acc.z += clamp01f(L.z * inv_norm);
                            sd += 512u;
                            rnd = rtm_bounce_rng(make_uint2(q, sd), p.rk);
                            end = !((rnd.x < thr) && (1u < (unsigned)p.max_depth));
                        }
#endif
                        if (!end) {
                            const float4 al = __ldg(mat.alb + id);
                            T.x *= al.x; T.y *= al.y; T.z *= al.z;                               // colorKD
                            d = cosine_dir(w, rnd.y, ((rnd.x & 0xFFFu) << 10) | (rnd.y & 0x3FFu));
                            eps_bits = rnd.x ^ (rnd.y * 0x9E3779B9u);
                            ++traced;
                            bool self = false;
                            if (ACNE) {
                                const float cosw = dot(w, d);
                                if (cpsn + cosw <= 0.f) {
                                    // SettingData.cpp:208-212: the far root of the sphere the ray stands (just) inside
                                    const float rr = fast_rcp(__ldg(mat.geo + id).w);
                                    const float qq = -cpsn * rr * RTM_ACNE_2TAU, bq = rr * cosw;
                                    end = apply_hit(id, qq * fast_rcp(bq + fast_sqrt(fmaf(bq, bq, qq))));
                                    self = true;   // stands on the next vertex: decide again
                                }
                            }
                            if (!self) {
                                ocode = __float_as_int(al.w);
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
                        sd = (sd | 511u) + 1u;   // next sample, depth 0
                        if (sd == sd_end) {
                            records[ql] = make_float4(acc.x, acc.y, acc.z, 0.f);
                            ++units;
                            flush_if_big();
                            have = false;
                            deciding = false;
                        } else {
                            stand_on_primary(records[ql]);
                            end = false;
                        }
                    }
                }
                if (!__any_sync(0xffffffffu, deciding)) break;
            }
            if (start) {
                bvh_ray_begin<NP, ST, true>(hot, sh_mask, sh_big, bvh, bvh_big_slot(ocode), o, d, key_mask, ray, stack, n_tests);
                sslot = bvh_leaf_slot(ocode);
                if (ray.node == RTM_BVH_NONE_DEV) pending = true;   // (no hierarchy to walk: cannot happen on this path)
            }
        }
        if (act) {
            ps.flags[slot] = (have ? kHave : 0u) | (fresh ? kFresh : 0u) | (pending ? kPending : 0u) | (done ? kDone : 0u);
            ps.node[slot] = have ? ray.node : RTM_BVH_NONE_DEV;
            if (have) {
                ps.ql[slot] = ql; ps.sd[slot] = sd; ps.id[slot] = id;
                ps.ax[slot] = acc.x; ps.ay[slot] = acc.y; ps.az[slot] = acc.z;
                ps.Tx[slot] = T.x; ps.Ty[slot] = T.y; ps.Tz[slot] = T.z;
                ps.Lx[slot] = L.x; ps.Ly[slot] = L.y; ps.Lz[slot] = L.z;
                ps.ox[slot] = o.x; ps.oy[slot] = o.y; ps.oz[slot] = o.z;
                ps.dx[slot] = d.x; ps.dy[slot] = d.y; ps.dz[slot] = d.z;
                ps.best[slot] = ray.best; ps.hit[slot] = ray.hit; ps.sp[slot] = ray.sp; ps.sslot[slot] = sslot;
                if (ACNE) ps.eps[slot] = eps_bits;
            }
        }
        __syncwarp();
    }

    unsigned long long rays_ref = (unsigned long long)units * ns_all + traced, traced64 = traced;
    unsigned long long tests = n_tests, nodes = n_nodes;
    for (int off = 16; off > 0; off >>= 1) {
        rays_ref += __shfl_down_sync(0xffffffffu, rays_ref, off);
        traced64 += __shfl_down_sync(0xffffffffu, traced64, off);
        tests += __shfl_down_sync(0xffffffffu, tests, off);
        nodes += __shfl_down_sync(0xffffffffu, nodes, off);
    }
    if (lane == 0) {
        atomicAdd(&counters->rays_ref, rays_ref);
        atomicAdd(&counters->rays_traced, traced64);
        atomicAdd(&counters->tests, tests);
        atomicAdd(&counters->nodes, nodes);
    }
}

template <int NP, bool ACNE>
void launch_pool2(const RtmHotScene& hot, const RtmBvhDev& bvh, const RtmMatDev& mat, const RtmPassParams& p, float4* records, RtmCounters* counters,
                  int sm_count, const float* cps0, int n_sph, cudaStream_t st) {
    auto kern = radiance_bvh_pool_kernel<NP, ACNE>;
    // 44 KB of ray state per block: ask for the shared-memory end of the L1 split, or fewer blocks than planned become resident
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, RTM_MEGA_THREADS, 0);
    if (per_sm < 1) per_sm = 1;
    kern<<<sm_count * per_sm, RTM_MEGA_THREADS, 0, st>>>(hot, p, records, counters, bvh, mat, cps0, n_sph);
}
template <int NP>
void launch_pool(const RtmHotScene& hot, const RtmBvhDev& bvh, const RtmMatDev& mat, const RtmPassParams& p, float4* records, RtmCounters* counters,
                 int sm_count, const float* cps0, int n_sph, cudaStream_t st) {
    if (cps0) launch_pool2<NP, true>(hot, bvh, mat, p, records, counters, sm_count, cps0, n_sph, st);
    else launch_pool2<NP, false>(hot, bvh, mat, p, records, counters, sm_count, nullptr, n_sph, st);
}

}  // namespace

bool rtm_pool_kernel_takes(const RtmBvhDev& bvh, int32_t n_pl) {
    return bvh.nodes16 != nullptr && n_pl == 0 && bvh.depth + 2 <= RTM_POOL_LEVELS;
}

void rtm_launch_radiance_bvh_pool(const RtmHotScene& hot, const RtmBvhDev& bvh, const RtmMatDev& mat, const RtmPassParams& p, float4* records,
                                  RtmCounters* counters, int32_t sm_count, const float* cps0, int32_t n_sph, cudaStream_t st) {
    switch ((bvh.n_big + 1) / 2) {
        case 0: launch_pool<0>(hot, bvh, mat, p, records, counters, sm_count, cps0, n_sph, st); break;
        case 1: launch_pool<1>(hot, bvh, mat, p, records, counters, sm_count, cps0, n_sph, st); break;
        case 2: launch_pool<2>(hot, bvh, mat, p, records, counters, sm_count, cps0, n_sph, st); break;
        case 3: launch_pool<3>(hot, bvh, mat, p, records, counters, sm_count, cps0, n_sph, st); break;
        default: launch_pool<4>(hot, bvh, mat, p, records, counters, sm_count, cps0, n_sph, st); break;
    }
}
