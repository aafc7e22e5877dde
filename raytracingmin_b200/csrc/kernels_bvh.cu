// kernels_bvh.cu — the FP32 bounce loop of the hierarchy path (scenes of more than 64 spheres, config C5).
#include "kernels_common.cuh"

namespace {

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
#define RTM_BVH_S_THRESH 12   // swept again on the compact-node walk (r02): 12 6.18, 16 6.11, 20 5.95, 24 5.60 Grays/s on C5
#endif
#ifndef RTM_BVH_BURST
#define RTM_BVH_BURST 8
#endif
#ifndef RTM_BVH_MINBLOCKS
#define RTM_BVH_MINBLOCKS 8   // 64 registers: reachable without spills since the S-only path state is parked in shared memory (7: 6.64, 8: 7.07 Grays/s on C5)
#endif
#ifndef RTM_BVH_HOLD
#define RTM_BVH_HOLD 0
#endif
#ifndef RTM_BVH_SHPRIM
#define RTM_BVH_SHPRIM 1   // the sub-pixel's primary-hit constants in shared memory (restart = 3 LDS.128, no global load)
#endif
// The S block's decision loop is left by every lane on its own (the lanes meet again at a __syncwarp before the ray start,
// which therefore still runs once): a lane that has its direction no longer sits through the re-decisions of the others
// (ACNE self-hits), and the warp-wide ballot per round is gone.  0 = the warp-uniform rounds (RTM_BVH_S_ROUND).  7.05 -> 7.30
// Grays/s on C5 (profiles/r02/c5_sblock_r02i.jsonl).
#ifndef RTM_BVH_S_PERLANE
#define RTM_BVH_S_PERLANE 1
#endif
#ifndef RTM_BVH_S_ROUND
#define RTM_BVH_S_ROUND 1   // lanes that must need another roulette decision for the S block to run another round (swept: 1 4.50, 2 4.46, 4 4.45, 8 4.34 Grays/s on C5)
#endif

// Material records (geo / emi / alb: 48 bytes per sphere, 480 KB at 10^4 spheres, one random sphere per hit) have no reuse
// worth an L1 line, and an LRU model of the SM's L1 (scripts/study_bvh_layout.py) puts the node hit rate at 62 % with them
// allocating (ncu measured 60 %) and at 75 % without.  Reading them with L1::no_allocate (RTM_BVH_MAT_NA=1) was measured:
// 6.30 against 6.38 Grays/s on C5 (same box, profiles/r02/c5_alternatives_r02g.jsonl) — the hit rate is not what the walk
// waits for (a warp waits for its slowest lane; with 17 lanes on 17 nodes one of them misses either way).  Left off.
#ifndef RTM_BVH_MAT_NA
#define RTM_BVH_MAT_NA 0
#endif
__device__ __forceinline__ float4 mat_ld(const float4* p) {
#if RTM_BVH_MAT_NA
    float4 v;
    asm("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
#else
    return __ldg(p);
#endif
}

// PL: n_pl planes (extension) behind the n_sph spheres in the material records, tested at the start of every ray
// SST: the per-lane traversal stack lives in shared memory, stack[level][thread], RTM_BVH_SSTACK levels (the host
// picks it when the tree is shallow enough: depth + 2 entries are the most a walk can hold); otherwise a private
// 64-entry array in local memory.
// WK: the nodes walked — 0 = 64-byte float nodes, 1 = compact half-precision two-child nodes (one 32-byte request per visit;
// bvh.h: RtmBvh16Host), 2 = compact 4-wide nodes (two requests per visit, half the visits; bvh.h: RtmBvh16x4Host)
#ifndef RTM_BVH4_MINBLOCKS
#define RTM_BVH4_MINBLOCKS 7
#endif
#ifndef RTM_BVH4_BURST
#define RTM_BVH4_BURST 4   // 4-wide steps per burst = the work of 8 two-child steps (measured: 4 6.42, 8 6.04, 12 5.31 Grays/s on C5)
#endif
template <int NP, bool ACNE, bool PL, bool SST, int WK>
__global__ void __launch_bounds__(RTM_MEGA_THREADS, WK == 2 ? RTM_BVH4_MINBLOCKS : RTM_BVH_MINBLOCKS)
radiance_bvh_kernel(const __grid_constant__ RtmHotScene hot, const __grid_constant__ RtmPassParams p, float4* __restrict__ records,
                    RtmCounters* __restrict__ counters, RtmBvhDev bvh, RtmMatDev mat, const float* __restrict__ cps0, int n_sph, int n_pl) {
    __shared__ SelfMask sh_mask;
    __shared__ SharedBig sh_big;
#if RTM_BVH_SHPRIM
    // What a sample restart needs of the lane's sub-pixel — primary hit point + id, oriented normal there + ACNE state, its
    // emission + roulette threshold — stays in shared memory ([k][thread], three LDS.128) for as long as the lane holds the
    // sub-pixel: a restart costs no global load, so a lane closes and restarts samples in a private loop inside the round
    // (kernels_mega.cu: SharedPrimary, the same idea).  6 KB a block.
    __shared__ float4 sh_ppos[RTM_MEGA_THREADS], sh_pnrm[RTM_MEGA_THREADS], sh_pemi[RTM_MEGA_THREADS];
#endif
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
    bool fresh = false;              // the lane stands on its sub-pixel's primary hit, and the S block has yet to load it
    // Path state that only the S block touches — throughput T, the sample's radiance L, the sub-pixel's sum acc, the object
    // the path stands on, sample << 9 | depth (the RNG counter word itself, as in radiance_mega_kernel), the record index ql
    // (the global sub-pixel index, the RNG counter's other word, is derived from it where it is needed) and the ACNE bits —
    // is PARKED in shared memory ([k][thread], 3 LDS.128 + 1 LDS at S entry, the same in stores at S exit) while the lane
    // walks: 13 registers the T block does not have to carry.
    __shared__ float4 sh_pT[RTM_MEGA_THREADS], sh_pL[RTM_MEGA_THREADS], sh_pA[RTM_MEGA_THREADS];   // T.xyz id | L.xyz sd | acc.xyz ql
    __shared__ uint32_t sh_peps[ACNE ? RTM_MEGA_THREADS : 1];
    f3 o = mk(0, 0, 0), d = mk(0, 0, 1);
    const unsigned sd_end = (unsigned)p.s_end << 9;
    // The oriented normal w and the ACNE state cpsn are NOT lane state: they are produced (apply_hit / stand_on_primary) and
    // consumed (roulette, direction sampling) inside one S block, so they are locals of that block — four registers that
    // the T block's |1/d| needs (72 registers = 7 resident blocks either way).
    int sslot = -1;                  // leaf slot of the sphere the current ray leaves (self test inside the leaves)
    BvhRay ray;
    ray.node = RTM_BVH_NONE_DEV;
    constexpr int ST = SST ? RTM_MEGA_THREADS : 1;
    __shared__ int sh_stack[SST ? RTM_BVH_SSTACK * RTM_MEGA_THREADS : 1];
    int lstack[SST ? 1 : RTM_BVH_STACK_DEV];
    int* const stack = SST ? (sh_stack + threadIdx.x) : lstack;
    // per-lane counters, all 32-bit (register pressure: the kernel sits at 72): sub-pixels done, bounce rays, sphere tests,
    // node visits.  They are flushed to the 64-bit device counters whenever one passes 2^31, checked when a sub-pixel ends.
    unsigned units = 0, traced = 0;
    unsigned n_tests = 0, n_nodes = 0;
    const unsigned ns_all = (unsigned)(p.s_end - p.s_begin);
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
        // ---- refill (warp-uniform): lanes without a sub-pixel take the next ones of the warp's chunk
        unsigned need = __ballot_sync(0xffffffffu, !have && !done);
        while (need) {
            if (w_next == w_end) {
                // guided self-scheduling (see radiance_mega_kernel): full chunks while plenty is left, smaller ones near the end
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
                    if (!have) done = true;
                    break;
                }
                w_next = (unsigned)start;
                w_end = (start + chunk < (unsigned long long)n_sub) ? (unsigned)start + chunk : n_sub;
            }
            const unsigned avail = w_end - w_next;
            const unsigned rank = __popc(need & lt_mask);
            const bool mine = ((need >> lane) & 1u) && rank < avail;
            if (mine) {
                const unsigned ql = w_next + rank;
                const float4 rec = records[ql];
                const unsigned wbits = __float_as_uint(rec.w);
                const unsigned ns = (unsigned)(s_end - s_begin);
                const bool miss = (wbits == RTM_REC_MISS);
                const int Pid = miss ? 0 : (int)(wbits & RTM_REC_IDMASK);
                RTM_ASSERT(ql < n_sub && (miss || Pid < n_sph + n_pl));
                const float4 e0 = miss ? make_float4(0.f, 0.f, 0.f, 0.f) : mat_ld(mat.emi + Pid);   // no record to read for a miss (n may be 0)
                if (miss || __float_as_uint(e0.w) == 0u) {
                    float4 out = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (!miss) {
                        out.x = clamp01f(e0.x * inv_norm) * (float)ns;
                        out.y = clamp01f(e0.y * inv_norm) * (float)ns;
                        out.z = clamp01f(e0.z * inv_norm) * (float)ns;
                    }
                    records[ql] = out;
                    ++units;
                } else {
                    have = true;
                    sh_pA[threadIdx.x] = make_float4(0.f, 0.f, 0.f, __uint_as_float(ql));
                    sh_pL[threadIdx.x] = make_float4(0.f, 0.f, 0.f, __uint_as_float((unsigned)s_begin << 9));
                    fresh = true;   // the S block of this very iteration stands the lane on its primary hit
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

        // ---- S block.  Entered by the whole warp (lanes outside it are predicated off).  Each lane decides on its own until it
        // has a direction (or its sub-pixel is finished): roulette, close-and-restart of ended samples in a private loop,
        // direction sampling, the ACNE self-hit re-decision.  The lanes meet again at a __syncwarp and the packed
        // big-sphere test + ray start run ONCE for every lane that has a direction.  (Round 1's per-lane loop had no such
        // meeting point: the ~140-instruction ray start ran once per group of lanes that left the loop together, 3.3
        // times per entry at 6 of 32 lanes — ncu_bvh_v3.  RTM_BVH_S_PERLANE = 0 is the intermediate form, warp-uniform
        // rounds that continue while RTM_BVH_S_ROUND or more lanes need one.)
        const bool in_s = (tr == 0u || __popc(sready) >= RTM_BVH_S_THRESH) && have && !tracing;
        if (__any_sync(0xffffffffu, in_s)) {
            __syncwarp();   // the refill's stores above
            const float4 pT = sh_pT[threadIdx.x], pL = sh_pL[threadIdx.x], pA = sh_pA[threadIdx.x];
            f3 T = mk(pT.x, pT.y, pT.z), L = mk(pL.x, pL.y, pL.z), acc = mk(pA.x, pA.y, pA.z);
            int id = __float_as_int(pT.w);
            unsigned sd = __float_as_uint(pL.w);
            const unsigned ql = __float_as_uint(pA.w);
            uint32_t eps_bits = ACNE ? sh_peps[threadIdx.x] : 0u;
            f3 w = mk(0, 0, 1);
            float cpsn = 0.f;
            unsigned thr = 0;   // roulette threshold kd * 2^32 of the surface the lane stands on (from the emission record read there)
    // the primary hit of the lane's sub-pixel is NOT lane state here: it is re-read from records[ql] (which holds it
            // until the sub-pixel's sum is written) at every sample restart — 8 registers less, which is what lets
            // RTM_BVH_MINBLOCKS blocks stay resident
            auto stand_on_primary = [&](const float4 rec) {
                const unsigned wbits = __float_as_uint(rec.w);
                const int pid = (int)(wbits & RTM_REC_IDMASK);
                const float4 g = mat_ld(mat.geo + pid), e0 = mat_ld(mat.emi + pid);
                const float ir = (wbits & RTM_REC_FLIP) ? -g.w : g.w;
                o = mk(rec.x, rec.y, rec.z);
                w = mk((o.x - g.x) * ir, (o.y - g.y) * ir, (o.z - g.z) * ir);
                if (PL && pid >= n_sph) w = (wbits & RTM_REC_FLIP) ? mk(-g.x, -g.y, -g.z) : mk(g.x, g.y, g.z);
                id = pid;
                T = mk(1, 1, 1); L = mk(e0.x, e0.y, e0.z); sd &= ~511u;
                thr = __float_as_uint(e0.w);
                if (ACNE) cpsn = (wbits & RTM_REC_FLIP) ? 0.f : cps0[ql] * g.w * RTM_ACNE_INV2TAU;
#if RTM_BVH_SHPRIM
                sh_ppos[threadIdx.x] = make_float4(o.x, o.y, o.z, __int_as_float(pid));
                sh_pnrm[threadIdx.x] = make_float4(w.x, w.y, w.z, cpsn);
                sh_pemi[threadIdx.x] = e0;
#endif
            };
#if RTM_BVH_SHPRIM
            auto restart_sample = [&]() {
                const float4 a = sh_ppos[threadIdx.x], b = sh_pnrm[threadIdx.x], c = sh_pemi[threadIdx.x];
                o = mk(a.x, a.y, a.z); id = __float_as_int(a.w);
                w = mk(b.x, b.y, b.z); cpsn = b.w;
                T = mk(1, 1, 1); L = mk(c.x, c.y, c.z); sd &= ~511u;
                thr = __float_as_uint(c.w);
            };
#else
            auto restart_sample = [&]() { stand_on_primary(records[ql]); };
#endif
            // a finished ray is applied to the path; returns true when the path ends there
            auto apply_hit = [&](int id2, float t) -> bool {
                if (id2 < 0) return true;                                            // miss: Renderer.cpp:116
                RTM_ASSERT(id2 < n_sph + n_pl);
                o = mk(fmaf(d.x, t, o.x), fmaf(d.y, t, o.y), fmaf(d.z, t, o.z));     // :79
                const float4 g = mat_ld(mat.geo + id2);
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
                ++sd;
                if (ACNE) cpsn = flip ? 0.f : t * t * ref_normalize_eps(eps_bits) * g.w * RTM_ACNE_INV2TAU;
                const float4 e2 = mat_ld(mat.emi + id2);
                L.x = fmaf(T.x, e2.x, L.x); L.y = fmaf(T.y, e2.y, L.y); L.z = fmaf(T.z, e2.z, L.z);
                thr = __float_as_uint(e2.w);
                return thr == 0u;                                  // kd == 0: the next roulette cannot continue
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
                    // RNG counter = GLOBAL sub-pixel index: the image does not depend on the partition (one register less
                    // than carrying it; the division only exists for sub-pixel-sliced multi-GPU handles)
                    const unsigned q = full_grid ? ql : (ql / sub_count) * grid + sub_begin + (ql % sub_count);
                    uint2 rnd = make_uint2(0u, 0u);
                    // Private loop: close samples until one survives its next roulette (or the sub-pixel is finished).  A
                    // restart is three LDS.128 (RTM_BVH_SHPRIM), so a lane whose path ended no longer asks the whole warp
                    // for another round — rounds remain for the ACNE self-hit re-decisions only.
#pragma unroll 1
                    for (;;) {
                        if (!end) {
                            rnd = rtm_bounce_rng(make_uint2(q, sd), p.rk);
                            end = !((rnd.x < thr) && ((sd & 511u) + 1u < (unsigned)p.max_depth));   // Renderer.cpp:78
                        }
                        if (!end) break;
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
                            break;
                        }
                        restart_sample();
                        end = false;
                    }
                    if (deciding) {
                        const float4 al = mat_ld(mat.alb + id);
                        T.x *= al.x; T.y *= al.y; T.z *= al.z;                               // colorKD
                        d = cosine_dir(w, rnd.y, ((rnd.x & 0xFFFu) << 10) | (rnd.y & 0x3FFu));
                        eps_bits = rnd.x ^ (rnd.y * 0x9E3779B9u);
                        ++traced;
                        bool self = false;
                        if (ACNE) {
                            const float cosw = dot(w, d);
                            if (cpsn + cosw <= 0.f && !(PL && id >= n_sph)) {
                                // SettingData.cpp:208-212: the far root of the sphere the ray stands (just) inside
                                const float rr = fast_rcp(mat_ld(mat.geo + id).w);
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
#if RTM_BVH_S_PERLANE
                if (!deciding) break;   // per-lane exit: lanes wait at the __syncwarp below
#else
                if (__popc(__ballot_sync(0xffffffffu, deciding)) < RTM_BVH_S_ROUND) break;
#endif
            }
#if RTM_BVH_S_PERLANE
            __syncwarp();
#endif
            if (deciding && have) fresh = true;   // (only with RTM_BVH_S_ROUND > 1) restarted in the last round: stands on its primary hit
            if (start) {
                // the packed big-sphere test (~140 instructions) and the planes: once, for all lanes that start a ray
                bvh_ray_begin<NP, ST, WK != 0>(hot, sh_mask, sh_big, bvh, bvh_big_slot(ocode), o, d, key_mask, ray, stack, n_tests);
                sslot = bvh_leaf_slot(ocode);
                if (PL) bvh_ray_planes(hot, n_sph, n_pl, id, o, d, ray);
                if (ray.node != RTM_BVH_NONE_DEV) tracing = true;
                else pending = true;   // empty hierarchy: the hit (big spheres / planes) is applied at the next entry
            }
            if (in_s) {
                sh_pT[threadIdx.x] = make_float4(T.x, T.y, T.z, __int_as_float(id));
                sh_pL[threadIdx.x] = make_float4(L.x, L.y, L.z, __uint_as_float(sd));
                sh_pA[threadIdx.x] = make_float4(acc.x, acc.y, acc.z, __uint_as_float(ql));
                if (ACNE) sh_peps[threadIdx.x] = eps_bits;
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
                    ray.node = bvh_pop<ST>(stack, ray);
                }
                if (!(tracing && bvh_is_inner(ray.node))) break;
                if (WK == 2) bvh16x4_inner_step<ST>(bvh, ray, stack);
                else if (WK == 1) bvh16_inner_step<ST>(bvh, ray, stack);
                else bvh_inner_step<ST>(bvh, ray, stack);
                ++n_nodes;
            }
            if (tracing && held != 0) bvh_leaf_test(bvh, held, sslot, o, d, ray, n_tests);
            else if (tracing && ray.node < 0) bvh_leaf_step<ST>(bvh, sslot, o, d, ray, stack, n_tests);
#else
            int k = 0;
#pragma unroll 1
            for (; k < (WK == 2 ? RTM_BVH4_BURST : RTM_BVH_BURST) && tracing && bvh_is_inner(ray.node); ++k) {
                if (WK == 2) bvh16x4_inner_step<ST>(bvh, ray, stack);
                else if (WK == 1) bvh16_inner_step<ST>(bvh, ray, stack);
                else bvh_inner_step<ST>(bvh, ray, stack);
            }
            n_nodes += (unsigned)k;   // visits counted once per burst, not once per step
            if (tracing && ray.node < 0) bvh_leaf_step<ST>(bvh, sslot, o, d, ray, stack, n_tests);
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

// the hierarchy variant of the hook: same trace function as radiance_mega_kernel<*, true>
template <int NP, int WK>
__global__ void closest_hit_bvh_f32_kernel(const __grid_constant__ RtmHotScene hot, RtmBvhDev bvh, RtmMatDev mat, unsigned key_mask, double rx, double ry, double rz, uint32_t n,
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
    const int ocode = (oid >= 0) ? __float_as_int(__ldg(mat.alb + oid).w) : -1;
    float t;
    unsigned a = 0, b = 0;
    hit[i] = trace_bvh_f32<NP, WK>(hot, sh_mask, sh_big, bvh, oid, ocode, o, d, t, key_mask, a, b, n_sph, n_pl);
    tout[i] = (double)t;
}

template <int NP, bool ACNE, bool PL, bool SST, int WK>
void launch_bvh4(const RtmHotScene& hot, const RtmBvhDev& bvh, const RtmMatDev& mat, const RtmPassParams& p, float4* records, RtmCounters* counters,
                 int sm_count, const float* cps0, int n_sph, int n_pl, cudaStream_t st) {
    // 26 KB of shared memory a block (traversal stacks, parked path state, primary-hit constants): RTM_BVH_MINBLOCKS of them
    // need the shared-memory end of the L1 split
    cudaFuncSetAttribute(radiance_bvh_kernel<NP, ACNE, PL, SST, WK>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, radiance_bvh_kernel<NP, ACNE, PL, SST, WK>, RTM_MEGA_THREADS, 0);
    if (per_sm < 1) per_sm = 1;
    radiance_bvh_kernel<NP, ACNE, PL, SST, WK><<<sm_count * per_sm, RTM_MEGA_THREADS, 0, st>>>(hot, p, records, counters, bvh, mat, cps0, n_sph, n_pl);
}
template <int NP, bool ACNE, bool PL, bool SST>
void launch_bvh3(const RtmHotScene& hot, const RtmBvhDev& bvh, const RtmMatDev& mat, const RtmPassParams& p, float4* records, RtmCounters* counters,
                 int sm_count, const float* cps0, int n_sph, int n_pl, cudaStream_t st) {
    if (bvh.nodes16x4) launch_bvh4<NP, ACNE, PL, SST, 2>(hot, bvh, mat, p, records, counters, sm_count, cps0, n_sph, n_pl, st);
    else if (bvh.nodes16) launch_bvh4<NP, ACNE, PL, SST, 1>(hot, bvh, mat, p, records, counters, sm_count, cps0, n_sph, n_pl, st);
    else launch_bvh4<NP, ACNE, PL, SST, 0>(hot, bvh, mat, p, records, counters, sm_count, cps0, n_sph, n_pl, st);
}
template <int NP, bool ACNE, bool PL>
void launch_bvh2(const RtmHotScene& hot, const RtmBvhDev& bvh, const RtmMatDev& mat, const RtmPassParams& p, float4* records, RtmCounters* counters,
                 int sm_count, const float* cps0, int n_sph, int n_pl, int walk, cudaStream_t st) {
    // a walk holds at most depth + 2 stack entries (one per level + the sentinel); walk & 1 (rtm_options.debug_bvh_walk,
    // A/B runs) keeps the stack in local memory
    const int need = bvh.nodes16x4 ? 3 * bvh.depth4 + 1 : bvh.depth + 2;
    if (!(walk & 1) && need <= RTM_BVH_SSTACK) launch_bvh3<NP, ACNE, PL, true>(hot, bvh, mat, p, records, counters, sm_count, cps0, n_sph, n_pl, st);
    else launch_bvh3<NP, ACNE, PL, false>(hot, bvh, mat, p, records, counters, sm_count, cps0, n_sph, n_pl, st);
}
template <int NP>
void launch_bvh(const RtmHotScene& hot, const RtmBvhDev& bvh, const RtmMatDev& mat, const RtmPassParams& p, float4* records, RtmCounters* counters,
                int sm_count, const float* cps0, int n_sph, int n_pl, int walk, cudaStream_t st) {
    if (n_pl > 0) {
        if (cps0) launch_bvh2<NP, true, true>(hot, bvh, mat, p, records, counters, sm_count, cps0, n_sph, n_pl, walk, st);
        else launch_bvh2<NP, false, true>(hot, bvh, mat, p, records, counters, sm_count, nullptr, n_sph, n_pl, walk, st);
    } else {
        if (cps0) launch_bvh2<NP, true, false>(hot, bvh, mat, p, records, counters, sm_count, cps0, n_sph, 0, walk, st);
        else launch_bvh2<NP, false, false>(hot, bvh, mat, p, records, counters, sm_count, nullptr, n_sph, 0, walk, st);
    }
}
}  // namespace

void rtm_launch_radiance_bvh(const RtmHotScene& hot, const RtmBvhDev& bvh, const RtmMatDev& mat, const RtmPassParams& p, float4* records,
                             RtmCounters* counters, int32_t sm_count, const float* cps0, int32_t n_sph, int32_t n_pl, int32_t walk,
                             cudaStream_t st) {
    // walk (rtm_options.debug_bvh_walk) bit 4: the ray-pool kernel (kernels_bvh_pool.cu), where it takes the scene.  An A/B
    // variant: bit-identical image, measured slower than this file's kernel (5.5-5.7 against 6.44 Grays/s on C5)
    if ((walk & 16) && !(walk & (1 | 8)) && rtm_pool_kernel_takes(bvh, n_pl)) {
        rtm_launch_radiance_bvh_pool(hot, bvh, mat, p, records, counters, sm_count, cps0, n_sph, st);
        return;
    }
    switch ((bvh.n_big + 1) / 2) {
        case 0: launch_bvh<0>(hot, bvh, mat, p, records, counters, sm_count, cps0, n_sph, n_pl, walk, st); break;
        case 1: launch_bvh<1>(hot, bvh, mat, p, records, counters, sm_count, cps0, n_sph, n_pl, walk, st); break;
        case 2: launch_bvh<2>(hot, bvh, mat, p, records, counters, sm_count, cps0, n_sph, n_pl, walk, st); break;
        case 3: launch_bvh<3>(hot, bvh, mat, p, records, counters, sm_count, cps0, n_sph, n_pl, walk, st); break;
        default: launch_bvh<4>(hot, bvh, mat, p, records, counters, sm_count, cps0, n_sph, n_pl, walk, st); break;
    }
}

void rtm_launch_closest_hit_bvh_f32(const RtmHotScene& hot, const RtmBvhDev& bvh, const RtmMatDev& mat, int32_t n_spheres, int32_t n_planes,
                                    const double* refd, uint32_t n, const double* org, const double* dir, const int32_t* origin_id,
                                    int32_t* hit, double* t, cudaStream_t st) {
    const unsigned g = (n + 127) / 128;
    switch ((bvh.n_big + 1) / 2) {
#define RTM_CASE(NP)                                                                                                                              \
    do {                                                                                                                                          \
        if (bvh.nodes16x4) closest_hit_bvh_f32_kernel<NP, 2><<<g, 128, 0, st>>>(hot, bvh, mat, 0xFFFFFFF8u, refd[0], refd[1], refd[2], n, org, dir, origin_id, n_spheres, n_planes, hit, t); \
        else if (bvh.nodes16) closest_hit_bvh_f32_kernel<NP, 1><<<g, 128, 0, st>>>(hot, bvh, mat, 0xFFFFFFF8u, refd[0], refd[1], refd[2], n, org, dir, origin_id, n_spheres, n_planes, hit, t); \
        else closest_hit_bvh_f32_kernel<NP, 0><<<g, 128, 0, st>>>(hot, bvh, mat, 0xFFFFFFF8u, refd[0], refd[1], refd[2], n, org, dir, origin_id, n_spheres, n_planes, hit, t);          \
    } while (0)
        case 0: RTM_CASE(0); break;
        case 1: RTM_CASE(1); break;
        case 2: RTM_CASE(2); break;
        case 3: RTM_CASE(3); break;
        default: RTM_CASE(4); break;
#undef RTM_CASE
    }
}
