// kernels_mega.cu — the FP32 bounce loop of the brute-force path (product path for scenes of up to 64 spheres,
// the headline configuration C4): a persistent megakernel.
//
// What the reference does per sample (src/Renderer.cpp:57-117, L-ABC): closest
// hit over all spheres, Russian roulette on kd, cosine-weighted bounce, recurse.
// Here each lane owns one sub-pixel at a time (its exact primary hit comes from the FP64 pass), walks that
// sub-pixel's samples one hit per loop iteration, and pulls the next sub-pixel from a warp-level
// work cursor as soon as it finishes, so a warp never waits for its longest path.
#include "kernels_common.cuh"

namespace {

#ifndef RTM_MEGA_RETRY
#define RTM_MEGA_RETRY 12   // lanes killed by the roulette in one round that trigger another round before the bounce block
#endif
// Sub-pixels a warp reserves per global atomic.  16, not round 1's 64: a warp keeps at most half a round of private
// backlog when the cursor runs dry (the persistent kernel's tail was 0.86 ms of a 45.9 ms step at 8 GPUs, SCALE_r01), and
// short frames gain (C3 166.0 -> 169.3 Grays/s, C4 unchanged: profiles/r02/mega_variants_ab2.jsonl).  A "guided" cursor
// (chunks shrinking near the end) was measured too: C3 +3.6 % but C4 -3.5 %, its refill code changes the hot loop's schedule.
#define RTM_MEGA_FETCH 16u
#ifndef RTM_MEGA_MINBLOCKS
#define RTM_MEGA_MINBLOCKS 6
#endif

// Per-lane constants of the sub-pixel a lane is working on, in SHARED memory ([k][thread]: conflict-free 16-byte
// accesses): a sample restart is three LDS.128 instead of eleven register moves, and the eleven registers are free
// (round 1 kept P, Pw, Pid, Pcpsn in registers: 80 registers, and the restart block cost 32 issue slots per loop
// iteration at 6 of 32 lanes — profiles/r01/ncu_megav6_c4_summary.txt).
struct SharedPrimary {
    float4 pos[RTM_MEGA_THREADS];   // primary hit point (relative), hit id (int bits)
    float4 nrm[RTM_MEGA_THREADS];   // oriented normal there, ACNE state c'/(2 tau r)
    float4 emi[RTM_MEGA_THREADS];   // emission of the primary hit, its roulette threshold (uint bits)
};

// ---- the radiance megakernel ---------------------------------------------------
// Persistent: grid = SMs x resident blocks.  A lane owns one sub-pixel (unit) at a time
// and walks its samples; one loop iteration = one roulette decision + one bounce ray.
// The lane state is "standing on a hit": origin o on sphere id with oriented normal w,
// throughput T, and radiance L that already includes the emission of this hit
// (Renderer.cpp:109 unrolled forward: L += T*e, T *= colorKD).  A path ends when the
// roulette kills it (Renderer.cpp:112), the bounce ray misses (:116), the depth cap is
// reached, or it lands on a surface with kd == 0 (the roulette there can never continue,
// so the sample is closed in the same iteration instead of spending another one).
// RNG: one rtm_bounce_rng call per hit (pcg3d hash, philox.cuh), counter = (sub-pixel, sample<<9 | depth) — the
// second word is ONE lane register `sd` (a hit adds 1, a sample restart rounds up to the next multiple of 512) —
// key from the seed: x -> roulette (32-bit compare against kd*2^32), y -> azimuth, the low bits of x and y -> height.
// PL: the scene also holds n_pl planes (extension); their records sit behind the spheres', ids n .. n+n_pl-1.
// NEE: next-event estimation (extension).  At every vertex that survives the roulette one light
// of the list (hot.full->lgeo / lemi) is chosen uniformly, a direction is drawn uniformly inside the cone the
// light sphere subtends (pdf 1 / (2 pi (1 - cos_max))), a shadow ray is traced by EVERY lane of the branch
// (lanes without a valid direction shoot along w and ignore the answer: no divergence around the trace) and
//     L += T * e_light * max(0, w.l) * 2 (1 - cos_max) * n_lights
// if it reaches the light.  A bounce ray that lands on a listed light from outside adds no emission (it was
// counted by the sampling at the previous vertex); from inside it does (no cone there).  Same expectation
// as the reference's estimator, ~8x less variance on the Cornell box.
// RT: roulette rounds (see the loop body) — chosen by the host for scenes whose visible surfaces kill most paths.
// CL: the per-sample clamp of Renderer.cpp:240-241 is evaluated.  The host clears it when the clamp provably never
// engages (every emission >= 0 and the largest radiance a path can carry, divided by spp, stays below 1 — true for
// every shipped scene); the lane then keeps ONE running sum over the sub-pixel's samples (L is not reset between
// samples) and multiplies by 1/(ss*ss*S) once at the end.
template <int N, bool ACNE, bool PL, bool NEE, bool RT, bool CL>
__global__ void __launch_bounds__(RTM_MEGA_THREADS, RTM_MEGA_MINBLOCKS)
radiance_mega_kernel(const __grid_constant__ RtmHotScene hot, const __grid_constant__ RtmPassParams p, float4* __restrict__ records,
                     RtmCounters* __restrict__ counters, int n_rt, const float* __restrict__ cps0, int n_pl) {
    __shared__ float4 sh_lgeo[NEE ? RTM_MAX_LIGHTS : 1], sh_lemi[NEE ? RTM_MAX_LIGHTS : 1];
    if (NEE && threadIdx.x < RTM_MAX_LIGHTS) {
        sh_lgeo[threadIdx.x] = hot.full->lgeo[threadIdx.x];
        sh_lemi[threadIdx.x] = hot.full->lemi[threadIdx.x];
    }
    unsigned long long shadow = 0;
    // N spheres (0 = runtime n_rt) with their records in shared memory
    __shared__ SharedScene sh;
    __shared__ SelfMask sh_mask;
    __shared__ SharedPrimary sh_prim;
    const int n = (N > 0) ? N : n_rt;
    init_selfmask(sh_mask);
    load_shared_scene(sh, hot.full, PL ? n + n_pl : n);

    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    const unsigned n_sub = (unsigned)p.n_sub;   // renderer_create guarantees n_sub < 2^32
    const float inv_norm = p.inv_norm;
    const unsigned sd_begin = (unsigned)p.s_begin << 9, sd_end = (unsigned)p.s_end << 9;
    const unsigned ns = (unsigned)(p.s_end - p.s_begin);
    const unsigned dcap = (unsigned)(p.max_depth - 1);

    unsigned w_next = 0, w_end = 0;  // warp-uniform slice of the work cursor
    bool have = false;
    bool wdone = false;              // warp-uniform: the work cursor is exhausted

    const unsigned sub_begin = (unsigned)p.sub_begin, sub_count = (unsigned)p.sub_count;
    const unsigned grid = (unsigned)(p.frame.ss * p.frame.ss);
    const bool full_grid = (sub_count == grid);
    // unit (sub-pixel) state: ql = record index in this handle's slice, q = global sub-pixel index
    unsigned q = 0, ql = 0;
    f3 acc = mk(0, 0, 0);            // CL only
    // path state
    f3 o = mk(0, 0, 0), w = mk(0, 0, 1), T = mk(1, 1, 1), L = mk(0, 0, 0);
    int id = 0;
    unsigned sd = 0;                 // sample << 9 | depth
    unsigned thr = 0;                // roulette threshold kd * 2^32 of the surface the lane stands on
    float cpsn = 0.f;                // ACNE: c' / (2 tau r) at the current origin

    unsigned long long rays_ref = 0, traced = 0;
    unsigned traced32 = 0;           // bounce rays of the current sub-pixel (<= samples * 511 < 2^32)
    const unsigned key_mask = p.key_mask;  // 0xFFFFFFF8 from the host: kept in a register for key_pack

    for (;;) {
        // ---- refill: lanes without a unit take the next sub-pixels of the warp's chunk
        if (__any_sync(0xffffffffu, !have)) {
            unsigned need = wdone ? 0u : __ballot_sync(0xffffffffu, !have);
            while (need) {
                if (w_next == w_end) {
                    unsigned long long start = 0;
                    if (lane == 0) start = atomicAdd(&counters->work_next, (unsigned long long)RTM_MEGA_FETCH);
                    start = __shfl_sync(0xffffffffu, start, 0);
                    if (start >= (unsigned long long)n_sub) {
                        wdone = true;
                        break;
                    }
                    w_next = (unsigned)start;
                    w_end = (start + RTM_MEGA_FETCH < (unsigned long long)n_sub) ? (unsigned)start + RTM_MEGA_FETCH : n_sub;
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
                    const bool miss = (wbits == RTM_REC_MISS);
                    const int Pid = miss ? 0 : (int)(wbits & RTM_REC_IDMASK);
                    RTM_ASSERT(ql < n_sub && (miss || Pid < n + (PL ? n_pl : 0)));
                    const float4 e0 = sh.emi[Pid];
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
                        const float4 g = sh.geo[Pid];
                        const float ir = (wbits & RTM_REC_FLIP) ? -g.w : g.w;
                        o = mk(rec.x, rec.y, rec.z);
                        w = mk((o.x - g.x) * ir, (o.y - g.y) * ir, (o.z - g.z) * ir);
                        if (PL && Pid >= n) w = (wbits & RTM_REC_FLIP) ? mk(-g.x, -g.y, -g.z) : mk(g.x, g.y, g.z);
                        cpsn = 0.f;
                        if (ACNE) cpsn = (wbits & RTM_REC_FLIP) ? 0.f : cps0[ql] * g.w * RTM_ACNE_INV2TAU;
                        sh_prim.pos[threadIdx.x] = make_float4(o.x, o.y, o.z, __int_as_float(Pid));
                        sh_prim.nrm[threadIdx.x] = make_float4(w.x, w.y, w.z, cpsn);
                        sh_prim.emi[threadIdx.x] = e0;
                        acc = mk(0, 0, 0);
                        sd = sd_begin;
                        id = Pid;
                        thr = __float_as_uint(e0.w);
                        T = mk(1, 1, 1); L = mk(e0.x, e0.y, e0.z);
                    }
                }
                const unsigned took = (__popc(need) < avail) ? __popc(need) : avail;
                w_next += took;
                need = __ballot_sync(0xffffffffu, !have && !mine);
            }
            if (!__any_sync(0xffffffffu, have)) {
                if (wdone) break;
                continue;
            }
        }

        // closes the sample (Renderer.cpp:240-242), then finishes the sub-pixel or restarts at the primary hit
        auto close_sample = [&]() {
            traced32 += sd & 511u;                 // every bounce ray of the path added 1 to the depth field (a miss too)
            if (CL) {
                acc.x += clamp01f(L.x * inv_norm);
                acc.y += clamp01f(L.y * inv_norm);
                acc.z += clamp01f(L.z * inv_norm);
            }
            sd = (sd | 511u) + 1u;                 // next sample, depth 0
            if (sd == sd_end) {
                records[ql] = CL ? make_float4(acc.x, acc.y, acc.z, 0.f) : make_float4(L.x * inv_norm, L.y * inv_norm, L.z * inv_norm, 0.f);
                rays_ref += ns;
                traced += traced32;
                traced32 = 0;
                have = false;
            } else {
                const float4 pp = sh_prim.pos[threadIdx.x], pn = sh_prim.nrm[threadIdx.x], pe = sh_prim.emi[threadIdx.x];
                o = mk(pp.x, pp.y, pp.z); id = __float_as_int(pp.w);
                w = mk(pn.x, pn.y, pn.z);
                if (ACNE) cpsn = pn.w;
                thr = __float_as_uint(pe.w);
                T = mk(1, 1, 1);
                if (CL) L = mk(pe.x, pe.y, pe.z);
                else L = mk(L.x + pe.x, L.y + pe.y, L.z + pe.z);   // running sum over the samples
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
        uint2 rnd = rtm_bounce_rng(make_uint2(q, sd), p.rk);
        bool go = have && (rnd.x < thr) && ((sd & 511u) < dcap);
        if (RT) {
            bool killed = have && !go;
            while (__popc(__ballot_sync(0xffffffffu, killed)) >= RTM_MEGA_RETRY) {   // warp-uniform
                if (killed) {
                    close_sample();
                    rnd = rtm_bounce_rng(make_uint2(q, sd), p.rk);
                    go = have && (rnd.x < thr) && ((sd & 511u) < dcap);
                    killed = have && !go;
                }
            }
        }
        // the common case closes a killed path's sample together with the paths the bounce block ends (one pass)
        bool end = have;
        if (go) {
            const float4 al = sh.alb[id];
            T.x *= al.x; T.y *= al.y; T.z *= al.z;   // colorKD, SettingData.h:11-13
            if (NEE) {
                const uint2 r2 = rtm_bounce_rng(make_uint2(q, sd), p.rk + 1);
                const unsigned nl = (unsigned)hot.n_lights;
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
                int ids = trace_f32<N>(hot, sh, sh_mask, n, (PL && id >= n) ? -1 : id, o, ls, ts, key_mask);
                if (PL) trace_planes_f32(hot, n, n_pl, id, o, ls, ts, ids);
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
                const float rr = fast_rcp(sh.geo[id].w);
                const float qq = -cpsn * rr * RTM_ACNE_2TAU, bq = rr * cosw;
                t = qq * fast_rcp(bq + fast_sqrt(fmaf(bq, bq, qq)));
                id2 = id;
            } else {
                // a ray leaving a plane has no origin sphere (the self mask only has rows for spheres)
                id2 = trace_f32<N>(hot, sh, sh_mask, n, (PL && id >= n) ? -1 : id, o, d, t, key_mask);
                if (PL) trace_planes_f32(hot, n, n_pl, id, o, d, t, id2);
            }
            ++sd;                                                                 // one more bounce ray (hit or miss)
            if (id2 >= 0) {
                RTM_ASSERT(id2 < n + (PL ? n_pl : 0) && (sd & 511u) != 0u);
                o = mk(fmaf(d.x, t, o.x), fmaf(d.y, t, o.y), fmaf(d.z, t, o.z));  // Renderer.cpp:79
                const float4 g = sh.geo[id2];
                f3 nn = mk((o.x - g.x) * g.w, (o.y - g.y) * g.w, (o.z - g.z) * g.w);
                if (PL && id2 >= n) nn = mk(g.x, g.y, g.z);
                const bool flip = !(dot(nn, d) < 0.f);                            // :82-83
                w = flip ? mk(-nn.x, -nn.y, -nn.z) : nn;
                id = id2;
                if (ACNE) cpsn = flip ? 0.f : t * t * ref_normalize_eps(rnd.x ^ (rnd.y * 0x9E3779B9u)) * g.w * RTM_ACNE_INV2TAU;
                const float4 e2 = sh.emi[id2];
                // NEE: a listed light reached from outside was already counted by the sampling above
                if (!NEE || flip || __float_as_int(sh.alb[id2].w) == 0) {
                    L.x = fmaf(T.x, e2.x, L.x); L.y = fmaf(T.y, e2.y, L.y); L.z = fmaf(T.z, e2.z, L.z);
                }
                thr = __float_as_uint(e2.w);
                end = (thr == 0u);   // kd == 0: the next roulette cannot continue
            }
            // else: miss, the recursion returns 0 (Renderer.cpp:116) and the sample ends
        }
        if (end) close_sample();
    }

    rays_ref += traced;
    traced += shadow;   // shadow rays are closest-hit queries too, but not rays of the reference's estimator
    unsigned long long tests = traced * (unsigned long long)n;
    for (int off = 16; off > 0; off >>= 1) {
        rays_ref += __shfl_down_sync(0xffffffffu, rays_ref, off);
        traced += __shfl_down_sync(0xffffffffu, traced, off);
        tests += __shfl_down_sync(0xffffffffu, tests, off);
    }
    if (lane == 0) {
        atomicAdd(&counters->rays_ref, rays_ref);
        atomicAdd(&counters->rays_traced, traced);
        atomicAdd(&counters->tests, tests);
    }
}

template <int N, bool ACNE, bool PL, bool NEE, bool RT, bool CL>
void launch_mega3(const RtmHotScene& hot, const RtmPassParams& p, float4* records, RtmCounters* counters, int n_rt, int n_pl, int sm_count,
                  cudaStream_t st, const float* cps0) {
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, radiance_mega_kernel<N, ACNE, PL, NEE, RT, CL>, RTM_MEGA_THREADS, 0);
    if (per_sm < 1) per_sm = 1;
    radiance_mega_kernel<N, ACNE, PL, NEE, RT, CL><<<sm_count * per_sm, RTM_MEGA_THREADS, 0, st>>>(hot, p, records, counters, n_rt, cps0, n_pl);
}
template <int N, bool ACNE, bool PL, bool NEE, bool RT>
void launch_mega2(const RtmHotScene& hot, const RtmPassParams& p, float4* records, RtmCounters* counters, int n_rt, int n_pl, bool clamp,
                  int sm_count, cudaStream_t st, const float* cps0) {
    // the no-clamp instantiation exists for the reference's own scene model and estimator (the headline kernels)
    if (!clamp && !PL && !NEE) launch_mega3<N, ACNE, PL, NEE, RT, false>(hot, p, records, counters, n_rt, n_pl, sm_count, st, cps0);
    else launch_mega3<N, ACNE, PL, NEE, RT, true>(hot, p, records, counters, n_rt, n_pl, sm_count, st, cps0);
}
// cps0 != null selects the ACNE instantiation (see ref_normalize_eps), n_pl > 0 the one with the plane loop
template <int N>
void launch_mega(const RtmHotScene& hot, const RtmPassParams& p, float4* records, RtmCounters* counters, int n_rt, int n_pl, bool nee,
                 bool retry, bool clamp, int sm_count, cudaStream_t st, const float* cps0) {
#define RTM_GO(ACNE, PL, NEE, RT) launch_mega2<N, ACNE, PL, NEE, RT>(hot, p, records, counters, n_rt, n_pl, clamp, sm_count, st, cps0)
    if (retry && !nee && n_pl == 0) {   // roulette rounds: instantiated for the reference's own scene model
        if (cps0) RTM_GO(true, false, false, true);
        else RTM_GO(false, false, false, true);
    } else if (nee) {
        if (n_pl > 0) {
            if (cps0) RTM_GO(true, true, true, false);
            else RTM_GO(false, true, true, false);
        } else {
            if (cps0) RTM_GO(true, false, true, false);
            else RTM_GO(false, false, true, false);
        }
    } else if (n_pl > 0) {
        if (cps0) RTM_GO(true, true, false, false);
        else RTM_GO(false, true, false, false);
    } else {
        if (cps0) RTM_GO(true, false, false, false);
        else RTM_GO(false, false, false, false);
    }
#undef RTM_GO
}

}  // namespace

void rtm_launch_radiance_mega(const RtmHotScene& hot, int32_t n, int32_t n_pl, bool nee, bool retry, bool clamp, const RtmPassParams& p,
                              float4* records, RtmCounters* counters, int32_t sm_count, const float* cps0, cudaStream_t st) {
    switch (n) {
        case 1: launch_mega<1>(hot, p, records, counters, n, n_pl, nee, retry, clamp, sm_count, st, cps0); break;
        case 2: launch_mega<2>(hot, p, records, counters, n, n_pl, nee, retry, clamp, sm_count, st, cps0); break;
        case 3: launch_mega<3>(hot, p, records, counters, n, n_pl, nee, retry, clamp, sm_count, st, cps0); break;
        case 4: launch_mega<4>(hot, p, records, counters, n, n_pl, nee, retry, clamp, sm_count, st, cps0); break;
        case 5: launch_mega<5>(hot, p, records, counters, n, n_pl, nee, retry, clamp, sm_count, st, cps0); break;
        case 6: launch_mega<6>(hot, p, records, counters, n, n_pl, nee, retry, clamp, sm_count, st, cps0); break;
        case 7: launch_mega<7>(hot, p, records, counters, n, n_pl, nee, retry, clamp, sm_count, st, cps0); break;
        case 8: launch_mega<8>(hot, p, records, counters, n, n_pl, nee, retry, clamp, sm_count, st, cps0); break;
        default: launch_mega<0>(hot, p, records, counters, n, n_pl, nee, retry, clamp, sm_count, st, cps0); break;
    }
}
