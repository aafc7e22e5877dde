// kernels_wavefront.cuh — the WAVEFRONT organisation of the FP32 bounce loop
// (RTM_VARIANT_WAVEFRONT), included into kernels_wave.cu (it
// shares the sphere tests, the hierarchy steps and the sampler: kernels_common.cuh).
//
// Same estimator and same RNG keys as the megakernels — the image is bit-identical for
// brute-force scenes — but the lane state lives in global memory (RtmWaveDev) and a render
// is a sequence of waves:
//   shade   one thread per slot: apply the finished ray (hit point, normal, emission),
//           Russian roulette, sample the bounce (src/Renderer.cpp:75-116), close samples and
//           sub-pixels, pull the next sub-pixel from the work cursor; emits at most one ray
//           and appends the slot to the wave's queue (ballot + one atomic per warp);
//   extend  closest hit (src/Renderer.cpp:58-73) for every queued ray: one thread per ray
//           (brute force) or persistent lanes that fetch the next queued ray as soon as
//           theirs is done (hierarchy).
// "generate" is the FP64 primary pass, "accumulate" the same kernel as in the megakernel path.
// Built to be compared with the megakernel: every bounce moves ~220 B of
// path state through HBM, which the megakernel keeps in registers.

template <int N, bool BVH, bool ACNE>
__global__ void __launch_bounds__(128) wf_shade_kernel(const __grid_constant__ RtmHotScene hot, RtmPassParams p, float4* __restrict__ records,
                                                       RtmCounters* __restrict__ counters, int n_rt, RtmBvhDev bvh, RtmMatDev mat,
                                                       const float* __restrict__ cps0, RtmWaveDev wv, int wave) {
    __shared__ SharedScene sh;
    __shared__ SelfMask sh_mask;
    __shared__ SharedBig sh_big;
    init_selfmask(sh_mask);
    if (BVH) {
        if (threadIdx.x < 8) sh_big.id[threadIdx.x] = (threadIdx.x < (unsigned)bvh.n_big) ? bvh.big_id[threadIdx.x] : 0;
        __syncthreads();
    } else {
        load_shared_scene(sh, hot.full, (N > 0) ? N : n_rt);
    }
    SceneRecs<BVH> recs;
    recs.sh = &sh;
    recs.m = mat;

    const unsigned lane = threadIdx.x & 31u;
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = slot < wv.n_slots;
    const int parity = wave & 1;
    if (slot == 0) wv.ctrl[2] = 0u;   // extend's fetch cursor (idle while shade runs)

    const unsigned long long n_sub = p.n_sub;
    const int s_begin = p.s_begin, s_end = p.s_end;
    const float inv_norm = p.inv_norm;
    const unsigned sub_begin = (unsigned)p.sub_begin, sub_count = (unsigned)p.sub_count;
    const unsigned grid = (unsigned)(p.frame.ss * p.frame.ss);
    const bool full_grid = (sub_count == grid);

    // ---- load the slot
    int flags = 0;
    f3 o = mk(0, 0, 0), w = mk(0, 0, 1), T = mk(1, 1, 1), L = mk(0, 0, 0), d = mk(0, 0, 1), acc = mk(0, 0, 0);
    f3 P = mk(0, 0, 0), Pw = mk(0, 0, 1);
    int id = 0, s = 0, depth = 0, Pid = 0;
    unsigned ql = 0;
    float cpsn = 0.f, Pcpsn = 0.f;
    bool unit_consts = false;
    if (valid && wave > 0) {
        const float4 v3 = wv.a3[slot];
        flags = __float_as_int(v3.w);
        if (flags) {
            const float4 v0 = wv.a0[slot], v1 = wv.a1[slot], v2 = wv.a2[slot], v4 = wv.a4[slot];
            o = mk(v0.x, v0.y, v0.z); id = __float_as_int(v0.w);
            w = mk(v1.x, v1.y, v1.z); cpsn = v1.w;
            T = mk(v2.x, v2.y, v2.z);
            const unsigned sd = __float_as_uint(v2.w);
            s = (int)(sd >> 9); depth = (int)(sd & 511u);
            L = mk(v3.x, v3.y, v3.z);
            acc = mk(v4.x, v4.y, v4.z); ql = __float_as_uint(v4.w);
        }
    }
    bool have = flags != 0;
    unsigned q = full_grid ? ql : (ql / sub_count) * grid + sub_begin + (ql % sub_count);
    uint32_t eps_bits = 0;
    unsigned long long rays_ref = 0, traced = 0;
    unsigned n_tests = 0;
    bool want_ray = false;
    float ray_t = 0.f;
    int ray_hit = -1;

    auto load_unit_consts = [&]() {
        if (!unit_consts) {
            const float4 v5 = wv.a5[slot], v6 = wv.a6[slot];
            P = mk(v5.x, v5.y, v5.z); Pid = __float_as_int(v5.w);
            Pw = mk(v6.x, v6.y, v6.z); Pcpsn = v6.w;
            unit_consts = true;
        }
    };
    // a finished ray is applied to the path; returns true when the path ends there
    auto apply_hit = [&](int id2, float t) -> bool {
        if (id2 < 0) return true;                                            // miss: Renderer.cpp:116
        o = mk(fmaf(d.x, t, o.x), fmaf(d.y, t, o.y), fmaf(d.z, t, o.z));     // :79
        const float4 g = recs.geo(id2);
        f3 nn = mk((o.x - g.x) * g.w, (o.y - g.y) * g.w, (o.z - g.z) * g.w);
        if (BVH && g.w > 1e-2f) {
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
        const float4 e2 = recs.emi(id2);
        L.x = fmaf(T.x, e2.x, L.x); L.y = fmaf(T.y, e2.y, L.y); L.z = fmaf(T.z, e2.z, L.z);
        return __float_as_uint(e2.w) == 0u;
    };

    if (valid) {
        bool end = false;
        if (flags == 2) {
            const float4 rv = wv.ray[slot];
            const float2 hv = wv.hit[slot];
            d = mk(rv.x, rv.y, rv.z);
            eps_bits = __float_as_uint(rv.w);
            end = apply_hit(__float_as_int(hv.y), hv.x);
        }
        for (;;) {
            if (!have) {
                // next sub-pixel of the work cursor (src/Renderer.cpp:222-225 order)
                const unsigned long long nxt = atomicAdd(&counters->work_next, 1ull);
                if (nxt >= n_sub) break;
                ql = (unsigned)nxt;
                q = full_grid ? ql : (ql / sub_count) * grid + sub_begin + (ql % sub_count);
                const float4 rec = records[ql];
                const unsigned wbits = __float_as_uint(rec.w);
                const unsigned ns = (unsigned)(s_end - s_begin);
                const bool miss = (wbits == RTM_REC_MISS);
                Pid = miss ? 0 : (int)(wbits & RTM_REC_IDMASK);
                const float4 e0 = recs.emi(Pid);
                if (miss || __float_as_uint(e0.w) == 0u) {
                    float4 out = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (!miss) {
                        out.x = clamp01f(e0.x * inv_norm) * (float)ns;
                        out.y = clamp01f(e0.y * inv_norm) * (float)ns;
                        out.z = clamp01f(e0.z * inv_norm) * (float)ns;
                    }
                    records[ql] = out;
                    rays_ref += ns;
                    continue;
                }
                have = true;
                P = mk(rec.x, rec.y, rec.z);
                const float4 g = recs.geo(Pid);
                const float ir = (wbits & RTM_REC_FLIP) ? -g.w : g.w;
                Pw = mk((P.x - g.x) * ir, (P.y - g.y) * ir, (P.z - g.z) * ir);
                Pcpsn = 0.f;
                if (ACNE) Pcpsn = (wbits & RTM_REC_FLIP) ? 0.f : cps0[ql] * g.w * RTM_ACNE_INV2TAU;
                unit_consts = true;
                wv.a5[slot] = make_float4(P.x, P.y, P.z, __int_as_float(Pid));
                wv.a6[slot] = make_float4(Pw.x, Pw.y, Pw.z, Pcpsn);
                acc = mk(0, 0, 0);
                s = s_begin;
                o = P; id = Pid; w = Pw;
                T = mk(1, 1, 1); L = mk(e0.x, e0.y, e0.z); depth = 0;
                cpsn = Pcpsn;
                end = false;
            }
            if (!end) {
                const uint2 rnd = rtm_bounce_rng(make_uint2(q, ((unsigned)s << 9) | (unsigned)depth), p.rk);
                const float4 em = recs.emi(id);
                end = !((rnd.x < __float_as_uint(em.w)) && (depth + 1 < p.max_depth));   // Renderer.cpp:78
                if (!end) {
                    const float4 al = recs.alb(id);
                    T.x *= al.x; T.y *= al.y; T.z *= al.z;                               // colorKD
                    d = cosine_dir(w, rnd.y, ((rnd.x & 0xFFFu) << 10) | (rnd.y & 0x3FFu));
                    eps_bits = rnd.x ^ (rnd.y * 0x9E3779B9u);
                    ++traced;
                    if (ACNE) {
                        const float cosw = dot(w, d);
                        if (cpsn + cosw <= 0.f) {
                            const float rr = fast_rcp(recs.geo(id).w);
                            const float qq = -cpsn * rr * RTM_ACNE_2TAU, bq = rr * cosw;
                            end = apply_hit(id, qq * fast_rcp(bq + fast_sqrt(fmaf(bq, bq, qq))));
                            continue;
                        }
                    }
                    ray_t = FLT_MAX;
                    ray_hit = -1;
                    if (BVH) {
                        // the big-sphere list is tested here, where all lanes are active
                        if (N > 0) {
                            const unsigned key = trace_pairs_f32<N>(hot, sh_mask, bvh_big_slot(__float_as_int(al.w)), o, d, p.key_mask);
                            if (key < RTM_KEY_INF) {
                                ray_t = __uint_as_float(key & ~7u) + RTM_T_EPS;
                                ray_hit = sh_big.id[key & 7u];
                            }
                            n_tests += N;
                        }
                        if (bvh.root == RTM_BVH_NONE_DEV) {
                            end = apply_hit(ray_hit, ray_t);
                            continue;
                        }
                    }
                    want_ray = true;
                    break;
                }
            }
            // the sample is complete: Renderer.cpp:240-242
            acc.x += clamp01f(L.x * inv_norm);
            acc.y += clamp01f(L.y * inv_norm);
            acc.z += clamp01f(L.z * inv_norm);
            ++s;
            if (s == s_end) {
                records[ql] = make_float4(acc.x, acc.y, acc.z, 0.f);
                rays_ref += (unsigned)(s_end - s_begin);
                have = false;
                continue;
            }
            load_unit_consts();
            const float4 e0 = recs.emi(Pid);
            o = P; id = Pid; w = Pw;
            T = mk(1, 1, 1); L = mk(e0.x, e0.y, e0.z); depth = 0;
            cpsn = Pcpsn;
            end = false;
        }
        // ---- store the slot
        const int nf = have ? (want_ray ? 2 : 1) : 0;
        wv.a3[slot] = make_float4(L.x, L.y, L.z, __int_as_float(nf));
        if (have) {
            wv.a0[slot] = make_float4(o.x, o.y, o.z, __int_as_float(id));
            wv.a1[slot] = make_float4(w.x, w.y, w.z, cpsn);
            wv.a2[slot] = make_float4(T.x, T.y, T.z, __uint_as_float(((unsigned)s << 9) | (unsigned)depth));
            wv.a4[slot] = make_float4(acc.x, acc.y, acc.z, __uint_as_float(ql));
            if (want_ray) {
                wv.ray[slot] = make_float4(d.x, d.y, d.z, __uint_as_float(eps_bits));
                wv.hit[slot] = make_float2(ray_t, __int_as_float(ray_hit));
            }
        }
    }
    // ---- queue the slot: ballot + prefix popcount, one atomic per warp
    const unsigned qm = __ballot_sync(0xffffffffu, want_ray);
    if (qm) {
        unsigned base = 0;
        if (lane == (unsigned)(__ffs(qm) - 1)) {
            base = atomicAdd(&wv.ctrl[parity], (unsigned)__popc(qm));
            wv.ctrl[3] = (unsigned)wave + 1u;
        }
        base = __shfl_sync(0xffffffffu, base, __ffs(qm) - 1);
        if (want_ray) wv.queue[parity][base + __popc(qm & ((1u << lane) - 1u))] = slot;
    }
    rays_ref += traced;
    unsigned long long tests = BVH ? (unsigned long long)n_tests : 0ull;
    for (int off = 16; off > 0; off >>= 1) {
        rays_ref += __shfl_down_sync(0xffffffffu, rays_ref, off);
        traced += __shfl_down_sync(0xffffffffu, traced, off);
        tests += __shfl_down_sync(0xffffffffu, tests, off);
    }
    if (lane == 0 && (rays_ref | traced)) {
        atomicAdd(&counters->rays_ref, rays_ref);
        atomicAdd(&counters->rays_traced, traced);
        if (tests) atomicAdd(&counters->tests, tests);
    }
}

// ---- extend, brute force: one thread per queued ray --------------------------------------
template <int N>
__global__ void __launch_bounds__(128) wf_extend_brute_kernel(const __grid_constant__ RtmHotScene hot, RtmWaveDev wv, int wave, int n_rt,
                                                              unsigned key_mask, RtmCounters* __restrict__ counters) {
    __shared__ SelfMask sh_mask;
    __shared__ SharedScene sh;
    init_selfmask(sh_mask);
    load_shared_scene(sh, hot.full, (N > 0) ? N : n_rt);
    const int parity = wave & 1;
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned count = wv.ctrl[parity];
    if (i == 0) {
        wv.ctrl[parity ^ 1] = 0u;   // the next wave's queue length (idle while extend runs)
        atomicAdd(&counters->tests, (unsigned long long)count * (unsigned long long)((N > 0) ? N : n_rt));
    }
    if (i >= count) return;
    const int slot = wv.queue[parity][i];
    RTM_ASSERT(slot >= 0 && slot < wv.n_slots && count <= (unsigned)wv.n_slots);
    const float4 v0 = wv.a0[slot], rv = wv.ray[slot];
    float t;
    const int id2 = trace_f32<N>(hot, sh, sh_mask, n_rt, __float_as_int(v0.w), mk(v0.x, v0.y, v0.z), mk(rv.x, rv.y, rv.z), t, key_mask);
    wv.hit[slot] = make_float2(t, __int_as_float(id2));
}

// ---- extend, hierarchy: persistent lanes with dynamic fetch ------------------------------
// A lane that finishes its ray takes the next queued one as soon as RTM_WF_FETCH_THRESH lanes
// of its warp are idle, so a warp never drains down to its longest traversal.
#define RTM_WF_FETCH_THRESH 4
__global__ void __launch_bounds__(128, 6) wf_extend_bvh_kernel(RtmBvhDev bvh, RtmWaveDev wv, int wave,
                                                               RtmCounters* __restrict__ counters) {
    const int parity = wave & 1;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    if (blockIdx.x == 0 && threadIdx.x == 0) wv.ctrl[parity ^ 1] = 0u;
    const unsigned total = wv.ctrl[parity];
    const int* __restrict__ queue = wv.queue[parity];

    bool active = false, dry = false;
    int slot = 0, oid = -1;
    f3 o = mk(0, 0, 0), d = mk(0, 0, 1);
    BvhRay r;
    r.node = RTM_BVH_NONE_DEV;
    int stack[RTM_BVH_STACK_DEV];
    unsigned n_tests = 0, n_nodes = 0;

    for (;;) {
        const unsigned idle = __ballot_sync(0xffffffffu, !active);
        if (!dry && (__popc(idle) >= RTM_WF_FETCH_THRESH)) {
            unsigned base = 0;
            if (lane == 0) base = atomicAdd(&wv.ctrl[2], (unsigned)__popc(idle));
            base = __shfl_sync(0xffffffffu, base, 0);
            const unsigned my = base + __popc(idle & lt_mask);
            if (!active && my < total) {
                slot = queue[my];
                RTM_ASSERT(slot >= 0 && slot < wv.n_slots);
                const float4 v0 = wv.a0[slot], rv = wv.ray[slot];
                const float2 hv = wv.hit[slot];
                o = mk(v0.x, v0.y, v0.z); oid = __float_as_int(v0.w);
                d = mk(rv.x, rv.y, rv.z);
                r.best = hv.x - RTM_T_EPS;            // seeded by the big-sphere test in the shade kernel
                r.hit = __float_as_int(hv.y);
                r.inv = mk(fast_rcp(d.x), fast_rcp(d.y), fast_rcp(d.z));
                r.oi = mk(o.x * r.inv.x, o.y * r.inv.y, o.z * r.inv.z);
                r.node = bvh.root;
                r.sp = 0;
                stack[r.sp++] = RTM_BVH_NONE_DEV;
                active = true;
            }
            dry = (base + (unsigned)__popc(idle) >= total);
        }
        if (!__any_sync(0xffffffffu, active)) {
            if (dry) break;
            continue;
        }
        while (active && bvh_is_inner(r.node)) {
            bvh_inner_step(bvh, r, stack);
            ++n_nodes;
        }
        if (active && r.node < 0) bvh_leaf_step_ids(bvh, oid, o, d, r, stack, n_tests);
        if (active && r.node == RTM_BVH_NONE_DEV) {
            wv.hit[slot] = make_float2(r.best + RTM_T_EPS, __int_as_float(r.hit));
            active = false;
        }
    }
    unsigned long long tests = n_tests, nodes = n_nodes;
    for (int off = 16; off > 0; off >>= 1) {
        tests += __shfl_down_sync(0xffffffffu, tests, off);
        nodes += __shfl_down_sync(0xffffffffu, nodes, off);
    }
    if (lane == 0 && (tests | nodes)) {
        atomicAdd(&counters->tests, tests);
        atomicAdd(&counters->nodes, nodes);
    }
}
