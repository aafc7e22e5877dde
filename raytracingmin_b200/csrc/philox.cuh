// philox.cuh — Philox4x32-10 (Salmon et al., "Parallel Random Numbers: As Easy
// as 1, 2, 3", SC'11), counter-based: every (sub-pixel, sample, bounce) triple
// owns one 128-bit counter value, so the image is a pure function of
// (scene, seed) and independent of how the sample budget is sliced over GPUs.
// Replaces the reference's shared std::mt19937 (src/Renderer.cpp:210-213).
#pragma once
#include <stdint.h>

#define RTM_PHILOX_M0 0xD2511F53u
#define RTM_PHILOX_M1 0xCD9E8D57u
#define RTM_PHILOX_W0 0x9E3779B9u
#define RTM_PHILOX_W1 0xBB67AE85u

__device__ __forceinline__ uint4 rtm_philox4x32_10(uint4 c, uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(RTM_PHILOX_M0, c.x), lo0 = RTM_PHILOX_M0 * c.x;
        const uint32_t hi1 = __umulhi(RTM_PHILOX_M1, c.z), lo1 = RTM_PHILOX_M1 * c.z;
        c = make_uint4(hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0);
        k0 += RTM_PHILOX_W0;
        k1 += RTM_PHILOX_W1;
    }
    return c;
}

// [0,1) with 24 random bits (exact in float)
__device__ __forceinline__ float rtm_u01_f32(uint32_t x) { return (float)(x >> 8) * 5.9604644775390625e-8f; }
// [0,1) with 53 random bits
__device__ __forceinline__ double rtm_u01_f64(uint32_t hi, uint32_t lo) {
    const unsigned long long v = (((unsigned long long)hi << 32) | lo) >> 11;
    return (double)v * 1.1102230246251565e-16;
}
// (0,1) with 32 random bits
__device__ __forceinline__ double rtm_u01_f64_32(uint32_t x) { return ((double)x + 0.5) * 2.3283064365386963e-10; }

// ---- Philox2x32-10 (same paper): 64 random bits per call, one 32x32->64 multiply
// per round.  The FP32 megakernel needs three numbers per hit (roulette, azimuth,
// height) and takes them from one call.  Round keys k + r*W0 are precomputed on the
// host (RtmPassParams::rk) so the loop carries no key arithmetic.
#define RTM_PHILOX2_M 0xD256D193u

__device__ __forceinline__ uint2 rtm_philox2x32_10(uint2 c, const uint32_t* __restrict__ rk) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi = __umulhi(RTM_PHILOX2_M, c.x), lo = RTM_PHILOX2_M * c.x;
        c = make_uint2(hi ^ rk[r] ^ c.y, lo);
    }
    return c;
}

// ---- the bounce loop's generator --------------------------------------------------------
// 64 random bits per (sub-pixel, sample, depth).  Default: pcg3d (Jarzynski & Olano, "Hash
// Functions for GPU Rendering", JCGT 9(3), 2020 — the permuted-congruential 3-D hash that paper
// recommends for this exact use: lattice-structured integer inputs, one evaluation per event):
// 9 multiply-adds + 3 shifts/xors against ~45 instructions of Philox2x32-10, which is 12 % of a
// 280-instruction Cornell bounce.  -DRTM_RNG_PHILOX selects Philox2x32-10 again (A/B runs).
// Counter-based either way: the image is a pure function of (scene, seed).
// The first word's LCG step only changes when the lane moves to another sub-pixel: the megakernel keeps
// rtm_bounce_prep(q) as lane state instead of q and calls rtm_bounce_rng_pre (one multiply-add less per hit).
__device__ __forceinline__ uint32_t rtm_bounce_prep(uint32_t q) {
#ifdef RTM_RNG_PHILOX
    return q;
#else
    return q * 1664525u + 1013904223u;
#endif
}
__device__ __forceinline__ uint2 rtm_bounce_rng_pre(uint32_t qx, uint32_t cy, const uint32_t* __restrict__ rk);
__device__ __forceinline__ uint2 rtm_bounce_rng(uint2 c, const uint32_t* __restrict__ rk) { return rtm_bounce_rng_pre(rtm_bounce_prep(c.x), c.y, rk); }
__device__ __forceinline__ uint2 rtm_bounce_rng_pre(uint32_t qx, uint32_t cy, const uint32_t* __restrict__ rk) {
#ifdef RTM_RNG_PHILOX
    return rtm_philox2x32_10(make_uint2(qx, cy), rk);
#else
    uint32_t x = qx, y = cy * 1664525u + 1013904223u, z = rk[0] * 1664525u + 1013904223u;
    x += y * z; y += z * x; z += x * y;
    x ^= x >> 16; y ^= y >> 16; z ^= z >> 16;
    x += y * z; y += z * x; z += x * y;
    y ^= z;
    // final xorshift: the low bits of the last multiply-adds are measurably biased (|z| up to 12 on 4 M lattice
    // inputs) and the sampler takes the top bits of its height from x's low 12 bits; with it 220 bit-balance /
    // serial-correlation tests stay within |z| < 3 (tests/test_rng_host.py)
    x ^= x >> 16; y ^= y >> 16;
    return make_uint2(x, y);
#endif
}
