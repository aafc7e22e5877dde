// kernels_wave.cu — the wavefront organisation of the FP32 bounce loop (RTM_VARIANT_WAVEFRONT): built to be compared
// with the megakernels, not the product path.
#include "kernels_common.cuh"

namespace {
#include "kernels_wavefront.cuh"

template <int N, bool BVH>
void launch_wf_shade(const RtmHotScene& hot, const RtmBvhDev& bvh, const RtmMatDev& mat, const RtmPassParams& p, float4* records, RtmCounters* counters,
                     int n_rt, const float* cps0, const RtmWaveDev& wv, int wave, cudaStream_t st) {
    const unsigned g = (unsigned)((wv.n_slots + 127) / 128);
    if (cps0) wf_shade_kernel<N, BVH, true><<<g, 128, 0, st>>>(hot, p, records, counters, n_rt, bvh, mat, cps0, wv, wave);
    else wf_shade_kernel<N, BVH, false><<<g, 128, 0, st>>>(hot, p, records, counters, n_rt, bvh, mat, nullptr, wv, wave);
}
}  // namespace

void rtm_launch_wf_shade(const RtmHotScene& hot, int32_t n, bool use_bvh, const RtmBvhDev& bvh, const RtmMatDev& mat, const RtmPassParams& p,
                         float4* records, RtmCounters* counters, const float* cps0, const RtmWaveDev& wv, int32_t wave,
                         cudaStream_t st) {
    if (use_bvh) {
        switch ((bvh.n_big + 1) / 2) {
            case 0: launch_wf_shade<0, true>(hot, bvh, mat, p, records, counters, 0, cps0, wv, wave, st); break;
            case 1: launch_wf_shade<2, true>(hot, bvh, mat, p, records, counters, 0, cps0, wv, wave, st); break;
            case 2: launch_wf_shade<4, true>(hot, bvh, mat, p, records, counters, 0, cps0, wv, wave, st); break;
            case 3: launch_wf_shade<6, true>(hot, bvh, mat, p, records, counters, 0, cps0, wv, wave, st); break;
            default: launch_wf_shade<8, true>(hot, bvh, mat, p, records, counters, 0, cps0, wv, wave, st); break;
        }
        return;
    }
    // brute force: the shade kernel only needs the per-sphere records (shared memory, any n <= 256)
    launch_wf_shade<0, false>(hot, bvh, mat, p, records, counters, n, cps0, wv, wave, st);
}

void rtm_launch_wf_extend(const RtmHotScene& hot, int32_t n, bool use_bvh, const RtmBvhDev& bvh, const RtmWaveDev& wv, int32_t wave, RtmCounters* counters,
                          int32_t sm_count, cudaStream_t st) {
    if (use_bvh) {
        int per_sm = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, wf_extend_bvh_kernel, 128, 0);
        if (per_sm < 1) per_sm = 1;
        wf_extend_bvh_kernel<<<sm_count * per_sm, 128, 0, st>>>(bvh, wv, wave, counters);
        return;
    }
    const unsigned g = (unsigned)((wv.n_slots + 127) / 128);
    switch (n) {
#define RTM_CASE(K) case K: wf_extend_brute_kernel<K><<<g, 128, 0, st>>>(hot, wv, wave, n, 0xFFFFFFF8u, counters); break
        RTM_CASE(1); RTM_CASE(2); RTM_CASE(3); RTM_CASE(4); RTM_CASE(5); RTM_CASE(6); RTM_CASE(7); RTM_CASE(8);
#undef RTM_CASE
        default: wf_extend_brute_kernel<0><<<g, 128, 0, st>>>(hot, wv, wave, n, 0xFFFFFFF8u, counters); break;
    }
}
