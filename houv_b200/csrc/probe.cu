// FP32 CUDA-core peak probe: the roofline denominator for K1/K3 (SURVEY.md
// section 8(d): "FP32_peak = measured FFMA-chain peak on the box").  Independent
// FFMA chains, no memory traffic; scalar (FFMA) or packed (FFMA2) flavour.
#include "common.cuh"

namespace houv {

constexpr int kProbeThreads = 256;
constexpr int kProbeCtas = 148 * 8;
constexpr int kChains = 16;  // independent accumulators per thread (latency 4 x 4 warps/SMSP)

template <bool PACKED>
__global__ void __launch_bounds__(kProbeThreads) probe_kernel(float *out, int iters) {
    float s = 1.0f + 1e-7f * threadIdx.x, c = 1e-9f * blockIdx.x;
    if (PACKED) {
        float2 acc[kChains / 2];
#pragma unroll
        for (int i = 0; i < kChains / 2; ++i) acc[i] = make_float2(i, -i);
        float2 m = make_float2(s, s), a = make_float2(c, c);
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int u = 0; u < 8; ++u)
#pragma unroll
                for (int i = 0; i < kChains / 2; ++i) acc[i] = __ffma2_rn(acc[i], m, a);
        }
        float r = 0;
#pragma unroll
        for (int i = 0; i < kChains / 2; ++i) r += acc[i].x + acc[i].y;
        if (r == 12345.678f) out[0] = r;  // keep the chains live
    } else {
        float acc[kChains];
#pragma unroll
        for (int i = 0; i < kChains; ++i) acc[i] = i;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int i = 0; i < kChains; ++i) acc[i] = __fmaf_rn(acc[i], s, c);
        }
        float r = 0;
#pragma unroll
        for (int i = 0; i < kChains; ++i) r += acc[i];
        if (r == 12345.678f) out[0] = r;
    }
}

}  // namespace houv

using namespace houv;

extern "C" const char *houv_version(void) { return "houv_b200 0.1 sm_100a"; }

extern "C" int houv_fp32_probe(float *scratch, int iters, int packed, double *flop_host,
                               void *stream) {
    if (!scratch || iters < 0) return HOUV_E_BADARG;
    if (packed)
        probe_kernel<true><<<kProbeCtas, kProbeThreads, 0, (cudaStream_t)stream>>>(scratch, iters);
    else
        probe_kernel<false><<<kProbeCtas, kProbeThreads, 0, (cudaStream_t)stream>>>(scratch, iters);
    // per thread and iteration: 64 FFMA (scalar) or 64 FFMA2 = 128 lane-FMAs (packed)
    if (flop_host)
        *flop_host = 2.0 * (packed ? 128.0 : 64.0) * (double)iters * kProbeThreads * kProbeCtas;
    return check_launch();
}
