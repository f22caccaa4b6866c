// Epilogue of calc_cd (/root/reference/registration/model_utils_completion.py:69-80 =
// completion/model_utils.py:69-80) and fscore (utils/metrics/CD/fscore.py:3-16): from the two
// distance lists of a Chamfer forward, per batch element
//     cd_p = (mean sqrt(dist1) + mean sqrt(dist2)) / 2,   cd_t = mean dist1 + mean dist2,
//     f1 = 2 p1 p2 / (p1 + p2) with p = mean(dist < threshold), 0 where that is NaN
// in ONE launch (one CTA per batch element) instead of ~12 torch reductions.  Forward only: a loss
// that needs gradients keeps the torch expression over chamfer_3DDist (houv_b200/loss.py).
// HBM-bound: reads 4*(N+M) bytes per batch element once.
#include "common.cuh"

namespace houv {

__global__ void __launch_bounds__(256) cd_reduce_kernel(const float *dist1, const float *dist2,
                                                        int N, int M, float threshold, float *cd_p,
                                                        float *cd_t, float *f1) {
    const int b = blockIdx.x, t = threadIdx.x;
    // per-thread partial sums in double: the result is then independent of the thread count to
    // well below float rounding (torch's own mean uses a different summation order anyway)
    double s[6] = {0, 0, 0, 0, 0, 0};  // sum sqrt d1, sum d1, count d1<th, then the same for d2
    const float *d1 = dist1 + (size_t)b * N, *d2 = dist2 + (size_t)b * M;
    for (int i = t; i < N; i += 256) {
        const float d = d1[i];
        s[0] += (double)sqrtf(d); s[1] += (double)d; s[2] += d < threshold ? 1.0 : 0.0;
    }
    for (int i = t; i < M; i += 256) {
        const float d = d2[i];
        s[3] += (double)sqrtf(d); s[4] += (double)d; s[5] += d < threshold ? 1.0 : 0.0;
    }
    __shared__ double red[8][6];
#pragma unroll
    for (int k = 0; k < 6; ++k)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s[k] += __shfl_xor_sync(0xffffffffu, s[k], o);
    if ((t & 31) == 0)
#pragma unroll
        for (int k = 0; k < 6; ++k) red[t >> 5][k] = s[k];
    __syncthreads();
    if (t == 0) {
        double r[6];
        for (int k = 0; k < 6; ++k) {
            r[k] = 0;
            for (int w = 0; w < 8; ++w) r[k] += red[w][k];
        }
        const float m1s = (float)(r[0] / N), m2s = (float)(r[3] / M);
        const float m1 = (float)(r[1] / N), m2 = (float)(r[4] / M);
        cd_p[b] = (m1s + m2s) / 2.f;
        cd_t[b] = m1 + m2;
        if (f1) {
            const float p1 = (float)(r[2] / N), p2 = (float)(r[5] / M);
            const float f = 2.f * p1 * p2 / (p1 + p2);
            f1[b] = f != f ? 0.f : f;  // fscore[isnan] = 0
        }
    }
}

}  // namespace houv

extern "C" int houv_cd_reduce(const float *dist1, const float *dist2, int B, int N, int M,
                              float threshold, float *cd_p, float *cd_t, float *f1, void *stream) {
    if (B < 0 || N <= 0 || M <= 0) return HOUV_E_BADARG;
    if (B == 0) return 0;
    if (!dist1 || !dist2 || !cd_p || !cd_t) return HOUV_E_BADARG;
    houv::cd_reduce_kernel<<<B, 256, 0, (cudaStream_t)stream>>>(dist1, dist2, N, M, threshold, cd_p,
                                                               cd_t, f1);
    return houv::check_launch();
}
