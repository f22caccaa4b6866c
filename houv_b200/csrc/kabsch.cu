// K4 -- warp-per-pair Kabsch for sm_100a.
//
// Replaces SVDHead.forward (/root/reference/registration/model_utils.py:220-255):
// centre both clouds, H = (src_c * w) corr_c^T, U S V^T = svd(H), R = V U^T with
// the det<0 reflection fix on V's last column, t = -R mean(src) + mean(corr)
// (weighted sums when weights are given, :251-254).  The reference loops over
// the batch in Python with one torch.svd + torch.det + host-synchronising `if`
// per sample; here one warp owns one pair: two coalesced passes over the
// [3,N] rows (means, then the 3x3 cross-covariance by warp-shuffle reduction)
// and a closed 3x3 solve on lane 0 (svd3.cuh).  Latency-bound: 24*N bytes per
// pair, a few hundred flops.
#include "common.cuh"
#include "svd3.cuh"

namespace houv {

template <int NV>
__device__ __forceinline__ void warp_sum(double (&v)[NV]) {
#pragma unroll
    for (int i = 0; i < NV; ++i)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
}

__global__ void __launch_bounds__(128) kabsch_kernel(const float *__restrict__ src,
                                                     const float *__restrict__ corr,
                                                     const float *__restrict__ weights, int B,
                                                     int N, float *__restrict__ Rout,
                                                     float *__restrict__ tout) {
    const int lane = threadIdx.x & 31;
    const int warps_per_cta = blockDim.x >> 5;
    for (int b = blockIdx.x * warps_per_cta + (threadIdx.x >> 5); b < B;
         b += gridDim.x * warps_per_cta) {
        const float *X = src + (size_t)b * 3 * N;   // rows x,y,z of length N
        const float *Y = corr + (size_t)b * 3 * N;
        const float *W = weights ? weights + (size_t)b * N : nullptr;
        // pass 1: plain means (always unweighted, :221-222) and weighted sums (:254)
        double s[12];
#pragma unroll
        for (int i = 0; i < 12; ++i) s[i] = 0.0;
        for (int n = lane; n < N; n += 32) {
            const float w = W ? W[n] : 1.f;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float x = X[c * N + n], y = Y[c * N + n];
                s[c] += x;
                s[3 + c] += y;
                s[6 + c] += (double)(w * x);
                s[9 + c] += (double)(w * y);
            }
        }
        warp_sum<12>(s);
        float mx[3], my[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            mx[c] = (float)(s[c] / N);
            my[c] = (float)(s[3 + c] / N);
        }
        // pass 2: H[i][j] = sum_n (x_i - mx_i) w (y_j - my_j)
        double h[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) h[i] = 0.0;
        for (int n = lane; n < N; n += 32) {
            const float w = W ? W[n] : 1.f;
            float xc[3], yc[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                xc[c] = (X[c * N + n] - mx[c]) * w;
                yc[c] = Y[c * N + n] - my[c];
            }
#pragma unroll
            for (int i = 0; i < 3; ++i)
#pragma unroll
                for (int j = 0; j < 3; ++j) h[i * 3 + j] += (double)(xc[i] * yc[j]);
        }
        warp_sum<9>(h);
        if (lane == 0) {
            double H[3][3], R[3][3];
#pragma unroll
            for (int i = 0; i < 3; ++i)
#pragma unroll
                for (int j = 0; j < 3; ++j) H[i][j] = h[i * 3 + j];
            kabsch_from_covariance(H, R);
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                double t;
                if (W)
                    t = -(R[i][0] * s[6] + R[i][1] * s[7] + R[i][2] * s[8]) + s[9 + i];
                else
                    t = -(R[i][0] * (double)mx[0] + R[i][1] * (double)mx[1] + R[i][2] * (double)mx[2]) +
                        (double)my[i];
                tout[(size_t)b * 3 + i] = (float)t;
#pragma unroll
                for (int j = 0; j < 3; ++j) Rout[(size_t)b * 9 + i * 3 + j] = (float)R[i][j];
            }
        }
    }
}

}  // namespace houv

using namespace houv;

extern "C" int houv_kabsch(const float *src, const float *src_corr, const float *weights, int B,
                           int N, float *R, float *t, void *stream) {
    if (B < 0 || N < 0) return HOUV_E_BADARG;
    if (B == 0) return 0;
    if (!src || !src_corr || !R || !t || N == 0) return HOUV_E_BADARG;
    const int warps = 4;
    int blocks = (B + warps - 1) / warps;
    if (blocks > 148 * 16) blocks = 148 * 16;
    kabsch_kernel<<<blocks, warps * 32, 0, (cudaStream_t)stream>>>(src, src_corr, weights, B, N, R, t);
    return check_launch();
}
