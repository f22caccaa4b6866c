// K6 -- one refinement step of IDAM's iterative weighted Kabsch, fused, for sm_100a.
//
// Replaces the tail of every iteration of IDAM.forward
// (/root/reference/registration/models/idam.py:317-342):
//   src_corr = tgt[:, :, corr_idx]                                  (:318)
//   w = sigmoid(logits); w *= (w >= median(w)); w /= (sum(w)+1e-8)   (:329-331, hybrid point elimination)
//   R_ab, t_ab = SVDHead(src, src_corr, w)                          (:335, idam.py:152-189)
//   src = R_ab src + t_ab;  R = R_ab R;  t = R_ab t + t_ab          (:338-340)
// which the reference runs as ~15 torch launches plus a per-sample Python loop of torch.svd with a
// host-synchronising `if`.  One CTA owns one pair; everything stays in shared memory / registers.
// The correspondence indices and weight logits come from the learned part of IDAM (similarity-matrix
// convolutions), which is outside this library.  Latency-bound: 40*N bytes per pair.
//
// torch.median of an even-sized row is the LOWER middle value, i.e. the element of rank (N-1)/2.
#include "common.cuh"
#include "svd3.cuh"

namespace houv {

constexpr int kStepThreads = 128;
constexpr int kStepMaxN = 8192;  // weights + the 6 coordinate rows stay in shared memory

template <int NV>
__device__ __forceinline__ void cta_sum(double (&v)[NV], double *scratch) {
#pragma unroll
    for (int i = 0; i < NV; ++i)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();  // scratch may still be read from the previous call
    if (lane == 0)
#pragma unroll
        for (int i = 0; i < NV; ++i) scratch[warp * NV + i] = v[i];
    __syncthreads();
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        double s = 0;
#pragma unroll
        for (int w = 0; w < kStepThreads / 32; ++w) s += scratch[w * NV + i];
        v[i] = s;
    }
}

struct StepArgs {
    const float *src, *tgt, *logits;
    const int *corr32;
    const long long *corr64;
    float *src_out, *R_acc, *t_acc, *R_ab, *t_ab, *w_out;
    int B, N, M;
};

__global__ void __launch_bounds__(kStepThreads) idam_step_kernel(const StepArgs A) {
    extern __shared__ __align__(16) float sm_f[];
    const int N = A.N, M = A.M;
    float *w = sm_f;            // [N]
    float *X = w + N;           // [3][N] src
    float *Y = X + 3 * N;       // [3][N] gathered correspondences
    __shared__ double scratch[(kStepThreads / 32) * 12];
    __shared__ float s_med, s_Rt[12];
    for (int b = blockIdx.x; b < A.B; b += gridDim.x) {
        const float *src = A.src + (size_t)b * 3 * N;
        const float *tgt = A.tgt + (size_t)b * 3 * M;
        __syncthreads();
        for (int n = threadIdx.x; n < N; n += kStepThreads) {
            const long long j = A.corr64 ? A.corr64[(size_t)b * N + n] : A.corr32[(size_t)b * N + n];
            w[n] = 1.f / (1.f + expf(-A.logits[(size_t)b * N + n]));
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                X[c * N + n] = src[c * N + n];
                Y[c * N + n] = tgt[c * M + j];
            }
        }
        __syncthreads();
        // median by rank counting: the value v with #(x < v) <= k < #(x <= v), k = (N-1)/2
        const int k = (N - 1) >> 1;
        for (int i = threadIdx.x; i < N; i += kStepThreads) {
            const float v = w[i];
            int less = 0, leq = 0;
            for (int j = 0; j < N; ++j) {
                const float u = w[j];
                less += u < v;
                leq += u <= v;
            }
            if (less <= k && k < leq) s_med = v;  // every writer holds the same value
        }
        __syncthreads();
        const float med = s_med;
        double s[12];
#pragma unroll
        for (int i = 0; i < 12; ++i) s[i] = 0.0;
        float wsum = 0.f;
        for (int n = threadIdx.x; n < N; n += kStepThreads) {
            const float v = w[n] >= med ? w[n] : 0.f;
            w[n] = v;
            wsum += v;
        }
        double ws[1] = {(double)wsum};
        cta_sum<1>(ws, scratch);
        const float inv = 1.f / ((float)ws[0] + 1e-8f);
        // pass 1: unweighted means (idam.py:160-161) and weighted sums (:188)
        for (int n = threadIdx.x; n < N; n += kStepThreads) {
            const float wn = w[n] * inv;
            w[n] = wn;
            if (A.w_out) A.w_out[(size_t)b * N + n] = wn;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float x = X[c * N + n], y = Y[c * N + n];
                s[c] += x;
                s[3 + c] += y;
                s[6 + c] += (double)(wn * x);
                s[9 + c] += (double)(wn * y);
            }
        }
        cta_sum<12>(s, scratch);
        float mx[3], my[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            mx[c] = (float)(s[c] / N);
            my[c] = (float)(s[3 + c] / N);
        }
        double h[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) h[i] = 0.0;
        for (int n = threadIdx.x; n < N; n += kStepThreads) {
            const float wn = w[n];
            float xc[3], yc[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                xc[c] = (X[c * N + n] - mx[c]) * wn;
                yc[c] = Y[c * N + n] - my[c];
            }
#pragma unroll
            for (int i = 0; i < 3; ++i)
#pragma unroll
                for (int j = 0; j < 3; ++j) h[i * 3 + j] += (double)(xc[i] * yc[j]);
        }
        cta_sum<9>(h, scratch);
        if (threadIdx.x == 0) {
            double H[3][3], R[3][3];
#pragma unroll
            for (int i = 0; i < 3; ++i)
#pragma unroll
                for (int j = 0; j < 3; ++j) H[i][j] = h[i * 3 + j];
            kabsch_from_covariance(H, R);
            float Rf[9], tf[3];
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                tf[i] = (float)(-(R[i][0] * s[6] + R[i][1] * s[7] + R[i][2] * s[8]) + s[9 + i]);
#pragma unroll
                for (int j = 0; j < 3; ++j) Rf[i * 3 + j] = (float)R[i][j];
            }
#pragma unroll
            for (int i = 0; i < 9; ++i) s_Rt[i] = Rf[i];
#pragma unroll
            for (int i = 0; i < 3; ++i) s_Rt[9 + i] = tf[i];
            if (A.R_ab)
#pragma unroll
                for (int i = 0; i < 9; ++i) A.R_ab[(size_t)b * 9 + i] = Rf[i];
            if (A.t_ab)
#pragma unroll
                for (int i = 0; i < 3; ++i) A.t_ab[(size_t)b * 3 + i] = tf[i];
            // R = R_ab R, t = R_ab t + t_ab  (idam.py:339-340), float32 like the reference
            if (A.R_acc && A.t_acc) {
                float Ro[9], to[3], Rn[9], tn[3];
#pragma unroll
                for (int i = 0; i < 9; ++i) Ro[i] = A.R_acc[(size_t)b * 9 + i];
#pragma unroll
                for (int i = 0; i < 3; ++i) to[i] = A.t_acc[(size_t)b * 3 + i];
#pragma unroll
                for (int i = 0; i < 3; ++i) {
#pragma unroll
                    for (int j = 0; j < 3; ++j)
                        Rn[i * 3 + j] = Rf[i * 3] * Ro[j] + Rf[i * 3 + 1] * Ro[3 + j] + Rf[i * 3 + 2] * Ro[6 + j];
                    tn[i] = Rf[i * 3] * to[0] + Rf[i * 3 + 1] * to[1] + Rf[i * 3 + 2] * to[2] + tf[i];
                }
#pragma unroll
                for (int i = 0; i < 9; ++i) A.R_acc[(size_t)b * 9 + i] = Rn[i];
#pragma unroll
                for (int i = 0; i < 3; ++i) A.t_acc[(size_t)b * 3 + i] = tn[i];
            }
        }
        __syncthreads();
        if (A.src_out) {
            float *out = A.src_out + (size_t)b * 3 * N;
            for (int n = threadIdx.x; n < N; n += kStepThreads) {
                const float x = X[n], y = X[N + n], z = X[2 * N + n];
#pragma unroll
                for (int i = 0; i < 3; ++i)
                    out[i * N + n] = s_Rt[i * 3] * x + s_Rt[i * 3 + 1] * y + s_Rt[i * 3 + 2] * z + s_Rt[9 + i];
            }
        }
    }
}

}  // namespace houv

using namespace houv;

extern "C" int houv_idam_step(const float *src, const float *tgt, const void *corr_idx,
                              int corr_is_int64, const float *weight_logits, int B, int N, int M,
                              float *src_out, float *R_acc, float *t_acc, float *R_ab, float *t_ab,
                              float *weights_out, void *stream) {
    if (B < 0 || N < 0 || M < 0) return HOUV_E_BADARG;
    if (B == 0) return 0;
    if (!src || !tgt || !corr_idx || !weight_logits || N == 0 || M == 0) return HOUV_E_BADARG;
    if ((R_acc == nullptr) != (t_acc == nullptr)) return HOUV_E_BADARG;
    if (N > kStepMaxN) return HOUV_E_TOOBIG;
    StepArgs a;
    a.src = src; a.tgt = tgt; a.logits = weight_logits;
    a.corr32 = corr_is_int64 ? nullptr : (const int *)corr_idx;
    a.corr64 = corr_is_int64 ? (const long long *)corr_idx : nullptr;
    a.src_out = src_out; a.R_acc = R_acc; a.t_acc = t_acc; a.R_ab = R_ab; a.t_ab = t_ab;
    a.w_out = weights_out;
    a.B = B; a.N = N; a.M = M;
    const size_t smem = sizeof(float) * 7 * (size_t)N;
    static unsigned long long smem_set = 0;  // per-device flag bits
    if (int e = ensure_dyn_smem(idam_step_kernel, (int)(sizeof(float) * 7 * kStepMaxN), smem_set)) return e;
    int blocks = B < 148 * 8 ? B : 148 * 8;
    idam_step_kernel<<<blocks, kStepThreads, smem, (cudaStream_t)stream>>>(a);
    return check_launch();
}
