// ICP -- point-to-point ICP with one CTA per pair resident for all iterations.
//
// Replaces the per-pair call site
//   o3d.registration.registration_icp(pcd, pcd2, 0.02, trans_init,
//       TransformationEstimationPointToPoint(), ICPConvergenceCriteria(max_iteration=500))
// of /root/reference/registration/train_ICP.py:137-153.  Open3D 0.9.0 is a
// third-party dependency that is not in the reference tree, so the loop below
// restates its published algorithm (see DESIGN.md "ICP": parity UNPINNED):
//   result = evaluate(init); repeat <= max_iteration times:
//     update = Kabsch/Umeyama(no scale) on the correspondences; T = update * T;
//     result' = evaluate(T); stop when |d fitness| and |d inlier_rmse| < 1e-6
//   evaluate: nearest target of every transformed source point, kept as a
//   correspondence when its squared distance < max_distance^2;
//   fitness = #corr / #source, inlier_rmse = sqrt(sum d^2 / #corr).
// Open3D runs this serially on the host with a KD-tree in float64; here the
// target cloud sits in shared memory (SoA), the NN search is K1's packed-fp32
// group-min scan, sums are reduced in float64, the 3x3 solve is svd3.cuh.
#include "common.cuh"
#include "svd3.cuh"

namespace houv {

constexpr int kIcpPts = 2048;
constexpr int kIcpThreads = 512;
constexpr int kIcpQ = 4;
constexpr int kIcpG = 8;
constexpr int kIcpSums = 17;  // count, sum d^2, sum y[3], sum t[3], sum y t^T [9]

struct IcpArgs {
    const float *src;   // [P,N,3]
    const float *tgt;   // [P,M,3]
    const float *init;  // [16] or [P,16] row-major 4x4
    float *T_out;       // [P,16]
    float *stats;       // [P,4] fitness, inlier_rmse, iterations run, correspondences
    int P, N, M, init_per_pair, max_iter;
    float max_dist;
    double rel_fitness, rel_rmse;
};

struct IcpSmem {
    float tx[kIcpPts], ty[kIcpPts], tz[kIcpPts];
    double red[kIcpThreads / 32][kIcpSums];
    double T[16];
    double sums[kIcpSums];
    int done;
};

__global__ void __launch_bounds__(kIcpThreads) icp_kernel(const IcpArgs A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    IcpSmem &S = *reinterpret_cast<IcpSmem *>(smem_raw);
    const int t = threadIdx.x, pair = blockIdx.x;
    const int N = A.N, M = A.M;
    const int Mpad = (M + kIcpG - 1) / kIcpG * kIcpG;
    const float *src = A.src + (size_t)pair * N * 3;
    const float *tgt = A.tgt + (size_t)pair * M * 3;
    const float inf = __int_as_float(0x7f800000);
    for (int j = t; j < Mpad; j += kIcpThreads) {
        const bool ok = j < M;
        S.tx[j] = ok ? __ldg(tgt + 3 * j) : inf;
        S.ty[j] = ok ? __ldg(tgt + 3 * j + 1) : 0.f;
        S.tz[j] = ok ? __ldg(tgt + 3 * j + 2) : 0.f;
    }
    if (t < 16) S.T[t] = (double)A.init[(A.init_per_pair ? (size_t)pair * 16 : 0) + t];
    if (t == 0) S.done = 0;
    __syncthreads();

    const float maxd2 = A.max_dist * A.max_dist;
    const float4 *x4 = reinterpret_cast<const float4 *>(S.tx);
    const float4 *y4 = reinterpret_cast<const float4 *>(S.ty);
    const float4 *z4 = reinterpret_cast<const float4 *>(S.tz);
    double fit = 0.0, rmse = 0.0;  // thread 0: result of the previous evaluation
    int iters = 0;

    for (int ev = 0; ev <= A.max_iter; ++ev) {
        // ---- evaluate the current transformation
        float Tf[12];
#pragma unroll
        for (int i = 0; i < 12; ++i) Tf[i] = (float)S.T[i];
        double acc[kIcpSums];
#pragma unroll
        for (int i = 0; i < kIcpSums; ++i) acc[i] = 0.0;
#pragma unroll 1
        for (int base = 0; base < N; base += kIcpThreads * kIcpQ) {
            float qx[kIcpQ], qy[kIcpQ], qz[kIcpQ], best[kIcpQ];
            int gb[kIcpQ];
#pragma unroll
            for (int r = 0; r < kIcpQ; ++r) {
                int q = base + r * kIcpThreads + t;
                q = q < N ? q : N - 1;
                const float x0 = __ldg(src + 3 * q), x1 = __ldg(src + 3 * q + 1),
                            x2 = __ldg(src + 3 * q + 2);
                qx[r] = __fmaf_rn(Tf[2], x2, __fmaf_rn(Tf[1], x1, __fmul_rn(Tf[0], x0))) + Tf[3];
                qy[r] = __fmaf_rn(Tf[6], x2, __fmaf_rn(Tf[5], x1, __fmul_rn(Tf[4], x0))) + Tf[7];
                qz[r] = __fmaf_rn(Tf[10], x2, __fmaf_rn(Tf[9], x1, __fmul_rn(Tf[8], x0))) + Tf[11];
                best[r] = inf;
                gb[r] = 0;
            }
            for (int g = 0; g < Mpad / kIcpG; ++g) {
                const float4 xa = x4[2 * g], xb = x4[2 * g + 1];
                const float4 ya = y4[2 * g], yb = y4[2 * g + 1];
                const float4 za = z4[2 * g], zb = z4[2 * g + 1];
#pragma unroll
                for (int r = 0; r < kIcpQ; ++r) {
                    const float2 nx = make_float2(-qx[r], -qx[r]), ny = make_float2(-qy[r], -qy[r]),
                                 nz = make_float2(-qz[r], -qz[r]);
                    float2 d0 = sqdist2(make_float2(xa.x, xa.y), make_float2(ya.x, ya.y),
                                        make_float2(za.x, za.y), nx, ny, nz);
                    float2 d1 = sqdist2(make_float2(xa.z, xa.w), make_float2(ya.z, ya.w),
                                        make_float2(za.z, za.w), nx, ny, nz);
                    float2 d2 = sqdist2(make_float2(xb.x, xb.y), make_float2(yb.x, yb.y),
                                        make_float2(zb.x, zb.y), nx, ny, nz);
                    float2 d3 = sqdist2(make_float2(xb.z, xb.w), make_float2(yb.z, yb.w),
                                        make_float2(zb.z, zb.w), nx, ny, nz);
                    float gm = min3(min3(d0.x, d0.y, d1.x), d1.y, d2.x);
                    gm = min3(min3(gm, d2.y, d3.x), d3.y, best[r]);
                    gb[r] = gm < best[r] ? g : gb[r];
                    best[r] = gm;
                }
            }
#pragma unroll
            for (int r = 0; r < kIcpQ; ++r) {
                const int q = base + r * kIcpThreads + t;
                if (q < N && best[r] < maxd2) {  // inside the correspondence radius
                    const int j0 = gb[r] * kIcpG;
                    int bj = j0;
#pragma unroll
                    for (int c = kIcpG - 1; c >= 0; --c)
                        if (sqdist_ref(S.tx[j0 + c], S.ty[j0 + c], S.tz[j0 + c], qx[r], qy[r], qz[r]) ==
                            best[r])
                            bj = j0 + c;
                    const double y[3] = {qx[r], qy[r], qz[r]};
                    const double w[3] = {S.tx[bj], S.ty[bj], S.tz[bj]};
                    acc[0] += 1.0;
                    acc[1] += (double)best[r];
#pragma unroll
                    for (int i = 0; i < 3; ++i) {
                        acc[2 + i] += y[i];
                        acc[5 + i] += w[i];
#pragma unroll
                        for (int j = 0; j < 3; ++j) acc[8 + i * 3 + j] += y[i] * w[j];
                    }
                }
            }
        }
        // ---- block reduction in float64
#pragma unroll
        for (int i = 0; i < kIcpSums; ++i)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], o);
        if ((t & 31) == 0)
#pragma unroll
            for (int i = 0; i < kIcpSums; ++i) S.red[t >> 5][i] = acc[i];
        __syncthreads();
        if (t == 0) {
            double s[kIcpSums];
            for (int i = 0; i < kIcpSums; ++i) {
                s[i] = 0.0;
                for (int w = 0; w < kIcpThreads / 32; ++w) s[i] += S.red[w][i];
            }
            const double cnt = s[0];
            const double nfit = cnt / (double)N;
            const double nrmse = cnt > 0 ? sqrt(s[1] / cnt) : 0.0;
            bool stop = false;
            if (ev > 0) {
                iters = ev;
                stop = fabs(fit - nfit) < A.rel_fitness && fabs(rmse - nrmse) < A.rel_rmse;
            }
            fit = nfit;
            rmse = nrmse;
            if (stop || ev == A.max_iter) {
                S.done = 1;
                for (int i = 0; i < 16; ++i) A.T_out[(size_t)pair * 16 + i] = (float)S.T[i];
                A.stats[(size_t)pair * 4 + 0] = (float)fit;
                A.stats[(size_t)pair * 4 + 1] = (float)rmse;
                A.stats[(size_t)pair * 4 + 2] = (float)iters;
                A.stats[(size_t)pair * 4 + 3] = (float)cnt;
            } else if (cnt > 0) {
                // update = argmin sum |R y + t - w|^2 over the correspondences; T <- update * T
                double H[3][3], R[3][3], ym[3], wm[3];
                for (int i = 0; i < 3; ++i) { ym[i] = s[2 + i] / cnt; wm[i] = s[5 + i] / cnt; }
                for (int i = 0; i < 3; ++i)
                    for (int j = 0; j < 3; ++j) H[i][j] = s[8 + i * 3 + j] - cnt * ym[i] * wm[j];
                kabsch_from_covariance(H, R);
                double tu[3], Tn[12];
                for (int i = 0; i < 3; ++i)
                    tu[i] = wm[i] - (R[i][0] * ym[0] + R[i][1] * ym[1] + R[i][2] * ym[2]);
                for (int i = 0; i < 3; ++i)
                    for (int j = 0; j < 4; ++j)
                        Tn[i * 4 + j] = R[i][0] * S.T[j] + R[i][1] * S.T[4 + j] + R[i][2] * S.T[8 + j] +
                                        (j == 3 ? tu[i] : 0.0);
                for (int i = 0; i < 12; ++i) S.T[i] = Tn[i];
            }
        }
        __syncthreads();
        if (S.done) break;
    }
}

}  // namespace houv

using namespace houv;

extern "C" int houv_icp(const float *src, const float *tgt, int P, int N, int M, const float *init,
                        int init_per_pair, float max_distance, int max_iteration,
                        double relative_fitness, double relative_rmse, float *T_out, float *stats,
                        void *stream) {
    if (P < 0 || N <= 0 || M <= 0 || max_iteration < 0) return HOUV_E_BADARG;
    if (P == 0) return 0;
    if (!src || !tgt || !init || !T_out || !stats) return HOUV_E_BADARG;
    if (N > kIcpPts || M > kIcpPts) return HOUV_E_TOOBIG;
    IcpArgs a;
    a.src = src; a.tgt = tgt; a.init = init; a.T_out = T_out; a.stats = stats;
    a.P = P; a.N = N; a.M = M; a.init_per_pair = init_per_pair; a.max_iter = max_iteration;
    a.max_dist = max_distance; a.rel_fitness = relative_fitness; a.rel_rmse = relative_rmse;
    static unsigned long long smem_set = 0;  // per-device flag bits
    if (int e = ensure_dyn_smem(icp_kernel, (int)sizeof(IcpSmem), smem_set)) return e;
    icp_kernel<<<P, kIcpThreads, sizeof(IcpSmem), (cudaStream_t)stream>>>(a);
    return check_launch();
}
