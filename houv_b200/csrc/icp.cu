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
// target cloud sits in shared memory (SoA), sums are reduced in float64, the 3x3
// solve is svd3.cuh.  The NN search only has to find neighbours inside the
// correspondence radius, so the target is binned once per pair into a uniform grid
// of cells no smaller than 1.01 x the radius (counting sort in shared memory, up to
// 24^3 cells) and a query looks at the 3 x 3 runs of cells around it: every target
// point within the radius is in one of them, and the nearest one (same float
// arithmetic as the brute-force scan, lowest index on equal distances) is the same
// point the full scan finds whenever that scan's minimum is inside the radius.
// Pairs whose radius is large against the cloud (fewer than 64 cells) keep K1's
// packed-fp32 group-min scan over all targets.
#include <stdlib.h>

#include <type_traits>

#include "common.cuh"
#include "svd3.cuh"

namespace houv {

constexpr int kIcpPts = 2048;
constexpr int kIcpThreads = 512;
constexpr int kIcpQ = 4;
constexpr int kIcpG = 8;
constexpr int kIcpSums = 17;  // count, sum d^2, sum y[3], sum t[3], sum y t^T [9]
constexpr int kIcpGridMax = 24;  // cells per axis at most
constexpr int kIcpCells = kIcpGridMax * kIcpGridMax * kIcpGridMax;

struct IcpArgs {
    const float *src;   // [P,N,3]
    const float *tgt;   // [P,M,3]
    const float *init;  // [16] or [P,16] row-major 4x4
    float *T_out;       // [P,16]
    float *stats;       // [P,4] fitness, inlier_rmse, iterations run, correspondences
    int P, N, M, init_per_pair, max_iter, allow_grid;
    float max_dist;
    double rel_fitness, rel_rmse;
};

struct IcpSmemCommon {
    double red[kIcpThreads / 32][kIcpSums];
    double T[16];
    double sums[kIcpSums];
    int done;
    float bb[6];                       // target bounding box: min xyz, max xyz
    float wmin[kIcpThreads / 32][3], wmax[kIcpThreads / 32][3];
};
struct IcpSmemScan : IcpSmemCommon {
    alignas(16) float tx[kIcpPts];  // target, original order (read as float4)
    alignas(16) float ty[kIcpPts];
    alignas(16) float tz[kIcpPts];
};
struct IcpSmemGrid : IcpSmemCommon {
    float4 pts[kIcpPts];               // target sorted by cell: x, y, z, original index (int bits)
    int cell_end[kIcpCells + 1];       // after the scatter: end of cell c (= start of cell c + 1)
    int scan[kIcpThreads];
};

// kGrid: the grid search for pairs whose grid has >= 64 cells; !kGrid: the full scan for the others.
// Both kernels are launched over all pairs; each pair is taken by exactly one of them (the choice needs
// the target's bounding box, which both compute).  The grid kernel needs half the registers of the scan
// and runs two CTAs per SM.
template <bool kGrid>
__global__ void __launch_bounds__(kIcpThreads, kGrid ? 2 : 1) icp_kernel(const IcpArgs A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    typedef typename std::conditional<kGrid, IcpSmemGrid, IcpSmemScan>::type Smem;
    Smem &S = *reinterpret_cast<Smem *>(smem_raw);
    IcpSmemGrid &SG = *reinterpret_cast<IcpSmemGrid *>(smem_raw);  // only touched when kGrid
    IcpSmemScan &SS = *reinterpret_cast<IcpSmemScan *>(smem_raw);  // only touched when !kGrid
    const int t = threadIdx.x, pair = blockIdx.x;
    const int N = A.N, M = A.M;
    const int Mpad = (M + kIcpG - 1) / kIcpG * kIcpG;
    const float *src = A.src + (size_t)pair * N * 3;
    const float *tgt = A.tgt + (size_t)pair * M * 3;
    const float inf = __int_as_float(0x7f800000);
    // ---- bounding box of the target, grid geometry
    float lo[3] = {inf, inf, inf}, hi[3] = {-inf, -inf, -inf};
    for (int j = t; j < M; j += kIcpThreads)
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float v = __ldg(tgt + 3 * j + c);
            lo[c] = fminf(lo[c], v);
            hi[c] = fmaxf(hi[c], v);
        }
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[c] = fminf(lo[c], __shfl_xor_sync(0xffffffffu, lo[c], o));
            hi[c] = fmaxf(hi[c], __shfl_xor_sync(0xffffffffu, hi[c], o));
        }
    if ((t & 31) == 0)
#pragma unroll
        for (int c = 0; c < 3; ++c) { S.wmin[t >> 5][c] = lo[c]; S.wmax[t >> 5][c] = hi[c]; }
    __syncthreads();
    if (t < 3) {
        float a = inf, b = -inf;
        for (int w = 0; w < kIcpThreads / 32; ++w) { a = fminf(a, S.wmin[w][t]); b = fmaxf(b, S.wmax[w][t]); }
        S.bb[t] = a; S.bb[3 + t] = b;
    }
    __syncthreads();
    const float ext = fmaxf(fmaxf(S.bb[3] - S.bb[0], S.bb[4] - S.bb[1]), S.bb[5] - S.bb[2]);
    // cell edge: at least 1.01 x the radius, and no more than kIcpGridMax cells along the longest axis
    const float cell = fmaxf(1.01f * A.max_dist, ext / (float)kIcpGridMax * 1.0001f);
    const float inv_cell = 1.f / cell;
    int gdim[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const int g = (int)floorf((S.bb[3 + c] - S.bb[c]) * inv_cell) + 1;
        gdim[c] = g < 1 ? 1 : (g > kIcpGridMax ? kIcpGridMax : g);
    }
    const int ncell = gdim[0] * gdim[1] * gdim[2];
    // non-finite coordinates or a radius that is large against the cloud: full scan
    const bool use_grid = A.allow_grid && ncell >= 64 && ext < 1e30f && cell > 0.f && cell < 1e30f;
    if (use_grid != kGrid) return;  // the other kernel's pair (uniform over the CTA)
    auto cell_of = [&](float v, int c) {  // clamped: points on the upper faces fall into the last cell
        int k = (int)floorf((v - S.bb[c]) * inv_cell);
        return k < 0 ? 0 : (k >= gdim[c] ? gdim[c] - 1 : k);
    };
    if (kGrid) {
        // counting sort of the target points by cell (x fastest)
        for (int c = t; c <= ncell; c += kIcpThreads) SG.cell_end[c] = 0;
        __syncthreads();
        for (int j = t; j < M; j += kIcpThreads) {
            const int c = (cell_of(__ldg(tgt + 3 * j + 2), 2) * gdim[1] + cell_of(__ldg(tgt + 3 * j + 1), 1)) * gdim[0] +
                          cell_of(__ldg(tgt + 3 * j), 0);
            atomicAdd(&SG.cell_end[c], 1);
        }
        __syncthreads();
        // exclusive scan over the cells: per-thread serial chunk + block scan of the chunk sums
        const int per = (ncell + kIcpThreads - 1) / kIcpThreads;
        const int c0 = t * per, c1 = min(ncell, c0 + per);
        int sum = 0;
        for (int c = c0; c < c1; ++c) sum += SG.cell_end[c];
        SG.scan[t] = sum;
        __syncthreads();
        for (int o = 1; o < kIcpThreads; o <<= 1) {
            const int v = t >= o ? SG.scan[t - o] : 0;
            __syncthreads();
            SG.scan[t] += v;
            __syncthreads();
        }
        int run = SG.scan[t] - sum;  // exclusive prefix of this thread's chunk
        for (int c = c0; c < c1; ++c) { const int n = SG.cell_end[c]; SG.cell_end[c] = run; run += n; }
        __syncthreads();
        // scatter: cell_end[c] walks from the start of cell c to its end
        for (int j = t; j < M; j += kIcpThreads) {
            const float x = __ldg(tgt + 3 * j), y = __ldg(tgt + 3 * j + 1), z = __ldg(tgt + 3 * j + 2);
            const int c = (cell_of(z, 2) * gdim[1] + cell_of(y, 1)) * gdim[0] + cell_of(x, 0);
            const int p = atomicAdd(&SG.cell_end[c], 1);
            SG.pts[p] = make_float4(x, y, z, __int_as_float(j));
        }
    } else {
        for (int j = t; j < Mpad; j += kIcpThreads) {
            const bool ok = j < M;
            SS.tx[j] = ok ? __ldg(tgt + 3 * j) : inf;
            SS.ty[j] = ok ? __ldg(tgt + 3 * j + 1) : 0.f;
            SS.tz[j] = ok ? __ldg(tgt + 3 * j + 2) : 0.f;
        }
    }
    if (t < 16) S.T[t] = (double)A.init[(A.init_per_pair ? (size_t)pair * 16 : 0) + t];
    if (t == 0) S.done = 0;
    __syncthreads();

    const float maxd2 = A.max_dist * A.max_dist;
    const float4 *x4 = reinterpret_cast<const float4 *>(SS.tx);
    const float4 *y4 = reinterpret_cast<const float4 *>(SS.ty);
    const float4 *z4 = reinterpret_cast<const float4 *>(SS.tz);
    double fit = 0.0, rmse = 0.0;  // thread 0: result of the previous evaluation
    int iters = 0;
    double Vwarm[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};  // thread 0: Jacobi warm start (svd3.cuh)

    for (int ev = 0; ev <= A.max_iter; ++ev) {
        // ---- evaluate the current transformation
        float Tf[12];
#pragma unroll
        for (int i = 0; i < 12; ++i) Tf[i] = (float)S.T[i];
        double acc[kIcpSums];
#pragma unroll
        for (int i = 0; i < kIcpSums; ++i) acc[i] = 0.0;
        if (kGrid) {
            // largest float below the squared radius: "d <= best" is then the whole acceptance test for the
            // first candidate, and one compare rejects the many candidates of the 27 cells that are farther
            const float start = __uint_as_float(__float_as_uint(maxd2) - 1u);
#pragma unroll 1
            for (int q = t; q < N; q += kIcpThreads) {
                const float x0 = __ldg(src + 3 * q), x1 = __ldg(src + 3 * q + 1), x2 = __ldg(src + 3 * q + 2);
                const float qx = __fmaf_rn(Tf[2], x2, __fmaf_rn(Tf[1], x1, __fmul_rn(Tf[0], x0))) + Tf[3];
                const float qy = __fmaf_rn(Tf[6], x2, __fmaf_rn(Tf[5], x1, __fmul_rn(Tf[4], x0))) + Tf[7];
                const float qz = __fmaf_rn(Tf[10], x2, __fmaf_rn(Tf[9], x1, __fmul_rn(Tf[8], x0))) + Tf[11];
                // unclamped cell of the query; a query more than one cell outside the box has no neighbour
                const int kx = (int)floorf((qx - S.bb[0]) * inv_cell), ky = (int)floorf((qy - S.bb[1]) * inv_cell),
                          kz = (int)floorf((qz - S.bb[2]) * inv_cell);
                float best = start;
                int bj = 0x7fffffff;
                float bx = 0.f, by = 0.f, bz = 0.f;
                const int xl = max(kx - 1, 0), xh = min(kx + 1, gdim[0] - 1);
                if (xl <= xh && maxd2 > 0.f)
                    for (int cz = max(kz - 1, 0); cz <= min(kz + 1, gdim[2] - 1); ++cz)
                        for (int cy = max(ky - 1, 0); cy <= min(ky + 1, gdim[1] - 1); ++cy) {
                            const int row = (cz * gdim[1] + cy) * gdim[0];
                            const int p0 = row + xl == 0 ? 0 : SG.cell_end[row + xl - 1], p1 = SG.cell_end[row + xh];
                            for (int p = p0; p < p1; ++p) {
                                const float4 c = SG.pts[p];
                                const float d = sqdist_ref(c.x, c.y, c.z, qx, qy, qz);
                                if (d <= best) {
                                    const int oj = __float_as_int(c.w);
                                    if (d < best || oj < bj) { best = d; bj = oj; bx = c.x; by = c.y; bz = c.z; }
                                }
                            }
                        }
                if (bj != 0x7fffffff) {
                    const double y[3] = {qx, qy, qz};
                    const double w[3] = {bx, by, bz};
                    acc[0] += 1.0;
                    acc[1] += (double)best;
#pragma unroll
                    for (int i = 0; i < 3; ++i) {
                        acc[2 + i] += y[i];
                        acc[5 + i] += w[i];
#pragma unroll
                        for (int j = 0; j < 3; ++j) acc[8 + i * 3 + j] += y[i] * w[j];
                    }
                }
            }
        } else
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
                        if (sqdist_ref(SS.tx[j0 + c], SS.ty[j0 + c], SS.tz[j0 + c], qx[r], qy[r], qz[r]) ==
                            best[r])
                            bj = j0 + c;
                    const double y[3] = {qx[r], qy[r], qz[r]};
                    const double w[3] = {SS.tx[bj], SS.ty[bj], SS.tz[bj]};
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
        if (t < kIcpSums) {  // one thread per quantity, warps in ascending order
            double v = 0.0;
            for (int w = 0; w < kIcpThreads / 32; ++w) v += S.red[w][t];
            S.sums[t] = v;
        }
        __syncthreads();
        if (t == 0) {
            double s[kIcpSums];
            for (int i = 0; i < kIcpSums; ++i) s[i] = S.sums[i];
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
                kabsch_from_covariance(H, R, Vwarm);
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
    a.allow_grid = !(getenv("HOUV_ICP_GRID") && atoi(getenv("HOUV_ICP_GRID")) == 0);  // 0: always scan
    static unsigned long long set_g = 0, set_s = 0;  // per-device flag bits
    if (int e = ensure_dyn_smem(icp_kernel<true>, (int)sizeof(IcpSmemGrid), set_g)) return e;
    if (int e = ensure_dyn_smem(icp_kernel<false>, (int)sizeof(IcpSmemScan), set_s)) return e;
    if (a.allow_grid) {
        icp_kernel<true><<<P, kIcpThreads, sizeof(IcpSmemGrid), (cudaStream_t)stream>>>(a);
        if (int e = check_launch()) return e;
    }
    icp_kernel<false><<<P, kIcpThreads, sizeof(IcpSmemScan), (cudaStream_t)stream>>>(a);
    return check_launch();
}
