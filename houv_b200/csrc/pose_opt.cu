// K3 -- fused HOUV pose optimisation for sm_100a: one CTA keeps one start (row)
// of the multi-start optimisation resident for ALL iterations.
//
// Replaces the hot loop of predict_model (/root/reference/registration/models/
// houv.py:120-126), i.e. per iteration and per row
//   HOUV.forward            houv.py:94-103   pose from 8 unconstrained scalars, transform
//   Predict_loss            houv.py:209-222  4 Chamfer calls (3-D + three axis-masked views)
//     calc_cd_percent       model_utils_completion.py:83-100  top-k trimmed sqrt-mean
//     chamfer_3DDist fwd    chamfer3D.cu:12-154   8 directed NN searches
//   .mean().backward()      autograd: min -> mean -> sqrt -> topk -> chamfer bwd (chamfer3D.cu:155-195)
//                           -> mask -> bmm -> Rodrigues -> normalisations -> sin re-parametrisation
//   Adam.step               torch.optim.Adam(lr=0.01), houv.py:118,126
// which the reference runs as ~120 kernel launches per iteration.
//
// Design (see DESIGN.md "K3"):
//  * target cloud and the transformed source live in shared memory as SoA; the
//    clouds are read from HBM once per row, everything else stays on chip;
//  * the four metrics share one evaluation: dx,dy,dz (3 FADD2), yy,xx (2 FMUL2),
//    xy=fma(dx,dx,yy), full=fma(dz,dz,xy), v0=fma(dz,dz,yy), v1=fma(dz,dz,xx)
//    (4 FFMA2) -- bit-identical to four separate reference calls on the masked
//    clouds because fma(0,0,c)==c and fma(a,a,+0)==a*a;
//  * arg-min tracking as in K1: FMNMX3 group minima + winning group id, exact
//    index recovered only where the gradient needs it (the direction that wins
//    the min(), and for the 3-D term only the k smallest);
//  * k-th smallest by an in-CTA 4x8-bit radix select on the float bit patterns;
//  * the gradient w.r.t. the transformed points is never materialised: only
//    G = sum_i g_i (x) x_i (3x3) and sum_i g_i (3) are reduced, then chained in
//    closed form to the 8 scalars and stepped with Adam by one thread.
#include <stdlib.h>

#include "pose_common.cuh"

namespace houv {

// Tunables of the scan are template parameters (variants are compiled side by side and
// picked with houv_set_pose_variant): Q = queries per thread per pass, H = 8-target halves
// per arg-min group (group = 8*H targets), U = unroll of the group loop.

struct Smem {
    float tx[kPts], ty[kPts], tz[kPts];   // target cloud (fixed)
    float yx[kPts], yy[kPts], yz[kPts];   // transformed source of this iteration
    float d3[2][kPts];                    // 3-D term distances: [0]=dist1 (tgt queries) [1]=dist2 (src queries)
    unsigned char gid[4][2][kPts];        // winning group per metric / direction / query
    int hist[2][256];
    float red[kThreads / 32][16];
    float pose[12];                       // R (row-major 9) + T (3)
    unsigned selu[8];                     // radix-select broadcast: prefix, rank, tie count
    double state[80];                     // PoseState of this row (touched by thread 0 only)
    double bt[2];                         // beta1^t, beta2^t
};

// block-wide sum of NV per-thread values; result valid in all threads
template <int NV>
__device__ __forceinline__ void block_sum(float (&v)[NV], Smem &S) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
    }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) {
#pragma unroll
        for (int i = 0; i < NV; ++i) S.red[w][i] = v[i];
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        float s = 0.f;
#pragma unroll
        for (int ww = 0; ww < kThreads / 32; ++ww) s += S.red[ww][i];
        v[i] = s;
    }
}

// One scan pass: kQ queries per thread against `ngroups` groups of 8 targets in
// SoA shared arrays.  best[m][r], gb[m][r]: running minimum / winning group of
// metric m (0 = 3-D, 1..3 = view with coordinate m-1 masked) for query r.
template <int kQ, int H, int U>
__device__ __forceinline__ void scan_pass(const float *__restrict__ sx, const float *__restrict__ sy,
                                          const float *__restrict__ sz, int ngroups,
                                          const float (&qx)[kQ], const float (&qy)[kQ],
                                          const float (&qz)[kQ], float (&best)[4][kQ],
                                          int (&gb)[4][kQ]) {
    const float4 *x4 = reinterpret_cast<const float4 *>(sx);
    const float4 *y4 = reinterpret_cast<const float4 *>(sy);
    const float4 *z4 = reinterpret_cast<const float4 *>(sz);
    float2 nqx[kQ], nqy[kQ], nqz[kQ];
#pragma unroll
    for (int r = 0; r < kQ; ++r) {
        nqx[r] = make_float2(-qx[r], -qx[r]);
        nqy[r] = make_float2(-qy[r], -qy[r]);
        nqz[r] = make_float2(-qz[r], -qz[r]);
#pragma unroll
        for (int m = 0; m < 4; ++m) { best[m][r] = kInf; gb[m][r] = 0; }
    }
#pragma unroll U
    for (int g = 0; g < ngroups; ++g) {
        float gm[4][kQ];
#pragma unroll
        for (int r = 0; r < kQ; ++r)
#pragma unroll
            for (int m = 0; m < 4; ++m) gm[m][r] = best[m][r];
#pragma unroll
        for (int hh = 0; hh < H; ++hh) {
            const int o = 2 * (g * H + hh);
            float4 X[2] = {x4[o], x4[o + 1]};
            float4 Y[2] = {y4[o], y4[o + 1]};
            float4 Z[2] = {z4[o], z4[o + 1]};
#pragma unroll
            for (int r = 0; r < kQ; ++r) {
                float2 f[4], a[4], b[4], c[4];  // full, view0, view1, view2 of the 4 target pairs
#pragma unroll
                for (int h = 0; h < 4; ++h) {
                    const float4 &xv = X[h >> 1], &yv = Y[h >> 1], &zv = Z[h >> 1];
                    float2 tx = (h & 1) ? make_float2(xv.z, xv.w) : make_float2(xv.x, xv.y);
                    float2 ty = (h & 1) ? make_float2(yv.z, yv.w) : make_float2(yv.x, yv.y);
                    float2 tz = (h & 1) ? make_float2(zv.z, zv.w) : make_float2(zv.x, zv.y);
                    float2 dx = __fadd2_rn(tx, nqx[r]), dy = __fadd2_rn(ty, nqy[r]),
                           dz = __fadd2_rn(tz, nqz[r]);
                    float2 yy = __fmul2_rn(dy, dy), xx = __fmul2_rn(dx, dx);
                    float2 xy = __ffma2_rn(dx, dx, yy);
                    f[h] = __ffma2_rn(dz, dz, xy);
                    a[h] = __ffma2_rn(dz, dz, yy);
                    b[h] = __ffma2_rn(dz, dz, xx);
                    c[h] = xy;
                }
#define HOUV_FOLD(M, V)                                                              \
    {                                                                                \
        float t_ = min3(min3(V[0].x, V[0].y, V[1].x), V[1].y, V[2].x);               \
        gm[M][r] = min3(min3(t_, V[2].y, V[3].x), V[3].y, gm[M][r]);                 \
    }
                HOUV_FOLD(0, f) HOUV_FOLD(1, a) HOUV_FOLD(2, b) HOUV_FOLD(3, c)
#undef HOUV_FOLD
            }
        }
        // strict '<' between groups in ascending order: the first minimum's group wins
#pragma unroll
        for (int r = 0; r < kQ; ++r)
#pragma unroll
            for (int m = 0; m < 4; ++m) {
                gb[m][r] = gm[m][r] < best[m][r] ? g : gb[m][r];
                best[m][r] = gm[m][r];
            }
    }
}

template <int kQ, int H, int U>
__global__ void __launch_bounds__(kThreads, 2) pose_kernel(const PoseArgs A) {
    constexpr int kG = 8 * H;  // targets per arg-min group
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem &S = *reinterpret_cast<Smem *>(smem_raw);
    const int t = threadIdx.x;
    const int row = blockIdx.x;
    const int pair = A.row_pair ? A.row_pair[row] : row / A.K;
    const int N = A.N, M = A.M;
    const int Npad = (N + kG - 1) / kG * kG, Mpad = (M + kG - 1) / kG * kG;
    const float *src = A.src + (size_t)pair * N * 3;
    const float *tgt = A.tgt + (size_t)pair * M * 3;

    // target cloud -> shared SoA, +inf padded (x and y: every metric sees inf)
    for (int j = t; j < Mpad; j += kThreads) {
        const bool ok = j < M;
        S.tx[j] = ok ? __ldg(tgt + 3 * j + 0) : kInf;
        S.ty[j] = ok ? __ldg(tgt + 3 * j + 1) : kInf;
        S.tz[j] = ok ? __ldg(tgt + 3 * j + 2) : 0.f;
    }

    // optimiser state of this row: shared memory, touched by thread 0 only (keeps ~70
    // values out of every thread's registers during the scans)
    static_assert(sizeof(PoseState) <= sizeof(S.state), "PoseState must fit Smem::state");
    PoseState &st = *reinterpret_cast<PoseState *>(S.state);
    double &b1t = S.bt[0], &b2t = S.bt[1];
    if (t == 0) {
        state_load(st, A, row);
        b1t = pow(A.beta1, (double)A.step0);
        b2t = pow(A.beta2, (double)A.step0);
    }
    // angle_base*pi/4 evaluated in double and rounded once, as python does (houv.py:96)
    const int abase = A.row_base ? A.row_base[row] : A.angle_base;
    const float boff = (float)((double)abase * 3.14159274101257324 / 4.0);
    const int k3d = A.k3d;
    // d(mean over the batch)/d(row): per row when rows of several batches share the launch
    const float gscale = A.row_scale ? A.row_scale[row] : A.grad_scale;

    for (int it = 0; it < A.iters; ++it) {
        // ---- (a) pose from the 8 scalars
        if (t == 0) pose_forward(st, boff, A.tran_scale, A.tran_offset, S.pose);
        __syncthreads();
        float Rm[9], Tv[3];
#pragma unroll
        for (int i = 0; i < 9; ++i) Rm[i] = S.pose[i];
#pragma unroll
        for (int i = 0; i < 3; ++i) Tv[i] = S.pose[9 + i];

        // ---- (b) transform: y = R x + T  (bmm(src, R^T) + T, houv.py:102)
        for (int i = t; i < Npad; i += kThreads) {
            if (i < N) {
                const float x0 = __ldg(src + 3 * i), x1 = __ldg(src + 3 * i + 1),
                            x2 = __ldg(src + 3 * i + 2);
                S.yx[i] = __fadd_rn(__fmaf_rn(x2, Rm[2], __fmaf_rn(x1, Rm[1], __fmul_rn(x0, Rm[0]))), Tv[0]);
                S.yy[i] = __fadd_rn(__fmaf_rn(x2, Rm[5], __fmaf_rn(x1, Rm[4], __fmul_rn(x0, Rm[3]))), Tv[1]);
                S.yz[i] = __fadd_rn(__fmaf_rn(x2, Rm[8], __fmaf_rn(x1, Rm[7], __fmul_rn(x0, Rm[6]))), Tv[2]);
            } else {
                S.yx[i] = kInf; S.yy[i] = kInf; S.yz[i] = 0.f;
            }
        }
        __syncthreads();

        // ---- (c,d) the 8 directed searches: dir 1 = dist2 (src queries over tgt),
        //            dir 0 = dist1 (tgt queries over transformed src)
        float vs[2][3] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};  // sum sqrt(d) of the view terms
#pragma unroll 1
        for (int dir = 1; dir >= 0; --dir) {
            const float *qX = dir ? S.yx : S.tx, *qY = dir ? S.yy : S.ty, *qZ = dir ? S.yz : S.tz;
            const float *tX = dir ? S.tx : S.yx, *tY = dir ? S.ty : S.yy, *tZ = dir ? S.tz : S.yz;
            const int nq = dir ? N : M, ngroups = (dir ? Mpad : Npad) / kG;
#pragma unroll 1
            for (int base = 0; base < nq; base += kThreads * kQ) {
                float qx[kQ], qy[kQ], qz[kQ], best[4][kQ];
                int gb[4][kQ];
#pragma unroll
                for (int r = 0; r < kQ; ++r) {
                    int q = base + r * kThreads + t;
                    q = q < nq ? q : nq - 1;
                    qx[r] = qX[q]; qy[r] = qY[q]; qz[r] = qZ[q];
                }
                scan_pass<kQ, H, U>(tX, tY, tZ, ngroups, qx, qy, qz, best, gb);
#pragma unroll
                for (int r = 0; r < kQ; ++r) {
                    const int q = base + r * kThreads + t;
                    if (q < nq) {
                        S.d3[dir][q] = best[0][r];
#pragma unroll
                        for (int m = 0; m < 4; ++m) S.gid[m][dir][q] = (unsigned char)gb[m][r];
#pragma unroll
                        for (int m = 1; m < 4; ++m) vs[dir][m - 1] += sqrtf(best[m][r]);
                    }
                }
            }
        }
        __syncthreads();

        // ---- (e) k-th smallest of d3[0][0..M) and d3[1][0..N): 4 x 8-bit radix select
        unsigned prefix[2] = {0u, 0u};
        int kk[2] = {k3d, k3d};
#pragma unroll 1
        for (int shift = 24; shift >= 0; shift -= 8) {
            for (int b = t; b < 512; b += kThreads) (&S.hist[0][0])[b] = 0;
            __syncthreads();
#pragma unroll
            for (int dir = 0; dir < 2; ++dir) {
                const int len = dir ? N : M;
                for (int q = t; q < len; q += kThreads) {
                    const unsigned bits = __float_as_uint(S.d3[dir][q]);
                    const bool match = (shift == 24) ||
                                       ((bits >> (shift + 8)) == (prefix[dir] >> (shift + 8)));
                    if (match) atomicAdd(&S.hist[dir][(bits >> shift) & 255u], 1);
                }
            }
            __syncthreads();
            if (t < 64) {
                const int dir = t >> 5, l = t & 31;
                int c[8], s = 0;
#pragma unroll
                for (int b = 0; b < 8; ++b) { c[b] = S.hist[dir][8 * l + b]; s += c[b]; }
                int incl = s;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    int n = __shfl_up_sync(0xffffffffu, incl, o);
                    if (l >= o) incl += n;
                }
                const int excl = incl - s;
                if (excl < kk[dir] && kk[dir] <= incl) {
                    int rem = kk[dir] - excl, bin = 0, cnt = 0;
                    bool done = false;
#pragma unroll
                    for (int b = 0; b < 8; ++b) {
                        if (!done) {
                            if (rem <= c[b]) { bin = 8 * l + b; cnt = c[b]; done = true; }
                            else rem -= c[b];
                        }
                    }
                    S.selu[dir * 2] = prefix[dir] | ((unsigned)bin << shift);
                    S.selu[dir * 2 + 1] = (unsigned)rem;
                    S.selu[4 + dir] = (unsigned)cnt;  // after the last round: number of ties
                }
            }
            __syncthreads();
#pragma unroll
            for (int dir = 0; dir < 2; ++dir) {
                prefix[dir] = S.selu[dir * 2];
                kk[dir] = (int)S.selu[dir * 2 + 1];
            }
        }
        float thr[2], tiew[2];
#pragma unroll
        for (int dir = 0; dir < 2; ++dir) {
            thr[dir] = __uint_as_float(prefix[dir]);
            tiew[dir] = (float)kk[dir] / (float)S.selu[4 + dir];
        }

        // ---- (f) the eight directional means and the four min() selections
        float sums[8];
        sums[0] = 0.f; sums[1] = 0.f;
#pragma unroll
        for (int dir = 0; dir < 2; ++dir) {
            const int len = dir ? N : M;
            for (int q = t; q < len; q += kThreads) {
                const float d = S.d3[dir][q];
                if (d < thr[dir]) sums[dir] += sqrtf(d);
            }
#pragma unroll
            for (int m = 0; m < 3; ++m) sums[2 + dir * 3 + m] = vs[dir][m];
        }
        block_sum<8>(sums, S);
        float cdv[4][2];  // [metric][dir]
#pragma unroll
        for (int dir = 0; dir < 2; ++dir) {
            cdv[0][dir] = (sums[dir] + (float)kk[dir] * sqrtf(thr[dir])) / (float)k3d;
#pragma unroll
            for (int m = 0; m < 3; ++m) cdv[1 + m][dir] = sums[2 + dir * 3 + m] / (float)N;
        }
        int side[4];
        float mins[4];
#pragma unroll
        for (int m = 0; m < 4; ++m) {  // torch.min over cat([dist1-side, dist2-side]): first wins ties
            side[m] = cdv[m][1] < cdv[m][0] ? 1 : 0;
            mins[m] = side[m] ? cdv[m][1] : cdv[m][0];
        }
        const float min_1 = mins[0];
        const float loss = __fadd_rn(__fmul_rn(min_1, A.w3d),
                                     __fmul_rn(A.wview, __fadd_rn(__fadd_rn(mins[1], mins[2]), mins[3])));
        if (t == 0) {
            if (A.trace) {
                A.trace[((size_t)it * A.rows + row) * 2 + 0] = loss;
                A.trace[((size_t)it * A.rows + row) * 2 + 1] = min_1;
            }
            if (it == A.iters - 1) {  // houv.py:134-138: results of the LAST forward
                A.min1_out[row] = min_1;
                if (A.loss_out) A.loss_out[row] = loss;
#pragma unroll
                for (int i = 0; i < 9; ++i) A.R_out[(size_t)row * 9 + i] = Rm[i];
#pragma unroll
                for (int i = 0; i < 3; ++i) A.T_out[(size_t)row * 3 + i] = Tv[i];
                if (A.tran_s_out) A.tran_s_out[row] = st.st;
            }
        }

        // ---- (h) gradient: only the side that won each min(), exact arg-min recovered
        //          from the winning group; reduce G = sum v (x) x and gT = sum v
        float acc[12];
#pragma unroll
        for (int i = 0; i < 12; ++i) acc[i] = 0.f;
#pragma unroll
        for (int m = 0; m < 4; ++m) {
            if (m > 0 && A.wview == 0.f) continue;  // loss = w3d*min_1 only (train_utils.py:439)
            const int dir = side[m];
            const float *qX = dir ? S.yx : S.tx, *qY = dir ? S.yy : S.ty, *qZ = dir ? S.yz : S.tz;
            const float *tX = dir ? S.tx : S.yx, *tY = dir ? S.ty : S.yy, *tZ = dir ? S.tz : S.yz;
            const int nq = dir ? N : M;
            // d(mean over rows)/d(row) * weight / k      houv.py:124,222; m_u_c.py:94-95
            const float coef = gscale * (m == 0 ? A.w3d : A.wview) / (float)(m == 0 ? k3d : N);
            for (int q = t; q < nq; q += kThreads) {
                float wt = 1.f;
                if (m == 0) {  // top-k membership of the trimmed 3-D term
                    const float d = S.d3[dir][q];
                    wt = d < thr[dir] ? 1.f : (d == thr[dir] ? tiew[dir] : 0.f);
                    if (wt == 0.f) continue;
                }
                const int j0 = (int)S.gid[m][dir][q] * kG;
                const float qx = qX[q], qy = qY[q], qz = qZ[q];
                float best = kInf;
                int bj = j0;
#pragma unroll
                for (int c = 0; c < kG; ++c) {  // ascending + strict '<': first minimum
                    const float d = metric_dist(m, tX[j0 + c], tY[j0 + c], tZ[j0 + c], qx, qy, qz);
                    if (d < best) { best = d; bj = j0 + c; }
                }
                // sqrt backward then chamfer backward (x2): d=0 gives inf*0 = NaN like autograd
                const float gg = (coef * wt) / (2.f * sqrtf(best)) * 2.f;
                const float sgn = dir ? 1.f : -1.f;  // (transformed src point) - (tgt point)
                float v0 = m == 1 ? 0.f : gg * (sgn * (qx - tX[bj]));
                float v1 = m == 2 ? 0.f : gg * (sgn * (qy - tY[bj]));
                float v2 = m == 3 ? 0.f : gg * (sgn * (qz - tZ[bj]));
                const int xi = dir ? q : bj;  // which source point the gradient lands on
                const float x0 = __ldg(src + 3 * xi), x1 = __ldg(src + 3 * xi + 1),
                            x2 = __ldg(src + 3 * xi + 2);
                acc[0] += v0 * x0; acc[1] += v0 * x1; acc[2] += v0 * x2;
                acc[3] += v1 * x0; acc[4] += v1 * x1; acc[5] += v1 * x2;
                acc[6] += v2 * x0; acc[7] += v2 * x1; acc[8] += v2 * x2;
                acc[9] += v0; acc[10] += v1; acc[11] += v2;
            }
        }
        block_sum<12>(acc, S);

        // ---- (i) closed-form pose backward + Adam on the 8 scalars
        if (t == 0) {
            float GR[3][3], gT[3], g[8];
#pragma unroll
            for (int i = 0; i < 3; ++i) {
#pragma unroll
                for (int j = 0; j < 3; ++j) GR[i][j] = acc[i * 3 + j];
                gT[i] = acc[9 + i];
            }
            pose_backward(st, GR, gT, A.tran_scale, g);
            adam_step(st, g, A, b1t, b2t);
        }
        // next iteration's (a) is ordered after this by block_sum's barriers + the one below
        __syncthreads();
    }

    if (t == 0) state_store(st, A, row);
}

static int g_pose_variant = -1;
constexpr int kDefaultPoseVariant = 20;  // 20: two rows per CTA (pose_duo.cu); 10..17: one row per CTA (pose_sym.cu); 0..5: two directed scans (this file)

template <int kQ, int H, int U>
static int launch_pose(const PoseArgs &a, cudaStream_t st) {
    static unsigned long long smem_set = 0;  // per-device flag bits
    if (int e = ensure_dyn_smem(pose_kernel<kQ, H, U>, (int)sizeof(Smem), smem_set)) return e;
    pose_kernel<kQ, H, U><<<a.rows, kThreads, sizeof(Smem), st>>>(a);
    return check_launch();
}

}  // namespace houv

using namespace houv;

extern "C" int houv_set_pose_variant(int v) {
    int old = g_pose_variant;
    g_pose_variant = v;
    return old;
}

extern "C" int houv_pose_optimise_ex(const HouvPoseProblem *q, void *stream) {
    if (!q) return HOUV_E_BADARG;
    int rows = q->rows;
    if (q->P < 0 || q->N <= 0 || q->M <= 0 || q->K <= 0 || q->iters < 0 || q->step0 < 0 || rows < 0)
        return HOUV_E_BADARG;
    if (rows == 0) rows = q->P * q->K;
    if (rows != q->P * q->K && !q->row_pair) return HOUV_E_BADARG;
    if (q->P == 0 || q->iters == 0 || rows == 0) return 0;
    if (!q->src || !q->tgt || (!q->params && !q->params64) || !q->min1_out || !q->R_out || !q->T_out)
        return HOUV_E_BADARG;
    if (q->N > kPts || q->M > kPts) return HOUV_E_TOOBIG;
    // calc_cd_percent takes k = int(points*percent) from BOTH lists and the view terms use
    // k = N (model_utils_completion.py:85-92); with N != M torch.topk would raise or trim.
    if (q->N != q->M) return HOUV_E_BADARG;
    PoseArgs a;
    a.src = q->src; a.tgt = q->tgt; a.params = q->params; a.adam_m = q->adam_m; a.adam_v = q->adam_v;
    a.params64 = q->params64; a.adam_m64 = q->adam_m64; a.adam_v64 = q->adam_v64;
    a.min1_out = q->min1_out; a.loss_out = q->loss_out; a.R_out = q->R_out; a.T_out = q->T_out;
    a.tran_s_out = q->tran_s_out; a.trace = q->trace; a.row_pair = q->row_pair; a.row_base = q->row_base;
    a.row_scale = q->row_scale;
    a.only_unpaired = 0;
    a.row_begin = 0;
    a.grid_rows = 0;
    a.P = q->P; a.N = q->N; a.M = q->M; a.K = q->K; a.rows = rows;
    a.step0 = q->step0; a.iters = q->iters; a.angle_base = q->angle_base;
    a.k3d = (int)((double)q->N * (double)q->percent);
    if (a.k3d < 1 || a.k3d > q->N) return HOUV_E_BADARG;
    a.lr = q->lr; a.beta1 = q->beta1; a.beta2 = q->beta2; a.eps = q->eps;
    a.grad_scale = q->grad_scale; a.w3d = q->w3d; a.wview = q->wview;
    a.tran_scale = q->tran_scale; a.tran_offset = q->tran_offset;
    cudaStream_t st = (cudaStream_t)stream;
    if (g_pose_variant < 0) {  // not set through the API: HOUV_POSE_VARIANT, else the default
        const char *e = getenv("HOUV_POSE_VARIANT");
        g_pose_variant = e ? atoi(e) : kDefaultPoseVariant;
    }
    switch (g_pose_variant) {
        case 20:  // two rows of a pair per CTA (pose_duo.cu); a loss without view terms keeps the
                  // one-metric kernel, which already runs two CTAs per SM
            return a.wview == 0.f ? launch_pose_sym(a, 84, st) : launch_pose_duo(a, false, st);
        case 21:  // the same with the column reduction pipelined under the sweep
            return a.wview == 0.f ? launch_pose_sym(a, 84, st) : launch_pose_duo(a, true, st);
        case 12: return launch_pose_sym(a, 8, st);     // one row per CTA, two rows evaluated before their folds
        case 13: return launch_pose_sym(a, 84, st);    // one row per CTA, four rows before their folds (round-1 default)
        case 15: return launch_pose_sym(a, 1004, st);  // four-metric sweep even when wview == 0
        case 16: return launch_pose_sym(a, 1008, st);  // variant 12 with the (distance, index) recovery
        case 17: return launch_pose_sym(a, 1009, st);  // ... with it only where no minimum is stored
        case 0: return launch_pose<4, 1, 1>(a, st);    // two directed scans
        // the other unroll / query-blocking variants measured in round 1 (1-5, 10, 11, 14:
        // profiles/r1_ncu_summary.md) lost and were pruned from the build
        default: return HOUV_E_BADARG;
    }
}

extern "C" int houv_pose_optimise(const float *src, const float *tgt, int P, int N, int M, int K,
                                  float *params, float *adam_m, float *adam_v, int step0,
                                  int iters, int angle_base, float lr, float beta1, float beta2,
                                  float eps, float grad_scale, float percent, float w3d,
                                  float wview, float *min1_out, float *loss_out, float *R_out,
                                  float *T_out, float *trace, const int *row_pair,
                                  const int *row_base, int rows, void *stream) {
    HouvPoseProblem q = {};
    q.src = src; q.tgt = tgt; q.P = P; q.N = N; q.M = M; q.K = K;
    q.params = params; q.adam_m = adam_m; q.adam_v = adam_v;
    q.step0 = step0; q.iters = iters; q.angle_base = angle_base;
    // the float arguments are roundings of the caller's doubles (0.9f etc.): widen the decimal
    // value back where it is one of torch's defaults so that 1-beta is what torch computes
    q.lr = (double)lr; q.beta1 = beta1 == 0.9f ? 0.9 : (double)beta1;
    q.beta2 = beta2 == 0.999f ? 0.999 : (double)beta2; q.eps = eps == 1e-8f ? 1e-8 : (double)eps;
    if (lr == 0.01f) q.lr = 0.01;
    q.grad_scale = grad_scale; q.percent = percent; q.w3d = w3d; q.wview = wview;
    q.tran_scale = 0.125f; q.tran_offset = 0.125f;  // houv.py:99
    q.min1_out = min1_out; q.loss_out = loss_out; q.R_out = R_out; q.T_out = T_out;
    q.trace = trace; q.row_pair = row_pair; q.row_base = row_base; q.rows = rows;
    return houv_pose_optimise_ex(&q, stream);
}
