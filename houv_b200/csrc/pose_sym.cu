// K3 (symmetric scan) -- fused HOUV pose optimisation for sm_100a, one CTA of 512 threads
// keeps one start (row of the multi-start batch) resident for ALL iterations, one CTA per SM.
//
// Same contract and same per-iteration semantics as pose_opt.cu (predict_model's hot loop,
// /root/reference/registration/models/houv.py:120-126: HOUV.forward :94-103, Predict_loss
// :209-222 = 4 x chamfer_3DDist forward [chamfer3D.cu:12-154] + top-k trimmed sqrt-means
// [model_utils_completion.py:83-100], autograd backward [chamfer3D.cu:155-195], Adam :118).
//
// What is different: the reference's two directed searches of one Chamfer call evaluate the SAME
// N x M matrix -- d(i,j) of "src_i queries tgt_j" is (t-y)^2 summed and of "tgt_j queries src_i"
// is (y-t)^2 summed in the same order, and fl(a-b) == -fl(b-a), so both are bit-identical.
// This kernel evaluates every d(i,j) ONCE (9 packed lane-ops for the four metrics) and folds
// it into a row minimum (src_i's nearest tgt) and a column minimum (tgt_j's nearest src):
//
//  * a thread owns 8 source rows in registers for the whole sweep; target columns stream past
//    the lanes in a rotation: in every step a lane evaluates its 8 rows against ONE pair of
//    columns (FADD2/FMUL2/FFMA2, the two columns are the two packed halves), and the running
//    column minima of that pair travel with it from lane to lane (8 SHFL per 16 evaluations);
//    the rotation runs inside 8-lane groups, so a finished column minimum covers 64 rows;
//  * row minima: one FMNMX3 per metric and row per step; column minima: FMNMX3 over row pairs.
//    Issue cost 72 packed + 64 FMNMX3 + 8 SHFL + 3 LDS.64 per 16 evaluations (~9.4 slots per
//    evaluation against 2 x 8.9 in the two-scan kernel);
//  * arg-min bookkeeping is by 64-wide groups, exact index recovered later only where the
//    gradient needs it: rows remember the 64-column superblock that last strictly improved
//    them (checked once per superblock against a shadow copy in shared memory); columns get
//    the FIRST 64-row sub-group (ascending, strict '<') holding their minimum from a per-
//    superblock reduction of the sub-groups' partial minima.  Ascending order + strict '<' at
//    every level = the reference's first-minimum rule (chamfer3D.cu:36,46,126);
//  * threads 0..255 sweep the lower half of the superblocks, threads 256..511 the upper half
//    with the same rows; the two row results are merged with the lower half winning ties.
//
// The one-metric instantiation (kNM = 1) uses 107 KB and 64 registers: two CTAs per SM.
//
// Everything after the scan (radix select of the k-th smallest, the min() side selection,
// gradient through the winning side, closed-form pose backward, Adam) is as in pose_opt.cu.
#include <type_traits>

#include "pose_common.cuh"
#include "pose_sweep.cuh"

namespace houv {

constexpr int kSymThreads = 512;
constexpr int kRowsPerThread = 8;

// kNM = metrics swept (4: Predict_loss, 1: the 3-D term alone).  The four-metric layout fills an SM
// (225 KB, one CTA); the one-metric layout is 107 KB so that two CTAs share an SM and one row's
// select / recovery / pose phases run under the other row's sweep.
template <int kNM>
struct SmemSymT {
    static constexpr int kSlots = 2 * kNM;  // float4 chunks of row minima per thread
    // NB the gradient phase addresses y and z as x + kPts and x + 2*kPts: keep the order
    float tx[kPts], ty[kPts], tz[kPts];   // target cloud (columns), fixed
    float yx[kPts], yy[kPts], yz[kPts];   // transformed source (rows) of this iteration
    float d3[2][kPts];                    // 3-D metric: [0] = column minima (dist1), [1] = row minima (dist2)
    unsigned char gid[kNM][2][kPts];      // winning group: [m][0][col] = 64-row sub-group, [m][1][row] = superblock
    unsigned char rgid1[kNM][kPts];       // rows: winning superblock seen by the upper column half
    float part[2][32][kNM][kSB];          // [column half][sub-group][metric][column]: partial column minima
    // row minima as of the last superblock boundary: per thread kSlots float4 chunks, chunk c = values
    // k = 4c..4c+3 (k = metric*8 + r), stored at position c ^ (thread & (kSlots-1)) so that the LDS.128 /
    // STS.128 of neighbouring threads spread over the bank groups
    float4 shadow[kSymThreads][kSlots];
    int hist[2][256];
    float red[kSymThreads / 32][16];
    float pose[12];
    unsigned selu[8];
    double state[80];                     // PoseState of this row (touched by thread 0 only)
    double bt[2];
};

template <int NV, class SmemSym>
__device__ __forceinline__ void block_sum_sym(float (&v)[NV], SmemSym &S) {
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
        for (int ww = 0; ww < kSymThreads / 32; ++ww) s += S.red[ww][i];
        v[i] = s;
    }
}

// kSymUnroll: unroll of the 8 rotation steps of a phase; kRecov: arg-min recovery -- 0 = running
// (distance, index) search everywhere; 1 = equality with the stored minimum where there is one;
// 2 = 1 + elsewhere candidates evaluated once into registers, minimum by FMNMX3, then equality
template <int kSymUnroll, int kRG, int kNM, int kRecov>
__global__ void __launch_bounds__(kSymThreads, kNM == 1 ? 2 : 1) pose_sym_kernel(const PoseArgs A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    using SmemSym = SmemSymT<kNM>;
    constexpr int kSwz = SmemSym::kSlots - 1;  // shadow chunk swizzle mask
    SmemSym &S = *reinterpret_cast<SmemSym *>(smem_raw);
    const int t = threadIdx.x;
    const int row = A.row_begin + blockIdx.x;
    if (A.only_unpaired && row_is_paired(A, row)) return;
    const int pair = A.row_pair ? A.row_pair[row] : row / A.K;
    const int N = A.N, M = A.M;  // rows = source points, columns = target points
    const float *src = A.src + (size_t)pair * N * 3;
    const float *tgt = A.tgt + (size_t)pair * M * 3;

    // geometry of the sweep
    const int half = t >> 8;            // which half of the superblocks this thread sweeps
    const int u = t & 255;              // thread within the half: owns rows 8u .. 8u+7
    const int w = u >> 5, l = u & 31;   // warp within the half, lane
    const int gq = l >> 3, ll = l & 7;  // 8-lane rotation group, lane within it
    const int sg = w * 4 + gq;          // 64-row sub-group: rows 64*sg .. 64*sg+63
    const int nsb = (M + kSB - 1) / kSB;          // superblocks holding real columns
    const int nsbh = (nsb + 1) / 2;               // rounds: superblocks per half
    const int nsg = (N + kSB - 1) / kSB;          // sub-groups holding real rows
    const bool scan_active = 256 * w < N;         // warp-uniform: this warp owns real rows
    const int row0 = 8 * u;
    int pad = 2 * nsbh * kSB;                     // columns touched by the sweep
    const int rpad = (N + 255) / 256 * 256;       // rows touched by the sweep
    pad = pad > rpad ? pad : rpad;                // (<= kPts)

    // target cloud -> shared SoA, +inf padded in x and y so every metric of a padded point is inf
    for (int j = t; j < pad; j += kSymThreads) {
        const bool ok = j < M;
        S.tx[j] = ok ? __ldg(tgt + 3 * j + 0) : kInf;
        S.ty[j] = ok ? __ldg(tgt + 3 * j + 1) : kInf;
        S.tz[j] = ok ? __ldg(tgt + 3 * j + 2) : 0.f;
    }

    static_assert(sizeof(PoseState) <= sizeof(S.state), "PoseState must fit SmemSym::state");
    PoseState &st = *reinterpret_cast<PoseState *>(S.state);
    double &b1t = S.bt[0], &b2t = S.bt[1];
    if (t == 0) {
        state_load(st, A, row);
        b1t = pow(A.beta1, (double)A.step0);
        b2t = pow(A.beta2, (double)A.step0);
    }
    const int abase = A.row_base ? A.row_base[row] : A.angle_base;
    const float boff = (float)((double)abase * 3.14159274101257324 / 4.0);
    const int k3d = A.k3d;
    // d(mean over the batch)/d(row): per row when rows of several batches share the launch
    const float gscale = A.row_scale ? A.row_scale[row] : A.grad_scale;

    // role of this thread in the per-superblock column reduction: metric rm, column rc of the
    // 128 columns (64 per half) finished in a round
    const int rm = t >> 7, rh = (t >> 6) & 1, rc = t & 63;

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
        for (int i = t; i < pad; i += kSymThreads) {
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
        // row bookkeeping of this sweep: shadow = +inf, winning superblock = 0
#pragma unroll
        for (int c = 0; c < 2 * kNM; ++c) S.shadow[t][c ^ (t & kSwz)] = make_float4(kInf, kInf, kInf, kInf);
        for (int i = t; i < kNM * 2 * kPts / 4; i += kSymThreads)
            reinterpret_cast<unsigned *>(&S.gid[0][0][0])[i] = 0u;
        for (int i = t; i < kNM * kPts / 4; i += kSymThreads)
            reinterpret_cast<unsigned *>(&S.rgid1[0][0])[i] = 0u;
        __syncthreads();

        // ---- (c) the symmetric sweep
        float nqx[8], nqy[8], nqz[8], best[kNM][8];
        if (scan_active) {
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                nqx[r] = -S.yx[row0 + r]; nqy[r] = -S.yy[row0 + r]; nqz[r] = -S.yz[row0 + r];
            }
        }
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int m = 0; m < kNM; ++m) best[m][r] = kInf;
        float vcol = 0.f;  // sum sqrt(column minimum) of view metric rm (reducer role)

#pragma unroll 1
        for (int rr = 0; rr < nsbh; ++rr) {
            const int sb = half * nsbh + rr;  // superblock swept by this half in this round
            if (scan_active) {
                const int sbbase = sb * kSB;
#pragma unroll 1
                for (int ph = 0; ph < 4; ++ph) {
                    const int cb = (gq + ph) & 3;  // 16-column block of this group in this phase
                    float cmx[kNM], cmy[kNM];  // running column minima of the pair in hand
#pragma unroll
                    for (int m = 0; m < kNM; ++m) { cmx[m] = kInf; cmy[m] = kInf; }
                    const int cbase = sbbase + 16 * cb;
#pragma unroll kSymUnroll
                    for (int s = 0; s < 8; ++s) {
                        const int c = cbase + 2 * ((ll + s) & 7);
                        const float2 X = *reinterpret_cast<const float2 *>(&S.tx[c]);
                        const float2 Y = *reinterpret_cast<const float2 *>(&S.ty[c]);
                        const float2 Z = *reinterpret_cast<const float2 *>(&S.tz[c]);
                        sym_step<kRG, kNM>(X, Y, Z, nqx, nqy, nqz, best, cmx, cmy);
                        if (s < 7) {  // the column pair moves on to lane ll-1; ours comes from ll+1
#pragma unroll
                            for (int m = 0; m < kNM; ++m) {
                                cmx[m] = __shfl_sync(0xffffffffu, cmx[m], ll + 1, 8);
                                cmy[m] = __shfl_sync(0xffffffffu, cmy[m], ll + 1, 8);
                            }
                        }
                    }
                    // after step 7 this lane holds the finished pair (ll+7)&7 of block cb
                    const int pc = 16 * cb + 2 * ((ll + 7) & 7);
#pragma unroll
                    for (int m = 0; m < kNM; ++m)
                        *reinterpret_cast<float2 *>(&S.part[half][sg][m][pc]) = make_float2(cmx[m], cmy[m]);
                }
                // rows: did this superblock strictly improve the running minimum?  One bit per
                // (metric, row); the shadow is refreshed unconditionally (best <= shadow always)
                unsigned improved = 0;
#pragma unroll
                for (int c = 0; c < 2 * kNM; ++c) {
                    const int m = c >> 1, r0 = 4 * (c & 1);
                    float4 &slot = S.shadow[t][c ^ (t & kSwz)];
                    const float4 pv = slot;
                    improved |= (best[m][r0 + 0] < pv.x ? 1u : 0u) << (4 * c + 0);
                    improved |= (best[m][r0 + 1] < pv.y ? 1u : 0u) << (4 * c + 1);
                    improved |= (best[m][r0 + 2] < pv.z ? 1u : 0u) << (4 * c + 2);
                    improved |= (best[m][r0 + 3] < pv.w ? 1u : 0u) << (4 * c + 3);
                    slot = make_float4(best[m][r0], best[m][r0 + 1], best[m][r0 + 2], best[m][r0 + 3]);
                }
                while (improved) {  // rarer and rarer as the sweep proceeds
                    const int k = __ffs(improved) - 1;
                    improved &= improved - 1;
                    if (half) S.rgid1[k >> 3][row0 + (k & 7)] = (unsigned char)sb;
                    else S.gid[k >> 3][1][row0 + (k & 7)] = (unsigned char)sb;
                }
            }
            __syncthreads();
            // columns: first sub-group (ascending rows) holding the minimum
            if (rm < kNM) {
                const int col = (rh * nsbh + rr) * kSB + rc;
                float bv = kInf;
                int bs = 0;
                for (int g = 0; g < nsg; ++g) {
                    const float v = S.part[rh][g][rm][rc];
                    if (v < bv) { bv = v; bs = g; }
                }
                S.gid[rm][0][col] = (unsigned char)bs;
                if (rm == 0) S.d3[0][col] = bv;
                else if (col < M) vcol += sqrtf(bv);
            }
            __syncthreads();
        }

        // ---- (d) rows: merge the two column halves (lower half wins ties = first minimum)
        float vs[2][3] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};  // sum sqrt(d) of the view terms [dir][view]
#pragma unroll
        for (int m = 0; m < 3; ++m) vs[0][m] = (rm == m + 1) ? vcol : 0.f;
        if (half == 0 && scan_active) {
#pragma unroll
            for (int m = 0; m < kNM; ++m)
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    const int i = row0 + r;
                    float b = best[m][r];
                    const float4 o4 = S.shadow[t + 256][(2 * m + (r >> 2)) ^ (t & kSwz)];  // (t+256)&7 == t&7
                    const float b1 = (r & 3) == 0 ? o4.x : (r & 3) == 1 ? o4.y : (r & 3) == 2 ? o4.z : o4.w;
                    if (b1 < b) {
                        b = b1;
                        S.gid[m][1][i] = S.rgid1[m][i];
                    }
                    if (i < N) {
                        if (m == 0) S.d3[1][i] = b;
                        else vs[1][m - 1] += sqrtf(b);
                    }
                }
        }
        __syncthreads();

        // ---- (e) k-th smallest of d3[0][0..M) and d3[1][0..N): 4 x 8-bit radix select
        unsigned prefix[2] = {0u, 0u};
        int kk[2] = {k3d, k3d};
#pragma unroll 1
        for (int shift = 24; shift >= 0; shift -= 8) {
            for (int b = t; b < 512; b += kSymThreads) (&S.hist[0][0])[b] = 0;
            __syncthreads();
#pragma unroll
            for (int dir = 0; dir < 2; ++dir) {
                const int len = dir ? N : M;
                if (shift == 24) {
                    // sign + 7 exponent bits: a cloud's distances share one or two values, so the
                    // lanes of a warp would queue on the same counter -- one add per distinct digit
                    for (int q0 = 0; q0 < len; q0 += kSymThreads) {
                        const int q = q0 + t;
                        const unsigned bin = q < len ? __float_as_uint(S.d3[dir][q]) >> 24 : 256u;
                        const unsigned peers = __match_any_sync(0xffffffffu, bin);
                        if (bin < 256u && (t & 31) == __ffs(peers) - 1)
                            atomicAdd(&S.hist[dir][bin], __popc(peers));
                    }
                    continue;
                }
                for (int q = t; q < len; q += kSymThreads) {
                    const unsigned bits = __float_as_uint(S.d3[dir][q]);
                    const bool match = (shift == 24) ||
                                       ((bits >> (shift + 8)) == (prefix[dir] >> (shift + 8)));
                    if (match) atomicAdd(&S.hist[dir][(bits >> shift) & 255u], 1);
                }
            }
            __syncthreads();
            if (t < 64) {
                const int dir = t >> 5, ln = t & 31;
                int c[8], s = 0;
#pragma unroll
                for (int b = 0; b < 8; ++b) { c[b] = S.hist[dir][8 * ln + b]; s += c[b]; }
                int incl = s;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    int n = __shfl_up_sync(0xffffffffu, incl, o);
                    if (ln >= o) incl += n;
                }
                const int excl = incl - s;
                if (excl < kk[dir] && kk[dir] <= incl) {
                    int rem = kk[dir] - excl, bin = 0, cnt = 0;
                    bool done = false;
#pragma unroll
                    for (int b = 0; b < 8; ++b) {
                        if (!done) {
                            if (rem <= c[b]) { bin = 8 * ln + b; cnt = c[b]; done = true; }
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
            for (int q = t; q < len; q += kSymThreads) {
                const float d = S.d3[dir][q];
                if (d < thr[dir]) sums[dir] += sqrtf(d);
            }
#pragma unroll
            for (int m = 0; m < 3; ++m) sums[2 + dir * 3 + m] = vs[dir][m];
        }
        block_sum_sym<8>(sums, S);
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

        // ---- (h) gradient: only the side that won each min(); the exact arg-min is recovered
        //          from the 64 candidates of the winning group.  Candidates are visited in a
        //          lane-skewed order (different lanes hold different groups: same offset = same
        //          bank), so the first minimum is taken lexicographically on (distance, index).
        float acc[12];
#pragma unroll
        for (int i = 0; i < 12; ++i) acc[i] = 0.f;
        const int lane = t & 31;
#pragma unroll
        for (int m = 0; m < kNM; ++m) {
            if (m > 0 && A.wview == 0.f) continue;  // loss = w3d*min_1 only (train_utils.py:439)
            auto side_grad = [&](auto dirc) {
            constexpr int dir = decltype(dirc)::value;  // compile-time: smem bases become immediates
            // x, y, z of a cloud are consecutive kPts-float arrays: one base pointer per cloud
            const float *qB = dir ? S.yx : S.tx, *tB = dir ? S.tx : S.yx;
            const int nq = dir ? N : M;
            // d(mean over rows)/d(row) * weight / k      houv.py:124,222; m_u_c.py:94-95
            const float coef = gscale * (m == 0 ? A.w3d : A.wview) / (float)(m == 0 ? k3d : N);
            // The minimum itself is still on chip for the 3-D metric (d3) and for every row metric
            // (the two halves' shadows hold their final row minima): there the candidates are
            // searched for EQUALITY with it (smallest matching index), which needs no running
            // (distance, index) pair.  Column minima of the view metrics were only summed.
            const bool known = kRecov >= 1 && (m == 0 || dir == 1);
            for (int q = t; q < nq; q += kSymThreads) {
                float wt = 1.f;
                float bd = kInf;
                if (m == 0) {  // top-k membership of the trimmed 3-D term
                    const float d = S.d3[dir][q];
                    wt = d < thr[dir] ? 1.f : (d == thr[dir] ? tiew[dir] : 0.f);
                    if (wt == 0.f) continue;
                    bd = d;
                } else if (known) {
                    const int uq = q >> 3, rq = q & 7;
                    const int slot = ((2 * m + (rq >> 2)) ^ (uq & kSwz)) * 4 + (rq & 3);
                    const float *sh = reinterpret_cast<const float *>(&S.shadow[0][0]);
                    bd = fminf(sh[uq * (8 * kNM) + slot], sh[(uq + 256) * (8 * kNM) + slot]);
                }
                const int j0 = (int)S.gid[m][dir][q] * kSB;
                const unsigned j0b = (unsigned)j0 * 4u, lane8 = (unsigned)lane * 8u;
                const char *tb = reinterpret_cast<const char *>(tB);
                const float qx = qB[q], qy = qB[kPts + q], qz = qB[2 * kPts + q];
                int bj = j0;
                if (known) {
                    unsigned mx = 0u, my = 0u;  // bit k: the pair visited at step k equals the minimum
#pragma unroll
                    for (int k = 0; k < kSB / 2; ++k) {
                        // byte offset of pair (k + lane) & 31 of the group: one add, one and-or
                        const unsigned off = ((lane8 + 8u * k) & 248u) | j0b;
                        const float2 X = *reinterpret_cast<const float2 *>(tb + off);
                        const float2 Y = *reinterpret_cast<const float2 *>(tb + 4 * kPts + off);
                        const float2 Z = *reinterpret_cast<const float2 *>(tb + 8 * kPts + off);
                        const float2 d = metric_dist2(m, X, Y, Z, qx, qy, qz);
                        if (d.x == bd) mx |= 1u << k;
                        if (d.y == bd) my |= 1u << k;
                    }
                    bj = j0 + first_match(mx, my, lane);  // a match always exists unless bd is NaN
                } else if (kRecov == 2) {
                    float2 dv[kSB / 2];
#pragma unroll
                    for (int k = 0; k < kSB / 2; ++k) {
                        const unsigned off = ((lane8 + 8u * k) & 248u) | j0b;
                        const float2 X = *reinterpret_cast<const float2 *>(tb + off);
                        const float2 Y = *reinterpret_cast<const float2 *>(tb + 4 * kPts + off);
                        const float2 Z = *reinterpret_cast<const float2 *>(tb + 8 * kPts + off);
                        dv[k] = metric_dist2(m, X, Y, Z, qx, qy, qz);
                    }
                    bd = kInf;
#pragma unroll
                    for (int k = 0; k < kSB / 2; ++k) bd = min3(bd, dv[k].x, dv[k].y);
                    unsigned mx = 0u, my = 0u;
#pragma unroll
                    for (int k = 0; k < kSB / 2; ++k) {
                        if (dv[k].x == bd) mx |= 1u << k;
                        if (dv[k].y == bd) my |= 1u << k;
                    }
                    bj = j0 + first_match(mx, my, lane);
                } else {
                    bd = kInf;
#pragma unroll 4
                    for (int k = 0; k < kSB / 2; ++k) {
                        const int c = j0 + 2 * ((k + lane) & (kSB / 2 - 1));
                        const float2 X = *reinterpret_cast<const float2 *>(&tB[c]);
                        const float2 Y = *reinterpret_cast<const float2 *>(&tB[kPts + c]);
                        const float2 Z = *reinterpret_cast<const float2 *>(&tB[2 * kPts + c]);
                        const float2 d = metric_dist2(m, X, Y, Z, qx, qy, qz);
                        // first minimum = lexicographic minimum of (distance, index)
                        if (d.x < bd || (d.x == bd && c < bj)) { bd = d.x; bj = c; }
                        if (d.y < bd || (d.y == bd && c + 1 < bj)) { bd = d.y; bj = c + 1; }
                    }
                }
                // sqrt backward then chamfer backward (x2): d=0 gives inf*0 = NaN like autograd
                const float gg = (coef * wt) / (2.f * sqrtf(bd)) * 2.f;
                const float sgn = dir ? 1.f : -1.f;  // (transformed src point) - (tgt point)
                float v0 = m == 1 ? 0.f : gg * (sgn * (qx - tB[bj]));
                float v1 = m == 2 ? 0.f : gg * (sgn * (qy - tB[kPts + bj]));
                float v2 = m == 3 ? 0.f : gg * (sgn * (qz - tB[2 * kPts + bj]));
                const int xi = dir ? q : bj;  // which source point the gradient lands on
                const float x0 = __ldg(src + 3 * xi), x1 = __ldg(src + 3 * xi + 1),
                            x2 = __ldg(src + 3 * xi + 2);
                acc[0] += v0 * x0; acc[1] += v0 * x1; acc[2] += v0 * x2;
                acc[3] += v1 * x0; acc[4] += v1 * x1; acc[5] += v1 * x2;
                acc[6] += v2 * x0; acc[7] += v2 * x1; acc[8] += v2 * x2;
                acc[9] += v0; acc[10] += v1; acc[11] += v2;
            }
            };
            if (side[m]) side_grad(std::integral_constant<int, 1>{});
            else side_grad(std::integral_constant<int, 0>{});
        }
        block_sum_sym<12>(acc, S);

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
        __syncthreads();
    }

    if (t == 0) state_store(st, A, row);
}

template <int kU, int kRG, int kNM, int kRecov = 2>
static int launch_sym(const PoseArgs &a, cudaStream_t st) {
    static unsigned long long smem_set = 0;  // per-device flag bits
    if (int e = ensure_dyn_smem(pose_sym_kernel<kU, kRG, kNM, kRecov>, (int)sizeof(SmemSymT<kNM>), smem_set))
        return e;
    // (the two-row launcher hands the rows it did not take to this kernel: grid_rows of them)
    pose_sym_kernel<kU, kRG, kNM, kRecov><<<a.grid_rows > 0 ? a.grid_rows : a.rows, kSymThreads,
                                            sizeof(SmemSymT<kNM>), st>>>(a);
    return check_launch();
}

int launch_pose_sym(const PoseArgs &a, int unroll, cudaStream_t st) {
    // a loss without view terms (wview == 0: train_utils.getPredict_angle) needs the 3-D metric only
    if (unroll == 1008)  // the (distance, index) recovery everywhere: kept to test the others against
        return a.wview == 0.f ? launch_sym<8, 2, 1, 0>(a, st) : launch_sym<8, 2, 4, 0>(a, st);
    if (unroll == 1009)  // equality search where the minimum is stored, (distance, index) elsewhere
        return a.wview == 0.f ? launch_sym<8, 2, 1, 1>(a, st) : launch_sym<8, 2, 4, 1>(a, st);
    if (a.wview == 0.f && unroll != 1004)
        return unroll == 84 ? launch_sym<8, 4, 1>(a, st) : launch_sym<8, 2, 1>(a, st);
    switch (unroll) {
        case 8: case 1004: return launch_sym<8, 2, 4>(a, st);
        case 84: return launch_sym<8, 4, 4>(a, st);
        default: return HOUV_E_BADARG;  // (unrolls 2 / 4 of round 1 were pruned)
    }
}

}  // namespace houv
