// Shared pieces of the K3 pose-optimisation kernels (pose_opt.cu: two directed scans per
// iteration; pose_sym.cu: one symmetric scan serving both directions).
#pragma once
#include "common.cuh"

namespace houv {

constexpr int kPts = 2048;       // max points per cloud handled by the fused kernel
constexpr int kThreads = 256;
constexpr float kInf = __builtin_huge_valf();

struct PoseState;
struct PoseArgs {
    const float *src;   // [P,N,3]
    const float *tgt;   // [P,M,3]
    float *params;      // [rows,8] in/out: V_c[3], angle_c, tran_c[3], tran_s
    float *adam_m;      // [rows,8] in/out or null
    float *adam_v;      // [rows,8] in/out or null
    float *min1_out;    // [rows]   3-D trimmed term of the LAST forward
    float *loss_out;    // [rows]   6*min_1 + sum of view terms of the last forward (or null)
    float *R_out;       // [rows,9]
    float *T_out;       // [rows,3]
    float *trace;       // [iters,rows,2] (loss, min_1) per forward, or null
    const int *row_pair;  // [rows] pair of each row, or null (= row / K)
    const int *row_base;  // [rows] angle_base of each row, or null (= angle_base)
    int P, N, M, K, rows;
    int step0, iters, angle_base, k3d;
    double lr, beta1, beta2, eps;  // Adam; rounded to float for float32 state as torch does
    float grad_scale, w3d, wview;
    // |T| = sin(tran_s*pi)*tran_scale + tran_offset: 0.125/0.125 in houv.py:99, 1/0 in
    // train_utils.getPredict_angle (train_utils.py:405)
    float tran_scale, tran_offset;
    // float64 parameters and Adam moments (getPredict_angle keeps float64 leaves on the host and
    // feeds their float32 roundings to the forward, train_utils.py:382-401); null = float32 state
    double *params64, *adam_m64, *adam_v64;
    float *tran_s_out;  // [rows] |T| of the last forward, or null
    const float *row_scale;  // [rows] per-row grad_scale, or null (= grad_scale)
    // one-row kernels only: skip every row that pose_duo_kernel took (rows 2b, 2b+1 of one pair)
    int only_unpaired;
    int row_begin;  // one-row kernels only: CTA b works on row row_begin + b
    int grid_rows;  // one-row kernels only: CTAs to launch (0 = rows)
};

__device__ __forceinline__ bool row_is_paired(const PoseArgs &A, int row) {
    const int other = row ^ 1;
    if (other >= A.rows) return false;
    return A.row_pair ? A.row_pair[row] == A.row_pair[other] : row / A.K == other / A.K;
}

// squared distance of metric m between (tx,ty,tz) and (qx,qy,qz), scalar, same
// arithmetic as the packed scan (and as the reference on masked clouds)
__device__ __forceinline__ float metric_dist(int m, float tx, float ty, float tz, float qx,
                                             float qy, float qz) {
    float dx = __fsub_rn(tx, qx), dy = __fsub_rn(ty, qy), dz = __fsub_rn(tz, qz);
    float yy = __fmul_rn(dy, dy);
    if (m == 0) return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, yy));
    if (m == 1) return __fmaf_rn(dz, dz, yy);
    if (m == 2) return __fmaf_rn(dz, dz, __fmul_rn(dx, dx));
    return __fmaf_rn(dx, dx, yy);
}

// the same for two targets at once (packed FADD2/FMUL2/FFMA2)
__device__ __forceinline__ float2 metric_dist2(int m, float2 tx, float2 ty, float2 tz, float qx,
                                               float qy, float qz) {
    const float2 dz = __fadd2_rn(tz, make_float2(-qz, -qz));
    if (m == 2) {
        const float2 dx = __fadd2_rn(tx, make_float2(-qx, -qx));
        return __ffma2_rn(dz, dz, __fmul2_rn(dx, dx));
    }
    const float2 dy = __fadd2_rn(ty, make_float2(-qy, -qy));
    const float2 yy = __fmul2_rn(dy, dy);
    if (m == 1) return __ffma2_rn(dz, dz, yy);
    const float2 dx = __fadd2_rn(tx, make_float2(-qx, -qx));
    const float2 xy = __ffma2_rn(dx, dx, yy);
    return m == 0 ? __ffma2_rn(dz, dz, xy) : xy;
}

// Per-row optimiser state, live in thread 0's registers for the whole kernel.
struct PoseState {
    double pd[8], md[8], vd[8];  // float64 state (only with PoseArgs::params64)
    float p[8], m[8], v[8];
    // forward intermediates kept for the closed-form backward
    float vh[3], n, A[3][3], A2[3][3], sa, ca, ang, u[3], nt, st;
};

__device__ __forceinline__ void pose_forward(PoseState &s, float boff, float tran_scale,
                                             float tran_offset, float *pose /*smem[12]*/) {
    const float pi = 3.14159274101257324f;  // float32 pi = acos(0)*2, houv.py:19
    // angle = sin(angle_c*pi)*pi/8 + pi/8 + angle_base*pi/4            houv.py:96
    s.ang = __fadd_rn(__fadd_rn(__fmul_rn(sinf(__fmul_rn(s.p[3], pi)), pi) / 8.f, pi / 8.f), boff);
    // V / sqrt(sum V^2)                                                 houv.py:71
    s.n = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(s.p[0], s.p[0]), __fmul_rn(s.p[1], s.p[1])),
                          __fmul_rn(s.p[2], s.p[2])));
#pragma unroll
    for (int i = 0; i < 3; ++i) s.vh[i] = s.p[i] / s.n;
    float (&A)[3][3] = s.A;
    A[0][0] = 0.f; A[0][1] = -s.vh[2]; A[0][2] = s.vh[1];
    A[1][0] = s.vh[2]; A[1][1] = 0.f; A[1][2] = -s.vh[0];
    A[2][0] = -s.vh[1]; A[2][1] = s.vh[0]; A[2][2] = 0.f;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            s.A2[i][j] = __fmaf_rn(A[i][2], A[2][j], __fmaf_rn(A[i][1], A[1][j], __fmul_rn(A[i][0], A[0][j])));
    s.sa = sinf(s.ang);
    s.ca = cosf(s.ang);
    const float omc = __fsub_rn(1.f, s.ca);
    // R = I + sin(a) A + (1-cos a) A^2                                   houv.py:85
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            pose[i * 3 + j] = __fadd_rn(__fadd_rn(i == j ? 1.f : 0.f, __fmul_rn(s.sa, A[i][j])),
                                        __fmul_rn(omc, s.A2[i][j]));
    // T = tran_c/|tran_c| * (sin(tran_s*pi)*0.125 + 0.125)               houv.py:88-100
    // (x*1 + 0 is exact, so scale 1 / offset 0 gives train_utils.py:405's sin(tran_s*pi)*1)
    s.st = __fadd_rn(__fmul_rn(sinf(__fmul_rn(s.p[7], pi)), tran_scale), tran_offset);
    s.nt = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(s.p[4], s.p[4]), __fmul_rn(s.p[5], s.p[5])),
                           __fmul_rn(s.p[6], s.p[6])));
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        s.u[i] = s.p[4 + i] / s.nt;
        pose[9 + i] = __fmul_rn(s.u[i], s.st);
    }
}

// Closed-form chain from dL/dR (GR), dL/dT (gT) to the 8 scalars.
__device__ __forceinline__ void pose_backward(const PoseState &s, const float (&GR)[3][3],
                                              const float (&gT)[3], float tran_scale,
                                              float (&g)[8]) {
    const float pi = 3.14159274101257324f;
    float dA = 0.f, dA2 = 0.f;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            dA += GR[i][j] * s.A[i][j];
            dA2 += GR[i][j] * s.A2[i][j];
        }
    const float dang = s.ca * dA + s.sa * dA2;  // d/da [sin a A + (1-cos a) A^2]
    const float omc = 1.f - s.ca;
    float GA[3][3];  // dL/dA = sin(a) GR + (1-cos a)(GR A^T + A^T GR)
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            float t = 0.f;
#pragma unroll
            for (int k = 0; k < 3; ++k) t += GR[i][k] * s.A[j][k] + s.A[k][i] * GR[k][j];
            GA[i][j] = s.sa * GR[i][j] + omc * t;
        }
    const float gv[3] = {GA[2][1] - GA[1][2], GA[0][2] - GA[2][0], GA[1][0] - GA[0][1]};
    const float vdot = s.vh[0] * gv[0] + s.vh[1] * gv[1] + s.vh[2] * gv[2];
#pragma unroll
    for (int i = 0; i < 3; ++i) g[i] = (gv[i] - s.vh[i] * vdot) / s.n;
    g[3] = dang * (pi / 8.f) * cosf(s.p[3] * pi) * pi;
    const float gs = gT[0] * s.u[0] + gT[1] * s.u[1] + gT[2] * s.u[2];
    const float udot = gs * s.st;  // u . (gT * st)
#pragma unroll
    for (int i = 0; i < 3; ++i) g[4 + i] = (gT[i] * s.st - s.u[i] * udot) / s.nt;
    g[7] = gs * tran_scale * cosf(s.p[7] * pi) * pi;
}

// torch.optim.Adam, no weight decay / amsgrad (houv.py:118, train_utils.py:390): bias corrections
// from the double hyper-parameters; the element-wise update in the precision of the state.
__device__ __forceinline__ void adam_step(PoseState &s, const float (&g)[8], const PoseArgs &A,
                                          double &b1t, double &b2t) {
    b1t *= A.beta1;
    b2t *= A.beta2;
    if (A.params64) {  // float64 leaves: gradients arrive as float32 and are widened exactly
        const double step_size = A.lr / (1.0 - b1t);
        const double bc2_sqrt = sqrt(1.0 - b2t);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const double gi = (double)g[i];
            s.md[i] = s.md[i] * A.beta1 + (1.0 - A.beta1) * gi;
            s.vd[i] = s.vd[i] * A.beta2 + (1.0 - A.beta2) * gi * gi;
            const double denom = sqrt(s.vd[i]) / bc2_sqrt + A.eps;
            s.pd[i] = s.pd[i] - step_size * (s.md[i] / denom);
            s.p[i] = (float)s.pd[i];  // .float() of the leaf, train_utils.py:397-401
        }
        return;
    }
    const float step_size = (float)(A.lr / (1.0 - b1t));
    const float bc2_sqrt = (float)sqrt(1.0 - b2t);
    const float beta1 = (float)A.beta1, beta2 = (float)A.beta2, eps = (float)A.eps;
    const float omb1 = (float)(1.0 - A.beta1), omb2 = (float)(1.0 - A.beta2);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        s.m[i] = __fadd_rn(__fmul_rn(s.m[i], beta1), __fmul_rn(omb1, g[i]));
        s.v[i] = __fadd_rn(__fmul_rn(s.v[i], beta2), __fmul_rn(omb2, __fmul_rn(g[i], g[i])));
        const float denom = __fadd_rn(sqrtf(s.v[i]) / bc2_sqrt, eps);
        s.p[i] = __fsub_rn(s.p[i], __fmul_rn(step_size, s.m[i] / denom));
    }
}

// load / store of the per-row state (thread 0)
__device__ __forceinline__ void state_load(PoseState &st, const PoseArgs &A, int row) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        if (A.params64) {
            st.pd[i] = A.params64[(size_t)row * 8 + i];
            st.md[i] = A.adam_m64 ? A.adam_m64[(size_t)row * 8 + i] : 0.0;
            st.vd[i] = A.adam_v64 ? A.adam_v64[(size_t)row * 8 + i] : 0.0;
            st.p[i] = (float)st.pd[i];
            st.m[i] = 0.f; st.v[i] = 0.f;
        } else {
            st.p[i] = A.params[(size_t)row * 8 + i];
            st.m[i] = A.adam_m ? A.adam_m[(size_t)row * 8 + i] : 0.f;
            st.v[i] = A.adam_v ? A.adam_v[(size_t)row * 8 + i] : 0.f;
        }
    }
}
__device__ __forceinline__ void state_store(const PoseState &st, const PoseArgs &A, int row) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        if (A.params64) {
            A.params64[(size_t)row * 8 + i] = st.pd[i];
            if (A.adam_m64) A.adam_m64[(size_t)row * 8 + i] = st.md[i];
            if (A.adam_v64) A.adam_v64[(size_t)row * 8 + i] = st.vd[i];
            if (A.params) A.params[(size_t)row * 8 + i] = st.p[i];
        } else {
            A.params[(size_t)row * 8 + i] = st.p[i];
            if (A.adam_m) A.adam_m[(size_t)row * 8 + i] = st.m[i];
            if (A.adam_v) A.adam_v[(size_t)row * 8 + i] = st.v[i];
        }
    }
}

int launch_pose_sym(const PoseArgs &a, int unroll, cudaStream_t st);  // pose_sym.cu
int launch_pose_duo(const PoseArgs &a, bool pipelined, cudaStream_t st);  // pose_duo.cu

}  // namespace houv
