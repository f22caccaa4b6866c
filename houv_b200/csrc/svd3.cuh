// Closed 3x3 Kabsch solve for one thread: R = V diag(1,1,det) U^T from the
// cross-covariance H = sum (x - xbar)(y - ybar)^T = U S V^T, i.e. what
// SVDHead.forward does with torch.svd + the reflection fix
// (/root/reference/registration/model_utils.py:232-240) and what Open3D's
// point-to-point estimation (Eigen::umeyama without scaling) computes.
//
// The SVD is a one-sided (Hestenes) Jacobi iteration on the 3 columns of H in
// double precision: no squaring of the condition number, a handful of sweeps,
// no library.  The work is a few hundred flops per pair; the kernels around it
// are latency-bound, so double costs nothing measurable.
#pragma once
#include <math.h>

namespace houv {

__device__ __forceinline__ void jacobi_pair(double (&A)[3][3], double (&V)[3][3], int p, int q,
                                            bool &rotated) {
    double alpha = 0, beta = 0, gamma = 0;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        alpha += A[i][p] * A[i][p];
        beta += A[i][q] * A[i][q];
        gamma += A[i][p] * A[i][q];
    }
    if (fabs(gamma) <= 1e-15 * sqrt(alpha * beta) || gamma == 0.0) return;
    rotated = true;
    const double zeta = (beta - alpha) / (2.0 * gamma);
    const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
    const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const double ap = A[i][p], aq = A[i][q];
        A[i][p] = c * ap - s * aq;
        A[i][q] = s * ap + c * aq;
        const double vp = V[i][p], vq = V[i][q];
        V[i][p] = c * vp - s * vq;
        V[i][q] = s * vp + c * vq;
    }
}

__device__ __forceinline__ double det3(const double (&M)[3][3]) {
    return M[0][0] * (M[1][1] * M[2][2] - M[1][2] * M[2][1]) -
           M[0][1] * (M[1][0] * M[2][2] - M[1][2] * M[2][0]) +
           M[0][2] * (M[1][0] * M[2][1] - M[1][1] * M[2][0]);
}

// H = U S V^T (S descending).  Returns R = V diag(1,1,sign) U^T with det R = +1.
// warm: optional orthogonal matrix to start the iteration from (in/out: the V found).  A caller that
// solves a slowly changing H again and again (ICP) then needs one or two sweeps instead of six or seven.
__device__ inline void kabsch_from_covariance(const double (&H)[3][3], double (&R)[3][3],
                                              double (*warm)[3] = nullptr) {
    double A[3][3], V[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            V[i][j] = warm ? warm[i][j] : (i == j ? 1.0 : 0.0);
            A[i][j] = H[i][j];
        }
    if (warm) {
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) A[i][j] = H[i][0] * V[0][j] + H[i][1] * V[1][j] + H[i][2] * V[2][j];
    }
    for (int sweep = 0; sweep < 30; ++sweep) {
        bool rotated = false;
        jacobi_pair(A, V, 0, 1, rotated);
        jacobi_pair(A, V, 0, 2, rotated);
        jacobi_pair(A, V, 1, 2, rotated);
        if (!rotated) break;
    }
    if (warm) {
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) warm[i][j] = V[i][j];
    }
    // A = U S (columns), sort columns by norm, descending
    double sg[3];
    int ord[3] = {0, 1, 2};
#pragma unroll
    for (int j = 0; j < 3; ++j) sg[j] = sqrt(A[0][j] * A[0][j] + A[1][j] * A[1][j] + A[2][j] * A[2][j]);
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < 2 - a; ++b)
            if (sg[ord[b]] < sg[ord[b + 1]]) { int tmp = ord[b]; ord[b] = ord[b + 1]; ord[b + 1] = tmp; }
    double U[3][3], W[3][3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        const int c = ord[j];
        const double s = sg[c];
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            W[i][j] = V[i][c];
            U[i][j] = s > 0 ? A[i][c] / s : 0.0;
        }
    }
    // rank deficiency: complete U to an orthonormal basis (direction of a null column is free)
    const double tiny = 1e-300;
    if (sg[ord[0]] <= tiny) {
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) U[i][j] = i == j ? 1.0 : 0.0;
    } else {
        if (sg[ord[1]] <= 1e-14 * sg[ord[0]]) {  // pick any unit vector orthogonal to u0
            int k = fabs(U[0][0]) < fabs(U[1][0]) ? (fabs(U[0][0]) < fabs(U[2][0]) ? 0 : 2)
                                                  : (fabs(U[1][0]) < fabs(U[2][0]) ? 1 : 2);
            double e[3] = {k == 0 ? 1.0 : 0.0, k == 1 ? 1.0 : 0.0, k == 2 ? 1.0 : 0.0};
            const double d = e[0] * U[0][0] + e[1] * U[1][0] + e[2] * U[2][0];
            double n = 0;
#pragma unroll
            for (int i = 0; i < 3; ++i) { e[i] -= d * U[i][0]; n += e[i] * e[i]; }
            n = sqrt(n);
#pragma unroll
            for (int i = 0; i < 3; ++i) U[i][1] = e[i] / n;
        }
        if (sg[ord[2]] <= 1e-14 * sg[ord[0]]) {  // u2 = u0 x u1
            U[0][2] = U[1][0] * U[2][1] - U[2][0] * U[1][1];
            U[1][2] = U[2][0] * U[0][1] - U[0][0] * U[2][1];
            U[2][2] = U[0][0] * U[1][1] - U[1][0] * U[0][1];
        }
    }
    // r = v u^T; if det r < 0: flip the last column of v (model_utils.py:236-239)
    const double sign = det3(W) * det3(U) < 0 ? -1.0 : 1.0;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            R[i][j] = W[i][0] * U[j][0] + W[i][1] * U[j][1] + sign * W[i][2] * U[j][2];
}

}  // namespace houv
