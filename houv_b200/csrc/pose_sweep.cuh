// The row-step of the symmetric sweep and the arg-min recovery helper shared by the K3 kernels
// pose_sym.cu (one row per CTA) and pose_duo.cu (two rows of one pair per CTA).
#pragma once
#include "pose_common.cuh"

namespace houv {

constexpr int kSB = 64;  // columns per superblock = rows per sub-group = arg-min group width

// The four metrics of one (row, column pair): f = 3-D, a/b/c = view with x/y/z masked.
// Same contraction as the reference on (masked) clouds: yy=dy*dy; xy=fma(dx,dx,yy);
// f=fma(dz,dz,xy); a=fma(dz,dz,yy); b=fma(dz,dz,dx*dx); c=xy.
__device__ __forceinline__ void eval4(const float2 X, const float2 Y, const float2 Z, float nqx,
                                      float nqy, float nqz, float2 &f, float2 &a, float2 &b,
                                      float2 &c) {
    const float2 dx = __fadd2_rn(X, make_float2(nqx, nqx));
    const float2 dy = __fadd2_rn(Y, make_float2(nqy, nqy));
    const float2 dz = __fadd2_rn(Z, make_float2(nqz, nqz));
    const float2 yy = __fmul2_rn(dy, dy), xx = __fmul2_rn(dx, dx);
    c = __ffma2_rn(dx, dx, yy);
    f = __ffma2_rn(dz, dz, c);
    a = __ffma2_rn(dz, dz, yy);
    b = __ffma2_rn(dz, dz, xx);
}

// One rotation step: 8 rows x one column pair, kRG rows evaluated before their folds.
// kNM = 4: the four metrics of Predict_loss; kNM = 1: the 3-D metric alone (a loss without the
// view terms, train_utils.py:439) -- 6 packed lane-ops and 2 min folds per evaluation.
template <int kRG, int kNM>
__device__ __forceinline__ void sym_step(const float2 X, const float2 Y, const float2 Z,
                                         const float (&nqx)[8], const float (&nqy)[8],
                                         const float (&nqz)[8], float (&best)[kNM][8],
                                         float (&cmx)[kNM], float (&cmy)[kNM]) {
#pragma unroll
    for (int r = 0; r < 8; r += kRG) {
        float2 v[kRG][4];
#pragma unroll
        for (int i = 0; i < kRG; ++i) {
            if (kNM == 4) {
                eval4(X, Y, Z, nqx[r + i], nqy[r + i], nqz[r + i], v[i][0], v[i][1], v[i][2], v[i][3]);
            } else {  // the same contraction as eval4's f: fma(dz,dz,fma(dx,dx,dy*dy))
                const float2 dx = __fadd2_rn(X, make_float2(nqx[r + i], nqx[r + i]));
                const float2 dy = __fadd2_rn(Y, make_float2(nqy[r + i], nqy[r + i]));
                const float2 dz = __fadd2_rn(Z, make_float2(nqz[r + i], nqz[r + i]));
                v[i][0] = __ffma2_rn(dz, dz, __ffma2_rn(dx, dx, __fmul2_rn(dy, dy)));
            }
        }
#pragma unroll
        for (int m = 0; m < kNM; ++m) {
#pragma unroll
            for (int i = 0; i < kRG; ++i) best[m][r + i] = min3(best[m][r + i], v[i][m].x, v[i][m].y);
#pragma unroll
            for (int i = 0; i < kRG; i += 2) {
                cmx[m] = min3(cmx[m], v[i][m].x, v[i + 1][m].x);
                cmy[m] = min3(cmy[m], v[i][m].y, v[i + 1][m].y);
            }
        }
    }
}

// Arg-min recovery: step k of a lane's walk over the 32 candidate pairs of a 64-wide group visits
// pair (k + lane) & 31; bit k of mx / my says the .x / .y candidate of that step equals the minimum.
// -> offset (0..63) of the smallest matching candidate inside the group, 0 when nothing matched.
__device__ __forceinline__ int first_match(unsigned mx, unsigned my, int lane) {
    mx = __funnelshift_l(mx, mx, lane);  // rotate: bit k -> bit (k + lane) & 31 = pair index
    my = __funnelshift_l(my, my, lane);
    const int fx = mx ? 2 * (__ffs(mx) - 1) : 128, fy = my ? 2 * (__ffs(my) - 1) + 1 : 128;
    const int f = min(fx, fy);
    return f < 128 ? f : 0;
}

}  // namespace houv
