// Instruction-order micro-benchmark for the K3 symmetric sweep step on sm_100a.
// One "row-step" = one thread row against one packed pair of columns, four metrics:
//   3 FADD2 (column pair - row scalar), 2 FMUL2 (yy, xx), 4 FFMA2 (xy, full, v0, v1), 8 FMNMX3
//   (4 row-min folds over the two halves, 4 column-min folds over two rows).
// The real kernel leaves the order to ptxas and measures 18.1 "pipe cycles" per 17; this tool issues the
// same dataflow through asm volatile in several fixed orders to see what the register file / operand
// reuse cache allow.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o k3step k3step.cu
#include <cstdio>
#include <cuda_runtime.h>

typedef unsigned long long u64;
#define ADD2S(d, a, s) asm volatile("{.reg .b64 t; mov.b64 t, {%2, %2}; add.rn.f32x2 %0, %1, t;}" : "=l"(d) : "l"(a), "f"(s))
#define MUL2(d, a, b) asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b))
#define FMA2(d, a, b, c) asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c))
#define MIN3(d, a, b, c) asm volatile("min.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c))
#define LO(p) __uint_as_float((unsigned)(p))
#define HI(p) __uint_as_float((unsigned)((p) >> 32))

constexpr int R = 8;  // rows per thread

struct Vals { u64 full, xy, v0, v1; };

// math of one row: 9 FMA-pipe instructions
template <int ORDER>
__device__ __forceinline__ void row_math(u64 cx, u64 cy, u64 cz, float nrx, float nry, float nrz, Vals &v) {
    u64 dx, dy, dz, yy, xx;
    ADD2S(dx, cx, nrx); ADD2S(dy, cy, nry); ADD2S(dz, cz, nrz);
    MUL2(yy, dy, dy); MUL2(xx, dx, dx);
    FMA2(v.xy, dx, dx, yy);
    FMA2(v.full, dz, dz, v.xy); FMA2(v.v0, dz, dz, yy); FMA2(v.v1, dz, dz, xx);
}

// folds of one row pair (a, b): 8 row folds + 8 column folds = 16 FMNMX3 for two rows
__device__ __forceinline__ void fold_pair(const Vals &a, const Vals &b, float (&rm)[R][4], int r, float (&cl)[4],
                                          float (&ch)[4]) {
    MIN3(rm[r][0], rm[r][0], LO(a.full), HI(a.full)); MIN3(rm[r][1], rm[r][1], LO(a.xy), HI(a.xy));
    MIN3(rm[r][2], rm[r][2], LO(a.v0), HI(a.v0));     MIN3(rm[r][3], rm[r][3], LO(a.v1), HI(a.v1));
    MIN3(rm[r + 1][0], rm[r + 1][0], LO(b.full), HI(b.full)); MIN3(rm[r + 1][1], rm[r + 1][1], LO(b.xy), HI(b.xy));
    MIN3(rm[r + 1][2], rm[r + 1][2], LO(b.v0), HI(b.v0));     MIN3(rm[r + 1][3], rm[r + 1][3], LO(b.v1), HI(b.v1));
    MIN3(cl[0], cl[0], LO(a.full), LO(b.full)); MIN3(ch[0], ch[0], HI(a.full), HI(b.full));
    MIN3(cl[1], cl[1], LO(a.xy), LO(b.xy));     MIN3(ch[1], ch[1], HI(a.xy), HI(b.xy));
    MIN3(cl[2], cl[2], LO(a.v0), LO(b.v0));     MIN3(ch[2], ch[2], HI(a.v0), HI(b.v0));
    MIN3(cl[3], cl[3], LO(a.v1), LO(b.v1));     MIN3(ch[3], ch[3], HI(a.v1), HI(b.v1));
}

// ORDER 0: math(r), math(r+1), folds(r, r+1)  -- blocks of 18 FMA-pipe then 16 ALU
// ORDER 1: software pipelined: math of pair k+1 interleaved instruction by instruction with folds of pair k
// ORDER 2: plain C++ (no volatile): ptxas chooses
template <int ORDER>
__global__ void __launch_bounds__(512, 1) k(float *out, long long *cyc, const u64 *cols, int iters) {
    __shared__ u64 sc[3][64];
    if (threadIdx.x < 192) sc[threadIdx.x / 64][threadIdx.x % 64] = cols[threadIdx.x];
    float rx[R], ry[R], rz[R], rm[R][4], cl[4], ch[4];
    for (int r = 0; r < R; ++r) {
        rx[r] = -0.01f * (threadIdx.x + r); ry[r] = -0.02f * (threadIdx.x + 2 * r); rz[r] = -0.03f * r;
        for (int m = 0; m < 4; ++m) rm[r][m] = 1e30f;
    }
    for (int m = 0; m < 4; ++m) cl[m] = ch[m] = 1e30f;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        const int c = (it + threadIdx.x) & 63;
        const u64 cx = sc[0][c], cy = sc[1][c], cz = sc[2][c];
        if (ORDER == 0) {
#pragma unroll
            for (int r = 0; r < R; r += 2) {
                Vals a, b;
                row_math<0>(cx, cy, cz, rx[r], ry[r], rz[r], a);
                row_math<0>(cx, cy, cz, rx[r + 1], ry[r + 1], rz[r + 1], b);
                fold_pair(a, b, rm, r, cl, ch);
            }
        }
        if (ORDER == 1) {
            Vals a, b, a2, b2;
            row_math<0>(cx, cy, cz, rx[0], ry[0], rz[0], a);
            row_math<0>(cx, cy, cz, rx[1], ry[1], rz[1], b);
#pragma unroll
            for (int r = 0; r < R; r += 2) {
                // interleave: 18 FMA-pipe instructions of the next pair with the 16 folds of this pair
                const int q = (r + 2) % R;  // last round recomputes pair 0 (keeps the count uniform)
                u64 dx, dy, dz, yy, xx, ex, ey, ez, y2, x2;
                ADD2S(dx, cx, rx[q]);          MIN3(rm[r][0], rm[r][0], LO(a.full), HI(a.full));
                ADD2S(dy, cy, ry[q]);          MIN3(rm[r][1], rm[r][1], LO(a.xy), HI(a.xy));
                ADD2S(dz, cz, rz[q]);          MIN3(rm[r][2], rm[r][2], LO(a.v0), HI(a.v0));
                MUL2(yy, dy, dy);              MIN3(rm[r][3], rm[r][3], LO(a.v1), HI(a.v1));
                MUL2(xx, dx, dx);              MIN3(rm[r + 1][0], rm[r + 1][0], LO(b.full), HI(b.full));
                FMA2(a2.xy, dx, dx, yy);       MIN3(rm[r + 1][1], rm[r + 1][1], LO(b.xy), HI(b.xy));
                FMA2(a2.full, dz, dz, a2.xy);  MIN3(rm[r + 1][2], rm[r + 1][2], LO(b.v0), HI(b.v0));
                FMA2(a2.v0, dz, dz, yy);       MIN3(rm[r + 1][3], rm[r + 1][3], LO(b.v1), HI(b.v1));
                FMA2(a2.v1, dz, dz, xx);       MIN3(cl[0], cl[0], LO(a.full), LO(b.full));
                ADD2S(ex, cx, rx[q + 1]);      MIN3(ch[0], ch[0], HI(a.full), HI(b.full));
                ADD2S(ey, cy, ry[q + 1]);      MIN3(cl[1], cl[1], LO(a.xy), LO(b.xy));
                ADD2S(ez, cz, rz[q + 1]);      MIN3(ch[1], ch[1], HI(a.xy), HI(b.xy));
                MUL2(y2, ey, ey);              MIN3(cl[2], cl[2], LO(a.v0), LO(b.v0));
                MUL2(x2, ex, ex);              MIN3(ch[2], ch[2], HI(a.v0), HI(b.v0));
                FMA2(b2.xy, ex, ex, y2);       MIN3(cl[3], cl[3], LO(a.v1), LO(b.v1));
                FMA2(b2.full, ez, ez, b2.xy);  MIN3(ch[3], ch[3], HI(a.v1), HI(b.v1));
                FMA2(b2.v0, ez, ez, y2);
                FMA2(b2.v1, ez, ez, x2);
                a = a2; b = b2;
            }
        }
        if (ORDER == 2) {
#pragma unroll
            for (int r = 0; r < R; r += 2) {
                float2 v[2][4];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const float2 X = *reinterpret_cast<const float2 *>(&cx), Y = *reinterpret_cast<const float2 *>(&cy),
                                 Z = *reinterpret_cast<const float2 *>(&cz);
                    const float2 dx = __fadd2_rn(X, make_float2(rx[r + h], rx[r + h]));
                    const float2 dy = __fadd2_rn(Y, make_float2(ry[r + h], ry[r + h]));
                    const float2 dz = __fadd2_rn(Z, make_float2(rz[r + h], rz[r + h]));
                    const float2 yy = __fmul2_rn(dy, dy), xx = __fmul2_rn(dx, dx);
                    v[h][1] = __ffma2_rn(dx, dx, yy);
                    v[h][0] = __ffma2_rn(dz, dz, v[h][1]);
                    v[h][2] = __ffma2_rn(dz, dz, yy);
                    v[h][3] = __ffma2_rn(dz, dz, xx);
                }
#pragma unroll
                for (int m = 0; m < 4; ++m) {
                    asm("min.f32 %0, %0, %1, %2;" : "+f"(rm[r][m]) : "f"(v[0][m].x), "f"(v[0][m].y));
                    asm("min.f32 %0, %0, %1, %2;" : "+f"(rm[r + 1][m]) : "f"(v[1][m].x), "f"(v[1][m].y));
                    asm("min.f32 %0, %0, %1, %2;" : "+f"(cl[m]) : "f"(v[0][m].x), "f"(v[1][m].x));
                    asm("min.f32 %0, %0, %1, %2;" : "+f"(ch[m]) : "f"(v[0][m].y), "f"(v[1][m].y));
                }
            }
        }
    }
    long long t1 = clock64();
    float s = 0;
    for (int r = 0; r < R; ++r)
        for (int m = 0; m < 4; ++m) s += rm[r][m];
    for (int m = 0; m < 4; ++m) s += cl[m] + ch[m];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int ORDER>
void run(const char *name, double rows_per_iter) {
    float *out; long long *cyc; u64 *cols;
    cudaMalloc(&out, 148 * 512 * 4); cudaMalloc(&cyc, 148 * 8); cudaMalloc(&cols, 192 * 8);
    float h[384];
    for (int i = 0; i < 384; ++i) h[i] = 0.001f * (i * 37 % 1000);
    cudaMemcpy(cols, h, sizeof(h), cudaMemcpyHostToDevice);
    const int iters = 2048;
    k<ORDER><<<148, 512>>>(out, cyc, cols, 16);
    k<ORDER><<<148, 512>>>(out, cyc, cols, iters);
    cudaDeviceSynchronize();
    long long hc[148]; cudaMemcpy(hc, cyc, sizeof(hc), cudaMemcpyDeviceToHost);
    // 4 warps per scheduler; a row-step is 17 half-rate instructions = "17 pipe cycles x 2"
    const double per_rowstep = (double)hc[0] / (iters * rows_per_iter * 4.0);
    printf("%-44s cycles per warp row-step = %6.2f  (= %5.2f pipe-cycle units; 17 = additive, 9 = FMA pipe only)\n",
           name, per_rowstep, per_rowstep / 2.0);
    cudaFree(out); cudaFree(cyc); cudaFree(cols);
}

int main() {
    run<0>("volatile: math pair, then 16 folds", 8);
    run<1>("volatile: math(k+1) interleaved with folds(k)", 8);
    run<2>("plain intrinsics, ptxas order", 8);
    return 0;
}
