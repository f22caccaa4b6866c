// Issue/pipe micro-benchmark for the K3 inner loop on sm_100a: how fast can one SM sub-partition
// issue FMNMX3 (ALU pipe), packed FFMA2 (FMA pipe) and mixes of the two, at 4 warps per scheduler?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu ; run: ./pipes
#include <cstdio>
#include <cuda_runtime.h>

#define MIN3(d, a, b, c) asm volatile("min.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c))
#define MIN2(d, a, b) asm volatile("min.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b))
// integer 3-input minimum on the float bit patterns (DPX; __vimin3_s32 -> VIMNMX3): non-negative floats order
// like signed integers, so for squared distances it folds the same values as FMNMX3
#define IMIN3(d, a, b, c)                                                                               \
    do {                                                                                                \
        int r_ = __vimin3_s32(__float_as_int(a), __float_as_int(b), __float_as_int(c));                 \
        asm volatile("" : "+r"(r_));                                                                    \
        d = __int_as_float(r_);                                                                         \
    } while (0)
#define UMIN3(d, a, b, c)                                                                               \
    do {                                                                                                \
        unsigned r_ = __vimin3_u32(__float_as_uint(a), __float_as_uint(b), __float_as_uint(c));         \
        asm volatile("" : "+r"(r_));                                                                    \
        d = __uint_as_float(r_);                                                                        \
    } while (0)
#define FMA2(d, a, b, c)                                                              \
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;"                                       \
                 : "=l"(d)                                                            \
                 : "l"(a), "l"(b), "l"(c))
#define FMA1(d, a, b, c) asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c))

constexpr int NCH = 8;
template <int MODE>
__global__ void __launch_bounds__(512, 1) k(float *out, long long *cyc, int iters) {
    float m[NCH], x[NCH], y[NCH];
    unsigned long long p[NCH], q[NCH];
    for (int i = 0; i < NCH; ++i) {
        m[i] = threadIdx.x + i; x[i] = threadIdx.x * 3.f + i; y[i] = i * 7.f + 1;
        p[i] = threadIdx.x + i; q[i] = 0x3f8000003f800000ull + i;
    }
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int rep = 0; rep < 4; ++rep) {
#pragma unroll
            for (int i = 0; i < NCH; ++i) {
                if (MODE == 0) { MIN3(m[i], m[i], x[i], y[i]); }
                if (MODE == 1) { FMA2(p[i], p[i], q[i], q[i]); }
                if (MODE == 2) { FMA2(p[i], p[i], q[i], q[i]); MIN3(m[i], m[i], x[i], y[i]); }
                if (MODE == 3) { MIN2(m[i], m[i], x[i]); }
                if (MODE == 4) { FMA1(x[i], x[i], y[i], y[i]); MIN3(m[i], m[i], x[(i + 1) % NCH], y[i]); }
                if (MODE == 5) { FMA1(x[i], x[i], y[i], y[i]); }
                if (MODE == 6) {  // 9 packed : 8 min3 (the K3 symmetric step)
                    FMA2(p[i], p[i], q[i], q[i]); MIN3(m[i], m[i], x[i], y[i]);
                    if (i == 0) FMA2(p[1], p[1], q[0], q[0]);
                }
                if (MODE == 8) {  // few distinct operands: is the register file the shared limiter?
                    FMA2(p[i], p[i], p[i], p[i]); MIN3(m[i], m[i], m[i], m[i]);
                }
                if (MODE == 9) {  // FFMA2 (3 distinct 64-bit operands) + FMNMX3 (3 distinct)
                    FMA2(p[i], p[i], q[i], q[(i + 1) % NCH]); MIN3(m[i], m[i], x[i], y[i]);
                }
                if (MODE == 10) {  // FFMA2 + 2 x FMNMX (2-input): same values folded as one FMNMX3
                    FMA2(p[i], p[i], q[i], q[i]); MIN2(m[i], m[i], x[i]); MIN2(m[i], m[i], y[i]);
                }
                if (MODE == 11) {  // FADD2 with scalar-broadcast operand + FMNMX3
                    asm volatile("{.reg .b64 t; mov.b64 t, {%1, %1}; add.rn.f32x2 %0, %0, t;}" : "+l"(p[i]) : "f"(y[i]));
                    MIN3(m[i], m[i], x[i], y[i]);
                }
                if (MODE == 12) {  // FFMA2 + IADD3-like integer op
                    FMA2(p[i], p[i], q[i], q[i]);
                    asm volatile("add.s32 %0, %0, %1;" : "+r"(*(int*)&x[i]) : "r"(*(int*)&y[i]));
                }
                if (MODE == 13) { IMIN3(m[i], m[i], x[i], y[i]); }
                if (MODE == 14) { FMA2(p[i], p[i], q[i], q[i]); IMIN3(m[i], m[i], x[i], y[i]); }
                if (MODE == 15) {  // 9 packed : 8 integer min3
                    FMA2(p[i], p[i], q[i], q[i]); IMIN3(m[i], m[i], x[i], y[i]);
                    if (i == 0) FMA2(p[1], p[1], q[0], q[0]);
                }
                if (MODE == 16) { UMIN3(m[i], m[i], x[i], y[i]); }
                if (MODE == 17) {  // 2-input integer min
                    int r_ = min(__float_as_int(m[i]), __float_as_int(x[i]));
                    asm volatile("" : "+r"(r_));
                    m[i] = __int_as_float(r_);
                }
                if (MODE == 7) {  // min3 on operands produced by the FMA pipe (cross-pipe forwarding)
                    FMA2(p[i], p[i], q[i], q[i]);
                    float lo = __uint_as_float((unsigned)p[i]), hi = __uint_as_float((unsigned)(p[i] >> 32));
                    MIN3(m[i], m[i], lo, hi);
                }
            }
        }
    }
    long long t1 = clock64();
    float s = 0;
    for (int i = 0; i < NCH; ++i) s += m[i] + x[i] + (float)p[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char *name, int per_iter_instr, int threads) {
    float *out; long long *cyc;
    cudaMalloc(&out, 148 * 512 * 4); cudaMalloc(&cyc, 148 * 8);
    int iters = 4096;
    k<MODE><<<148, threads>>>(out, cyc, 16);
    k<MODE><<<148, threads>>>(out, cyc, iters);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double c = h[0];
    double warps_per_smsp = threads / 32 / 4.0;
    double instr = (double)iters * per_iter_instr * warps_per_smsp;  // per scheduler
    printf("%-34s threads=%3d cycles=%9.0f  IPC/scheduler=%.3f\n", name, threads, c, instr / c);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    for (int th : {512}) {
        run<0>("FMNMX3", 4 * NCH, th);
        run<3>("FMNMX(2-input)", 4 * NCH, th);
        run<1>("FFMA2", 4 * NCH, th);
        run<5>("FFMA", 4 * NCH, th);
        run<2>("FFMA2+FMNMX3 1:1", 8 * NCH, th);
        run<4>("FFMA+FMNMX3 1:1", 8 * NCH, th);
        run<6>("FFMA2:FMNMX3 9:8", 4 * (2 * NCH + 1), th);
        run<7>("FFMA2->FMNMX3 dependent 1:1", 8 * NCH, th);
        run<8>("FFMA2(p,p,p)+FMNMX3(m,m,m)", 8 * NCH, th);
        run<9>("FFMA2 3 distinct + FMNMX3", 8 * NCH, th);
        run<10>("FFMA2 + 2xFMNMX", 12 * NCH, th);
        run<11>("FADD2 bcast + FMNMX3", 8 * NCH, th);
        run<12>("FFMA2 + IADD", 8 * NCH, th);
        run<13>("VIMNMX3 (s32)", 4 * NCH, th);
        run<16>("VIMNMX3 (u32)", 4 * NCH, th);
        run<17>("IMNMX 2-input (s32)", 4 * NCH, th);
        run<14>("FFMA2+VIMNMX3 1:1", 8 * NCH, th);
        run<15>("FFMA2:VIMNMX3 9:8", 4 * (2 * NCH + 1), th);
    }
    return 0;
}
