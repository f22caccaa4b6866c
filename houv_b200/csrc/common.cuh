// Shared device helpers for the houv_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/houv_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "houv_b200 is written for sm_100a (B200) only"
#endif

namespace houv {

// 3-input minimum (SASS FMNMX3, one ALU-pipe slot for two candidates).
// IEEE minNum semantics: a NaN operand is ignored, as `d < best` ignores it in
// the reference (chamfer3D.cu:46).
__device__ __forceinline__ float min3(float a, float b, float c) {
    float r;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}

// The reference's squared distance with nvcc's contraction spelled out
// (chamfer3D.cu:32-35 -> FADD,FADD,FADD,FMUL(y),FFMA(x),FFMA(z)); the _rn
// intrinsics are never re-contracted by the compiler.
__device__ __forceinline__ float sqdist_ref(float tx, float ty, float tz, float qx, float qy,
                                            float qz) {
    float dx = __fsub_rn(tx, qx), dy = __fsub_rn(ty, qy), dz = __fsub_rn(tz, qz);
    return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
}

// Two evaluations per issue slot (FADD2/FMUL2/FFMA2). nq* hold the NEGATED
// query coordinate in both halves: t + (-q) == t - q bit for bit.
__device__ __forceinline__ float2 sqdist2(float2 tx, float2 ty, float2 tz, float2 nqx, float2 nqy,
                                          float2 nqz) {
    float2 dx = __fadd2_rn(tx, nqx), dy = __fadd2_rn(ty, nqy), dz = __fadd2_rn(tz, nqz);
    return __ffma2_rn(dz, dz, __ffma2_rn(dx, dx, __fmul2_rn(dy, dy)));
}

// Arguments of the symmetric Chamfer forward (chamfer_sym.cu), launched from chamfer_fwd.cu.
struct SymArgs {
    const float *rows;   // [B,nr,3]
    const float *cols;   // [B,nc,3]
    float *dist_r;       // [B,nr]
    int *idx_r;          // [B,nr]  index into cols
    float *dist_c;       // [B,nc]
    int *idx_c;          // [B,nc]  index into rows
    unsigned long long *keys;  // [B,nc] workspace when gridDim.x > 1, else null
    unsigned long long *row_keys;  // [B,nr] workspace when the columns are split over gridDim.z > 1
    int B, nr, nc;
    int cols_per_split;  // multiple of 64; nc when there is one split
};

constexpr int kSymChunkRows = 2048;  // rows of `rows` handled by one CTA
int launch_chamfer_sym(const SymArgs &a, cudaStream_t st);
// Column splits the symmetric kernel uses for this shape (1 = none): small batches are cut along the
// columns as well so that the grid covers the GPU.  Shared by the launcher and the workspace size.
int sym_col_splits(int B, int nr, int nc);

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device attribute: set it once per device and
// call site (`done` = the call site's bit mask of devices already configured).
template <class F>
inline int ensure_dyn_smem(F *func, int bytes, unsigned long long &done) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 64 && ((done >> dev) & 1ull)) return 0;
    e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return (int)e;
    if (dev < 64) done |= 1ull << dev;
    return 0;
}

inline int check_launch() {
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : (int)e;
}

}  // namespace houv
