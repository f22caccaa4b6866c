// K2 -- fused Chamfer gradient scatter for sm_100a.
//
// Replaces the two NmDistanceGradKernel launches of chamfer_cuda_backward
// (/root/reference/utils/metrics/CD/chamfer3D/chamfer3D.cu:155-195).  The
// reference runs 16 CTAs x 256 threads that walk the batch serially; here ONE
// launch covers both directions with one thread per point, sized to the data.
// Memory-bound: 32 bytes of algorithmic traffic per point
// (12 xyz + 4 idx + 4 graddist + 12 grad), plus the gathered partner point.
//
// Arithmetic is the reference's: g = graddist*2; v = g*(a-b) rounded (FMUL),
// then float atomics (RED.ADD.F32.FTZ.RN) into both gradient buffers, which the
// caller zero-fills.  Like the reference the summation ORDER of colliding
// scatters is not deterministic.
#include <stdlib.h>

#include "common.cuh"

namespace houv {

struct GradDir {
    const float *a;       // [B,n,3] points that own graddist / idx
    const float *b;       // [B,m,3] points indexed by idx
    const float *gd;      // [B,n]
    const int *idx;       // [B,n]
    float *ga;            // [B,n,3]
    float *gb;            // [B,m,3]
    int n, m;
};
struct GradArgs {
    GradDir dir[2];
    long long split;  // elements of direction 0 (= B*N)
    long long total;  // B*(N+M)
    bool vec;         // both gradient buffers are 4-byte aligned at least (always) and the
                      // vector form is allowed (HOUV_K2_SCALAR unset)
};

// Three float atomics on an xyz triple as one 8-byte vector RED (red.global.add.v2.f32, sm_90+)
// plus one scalar RED: the triple starts 8-byte aligned for even points (x,y | z) and 4 mod 8 for
// odd ones (x | y,z).  Element-wise the additions are the same float atomics as before.
__device__ __forceinline__ void red3(float *o, float a, float b, float c, bool vec) {
    if (!vec) {
        atomicAdd(o, a); atomicAdd(o + 1, b); atomicAdd(o + 2, c);
    } else if ((reinterpret_cast<uintptr_t>(o) & 7) == 0) {
        atomicAdd(reinterpret_cast<float2 *>(o), make_float2(a, b));
        atomicAdd(o + 2, c);
    } else {
        atomicAdd(o, a);
        atomicAdd(reinterpret_cast<float2 *>(o + 1), make_float2(b, c));
    }
}

__global__ void __launch_bounds__(256) grad_kernel(const GradArgs args) {
    for (long long e = blockIdx.x * 256LL + threadIdx.x; e < args.total;
         e += (long long)gridDim.x * 256) {
        const bool second = e >= args.split;
        GradDir D;  // field-wise select keeps the argument struct out of local memory
        D.a = second ? args.dir[1].a : args.dir[0].a;
        D.b = second ? args.dir[1].b : args.dir[0].b;
        D.gd = second ? args.dir[1].gd : args.dir[0].gd;
        D.idx = second ? args.dir[1].idx : args.dir[0].idx;
        D.ga = second ? args.dir[1].ga : args.dir[0].ga;
        D.gb = second ? args.dir[1].gb : args.dir[0].gb;
        D.n = second ? args.dir[1].n : args.dir[0].n;
        D.m = second ? args.dir[1].m : args.dir[0].m;
        const long long w = second ? e - args.split : e;  // flat (batch, point)
        const long long bi = w / D.n;
        const int j2 = D.idx[w];
        const float *pa = D.a + w * 3;
        const float *pb = D.b + (bi * D.m + j2) * 3;
        const float g = __fmul_rn(D.gd[w], 2.f);
        float *oa = D.ga + w * 3;
        float *ob = D.gb + (bi * D.m + j2) * 3;
        float v[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) v[k] = __fmul_rn(g, __fsub_rn(__ldg(pa + k), __ldg(pb + k)));
        red3(oa, v[0], v[1], v[2], args.vec);
        red3(ob, -v[0], -v[1], -v[2], args.vec);
    }
}

}  // namespace houv

using namespace houv;

extern "C" int houv_chamfer3d_backward(const float *xyz1, const float *xyz2,
                                       const float *graddist1, const float *graddist2,
                                       const int *idx1, const int *idx2, int B, int N, int M,
                                       float *gradxyz1, float *gradxyz2, void *stream) {
    if (B < 0 || N < 0 || M < 0) return HOUV_E_BADARG;
    if (B == 0 || N == 0 || M == 0) return 0;  // nothing to gather from: grads stay zero
    if (!xyz1 || !xyz2 || !graddist1 || !graddist2 || !idx1 || !idx2 || !gradxyz1 || !gradxyz2)
        return HOUV_E_BADARG;
    GradArgs a;
    a.dir[0] = GradDir{xyz1, xyz2, graddist1, idx1, gradxyz1, gradxyz2, N, M};
    a.dir[1] = GradDir{xyz2, xyz1, graddist2, idx2, gradxyz2, gradxyz1, M, N};
    a.split = (long long)B * N;
    a.total = (long long)B * ((long long)N + M);
    static const bool scalar_only = getenv("HOUV_K2_SCALAR") != nullptr;
    a.vec = !scalar_only;
    long long blocks = (a.total + 255) / 256;
    const long long cap = 148LL * 8 * 4;  // grid-stride beyond a few waves
    if (blocks > cap) blocks = cap;
    grad_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(a);
    return check_launch();
}
