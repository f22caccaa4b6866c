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
    long long items;  // B * ceil((N+M)/256) chunks of 256 points
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

// Work is ordered batch by batch: one item = 256 consecutive points of one batch element, direction
// 0's N points first and direction 1's M points after them.  Both directions of a batch element hit
// the same two gradient buffers (once as the owner's contiguous rows, once as the scatter target),
// so keeping them adjacent in time makes the second touch an L2 hit instead of a second DRAM
// read-modify-write (ncu, 6400 x 2048^2: 1.91 GB of DRAM traffic per launch before, see
// profiles/r1_ncu_summary.md).
__global__ void __launch_bounds__(256) grad_kernel(const GradArgs args) {
    const int per = args.dir[0].n + args.dir[0].m;  // points per batch element, both directions
    const int chunks = (per + 255) >> 8;
    for (long long item = blockIdx.x; item < args.items; item += gridDim.x) {
        const long long bi = item / chunks;
        const int r = (int)(item - bi * chunks) * 256 + threadIdx.x;
        if (r >= per) continue;
        const bool second = r >= args.dir[0].n;
        GradDir D;  // field-wise select keeps the argument struct out of local memory
        D.a = second ? args.dir[1].a : args.dir[0].a;
        D.b = second ? args.dir[1].b : args.dir[0].b;
        D.gd = second ? args.dir[1].gd : args.dir[0].gd;
        D.idx = second ? args.dir[1].idx : args.dir[0].idx;
        D.ga = second ? args.dir[1].ga : args.dir[0].ga;
        D.gb = second ? args.dir[1].gb : args.dir[0].gb;
        D.n = second ? args.dir[1].n : args.dir[0].n;
        D.m = second ? args.dir[1].m : args.dir[0].m;
        const long long w = bi * D.n + (second ? r - args.dir[0].n : r);  // flat (batch, point)
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

// Batch-element-resident form: one CTA owns one batch element, stages both clouds in shared memory
// (coalesced, 16-byte loads where the element is aligned), builds both gradient buffers there, and
// adds the finished rows to the caller's buffers with one coalesced vector RED per 16 bytes.  Per
// point that is 20 B read + 12 B of RED instead of 4 scattered RED packets through the L1->XBAR
// port, which is what bounded grad_kernel (ncu: l1tex2xbar request cycles 78 % busy, DRAM 26 %).
// Shared memory has no native float add (atomicAdd compiles to an LDS / FADD / ATOMS.CAST.SPIN loop),
// which makes randomly addressed shared-memory wavefronts the next bound (ncu: l1tex data pipe 60-70 %).
// Measured 6400 x 2048^2 on one B200: 305 us against 527 us streaming (DRAM traffic 1.12 GB per launch,
// the 32 B/point floor plus the read half of the gradient read-modify-write).  A variant with the three
// compare-and-swap chains of a point issued together was slower (450 us) and is not kept.  Needs 24*(N+M) bytes of shared memory.
constexpr int kResThreads = 512;
constexpr int kResPer = 4;  // points per thread and direction held in registers (clouds <= 2048)

__device__ __forceinline__ void res_stage(float *dst, const float *src, int n3) {
    if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
        const int q = n3 >> 2;
        for (int i = threadIdx.x; i < q; i += kResThreads)
            reinterpret_cast<float4 *>(dst)[i] = __ldg(reinterpret_cast<const float4 *>(src) + i);
        for (int i = 4 * q + threadIdx.x; i < n3; i += kResThreads) dst[i] = __ldg(src + i);
    } else {
        for (int i = threadIdx.x; i < n3; i += kResThreads) dst[i] = __ldg(src + i);
    }
}

// One point: v = 2*gd*(a_i - b_j) goes to the owner's row (+v) and to the partner's row (-v).
__device__ __forceinline__ void res_point(const float *sa, const float *sb, float *ga, float *gb,
                                          int i, int j, float g) {
    const float v0 = __fmul_rn(g, __fsub_rn(sa[i * 3 + 0], sb[j * 3 + 0]));
    const float v1 = __fmul_rn(g, __fsub_rn(sa[i * 3 + 1], sb[j * 3 + 1]));
    const float v2 = __fmul_rn(g, __fsub_rn(sa[i * 3 + 2], sb[j * 3 + 2]));
    atomicAdd(ga + i * 3 + 0, v0); atomicAdd(gb + j * 3 + 0, -v0);
    atomicAdd(ga + i * 3 + 1, v1); atomicAdd(gb + j * 3 + 1, -v1);
    atomicAdd(ga + i * 3 + 2, v2); atomicAdd(gb + j * 3 + 2, -v2);
}

__device__ __forceinline__ void res_dir(const float *sa, const float *sb, float *ga, float *gb,
                                        const float *gd, const int *idx, int n) {
    for (int i = threadIdx.x; i < n; i += kResThreads)
        res_point(sa, sb, ga, gb, i, __ldg(idx + i), __fmul_rn(__ldg(gd + i), 2.f));
}

// assign = false: out += acc (the reference's accumulate-into contract); true: out = acc, which
// spares the caller's zero-fill and the read half of the read-modify-write.
__device__ __forceinline__ void res_flush(float *out, const float *acc, int n3, bool assign) {
    if ((reinterpret_cast<uintptr_t>(out) & 15) == 0) {
        const int q = n3 >> 2;
        for (int i = threadIdx.x; i < q; i += kResThreads) {
            const float4 v = *reinterpret_cast<const float4 *>(acc + 4 * i);
            if (assign) reinterpret_cast<float4 *>(out)[i] = v;
            else atomicAdd(reinterpret_cast<float4 *>(out) + i, v);
        }
        for (int i = 4 * q + threadIdx.x; i < n3; i += kResThreads) {
            if (assign) out[i] = acc[i];
            else atomicAdd(out + i, acc[i]);
        }
    } else {
        for (int i = threadIdx.x; i < n3; i += kResThreads) {
            if (assign) out[i] = acc[i];
            else atomicAdd(out + i, acc[i]);
        }
    }
}

__global__ void __launch_bounds__(kResThreads, 2) grad_resident_kernel(const GradArgs args, int B,
                                                                       bool assign) {
    extern __shared__ __align__(16) float sm[];
    const int n = args.dir[0].n, m = args.dir[0].m;
    const int n3 = n * 3, m3 = m * 3;
    const int n3p = (n3 + 3) & ~3, m3p = (m3 + 3) & ~3;  // keep every array 16-byte aligned
    float *s1 = sm, *s2 = s1 + n3p, *g1 = s2 + m3p, *g2 = g1 + n3p;
    const bool in_regs = n <= kResThreads * kResPer && m <= kResThreads * kResPer;
    for (int bi = blockIdx.x; bi < B; bi += gridDim.x) {
        const float *gd1 = args.dir[0].gd + (long long)bi * n, *gd2 = args.dir[1].gd + (long long)bi * m;
        const int *ix1 = args.dir[0].idx + (long long)bi * n, *ix2 = args.dir[1].idx + (long long)bi * m;
        // index / weight loads go out with the staging loads: one global round trip per element
        int j1[kResPer], j2[kResPer];
        float w1[kResPer], w2[kResPer];
        if (in_regs) {
#pragma unroll
            for (int q = 0; q < kResPer; ++q) {
                const int i = threadIdx.x + q * kResThreads;
                j1[q] = i < n ? __ldg(ix1 + i) : 0;
                w1[q] = i < n ? __ldg(gd1 + i) : 0.f;
                j2[q] = i < m ? __ldg(ix2 + i) : 0;
                w2[q] = i < m ? __ldg(gd2 + i) : 0.f;
            }
        }
        res_stage(s1, args.dir[0].a + (long long)bi * n3, n3);
        res_stage(s2, args.dir[0].b + (long long)bi * m3, m3);
        for (int i = threadIdx.x; i < (n3p + m3p) >> 2; i += kResThreads)
            reinterpret_cast<float4 *>(g1)[i] = make_float4(0.f, 0.f, 0.f, 0.f);  // g1 and g2
        __syncthreads();
        if (in_regs) {
#pragma unroll
            for (int q = 0; q < kResPer; ++q) {
                const int i = threadIdx.x + q * kResThreads;
                if (i < n) res_point(s1, s2, g1, g2, i, j1[q], __fmul_rn(w1[q], 2.f));
                if (i < m) res_point(s2, s1, g2, g1, i, j2[q], __fmul_rn(w2[q], 2.f));
            }
        } else {
            res_dir(s1, s2, g1, g2, gd1, ix1, n);
            res_dir(s2, s1, g2, g1, gd2, ix2, m);
        }
        __syncthreads();
        res_flush(args.dir[0].ga + (long long)bi * n3, g1, n3, assign);
        res_flush(args.dir[0].gb + (long long)bi * m3, g2, m3, assign);
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------
// Deterministic form: gather-side accumulation through the inverse index.
//
// The reference adds colliding scatters with float atomics in whatever order the GPU schedules them
// (chamfer3D.cu:166-171; registration/README.md:82-90 concedes the run-to-run drift).  Here one CTA
// owns one batch element, inverts idx1 / idx2 in shared memory (counting sort: integer atomics for
// the bucket sizes, block scan, fill, then every bucket sorted by point index) and every gradient
// row is summed by ONE thread in the order a single-thread run of the reference kernels would use:
//   gradxyz1[i] = (0 + own term of direction 1) - scattered terms of direction 2, ascending j
//   gradxyz2[k] = (0 - scattered terms of direction 1, ascending i) + own term of direction 2
// -- bit-identical to oracle/chamfer_oracle.c::oracle_chamfer_backward_f32 (= the reference kernel
// bodies run by one host thread), run to run and GPU to GPU.  No floating-point atomics anywhere:
// shared memory sees native integer atomics only, the rows leave with plain coalesced stores.
// Clouds up to 2048 points (indices held as 16-bit), 104 KB of shared memory, two CTAs per SM.
constexpr int kGThreads = 512;
constexpr int kGMax = 2048;      // points per cloud
constexpr int kGSortMax = 32;    // buckets up to this size are sorted in place; larger ones rescan idx

struct GatherSmem {
    float s1[kGMax * 3], s2[kGMax * 3];   // the two clouds (AoS)
    float w1[kGMax], w2[kGMax];           // 2 * graddist
    unsigned short ix1[kGMax], ix2[kGMax];
    unsigned short start1[kGMax + 2], start2[kGMax + 2];  // bucket offsets (exclusive scan), +1 = end
    unsigned short list1[kGMax], list2[kGMax];            // points grouped by the bucket they hit
    int cnt1[kGMax], cnt2[kGMax];         // bucket sizes / fill cursors
    int wsum[2][kGThreads / 32];
};

// exclusive scan of cnt[0..len) into start[0..len], cnt reset to 0; every thread owns 4 entries
__device__ __forceinline__ void gather_scan(int *cnt, unsigned short *start, int len, int *wsum) {
    const int t = threadIdx.x, base = 4 * t;
    int c[4], s = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) { c[q] = base + q < len ? cnt[base + q] : 0; s += c[q]; }
    int incl = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if ((t & 31) >= o) incl += v;
    }
    if ((t & 31) == 31) wsum[t >> 5] = incl;
    __syncthreads();
    int off = 0;
    for (int ww = 0; ww < (t >> 5); ++ww) off += wsum[ww];
    int run = off + incl - s;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        if (base + q < len) { start[base + q] = (unsigned short)run; cnt[base + q] = 0; }
        run += c[q];
    }
    if (t == kGThreads - 1) start[len] = (unsigned short)run;  // = number of points scattered
}

// Sum over the points of bucket [p0, p0+L) of `list` in ascending point order of
// -w[i] * (sa[i] - b_k): the scattered terms a single-thread reference run adds to row k.
__device__ __forceinline__ void gather_bucket(unsigned short *list, int p0, int L, const unsigned short *ix,
                                              int n_src, int k, const float *sa, const float *w,
                                              float bx, float by, float bz, float (&acc)[3]) {
    auto add = [&](int i) {
        const float g = w[i];
        acc[0] = __fadd_rn(acc[0], -__fmul_rn(g, __fsub_rn(sa[3 * i + 0], bx)));
        acc[1] = __fadd_rn(acc[1], -__fmul_rn(g, __fsub_rn(sa[3 * i + 1], by)));
        acc[2] = __fadd_rn(acc[2], -__fmul_rn(g, __fsub_rn(sa[3 * i + 2], bz)));
    };
    if (L <= kGSortMax) {
        for (int a = 1; a < L; ++a) {  // insertion sort: buckets hold one or two points as a rule
            const unsigned short v = list[p0 + a];
            int b = a - 1;
            while (b >= 0 && list[p0 + b] > v) { list[p0 + b + 1] = list[p0 + b]; --b; }
            list[p0 + b + 1] = v;
        }
        for (int a = 0; a < L; ++a) add(list[p0 + a]);
    } else {
        for (int i = 0; i < n_src; ++i)
            if (ix[i] == k) add(i);
    }
}

__global__ void __launch_bounds__(kGThreads, 2) grad_gather_kernel(const GradArgs args, int B, bool assign) {
    extern __shared__ __align__(16) unsigned char gsm_raw[];
    GatherSmem &S = *reinterpret_cast<GatherSmem *>(gsm_raw);
    const int n = args.dir[0].n, m = args.dir[0].m, t = threadIdx.x;
    for (int bi = blockIdx.x; bi < B; bi += gridDim.x) {
        const float *gd1 = args.dir[0].gd + (long long)bi * n, *gd2 = args.dir[1].gd + (long long)bi * m;
        const int *ix1 = args.dir[0].idx + (long long)bi * n, *ix2 = args.dir[1].idx + (long long)bi * m;
        res_stage(S.s1, args.dir[0].a + (long long)bi * n * 3, n * 3);
        res_stage(S.s2, args.dir[0].b + (long long)bi * m * 3, m * 3);
        for (int i = t; i < n; i += kGThreads) {
            S.w1[i] = __fmul_rn(__ldg(gd1 + i), 2.f);
            S.ix1[i] = (unsigned short)__ldg(ix1 + i);
            S.cnt2[i] = 0;
        }
        for (int j = t; j < m; j += kGThreads) {
            S.w2[j] = __fmul_rn(__ldg(gd2 + j), 2.f);
            S.ix2[j] = (unsigned short)__ldg(ix2 + j);
            S.cnt1[j] = 0;
        }
        __syncthreads();
        // bucket sizes: direction 1 scatters into the m points of cloud 2, direction 2 into cloud 1
        for (int i = t; i < n; i += kGThreads) atomicAdd(&S.cnt1[S.ix1[i]], 1);
        for (int j = t; j < m; j += kGThreads) atomicAdd(&S.cnt2[S.ix2[j]], 1);
        __syncthreads();
        gather_scan(S.cnt1, S.start1, m, S.wsum[0]);
        gather_scan(S.cnt2, S.start2, n, S.wsum[1]);
        __syncthreads();
        for (int i = t; i < n; i += kGThreads) {
            const int k = S.ix1[i];
            S.list1[S.start1[k] + atomicAdd(&S.cnt1[k], 1)] = (unsigned short)i;
        }
        for (int j = t; j < m; j += kGThreads) {
            const int k = S.ix2[j];
            S.list2[S.start2[k] + atomicAdd(&S.cnt2[k], 1)] = (unsigned short)j;
        }
        __syncthreads();
        // rows of cloud 1: own term of direction 1 first, then what direction 2 scatters, ascending
        float *o1 = args.dir[0].ga + (long long)bi * n * 3, *o2 = args.dir[0].gb + (long long)bi * m * 3;
        for (int i = t; i < n; i += kGThreads) {
            const float ax = S.s1[3 * i], ay = S.s1[3 * i + 1], az = S.s1[3 * i + 2];
            const int j = S.ix1[i];
            const float g = S.w1[i];
            float acc[3];
            acc[0] = __fadd_rn(0.f, __fmul_rn(g, __fsub_rn(ax, S.s2[3 * j + 0])));
            acc[1] = __fadd_rn(0.f, __fmul_rn(g, __fsub_rn(ay, S.s2[3 * j + 1])));
            acc[2] = __fadd_rn(0.f, __fmul_rn(g, __fsub_rn(az, S.s2[3 * j + 2])));
            gather_bucket(S.list2, S.start2[i], S.cnt2[i], S.ix2, m, i, S.s2, S.w2, ax, ay, az, acc);
#pragma unroll
            for (int c = 0; c < 3; ++c) o1[3 * i + c] = assign ? acc[c] : __fadd_rn(o1[3 * i + c], acc[c]);
        }
        // rows of cloud 2: what direction 1 scatters (ascending), then the own term of direction 2
        for (int k = t; k < m; k += kGThreads) {
            const float bx = S.s2[3 * k], by = S.s2[3 * k + 1], bz = S.s2[3 * k + 2];
            float acc[3] = {0.f, 0.f, 0.f};
            gather_bucket(S.list1, S.start1[k], S.cnt1[k], S.ix1, n, k, S.s1, S.w1, bx, by, bz, acc);
            const int i = S.ix2[k];
            const float g = S.w2[k];
            acc[0] = __fadd_rn(acc[0], __fmul_rn(g, __fsub_rn(bx, S.s1[3 * i + 0])));
            acc[1] = __fadd_rn(acc[1], __fmul_rn(g, __fsub_rn(by, S.s1[3 * i + 1])));
            acc[2] = __fadd_rn(acc[2], __fmul_rn(g, __fsub_rn(bz, S.s1[3 * i + 2])));
#pragma unroll
            for (int c = 0; c < 3; ++c) o2[3 * k + c] = assign ? acc[c] : __fadd_rn(o2[3 * k + c], acc[c]);
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------
// The deterministic gather kernel, PIPELINED (round 2): one persistent CTA of 1024 threads per SM.
// A batch element's six arrays (xyz1, xyz2, graddist1, graddist2, idx1, idx2: 80 KB at 2048 points)
// are contiguous in global memory: ONE thread brings the next element in with six TMA bulk copies
// (cp.async.bulk + mbarrier, two buffers) while all threads invert / gather the current one, and
// the finished gradient rows leave from shared memory by bulk store (assign) or bulk reduce-add
// (accumulate: one float add per address, still deterministic).  Global traffic is in flight all
// the time instead of between the phases of one element.  Same summation order, same bits as
// grad_gather_kernel.  Needs N, M multiples of 4 (16-byte rows) and 16-byte aligned arrays; other
// shapes take grad_gather_kernel.
constexpr int kG2Threads = 1024;

struct Gather2Stage {
    float s1[kGMax * 3], s2[kGMax * 3];  // clouds in; gradient rows out (same place)
    float gd1[kGMax], gd2[kGMax];
    int ix1[kGMax], ix2[kGMax];
};
struct Gather2Smem {
    Gather2Stage st[2];
    unsigned short start1[kGMax + 2], start2[kGMax + 2];
    unsigned short list1[kGMax], list2[kGMax];
    int cnt1[kGMax], cnt2[kGMax];
    int wsum[2][kG2Threads / 32];
    unsigned long long full[2];
};

__device__ __forceinline__ unsigned g2_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void g2_load(void *dst, const void *src, unsigned bytes, unsigned long long *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
                     "r"(g2_u32(dst)), "l"(src), "r"(bytes), "r"(g2_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void g2_store(void *dst, const void *src, unsigned bytes, bool add) {
    if (add)
        asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(dst),
                     "r"(g2_u32(src)), "r"(bytes)
                     : "memory");
    else
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(g2_u32(src)),
                     "r"(bytes)
                     : "memory");
}
__device__ __forceinline__ void g2_wait(unsigned long long *bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(g2_u32(bar)),
        "r"(parity)
        : "memory");
}

// the scattered terms of bucket [p0, p0+L), ascending point index (as gather_bucket, int indices).
// Buckets of up to 4 entries (nearly all: a point is the nearest neighbour of one other point on
// average): entries loaded together, sorted in registers by a 5-exchange network, their data loaded
// independently, then added in order -- no LDS -> LDS -> FADD chain per entry.
__device__ __forceinline__ void g2_bucket(unsigned short *list, int p0, int L, const int *ix, int n_src, int k,
                                          const float *sa, const float *gd, float bx, float by, float bz,
                                          float (&acc)[3]) {
    auto add = [&](int i) {
        const float g = __fmul_rn(gd[i], 2.f);
        acc[0] = __fadd_rn(acc[0], -__fmul_rn(g, __fsub_rn(sa[3 * i + 0], bx)));
        acc[1] = __fadd_rn(acc[1], -__fmul_rn(g, __fsub_rn(sa[3 * i + 1], by)));
        acc[2] = __fadd_rn(acc[2], -__fmul_rn(g, __fsub_rn(sa[3 * i + 2], bz)));
    };
    if (L == 0) return;
    if (L <= 4) {
        unsigned e[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) e[a] = a < L ? (unsigned)list[p0 + a] : 0xffffu;  // sentinels sort last
        auto cswap = [](unsigned &x, unsigned &y) { const unsigned lo = min(x, y), hi = max(x, y); x = lo; y = hi; };
        cswap(e[0], e[1]); cswap(e[2], e[3]); cswap(e[0], e[2]); cswap(e[1], e[3]); cswap(e[1], e[2]);
        float t0[4], t1[4], t2[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            const int i = a < L ? (int)e[a] : 0;
            const float g = __fmul_rn(gd[i], 2.f);
            t0[a] = -__fmul_rn(g, __fsub_rn(sa[3 * i + 0], bx));
            t1[a] = -__fmul_rn(g, __fsub_rn(sa[3 * i + 1], by));
            t2[a] = -__fmul_rn(g, __fsub_rn(sa[3 * i + 2], bz));
        }
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            if (a < L) {
                acc[0] = __fadd_rn(acc[0], t0[a]); acc[1] = __fadd_rn(acc[1], t1[a]); acc[2] = __fadd_rn(acc[2], t2[a]);
            }
        }
        return;
    }
    if (L <= kGSortMax) {
        for (int a = 1; a < L; ++a) {
            const unsigned short v = list[p0 + a];
            int b = a - 1;
            while (b >= 0 && list[p0 + b] > v) { list[p0 + b + 1] = list[p0 + b]; --b; }
            list[p0 + b + 1] = v;
        }
        for (int a = 0; a < L; ++a) add(list[p0 + a]);
    } else {
        for (int i = 0; i < n_src; ++i)
            if (ix[i] == k) add(i);
    }
}

// exclusive scans of cnt1[0..m) -> start1[0..m] and cnt2[0..n) -> start2[0..n]; both cnt arrays reset to 0
__device__ __forceinline__ void g2_scan2(Gather2Smem &S, int n, int m) {
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int a0 = 2 * t < m ? S.cnt1[2 * t] : 0, a1 = 2 * t + 1 < m ? S.cnt1[2 * t + 1] : 0;
    const int b0 = 2 * t < n ? S.cnt2[2 * t] : 0, b1 = 2 * t + 1 < n ? S.cnt2[2 * t + 1] : 0;
    const int sa = a0 + a1, sb = b0 + b1;
    int ia = sa, ib = sb;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int va = __shfl_up_sync(0xffffffffu, ia, o), vb = __shfl_up_sync(0xffffffffu, ib, o);
        if (lane >= o) { ia += va; ib += vb; }
    }
    if (lane == 31) { S.wsum[0][warp] = ia; S.wsum[1][warp] = ib; }
    __syncthreads();
    const int wa = S.wsum[0][lane], wb = S.wsum[1][lane];  // 32 warps: one per lane
    int ja = wa, jb = wb;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int va = __shfl_up_sync(0xffffffffu, ja, o), vb = __shfl_up_sync(0xffffffffu, jb, o);
        if (lane >= o) { ja += va; jb += vb; }
    }
    const int offa = __shfl_sync(0xffffffffu, ja - wa, warp), offb = __shfl_sync(0xffffffffu, jb - wb, warp);
    const int tota = __shfl_sync(0xffffffffu, ja, 31), totb = __shfl_sync(0xffffffffu, jb, 31);
    const int ra = offa + ia - sa, rb = offb + ib - sb;
    if (2 * t < m) { S.start1[2 * t] = (unsigned short)ra; S.cnt1[2 * t] = 0; }
    if (2 * t + 1 < m) { S.start1[2 * t + 1] = (unsigned short)(ra + a0); S.cnt1[2 * t + 1] = 0; }
    if (2 * t < n) { S.start2[2 * t] = (unsigned short)rb; S.cnt2[2 * t] = 0; }
    if (2 * t + 1 < n) { S.start2[2 * t + 1] = (unsigned short)(rb + b0); S.cnt2[2 * t + 1] = 0; }
    if (t == 0) { S.start1[m] = (unsigned short)tota; S.start2[n] = (unsigned short)totb; }
}

__global__ void __launch_bounds__(kG2Threads, 1) grad_gather2_kernel(const GradArgs args, int B, bool assign) {
    extern __shared__ __align__(128) unsigned char g2_raw[];
    Gather2Smem &S = *reinterpret_cast<Gather2Smem *>(g2_raw);
    const int n = args.dir[0].n, m = args.dir[0].m, t = threadIdx.x;
    const unsigned bytes = (unsigned)(n + m) * 20u;  // 12 xyz + 4 graddist + 4 idx per point
    if (t == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(g2_u32(&S.full[0])));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(g2_u32(&S.full[1])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = t; i < kGMax; i += kG2Threads) { S.cnt1[i] = 0; S.cnt2[i] = 0; }
    __syncthreads();
    auto issue = [&](int b, int p) {  // thread 0: the six arrays of element b into buffer p
        Gather2Stage &T = S.st[p];
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(g2_u32(&S.full[p])), "r"(bytes)
                     : "memory");
        g2_load(T.s1, args.dir[0].a + (size_t)b * n * 3, (unsigned)n * 12u, &S.full[p]);
        g2_load(T.s2, args.dir[0].b + (size_t)b * m * 3, (unsigned)m * 12u, &S.full[p]);
        g2_load(T.gd1, args.dir[0].gd + (size_t)b * n, (unsigned)n * 4u, &S.full[p]);
        g2_load(T.gd2, args.dir[1].gd + (size_t)b * m, (unsigned)m * 4u, &S.full[p]);
        g2_load(T.ix1, args.dir[0].idx + (size_t)b * n, (unsigned)n * 4u, &S.full[p]);
        g2_load(T.ix2, args.dir[1].idx + (size_t)b * m, (unsigned)m * 4u, &S.full[p]);
    };
    if (t == 0 && (int)blockIdx.x < B) issue(blockIdx.x, 0);
    int k = 0;
    for (int b = blockIdx.x; b < B; b += gridDim.x, ++k) {
        const int p = k & 1;
        Gather2Stage &T = S.st[p];
        if (t == 0 && b + (int)gridDim.x < B) {
            // buffer p^1 still feeds the bulk stores of the previous element: wait until they have READ it
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            issue(b + gridDim.x, p ^ 1);
        }
        g2_wait(&S.full[p], (unsigned)(k >> 1) & 1u);
        // bucket sizes: direction 1 scatters into the m points of cloud 2, direction 2 into cloud 1
        for (int i = t; i < n; i += kG2Threads) atomicAdd(&S.cnt1[T.ix1[i]], 1);
        for (int j = t; j < m; j += kG2Threads) atomicAdd(&S.cnt2[T.ix2[j]], 1);
        __syncthreads();
        g2_scan2(S, n, m);
        __syncthreads();
        for (int i = t; i < n; i += kG2Threads) {
            const int kk = T.ix1[i];
            S.list1[S.start1[kk] + atomicAdd(&S.cnt1[kk], 1)] = (unsigned short)i;
        }
        for (int j = t; j < m; j += kG2Threads) {
            const int kk = T.ix2[j];
            S.list2[S.start2[kk] + atomicAdd(&S.cnt2[kk], 1)] = (unsigned short)j;
        }
        __syncthreads();
        float r1[2][3], r2[2][3];
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int i = t + q * kG2Threads;
            if (i < n) {  // row of cloud 1: own term of direction 1, then what direction 2 scatters
                const float ax = T.s1[3 * i], ay = T.s1[3 * i + 1], az = T.s1[3 * i + 2];
                const int j = T.ix1[i];
                const float g = __fmul_rn(T.gd1[i], 2.f);
                r1[q][0] = __fadd_rn(0.f, __fmul_rn(g, __fsub_rn(ax, T.s2[3 * j + 0])));
                r1[q][1] = __fadd_rn(0.f, __fmul_rn(g, __fsub_rn(ay, T.s2[3 * j + 1])));
                r1[q][2] = __fadd_rn(0.f, __fmul_rn(g, __fsub_rn(az, T.s2[3 * j + 2])));
                g2_bucket(S.list2, S.start2[i], S.cnt2[i], T.ix2, m, i, T.s2, T.gd2, ax, ay, az, r1[q]);
                S.cnt2[i] = 0;  // ready for the next element's count
            }
            const int kk = t + q * kG2Threads;
            if (kk < m) {  // row of cloud 2: what direction 1 scatters, then the own term of direction 2
                const float bx = T.s2[3 * kk], by = T.s2[3 * kk + 1], bz = T.s2[3 * kk + 2];
                r2[q][0] = 0.f; r2[q][1] = 0.f; r2[q][2] = 0.f;
                g2_bucket(S.list1, S.start1[kk], S.cnt1[kk], T.ix1, n, kk, T.s1, T.gd1, bx, by, bz, r2[q]);
                S.cnt1[kk] = 0;
                const int i2 = T.ix2[kk];
                const float g = __fmul_rn(T.gd2[kk], 2.f);
                r2[q][0] = __fadd_rn(r2[q][0], __fmul_rn(g, __fsub_rn(bx, T.s1[3 * i2 + 0])));
                r2[q][1] = __fadd_rn(r2[q][1], __fmul_rn(g, __fsub_rn(by, T.s1[3 * i2 + 1])));
                r2[q][2] = __fadd_rn(r2[q][2], __fmul_rn(g, __fsub_rn(bz, T.s1[3 * i2 + 2])));
            }
        }
        __syncthreads();  // everybody is done reading the clouds: the gradient rows take their place
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int i = t + q * kG2Threads;
            if (i < n) { T.s1[3 * i] = r1[q][0]; T.s1[3 * i + 1] = r1[q][1]; T.s1[3 * i + 2] = r1[q][2]; }
            if (i < m) { T.s2[3 * i] = r2[q][0]; T.s2[3 * i + 1] = r2[q][1]; T.s2[3 * i + 2] = r2[q][2]; }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (t == 0) {
            g2_store(args.dir[0].ga + (size_t)b * n * 3, T.s1, (unsigned)n * 12u, !assign);
            g2_store(args.dir[0].gb + (size_t)b * m * 3, T.s2, (unsigned)m * 12u, !assign);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
    }
    if (t == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

}  // namespace houv

using namespace houv;

static int backward_impl(const float *xyz1, const float *xyz2, const float *graddist1,
                         const float *graddist2, const int *idx1, const int *idx2, int B, int N,
                         int M, float *gradxyz1, float *gradxyz2, bool assign, void *stream,
                         bool deterministic = false) {
    if (B < 0 || N < 0 || M < 0) return HOUV_E_BADARG;
    if (assign && B > 0) {  // an empty partner cloud leaves a zero gradient
        if ((N > 0 && !gradxyz1) || (M > 0 && !gradxyz2)) return HOUV_E_BADARG;
        if (N == 0 || M == 0) {
            cudaError_t e = cudaSuccess;
            if (N > 0) e = cudaMemsetAsync(gradxyz1, 0, sizeof(float) * 3 * (size_t)B * N, (cudaStream_t)stream);
            if (M > 0) e = cudaMemsetAsync(gradxyz2, 0, sizeof(float) * 3 * (size_t)B * M, (cudaStream_t)stream);
            return (int)e;
        }
    }
    if (B == 0 || N == 0 || M == 0) return 0;  // nothing to gather from: grads stay zero
    if (!xyz1 || !xyz2 || !graddist1 || !graddist2 || !idx1 || !idx2 || !gradxyz1 || !gradxyz2)
        return HOUV_E_BADARG;
    GradArgs a;
    a.dir[0] = GradDir{xyz1, xyz2, graddist1, idx1, gradxyz1, gradxyz2, N, M};
    a.dir[1] = GradDir{xyz2, xyz1, graddist2, idx2, gradxyz2, gradxyz1, M, N};
    if ((long long)N + M > 0x7fffff00LL) return HOUV_E_BADARG;
    a.items = (long long)B * (((long long)N + M + 255) / 256);
    static const bool scalar_only = getenv("HOUV_K2_SCALAR") != nullptr;
    a.vec = !scalar_only;
    // HOUV_K2_VARIANT: 0 = streaming grad_kernel, 1 = batch-element-resident kernel where it fits,
    // unset = resident when there are enough batch elements to fill the GPU.
    const int forced = getenv("HOUV_K2_VARIANT") ? atoi(getenv("HOUV_K2_VARIANT")) : -1;
    const size_t res_bytes = 2 * sizeof(float) * ((((size_t)N * 3 + 3) & ~(size_t)3) +
                                                  (((size_t)M * 3 + 3) & ~(size_t)3));
    const bool fits = res_bytes <= 200 * 1024;
    // The deterministic gather kernel (clouds up to 2048 points) on request -- the _deterministic
    // entry point or HOUV_K2_VARIANT=2 / 3; it costs ~1.4x the float-atomics kernels (measured, B200:
    // 6400 x 2048^2 448 us, pipelined 407 us, against 290 us resident; 32 x 2048^2 16.4 us against 14 us).
    // HOUV_K2_VARIANT: 1 = the float-atomics resident kernel, 0 = streaming.
    if (deterministic && (N > kGMax || M > kGMax)) return HOUV_E_TOOBIG;
    // 3 (and the deterministic entry point when the layout allows): the same sums, TMA-pipelined
    const bool al16 = ((reinterpret_cast<uintptr_t>(xyz1) | reinterpret_cast<uintptr_t>(xyz2) |
                        reinterpret_cast<uintptr_t>(graddist1) | reinterpret_cast<uintptr_t>(graddist2) |
                        reinterpret_cast<uintptr_t>(idx1) | reinterpret_cast<uintptr_t>(idx2) |
                        reinterpret_cast<uintptr_t>(gradxyz1) | reinterpret_cast<uintptr_t>(gradxyz2)) & 15) == 0;
    const bool can_pipe = N <= kGMax && M <= kGMax && N % 4 == 0 && M % 4 == 0 && al16;
    if (can_pipe && (forced == 3 || (forced < 0 && deterministic))) {
        int dev = 0, sms = 148;
        if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        const int grid = B < sms ? B : sms;
        static unsigned long long g2smem_set = 0;
        if (int e = ensure_dyn_smem(grad_gather2_kernel, (int)sizeof(Gather2Smem), g2smem_set)) return e;
        grad_gather2_kernel<<<grid, kG2Threads, sizeof(Gather2Smem), (cudaStream_t)stream>>>(a, B, assign);
        return check_launch();
    }
    if (N <= kGMax && M <= kGMax && (forced == 2 || (forced < 0 && deterministic))) {
        int grid = 2 * 148;
        if (grid > B) grid = B;
        static unsigned long long gsmem_set = 0;
        if (int e = ensure_dyn_smem(grad_gather_kernel, (int)sizeof(GatherSmem), gsmem_set)) return e;
        grad_gather_kernel<<<grid, kGThreads, sizeof(GatherSmem), (cudaStream_t)stream>>>(a, B, assign);
        return check_launch();
    }
    if (fits && (forced == 1 || (forced < 0 && B >= 148))) {
        const int per_sm = (int)((227 * 1024) / (res_bytes + 1024));
        int grid = 148 * (per_sm < 1 ? 1 : per_sm > 2 ? 2 : per_sm);  // registers allow 2 x 512 threads
        if (grid > B) grid = B;
        static unsigned long long smem_set = 0;  // per-device flag bits
        if (int e = ensure_dyn_smem(grad_resident_kernel, 200 * 1024, smem_set)) return e;
        grad_resident_kernel<<<grid, kResThreads, res_bytes, (cudaStream_t)stream>>>(a, B, assign);
        return check_launch();
    }
    if (assign) {  // the streaming kernel accumulates: zero-fill here, on the same stream
        cudaError_t e = cudaMemsetAsync(gradxyz1, 0, sizeof(float) * 3 * (size_t)B * N, (cudaStream_t)stream);
        if (e == cudaSuccess)
            e = cudaMemsetAsync(gradxyz2, 0, sizeof(float) * 3 * (size_t)B * M, (cudaStream_t)stream);
        if (e != cudaSuccess) return (int)e;
    }
    long long blocks = a.items;
    const long long cap = 148LL * 8 * 4;  // grid-stride beyond a few waves
    if (blocks > cap) blocks = cap;
    grad_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(a);
    return check_launch();
}

extern "C" int houv_chamfer3d_backward(const float *xyz1, const float *xyz2,
                                       const float *graddist1, const float *graddist2,
                                       const int *idx1, const int *idx2, int B, int N, int M,
                                       float *gradxyz1, float *gradxyz2, void *stream) {
    return backward_impl(xyz1, xyz2, graddist1, graddist2, idx1, idx2, B, N, M, gradxyz1, gradxyz2,
                         false, stream);
}

extern "C" int houv_chamfer3d_backward_deterministic(const float *xyz1, const float *xyz2,
                                                     const float *graddist1, const float *graddist2,
                                                     const int *idx1, const int *idx2, int B, int N,
                                                     int M, float *gradxyz1, float *gradxyz2,
                                                     int assign, void *stream) {
    return backward_impl(xyz1, xyz2, graddist1, graddist2, idx1, idx2, B, N, M, gradxyz1, gradxyz2,
                         assign != 0, stream, true);
}

extern "C" int houv_chamfer3d_backward_assign(const float *xyz1, const float *xyz2,
                                              const float *graddist1, const float *graddist2,
                                              const int *idx1, const int *idx2, int B, int N, int M,
                                              float *gradxyz1, float *gradxyz2, void *stream) {
    return backward_impl(xyz1, xyz2, graddist1, graddist2, idx1, idx2, B, N, M, gradxyz1, gradxyz2,
                         true, stream);
}
