// K1s -- symmetric brute-force Chamfer forward for sm_100a: every distance d(i,j) between
// xyz1_i and xyz2_j is evaluated ONCE and folded into both results of
// chamfer_cuda_forward (/root/reference/utils/metrics/CD/chamfer3D/chamfer3D.cu:136-154):
// the row minimum (dist1/idx1 of NmDistanceKernel(xyz1 -> xyz2), :142) and the column minimum
// (dist2/idx2 of NmDistanceKernel(xyz2 -> xyz1), :143).  The two launches of the reference
// compute the same matrix: (b-a) and (a-b) square to the same float, and the contraction
// fma(dz,dz,fma(dx,dx,dy*dy)) (chamfer3D.cu:32-35 as nvcc compiles it) is sign-symmetric,
// so sharing the evaluation changes no bit of dist or idx.
//
// Schedule (same rotation as the K3 symmetric sweep, csrc/pose_sym.cu):
//  * one CTA of 256 threads owns a chunk of 2048 rows of one batch element, 8 rows per thread
//    in registers, and sweeps ALL columns; grid = (row chunks, batch);
//  * columns are staged per 2048-column tile: the AoS tile comes in with ONE TMA bulk copy
//    (cp.async.bulk + mbarrier) into a landing buffer and is transposed to SoA x[] y[] z[];
//  * in every step a lane evaluates its 8 rows against one PAIR of columns with packed fp32
//    (FADD2/FMUL2/FFMA2: 6 lane-ops per evaluation); the pair's running column minima travel
//    with it from lane to lane inside 8-lane groups (2 SHFL per 16 evaluations);
//  * minima only are tracked (one FMNMX3 per row per step, one per row pair and column);
//    arg-mins are recovered afterwards from 64 candidates: rows remember the superblock (64
//    columns) that last strictly improved them, columns get the first 64-row sub-group holding
//    their minimum from a per-superblock reduction; the candidate equal to the known minimum
//    with the smallest index is the reference's first minimum (chamfer3D.cu:36,46,126);
//  * with more than one row chunk the column results of the chunks are combined with a 64-bit
//    atomicMin on (distance bits << 32 | index) keys in a caller-provided workspace;
//  * small batches (fewer CTAs than SMs) are also split along the columns (gridDim.z): each CTA sweeps
//    its rows against a slice of the columns and the row results are combined through the same kind
//    of keys (smallest distance, then smallest index = the reference's first minimum).
// Cost per evaluation: 6 FMA-pipe + 2 ALU-pipe lane-cycles for BOTH directions, against
// 2 x (6 + 1) in the two-scan kernel (chamfer_fwd.cu).
#include <stdlib.h>

#include "common.cuh"

namespace houv {

constexpr int kSRows = kSymChunkRows;  // rows per CTA (2048)
constexpr int kSThreads = 256;
constexpr int kSTile = 2048;     // columns resident per tile
constexpr int kSGrp = 64;        // superblock / sub-group width
constexpr int kPartStride = 66;  // floats per sub-group row of the partial-minimum buffer

struct SmemK1s {
    float cx[kSTile], cy[kSTile], cz[kSTile];  // column tile, SoA (y = x + kSTile, z = x + 2 kSTile)
    float rx[kSRows], ry[kSRows], rz[kSRows];  // this CTA's rows, SoA
    float stage[kSTile * 3];                   // TMA landing buffer (AoS tile)
    float part[2][32][kPartStride];            // [buffer][sub-group][column of the superblock]
    float cmin[kSTile];                        // column minima of the tile
    unsigned char csg[kSTile];                 // first sub-group holding the column minimum
    unsigned long long mbar;
};

__device__ __forceinline__ unsigned smem_u32(const void *p) {
    return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes,
                                         unsigned long long *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

// squared distances of one row against a column pair; nq* = negated row coordinates
__device__ __forceinline__ float2 sqdist2b(float2 X, float2 Y, float2 Z, float nqx, float nqy,
                                           float nqz) {
    const float2 dx = __fadd2_rn(X, make_float2(nqx, nqx));
    const float2 dy = __fadd2_rn(Y, make_float2(nqy, nqy));
    const float2 dz = __fadd2_rn(Z, make_float2(nqz, nqz));
    return __ffma2_rn(dz, dz, __ffma2_rn(dx, dx, __fmul2_rn(dy, dy)));
}

// smallest index j0+c (c < 64) whose distance to (qx,qy,qz) equals `target`, candidates read
// from SoA arrays base[], base[stride + ...], base[2 stride + ...] in a lane-skewed order
__device__ __forceinline__ int first_equal(const float *base, int stride, int j0, float qx,
                                           float qy, float qz, float target, int lane) {
    int bi = 0x7fffffff;
#pragma unroll 4
    for (int k = 0; k < kSGrp / 2; ++k) {
        const int c = j0 + 2 * ((k + lane) & (kSGrp / 2 - 1));
        const float2 X = *reinterpret_cast<const float2 *>(base + c);
        const float2 Y = *reinterpret_cast<const float2 *>(base + stride + c);
        const float2 Z = *reinterpret_cast<const float2 *>(base + 2 * stride + c);
        const float2 d = sqdist2b(X, Y, Z, -qx, -qy, -qz);
        if (d.y == target) bi = min(bi, c + 1);
        if (d.x == target) bi = min(bi, c);
    }
    return bi == 0x7fffffff ? j0 : bi;
}

template <int kU>
__global__ void __launch_bounds__(kSThreads, 2) chamfer_sym_kernel(const SymArgs A) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    SmemK1s &S = *reinterpret_cast<SmemK1s *>(smem_raw);
    const int t = threadIdx.x;
    const int w = t >> 5, l = t & 31;
    const int gq = l >> 3, ll = l & 7;  // 8-lane rotation group, lane within it
    const int sg = w * 4 + gq;          // 64-row sub-group: rows 64*sg .. 64*sg+63 of the chunk
    const int row0 = 8 * t;
    const int rbase = blockIdx.x * kSRows;
    const int nr = A.nr, nc = A.nc;
    const int c_lo = blockIdx.z * A.cols_per_split;          // this CTA's slice of the columns
    const int c_hi = min(nc, c_lo + A.cols_per_split);
    const int nrows = min(kSRows, nr - rbase);  // real rows of this chunk (> 0 by grid size)
    if (c_lo >= nc) {  // a slice past the last column (slices are rounded up to 64): its plane of row keys is +inf
        for (int b = blockIdx.y; b < A.B; b += gridDim.y)
            for (int i = t; i < nrows; i += kSThreads)
                A.row_keys[((size_t)blockIdx.z * A.B + b) * nr + rbase + i] = ~0ull;
        return;
    }
    const bool scan_active = 256 * w < nrows;   // warp-uniform
    const int nsg = (nrows + kSGrp - 1) / kSGrp;
    const float inf = __int_as_float(0x7f800000);

    if (t == 0) {
        mbar_init(&S.mbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    unsigned phase = 0;
    __syncthreads();

    for (int b = blockIdx.y; b < A.B; b += gridDim.y) {
        __syncthreads();  // the previous batch element's column recovery still reads rx/ry/rz
        const float *rp = A.rows + ((size_t)b * nr + rbase) * 3;
        const float *cp = A.cols + (size_t)b * nc * 3;

        // ---- this thread's 8 rows: registers (negated) + shared SoA copy for the column side
        float nqx[8], nqy[8], nqz[8], best[8], prev[8];
        int gid[8], idxr[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            const int i = row0 + r;
            const bool ok = i < nrows;
            const float x = ok ? __ldg(rp + 3 * i + 0) : inf;  // padded row: every distance inf/NaN
            const float y = ok ? __ldg(rp + 3 * i + 1) : inf;
            const float z = ok ? __ldg(rp + 3 * i + 2) : 0.f;
            S.rx[i] = x; S.ry[i] = y; S.rz[i] = z;
            nqx[r] = -x; nqy[r] = -y; nqz[r] = -z;
            best[r] = inf; prev[r] = inf; gid[r] = 0; idxr[r] = 0;
        }

        for (int t0 = c_lo; t0 < c_hi; t0 += kSTile) {
            const int cnt = min(kSTile, c_hi - t0);
            const int cnt_pad = (cnt + kSGrp - 1) / kSGrp * kSGrp;
            const int nsb = cnt_pad / kSGrp;
            const int sb0 = t0 / kSGrp;  // global number of this tile's first superblock
            const float *src = cp + (size_t)t0 * 3;
            __syncthreads();  // previous tile / batch element fully consumed (cx.., stage, cmin, rx..)

            // ---- stage the tile: one TMA bulk copy of the 16-byte-aligned part, AoS -> SoA
            const bool al16 = (reinterpret_cast<uintptr_t>(src) & 15) == 0;
            const int nb = al16 ? (cnt & ~3) : 0;  // points brought in by the bulk copy
            if (nb > 0) {
                if (t == 0) {
                    // order the generic-proxy reads of `stage` (previous tile) before the async write
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    mbar_expect_tx(&S.mbar, (unsigned)nb * 12u);
                    bulk_g2s(S.stage, src, (unsigned)nb * 12u, &S.mbar);
                }
                mbar_wait(&S.mbar, phase);
                phase ^= 1u;
                for (int p = t; p < nb / 4; p += kSThreads) {
                    const float4 *s4 = reinterpret_cast<const float4 *>(S.stage) + 3 * p;
                    const float4 a = s4[0], bb = s4[1], c = s4[2];
                    reinterpret_cast<float4 *>(S.cx)[p] = make_float4(a.x, a.w, bb.z, c.y);
                    reinterpret_cast<float4 *>(S.cy)[p] = make_float4(a.y, bb.x, bb.w, c.z);
                    reinterpret_cast<float4 *>(S.cz)[p] = make_float4(a.z, bb.y, c.x, c.w);
                }
            }
            for (int p = nb + t; p < cnt_pad; p += kSThreads) {  // unaligned tail and +inf padding
                const bool ok = p < cnt;
                S.cx[p] = ok ? __ldg(src + 3 * p + 0) : inf;
                S.cy[p] = ok ? __ldg(src + 3 * p + 1) : 0.f;
                S.cz[p] = ok ? __ldg(src + 3 * p + 2) : 0.f;
            }
            __syncthreads();

            // ---- symmetric sweep of the tile, one superblock (64 columns) per round
            unsigned dirty = 0;  // rows improved in this tile
            int buf = 0;
#pragma unroll 1
            for (int sb = 0; sb < nsb; ++sb) {
                if (scan_active) {
#pragma unroll 1
                    for (int ph = 0; ph < 4; ++ph) {
                        const int cb = (gq + ph) & 3;
                        const int cbase = sb * kSGrp + 16 * cb;
                        float cmx = inf, cmy = inf;  // running minima of the column pair in hand
#pragma unroll kU
                        for (int s = 0; s < 8; ++s) {
                            const int c = cbase + 2 * ((ll + s) & 7);
                            const float2 X = *reinterpret_cast<const float2 *>(&S.cx[c]);
                            const float2 Y = *reinterpret_cast<const float2 *>(&S.cy[c]);
                            const float2 Z = *reinterpret_cast<const float2 *>(&S.cz[c]);
#pragma unroll
                            for (int r = 0; r < 8; r += 2) {
                                const float2 d0 = sqdist2b(X, Y, Z, nqx[r], nqy[r], nqz[r]);
                                const float2 d1 = sqdist2b(X, Y, Z, nqx[r + 1], nqy[r + 1], nqz[r + 1]);
                                best[r] = min3(best[r], d0.x, d0.y);
                                best[r + 1] = min3(best[r + 1], d1.x, d1.y);
                                cmx = min3(cmx, d0.x, d1.x);
                                cmy = min3(cmy, d0.y, d1.y);
                            }
                            if (s < 7) {  // the pair moves on to lane ll-1, ours arrives from ll+1
                                cmx = __shfl_sync(0xffffffffu, cmx, ll + 1, 8);
                                cmy = __shfl_sync(0xffffffffu, cmy, ll + 1, 8);
                            }
                        }
                        const int pc = 16 * cb + 2 * ((ll + 7) & 7);  // the pair finished here
                        *reinterpret_cast<float2 *>(&S.part[buf][sg][pc]) = make_float2(cmx, cmy);
                    }
                    // rows: strict improvement during this superblock => it holds the first minimum
#pragma unroll
                    for (int r = 0; r < 8; ++r) {
                        if (best[r] < prev[r]) {
                            prev[r] = best[r];
                            gid[r] = sb0 + sb;
                            dirty |= 1u << r;
                        }
                    }
                }
                __syncthreads();
                // columns: first sub-group (ascending rows, strict '<') holding the minimum.
                // 4 lanes per column, 8 sub-groups each; part[buf] is rewritten two rounds later,
                // after the next barrier.
                {
                    const int col = t >> 2, q = t & 3;
                    float bv = inf;
                    int bs = 0;
                    const int g1 = min(nsg, 8 * q + 8);
                    for (int g = 8 * q; g < g1; ++g) {
                        const float v = S.part[buf][g][col];
                        if (v < bv) { bv = v; bs = g; }
                    }
#pragma unroll
                    for (int o = 1; o < 4; o <<= 1) {
                        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
                        const int os = __shfl_xor_sync(0xffffffffu, bs, o);
                        if (ov < bv || (ov == bv && os < bs)) { bv = ov; bs = os; }
                    }
                    if (q == 0) {
                        S.cmin[sb * kSGrp + col] = bv;
                        S.csg[sb * kSGrp + col] = (unsigned char)bs;
                    }
                }
                buf ^= 1;
            }
            __syncthreads();

            // ---- columns of this tile: exact first arg-min among the 64 rows of the winning
            //      sub-group, then write (single row chunk) or combine across chunks
            for (int col = t; col < cnt; col += kSThreads) {
                const float target = S.cmin[col];
                const int j0 = (int)S.csg[col] * kSGrp;
                const int bi = first_equal(S.rx, kSRows, j0, S.cx[col], S.cy[col], S.cz[col], target, l);
                const size_t o = (size_t)b * nc + t0 + col;
                if (A.keys) {
                    const unsigned long long key =
                        ((unsigned long long)__float_as_uint(target) << 32) | (unsigned)(rbase + bi);
                    atomicMin(A.keys + o, key);
                } else {
                    A.dist_c[o] = target;
                    A.idx_c[o] = rbase + bi;
                }
            }
            // ---- rows improved in this tile: exact first arg-min inside the winning superblock
            //      (with column slices the recovery is deferred: only the winning slice's superblock
            //      has to be searched, once, by chamfer_sym_unpack)
            if (scan_active && !A.row_keys) {
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    if (dirty & (1u << r)) {
                        const int j0 = (gid[r] - sb0) * kSGrp;
                        idxr[r] = t0 + first_equal(S.cx, kSTile, j0, -nqx[r], -nqy[r], -nqz[r], best[r], l);
                    }
                }
            }
        }

        // ---- row results
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            const int i = row0 + r;
            if (i < nrows) {
                const size_t o = (size_t)b * nr + rbase + i;
                if (A.row_keys) {
                    // column slices: every slice has its own plane of (distance, superblock) keys --
                    // plain stores, nothing to initialise; the unpack takes the smallest key over the
                    // planes (smallest distance, then earliest superblock: it holds the first minimum)
                    A.row_keys[(size_t)blockIdx.z * A.B * nr + o] =
                        ((unsigned long long)__float_as_uint(best[r]) << 32) | (unsigned)gid[r];
                } else {
                    A.dist_r[o] = best[r];
                    A.idx_r[o] = idxr[r];
                }
            }
        }
    }
}

// Column keys hold (distance, row index); row keys hold (distance, superblock of 64 columns): the row's
// first minimum is the first column of that superblock whose distance equals the minimum.  Blocks
// [0, col_blocks) unpack 256 column keys each, the others recover 32 rows each, eight lanes per row.
__global__ void __launch_bounds__(256) chamfer_sym_unpack(const unsigned long long *keys, float *dist, int *idx,
                                                         size_t n, unsigned col_blocks,
                                                         const unsigned long long *row_keys, int planes,
                                                         float *dist_r, int *idx_r, size_t n_r,
                                                         const float *rows, const float *cols, int nr, int nc) {
    if (blockIdx.x < col_blocks) {
        const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
        if (i < n) {
            const unsigned long long k = keys[i];
            dist[i] = __uint_as_float((unsigned)(k >> 32));
            idx[i] = (int)(unsigned)(k & 0xffffffffull);
        }
        return;
    }
    // eight lanes per row: lane `sub` reads the planes sub, sub+8, .. of the row's key and then the
    // candidates 8 sub .. 8 sub + 7 of the winning superblock (96 contiguous bytes, LDG.128 when the
    // cloud is 16-byte aligned) -- two dependent round trips per row, every load of a stage in flight
    // at once; the lowest matching candidate over the eight lanes is the first minimum
    const size_t i = (size_t)(blockIdx.x - col_blocks) * 32 + (threadIdx.x >> 3);
    const int sub = threadIdx.x & 7;
    const bool active = i < n_r;
    unsigned long long k = ~0ull;
    if (active)
        for (int z = sub; z < planes; z += 8) {
            const unsigned long long kz = row_keys[(size_t)z * n_r + i];
            k = kz < k ? kz : k;
        }
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
        const unsigned long long ko = __shfl_xor_sync(0xffffffffu, k, o);
        k = ko < k ? ko : k;
    }
    const float target = __uint_as_float((unsigned)(k >> 32));
    const int j0 = (int)(unsigned)(k & 0xffffffffull) * kSGrp;
    int first = kSGrp;
    if (active) {
        const size_t b = i / nr;
        const float qx = __ldg(rows + 3 * i), qy = __ldg(rows + 3 * i + 1), qz = __ldg(rows + 3 * i + 2);
        const float *c = cols + (b * nc + j0 + 8 * sub) * 3;
        const int m = min(8, nc - j0 - 8 * sub);
        if (m == 8 && (reinterpret_cast<uintptr_t>(c) & 15) == 0) {
            float4 v[6];
#pragma unroll
            for (int q = 0; q < 6; ++q) v[q] = __ldg(reinterpret_cast<const float4 *>(c) + q);
            const float *f = reinterpret_cast<const float *>(v);
#pragma unroll
            for (int pnt = 7; pnt >= 0; --pnt)
                if (sqdist_ref(f[3 * pnt], f[3 * pnt + 1], f[3 * pnt + 2], qx, qy, qz) == target)
                    first = 8 * sub + pnt;
        } else {
            for (int pnt = m - 1; pnt >= 0; --pnt)
                if (sqdist_ref(__ldg(c + 3 * pnt), __ldg(c + 3 * pnt + 1), __ldg(c + 3 * pnt + 2), qx, qy, qz) == target)
                    first = 8 * sub + pnt;
        }
    }
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) first = min(first, __shfl_xor_sync(0xffffffffu, first, o));
    if (active && sub == 0) {
        dist_r[i] = target;
        // a row that never improved (no finite distance) keeps the superblock's first index
        idx_r[i] = first >= kSGrp ? j0 : j0 + first;
    }
}

int sym_col_splits(int B, int nr, int nc) {
    static const int forced = getenv("HOUV_K1S_SPLIT") ? atoi(getenv("HOUV_K1S_SPLIT")) : 0;
    const long long ctas = (long long)((nr + kSRows - 1) / kSRows) * B;
    // two CTAs per SM (measured, C2 = 32 x 2048^2: 8 slices 66.6 us, 4 slices 68.6 us, directed kernel
    // 74.7 us; batches that already fill the GPU lose from slicing, C5: 0.71 ms with 4 slices, 0.665 without)
    int s = forced > 0 ? forced : (ctas >= 148 ? 1 : (int)(296 / ctas));
    const int most = forced > 0 ? nc / 64 : nc / 256;  // at least 256 columns per slice
    if (s > most) s = most;
    return s < 1 ? 1 : s;
}

// Launch the symmetric forward with `rows` = the cloud chunked over CTAs.  a.keys: workspace of
// B*nc 64-bit words, required when nr > kSRows; a.row_keys: sym_col_splits() planes of B*nr words,
// required when sym_col_splits() > 1.
int launch_chamfer_sym(const SymArgs &a, cudaStream_t st) {
    static unsigned long long smem_set = 0;  // per-device flag bits
    if (int e = ensure_dyn_smem(chamfer_sym_kernel<8>, (int)sizeof(SmemK1s), smem_set)) return e;
    const int chunks = (a.nr + kSRows - 1) / kSRows;
    const int splits = sym_col_splits(a.B, a.nr, a.nc);
    SymArgs k = a;
    if (chunks == 1) k.keys = nullptr;
    if (splits == 1) k.row_keys = nullptr;
    if ((chunks > 1 && !k.keys) || (splits > 1 && !k.row_keys)) return HOUV_E_BADARG;
    k.cols_per_split = splits == 1 ? a.nc : ((a.nc + splits - 1) / splits + kSGrp - 1) / kSGrp * kSGrp;
    const size_t n_c = chunks > 1 ? (size_t)a.B * a.nc : 0, n_r = splits > 1 ? (size_t)a.B * a.nr : 0;
    cudaError_t e = cudaSuccess;
    if (n_c) e = cudaMemsetAsync(k.keys, 0xff, sizeof(unsigned long long) * n_c, st);
    if (e != cudaSuccess) return (int)e;  // (the row-key planes are written whole: no initialisation)
    dim3 grid(chunks, a.B < 65535 ? a.B : 65535, splits);
    chamfer_sym_kernel<8><<<grid, kSThreads, sizeof(SmemK1s), st>>>(k);
    if (int err = check_launch()) return err;
    if (n_c + n_r) {
        const unsigned col_blocks = (unsigned)((n_c + 255) / 256), row_blocks = (unsigned)((n_r + 31) / 32);
        chamfer_sym_unpack<<<col_blocks + row_blocks, 256, 0, st>>>(k.keys, a.dist_c, a.idx_c, n_c, col_blocks,
                                                                   k.row_keys, splits, a.dist_r, a.idx_r, n_r,
                                                                   a.rows, a.cols, a.nr, a.nc);
        return check_launch();
    }
    return 0;
}

}  // namespace houv
