// K3 (symmetric scan, TWO rows per CTA) -- fused HOUV pose optimisation for sm_100a.
//
// Same contract, same per-iteration semantics and the SAME BITS as pose_sym_kernel<.,.,4,2>
// (pose_sym.cu; predict_model's hot loop, /root/reference/registration/models/houv.py:120-126:
// HOUV.forward :94-103, Predict_loss :209-222 = 4 x chamfer_3DDist forward [chamfer3D.cu:12-154]
// + top-k trimmed sqrt-means [model_utils_completion.py:83-100], autograd backward
// [chamfer3D.cu:155-195], Adam :118).
//
// Why: pose_sym_kernel owns an SM with one row (start), and 18 % of every iteration is spent
// outside the sweep (arg-min recovery 8.7 %, per-superblock column reduction 4.4 %, radix select,
// sums, pose chain and Adam on one thread 4.9 %) with all 16 warps waiting on __syncthreads()
// while the FMA and ALU pipes idle.  Here a CTA of 512 threads is two TEAMS of 256 threads, each
// team optimising its own start of the SAME pair: the target cloud is shared (one 24 KB copy),
// everything else is per team, and the teams synchronise with named barriers (bar.sync id, 256)
// instead of __syncthreads(), half an iteration out of phase -- one team's recovery / select /
// pose chain runs under the other team's sweep.
//
// What makes two rows fit in 227 KB (207 KB used):
//  * one team sweeps ALL superblocks in ascending order with its 256 threads x 8 rows (instead of
//    two halves of the CTA taking 16 superblocks each), so there is ONE shadow of the row minima
//    (32 KB) and no second winning-superblock map;
//  * the partial column minima part[32 sub-groups][4 metrics][64 columns] (32 KB) are live only
//    inside the sweep, where the transformed source is dead (every thread holds its 8 rows in
//    registers): both share one 32 KB region.  After the sweep the rows are written back from the
//    registers (-(-y) == y) next to the 3-D row minima.
//
// Bit-identity with pose_sym_kernel (tests/test_pose_gpu.py::test_duo_equals_single_row_kernel):
// minima, arg-min groups and the radix select are order-free; every floating-point SUM is taken in
// pose_sym_kernel's order by letting thread u of a team play its virtual threads u and u + 256
// (two partial sums, two shuffle trees, sixteen virtual-warp slots summed in order).
#include <type_traits>

#include "pose_common.cuh"
#include "pose_sweep.cuh"

namespace houv {

constexpr int kDuoThreads = 512;
constexpr int kTeam = 256;  // threads per team = per row
// pipelined column reduction (kPipe): four buffers of partial column minima, one per 16-column
// phase: [metric][sub-group][16 columns], metric stride padded so that the reducer's lanes (two
// metrics x 16 columns) fall into 32 different banks
constexpr int kPartM = 32 * 16 + 16;   // floats per metric
constexpr int kPartBuf = 4 * kPartM;   // floats per buffer

struct TeamSmem {
    // inside the sweep: part[sub-group][metric][column], partial column minima of one superblock;
    // outside: yx, yy, yz (transformed source, consecutive kPts-float arrays: the recovery addresses
    // y and z as x + kPts, x + 2*kPts) and d3r (3-D row minima, dist2)
    float region[4 * kPartBuf > 4 * kPts ? 4 * kPartBuf : 4 * kPts];
    float d3c[kPts];                  // 3-D column minima (dist1)
    unsigned char gid[4][2][kPts];    // winning group: [m][0][col] = 64-row sub-group, [m][1][row] = superblock
    // row minima as of the last superblock boundary: per thread 8 float4 chunks, chunk c = values
    // k = 4c..4c+3 (k = metric*8 + r) at position c ^ (thread & 7)
    float4 shadow[kTeam][8];
    int hist[2][256];
    float red[16][16];                // [virtual warp of pose_sym_kernel][value]
    float pose[12];
    unsigned selu[8];
    double state[80];                 // PoseState of this row (touched by the team's thread 0 only)
    double bt[2];
    // kPipe: full[b] -- all 8 warps stored their partial minima of the phase in buffer b;
    // empty[b] -- its two reducer warps are done with it
    unsigned long long full[4], empty[4];
    float vbuf[2][4][kSB];            // kPipe: view column sums [upper half][metric][column] on their way to team_sum
};

struct DuoSmem {
    float tx[kPts], ty[kPts], tz[kPts];  // target cloud (columns) of the pair, shared by both teams
    TeamSmem team[2];
};

__device__ __forceinline__ unsigned duo_smem_u32(const void *p) {
    return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void duo_mbar_init(unsigned long long *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(duo_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void duo_mbar_arrive(unsigned long long *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(duo_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void duo_mbar_wait(unsigned long long *bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(duo_smem_u32(bar)),
        "r"(parity)
        : "memory");
}

__device__ __forceinline__ void team_sync(int team) {
    asm volatile("bar.sync %0, %1;" ::"r"(team + 1), "r"(kTeam) : "memory");
}

// Sum of NV values over the 512 virtual threads of pose_sym_kernel's block_sum_sym, bit for bit:
// a[] / b[] are this thread's partial sums as virtual thread u / u + 256 (virtual warps tw and
// tw + 8); values whose bit is set in `colmask` are the column-side view sums, whose reducer role
// (metric m, column rc) belongs to virtual thread 128 m + 64 h + rc (h = upper half of the
// superblocks): virtual warps 4 m + (rc >> 5) and that + 2.
template <int NV>
__device__ __forceinline__ void team_sum(float (&a)[NV], float (&b)[NV], unsigned colmask,
                                         TeamSmem &T, int team, int u) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            a[i] += __shfl_xor_sync(0xffffffffu, a[i], o);
            b[i] += __shfl_xor_sync(0xffffffffu, b[i], o);
        }
    }
    const int tw = u >> 5, l = u & 31;
    team_sync(team);
    if (l == 0) {
        const int cw = 4 * (tw >> 1) + (tw & 1);
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const bool col = (colmask >> i) & 1u;
            T.red[col ? cw : tw][i] = a[i];
            T.red[col ? cw + 2 : tw + 8][i] = b[i];
        }
    }
    team_sync(team);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        float s = 0.f;
#pragma unroll
        for (int ww = 0; ww < 16; ++ww) s += T.red[ww][i];
        a[i] = s;
    }
}

template <int kSymUnroll, int kRG, bool kPipe>
__global__ void __launch_bounds__(kDuoThreads, 1) pose_duo_kernel(const PoseArgs A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    DuoSmem &S = *reinterpret_cast<DuoSmem *>(smem_raw);
    const int team = threadIdx.x >> 8, u = threadIdx.x & (kTeam - 1);
    const int N = A.N, M = A.M;  // rows = source points, columns = target points

    // rows 2b and 2b+1 share a CTA when they belong to the same pair (always, for the row layouts
    // solver.py builds: K starts per pair, K even); rows that cannot be paired are left to the
    // one-row kernel, which launch_pose_duo runs afterwards on exactly those (PoseArgs::only_unpaired)
    const int rowa = 2 * blockIdx.x, rowb = rowa + 1;
    if (rowb >= A.rows) return;
    const int pair = A.row_pair ? A.row_pair[rowa] : rowa / A.K;
    if ((A.row_pair ? A.row_pair[rowb] : rowb / A.K) != pair) return;
    const int row = rowa + team;

    TeamSmem &T = S.team[team];
    float *const yx = T.region, *const yy = T.region + kPts, *const yz = T.region + 2 * kPts;
    float *const d3r = T.region + 3 * kPts;
    float (*const part)[4][kSB] = reinterpret_cast<float (*)[4][kSB]>(T.region);
    const float *const d3p[2] = {T.d3c, d3r};
    const float *const sh = reinterpret_cast<const float *>(&T.shadow[0][0]);

    // geometry of the sweep
    const int w = u >> 5, l = u & 31;   // warp within the team, lane
    const int gq = l >> 3, ll = l & 7;  // 8-lane rotation group, lane within it
    const int sg = w * 4 + gq;          // 64-row sub-group: rows 64*sg .. 64*sg+63
    const int nsb = (M + kSB - 1) / kSB;          // superblocks holding real columns
    const int nsbh = (nsb + 1) / 2;               // pose_sym_kernel's split of the superblocks
    const int nsg = (N + kSB - 1) / kSB;          // sub-groups holding real rows
    const bool scan_active = 256 * w < N;         // warp-uniform: this warp owns real rows
    const int row0 = 8 * u;
    int pad = 2 * nsbh * kSB;                     // columns touched by the sweep
    const int rpad = (N + 255) / 256 * 256;       // rows touched by the sweep
    pad = pad > rpad ? pad : rpad;                // (<= kPts)
    // role in the per-superblock column reduction: metric rm, column rc of the superblock
    const int rm = u >> 6, rc = u & 63;
    const int k3d = A.k3d;
    const int lane = u & 31;

    static_assert(sizeof(PoseState) <= sizeof(T.state), "PoseState must fit TeamSmem::state");
    PoseState &st = *reinterpret_cast<PoseState *>(T.state);
    double &b1t = T.bt[0], &b2t = T.bt[1];

    {
        const float *src = A.src + (size_t)pair * N * 3;
        const float *tgt = A.tgt + (size_t)pair * M * 3;

        // target cloud -> shared SoA, +inf padded in x and y so every metric of a padded point is inf
        for (int j = threadIdx.x; j < pad; j += kDuoThreads) {
            const bool ok = j < M;
            S.tx[j] = ok ? __ldg(tgt + 3 * j + 0) : kInf;
            S.ty[j] = ok ? __ldg(tgt + 3 * j + 1) : kInf;
            S.tz[j] = ok ? __ldg(tgt + 3 * j + 2) : 0.f;
        }
        if (u == 0) {
            state_load(st, A, row);
            b1t = pow(A.beta1, (double)A.step0);
            b2t = pow(A.beta2, (double)A.step0);
            if (kPipe) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    duo_mbar_init(&T.full[i], kTeam / 32);  // one arrival per warp
                    duo_mbar_init(&T.empty[i], 2);          // the two reducer warps of the phase
                }
            }
        }
        int gs = 0;  // kPipe: phases swept by this team so far (all iterations), buffer = gs & 3
        __syncthreads();
        // team 1 starts half a sweep late (released from inside team 0's first sweep), so that from
        // then on one team's phases outside the sweep fall into the other team's sweep
        if (team == 1) asm volatile("bar.sync 3, %0;" ::"r"(kDuoThreads) : "memory");

        const int abase = A.row_base ? A.row_base[row] : A.angle_base;
        const float boff = (float)((double)abase * 3.14159274101257324 / 4.0);
        // d(mean over the batch)/d(row): per row when rows of several batches share the launch
        const float gscale = A.row_scale ? A.row_scale[row] : A.grad_scale;

        for (int it = 0; it < A.iters; ++it) {
            // ---- (a) pose from the 8 scalars
            if (u == 0) pose_forward(st, boff, A.tran_scale, A.tran_offset, T.pose);
            team_sync(team);
            float Rm[9], Tv[3];
#pragma unroll
            for (int i = 0; i < 9; ++i) Rm[i] = T.pose[i];
#pragma unroll
            for (int i = 0; i < 3; ++i) Tv[i] = T.pose[9 + i];

            // ---- (b) transform: y = R x + T  (bmm(src, R^T) + T, houv.py:102)
            for (int i = u; i < pad; i += kTeam) {
                if (i < N) {
                    const float x0 = __ldg(src + 3 * i), x1 = __ldg(src + 3 * i + 1),
                                x2 = __ldg(src + 3 * i + 2);
                    yx[i] = __fadd_rn(__fmaf_rn(x2, Rm[2], __fmaf_rn(x1, Rm[1], __fmul_rn(x0, Rm[0]))), Tv[0]);
                    yy[i] = __fadd_rn(__fmaf_rn(x2, Rm[5], __fmaf_rn(x1, Rm[4], __fmul_rn(x0, Rm[3]))), Tv[1]);
                    yz[i] = __fadd_rn(__fmaf_rn(x2, Rm[8], __fmaf_rn(x1, Rm[7], __fmul_rn(x0, Rm[6]))), Tv[2]);
                } else {
                    yx[i] = kInf; yy[i] = kInf; yz[i] = 0.f;
                }
            }
            // row bookkeeping of this sweep: shadow = +inf, winning groups = 0
#pragma unroll
            for (int c = 0; c < 8; ++c) T.shadow[u][c ^ (u & 7)] = make_float4(kInf, kInf, kInf, kInf);
            for (int i = u; i < 4 * 2 * kPts / 4; i += kTeam)
                reinterpret_cast<unsigned *>(&T.gid[0][0][0])[i] = 0u;
            team_sync(team);

            // ---- (c) the symmetric sweep: rows into registers, then the region becomes `part`
            float nqx[8], nqy[8], nqz[8], best[4][8];
            if (scan_active) {
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    nqx[r] = -yx[row0 + r]; nqy[r] = -yy[row0 + r]; nqz[r] = -yz[row0 + r];
                }
            }
#pragma unroll
            for (int r = 0; r < 8; ++r)
#pragma unroll
                for (int m = 0; m < 4; ++m) best[m][r] = kInf;
            float vcolA = 0.f, vcolB = 0.f;  // sum sqrt(column minimum) of view metric rm, lower / upper superblocks
            team_sync(team);

            if constexpr (!kPipe) {
#pragma unroll 1
            for (int sb = 0; sb < nsb; ++sb) {
                if (scan_active) {
                    const int sbbase = sb * kSB;
#pragma unroll 1
                    for (int ph = 0; ph < 4; ++ph) {
                        const int cb = (gq + ph) & 3;  // 16-column block of this group in this phase
                        float cmx[4], cmy[4];  // running column minima of the pair in hand
#pragma unroll
                        for (int m = 0; m < 4; ++m) { cmx[m] = kInf; cmy[m] = kInf; }
                        const int cbase = sbbase + 16 * cb;
#pragma unroll kSymUnroll
                        for (int s = 0; s < 8; ++s) {
                            const int c = cbase + 2 * ((ll + s) & 7);
                            const float2 X = *reinterpret_cast<const float2 *>(&S.tx[c]);
                            const float2 Y = *reinterpret_cast<const float2 *>(&S.ty[c]);
                            const float2 Z = *reinterpret_cast<const float2 *>(&S.tz[c]);
                            sym_step<kRG, 4>(X, Y, Z, nqx, nqy, nqz, best, cmx, cmy);
                            if (s < 7) {  // the column pair moves on to lane ll-1; ours comes from ll+1
#pragma unroll
                                for (int m = 0; m < 4; ++m) {
                                    cmx[m] = __shfl_sync(0xffffffffu, cmx[m], ll + 1, 8);
                                    cmy[m] = __shfl_sync(0xffffffffu, cmy[m], ll + 1, 8);
                                }
                            }
                        }
                        // after step 7 this lane holds the finished pair (ll+7)&7 of block cb
                        const int pc = 16 * cb + 2 * ((ll + 7) & 7);
#pragma unroll
                        for (int m = 0; m < 4; ++m)
                            *reinterpret_cast<float2 *>(&part[sg][m][pc]) = make_float2(cmx[m], cmy[m]);
                    }
                    // rows: did this superblock strictly improve the running minimum?  One bit per
                    // (metric, row); the shadow is refreshed unconditionally (best <= shadow always)
                    unsigned improved = 0;
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        const int m = c >> 1, r0 = 4 * (c & 1);
                        float4 &slot = T.shadow[u][c ^ (u & 7)];
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
                        T.gid[k >> 3][1][row0 + (k & 7)] = (unsigned char)sb;
                    }
                }
                if (team == 0 && it == 0 && sb == (nsb >> 1))
                    asm volatile("bar.arrive 3, %0;" ::"r"(kDuoThreads) : "memory");
                team_sync(team);
                // columns: first sub-group (ascending rows) holding the minimum
                {
                    const int col = sb * kSB + rc;
                    float bv = kInf;
                    int bs = 0;
                    for (int g = 0; g < nsg; ++g) {
                        const float v = part[g][rm][rc];
                        if (v < bv) { bv = v; bs = g; }
                    }
                    T.gid[rm][0][col] = (unsigned char)bs;
                    if (rm == 0) T.d3c[col] = bv;
                    else if (col < M) {
                        if (sb < nsbh) vcolA += sqrtf(bv);
                        else vcolB += sqrtf(bv);
                    }
                }
                team_sync(team);
            }
            } else {
                // ---- pipelined form: every 16-column phase has its own buffer of partial column
                // minima (4 buffers); the warps never stop at a barrier inside the sweep.  After
                // its phase gs a warp arrives on full[gs & 3]; the two warps whose turn it is
                // (warps 2*ph, 2*ph+1 for the phase ph = (gs-1) & 3 just behind) wait for that
                // buffer to be complete, reduce its 4 metrics x 16 columns over the 32 sub-groups
                // and release it (empty), which every warp checks before storing phase gs + 3.
                // All groups of a warp take the SAME 16-column block in a phase (ascending), so a
                // buffer is complete after one phase.
                const int gs0 = gs;  // a multiple of 4: (gs & 3) is the phase inside the superblock
                auto reduce_phase = [&](int p) {
                    const int b = p & 3, pl = p - gs0, sbp = pl >> 2, php = pl & 3;
                    duo_mbar_wait(&T.full[b], (unsigned)(p >> 2) & 1u);
                    const int m = (u >> 4) & 3, c16 = u & 15;
                    const float *pp = T.region + b * kPartBuf + m * kPartM + c16;
                    float bv = kInf;
                    int bs = 0;
                    if (nsg == 32) {
                        // minimum of each chunk of 8 sub-groups (FMNMX3 tree), first chunk holding
                        // the overall minimum (ascending, strict '<'), then its first sub-group
                        int bc = 0;
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            float v[8];
#pragma unroll
                            for (int k = 0; k < 8; ++k) v[k] = pp[(8 * c + k) * 16];
                            const float cm = min3(min3(v[0], v[1], v[2]), min3(v[3], v[4], v[5]),
                                                  fminf(v[6], v[7]));
                            if (cm < bv) { bv = cm; bc = c; }
                        }
                        const float *pc8 = pp + bc * 128;
                        bs = 8 * bc;
#pragma unroll
                        for (int k = 7; k >= 0; --k)
                            if (pc8[k * 16] == bv) bs = 8 * bc + k;
                    } else {
                        for (int g = 0; g < nsg; ++g) {
                            const float v = pp[g * 16];
                            if (v < bv) { bv = v; bs = g; }
                        }
                    }
                    const int col = sbp * kSB + 16 * php + c16;
                    T.gid[m][0][col] = (unsigned char)bs;
                    if (m == 0) T.d3c[col] = bv;
                    else if (col < M) {
                        if (sbp < nsbh) vcolA += sqrtf(bv);
                        else vcolB += sqrtf(bv);
                    }
                    __syncwarp();
                    if (lane == 0) duo_mbar_arrive(&T.empty[b]);
                };
#pragma unroll 1
                for (int sb = 0; sb < nsb; ++sb) {
                    const int sbbase = sb * kSB;
#pragma unroll 1
                    for (int ph = 0; ph < 4; ++ph) {
                        const int b = gs & 3;
                        if (scan_active) {
                            float cmx[4], cmy[4];  // running column minima of the pair in hand
#pragma unroll
                            for (int m = 0; m < 4; ++m) { cmx[m] = kInf; cmy[m] = kInf; }
                            const int cbase = sbbase + 16 * ph;
#pragma unroll kSymUnroll
                            for (int s = 0; s < 8; ++s) {
                                const int c = cbase + 2 * ((ll + s) & 7);
                                const float2 X = *reinterpret_cast<const float2 *>(&S.tx[c]);
                                const float2 Y = *reinterpret_cast<const float2 *>(&S.ty[c]);
                                const float2 Z = *reinterpret_cast<const float2 *>(&S.tz[c]);
                                sym_step<kRG, 4>(X, Y, Z, nqx, nqy, nqz, best, cmx, cmy);
                                if (s < 7) {
#pragma unroll
                                    for (int m = 0; m < 4; ++m) {
                                        cmx[m] = __shfl_sync(0xffffffffu, cmx[m], ll + 1, 8);
                                        cmy[m] = __shfl_sync(0xffffffffu, cmy[m], ll + 1, 8);
                                    }
                                }
                            }
                            // buffer b was last used by phase gs - 4: reduced by now?
                            if (gs >= 4) duo_mbar_wait(&T.empty[b], (unsigned)((gs >> 2) - 1) & 1u);
                            float *pw = T.region + b * kPartBuf + sg * 16 + 2 * ((ll + 7) & 7);
#pragma unroll
                            for (int m = 0; m < 4; ++m)
                                *reinterpret_cast<float2 *>(pw + m * kPartM) = make_float2(cmx[m], cmy[m]);
                        } else if (gs >= 4) {
                            // warps without rows store nothing but keep in step: an early arrival
                            // would be counted towards the buffer's previous use
                            duo_mbar_wait(&T.empty[b], (unsigned)((gs >> 2) - 1) & 1u);
                        }
                        __syncwarp();
                        if (lane == 0) duo_mbar_arrive(&T.full[b]);
                        if (gs > gs0 && ((gs - 1) & 3) == (u >> 6)) reduce_phase(gs - 1);
                        ++gs;
                    }
                    if (scan_active) {
                        unsigned improved = 0;
#pragma unroll
                        for (int c = 0; c < 8; ++c) {
                            const int m = c >> 1, r0 = 4 * (c & 1);
                            float4 &slot = T.shadow[u][c ^ (u & 7)];
                            const float4 pv = slot;
                            improved |= (best[m][r0 + 0] < pv.x ? 1u : 0u) << (4 * c + 0);
                            improved |= (best[m][r0 + 1] < pv.y ? 1u : 0u) << (4 * c + 1);
                            improved |= (best[m][r0 + 2] < pv.z ? 1u : 0u) << (4 * c + 2);
                            improved |= (best[m][r0 + 3] < pv.w ? 1u : 0u) << (4 * c + 3);
                            slot = make_float4(best[m][r0], best[m][r0 + 1], best[m][r0 + 2], best[m][r0 + 3]);
                        }
                        while (improved) {
                            const int k = __ffs(improved) - 1;
                            improved &= improved - 1;
                            T.gid[k >> 3][1][row0 + (k & 7)] = (unsigned char)sb;
                        }
                    }
                    if (team == 0 && it == 0 && sb == (nsb >> 1))
                        asm volatile("bar.arrive 3, %0;" ::"r"(kDuoThreads) : "memory");
                }
                if (((gs - 1) & 3) == (u >> 6)) reduce_phase(gs - 1);  // the last phase
                // the view column sums go to the threads team_sum expects them in (metric u >> 6,
                // column u & 63); this thread reduced metric (u >> 4) & 3, column 16 (u >> 6) + (u & 15)
                {
                    const int m = (u >> 4) & 3, col = 16 * (u >> 6) + (u & 15);
                    T.vbuf[0][m][col] = vcolA;
                    T.vbuf[1][m][col] = vcolB;
                }
                team_sync(team);
                vcolA = T.vbuf[0][rm][rc];
                vcolB = T.vbuf[1][rm][rc];
            }

            // ---- (d) the region becomes the transformed source again (+ the 3-D row minima)
            float sA[8], sB[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) { sA[i] = 0.f; sB[i] = 0.f; }
            if (scan_active) {
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    yx[row0 + r] = -nqx[r]; yy[row0 + r] = -nqy[r]; yz[row0 + r] = -nqz[r];
                }
#pragma unroll
                for (int m = 0; m < 4; ++m)
#pragma unroll
                    for (int r = 0; r < 8; ++r) {
                        const int i = row0 + r;
                        if (i < N) {
                            if (m == 0) d3r[i] = best[0][r];
                            else sA[5 + m - 1] += sqrtf(best[m][r]);  // sums[2 + 1*3 + (m-1)]
                        }
                    }
            }
            team_sync(team);

            // ---- (e) k-th smallest of d3c[0..M) and d3r[0..N): 4 x 8-bit radix select
            unsigned prefix[2] = {0u, 0u};
            int kk[2] = {k3d, k3d};
#pragma unroll 1
            for (int shift = 24; shift >= 0; shift -= 8) {
                for (int b = u; b < 512; b += kTeam) (&T.hist[0][0])[b] = 0;
                team_sync(team);
#pragma unroll
                for (int dir = 0; dir < 2; ++dir) {
                    const int len = dir ? N : M;
                    if (shift == 24) {
                        // sign + 7 exponent bits: a cloud's distances share one or two values, so the
                        // lanes of a warp would queue on the same counter -- one add per distinct digit
                        for (int q0 = 0; q0 < len; q0 += kTeam) {
                            const int q = q0 + u;
                            const unsigned bin = q < len ? __float_as_uint(d3p[dir][q]) >> 24 : 256u;
                            const unsigned peers = __match_any_sync(0xffffffffu, bin);
                            if (bin < 256u && lane == __ffs(peers) - 1)
                                atomicAdd(&T.hist[dir][bin], __popc(peers));
                        }
                        continue;
                    }
                    for (int q = u; q < len; q += kTeam) {
                        const unsigned bits = __float_as_uint(d3p[dir][q]);
                        if ((bits >> (shift + 8)) == (prefix[dir] >> (shift + 8)))
                            atomicAdd(&T.hist[dir][(bits >> shift) & 255u], 1);
                    }
                }
                team_sync(team);
                if (u < 64) {
                    const int dir = u >> 5, ln = u & 31;
                    int c[8], s = 0;
#pragma unroll
                    for (int b = 0; b < 8; ++b) { c[b] = T.hist[dir][8 * ln + b]; s += c[b]; }
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
                        T.selu[dir * 2] = prefix[dir] | ((unsigned)bin << shift);
                        T.selu[dir * 2 + 1] = (unsigned)rem;
                        T.selu[4 + dir] = (unsigned)cnt;  // after the last round: number of ties
                    }
                }
                team_sync(team);
#pragma unroll
                for (int dir = 0; dir < 2; ++dir) {
                    prefix[dir] = T.selu[dir * 2];
                    kk[dir] = (int)T.selu[dir * 2 + 1];
                }
            }
            float thr[2], tiew[2];
#pragma unroll
            for (int dir = 0; dir < 2; ++dir) {
                thr[dir] = __uint_as_float(prefix[dir]);
                tiew[dir] = (float)kk[dir] / (float)T.selu[4 + dir];
            }

            // ---- (f) the eight directional means and the four min() selections
#pragma unroll
            for (int dir = 0; dir < 2; ++dir) {
                const int len = dir ? N : M;
                for (int q = u; q < len; q += kTeam) {
                    const float d = d3p[dir][q];
                    if (d < thr[dir]) {
                        if (q & kTeam) sB[dir] += sqrtf(d);
                        else sA[dir] += sqrtf(d);
                    }
                }
            }
#pragma unroll
            for (int m = 0; m < 3; ++m) {
                sA[2 + m] = (rm == m + 1) ? vcolA : 0.f;
                sB[2 + m] = (rm == m + 1) ? vcolB : 0.f;
            }
            team_sum<8>(sA, sB, 0x1cu, T, team, u);
            float cdv[4][2];  // [metric][dir]
#pragma unroll
            for (int dir = 0; dir < 2; ++dir) {
                cdv[0][dir] = (sA[dir] + (float)kk[dir] * sqrtf(thr[dir])) / (float)k3d;
#pragma unroll
                for (int m = 0; m < 3; ++m) cdv[1 + m][dir] = sA[2 + dir * 3 + m] / (float)N;
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
            if (u == 0) {
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
            //          from the 64 candidates of the winning group (as pose_sym_kernel, kRecov = 2)
            float accA[12], accB[12];
#pragma unroll
            for (int i = 0; i < 12; ++i) { accA[i] = 0.f; accB[i] = 0.f; }
#pragma unroll
            for (int m = 0; m < 4; ++m) {
                if (m > 0 && A.wview == 0.f) continue;  // loss = w3d*min_1 only (train_utils.py:439)
                auto side_grad = [&](auto dirc) {
                constexpr int dir = decltype(dirc)::value;  // compile-time: smem bases become immediates
                const float *qB = dir ? yx : S.tx, *tB = dir ? S.tx : yx;
                const int nq = dir ? N : M;
                // d(mean over rows)/d(row) * weight / k      houv.py:124,222; m_u_c.py:94-95
                const float coef = gscale * (m == 0 ? A.w3d : A.wview) / (float)(m == 0 ? k3d : N);
                const bool known = m == 0 || dir == 1;  // the minimum itself is still on chip
                for (int q = u; q < nq; q += kTeam) {
                    float wt = 1.f;
                    float bd = kInf;
                    if (m == 0) {  // top-k membership of the trimmed 3-D term
                        const float d = d3p[dir][q];
                        wt = d < thr[dir] ? 1.f : (d == thr[dir] ? tiew[dir] : 0.f);
                        if (wt == 0.f) continue;
                        bd = d;
                    } else if (known) {
                        const int uq = q >> 3, rq = q & 7;
                        bd = sh[uq * 32 + (((2 * m + (rq >> 2)) ^ (uq & 7)) * 4 + (rq & 3))];
                    }
                    const int j0 = (int)T.gid[m][dir][q] * kSB;
                    const unsigned j0b = (unsigned)j0 * 4u, lane8 = (unsigned)lane * 8u;
                    const char *tb = reinterpret_cast<const char *>(tB);
                    const float qx = qB[q], qy = qB[kPts + q], qz = qB[2 * kPts + q];
                    int bj = j0;
                    if (known) {
                        unsigned mx = 0u, my = 0u;  // bit k: the pair visited at step k equals the minimum
#pragma unroll
                        for (int k = 0; k < kSB / 2; ++k) {
                            const unsigned off = ((lane8 + 8u * k) & 248u) | j0b;
                            const float2 X = *reinterpret_cast<const float2 *>(tb + off);
                            const float2 Y = *reinterpret_cast<const float2 *>(tb + 4 * kPts + off);
                            const float2 Z = *reinterpret_cast<const float2 *>(tb + 8 * kPts + off);
                            const float2 d = metric_dist2(m, X, Y, Z, qx, qy, qz);
                            if (d.x == bd) mx |= 1u << k;
                            if (d.y == bd) my |= 1u << k;
                        }
                        bj = j0 + first_match(mx, my, lane);  // a match always exists unless bd is NaN
                    } else {
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
                    }
                    // sqrt backward then chamfer backward (x2): d=0 gives inf*0 = NaN like autograd
                    const float gg = (coef * wt) / (2.f * sqrtf(bd)) * 2.f;
                    const float sgn = dir ? 1.f : -1.f;  // (transformed src point) - (tgt point)
                    const float v0 = m == 1 ? 0.f : gg * (sgn * (qx - tB[bj]));
                    const float v1 = m == 2 ? 0.f : gg * (sgn * (qy - tB[kPts + bj]));
                    const float v2 = m == 3 ? 0.f : gg * (sgn * (qz - tB[2 * kPts + bj]));
                    const int xi = dir ? q : bj;  // which source point the gradient lands on
                    const float x0 = __ldg(src + 3 * xi), x1 = __ldg(src + 3 * xi + 1),
                                x2 = __ldg(src + 3 * xi + 2);
                    const float c[12] = {v0 * x0, v0 * x1, v0 * x2, v1 * x0, v1 * x1, v1 * x2,
                                         v2 * x0, v2 * x1, v2 * x2, v0, v1, v2};
                    if (q & kTeam) {  // virtual thread u + 256 of pose_sym_kernel
#pragma unroll
                        for (int i = 0; i < 12; ++i) accB[i] += c[i];
                    } else {
#pragma unroll
                        for (int i = 0; i < 12; ++i) accA[i] += c[i];
                    }
                }
                };
                if (side[m]) side_grad(std::integral_constant<int, 1>{});
                else side_grad(std::integral_constant<int, 0>{});
            }
            team_sum<12>(accA, accB, 0u, T, team, u);

            // ---- (i) closed-form pose backward + Adam on the 8 scalars
            if (u == 0) {
                float GR[3][3], gT[3], g[8];
#pragma unroll
                for (int i = 0; i < 3; ++i) {
#pragma unroll
                    for (int j = 0; j < 3; ++j) GR[i][j] = accA[i * 3 + j];
                    gT[i] = accA[9 + i];
                }
                pose_backward(st, GR, gT, A.tran_scale, g);
                adam_step(st, g, A, b1t, b2t);
            }
            team_sync(team);
        }

        if (u == 0) state_store(st, A, row);
    }
}

static int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev >= 64) return 148;
    if (!cached[dev]) {
        int n = 0;
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        cached[dev] = n > 0 ? n : 148;
    }
    return cached[dev];
}

template <int kU, int kRG, bool kPipe>
static int launch_duo(const PoseArgs &a, cudaStream_t st) {
    static unsigned long long smem_set = 0;  // per-device flag bits
    if (int e = ensure_dyn_smem(pose_duo_kernel<kU, kRG, kPipe>, (int)sizeof(DuoSmem), smem_set)) return e;
    // One CTA per SM, two rows per CTA: whole waves of two-row CTAs, then -- when what is left fits
    // one wave of ONE-row CTAs (which lasts ~0.55 of a two-row wave) -- the rest, and an odd last
    // row, through pose_sym_kernel.  Same bits either way.
    const int sms = sm_count();
    int nduo = a.rows / 2;
    const int whole = nduo / sms * sms;
    if (whole > 0 && a.rows - 2 * whole <= sms) nduo = whole;
    if (nduo > 0) {
        pose_duo_kernel<kU, kRG, kPipe><<<nduo, kDuoThreads, sizeof(DuoSmem), st>>>(a);
        if (int e = check_launch()) return e;
        // rows without a partner of the same pair (neighbours of different pairs in a caller-made
        // row_pair, odd K): the one-row kernel on exactly those, every other CTA exits at once
        if (a.row_pair || (a.K & 1)) {
            PoseArgs b = a;
            b.only_unpaired = 1;
            b.grid_rows = 2 * nduo;
            if (int e = launch_pose_sym(b, 84, st)) return e;
        }
    }
    if (a.rows > 2 * nduo) {
        PoseArgs b = a;
        b.row_begin = 2 * nduo;
        b.grid_rows = a.rows - 2 * nduo;
        return launch_pose_sym(b, 84, st);
    }
    return 0;
}

// pipelined = column reduction overlapped with the sweep (variant 21); else one barrier pair per
// superblock (variant 20)
int launch_pose_duo(const PoseArgs &a, bool pipelined, cudaStream_t st) {
    return pipelined ? launch_duo<8, 4, true>(a, st) : launch_duo<8, 4, false>(a, st);
}

}  // namespace houv
