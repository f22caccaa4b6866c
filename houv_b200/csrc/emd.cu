// K7 -- EMD approximation (auction algorithm), all iterations in ONE launch, for sm_100a.
//
// Replaces emd_cuda_forward / emd_cuda_backward
// (/root/reference/utils/metrics/EMD/emd_cuda.cu:230-316; wrapper emd_module.py:40-90).  The reference
// launches seven kernels per iteration (clear, calc_unass_cnt, calc_unass_cnt_sum, calc_unass_idx, Bid,
// GetMax, Assign; :253-264) with every array in global memory; evaluation runs use 3000 iterations
// (registration/models/pcn.py:169), almost all of them with a handful of unassigned points, so the
// reference's cost is launch latency x 7 x iters.  Here one CTA owns one batch element for the whole
// auction: prices, bids, both assignment maps, the unassigned list and the target cloud (SoA) stay in
// shared memory (48 B per point, n <= 4096; larger clouds use a caller-provided global workspace through
// the same code, compiled a second time for generic addresses), phases are separated by __syncthreads() instead of launches, and the loop ends as
// soon as nothing is unassigned (the reference's remaining iterations are no-ops).
//
// Arithmetic is the reference's as nvcc compiles it (SASS of the reference extension, sm_100a):
//   value(k) = (float)((3.0 - (double)sqrtf(fma(dz,dz,fma(dx,dx,dy*dy)))) - (double)price[k])   (:140-144)
//   increment = (best - better) + eps (float, :175);  GetMax compares in double with +-1e-6 (:187)
//   dist = fma(dz,dz,fma(dx,dx,dy*dy)) (:222-226)
// The two places where the reference's result depends on scheduling are fixed the same way as in
// oracle/emd_oracle.c: equal values -> lowest k; bidders within 1e-6 of an object's maximum -> lowest j.
#include <cooperative_groups.h>
#include <limits.h>
#include <stdlib.h>

#include "common.cuh"

namespace houv {

constexpr int kEmdThreads = 1024;
constexpr int kEmdSmall = 32;    // bidders resolved inside one warp
constexpr int kEmdWords = 12;    // state words per point
constexpr int kEmdPtsN = 12288;   // candidates (16 B/point) still in shared memory up to here
constexpr int kEmdSmemN = 4096;   // 48 B/point of state in shared memory up to here

struct EmdArgs {
    const float *xyz1, *xyz2;
    float *dist;
    int *assignment;
    int *iters_used;   // [B] or null
    unsigned *ws;      // global state, 12*n words per batch element, when n > kEmdSmemN
    int B, n, iters;
    float eps;
};

struct Top2 {
    float best, better;
    int i, i2;  // indices of best and of better (i2 is only a hint for the next sweep's bound)
};

// winner = larger value, lower index on equal values; better = max(winner's better, loser's best)
__device__ __forceinline__ Top2 merge(const Top2 &a, const Top2 &b) {
    const bool b_wins = b.best > a.best || (b.best == a.best && (unsigned)b.i < (unsigned)a.i);
    const Top2 &w = b_wins ? b : a, &l = b_wins ? a : b;
    Top2 r;
    r.best = w.best;
    r.i = w.i;
    const bool from_loser = l.best > w.better;
    r.better = from_loser ? l.best : w.better;
    r.i2 = from_loser ? l.i : w.i2;
    return r;
}

__device__ __forceinline__ Top2 shfl_xor(const Top2 &t, int o) {
    Top2 u;
    u.best = __shfl_xor_sync(0xffffffffu, t.best, o);
    u.better = __shfl_xor_sync(0xffffffffu, t.better, o);
    u.i = __shfl_xor_sync(0xffffffffu, t.i, o);
    u.i2 = __shfl_xor_sync(0xffffffffu, t.i2, o);
    return u;
}

// The reference's value of candidate c = (x, y, z, price) for the bidder at (px, py, pz)   (:140-144)
__device__ __forceinline__ float emd_value(const float4 c, float px, float py, float pz) {
    const float dx = __fsub_rn(c.x, px), dy = __fsub_rn(c.y, py), dz = __fsub_rn(c.z, pz);
    const float sq = __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
    return (float)((3.0 - (double)__fsqrt_rn(sq)) - (double)c.w);
}

// One bidder's sweep over the candidates k = lane_g, lane_g + G, ... of pts[0..n): the exact two best values
// (reference arithmetic and tie rule) merged over the w = min(G, 32) lanes of the warp that serve it.
// o1, o2 = the bidder's best and second-best objects of its previous bid (-1: none).  Warp-convergent.
__device__ __forceinline__ Top2 sweep_bidder(bool valid, int j, int o1, int o2, const float *x1,
                                             const float4 *pts, int n, int lane_g, int G, int w) {
    Top2 t{-1e9f, -1e9f, -1, -1};
    float px = 0.f, py = 0.f, pz = 0.f;
    float bound = -1e9f;  // a lower bound of the bidder's final second best; -1e9 keeps everything
    // Two cheap lower bounds before the sweep.  (1) The bidder's best and second-best objects
    // of its previous bid, at today's prices: few prices moved since, so the smaller of the two
    // values is usually within a few eps of the final second best.  (2) The first candidate of
    // every lane, exactly; the second best of those w values.  Without them a lane that sees
    // only n/G candidates never builds a useful bound of its own.
    if (valid) {
        px = __ldg(x1 + 3 * j); py = __ldg(x1 + 3 * j + 1); pz = __ldg(x1 + 3 * j + 2);
        if (o1 >= 0 && o2 >= 0)
            bound = fminf(emd_value(pts[o1], px, py, pz), emd_value(pts[o2], px, py, pz));
        if (lane_g < n) {
            t.best = emd_value(pts[lane_g], px, py, pz);
            t.i = lane_g;
        }
    }
    if (w > 1) {
        Top2 g = t;
        for (int o = w >> 1; o > 0; o >>= 1) g = merge(g, shfl_xor(g, o));
        bound = fmaxf(bound, g.better);
    }
    if (valid) {
        // A candidate whose value is certainly below a lower bound of the bidder's final second
        // best cannot be among its two best, so it is dropped on the squared distance alone:
        // value < bound  <=>  sqrt(sq) > 3 - bound - price, tested with a margin 16x the rounding
        // of the quantities involved.  That keeps the square root, the two float->double
        // conversions, the double subtractions and the conversion back (XU / FP64 pipes) off all
        // but the few candidates that can matter.
        float fa = 3.f - bound;
        float fb = fa + 0x1p-20f * (fabsf(fa) + fabsf(bound) + 4.f);
#pragma unroll 2
        for (int k = lane_g + G; k < n; k += G) {
            const float4 c = pts[k];
            const float dx = __fsub_rn(c.x, px), dy = __fsub_rn(c.y, py), dz = __fsub_rn(c.z, pz);
            const float sq = __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
            const float th = fmaf(c.w, -(1.f - 0x1p-20f), fb);  // 3 - bound - price + margin
            if (th <= 0.f || sq > th * th) continue;
            const float d = (float)((3.0 - (double)__fsqrt_rn(sq)) - (double)c.w);
            if (d > t.best) { t.better = t.best; t.i2 = t.i; t.best = d; t.i = k; }
            else if (d > t.better) { t.better = d; t.i2 = k; }
            else continue;
            if (t.better > bound) {
                bound = t.better;
                fa = 3.f - bound;
                fb = fa + 0x1p-20f * (fabsf(fa) + fabsf(bound) + 4.f);
            }
        }
    }
    for (int o = w >> 1; o > 0; o >>= 1) t = merge(t, shfl_xor(t, o));
    return t;
}

// State of one batch element, 12 words per point: pts[n] = (x, y, z, price) of the target cloud and its
// prices packed so that one 16-byte load feeds a candidate evaluation, then max_inc, bid_inc, max_idx,
// ainv, asg, bid, sec, list.  kSmem: the block lives in dynamic shared memory (pointers provably shared,
// so LDS/STS/ATOMS); otherwise in the caller's workspace.
// kPts (only meaningful when !kSmem): the state is in the workspace but pts[] still fits shared memory
// (16 B per point, n <= 12288), which keeps the sweep's candidate loads on chip.
template <bool kSmem, bool kPts = kSmem>
__global__ void __launch_bounds__(kEmdThreads, 1) emd_auction_kernel(const EmdArgs A) {
    extern __shared__ __align__(16) unsigned sm_u[];
    __shared__ int s_cnt;
    __shared__ Top2 s_merge[kEmdThreads / 32];
    const int n = A.n, tid = threadIdx.x;
    for (int b = blockIdx.x; b < A.B; b += gridDim.x) {
        unsigned *base = kSmem ? sm_u : A.ws + (size_t)b * kEmdWords * n;  // 16-byte aligned (12 words)
        float4 *pts = reinterpret_cast<float4 *>(kPts ? sm_u : base);
        float *max_inc = reinterpret_cast<float *>(base + 4 * (size_t)n);
        float *bid_inc = max_inc + n;
        int *max_idx = reinterpret_cast<int *>(bid_inc + n);
        int *ainv = max_idx + n;
        int *asg = ainv + n;
        int *bid = asg + n;   // last best object of a point (its bid)
        int *sec = bid + n;   // last second-best object of a point
        int *list = sec + n;
        const float *x1 = A.xyz1 + (size_t)b * n * 3;
        const float *x2g = A.xyz2 + (size_t)b * n * 3;
        __syncthreads();
        for (int k = tid; k < n; k += kEmdThreads) {
            pts[k] = make_float4(x2g[3 * k], x2g[3 * k + 1], x2g[3 * k + 2], 0.f);
            max_inc[k] = 0.f; max_idx[k] = INT_MAX; ainv[k] = -1; asg[k] = -1; bid[k] = -1; sec[k] = -1;
        }
        int used = 0;
        bool have_list = false;  // the small path below leaves the next iteration's list behind
        for (int it = 0; it < A.iters; ++it) {
            const bool last = it == A.iters - 1;
            if (!have_list) {
                if (tid == 0) s_cnt = 0;
                __syncthreads();
                for (int j = tid; j < n; j += kEmdThreads)
                    if (asg[j] == -1) list[atomicAdd(&s_cnt, 1)] = j;
                __syncthreads();
            }
            const int U = s_cnt;
            if (U == 0) break;
            ++used;
            // ---- Bid: G threads per unassigned point (power of two), strided over the candidates
            int G = 1;
            while (G * 2 * U <= kEmdThreads) G *= 2;
            const int groups = kEmdThreads / G;
            const int lane_g = tid & (G - 1);
            const int w = G < 32 ? G : 32;  // lanes of a warp that serve the same bidder
            for (int p0 = 0; p0 < U; p0 += groups) {
                const int p = p0 + tid / G;
                const bool valid = p < U;
                const int j = valid ? list[p] : 0;
                const Top2 t = sweep_bidder(valid, j, valid ? bid[j] : -1, valid ? sec[j] : -1, x1, pts, n, lane_g, G, w);
                if (U <= kEmdSmall) {
                    // at most 32 bidders: one sweep round; hand the per-warp (G > 32) or per-bidder results
                    // to warp 0 through shared memory
                    if (G > 32) { if ((tid & 31) == 0) s_merge[tid >> 5] = t; }
                    else if (valid && lane_g == 0) s_merge[p] = t;
                } else if (valid && lane_g == 0) {
                    const float inc = __fadd_rn(__fsub_rn(t.best, t.better), A.eps);
                    bid[j] = t.i;
                    sec[j] = t.i2;
                    bid_inc[j] = inc;
                    // float max through the signed-int order: increments are >= 0, the cleared
                    // value -1e9 is negative (emd_cuda.cu:10-20 is the same max)
                    atomicMax(reinterpret_cast<int *>(max_inc) + t.i, __float_as_int(inc));
                }
            }
            __syncthreads();
            if (U <= kEmdSmall) {
                // ---- GetMax + Assign for <= 32 bidders inside warp 0, lane = bidder: no scratch arrays, no
                // block barriers between the phases, and the next iteration's unassigned list (losers and
                // displaced owners) is written directly instead of rescanning all n points.
                if (tid < 32) {
                    const int lane = tid;
                    Top2 t = s_merge[lane];
                    if (G > 32) {
                        const int wb = G >> 5;  // warps (= lanes here) per bidder, a power of two
                        for (int o = wb >> 1; o > 0; o >>= 1) t = merge(t, shfl_xor(t, o));
                        const int src = (lane * wb) & 31;
                        t.best = __shfl_sync(0xffffffffu, t.best, src);
                        t.better = __shfl_sync(0xffffffffu, t.better, src);
                        t.i = __shfl_sync(0xffffffffu, t.i, src);
                        t.i2 = __shfl_sync(0xffffffffu, t.i2, src);
                    }
                    const bool bidder = lane < U;
                    const int j = bidder ? list[lane] : -1;
                    const int o = bidder ? t.i : -1 - lane;  // distinct dummies never match
                    const float inc = __fadd_rn(__fsub_rn(t.best, t.better), A.eps);
                    float mx = inc;
                    for (int q = 0; q < U; ++q) {
                        const int oq = __shfl_sync(0xffffffffu, o, q);
                        const float iq = __shfl_sync(0xffffffffu, inc, q);
                        if (oq == o && iq > mx) mx = iq;
                    }
                    const bool qual = (double)inc - 1e-6 <= (double)mx && (double)mx <= (double)inc + 1e-6;
                    bool win = qual;
                    for (int q = 0; q < U; ++q) {
                        const int oq = __shfl_sync(0xffffffffu, o, q);
                        const int jq = __shfl_sync(0xffffffffu, j, q);
                        const bool qq = __shfl_sync(0xffffffffu, (int)qual, q) != 0;
                        if (oq == o && qq && jq < j) win = false;
                    }
                    win = bidder && (win || last);
                    int item = -1;  // this lane's contribution to the next list
                    if (bidder) { bid[j] = t.i; sec[j] = t.i2; }
                    if (win) {
                        const int prev = ainv[o];
                        if (!last && prev != -1) { asg[prev] = -1; item = prev; }
                        ainv[o] = j;
                        asg[j] = o;
                        pts[o].w = __fadd_rn(pts[o].w, inc);
                    } else if (bidder) {
                        item = j;
                    }
                    const unsigned has = __ballot_sync(0xffffffffu, item != -1);
                    __syncwarp();  // every lane has read list[] before it is overwritten
                    if (item != -1) list[__popc(has & ((1u << lane) - 1))] = item;
                    if (lane == 0) s_cnt = __popc(has);
                }
                have_list = true;
                __syncthreads();
                continue;
            }
            have_list = false;
            // ---- GetMax (:181-194): lowest j among the bidders within 1e-6 of the object's maximum
            for (int p = tid; p < U; p += kEmdThreads) {
                const int j = list[p], o = bid[j];
                const double bi = (double)bid_inc[j], mx = (double)max_inc[o];
                if (bi - 1e-6 <= mx && mx <= bi + 1e-6) atomicMin(max_idx + o, j);
            }
            __syncthreads();
            // ---- Assign (:196-216)
            for (int p = tid; p < U; p += kEmdThreads) {
                const int j = list[p], o = bid[j];
                if (last || max_idx[o] == j) {
                    const int prev = ainv[o];
                    if (!last && prev != -1) asg[prev] = -1;
                    ainv[o] = j;
                    asg[j] = o;
                    pts[o].w = __fadd_rn(pts[o].w, bid_inc[j]);
                    max_inc[o] = -1e9f;
                    max_idx[o] = INT_MAX;
                }
            }
            __syncthreads();
        }
        __syncthreads();
        // ---- CalcDist (:218-228)
        for (int j = tid; j < n; j += kEmdThreads) {
            const int k = asg[j];
            const float dx = __fsub_rn(x1[3 * j], x2g[3 * k]), dy = __fsub_rn(x1[3 * j + 1], x2g[3 * k + 1]),
                        dz = __fsub_rn(x1[3 * j + 2], x2g[3 * k + 2]);
            A.dist[(size_t)b * n + j] = __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
            A.assignment[(size_t)b * n + j] = k;
        }
        if (A.iters_used && tid == 0) A.iters_used[b] = used;
    }
}

// The same auction with kC CTAs (a thread-block cluster) per batch element, for batches too small to give
// every SM an element of its own.  The bidders of an iteration are dealt round-robin to the CTAs; each
// CTA keeps its OWN copy of pts[] (target cloud + prices) in shared memory and sweeps all n candidates for
// its bidders with its 1024 threads; everything else (list, bids, assignment maps) lives in rank 0's shared
// memory and is read / written by the other ranks through distributed shared memory.  Rank 0 resolves the
// iteration (GetMax + Assign) and broadcasts the new prices into every copy.  Two cluster barriers per
// iteration (~0.2 us each) replace nothing in the reference: its seven launches per iteration stay gone.
// Results are bit-identical to the single-CTA kernel (same per-bidder arithmetic, same tie rules).
// kState: the master state sits in rank 0's shared memory (n <= 4096); otherwise in the caller's workspace
// (global memory, visible to every rank after the cluster barriers) while every CTA still keeps its own
// copy of pts[] in shared memory (n <= 12288).
template <int kC, bool kState>
__global__ void __launch_bounds__(kEmdThreads, 1) emd_auction_cluster_kernel(const EmdArgs A) {
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    extern __shared__ __align__(16) unsigned sm_u[];
    __shared__ int s_cnt;
    __shared__ Top2 s_merge[kEmdThreads / 32];  // per-warp partials of this CTA
    __shared__ Top2 s_fin[kEmdSmall];           // rank 0: final per-bidder results of the small path
    const int n = A.n, tid = threadIdx.x;
    const int rank = (int)cluster.block_rank();
    const int n_clusters = gridDim.x / kC;
    // local copy of the candidates; master state on rank 0 (generic pointers into its shared memory)
    float4 *pts = reinterpret_cast<float4 *>(sm_u);
    int *m_cnt = cluster.map_shared_rank(&s_cnt, 0);
    Top2 *m_fin = cluster.map_shared_rank(s_fin, 0);
    for (int b = blockIdx.x / kC; b < A.B; b += n_clusters) {
        unsigned *base0 = kState ? cluster.map_shared_rank(sm_u, 0) : A.ws + (size_t)b * kEmdWords * n;
        float *max_inc = reinterpret_cast<float *>(base0 + 4 * (size_t)n);
        float *bid_inc = max_inc + n;
        int *max_idx = reinterpret_cast<int *>(bid_inc + n);
        int *ainv = max_idx + n;
        int *asg = ainv + n;
        int *bid = asg + n;
        int *sec = bid + n;
        int *list = sec + n;
        const float *x1 = A.xyz1 + (size_t)b * n * 3;
        const float *x2g = A.xyz2 + (size_t)b * n * 3;
        for (int k = tid; k < n; k += kEmdThreads) {
            pts[k] = make_float4(x2g[3 * k], x2g[3 * k + 1], x2g[3 * k + 2], 0.f);
            if (rank == 0) {
                max_inc[k] = 0.f; max_idx[k] = INT_MAX; ainv[k] = -1; asg[k] = -1; bid[k] = -1; sec[k] = -1;
            }
        }
        int used = 0;
        bool have_list = false;
        for (int it = 0; it < A.iters; ++it) {
            const bool last = it == A.iters - 1;
            if (!have_list && rank == 0) {
                if (tid == 0) s_cnt = 0;
                __syncthreads();
                for (int j = tid; j < n; j += kEmdThreads)
                    if (asg[j] == -1) list[atomicAdd(&s_cnt, 1)] = j;
            }
            cluster.sync();  // list, count, prices of every copy and the hints are in place
            const int U = *m_cnt;
            if (U == 0) break;
            ++used;
            // ---- Bid: this CTA's bidders are p = rank, rank + kC, ...
            const int Ur = U > rank ? (U - rank + kC - 1) / kC : 0;
            const bool small = U <= kEmdSmall;  // then Ur <= 8 and G >= 128: one round, warp partials
            int G = kEmdThreads;
            if (Ur > 0) { G = 1; while (G * 2 * Ur <= kEmdThreads) G *= 2; }
            if (!small && G > 32) G = 32;  // the array path finishes a bidder inside one warp
            const int groups = kEmdThreads / G;
            const int lane_g = tid & (G - 1);
            const int w = G < 32 ? G : 32;
            for (int q0 = 0; q0 < Ur; q0 += groups) {
                const int q = q0 + tid / G;
                const bool valid = q < Ur;
                const int j = valid ? list[q * kC + rank] : 0;
                const Top2 t = sweep_bidder(valid, j, valid ? bid[j] : -1, valid ? sec[j] : -1, x1, pts, n,
                                            lane_g, G, w);
                if (small) {
                    if ((tid & 31) == 0) s_merge[tid >> 5] = t;
                } else if (valid && lane_g == 0) {
                    const float inc = __fadd_rn(__fsub_rn(t.best, t.better), A.eps);
                    bid[j] = t.i;
                    sec[j] = t.i2;
                    bid_inc[j] = inc;
                    atomicMax(reinterpret_cast<int *>(max_inc) + t.i, __float_as_int(inc));
                }
            }
            if (small && Ur > 0) {
                // this CTA's warp 0 finishes its bidders (G / 32 warp partials each) and posts them to rank 0
                __syncthreads();
                if (tid < 32) {
                    Top2 t = s_merge[tid];
                    const int wb = G >> 5;
                    for (int o = wb >> 1; o > 0; o >>= 1) t = merge(t, shfl_xor(t, o));
                    if (tid % wb == 0 && tid / wb < Ur) m_fin[(tid / wb) * kC + rank] = t;
                }
            }
            cluster.sync();  // every bid of the iteration has reached rank 0
            have_list = small;
            if (rank != 0) continue;
            if (small) {
                // ---- GetMax + Assign inside warp 0 of rank 0, lane = bidder (as in the single-CTA kernel)
                if (tid < 32) {
                    const int lane = tid;
                    const Top2 t = s_fin[lane];
                    const bool bidder = lane < U;
                    const int j = bidder ? list[lane] : -1;
                    const int o = bidder ? t.i : -1 - lane;
                    const float inc = __fadd_rn(__fsub_rn(t.best, t.better), A.eps);
                    float mx = inc;
                    for (int q = 0; q < U; ++q) {
                        const int oq = __shfl_sync(0xffffffffu, o, q);
                        const float iq = __shfl_sync(0xffffffffu, inc, q);
                        if (oq == o && iq > mx) mx = iq;
                    }
                    const bool qual = (double)inc - 1e-6 <= (double)mx && (double)mx <= (double)inc + 1e-6;
                    bool win = qual;
                    for (int q = 0; q < U; ++q) {
                        const int oq = __shfl_sync(0xffffffffu, o, q);
                        const int jq = __shfl_sync(0xffffffffu, j, q);
                        const bool qq = __shfl_sync(0xffffffffu, (int)qual, q) != 0;
                        if (oq == o && qq && jq < j) win = false;
                    }
                    win = bidder && (win || last);
                    int item = -1;
                    if (bidder) { bid[j] = t.i; sec[j] = t.i2; }
                    if (win) {
                        const int prev = ainv[o];
                        if (!last && prev != -1) { asg[prev] = -1; item = prev; }
                        ainv[o] = j;
                        asg[j] = o;
                        const float np = __fadd_rn(pts[o].w, inc);
#pragma unroll
                        for (int r = 0; r < kC; ++r) cluster.map_shared_rank(pts, r)[o].w = np;
                    } else if (bidder) {
                        item = j;
                    }
                    const unsigned has = __ballot_sync(0xffffffffu, item != -1);
                    __syncwarp();
                    if (item != -1) list[__popc(has & ((1u << lane) - 1))] = item;
                    if (lane == 0) s_cnt = __popc(has);
                }
                continue;
            }
            // ---- GetMax / Assign over the arrays (more than 32 bidders), rank 0 only
            for (int p = tid; p < U; p += kEmdThreads) {
                const int j = list[p], o = bid[j];
                const double bi = (double)bid_inc[j], mx = (double)max_inc[o];
                if (bi - 1e-6 <= mx && mx <= bi + 1e-6) atomicMin(max_idx + o, j);
            }
            __syncthreads();
            for (int p = tid; p < U; p += kEmdThreads) {
                const int j = list[p], o = bid[j];
                if (last || max_idx[o] == j) {
                    const int prev = ainv[o];
                    if (!last && prev != -1) asg[prev] = -1;
                    ainv[o] = j;
                    asg[j] = o;
                    const float np = __fadd_rn(pts[o].w, bid_inc[j]);
#pragma unroll
                    for (int r = 0; r < kC; ++r) cluster.map_shared_rank(pts, r)[o].w = np;
                    max_inc[o] = -1e9f;
                    max_idx[o] = INT_MAX;
                }
            }
            __syncthreads();
        }
        cluster.sync();  // rank 0 is done writing into the other copies; its state is final
        if (rank == 0) {
            for (int j = tid; j < n; j += kEmdThreads) {
                const int k = asg[j];
                const float dx = __fsub_rn(x1[3 * j], x2g[3 * k]), dy = __fsub_rn(x1[3 * j + 1], x2g[3 * k + 1]),
                            dz = __fsub_rn(x1[3 * j + 2], x2g[3 * k + 2]);
                A.dist[(size_t)b * n + j] = __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
                A.assignment[(size_t)b * n + j] = k;
            }
            if (A.iters_used && tid == 0) A.iters_used[b] = used;
        }
        cluster.sync();  // nobody re-initialises its copy while rank 0 still reads or others still write
    }
}

// emd_cuda_backward (:270-305): only xyz1 receives a gradient; every point owns its row.
__global__ void __launch_bounds__(256) emd_grad_kernel(const float *__restrict__ xyz1,
                                                       const float *__restrict__ xyz2,
                                                       const float *__restrict__ graddist,
                                                       const int *__restrict__ assignment, long long total,
                                                       int n, float *__restrict__ gradxyz1) {
    for (long long e = blockIdx.x * 256LL + threadIdx.x; e < total; e += 256LL * gridDim.x) {
        const long long b = e / n;
        const float g = __fmul_rn(graddist[e], 2.f);
        const float *p = xyz1 + 3 * e, *q = xyz2 + 3 * (b * n + assignment[e]);
#pragma unroll
        for (int c = 0; c < 3; ++c)
            gradxyz1[3 * e + c] = __fadd_rn(gradxyz1[3 * e + c], __fmul_rn(g, __fsub_rn(p[c], q[c])));
    }
}

}  // namespace houv

using namespace houv;

extern "C" size_t houv_emd_workspace_bytes(int B, int n) {
    if (B <= 0 || n <= kEmdSmemN) return 0;
    return sizeof(unsigned) * kEmdWords * (size_t)n * (size_t)B;
}

extern "C" int houv_emd_forward(const float *xyz1, const float *xyz2, int B, int n, float eps, int iters,
                                float *dist, int *assignment, int *iters_used, void *workspace,
                                size_t workspace_bytes, void *stream) {
    if (B < 0 || n < 0 || iters <= 0 || !(eps >= 0.f)) return HOUV_E_BADARG;
    if (B == 0 || n == 0) return 0;
    if (!xyz1 || !xyz2 || !dist || !assignment) return HOUV_E_BADARG;
    const size_t need = houv_emd_workspace_bytes(B, n);
    if (need && (!workspace || workspace_bytes < need)) return HOUV_E_BADARG;
    EmdArgs a;
    a.xyz1 = xyz1; a.xyz2 = xyz2; a.dist = dist; a.assignment = assignment; a.iters_used = iters_used;
    a.ws = (unsigned *)workspace; a.B = B; a.n = n; a.iters = iters; a.eps = eps;
    const int grid = B < 148 * 4 ? B : 148 * 4;
    // small batches: a cluster of 4 CTAs per element (HOUV_EMD_CLUSTER=0 turns it off, =1 forces it)
    const int cl = getenv("HOUV_EMD_CLUSTER") ? atoi(getenv("HOUV_EMD_CLUSTER")) : -1;
    constexpr int kC = 4;
    if (n <= kEmdPtsN && n >= 256 && (cl == 1 || (cl < 0 && B * kC <= 148))) {
        const bool state_smem = n <= kEmdSmemN;
        static unsigned long long set_a = 0, set_b = 0;  // per-device flag bits
        if (int e = ensure_dyn_smem(emd_auction_cluster_kernel<kC, true>,
                                    (int)(sizeof(unsigned) * kEmdWords * kEmdSmemN), set_a))
            return e;
        if (int e = ensure_dyn_smem(emd_auction_cluster_kernel<kC, false>, (int)(sizeof(float4) * kEmdPtsN), set_b))
            return e;
        cudaLaunchConfig_t cfg = {};
        const int clusters = B < 32 ? B : 32;
        cfg.gridDim = dim3(clusters * kC);
        cfg.blockDim = dim3(kEmdThreads);
        cfg.dynamicSmemBytes = state_smem ? sizeof(unsigned) * kEmdWords * (size_t)n : sizeof(float4) * (size_t)n;
        cfg.stream = (cudaStream_t)stream;
        cudaLaunchAttribute attr;
        attr.id = cudaLaunchAttributeClusterDimension;
        attr.val.clusterDim.x = kC; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
        cfg.attrs = &attr;
        cfg.numAttrs = 1;
        cudaError_t e = state_smem ? cudaLaunchKernelEx(&cfg, emd_auction_cluster_kernel<kC, true>, a)
                                   : cudaLaunchKernelEx(&cfg, emd_auction_cluster_kernel<kC, false>, a);
        if (e == cudaSuccess) return check_launch();
        // a partition that cannot co-schedule 4 such CTAs (MIG slice, MPS limit): the one-CTA kernels below
        (void)cudaGetLastError();
        if (cl == 1) return (int)e;
    }
    if (n <= kEmdSmemN) {
        static unsigned long long smem_set = 0;  // per-device flag bits
        if (int e = ensure_dyn_smem(emd_auction_kernel<true>,
                                    (int)(sizeof(unsigned) * kEmdWords * kEmdSmemN), smem_set))
            return e;
        emd_auction_kernel<true><<<grid, kEmdThreads, sizeof(unsigned) * kEmdWords * (size_t)n, (cudaStream_t)stream>>>(a);
    } else if (n <= kEmdPtsN) {
        static unsigned long long smem_set = 0;  // per-device flag bits
        if (int e = ensure_dyn_smem(emd_auction_kernel<false, true>, (int)(sizeof(float4) * kEmdPtsN), smem_set))
            return e;
        emd_auction_kernel<false, true><<<grid, kEmdThreads, sizeof(float4) * (size_t)n, (cudaStream_t)stream>>>(a);
    } else {
        emd_auction_kernel<false, false><<<grid, kEmdThreads, 0, (cudaStream_t)stream>>>(a);
    }
    return check_launch();
}

extern "C" int houv_emd_backward(const float *xyz1, const float *xyz2, const float *graddist,
                                 const int *assignment, int B, int n, float *gradxyz1, void *stream) {
    if (B < 0 || n < 0) return HOUV_E_BADARG;
    if (B == 0 || n == 0) return 0;
    if (!xyz1 || !xyz2 || !graddist || !assignment || !gradxyz1) return HOUV_E_BADARG;
    const long long total = (long long)B * n;
    long long blocks = (total + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    emd_grad_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(xyz1, xyz2, graddist, assignment, total,
                                                                       n, gradxyz1);
    return check_launch();
}
