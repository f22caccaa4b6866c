/* ORACLE -- TEST INFRASTRUCTURE ONLY (never linked into houv_b200).
 *
 * CPU restatement of the reference's EMD approximation (auction algorithm),
 * /root/reference/utils/metrics/EMD/emd_cuda.cu:
 *   per iteration (emd_cuda_forward :253-266): list the unassigned points of xyz1
 *   (:30-94), Bid (:95-179), GetMax (:181-194), Assign (:196-216); CalcDist at the end (:218-228).
 * Arithmetic as compiled by nvcc for the reference (SASS of oracle/_ref/emd.so, sm_100a):
 *   value(k) = (float)((3.0 - (double)sqrtf(fma(dz,dz,fma(dx,dx,dy*dy)))) - (double)price[k]),
 *              d = xyz2[k] - xyz1[j]                                  (:140-144, `3.0` is a double)
 *   increment = (best - better) + eps in float                         (:175)
 *   dist      = fma(dz,dz,fma(dx,dx,dy*dy)), d = xyz1[j] - xyz2[k]     (:222-226)
 * The reference leaves two things to the hardware's scheduling; this restatement (and the CUDA
 * kernel it checks) fixes them:
 *   - equal values in Bid: the lowest index k wins (what the reference's thread layout gives for
 *     n <= 2048, one shared-memory batch; :125-157, :164-173);
 *   - several bidders for one object whose increments lie within 1e-6 of the maximum (GetMax
 *     :187-191, last writer wins in the reference): the lowest j wins.
 * Iterations after every point is assigned do nothing in the reference; the loop stops there.
 */
#include <math.h>
#include <stdlib.h>

static float value_of(const float *p1, const float *p2, float price) {
    float dx = p2[0] - p1[0], dy = p2[1] - p1[1], dz = p2[2] - p1[2];
    float sq = fmaf(dz, dz, fmaf(dx, dx, dy * dy));
    return (float)((3.0 - (double)sqrtf(sq)) - (double)price);
}

/* xyz1, xyz2: [B,n,3]; dist [B,n]; assignment [B,n]; price_out [B,n] or NULL;
 * iters_used [B] or NULL = iterations that had at least one unassigned point. */
int emd_oracle_forward(const float *xyz1, const float *xyz2, int B, int n, float eps, int iters,
                       float *dist, int *assignment, float *price_out, int *iters_used) {
    if (B < 0 || n <= 0 || iters <= 0) return -1;
    float *price = (float *)malloc(sizeof(float) * n);
    float *max_inc = (float *)malloc(sizeof(float) * n);
    float *bid_inc = (float *)malloc(sizeof(float) * n);
    int *bid = (int *)malloc(sizeof(int) * n);
    int *max_idx = (int *)malloc(sizeof(int) * n);
    int *ainv = (int *)malloc(sizeof(int) * n);
    int *list = (int *)malloc(sizeof(int) * n);
    for (int b = 0; b < B; ++b) {
        const float *x1 = xyz1 + (size_t)b * n * 3, *x2 = xyz2 + (size_t)b * n * 3;
        int *as = assignment + (size_t)b * n;
        for (int k = 0; k < n; ++k) {
            price[k] = 0.f; max_inc[k] = 0.f; max_idx[k] = 0; ainv[k] = -1; as[k] = -1; /* emd_module.py:57-66 */
        }
        int used = 0;
        for (int it = 0; it < iters; ++it) {
            const int last = it == iters - 1;
            int U = 0;
            for (int j = 0; j < n; ++j)
                if (as[j] == -1) list[U++] = j;
            if (U == 0) break;
            ++used;
            for (int p = 0; p < U; ++p) { /* Bid */
                const int j = list[p];
                float best = -1e9f, better = -1e9f;
                int best_i = -1;
                for (int k = 0; k < n; ++k) {
                    const float d = value_of(x1 + 3 * j, x2 + 3 * k, price[k]);
                    if (d > best) { better = best; best = d; best_i = k; }
                    else if (d > better) better = d;
                }
                const float inc = (best - better) + eps;
                bid[j] = best_i;
                bid_inc[j] = inc;
                if (inc > max_inc[best_i]) max_inc[best_i] = inc;
            }
            for (int p = 0; p < U; ++p) max_idx[bid[list[p]]] = 0x7fffffff;
            for (int p = 0; p < U; ++p) { /* GetMax, lowest j among the near-maximal bidders */
                const int j = list[p], o = bid[j];
                const double bi = (double)bid_inc[j], mx = (double)max_inc[o];
                if (bi - 1e-6 <= mx && mx <= bi + 1e-6 && j < max_idx[o]) max_idx[o] = j;
            }
            for (int p = 0; p < U; ++p) { /* Assign */
                const int j = list[p], o = bid[j];
                if (last || max_idx[o] == j) {
                    const int prev = ainv[o];
                    if (!last && prev != -1) as[prev] = -1;
                    ainv[o] = j;
                    as[j] = o;
                    price[o] += bid_inc[j];
                    max_inc[o] = -1e9f;
                }
            }
        }
        for (int j = 0; j < n; ++j) { /* CalcDist */
            const float *p = x1 + 3 * j, *q = x2 + 3 * as[j];
            float dx = p[0] - q[0], dy = p[1] - q[1], dz = p[2] - q[2];
            dist[(size_t)b * n + j] = fmaf(dz, dz, fmaf(dx, dx, dy * dy));
        }
        if (price_out)
            for (int k = 0; k < n; ++k) price_out[(size_t)b * n + k] = price[k];
        if (iters_used) iters_used[b] = used;
    }
    free(price); free(max_inc); free(bid_inc); free(bid); free(max_idx); free(ainv); free(list);
    return 0;
}

/* emd_cuda_backward (:270-305): gradxyz1[j] += 2*graddist[j]*(xyz1[j] - xyz2[assignment[j]]). */
int emd_oracle_backward(const float *xyz1, const float *xyz2, const float *graddist,
                        const int *assignment, int B, int n, float *gradxyz1) {
    for (size_t e = 0; e < (size_t)B * n; ++e) {
        const size_t b = e / n;
        const float *p = xyz1 + 3 * e, *q = xyz2 + 3 * (b * n + assignment[e]);
        const float g = graddist[e] * 2;
        for (int c = 0; c < 3; ++c) gradxyz1[3 * e + c] += g * (p[c] - q[c]);
    }
    return 0;
}
