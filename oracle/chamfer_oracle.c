/* ORACLE -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement of the reference's Chamfer operator.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this library; the product (houv_b200/) never does.
 *
 * Parity status: PINNED.  tests/test_oracle_cpu.py checks this file against
 *   (1) oracle/_ref/libref_chamfer.so = the reference's own kernel bodies
 *       (chamfer3D.cu:12-134, 155-174) compiled for the host (oracle/Makefile),
 *   (2) the reference's chamfer_python.distChamfer golden vectors
 *       (tests/golden/chamfer_unit.npz, made by tests/golden/make_golden.py),
 *   with idx bit-exact, as the reference's own unit test demands
 *   (utils/metrics/CD/unit_test.py:29-33).
 *
 * Arithmetic follows what nvcc 12.9 emits for chamfer3D.cu:32-35 at sm_100a
 * (SASS: FADD x3, FMUL(y), FFMA(x), FFMA(z)):
 *        d = fma(dz, dz, fma(dx, dx, dy*dy)),  dx = tgt.x - qry.x ...
 * Build with -ffp-contract=off so that only the explicit fmaf() calls fuse.
 */
#include <math.h>
#include <pthread.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#define ORACLE_TILE 512 /* chamfer3D.cu:13 */

static inline float sqdist(float qx, float qy, float qz, const float *t) {
    float dx = t[0] - qx, dy = t[1] - qy, dz = t[2] - qz;
    return fmaf(dz, dz, fmaf(dx, dx, dy * dy));
}

/* One directed search, chamfer3D.cu:12-134: for every query of xyz[b,n,3] the
 * squared distance to, and index of, the first nearest point of xyz2[b,m,3].
 * The tile loop is kept because it is observable through NaNs (:36,:126). */
typedef struct {
    int n, m;
    const float *xyz, *xyz2;
    float *result;
    int *result_i;
    int64_t w0, w1;
} nn_job;

/* Eight queries of one batch element at a time (AVX2 + FMA), targets broadcast: per lane exactly
 * the scalar sequence above -- vsubps x3, vmulps (dy*dy), vfmadd (dx), vfmadd (dz), ordered strict
 * '<' (a NaN distance never wins except at tile position 0, chamfer3D.cu:36), blend -- so the
 * results are bit-identical to the scalar loop (tests/test_oracle_cpu.py compares both with
 * oracle/_ref).  Only the host time changes: the full-size goldens and the timed CPU baseline
 * need 2048 x 2048 x 64 searches per iteration. */
#if defined(__x86_64__)
#include <immintrin.h>
__attribute__((target("avx2,fma"))) static void nn_rows8(const float *q /*8 queries, AoS*/,
                                                          int nq, const float *tg, int m,
                                                          float *result, int *result_i) {
    float qx[8], qy[8], qz[8];
    for (int l = 0; l < 8; ++l) {
        int s = l < nq ? l : 0;
        qx[l] = q[3 * s]; qy[l] = q[3 * s + 1]; qz[l] = q[3 * s + 2];
    }
    const __m256 QX = _mm256_loadu_ps(qx), QY = _mm256_loadu_ps(qy), QZ = _mm256_loadu_ps(qz);
    __m256 res = _mm256_setzero_ps();
    __m256i resi = _mm256_setzero_si256();
    for (int k2 = 0; k2 < m; k2 += ORACLE_TILE) {
        int end_k = (m < k2 + ORACLE_TILE ? m : k2 + ORACLE_TILE) - k2;
        __m256 best = _mm256_setzero_ps();
        __m256i besti = _mm256_setzero_si256();
        for (int k = 0; k < end_k; ++k) {
            const float *t = tg + 3 * (size_t)(k2 + k);
            __m256 dx = _mm256_sub_ps(_mm256_broadcast_ss(t), QX);
            __m256 dy = _mm256_sub_ps(_mm256_broadcast_ss(t + 1), QY);
            __m256 dz = _mm256_sub_ps(_mm256_broadcast_ss(t + 2), QZ);
            __m256 d = _mm256_fmadd_ps(dz, dz, _mm256_fmadd_ps(dx, dx, _mm256_mul_ps(dy, dy)));
            __m256 lt = k == 0 ? _mm256_castsi256_ps(_mm256_set1_epi32(-1))
                               : _mm256_cmp_ps(d, best, _CMP_LT_OQ);
            best = _mm256_blendv_ps(best, d, lt);
            besti = _mm256_castps_si256(_mm256_blendv_ps(
                _mm256_castsi256_ps(besti), _mm256_castsi256_ps(_mm256_set1_epi32(k + k2)), lt));
        }
        /* chamfer3D.cu:126: result > best (false when either is NaN), first tile unconditional */
        __m256 gt = k2 == 0 ? _mm256_castsi256_ps(_mm256_set1_epi32(-1))
                            : _mm256_cmp_ps(res, best, _CMP_GT_OQ);
        res = _mm256_blendv_ps(res, best, gt);
        resi = _mm256_castps_si256(
            _mm256_blendv_ps(_mm256_castsi256_ps(resi), _mm256_castsi256_ps(besti), gt));
    }
    float r[8];
    int ri[8];
    _mm256_storeu_ps(r, res);
    _mm256_storeu_si256((__m256i *)ri, resi);
    for (int l = 0; l < nq; ++l) { result[l] = r[l]; result_i[l] = ri[l]; }
}
static int g_simd = -1; /* -1 = decide at first use; oracle_set_simd(0) forces the scalar loop */
void oracle_set_simd(int on) { g_simd = on ? -1 : 0; }
static int use_simd(void) {
    if (g_simd < 0) g_simd = __builtin_cpu_supports("avx2") && __builtin_cpu_supports("fma");
    return g_simd;
}
#else
void oracle_set_simd(int on) { (void)on; }
static int use_simd(void) { return 0; }
#endif

static void *nn_worker(void *arg) {
    nn_job *J = (nn_job *)arg;
    int n = J->n, m = J->m;
    int64_t w = J->w0;
#if defined(__x86_64__)
    if (use_simd() && m > 0) {
        while (w < J->w1) {
            int i = (int)(w / n);
            int64_t row_end = (int64_t)(i + 1) * n; /* stay inside one batch element */
            int64_t lim = row_end < J->w1 ? row_end : J->w1;
            int nq = (int)(lim - w < 8 ? lim - w : 8);
            nn_rows8(J->xyz + 3 * w, nq, J->xyz2 + (size_t)i * m * 3, m, J->result + w,
                     J->result_i + w);
            w += nq;
        }
        return NULL;
    }
#endif
    for (; w < J->w1; ++w) {
        int i = (int)(w / n);
        const float *q = J->xyz + 3 * w;
        const float *tg = J->xyz2 + (size_t)i * m * 3;
        for (int k2 = 0; k2 < m; k2 += ORACLE_TILE) {
            int end_k = (m < k2 + ORACLE_TILE ? m : k2 + ORACLE_TILE) - k2;
            float best = 0.f;
            int best_i = 0;
            for (int k = 0; k < end_k; ++k) {
                float d = sqdist(q[0], q[1], q[2], tg + 3 * (size_t)(k2 + k));
                if (k == 0 || d < best) { best = d; best_i = k + k2; }
            }
            if (k2 == 0 || J->result[w] > best) { J->result[w] = best; J->result_i[w] = best_i; }
        }
    }
    return NULL;
}

static int g_threads = 1;
/* number of host threads the directed search uses (queries are independent) */
void oracle_set_threads(int t) { g_threads = t < 1 ? 1 : (t > 256 ? 256 : t); }
int oracle_get_threads(void) { return g_threads; }

void oracle_nn_forward(int b, int n, const float *xyz, int m, const float *xyz2,
                       float *result, int *result_i) {
    int64_t total = (int64_t)b * n;
    int T = g_threads;
    if (total < T) T = total > 0 ? (int)total : 1;
    nn_job jobs[256];
    pthread_t th[256];
    for (int t = 0; t < T; ++t) {
        nn_job J = {n, m, xyz, xyz2, result, result_i, total * t / T, total * (t + 1) / T};
        jobs[t] = J;
    }
    for (int t = 1; t < T; ++t) pthread_create(&th[t], NULL, nn_worker, &jobs[t]);
    nn_worker(&jobs[0]);
    for (int t = 1; t < T; ++t) pthread_join(th[t], NULL);
}

/* chamfer_cuda_forward, chamfer3D.cu:136-154 (two directed launches). */
void oracle_chamfer_forward(int b, int n, const float *xyz1, int m, const float *xyz2,
                            float *dist1, int *idx1, float *dist2, int *idx2) {
    oracle_nn_forward(b, n, xyz1, m, xyz2, dist1, idx1);
    oracle_nn_forward(b, m, xyz2, n, xyz1, dist2, idx2);
}

/* NmDistanceGradKernel, chamfer3D.cu:155-174, float32, accumulated in plain
 * index order (what a single-thread launch of the reference does).  Buffers
 * must arrive zero-filled (dist_chamfer_3D.py:56-60). */
static void grad_dir_f32(int b, int n, const float *xyz1, int m, const float *xyz2,
                         const float *gd1, const int *idx1, float *g1, float *g2) {
    for (int i = 0; i < b; ++i)
        for (int j = 0; j < n; ++j) {
            size_t a = ((size_t)i * n + j) * 3;
            int j2 = idx1[(size_t)i * n + j];
            size_t c = ((size_t)i * m + j2) * 3;
            float g = gd1[(size_t)i * n + j] * 2;
            for (int k = 0; k < 3; ++k) {
                float v = g * (xyz1[a + k] - xyz2[c + k]);
                g1[a + k] += v;
                g2[c + k] += -v;
            }
        }
}

void oracle_chamfer_backward_f32(int b, int n, const float *xyz1, int m, const float *xyz2,
                                 const float *gd1, const int *idx1, const float *gd2,
                                 const int *idx2, float *gxyz1, float *gxyz2) {
    grad_dir_f32(b, n, xyz1, m, xyz2, gd1, idx1, gxyz1, gxyz2); /* :184 */
    grad_dir_f32(b, m, xyz2, n, xyz1, gd2, idx2, gxyz2, gxyz1); /* :185 */
}

/* Same sums in float64: the order-independent yardstick for the CUDA scatter,
 * whose float atomics make the reference itself order-dependent. */
static void grad_dir_f64(int b, int n, const float *xyz1, int m, const float *xyz2,
                         const float *gd1, const int *idx1, double *g1, double *g2) {
    for (int i = 0; i < b; ++i)
        for (int j = 0; j < n; ++j) {
            size_t a = ((size_t)i * n + j) * 3;
            int j2 = idx1[(size_t)i * n + j];
            size_t c = ((size_t)i * m + j2) * 3;
            double g = (double)gd1[(size_t)i * n + j] * 2.0;
            for (int k = 0; k < 3; ++k) {
                double v = g * ((double)xyz1[a + k] - (double)xyz2[c + k]);
                g1[a + k] += v;
                g2[c + k] -= v;
            }
        }
}

void oracle_chamfer_backward_f64(int b, int n, const float *xyz1, int m, const float *xyz2,
                                 const float *gd1, const int *idx1, const float *gd2,
                                 const int *idx2, double *gxyz1, double *gxyz2) {
    grad_dir_f64(b, n, xyz1, m, xyz2, gd1, idx1, gxyz1, gxyz2);
    grad_dir_f64(b, m, xyz2, n, xyz1, gd2, idx2, gxyz2, gxyz1);
}
