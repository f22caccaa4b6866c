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

static void *nn_worker(void *arg) {
    nn_job *J = (nn_job *)arg;
    int n = J->n, m = J->m;
    for (int64_t w = J->w0; w < J->w1; ++w) {
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
