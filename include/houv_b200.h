/* houv_b200 -- C ABI of the B200-native (sm_100a) HOUV registration hot path.
 *
 * This is the drop-in boundary: plain pointers and sizes, no torch types.
 * Every pointer is a DEVICE pointer unless its name ends in _host.  All
 * buffers are caller-owned; the library never allocates or frees behind the
 * caller's back and keeps no global state besides the tuning knobs.
 * `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 * every call only ENQUEUES work on it and is CUDA-graph capturable.
 * Return value: 0 = ok, otherwise a cudaError_t (or HOUV_E_* below) --
 * the reference's pybind entry points return 1/0 and printf
 * (chamfer3D.cu:145-151, 187-193); the Python binding maps our codes to
 * RuntimeError.
 *
 * Reference interfaces replaced (paths relative to /root/reference):
 *   houv_chamfer3d_forward   <- chamfer_3D.forward  utils/metrics/CD/chamfer3D/chamfer_cuda.cpp:17-19,31
 *                               = chamfer_cuda_forward            chamfer3D.cu:136-154
 *   houv_chamfer3d_backward  <- chamfer_3D.backward chamfer_cuda.cpp:22-27,32
 *                               = chamfer_cuda_backward           chamfer3D.cu:176-195
 *   houv_chamfer3d_backward_assign <- chamfer_3DFunction.backward's zero-fill + backward  dist_chamfer_3D.py:56-63
 *   houv_nn_search           <- one NmDistanceKernel launch       chamfer3D.cu:12-134,142
 *   houv_cd_reduce           <- calc_cd / fscore reductions      registration/model_utils_completion.py:69-80
 *   houv_kabsch              <- SVDHead.forward     registration/model_utils.py:220-255
 *   houv_idam_step           <- IDAM refinement iteration tail     registration/models/idam.py:317-342
 *   houv_emd_forward/backward <- emd.forward / emd.backward          utils/metrics/EMD/emd.cpp:14-29, emd_cuda.cu:230-316
 *   houv_icp                 <- o3d registration_icp call site    registration/train_ICP.py:148-151
 *   houv_pose_optimise       <- predict_model hot loop            registration/models/houv.py:106-138
 */
#ifndef HOUV_B200_H
#define HOUV_B200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HOUV_E_BADARG 100001 /* null pointer / negative size */
#define HOUV_E_TOOBIG 100002 /* B*N*3 or B*M*3 would overflow the index math */
#define HOUV_E_SCRATCH 100003 /* houv_chamfer3d_forward_static: the problem needs more scratch than
                                 HOUV_STATIC_SCRATCH_BYTES (use houv_chamfer3d_forward_ws) */
#define HOUV_STATIC_SCRATCH_BYTES (8u << 20) /* size of the library's built-in scratch array */

/* Library identification, e.g. "houv_b200 0.1 sm_100a". */
const char *houv_version(void);

/* Bidirectional Chamfer NN search.
 *   xyz1[B,N,3], xyz2[B,M,3] float32 row-major (AoS xyz)
 *   dist1[B,N] = min_j |xyz1_i - xyz2_j|^2, idx1[B,N] = FIRST arg-min j (int32)
 *   dist2[B,M], idx2[B,M] likewise with the roles swapped.
 * Arithmetic is the reference's, bit for bit: d = fma(dz,dz,fma(dx,dx,dy*dy)),
 * dx = target.x - query.x, strict '<' in ascending j  (chamfer3D.cu:32-36,126).
 * Outputs need not be pre-zeroed.  M == 0 (or N == 0) zero-fills the affected
 * outputs, which is what the reference's zero-initialised buffers end up as.
 * FINITE INPUTS ONLY.  Every finite float32 is handled bit for bit like the reference,
 * denormals and coordinates whose differences overflow to +inf included
 * (tests/test_chamfer_gpu.py::test_extreme_finite_inputs).  NaN or +-inf COORDINATES are
 * unsupported input: the reference's outcome for them depends on where the NaN sits inside its
 * 512-point tiles (a NaN at tile position 0 blocks that tile, and the first tile for good:
 * chamfer3D.cu:36,79,121,126 -- restated and pinned in oracle/chamfer_oracle.c,
 * tests/test_oracle_cpu.py::test_nan_semantics_follow_the_reference_kernel); these kernels skip
 * NaN distances in every minimum instead and do not reproduce that. */
int houv_chamfer3d_forward(const float *xyz1, const float *xyz2, int B, int N, int M,
                           float *dist1, float *dist2, int *idx1, int *idx2, void *stream);

/* The same with a caller-provided scratch buffer.  With it the library may pick the symmetric
 * kernel (csrc/chamfer_sym.cu: every distance evaluated once for both directions) also when the
 * larger cloud has more than 2048 points; results are identical bit for bit either way.
 * workspace: device memory of at least houv_chamfer3d_workspace_bytes(B,N,M) bytes, 8-byte
 * aligned (may be NULL when that is 0); its contents are scratch. */
size_t houv_chamfer3d_workspace_bytes(int B, int N, int M);
int houv_chamfer3d_forward_ws(const float *xyz1, const float *xyz2, int B, int N, int M,
                              float *dist1, float *dist2, int *idx1, int *idx2,
                              void *workspace, size_t workspace_bytes, void *stream);

/* The same call for callers that cannot allocate (CUDA graph capture, serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * The scratch lives in a static device array of HOUV_STATIC_SCRATCH_BYTES (8 MiB)
 * inside the library -- every GPU that loads the library holds it until the process exits -- and
 * the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).  Inputs
 * and outputs are the caller's, as above; results are identical bit for bit to
 * houv_chamfer3d_forward_ws.  A problem whose scratch (houv_chamfer3d_workspace_bytes) exceeds
 * the array is rejected with HOUV_E_SCRATCH before anything is launched: e.g. 32 x 2048 x 2048
 * needs 4 MiB, 64 x 16384 x 2048 1 MiB, 200 x 8192 x 8192 12.5 MiB (rejected).  All calls share
 * the one array (it keeps the last call's contents; the library re-initialises what it must):
 * calls that overlap on the GPU -- two streams, two host threads on one device -- RACE on it;
 * serialise them or use houv_chamfer3d_forward_ws with a workspace per stream. */
int houv_chamfer3d_forward_static(const float *xyz1, const float *xyz2, int B, int N, int M,
                                  float *dist1, float *dist2, int *idx1, int *idx2, void *stream);
size_t houv_chamfer3d_static_scratch_bytes(void);

/* One directed search: query[B,N,3] against target[B,M,3] -> dist[B,N], idx[B,N]. */
int houv_nn_search(const float *query, const float *target, int B, int N, int M,
                   float *dist, int *idx, void *stream);

/* Gradient scatter of both directions, fused in one launch.
 *   g = 2*graddist1[b,i];  gradxyz1[b,i]      += g*(xyz1[b,i]-xyz2[b,idx1[b,i]])
 *                          gradxyz2[b,idx1..] -= the same            (chamfer3D.cu:155-174)
 *   and symmetrically for graddist2/idx2.
 * gradxyz1/gradxyz2 are ACCUMULATED INTO (float atomics, like the reference) and
 * must arrive zero-filled for a plain backward (dist_chamfer_3D.py:56-60). */
int houv_chamfer3d_backward(const float *xyz1, const float *xyz2, const float *graddist1,
                            const float *graddist2, const int *idx1, const int *idx2, int B,
                            int N, int M, float *gradxyz1, float *gradxyz2, void *stream);

/* The same gradient WRITTEN to gradxyz1/gradxyz2 (buffers may arrive uninitialised): what
 * chamfer_3DFunction.backward computes after its own zero-fill (dist_chamfer_3D.py:56-63), without the
 * separate fill pass.  Large batches of clouds that fit one SM's shared memory (24*(N+M) bytes <= 200 KB)
 * are stored directly; other shapes are zero-filled on `stream` and accumulated as above. */
int houv_chamfer3d_backward_assign(const float *xyz1, const float *xyz2, const float *graddist1,
                                   const float *graddist2, const int *idx1, const int *idx2, int B,
                                   int N, int M, float *gradxyz1, float *gradxyz2, void *stream);

/* The same gradient, DETERMINISTIC: every row is summed by one thread in the order a single-thread
 * run of the reference kernels adds it (direction 1 then direction 2, ascending point index), so
 * the result is identical run to run and bit-identical to the reference kernel bodies executed
 * sequentially (oracle/chamfer_oracle.c) -- the reference's float atomics are order-dependent
 * (chamfer3D.cu:166-171, registration/README.md:82-90).  assign != 0 writes the buffers, 0 adds
 * into them.  Clouds up to 2048 points (HOUV_E_TOOBIG above); ~1.4x the time of the default
 * (TMA-pipelined when the arrays are 16-byte aligned and N, M are multiples of 4).
 * The Python layer calls this under torch.use_deterministic_algorithms(True). */
int houv_chamfer3d_backward_deterministic(const float *xyz1, const float *xyz2,
                                          const float *graddist1, const float *graddist2,
                                          const int *idx1, const int *idx2, int B, int N, int M,
                                          float *gradxyz1, float *gradxyz2, int assign,
                                          void *stream);

/* Epilogue of calc_cd + fscore (model_utils_completion.py:69-80, utils/metrics/CD/fscore.py:3-16)
 * on the outputs of a Chamfer forward: cd_p[B] = (mean sqrt dist1 + mean sqrt dist2)/2,
 * cd_t[B] = mean dist1 + mean dist2, f1[B] (or NULL) = F-score at `threshold` (1e-4 in the
 * reference), 0 where it is NaN.  Forward only. */
int houv_cd_reduce(const float *dist1, const float *dist2, int B, int N, int M, float threshold,
                   float *cd_p, float *cd_t, float *f1, void *stream);

/* Fused multi-start pose optimisation (K3): `iters` iterations of
 *   HOUV.forward -> Predict_loss -> backward -> Adam.step      (houv.py:120-126)
 * for rows = P*K independent starts; row r optimises pair r / K (the K starts of a
 * pair are consecutive rows, as predict_model lays them out, houv.py:111-114).
 *   src[P,N,3], tgt[P,M,3]   the pairs (N == M <= 2048)
 *   params[rows,8]           in: initial V_c[3], angle_c, tran_c[3], tran_s (reset_weight,
 *                            houv.py:40-61, drawn on the host); out: values after `iters` steps
 *   adam_m, adam_v[rows,8]   in/out Adam moments, or NULL (= start from zero, not returned)
 *   step0                    Adam steps already taken (0 for a fresh run; allows resume)
 *   angle_base               0..3: 45-degree window offset of the rotation angle (houv.py:96)
 *   lr, beta1, beta2, eps    torch.optim.Adam hyper-parameters (reference: 0.01, .9, .999, 1e-8)
 *   grad_scale               d(mean over the batch)/d(row) = 1/rows_in_batch (houv.py:124)
 *   percent, w3d, wview      trimmed fraction of the 3-D term (0.5) and the 6 / 1 weights (houv.py:222)
 *   min1_out[rows], R_out[rows,9], T_out[rows,3]  of the LAST forward, i.e. after iters-1
 *                            steps, exactly what predict_model returns (houv.py:134-138)
 *   loss_out[rows] or NULL   full loss 6*min_1 + views of the last forward
 *   trace[iters,rows,2] or NULL  (loss, min_1) of every forward
 *   row_pair[rows] or NULL   pair optimised by each row (default row / K); with it `rows` may
 *                            differ from P*K, e.g. the three retry passes of solve_model
 *                            (houv.py:168-169) in one launch
 *   row_base[rows] or NULL   per-row angle_base overriding the scalar
 *   rows                     number of rows; 0 = P*K */
int houv_pose_optimise(const float *src, const float *tgt, int P, int N, int M, int K,
                       float *params, float *adam_m, float *adam_v, int step0, int iters,
                       int angle_base, float lr, float beta1, float beta2, float eps,
                       float grad_scale, float percent, float w3d, float wview,
                       float *min1_out, float *loss_out, float *R_out, float *T_out,
                       float *trace, const int *row_pair, const int *row_base, int rows,
                       void *stream);

/* The same kernel with every knob in a struct (zero-initialise, then fill).  Adds what the older
 * functional solver of the reference needs -- train_utils.getPredict_angle / solve
 * (registration/train_utils.py:359-572), the path behind test.py:64: float64 parameters and
 * Adam state (params64 != NULL: the forward sees their float32 roundings, :397-401),
 * |T| = sin(tran_s*pi)*tran_scale + tran_offset (1 and 0 at :405; 0.125/0.125 in houv.py:99),
 * double Adam hyper-parameters (lr 0.1 at :390), wview = 0 (loss 6*min_1 only, :439: the view
 * terms then take no part in the gradient), and tran_s_out[rows] = |T| of the last forward
 * (returned by getPredict_angle, :456). */
typedef struct HouvPoseProblem {
    const float *src, *tgt;          /* [P,N,3], [P,M,3] */
    int P, N, M, K;
    float *params, *adam_m, *adam_v; /* [rows,8] float32 state (HOUV); params may be NULL with params64 */
    double *params64, *adam_m64, *adam_v64; /* [rows,8] float64 state, or NULL */
    int step0, iters, angle_base;
    double lr, beta1, beta2, eps;
    float grad_scale, percent, w3d, wview;
    float tran_scale, tran_offset;
    float *min1_out, *loss_out, *R_out, *T_out, *tran_s_out, *trace;
    const int *row_pair, *row_base;
    int rows;                        /* 0 = P*K */
    const float *row_scale;          /* [rows] per-row grad_scale (rows of several batches in one
                                        launch, each batch with its own 1/rows_in_batch), or NULL */
} HouvPoseProblem;
int houv_pose_optimise_ex(const HouvPoseProblem *problem, void *stream);

/* Batched Kabsch (K4), SVDHead.forward semantics (model_utils.py:220-255).
 *   src[B,3,N], src_corr[B,3,N]  channel-first clouds, corresponding columns
 *   weights[B,N] or NULL         H = (src_c * w) corr_c^T; means stay unweighted (:221-227),
 *                                t uses the weighted sums (:254)
 *   R[B,9] row-major, t[B,3]     R = V U^T of H = U S V^T, last column of V negated when
 *                                det(V U^T) < 0 (:236-239); t = -R mean(src) + mean(corr) */
int houv_kabsch(const float *src, const float *src_corr, const float *weights, int B, int N,
                float *R, float *t, void *stream);

/* One refinement step of IDAM's iterative weighted Kabsch, fused (K6; registration/models/idam.py:317-342):
 *   src[B,3,N], tgt[B,3,M] channel-first; corr_idx[B,N] (int32, or int64 when corr_is_int64 != 0) indexes
 *   tgt columns and must lie in [0,M); weight_logits[B,N] are the soft-elimination scores BEFORE the sigmoid.
 *   w = sigmoid(logits); w *= (w >= median(w)) [lower median, as torch.median]; w /= sum(w) + 1e-8   (:329-331)
 *   R_ab, t_ab = SVDHead(src, tgt[:, :, corr_idx], w)                              (:318, :335, idam.py:157-189)
 *   src_out = R_ab src + t_ab;  R_acc <- R_ab R_acc;  t_acc <- R_ab t_acc + t_ab   (:338-340)
 * src_out[B,3,N], R_acc[B,9]/t_acc[B,3] (updated in place, both or neither), R_ab[B,9], t_ab[B,3] and
 * weights_out[B,N] may each be NULL.  N <= 8192 (HOUV_E_TOOBIG above). */
int houv_idam_step(const float *src, const float *tgt, const void *corr_idx, int corr_is_int64,
                   const float *weight_logits, int B, int N, int M, float *src_out, float *R_acc,
                   float *t_acc, float *R_ab, float *t_ab, float *weights_out, void *stream);

/* EMD approximation by the auction algorithm (K7; utils/metrics/EMD/emd_cuda.cu:230-268, emd_module.py:40-75):
 *   xyz1[B,n,3] (prediction), xyz2[B,n,3] (ground truth), coordinates normalised to [0,1]; eps > 0 trades
 *   accuracy for convergence, iters = iteration cap (the last one assigns every remaining point to its bid).
 *   dist[B,n] = squared distance of xyz1[b,j] to xyz2[b,assignment[b,j]]; assignment[B,n] int32 (not
 *   guaranteed to be a bijection, like the reference); iters_used[B] or NULL = iterations that still had
 *   unassigned points.  All iterations run in one launch, one CTA per batch element.  Any n >= 1 and any B
 *   (the reference needs n % 1024 == 0 and B <= 512); n > 4096 needs houv_emd_workspace_bytes(B,n) bytes of
 *   device scratch.  Ties the reference leaves to thread scheduling are resolved towards the lowest index. */
size_t houv_emd_workspace_bytes(int B, int n);
int houv_emd_forward(const float *xyz1, const float *xyz2, int B, int n, float eps, int iters, float *dist,
                     int *assignment, int *iters_used, void *workspace, size_t workspace_bytes, void *stream);
/* gradxyz1[b,j] += 2*graddist[b,j]*(xyz1[b,j] - xyz2[b,assignment[b,j]]); xyz2 gets no gradient
 * (emd_cuda.cu:270-305, emd_module.py:77-86). */
int houv_emd_backward(const float *xyz1, const float *xyz2, const float *graddist, const int *assignment,
                      int B, int n, float *gradxyz1, void *stream);

/* Point-to-point ICP, one CTA per pair (train_ICP.py:148-151 call site; Open3D 0.9.0
 * algorithm restated, see csrc/icp.cu).
 *   src[P,N,3] -> tgt[P,M,3] (N, M <= 2048); init: one row-major 4x4 (init_per_pair = 0) or
 *   [P,16]; max_distance = correspondence radius (0.02), max_iteration (500),
 *   relative_fitness / relative_rmse = convergence criteria (Open3D defaults 1e-6).
 *   T_out[P,16] row-major 4x4 (bottom row 0 0 0 1); stats[P,4] = fitness, inlier_rmse,
 *   iterations run, number of correspondences of the final evaluation. */
int houv_icp(const float *src, const float *tgt, int P, int N, int M, const float *init,
             int init_per_pair, float max_distance, int max_iteration, double relative_fitness,
             double relative_rmse, float *T_out, float *stats, void *stream);

/* Tuning knob for the NN kernel: variant 0..3 = two directed scans with (queries/thread,
 * threads/CTA) (2,128) (4,128) (4,64) (2,256); 10 = the symmetric kernel whenever its workspace
 * requirement is met; -1 = automatic (symmetric when it fills the GPU).  Returns the old value. */
int houv_set_nn_variant(int variant);

/* Tuning knob for K3: 0..5 = the two-directed-scans kernel (csrc/pose_opt.cu) with (queries/
 * thread, 8-target halves per arg-min group, unroll) (4,1,1) (4,1,2) (4,2,1) (2,1,2) (2,2,2)
 * (4,2,2), identical results; 10..14 = the symmetric-scan kernel (csrc/pose_sym.cu) with
 * different unrolls, identical results among themselves and equal to 0..5 up to the summation
 * order of the loss terms; 15 = 12 with the four-metric sweep forced when wview == 0; 16 = 12
 * with the (distance, index) arg-min recovery everywhere instead of the equality search (both
 * bit-identical to 12); -1 = default (13, or $HOUV_POSE_VARIANT).  Returns the old value. */
int houv_set_pose_variant(int variant);

/* FP32 CUDA-core peak probe: runs `iters` x 64 FFMA (or FFMA2 when packed) per thread on a
 * full-chip grid; *flop_host receives the FLOP it executed (2 per FFMA lane-op).
 * Time it with events around the call to get the roofline denominator. */
int houv_fp32_probe(float *scratch /* >= 4 floats */, int iters, int packed,
                    double *flop_host, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* HOUV_B200_H */
