/* libpe_rotation.so - C ABI of the B200-native rotation-mapping path.
 *
 * The reference (henrikgruner/PoseEstimation) has no FFI: its boundary is the Python function names of
 * rotation_representation.py, loss.py and point_cloud/main.py.  Each entry point below is what a binding for one
 * of those functions calls; the comment above it names the reference function it replaces (paths relative to
 * /root/reference).  INTEGRATION.md shows the ctypes stub for each.
 *
 * Conventions (all entry points):
 *   - every pointer is a DEVICE pointer into memory the caller owns (PyTorch's allocator).  The library owns one piece of
 *     device memory: 128 KB per device of 32-byte launch slots (tile counter + two float64 accumulators), allocated and
 *     zeroed by pe_init(device).  An eager launch uses the slot of its (device, stream) - launches on one stream never
 *     overlap, so they can share it for ever; a launch recorded by a stream capture gets a slot of its own that no eager
 *     launch and no other captured launch ever uses, so graphs may replay next to eager work.  Call pe_init before
 *     capturing (an un-initialised device is initialised by its first eager launch; inside a capture that is refused
 *     with PE_ERR_BAD_ARG because it would allocate).  pe_svd_orth_frobenius_step_metrics also takes caller-owned
 *     scratch instead of a slot;
 *   - a "matrix" is 9 consecutive floats, row-major: m[3*i + j] = M[i][j]  (x.view(-1, 3, 3));
 *   - `n` is the number of records (matrices / vectors); n == 0 is a no-op that still zeroes accumulators;
 *   - `*_stride` is the distance in FLOATS between consecutive input records (9 for a dense (n, 9) array, 12 for
 *     the out[:, :9] view of a (n, 12) network output); outputs are always dense;
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it, nothing synchronises;
 *   - accumulators: `double* sum` and `int* flag` are WRITTEN by the call (previous content ignored; the partial sums
 *     and flag bits meet in the launch slot and the last thread block to finish stores the totals): a reducing call is
 *     ONE stream operation, no memset in front of it;
 *   - return value: 0 = success, > 0 = a cudaError_t from the launch, < 0 = PE_ERR_* below;
 *     pe_last_error() gives a message for the calling thread's last non-zero return;
 *   - there is no CPU path: without an sm_100 device every compute entry point fails.
 */
#ifndef PE_ROTATION_H
#define PE_ROTATION_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PE_VERSION 200 /* 0.2.0 */

#define PE_ERR_BAD_ARG (-1)   /* null pointer, negative count, stride smaller than the record, array not naturally
                                 aligned (float: 4 bytes, double: 8 bytes) */
#define PE_ERR_NO_DEVICE (-2) /* current device is not sm_100 */
#define PE_ERR_INTERNAL (-3)

int pe_version(void);
const char* pe_last_error(void);
/* kernels launched by this process through the library so far (bench.py's gpu_launches) */
long long pe_launch_count(void);
/* Allocate and zero the launch slots of `device` (a CUDA device ordinal; < 0 = the current device).  Idempotent,
 * thread-safe, synchronises the device once.  Nothing else in the library allocates. */
int pe_init(int device);

/* ---- a1  symmetric_orthogonalization(x)                 rotation_representation.py:192-206 ------------------
 * r[n][9] = U diag(1, 1, det(U V^T)) V^T with m = U S V^T.  det_sign (nullable, n floats) receives the factor
 * det(U V^T) = +-1 of line 202. */
int pe_svd_orth_fwd(const float* m, int64_t m_stride, float* r, float* det_sign, int64_t n, void* stream);
/* autograd through lines 199-205 (torch's svd_backward, det_backward, ...), as one closed-form pullback:
 * dm = dL/dm given g = dL/dr.  `r` is the forward output; pass NULL to have it recomputed from m. */
int pe_svd_orth_bwd(const float* m, int64_t m_stride, const float* r, const float* g, float* dm, int64_t n,
                    void* stream);

/* ---- a2  compute_rotation_matrix_from_ortho6d(poses)    rotation_representation.py:21-36 (= 174-189) -------- */
int pe_ortho6d_fwd(const float* p, int64_t p_stride, float* r, int64_t n, void* stream);
int pe_ortho6d_bwd(const float* p, int64_t p_stride, const float* g, float* dp, int64_t n, void* stream);

/* ---- a3  compute_rotation_matrix_from_quaternion(q)     rotation_representation.py:139-172 (+ 38-50) -------- */
int pe_quat_fwd(const float* q, int64_t q_stride, float* r, int64_t n, void* stream);
int pe_quat_bwd(const float* q, int64_t q_stride, const float* g, float* dq, int64_t n, void* stream);

/* ---- a4  compute_rotation_matrix_from_euler(euler)      rotation_representation.py:92-115 -------------------- */
int pe_euler_fwd(const float* e, int64_t e_stride, float* r, int64_t n, void* stream);
int pe_euler_bwd(const float* e, int64_t e_stride, const float* g, float* de, int64_t n, void* stream);

/* ---- 8(f).1  compute_rotation_matrix_from_ortho5d(a)    rotation_representation.py:72-90, 120-136 ------------ */
int pe_ortho5d_fwd(const float* a, int64_t a_stride, float* r, int64_t n, void* stream);
int pe_ortho5d_bwd(const float* a, int64_t a_stride, const float* g, float* da, int64_t n, void* stream);

/* ---- 8(f).1  vec_3d_to_SO3(x) = so3_exp_map(x)          rotation_representation.py:245-320 ------------------- */
int pe_expmap_fwd(const float* v, int64_t v_stride, float* r, int64_t n, void* stream);
int pe_expmap_bwd(const float* v, int64_t v_stride, const float* g, float* dv, int64_t n, void* stream);
/* so3_exp_map(log_rot, eps) with the caller's clamp bound on |v|^2 (:245-272; the two above use the default 1e-4). */
int pe_expmap_eps_fwd(const float* v, int64_t v_stride, float eps, float* r, int64_t n, void* stream);
int pe_expmap_eps_bwd(const float* v, int64_t v_stride, const float* g, float eps, float* dv, int64_t n, void* stream);

/* ---- a5  loss_frobenius(R_pred, R_true)                 Comparison/main.py:12-16 (+ 7 copies) ----------------
 * *sum = sum_b |a_b - b_b|_F (the caller divides by n: the mean of line 16; across ranks the sums all-reduce). */
int pe_frobenius_fwd(const float* a, const float* b, double* sum, int64_t n, void* stream);
/* the same, and *mean = (float)(*sum / n) written by the same launch: the 0-d float32 tensor loss_frobenius returns
 * (line 16), without a second launch that divides and casts.  n == 0: *sum = 0, *mean = NaN (torch's empty mean). */
int pe_frobenius_mean_fwd(const float* a, const float* b, double* sum, float* mean, int64_t n, void* stream);
/* da = gscale * (a - b)/|a - b|_F, db = -da, with gscale = (*gout) * inv_n; 0 where the norm is 0.
 * gout is a device scalar (the upstream gradient of the 0-d loss); da or db may be NULL. */
int pe_frobenius_bwd(const float* a, const float* b, const float* gout, float inv_n, float* da, float* db,
                     int64_t n, void* stream);

/* ---- pe_frobenius_fwd without any allocation ---------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  Same inputs, output and
 * return convention as pe_frobenius_fwd; *sum agrees with it to float64 round-off (same float32 terms, other order) and
 * is bitwise reproducible from call to call.  Its scratch - one float64 partial per 256 records and a ticket word - is a
 * file-scope static __device__ array of 256 KB + 256 B that every GPU loading the library holds until the process exits;
 * that is the limit: n <= PE_STATIC_MAX_RECORDS, larger n returns PE_ERR_BAD_ARG.  The call makes no cudaMalloc /
 * cudaFree / allocating library call, and does not need pe_init.  ALL calls share the one array: calls that overlap on
 * the GPU (two streams, two host threads) race on it - serialise them.  The first call on a device resolves the array's
 * device address (cudaGetSymbolAddress); make one call outside a stream capture before capturing it. */
#define PE_STATIC_MAX_RECORDS 8388608 /* 2^23 matrices = 32768 tiles of 256 */
int pe_frobenius_fwd_static(const float* a, const float* b, double* sum, int64_t n, void* stream);

/* ---- a6  compute_geodesic_distance_from_two_matrices    rotation_representation.py:209-227  (lo, hi = -1, 1)
 * ---- a8  geodesic(R1, R2, reduction)                    point_cloud/main.py:61-73   (lo, hi = -1+1e-7, 1-1e-7)
 * theta_b = acos(clamp((tr(m1_b m2_b^T) - 1)/2, lo, hi)) in radians.  theta (n floats) and sum are each
 * nullable; sum receives sum_b theta_b. */
int pe_geodesic_fwd(const float* m1, const float* m2, float lo, float hi, float* theta, double* sum, int64_t n,
                    void* stream);
/* geodesic(R1, R2, reduction='mean' | 'sum') in ONE launch (point_cloud/main.py:70-73): sum as above and
 * *out = (float)(sum * scale), scale = 1/n for the mean, 1 for the sum.  n == 0 stores NaN (the mean of an empty
 * batch, as torch's); the caller owns the empty 'sum'. */
int pe_geodesic_reduce_fwd(const float* m1, const float* m2, float lo, float hi, double* sum, float* out,
                           double scale, int64_t n, void* stream);
/* g: n upstream gradients (g_is_scalar = 0) or one device scalar broadcast to all records, times gmul
 * (reduction='mean': g_is_scalar = 1, gmul = 1/n).  d1 or d2 may be NULL.  Gradient is 0 where the clamp is
 * active, as torch.clamp's is. */
int pe_geodesic_bwd(const float* m1, const float* m2, float lo, float hi, const float* g, int g_is_scalar,
                    float gmul, float* d1, float* d2, int64_t n, void* stream);

/* ---- a7  angle_error(R1, R2)                            rotation_representation.py:230-242 -------------------
 * float64 arithmetic on float32 inputs: deg_b = acos(clamp((tr(r1_b^T r2_b) - 1)/2, -1, 1)) * 180/pi.
 * deg (n doubles) and sum are each nullable.  *flag becomes non-zero if any cosine lies outside [-1.1, 1.1]:
 * the binding raises ValueError (line 238). */
int pe_angle_error_fwd(const float* r1, const float* r2, double* deg, double* sum, int* flag, int64_t n,
                       void* stream);

/* ---- fused training step: loss_frobenius(symmetric_orthogonalization(x), R_true) -------------------------------
 * Comparison/main.py:59-62, UPNA/main.py:54-58, 3D-Pose/main.py:60,85-86.
 * fwd: r (nullable) as pe_svd_orth_fwd, *sum = sum_b |r_true_b - r_b|_F.
 * bwd: dm = pullback of gscale * (r - r_true)/|r - r_true|_F through the orthogonalisation, recomputing r from
 *      m, with gscale = (*gout) * inv_n. */
int pe_svd_orth_frobenius_fwd(const float* m, int64_t m_stride, const float* r_true, float* r, double* sum,
                              int64_t n, void* stream);
int pe_svd_orth_frobenius_bwd(const float* m, int64_t m_stride, const float* r_true, const float* gout,
                              float inv_n, float* dm, int64_t n, void* stream);

/* single pass of the same step: ONE decomposition per matrix yields the loss and the gradient (108 B/matrix).
 * *sum = sum_b |r_true_b - r_b|_F and dm = inv_n * pullback((r - r_true)/|r - r_true|_F), i.e. the gradient of the
 * batch-mean loss for an upstream gradient of 1.  The autograd backward is then pe_scale_by_scalar(dm, gout). */
int pe_svd_orth_frobenius_step(const float* m, int64_t m_stride, const float* r_true, float inv_n, float* dm,
                               double* sum, int64_t n, void* stream);
/* the same, and *mean = (float)(*sum / n): the loss tensor itself, written by the launch's last CTA. */
int pe_svd_orth_frobenius_step_mean(const float* m, int64_t m_stride, const float* r_true, float inv_n, float* dm,
                                    double* sum, float* mean, int64_t n, void* stream);
/* The training step with its metric (Comparison/main.py:59-70, 3D-Pose/main.py:60-62,85-91: loss_frobenius(outR, R)
 * and angle_error(outR, R) from the same outR):  as pe_svd_orth_frobenius_step, and additionally
 *   sums[0] = sum_b |r_true_b - r_b|_F,  sums[1] = sum_b angle_error(r_b, r_true_b) in degrees (float64 trace and acos
 *   of the fp32 r the step would have written, rotation_representation.py:230-242),
 *   sums[2] = 1.0 if a cosine left [-1.1, 1.1] (:237; the binding raises ValueError), else 0.0; *flag (nullable) the
 *   same as an int.  sums[0..2] is the 24-byte vector the ranks all-reduce.
 *   *loss_mean (nullable) = (float)(sums[0] / n),  *angle_mean (nullable) = sums[1] / n.
 * scratch: NULL, or 32 bytes of caller-owned device memory, 8-byte aligned, zero before its first use (the launch
 * leaves it zero again): then the launch touches no library-owned memory at all - one scratch per launch that can be
 * in flight at a time (e.g. per captured graph node). */
int pe_svd_orth_frobenius_step_metrics(const float* m, int64_t m_stride, const float* r_true, float inv_n, float* dm,
                                       double* sums, float* loss_mean, double* angle_mean, int* flag, void* scratch,
                                       int64_t n, void* stream);
/* y[i] = x[i] * (*g) for nwords floats; g is a DEVICE scalar (no host sync); x, y 16-byte aligned, y may alias x. */
int pe_scale_by_scalar(const float* x, const float* g, float* y, int64_t nwords, void* stream);
/* x[i] *= (*g) in place; does nothing when *g == 1 (tested on the device, no host sync): the backward of the
 * single-pass step under `loss.backward()`, whose upstream gradient is exactly 1 at every call site of the reference. */
int pe_scale_by_scalar_inplace(float* x, const float* g, int64_t nwords, void* stream);

/* ---- fused map + metric: sum_b metric(map(x_b), r_true_b), the rotation is never written ---
 * What the evaluation loops of the reference compute in two steps - R = transform_output[rot_rep](out), then
 * angle_error(R, R_true).mean() (UPNA/main.py:93-96, 3D-Pose/main.py:116-120) or a loss of R - in one pass of
 * 4 W + 36 bytes per record.  `map` selects the representation (x has W = 9, 6, 4, 3, 5, 3 words per record, read with
 * x_stride), `metric` the distance: Frobenius norm of the difference, geodesic angle in radians (fp32 acos, the cosine
 * clamped to [lo, hi]: -1, 1 for a6, -+(1 - 1e-7) for a8), or angle_error in degrees (float64; *flag, required, gets
 * bit 0 if a cosine left [-1.1, 1.1]; lo, hi unused).  Divide *sum by n for the mean. */
#define PE_MAP_SVD 0
#define PE_MAP_ORTHO6D 1
#define PE_MAP_QUAT 2
#define PE_MAP_EULER 3
#define PE_MAP_ORTHO5D 4
#define PE_MAP_EXPMAP 5
#define PE_METRIC_FROBENIUS 0
#define PE_METRIC_GEODESIC 1
#define PE_METRIC_ANGLE_DEG 2
int pe_map_metric_fwd(int map, int metric, const float* x, int64_t x_stride, const float* r_true, float lo, float hi,
                      double* sum, int* flag, int64_t n, void* stream);

/* ---- a9  Kabsch alignment - no reference function (site: point_cloud/main.py:171-181) --------------------------
 * p, q: (b, npts, 3) dense.  r[b][9] = the rotation minimising sum_k |R p_k + t - q_k|^2, i.e.
 * pe_svd_orth of sum_k (q_k - qbar)(p_k - pbar)^T;  t (nullable, b x 3) = qbar - R pbar. */
int pe_kabsch_fwd(const float* p, const float* q, int64_t b, int64_t npts, float* r, float* t, void* stream);

/* as pe_kabsch_fwd, and additionally ctx[b][15] = (H row-major 9, pbar 3, qbar 3): what pe_kabsch_bwd_coeffs needs. */
int pe_kabsch_fwd_ctx(const float* p, const float* q, int64_t b, int64_t npts, float* r, float* t, float* ctx,
                      void* stream);
/* Kabsch backward, per-pair part.  g[b][12] = (dL/dR 9, dL/dt 3).  Writes the affine coefficients
 * (A row-major 9, offset 3) of the two point gradients:  dL/dq_k = A_q p_k + b_q,  dL/dp_k = A_p q_k + b_p,
 * which two pe_affine_points launches then apply (48 B per point pair in total). */
int pe_kabsch_bwd_coeffs(const float* ctx, const float* r, const float* g, int64_t npts, float* coef_q, float* coef_p,
                         int64_t b, void* stream);
/* Weighted Kabsch ("weighted / with-translation variants"; no reference function - the site is
 * point_cloud/main.py:171-197).  w: (b, npts) weights >= 0.  With W = sum_k w_k, pbar = sum w p / W, qbar likewise:
 * r = pe_svd_orth of sum_k w_k (q_k - qbar)(p_k - pbar)^T, t (nullable) = qbar - R pbar, and ctx (nullable)
 * [b][16] = (that matrix 9, pbar 3, qbar 3, W 1).  A pair whose weights are all zero gets the rotation of the zero
 * matrix, not NaN.  28 B per point pair. */
int pe_kabsch_weighted_fwd(const float* p, const float* q, const float* w, int64_t b, int64_t npts, float* r, float* t,
                           float* ctx, void* stream);
/* per-pair part of its backward: as pe_kabsch_bwd_coeffs with the 16-word context (1/N becomes 1/W). */
int pe_kabsch_weighted_bwd_coeffs(const float* ctx, const float* r, const float* g, float* coef_q, float* coef_p,
                                  int64_t b, void* stream);
/* point part of its backward, one pass over p, q, w:  dq_k = w_k (A_q p_k + b_q),  dp_k = w_k (A_p q_k + b_p),
 * dw_k = (q_k - qbar).(A_q p_k + b_q) - (R^T g_t / W).(p_k - pbar).  Each of dp, dq, dw is nullable (not all). */
int pe_kabsch_weighted_bwd_points(const float* p, const float* q, const float* w, const float* coef_q,
                                  const float* coef_p, const float* ctx, int64_t b, int64_t npts, float* dp, float* dq,
                                  float* dw, void* stream);
/* m[b][9] = sum_k g[b][k] p[b][k]^T (uncentred), gsum[b][3] = sum_k g[b][k] (nullable): the reduction behind the
 * gradient of pe_affine_points w.r.t. its matrix and offset. */
int pe_cloud_outer_sum(const float* g, const float* p, int64_t b, int64_t npts, float* m, float* gsum, void* stream);

/* ---- per-cloud affine map of xyz points ---------------------------------------------------------------------------
 * out[b][k] = A_b p[b][k] + t_b.  Serves pc2 = gt_rmat @ pc1 (point_cloud/main.py:173-181, without the (B*N,3,3)
 * expansion), transform_pts (Iterative/loss.py:51-69: a = T, a_stride 16, a_row 4, t = T + 3, t_stride 16, t_elem 4)
 * and the Kabsch backward.  a_row = floats between rows of A; transpose != 0 applies A^T; t nullable. */
int pe_affine_points(const float* a, int64_t a_stride, int a_row, int transpose, const float* t, int64_t t_stride,
                     int t_elem, const float* p, int64_t b, int64_t npts, float* out, void* stream);

/* ---- cloud-pair rotation sampler, deterministic part      point_cloud/prepare.py:21-49 ---------------------------
 * r_b = matrix of the quaternion (cos theta_b, normalize(axis_b) sin theta_b); |axis| floored at 1e-8 (:12-18).
 * theta (n), axis (n, 3) come from the caller's RNG (the reference: uniform(-pi, pi) and randn). */
int pe_axis_angle_fwd(const float* theta, const float* axis, float* r, int64_t n, void* stream);

/* ---- CosyPose pose update                                 Iterative/utility.py:90-128 ----------------------------
 * out[b][>=12] = network output (9 rotation entries, vx, vy, vz), t_init/t_pred = 4x4 row-major.
 * dR = pe_svd_orth(out[:9]); R' = dR R_k; z' = vz z_k; x' = (vx/fx + x_k/z_k) z'; y' likewise; last row (0,0,0,1).
 * bwd: d_out[b][12] from g_t[b][16] = dL/dT_pred (T_init is a detached input, Iterative/main.py:93). */
int pe_pose_update_fwd(const float* out, int64_t out_stride, const float* t_init, float fx, float fy, float* t_pred,
                       int64_t n, void* stream);
int pe_pose_update_bwd(const float* out, int64_t out_stride, const float* t_init, const float* g_t, float fx, float fy,
                       float* d_out, int64_t n, void* stream);

/* ---- ADD-L1 mesh losses                                   Iterative/loss.py:10-69 --------------------------------
 * t_gt, t_pred: (b, 16) poses; pts: (b, npts, 3) vertices, read ONCE (12 B/vertex).
 * disentangled == 0: item_b = mean_{k,i} |(T_gt p_k - T_pred p_k)_i|                       (compute_ADD_L1_loss)
 * disentangled != 0: rotation + xy-translation + depth terms of compute_disentangled_ADD_L1_loss.
 * item (b, nullable), *sum = sum_b item_b (nullable), grad[b][12] = d item_b / d T_pred[:3, :4] (nullable). */
int pe_add_l1(const float* t_gt, const float* t_pred, const float* pts, int64_t b, int64_t npts, int disentangled,
              float* item, double* sum, float* grad, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PE_ROTATION_H */
