/* libcalib_b200.so -- C-ABI of the B200-native (sm_100a) Levenberg-Marquardt refinement
 * stage of Zhang's-method camera calibration.
 *
 * The reference (pvphan/camera-calibration) is pure Python and has no FFI; the boundary it
 * offers for this path is the method
 *     Calibrator.refineCalibrationParameters(A, W, k, allDetections, maxIters, shouldPrint)
 *         -> (sse, A, W, k)                                   src/calibrate.py:117-171
 * plus, for parity, ProjectionJacobian.compute (src/jacobian.py:48-85) and
 * Calibrator.projectAllPoints (src/calibrate.py:190-197).  Each entry point below names the
 * reference code it replaces.  INTEGRATION.md shows the ctypes stub a maintainer of the
 * reference would add.
 *
 * Conventions
 *   - plain C, no exceptions or aborts cross the ABI; every call returns 0 on success or a
 *     negative CALIB_E_* code, and calib_last_error() gives the text;
 *   - the caller owns every host buffer (C-contiguous float64 / int64, as numpy makes them);
 *     the library owns the handle and all device memory;
 *   - calls are synchronous (they return when the handle's stream is idle);
 *   - a handle is bound to one GPU and is not thread-safe; distinct handles are independent;
 *   - there is NO CPU fallback: without a CUDA device calib_create fails with CALIB_E_CUDA.
 *
 * Parameter vector P (src/calibrate.py:199-229, src/symbolic.py:8-9), L = 10 | 9:
 *     [alpha, beta, gamma, uc, vc, k...(radtan k1,k2,p1,p2,k3 | fisheye k1..k4),
 *      (rho_x, rho_y, rho_z [DEGREES], tx, ty, tz) for view 0, ..., view M-1]      K = L + 6M
 * Points: view-major CSR.  view_offsets[M+1]; model_pts N x 3 (X,Y,Z); sensor_pts N x 2 (u,v).
 * Jacobian rows: view-major, point order as given, (u,v) interleaved (src/jacobian.py:175-186).
 */
#ifndef CALIB_B200_H
#define CALIB_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CALIB_MODEL_RADTAN 0  /* src/distortion.py:74-108  RadialTangentialModel */
#define CALIB_MODEL_FISHEYE 1 /* src/distortion.py:194-220 FisheyeModel */

#define CALIB_OK 0
#define CALIB_E_ARG -1    /* bad argument (NULL, negative size, unknown model) */
#define CALIB_E_CUDA -2   /* CUDA runtime error or no device */
#define CALIB_E_STATE -3  /* call not valid in this state (e.g. maxIters <= 0) */
#define CALIB_E_COMM -4   /* NCCL error / communicator missing */

typedef struct calib_handle calib_t;
typedef struct calib_batch_handle calib_batch_t;

/* Last error text of a handle (or of the last failed create when h == NULL). */
const char* calib_last_error(const calib_t* h);

/* Copies the detections to device memory once (they are constant over the LM loop).
 * Replaces the per-iteration Python lists of src/calibrate.py:138-139,277-282. */
int calib_create(calib_t** out, int model, int64_t M, const int64_t* view_offsets,
                 const double* model_pts, const double* sensor_pts, int device_id);
/* The same from the reference's own container: allDetections is a list of per-view (sensorPoints
 * (N_i,2), modelPoints (N_i,3)) arrays (src/dataset.py:31-34, src/calibrate.py:277-282), each its own
 * allocation.  counts[i] = N_i; model_ptrs[i] / sensor_ptrs[i] = the C-contiguous float64 data of view i
 * (may be NULL where N_i = 0).  The views are gathered into page-locked staging owned by the library by a
 * few host threads, chunk by chunk, while the previous chunks cross PCIe -- the caller never builds the
 * concatenated copy.  Results are identical to calib_create on the concatenated arrays. */
int calib_create_from_views(calib_t** out, int model, int64_t M, const int64_t* counts,
                            const double* const* model_ptrs, const double* const* sensor_ptrs, int device_id);
void calib_destroy(calib_t* h);
/* Device memory of destroyed handles (and the scratch of the per-call entry points) stays in a stream-ordered
 * pool that the library owns, one per device -- never the device's default pool -- so that the next calib_create
 * pays no cudaMalloc.  calib_trim hands what is not in use back to the driver (e.g. before another framework in
 * the process needs the memory).  No counterpart in the reference (its arrays are numpy's). */
int calib_trim(int device_id);

int64_t calib_num_params(const calib_t* h);  /* K = L + 6M */
int64_t calib_num_points(const calib_t* h);  /* N */

/* Tuning: points per warp of the sweeps (0 = automatic: one contiguous chunk per resident warp
 * of the persistent grids).  Rebuilds the partition; results change only in summation order. */
int calib_set_segment(calib_t* h, int points_per_task);

/* How the LM loop (src/calibrate.py:143-168) is driven.  2: one CUDA graph whose WHILE node
 * repeats the iteration until the device-side accept step clears the condition -- a whole
 * refinement is one launch and one read-back.  1: the host launches iteration graphs in small
 * batches and looks at the device state between them.  0 (default): 2 on one GPU and with the
 * peer-memory transport, 1 with NCCL (a WHILE body cannot hold NCCL nodes: 2 then fails with
 * CALIB_E_STATE). */
int calib_set_loop_mode(calib_t* h, int mode);

/* Damping ladder.  After a rejected step the reference multiplies lambda by 10 and tries again from
 * the same point with the same normal equations (src/calibrate.py:161-166).  The library may take
 * the next few of those trial steps ahead of the decisions -- one pass of the Schur, back-substitution
 * and cost kernels for up to max_candidates dampings lambda, 10 lambda, ... (one candidate after an
 * accepted step, doubling after rejected ones) -- and then replays the reference's accept/reject rule
 * candidate by candidate.  Iterates, costs, lambdas, trace rows and the iteration count are those of
 * the one-at-a-time loop, bit for bit; only the number of launches changes.  1, 2 or 4 (default 4). */
int calib_set_ladder(calib_t* h, int max_candidates);
/* Candidates of the pass that follows an ACCEPTED step: 1; 2 (the retry at 10 lambda rides along: LM runs
 * bounce, the step tried right after an accepted one is often rejected); 0 (default) = adaptive, 2 while that
 * keeps happening, else 1.  Bit-identical results. */
int calib_set_ladder_start(calib_t* h, int candidates);

/* How the trial-cost sweep brings the detections in: 1 = bulk copies (TMA) into a per-warp
 * shared-memory ring, 0 = coalesced loads with register prefetch, -1 (default) = by size (the ring
 * when the detections exceed what L2 keeps between sweeps).  Same points, same code; a lane's share of a
 * warp's chunk differs between the two, so the costs agree to rounding (1e-13), not bit for bit. */
int calib_set_cost_staging(calib_t* h, int mode);

/* Page-locked host memory for detections / parameters handed to calib_create / calib_lm_run
 * (optional: any host pointer works, pinned ones copy at full PCIe rate).  NULL on failure. */
void* calib_host_alloc(size_t bytes);
void calib_host_free(void* p);
/* Concatenates n separately allocated host arrays (srcs[i], bytes[i]) into dst with a few threads: the
 * per-view arrays of the reference's allDetections list (src/dataset.py:31-34) -> the CSR buffers of
 * calib_create.  nthreads <= 0: automatic.  Host only; no device needed. */
int calib_host_gather(void* dst, const void* const* srcs, const int64_t* bytes, int64_t n, int nthreads);

/* Calibrator.projectAllPoints, src/calibrate.py:190-197.  y: N x 2. */
int calib_project(calib_t* h, const double* P, double* y);

/* Calibrator._computeReprojectionError, src/calibrate.py:178-188. */
int calib_cost(calib_t* h, const double* P, double* E);

/* Parity/debug: r = measured - projected (2N, src/calibrate.py:151) and the two non-zero
 * blocks of ProjectionJacobian.compute (src/jacobian.py:76-83): Jc 2N x L, Jv 2N x 6.
 * Any output may be NULL. */
int calib_eval_residual_jacobian(calib_t* h, const double* P, double* r, double* Jc, double* Jv);

/* Parity/debug: the blocks of J^T J and J^T r (src/calibrate.py:146,152) as the fused
 * kernels accumulate them, never forming J: A L x L, B M x L x 6, D M x 6 x 6, g K, E. */
int calib_eval_normal_equations(calib_t* h, const double* P, double* A, double* B, double* D,
                                double* g, double* E);

/* One damped step from P without moving: delta (K) solves
 * (JTJ + lam diag JTJ) delta = JTr  (src/calibrate.py:147,152) by Schur complement. */
int calib_lm_step(calib_t* h, const double* P, double lam, double* delta);

/* The LM loop of Calibrator.refineCalibrationParameters, src/calibrate.py:141-171, entirely on
 * the device.  P_inout: K values, refined in place.  lam0/lam_min/lam_max/err_min are the
 * Calibrator class constants (src/calibrate.py:13-16).  sse_out = cost before the last
 * executed iteration's step (the reference's stale Pt_error, :155,171).  trace (nullable):
 * maxIters x 4 doubles, rows (E, E', lambda, accepted) for the executed iterations.
 * maxIters <= 0 returns CALIB_E_STATE (the reference raises UnboundLocalError). */
int calib_lm_run(calib_t* h, double* P_inout, int maxIters, double lam0, double lam_min,
                 double lam_max, double err_min, double* sse_out, int* iters_out, double* trace);

/* Same loop, but P stays on the device between calls: upload / run / download separately so a
 * caller can time the device part alone (bench.py "value"). */
int calib_upload_params(calib_t* h, const double* P);
int calib_lm_run_resident(calib_t* h, int maxIters, double lam0, double lam_min, double lam_max,
                          double err_min, double* sse_out, int* iters_out);
int calib_download_params(calib_t* h, double* P);
int calib_download_trace(calib_t* h, int iters, double* trace);
/* What the reference prints per iteration when shouldPrint is set (src/calibrate.py:158-159, 269-274):
 * A and k of the iterate BEFORE the update.  calib_set_param_trace(h, 1): the runs that follow record the
 * L shared parameters (alpha, beta, gamma, uc, vc, k...) of the iterate each trace row started from (written
 * by the accept step: no extra launch); calib_download_param_trace: out = iters x L. */
int calib_set_param_trace(calib_t* h, int on);
int calib_download_param_trace(calib_t* h, int iters, double* out);

/* Device time of the last calib_lm_run* (CUDA events on the handle's stream), and the number
 * of kernels it launched.  Per-kernel timing: calib_set_profiling(h,1) brackets every launch of
 * the fused passes with events (slower; not for headline numbers). */
int calib_last_run_stats(const calib_t* h, double* device_ms, int64_t* kernel_launches,
                         int64_t* jacobian_evals);
int calib_set_profiling(calib_t* h, int on);
/* which: 0 = fused evaluation pass (k_eval), 1 = per-view Schur + reduced system (k_schur),
 * 2 = back-substitution (k_backsub), 3 = cost sweep (k_cost), 4 = reduced solve as its own
 * launch (sharded views only).  Sums over the last run. */
int calib_kernel_time(const calib_t* h, int which, double* total_ms, int64_t* launches);

/* Average device time (microseconds) of one kernel of the iteration at P with damping lam:
 * which 0 = k_eval, 1 = k_schur (+ reduced solve), 2 = k_backsub, 3 = k_cost; `reps` launches
 * back to back between two CUDA events on the handle's stream.  flush = 0: detections stay in L2
 * between launches, as between the iterations of a run; flush = 1: 256 MB L2 flush before every
 * launch (its own time subtracted).  This is what bench.py's roofline divides by. */
int calib_time_kernel(calib_t* h, const double* P, double lam, int which, int reps, int flush,
                      double* avg_us);

/* Micro-benchmarks used for the roofline denominators (FP64 FMA pipe; device copy). */
int calib_measure_fp64_peak(int device_id, double* tflops);
int calib_flush_l2(calib_t* h);

/* Calibrator._refineHomographies / _refineHomography, src/calibrate.py:60-115 (the step before the
 * refinement in estimateCalibrationParameters, :53): M independent Levenberg-Marquardt fits of the 9
 * homography entries to the handle's detections (model X, Y -> sensor u, v), HomographyJacobian
 * (src/jacobian.py:88-121) in analytic form, one warp per view, the whole loop in one launch.
 * H_inout: M x 9 row-major (h11 h12 h13 h21 ... h33) initial homographies, replaced by the refined
 * ones divided by h33 (:108-109).  The reference hard-codes maxIters = 20 and lambda0 = 1e-3 (:77-78)
 * and uses the class constants for the rest.  iters_out (nullable): M; trace (nullable): M x maxIters
 * x 4 rows (E, E', lambda, accepted), zero after a view's last iteration. */
int calib_refine_homographies(calib_t* h, double* H_inout, int maxIters, double lam0, double lam_min,
                              double lam_max, double err_min, int* iters_out, double* trace);

/* Per-point parts of the closed-form initial estimate, Calibrator.estimateCalibrationParameters
 * (src/calibrate.py:41-58), on the handle's detections; what is left for the caller is O(M) small
 * dense algebra (see camera_calibration_b200/linearcalibrate.py and INTEGRATION.md).
 *
 * calib_dlt_gram: linearcalibrate.estimateHomography, src/linearcalibrate.py:24-58.  Per view the point
 * normalisations of computeNormalizationMatrix (:61-91) -- norm: M x 6 = (mean_u, mean_v, scale_a,
 * mean_X, mean_Y, scale_b) -- and the Gram matrix M^T M of the 2N x 9 DLT matrix of :46-52 in packed
 * form -- gram: M x 24 = the six entries (XX, XY, X, YY, Y, 1) of p p^T, p = (X, Y, 1) normalised,
 * summed with weights 1, u, v, u^2 + v^2: M^T M = [[P1, 0, -Pu], [0, P1, -Pv], [-Pu, -Pv, Pq]].  The
 * homography is its eigenvector of the smallest eigenvalue, de-normalised (:55-57).
 *
 * calib_distortion_rows: DistortionModel.estimateDistortion, src/distortion.py:110-191 (radial-
 * tangential) / :222-271 (fisheye).  A: 3 x 3 intrinsic matrix, W: M x 12 board poses in the camera
 * (R row-major, then t).  rows: M x 20 per-view sums: the upper triangle of D^T D (nd (nd+1) / 2,
 * row-major) then D^T Ddot (nd), nd = 5 | 4, zero padded; k solves (sum D^T D) k = sum D^T Ddot. */
int calib_dlt_gram(calib_t* h, double* gram, double* norm);
/* linearcalibrate.estimateHomographies, src/linearcalibrate.py:7-21, complete: the Gram sums above and,
 * by the same warp, the eigenvector of the smallest eigenvalue of the 9 x 9 matrix (cyclic Jacobi
 * rotations; the reference: last right-singular vector of the 2N x 9 matrix, :52-56), de-normalised and
 * divided by H[2,2] (:57-58).  H: M x 9 row-major. */
int calib_dlt_homographies(calib_t* h, double* H);
int calib_distortion_rows(calib_t* h, const double* A, const double* W, double* rows);

/* Multi-GPU, one process per GPU: views are sharded, every rank holds the L shared parameters
 * and its own views' extrinsics.  The reduced system (L x L Schur complement, its right-hand
 * side and the cost) is all-reduced over NCCL each iteration; everything else is local.
 * id128: the 128-byte ncclUniqueId made by rank 0 and broadcast by the host. */
int calib_comm_unique_id(void* id128);
int calib_comm_init(calib_t* h, int nranks, int rank, const void* id128);
/* Borrow the communicator of another handle of this process (it must outlive h): lets a
 * long-lived communicator serve one calib_create/calib_lm_run/calib_destroy per calibration. */
int calib_comm_share(calib_t* h, const calib_t* owner);
/* Every handle created on device_id from now on borrows owner's communicators (as calib_comm_share
 * would, but before calib_create builds its loop graph); owner = NULL clears it.  Destroying the owner
 * clears it too. */
int calib_comm_set_default(const calib_t* owner, int device_id);

/* The same exchange WITHOUT a library collective: the last thread block of the Schur kernel / the
 * cost kernel writes its 161-double reduced system (resp. the scalar trial cost) straight into every
 * peer's inbox through the NVLink mapping of peer memory, waits for the peers' flags and sums in rank
 * order (bit-identical on all ranks).  The iteration keeps its four launches and the whole LM loop
 * stays one WHILE-node graph launch per rank.  At most 8 ranks, all on one node.
 *   calib_peer_export : allocates this rank's inbox and returns its 64-byte CUDA IPC handle;
 *   (the host all-gathers the handles, e.g. with torch.distributed)
 *   calib_peer_attach : maps the peers' inboxes; handles = nranks x 64 bytes, in rank order.
 * Every rank must then make the same sequence of calib_* compute calls (each exchange is a
 * rendezvous); a rank that never arrives makes the others return CALIB_E_COMM after a timeout.
 * Keep the exporting handle alive until all ranks are done (a host barrier before calib_destroy).
 * calib_comm_share also lends the peer communicator. */
int calib_peer_export(calib_t* h, void* handle64);
int calib_peer_attach(calib_t* h, int nranks, int rank, const void* handles);
/* 0 = automatic (peer memory when attached, else NCCL), 1 = NCCL, 2 = peer memory. */
int calib_set_transport(calib_t* h, int mode);
int calib_transport(const calib_t* h);   /* in use: 0 none (one GPU), 1 NCCL, 2 peer memory */

/* ONE process, SEVERAL GPUs: what the reference's boundary needs to use more than one device, since its
 * call is one synchronous method of one Python process (Calibrator.refineCalibrationParameters,
 * src/calibrate.py:117-118; main.calibrateCamera, src/main.py:34-35).  The library shards the views over
 * device_ids[0 .. n_gpus) (contiguous ranges balanced by point count; never more devices than views), keeps
 * one ordinary handle per device -- created by one host thread each, so the uploads use all the PCIe links
 * at once -- and joins them with the in-kernel exchange described above: the peers' inboxes are reached
 * through cudaDeviceEnablePeerAccess (one address space; no IPC handles, no torch, no launcher).  The
 * caller passes and gets back ONE full-length parameter vector (K = L + 6 M); results are those of the
 * one-process-per-GPU path, bit for bit.  calib_multi_lm_run: arguments of calib_lm_run plus device_ms
 * (nullable: device time of the slowest rank, CUDA events).  One host thread enqueues every device's
 * whole-loop launch, then waits for all of them.  The devices must be peer-accessible (NVLink / NVSwitch). */
typedef struct calib_multi_handle calib_multi_t;
int calib_multi_create(calib_multi_t** out, int model, int64_t M, const int64_t* view_offsets,
                       const double* model_pts, const double* sensor_pts, int n_gpus, const int* device_ids);
int calib_multi_create_from_views(calib_multi_t** out, int model, int64_t M, const int64_t* counts,
                                  const double* const* model_ptrs, const double* const* sensor_ptrs,
                                  int n_gpus, const int* device_ids);
void calib_multi_destroy(calib_multi_t* h);
const char* calib_multi_last_error(const calib_multi_t* h);
int calib_multi_num_gpus(const calib_multi_t* h);        /* devices in use (<= n_gpus asked for) */
int64_t calib_multi_num_params(const calib_multi_t* h);  /* K = L + 6M */
int calib_multi_view_bounds(const calib_multi_t* h, int64_t* bounds /* num_gpus + 1 */);
int calib_multi_set_loop_mode(calib_multi_t* h, int mode);   /* 0 auto, 2 WHILE graph */
int calib_multi_set_ladder(calib_multi_t* h, int max_candidates);
int calib_multi_flush_l2(calib_multi_t* h);                  /* calib_flush_l2 on every device */
int calib_multi_set_param_trace(calib_multi_t* h, int on);
int calib_multi_download_param_trace(calib_multi_t* h, int iters, double* out);
int calib_multi_lm_run(calib_multi_t* h, double* P_inout, int maxIters, double lam0, double lam_min,
                       double lam_max, double err_min, double* sse_out, int* iters_out, double* trace,
                       double* device_ms);

/* Batch of independent calibrations (a Python `for` over refineCalibrationParameters calls in the
 * reference; BASELINE config 4): one thread block per problem, the whole LM loop in one launch,
 * problems up to about 100 views.  problem_view_offsets[n_problems+1] indexes view_offsets
 * (total_views+1, point offsets over ALL problems); P is the concatenation of the problems'
 * parameter vectors (problem p starts at L p + 6 problem_view_offsets[p]); sse / iters have
 * n_problems entries; trace (nullable): n_problems x maxIters x 4, rows (E, E', lambda,
 * accepted), zero after a problem's last iteration.  Split the problems over GPUs by creating
 * one handle per device: there is no communication between problems. */
int calib_batch_create(calib_batch_t** out, int model, int64_t n_problems,
                       const int64_t* problem_view_offsets, const int64_t* view_offsets,
                       const double* model_pts, const double* sensor_pts, int device_id);
void calib_batch_destroy(calib_batch_t* h);
const char* calib_batch_last_error(const calib_batch_t* h);
int calib_batch_lm_run(calib_batch_t* h, double* P_inout, int maxIters, double lam0,
                       double lam_min, double lam_max, double err_min, double* sse_out,
                       int* iters_out, double* trace, double* device_ms);

/* ---- allocation-free entry point --------------------------------------------------------------------
 * NO counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Total reprojection error E(P) = sum |y_observed - project(P)|^2 (what calib_cost returns) of detections that are
 * ALREADY ON THE DEVICE; every pointer is a DEVICE pointer of device_id:
 *   view_offsets_dev  int64[M + 1]  first point of every view (view_offsets_dev[0] = 0, [M] = N), ragged / empty views allowed
 *   model_pts_dev     double[N][3], sensor_pts_dev  double[N][2]   (the layout calib_create takes from the host)
 *   P_dev             double[L + 6 M]  parameter vector (L = 10 radtan, 9 fisheye), E_dev  double[1]  the result
 * The kernels are enqueued on cuda_stream (a cudaStream_t; NULL = the default stream) and the call returns without
 * synchronising: E_dev is valid in stream order.  The call allocates nothing (no cudaMalloc / cudaFree / host
 * allocation, no stream or event): its scratch -- view constants, per-block partial sums, the reduction's ticket --
 * is one static device array of 206 336 bytes that every GPU loading the library holds until the process exits.
 * Limit: M <= CALIB_NOALLOC_MAX_VIEWS (1024); more is refused with CALIB_E_ARG and a message (calib_last_error(NULL)).
 * All calls on a GPU share that array: calls in stream order are independent (each rewrites what it reads and clears
 * the ticket first), calls that OVERLAP on one GPU (two streams, two threads) race -- serialise them.
 * Agrees with calib_cost to rounding (same per-point arithmetic, another summation order): 1e-12 relative. */
#define CALIB_NOALLOC_MAX_VIEWS 1024
int calib_cost_noalloc(int model, int64_t M, const int64_t* view_offsets_dev, int64_t N,
                       const double* model_pts_dev, const double* sensor_pts_dev, const double* P_dev,
                       double* E_dev, int device_id, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif
