/* ctba.h — C ABI of the B200-native continuous-time bundle adjuster.
 *
 * Drop-in boundary: everything the reference executes inside
 *     ceres::Solver::Summary BSplineSO3R3<N>::optimize(int max_num_iterations)
 *         /root/reference/include/spline/bspline_so3r3.h:1768-1782   (-> ceres::Solve, :1778)
 * plus the problem-assembly calls that feed it.  Each entry point names the reference interface it
 * replaces.  Plain pointers and sizes only; no C++ types, no exceptions, no abort() across the ABI.
 * All floating point data is IEEE binary64.  Quaternions are stored [x y z w] (so3.hpp:185-187).
 *
 * The library is CUDA-only (sm_100a): every compute entry point returns CTBA_E_CUDA when no device
 * is usable.  There is no CPU fallback.
 */
#ifndef CTBA_H_
#define CTBA_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ctba_problem ctba_problem;

enum {
  CTBA_OK = 0,
  CTBA_E_ARG = -1,      /* bad argument (replaces GVI_FUSION_ASSERT_STREAM aborts, util/assert.h:65-70) */
  CTBA_E_CUDA = -2,     /* CUDA runtime failure / no device */
  CTBA_E_NCCL = -3,
  CTBA_E_NUMERIC = -4,  /* non-finite values / factorisation breakdown */
  CTBA_E_OOM = -5,
  CTBA_E_STATE = -6     /* call order violation (e.g. measurements before knots) */
};

/* residual kinds (ctba_evaluate) */
enum { CTBA_REPROJ = 0, CTBA_ACCEL = 1, CTBA_GYRO = 2, CTBA_GP = 3, CTBA_BIAS = 4, CTBA_GRAVITY = 5, CTBA_PRIOR = 6, CTBA_ALIGN = 7 };

/* BSplineSO3R3<N>::BSplineSO3R3(time_interval_ns, start_time_ns)      bspline_so3r3.h:75-117
 * order N in {4,6}; `device` = CUDA ordinal (-1: planning-only handle — host problem assembly and ctba_shard_plan work,
 * every compute entry point returns CTBA_E_CUDA). */
int ctba_create(int order, int64_t dt_ns, int64_t start_t_ns, int device, ctba_problem** out);
int ctba_destroy(ctba_problem* p);
const char* ctba_last_error(const ctba_problem* p);
/* Use an existing CUDA stream (cudaStream_t passed as void*) for all work of this handle; 0 = own stream. */
int ctba_set_stream(ctba_problem* p, void* cuda_stream);

/* initializeFromNumTraj / so3_knots_, r3_knots_                        bspline_so3r3.h:336-353, 1936-1939 */
int ctba_set_knots(ctba_problem* p, int n, const double* quat_xyzw /*4n*/, const double* pos /*3n*/);
int ctba_num_knots(const ctba_problem* p);
int ctba_get_knots(ctba_problem* p, double* quat_xyzw, double* pos);

/* initializeBiases(dt_s_bias, init_acc_bias, init_gyro_bias)            bspline_so3r3.h:355-369
 * returns the number of bias knots (>=0) or an error (<0). */
int ctba_initialize_biases(ctba_problem* p, double dt_s_bias, const double* acc0, const double* gyr0);
int ctba_num_bias_knots(const ctba_problem* p);
int ctba_set_biases(ctba_problem* p, const double* acc /*3nb*/, const double* gyr /*3nb*/);
int ctba_get_biases(ctba_problem* p, double* acc, double* gyr);

/* set3DPoints (points3D_[id].p_wl); landmark identity is the array index  bspline_so3r3.h:1907-1908, 1982 */
int ctba_set_landmarks(ctba_problem* p, int n, const double* xyz /*3n*/);
int ctba_get_landmarks(ctba_problem* p, double* xyz);

/* setCameras (first camera only): COLMAP model id 4|5|7, intr = fx fy cx cy, dist = 4 coefficients,
 * T_b_c as quaternion + translation, t_offset_cam_imu                   bspline_so3r3.h:1899-1903, sensor/camera.h:41-90 */
int ctba_set_camera(ctba_problem* p, int model, const double* intr, const double* dist, const double* q_bc, const double* p_bc,
                    double t_offset_cam_imu_s);
/* setScale / setPbp / setGravity / setInitSplineGlobSensTimeOffset      bspline_so3r3.h:1792-1798, 1841-1845, 1880 */
int ctba_set_globals(ctba_problem* p, double scale, const double* p_bp, const double* g, double t_s_gp_s);
/* corr[10] = scale_corr, R_b_c_corr (xyzw), p_b_c_corr, t_offset_cam_imu_corr_s, t_s_gp_corr_s (members :1955-1973);
 * only needed to evaluate at a non-default linearisation point. */
int ctba_set_corrections(ctba_problem* p, const double* corr10);
/* out[19] = corr[0..9], p_bp[10..12], g[13..15], t_offset_cam_imu_s[16], t_s_gp_s[17], scale_[18]
 * (getScale/getPbp/getGravity/getImuCamRelTransformationCorr/getSplineImuTimeOffset...  :1792-1881) */
int ctba_get_globals(ctba_problem* p, double* out19);
/* setOptimizeScale / setOptimizePbp / setOptimizeGravity / setOptimizeCamImuExtrinsics /
 * setOptimizeTimeOffsetSplineGlobalSensor                                bspline_so3r3.h:1794-1897 */
int ctba_set_flags(ctba_problem* p, int optimize_scale, int optimize_pbp, int optimize_gravity, int optimize_camimu_extrinsics,
                   int optimize_t_offset_s_gs);

/* addImagePointsMeasurement(frame, t_ns, noise) for F frames at once: CSR obs_begin[F+1] over the frames'
 * VALID reprojections (is_reproj_valid), uv pixels, landmark indices.  CauchyLoss(3.0) per frame.
 * Returns the number of frames accepted (skip rules :962-973).           bspline_so3r3.h:960-1046 */
int ctba_add_frames(ctba_problem* p, int F, const int64_t* t_ns, const int32_t* obs_begin, const double* uv /*2M*/,
                    const int32_t* lm_index /*M*/, double sigma_px);
/* addAccelMeasurement + addGyroMeasurement for n samples                 bspline_so3r3.h:523-702 */
int ctba_add_imu(ctba_problem* p, int n, const int64_t* t_ns, const double* acc /*3n*/, const double* gyr /*3n*/,
                 double acc_noise_density, double gyro_noise_density);
/* addGlobalPosMeasurement for n samples                                  bspline_so3r3.h:880-958 */
int ctba_add_gp(ctba_problem* p, int n, const int64_t* t_ns, const double* pos /*3n*/, double noise);
/* addPositionMeasurement / addOrientationMeasurement for n samples (the spline-fitting tools: src/fit_spline_to_colmap.cpp:182-192,
 * fit_spline_to_gp_measurements)                                         bspline_so3r3.h:1048-1085, 1087-1123
 * residuals: PositionCostFunctor (optimization/continuous_time/position_cost_functor.h:39-53)  r = (p(u) - meas) / noise,
 *            OrientationCostFunctor (orientation_cost_functor.h:39-57)   r = 2 vec(R(u) meas^-1) / noise;  quaternions as x y z w.
 * Returns the number of samples accepted (same skip rules as the reference).  Problems with pose priors run on one GPU. */
int ctba_add_position_measurements(ctba_problem* p, int n, const int64_t* t_ns, const double* pos /*3n*/, double noise);
int ctba_add_orientation_measurements(ctba_problem* p, int n, const int64_t* t_ns, const double* quat_xyzw /*4n*/, double noise);
int ctba_num_priors(const ctba_problem* p);
/* ---- align_spline_to_global_frame (src/align_spline_to_global_frame.cpp:153-226): a FROZEN spline expressed in a gauge frame g is aligned
 * to the world frame of the global position sensor.  Unknowns: scale_corr, R_w_g_corr, p_w_g_corr, and p_bp / t_s_gp_corr when their
 * optimise flags are set; residual GlobalPositionSensorWithAlignmentCorrectionCostFunctor (global_position_cost_functor.h:82-150),
 * CauchyLoss(0.10) per measurement.  A handle that holds alignment measurements solves THIS problem in ctba_solve (every knot constant,
 * bspline_so3r3.h:1191-1199); mixing them with other measurement kinds is CTBA_E_STATE.
 * setWorldColmapTransformation(T_w_g)                                    bspline_so3r3.h:1825     T = qx qy qz qw tx ty tz;
 * corr7 = R_w_g_corr | p_w_g_corr (NULL: unchanged; only needed to evaluate at a non-default linearisation point) */
int ctba_set_world_alignment(ctba_problem* p, const double* T_wg7, const double* corr7);
/* getWorldColmapTransformation / getWorldColmapTransformationCorr       bspline_so3r3.h:1827-1840  (either pointer may be NULL) */
int ctba_get_world_alignment(ctba_problem* p, double* T_wg7, double* corr7);
/* addWorldColmapAlignmentMeasurement for n samples; returns the number accepted   bspline_so3r3.h:1121-1211 */
int ctba_add_alignment_measurements(ctba_problem* p, int n, const int64_t* t_ns, const double* pos /*3n*/, double noise);
int ctba_num_alignment_measurements(const ctba_problem* p);
/* addImuBiasCost(bias_err_freq_hz, bias_err_dt_s); random walks from ImuCalibration   bspline_so3r3.h:706-792 */
int ctba_add_bias_cost(ctba_problem* p, double freq_hz, double dt_s, double acc_random_walk, double gyro_random_walk);
/* addGravityNormError()                                                  bspline_so3r3.h:1213-1227 */
int ctba_add_gravity_norm(ctba_problem* p);

/* counts[8] = frames, observations, accel, gyro, gp, bias blocks, gravity, reduced-system unknowns */
int ctba_counts(ctba_problem* p, int64_t* counts);

/* Raw residuals and local (tangent-space) Jacobians of one residual kind at the current state, row-major,
 * without the robust-loss correction (what DynamicAutoDiffCostFunction::Evaluate + LocalParameterization give):
 *   CTBA_REPROJ  per observation 2 x (6N+10): [3N rot | 3N pos | scale_corr | R_bc_corr 3 | p_bc_corr 3 | landmark 3]
 *   CTBA_ACCEL   per sample      3 x (6N+16): [3N rot | 3N pos | 4 acc-bias knots | g 3 | t_offset_cam_imu_corr]
 *   CTBA_GYRO    per sample      3 x (3N+13): [3N rot | 4 gyro-bias knots | t_offset_cam_imu_corr]
 *   CTBA_GP      per sample      3 x (6N+5) : [3N rot | 3N pos | p_bp 3 | t_s_gp_corr | scale_corr]
 *   CTBA_BIAS    per block       3 x 24     : [4 knots @s0 | 4 knots @s1]
 *   CTBA_GRAVITY                 1 x 3
 *   CTBA_PRIOR                   3 x 3N   (position prior: the position knots; orientation prior: the right tangents of the rotation knots)
 *   CTBA_ALIGN   per sample      3 x 11     : [t_s_gp_corr | scale_corr | R_w_g_corr 3 | p_w_g_corr 3 | p_bp 3]   (the knots are constant)
 * residuals / jacobians may be NULL.  Returns the number of residuals written (>=0) or an error.
 * Also serves evaluate*Errors (bspline_so3r3.h:1229-1579). */
int ctba_evaluate(ctba_problem* p, int kind, double* residuals, double* jacobians);
/* cost[7]: 1/2 sum rho(|r|^2) per kind (Cauchy on reprojection frames), cost[6] = total */
int ctba_cost(ctba_problem* p, double* cost7);
/* the same with cost[7] = the pose priors (position + orientation measurements; included in the total cost[6]) */
int ctba_cost_terms(ctba_problem* p, double* cost8);

typedef struct ctba_iteration {
  double cost, cost_change, gradient_max_norm, step_norm, tr_ratio, tr_radius;
  int successful;
} ctba_iteration;

typedef struct ctba_summary {
  int termination_type; /* 0 CONVERGENCE, 1 NO_CONVERGENCE, 2 FAILURE  (ceres::TerminationType; drivers test == 0) */
  double initial_cost, final_cost;
  int num_successful_steps, num_unsuccessful_steps;
  int num_iterations;            /* entries written to `iterations` (iteration 0 = initial evaluation) */
  ctba_iteration* iterations;    /* caller-provided array of `iterations_capacity` entries, may be NULL */
  int iterations_capacity;
  double time_total_s, time_eval_s, time_build_s, time_solve_s; /* device-side breakdown */
  int64_t kernel_launches;       /* kernels launched by this call */
} ctba_summary;

/* BSplineSO3R3::optimize(max_num_iterations) == ceres::Solve            bspline_so3r3.h:1768-1782
 * Runs trust-region LM entirely on the device and writes the optimised state back into the handle
 * (ctba_get_* to read).  Non-convergence is reported in the summary, not as an error. */
int ctba_solve(ctba_problem* p, int max_num_iterations, ctba_summary* summary);

/* computeValidReprojections(frames): the grid-based choice of the reprojections that enter the problem        bspline_so3r3.h:406-521
 * (per-frame feature grid of Frame::addPoint2D / setGrid, sensor/camera.h:155-184).  Input = every 2-D point that has a 3-D point,
 * frame-major CSR pt_begin[F+1] in Frame::points2D order (each landmark at most once per frame, camera.h:372-380).  Output:
 * valid[M] = Point2D::is_reproj_valid, n_valid[F] = Frame::n_valid_reprojections.  Evaluated at the handle's CURRENT knots, scale and
 * camera extrinsics (the drivers call it once, before the first Solve).  Sequential over frames, on the host, like the reference.
 * The caller then passes the valid points to ctba_add_frames. */
int ctba_compute_valid_reprojections(ctba_problem* p, int F, const int64_t* t_ns, const int32_t* pt_begin, const double* uv /*2M*/,
                                     const int32_t* lm_index /*M*/, int width, int height, int cell_size_px, uint8_t* valid /*M*/,
                                     int32_t* n_valid /*F*/);
/* evaluateReprojectionErrors / evaluateAccelErrors / evaluateGyroErrors / evaluateGlobalPosErrors               bspline_so3r3.h:1229-1579
 * norm of every measurement's residual with unit weight (inv_std = 1), in the order the measurements were added;
 * kind = CTBA_REPROJ | CTBA_ACCEL | CTBA_GYRO | CTBA_GP | CTBA_PRIOR (evaluatePositionErrors / evaluateOrientationErrors, :1581-1680) |
 * CTBA_ALIGN (evaluateWorldColmapAlignmentErrors, :1682-1790).  Returns the number of errors written (errors may be NULL to query it). */
int ctba_evaluate_errors(ctba_problem* p, int kind, double* errors);
/* getPose / getAccel / getGyro / getAccelBias / getGyroBias at n timestamps, evaluated on the device at the CURRENT state
 *                                                                       bspline_so3r3.h:221-264, 122-174, 176-213, 266-295, 297-327
 * — the loops of traceSpline / traceAccBiasSpline / traceGyroBiasSpline (include/util/save.h:99-153) and of the drivers' predicted-IMU
 * dumps (src/optimize_continuous_time.cpp:88-130) as one call.  out: 7 doubles per pose in the trace file order tx ty tz qx qy qz qw,
 * 3 doubles otherwise.  GYRO_WORLD applies R_wb^-1 to the body rate exactly like the reference (:203-207); the bias getters evaluate
 * the bias knots with the order-N helper (:285-292).  A time before the spline start or a window beyond the last knot is where the
 * reference asserts: CTBA_E_ARG, nothing written. */
enum { CTBA_SAMPLE_POSE = 0, CTBA_SAMPLE_ACCEL_BODY = 1, CTBA_SAMPLE_ACCEL_WORLD = 2, CTBA_SAMPLE_GYRO_BODY = 3, CTBA_SAMPLE_GYRO_WORLD = 4,
       CTBA_SAMPLE_ACCEL_BIAS = 5, CTBA_SAMPLE_GYRO_BIAS = 6 };
int ctba_sample(ctba_problem* p, int what, int64_t n, const int64_t* t_ns, double* out);

/* relinearizeSplineImuTimeOffset + relinearizeSplineGlobSensTimeOffset   bspline_so3r3.h:1807-1811, 1853-1857
 * (option "relinearize_mask": 1 = only the spline-imu offset, 2 = only the spline / global-sensor offset, 3 = both, the default) */
int ctba_relinearize_time_offsets(ctba_problem* p);

/* Options: "bench_no_termination" (1 = ctba_solve runs exactly max_num_iterations LM iterations, convergence tests
 * disabled; measurement only), "cauchy_a" (CauchyLoss parameter of the reprojection blocks, default 3.0, bspline_so3r3.h:1032). */
int ctba_set_option(ctba_problem* p, const char* name, double value);
/* More options: "linear_solver" (0 = bordered band Cholesky, the exact step of the reference's SPARSE_NORMAL_CHOLESKY, bspline_so3r3.h:1771;
 * 1 = block-Jacobi PCG on the same reduced system, the ITERATIVE_SCHUR-style alternative for bands too wide to factor), "pcg_tol"
 * (|r| <= tol |b|, default 1e-10), "pcg_max_iters", "band_partitions", "band_nested", "bench_accept_all".
 * Implementation switches of the landmark Schur gather, all exact and bit-reproducible, kept for A/B measurements:
 * "schur_mask" (1 = per-observation track bitmask + warp compaction, default; 0 = list search), "schur_threads" (128 / 192 / 256 threads
 * per frame, default 192), "schur_list_a" (partner blocks from the list-ordered copy, default 1), "schur_compact", "schur_two_stage",
 * "schur_thread_below".
 * ctba_get_stat: read-only facts about the last solve — "pcg_iterations", "pcg_rel_residual", "band_block_columns",
 * "band_half_width_blocks", "band_launches". */
int ctba_get_stat(ctba_problem* p, const char* name, double* value);
/* Forget the device mirrors: the next evaluate/solve uploads state and measurements from the host copies again
 * (used to time the end-to-end path with host buffers). */
int ctba_invalidate_device(ctba_problem* p);
/* bytes[2] = host->device bytes of a full upload (state + measurements), device->host bytes of a state read-back */
int ctba_transfer_bytes(ctba_problem* p, int64_t* bytes2);

/* Device-only benchmarking hooks (inputs already resident): run `reps` passes of the residual+Jacobian
 * kernels (which != 0) or of a full LM iteration, return the kernel launch count. */
int ctba_bench_evaluate(ctba_problem* p, int reps, int64_t* launches);
/* The per-observation reprojection kernel alone, every launch timed with CUDA events on the handle's stream; flush_l2 != 0 overwrites a
 * 256 MiB buffer between launches so that inputs come from HBM.  ms_avg = average launch duration. */
int ctba_bench_reproj(ctba_problem* p, int reps, int flush_l2, double* ms_avg);

/* ---- multi-GPU: ONE problem sharded by time segment over the GPUs of a node --------------------------------
 * The reference has no counterpart (Ceres runs `num_threads` CPU threads, bspline_so3r3.h:1773).  Every rank (one process per GPU)
 * makes the SAME calls as a single-GPU caller, with the COMPLETE problem, plus one ctba_comm_init_*; the first ctba_solve /
 * ctba_evaluate / ctba_cost then keeps only the rank's units (landmarks with all their observations, IMU / GP / bias samples whose
 * first control point lies in the rank's time segment) and runs the LM loop with five small collectives per iteration.  After
 * ctba_solve every rank holds the whole optimised state (ctba_get_* as usual).  On a sharded handle ctba_counts / ctba_evaluate
 * describe the rank's shard; ctba_cost and the summary of ctba_solve describe the whole problem. */
/* rank 0: 128-byte NCCL unique id, to be broadcast to the other ranks by the caller's own means */
int ctba_comm_unique_id(char* out128);
int ctba_comm_init_nccl(ctba_problem* p, int world, int rank, const char* id128);
/* In-process group for testing on ONE device: `world` handles of one process, each driven by its own host thread (calls that
 * evaluate or solve are collective: every handle's thread must make them). */
int ctba_comm_init_local(ctba_problem** handles, int world);
/* Host-only shard plan of the problem currently held by `p` (works on a planning handle created with device = -1):
 *   plan[0..15] = P, nbc, nbw, q_lo, q_hi, bc_lo, bc_hi, own_lo, own_hi, F, f_first, f_cost_begin, nodes, hbn, T, KD
 *   plan[16+2q], plan[17+2q] = first block column and width of the interior of partition q   (plan: 16 + 64 entries)
 *   *_owner (optional) = owning rank of every landmark / IMU sample / GP sample / bias block */
int ctba_shard_plan(ctba_problem* p, int world, int rank, int64_t* plan, int32_t* lm_owner, int32_t* imu_owner, int32_t* gp_owner,
                    int32_t* bias_owner);

/* ---- allocation-free entry point ----------------------------------------------------------------------------------------
 * NO counterpart in the reference: added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  It calls no cudaMalloc / cudaFree / allocating library routine.
 *
 * Normal equations of the world-alignment problem (what one LM iteration of ctba_solve accumulates on a handle that holds alignment
 * measurements: GlobalPositionSensorWithAlignmentCorrectionCostFunctor under CauchyLoss(cauchy_a), every knot constant) at the
 * alignment state x12 = [t_s_gp_corr | scale_corr | R_w_g_corr x y z w | p_w_g_corr | p_bp], 11 tangent unknowns in the same order
 * (rotation: right tangent):
 *     d_sys78 = [ upper triangle of J^T J, row-major (66) | J^T r (11) | cost (1) ]      (DEVICE memory, written on `cuda_stream`)
 * T_wg7 ([q x y z w | p]) and x12 are HOST arrays read before the call returns; d_knot_q (4 n_knots), d_knot_p (3 n_knots), d_t_ns (n),
 * d_pos (3n) and d_sys78 are DEVICE pointers on `device`, owned by the caller.  Measurements before the spline start or beyond the last
 * window are skipped like ctba_add_alignment_measurements does.  The call only enqueues two kernels on `cuda_stream` (no
 * synchronisation): the caller synchronises before reading d_sys78.
 * Scratch: per-CTA partial sums live in ONE file-scope static __device__ array of 296 x 78 doubles (184 704 bytes per GPU that loads the
 * library, held until exit).  Limit: n <= CTBA_NOALLOC_MAX_MEASUREMENTS, larger n is CTBA_E_ARG with a message in `err` (optional
 * buffer of err_len bytes).  Every call shares the array (it needs no clearing: a call reads only what it wrote), so calls that OVERLAP
 * on the GPU — two streams, two host threads — race: serialise them. */
#define CTBA_NOALLOC_MAX_MEASUREMENTS (1 << 20)
int ctba_align_system_noalloc(int order, int64_t dt_ns, int64_t start_t_ns, int device, void* cuda_stream, double scale, double t_s_gp_s,
                              const double* T_wg7, const double* x12, double cauchy_a, int n_knots, const double* d_knot_q,
                              const double* d_knot_p, int n, const int64_t* d_t_ns, const double* d_pos, double noise, double* d_sys78,
                              char* err, int err_len);

#ifdef __cplusplus
}
#endif
#endif /* CTBA_H_ */
