/* ctba_dt.h — C ABI of the discrete-time estimator on the GPU (BASELINE.json configs[3]).
 *
 * Drop-in boundary: what the reference executes inside
 *         /root/reference/include/discrete_time_estimator/estimator.h:699-713   DiscreteTimeEstimator::optimize() -> ceres::Solve
 * plus the problem-assembly calls that feed it (estimator.h:88-165, 373-519) and the error evaluators the drivers print (:521-681).
 * One SE3 state [qx qy qz qw tx ty tz] (Sophus layout, se3.hpp:511-520) and one 9-vector [v, b_g, b_a] per frame; three cost terms:
 *   reprojection of a frame's valid points   include/optimization/discrete_time/reprojection_cost_functor.h:36-116   (CauchyLoss(3) per frame)
 *   pre-integrated IMU factor (15 residuals) include/optimization/discrete_time/imu_cost_functor.h:36-942
 *   interpolated global position             include/optimization/discrete_time/global_position_cost_functor.h:24-51
 * Same conventions as ctba.h: plain pointers and sizes, int return codes (CTBA_OK / CTBA_E_*), every array row-major, host pointers.
 * The residuals, Jacobians, normal equations, the bordered band Cholesky and the LM loop all run on the device; there is no CPU fallback.
 *
 * Not carried over (no driver of the reference calls them): addPosePrior / addVelPrior (estimator.h:327-371), setState*Constant (:683-697).
 */
#ifndef CTBA_DT_H_
#define CTBA_DT_H_
#include "ctba.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ctba_dt_problem ctba_dt_problem;

/* residual kinds of ctba_dt_evaluate / ctba_dt_evaluate_errors */
enum { CTBA_DT_REPROJ = 0, CTBA_DT_IMU = 1, CTBA_DT_GP = 2 };

/* DiscreteTimeEstimator::DiscreteTimeEstimator()                       estimator.h:52-86 */
int ctba_dt_create(int device, ctba_dt_problem** out);
int ctba_dt_destroy(ctba_dt_problem* p);
const char* ctba_dt_last_error(const ctba_dt_problem* p);
/* Use an existing CUDA stream (cudaStream_t passed as void*) for all work of this handle; 0 = own stream. */
int ctba_dt_set_stream(ctba_dt_problem* p, void* cuda_stream);
/* The host-side copies become authoritative again: the next compute call uploads state and measurements (bench.py's end-to-end leg). */
int ctba_dt_invalidate_device(ctba_dt_problem* p);

/* setCameras (one camera; the functor reads cameras_[0])               estimator.h:718, sensor/camera.h
 * model: 4 radtan | 5 equidistant | 7 FOV (colmap ids, as ctba_set_camera); intr = fx fy cx cy; dist = 4; q_bc = x y z w; width, height = image size
 * used by the out-of-image gate of the shifted feature (reprojection_cost_functor.h:87-94). */
int ctba_dt_set_camera(ctba_dt_problem* p, int model, const double* intr, const double* dist, const double* q_bc, const double* p_bc,
                       double t_offset_cam_imu_s, double width, double height);
/* setImu (noise densities, random walks, saturation limits) + setGravity    estimator.h:715, 728; sensor/imu_calib.h:13-43 */
int ctba_dt_set_imu(ctba_dt_problem* p, double gyro_noise_density, double acc_noise_density, double gyro_random_walk, double acc_random_walk,
                    double g_max, double a_max, const double* g_W);
/* initializeStates(init_poses, cam_ts): F states, ts_imu = ts_cam + t_offset_cam_imu, speed-and-bias zero     estimator.h:88-103 */
int ctba_dt_initialize_states(ctba_dt_problem* p, int F, const double* poses /*7F*/, const double* ts_cam /*F*/);
/* initializeLinearVelocity: finite differences of the positions        estimator.h:105-125 */
int ctba_dt_initialize_linear_velocity(ctba_dt_problem* p);
/* initializeBiases / setStateVelocity: the whole [v, b_g, b_a] table    estimator.h:127-139, 723-726 */
int ctba_dt_set_speed_and_bias(ctba_dt_problem* p, const double* sb /*9F*/);
/* overwrite poses and speed-and-biases of the existing states (measurements and IMU caches stay) */
int ctba_dt_set_states(ctba_dt_problem* p, const double* poses /*7F*/, const double* sb /*9F*/);
/* getStates                                                            estimator.h:141-147 */
int ctba_dt_get_states(ctba_dt_problem* p, double* poses /*7F*/, double* sb /*9F*/);
/* set3DPoints: dense table, the caller maps ids to rows                estimator.h:720 */
int ctba_dt_set_landmarks(ctba_dt_problem* p, int L, const double* xyz);
int ctba_dt_get_landmarks(ctba_dt_problem* p, double* xyz);
/* setScale, setPbp, setGlobSensImuTimeOffset                           estimator.h:731-743 */
int ctba_dt_set_globals(ctba_dt_problem* p, double scale, const double* p_bp, double t_offset_globsens_imu_s);
/* corr9 = scale_corr, R_BC_corr (x y z w), p_BC_corr (3), t_offset_cam_imu_corr ; t_offset_globsens_imu_corr separately    estimator.h:793-811 */
int ctba_dt_set_corrections(ctba_dt_problem* p, const double* corr9, double t_gs_corr);
/* out13 = scale_corr, R_BC_corr(4), p_BC_corr(3), t_offset_cam_imu_corr, t_offset_globsens_imu_corr, p_BP(3)   (getScaleCorr etc., :744-770) */
int ctba_dt_get_globals(ctba_dt_problem* p, double* out13);
/* setOptimizeScale / setOptimizePbp / setOptimizeCamImuExtrinsics / setOptimizeTimeOffsetGlobSensImu           estimator.h:733-750 */
int ctba_dt_set_flags(ctba_dt_problem* p, int optimize_scale, int optimize_pbp, int optimize_camimu_extrinsics, int optimize_t_offset_globsens_imu);

/* addImagePointsMeasurement for nF frames at once: the VALID reprojections only, frame-major CSR obs_begin[nF+1];
 * vel = Point2D::velocity (computeFeatureVelocity, estimator.h:293-325).  Frames without valid points are skipped.        estimator.h:373-441 */
int ctba_dt_add_frames(ctba_dt_problem* p, int nF, const int32_t* state_id, const int32_t* obs_begin, const double* uv, const double* vel,
                       const int32_t* lm_index, double sigma_px);
/* addImuMeasurements(meas, frame_0, frame_0 + 1): the samples containing both edges (Imu::getMeasurementsContainingEdges)  estimator.h:443-459 */
int ctba_dt_add_imu(ctba_dt_problem* p, int frame0, int n, const double* t_s, const double* gyr /*3n*/, const double* acc /*3n*/);
/* addGlobalSensMeasurement for n measurements: frame0 = GlobalPosMeasurement::closest_frame_id                            estimator.h:461-519 */
int ctba_dt_add_gp(ctba_dt_problem* p, int n, const double* t_s, const double* pos /*3n*/, const int32_t* frame0, double sigma_m);

/* computeValidReprojections(frames): the grid-based choice of the reprojections that enter the problem, frame k judged at state k's
 * CURRENT pose (30 cm / behind-the-camera test).  Same arrays as ctba_compute_valid_reprojections.  Host pass, sequential over frames.   estimator.h:167-291 */
int ctba_dt_compute_valid_reprojections(ctba_dt_problem* p, int F, const int32_t* pt_begin, const double* uv /*2M*/, const int32_t* lm_index /*M*/,
                                        int width, int height, int cell_size_px, uint8_t* valid /*M*/, int32_t* n_valid /*F*/);

/* counts5 = frames with points, valid reprojections, IMU factors, global positions, unknowns (tangent size of the free blocks) */
int ctba_dt_counts(ctba_dt_problem* p, int64_t* counts5);

/* Raw residuals (x 1/noise, before the loss) and LOCAL Jacobians of one kind at the current state, row-major:
 *   CTBA_DT_REPROJ  2 per point,   2 x 17 [pose 6 | scale 1 | R_BC 3 | p_BC 3 | t_cam_imu 1 | landmark 3]
 *   CTBA_DT_IMU     15 per factor, 15 x 30 [pose k 6 | speed-and-bias k 9 | pose k+1 6 | speed-and-bias k+1 9]
 *   CTBA_DT_GP      3 per sample,  3 x 17 [pose0 6 | pose1 6 | scale 1 | t_gs 1 | p_BP 3]
 * pose columns = SE3 tangent [upsilon, omega] of T * exp(delta).  Like every Evaluate of the reference, a call evaluates the IMU factors
 * and may redo their cached pre-integration (imu_cost_functor.h:693-704).  j may be NULL. */
int ctba_dt_evaluate(ctba_dt_problem* p, int kind, double* r, double* j);
/* cost4 = 1/2 sum rho(|r|^2): reprojection, IMU, global position, total */
int ctba_dt_cost(ctba_dt_problem* p, double* cost4);
/* evaluateReprojectionErrors / evaluateImuErrors / evaluateGlobalPosErrors: unit-noise norms per point / 15-vector per factor from a FRESH
 * pre-integration at the current state / unit-noise norms per sample                                                       estimator.h:521-681 */
int ctba_dt_evaluate_errors(ctba_dt_problem* p, int kind, double* out);
/* re-integration counter of IMU factor k (tests) */
int ctba_dt_imu_redo_count(ctba_dt_problem* p, int k, int* count);

/* DiscreteTimeEstimator::optimize() == ceres::Solve (trust-region LM, SPARSE_NORMAL_CHOLESKY, defaults)                     estimator.h:699-713
 * Landmark Schur complement + bordered block-banded Cholesky over the 15-scalar frame nodes (ctba_band.cu), all on the device. */
int ctba_dt_solve(ctba_dt_problem* p, int max_num_iterations, ctba_summary* summary);
/* "bench_no_termination", "bench_accept_all", "band_partitions", "band_graph" as ctba_set_option */
int ctba_dt_set_option(ctba_dt_problem* p, const char* name, double value);
/* "band_block_columns", "band_half_width_blocks", "band_launches" */
int ctba_dt_get_stat(ctba_dt_problem* p, const char* name, double* value);

#ifdef __cplusplus
}
#endif
#endif
