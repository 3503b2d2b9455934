// ctba_internal.h — host-side problem object shared by the C-ABI (ctba_api.cu), the evaluation kernels
// (ctba_eval.cu) and the LM solver (ctba_solve.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/ctba.h"
#include "ctba_terms.h"
#include "ctba_shard.h"

namespace ctba {

// Host arrays that are uploaded on every solve live in page-locked memory once they are large enough to matter (>= 64 KiB), so that the
// host -> device copies run at the link rate instead of through the driver's staging buffers.  Without a CUDA device (planning-only
// handles on a CPU box) the allocator is plain malloc.
void* host_alloc(size_t bytes);
void host_free(void* p);
template <class T>
struct HostAlloc {
  using value_type = T;
  HostAlloc() = default;
  template <class U> HostAlloc(const HostAlloc<U>&) {}
  T* allocate(size_t n) { return static_cast<T*>(host_alloc(n * sizeof(T))); }
  void deallocate(T* p, size_t) { host_free(p); }
  template <class U> bool operator==(const HostAlloc<U>&) const { return true; }
  template <class U> bool operator!=(const HostAlloc<U>&) const { return false; }
};
template <class T> using HVec = std::vector<T, HostAlloc<T>>;

template <class T>
struct DBuf {
  T* p = nullptr;
  size_t n = 0;
  cudaError_t resize(size_t m) {
    if (m <= n && p) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; n = 0;
    if (m == 0) return cudaSuccess;
    cudaError_t e = cudaMalloc((void**)&p, m * sizeof(T));
    if (e == cudaSuccess) n = m;
    return e;
  }
  template <class A>
  cudaError_t upload(const std::vector<T, A>& h, cudaStream_t s) {
    cudaError_t e = resize(h.size());
    if (e != cudaSuccess || h.empty()) return e;
    return cudaMemcpyAsync(p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, s);
  }
  void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
};

struct StatePtrs {
  double *knot_q, *knot_p, *bias_a, *bias_g, *lm, *glb;
};

struct DeviceState {
  DBuf<double> knot_q, knot_p, bias_a, bias_g, lm, glb;
  StatePtrs ptrs() const { return {knot_q.p, knot_p.p, bias_a.p, bias_g.p, lm.p, glb.p}; }
};

// Evaluation outputs (device)
struct EvalBuffers {
  DBuf<double> frame_pose;        // F * sizeof(FramePose<N>)/8
  DBuf<double> r_reproj;          // 2M   (double2 per observation)
  DBuf<double> Jl;                // 6M   planar: Jl[c*M + i]
  DBuf<double> f_sq, f_w, f_cost; // per frame: |r|^2, sqrt(rho'), 1/2 rho
  DBuf<double> r_acc, r_gyr, J_acc, J_gyr;
  DBuf<double> r_gp, J_gp, r_bias, J_bias, r_grav, J_grav, r_prior, J_prior;
  DBuf<double> partial;           // reduction scratch
  DBuf<double> cost;              // 8 doubles
};

struct SolverData;  // ctba_solve.cu
struct AlignData;   // ctba_align.cu

}  // namespace ctba

struct ctba_problem {
  int N = 6;
  int64_t dt_ns = 0, start_t_ns = 0;
  double dt_s = 0, inv_dt = 0;
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  std::string err;

  // ---- host state (bspline_so3r3.h:1936-1973) ----
  ctba::HVec<double> knot_q, knot_p, bias_a, bias_g, lm;
  double glb[ctba::GLB_STATE_SIZE] = {0, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  double scale = 0, t_offset_cam_imu_s = 0, cam_t_offset_cam_imu = 0, t_s_gp_s = 0, t_s_gp_init_s = 0;
  int64_t dt_ns_bias = 1000000000;
  double inv_dt_bias = 1.0;
  bool optimize_p_bp = true, optimize_gravity = false, optimize_scale = false, optimize_camimu_extrinsics = true,
       optimize_t_offset_s_gs = true;
  ctba::Camera cam{4, 1, 1, 0, 0, {0, 0, 0, 0}};
  double q_bc0[4] = {0, 0, 0, 1}, p_bc0[3] = {0, 0, 0};
  double cauchy_a = 3.0;
  bool bench_no_termination = false;
  bool bench_accept_all = false;   // timing only: every valid candidate is taken, so that each iteration does the full work of an accepted step
  int relinearize_mask = 3;  // ctba_relinearize_time_offsets folds: bit 0 the spline-imu offset, bit 1 the spline / global-sensor offset
  int band_partitions = 0; // time partitions of the band factorisation: 0 = automatic, k = force k (1 = plain sequential)
  int band_nested = 0;     // 1: interface systems are reduced recursively (odd-even over the separators); 0: one flat interface system
  bool schur_two_stage = true;   // knot-space expansion of the Schur blocks through U(f, kc) (false: one pass over all frame pairs per knot pair)
  bool schur_sorted = true;    // compact gather with per-observation track bitmasks + warp compaction (k_frame_schur_m); option name "schur_mask"
  double schur_thread_below = 16.0;   // (frames per knot)^2 below which the knot-space expansion runs one thread per (knot, distance) instead of the two-stage gather
  int schur_threads = 192;     // CTA size of k_frame_schur_m (128 / 192 / 256)
  bool schur_list_a = true;    // compact gather: partner blocks fetched from a copy in landmark-list order (one dependent load less per partner)
  bool schur_compact = true;   // landmark Schur gather from the 6-double A blocks (false: from the 18-double Y / Z blocks)
  bool band_graph = true;  // replay the factor+solve launch sequence as a CUDA graph
  int linear_solver = 0;   // 0: bordered band Cholesky (exact step); 1: block-Jacobi PCG on the same reduced system (ctba_pcg.cu)
  double pcg_tol = 1e-10;  // PCG stops at |r| <= pcg_tol |b|
  int pcg_max_iters = 20000;

  // ---- host work items ----
  ctba::HVec<int32_t> f_s, f_obs_begin{0};
  ctba::HVec<double> f_u, f_inv_std;
  ctba::HVec<double> o_uv;
  ctba::HVec<int32_t> o_lm, o_frame;
  ctba::HVec<int32_t> i_s, i_sb;
  ctba::HVec<double> i_st_s, i_k, i_ub, i_inv_a, i_inv_g, i_acc, i_gyr;
  ctba::HVec<int32_t> g_s;
  ctba::HVec<double> g_st_s, g_k, g_inv_std, g_meas;
  // every IMU / global-position measurement the caller passed, skipped ones included: evaluate*Errors re-derives the windows from the CURRENT
  // (folded) time offsets and re-applies the skip rules, like the reference (bspline_so3r3.h:1239-1256, 1692-1707)
  std::vector<int64_t> raw_imu_t, raw_gp_t;
  std::vector<double> raw_imu_acc, raw_imu_gyr, raw_gp_pos;
  double raw_imu_inv_a = 1.0, raw_imu_inv_g = 1.0, raw_gp_inv = 1.0;
  ctba::HVec<int32_t> b_s0, b_s1, b_which;
  ctba::HVec<double> b_u0, b_u1, b_inv_std;
  ctba::HVec<int32_t> q_s, q_which;             // pose priors (addPositionMeasurement / addOrientationMeasurement): window, 0 position | 1 orientation
  ctba::HVec<double> q_u, q_inv_std, q_meas;    // u, 1 / noise, measurement (4 per prior: x y z - | qx qy qz qw)
  bool gravity_norm = false;
  // world alignment of a frozen spline (addWorldColmapAlignmentMeasurement, bspline_so3r3.h:1121-1211; members :1966-1970)
  ctba::HVec<int32_t> al_s;
  ctba::HVec<double> al_st_s, al_k, al_inv_std, al_meas;
  double T_wg[7] = {0, 0, 0, 1, 0, 0, 0}, R_wg_corr[4] = {0, 0, 0, 1}, p_wg_corr[3] = {0, 0, 0};
  double align_cauchy_a = 0.10;          // CauchyLoss(0.10), :1189
  ctba::AlignData* align = nullptr;      // device side (ctba_align.cu)
  bool align_dirty = true;

  // ---- device ----
  bool meas_dirty = true, state_dirty = true, host_state_stale = false, struct_dirty = true;
  ctba::DeviceState dstate;
  ctba::DBuf<int32_t> d_f_s, d_f_obs_begin, d_o_lm, d_o_frame, d_i_s, d_i_sb, d_g_s, d_b_s0, d_b_s1, d_b_which;
  ctba::DBuf<int32_t> d_q_s, d_q_which;
  ctba::DBuf<double> d_q_u, d_q_inv_std, d_q_meas;
  ctba::DBuf<double> d_f_u, d_f_inv_std, d_o_uv, d_i_st_s, d_i_k, d_i_ub, d_i_inv_a, d_i_inv_g, d_i_acc, d_i_gyr, d_g_st_s, d_g_k,
      d_g_inv_std, d_g_meas, d_b_u0, d_b_u1, d_b_inv_std;
  ctba::EvalBuffers ev;      // evaluation at the current state (Jacobians used by the assembly kernels)
  ctba::EvalBuffers ev_alt;  // evaluation at the LM candidate; swapped with `ev` when the step is accepted
  ctba::SolverData* solver = nullptr;
  int64_t launches = 0;  // kernels launched by this handle (bench bookkeeping)

  // ---- multi-GPU: time-segment sharding (ctba_shard.h) ----
  ctba::Comm* comm = nullptr;            // set by ctba_comm_init_*; world > 1 => the next solve shards the problem
  bool comm_owned = true;
  bool sharded = false;                  // the host work items above hold only this rank's units
  ctba::ShardInfo* shard = nullptr;
  ctba::GlobalStructure* gstruct = nullptr;
  ctba::DBuf<double> d_fsq_glob;         // per GLOBAL frame: squared reprojection norm (all-reduced)
  ctba::DBuf<double> d_gather;           // all-gather landing zone: world * 8 doubles
  ctba::DBuf<double> d_stage;            // masked state for the final re-assembly

  size_t numKnots() const { return knot_q.size() / 4; }
  size_t numBiasKnots() const { return bias_a.size() / 3; }
  size_t numFrames() const { return f_s.size(); }
  size_t numObs() const { return o_lm.size(); }
};

namespace ctba {
Globals make_globals(const ctba_problem* p);
int sync_to_device(ctba_problem* p);                 // upload dirty measurements/state
int download_state(ctba_problem* p);                 // device state -> host vectors
// residual (+Jacobian) evaluation of every term at state S into p->ev; cost7 (device) filled when want_cost
int launch_evaluate(ctba_problem* p, EvalBuffers& ev, const StatePtrs& S, bool jac);
int bench_reproj(ctba_problem* p, int reps, int flush_l2, double* ms_avg);
int launch_cost(ctba_problem* p, EvalBuffers& ev);   // ev.cost[0..6] from the residual buffers
int expand_reproj_dense(ctba_problem* p, const StatePtrs& S, double* d_out);  // 2M x (6N+10)
void solver_free(SolverData* s);
int solver_solve(ctba_problem* p, int max_iters, ctba_summary* out);
int solver_num_unknowns(ctba_problem* p);
// ctba_align.cu: the frozen-spline alignment problem (handles that hold alignment measurements)
int align_solve(ctba_problem* p, int max_iters, ctba_summary* out);
int align_evaluate(ctba_problem* p, double* r_out, double* j_out, double* err_out);   // 3n residuals (x 1/noise), 3n x 11 Jacobians, n unit-weight norms
void align_free(AlignData* a);
// evaluate*Errors: swap in IMU / GP work items rebuilt from the raw measurements with the current time offsets (begin), swap the solver's
// work items back (end).  The solver structure is not touched.
struct MeasBackup {
  HVec<int32_t> i_s, i_sb, g_s;
  HVec<double> i_st_s, i_k, i_ub, i_inv_a, i_inv_g, i_acc, i_gyr, g_st_s, g_k, g_inv_std, g_meas;
  bool active = false;
};
int remap_measurements_begin(ctba_problem* p, MeasBackup* b);
void remap_measurements_end(ctba_problem* p, MeasBackup* b);
int shard_build(ctba_problem* p);                    // ctba_shard.cu: filter the host work items to this rank's units
int shard_reduce_cost(ctba_problem* p, EvalBuffers& ev, double* host_cost8);   // cost terms of all ranks, summed in rank order
int shard_gather_state(ctba_problem* p);             // owners' knots / biases / landmarks -> every rank
}  // namespace ctba

#define CTBA_CUDA(p, call)                                                                                   \
  do {                                                                                                       \
    cudaError_t e__ = (call);                                                                                \
    if (e__ != cudaSuccess) {                                                                                \
      (p)->err = std::string(#call) + ": " + cudaGetErrorString(e__);                                        \
      return CTBA_E_CUDA;                                                                                    \
    }                                                                                                        \
  } while (0)
