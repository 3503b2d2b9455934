// ctba_solver.h — data layout of the reduced (landmark-eliminated) normal equations on the device.
//
// Unknown ordering of the reduced system ("nodes"): spline knots and bias knots merged in time order,
// 6 scalars per node (knot: rot 3 + pos 3; bias node: acc 3 + gyro 3)  -> block-banded matrix;
// the <= 15 global parameters form a dense border; landmarks are eliminated per landmark (3x3 Schur).
//   band    : lower block band, 32x32 scalar blocks, block (ib,jb), 0 <= ib-jb <= nbw, at ((jb*(nbw+1)+(ib-jb))*1024 + (r%32)*32 + (c%32))
//   border  : 16 rows (15 globals + the right-hand side as a 16th row) x n_pad, row-major
//   corner  : 16 x 16 (lower part used)
// This replaces Ceres' SPARSE_NORMAL_CHOLESKY on the full Jacobian (bspline_so3r3.h:1771); the step it
// produces is the same exact LM step.
#pragma once
#include "ctba_internal.h"

namespace ctba {

constexpr int kNB = 32;          // scalar block size of the band
constexpr int kBorder = 16;      // border rows: GT_SIZE globals + rhs
constexpr int kRhsRow = 15;

struct BandLayout {
  int n_band;   // 6 * nodes
  int n_pad;    // multiple of 32
  int nbc;      // block columns
  int nbw;      // block sub-diagonals kept
  double* band; // device
  double* border;
  double* corner;
  __host__ __device__ size_t band_elems() const { return (size_t)nbc * (nbw + 1) * kNB * kNB; }
};

__device__ __forceinline__ size_t band_addr(const BandLayout& B, int r, int c) {
  const int jb = c >> 5, ib = r >> 5;
  return (((size_t)jb * (B.nbw + 1) + (ib - jb)) << 10) + ((r & 31) << 5) + (c & 31);
}

// symmetric accumulate of entry (i,j), unified index space: [0,n_pad) band scalars, n_pad + t globals
__device__ __forceinline__ void add_sym(const BandLayout& B, int i, int j, double v) {
  int r = i > j ? i : j, c = i > j ? j : i;
  if (r < B.n_pad) atomicAdd(B.band + band_addr(B, r, c), v);
  else if (c < B.n_pad) atomicAdd(B.border + (size_t)(r - B.n_pad) * B.n_pad + c, v);
  else atomicAdd(B.corner + (r - B.n_pad) * kBorder + (c - B.n_pad), v);
}

// host analysis of the problem structure (assemble_structure; also used on the whole problem before sharding)
struct HostStructure {
  int K = 0, Kb = 0, L = 0, F = 0, M = 0, nodes = 0, hbn = 0, T = 0, KD = 0, max_frame_obs = 1, n_band = 0, n_pad = 0, nbw = 0;
  bool has_imu = false;
  std::vector<int32_t> node_of_knot, node_of_bias, lob, lobs, opos, oend, imu_seg, gp_seg, prior_seg, lmf, kfb, kfl, lm_smin;
  std::vector<unsigned long long> omask;   // per observation: frames ahead in which its landmark is seen again
  std::vector<uint8_t> tk, tb, tl;
  int gactive[GT_SIZE] = {0};
};
void analyse_structure(const ctba_problem* p, const GlobalStructure* gs, HostStructure* hs);

struct SolverData {
  // ---- structure ----
  int K = 0, Kb = 0, nodes = 0, hbn = 0, L = 0, F = 0, T = 0, KD = 0;   // KD: largest knot distance coupled through a landmark
  bool has_bias = false;
  std::vector<int32_t> node_of_knot, node_of_bias;
  DBuf<int32_t> d_node_of_knot, d_node_of_bias, d_lm_obs_begin, d_lm_obs;
  DBuf<int32_t> d_lm_frame;   // frame of every entry of the landmark observation lists
  DBuf<int32_t> d_lm_owner;   // sharded: owning rank of every landmark
  int max_frame_obs = 1;
  double frames_per_knot = 0.0;   // average number of frames whose window contains a knot (over the knots that have any)
  DBuf<int32_t> d_imu_seg, d_gp_seg, d_prior_seg;   // runs of samples with the same unknowns (CSR offsets)
  int n_imu_seg = 0, n_gp_seg = 0, n_prior_seg = 0;
  DBuf<int32_t> d_o_pos, d_o_end, d_kf_begin, d_kf_list;
  DBuf<unsigned long long> d_o_mask;   // track bitmask per observation (k_frame_schur_m)   // observation -> slot in its landmark's list; knot -> frames (CSR)
  DBuf<double> Vi, vig, ViYg, Yo, Zo, lm_corner_partial;   // Schur gather scratch (ctba_assemble.cu)
  DBuf<uint8_t> d_touch_knot, d_touch_bias, d_touch_lm;  // bit0 rot/acc, bit1 pos/gyr
  int gactive[GT_SIZE] = {0};
  BandLayout H0{}, Hw{};
  DBuf<double> band0, bandw, border0, borderw, corner0, cornerw;
  DBuf<double> grad;          // n_pad + 16 : J^T r of the band + global unknowns (no Schur); sharded: this rank's partial sum
  DBuf<double> tot;           // [gradient | diag(J^T J)] (n_pad + 16 each) of the WHOLE problem (sharded: all-reduced)
  int bc_lo = 0, bc_hi = 0;   // block columns this handle's partial system touches
  int own_lo = 0, own_hi = 0; // scalar columns it owns (damping, step scalars)
  bool is_root = true;        // owns the global parameters
  DBuf<double> V0, gl;        // per landmark: 6 (xx xy xz yy yz zz), 3
  DBuf<double> Sp, Sb, rp;    // pose-space Schur blocks: F*(T+1)*36, F*42, F*6
  DBuf<double> Uk;            // half-expanded blocks U(f, kc) = sum_f2 S(f, f2) D_f2[kc]: F*(KD+N)*36 (dense frame rates only)
  DBuf<double> jscale, jscale_lm;   // Jacobi scaling (fixed per Solve): n_pad+16, 3L
  DBuf<double> delta, delta_lm;     // step: n_pad+16, 3L
  DBuf<double> dpose;               // per frame 6: pose-space step
  void* band_stream2 = nullptr; void* band_stream3 = nullptr; void* band_evA = nullptr; void* band_evB = nullptr; void* band_evC = nullptr;   // side streams + events of the step launches
  void* band_plan = nullptr;        // BandPlan (ctba_band.cu): time partitions, interface system, scratch
  void* pcg = nullptr;              // PcgData (ctba_pcg.cu)
  int pcg_iterations = 0; double pcg_rel_residual = 0.0;   // of the last PCG solve
  int64_t band_launches = 0;        // kernels in one factor+solve
  void* band_graph_exec = nullptr;  // cudaGraphExec_t of the factor+solve launch sequence (band_factor_solve)
  void* band_graph_exec1 = nullptr; void* band_graph_exec2 = nullptr;   // sharded: the halves before / after the interface all-reduce
  int64_t band_launches1 = 0, band_launches2 = 0;
  int band_graph_key = -1;
  DBuf<double> scal;                // small scalar outputs (see SC_* in ctba_solve.cu)
  DBuf<int32_t> flag;               // numeric failure flag
  DeviceState cand;
  bool built = false;
};

// assembly kernels (ctba_assemble.cu)
int assemble_structure(ctba_problem* p);                        // host: nodes, bandwidth, landmark CSR, buffers
int assemble_direct(ctba_problem* p);                           // H0, grad, V0, gl from the evaluation buffers
int assemble_schur(ctba_problem* p, double radius);             // Hw = H0 + D^2 - Schur ; rhs row ; uses jscale
int compute_jscale(ctba_problem* p);
int backsub_and_candidate(ctba_problem* p, const StatePtrs& S, const StatePtrs& C, double radius);
int step_scalars(ctba_problem* p, const StatePtrs& S, const StatePtrs& C, double radius);
int band_factor_solve(ctba_problem* p);                          // ctba_band.cu
int band_pcg_solve(ctba_problem* p);                             // ctba_pcg.cu (option linear_solver = 1)
void pcg_release(SolverData* sd);
void band_graph_release(SolverData* sd);
void band_plan_release(SolverData* sd);

enum { SC_MODEL = 0, SC_STEP2 = 1, SC_XNORM2 = 2, SC_GMAX = 3, SC_COUNT = 8 };

}  // namespace ctba
