// ctba_assemble.cu — normal-equation assembly on the device (sm_100a, FP64).
//
// The reprojection Jacobian of an observation factors as J_o = Jl_o * [ I3 | X_o ], Jl_o = dr/dp_w (2x3, stored by
// k_reproj) and X_o (3x13) = d p_w-equivalent / d(frame pose 6, reprojection globals 7), computed on the fly from the
// frame pose record and the landmark.  With A_o = w_f^2 Jl^T Jl (3x3), b_o = w_f^2 Jl^T r:
//   landmark block   V_l = sum_o A_o,             g_l = sum_o b_o
//   pose-space block Q_f = sum_o X_o^T A_o X_o,   q_f = sum_o X_o^T b_o          (k_frame_direct -> knot space via the chain D_f)
//   Schur            S(f,f') = - sum_l (A_o X_o)^T (V_l + D_l)^-1 (A_o' X_o')        (k_lm_prepare, k_obs_yz, k_frame_schur: gather, no atomics)
//   knot space       H(k, k') += sum_{f ni k, f' ni k'} D_f[k]^T S(f,f') D_f'[k']   (k_knot_schur, k_knot_border: one writer per entry)
// IMU / GP / bias / gravity rows are dense local Jacobians and go straight into the band (k_rows_accumulate).
// Replaces Ceres' J^T J formation + elimination inside SPARSE_NORMAL_CHOLESKY; LossFunction corrector
// (CauchyLoss(3.0) per frame, bspline_so3r3.h:1032) enters as the per-frame weight w_f = sqrt(rho').
#include <algorithm>
#include "ctba_solver.h"

namespace ctba {

static inline int cdiv(size_t a, int b) { return (int)((a + b - 1) / b); }

struct GMask { int a[GT_SIZE]; };

// X_o (3x13): columns [phi_world 3 | p_wb 3 | scale 1 | R_bc_corr 3 | p_bc_corr 3]
template <int N>
__device__ __forceinline__ void obs_X(const Globals& G, const double* glb, const FramePose<N>& fp, v3 pw, double X[3][13]) {
  v3 pwb{fp.p_wb[0], fp.p_wb[1], fp.p_wb[2]};
  v3 d = pw - pwb;
  m3 hd = hat(d);
  m3 Rwb, Rcw;
#pragma unroll
  for (int e = 0; e < 9; ++e) { Rwb.a[e] = fp.R_wb[e]; Rcw.a[e] = fp.R_cw[e]; }
  m3 Rwc = transpose(Rcw);
  const double s = G.scale + glb[GLB_SCALE_CORR];
  m3 R0 = quat_matrix(quat{G.q_bc0[0], G.q_bc0[1], G.q_bc0[2], G.q_bc0[3]});
  v3 pbc = mul(R0, v3{glb[GLB_PBC_CORR], glb[GLB_PBC_CORR + 1], glb[GLB_PBC_CORR + 2]}) + v3{G.p_bc0[0], G.p_bc0[1], G.p_bc0[2]};
  v3 Rp = mul(Rwb, pbc);
  v3 pwc = s * Rp + pwb;
  m3 hc = mul(hat(pw - pwc), Rwc);
  m3 RR0 = mul(Rwb, R0);
  const double rp[3] = {Rp.x, Rp.y, Rp.z};
#pragma unroll
  for (int r = 0; r < 3; ++r) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      X[r][c] = hd.a[3 * r + c];
      X[r][3 + c] = (r == c) ? -1.0 : 0.0;
      X[r][7 + c] = hc.a[3 * r + c];
      X[r][10 + c] = -s * RR0.a[3 * r + c];
    }
    X[r][6] = -rp[r];
  }
}

// per observation: A (3x3 sym as 6: xx xy xz yy yz zz) and b (3), both including the frame's robust weight
__device__ __forceinline__ void obs_Ab(const double* __restrict__ Jl, const double2* __restrict__ r, const double* __restrict__ f_w, int M, int o,
                                       int f, double A[6], double b[3]) {
  double j[6];
#pragma unroll
  for (int c = 0; c < 6; ++c) j[c] = Jl[(size_t)c * M + o];
  const double w = f_w[f], w2 = w * w;
  const double2 rr = r[o];
  A[0] = w2 * (j[0] * j[0] + j[3] * j[3]); A[1] = w2 * (j[0] * j[1] + j[3] * j[4]); A[2] = w2 * (j[0] * j[2] + j[3] * j[5]);
  A[3] = w2 * (j[1] * j[1] + j[4] * j[4]); A[4] = w2 * (j[1] * j[2] + j[4] * j[5]); A[5] = w2 * (j[2] * j[2] + j[5] * j[5]);
  b[0] = w2 * (j[0] * rr.x + j[3] * rr.y); b[1] = w2 * (j[1] * rr.x + j[4] * rr.y); b[2] = w2 * (j[2] * rr.x + j[5] * rr.y);
}
__device__ __forceinline__ void symmul3(const double A[6], const double v[3], double out[3]) {
  out[0] = A[0] * v[0] + A[1] * v[1] + A[2] * v[2];
  out[1] = A[1] * v[0] + A[3] * v[1] + A[4] * v[2];
  out[2] = A[2] * v[0] + A[4] * v[1] + A[5] * v[2];
}

// ---- landmark blocks: V0_l, g_l (thread per landmark; obs lists are short) ------------------------------------
__global__ void k_landmark_blocks(int L, const int32_t* __restrict__ lm_obs_begin, const int32_t* __restrict__ lm_obs,
                                  const int32_t* __restrict__ o_frame, const double* __restrict__ Jl, const double2* __restrict__ r,
                                  const double* __restrict__ f_w, int M, double* __restrict__ V0, double* __restrict__ gl) {
  int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= L) return;
  double V[6] = {0, 0, 0, 0, 0, 0}, g[3] = {0, 0, 0};
  for (int i = lm_obs_begin[l]; i < lm_obs_begin[l + 1]; ++i) {
    const int o = lm_obs[i];
    double A[6], b[3];
    obs_Ab(Jl, r, f_w, M, o, o_frame[o], A, b);
#pragma unroll
    for (int c = 0; c < 6; ++c) V[c] += A[c];
#pragma unroll
    for (int c = 0; c < 3; ++c) g[c] += b[c];
  }
#pragma unroll
  for (int c = 0; c < 6; ++c) V0[6 * (size_t)l + c] = V[c];
#pragma unroll
  for (int c = 0; c < 3; ++c) gl[3 * (size_t)l + c] = g[c];
}

// ---- pose chain helpers ---------------------------------------------------------------------------------------
// D_f maps the 6N knot tangents of frame f's window to its pose tangent (phi_world, p_wb):
//   column a = 6*i + ca (knot i, component ca): ca<3 -> phi rows get Crot_i[:,ca]; ca>=3 -> p row (ca-3) gets cpos_i.
template <int N>
__device__ __forceinline__ double chain_apply_T(const FramePose<N>& fp, int a, const double* v6) {  // (D_f^T v)[a]
  const int i = a / 6, ca = a % 6;
  if (ca < 3) return fp.Crot[i][ca] * v6[0] + fp.Crot[i][3 + ca] * v6[1] + fp.Crot[i][6 + ca] * v6[2];
  return fp.cpos[i] * v6[ca];
}

// ---- direct (non-Schur) reprojection part: warp per frame ------------------------------------------------------
template <int N>
__global__ void __launch_bounds__(128) k_frame_direct(const __grid_constant__ Globals G, const double* __restrict__ glb,
                                                      const FramePose<N>* __restrict__ fps, const double* __restrict__ lm, int F,
                                                      const int32_t* __restrict__ f_s, const int32_t* __restrict__ obs_begin,
                                                      const int32_t* __restrict__ o_lm, const double* __restrict__ Jl,
                                                      const double2* __restrict__ r, const double* __restrict__ f_w, int M,
                                                      const int32_t* __restrict__ node_of_knot, GMask gm, BandLayout H, double* __restrict__ grad) {
  __shared__ double shQ[4][13 * 13 + 13];
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int f = blockIdx.x * 4 + wib;
  if (f >= F) return;
  const FramePose<N>& fp = fps[f];
  double Q[91], q[13];
#pragma unroll
  for (int e = 0; e < 91; ++e) Q[e] = 0.0;
#pragma unroll
  for (int e = 0; e < 13; ++e) q[e] = 0.0;
  for (int o = obs_begin[f] + lane; o < obs_begin[f + 1]; o += 32) {
    const int l = o_lm[o];
    v3 pw{lm[3 * (size_t)l], lm[3 * (size_t)l + 1], lm[3 * (size_t)l + 2]};
    double X[3][13], A[6], b[3];
    obs_X<N>(G, glb, fp, pw, X);
    obs_Ab(Jl, r, f_w, M, o, f, A, b);
    double Y[3][13];  // A X
#pragma unroll
    for (int c = 0; c < 13; ++c) {
      const double x[3] = {X[0][c], X[1][c], X[2][c]};
      double y[3];
      symmul3(A, x, y);
      Y[0][c] = y[0]; Y[1][c] = y[1]; Y[2][c] = y[2];
      q[c] += X[0][c] * b[0] + X[1][c] * b[1] + X[2][c] * b[2];
    }
    int e = 0;
#pragma unroll
    for (int a = 0; a < 13; ++a)
#pragma unroll
      for (int c = 0; c <= a; ++c) { Q[e] += X[0][a] * Y[0][c] + X[1][a] * Y[1][c] + X[2][a] * Y[2][c]; ++e; }
  }
#pragma unroll
  for (int e = 0; e < 91; ++e)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) Q[e] += __shfl_xor_sync(0xffffffffu, Q[e], o);
#pragma unroll
  for (int e = 0; e < 13; ++e)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) q[e] += __shfl_xor_sync(0xffffffffu, q[e], o);
  double* sQ = shQ[wib];
  if (lane == 0) {
    int e = 0;
#pragma unroll
    for (int a = 0; a < 13; ++a)
#pragma unroll
      for (int c = 0; c <= a; ++c) { sQ[a * 13 + c] = Q[e]; sQ[c * 13 + a] = Q[e]; ++e; }
#pragma unroll
    for (int c = 0; c < 13; ++c) sQ[169 + c] = q[c];
  }
  __syncwarp();
  // expand to knot space: unknown a in [0, 6N) -> knot window, a in [6N, 6N+7) -> reprojection globals
  const int s = f_s[f];
  constexpr int W = 6 * N + 7;
  auto gidx = [&](int a) -> int {
    if (a < 6 * N) return 6 * node_of_knot[s + a / 6] + a % 6;
    return H.n_pad + (a - 6 * N);  // GT_SCALE..GT_PBC+2 are border rows 0..6
  };
  auto active = [&](int a) -> bool { return a < 6 * N ? true : gm.a[a - 6 * N] != 0; };
  for (int e = lane; e < W * (W + 1) / 2; e += 32) {
    // unrank lower-triangular index
    int a = (int)((sqrt(8.0 * e + 1.0) - 1.0) * 0.5);
    while ((a + 1) * (a + 2) / 2 <= e) ++a;
    while (a * (a + 1) / 2 > e) --a;
    const int c = e - a * (a + 1) / 2;
    if (!active(a) || !active(c)) continue;
    // t[m] = sum_n Q[m][n] Dc[n]  (m over 13 pose/global rows)
    double val;
    if (c < 6 * N) {
      double t6[6];
      const int ic = c / 6, cc = c % 6;
#pragma unroll
      for (int m = 0; m < 6; ++m) {
        if (cc < 3) t6[m] = sQ[m * 13 + 0] * fp.Crot[ic][cc] + sQ[m * 13 + 1] * fp.Crot[ic][3 + cc] + sQ[m * 13 + 2] * fp.Crot[ic][6 + cc];
        else t6[m] = sQ[m * 13 + cc] * fp.cpos[ic];
      }
      if (a < 6 * N) val = chain_apply_T<N>(fp, a, t6);
      else {
        const int ga = 6 + (a - 6 * N);
        if (cc < 3) val = sQ[ga * 13 + 0] * fp.Crot[ic][cc] + sQ[ga * 13 + 1] * fp.Crot[ic][3 + cc] + sQ[ga * 13 + 2] * fp.Crot[ic][6 + cc];
        else val = sQ[ga * 13 + cc] * fp.cpos[ic];
      }
    } else {
      val = sQ[(6 + a - 6 * N) * 13 + (6 + c - 6 * N)];  // both global (a >= c >= 6N)
    }
    add_sym(H, gidx(a), gidx(c), val);
  }
  for (int a = lane; a < W; a += 32) {
    if (!active(a)) continue;
    double v;
    if (a < 6 * N) v = chain_apply_T<N>(fp, a, sQ + 169);
    else v = sQ[169 + 6 + (a - 6 * N)];
    atomicAdd(grad + gidx(a), v);
  }
}

// ---- dense local-Jacobian rows (IMU / GP / bias / gravity) into the band: warp per residual block ----------------
// kind: 1 accel, 2 gyro, 3 gp, 4 bias, 5 gravity.  cols = width of the row.
template <int N>
__global__ void __launch_bounds__(128) k_rows_accumulate(int kind, int n, int width, const double* __restrict__ J, const double* __restrict__ r,
                                                         const int32_t* __restrict__ i_s, const int32_t* __restrict__ i_sb,
                                                         const int32_t* __restrict__ b_s1, const int32_t* __restrict__ b_which,
                                                         const int32_t* __restrict__ node_of_knot, const int32_t* __restrict__ node_of_bias,
                                                         GMask gm, BandLayout H, double* __restrict__ grad) {
  __shared__ double shJ[4][3 * 64];
  __shared__ int shI[4][64];
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int i = blockIdx.x * 4 + wib;
  if (i >= n) return;
  const int nres = kind == 5 ? 1 : 3;
  double* sJ = shJ[wib];
  int* sI = shI[wib];
  for (int c = lane; c < width; c += 32) {
    int idx = -1;
    if (kind == 1) {  // [3N rot | 3N pos | 12 acc-bias | 3 g | 1 t]
      if (c < 3 * N) idx = 6 * node_of_knot[i_s[i] + c / 3] + c % 3;
      else if (c < 6 * N) idx = 6 * node_of_knot[i_s[i] + (c - 3 * N) / 3] + 3 + c % 3;
      else if (c < 6 * N + 12) idx = 6 * node_of_bias[i_sb[i] + (c - 6 * N) / 3] + (c - 6 * N) % 3;
      else if (c < 6 * N + 15) idx = gm.a[GT_G] ? H.n_pad + GT_G + (c - 6 * N - 12) : -1;
      else idx = gm.a[GT_T_CAM_IMU] ? H.n_pad + GT_T_CAM_IMU : -1;
    } else if (kind == 2) {  // [3N rot | 12 gyro-bias | 1 t]
      if (c < 3 * N) idx = 6 * node_of_knot[i_s[i] + c / 3] + c % 3;
      else if (c < 3 * N + 12) idx = 6 * node_of_bias[i_sb[i] + (c - 3 * N) / 3] + 3 + (c - 3 * N) % 3;
      else idx = gm.a[GT_T_CAM_IMU] ? H.n_pad + GT_T_CAM_IMU : -1;
    } else if (kind == 3) {  // [3N rot | 3N pos | p_bp 3 | t 1 | scale 1]
      if (c < 3 * N) idx = 6 * node_of_knot[i_s[i] + c / 3] + c % 3;
      else if (c < 6 * N) idx = 6 * node_of_knot[i_s[i] + (c - 3 * N) / 3] + 3 + c % 3;
      else if (c < 6 * N + 3) idx = gm.a[GT_PBP] ? H.n_pad + GT_PBP + (c - 6 * N) : -1;
      else if (c == 6 * N + 3) idx = gm.a[GT_T_GP] ? H.n_pad + GT_T_GP : -1;
      else idx = gm.a[GT_SCALE] ? H.n_pad + GT_SCALE : -1;
    } else if (kind == 4) {  // [4 knots @s0 | 4 knots @s1], acc (which 0) or gyro (which 1)
      const int off = b_which[i] ? 3 : 0;
      if (c < 12) idx = 6 * node_of_bias[i_s[i] + c / 3] + off + c % 3;
      else idx = 6 * node_of_bias[b_s1[i] + (c - 12) / 3] + off + c % 3;
    } else {  // gravity: g 3
      idx = gm.a[GT_G] ? H.n_pad + GT_G + c : -1;
    }
    sI[c] = idx;
    for (int rr = 0; rr < nres; ++rr) sJ[rr * 64 + c] = J[((size_t)i * nres + rr) * width + c];
  }
  __syncwarp();
  double res[3] = {0, 0, 0};
  for (int rr = 0; rr < nres; ++rr) res[rr] = r[(size_t)i * nres + rr];
  const int npairs = width * (width + 1) / 2;
  for (int e = lane; e < npairs; e += 32) {
    int a = (int)((sqrt(8.0 * e + 1.0) - 1.0) * 0.5);
    while ((a + 1) * (a + 2) / 2 <= e) ++a;
    while (a * (a + 1) / 2 > e) --a;
    const int c = e - a * (a + 1) / 2;
    const int ia = sI[a], ic = sI[c];
    if (ia < 0 || ic < 0) continue;
    double v = 0.0;
    for (int rr = 0; rr < nres; ++rr) v += sJ[rr * 64 + a] * sJ[rr * 64 + c];
    if (v != 0.0) add_sym(H, ia, ic, v);
  }
  for (int c = lane; c < width; c += 32) {
    if (sI[c] < 0) continue;
    double v = 0.0;
    for (int rr = 0; rr < nres; ++rr) v += sJ[rr * 64 + c] * res[rr];
    if (v != 0.0) atomicAdd(grad + sI[c], v);
  }
}

// Same accumulation for the sample kinds that arrive at a high rate (1 accel, 2 gyro, 3 global position): consecutive samples inside one knot
// interval (and one bias-knot interval) address the SAME unknowns, so a CTA sums J^T J and J^T r over such a run of samples in registers
// and issues one atomic per entry and run instead of one per entry and sample (200 Hz IMU, 0.1 s knots: 20 x fewer atomics).
constexpr int kSegThreads = 128, kSegBatch = 8;   // widest row: 6 N + 16 = 52 columns = 13 blocks of 4 -> 91 block pairs <= 128 threads
template <int N>
__global__ void __launch_bounds__(kSegThreads) k_rows_accumulate_seg(int kind, int nseg, const int32_t* __restrict__ seg_begin, int width, int planar_n,
                                                                     const double* __restrict__ J, const double* __restrict__ r,
                                                                     const int32_t* __restrict__ i_s, const int32_t* __restrict__ i_sb,
                                                                     const int32_t* __restrict__ node_of_knot,
                                                                     const int32_t* __restrict__ node_of_bias, GMask gm, BandLayout H,
                                                                     double* __restrict__ grad) {
  __shared__ __align__(32) double sJ[kSegBatch][3 * 64];
  __shared__ double sR[kSegBatch][3];
  __shared__ int sI[64];
  const int seg = blockIdx.x, tid = threadIdx.x;
  if (seg >= nseg) return;
  const int i0 = seg_begin[seg], i1 = seg_begin[seg + 1];
  for (int c = tid; c < width; c += kSegThreads) {
    int idx = -1;
    if (kind == 1) {  // [3N rot | 3N pos | 12 acc-bias | 3 g | 1 t]
      if (c < 3 * N) idx = 6 * node_of_knot[i_s[i0] + c / 3] + c % 3;
      else if (c < 6 * N) idx = 6 * node_of_knot[i_s[i0] + (c - 3 * N) / 3] + 3 + c % 3;
      else if (c < 6 * N + 12) idx = 6 * node_of_bias[i_sb[i0] + (c - 6 * N) / 3] + (c - 6 * N) % 3;
      else if (c < 6 * N + 15) idx = gm.a[GT_G] ? H.n_pad + GT_G + (c - 6 * N - 12) : -1;
      else idx = gm.a[GT_T_CAM_IMU] ? H.n_pad + GT_T_CAM_IMU : -1;
    } else if (kind == 2) {  // [3N rot | 12 gyro-bias | 1 t]
      if (c < 3 * N) idx = 6 * node_of_knot[i_s[i0] + c / 3] + c % 3;
      else if (c < 3 * N + 12) idx = 6 * node_of_bias[i_sb[i0] + (c - 3 * N) / 3] + 3 + (c - 3 * N) % 3;
      else idx = gm.a[GT_T_CAM_IMU] ? H.n_pad + GT_T_CAM_IMU : -1;
    } else if (kind == 3) {  // [3N rot | 3N pos | p_bp 3 | t 1 | scale 1]
      if (c < 3 * N) idx = 6 * node_of_knot[i_s[i0] + c / 3] + c % 3;
      else if (c < 6 * N) idx = 6 * node_of_knot[i_s[i0] + (c - 3 * N) / 3] + 3 + c % 3;
      else if (c < 6 * N + 3) idx = gm.a[GT_PBP] ? H.n_pad + GT_PBP + (c - 6 * N) : -1;
      else if (c == 6 * N + 3) idx = gm.a[GT_T_GP] ? H.n_pad + GT_T_GP : -1;
      else idx = gm.a[GT_SCALE] ? H.n_pad + GT_SCALE : -1;
    } else {  // pose prior: 3N columns over the position knots (i_sb = 0) or the rotation knots (i_sb = 1)
      idx = 6 * node_of_knot[i_s[i0] + c / 3] + (i_sb[i0] ? 0 : 3) + c % 3;
    }
    sI[c] = idx;
  }
  // thread = one 4 x 4 block (ba >= bc) of J^T J: 8 shared-memory loads per 16 FMAs and residual row
  const int wpad = (width + 3) & ~3, nblk = wpad >> 2, nblocks = nblk * (nblk + 1) / 2;
  for (int c = width + tid; c < wpad; c += kSegThreads) sI[c] = -1;
  int ba = 0, bc = 0;
  if (tid < nblocks) {
    ba = (int)((sqrtf(8.0f * tid + 1.0f) - 1.0f) * 0.5f);
    while ((ba + 1) * (ba + 2) / 2 <= tid) ++ba;
    while (ba * (ba + 1) / 2 > tid) --ba;
    bc = tid - ba * (ba + 1) / 2;
  }
  double acc[16];
#pragma unroll
  for (int q = 0; q < 16; ++q) acc[q] = 0.0;
  double gacc = 0.0;
  for (int b0 = i0; b0 < i1; b0 += kSegBatch) {
    const int nb = min(kSegBatch, i1 - b0);
    __syncthreads();
    if (planar_n > 0) {   // J[entry][sample] with planar_n samples (the IMU kernels' layout): sample index fastest, 64-byte pieces per entry
      for (int e = tid; e < nb * 3 * wpad; e += kSegThreads) {
        const int smp = e % nb, rc = e / nb, rr = rc / wpad, c = rc - rr * wpad;
        sJ[smp][rr * 64 + c] = c < width ? J[(size_t)(rr * width + c) * planar_n + (b0 + smp)] : 0.0;
      }
    } else {
      for (int e = tid; e < nb * 3 * wpad; e += kSegThreads) {
        const int smp = e / (3 * wpad), rem = e - smp * 3 * wpad, rr = rem / wpad, c = rem - rr * wpad;
        sJ[smp][rr * 64 + c] = c < width ? J[((size_t)(b0 + smp) * 3 + rr) * width + c] : 0.0;
      }
    }
    if (tid < nb * 3) sR[tid / 3][tid % 3] = r[(size_t)b0 * 3 + tid];
    __syncthreads();
    for (int smp = 0; smp < nb; ++smp) {
      const double* j0 = sJ[smp];
      if (tid < nblocks) {
#pragma unroll
        for (int rr = 0; rr < 3; ++rr) {
          const double4 va = *reinterpret_cast<const double4*>(j0 + rr * 64 + 4 * ba);
          const double4 vc = *reinterpret_cast<const double4*>(j0 + rr * 64 + 4 * bc);
          const double xa[4] = {va.x, va.y, va.z, va.w}, xc[4] = {vc.x, vc.y, vc.z, vc.w};
#pragma unroll
          for (int ia = 0; ia < 4; ++ia)
#pragma unroll
            for (int ic = 0; ic < 4; ++ic) acc[4 * ia + ic] = fma(xa[ia], xc[ic], acc[4 * ia + ic]);
        }
      }
      if (tid < width) gacc += j0[tid] * sR[smp][0] + j0[64 + tid] * sR[smp][1] + j0[128 + tid] * sR[smp][2];
    }
  }
  if (tid < nblocks) {
#pragma unroll
    for (int ia = 0; ia < 4; ++ia)
#pragma unroll
      for (int ic = 0; ic < 4; ++ic) {
        if (ba == bc && ic > ia) continue;
        const int ua = sI[4 * ba + ia], uc = sI[4 * bc + ic];
        if (ua < 0 || uc < 0 || acc[4 * ia + ic] == 0.0) continue;
        add_sym(H, ua, uc, acc[4 * ia + ic]);
      }
  }
  if (tid < width && sI[tid] >= 0 && gacc != 0.0) atomicAdd(grad + sI[tid], gacc);
}

// ---- Schur complement of the landmarks, gather formulation (no atomics, bitwise reproducible) ------------------------------------
//   k_lm_prepare   thread per landmark : damped V_l, Vi = V_l^-1, vig = Vi g_l; with free reprojection globals also Yg_l = sum_o A_o Xg_o,
//                                        ViYg_l, and per-block partials of the global x global corner / right-hand side
//   k_obs_yz       thread per obs      : Y_o = A_o X_o[:, pose 6] (3x6) and Z_o = Vi_l Y_o, AoS 18 doubles each
//   k_frame_schur  CTA per frame f     : Sp(f, df) = - sum over landmarks seen in f and f+df of Y_a^T Z_b   (warp per df, lanes over the
//                                        frame's observations, partner found through the landmark's observation list);
//                                        rp(f) = - sum_a Y_a^T vig, Sb(f) = - sum_a Y_a^T ViYg   (pose x reprojection globals)
//   k_knot_schur   warp per (knot, dk) : 6x6 knot block  += sum_{f ni k, f' ni k-dk} D_f[k]^T Sfull(f,f') D_f'[k-dk], written by one thread
//   k_knot_border  thread per knot dof : border rows and right-hand side  D_f[k]^T Sb(f), D_f[k]^T rp(f)
// Layouts: Sp[(f*(T+1)+df)*36 + 6*m + n] rows pose f, cols pose f+df (df = 0: full symmetric block); Sb[f*42 + 7*m + g]; rp[f*6 + m].
struct LmPrep { double Vi[6], vig[3]; };

__device__ __forceinline__ void landmark_inverse(const double* __restrict__ V0, const double* __restrict__ jscale_lm, double radius, int l, double Vi[6]) {
  double V[6];
#pragma unroll
  for (int c = 0; c < 6; ++c) V[c] = V0[6 * (size_t)l + c];
  const int di[3] = {0, 3, 5};
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const double sc = jscale_lm[3 * (size_t)l + c];
    double dd = V[di[c]] * sc * sc;
    dd = fmin(fmax(dd, 1e-6), 1e32);
    V[di[c]] += dd / (radius * sc * sc);
  }
  const double c00 = V[3] * V[5] - V[4] * V[4], c01 = V[2] * V[4] - V[1] * V[5], c02 = V[1] * V[4] - V[2] * V[3];
  const double det = V[0] * c00 + V[1] * c01 + V[2] * c02, id = 1.0 / det;
  Vi[0] = c00 * id; Vi[1] = c01 * id; Vi[2] = c02 * id;
  Vi[3] = (V[0] * V[5] - V[2] * V[2]) * id; Vi[4] = (V[1] * V[2] - V[0] * V[4]) * id; Vi[5] = (V[0] * V[3] - V[1] * V[1]) * id;
}

constexpr int kCornerVals = 28 + 7;   // lower triangle of the 7x7 global block + 7 right-hand-side entries
template <int N>
__global__ void __launch_bounds__(128) k_lm_prepare(const __grid_constant__ Globals G, const double* __restrict__ glb,
                                                    const FramePose<N>* __restrict__ fps, const double* __restrict__ lm, int L,
                                                    const int32_t* __restrict__ lm_obs_begin, const int32_t* __restrict__ lm_obs,
                                                    const int32_t* __restrict__ o_frame, const double* __restrict__ Jl,
                                                    const double2* __restrict__ r, const double* __restrict__ f_w, int M,
                                                    const double* __restrict__ V0, const double* __restrict__ gl,
                                                    const double* __restrict__ jscale_lm, double radius, GMask gm, bool any_g,
                                                    double* __restrict__ Vi_out /*6L*/, double* __restrict__ vig_out /*3L*/,
                                                    double* __restrict__ ViYg_out /*21L*/, double* __restrict__ corner_partial /*grid x 35*/) {
  __shared__ double red[4][kCornerVals];
  const int l = blockIdx.x * blockDim.x + threadIdx.x;
  double cv[kCornerVals];
#pragma unroll
  for (int e = 0; e < kCornerVals; ++e) cv[e] = 0.0;
  if (l < L && lm_obs_begin[l + 1] > lm_obs_begin[l]) {
    double Vi[6], g[3], vig[3];
    landmark_inverse(V0, jscale_lm, radius, l, Vi);
#pragma unroll
    for (int c = 0; c < 3; ++c) g[c] = gl[3 * (size_t)l + c];
    symmul3(Vi, g, vig);
#pragma unroll
    for (int c = 0; c < 6; ++c) Vi_out[6 * (size_t)l + c] = Vi[c];
#pragma unroll
    for (int c = 0; c < 3; ++c) vig_out[3 * (size_t)l + c] = vig[c];
    if (any_g) {
      v3 pw{lm[3 * (size_t)l], lm[3 * (size_t)l + 1], lm[3 * (size_t)l + 2]};
      double Yg[3][7];
#pragma unroll
      for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int c = 0; c < 7; ++c) Yg[a][c] = 0.0;
      for (int i = lm_obs_begin[l]; i < lm_obs_begin[l + 1]; ++i) {
        const int o = lm_obs[i], f = o_frame[o];
        double X[3][13], A[6], b[3];
        obs_X<N>(G, glb, fps[f], pw, X);
        obs_Ab(Jl, r, f_w, M, o, f, A, b);
#pragma unroll
        for (int c = 0; c < 7; ++c) {
          const double x[3] = {X[0][6 + c], X[1][6 + c], X[2][6 + c]};
          double y[3];
          symmul3(A, x, y);
          Yg[0][c] += y[0]; Yg[1][c] += y[1]; Yg[2][c] += y[2];
        }
      }
      double ViYg[3][7];
#pragma unroll
      for (int c = 0; c < 7; ++c) {
        const double y[3] = {Yg[0][c], Yg[1][c], Yg[2][c]};
        double z[3];
        symmul3(Vi, y, z);
        ViYg[0][c] = gm.a[c] ? z[0] : 0.0; ViYg[1][c] = gm.a[c] ? z[1] : 0.0; ViYg[2][c] = gm.a[c] ? z[2] : 0.0;
#pragma unroll
        for (int a = 0; a < 3; ++a) ViYg_out[21 * (size_t)l + 7 * a + c] = ViYg[a][c];
      }
      int e = 0;
#pragma unroll
      for (int a = 0; a < 7; ++a) {
#pragma unroll
        for (int c = 0; c <= a; ++c) { cv[e] = gm.a[a] ? -(Yg[0][a] * ViYg[0][c] + Yg[1][a] * ViYg[1][c] + Yg[2][a] * ViYg[2][c]) : 0.0; ++e; }
      }
#pragma unroll
      for (int a = 0; a < 7; ++a) cv[28 + a] = gm.a[a] ? -(Yg[0][a] * vig[0] + Yg[1][a] * vig[1] + Yg[2][a] * vig[2]) : 0.0;
    }
  }
  if (!any_g) return;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
#pragma unroll
  for (int e = 0; e < kCornerVals; ++e) {
    double v = cv[e];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[wib][e] = v;
  }
  __syncthreads();
  if (threadIdx.x < kCornerVals) corner_partial[(size_t)blockIdx.x * kCornerVals + threadIdx.x] = (red[0][threadIdx.x] + red[1][threadIdx.x]) + (red[2][threadIdx.x] + red[3][threadIdx.x]);
}

// corner / global right-hand side: deterministic sum of the per-block partials (one CTA)
// one warp (= one CTA) per value: lanes take the per-block partials in a fixed stride, then a fixed shuffle tree (bit-reproducible)
__global__ void __launch_bounds__(32) k_corner_reduce(int nblk, const double* __restrict__ partial, BandLayout Hw, double* __restrict__ rhs_glob) {
  const int e = blockIdx.x, lane = threadIdx.x;
  if (e >= kCornerVals) return;
  double s = 0.0;
  for (int b = lane; b < nblk; b += 32) s += partial[(size_t)b * kCornerVals + e];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane != 0) return;
  if (e < 28) {
    int a = 0;
    while ((a + 1) * (a + 2) / 2 <= e) ++a;
    const int c = e - a * (a + 1) / 2;
    Hw.corner[a * kBorder + c] += s;
  } else {
    rhs_glob[e - 28] += s;
  }
}

template <int N>
__global__ void __launch_bounds__(128) k_obs_yz(const __grid_constant__ Globals G, const double* __restrict__ glb, const FramePose<N>* __restrict__ fps,
                                                const double* __restrict__ lm, int M, const int32_t* __restrict__ o_frame,
                                                const int32_t* __restrict__ o_lm, const double* __restrict__ Jl, const double2* __restrict__ r,
                                                const double* __restrict__ f_w, const double* __restrict__ Vi_all, double* __restrict__ Y_out,
                                                double* __restrict__ Z_out) {
  const int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= M) return;
  const int f = o_frame[o], l = o_lm[o];
  v3 pw{lm[3 * (size_t)l], lm[3 * (size_t)l + 1], lm[3 * (size_t)l + 2]};
  double X[3][13], A[6], b[3], Vi[6];
  obs_X<N>(G, glb, fps[f], pw, X);
  obs_Ab(Jl, r, f_w, M, o, f, A, b);
#pragma unroll
  for (int c = 0; c < 6; ++c) Vi[c] = Vi_all[6 * (size_t)l + c];
  double2* Yo = reinterpret_cast<double2*>(Y_out + 18 * (size_t)o);
  double2* Zo = reinterpret_cast<double2*>(Z_out + 18 * (size_t)o);
  double Y[18], Z[18];
#pragma unroll
  for (int c = 0; c < 6; ++c) {
    const double x[3] = {X[0][c], X[1][c], X[2][c]};
    double y[3], z[3];
    symmul3(A, x, y);
    symmul3(Vi, y, z);
    Y[c] = y[0]; Y[6 + c] = y[1]; Y[12 + c] = y[2];     // Y[a*6 + c]
    Z[c] = z[0]; Z[6 + c] = z[1]; Z[12 + c] = z[2];
  }
#pragma unroll
  for (int e = 0; e < 9; ++e) { Yo[e] = make_double2(Y[2 * e], Y[2 * e + 1]); Zo[e] = make_double2(Z[2 * e], Z[2 * e + 1]); }
}

constexpr int kFsThreads = 256;
__global__ void __launch_bounds__(kFsThreads) k_frame_schur(int F, int T, const int32_t* __restrict__ obs_begin, const int32_t* __restrict__ o_lm,
                                                            const int32_t* __restrict__ o_frame, const int32_t* __restrict__ o_pos,
                                                            const int32_t* __restrict__ lm_obs_begin, const int32_t* __restrict__ lm_obs,
                                                            const int32_t* __restrict__ lm_frame, const double* __restrict__ Yall,
                                                            const double* __restrict__ Zall,
                                                            const double* __restrict__ vig_all, const double* __restrict__ ViYg_all, bool any_g,
                                                            double* __restrict__ Sp, double* __restrict__ Sb, double* __restrict__ rp) {
  __shared__ double red[kFsThreads / 32][48];
  extern __shared__ double sY[];   // the frame's Y blocks: n x 18
  const int f = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b0 = obs_begin[f], n = obs_begin[f + 1] - b0;
  for (int e = threadIdx.x; e < 9 * n; e += kFsThreads)
    reinterpret_cast<double2*>(sY)[e] = reinterpret_cast<const double2*>(Yall + 18 * (size_t)b0)[e];
  __syncthreads();
  // ---- pose right-hand side and pose x globals block: threads over the frame's observations ----
  {
    double acc[48];
#pragma unroll
    for (int e = 0; e < 48; ++e) acc[e] = 0.0;
    for (int i = threadIdx.x; i < n; i += kFsThreads) {
      const int a = b0 + i, l = o_lm[a];
      const double* Y = sY + 18 * i;
      const double v0 = vig_all[3 * (size_t)l], v1 = vig_all[3 * (size_t)l + 1], v2 = vig_all[3 * (size_t)l + 2];
      double y[18];
#pragma unroll
      for (int e = 0; e < 18; ++e) y[e] = Y[e];
#pragma unroll
      for (int m = 0; m < 6; ++m) acc[m] -= y[m] * v0 + y[6 + m] * v1 + y[12 + m] * v2;
      if (any_g) {
        const double* W = ViYg_all + 21 * (size_t)l;
#pragma unroll
        for (int g = 0; g < 7; ++g) {
          const double w0 = W[g], w1 = W[7 + g], w2 = W[14 + g];
#pragma unroll
          for (int m = 0; m < 6; ++m) acc[6 + 7 * m + g] -= y[m] * w0 + y[6 + m] * w1 + y[12 + m] * w2;
        }
      }
    }
    const int nv = any_g ? 48 : 6;
#pragma unroll
    for (int e = 0; e < 48; ++e) {
      if (e < nv) {
        double v = acc[e];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) red[warp][e] = v;
      }
    }
    __syncthreads();
    if (threadIdx.x < nv) {
      double v = 0.0;
#pragma unroll
      for (int w = 0; w < kFsThreads / 32; ++w) v += red[w][threadIdx.x];
      if (threadIdx.x < 6) rp[6 * (size_t)f + threadIdx.x] = v;
      else Sb[42 * (size_t)f + (threadIdx.x - 6)] = v;
    }
    if (!any_g && threadIdx.x < 42) Sb[42 * (size_t)f + threadIdx.x] = 0.0;
  }
  // ---- pose-space blocks: warp per frame distance ----
  for (int df = warp; df <= T; df += kFsThreads / 32) {
    const int f2 = f + df;
    double acc[36];
#pragma unroll
    for (int e = 0; e < 36; ++e) acc[e] = 0.0;
    if (f2 < F) {
      for (int i = lane; i < n; i += 32) {
        const int a = b0 + i, l = o_lm[a];
        const int lend = lm_obs_begin[l + 1];
        int pos = o_pos[a];                       // position of a in the landmark's (frame-ordered) observation list
        // tracks are mostly gap-free: try the slot df entries ahead first, fall back to a scan from a's own slot
        if (df > 0 && pos + df < lend && lm_frame[pos + df] == f2 && lm_frame[pos + df - 1] < f2) pos += df;
        bool loaded = false;
        double y[18];
        for (; pos < lend; ++pos) {
          const int fb = lm_frame[pos];
          if (fb > f2) break;
          if (fb < f2) continue;
          const int b = lm_obs[pos];
          if (!loaded) {
            const double* Y = sY + 18 * i;
#pragma unroll
            for (int e = 0; e < 18; ++e) y[e] = Y[e];
            loaded = true;
          }
          const double2* Zp = reinterpret_cast<const double2*>(Zall + 18 * (size_t)b);
          double z[18];
#pragma unroll
          for (int e = 0; e < 9; ++e) { const double2 v = Zp[e]; z[2 * e] = v.x; z[2 * e + 1] = v.y; }
          const bool dup = (b != a) && df == 0;    // a second observation of the landmark in the same frame: add the transposed product too
#pragma unroll
          for (int m = 0; m < 6; ++m)
#pragma unroll
            for (int c = 0; c < 6; ++c) {
              double v = y[m] * z[c] + y[6 + m] * z[6 + c] + y[12 + m] * z[12 + c];
              if (dup) v += y[c] * z[m] + y[6 + c] * z[6 + m] + y[12 + c] * z[12 + m];
              acc[6 * m + c] += v;
            }
        }
      }
    }
#pragma unroll
    for (int e = 0; e < 36; ++e) {
      double v = acc[e];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      acc[e] = v;
    }
    double* dst = Sp + ((size_t)f * (T + 1) + df) * 36;
#pragma unroll
    for (int e = 0; e < 36; ++e) if (lane == e % 32 && (e < 32 || lane < 4)) dst[e] = -acc[e];
  }
}

// ---- compact gather (default): per observation only A_o = w^2 Jl^T Jl (6 doubles) is stored.  With X_o[:, pose] = [hat(d) | -I],
// d = p_w - p_wb(f):   Y_a^T Z_b = [H_a^T ; -I] M [H_b | -I],  M = A_a Vi A_b  (3x3),  d_b = d_a + p_wb(f) - p_wb(f2)  (same landmark),
// so a partner observation costs 48 B of gathered L2 traffic instead of the 144 B of its Z block, and the frame's own data is
// 15 doubles per observation (A, d, Vi) in shared memory.
// Al (optional): the same blocks in the order of the landmarks' observation lists (slot o_pos[o]), so that k_frame_schur_c can fetch a
// partner's block from the list slot it already knows instead of through the slot's observation index (one dependent load less).
__global__ void k_obs_a(int M, const int32_t* __restrict__ o_frame, const double* __restrict__ Jl, const double2* __restrict__ r,
                        const double* __restrict__ f_w, double* __restrict__ Ao, const int32_t* __restrict__ o_pos, double* __restrict__ Al) {
  const int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= M) return;
  double A[6], b[3];
  obs_Ab(Jl, r, f_w, M, o, o_frame[o], A, b);
  double2* dst = reinterpret_cast<double2*>(Ao + 6 * (size_t)o);
  dst[0] = make_double2(A[0], A[1]); dst[1] = make_double2(A[2], A[3]); dst[2] = make_double2(A[4], A[5]);
  if (Al) {
    double2* dl = reinterpret_cast<double2*>(Al + 6 * (size_t)o_pos[o]);
    dl[0] = make_double2(A[0], A[1]); dl[1] = make_double2(A[2], A[3]); dl[2] = make_double2(A[4], A[5]);
  }
}

__device__ __forceinline__ void sym_mul_sym(const double S[6], const double B[6], double W[9]) {   // W = S * B, both symmetric (xx xy xz yy yz zz)
  const double s[9] = {S[0], S[1], S[2], S[1], S[3], S[4], S[2], S[4], S[5]};
  const double b[9] = {B[0], B[1], B[2], B[1], B[3], B[4], B[2], B[4], B[5]};
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) W[3 * i + j] = s[3 * i] * b[j] + s[3 * i + 1] * b[3 + j] + s[3 * i + 2] * b[6 + j];
}

// one (own observation, partner observation) product accumulated into the 6 x 6 pose block:
//   acc += [H_a^T ; -I] (A_a Vi A_b) [H_b | -I]        with  H^T n = n x d,   (row of M) H = (row of M) x d
__device__ __forceinline__ void schur_pair_m(const double Mm[9], double dax, double day, double daz, double dbx, double dby, double dbz,
                                             double acc[36]);
__device__ __forceinline__ void schur_pair(const double Aa[6], const double Vi[6], const double Ab[6], double dax, double day, double daz,
                                           double dbx, double dby, double dbz, double acc[36]) {
  double W[9], Mm[9];
  sym_mul_sym(Vi, Ab, W);
  {
    const double s[9] = {Aa[0], Aa[1], Aa[2], Aa[1], Aa[3], Aa[4], Aa[2], Aa[4], Aa[5]};
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 3; ++c) Mm[3 * r + c] = s[3 * r] * W[c] + s[3 * r + 1] * W[3 + c] + s[3 * r + 2] * W[6 + c];
  }
  schur_pair_m(Mm, dax, day, daz, dbx, dby, dbz, acc);
}
// the common case: C_a = A_a Vi (3 x 3, precomputed once per own observation) so that M = C_a A_b costs one 3 x 3 product per partner
__device__ __forceinline__ void schur_pair_c(const double Ca[9], const double Ab[6], double dax, double day, double daz, double dbx, double dby,
                                             double dbz, double acc[36]) {
  const double b[9] = {Ab[0], Ab[1], Ab[2], Ab[1], Ab[3], Ab[4], Ab[2], Ab[4], Ab[5]};
  double Mm[9];
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c) Mm[3 * r + c] = Ca[3 * r] * b[c] + Ca[3 * r + 1] * b[3 + c] + Ca[3 * r + 2] * b[6 + c];
  schur_pair_m(Mm, dax, day, daz, dbx, dby, dbz, acc);
}
__device__ __forceinline__ void schur_pair_m(const double Mm[9], double dax, double day, double daz, double dbx, double dby, double dbz,
                                             double acc[36]) {
  double Nn[18];   // [M H_b | M]
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const double m0 = Mm[3 * r], m1 = Mm[3 * r + 1], m2 = Mm[3 * r + 2];
    Nn[6 * r] = m1 * dbz - m2 * dby; Nn[6 * r + 1] = m2 * dbx - m0 * dbz; Nn[6 * r + 2] = m0 * dby - m1 * dbx;
    Nn[6 * r + 3] = m0; Nn[6 * r + 4] = m1; Nn[6 * r + 5] = m2;
  }
#pragma unroll
  for (int c = 0; c < 6; ++c) {
    const double n0 = Nn[c], n1 = Nn[6 + c], n2 = Nn[12 + c];
    if (c < 3) {
      acc[c] += n1 * daz - n2 * day; acc[6 + c] += n2 * dax - n0 * daz; acc[12 + c] += n0 * day - n1 * dax;
      acc[18 + c] -= n0; acc[24 + c] -= n1; acc[30 + c] -= n2;
    } else {
      acc[c] -= n1 * daz - n2 * day; acc[6 + c] -= n2 * dax - n0 * daz; acc[12 + c] -= n0 * day - n1 * dax;
      acc[18 + c] += n0; acc[24 + c] += n1; acc[30 + c] += n2;
    }
  }
}

// CTA per frame.  Shared memory (structure of arrays over the frame's observations): C = A Vi (9), d (3), {slot in the landmark's list,
// end of the list (negative: the list has two observations in one frame)}, frame of the landmark's last observation.
constexpr int kFcThreads = 128;
constexpr int kFcRows = 14;
template <int N, bool LISTA>
__global__ void __launch_bounds__(kFcThreads, 3) k_frame_schur_c(int F, int T, const FramePose<N>* __restrict__ fps, const double* __restrict__ lm,
                                                                 const int32_t* __restrict__ obs_begin, const int32_t* __restrict__ o_lm,
                                                                 const int32_t* __restrict__ o_pos, const int32_t* __restrict__ o_end,
                                                                 const int32_t* __restrict__ lm_obs, const int32_t* __restrict__ lm_frame,
                                                                 const double* __restrict__ Ao, const double* __restrict__ Al,
                                                                 const double* __restrict__ Vi_all,
                                                                 const double* __restrict__ vig_all, const double* __restrict__ ViYg_all, bool any_g,
                                                                 int n_pad, double* __restrict__ Sp, double* __restrict__ Sb, double* __restrict__ rp) {
  __shared__ double red[kFcThreads / 32][48];
  extern __shared__ double sO[];
  const int f = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b0 = obs_begin[f], n = obs_begin[f + 1] - b0;
  const double px = fps[f].p_wb[0], py = fps[f].p_wb[1], pz = fps[f].p_wb[2];
  int2* sPE = reinterpret_cast<int2*>(sO + 12 * (size_t)n_pad);
  int* sLast = reinterpret_cast<int*>(sO + 13 * (size_t)n_pad);
  for (int i = threadIdx.x; i < n; i += kFcThreads) {
    const int a = b0 + i, l = o_lm[a];
    const double2* Ap = reinterpret_cast<const double2*>(Ao + 6 * (size_t)a);
    const double2* Vp = reinterpret_cast<const double2*>(Vi_all + 6 * (size_t)l);
    const double2 a0 = Ap[0], a1 = Ap[1], a2 = Ap[2], v0 = Vp[0], v1 = Vp[1], v2 = Vp[2];
    const int e = o_end[a];
    {
      const double A6[6] = {a0.x, a0.y, a1.x, a1.y, a2.x, a2.y}, V6[6] = {v0.x, v0.y, v1.x, v1.y, v2.x, v2.y};
      double C9[9];
      sym_mul_sym(A6, V6, C9);
#pragma unroll
      for (int c = 0; c < 9; ++c) sO[c * n_pad + i] = C9[c];
    }
    sO[9 * n_pad + i] = lm[3 * (size_t)l] - px; sO[10 * n_pad + i] = lm[3 * (size_t)l + 1] - py; sO[11 * n_pad + i] = lm[3 * (size_t)l + 2] - pz;
    sPE[i] = make_int2(o_pos[a], e);
    sLast[i] = lm_frame[(e < 0 ? -e : e) - 1];
  }
  __syncthreads();
  // ---- pose right-hand side and pose x globals block:  Y_a^T v = [ (A v) x d ; -(A v) ]
  {
    double acc[48];
#pragma unroll
    for (int e = 0; e < 48; ++e) acc[e] = 0.0;
    for (int i = threadIdx.x; i < n; i += kFcThreads) {
      const int l = o_lm[b0 + i];
      const double2* Ap = reinterpret_cast<const double2*>(Ao + 6 * (size_t)(b0 + i));
      const double2 q0 = Ap[0], q1 = Ap[1], q2 = Ap[2];
      const double A[6] = {q0.x, q0.y, q1.x, q1.y, q2.x, q2.y};
      const double dx = sO[9 * n_pad + i], dy = sO[10 * n_pad + i], dz = sO[11 * n_pad + i];
      auto apply = [&](const double v[3], double out[6]) {
        double w[3];
        symmul3(A, v, w);
        out[0] = w[1] * dz - w[2] * dy; out[1] = w[2] * dx - w[0] * dz; out[2] = w[0] * dy - w[1] * dx;
        out[3] = -w[0]; out[4] = -w[1]; out[5] = -w[2];
      };
      const double v[3] = {vig_all[3 * (size_t)l], vig_all[3 * (size_t)l + 1], vig_all[3 * (size_t)l + 2]};
      double o6[6];
      apply(v, o6);
#pragma unroll
      for (int m = 0; m < 6; ++m) acc[m] -= o6[m];
      if (any_g) {
        const double* W = ViYg_all + 21 * (size_t)l;
#pragma unroll
        for (int g = 0; g < 7; ++g) {
          const double vg[3] = {W[g], W[7 + g], W[14 + g]};
          apply(vg, o6);
#pragma unroll
          for (int m = 0; m < 6; ++m) acc[6 + 7 * m + g] -= o6[m];
        }
      }
    }
    const int nv = any_g ? 48 : 6;
#pragma unroll
    for (int e = 0; e < 48; ++e) {
      if (e < nv) {
        double v = acc[e];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) red[warp][e] = v;
      }
    }
    __syncthreads();
    if (threadIdx.x < nv) {
      double v = 0.0;
#pragma unroll
      for (int w = 0; w < kFcThreads / 32; ++w) v += red[w][threadIdx.x];
      if (threadIdx.x < 6) rp[6 * (size_t)f + threadIdx.x] = v;
      else Sb[42 * (size_t)f + (threadIdx.x - 6)] = v;
    }
    if (!any_g && threadIdx.x < 42) Sb[42 * (size_t)f + threadIdx.x] = 0.0;
  }
  // ---- pose-space blocks: warp per frame distance; a lane looks the partner of two observations up at once (independent loads) ----
  for (int df = warp; df <= T; df += kFcThreads / 32) {
    const int f2 = f + df;
    double acc[36];
#pragma unroll
    for (int e = 0; e < 36; ++e) acc[e] = 0.0;
    if (f2 < F) {
      const double ex = px - fps[f2].p_wb[0], ey = py - fps[f2].p_wb[1], ez = pz - fps[f2].p_wb[2];   // d_b = d_a + e
      auto own = [&](int i, double Ca[9], double d[3]) {
#pragma unroll
        for (int c = 0; c < 9; ++c) Ca[c] = sO[c * n_pad + i];
        d[0] = sO[9 * n_pad + i]; d[1] = sO[10 * n_pad + i]; d[2] = sO[11 * n_pad + i];
      };
      // general search (gaps in the track, or two observations of the landmark in one frame)
      auto slow = [&](int i, int pos, int lend) {
        const int a = b0 + i;
        double Aa[6], Vi[6];
        const double d[3] = {sO[9 * n_pad + i], sO[10 * n_pad + i], sO[11 * n_pad + i]};
        {
          const int l = o_lm[a];
#pragma unroll
          for (int c = 0; c < 6; ++c) { Aa[c] = Ao[6 * (size_t)a + c]; Vi[c] = Vi_all[6 * (size_t)l + c]; }
        }
        for (; pos < lend; ++pos) {
          const int fb = lm_frame[pos];
          if (fb > f2) break;
          if (fb < f2) continue;
          const int b = lm_obs[pos];
          const double2* Ap = reinterpret_cast<const double2*>(Ao + 6 * (size_t)b);
          const double2 a0 = Ap[0], a1 = Ap[1], a2 = Ap[2];
          const double Ab[6] = {a0.x, a0.y, a1.x, a1.y, a2.x, a2.y};
          schur_pair(Aa, Vi, Ab, d[0], d[1], d[2], d[0] + ex, d[1] + ey, d[2] + ez, acc);
          if (b != a && df == 0) schur_pair(Ab, Vi, Aa, d[0], d[1], d[2], d[0], d[1], d[2], acc);   // same frame, same landmark: the transposed product too
        }
      };
      for (int i0 = lane; i0 < n; i0 += 64) {
        int bsel[2] = {-1, -1};
        double Ab[2][6];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int i = i0 + 32 * u;
          if (i >= n) continue;
          const int2 pe = sPE[i];
          if (pe.y < 0) { slow(i, pe.x, -pe.y); continue; }
          if (f2 > sLast[i]) continue;
          const int j = min(pe.x + df, pe.y - 1);
          if (LISTA) {   // the block of list slot j travels together with the slot's frame index: ONE L2 round trip per partner
            const int fj = lm_frame[j];
            const double2* Ap = reinterpret_cast<const double2*>(Al + 6 * (size_t)j);
            const double2 a0 = Ap[0], a1 = Ap[1], a2 = Ap[2];
            Ab[u][0] = a0.x; Ab[u][1] = a0.y; Ab[u][2] = a1.x; Ab[u][3] = a1.y; Ab[u][4] = a2.x; Ab[u][5] = a2.y;
            if (fj == f2) bsel[u] = j;
            else slow(i, pe.x + 1, j);
          } else {
            const int fj = lm_frame[j], bj = lm_obs[j];
            if (fj == f2) bsel[u] = bj;              // strictly increasing frames: the only partner
            else slow(i, pe.x + 1, j);               // a gap in the track: the partner, if any, sits before slot j
          }
        }
        if (!LISTA) {
#pragma unroll
          for (int u = 0; u < 2; ++u)
            if (bsel[u] >= 0) {
              const double2* Ap = reinterpret_cast<const double2*>(Ao + 6 * (size_t)bsel[u]);
              const double2 a0 = Ap[0], a1 = Ap[1], a2 = Ap[2];
              Ab[u][0] = a0.x; Ab[u][1] = a0.y; Ab[u][2] = a1.x; Ab[u][3] = a1.y; Ab[u][4] = a2.x; Ab[u][5] = a2.y;
            }
        }
#pragma unroll
        for (int u = 0; u < 2; ++u)
          if (bsel[u] >= 0) {
            double Ca[9], d[3];
            own(i0 + 32 * u, Ca, d);
            schur_pair_c(Ca, Ab[u], d[0], d[1], d[2], d[0] + ex, d[1] + ey, d[2] + ez, acc);
          }
      }
    }
#pragma unroll
    for (int e = 0; e < 36; ++e) {
      double v = acc[e];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      acc[e] = v;
    }
    double* dst = Sp + ((size_t)f * (T + 1) + df) * 36;
#pragma unroll
    for (int e = 0; e < 36; ++e) if (lane == e % 32 && (e < 32 || lane < 4)) dst[e] = -acc[e];
  }
}

// Sum of 36 per-lane values over the 32 lanes of a warp by recursive halving: in every round a lane keeps one half of its values and hands
// the other half to its partner (37 exchanges instead of the 36 x 5 of a butterfly per value).  On return the lane holds the warp totals
// of entries idx0 .. idx0 + nv - 1 (nv in {0, 1, 2}) in a[0], a[1]; every entry ends on exactly one lane.  Fixed order: bit-reproducible.
__device__ __forceinline__ void warp_sum36(double (&a)[36], int lane, int& idx0, int& nv) {
  const unsigned full = 0xffffffffu;
  int base = 0;
  {
    const bool up = lane & 16;
#pragma unroll
    for (int k = 0; k < 18; ++k) {
      const double keep = up ? a[k + 18] : a[k], send = up ? a[k] : a[k + 18];
      a[k] = keep + __shfl_xor_sync(full, send, 16);
    }
    base += up ? 18 : 0;
  }
  {
    const bool up = lane & 8;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const double keep = up ? a[k + 9] : a[k], send = up ? a[k] : a[k + 9];
      a[k] = keep + __shfl_xor_sync(full, send, 8);
    }
    base += up ? 9 : 0;
  }
  int n;
  {   // 9 -> 5 | 4
    const bool up = lane & 4;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const double keep = up ? a[k + 5] : a[k], send = up ? a[k] : a[k + 5];
      a[k] = keep + __shfl_xor_sync(full, send, 4);
    }
    const double r4 = __shfl_xor_sync(full, a[4], 4);
    a[4] = up ? 0.0 : a[4] + r4;
    base += up ? 5 : 0; n = up ? 4 : 5;
  }
  {   // 5 (or 4) -> 3 | 2 (or 1)
    const bool up = lane & 2;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const double keep = up ? a[k + 3] : a[k], send = up ? a[k] : a[k + 3];
      a[k] = keep + __shfl_xor_sync(full, send, 2);
    }
    const double r2 = __shfl_xor_sync(full, a[2], 2);
    a[2] = up ? 0.0 : a[2] + r2;
    base += up ? 3 : 0; n = up ? n - 3 : 3;
  }
  {   // 3 (or 2, 1) -> 2 | 1 (or 0)
    const bool up = lane & 1;
    const double keep = up ? a[2] : a[0], send = up ? a[0] : a[2];
    a[0] = keep + __shfl_xor_sync(full, send, 1);
    const double r1 = __shfl_xor_sync(full, a[1], 1);
    a[1] = up ? 0.0 : a[1] + r1;
    base += up ? 2 : 0; n = up ? n - 2 : (n < 2 ? n : 2);
  }
  idx0 = base; nv = n < 0 ? 0 : n;
}

// ---- the same gather with a TRACK BITMASK per observation and warp-level compaction (default) -------------------------------------
// k_frame_schur_c walks all n observations of the frame for every frame distance df; a lane whose landmark is not seen in frame f + df
// idles, and a landmark whose track has gaps sends its lane into a serial search of the observation list (ncu, config 3: 17.8 of 32
// threads active per instruction, long-scoreboard stalls 3.6 per issue, 5.4 ms).  Here every observation carries a 64-bit mask, bit d set
// iff its landmark is seen again d frames later (bit 0 = itself; host-built, fixed per structure), so that
//     partner slot of (observation, df) = own slot + popc(mask & (2 << df) - 1) - 1
// needs no look-up at all, and the partner's block comes from the copy in landmark-list order (Al): ONE L2 round trip per pair, no
// search.  Each warp owns a frame distance, scans the frame's masks 32 at a time and queues the observations that have a partner
// (ballot + popc into a warp-private shared-memory queue); whenever 32 are queued they are processed by a full warp.  Landmarks seen twice
// in one frame (o_end < 0) keep the general search of k_frame_schur_c.  Fixed summation order: bit-reproducible.
template <int N, int kFsmThreads>
__global__ void __launch_bounds__(kFsmThreads, 2) k_frame_schur_m(int F, int T, const FramePose<N>* __restrict__ fps, const double* __restrict__ lm,
                                                                  const int32_t* __restrict__ obs_begin, const int32_t* __restrict__ o_lm,
                                                                  const unsigned long long* __restrict__ o_mask, const int32_t* __restrict__ o_pos,
                                                                  const int32_t* __restrict__ o_end, const int32_t* __restrict__ lm_obs,
                                                                  const int32_t* __restrict__ lm_frame, const double* __restrict__ Ao,
                                                                  const double* __restrict__ Al, const double* __restrict__ Vi_all,
                                                                  const double* __restrict__ vig_all, const double* __restrict__ ViYg_all, bool any_g,
                                                                  int n_pad, double* __restrict__ Sp, double* __restrict__ Sb, double* __restrict__ rp) {
  constexpr int kWarps = kFsmThreads / 32;
  __shared__ double red[kWarps][48];
  __shared__ int queue[kWarps][64];
  __shared__ int s_dup;
  extern __shared__ double sO[];
  const int f = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b0 = obs_begin[f], n = obs_begin[f + 1] - b0;
  const double px = fps[f].p_wb[0], py = fps[f].p_wb[1], pz = fps[f].p_wb[2];
  unsigned long long* sMask = reinterpret_cast<unsigned long long*>(sO + 12 * (size_t)n_pad);   // 0 for landmarks seen twice in one frame
  int* sPos = reinterpret_cast<int*>(sO + 13 * (size_t)n_pad);
  if (threadIdx.x == 0) s_dup = 0;
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += kFsmThreads) {
    const int a = b0 + i, l = o_lm[a];
    const double2* Ap = reinterpret_cast<const double2*>(Ao + 6 * (size_t)a);
    const double2* Vp = reinterpret_cast<const double2*>(Vi_all + 6 * (size_t)l);
    const double2 a0 = Ap[0], a1 = Ap[1], a2 = Ap[2], v0 = Vp[0], v1 = Vp[1], v2 = Vp[2];
    {
      const double A6[6] = {a0.x, a0.y, a1.x, a1.y, a2.x, a2.y}, V6[6] = {v0.x, v0.y, v1.x, v1.y, v2.x, v2.y};
      double C9[9];
      sym_mul_sym(A6, V6, C9);
#pragma unroll
      for (int c = 0; c < 9; ++c) sO[c * n_pad + i] = C9[c];
    }
    sO[9 * n_pad + i] = lm[3 * (size_t)l] - px; sO[10 * n_pad + i] = lm[3 * (size_t)l + 1] - py; sO[11 * n_pad + i] = lm[3 * (size_t)l + 2] - pz;
    const bool dup = o_end[a] < 0;
    sMask[i] = dup ? 0ull : o_mask[a];
    sPos[i] = o_pos[a];
    if (dup) s_dup = 1;
  }
  __syncthreads();
  // ---- pose right-hand side and pose x globals block:  Y_a^T v = [ (A v) x d ; -(A v) ]
  {
    double acc[48];
#pragma unroll
    for (int e = 0; e < 48; ++e) acc[e] = 0.0;
    for (int i = threadIdx.x; i < n; i += kFsmThreads) {
      const int l = o_lm[b0 + i];
      const double2* Ap = reinterpret_cast<const double2*>(Ao + 6 * (size_t)(b0 + i));
      const double2 q0 = Ap[0], q1 = Ap[1], q2 = Ap[2];
      const double A[6] = {q0.x, q0.y, q1.x, q1.y, q2.x, q2.y};
      const double dx = sO[9 * n_pad + i], dy = sO[10 * n_pad + i], dz = sO[11 * n_pad + i];
      auto apply = [&](const double v[3], double out[6]) {
        double w[3];
        symmul3(A, v, w);
        out[0] = w[1] * dz - w[2] * dy; out[1] = w[2] * dx - w[0] * dz; out[2] = w[0] * dy - w[1] * dx;
        out[3] = -w[0]; out[4] = -w[1]; out[5] = -w[2];
      };
      const double v[3] = {vig_all[3 * (size_t)l], vig_all[3 * (size_t)l + 1], vig_all[3 * (size_t)l + 2]};
      double o6[6];
      apply(v, o6);
#pragma unroll
      for (int m = 0; m < 6; ++m) acc[m] -= o6[m];
      if (any_g) {
        const double* W = ViYg_all + 21 * (size_t)l;
#pragma unroll
        for (int g = 0; g < 7; ++g) {
          const double vg[3] = {W[g], W[7 + g], W[14 + g]};
          apply(vg, o6);
#pragma unroll
          for (int m = 0; m < 6; ++m) acc[6 + 7 * m + g] -= o6[m];
        }
      }
    }
    const int nv = any_g ? 48 : 6;
#pragma unroll
    for (int e = 0; e < 48; ++e) {
      if (e < nv) {
        double v = acc[e];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) red[warp][e] = v;
      }
    }
    __syncthreads();
    if (threadIdx.x < nv) {
      double v = 0.0;
#pragma unroll
      for (int w = 0; w < kWarps; ++w) v += red[w][threadIdx.x];
      if (threadIdx.x < 6) rp[6 * (size_t)f + threadIdx.x] = v;
      else Sb[42 * (size_t)f + (threadIdx.x - 6)] = v;
    }
    if (!any_g && threadIdx.x < 42) Sb[42 * (size_t)f + threadIdx.x] = 0.0;
  }
  const bool have_dup = s_dup != 0;
  int* q = queue[warp];
  // ---- pose-space blocks: warp per frame distance ----
  for (int df = warp; df <= T; df += kWarps) {
    const int f2 = f + df;
    double acc[36];
#pragma unroll
    for (int e = 0; e < 36; ++e) acc[e] = 0.0;
    if (f2 < F) {
      const double ex = px - fps[f2].p_wb[0], ey = py - fps[f2].p_wb[1], ez = pz - fps[f2].p_wb[2];   // d_b = d_a + e
      const unsigned long long below = (2ull << df) - 1ull;     // bits 0 .. df
      // One pair per lane is kept PENDING: the partner block of a batch is requested as soon as the batch is full and consumed when the
      // next one is (or at the end), so its L2 round trip overlaps the scan and the arithmetic of the previous batch.
      auto request = [&](int i, double Ab[6]) {                    // partner block of observation i in frame f2 (list order)
        const int slot = sPos[i] + __popcll(sMask[i] & below) - 1;
        const double2* Ap = reinterpret_cast<const double2*>(Al + 6 * (size_t)slot);
        const double2 a0 = Ap[0], a1 = Ap[1], a2 = Ap[2];
        Ab[0] = a0.x; Ab[1] = a0.y; Ab[2] = a1.x; Ab[3] = a1.y; Ab[4] = a2.x; Ab[5] = a2.y;
      };
      auto consume = [&](int i, const double Ab[6]) {
        double Ca[9];
#pragma unroll
        for (int c = 0; c < 9; ++c) Ca[c] = sO[c * n_pad + i];
        const double d0 = sO[9 * n_pad + i], d1 = sO[10 * n_pad + i], d2 = sO[11 * n_pad + i];
        schur_pair_c(Ca, Ab, d0, d1, d2, d0 + ex, d1 + ey, d2 + ez, acc);
      };
      int qn = 0, pend = -1;
      double pA[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
      for (int i0 = 0; i0 < n; i0 += 32) {
        const int i = i0 + lane;
        const bool act = i < n && ((sMask[i] >> df) & 1ull);
        const unsigned bal = __ballot_sync(0xffffffffu, act);
        if (act) q[qn + __popc(bal & ((1u << lane) - 1u))] = i;
        qn += __popc(bal);
        __syncwarp();
        if (qn >= 32) {
          const int mine = q[lane];
          const int carry = (lane + 32 < qn) ? q[lane + 32] : 0;
          __syncwarp();
          if (lane + 32 < qn) q[lane] = carry;
          qn -= 32;
          __syncwarp();
          double nA[6];
          request(mine, nA);
          if (pend >= 0) consume(pend, pA);
          pend = mine;
#pragma unroll
          for (int c = 0; c < 6; ++c) pA[c] = nA[c];
        }
      }
      {
        double nA[6];
        const int last = lane < qn ? q[lane] : -1;
        if (last >= 0) request(last, nA);
        if (pend >= 0) consume(pend, pA);
        if (last >= 0) consume(last, nA);
      }
      __syncwarp();
      if (have_dup) {   // landmarks seen twice in one frame: the general search over the landmark's list
        for (int i = lane; i < n; i += 32) {
          const int a = b0 + i;
          const int e = o_end[a];
          if (e >= 0) continue;
          double Aa[6], Vi[6];
          const double d[3] = {sO[9 * n_pad + i], sO[10 * n_pad + i], sO[11 * n_pad + i]};
          {
            const int l = o_lm[a];
#pragma unroll
            for (int c = 0; c < 6; ++c) { Aa[c] = Ao[6 * (size_t)a + c]; Vi[c] = Vi_all[6 * (size_t)l + c]; }
          }
          for (int pos = sPos[i]; pos < -e; ++pos) {
            const int fb = lm_frame[pos];
            if (fb > f2) break;
            if (fb < f2) continue;
            const int b = lm_obs[pos];
            const double2* Ap = reinterpret_cast<const double2*>(Al + 6 * (size_t)pos);
            const double2 a0 = Ap[0], a1 = Ap[1], a2 = Ap[2];
            const double Ab[6] = {a0.x, a0.y, a1.x, a1.y, a2.x, a2.y};
            schur_pair(Aa, Vi, Ab, d[0], d[1], d[2], d[0] + ex, d[1] + ey, d[2] + ez, acc);
            if (b != a && df == 0) schur_pair(Ab, Vi, Aa, d[0], d[1], d[2], d[0], d[1], d[2], acc);   // same frame, same landmark: the transposed product too
          }
        }
      }
    }
    int idx0, nv;
    warp_sum36(acc, lane, idx0, nv);
    double* dst = Sp + ((size_t)f * (T + 1) + df) * 36;
    if (nv > 0) dst[idx0] = -acc[0];
    if (nv > 1) dst[idx0 + 1] = -acc[1];
  }
}

// D_f[k]^T S D_f'[k'] for one frame pair; i / i2 = local knot indices inside the two windows
template <int N>
__device__ __forceinline__ void chain_block_acc(const FramePose<N>& fa, int i, const FramePose<N>& fb, int i2, const double* S /*36, [m][n]*/, bool transpose,
                                                double C[36]) {
  const double* Ca = fa.Crot[i];
  const double* Cb = fb.Crot[i2];
  const double ca = fa.cpos[i], cb = fb.cpos[i2];
  double Sm[36];
#pragma unroll
  for (int m = 0; m < 6; ++m)
#pragma unroll
    for (int n = 0; n < 6; ++n) Sm[6 * m + n] = transpose ? S[6 * n + m] : S[6 * m + n];
  // T = S * Db : columns 0..2 = S[:,0:3] * Cb, columns 3..5 = cb * S[:,3:6]
  double Tm[36];
#pragma unroll
  for (int m = 0; m < 6; ++m) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      Tm[6 * m + c] = Sm[6 * m] * Cb[c] + Sm[6 * m + 1] * Cb[3 + c] + Sm[6 * m + 2] * Cb[6 + c];
      Tm[6 * m + 3 + c] = cb * Sm[6 * m + 3 + c];
    }
  }
  // C += Da^T * T : rows 0..2 = Ca^T T[0:3,:], rows 3..5 = ca * T[3:6,:]
#pragma unroll
  for (int a = 0; a < 3; ++a)
#pragma unroll
    for (int n = 0; n < 6; ++n) {
      C[6 * a + n] += Ca[a] * Tm[n] + Ca[3 + a] * Tm[6 + n] + Ca[6 + a] * Tm[12 + n];
      C[6 * (3 + a) + n] += ca * Tm[6 * (3 + a) + n];
    }
}

// Same product with S streamed from memory one row (or, transposed, one column) at a time: only the 6 x 6 accumulator and one row of
// S and of S * Db are live, so the kernel keeps ~110 registers instead of ~210 and twice as many warps are resident.
template <int N>
__device__ __forceinline__ void chain_block_acc_stream(const FramePose<N>& fa, int i, const FramePose<N>& fb, int i2, const double* __restrict__ S,
                                                       bool transpose, double C[36]) {
  const double* Ca = fa.Crot[i];
  const double* Cb = fb.Crot[i2];
  const double ca = fa.cpos[i], cb = fb.cpos[i2];
  double cbv[9];
#pragma unroll
  for (int e = 0; e < 9; ++e) cbv[e] = Cb[e];
#pragma unroll
  for (int m = 0; m < 6; ++m) {
    double sr[6];
    if (!transpose) {
      const double2* p2 = reinterpret_cast<const double2*>(S + 6 * m);
      const double2 v0 = p2[0], v1 = p2[1], v2 = p2[2];
      sr[0] = v0.x; sr[1] = v0.y; sr[2] = v1.x; sr[3] = v1.y; sr[4] = v2.x; sr[5] = v2.y;
    } else {
#pragma unroll
      for (int n = 0; n < 6; ++n) sr[n] = S[6 * n + m];
    }
    double tr[6];   // row m of S * Db
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      tr[c] = sr[0] * cbv[c] + sr[1] * cbv[3 + c] + sr[2] * cbv[6 + c];
      tr[3 + c] = cb * sr[3 + c];
    }
    if (m < 3) {
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        const double w = Ca[3 * m + a];
#pragma unroll
        for (int n = 0; n < 6; ++n) C[6 * a + n] += w * tr[n];
      }
    } else {
#pragma unroll
      for (int n = 0; n < 6; ++n) C[6 * m + n] += ca * tr[n];
    }
  }
}

template <int N>
__global__ void __launch_bounds__(128, 4) k_knot_schur(const FramePose<N>* __restrict__ fps, int K, int KD, int T, const int32_t* __restrict__ f_s,
                                                    const int32_t* __restrict__ kf_begin, const int32_t* __restrict__ kf_list,
                                                    const int32_t* __restrict__ node_of_knot, const double* __restrict__ Sp, BandLayout Hw) {
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (w >= K * (KD + 1)) return;
  const int kr = w / (KD + 1), dk = w % (KD + 1);
  const int kc = kr - dk;
  if (kc < 0) return;
  const int rb = kf_begin[kr], nr = kf_begin[kr + 1] - rb;
  const int cb = kf_begin[kc], nc = kf_begin[kc + 1] - cb;
  if (nr == 0 || nc == 0) return;
  double C[36];
#pragma unroll
  for (int e = 0; e < 36; ++e) C[e] = 0.0;
  bool any = false;
  for (int idx = lane; idx < nr * nc; idx += 32) {
    const int f = kf_list[rb + idx / nc], f2 = kf_list[cb + idx % nc];
    const int d = f2 - f;
    if (d > T || -d > T) continue;
    const double* S = Sp + ((size_t)(d >= 0 ? f : f2) * (T + 1) + (d >= 0 ? d : -d)) * 36;
    chain_block_acc_stream<N>(fps[f], kr - f_s[f], fps[f2], kc - f_s[f2], S, d < 0, C);
    any = true;
  }
  if (!__any_sync(0xffffffffu, any)) return;
#pragma unroll
  for (int e = 0; e < 36; ++e) {
    double v = C[e];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    C[e] = v;
  }
  const int r0 = 6 * node_of_knot[kr], c0 = 6 * node_of_knot[kc];
#pragma unroll
  for (int e = 0; e < 36; ++e) {
    if (lane == e % 32 && (e < 32 || lane < 4)) {
      const int a = e / 6, c = e % 6;
      if (dk > 0 || c <= a) Hw.band[band_addr(Hw, r0 + a, c0 + c)] += C[e];
    }
  }
}

// Two-stage expansion for dense frame rates: sum_{f in F(kr)} sum_{f2 in F(kc)} D_f[kr]^T S(f,f2) D_f2[kc]
//   = sum_{f in F(kr)} D_f[kr]^T U(f, kc),   U(f, kc) = sum_{f2 in F(kc), |f - f2| <= T} S(f,f2) D_f2[kc].
// U(f, kc) is shared by the N row knots of frame f's window, so the frame-pair products are formed once instead of N times
// (config 2: 0.13 instead of 0.8 GFMA).  Column knots of frame f: kc = s_f - KD + j, j in [0, KD + N).
template <int N>
__global__ void __launch_bounds__(128) k_pose_to_knot_cols(const FramePose<N>* __restrict__ fps, int F, int K, int KD, int T,
                                                           const int32_t* __restrict__ f_s, const int32_t* __restrict__ kf_begin,
                                                           const int32_t* __restrict__ kf_list, const double* __restrict__ Sp,
                                                           double* __restrict__ Uk) {
  const int W = KD + N;
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (size_t)F * W) return;
  const int f = (int)(t / W), j = (int)(t % W);
  const int kc = f_s[f] - KD + j;
  double U[36];
#pragma unroll
  for (int e = 0; e < 36; ++e) U[e] = 0.0;
  if (kc >= 0 && kc < K) {
    for (int q = kf_begin[kc]; q < kf_begin[kc + 1]; ++q) {
      const int f2 = kf_list[q];
      const int d = f2 - f;
      if (d > T || -d > T) continue;
      const double* __restrict__ S = Sp + ((size_t)(d >= 0 ? f : f2) * (T + 1) + (d >= 0 ? d : -d)) * 36;
      const FramePose<N>& fb = fps[f2];
      const int i2 = kc - f_s[f2];
      double cbv[9];
#pragma unroll
      for (int e = 0; e < 9; ++e) cbv[e] = fb.Crot[i2][e];
      const double cb = fb.cpos[i2];
#pragma unroll
      for (int m = 0; m < 6; ++m) {
        double sr[6];
        if (d >= 0) {
          const double2* p2 = reinterpret_cast<const double2*>(S + 6 * m);
          const double2 v0 = p2[0], v1 = p2[1], v2 = p2[2];
          sr[0] = v0.x; sr[1] = v0.y; sr[2] = v1.x; sr[3] = v1.y; sr[4] = v2.x; sr[5] = v2.y;
        } else {
#pragma unroll
          for (int n = 0; n < 6; ++n) sr[n] = S[6 * n + m];     // S(f, f2) = S(f2, f)^T
        }
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          U[6 * m + c] += sr[0] * cbv[c] + sr[1] * cbv[3 + c] + sr[2] * cbv[6 + c];
          U[6 * m + 3 + c] += cb * sr[3 + c];
        }
      }
    }
  }
  double2* out = reinterpret_cast<double2*>(Uk + t * 36);
#pragma unroll
  for (int e = 0; e < 18; ++e) out[e] = make_double2(U[2 * e], U[2 * e + 1]);
}

// second stage: thread per (row knot, knot distance); one writer per 6 x 6 block of the band
template <int N>
__global__ void __launch_bounds__(128) k_knot_schur_cols(const FramePose<N>* __restrict__ fps, int K, int KD, const int32_t* __restrict__ f_s,
                                                         const int32_t* __restrict__ kf_begin, const int32_t* __restrict__ kf_list,
                                                         const int32_t* __restrict__ node_of_knot, const double* __restrict__ Uk, BandLayout Hw) {
  const size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= (size_t)K * (KD + 1)) return;
  const int kr = (int)(w / (KD + 1)), dk = (int)(w % (KD + 1));
  const int kc = kr - dk;
  if (kc < 0) return;
  if (kf_begin[kc + 1] == kf_begin[kc]) return;
  const int W = KD + N;
  double C[36];
#pragma unroll
  for (int e = 0; e < 36; ++e) C[e] = 0.0;
  bool any = false;
  for (int q = kf_begin[kr]; q < kf_begin[kr + 1]; ++q) {
    const int f = kf_list[q];
    const int i = kr - f_s[f];
    const int j = kc - (f_s[f] - KD);                 // 0 <= j < W because 0 <= i < N and 0 <= dk <= KD
    const double2* U2 = reinterpret_cast<const double2*>(Uk + ((size_t)f * W + j) * 36);
    const FramePose<N>& fa = fps[f];
    const double ca = fa.cpos[i];
    double cav[9];
#pragma unroll
    for (int e = 0; e < 9; ++e) cav[e] = fa.Crot[i][e];
#pragma unroll
    for (int m = 0; m < 6; ++m) {
      const double2 v0 = U2[3 * m], v1 = U2[3 * m + 1], v2 = U2[3 * m + 2];
      const double ur[6] = {v0.x, v0.y, v1.x, v1.y, v2.x, v2.y};
      if (m < 3) {
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
          for (int n = 0; n < 6; ++n) C[6 * a + n] += cav[3 * m + a] * ur[n];
      } else {
#pragma unroll
        for (int n = 0; n < 6; ++n) C[6 * m + n] += ca * ur[n];
      }
    }
    any = true;
  }
  if (!any) return;
  const int r0 = 6 * node_of_knot[kr], c0 = 6 * node_of_knot[kc];
  if (dk > 0) {   // off-diagonal 6 x 6 block: whole rows, two columns at a time (c0 is even, so a pair never straddles a 32-column block)
#pragma unroll
    for (int a = 0; a < 6; ++a)
#pragma unroll
      for (int q = 0; q < 3; ++q) {
        double2* d = reinterpret_cast<double2*>(Hw.band + band_addr(Hw, r0 + a, c0 + 2 * q));
        double2 v = *d;
        v.x += C[6 * a + 2 * q]; v.y += C[6 * a + 2 * q + 1];
        *d = v;
      }
  } else {        // diagonal block: lower triangle only
#pragma unroll
    for (int e = 0; e < 36; ++e) {
      const int a = e / 6, c = e % 6;
      if (c <= a) Hw.band[band_addr(Hw, r0 + a, c0 + c)] += C[e];
    }
  }
}

// Same gather with one THREAD per (knot, dk): for sparse frame rates (fewer than ~4 frames per knot, e.g. the 2.8 Hz flight of config 3) a
// warp per task leaves 29 of 32 lanes idle and spends its time in the 36 x 5 reduction shuffles.
template <int N>
__global__ void __launch_bounds__(128) k_knot_schur_thread(const FramePose<N>* __restrict__ fps, int K, int KD, int T, const int32_t* __restrict__ f_s,
                                                           const int32_t* __restrict__ kf_begin, const int32_t* __restrict__ kf_list,
                                                           const int32_t* __restrict__ node_of_knot, const double* __restrict__ Sp, BandLayout Hw) {
  const size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= (size_t)K * (KD + 1)) return;
  const int kr = (int)(w / (KD + 1)), dk = (int)(w % (KD + 1));
  const int kc = kr - dk;
  if (kc < 0) return;
  const int rb = kf_begin[kr], nr = kf_begin[kr + 1] - rb;
  const int cb = kf_begin[kc], nc = kf_begin[kc + 1] - cb;
  if (nr == 0 || nc == 0) return;
  double C[36];
#pragma unroll
  for (int e = 0; e < 36; ++e) C[e] = 0.0;
  bool any = false;
  for (int a = 0; a < nr; ++a) {
    const int f = kf_list[rb + a];
    for (int b = 0; b < nc; ++b) {
      const int f2 = kf_list[cb + b];
      const int d = f2 - f;
      if (d > T || -d > T) continue;
      chain_block_acc_stream<N>(fps[f], kr - f_s[f], fps[f2], kc - f_s[f2], Sp + ((size_t)(d >= 0 ? f : f2) * (T + 1) + (d >= 0 ? d : -d)) * 36, d < 0, C);
      any = true;
    }
  }
  if (!any) return;
  const int r0 = 6 * node_of_knot[kr], c0 = 6 * node_of_knot[kc];
  if (dk > 0) {   // off-diagonal 6 x 6 block: whole rows, two columns at a time (c0 is even, so a pair never straddles a 32-column block)
#pragma unroll
    for (int a = 0; a < 6; ++a)
#pragma unroll
      for (int q = 0; q < 3; ++q) {
        double2* d = reinterpret_cast<double2*>(Hw.band + band_addr(Hw, r0 + a, c0 + 2 * q));
        double2 v = *d;
        v.x += C[6 * a + 2 * q]; v.y += C[6 * a + 2 * q + 1];
        *d = v;
      }
  } else {        // diagonal block: lower triangle only
#pragma unroll
    for (int e = 0; e < 36; ++e) {
      const int a = e / 6, c = e % 6;
      if (c <= a) Hw.band[band_addr(Hw, r0 + a, c0 + c)] += C[e];
    }
  }
}

// border rows (pose x globals mapped to knot dofs) and the Schur right-hand-side correction, thread per (knot, dof)
template <int N>
__global__ void k_knot_border(const FramePose<N>* __restrict__ fps, int K, const int32_t* __restrict__ f_s, const int32_t* __restrict__ kf_begin,
                              const int32_t* __restrict__ kf_list, const int32_t* __restrict__ node_of_knot, const double* __restrict__ Sb,
                              const double* __restrict__ rp, GMask gm, bool any_g, BandLayout Hw, double* __restrict__ rhs) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= 6 * K) return;
  const int k = t / 6, a = t % 6;
  double acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.0;
  for (int j = kf_begin[k]; j < kf_begin[k + 1]; ++j) {
    const int f = kf_list[j];
    const FramePose<N>& fp = fps[f];
    const int i = k - f_s[f];
    // row a of D_f[k]^T : rot dof a<3 -> Crot[i][:,a] over pose rows 0..2 ; pos dof -> cpos over pose row a
    if (a < 3) {
      const double w0 = fp.Crot[i][a], w1 = fp.Crot[i][3 + a], w2 = fp.Crot[i][6 + a];
      acc[7] += w0 * rp[6 * (size_t)f] + w1 * rp[6 * (size_t)f + 1] + w2 * rp[6 * (size_t)f + 2];
      if (any_g) {
#pragma unroll
        for (int g = 0; g < 7; ++g) acc[g] += w0 * Sb[42 * (size_t)f + g] + w1 * Sb[42 * (size_t)f + 7 + g] + w2 * Sb[42 * (size_t)f + 14 + g];
      }
    } else {
      const double c = fp.cpos[i];
      acc[7] += c * rp[6 * (size_t)f + a];
      if (any_g) {
#pragma unroll
        for (int g = 0; g < 7; ++g) acc[g] += c * Sb[42 * (size_t)f + 7 * a + g];
      }
    }
  }
  const int idx = 6 * node_of_knot[k] + a;
  rhs[idx] += acc[7];
  if (any_g) {
#pragma unroll
    for (int g = 0; g < 7; ++g) if (gm.a[g]) Hw.border[(size_t)g * Hw.n_pad + idx] += acc[g];
  }
}

// ---- Jacobi scaling (iteration 0 of each Solve): 1 / (1 + sqrt(diag J^T J)) ---------------------------------------
// [gradient | diag(J^T J)] of this handle's normal equations, packed for the all-reduce of the sharded solve
__global__ void k_pack_tot(BandLayout H, int n_band, int col_lo, int col_hi, GMask gm, const double* __restrict__ grad, double* __restrict__ tot) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = H.n_pad + kBorder;
  if (i >= n) return;
  tot[i] = grad[i];
  double d = 0.0;
  if (i < n_band) d = (i >= col_lo && i < col_hi) ? H.band[band_addr(H, i, i)] : 0.0;   // outside the handle's column range the band is never initialised
  else if (i >= H.n_pad && i < H.n_pad + GT_SIZE) { const int t = i - H.n_pad; d = gm.a[t] ? H.corner[t * kBorder + t] : 0.0; }
  tot[n + i] = d;
}

__global__ void k_jscale(int n_pad, const double* __restrict__ diag /* n_pad + 16 */, const double* __restrict__ V0, int L, double* __restrict__ jscale,
                         double* __restrict__ jscale_lm) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_pad + GT_SIZE) jscale[i] = 1.0 / (1.0 + sqrt(fmax(diag[i], 0.0)));
  else if (i < n_pad + kBorder) jscale[i] = 1.0;
  const int j = i;
  if (j < 3 * L) {
    const int di[3] = {0, 3, 5};
    jscale_lm[j] = 1.0 / (1.0 + sqrt(fmax(V0[6 * (size_t)(j / 3) + di[j % 3]], 0.0)));
  }
}

// Hw = H0 + LM diagonal (in unscaled space: clamp(H_jj s_j^2, 1e-6, 1e32) / (radius s_j^2)); padding rows get identity;
// inactive globals get identity; rhs row of the border = -grad (band part) ; corner rhs column = -grad (global part)
__global__ void k_add_damping(BandLayout H0, BandLayout Hw, int n_band, const double* __restrict__ jscale, const double* __restrict__ diag,
                              double radius, GMask gm, int own_lo, int own_hi, bool is_root, double* __restrict__ dvec /* n_pad+16: D^2 */) {
  // diag = diagonal of the WHOLE problem's J^T J; a sharded handle adds the damping only to the columns it owns (H0 holds its partial sums)
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < Hw.n_pad) {
    const bool own = i >= own_lo && i < own_hi;
    const size_t ad = band_addr(Hw, i, i);
    if (i < n_band) {
      const double sc = jscale[i];
      const double dd = fmin(fmax(diag[i] * sc * sc, 1e-6), 1e32) / (radius * sc * sc);
      if (own) Hw.band[ad] = H0.band[ad] + dd;
      dvec[i] = dd;
    } else {
      if (own) Hw.band[ad] = 1.0;
      dvec[i] = 0.0;
    }
  } else if (i < Hw.n_pad + GT_SIZE) {
    const int t = i - Hw.n_pad;
    if (gm.a[t]) {
      const double sc = jscale[i];
      const double dd = fmin(fmax(diag[i] * sc * sc, 1e-6), 1e32) / (radius * sc * sc);
      if (is_root) atomicAdd(Hw.corner + t * kBorder + t, dd);
      dvec[i] = dd;
    } else {
      Hw.corner[t * kBorder + t] = is_root ? 1.0 : 0.0;
      dvec[i] = 0.0;
    }
  }
}

// rhs row of the border (row 15) and rhs column of the corner: -(grad + schur correction accumulated in rhs_acc)
__global__ void k_set_rhs(BandLayout Hw, const double* __restrict__ grad, const double* __restrict__ rhs_acc, GMask gm) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < Hw.n_pad) {
    Hw.border[(size_t)kRhsRow * Hw.n_pad + i] = -(grad[i] + rhs_acc[i]);
  } else if (i < Hw.n_pad + GT_SIZE) {
    const int t = i - Hw.n_pad;
    Hw.corner[kRhsRow * kBorder + t] = gm.a[t] ? -(grad[i] + rhs_acc[i]) : 0.0;
  }
}

// ---- step application ------------------------------------------------------------------------------------------------
// pose-space step of each frame: dpose_f = D_f * delta(window f)
template <int N>
__global__ void k_frame_dpose(const FramePose<N>* __restrict__ fps, int F, const int32_t* __restrict__ f_s, const int32_t* __restrict__ node_of_knot,
                              const double* __restrict__ delta, double* __restrict__ dpose) {
  int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= F) return;
  const FramePose<N>& fp = fps[f];
  double out[6] = {0, 0, 0, 0, 0, 0};
  const int s = f_s[f];
  for (int i = 0; i < N; ++i) {
    const double* d = delta + 6 * (size_t)node_of_knot[s + i];
    for (int m = 0; m < 3; ++m) {
      out[m] += fp.Crot[i][3 * m] * d[0] + fp.Crot[i][3 * m + 1] * d[1] + fp.Crot[i][3 * m + 2] * d[2];
      out[3 + m] += fp.cpos[i] * d[3 + m];
    }
  }
  for (int m = 0; m < 6; ++m) dpose[6 * (size_t)f + m] = out[m];
}

// landmark back-substitution: delta_l = -Vi (g_l + sum_o Y_o [dpose_f ; dglob])   (thread per landmark)
template <int N>
__global__ void k_landmark_backsub(const __grid_constant__ Globals G, const double* __restrict__ glb, const FramePose<N>* __restrict__ fps,
                                   const double* __restrict__ lm, int L, const int32_t* __restrict__ lm_obs_begin,
                                   const int32_t* __restrict__ lm_obs, const int32_t* __restrict__ o_frame, const double* __restrict__ Jl,
                                   const double2* __restrict__ r, const double* __restrict__ f_w, int M, const double* __restrict__ V0,
                                   const double* __restrict__ gl, const double* __restrict__ jscale_lm, double radius, GMask gm,
                                   const double* __restrict__ dpose, const double* __restrict__ dglob /* 15 */, double* __restrict__ delta_lm,
                                   double* __restrict__ dvec_lm) {
  int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= L) return;
  const int b0 = lm_obs_begin[l], e0 = lm_obs_begin[l + 1];
  if (b0 == e0) { for (int c = 0; c < 3; ++c) { delta_lm[3 * (size_t)l + c] = 0.0; dvec_lm[3 * (size_t)l + c] = 0.0; } return; }
  double V[6], Vi[6], acc[3];
  for (int c = 0; c < 6; ++c) V[c] = V0[6 * (size_t)l + c];
  for (int c = 0; c < 3; ++c) acc[c] = gl[3 * (size_t)l + c];
  const int di[3] = {0, 3, 5};
  for (int c = 0; c < 3; ++c) {
    const double sc = jscale_lm[3 * (size_t)l + c];
    double dd = fmin(fmax(V[di[c]] * sc * sc, 1e-6), 1e32) / (radius * sc * sc);
    V[di[c]] += dd;
    dvec_lm[3 * (size_t)l + c] = dd;
  }
  {
    const double c00 = V[3] * V[5] - V[4] * V[4], c01 = V[2] * V[4] - V[1] * V[5], c02 = V[1] * V[4] - V[2] * V[3];
    const double det = V[0] * c00 + V[1] * c01 + V[2] * c02, id = 1.0 / det;
    Vi[0] = c00 * id; Vi[1] = c01 * id; Vi[2] = c02 * id;
    Vi[3] = (V[0] * V[5] - V[2] * V[2]) * id; Vi[4] = (V[1] * V[2] - V[0] * V[4]) * id; Vi[5] = (V[0] * V[3] - V[1] * V[1]) * id;
  }
  v3 pw{lm[3 * (size_t)l], lm[3 * (size_t)l + 1], lm[3 * (size_t)l + 2]};
  for (int i = b0; i < e0; ++i) {
    const int o = lm_obs[i], f = o_frame[o];
    double X[3][13], A[6], b[3];
    obs_X<N>(G, glb, fps[f], pw, X);
    obs_Ab(Jl, r, f_w, M, o, f, A, b);
    double xd[3] = {0, 0, 0};  // X * [dpose ; dglob(active)]
    for (int c = 0; c < 6; ++c) { const double dv = dpose[6 * (size_t)f + c]; xd[0] += X[0][c] * dv; xd[1] += X[1][c] * dv; xd[2] += X[2][c] * dv; }
    for (int c = 0; c < 7; ++c) if (gm.a[c]) { const double dv = dglob[c]; xd[0] += X[0][6 + c] * dv; xd[1] += X[1][6 + c] * dv; xd[2] += X[2][6 + c] * dv; }
    double y[3];
    symmul3(A, xd, y);
    acc[0] += y[0]; acc[1] += y[1]; acc[2] += y[2];
  }
  double d[3];
  symmul3(Vi, acc, d);
  for (int c = 0; c < 3; ++c) delta_lm[3 * (size_t)l + c] = -d[c];
}

// candidate state = state (+) delta   (knots / biases: thread per node; landmarks; globals by one thread)
__global__ void k_apply_nodes(int K, int Kb, const int32_t* __restrict__ node_of_knot, const int32_t* __restrict__ node_of_bias,
                              const double* __restrict__ delta, StatePtrs S, StatePtrs C) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < K) {
    const double* d = delta + 6 * (size_t)node_of_knot[i];
    quat q{S.knot_q[4 * (size_t)i], S.knot_q[4 * (size_t)i + 1], S.knot_q[4 * (size_t)i + 2], S.knot_q[4 * (size_t)i + 3]};
    quat qn = quat_mul(q, so3_exp(v3{d[0], d[1], d[2]}));  // lie_local_parametrization.h:34-41
    C.knot_q[4 * (size_t)i] = qn.x; C.knot_q[4 * (size_t)i + 1] = qn.y; C.knot_q[4 * (size_t)i + 2] = qn.z; C.knot_q[4 * (size_t)i + 3] = qn.w;
    for (int c = 0; c < 3; ++c) C.knot_p[3 * (size_t)i + c] = S.knot_p[3 * (size_t)i + c] + d[3 + c];
  } else if (i < K + Kb) {
    const int j = i - K;
    const double* d = delta + 6 * (size_t)node_of_bias[j];
    for (int c = 0; c < 3; ++c) {
      C.bias_a[3 * (size_t)j + c] = S.bias_a[3 * (size_t)j + c] + d[c];
      C.bias_g[3 * (size_t)j + c] = S.bias_g[3 * (size_t)j + c] + d[3 + c];
    }
  }
}
__global__ void k_apply_landmarks_globals(int L, const double* __restrict__ delta_lm, const double* __restrict__ dglob, GMask gm, StatePtrs S, StatePtrs C) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < 3 * L) C.lm[i] = S.lm[i] + delta_lm[i];
  if (i == 0) {
    double g[GLB_STATE_SIZE];
    for (int c = 0; c < GLB_STATE_SIZE; ++c) g[c] = S.glb[c];
    if (gm.a[GT_SCALE]) g[GLB_SCALE_CORR] += dglob[GT_SCALE];
    if (gm.a[GT_RBC]) {
      quat q{g[GLB_RBC_CORR], g[GLB_RBC_CORR + 1], g[GLB_RBC_CORR + 2], g[GLB_RBC_CORR + 3]};
      quat qn = quat_mul(q, so3_exp(v3{dglob[GT_RBC], dglob[GT_RBC + 1], dglob[GT_RBC + 2]}));
      g[GLB_RBC_CORR] = qn.x; g[GLB_RBC_CORR + 1] = qn.y; g[GLB_RBC_CORR + 2] = qn.z; g[GLB_RBC_CORR + 3] = qn.w;
    }
    for (int c = 0; c < 3; ++c) {
      if (gm.a[GT_PBC + c]) g[GLB_PBC_CORR + c] += dglob[GT_PBC + c];
      if (gm.a[GT_PBP + c]) g[GLB_PBP + c] += dglob[GT_PBP + c];
      if (gm.a[GT_G + c]) g[GLB_G + c] += dglob[GT_G + c];
    }
    if (gm.a[GT_T_CAM_IMU]) g[GLB_T_CAM_IMU_CORR] += dglob[GT_T_CAM_IMU];
    if (gm.a[GT_T_GP]) g[GLB_T_GP_CORR] += dglob[GT_T_GP];
    for (int c = 0; c < GLB_STATE_SIZE; ++c) C.glb[c] = g[c];
  }
}

// ---- scalars of an LM step (deterministic two-stage reductions) ---------------------------------------------------------
// item space: [0, nK) knots, [nK, nK+nKb) bias nodes, then landmarks (3 scalars each handled per landmark), then 1 item for globals
// out partials: per block 4 doubles: model (delta^T D^2 delta - delta^T g), step^2, xnorm^2, gmax
__device__ __forceinline__ double gmax_quat(quat q, v3 g) {  // max | q - q*exp(-g) |
  quat qn = quat_mul(q, so3_exp(v3{-g.x, -g.y, -g.z}));
  return fmax(fmax(fabs(qn.x - q.x), fabs(qn.y - q.y)), fmax(fabs(qn.z - q.z), fabs(qn.w - q.w)));
}
__global__ void __launch_bounds__(256) k_step_scalars(int K, int Kb, int L, const int32_t* __restrict__ node_of_knot,
                                                      const int32_t* __restrict__ node_of_bias, const uint8_t* __restrict__ touch_knot,
                                                      const uint8_t* __restrict__ touch_bias, const uint8_t* __restrict__ touch_lm, GMask gm, int n_pad,
                                                      const double* __restrict__ delta, const double* __restrict__ delta_lm,
                                                      const double* __restrict__ dvec, const double* __restrict__ dvec_lm,
                                                      const double* __restrict__ grad, const double* __restrict__ gl, StatePtrs S, StatePtrs C,
                                                      bool have_step, int own_lo, int own_hi, bool is_root, double* __restrict__ partial) {
  __shared__ double sh[4][256];
  double model = 0, step2 = 0, xn2 = 0, gmx = 0;
  const int total = K + Kb + L + 1;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    if (i < K) {
      const int nd = 6 * node_of_knot[i];
      if (nd < own_lo || nd >= own_hi) continue;   // sharded: a node is counted by the rank owning its first column
      const uint8_t tm = touch_knot[i];
      if (have_step) for (int c = 0; c < 6; ++c) model += delta[nd + c] * (dvec[nd + c] * delta[nd + c] - grad[nd + c]);
      if (tm & 1) {
        quat q{S.knot_q[4 * (size_t)i], S.knot_q[4 * (size_t)i + 1], S.knot_q[4 * (size_t)i + 2], S.knot_q[4 * (size_t)i + 3]};
        xn2 += q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w;
        gmx = fmax(gmx, gmax_quat(q, v3{grad[nd], grad[nd + 1], grad[nd + 2]}));
        if (have_step) for (int c = 0; c < 4; ++c) { double d = C.knot_q[4 * (size_t)i + c] - S.knot_q[4 * (size_t)i + c]; step2 += d * d; }
      }
      if (tm & 2) {
        for (int c = 0; c < 3; ++c) {
          const double x = S.knot_p[3 * (size_t)i + c];
          xn2 += x * x; gmx = fmax(gmx, fabs(grad[nd + 3 + c]));
          if (have_step) { double d = C.knot_p[3 * (size_t)i + c] - x; step2 += d * d; }
        }
      }
    } else if (i < K + Kb) {
      const int j = i - K, nd = 6 * node_of_bias[j];
      if (nd < own_lo || nd >= own_hi) continue;
      const uint8_t tm = touch_bias[j];
      if (have_step) for (int c = 0; c < 6; ++c) model += delta[nd + c] * (dvec[nd + c] * delta[nd + c] - grad[nd + c]);
      for (int c = 0; c < 3; ++c) {
        if (tm & 1) { const double x = S.bias_a[3 * (size_t)j + c]; xn2 += x * x; gmx = fmax(gmx, fabs(grad[nd + c])); if (have_step) { double d = C.bias_a[3 * (size_t)j + c] - x; step2 += d * d; } }
        if (tm & 2) { const double x = S.bias_g[3 * (size_t)j + c]; xn2 += x * x; gmx = fmax(gmx, fabs(grad[nd + 3 + c])); if (have_step) { double d = C.bias_g[3 * (size_t)j + c] - x; step2 += d * d; } }
      }
    } else if (i < K + Kb + L) {
      const int l = i - K - Kb;
      if (!touch_lm[l]) continue;
      for (int c = 0; c < 3; ++c) {
        const double x = S.lm[3 * (size_t)l + c], g = gl[3 * (size_t)l + c];
        xn2 += x * x; gmx = fmax(gmx, fabs(g));
        if (have_step) {
          const double dl = delta_lm[3 * (size_t)l + c];
          model += dl * (dvec_lm[3 * (size_t)l + c] * dl - g);
          double d = C.lm[3 * (size_t)l + c] - x; step2 += d * d;
        }
      }
    } else if (is_root) {
      // globals
      const double* dg = delta + n_pad;
      const double* gg = grad + n_pad;
      const double* dv = dvec + n_pad;
      auto eu = [&](int gt, int gs) {
        if (!gm.a[gt]) return;
        const double x = S.glb[gs];
        xn2 += x * x; gmx = fmax(gmx, fabs(gg[gt]));
        if (have_step) { model += dg[gt] * (dv[gt] * dg[gt] - gg[gt]); double d = C.glb[gs] - x; step2 += d * d; }
      };
      eu(GT_SCALE, GLB_SCALE_CORR);
      if (gm.a[GT_RBC]) {
        quat q{S.glb[GLB_RBC_CORR], S.glb[GLB_RBC_CORR + 1], S.glb[GLB_RBC_CORR + 2], S.glb[GLB_RBC_CORR + 3]};
        xn2 += q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w;
        gmx = fmax(gmx, gmax_quat(q, v3{gg[GT_RBC], gg[GT_RBC + 1], gg[GT_RBC + 2]}));
        if (have_step) {
          for (int c = 0; c < 3; ++c) model += dg[GT_RBC + c] * (dv[GT_RBC + c] * dg[GT_RBC + c] - gg[GT_RBC + c]);
          for (int c = 0; c < 4; ++c) { double d = C.glb[GLB_RBC_CORR + c] - S.glb[GLB_RBC_CORR + c]; step2 += d * d; }
        }
      }
      for (int c = 0; c < 3; ++c) { eu(GT_PBC + c, GLB_PBC_CORR + c); eu(GT_PBP + c, GLB_PBP + c); eu(GT_G + c, GLB_G + c); }
      eu(GT_T_CAM_IMU, GLB_T_CAM_IMU_CORR);
      eu(GT_T_GP, GLB_T_GP_CORR);
    }
  }
  sh[0][threadIdx.x] = model; sh[1][threadIdx.x] = step2; sh[2][threadIdx.x] = xn2; sh[3][threadIdx.x] = gmx;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) {
      sh[0][threadIdx.x] += sh[0][threadIdx.x + o]; sh[1][threadIdx.x] += sh[1][threadIdx.x + o];
      sh[2][threadIdx.x] += sh[2][threadIdx.x + o]; sh[3][threadIdx.x] = fmax(sh[3][threadIdx.x], sh[3][threadIdx.x + o]);
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) for (int k = 0; k < 4; ++k) partial[4 * blockIdx.x + k] = sh[k][0];
}
// one warp: lanes take the per-block partials in a fixed stride, then a fixed shuffle tree (bit-reproducible)
__global__ void k_step_scalars_final(int nblocks, const double* __restrict__ partial, double* __restrict__ out) {
  if (blockIdx.x != 0 || threadIdx.x >= 32) return;
  double model = 0, step2 = 0, xn2 = 0, gmx = 0;
  for (int b = threadIdx.x; b < nblocks; b += 32) { model += partial[4 * b]; step2 += partial[4 * b + 1]; xn2 += partial[4 * b + 2]; gmx = fmax(gmx, partial[4 * b + 3]); }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    model += __shfl_xor_sync(0xffffffffu, model, o); step2 += __shfl_xor_sync(0xffffffffu, step2, o);
    xn2 += __shfl_xor_sync(0xffffffffu, xn2, o); gmx = fmax(gmx, __shfl_xor_sync(0xffffffffu, gmx, o));
  }
  if (threadIdx.x == 0) { out[SC_MODEL] = 0.5 * model; out[SC_STEP2] = step2; out[SC_XNORM2] = xn2; out[SC_GMAX] = gmx; }
}

// ================================================== host side =====================================================
static GMask make_gmask(const SolverData* sd) { GMask g; for (int i = 0; i < GT_SIZE; ++i) g.a[i] = sd->gactive[i]; return g; }

// Host analysis of the problem structure (no device work): node order, half bandwidth, touch masks, landmark / knot CSR lists.
void analyse_structure(const ctba_problem* p, const GlobalStructure* gs, HostStructure* hs) {
  const int N = p->N;
  const int K = (int)p->numKnots(), F = (int)p->numFrames(), M = (int)p->numObs(), L = (int)(p->lm.size() / 3);
  const bool has_imu = gs ? gs->has_imu : (!p->i_s.empty() || !p->b_s0.empty());
  const int Kb = has_imu ? (int)p->numBiasKnots() : 0;
  hs->K = K; hs->Kb = Kb; hs->L = L; hs->F = F; hs->M = M; hs->has_imu = has_imu;
  // ---- node ordering: knots and bias knots merged by time (knot k at k*dt, bias knot j at j*dt_bias; ties: knot first)
  hs->node_of_knot.assign(K, 0); hs->node_of_bias.assign(std::max(Kb, 1), 0);
  {
    int ik = 0, ib = 0, node = 0;
    while (ik < K || ib < Kb) {
      const bool take_knot = ib >= Kb || (ik < K && (int64_t)ik * p->dt_ns <= (int64_t)ib * p->dt_ns_bias);
      if (take_knot) hs->node_of_knot[ik++] = node++; else hs->node_of_bias[ib++] = node++;
    }
    hs->nodes = node;
  }
  const std::vector<int32_t>&nk = hs->node_of_knot, &nb = hs->node_of_bias;
  // ---- half bandwidth in nodes over every coupling group
  int hbn = N - 1;
  auto span = [&](int lo, int hi) { hbn = std::max(hbn, hi - lo); };
  std::vector<uint8_t>&tk = hs->tk, &tb = hs->tb, &tl = hs->tl;
  tk.assign(std::max(K, 1), 0); tb.assign(std::max(Kb, 1), 0); tl.assign(std::max(L, 1), 0);
  for (size_t i = 0; i < p->i_s.size(); ++i) {
    const int s = p->i_s[i], sb = p->i_sb[i];
    const int lo = std::min(nk[s], nb[sb]), hi = std::max(nk[s + N - 1], nb[sb + 3]);
    span(lo, hi);
    for (int j = 0; j < N; ++j) tk[s + j] |= 3;
    for (int j = 0; j < 4; ++j) tb[sb + j] |= 3;
    // the reference registers N_ bias blocks per IMU residual (bspline_so3r3.h:585-588): blocks 4..N-1 are in the problem too
    for (int j = 4; j < N; ++j) tb[sb + j] |= 3;
  }
  for (size_t i = 0; i < p->g_s.size(); ++i) { const int s = p->g_s[i]; span(nk[s], nk[s + N - 1]); for (int j = 0; j < N; ++j) tk[s + j] |= 3; }
  for (size_t i = 0; i < p->b_s0.size(); ++i) {
    const int a = p->b_s0[i], b = p->b_s1[i];
    span(nb[std::min(a, b)], nb[std::max(a, b) + 3]);
    for (int j = 0; j < 4; ++j) { tb[a + j] |= (p->b_which[i] ? 2 : 1); tb[b + j] |= (p->b_which[i] ? 2 : 1); }
  }
  for (size_t i = 0; i < p->q_s.size(); ++i) {
    const int s = p->q_s[i];
    span(nk[s], nk[s + N - 1]);
    for (int j = 0; j < N; ++j) tk[s + j] |= (p->q_which[i] ? 1 : 2);
  }
  // runs of consecutive samples that address the same unknowns (k_rows_accumulate_seg)
  hs->prior_seg.assign(1, 0);
  for (size_t i = 1; i < p->q_s.size(); ++i)
    if (p->q_s[i] != p->q_s[i - 1] || p->q_which[i] != p->q_which[i - 1]) hs->prior_seg.push_back((int32_t)i);
  if (!p->q_s.empty()) hs->prior_seg.push_back((int32_t)p->q_s.size());
  hs->imu_seg.assign(1, 0);
  for (size_t i = 1; i < p->i_s.size(); ++i)
    if (p->i_s[i] != p->i_s[i - 1] || p->i_sb[i] != p->i_sb[i - 1]) hs->imu_seg.push_back((int32_t)i);
  if (!p->i_s.empty()) hs->imu_seg.push_back((int32_t)p->i_s.size());
  hs->gp_seg.assign(1, 0);
  for (size_t i = 1; i < p->g_s.size(); ++i)
    if (p->g_s[i] != p->g_s[i - 1]) hs->gp_seg.push_back((int32_t)i);
  if (!p->g_s.empty()) hs->gp_seg.push_back((int32_t)p->g_s.size());
  // landmark -> observation CSR (observations of a landmark in frame order, as they are stored)
  std::vector<int32_t>&lob = hs->lob, &lobs = hs->lobs;
  lob.assign(L + 1, 0); lobs.assign(M, 0);
  for (int o = 0; o < M; ++o) lob[p->o_lm[o] + 1]++;
  for (int l = 0; l < L; ++l) lob[l + 1] += lob[l];
  {
    std::vector<int32_t> cur(lob.begin(), lob.end() - 1);
    for (int o = 0; o < M; ++o) lobs[cur[p->o_lm[o]]++] = o;
  }
  hs->opos.assign(std::max(M, 1), 0); hs->lmf.assign(std::max(M, 1), 0);
  for (int l = 0; l < L; ++l) for (int i = lob[l]; i < lob[l + 1]; ++i) { hs->opos[lobs[i]] = i; hs->lmf[i] = p->o_frame[lobs[i]]; }
  // end of the landmark's list per observation; negative when the landmark is seen twice in one frame (k_frame_schur_c's general search)
  hs->oend.assign(std::max(M, 1), 1);
  for (int l = 0; l < L; ++l) {
    bool strictly = true;
    for (int i = lob[l] + 1; i < lob[l + 1]; ++i) strictly = strictly && hs->lmf[i] > hs->lmf[i - 1];
    for (int i = lob[l]; i < lob[l + 1]; ++i) hs->oend[lobs[i]] = strictly ? lob[l + 1] : -lob[l + 1];
  }
  hs->max_frame_obs = 1;
  for (int f = 0; f < F; ++f) hs->max_frame_obs = std::max(hs->max_frame_obs, (int)(p->f_obs_begin[f + 1] - p->f_obs_begin[f]));
  // per observation: bit d set iff its landmark is seen again d frames later (bit 0 = itself), d < 64 (k_frame_schur_m)
  hs->omask.assign(std::max(M, 1), 0ull);
  for (int l = 0; l < L; ++l)
    for (int i = lob[l]; i < lob[l + 1]; ++i) {
      unsigned long long m = 0ull;
      for (int k = i; k < lob[l + 1]; ++k) {
        const int d = hs->lmf[k] - hs->lmf[i];
        if (d >= 64) break;
        m |= 1ull << d;
      }
      hs->omask[lobs[i]] = m;
    }
  int T = 0, KD = N - 1;
  hs->lm_smin.assign(std::max(L, 1), -1);
  for (int l = 0; l < L; ++l) {
    if (lob[l + 1] == lob[l]) continue;
    tl[l] = 1;
    int fmin = 1 << 30, fmax_ = -1, smin = 1 << 30, smax = -1;
    for (int i = lob[l]; i < lob[l + 1]; ++i) {
      const int f = p->o_frame[lobs[i]];
      fmin = std::min(fmin, f); fmax_ = std::max(fmax_, f);
      smin = std::min(smin, (int)p->f_s[f]); smax = std::max(smax, (int)p->f_s[f]);
    }
    hs->lm_smin[l] = smin;
    T = std::max(T, fmax_ - fmin);
    KD = std::max(KD, smax + N - 1 - smin);
    span(nk[smin], nk[smax + N - 1]);
  }
  for (int f = 0; f < F; ++f) { const int s = p->f_s[f]; span(nk[s], nk[s + N - 1]); for (int j = 0; j < N; ++j) tk[s + j] |= 3; }
  // every SO3 knot carries a local parameterisation in the reference and is therefore part of the program (:378-399)
  for (int k = 0; k < K; ++k) tk[k] |= 1;
  hs->T = T; hs->hbn = hbn; hs->KD = KD;
  // knot -> frames whose window contains it (CSR), for the gather kernels
  std::vector<int32_t>&kfb = hs->kfb, &kfl = hs->kfl;
  kfb.assign(K + 1, 0); kfl.assign((size_t)std::max(F, 1) * N, 0);
  for (int f = 0; f < F; ++f) for (int j = 0; j < N; ++j) kfb[p->f_s[f] + j + 1]++;
  for (int k = 0; k < K; ++k) kfb[k + 1] += kfb[k];
  {
    std::vector<int32_t> cur(kfb.begin(), kfb.end() - 1);
    for (int f = 0; f < F; ++f) for (int j = 0; j < N; ++j) kfl[cur[p->f_s[f] + j]++] = f;
  }
  // ---- which globals are free (constness set inside add*Measurement)
  const bool anyF = gs ? gs->any_frames : F > 0, anyI = gs ? gs->any_imu : !p->i_s.empty(), anyG = gs ? gs->any_gp : !p->g_s.empty();
  const bool grav = gs ? gs->gravity_norm : p->gravity_norm;
  for (int i = 0; i < GT_SIZE; ++i) hs->gactive[i] = 0;
  hs->gactive[GT_SCALE] = p->optimize_scale && (anyF || anyG);
  for (int c = 0; c < 3; ++c) {
    hs->gactive[GT_RBC + c] = p->optimize_camimu_extrinsics && anyF;
    hs->gactive[GT_PBC + c] = p->optimize_camimu_extrinsics && anyF;
    hs->gactive[GT_PBP + c] = p->optimize_p_bp && anyG;
    hs->gactive[GT_G + c] = (anyI && p->optimize_gravity) || (!anyI && grav);
  }
  hs->gactive[GT_T_CAM_IMU] = p->optimize_camimu_extrinsics && anyI;
  hs->gactive[GT_T_GP] = p->optimize_t_offset_s_gs && anyG;
  if (gs) { hs->hbn = gs->hbn; hs->tk = gs->tk; hs->tb = gs->tb; }   // sharded: band width and touched unknowns of the WHOLE problem
  hs->n_band = 6 * hs->nodes;
  hs->n_pad = ((hs->n_band + kNB - 1) / kNB) * kNB;
  const int bw = 6 * hs->hbn + 5;
  hs->nbw = std::min((bw + 2 * kNB - 2) / kNB, std::max(hs->n_pad / kNB - 1, 0));
}

int assemble_structure(ctba_problem* p) {
  if (!p->solver) p->solver = new SolverData();
  SolverData* sd = p->solver;
  band_graph_release(sd);   // the captured launch sequence refers to the old structure / buffers
  HostStructure hs;
  analyse_structure(p, p->gstruct, &hs);
  const int K = hs.K, Kb = hs.Kb, F = hs.F, M = hs.M, L = hs.L, T = hs.T;
  sd->K = K; sd->Kb = Kb; sd->L = L; sd->F = F; sd->has_bias = Kb > 0; sd->nodes = hs.nodes;
  sd->node_of_knot = hs.node_of_knot; sd->node_of_bias = hs.node_of_bias;
  sd->max_frame_obs = hs.max_frame_obs;
  if ((size_t)sd->max_frame_obs * 18 * sizeof(double) > 200 * 1024) { p->err = "assemble_structure: more than 1422 valid observations in one frame"; return CTBA_E_ARG; }
  sd->T = T; sd->hbn = hs.hbn; sd->KD = hs.KD;
  {
    int covered = 0;
    for (int k = 0; k < K; ++k) covered += hs.kfb[k + 1] > hs.kfb[k];
    sd->frames_per_knot = covered ? (double)F * p->N / covered : 0.0;
  }
  for (int i = 0; i < GT_SIZE; ++i) sd->gactive[i] = hs.gactive[i];
  const int n_band = hs.n_band, n_pad = hs.n_pad, nbw = hs.nbw;
  BandLayout B{n_band, n_pad, n_pad / kNB, nbw, nullptr, nullptr, nullptr};
  // column range of the band this handle ever touches (sharded: left separator .. right separator of the rank's partitions)
  sd->bc_lo = 0; sd->bc_hi = B.nbc; sd->own_lo = 0; sd->own_hi = n_pad; sd->is_root = true;
  if (p->sharded) { sd->bc_lo = p->shard->bc_lo; sd->bc_hi = p->shard->bc_hi; sd->own_lo = p->shard->own_lo; sd->own_hi = p->shard->own_hi; sd->is_root = p->shard->rank == 0; }
  const size_t be = B.band_elems();
  CTBA_CUDA(p, sd->band0.resize(be)); CTBA_CUDA(p, sd->bandw.resize(be));
  CTBA_CUDA(p, sd->border0.resize((size_t)kBorder * n_pad)); CTBA_CUDA(p, sd->borderw.resize((size_t)kBorder * n_pad));
  CTBA_CUDA(p, sd->corner0.resize(kBorder * kBorder)); CTBA_CUDA(p, sd->cornerw.resize(kBorder * kBorder));
  sd->H0 = B; sd->H0.band = sd->band0.p; sd->H0.border = sd->border0.p; sd->H0.corner = sd->corner0.p;
  sd->Hw = B; sd->Hw.band = sd->bandw.p; sd->Hw.border = sd->borderw.p; sd->Hw.corner = sd->cornerw.p;
  CTBA_CUDA(p, sd->grad.resize(n_pad + kBorder)); CTBA_CUDA(p, sd->delta.resize(n_pad + kBorder));
  CTBA_CUDA(p, sd->tot.resize(2 * (size_t)(n_pad + kBorder)));     // [gradient | diag(J^T J)] of the whole problem
  CTBA_CUDA(p, sd->jscale.resize(2 * (size_t)(n_pad + kBorder)));  // second half: D^2 vector
  CTBA_CUDA(p, sd->V0.resize(6 * (size_t)std::max(L, 1))); CTBA_CUDA(p, sd->gl.resize(3 * (size_t)std::max(L, 1)));
  CTBA_CUDA(p, sd->jscale_lm.resize(6 * (size_t)std::max(L, 1)));  // second half: D^2 of landmarks
  CTBA_CUDA(p, sd->delta_lm.resize(3 * (size_t)std::max(L, 1)));
  CTBA_CUDA(p, sd->Sp.resize((size_t)std::max(F, 1) * (T + 1) * 36)); CTBA_CUDA(p, sd->Sb.resize((size_t)std::max(F, 1) * 42));
  CTBA_CUDA(p, sd->rp.resize((size_t)std::max(F, 1) * 6 + n_pad + kBorder));  // + rhs accumulator (n_pad+16)
  CTBA_CUDA(p, sd->dpose.resize((size_t)std::max(F, 1) * 6));
  CTBA_CUDA(p, sd->scal.resize(SC_COUNT + 4 * 256));
  CTBA_CUDA(p, sd->flag.resize(4));
  cudaStream_t st = p->stream;
  CTBA_CUDA(p, cudaMemsetAsync(sd->delta.p, 0, (n_pad + kBorder) * sizeof(double), st));
  CTBA_CUDA(p, sd->d_node_of_knot.upload(sd->node_of_knot, st)); CTBA_CUDA(p, sd->d_node_of_bias.upload(sd->node_of_bias, st));
  CTBA_CUDA(p, sd->d_lm_obs_begin.upload(hs.lob, st)); CTBA_CUDA(p, sd->d_lm_obs.upload(hs.lobs, st));
  CTBA_CUDA(p, sd->d_lm_frame.upload(hs.lmf, st));
  CTBA_CUDA(p, sd->d_o_pos.upload(hs.opos, st)); CTBA_CUDA(p, sd->d_o_end.upload(hs.oend, st)); CTBA_CUDA(p, sd->d_o_mask.upload(hs.omask, st));
  CTBA_CUDA(p, sd->d_imu_seg.upload(hs.imu_seg, st)); CTBA_CUDA(p, sd->d_gp_seg.upload(hs.gp_seg, st));
  CTBA_CUDA(p, sd->d_prior_seg.upload(hs.prior_seg, st));
  sd->n_imu_seg = (int)hs.imu_seg.size() - 1; sd->n_gp_seg = (int)hs.gp_seg.size() - 1; sd->n_prior_seg = (int)hs.prior_seg.size() - 1; CTBA_CUDA(p, sd->d_kf_begin.upload(hs.kfb, st)); CTBA_CUDA(p, sd->d_kf_list.upload(hs.kfl, st));
  CTBA_CUDA(p, sd->Vi.resize(6 * (size_t)std::max(L, 1))); CTBA_CUDA(p, sd->vig.resize(3 * (size_t)std::max(L, 1)));
  CTBA_CUDA(p, sd->ViYg.resize(21 * (size_t)std::max(L, 1)));
  CTBA_CUDA(p, sd->Yo.resize(18 * (size_t)std::max(M, 1))); CTBA_CUDA(p, sd->Zo.resize(18 * (size_t)std::max(M, 1)));
  CTBA_CUDA(p, sd->lm_corner_partial.resize((size_t)cdiv(std::max(L, 1), 128) * 35));
  CTBA_CUDA(p, sd->d_touch_knot.upload(hs.tk, st)); CTBA_CUDA(p, sd->d_touch_bias.upload(hs.tb, st)); CTBA_CUDA(p, sd->d_touch_lm.upload(hs.tl, st));
  // candidate state buffers
  CTBA_CUDA(p, sd->cand.knot_q.resize(p->dstate.knot_q.n)); CTBA_CUDA(p, sd->cand.knot_p.resize(p->dstate.knot_p.n));
  CTBA_CUDA(p, sd->cand.bias_a.resize(p->dstate.bias_a.n)); CTBA_CUDA(p, sd->cand.bias_g.resize(p->dstate.bias_g.n));
  CTBA_CUDA(p, sd->cand.lm.resize(p->dstate.lm.n)); CTBA_CUDA(p, sd->cand.glb.resize(p->dstate.glb.n));
  CTBA_CUDA(p, cudaStreamSynchronize(st));
  sd->built = true;
  return CTBA_OK;
}

template <int N>
static int assemble_direct_T(ctba_problem* p) {
  SolverData* sd = p->solver;
  cudaStream_t st = p->stream;
  const Globals G = make_globals(p);
  const GMask gm = make_gmask(sd);
  const int F = sd->F, M = (int)p->numObs(), L = sd->L;
  BandLayout& H = sd->H0;
  const size_t col_elems = (size_t)(H.nbw + 1) * kNB * kNB;   // one block column of the band
  CTBA_CUDA(p, cudaMemsetAsync(H.band + sd->bc_lo * col_elems, 0, (size_t)(sd->bc_hi - sd->bc_lo) * col_elems * sizeof(double), st));
  CTBA_CUDA(p, cudaMemsetAsync(H.border, 0, (size_t)kBorder * H.n_pad * sizeof(double), st));
  CTBA_CUDA(p, cudaMemsetAsync(H.corner, 0, kBorder * kBorder * sizeof(double), st));
  CTBA_CUDA(p, cudaMemsetAsync(sd->grad.p, 0, (H.n_pad + kBorder) * sizeof(double), st));
  const StatePtrs S = p->dstate.ptrs();
  const FramePose<N>* fps = (const FramePose<N>*)p->ev.frame_pose.p;
  if (F > 0) {
    if (L > 0) {
      k_landmark_blocks<<<cdiv(L, 128), 128, 0, st>>>(L, sd->d_lm_obs_begin.p, sd->d_lm_obs.p, p->d_o_frame.p, p->ev.Jl.p,
                                                      (const double2*)p->ev.r_reproj.p, p->ev.f_w.p, M, sd->V0.p, sd->gl.p);
      p->launches++;
    }
    k_frame_direct<N><<<cdiv(F, 4), 128, 0, st>>>(G, S.glb, fps, S.lm, F, p->d_f_s.p, p->d_f_obs_begin.p, p->d_o_lm.p, p->ev.Jl.p,
                                                  (const double2*)p->ev.r_reproj.p, p->ev.f_w.p, M, sd->d_node_of_knot.p, gm, H, sd->grad.p);
    p->launches++;
  }
  const int nI = (int)p->i_s.size(), nG = (int)p->g_s.size(), nB = (int)p->b_s0.size();
  if (nI > 0) {
    k_rows_accumulate_seg<N><<<sd->n_imu_seg, kSegThreads, 0, st>>>(1, sd->n_imu_seg, sd->d_imu_seg.p, 6 * N + 16, nI, p->ev.J_acc.p, p->ev.r_acc.p,
                                                                    p->d_i_s.p, p->d_i_sb.p, sd->d_node_of_knot.p, sd->d_node_of_bias.p, gm, H, sd->grad.p);
    k_rows_accumulate_seg<N><<<sd->n_imu_seg, kSegThreads, 0, st>>>(2, sd->n_imu_seg, sd->d_imu_seg.p, 3 * N + 13, nI, p->ev.J_gyr.p, p->ev.r_gyr.p,
                                                                    p->d_i_s.p, p->d_i_sb.p, sd->d_node_of_knot.p, sd->d_node_of_bias.p, gm, H, sd->grad.p);
    p->launches += 2;
  }
  if (nG > 0) {
    // runs of samples per knot interval only pay when there are several samples per interval (a 10 Hz receiver on 0.1 s knots has one:
    // config 3 spent 0.49 ms on 36 k single-sample CTAs); otherwise one warp per sample
    if (2 * sd->n_gp_seg > nG)
      k_rows_accumulate<N><<<cdiv(nG, 4), 128, 0, st>>>(3, nG, 6 * N + 5, p->ev.J_gp.p, p->ev.r_gp.p, p->d_g_s.p, nullptr, nullptr, nullptr,
                                                        sd->d_node_of_knot.p, sd->d_node_of_bias.p, gm, H, sd->grad.p);
    else
      k_rows_accumulate_seg<N><<<sd->n_gp_seg, kSegThreads, 0, st>>>(3, sd->n_gp_seg, sd->d_gp_seg.p, 6 * N + 5, 0, p->ev.J_gp.p, p->ev.r_gp.p, p->d_g_s.p,
                                                                     nullptr, sd->d_node_of_knot.p, sd->d_node_of_bias.p, gm, H, sd->grad.p);
    p->launches++;
  }
  if (!p->q_s.empty()) {
    k_rows_accumulate_seg<N><<<sd->n_prior_seg, kSegThreads, 0, st>>>(6, sd->n_prior_seg, sd->d_prior_seg.p, 3 * N, 0, p->ev.J_prior.p, p->ev.r_prior.p,
                                                                      p->d_q_s.p, p->d_q_which.p, sd->d_node_of_knot.p, sd->d_node_of_bias.p, gm, H, sd->grad.p);
    p->launches++;
  }
  if (nB > 0) {
    k_rows_accumulate<N><<<cdiv(nB, 4), 128, 0, st>>>(4, nB, 24, p->ev.J_bias.p, p->ev.r_bias.p, p->d_b_s0.p, nullptr, p->d_b_s1.p, p->d_b_which.p,
                                                      sd->d_node_of_knot.p, sd->d_node_of_bias.p, gm, H, sd->grad.p);
    p->launches++;
  }
  if (p->gravity_norm) {
    k_rows_accumulate<N><<<1, 128, 0, st>>>(5, 1, 3, p->ev.J_grav.p, p->ev.r_grav.p, nullptr, nullptr, nullptr, nullptr, sd->d_node_of_knot.p,
                                            sd->d_node_of_bias.p, gm, H, sd->grad.p);
    p->launches++;
  }
  k_pack_tot<<<cdiv(H.n_pad + kBorder, 256), 256, 0, st>>>(H, H.n_band, kNB * sd->bc_lo, kNB * sd->bc_hi, gm, sd->grad.p, sd->tot.p);
  p->launches++;
  CTBA_CUDA(p, cudaGetLastError());
  if (p->sharded && p->comm->allreduce_sum(sd->tot.p, 2 * (size_t)(H.n_pad + kBorder), st)) { p->err = p->comm->error(); return CTBA_E_NCCL; }
  return CTBA_OK;
}
int assemble_direct(ctba_problem* p) { return p->N == 4 ? assemble_direct_T<4>(p) : assemble_direct_T<6>(p); }

int compute_jscale(ctba_problem* p) {
  SolverData* sd = p->solver;
  const int n = std::max(sd->H0.n_pad + kBorder, 3 * sd->L);
  k_jscale<<<cdiv(n, 256), 256, 0, p->stream>>>(sd->H0.n_pad, sd->tot.p + (sd->H0.n_pad + kBorder), sd->V0.p, sd->L, sd->jscale.p, sd->jscale_lm.p);
  p->launches++;
  CTBA_CUDA(p, cudaGetLastError());
  return CTBA_OK;
}

template <int N>
static int assemble_schur_T(ctba_problem* p, double radius) {
  SolverData* sd = p->solver;
  cudaStream_t st = p->stream;
  const Globals G = make_globals(p);
  const GMask gm = make_gmask(sd);
  const int F = sd->F, M = (int)p->numObs(), L = sd->L, T = sd->T;
  BandLayout &H0 = sd->H0, &Hw = sd->Hw;
  const int n_pad = Hw.n_pad;
  const size_t col_elems = (size_t)(H0.nbw + 1) * kNB * kNB;
  CTBA_CUDA(p, cudaMemcpyAsync(Hw.band + sd->bc_lo * col_elems, H0.band + sd->bc_lo * col_elems, (size_t)(sd->bc_hi - sd->bc_lo) * col_elems * sizeof(double),
                               cudaMemcpyDeviceToDevice, st));
  CTBA_CUDA(p, cudaMemcpyAsync(Hw.border, H0.border, (size_t)kBorder * n_pad * sizeof(double), cudaMemcpyDeviceToDevice, st));
  CTBA_CUDA(p, cudaMemcpyAsync(Hw.corner, H0.corner, kBorder * kBorder * sizeof(double), cudaMemcpyDeviceToDevice, st));
  double* dvec = sd->jscale.p + (n_pad + kBorder);
  double* rhs_acc = sd->rp.p + (size_t)std::max(F, 1) * 6;  // n_pad + 16 accumulator of Schur right-hand-side corrections
  CTBA_CUDA(p, cudaMemsetAsync(rhs_acc, 0, (n_pad + kBorder) * sizeof(double), st));
  // LM diagonal first (it reads the undamped diagonal of H0), Schur terms are then added atomically on top
  k_add_damping<<<cdiv(n_pad + kBorder, 256), 256, 0, st>>>(H0, Hw, H0.n_band, sd->jscale.p, sd->tot.p + (n_pad + kBorder), radius, gm, sd->own_lo, sd->own_hi,
                                                            sd->is_root, dvec);
  p->launches++;
  const StatePtrs S = p->dstate.ptrs();
  if (F > 0 && L > 0) {
    const FramePose<N>* fps = (const FramePose<N>*)p->ev.frame_pose.p;
    bool any_g = false;
    for (int c = 0; c < 7; ++c) any_g |= gm.a[c] != 0;
    const int nlb = cdiv(L, 128);
    k_lm_prepare<N><<<nlb, 128, 0, st>>>(G, S.glb, fps, S.lm, L, sd->d_lm_obs_begin.p, sd->d_lm_obs.p, p->d_o_frame.p, p->ev.Jl.p,
                                         (const double2*)p->ev.r_reproj.p, p->ev.f_w.p, M, sd->V0.p, sd->gl.p, sd->jscale_lm.p, radius, gm, any_g,
                                         sd->Vi.p, sd->vig.p, sd->ViYg.p, sd->lm_corner_partial.p);
    if (any_g) { k_corner_reduce<<<kCornerVals, 32, 0, st>>>(nlb, sd->lm_corner_partial.p, Hw, rhs_acc + n_pad); p->launches++; }
    if (p->schur_compact) {
      double* Al = p->schur_list_a ? sd->Zo.p : nullptr;   // Zo (18 M doubles) is free in the compact path
      k_obs_a<<<cdiv(M, 256), 256, 0, st>>>(M, p->d_o_frame.p, p->ev.Jl.p, (const double2*)p->ev.r_reproj.p, p->ev.f_w.p, sd->Yo.p, sd->d_o_pos.p, Al);
      const int n_pad = (sd->max_frame_obs + 1) & ~1;
      const size_t fs_smem = (size_t)n_pad * kFcRows * sizeof(double);
      if (Al && p->schur_sorted && T + 1 <= 64) {
        const size_t sm_smem = (size_t)n_pad * 14 * sizeof(double);   // C (9) + d (3) + track mask (1) rows of doubles, list slot (int)
        auto launch = [&](auto kern, int threads) -> int {
          if (sm_smem > 48 * 1024) CTBA_CUDA(p, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_smem));
          kern<<<F, threads, sm_smem, st>>>(F, T, fps, S.lm, p->d_f_obs_begin.p, p->d_o_lm.p, sd->d_o_mask.p, sd->d_o_pos.p, sd->d_o_end.p, sd->d_lm_obs.p,
                                            sd->d_lm_frame.p, sd->Yo.p, Al, sd->Vi.p, sd->vig.p, sd->ViYg.p, any_g, n_pad, sd->Sp.p, sd->Sb.p, sd->rp.p);
          return CTBA_OK;
        };
        int rc;
        if (p->schur_threads == 128) rc = launch(k_frame_schur_m<N, 128>, 128);
        else if (p->schur_threads == 192) rc = launch(k_frame_schur_m<N, 192>, 192);
        else rc = launch(k_frame_schur_m<N, 256>, 256);
        if (rc) return rc;
      } else {
        auto kfc = Al ? k_frame_schur_c<N, true> : k_frame_schur_c<N, false>;
        if (fs_smem > 48 * 1024) CTBA_CUDA(p, cudaFuncSetAttribute(kfc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fs_smem));
        kfc<<<F, kFcThreads, fs_smem, st>>>(F, T, fps, S.lm, p->d_f_obs_begin.p, p->d_o_lm.p, sd->d_o_pos.p, sd->d_o_end.p, sd->d_lm_obs.p,
                                            sd->d_lm_frame.p, sd->Yo.p, Al, sd->Vi.p, sd->vig.p, sd->ViYg.p, any_g, n_pad, sd->Sp.p, sd->Sb.p, sd->rp.p);
      }
    } else {
      k_obs_yz<N><<<cdiv(M, 128), 128, 0, st>>>(G, S.glb, fps, S.lm, M, p->d_o_frame.p, p->d_o_lm.p, p->ev.Jl.p, (const double2*)p->ev.r_reproj.p,
                                                p->ev.f_w.p, sd->Vi.p, sd->Yo.p, sd->Zo.p);
      const size_t fs_smem = (size_t)sd->max_frame_obs * 18 * sizeof(double);
      if (fs_smem > 48 * 1024) CTBA_CUDA(p, cudaFuncSetAttribute(k_frame_schur, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fs_smem));
      k_frame_schur<<<F, kFsThreads, fs_smem, st>>>(F, T, p->d_f_obs_begin.p, p->d_o_lm.p, p->d_o_frame.p, sd->d_o_pos.p, sd->d_lm_obs_begin.p,
                                              sd->d_lm_obs.p, sd->d_lm_frame.p, sd->Yo.p, sd->Zo.p, sd->vig.p, sd->ViYg.p, any_g, sd->Sp.p, sd->Sb.p, sd->rp.p);
    }
    // frames per knot = F N / K; a task pairs the frames of two knots
    const double fpk = sd->frames_per_knot;
    if (fpk * fpk < p->schur_thread_below)
      k_knot_schur_thread<N><<<cdiv((size_t)sd->K * (sd->KD + 1), 128), 128, 0, st>>>(fps, sd->K, sd->KD, T, p->d_f_s.p, sd->d_kf_begin.p,
                                                                                       sd->d_kf_list.p, sd->d_node_of_knot.p, sd->Sp.p, Hw);
    else if (p->schur_two_stage) {
      const size_t nU = (size_t)F * (sd->KD + N);
      CTBA_CUDA(p, sd->Uk.resize(nU * 36));
      k_pose_to_knot_cols<N><<<cdiv(nU, 128), 128, 0, st>>>(fps, F, sd->K, sd->KD, T, p->d_f_s.p, sd->d_kf_begin.p, sd->d_kf_list.p, sd->Sp.p, sd->Uk.p);
      k_knot_schur_cols<N><<<cdiv((size_t)sd->K * (sd->KD + 1), 128), 128, 0, st>>>(fps, sd->K, sd->KD, p->d_f_s.p, sd->d_kf_begin.p, sd->d_kf_list.p,
                                                                                      sd->d_node_of_knot.p, sd->Uk.p, Hw);
      p->launches++;
    } else
      k_knot_schur<N><<<cdiv((size_t)sd->K * (sd->KD + 1) * 32, 128), 128, 0, st>>>(fps, sd->K, sd->KD, T, p->d_f_s.p, sd->d_kf_begin.p,
                                                                                     sd->d_kf_list.p, sd->d_node_of_knot.p, sd->Sp.p, Hw);
    k_knot_border<N><<<cdiv((size_t)6 * sd->K, 128), 128, 0, st>>>(fps, sd->K, p->d_f_s.p, sd->d_kf_begin.p, sd->d_kf_list.p,
                                                                    sd->d_node_of_knot.p, sd->Sb.p, sd->rp.p, gm, any_g, Hw, rhs_acc);
    p->launches += 4;
  }
  k_set_rhs<<<cdiv(n_pad + kBorder, 256), 256, 0, st>>>(Hw, sd->grad.p, rhs_acc, gm);
  p->launches++;
  CTBA_CUDA(p, cudaGetLastError());
  return CTBA_OK;
}
int assemble_schur(ctba_problem* p, double radius) { return p->N == 4 ? assemble_schur_T<4>(p, radius) : assemble_schur_T<6>(p, radius); }

template <int N>
static int backsub_T(ctba_problem* p, const StatePtrs& S, const StatePtrs& C, double radius) {
  SolverData* sd = p->solver;
  cudaStream_t st = p->stream;
  const Globals G = make_globals(p);
  const GMask gm = make_gmask(sd);
  const int F = sd->F, M = (int)p->numObs(), L = sd->L;
  const int n_pad = sd->Hw.n_pad;
  const FramePose<N>* fps = (const FramePose<N>*)p->ev.frame_pose.p;
  double* dvec_lm = sd->jscale_lm.p + 3 * (size_t)std::max(L, 1);
  if (F > 0 && L > 0) {
    k_frame_dpose<N><<<cdiv(F, 128), 128, 0, st>>>(fps, F, p->d_f_s.p, sd->d_node_of_knot.p, sd->delta.p, sd->dpose.p);
    k_landmark_backsub<N><<<cdiv(L, 128), 128, 0, st>>>(G, S.glb, fps, S.lm, L, sd->d_lm_obs_begin.p, sd->d_lm_obs.p, p->d_o_frame.p, p->ev.Jl.p,
                                                        (const double2*)p->ev.r_reproj.p, p->ev.f_w.p, M, sd->V0.p, sd->gl.p, sd->jscale_lm.p,
                                                        radius, gm, sd->dpose.p, sd->delta.p + n_pad, sd->delta_lm.p, dvec_lm);
    p->launches += 2;
  } else if (L > 0) {
    CTBA_CUDA(p, cudaMemsetAsync(sd->delta_lm.p, 0, 3 * (size_t)L * sizeof(double), st));
    CTBA_CUDA(p, cudaMemsetAsync(dvec_lm, 0, 3 * (size_t)L * sizeof(double), st));
  }
  k_apply_nodes<<<cdiv(sd->K + sd->Kb, 128), 128, 0, st>>>(sd->K, sd->Kb, sd->d_node_of_knot.p, sd->d_node_of_bias.p, sd->delta.p, S, C);
  k_apply_landmarks_globals<<<cdiv(std::max(3 * L, 1), 256), 256, 0, st>>>(L, sd->delta_lm.p, sd->delta.p + n_pad, gm, S, C);
  p->launches += 2;
  if (sd->Kb == 0) {  // bias arrays are unused but keep the candidate mirror defined
    CTBA_CUDA(p, cudaMemcpyAsync(C.bias_a, S.bias_a, p->dstate.bias_a.n * sizeof(double), cudaMemcpyDeviceToDevice, st));
    CTBA_CUDA(p, cudaMemcpyAsync(C.bias_g, S.bias_g, p->dstate.bias_g.n * sizeof(double), cudaMemcpyDeviceToDevice, st));
  }
  CTBA_CUDA(p, cudaGetLastError());
  return CTBA_OK;
}
int backsub_and_candidate(ctba_problem* p, const StatePtrs& S, const StatePtrs& C, double radius) {
  return p->N == 4 ? backsub_T<4>(p, S, C, radius) : backsub_T<6>(p, S, C, radius);
}

// have_step = (radius > 0): model cost change, step norm; always: |x|^2 and gradient max norm
int step_scalars(ctba_problem* p, const StatePtrs& S, const StatePtrs& C, double radius) {
  SolverData* sd = p->solver;
  const int n_pad = sd->Hw.n_pad;
  const int L = sd->L;
  double* dvec = sd->jscale.p + (n_pad + kBorder);
  double* dvec_lm = sd->jscale_lm.p + 3 * (size_t)std::max(L, 1);
  const int total = sd->K + sd->Kb + L + 1;
  const int nb = std::min(256, cdiv(total, 256));
  k_step_scalars<<<nb, 256, 0, p->stream>>>(sd->K, sd->Kb, L, sd->d_node_of_knot.p, sd->d_node_of_bias.p, sd->d_touch_knot.p, sd->d_touch_bias.p,
                                            sd->d_touch_lm.p, make_gmask(sd), n_pad, sd->delta.p, sd->delta_lm.p, dvec, dvec_lm, sd->tot.p, sd->gl.p, S, C,
                                            radius > 0, sd->own_lo, sd->own_hi, sd->is_root, sd->scal.p + SC_COUNT);
  k_step_scalars_final<<<1, 32, 0, p->stream>>>(nb, sd->scal.p + SC_COUNT, sd->scal.p);
  p->launches += 2;
  CTBA_CUDA(p, cudaGetLastError());
  return CTBA_OK;
}

}  // namespace ctba
