// ctba_select.cu — host-side measurement bookkeeping that the reference performs around Solve():
//   ctba_compute_valid_reprojections   BSplineSO3R3::computeValidReprojections   include/spline/bspline_so3r3.h:406-521
//                                      (+ the per-frame feature grid of Frame::addPoint2D / setGrid, include/sensor/camera.h:155-184)
//   ctba_evaluate_errors               evaluate{Accel,Gyro,GlobalPos,Reprojection}Errors   bspline_so3r3.h:1229-1579
//   ctba_sample                        getPose / getAccel / getGyro / get*Bias at many timestamps   bspline_so3r3.h:122-327
//                                      (the loops of traceSpline & co., include/util/save.h:99-153), thread per timestamp
// The selection is sequential from frame to frame (the grid of frame k-1 seeds frame k), so it stays on the host like in the
// reference; it is written over flat arrays (CSR frames, one stamp array for the landmark -> point lookup instead of std::find).
#include <cmath>
#include <vector>
#include "ctba_solver.h"
#include "ctba_select_core.h"

using namespace ctba;

namespace {

// R_wb, p_wb of the spline at (s, u): getPose(), bspline_so3r3.h:221-270
template <int N>
void host_pose(const ctba_problem* p, const SplineTables& tb, int64_t s, double u, m3* R, v3* t) {
  So3SplineResult<N> sr;
  so3_spline_eval<N, false, false, false>(tb, &p->knot_q[4 * (size_t)s], u, p->inv_dt, &sr);
  *R = quat_matrix(sr.q);
  double c[N];
  spline_coeffs<N>(tb.blend, tb.base, 0, 1.0, u, c);
  v3 a{0, 0, 0};
  for (int i = 0; i < N; ++i) { a.x += c[i] * p->knot_p[3 * (s + i)]; a.y += c[i] * p->knot_p[3 * (s + i) + 1]; a.z += c[i] * p->knot_p[3 * (s + i) + 2]; }
  *t = a;
}

__global__ void k_error_norms(int n, int nres, const double* __restrict__ r, const double* __restrict__ inv_std, const int32_t* __restrict__ idx,
                              double* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double s = 0.0;
  for (int c = 0; c < nres; ++c) { const double v = r[(size_t)i * nres + c]; s += v * v; }
  const double w = inv_std[idx ? idx[i] : i];
  out[i] = sqrt(s) / w;   // the reference evaluates the functor with inv_std = 1
}

// thread per timestamp; (window, u) from the int64 nanosecond arithmetic of the reference getters
template <int N>
__global__ void k_sample(const __grid_constant__ Globals G, StatePtrs S, int what, int64_t n, const int64_t* __restrict__ t_ns, int64_t start_t_ns,
                         int64_t dt_ns, int64_t dt_ns_bias, double* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int64_t st_ns = t_ns[i] - start_t_ns;
  if (what == CTBA_SAMPLE_ACCEL_BIAS || what == CTBA_SAMPLE_GYRO_BIAS) {
    const int64_t s = st_ns / dt_ns_bias;
    const double u = double(st_ns % dt_ns_bias) / double(dt_ns_bias);
    const double* b = (what == CTBA_SAMPLE_ACCEL_BIAS ? S.bias_a : S.bias_g) + 3 * s;
    double c[N];
    spline_coeffs<N>(G.tb.blend, G.tb.base, 0, 1.0, u, c);     // order N on the bias knots, like the reference getter
    double v[3] = {0, 0, 0};
    for (int j = 0; j < N; ++j) for (int k = 0; k < 3; ++k) v[k] += c[j] * b[3 * j + k];
    for (int k = 0; k < 3; ++k) out[3 * i + k] = v[k];
    return;
  }
  const int64_t s = st_ns / dt_ns;
  const double u = double(st_ns % dt_ns) / double(dt_ns);
  const double* kq = S.knot_q + 4 * s;
  const double* kp = S.knot_p + 3 * s;
  auto r3 = [&](int D, double sc) {
    double c[N];
    spline_coeffs<N>(G.tb.blend, G.tb.base, D, sc, u, c);
    v3 a{0, 0, 0};
    for (int j = 0; j < N; ++j) { a.x += c[j] * kp[3 * j]; a.y += c[j] * kp[3 * j + 1]; a.z += c[j] * kp[3 * j + 2]; }
    return a;
  };
  if (what == CTBA_SAMPLE_POSE) {
    So3SplineResult<N> sr;
    so3_spline_eval<N, false, false, false>(G.tb, kq, u, G.inv_dt, &sr);
    const v3 t = r3(0, 1.0);
    double* o = out + 7 * i;
    o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = sr.q.x; o[4] = sr.q.y; o[5] = sr.q.z; o[6] = sr.q.w;
  } else if (what == CTBA_SAMPLE_ACCEL_BODY || what == CTBA_SAMPLE_ACCEL_WORLD) {
    So3SplineResult<N> sr;
    so3_spline_eval<N, false, false, false>(G.tb, kq, u, G.inv_dt, &sr);
    v3 a = r3(2, G.inv_dt * G.inv_dt);
    a.x += S.glb[GLB_G]; a.y += S.glb[GLB_G + 1]; a.z += S.glb[GLB_G + 2];
    if (what == CTBA_SAMPLE_ACCEL_BODY) a = quat_act(quat_inv(sr.q), a);
    out[3 * i] = a.x; out[3 * i + 1] = a.y; out[3 * i + 2] = a.z;
  } else {
    So3SplineResult<N> sr;
    so3_spline_eval<N, true, false, false>(G.tb, kq, u, G.inv_dt, &sr);
    v3 w = sr.w;
    if (what == CTBA_SAMPLE_GYRO_WORLD) w = quat_act(quat_inv(sr.q), w);   // sic (bspline_so3r3.h:203-207)
    out[3 * i] = w.x; out[3 * i + 1] = w.y; out[3 * i + 2] = w.z;
  }
}

}  // namespace

extern "C" {

int ctba_compute_valid_reprojections(ctba_problem* p, int F, const int64_t* t_ns, const int32_t* pt_begin, const double* uv, const int32_t* lm_index,
                                     int width, int height, int cell_size_px, uint8_t* valid, int32_t* n_valid) {
  if (!p) return CTBA_E_ARG;
  if (F < 0 || !t_ns || !pt_begin || !valid || !n_valid || width <= 0 || height <= 0 || cell_size_px <= 1) { p->err = "compute_valid_reprojections: bad arguments"; return CTBA_E_ARG; }
  if (p->host_state_stale) { p->err = "compute_valid_reprojections: read the state back first (ctba_get_knots)"; return CTBA_E_STATE; }
  const int N = p->N;
  SplineTables tb;
  fill_tables(N, &tb);
  const m3 R_bc = quat_matrix(quat{p->q_bc0[0], p->q_bc0[1], p->q_bc0[2], p->q_bc0[3]});
  const v3 p_bc{p->p_bc0[0], p->p_bc0[1], p->p_bc0[2]};
  const double scale = p->scale + p->glb[GLB_SCALE_CORR];   // getScale() :1881
  // frame -> pose of the spline at its time; frames outside the spline are skipped (:418, :422)
  auto pose = [&](int f, m3* R_wb, v3* p_wb) -> bool {
    const int64_t st_ns = t_ns[f] - p->start_t_ns;
    if (st_ns < 0) return false;
    const int64_t s = st_ns / p->dt_ns;
    if ((size_t)(s + N) > p->numKnots()) return false;
    const double u = double(st_ns % p->dt_ns) / double(p->dt_ns);
    if (N == 4) host_pose<4>(p, tb, s, u, R_wb, p_wb); else host_pose<6>(p, tb, s, u, R_wb, p_wb);
    return true;
  };
  return select_valid_reprojections(F, pt_begin, uv, lm_index, (int64_t)p->lm.size() / 3, p->lm.data(), width, height, cell_size_px, R_bc, p_bc, scale, pose, valid, n_valid,
                                    &p->err);
}

int ctba_evaluate_errors(ctba_problem* p, int kind, double* errors) {
  if (!p) return CTBA_E_ARG;
  if (kind == CTBA_ALIGN) {   // evaluateWorldColmapAlignmentErrors :1682-1790
    if (p->device < 0) { p->err = "planning-only handle (device -1): no compute entry points"; return CTBA_E_CUDA; }
    if (p->al_s.empty() || !errors) return (int)p->al_s.size();
    CTBA_CUDA(p, cudaSetDevice(p->device));
    return align_evaluate(p, nullptr, nullptr, errors);
  }
  if ((kind < CTBA_REPROJ || kind > CTBA_GP) && kind != CTBA_PRIOR) { p->err = "evaluate_errors: kind must be reprojection, accel, gyro, global position or pose prior"; return CTBA_E_ARG; }
  if (p->device < 0) { p->err = "planning-only handle (device -1): no compute entry points"; return CTBA_E_CUDA; }
  CTBA_CUDA(p, cudaSetDevice(p->device));
  // accel / gyro / global position: the reference re-derives every sample's window from the current (possibly folded) time offsets and
  // re-applies the skip rules (bspline_so3r3.h:1239-1256, 1346-1363, 1692-1707); the solver's own work items keep the windows they were
  // added with, so the error pass runs on rebuilt work items and the solver's are swapped back afterwards.
  MeasBackup backup;
  const bool remap = (kind == CTBA_ACCEL || kind == CTBA_GYRO || kind == CTBA_GP) && !p->sharded;
  struct Restore { ctba_problem* p; MeasBackup* b; ~Restore() { remap_measurements_end(p, b); } } restore{p, &backup};
  if (remap) remap_measurements_begin(p, &backup);
  int rc = sync_to_device(p); if (rc) return rc;
  rc = launch_evaluate(p, p->ev, p->dstate.ptrs(), false); if (rc) return rc;
  cudaStream_t st = p->stream;
  int n = 0, nres = 3;
  const double *r = nullptr, *w = nullptr;
  const int32_t* idx = nullptr;
  switch (kind) {
    case CTBA_REPROJ: n = (int)p->numObs(); nres = 2; r = p->ev.r_reproj.p; w = p->d_f_inv_std.p; idx = p->d_o_frame.p; break;
    case CTBA_ACCEL: n = (int)p->i_s.size(); r = p->ev.r_acc.p; w = p->d_i_inv_a.p; break;
    case CTBA_GYRO: n = (int)p->i_s.size(); r = p->ev.r_gyr.p; w = p->d_i_inv_g.p; break;
    case CTBA_PRIOR: n = (int)p->q_s.size(); r = p->ev.r_prior.p; w = p->d_q_inv_std.p; break;   // evaluatePositionErrors / evaluateOrientationErrors :1581-1680
    default: n = (int)p->g_s.size(); r = p->ev.r_gp.p; w = p->d_g_inv_std.p; break;
  }
  if (n == 0 || !errors) return n;
  DBuf<double> out;
  CTBA_CUDA(p, out.resize(n));
  k_error_norms<<<(n + 255) / 256, 256, 0, st>>>(n, nres, r, w, idx, out.p);
  p->launches++;
  cudaError_t e = cudaMemcpyAsync(errors, out.p, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  out.release();
  if (e != cudaSuccess) { p->err = cudaGetErrorString(e); return CTBA_E_CUDA; }
  return n;
}

int ctba_sample(ctba_problem* p, int what, int64_t n, const int64_t* t_ns, double* out) {
  if (!p) return CTBA_E_ARG;
  if (what < CTBA_SAMPLE_POSE || what > CTBA_SAMPLE_GYRO_BIAS || n < 0 || (n > 0 && (!t_ns || !out))) { p->err = "sample: bad arguments"; return CTBA_E_ARG; }
  if (p->device < 0) { p->err = "planning-only handle (device -1): no compute entry points"; return CTBA_E_CUDA; }
  if (n == 0) return CTBA_OK;
  const bool bias = what == CTBA_SAMPLE_ACCEL_BIAS || what == CTBA_SAMPLE_GYRO_BIAS;
  const int64_t dt = bias ? p->dt_ns_bias : p->dt_ns;
  const size_t nk = bias ? p->numBiasKnots() : p->numKnots();
  for (int64_t i = 0; i < n; ++i) {                        // the reference getters assert on these (bspline_so3r3.h:126-136, 225-235, 271-280)
    const int64_t st_ns = t_ns[i] - p->start_t_ns;
    if (st_ns < 0 || (size_t)(st_ns / dt + p->N) > nk) { p->err = "sample: time outside the spline"; return CTBA_E_ARG; }
  }
  CTBA_CUDA(p, cudaSetDevice(p->device));
  int rc = sync_to_device(p); if (rc) return rc;
  cudaStream_t st = p->stream;
  const int w = what == CTBA_SAMPLE_POSE ? 7 : 3;
  DBuf<double> d_out;
  DBuf<int64_t> d_t;
  CTBA_CUDA(p, d_out.resize((size_t)n * w));
  CTBA_CUDA(p, d_t.resize((size_t)n));
  cudaError_t e = cudaMemcpyAsync(d_t.p, t_ns, (size_t)n * sizeof(int64_t), cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) {
    const Globals G = make_globals(p);
    const int grid = (int)((n + 127) / 128);
    if (p->N == 4) k_sample<4><<<grid, 128, 0, st>>>(G, p->dstate.ptrs(), what, n, d_t.p, p->start_t_ns, p->dt_ns, p->dt_ns_bias, d_out.p);
    else k_sample<6><<<grid, 128, 0, st>>>(G, p->dstate.ptrs(), what, n, d_t.p, p->start_t_ns, p->dt_ns, p->dt_ns_bias, d_out.p);
    p->launches++;
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_out.p, (size_t)n * w * sizeof(double), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  d_out.release(); d_t.release();
  if (e != cudaSuccess) { p->err = cudaGetErrorString(e); return CTBA_E_CUDA; }
  return CTBA_OK;
}

}  // extern "C"
