// ctba_api.cu — the C ABI (include/ctba.h): problem assembly on the host, device mirrors, evaluate/solve entry points.
// Problem-assembly semantics (timestamp -> spline window, skip rules, clamps) restate
// /root/reference/include/spline/bspline_so3r3.h:355-369, 523-1046, 1807-1875 (cited per function in ctba.h).
#include <cmath>
#include <cstdlib>
#include <cstring>
#include "ctba_solver.h"

using namespace ctba;

namespace ctba {

int sync_to_device(ctba_problem* p) {
  if (p->device < 0) { p->err = "planning-only handle (device -1): no compute entry points"; return CTBA_E_CUDA; }
  if (p->comm && p->comm->world > 1 && !p->sharded) { const int rc = shard_build(p); if (rc) return rc; }
  cudaStream_t st = p->stream;
  if (p->meas_dirty) {
    CTBA_CUDA(p, p->d_f_s.upload(p->f_s, st)); CTBA_CUDA(p, p->d_f_u.upload(p->f_u, st)); CTBA_CUDA(p, p->d_f_inv_std.upload(p->f_inv_std, st));
    CTBA_CUDA(p, p->d_f_obs_begin.upload(p->f_obs_begin, st));
    CTBA_CUDA(p, p->d_o_uv.upload(p->o_uv, st)); CTBA_CUDA(p, p->d_o_lm.upload(p->o_lm, st)); CTBA_CUDA(p, p->d_o_frame.upload(p->o_frame, st));
    CTBA_CUDA(p, p->d_i_s.upload(p->i_s, st)); CTBA_CUDA(p, p->d_i_sb.upload(p->i_sb, st)); CTBA_CUDA(p, p->d_i_st_s.upload(p->i_st_s, st));
    CTBA_CUDA(p, p->d_i_k.upload(p->i_k, st)); CTBA_CUDA(p, p->d_i_ub.upload(p->i_ub, st)); CTBA_CUDA(p, p->d_i_inv_a.upload(p->i_inv_a, st));
    CTBA_CUDA(p, p->d_i_inv_g.upload(p->i_inv_g, st)); CTBA_CUDA(p, p->d_i_acc.upload(p->i_acc, st)); CTBA_CUDA(p, p->d_i_gyr.upload(p->i_gyr, st));
    CTBA_CUDA(p, p->d_g_s.upload(p->g_s, st)); CTBA_CUDA(p, p->d_g_st_s.upload(p->g_st_s, st)); CTBA_CUDA(p, p->d_g_k.upload(p->g_k, st));
    CTBA_CUDA(p, p->d_g_inv_std.upload(p->g_inv_std, st)); CTBA_CUDA(p, p->d_g_meas.upload(p->g_meas, st));
    CTBA_CUDA(p, p->d_b_s0.upload(p->b_s0, st)); CTBA_CUDA(p, p->d_b_s1.upload(p->b_s1, st)); CTBA_CUDA(p, p->d_b_which.upload(p->b_which, st));
    CTBA_CUDA(p, p->d_b_u0.upload(p->b_u0, st)); CTBA_CUDA(p, p->d_b_u1.upload(p->b_u1, st)); CTBA_CUDA(p, p->d_b_inv_std.upload(p->b_inv_std, st));
    CTBA_CUDA(p, p->d_q_s.upload(p->q_s, st)); CTBA_CUDA(p, p->d_q_which.upload(p->q_which, st)); CTBA_CUDA(p, p->d_q_u.upload(p->q_u, st));
    CTBA_CUDA(p, p->d_q_inv_std.upload(p->q_inv_std, st)); CTBA_CUDA(p, p->d_q_meas.upload(p->q_meas, st));
    p->meas_dirty = false;
  }
  if (p->state_dirty) {
    // zero padding so that 12-element bias window reads never leave the allocation even with zero bias knots
    auto up_padded = [&](DBuf<double>& d, const HVec<double>& h, size_t min_n) -> cudaError_t {
      cudaError_t e = d.resize(std::max(h.size(), min_n));
      if (e == cudaSuccess && h.size() < min_n) e = cudaMemsetAsync(d.p, 0, min_n * sizeof(double), st);
      if (e == cudaSuccess && !h.empty()) e = cudaMemcpyAsync(d.p, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice, st);
      return e;
    };
    CTBA_CUDA(p, p->dstate.knot_q.upload(p->knot_q, st)); CTBA_CUDA(p, p->dstate.knot_p.upload(p->knot_p, st));
    CTBA_CUDA(p, up_padded(p->dstate.bias_a, p->bias_a, 12)); CTBA_CUDA(p, up_padded(p->dstate.bias_g, p->bias_g, 12));
    CTBA_CUDA(p, up_padded(p->dstate.lm, p->lm, 3));
    std::vector<double> g(p->glb, p->glb + GLB_STATE_SIZE);
    CTBA_CUDA(p, p->dstate.glb.upload(g, st));
    CTBA_CUDA(p, cudaStreamSynchronize(st));  // the staging vectors above are temporaries
    p->state_dirty = false;
    p->host_state_stale = false;
  }
  return CTBA_OK;
}

int download_state(ctba_problem* p) {
  if (!p->host_state_stale || p->device < 0) return CTBA_OK;
  cudaStream_t st = p->stream;
  auto dl = [&](HVec<double>& h, const DBuf<double>& d) -> cudaError_t {
    if (h.empty()) return cudaSuccess;
    return cudaMemcpyAsync(h.data(), d.p, h.size() * sizeof(double), cudaMemcpyDeviceToHost, st);
  };
  CTBA_CUDA(p, dl(p->knot_q, p->dstate.knot_q)); CTBA_CUDA(p, dl(p->knot_p, p->dstate.knot_p));
  CTBA_CUDA(p, dl(p->bias_a, p->dstate.bias_a)); CTBA_CUDA(p, dl(p->bias_g, p->dstate.bias_g));
  CTBA_CUDA(p, dl(p->lm, p->dstate.lm));
  CTBA_CUDA(p, cudaMemcpyAsync(p->glb, p->dstate.glb.p, GLB_STATE_SIZE * sizeof(double), cudaMemcpyDeviceToHost, st));
  CTBA_CUDA(p, cudaStreamSynchronize(st));
  p->host_state_stale = false;
  return CTBA_OK;
}

}  // namespace ctba

#define CHECK_P(p) do { if (!(p)) return CTBA_E_ARG; } while (0)
static int fail(ctba_problem* p, int code, const char* msg) { p->err = msg; return code; }

namespace ctba {
namespace {
constexpr size_t kPinMin = 64 * 1024, kHdr = 64;
constexpr uint64_t kMagicPinned = 0x50494e4e45444d45ull, kMagicPlain = 0x504c41494e4d454dull;
bool can_pin() {
  static const bool ok = [] { int n = 0; return cudaGetDeviceCount(&n) == cudaSuccess && n > 0; }();
  return ok;
}
}  // namespace
void* host_alloc(size_t bytes) {
  char* base = nullptr;
  bool pinned = false;
  if (bytes >= kPinMin && can_pin()) {
    void* q = nullptr;
    if (cudaHostAlloc(&q, bytes + kHdr, cudaHostAllocPortable) == cudaSuccess) { base = static_cast<char*>(q); pinned = true; }
    else cudaGetLastError();
  }
  if (!base) base = static_cast<char*>(std::malloc(bytes + kHdr));
  if (!base) throw std::bad_alloc();
  *reinterpret_cast<uint64_t*>(base) = pinned ? kMagicPinned : kMagicPlain;
  return base + kHdr;
}
void host_free(void* p) {
  if (!p) return;
  char* base = static_cast<char*>(p) - kHdr;
  if (*reinterpret_cast<uint64_t*>(base) == kMagicPinned) cudaFreeHost(base); else std::free(base);
}
}  // namespace ctba

namespace ctba {
// planar [entry][sample] -> row-major [sample][entry] for ctba_evaluate's documented layout
__global__ void k_planar_to_rows(int n, int per, const double* __restrict__ in, double* __restrict__ out) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (size_t)n * per) return;
  const size_t i = t / per, e = t % per;
  out[t] = in[e * n + i];
}
}  // namespace ctba

extern "C" {

int ctba_create(int order, int64_t dt_ns, int64_t start_t_ns, int device, ctba_problem** out) {
  if (!out) return CTBA_E_ARG;
  *out = nullptr;
  if ((order != 4 && order != 6) || dt_ns <= 0) return CTBA_E_ARG;
  int ndev = 0;
  // device == -1: planning-only handle (host problem assembly + ctba_shard_plan; every compute entry point returns CTBA_E_CUDA)
  if (device != -1 && (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev)) return CTBA_E_CUDA;
  if (device != -1 && cudaSetDevice(device) != cudaSuccess) return CTBA_E_CUDA;
  ctba_problem* p = new ctba_problem();
  p->N = order; p->dt_ns = dt_ns; p->start_t_ns = start_t_ns; p->device = device;
  p->dt_s = dt_ns * 1e-9;
  p->inv_dt = 1e9 / dt_ns;
  if (device != -1) {
    if (cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking) != cudaSuccess) { delete p; return CTBA_E_CUDA; }
    p->own_stream = true;
  }
  *out = p;
  return CTBA_OK;
}

int ctba_destroy(ctba_problem* p) {
  CHECK_P(p);
  if (p->device < 0) { delete p->shard; delete p->gstruct; delete p; return CTBA_OK; }
  if (p->device >= 0) cudaSetDevice(p->device);
  cudaStreamSynchronize(p->stream);
  if (p->comm && p->comm_owned) delete p->comm;
  delete p->shard; delete p->gstruct;
  p->d_fsq_glob.release(); p->d_gather.release(); p->d_stage.release();
  if (p->solver) solver_free(p->solver);
  align_free(p->align);
  DBuf<double>* dd[] = {&p->dstate.knot_q, &p->dstate.knot_p, &p->dstate.bias_a, &p->dstate.bias_g, &p->dstate.lm, &p->dstate.glb,
                        &p->d_f_u, &p->d_f_inv_std, &p->d_o_uv, &p->d_i_st_s, &p->d_i_k, &p->d_i_ub, &p->d_i_inv_a, &p->d_i_inv_g, &p->d_i_acc,
                        &p->d_i_gyr, &p->d_g_st_s, &p->d_g_k, &p->d_g_inv_std, &p->d_g_meas, &p->d_b_u0, &p->d_b_u1, &p->d_b_inv_std,
                        &p->ev.frame_pose, &p->ev.r_reproj, &p->ev.Jl, &p->ev.f_sq, &p->ev.f_w, &p->ev.f_cost, &p->ev.r_acc, &p->ev.r_gyr,
                        &p->ev.J_acc, &p->ev.J_gyr, &p->ev.r_gp, &p->ev.J_gp, &p->ev.r_bias, &p->ev.J_bias, &p->ev.r_grav, &p->ev.J_grav,
                        &p->ev.partial, &p->ev.cost,
                        &p->ev_alt.frame_pose, &p->ev_alt.r_reproj, &p->ev_alt.Jl, &p->ev_alt.f_sq, &p->ev_alt.f_w, &p->ev_alt.f_cost, &p->ev_alt.r_acc,
                        &p->ev_alt.r_gyr, &p->ev_alt.J_acc, &p->ev_alt.J_gyr, &p->ev_alt.r_gp, &p->ev_alt.J_gp, &p->ev_alt.r_bias, &p->ev_alt.J_bias,
                        &p->ev_alt.r_grav, &p->ev_alt.J_grav, &p->ev_alt.partial, &p->ev_alt.cost,
                        &p->ev.r_prior, &p->ev.J_prior, &p->ev_alt.r_prior, &p->ev_alt.J_prior, &p->d_q_u, &p->d_q_inv_std, &p->d_q_meas};
  for (auto* d : dd) d->release();
  DBuf<int32_t>* di[] = {&p->d_f_s, &p->d_f_obs_begin, &p->d_o_lm, &p->d_o_frame, &p->d_i_s, &p->d_i_sb, &p->d_g_s, &p->d_b_s0, &p->d_b_s1, &p->d_b_which,
                         &p->d_q_s, &p->d_q_which};
  for (auto* d : di) d->release();
  if (p->own_stream) cudaStreamDestroy(p->stream);
  delete p;
  return CTBA_OK;
}

const char* ctba_last_error(const ctba_problem* p) { return p ? p->err.c_str() : "null handle"; }

int ctba_set_stream(ctba_problem* p, void* s) {
  CHECK_P(p);
  cudaStreamSynchronize(p->stream);
  if (s) {
    if (p->own_stream) cudaStreamDestroy(p->stream);
    p->stream = (cudaStream_t)s; p->own_stream = false;
  } else if (!p->own_stream) {
    CTBA_CUDA(p, cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking));
    p->own_stream = true;
  }
  return CTBA_OK;
}

int ctba_set_knots(ctba_problem* p, int n, const double* q, const double* pos) {
  CHECK_P(p);
  if (n < 0 || !q || !pos) return fail(p, CTBA_E_ARG, "set_knots: bad arguments");
  if (!p->f_s.empty() || !p->i_s.empty() || !p->g_s.empty() || !p->q_s.empty() || !p->al_s.empty() || !p->b_s0.empty()) {
    if ((size_t)n != p->numKnots()) return fail(p, CTBA_E_STATE, "set_knots: knot count cannot change after measurements were added");
  }
  // after a solve the device holds the optimised landmarks / biases / globals: bring them to the host before the host copy becomes the
  // master again (otherwise the next upload would overwrite them with the pre-solve values, or a getter would overwrite the new knots)
  if (p->device >= 0) cudaSetDevice(p->device);
  { int rc = download_state(p); if (rc) return rc; }
  if ((size_t)n != p->numKnots()) p->struct_dirty = true;
  p->knot_q.assign(q, q + 4 * (size_t)n); p->knot_p.assign(pos, pos + 3 * (size_t)n);
  p->state_dirty = true;
  return CTBA_OK;
}
int ctba_num_knots(const ctba_problem* p) { return p ? (int)p->numKnots() : CTBA_E_ARG; }
int ctba_get_knots(ctba_problem* p, double* q, double* pos) {
  CHECK_P(p);
  if (p->device >= 0) cudaSetDevice(p->device);
  int rc = download_state(p); if (rc) return rc;
  if (q) std::memcpy(q, p->knot_q.data(), p->knot_q.size() * sizeof(double));
  if (pos) std::memcpy(pos, p->knot_p.data(), p->knot_p.size() * sizeof(double));
  return CTBA_OK;
}

int ctba_initialize_biases(ctba_problem* p, double dt_s_bias, const double* a0, const double* g0) {
  CHECK_P(p);
  if (!(dt_s_bias > 0) || !a0 || !g0) return fail(p, CTBA_E_ARG, "initialize_biases: bad arguments");
  if (!p->i_s.empty() || !p->b_s0.empty()) return fail(p, CTBA_E_STATE, "initialize_biases: IMU measurements already added");
  if (p->device >= 0) cudaSetDevice(p->device);
  { int rc = download_state(p); if (rc) return rc; }
  p->dt_ns_bias = (int64_t) static_cast<uint64_t>(dt_s_bias / 1e-9);  // :363
  p->inv_dt_bias = 1e9 / p->dt_ns_bias;
  double d = p->dt_s * (double)(p->knot_p.size() / 3);
  int num_knots = (int)(d / dt_s_bias);
  p->struct_dirty = true;
  p->bias_a.resize(3 * (size_t)num_knots); p->bias_g.resize(3 * (size_t)num_knots);
  for (int i = 0; i < num_knots; ++i) for (int c = 0; c < 3; ++c) { p->bias_a[3 * i + c] = a0[c]; p->bias_g[3 * i + c] = g0[c]; }
  p->state_dirty = true;
  return num_knots;
}
int ctba_num_bias_knots(const ctba_problem* p) { return p ? (int)p->numBiasKnots() : CTBA_E_ARG; }
int ctba_set_biases(ctba_problem* p, const double* a, const double* g) {
  CHECK_P(p);
  if (!a || !g) return CTBA_E_ARG;
  if (p->device >= 0) cudaSetDevice(p->device);
  int rc = download_state(p); if (rc) return rc;
  std::memcpy(p->bias_a.data(), a, p->bias_a.size() * sizeof(double)); std::memcpy(p->bias_g.data(), g, p->bias_g.size() * sizeof(double));
  p->state_dirty = true;
  return CTBA_OK;
}
int ctba_get_biases(ctba_problem* p, double* a, double* g) {
  CHECK_P(p);
  if (p->device >= 0) cudaSetDevice(p->device);
  int rc = download_state(p); if (rc) return rc;
  if (a) std::memcpy(a, p->bias_a.data(), p->bias_a.size() * sizeof(double));
  if (g) std::memcpy(g, p->bias_g.data(), p->bias_g.size() * sizeof(double));
  return CTBA_OK;
}

int ctba_set_landmarks(ctba_problem* p, int n, const double* xyz) {
  CHECK_P(p);
  if (n < 0 || !xyz) return fail(p, CTBA_E_ARG, "set_landmarks: bad arguments");
  for (int32_t l : p->o_lm) if (l >= n) return fail(p, CTBA_E_ARG, "set_landmarks: existing observations reference a landmark beyond n");
  if (p->device >= 0) cudaSetDevice(p->device);
  int rc = download_state(p); if (rc) return rc;
  if ((size_t)n * 3 != p->lm.size()) p->struct_dirty = true;
  p->lm.assign(xyz, xyz + 3 * (size_t)n);
  p->state_dirty = true;
  return CTBA_OK;
}
int ctba_get_landmarks(ctba_problem* p, double* xyz) {
  CHECK_P(p);
  if (p->device >= 0) cudaSetDevice(p->device);
  int rc = download_state(p); if (rc) return rc;
  if (xyz) std::memcpy(xyz, p->lm.data(), p->lm.size() * sizeof(double));
  return CTBA_OK;
}

int ctba_set_camera(ctba_problem* p, int model, const double* intr, const double* dist, const double* q_bc, const double* p_bc, double t_off) {
  CHECK_P(p);
  if (!intr || !dist || !q_bc || !p_bc) return fail(p, CTBA_E_ARG, "set_camera: null pointer");
  if (model != 4 && model != 5 && model != 7) return fail(p, CTBA_E_ARG, "set_camera: camera model id not supported (camera_calib.h:241)");
  p->cam.model = model; p->cam.fx = intr[0]; p->cam.fy = intr[1]; p->cam.cx = intr[2]; p->cam.cy = intr[3];
  for (int i = 0; i < 4; ++i) { p->cam.d[i] = dist[i]; p->q_bc0[i] = q_bc[i]; }
  for (int i = 0; i < 3; ++i) p->p_bc0[i] = p_bc[i];
  p->cam_t_offset_cam_imu = t_off;
  p->t_offset_cam_imu_s = t_off;
  return CTBA_OK;
}

int ctba_set_globals(ctba_problem* p, double scale, const double* p_bp, const double* g, double t_s_gp_s) {
  CHECK_P(p);
  if (!p_bp || !g) return CTBA_E_ARG;
  if (p->device >= 0) cudaSetDevice(p->device);
  int rc = download_state(p); if (rc) return rc;
  p->scale = scale;
  for (int i = 0; i < 3; ++i) { p->glb[GLB_PBP + i] = p_bp[i]; p->glb[GLB_G + i] = g[i]; }
  p->t_s_gp_init_s = t_s_gp_s; p->t_s_gp_s = t_s_gp_s;
  p->state_dirty = true;
  return CTBA_OK;
}
int ctba_set_corrections(ctba_problem* p, const double* c) {
  CHECK_P(p);
  if (!c) return CTBA_E_ARG;
  if (p->device >= 0) cudaSetDevice(p->device);
  int rc = download_state(p); if (rc) return rc;
  p->glb[GLB_SCALE_CORR] = c[0];
  for (int i = 0; i < 4; ++i) p->glb[GLB_RBC_CORR + i] = c[1 + i];
  for (int i = 0; i < 3; ++i) p->glb[GLB_PBC_CORR + i] = c[5 + i];
  p->glb[GLB_T_CAM_IMU_CORR] = c[8]; p->glb[GLB_T_GP_CORR] = c[9];
  p->state_dirty = true;
  return CTBA_OK;
}
int ctba_get_globals(ctba_problem* p, double* o) {
  CHECK_P(p);
  if (!o) return CTBA_E_ARG;
  if (p->device >= 0) cudaSetDevice(p->device);
  int rc = download_state(p); if (rc) return rc;
  o[0] = p->glb[GLB_SCALE_CORR];
  for (int i = 0; i < 4; ++i) o[1 + i] = p->glb[GLB_RBC_CORR + i];
  for (int i = 0; i < 3; ++i) o[5 + i] = p->glb[GLB_PBC_CORR + i];
  o[8] = p->glb[GLB_T_CAM_IMU_CORR]; o[9] = p->glb[GLB_T_GP_CORR];
  for (int i = 0; i < 3; ++i) { o[10 + i] = p->glb[GLB_PBP + i]; o[13 + i] = p->glb[GLB_G + i]; }
  o[16] = p->t_offset_cam_imu_s; o[17] = p->t_s_gp_s; o[18] = p->scale;
  return CTBA_OK;
}
int ctba_set_flags(ctba_problem* p, int s, int pbp, int g, int ci, int tgs) {
  CHECK_P(p);
  p->struct_dirty = true;
  p->optimize_scale = s; p->optimize_p_bp = pbp; p->optimize_gravity = g; p->optimize_camimu_extrinsics = ci; p->optimize_t_offset_s_gs = tgs;
  return CTBA_OK;
}

int ctba_add_frames(ctba_problem* p, int F, const int64_t* t_ns, const int32_t* obs_begin, const double* uv, const int32_t* lmi, double noise) {
  CHECK_P(p);
  if (F < 0 || !t_ns || !obs_begin || (!uv && F) || (!lmi && F) || !(noise > 0)) return fail(p, CTBA_E_ARG, "add_frames: bad arguments");
  const int N = p->N;
  const int64_t L = (int64_t)p->lm.size() / 3;
  int added = 0;
  for (int f = 0; f < F; ++f) {
    if (obs_begin[f + 1] < obs_begin[f]) return fail(p, CTBA_E_ARG, "add_frames: obs_begin not monotone");
    int64_t st_ns = t_ns[f] - p->start_t_ns;
    if (st_ns < 0) continue;                                    // :964
    const int n = obs_begin[f + 1] - obs_begin[f];
    if (n == 0) continue;                                       // :967
    int64_t s = st_ns / p->dt_ns;
    double u = double(st_ns % p->dt_ns) / double(p->dt_ns);     // :969-970
    if (size_t(s + N) > p->numKnots()) continue;                // :973
    for (int i = obs_begin[f]; i < obs_begin[f + 1]; ++i)
      if (lmi[i] < 0 || lmi[i] >= L) return fail(p, CTBA_E_ARG, "add_frames: Point 3D not found (bspline_so3r3.h:1024)");
    const int fid = (int)p->f_s.size();
    p->f_s.push_back((int32_t)s); p->f_u.push_back(u); p->f_inv_std.push_back(1.0 / noise);
    for (int i = obs_begin[f]; i < obs_begin[f + 1]; ++i) {
      p->o_uv.push_back(uv[2 * (size_t)i]); p->o_uv.push_back(uv[2 * (size_t)i + 1]);
      p->o_lm.push_back(lmi[i]); p->o_frame.push_back(fid);
    }
    p->f_obs_begin.push_back((int32_t)p->o_lm.size());
    ++added;
  }
  p->meas_dirty = true; p->struct_dirty = true;
  return added;
}

static int64_t end_t_ns_bias(const ctba_problem* p) {  // getEndTnsBiasSpline :1870-1875 (N_, not Nb_)
  int a = (int)p->numBiasKnots() - p->N + 1;
  int64_t t_end_ns = (int64_t)a * p->dt_ns_bias + p->start_t_ns;
  return (int64_t)((double)t_end_ns - 1e6);   // the reference subtracts the double literal 1e6: the result is rounded through double (:1874)
}

// window mapping of one IMU sample (bspline_so3r3.h:525-563); false = skipped
static bool map_imu_sample(const ctba_problem* p, int64_t t_ns, int32_t* s_out, int32_t* sb_out, double* st_s_out, double* k_out, double* ub_out) {
  const int N = p->N;
  int64_t t_offset_imu_cam_ns = static_cast<int64_t>(-p->t_offset_cam_imu_s * 1e9);  // :525-526
  int64_t st_ns = (t_ns + t_offset_imu_cam_ns - p->start_t_ns);
  double st_s = st_ns * 1e-9;
  if (st_ns < 0) return false;
  int64_t s = st_ns / p->dt_ns;
  int k_int = (int)(st_ns / p->dt_ns);
  if (size_t(s + N) > p->numKnots()) return false;
  int64_t st_bias_ns = (t_ns - p->start_t_ns);
  int64_t s_bias = st_bias_ns / p->dt_ns_bias;
  double u_bias = double(st_bias_ns % p->dt_ns_bias) / double(p->dt_ns_bias);
  if (s_bias < 0) { *sb_out = -1; return true; }
  if (size_t(s_bias + N) > p->numBiasKnots()) {  // :557-563
    st_bias_ns = end_t_ns_bias(p) - p->start_t_ns;
    s_bias = st_bias_ns / p->dt_ns_bias;
    u_bias = double(st_bias_ns % p->dt_ns_bias) / double(p->dt_ns_bias);
  }
  *s_out = (int32_t)s; *sb_out = (int32_t)s_bias; *st_s_out = st_s; *k_out = (double)k_int; *ub_out = u_bias;
  return true;
}
// window mapping of one global-position sample (:882-895)
static bool map_gp_sample(const ctba_problem* p, int64_t t_ns, int32_t* s_out, double* st_s_out, double* k_out) {
  int64_t t_s_gp_init_ns = static_cast<int64_t>(p->t_s_gp_s * 1e9);  // :882
  int64_t st_ns = (t_ns - p->start_t_ns - t_s_gp_init_ns);
  double st_s = st_ns * 1e-9;
  if (st_ns < 0) return false;
  int64_t s = st_ns / p->dt_ns;
  int k_int = (int)(st_ns / p->dt_ns);
  if (size_t(s + p->N) > p->numKnots()) return false;
  *s_out = (int32_t)s; *st_s_out = st_s; *k_out = (double)k_int;
  return true;
}

int ctba_add_imu(ctba_problem* p, int n, const int64_t* t_ns, const double* acc, const double* gyr, double acc_nd, double gyr_nd) {
  CHECK_P(p);
  if (n < 0 || !t_ns || !acc || !gyr || !(acc_nd > 0) || !(gyr_nd > 0)) return fail(p, CTBA_E_ARG, "add_imu: bad arguments");
  if ((int)p->numBiasKnots() < p->N) return fail(p, CTBA_E_STATE, "add_imu: initialize_biases first (needs >= N bias knots)");
  int added = 0;
  for (int i = 0; i < n; ++i) {
    int32_t s = 0, sb = 0; double st_s = 0, k = 0, ub = 0;
    p->raw_imu_t.push_back(t_ns[i]);
    for (int c = 0; c < 3; ++c) { p->raw_imu_acc.push_back(acc[3 * (size_t)i + c]); p->raw_imu_gyr.push_back(gyr[3 * (size_t)i + c]); }
    if (!map_imu_sample(p, t_ns[i], &s, &sb, &st_s, &k, &ub)) continue;
    if (sb < 0) return fail(p, CTBA_E_ARG, "add_imu: s_bias < 0 (bspline_so3r3.h:548)");
    p->i_s.push_back(s); p->i_sb.push_back(sb);
    p->i_st_s.push_back(st_s); p->i_k.push_back(k); p->i_ub.push_back(ub);
    p->i_inv_a.push_back(1.0 / acc_nd); p->i_inv_g.push_back(1.0 / gyr_nd);
    for (int c = 0; c < 3; ++c) { p->i_acc.push_back(acc[3 * (size_t)i + c]); p->i_gyr.push_back(gyr[3 * (size_t)i + c]); }
    ++added;
  }
  p->raw_imu_inv_a = 1.0 / acc_nd; p->raw_imu_inv_g = 1.0 / gyr_nd;
  p->meas_dirty = true; p->struct_dirty = true;
  return added;
}

int ctba_add_gp(ctba_problem* p, int n, const int64_t* t_ns, const double* pos, double noise) {
  CHECK_P(p);
  if (n < 0 || !t_ns || !pos || !(noise > 0)) return fail(p, CTBA_E_ARG, "add_gp: bad arguments");
  int added = 0;
  for (int i = 0; i < n; ++i) {
    int32_t s = 0; double st_s = 0, k = 0;
    p->raw_gp_t.push_back(t_ns[i]);
    for (int c = 0; c < 3; ++c) p->raw_gp_pos.push_back(pos[3 * (size_t)i + c]);
    if (!map_gp_sample(p, t_ns[i], &s, &st_s, &k)) continue;
    p->g_s.push_back(s); p->g_st_s.push_back(st_s); p->g_k.push_back(k); p->g_inv_std.push_back(1.0 / noise);
    for (int c = 0; c < 3; ++c) p->g_meas.push_back(pos[3 * (size_t)i + c]);
    ++added;
  }
  p->raw_gp_inv = 1.0 / noise;
  p->meas_dirty = true; p->struct_dirty = true;
  return added;
}

}  // extern "C"
namespace ctba {
int remap_measurements_begin(ctba_problem* p, MeasBackup* b) {
  b->i_s.swap(p->i_s); b->i_sb.swap(p->i_sb); b->g_s.swap(p->g_s);
  b->i_st_s.swap(p->i_st_s); b->i_k.swap(p->i_k); b->i_ub.swap(p->i_ub); b->i_inv_a.swap(p->i_inv_a); b->i_inv_g.swap(p->i_inv_g);
  b->i_acc.swap(p->i_acc); b->i_gyr.swap(p->i_gyr); b->g_st_s.swap(p->g_st_s); b->g_k.swap(p->g_k); b->g_inv_std.swap(p->g_inv_std); b->g_meas.swap(p->g_meas);
  b->active = true;
  for (size_t i = 0; i < p->raw_imu_t.size(); ++i) {
    int32_t s = 0, sb = 0; double st_s = 0, k = 0, ub = 0;
    if (!map_imu_sample(p, p->raw_imu_t[i], &s, &sb, &st_s, &k, &ub) || sb < 0) continue;
    p->i_s.push_back(s); p->i_sb.push_back(sb); p->i_st_s.push_back(st_s); p->i_k.push_back(k); p->i_ub.push_back(ub);
    p->i_inv_a.push_back(p->raw_imu_inv_a); p->i_inv_g.push_back(p->raw_imu_inv_g);
    for (int c = 0; c < 3; ++c) { p->i_acc.push_back(p->raw_imu_acc[3 * i + c]); p->i_gyr.push_back(p->raw_imu_gyr[3 * i + c]); }
  }
  for (size_t i = 0; i < p->raw_gp_t.size(); ++i) {
    int32_t s = 0; double st_s = 0, k = 0;
    if (!map_gp_sample(p, p->raw_gp_t[i], &s, &st_s, &k)) continue;
    p->g_s.push_back(s); p->g_st_s.push_back(st_s); p->g_k.push_back(k); p->g_inv_std.push_back(p->raw_gp_inv);
    for (int c = 0; c < 3; ++c) p->g_meas.push_back(p->raw_gp_pos[3 * i + c]);
  }
  p->meas_dirty = true;
  return CTBA_OK;
}
void remap_measurements_end(ctba_problem* p, MeasBackup* b) {
  if (!b->active) return;
  b->i_s.swap(p->i_s); b->i_sb.swap(p->i_sb); b->g_s.swap(p->g_s);
  b->i_st_s.swap(p->i_st_s); b->i_k.swap(p->i_k); b->i_ub.swap(p->i_ub); b->i_inv_a.swap(p->i_inv_a); b->i_inv_g.swap(p->i_inv_g);
  b->i_acc.swap(p->i_acc); b->i_gyr.swap(p->i_gyr); b->g_st_s.swap(p->g_st_s); b->g_k.swap(p->g_k); b->g_inv_std.swap(p->g_inv_std); b->g_meas.swap(p->g_meas);
  b->active = false;
  p->meas_dirty = true;
}
}  // namespace ctba
extern "C" {

// addWorldColmapAlignmentMeasurement (bspline_so3r3.h:1121-1211): the window arithmetic and skip rules of addGlobalPosMeasurement
int ctba_add_alignment_measurements(ctba_problem* p, int n, const int64_t* t_ns, const double* pos, double noise) {
  CHECK_P(p);
  if (n < 0 || !t_ns || !pos || !(noise > 0)) return fail(p, CTBA_E_ARG, "add_alignment_measurements: bad arguments");
  const int N = p->N;
  int added = 0;
  for (int i = 0; i < n; ++i) {
    int64_t t_s_gp_init_ns = static_cast<int64_t>(p->t_s_gp_s * 1e9);  // :1127
    int64_t st_ns = (t_ns[i] - p->start_t_ns - t_s_gp_init_ns);
    double st_s = st_ns * 1e-9;
    if (st_ns < 0) continue;
    int64_t s = st_ns / p->dt_ns;
    int k_int = (int)(st_ns / p->dt_ns);
    if (size_t(s + N) > p->numKnots()) continue;
    p->al_s.push_back((int32_t)s); p->al_st_s.push_back(st_s); p->al_k.push_back((double)k_int); p->al_inv_std.push_back(1.0 / noise);
    for (int c = 0; c < 3; ++c) p->al_meas.push_back(pos[3 * (size_t)i + c]);
    ++added;
  }
  p->align_dirty = true;
  return added;
}
int ctba_num_alignment_measurements(const ctba_problem* p) { return p ? (int)p->al_s.size() : CTBA_E_ARG; }

// setWorldColmapTransformation (:1825); corr7 (tests: a non-identity linearisation point) = R_w_g_corr (xyzw) | p_w_g_corr, NULL keeps it
int ctba_set_world_alignment(ctba_problem* p, const double* T_wg7, const double* corr7) {
  CHECK_P(p);
  if (!T_wg7) return fail(p, CTBA_E_ARG, "set_world_alignment: T_wg is NULL");
  for (int i = 0; i < 7; ++i) p->T_wg[i] = T_wg7[i];
  if (corr7) { for (int i = 0; i < 4; ++i) p->R_wg_corr[i] = corr7[i]; for (int i = 0; i < 3; ++i) p->p_wg_corr[i] = corr7[4 + i]; }
  return CTBA_OK;
}
// getWorldColmapTransformation (:1827-1834): (R(T_w_g) R_w_g_corr, t(T_w_g) + p_w_g_corr); getWorldColmapTransformationCorr (:1836-1840)
int ctba_get_world_alignment(ctba_problem* p, double* T_wg7, double* corr7) {
  CHECK_P(p);
  if (T_wg7) {
    const quat q = quat_mul(quat{p->T_wg[0], p->T_wg[1], p->T_wg[2], p->T_wg[3]}, quat{p->R_wg_corr[0], p->R_wg_corr[1], p->R_wg_corr[2], p->R_wg_corr[3]});
    T_wg7[0] = q.x; T_wg7[1] = q.y; T_wg7[2] = q.z; T_wg7[3] = q.w;
    for (int i = 0; i < 3; ++i) T_wg7[4 + i] = p->T_wg[4 + i] + p->p_wg_corr[i];
  }
  if (corr7) { for (int i = 0; i < 4; ++i) corr7[i] = p->R_wg_corr[i]; for (int i = 0; i < 3; ++i) corr7[4 + i] = p->p_wg_corr[i]; }
  return CTBA_OK;
}

// addPositionMeasurement (bspline_so3r3.h:1048-1085) / addOrientationMeasurement (:1087-1123): same window arithmetic and skip rules
static int add_priors(ctba_problem* p, int which, int n, const int64_t* t_ns, const double* meas, double noise) {
  const int N = p->N, w = which ? 4 : 3;
  int added = 0;
  for (int i = 0; i < n; ++i) {
    const int64_t st_ns = t_ns[i] - p->start_t_ns;
    if (st_ns < 0) continue;
    const int64_t s = st_ns / p->dt_ns;
    const double u = double(st_ns % p->dt_ns) / double(p->dt_ns);
    if (size_t(s + N) > p->numKnots()) continue;
    p->q_s.push_back((int32_t)s); p->q_which.push_back(which); p->q_u.push_back(u); p->q_inv_std.push_back(1.0 / noise);
    for (int c = 0; c < 4; ++c) p->q_meas.push_back(c < w ? meas[(size_t)w * i + c] : 0.0);
    ++added;
  }
  p->meas_dirty = true; p->struct_dirty = true;
  return added;
}
int ctba_add_position_measurements(ctba_problem* p, int n, const int64_t* t_ns, const double* pos, double noise) {
  CHECK_P(p);
  if (n < 0 || !t_ns || !pos || !(noise > 0)) return fail(p, CTBA_E_ARG, "add_position_measurements: bad arguments");
  return add_priors(p, 0, n, t_ns, pos, noise);
}
int ctba_add_orientation_measurements(ctba_problem* p, int n, const int64_t* t_ns, const double* quat_xyzw, double noise) {
  CHECK_P(p);
  if (n < 0 || !t_ns || !quat_xyzw || !(noise > 0)) return fail(p, CTBA_E_ARG, "add_orientation_measurements: bad arguments");
  return add_priors(p, 1, n, t_ns, quat_xyzw, noise);
}

int ctba_num_priors(const ctba_problem* p) { return p ? (int)p->q_s.size() : CTBA_E_ARG; }

int ctba_add_bias_cost(ctba_problem* p, double freq_hz, double bias_err_dt_s, double acc_rw, double gyr_rw) {
  CHECK_P(p);
  if (!(freq_hz > 0) || !(acc_rw > 0) || !(gyr_rw > 0)) return fail(p, CTBA_E_ARG, "add_bias_cost: bad arguments");
  // the two 4-knot windows of a block must be disjoint (Ceres rejects duplicate parameter blocks;
  // optimize_vi_continuous_time.cpp:489-490 asserts bias_err_dt_s >= 4*dt_bias)
  if (bias_err_dt_s * 1e9 < 4.0 * (double)p->dt_ns_bias) return fail(p, CTBA_E_ARG, "add_bias_cost: bias_err_dt_s must be >= 4 * dt_bias");
  const int Nb = 4;
  double err_freq_s = 1.0 / freq_hz;
  double err_freq_ns = (double) static_cast<uint64_t>(err_freq_s / 1e-9);
  int64_t t_start_ns = p->start_t_ns;
  int64_t t_end_ns = end_t_ns_bias(p);
  int64_t bias_err_dt_ns = static_cast<int64_t>(bias_err_dt_s * 1e9);
  int added = 0;
  for (int64_t t_ns = t_start_ns; t_ns < t_end_ns; t_ns += err_freq_ns) {
    int64_t st0 = t_ns - p->start_t_ns, st1 = t_ns + bias_err_dt_ns - p->start_t_ns;
    int64_t s0 = st0 / p->dt_ns_bias, s1 = st1 / p->dt_ns_bias;
    double u0 = double(st0 % p->dt_ns_bias) / double(p->dt_ns_bias), u1 = double(st1 % p->dt_ns_bias) / double(p->dt_ns_bias);
    if (size_t(s0 + Nb) > p->numBiasKnots()) continue;
    if (size_t(s1 + Nb) > p->numBiasKnots()) continue;
    for (int which = 0; which < 2; ++which) {
      p->b_s0.push_back((int32_t)s0); p->b_s1.push_back((int32_t)s1); p->b_which.push_back(which);
      p->b_u0.push_back(u0); p->b_u1.push_back(u1);
      p->b_inv_std.push_back(1.0 / (0.001 * (which == 0 ? acc_rw : gyr_rw)));  // :736, :765
    }
    ++added;
  }
  p->meas_dirty = true; p->struct_dirty = true;
  return added;
}

int ctba_add_gravity_norm(ctba_problem* p) { CHECK_P(p); p->gravity_norm = true; p->struct_dirty = true; return CTBA_OK; }

int ctba_counts(ctba_problem* p, int64_t* c) {
  CHECK_P(p);
  if (!c) return CTBA_E_ARG;
  c[0] = (int64_t)p->numFrames(); c[1] = (int64_t)p->numObs(); c[2] = (int64_t)p->i_s.size(); c[3] = (int64_t)p->i_s.size();
  c[4] = (int64_t)p->g_s.size(); c[5] = (int64_t)p->b_s0.size(); c[6] = p->gravity_norm ? 1 : 0;
  c[7] = solver_num_unknowns(p);
  return CTBA_OK;
}

int ctba_evaluate(ctba_problem* p, int kind, double* r_out, double* j_out) {
  CHECK_P(p);
  if (kind < 0 || kind > CTBA_ALIGN) return fail(p, CTBA_E_ARG, "evaluate: unknown residual kind");
  if (p->device < 0) return fail(p, CTBA_E_CUDA, "planning-only handle (device -1): no compute entry points");
  CTBA_CUDA(p, cudaSetDevice(p->device));
  if (kind == CTBA_ALIGN) {
    if (p->al_s.empty()) return 0;
    const int n = align_evaluate(p, r_out, j_out, nullptr);
    return n < 0 ? n : 3 * n;
  }
  int rc = sync_to_device(p); if (rc) return rc;
  const StatePtrs S = p->dstate.ptrs();
  rc = launch_evaluate(p, p->ev, S, true); if (rc) return rc;
  const int N = p->N;
  cudaStream_t st = p->stream;
  size_t nres = 0, width = 0;
  const double *d_r = nullptr, *d_J = nullptr;
  DBuf<double> dense;
  switch (kind) {
    case CTBA_REPROJ: {
      nres = 2 * p->numObs(); width = 6 * N + 10; d_r = p->ev.r_reproj.p;
      if (j_out && nres) {
        CTBA_CUDA(p, dense.resize(nres * width));
        rc = expand_reproj_dense(p, S, dense.p); if (rc) { dense.release(); return rc; }
        d_J = dense.p;
      }
      break;
    }
    case CTBA_ACCEL:
    case CTBA_GYRO: {
      const bool acc = kind == CTBA_ACCEL;
      nres = 3 * p->i_s.size(); width = acc ? 6 * N + 16 : 3 * N + 13; d_r = acc ? p->ev.r_acc.p : p->ev.r_gyr.p;
      if (j_out && nres) {     // the IMU Jacobians are stored planar on the device
        CTBA_CUDA(p, dense.resize(nres * width));
        const int n = (int)p->i_s.size(), per = 3 * (int)width;
        k_planar_to_rows<<<(unsigned)(((size_t)n * per + 255) / 256), 256, 0, st>>>(n, per, acc ? p->ev.J_acc.p : p->ev.J_gyr.p, dense.p);
        p->launches++;
        d_J = dense.p;
      }
      break;
    }
    case CTBA_GP: nres = 3 * p->g_s.size(); width = 6 * N + 5; d_r = p->ev.r_gp.p; d_J = p->ev.J_gp.p; break;
    case CTBA_BIAS: nres = 3 * p->b_s0.size(); width = 24; d_r = p->ev.r_bias.p; d_J = p->ev.J_bias.p; break;
    case CTBA_GRAVITY: nres = p->gravity_norm ? 1 : 0; width = 3; d_r = p->ev.r_grav.p; d_J = p->ev.J_grav.p; break;
    case CTBA_PRIOR: nres = 3 * p->q_s.size(); width = 3 * N; d_r = p->ev.r_prior.p; d_J = p->ev.J_prior.p; break;
  }
  cudaError_t e = cudaSuccess;
  if (nres && r_out) e = cudaMemcpyAsync(r_out, d_r, nres * sizeof(double), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess && nres && j_out) e = cudaMemcpyAsync(j_out, d_J, nres * width * sizeof(double), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  dense.release();
  if (e != cudaSuccess) { p->err = cudaGetErrorString(e); return CTBA_E_CUDA; }
  return (int)nres;
}

int ctba_cost(ctba_problem* p, double* cost7) {
  CHECK_P(p);
  if (!cost7) return CTBA_E_ARG;
  if (p->device < 0) return fail(p, CTBA_E_CUDA, "planning-only handle (device -1): no compute entry points");
  CTBA_CUDA(p, cudaSetDevice(p->device));
  int rc = sync_to_device(p); if (rc) return rc;
  rc = launch_evaluate(p, p->ev, p->dstate.ptrs(), false); if (rc) return rc;
  rc = launch_cost(p, p->ev); if (rc) return rc;
  double c8[8];
  rc = shard_reduce_cost(p, p->ev, c8); if (rc) return rc;   // sharded handles: the cost terms of the whole problem
  for (int i = 0; i < 7; ++i) cost7[i] = c8[i];
  return CTBA_OK;
}

int ctba_cost_terms(ctba_problem* p, double* cost8) {
  double c7[7];
  int rc = ctba_cost(p, c7); if (rc) return rc;
  if (!cost8) return CTBA_E_ARG;
  double c8[8];
  rc = shard_reduce_cost(p, p->ev, c8); if (rc) return rc;
  for (int i = 0; i < 8; ++i) cost8[i] = c8[i];
  return CTBA_OK;
}

int ctba_solve(ctba_problem* p, int max_num_iterations, ctba_summary* summary) {
  CHECK_P(p);
  if (max_num_iterations < 0 || !summary) return fail(p, CTBA_E_ARG, "solve: bad arguments");
  if (p->device < 0) return fail(p, CTBA_E_CUDA, "planning-only handle (device -1): no compute entry points");
  CTBA_CUDA(p, cudaSetDevice(p->device));
  if (!p->al_s.empty()) return align_solve(p, max_num_iterations, summary);   // align_spline_to_global_frame: frozen spline, <= 11 unknowns
  int rc = sync_to_device(p); if (rc) return rc;
  return solver_solve(p, max_num_iterations, summary);
}

int ctba_relinearize_time_offsets(ctba_problem* p) {
  CHECK_P(p);
  if (p->device >= 0) cudaSetDevice(p->device);
  int rc = download_state(p); if (rc) return rc;
  // :1807-1811, :1853-1857 — bookkeeping only: the work items keep the st_s, k captured when they were added
  if (p->relinearize_mask & 1) { p->t_offset_cam_imu_s += p->glb[GLB_T_CAM_IMU_CORR]; p->glb[GLB_T_CAM_IMU_CORR] = 0.0; }   // relinearizeSplineImuTimeOffset
  if (p->relinearize_mask & 2) { p->t_s_gp_s += p->glb[GLB_T_GP_CORR]; p->glb[GLB_T_GP_CORR] = 0.0; }                        // relinearizeSplineGlobSensTimeOffset
  p->state_dirty = true;
  return CTBA_OK;
}

int ctba_set_option(ctba_problem* p, const char* name, double value) {
  CHECK_P(p);
  if (!name) return CTBA_E_ARG;
  if (!std::strcmp(name, "bench_no_termination")) { p->bench_no_termination = value != 0.0; return CTBA_OK; }
  if (!std::strcmp(name, "linear_solver")) { p->linear_solver = value != 0.0 ? 1 : 0; return CTBA_OK; }
  if (!std::strcmp(name, "pcg_tol")) { if (!(value > 0)) return fail(p, CTBA_E_ARG, "pcg_tol must be > 0"); p->pcg_tol = value; return CTBA_OK; }
  if (!std::strcmp(name, "pcg_max_iters")) { p->pcg_max_iters = (int)value; return CTBA_OK; }
  if (!std::strcmp(name, "bench_accept_all")) { p->bench_accept_all = value != 0.0; return CTBA_OK; }
  if (!std::strcmp(name, "relinearize_mask")) { p->relinearize_mask = (int)value & 3; return CTBA_OK; }
  if (!std::strcmp(name, "band_partitions")) { p->band_partitions = (int)value; return CTBA_OK; }
  if (!std::strcmp(name, "schur_two_stage")) { p->schur_two_stage = value != 0.0; return CTBA_OK; }
  if (!std::strcmp(name, "schur_mask")) { p->schur_sorted = value != 0.0; return CTBA_OK; }
  if (!std::strcmp(name, "schur_thread_below")) { p->schur_thread_below = value; return CTBA_OK; }
  if (!std::strcmp(name, "schur_threads")) { p->schur_threads = (int)value; return CTBA_OK; }
  if (!std::strcmp(name, "schur_list_a")) { p->schur_list_a = value != 0.0; return CTBA_OK; }
  if (!std::strcmp(name, "schur_compact")) { p->schur_compact = value != 0.0; return CTBA_OK; }
  if (!std::strcmp(name, "band_nested")) { p->band_nested = value != 0.0; return CTBA_OK; }
  if (!std::strcmp(name, "band_graph")) { p->band_graph = value != 0.0; return CTBA_OK; }
  if (!std::strcmp(name, "cauchy_a")) { if (!(value > 0)) return fail(p, CTBA_E_ARG, "cauchy_a must be > 0"); p->cauchy_a = value; return CTBA_OK; }
  return fail(p, CTBA_E_ARG, "set_option: unknown option");
}

// read-only facts about the last solve / the solver's structure (bench and sweep reports)
int ctba_get_stat(ctba_problem* p, const char* name, double* value) {
  CHECK_P(p);
  if (!name || !value) return CTBA_E_ARG;
  SolverData* sd = p->solver;
  if (!sd || !sd->built) return fail(p, CTBA_E_STATE, "get_stat: no solve has run on this handle");
  if (!std::strcmp(name, "pcg_iterations")) { *value = sd->pcg_iterations; return CTBA_OK; }
  if (!std::strcmp(name, "pcg_rel_residual")) { *value = sd->pcg_rel_residual; return CTBA_OK; }
  if (!std::strcmp(name, "band_block_columns")) { *value = sd->Hw.nbc; return CTBA_OK; }
  if (!std::strcmp(name, "band_half_width_blocks")) { *value = sd->Hw.nbw; return CTBA_OK; }
  if (!std::strcmp(name, "band_launches")) { *value = (double)sd->band_launches; return CTBA_OK; }
  return fail(p, CTBA_E_ARG, "get_stat: unknown name");
}

int ctba_invalidate_device(ctba_problem* p) {
  CHECK_P(p);
  if (p->device >= 0) cudaSetDevice(p->device);
  int rc = download_state(p); if (rc) return rc;
  p->meas_dirty = true; p->state_dirty = true; p->align_dirty = true;
  return CTBA_OK;
}

int ctba_transfer_bytes(ctba_problem* p, int64_t* b) {
  CHECK_P(p);
  if (!b) return CTBA_E_ARG;
  const int64_t state = (int64_t)(p->knot_q.size() + p->knot_p.size() + p->bias_a.size() + p->bias_g.size() + p->lm.size() + GLB_STATE_SIZE) * 8;
  int64_t meas = (int64_t)(p->q_s.size() + p->q_which.size()) * 4 + (int64_t)(p->q_u.size() + p->q_inv_std.size() + p->q_meas.size()) * 8;
  meas += (int64_t)(p->f_s.size() + p->f_obs_begin.size() + p->o_lm.size() + p->o_frame.size() + p->i_s.size() + p->i_sb.size() + p->g_s.size() +
                    p->b_s0.size() + p->b_s1.size() + p->b_which.size()) * 4;
  meas += (int64_t)(p->f_u.size() + p->f_inv_std.size() + p->o_uv.size() + p->i_st_s.size() + p->i_k.size() + p->i_ub.size() + p->i_inv_a.size() +
                    p->i_inv_g.size() + p->i_acc.size() + p->i_gyr.size() + p->g_st_s.size() + p->g_k.size() + p->g_inv_std.size() + p->g_meas.size() +
                    p->b_u0.size() + p->b_u1.size() + p->b_inv_std.size()) * 8;
  b[0] = state + meas; b[1] = state;
  return CTBA_OK;
}

int ctba_bench_evaluate(ctba_problem* p, int reps, int64_t* launches) {
  CHECK_P(p);
  if (p->device < 0) return fail(p, CTBA_E_CUDA, "planning-only handle (device -1): no compute entry points");
  CTBA_CUDA(p, cudaSetDevice(p->device));
  int rc = sync_to_device(p); if (rc) return rc;
  const int64_t l0 = p->launches;
  for (int i = 0; i < reps; ++i) { rc = launch_evaluate(p, p->ev, p->dstate.ptrs(), true); if (rc) return rc; }
  if (launches) *launches = p->launches - l0;
  return CTBA_OK;
}

int ctba_bench_reproj(ctba_problem* p, int reps, int flush_l2, double* ms_avg) {
  CHECK_P(p);
  if (!ms_avg) return CTBA_E_ARG;
  if (p->device < 0) return fail(p, CTBA_E_CUDA, "planning-only handle (device -1): no compute entry points");
  CTBA_CUDA(p, cudaSetDevice(p->device));
  int rc = sync_to_device(p); if (rc) return rc;
  return bench_reproj(p, reps, flush_l2, ms_avg);
}

// ---- multi-GPU: time-segment sharding (ctba_shard.h) ----
int ctba_comm_unique_id(char* out128) {
  if (!out128) return CTBA_E_ARG;
  char err[256];
  return nccl_unique_id(out128, err, sizeof(err)) ? CTBA_E_NCCL : CTBA_OK;
}

int ctba_comm_init_nccl(ctba_problem* p, int world, int rank, const char* id128) {
  CHECK_P(p);
  if (world < 1 || rank < 0 || rank >= world || !id128) return fail(p, CTBA_E_ARG, "comm_init_nccl: bad arguments");
  if (p->comm || p->sharded) return fail(p, CTBA_E_STATE, "comm_init: the handle already has a communicator");
  char err[256] = "";
  Comm* c = make_nccl_comm(world, rank, id128, p->device, err, sizeof(err));
  if (!c) { p->err = err; return CTBA_E_NCCL; }
  p->comm = c; p->comm_owned = true; p->struct_dirty = true;
  return CTBA_OK;
}

int ctba_comm_init_local(ctba_problem** handles, int world) {
  if (!handles || world < 1) return CTBA_E_ARG;
  for (int r = 0; r < world; ++r) {
    if (!handles[r]) return CTBA_E_ARG;
    if (handles[r]->comm || handles[r]->sharded) return fail(handles[r], CTBA_E_STATE, "comm_init: the handle already has a communicator");
    if (handles[r]->device != handles[0]->device || handles[r]->device < 0) return fail(handles[r], CTBA_E_ARG, "comm_init_local: all handles must live on one device");
  }
  std::vector<Comm*> cs;
  if (!make_local_comms(world, &cs)) return fail(handles[0], CTBA_E_ARG, "comm_init_local: world size not supported");
  for (int r = 0; r < world; ++r) { handles[r]->comm = cs[r]; handles[r]->comm_owned = true; handles[r]->struct_dirty = true; }
  return CTBA_OK;
}

int ctba_shard_plan(ctba_problem* p, int world, int rank, int64_t* plan, int32_t* lm_owner, int32_t* imu_owner, int32_t* gp_owner, int32_t* bias_owner) {
  CHECK_P(p);
  if (world < 1 || rank < 0 || rank >= world || !plan) return fail(p, CTBA_E_ARG, "shard_plan: bad arguments");
  if (p->sharded) return fail(p, CTBA_E_STATE, "shard_plan: the handle already holds one shard only");
  ShardInfo sh; GlobalStructure gs; HostStructure hs;
  if (shard_plan(p, world, rank, &sh, &gs, &hs)) return fail(p, CTBA_E_ARG, "shard_plan: the reduced system is too short to cut into one time partition per rank");
  const int64_t v[16] = {sh.layout.P, sh.nbc, sh.nbw, sh.q_lo, sh.q_hi, sh.bc_lo, sh.bc_hi, sh.own_lo, sh.own_hi, sh.F_global, sh.f_first, sh.f_cost_begin,
                         hs.nodes, hs.hbn, hs.T, hs.KD};
  for (int i = 0; i < 16; ++i) plan[i] = v[i];
  for (int q = 0; q < sh.layout.P && q < 32; ++q) { plan[16 + 2 * q] = sh.layout.c0[q]; plan[17 + 2 * q] = sh.layout.nI[q]; }
  if (lm_owner) for (size_t l = 0; l < p->lm.size() / 3; ++l) lm_owner[l] = sh.lm_owner[l];
  if (imu_owner) for (size_t i = 0; i < p->i_s.size(); ++i) imu_owner[i] = sh.imu_owner[i];
  if (gp_owner) for (size_t i = 0; i < p->g_s.size(); ++i) gp_owner[i] = sh.gp_owner[i];
  if (bias_owner) for (size_t i = 0; i < p->b_s0.size(); ++i) bias_owner[i] = sh.bias_owner[i];
  return CTBA_OK;
}

}  // extern "C"
