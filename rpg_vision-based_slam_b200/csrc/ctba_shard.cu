// ctba_shard.cu — time-segment sharding of one problem over several GPUs: communicators, the host-side shard planner,
// and the two state/cost exchanges that are not part of a kernel pipeline.  Design: ctba_shard.h, DESIGN.md §7.
//
// Communicators:
//   NcclComm   one process per GPU (torchrun); libnccl.so.2 is resolved at run time with dlopen so that the library links and loads
//              on boxes without NCCL and shares the copy a host framework (PyTorch) may already have loaded.
//   LocalComm  all ranks are host threads of ONE process on ONE device.  Collectives are host barriers around a reduction kernel; no
//              kernel ever waits for another kernel, so this is safe on a single GPU.  It exists to test the sharded pipeline on a
//              one-GPU box against the unsharded solve, bit for bit the same kernels as the NCCL path.
#include <dlfcn.h>
#include <algorithm>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include "ctba_solver.h"

namespace ctba {

static inline int cdiv(size_t a, int b) { return (int)((a + b - 1) / b); }

// ------------------------------------------------------------------------------------------------------------------
// LocalComm
// ------------------------------------------------------------------------------------------------------------------
constexpr int kMaxLocalRanks = 16;
struct PtrList { double* p[kMaxLocalRanks]; };
// out_r[i] = ((b_0[i] + b_1[i]) + b_2[i]) + ...  written to every rank's buffer (fixed order: every rank sees the same bits)
__global__ void k_local_allreduce(const __grid_constant__ PtrList L, int world, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    double s = L.p[0][i];
    for (int r = 1; r < world; ++r) s += L.p[r][i];
    for (int r = 0; r < world; ++r) L.p[r][i] = s;
  }
}

struct LocalGroup {
  int world;
  std::mutex m;
  std::condition_variable cv;
  int count = 0;
  uint64_t gen = 0;
  PtrList ptrs{};
  explicit LocalGroup(int w) : world(w) {}
  void barrier() {
    std::unique_lock<std::mutex> lk(m);
    const uint64_t g = gen;
    if (++count == world) { count = 0; ++gen; cv.notify_all(); }
    else cv.wait(lk, [&] { return gen != g; });
  }
};

struct LocalComm : Comm {
  std::shared_ptr<LocalGroup> g;
  std::string err;
  const char* kind() const override { return "local"; }
  const char* error() const override { return err.c_str(); }
  int allreduce_sum(double* buf, size_t n, cudaStream_t st) override {
    if (cudaStreamSynchronize(st) != cudaSuccess) { err = "local allreduce: stream sync failed"; g->barrier(); g->barrier(); return -1; }
    g->ptrs.p[rank] = buf;
    g->barrier();
    int rc = 0;
    if (rank == 0 && n > 0) {
      k_local_allreduce<<<std::min(cdiv(n, 256), 1184), 256, 0, st>>>(g->ptrs, world, n);
      if (cudaStreamSynchronize(st) != cudaSuccess) { err = "local allreduce: kernel failed"; rc = -1; }
    }
    g->barrier();
    return rc;
  }
  int allgather(const double* send, double* recv, size_t n, cudaStream_t st) override {
    if (cudaStreamSynchronize(st) != cudaSuccess) { err = "local allgather: stream sync failed"; g->barrier(); g->barrier(); return -1; }
    g->ptrs.p[rank] = const_cast<double*>(send);
    g->barrier();
    int rc = 0;
    for (int r = 0; r < world; ++r)
      if (cudaMemcpyAsync(recv + (size_t)r * n, g->ptrs.p[r], n * sizeof(double), cudaMemcpyDeviceToDevice, st) != cudaSuccess) rc = -1;
    if (cudaStreamSynchronize(st) != cudaSuccess) rc = -1;
    if (rc) err = "local allgather: copy failed";
    g->barrier();
    return rc;
  }
};

Comm* make_local_comms(int world, std::vector<Comm*>* out) {
  if (world < 1 || world > kMaxLocalRanks) return nullptr;
  auto g = std::make_shared<LocalGroup>(world);
  out->clear();
  for (int r = 0; r < world; ++r) {
    LocalComm* c = new LocalComm();
    c->rank = r; c->world = world; c->g = g;
    out->push_back(c);
  }
  return (*out)[0];
}

// ------------------------------------------------------------------------------------------------------------------
// NcclComm (dlopen)
// ------------------------------------------------------------------------------------------------------------------
namespace {
// the subset of nccl.h this file needs (ABI-stable since NCCL 2.0)
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { kNcclSum = 0, kNcclFloat64 = 8 };
struct NcclApi {
  void* h = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
};
NcclApi& nccl_api() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) { api.h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL); if (api.h) break; }
    if (!api.h) return;
    api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(api.h, "ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))dlsym(api.h, "ncclCommInitRank");
    api.CommDestroy = (decltype(api.CommDestroy))dlsym(api.h, "ncclCommDestroy");
    api.AllReduce = (decltype(api.AllReduce))dlsym(api.h, "ncclAllReduce");
    api.AllGather = (decltype(api.AllGather))dlsym(api.h, "ncclAllGather");
    api.GetErrorString = (decltype(api.GetErrorString))dlsym(api.h, "ncclGetErrorString");
    api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllReduce && api.AllGather && api.GetErrorString;
  });
  return api;
}
}  // namespace

struct NcclComm : Comm {
  ncclComm_t comm = nullptr;
  std::string err;
  ~NcclComm() override { if (comm) nccl_api().CommDestroy(comm); }
  const char* kind() const override { return "nccl"; }
  const char* error() const override { return err.c_str(); }
  int allreduce_sum(double* buf, size_t n, cudaStream_t st) override {
    const ncclResult_t r = nccl_api().AllReduce(buf, buf, n, kNcclFloat64, kNcclSum, comm, st);
    if (r) { err = std::string("ncclAllReduce: ") + nccl_api().GetErrorString(r); return -1; }
    return 0;
  }
  int allgather(const double* send, double* recv, size_t n, cudaStream_t st) override {
    const ncclResult_t r = nccl_api().AllGather(send, recv, n, kNcclFloat64, comm, st);
    if (r) { err = std::string("ncclAllGather: ") + nccl_api().GetErrorString(r); return -1; }
    return 0;
  }
};

int nccl_unique_id(char* out128, char* err, size_t errlen) {
  NcclApi& a = nccl_api();
  if (!a.ok) { snprintf(err, errlen, "libnccl.so.2 not found or incomplete"); return -1; }
  ncclUniqueId id;
  const ncclResult_t r = a.GetUniqueId(&id);
  if (r) { snprintf(err, errlen, "ncclGetUniqueId: %s", a.GetErrorString(r)); return -1; }
  std::memcpy(out128, id.internal, 128);
  return 0;
}

Comm* make_nccl_comm(int world, int rank, const char* unique_id128, int device, char* err, size_t errlen) {
  NcclApi& a = nccl_api();
  if (!a.ok) { snprintf(err, errlen, "libnccl.so.2 not found or incomplete"); return nullptr; }
  if (cudaSetDevice(device) != cudaSuccess) { snprintf(err, errlen, "cudaSetDevice(%d) failed", device); return nullptr; }
  ncclUniqueId id;
  std::memcpy(id.internal, unique_id128, 128);
  NcclComm* c = new NcclComm();
  c->rank = rank; c->world = world;
  const ncclResult_t r = a.CommInitRank(&c->comm, world, id, rank);
  if (r) { snprintf(err, errlen, "ncclCommInitRank: %s", a.GetErrorString(r)); c->comm = nullptr; delete c; return nullptr; }
  return c;
}

// ------------------------------------------------------------------------------------------------------------------
// shard planner (host only)
// ------------------------------------------------------------------------------------------------------------------
// Decides, from the WHOLE problem held in p's host vectors, the partition layout and the owner rank of every unit; then replaces the
// host work items by this rank's units.  Deterministic: every rank computes the same plan from the same inputs.
int shard_plan(const ctba_problem* p, int world, int rank, ShardInfo* sh, GlobalStructure* gs, HostStructure* hs_out) {
  HostStructure hs;
  analyse_structure(p, nullptr, &hs);
  gs->has_imu = hs.has_imu; gs->any_frames = hs.F > 0; gs->any_imu = !p->i_s.empty(); gs->any_gp = !p->g_s.empty();
  gs->gravity_norm = p->gravity_norm; gs->hbn = hs.hbn; gs->tk = hs.tk; gs->tb = hs.tb;
  sh->world = world; sh->rank = rank;
  sh->nbc = hs.n_pad / kNB; sh->nbw = hs.nbw;
  if (band_partition_layout(sh->nbc, sh->nbw, p->band_partitions, world, p->band_nested, &sh->layout)) return -1;
  const int per = sh->layout.P / world;
  sh->q_lo = rank * per; sh->q_hi = (rank + 1) * per;
  sh->bc_lo = sh->layout.sep_start(sh->q_lo - 1);
  sh->bc_hi = sh->q_hi == sh->layout.P ? sh->nbc : sh->layout.c0[sh->q_hi];
  sh->own_lo = kNB * sh->layout.sep_start(sh->q_lo - 1);
  sh->own_hi = kNB * sh->layout.sep_start(sh->q_hi - 1);
  // landmark owners: rank of the block column of the first knot any of its frames touches
  const int L = hs.L;
  sh->lm_owner.assign(std::max(L, 1), 0);
  for (int l = 0; l < L; ++l)
    if (hs.lm_smin[l] >= 0) sh->lm_owner[l] = sh->rank_of_block(6 * hs.node_of_knot[hs.lm_smin[l]] / kNB);
  const std::vector<int32_t>&nk = hs.node_of_knot, &nb = hs.node_of_bias;
  sh->imu_owner.resize(p->i_s.size()); sh->gp_owner.resize(p->g_s.size()); sh->bias_owner.resize(p->b_s0.size());
  for (size_t i = 0; i < p->i_s.size(); ++i) sh->imu_owner[i] = sh->rank_of_block(6 * std::min(nk[p->i_s[i]], nb[p->i_sb[i]]) / kNB);
  for (size_t i = 0; i < p->g_s.size(); ++i) sh->gp_owner[i] = sh->rank_of_block(6 * nk[p->g_s[i]] / kNB);
  for (size_t i = 0; i < p->b_s0.size(); ++i) sh->bias_owner[i] = sh->rank_of_block(6 * nb[std::min(p->b_s0[i], p->b_s1[i])] / kNB);
  sh->F_global = hs.F;
  // frame ranges of every rank (contiguous, covering the observations of the rank's landmarks)
  std::vector<int> f_lo(world, 1 << 30), f_hi(world, -1);
  for (int o = 0; o < hs.M; ++o) {
    const int r = sh->lm_owner[p->o_lm[o]], f = p->o_frame[o];
    f_lo[r] = std::min(f_lo[r], f); f_hi[r] = std::max(f_hi[r], f);
  }
  sh->f_first = f_hi[rank] >= 0 ? f_lo[rank] : 0;
  int covered = -1;   // last global frame whose robust cost a lower rank already counts
  for (int r = 0; r < rank; ++r) covered = std::max(covered, f_hi[r]);
  sh->f_cost_begin = std::max(0, covered + 1 - sh->f_first);
  if (hs_out) *hs_out = std::move(hs);
  return 0;
}

template <class T, class A>
static void keep(std::vector<T, A>& v, const std::vector<char>& mask, int width = 1) {
  size_t w = 0;
  for (size_t i = 0; i < mask.size(); ++i)
    if (mask[i]) { for (int c = 0; c < width; ++c) v[w * width + c] = v[i * width + c]; ++w; }
  v.resize(w * width);
}

int shard_build(ctba_problem* p) {
  if (!p->comm || p->comm->world <= 1 || p->sharded) return CTBA_OK;
  const int world = p->comm->world, rank = p->comm->rank;
  if (!p->q_s.empty()) { p->err = "shard: pose priors (spline-fitting problems) are solved on one GPU"; return CTBA_E_ARG; }
  if (!p->shard) p->shard = new ShardInfo();
  if (!p->gstruct) p->gstruct = new GlobalStructure();
  ShardInfo& sh = *p->shard;
  HostStructure hs;
  if (shard_plan(p, world, rank, &sh, p->gstruct, &hs)) {
    p->err = "shard: the reduced system is too short (or its band too wide) to cut into one time partition per rank";
    return CTBA_E_ARG;
  }
  const int N = p->N;
  (void)N;
  // ---- frames / observations: the rank's landmarks with all their observations, frames = the contiguous range they span
  {
    const int F = hs.F, M = hs.M;
    int f_lo = 1 << 30, f_hi = -1;
    std::vector<char> ko(M, 0);
    for (int o = 0; o < M; ++o)
      if (sh.lm_owner[p->o_lm[o]] == rank) { ko[o] = 1; f_lo = std::min(f_lo, (int)p->o_frame[o]); f_hi = std::max(f_hi, (int)p->o_frame[o]); }
    HVec<int32_t> fs, fob{0}, olm, ofr;
    HVec<double> fu, fis, ouv;
    {
      size_t nk = 0;
      for (int o = 0; o < M; ++o) nk += ko[o];
      olm.reserve(nk); ofr.reserve(nk); ouv.reserve(2 * nk);   // page-locked storage: allocate once
    }
    for (int f = (f_hi >= 0 ? f_lo : 0); f <= f_hi && f < F; ++f) {
      fs.push_back(p->f_s[f]); fu.push_back(p->f_u[f]); fis.push_back(p->f_inv_std[f]);
      for (int o = p->f_obs_begin[f]; o < p->f_obs_begin[f + 1]; ++o) {
        if (!ko[o]) continue;
        olm.push_back(p->o_lm[o]); ofr.push_back(f - f_lo); ouv.push_back(p->o_uv[2 * (size_t)o]); ouv.push_back(p->o_uv[2 * (size_t)o + 1]);
      }
      fob.push_back((int32_t)olm.size());
    }
    p->f_s.swap(fs); p->f_u.swap(fu); p->f_inv_std.swap(fis); p->f_obs_begin.swap(fob); p->o_lm.swap(olm); p->o_frame.swap(ofr); p->o_uv.swap(ouv);
  }
  {  // IMU samples
    const size_t n = p->i_s.size();
    std::vector<char> k(n, 0);
    for (size_t i = 0; i < n; ++i) k[i] = sh.imu_owner[i] == rank;
    keep(p->i_s, k); keep(p->i_sb, k); keep(p->i_st_s, k); keep(p->i_k, k); keep(p->i_ub, k); keep(p->i_inv_a, k); keep(p->i_inv_g, k);
    keep(p->i_acc, k, 3); keep(p->i_gyr, k, 3);
  }
  {  // global-position samples
    const size_t n = p->g_s.size();
    std::vector<char> k(n, 0);
    for (size_t i = 0; i < n; ++i) k[i] = sh.gp_owner[i] == rank;
    keep(p->g_s, k); keep(p->g_st_s, k); keep(p->g_k, k); keep(p->g_inv_std, k); keep(p->g_meas, k, 3);
  }
  {  // bias random-walk blocks
    const size_t n = p->b_s0.size();
    std::vector<char> k(n, 0);
    for (size_t i = 0; i < n; ++i) k[i] = sh.bias_owner[i] == rank;
    keep(p->b_s0, k); keep(p->b_s1, k); keep(p->b_which, k); keep(p->b_u0, k); keep(p->b_u1, k); keep(p->b_inv_std, k);
  }
  if (rank != 0) p->gravity_norm = false;   // the gravity-norm residual touches only globals: rank 0 (gactive comes from the global structure)
  p->sharded = true;
  p->meas_dirty = true; p->struct_dirty = true;
  return CTBA_OK;
}

// ------------------------------------------------------------------------------------------------------------------
// exchanges outside the kernel pipelines
// ------------------------------------------------------------------------------------------------------------------
int shard_reduce_cost(ctba_problem* p, EvalBuffers& ev, double* host_cost8) {
  cudaStream_t st = p->stream;
  if (!p->sharded) {
    CTBA_CUDA(p, cudaMemcpyAsync(host_cost8, ev.cost.p, 8 * sizeof(double), cudaMemcpyDeviceToHost, st));
    CTBA_CUDA(p, cudaStreamSynchronize(st));
    return CTBA_OK;
  }
  const int W = p->comm->world;
  CTBA_CUDA(p, p->d_gather.resize((size_t)W * 8));
  if (p->comm->allgather(ev.cost.p, p->d_gather.p, 8, st)) { p->err = p->comm->error(); return CTBA_E_NCCL; }
  std::vector<double> h((size_t)W * 8);
  CTBA_CUDA(p, cudaMemcpyAsync(h.data(), p->d_gather.p, h.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
  CTBA_CUDA(p, cudaStreamSynchronize(st));
  for (int c = 0; c < 8; ++c) { double s = 0.0; for (int r = 0; r < W; ++r) s += h[(size_t)r * 8 + c]; host_cost8[c] = s; }
  return CTBA_OK;
}

// stage = owned entries of the state, zero elsewhere; all-reduced; written back: every rank ends with the whole optimised state
__global__ void k_stage_state(int K, int Kb, int L, const int32_t* __restrict__ node_of_knot, const int32_t* __restrict__ node_of_bias,
                              const int32_t* __restrict__ lm_owner, int rank, int own_lo, int own_hi, StatePtrs S, double* __restrict__ stage) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  double* sq = stage; double* sp = sq + 4 * (size_t)K; double* sa = sp + 3 * (size_t)K; double* sg = sa + 3 * (size_t)Kb; double* sl = sg + 3 * (size_t)Kb;
  if (i < K) {
    const int nd = 6 * node_of_knot[i];
    const bool own = nd >= own_lo && nd < own_hi;
    for (int c = 0; c < 4; ++c) sq[4 * (size_t)i + c] = own ? S.knot_q[4 * (size_t)i + c] : 0.0;
    for (int c = 0; c < 3; ++c) sp[3 * (size_t)i + c] = own ? S.knot_p[3 * (size_t)i + c] : 0.0;
  } else if (i < K + Kb) {
    const int j = i - K, nd = 6 * node_of_bias[j];
    const bool own = nd >= own_lo && nd < own_hi;
    for (int c = 0; c < 3; ++c) { sa[3 * (size_t)j + c] = own ? S.bias_a[3 * (size_t)j + c] : 0.0; sg[3 * (size_t)j + c] = own ? S.bias_g[3 * (size_t)j + c] : 0.0; }
  } else if (i < K + Kb + L) {
    const int l = i - K - Kb;
    const bool own = lm_owner[l] == rank;
    for (int c = 0; c < 3; ++c) sl[3 * (size_t)l + c] = own ? S.lm[3 * (size_t)l + c] : 0.0;
  }
}
__global__ void k_unstage_state(int K, int Kb, int L, const double* __restrict__ stage, StatePtrs S) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t nq = 4 * (size_t)K, np = 3 * (size_t)K, nb = 3 * (size_t)Kb, nl = 3 * (size_t)L;
  if (i < nq) S.knot_q[i] = stage[i];
  else if (i < nq + np) S.knot_p[i - nq] = stage[i];
  else if (i < nq + np + nb) S.bias_a[i - nq - np] = stage[i];
  else if (i < nq + np + 2 * nb) S.bias_g[i - nq - np - nb] = stage[i];
  else if (i < nq + np + 2 * nb + nl) S.lm[i - nq - np - 2 * nb] = stage[i];
}

int shard_gather_state(ctba_problem* p) {
  if (!p->sharded) return CTBA_OK;
  SolverData* sd = p->solver;
  cudaStream_t st = p->stream;
  const int K = sd->K, Kb = sd->Kb, L = sd->L;
  const size_t n = 7 * (size_t)K + 6 * (size_t)Kb + 3 * (size_t)L;
  CTBA_CUDA(p, p->d_stage.resize(std::max<size_t>(n, 1)));
  if (!sd->d_lm_owner.p || sd->d_lm_owner.n < (size_t)std::max(L, 1)) CTBA_CUDA(p, sd->d_lm_owner.upload(p->shard->lm_owner, st));
  const StatePtrs S = p->dstate.ptrs();
  k_stage_state<<<cdiv(K + Kb + L, 256), 256, 0, st>>>(K, Kb, L, sd->d_node_of_knot.p, sd->d_node_of_bias.p, sd->d_lm_owner.p, p->shard->rank, sd->own_lo,
                                                       sd->own_hi, S, p->d_stage.p);
  if (p->comm->allreduce_sum(p->d_stage.p, n, st)) { p->err = p->comm->error(); return CTBA_E_NCCL; }
  k_unstage_state<<<cdiv(n, 256), 256, 0, st>>>(K, Kb, L, p->d_stage.p, S);
  p->launches += 2;
  p->host_state_stale = true;
  CTBA_CUDA(p, cudaGetLastError());
  return CTBA_OK;
}

}  // namespace ctba
