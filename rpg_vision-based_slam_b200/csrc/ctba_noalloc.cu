// ctba_noalloc.cu — ctba_align_system_noalloc: the one entry point of the library that allocates nothing.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc (code under CUDA
// graph capture, tight serving loops, harnesses that forbid it).  It is a variant of k_align_accumulate + k_align_reduce (ctba_align.cu):
// the normal equations of the world-alignment problem (GlobalPositionSensorWithAlignmentCorrectionCostFunctor under CauchyLoss, every
// knot constant) at a caller-given alignment state, on caller-owned DEVICE buffers.  The per-CTA partial sums — the scratch that
// ctba_align.cu keeps in a cudaMalloc'ed buffer of the handle — live in a file-scope static __device__ array of fixed size:
//   kNoAllocMaxCtas x AS_SIZE doubles = 296 x 78 x 8 B = 184 704 B, held by every GPU that loads the library until the process exits.
// In host code the array's NAME is a host-side stand-in, not a device address; the address is taken once per device, with that device
// current, by cudaGetSymbolAddress and passed to the kernels.
//
// Every call shares the array.  Nothing in it has to start at zero: each CTA of the accumulate kernel overwrites its whole AS_SIZE row
// and the reduce kernel reads exactly the rows of the CTAs of the same call, so the contents left by the previous call are never read.
// Calls that overlap on the GPU (two streams, two host threads, two processes do NOT share: the array is per process and per device)
// race on the array: callers serialise them on one stream or with events.
//
// The timestamp -> spline window arithmetic and the skip rules of addWorldColmapAlignmentMeasurement (bspline_so3r3.h:1121-1211), which
// ctba_add_alignment_measurements does on the host, are done per measurement in the kernel with the same int64 arithmetic.
#include <cstdio>
#include <mutex>
#include "ctba_internal.h"
#include "ctba_align.h"

namespace ctba {
namespace {

constexpr int kNaThreads = 128, kNaWarps = kNaThreads / 32;
constexpr int kNoAllocMaxCtas = 296;                        // 2 x 148 SMs
constexpr int kNoAllocMaxMeasurements = 1 << 20;            // stated limit of the entry point (CTBA_NOALLOC_MAX_MEASUREMENTS)
constexpr int kMaxDevices = 64;

static __device__ __align__(256) double g_noalloc_scratch[kNoAllocMaxCtas * AS_SIZE];

struct NaState { double T_wg[7]; double x[AL_STATE]; };

template <int N>
__global__ void __launch_bounds__(kNaThreads) k_noalloc_align_accumulate(const __grid_constant__ Globals G, const __grid_constant__ NaState S,
                                                                        const double* __restrict__ knot_q, const double* __restrict__ knot_p,
                                                                        int n_knots, int n, const int64_t* __restrict__ t_ns,
                                                                        const double* __restrict__ pos, int64_t t0_ns, int64_t dt_ns, double inv_std,
                                                                        double cauchy_b, double* __restrict__ partial) {
  __shared__ double sh[kNaWarps][AS_SIZE];
  double acc[AS_SIZE];
#pragma unroll
  for (int v = 0; v < AS_SIZE; ++v) acc[v] = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int64_t st_ns = t_ns[i] - t0_ns;                  // t - start_t_ns - int64(t_s_gp_s * 1e9)        bspline_so3r3.h:1127-1131
    if (st_ns < 0) continue;
    const int64_t s = st_ns / dt_ns;
    if (s + N > (int64_t)n_knots) continue;
    const double st_s = st_ns * 1e-9, k = (double)(int)s;
    double kq[4 * N], kp[3 * N];
#pragma unroll
    for (int e = 0; e < 4 * N; ++e) kq[e] = knot_q[4 * (size_t)s + e];
#pragma unroll
    for (int e = 0; e < 3 * N; ++e) kp[e] = knot_p[3 * (size_t)s + e];
    const double m[3] = {pos[3 * (size_t)i], pos[3 * (size_t)i + 1], pos[3 * (size_t)i + 2]};
    align_sample_accumulate<N>(G, S.T_wg, S.x, kq, kp, st_s, k, m, inv_std, cauchy_b, true, acc);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int v = 0; v < AS_SIZE; ++v) {
    double val = acc[v];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) val += __shfl_down_sync(0xffffffffu, val, o);
    if (lane == 0) sh[warp][v] = val;
  }
  __syncthreads();
  if (threadIdx.x < AS_SIZE) {   // the CTA's whole row is overwritten: nothing of an earlier call survives in the rows this call reads
    double sum = 0.0;
#pragma unroll
    for (int w = 0; w < kNaWarps; ++w) sum += sh[w][threadIdx.x];
    partial[(size_t)blockIdx.x * AS_SIZE + threadIdx.x] = sum;
  }
}

__global__ void __launch_bounds__(kNaThreads) k_noalloc_align_reduce(int nblocks, const double* __restrict__ partial, double* __restrict__ sys) {
  if (threadIdx.x < AS_SIZE) {
    double sum = 0.0;
    for (int b = 0; b < nblocks; ++b) sum += partial[(size_t)b * AS_SIZE + threadIdx.x];   // CTA order: bit-reproducible
    sys[threadIdx.x] = sum;
  }
}

std::mutex g_na_mutex;
double* g_na_ptr[kMaxDevices] = {};   // device address of g_noalloc_scratch, per device ordinal

int na_fail(char* err, int err_len, int code, const char* msg) {
  if (err && err_len > 0) snprintf(err, (size_t)err_len, "%s", msg);
  return code;
}

}  // namespace
}  // namespace ctba

using namespace ctba;

extern "C" int ctba_align_system_noalloc(int order, int64_t dt_ns, int64_t start_t_ns, int device, void* cuda_stream, double scale,
                                         double t_s_gp_s, const double* T_wg7, const double* x12, double cauchy_a, int n_knots,
                                         const double* d_knot_q, const double* d_knot_p, int n, const int64_t* d_t_ns, const double* d_pos,
                                         double noise, double* d_sys78, char* err, int err_len) {
  if (err && err_len > 0) err[0] = 0;
  if (order != 4 && order != 6) return na_fail(err, err_len, CTBA_E_ARG, "ctba_align_system_noalloc: order must be 4 or 6");
  if (dt_ns <= 0 || !T_wg7 || !x12 || !d_knot_q || !d_knot_p || !d_sys78 || n_knots < order || n < 0 || (n > 0 && (!d_t_ns || !d_pos)) ||
      !(noise > 0) || !(cauchy_a >= 0))
    return na_fail(err, err_len, CTBA_E_ARG, "ctba_align_system_noalloc: bad arguments");
  if (n > kNoAllocMaxMeasurements) {
    char msg[200];
    snprintf(msg, sizeof msg, "ctba_align_system_noalloc: %d measurements exceed the fixed limit of %d (static device scratch of %d x %d doubles)",
             n, kNoAllocMaxMeasurements, kNoAllocMaxCtas, (int)AS_SIZE);
    return na_fail(err, err_len, CTBA_E_ARG, msg);
  }
  if (device < 0 || device >= kMaxDevices) return na_fail(err, err_len, CTBA_E_CUDA, "ctba_align_system_noalloc: no such device");
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return na_fail(err, err_len, CTBA_E_CUDA, cudaGetErrorString(e));
  double* scratch = nullptr;
  {   // set-up, once per device, with that device current: the symbol's device address (its name is not one in host code)
    std::lock_guard<std::mutex> lock(g_na_mutex);
    if (!g_na_ptr[device]) {
      double* ptr = nullptr;
      e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
      if (e != cudaSuccess || !ptr) return na_fail(err, err_len, CTBA_E_CUDA, e != cudaSuccess ? cudaGetErrorString(e) : "cudaGetSymbolAddress: null");
      g_na_ptr[device] = ptr;
    }
    scratch = g_na_ptr[device];
  }
  Globals G;   // the fields align_eval reads (ctba_terms.h): tables, knot spacing, scale; the rest zeroed
  memset(&G, 0, sizeof G);
  fill_tables(order, &G.tb);
  fill_tables(4, &G.tb4);
  G.dt_s = dt_ns * 1e-9; G.inv_dt = 1e9 / dt_ns; G.scale = scale; G.gate_sq = 100.0;
  NaState S;
  for (int i = 0; i < 7; ++i) S.T_wg[i] = T_wg7[i];
  for (int i = 0; i < AL_STATE; ++i) S.x[i] = x12[i];
  const int64_t t0_ns = start_t_ns + static_cast<int64_t>(t_s_gp_s * 1e9);   // :1127
  const int grid = std::max(1, std::min((n + kNaThreads - 1) / kNaThreads, kNoAllocMaxCtas));
  cudaStream_t st = (cudaStream_t)cuda_stream;
  const double inv_std = 1.0 / noise, cauchy_b = cauchy_a * cauchy_a;
  if (order == 4)
    k_noalloc_align_accumulate<4><<<grid, kNaThreads, 0, st>>>(G, S, d_knot_q, d_knot_p, n_knots, n, d_t_ns, d_pos, t0_ns, dt_ns, inv_std, cauchy_b,
                                                               scratch);
  else
    k_noalloc_align_accumulate<6><<<grid, kNaThreads, 0, st>>>(G, S, d_knot_q, d_knot_p, n_knots, n, d_t_ns, d_pos, t0_ns, dt_ns, inv_std, cauchy_b,
                                                               scratch);
  k_noalloc_align_reduce<<<1, kNaThreads, 0, st>>>(grid, scratch, d_sys78);
  e = cudaGetLastError();
  if (e != cudaSuccess) return na_fail(err, err_len, CTBA_E_CUDA, cudaGetErrorString(e));
  return CTBA_OK;
}
