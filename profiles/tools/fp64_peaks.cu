// fp64_peaks.cu — FP64 roofline denominators of this pool's B200 (MEASURED_PEAKS.json holds HBM and bf16 only) and the
// operand-feed experiments behind the band kernels' design:
//   dfma      : independent DFMA chains in registers                         -> FP64 vector peak
//   dmma_reg  : mma.sync.m8n8k4.f64 (SASS DMMA) chains in registers          -> FP64 tensor-pipe peak
//   dmma_lat  : ONE dependent DMMA chain per warp, one warp per SM           -> dependent-issue latency of DMMA
//   dmma_l2   : 32x32x32 block products accumulated over a K loop, fragments loaded straight from L2 (two 256-bit loads per fragment)
//   dmma_tma  : the same products, operand blocks staged by cp.async.bulk + mbarrier (1-D TMA) into shared memory, per-warp double
//               buffer; variant 0 = whole 8 KB blocks (2-way bank conflicts on the fragment reads), variant 1 = one 256 B copy per row
//               into rows padded to 320 B (conflict-free reads, 64 small copies per stage)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o fp64_peaks fp64_peaks.cu ; run: ./fp64_peaks out.json
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

__global__ void __launch_bounds__(256) k_dfma(double* out, int iters) {
  double a[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 1e-3 + i;
  const double x = 1.0000001, y = 1e-9;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = fma(a[i], x, y);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int CHAINS>
__global__ void __launch_bounds__(256) k_dmma_reg(double* out, int iters) {
  double c[CHAINS][2];
#pragma unroll
  for (int i = 0; i < CHAINS; ++i) c[i][0] = c[i][1] = 0.0;
  double a = 1.0 + threadIdx.x * 1e-6, b = 1.0 - threadIdx.x * 1e-6;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) dmma884(c[i][0], c[i][1], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < CHAINS; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// ---- block products from L2: CTA of 8 warps, warp (wr, wc) owns 16 rows x 32 columns of a 64 x 64 tile; K loop over nk blocks ----
__device__ __forceinline__ void load_frag8(const double* __restrict__ p, double v[8]) {
  asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3]) : "l"(p));
  asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v[4]), "=d"(v[5]), "=d"(v[6]), "=d"(v[7]) : "l"(p + 4));
}
__global__ void __launch_bounds__(256, 2) k_dmma_l2(const double* __restrict__ blocks, size_t nblocks, int nk, double* out) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int wr = warp >> 1, wc = warp & 1;
  double acc[2][4][2];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int h = 0; h < 4; ++h) acc[i][h][0] = acc[i][h][1] = 0.0;
  size_t base = ((size_t)blockIdx.x * 97) % nblocks;
  for (int k = 0; k < nk; ++k) {
    const double* A = blocks + (((base + 4 * (size_t)k + (wr >> 1)) % nblocks) << 10) + (size_t)(16 * (wr & 1) + g) * 32 + 8 * t;
    const double* B = blocks + (((base + 4 * (size_t)k + 2 + wc) % nblocks) << 10) + (size_t)g * 32 + 8 * t;
    double a0[8], a1[8], b0[8], b1[8], b2[8], b3[8];
    load_frag8(A, a0); load_frag8(A + 8 * 32, a1);
    load_frag8(B, b0); load_frag8(B + 8 * 32, b1); load_frag8(B + 16 * 32, b2); load_frag8(B + 24 * 32, b3);
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
      dmma884(acc[0][0][0], acc[0][0][1], a0[ks], b0[ks]); dmma884(acc[1][0][0], acc[1][0][1], a1[ks], b0[ks]);
      dmma884(acc[0][1][0], acc[0][1][1], a0[ks], b1[ks]); dmma884(acc[1][1][0], acc[1][1][1], a1[ks], b1[ks]);
      dmma884(acc[0][2][0], acc[0][2][1], a0[ks], b2[ks]); dmma884(acc[1][2][0], acc[1][2][1], a1[ks], b2[ks]);
      dmma884(acc[0][3][0], acc[0][3][1], a0[ks], b3[ks]); dmma884(acc[1][3][0], acc[1][3][1], a1[ks], b3[ks]);
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int h = 0; h < 4; ++h) s += acc[i][h][0] + acc[i][h][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// ---- block products from shared memory staged by cp.async.bulk ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count)); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes),
               "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n .reg .pred p;\n WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void lds128(const double* p, double& x, double& y) {
  asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(x), "=d"(y) : "r"(smem_u32(p)));
}
// One warp = one 32x32x32 product per k, its own 2-stage ring of (A, B).  Fragment element e = 2 m + h of lane (g, t) is column 2 t + 8 m + h.
template <int VAR, int WARPS>
__global__ void __launch_bounds__(32 * WARPS, 1) k_dmma_tma(const double* __restrict__ blocks, size_t nblocks, int nk, double* out) {
  constexpr int LD = VAR ? 40 : 32;
  constexpr int BLK = 32 * LD;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  double* ring = reinterpret_cast<double*>(smem_raw) + (size_t)warp * 4 * BLK;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + (size_t)WARPS * 4 * BLK * sizeof(double)) + 2 * warp;
  if (lane == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); }
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncwarp();
  const size_t base = (((size_t)blockIdx.x * WARPS + warp) * 131) % nblocks;
  auto issue = [&](int k, int s) {
    const double* A = blocks + (((base + 2 * (size_t)k) % nblocks) << 10);
    const double* B = blocks + (((base + 2 * (size_t)k + 1) % nblocks) << 10);
    double* dA = ring + (size_t)(2 * s) * BLK; double* dB = dA + BLK;
    if (lane == 0) mbar_expect_tx(&bars[s], 2 * 8192);
    __syncwarp();
    if (VAR == 0) {
      if (lane == 0) { bulk_g2s(dA, A, 8192, &bars[s]); bulk_g2s(dB, B, 8192, &bars[s]); }
    } else {
      bulk_g2s(dA + lane * LD, A + lane * 32, 256, &bars[s]);
      bulk_g2s(dB + lane * LD, B + lane * 32, 256, &bars[s]);
    }
  };
  double acc[4][4][2];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int h = 0; h < 4; ++h) acc[i][h][0] = acc[i][h][1] = 0.0;
  issue(0, 0);
  if (nk > 1) issue(1, 1);
  for (int k = 0; k < nk; ++k) {
    const int s = k & 1;
    mbar_wait(&bars[s], (k >> 1) & 1);
    const double* sA = ring + (size_t)(2 * s) * BLK + g * LD + 2 * t;
    const double* sB = sA + BLK;
#pragma unroll
    for (int m = 0; m < 4; ++m) {
      double a[4][2], b[4][2];
#pragma unroll
      for (int i = 0; i < 4; ++i) { lds128(sA + 8 * i * LD + 8 * m, a[i][0], a[i][1]); lds128(sB + 8 * i * LD + 8 * m, b[i][0], b[i][1]); }
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i][h], b[j][h]);
    }
    __syncwarp();
    if (k + 2 < nk) issue(k + 2, s);
  }
  double sum = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int h = 0; h < 4; ++h) sum += acc[i][h][0] + acc[i][h][1];
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = sum;
}

template <class F>
static double time_ms(F&& f, int reps) {
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  f(); CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < reps; ++r) {
    CK(cudaEventRecord(a)); f(); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    float ms; CK(cudaEventElapsedTime(&ms, a, b)); if (ms < best) best = ms;
  }
  CK(cudaGetLastError());
  return best;
}

int main(int argc, char** argv) {
  cudaDeviceProp pr; CK(cudaGetDeviceProperties(&pr, 0));
  const int sms = pr.multiProcessorCount;
  double* out; CK(cudaMalloc(&out, (size_t)sms * 64 * 256 * sizeof(double)));
  const size_t nblocks = 6144;   // 48 MB of 8 KB blocks: resident in the 126 MB L2 after the first pass
  double* blocks; CK(cudaMalloc(&blocks, nblocks * 8192));
  { std::vector<double> h(nblocks * 1024); for (size_t i = 0; i < h.size(); ++i) h[i] = 1e-3 * (double)((i * 2654435761u) % 1000); CK(cudaMemcpy(blocks, h.data(), h.size() * 8, cudaMemcpyHostToDevice)); }
  FILE* f = fopen(argc > 1 ? argv[1] : "fp64_peaks.json", "w");
  fprintf(f, "{\n \"gpu\": \"%s\", \"sms\": %d, \"clock_khz_max\": %d,\n", pr.name, sms, pr.clockRate);
  {
    const int iters = 1 << 15, ctas = sms * 8;
    double ms = time_ms([&] { k_dfma<<<ctas, 256>>>(out, iters); }, 5);
    double tf = 2.0 * 16 * iters * 256.0 * ctas / (ms * 1e-3) / 1e12;
    printf("dfma      : %.3f ms  %.2f TFLOP/s\n", ms, tf); fprintf(f, " \"dfma_tflops\": %.3f,\n", tf);
  }
  {
    const int iters = 1 << 13, ctas = sms * 8;
    double ms = time_ms([&] { k_dmma_reg<16><<<ctas, 256>>>(out, iters); }, 5);
    double tf = 2.0 * 8 * 8 * 4 * 16.0 * iters * 8 * ctas / (ms * 1e-3) / 1e12;
    printf("dmma_reg  : %.3f ms  %.2f TFLOP/s\n", ms, tf); fprintf(f, " \"dmma_tflops\": %.3f,\n", tf);
  }
  {   // one warp per SM, one dependent chain: time per DMMA = its dependent-issue latency
    const int iters = 1 << 14;
    double ms = time_ms([&] { k_dmma_reg<1><<<sms, 32>>>(out, iters); }, 5);
    double cyc = ms * 1e-3 * pr.clockRate * 1e3 / iters;
    printf("dmma_lat  : %.3f ms  %.1f cycles per dependent DMMA (at the max clock)\n", ms, cyc); fprintf(f, " \"dmma_dependent_latency_cycles\": %.1f,\n", cyc);
    for (int ch : {2, 4, 8}) {
      double m2 = ch == 2 ? time_ms([&] { k_dmma_reg<2><<<sms, 32>>>(out, iters); }, 5) : ch == 4 ? time_ms([&] { k_dmma_reg<4><<<sms, 32>>>(out, iters); }, 5)
                                                                                                   : time_ms([&] { k_dmma_reg<8><<<sms, 32>>>(out, iters); }, 5);
      double c2 = m2 * 1e-3 * pr.clockRate * 1e3 / (iters * (double)ch);
      printf("dmma 1 warp/SM, %d chains: %.1f cycles per DMMA issued\n", ch, c2); fprintf(f, " \"dmma_cycles_per_issue_1warp_%dchains\": %.1f,\n", ch, c2);
    }
  }
  for (int nk : {4, 16, 64}) {
    const int ctas = sms * 2 * 8;
    double ms = time_ms([&] { k_dmma_l2<<<ctas, 256>>>(blocks, nblocks, nk, out); }, 5);
    double tf = 2.0 * 64 * 64 * 32 * (double)nk * ctas / (ms * 1e-3) / 1e12;
    printf("dmma_l2   nk=%2d: %.3f ms  %.2f TFLOP/s  (64x64 CTA tile, 32 KB of operands per k)\n", nk, ms, tf); fprintf(f, " \"dmma_l2_nk%d_tflops\": %.3f,\n", nk, tf);
  }
  auto run_tma = [&](auto kern, int var, int warps, const char* name) {
    const int ld = var ? 40 : 32;
    const size_t smem = (size_t)warps * 4 * 32 * ld * 8 + warps * 16;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    for (int nk : {4, 16, 64}) {
      const int ctas = sms * 4;
      double ms = time_ms([&] { kern<<<ctas, 32 * warps, smem>>>(blocks, nblocks, nk, out); }, 5);
      double tf = 2.0 * 32 * 32 * 32 * (double)nk * warps * ctas / (ms * 1e-3) / 1e12;
      printf("%s nk=%2d: %.3f ms  %.2f TFLOP/s  (%d warps x 32x32 tile, 16 KB of operands per warp per k, smem %zu B)\n", name, nk, ms, tf, warps, smem);
      fprintf(f, " \"%s_nk%d_tflops\": %.3f,\n", name, nk, tf);
    }
  };
  run_tma(k_dmma_tma<0, 4>, 0, 4, "dmma_tma_block_w4");
  run_tma(k_dmma_tma<0, 6>, 0, 6, "dmma_tma_block_w6");
  run_tma(k_dmma_tma<1, 4>, 1, 4, "dmma_tma_rows_w4");
  fprintf(f, " \"how\": \"profiles/tools/fp64_peaks.cu, best of 5, CUDA events\"\n}\n");
  fclose(f);
  return 0;
}
