// vx_peaks.cu — FP64 pipe peaks measured on the device the library runs on.
// MEASURED_PEAKS.json carries no FP64 figure and the BA Hessian build (B1) is FP64-bound for k >= 3
// occupied frames per voxel, so bench.py measures the denominators itself:
//   mode 0: dependent-free DFMA chains on the CUDA cores (8 independent accumulators per thread)
//   mode 1: mma.sync.m8n8k4.f64 (DMMA) chains on the tensor pipe (4 independent accumulators per warp)
#include "vx_common.cuh"

namespace {
__global__ void __launch_bounds__(256) k_peak_dfma(double *out, int iters, double a, double b) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
#pragma unroll 1
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int u = 0; u < 8; u++) {
      x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
      x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
    }
  }
  out[blockIdx.x * (size_t)blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}

__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__global__ void __launch_bounds__(256) k_peak_dmma(double *out, int iters, double a, double b) {
  double c[8][2];
#pragma unroll
  for (int k = 0; k < 8; k++) { c[k][0] = threadIdx.x + k; c[k][1] = k; }
#pragma unroll 1
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int u = 0; u < 4; u++) {
#pragma unroll
      for (int k = 0; k < 8; k++) dmma(c[k][0], c[k][1], a, b);
    }
  }
  double s = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) s += c[k][0] + c[k][1];
  out[blockIdx.x * (size_t)blockDim.x + threadIdx.x] = s;
}
// mode 2: even warps run the DFMA body, odd warps the DMMA body, same iteration count: if the two instruction classes had
// separate pipes the combined rate would approach the sum of modes 0 and 1; if they share the FP64 datapath it stays near one.
__global__ void __launch_bounds__(256) k_peak_mixed(double *out, int iters, double a, double b) {
  double r = 0.0;
  if (((threadIdx.x >> 5) & 1) == 0) {
    double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
#pragma unroll 1
    for (int i = 0; i < iters; i++) {
#pragma unroll
      for (int u = 0; u < 8; u++) {
        x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
        x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
      }
    }
    r = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
  } else {
    double c[8][2];
#pragma unroll
    for (int k = 0; k < 8; k++) { c[k][0] = threadIdx.x + k; c[k][1] = k; }
#pragma unroll 1
    for (int i = 0; i < iters; i++) {
#pragma unroll
      for (int u = 0; u < 4; u++) {
#pragma unroll
        for (int k = 0; k < 8; k++) dmma(c[k][0], c[k][1], a, b);
      }
    }
#pragma unroll
    for (int k = 0; k < 8; k++) r += c[k][0] + c[k][1];
  }
  out[blockIdx.x * (size_t)blockDim.x + threadIdx.x] = r;
}
// modes 16 + ilp (ilp = 1, 2, 4, 8): ONE warp, `ilp` independent dependent-DFMA chains; returns cycles per DFMA issued
// (ilp = 1: the dependent-issue latency of the FP64 pipe).  modes 32 + ilp: the same for DMMA chains.
template <int ILP, bool MMA>
__global__ void k_latency(double *out, long long *cyc, int iters, double a, double b) {
  double x[8][2];
#pragma unroll
  for (int k = 0; k < 8; k++) { x[k][0] = threadIdx.x + k; x[k][1] = k; }
  const long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int u = 0; u < 8; u++) {
#pragma unroll
      for (int k = 0; k < ILP; k++) {
        if (MMA) dmma(x[k][0], x[k][1], a, b);
        else x[k][0] = fma(x[k][0], a, b);
      }
    }
  }
  const long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) s += x[k][0] + x[k][1];
  out[threadIdx.x] = s;
  if (threadIdx.x == 0) *cyc = t1 - t0;
}
}  // namespace

extern "C" int vx_bench_fp64_latency(vx_ctx *ctx, int mma, int ilp, double *cycles_per_instr) {
  VX_TRY(vx::bind(ctx));
  VX_ARG(cycles_per_instr != nullptr && (ilp == 1 || ilp == 2 || ilp == 4 || ilp == 8));
  const int iters = 2048;
  VX_TRY(ctx->scratch.reserve(sizeof(double) * 64));
  double *out = ctx->scratch.as<double>();
  long long *cyc = reinterpret_cast<long long *>(out + 32);
  for (int rep = 0; rep < 2; rep++) {
#define VX_LAT(I)                                                                                         \
    if (ilp == I) {                                                                                       \
      if (mma) k_latency<I, true><<<1, 32, 0, ctx->stream>>>(out, cyc, iters, 0.999999, 1e-9);            \
      else k_latency<I, false><<<1, 32, 0, ctx->stream>>>(out, cyc, iters, 0.999999, 1e-9);               \
    }
    VX_LAT(1) VX_LAT(2) VX_LAT(4) VX_LAT(8)
#undef VX_LAT
    VX_LAUNCH_CHECK(ctx);
    VX_CUDA(cudaStreamSynchronize(ctx->stream));
  }
  long long h = 0;
  VX_CUDA(cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost));
  *cycles_per_instr = (double)h / ((double)iters * 8 * ilp);
  return VX_OK;
}

extern "C" int vx_bench_fp64(vx_ctx *ctx, int mode, double *tflops) {
  VX_TRY(vx::bind(ctx));
  VX_ARG(tflops != nullptr && mode >= 0 && mode <= 2);
  const int blocks = ctx->sm_count * 8, threads = 256, iters = 4096;
  VX_TRY(ctx->scratch.reserve(sizeof(double) * (size_t)blocks * threads));
  cudaEvent_t e0, e1;
  VX_CUDA(cudaEventCreate(&e0));
  VX_CUDA(cudaEventCreate(&e1));
  double best = 0.0;
  for (int rep = 0; rep < 4; rep++) {
    VX_CUDA(cudaEventRecord(e0, ctx->stream));
    if (mode == 0) k_peak_dfma<<<blocks, threads, 0, ctx->stream>>>(ctx->scratch.as<double>(), iters, 0.999999, 1e-9);
    else if (mode == 1) k_peak_dmma<<<blocks, threads, 0, ctx->stream>>>(ctx->scratch.as<double>(), iters, 0.999999, 1e-9);
    else k_peak_mixed<<<blocks, threads, 0, ctx->stream>>>(ctx->scratch.as<double>(), iters, 0.999999, 1e-9);
    VX_LAUNCH_CHECK(ctx);
    VX_CUDA(cudaEventRecord(e1, ctx->stream));
    VX_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    VX_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    // mode 0: 64 FMA per thread-iteration = 128 flop; mode 1: 32 DMMA per warp-iteration x 512 flop
    // mode 2: half the warps of each kind
    const double f0 = (double)blocks * threads * iters * 128.0, f1 = (double)blocks * (threads / 32) * iters * 32.0 * 512.0;
    double flop = mode == 0 ? f0 : (mode == 1 ? f1 : 0.5 * (f0 + f1));
    double tf = flop / (ms * 1e-3) / 1e12;
    if (rep > 0 && tf > best) best = tf;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  *tflops = best;
  return VX_OK;
}
