// vx_common.cuh — shared internals of libvoxelba (context, error plumbing, device buffers).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdarg>
#include <cstring>
#include <cstdlib>
#include <ctime>
#include <string>
#include <vector>
#include "../../include/voxelba.h"

namespace vx {

void set_error(const char *fmt, ...);

#define VX_CUDA(call)                                                                        \
  do {                                                                                       \
    cudaError_t _e = (call);                                                                 \
    if (_e != cudaSuccess) {                                                                 \
      vx::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e));    \
      return VX_ERR_CUDA;                                                                    \
    }                                                                                        \
  } while (0)

#define VX_TRY(call)                 \
  do {                               \
    int _s = (call);                 \
    if (_s != VX_OK) return _s;      \
  } while (0)

#define VX_ARG(cond)                                                       \
  do {                                                                     \
    if (!(cond)) {                                                         \
      vx::set_error("%s:%d bad argument: %s", __FILE__, __LINE__, #cond);  \
      return VX_ERR_ARG;                                                   \
    }                                                                      \
  } while (0)

// Device allocations go through a per-process cache of freed blocks (vx_ctx.cu): the per-call work buffers of the sweep-rate
// entry points (LIO update, map insertion, down-sampling) would otherwise pay cudaMalloc + cudaFree (~0.1-0.3 ms each, the
// free with a device-wide synchronisation) on every call — 1.3 ms of a 1.46 ms LIO update before the cache.  A cached block is
// reused only by work issued later; every entry point issues on its context stream and synchronises it before it releases, so
// a reused block is never still in use.  vx_ctx_destroy returns the device's cached blocks to the driver.
void *cached_malloc(size_t bytes, size_t *got);
void cached_free(void *p, size_t bytes);
void cached_trim(int device);

// Grow-only device buffer (byte-addressed).  Freed with the owner.
struct DevBuf {
  void *p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes) {
    if (bytes <= cap) return VX_OK;
    // a LIVE block is being replaced: kernels still queued may read it, and the cache would hand it to another context on
    // another stream at once (the cudaFree this cache replaced synchronised implicitly).  Growth is rare: wait for the device.
    if (p) cudaDeviceSynchronize();
    release();
    size_t want = bytes + bytes / 4 + 256, got = 0;
    p = cached_malloc(want, &got);
    if (!p) { set_error("cudaMalloc(%zu) -> %s", want, cudaGetErrorString(cudaGetLastError())); return VX_ERR_NOMEM; }
    cap = got;
    return VX_OK;
  }
  void release() { if (p) cached_free(p, cap); p = nullptr; cap = 0; }
  template <class T> T *as() const { return (T *)p; }
};

// Wall-clock phase timer for the drivers (VX_TIMING=1 in the environment: synchronises the stream at every lap and prints
// "[tag] phase ms" on stderr; off: no synchronisation, no output).
struct PhaseTimer {
  const char *tag; cudaStream_t st; bool on; double t0;
  static double now() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return 1e3 * ts.tv_sec + 1e-6 * ts.tv_nsec; }
  PhaseTimer(const char *tag_, cudaStream_t st_) : tag(tag_), st(st_) { static const bool e = getenv("VX_TIMING") != nullptr; on = e; t0 = on ? now() : 0.0; }
  void lap(const char *what) {
    if (!on) return;
    cudaStreamSynchronize(st);
    const double t = now();
    fprintf(stderr, "[%s] %s %.3f ms\n", tag, what, t - t0);
    t0 = t;
  }
};

// Optional per-kernel device timing (bench.py's roofline leg): when enabled every VX_LAUNCH_CHECK'd
// launch is bracketed by two events on the context stream; vx_ctx_profile_read sums them by name.
struct ProfSpan { const char *name; cudaEvent_t a, b; };
}  // namespace vx

struct vx_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  // vx_factor_hessian on a batched factor: window groups are evaluated on their own streams (the CTAs of a later group fill
  // the SM slots an earlier one leaves idle) and each group's Hess/JacT go to the host behind its last kernel
  cudaStream_t group_stream[8] = {};
  cudaEvent_t group_ev[8] = {};
  cudaEvent_t fork_ev = nullptr;
  int sm_count = 148;
  int64_t launches = 0;
  bool profiling = false;
  std::vector<vx::ProfSpan> spans;
  cudaEvent_t prof_open = nullptr;
  vx::DevBuf scratch;        // CUB temp storage and short-lived work arrays
  vx::DevBuf scratch2;
  void *pinned = nullptr;    // small pinned staging area for scalar read-backs
  // vx_ldlt.cu: factored diagonal blocks + the captured panel sweep (replayed while n and the buffers stay the same)
  vx::DevBuf ldlt_side, ldlt_mat;
  cudaGraphExec_t ldlt_exec = nullptr;
  int ldlt_n = 0;
  int64_t ldlt_launches = 0;
  double *ldlt_B = nullptr, *ldlt_F = nullptr;
  cudaStream_t ldlt_aux = nullptr;                  // second branch of the sweep (the next diagonal block, beside the trailing update)
  cudaEvent_t ldlt_ev[2] = {nullptr, nullptr};
  size_t pinned_cap = 0;
  struct vx_factor *kept[2] = {nullptr, nullptr};   // vx_hba_global(keep_factors): bottom batch, top level
};

namespace vx {
inline int bind(vx_ctx *ctx) {
  VX_ARG(ctx != nullptr);
  VX_CUDA(cudaSetDevice(ctx->device));
  return VX_OK;
}
inline int blocks_for(int64_t n, int threads) { return (int)((n + threads - 1) / threads); }
}  // namespace vx

#define VX_LAUNCH_CHECK(ctx)                    \
  do {                                          \
    (ctx)->launches++;                          \
    VX_CUDA(cudaGetLastError());                \
  } while (0)

// Launch a kernel under a name: VX_KERNEL(ctx, "k_name", k_name<<<g, b, s, st>>>(args...));
#define VX_KERNEL(ctx, name, ...)                                           \
  do {                                                                      \
    cudaEvent_t _ea = nullptr, _eb = nullptr;                               \
    if ((ctx)->profiling) {                                                 \
      VX_CUDA(cudaEventCreate(&_ea));                                       \
      VX_CUDA(cudaEventCreate(&_eb));                                       \
      VX_CUDA(cudaEventRecord(_ea, (ctx)->stream));                         \
    }                                                                       \
    __VA_ARGS__;                                                            \
    (ctx)->launches++;                                                      \
    VX_CUDA(cudaGetLastError());                                            \
    if ((ctx)->profiling) {                                                 \
      VX_CUDA(cudaEventRecord(_eb, (ctx)->stream));                         \
      (ctx)->spans.push_back(vx::ProfSpan{name, _ea, _eb});                 \
    }                                                                       \
  } while (0)
