// vx_ctx.cu — context, error text, version.
#include <algorithm>
#include "vx_common.cuh"

namespace vx {
static thread_local char g_err[512] = "";
void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace vx

extern "C" {

const char *vx_last_error(void) { return vx::g_err; }
int vx_version(void) { return 100; }

int vx_ctx_create(int device, vx_ctx **out) {
  VX_ARG(out != nullptr);
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev <= 0) {
    vx::set_error("no CUDA device (cudaGetDeviceCount: %s); libvoxelba has no CPU fallback", cudaGetErrorString(e));
    return VX_ERR_CUDA;
  }
  VX_ARG(device >= 0 && device < ndev);
  VX_CUDA(cudaSetDevice(device));
  vx_ctx *c = new vx_ctx();
  c->device = device;
  cudaDeviceProp prop;
  VX_CUDA(cudaGetDeviceProperties(&prop, device));
  c->sm_count = prop.multiProcessorCount;
  VX_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  {
    // earlier groups get the higher priority: the block scheduler then drains them first, so their copies start early
    int least = 0, greatest = 0;
    VX_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    for (int g = 0; g < 8; g++)
      VX_CUDA(cudaStreamCreateWithPriority(&c->group_stream[g], cudaStreamNonBlocking, std::min(least, greatest + g)));
  }
  for (auto &e : c->group_ev) VX_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  VX_CUDA(cudaEventCreateWithFlags(&c->fork_ev, cudaEventDisableTiming));
  c->pinned_cap = 1 << 16;
  VX_CUDA(cudaMallocHost(&c->pinned, c->pinned_cap));
  *out = c;
  return VX_OK;
}

int vx_ctx_destroy(vx_ctx *ctx) {
  if (!ctx) return VX_OK;
  cudaSetDevice(ctx->device);
  if (ctx->stream) { cudaStreamSynchronize(ctx->stream); cudaStreamDestroy(ctx->stream); }
  for (auto &g : ctx->group_stream) if (g) { cudaStreamSynchronize(g); cudaStreamDestroy(g); }
  for (auto &e : ctx->group_ev) if (e) cudaEventDestroy(e);
  if (ctx->fork_ev) cudaEventDestroy(ctx->fork_ev);
  for (auto &s : ctx->spans) { cudaEventDestroy(s.a); cudaEventDestroy(s.b); }
  ctx->scratch.release();
  ctx->scratch2.release();
  if (ctx->pinned) cudaFreeHost(ctx->pinned);
  delete ctx;
  return VX_OK;
}

int vx_ctx_sync(vx_ctx *ctx) {
  VX_TRY(vx::bind(ctx));
  VX_CUDA(cudaStreamSynchronize(ctx->stream));
  return VX_OK;
}

int vx_ctx_profile(vx_ctx *ctx, int enable) {
  VX_TRY(vx::bind(ctx));
  VX_CUDA(cudaStreamSynchronize(ctx->stream));
  for (auto &s : ctx->spans) { cudaEventDestroy(s.a); cudaEventDestroy(s.b); }
  ctx->spans.clear();
  ctx->profiling = enable != 0;
  return VX_OK;
}

int vx_ctx_profile_read(vx_ctx *ctx, char *names, size_t names_cap, double *ms, int64_t *counts, int max_entries, int *n_out) {
  VX_TRY(vx::bind(ctx));
  VX_ARG(n_out != nullptr);
  VX_CUDA(cudaStreamSynchronize(ctx->stream));
  std::vector<const char *> nm;
  std::vector<double> tot;
  std::vector<int64_t> cnt;
  for (auto &s : ctx->spans) {
    float t = 0.f;
    VX_CUDA(cudaEventElapsedTime(&t, s.a, s.b));
    size_t k = 0;
    for (; k < nm.size(); k++) if (strcmp(nm[k], s.name) == 0) break;
    if (k == nm.size()) { nm.push_back(s.name); tot.push_back(0.0); cnt.push_back(0); }
    tot[k] += t; cnt[k] += 1;
  }
  int n = 0;
  size_t used = 0;
  if (names && names_cap) names[0] = 0;
  for (size_t k = 0; k < nm.size() && n < max_entries; k++) {
    size_t len = strlen(nm[k]);
    if (names) {
      if (used + len + 2 > names_cap) break;
      memcpy(names + used, nm[k], len); used += len; names[used++] = '\n'; names[used] = 0;
    }
    if (ms) ms[n] = tot[k];
    if (counts) counts[n] = cnt[k];
    n++;
  }
  *n_out = n;
  return VX_OK;
}

void *vx_ctx_stream(vx_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }
int64_t vx_ctx_launch_count(vx_ctx *ctx) { return ctx ? ctx->launches : 0; }

}  // extern "C"
