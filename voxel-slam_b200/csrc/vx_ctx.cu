// vx_ctx.cu — context, error text, version.
#include <algorithm>
#include <mutex>
#include <cstdlib>
#include "vx_common.cuh"

namespace vx {
// ---- cache of freed device blocks (see DevBuf, vx_common.cuh) ---------------------------------------------------
namespace {
struct FreeBlock { void *p; size_t bytes; int device; };
std::mutex g_cache_mu;
std::vector<FreeBlock> g_cache;
size_t g_cache_bytes = 0;
// retained at most (per process, all devices); beyond it blocks go back to the driver.  A 2,000-keyframe global BA cycles through
// ~6 GB of work buffers per map build at two very different sizes (80 M point instances at the bottom level, 35 M at the top):
// with 8 GiB retained every build paid cudaFree (device-wide synchronisation) + cudaMalloc for most of them — 100 of the 135 ms
// of a bottom-level build.  B200 has 180 GB; VX_CACHE_MAX_GB overrides.
const size_t CACHE_MAX_BYTES = [] { const char *e = std::getenv("VX_CACHE_MAX_GB"); return (size_t)(e ? std::atoi(e) : 48) << 30; }();
constexpr size_t CACHE_MAX_BLOCKS = 2048;
}  // namespace

void *cached_malloc(size_t bytes, size_t *got) {
  int dev = 0;
  cudaGetDevice(&dev);
  {
    std::lock_guard<std::mutex> lk(g_cache_mu);
    int best = -1;
    for (int i = 0; i < (int)g_cache.size(); i++) {
      const FreeBlock &b = g_cache[i];
      if (b.device != dev || b.bytes < bytes || b.bytes > 2 * bytes + 4096) continue;   // best fit, at most 2x oversized
      if (best < 0 || b.bytes < g_cache[best].bytes) best = i;
    }
    if (best >= 0) {
      FreeBlock b = g_cache[best];
      g_cache.erase(g_cache.begin() + best);
      g_cache_bytes -= b.bytes;
      *got = b.bytes;
      return b.p;
    }
  }
  void *p = nullptr;
  if (cudaMalloc(&p, bytes) != cudaSuccess) {
    cudaGetLastError();
    cached_trim(dev);                                   // out of memory: give the cached blocks back and try once more
    if (cudaMalloc(&p, bytes) != cudaSuccess) return nullptr;
  }
  *got = bytes;
  return p;
}

void cached_free(void *p, size_t bytes) {
  if (!p) return;
  cudaPointerAttributes at;
  int dev = 0;
  if (cudaPointerGetAttributes(&at, p) == cudaSuccess) dev = at.device; else { cudaGetLastError(); cudaGetDevice(&dev); }
  {
    std::lock_guard<std::mutex> lk(g_cache_mu);
    if (g_cache_bytes + bytes <= CACHE_MAX_BYTES && g_cache.size() < CACHE_MAX_BLOCKS) {
      g_cache.push_back({p, bytes, dev});
      g_cache_bytes += bytes;
      return;
    }
  }
  cudaFree(p);
}

void cached_trim(int device) {
  std::vector<FreeBlock> out;
  {
    std::lock_guard<std::mutex> lk(g_cache_mu);
    for (size_t i = 0; i < g_cache.size();) {
      if (g_cache[i].device == device) { out.push_back(g_cache[i]); g_cache_bytes -= g_cache[i].bytes; g_cache.erase(g_cache.begin() + i); }
      else i++;
    }
  }
  int cur = 0;
  cudaGetDevice(&cur);
  if (!out.empty()) cudaSetDevice(device);
  for (auto &b : out) cudaFree(b.p);
  if (!out.empty()) cudaSetDevice(cur);
}

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

int vx_ctx_destroy(vx_ctx *ctx);
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
  auto init = [&]() -> int {
    cudaDeviceProp prop;
    VX_CUDA(cudaGetDeviceProperties(&prop, device));
    c->sm_count = prop.multiProcessorCount;
    VX_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    // earlier groups get the higher priority: the block scheduler then drains them first, so their copies start early
    int least = 0, greatest = 0;
    VX_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    for (int g = 0; g < 8; g++)
      VX_CUDA(cudaStreamCreateWithPriority(&c->group_stream[g], cudaStreamNonBlocking, std::min(least, greatest + g)));
    for (auto &e : c->group_ev) VX_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    VX_CUDA(cudaEventCreateWithFlags(&c->fork_ev, cudaEventDisableTiming));
    c->pinned_cap = 1 << 16;
    VX_CUDA(cudaMallocHost(&c->pinned, c->pinned_cap));
    return VX_OK;
  };
  const int rc = init();
  if (rc != VX_OK) { vx_ctx_destroy(c); return rc; }   // no leak of the context and its streams on a failed set-up
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
  for (auto &k : ctx->kept) if (k) { vx_factor_free(k); k = nullptr; }
  ctx->scratch.release();
  ctx->scratch2.release();
  if (ctx->ldlt_exec) cudaGraphExecDestroy(ctx->ldlt_exec);
  if (ctx->ldlt_aux) cudaStreamDestroy(ctx->ldlt_aux);
  for (cudaEvent_t e : ctx->ldlt_ev) if (e) cudaEventDestroy(e);
  ctx->ldlt_side.release();
  ctx->ldlt_mat.release();
  if (ctx->pinned) cudaFreeHost(ctx->pinned);
  vx::cached_trim(ctx->device);
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
