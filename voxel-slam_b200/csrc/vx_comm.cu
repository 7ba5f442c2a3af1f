// vx_comm.cu — NCCL communicator behind the C ABI (vx_comm_*): SURVEY §8e, VERDICT r01 "multi-GPU behind the C ABI".
// The collectives replace the reference's host-thread partial sums (voxel_map.hpp:382-384) across GPUs.
#include "vx_comm.cuh"
#include <nccl.h>
#include <dlfcn.h>
#include <mutex>

namespace {
struct NcclApi {
  void *lib = nullptr;
  decltype(&ncclGetVersion) GetVersion = nullptr;
  decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
  decltype(&ncclCommInitRank) CommInitRank = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
  decltype(&ncclAllReduce) AllReduce = nullptr;
  decltype(&ncclAllGather) AllGather = nullptr;
  decltype(&ncclSend) Send = nullptr;
  decltype(&ncclRecv) Recv = nullptr;
  decltype(&ncclGroupStart) GroupStart = nullptr;
  decltype(&ncclGroupEnd) GroupEnd = nullptr;
  bool ok = false;
};
NcclApi g_nccl;
std::once_flag g_nccl_once;

void nccl_bind() {
  // a copy the process already holds (PyTorch bundles its own libnccl.so.2) wins over the system one: two NCCL builds in
  // one process would each keep their own topology caches and proxy threads
  const char *names[] = {"libnccl.so.2", "libnccl.so"};
  void *h = nullptr;
  for (const char *n : names) { h = dlopen(n, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL); if (h) break; }
  if (!h) for (const char *n : names) { h = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
  if (!h) return;
  g_nccl.lib = h;
#define VX_NCCL_SYM(name) g_nccl.name = (decltype(g_nccl.name))dlsym(h, "nccl" #name); if (!g_nccl.name) return
  VX_NCCL_SYM(GetVersion); VX_NCCL_SYM(GetUniqueId); VX_NCCL_SYM(CommInitRank); VX_NCCL_SYM(CommDestroy); VX_NCCL_SYM(GetErrorString);
  VX_NCCL_SYM(AllReduce); VX_NCCL_SYM(AllGather); VX_NCCL_SYM(Send); VX_NCCL_SYM(Recv); VX_NCCL_SYM(GroupStart); VX_NCCL_SYM(GroupEnd);
#undef VX_NCCL_SYM
  g_nccl.ok = true;
}
int nccl_ready() {
  std::call_once(g_nccl_once, nccl_bind);
  if (!g_nccl.ok) { vx::set_error("NCCL unavailable: libnccl.so.2 not found (%s)", dlerror() ? dlerror() : "missing symbol"); return VX_ERR_NCCL; }
  return VX_OK;
}
}  // namespace

#define VX_NCCL(call)                                                                                   \
  do {                                                                                                  \
    ncclResult_t _r = (call);                                                                           \
    if (_r != ncclSuccess) {                                                                            \
      vx::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(_r));            \
      return VX_ERR_NCCL;                                                                               \
    }                                                                                                   \
  } while (0)

namespace vx {
int comm_allreduce_f64(vx_comm *c, double *d_buf, int64_t n) {
  if (!c || c->size == 1 || n == 0) return VX_OK;
  VX_NCCL(g_nccl.AllReduce(d_buf, d_buf, (size_t)n, ncclDouble, ncclSum, (ncclComm_t)c->nccl, c->ctx->stream));
  c->bytes_allreduced += 8 * n; c->n_collectives++;
  return VX_OK;
}
int comm_allreduce_i64(vx_comm *c, int64_t *d_buf, int64_t n) {
  if (!c || c->size == 1 || n == 0) return VX_OK;
  VX_NCCL(g_nccl.AllReduce(d_buf, d_buf, (size_t)n, ncclInt64, ncclSum, (ncclComm_t)c->nccl, c->ctx->stream));
  c->bytes_allreduced += 8 * n; c->n_collectives++;
  return VX_OK;
}
int comm_allgather_bytes(vx_comm *c, const void *d_send, void *d_recv, size_t bytes) {
  if (bytes == 0) return VX_OK;
  if (!c || c->size == 1) {
    if (d_send != d_recv) VX_CUDA(cudaMemcpyAsync(d_recv, d_send, bytes, cudaMemcpyDeviceToDevice, c ? c->ctx->stream : nullptr));
    return VX_OK;
  }
  VX_NCCL(g_nccl.AllGather(d_send, d_recv, bytes, ncclChar, (ncclComm_t)c->nccl, c->ctx->stream));
  c->bytes_exchanged += (int64_t)bytes * c->size; c->n_collectives++;
  return VX_OK;
}
int comm_alltoallv(vx_comm *c, const void *d_send, const int64_t *send_off, const int64_t *send_cnt, void *d_recv,
                   const int64_t *recv_off, const int64_t *recv_cnt, size_t rec) {
  VX_ARG(c != nullptr);
  cudaStream_t st = c->ctx->stream;
  // the rank's own share never leaves the device
  if (send_cnt[c->rank] > 0)
    VX_CUDA(cudaMemcpyAsync((char *)d_recv + (size_t)recv_off[c->rank] * rec, (const char *)d_send + (size_t)send_off[c->rank] * rec,
                            (size_t)send_cnt[c->rank] * rec, cudaMemcpyDeviceToDevice, st));
  if (c->size == 1) return VX_OK;
  VX_NCCL(g_nccl.GroupStart());
  for (int p = 0; p < c->size; p++) {
    if (p == c->rank) continue;
    if (send_cnt[p] > 0) {
      VX_NCCL(g_nccl.Send((const char *)d_send + (size_t)send_off[p] * rec, (size_t)send_cnt[p] * rec, ncclChar, p, (ncclComm_t)c->nccl, st));
      c->bytes_exchanged += send_cnt[p] * (int64_t)rec;
    }
    if (recv_cnt[p] > 0) VX_NCCL(g_nccl.Recv((char *)d_recv + (size_t)recv_off[p] * rec, (size_t)recv_cnt[p] * rec, ncclChar, p, (ncclComm_t)c->nccl, st));
  }
  VX_NCCL(g_nccl.GroupEnd());
  c->n_collectives++;
  return VX_OK;
}
}  // namespace vx

extern "C" {

int vx_comm_unique_id(void *id128) {
  VX_ARG(id128 != nullptr);
  VX_TRY(nccl_ready());
  static_assert(sizeof(ncclUniqueId) == VX_COMM_ID_BYTES, "ncclUniqueId is 128 bytes");
  ncclUniqueId id;
  VX_NCCL(g_nccl.GetUniqueId(&id));
  memcpy(id128, &id, sizeof(id));
  return VX_OK;
}

int vx_comm_create(vx_ctx *ctx, const void *id128, int rank, int n_ranks, vx_comm **out) {
  VX_ARG(out != nullptr);
  *out = nullptr;
  VX_ARG(ctx != nullptr && n_ranks >= 1 && rank >= 0 && rank < n_ranks && (n_ranks == 1 || id128 != nullptr));
  VX_TRY(vx::bind(ctx));
  vx_comm *c = new vx_comm();
  c->ctx = ctx; c->rank = rank; c->size = n_ranks;
  if (n_ranks > 1) {
    int s = nccl_ready();
    if (s != VX_OK) { delete c; return s; }
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    ncclComm_t nc = nullptr;
    ncclResult_t r = g_nccl.CommInitRank(&nc, n_ranks, id, rank);
    if (r != ncclSuccess) {
      vx::set_error("ncclCommInitRank(rank %d of %d) -> %s", rank, n_ranks, g_nccl.GetErrorString(r));
      delete c;
      return VX_ERR_NCCL;
    }
    c->nccl = nc;
  }
  *out = c;
  return VX_OK;
}

int vx_comm_free(vx_comm *c) {
  if (!c) return VX_OK;
  if (c->ctx) { cudaSetDevice(c->ctx->device); cudaStreamSynchronize(c->ctx->stream); }
  if (c->nccl) g_nccl.CommDestroy((ncclComm_t)c->nccl);
  c->pack.release(); c->counts.release();
  delete c;
  return VX_OK;
}

int vx_comm_rank(const vx_comm *c) { return c ? c->rank : 0; }
int vx_comm_size(const vx_comm *c) { return c ? c->size : 1; }
int vx_comm_nccl_version(void) {
  if (nccl_ready() != VX_OK) return 0;
  int v = 0;
  return g_nccl.GetVersion(&v) == ncclSuccess ? v : 0;
}

int vx_comm_allreduce_f64(vx_comm *c, void *d_buf, int64_t n) {
  VX_ARG(c != nullptr && n >= 0 && (n == 0 || d_buf != nullptr));
  VX_TRY(vx::bind(c->ctx));
  return vx::comm_allreduce_f64(c, (double *)d_buf, n);
}

int vx_comm_stats(const vx_comm *c, int64_t *bytes_allreduced, int64_t *bytes_exchanged, int64_t *n_collectives) {
  VX_ARG(c != nullptr);
  if (bytes_allreduced) *bytes_allreduced = c->bytes_allreduced;
  if (bytes_exchanged) *bytes_exchanged = c->bytes_exchanged;
  if (n_collectives) *n_collectives = c->n_collectives;
  return VX_OK;
}

}  // extern "C"
