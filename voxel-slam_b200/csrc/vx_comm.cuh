// vx_comm.cuh — the multi-GPU communicator of libvoxelba: one process per GPU, NCCL over NVLink / NVSwitch.
//
// The reference parallelises this path over host threads and sums the per-thread partial Hess / JacT / cost
// (voxel_map.hpp:373-386); across GPUs the same sum is an all-reduce, and the voxelization of a map that does
// not fit one rank's share is one exchange by root-voxel owner.  NCCL is bound at run time
// (dlopen of libnccl.so.2, or the copy a host process such as PyTorch already loaded), so the library loads
// and runs single-GPU on a machine without NCCL; vx_comm_create then returns VX_ERR_NCCL.
#pragma once
#include "vx_common.cuh"

struct vx_comm {
  vx_ctx *ctx = nullptr;
  void *nccl = nullptr;       // ncclComm_t
  int rank = 0, size = 1;
  vx::DevBuf pack;            // packed [lower triangle | JacT | cost] of an evaluation
  vx::DevBuf counts;          // size x size int64 exchange counts
  int64_t bytes_allreduced = 0, bytes_exchanged = 0, n_collectives = 0;
};

namespace vx {
// in-place SUM over ranks of n doubles on the context stream (no synchronisation)
int comm_allreduce_f64(vx_comm *c, double *d_buf, int64_t n);
int comm_allreduce_i64(vx_comm *c, int64_t *d_buf, int64_t n);
// every rank contributes `bytes` bytes; recv holds size * bytes, rank-major
int comm_allgather_bytes(vx_comm *c, const void *d_send, void *d_recv, size_t bytes);
// grouped ncclSend / ncclRecv all-to-all: rank r receives recv_cnt[s] records of `rec` bytes from every s (recv_off[s] =
// record offset in d_recv) and sends send_cnt[d] records from d_send + send_off[d] * rec to every d
int comm_alltoallv(vx_comm *c, const void *d_send, const int64_t *send_off, const int64_t *send_cnt, void *d_recv,
                   const int64_t *recv_off, const int64_t *recv_cnt, size_t rec);
}  // namespace vx
