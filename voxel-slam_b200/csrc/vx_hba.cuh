// vx_hba.cuh — internals shared by the HBA entry points (vx_hba.cu, vx_hba_global.cu).
#pragma once
#include <functional>
#include "vx_factor.cuh"
#include "vx_comm.cuh"

namespace vx {
int gba_build_impl(vx_ctx *ctx, const float *d_xyz_in, const int64_t *d_coff_in, const int64_t *cloud_off, int n_clouds, const int32_t *frame_cloud,
                   const vx_pose *poses, const vx_gba_params *params, vx_factor *f, const int32_t *d_pt_frame, int64_t n_pts);
int gba_build_sharded_impl(vx_ctx *ctx, vx_comm *comm, const float *d_xyz, const int64_t *cloud_off, int n_clouds, const int32_t *cloud_frame,
                           const vx_pose *poses, const vx_gba_params *params, vx_factor *f, int64_t *n_owned);
int ba_lidar_impl(vx_ctx *ctx, vx_factor *f, vx_pose *poses, double *hess_out, double *resis, int max_iter, int thd_num,
                  int *converged, int *accept_trace, int *n_iters, int *win_status, vx_comm *comm, double *edge_diag);
using HbaBuildFn = std::function<int(const std::vector<int> &ws, const vx_pose *ps, const vx_gba_params &gp, vx_factor *fac)>;
int hba_core(vx_ctx *ctx, vx_comm *comm, const HbaBuildFn &build, const int32_t *frame_id, const int32_t *frame_mp, vx_pose *poses, int W, int NW,
             const vx_hba_params *prm, int max_iter, int thread_num, vx_pgo_edge *edges, int64_t *n_edges, double *hess_out, double *resis_out,
             int32_t *n_voxels_out, int32_t *status_out, int64_t *n_voxel_frames_evaluated, vx_factor **keep = nullptr);
int allreduce_evaluation(vx_ctx *ctx, vx_comm *comm, vx_factor *f);
// Merged submap clouds of a batch of windows on the device (voxelslam.cpp:2954-2985): window w = frames [wf_off[w], wf_off[w+1]),
// every frame's cloud moved into the frame of the window's first pose, down_sampling_voxel(leaf) per window.
//   d_out_xyz   float triples of the cells, windows back to back (ascending packed cell key inside a window); reserved here
//   win_off     host, n_windows + 1
//   d_out48     optional vx_point48 records (intensity = frame_mp of the cell's first point, curvature = count)
// Returns VX_ERR_ARG with *overflow = 1 when a cell index does not fit the 18-bit batched key (caller falls back per window).
int cloud_bboxes(vx_ctx *ctx, const float *d_xyz, const int64_t *d_coff, int n_clouds, float *d_bbox);   // vx_gba.cu
int merge_clouds_batched(vx_ctx *ctx, const float *d_xyz, const int64_t *cloud_off, int n_clouds, const int32_t *frame_cloud, const int32_t *frame_mp,
                         const vx_pose *poses, const int32_t *wf_off, int n_windows, double leaf, DevBuf &d_out_xyz, int64_t *win_off, DevBuf *d_out48,
                         int *overflow);
}  // namespace vx
