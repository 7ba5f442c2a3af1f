// vx_factor.cuh — device-resident LidarFactor (reference: class LidarFactor, voxel_map.hpp:124-339).
//
// HBM layout (all SoA, "planes" = component-major so a warp reading one component of 32 consecutive
// records issues one coalesced request):
//   voxel-major CSR over voxel-frames (a voxel-frame = a (voxel, frame) pair with N != 0):
//     off[V+1]            int32   first voxel-frame of each voxel
//     frame[nnz]          int32   global frame id (window*W + frame) of each voxel-frame, ascending per voxel
//     vox[nnz]            int32   owning voxel
//     mom[9][nnz]         double  P00 P01 P02 P11 P12 P22 v0 v1 v2 of the scan-local PointCluster
//     cnt[nnz]            int32   N of the cluster                      -> 76 B per voxel-frame (+8 B index)
//   per voxel:
//     fix[10][V]          double  sig_vecs (P6, v3, N)
//     coe[V]              double
//     lam[3][V] U[9][V]   double  cached eig_values / eig_vectors (column-major U)
//     mrg[10][V]          double  cached pcr_adds (P6, v3, N)
//     win[V]              int32   window of the voxel (batch mode), voxels sorted by window
#pragma once
#include "vx_common.cuh"

struct vx_factor {
  vx_ctx *ctx = nullptr;
  int W = 0, n_win = 1;
  int64_t V = 0, nnz = 0;
  int F() const { return W * n_win; }

  // content
  vx::DevBuf off, frame, vox, mom, cnt, fix, coe, lam, U, mrg, win, wv_off;
  // evaluation state
  vx::DevBuf poses;                 // F x 12 doubles
  vx::DevBuf H, J, cost_h, cost_r;  // outputs
  vx::DevBuf cS;                    // [3][V] coe * {2/(l0-l1), 2/(l0-l2), -2/NN^2}; the fused path adds [3][V] vBar and [V] 1/NN
  vx::DevBuf res_v;                 // [V] coe * lambda_0
  vx::DevBuf G;                     // [nnz][18] per voxel-frame rows g1 g2 w
  // reduction plans (built on device at upload)
  bool plan_ready = false;
  // voxel range whose division planes cS[3..6] (and cS[0..2]) match the cache lam / mrg / coe: written by k_residual with the
  // cache, rebuilt by k_cost_prep when an evaluation finds its range uncovered.  Every other writer of the cache empties it.
  int64_t prep_lo = 0, prep_hi = 0;
  vx::DevBuf fm_order, fm_off;                          // frame-major order of voxel-frames
  vx::DevBuf dseg_frame, dseg_beg, dseg_end, dseg_off;  // diagonal-tile segments
  vx::DevBuf pair_ab;                                   // packed (a<<32 | b), sorted by tile
  vx::DevBuf tile_key, tile_off;                        // tile = (frame_i * F + frame_j), i<j
  vx::DevBuf oseg_tile, oseg_beg, oseg_end, oseg_off;   // off-diagonal tile segments
  vx::DevBuf dpart, opart;                              // per-segment partial sums
  vx::DevBuf part_t, part_d;                            // fused small-window path: per-item tile / diagonal partials
  vx::DevBuf lm_H, lm_J, lm_dx, lm_q1, lm_u, lm_mask;   // vx_ba_lidar: Hess / JacT each window last accepted, step, damping, masks
  int64_t n_pairs = 0, n_tiles = 0, n_dseg = 0, n_oseg = 0;
  int seg_len = 1024;
  int64_t n_hess_evals = 0, n_res_evals = 0;   // acc_evaluate2 / evaluate_only_residual passes launched on this factor
  bool force_pairlist = false;   // tests: evaluate through the generic pair-list path even when a fused one exists
};

namespace vx {
// (Re)builds vox[], the frame-major order and the tile/segment plans from off[]/frame[] on device.
int factor_build_plan(vx_factor *f);
// Fills vox[] after off[]/frame[] changed and invalidates the plans (rebuilt lazily by the generic path).
int factor_finish_upload(vx_factor *f);
// Allocates per-voxel / per-voxel-frame arrays for V voxels and nnz voxel-frames.
int factor_alloc(vx_factor *f, int64_t V, int64_t nnz);
// Fused tensor-core path for small windows (vx_hess_small.cu); *handled=false when W has no instantiation.
// with_cost: the finaliser also writes the windows' cost (cost_h) from the cached eigenvalues, bitwise what k_cost gives.
int hessian_small(vx_ctx *ctx, vx_factor *f, int64_t head, int64_t end, bool *handled, int w_lo = 0, int w_hi = -1, bool with_cost = false);
inline bool hessian_small_handles(int W) { return W == 5 || W == 8 || W == 10 || W == 12 || W == 15 || W == 16 || W == 20; }
}  // namespace vx
