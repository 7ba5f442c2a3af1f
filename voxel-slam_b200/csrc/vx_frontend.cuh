// vx_frontend.cuh — per-point covariance front-end (A5) shared by the C entry points and the LIO driver.
#pragma once
#include "vx_common.cuh"

namespace vx {
// calcBodyVar + var_init (voxelslam.hpp:180-234) for n device-resident float xyz triples -> n vx_point_var records.
int var_init_device(vx_ctx *ctx, const float *d_xyz, int64_t n, const vx_pose &ext, double dept_err, double beam_err, vx_point_var *d_out);
// pvec_update (voxelslam.hpp:242-265): d_var_out / d_pwld may alias nothing; d_pv_out may equal d_pv_in.
int pvec_update_device(vx_ctx *ctx, const vx_point_var *d_pv_in, int64_t n, const vx_pose &x, const double *rot_var9, const double *tsl_var9,
                       vx_point_var *d_pv_out, double *d_pwld);
}  // namespace vx
