// vx_frontend.cu — A5: calcBodyVar (voxelslam.hpp:180-200), var_init (:210-234), pvec_update (:242-265).
// One thread per point; 12 B in / 96 B out (var_init), 96 B in / 96 + 24 B out (pvec_update): pure HBM streaming.
// Records are staged through shared memory so the 96-byte AoS rows are read and written with coalesced 8-byte lanes.
#include "vx_frontend.cuh"
#include "vx_math.cuh"

using namespace vx;

namespace {

constexpr int FE_THREADS = 128;

struct Pose12 { double R[9]; double p[3]; };   // column-major R
struct Mat9 { double m[9]; };                   // row-major 3x3 kernel argument

__device__ __forceinline__ void store_records(const double (*rec)[13], vx_point_var *out, int64_t base, int64_t n) {
  // rec: [FE_THREADS][12 (+1 pad)] in shared memory; cooperative coalesced write of the block's rows
  double *o = reinterpret_cast<double *>(out + base);
  const int64_t rows = min((int64_t)FE_THREADS, n - base);
  for (int64_t k = threadIdx.x; k < rows * 12; k += FE_THREADS) o[k] = rec[k / 12][k % 12];
}

// R (col-major) * M (col-major, symmetric or not) * R^T
__device__ __forceinline__ void rot_sandwich(const double *R, const double *M, double *out) {
  double T[9];
  for (int c = 0; c < 3; c++)
    for (int r = 0; r < 3; r++) T[c * 3 + r] = R[0 * 3 + r] * M[c * 3 + 0] + R[1 * 3 + r] * M[c * 3 + 1] + R[2 * 3 + r] * M[c * 3 + 2];
  for (int c = 0; c < 3; c++)
    for (int r = 0; r < 3; r++) out[c * 3 + r] = T[0 * 3 + r] * R[0 * 3 + c] + T[1 * 3 + r] * R[1 * 3 + c] + T[2 * 3 + r] * R[2 * 3 + c];
}

__global__ void __launch_bounds__(FE_THREADS) k_var_init(const float *__restrict__ xyz, int64_t n, Pose12 ext, float range_inc, double dir_var,
                                                        vx_point_var *__restrict__ out) {
  __shared__ double rec[FE_THREADS][13];
  const int64_t base = blockIdx.x * (int64_t)FE_THREADS, i = base + threadIdx.x;
  if (i < n) {
    double pb[3] = {(double)xyz[3 * i], (double)xyz[3 * i + 1], (double)xyz[3 * i + 2]};
    if (pb[2] == 0) pb[2] = 0.0001;                                         // :182-183 (the point itself is modified)
    const float range = (float)sqrt(pb[0] * pb[0] + pb[1] * pb[1] + pb[2] * pb[2]);
    const float range_var = range_inc * range_inc;
    const double nrm = sqrt(pb[0] * pb[0] + pb[1] * pb[1] + pb[2] * pb[2]);
    const double d[3] = {pb[0] / nrm, pb[1] / nrm, pb[2] / nrm};
    double b1[3] = {1.0, 1.0, -(d[0] + d[1]) / d[2]};
    const double n1 = sqrt(b1[0] * b1[0] + b1[1] * b1[1] + b1[2] * b1[2]);
    b1[0] /= n1; b1[1] /= n1; b1[2] /= n1;
    double b2[3] = {b1[1] * d[2] - b1[2] * d[1], b1[2] * d[0] - b1[0] * d[2], b1[0] * d[1] - b1[1] * d[0]};
    const double n2 = sqrt(b2[0] * b2[0] + b2[1] * b2[1] + b2[2] * b2[2]);
    b2[0] /= n2; b2[1] /= n2; b2[2] /= n2;
    // A = range * hat(d) * [b1 b2]  (3x2): hat(d) b = d x b
    const double rg = (double)range;
    const double a1[3] = {rg * (d[1] * b1[2] - d[2] * b1[1]), rg * (d[2] * b1[0] - d[0] * b1[2]), rg * (d[0] * b1[1] - d[1] * b1[0])};
    const double a2[3] = {rg * (d[1] * b2[2] - d[2] * b2[1]), rg * (d[2] * b2[0] - d[0] * b2[2]), rg * (d[0] * b2[1] - d[1] * b2[0])};
    double var[9], wv[9];
    for (int c = 0; c < 3; c++)
      for (int r = 0; r < 3; r++) var[c * 3 + r] = d[r] * (double)range_var * d[c] + (a1[r] * dir_var * a1[c] + a2[r] * dir_var * a2[c]);
    rot_sandwich(ext.R, var, wv);
    double *o = rec[threadIdx.x];
    // the point decides voxel keys downstream: no FMA contraction, the reference's operation order
    for (int r = 0; r < 3; r++)
      o[r] = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(ext.R[0 * 3 + r], pb[0]), __dmul_rn(ext.R[1 * 3 + r], pb[1])), __dmul_rn(ext.R[2 * 3 + r], pb[2])), ext.p[r]);
    for (int k = 0; k < 9; k++) o[3 + k] = wv[k];
  }
  __syncthreads();
  store_records(rec, out, base, n);
}

__global__ void __launch_bounds__(FE_THREADS) k_pvec_update(const vx_point_var *__restrict__ in, int64_t n, Pose12 x, Mat9 rot_var, Mat9 tsl_var,
                                                           vx_point_var *__restrict__ out, double *__restrict__ pwld) {
  __shared__ double rec[FE_THREADS][13];
  const int64_t base = blockIdx.x * (int64_t)FE_THREADS, i = base + threadIdx.x;
  const int64_t rows = min((int64_t)FE_THREADS, n - base);
  {
    const double *src = reinterpret_cast<const double *>(in + base);
    for (int64_t k = threadIdx.x; k < rows * 12; k += FE_THREADS) rec[k / 12][k % 12] = src[k];
  }
  __syncthreads();
  double w[3] = {0, 0, 0};
  if (i < n) {
    double *o = rec[threadIdx.x];
    const double p[3] = {o[0], o[1], o[2]};
    double var[9], rv[9];
    for (int k = 0; k < 9; k++) var[k] = o[3 + k];
    rot_sandwich(x.R, var, rv);
    // hat(p) * rot_var * hat(p)^T, rot_var row-major: .m[r*3+c]
    const double ph[9] = {0, p[2], -p[1], -p[2], 0, p[0], p[1], -p[0], 0};   // column-major hat(p)
    double T[9];
    for (int c = 0; c < 3; c++)
      for (int r = 0; r < 3; r++) T[c * 3 + r] = ph[0 * 3 + r] * rot_var.m[0 * 3 + c] + ph[1 * 3 + r] * rot_var.m[1 * 3 + c] + ph[2 * 3 + r] * rot_var.m[2 * 3 + c];
    for (int c = 0; c < 3; c++)
      for (int r = 0; r < 3; r++) {
        const double hv = T[0 * 3 + r] * ph[0 * 3 + c] + T[1 * 3 + r] * ph[1 * 3 + c] + T[2 * 3 + r] * ph[2 * 3 + c];
        o[3 + c * 3 + r] = rv[c * 3 + r] + hv + tsl_var.m[r * 3 + c];
      }
    for (int r = 0; r < 3; r++)   // pwld feeds the key rule (A1): exact operation order, no contraction
      w[r] = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(x.R[0 * 3 + r], p[0]), __dmul_rn(x.R[1 * 3 + r], p[1])), __dmul_rn(x.R[2 * 3 + r], p[2])), x.p[r]);
  }
  __syncthreads();
  store_records(rec, out, base, n);
  if (pwld && i < n) { pwld[3 * i] = w[0]; pwld[3 * i + 1] = w[1]; pwld[3 * i + 2] = w[2]; }
}

Pose12 pose12(const vx_pose &p) { Pose12 q; for (int k = 0; k < 9; k++) q.R[k] = p.R[k]; for (int k = 0; k < 3; k++) q.p[k] = p.p[k]; return q; }
Mat9 rowmajor(const double *cm) { Mat9 m; for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) m.m[r * 3 + c] = cm[c * 3 + r]; return m; }

}  // namespace

namespace vx {

int var_init_device(vx_ctx *ctx, const float *d_xyz, int64_t n, const vx_pose &ext, double dept_err, double beam_err, vx_point_var *d_out) {
  if (n == 0) return VX_OK;
  // DEG2RAD is PCL 1.10's macro ((x)*0.017453293); degree_inc arrives as float (voxelslam.hpp:180)
  const double s = std::sin((double)(float)beam_err * 0.017453293);
  const double dir_var = std::pow(s, 2);
  VX_KERNEL(ctx, "k_var_init", k_var_init<<<blocks_for(n, FE_THREADS), FE_THREADS, 0, ctx->stream>>>(d_xyz, n, pose12(ext), (float)dept_err, dir_var, d_out));
  return VX_OK;
}

int pvec_update_device(vx_ctx *ctx, const vx_point_var *d_pv_in, int64_t n, const vx_pose &x, const double *rot_var9, const double *tsl_var9,
                       vx_point_var *d_pv_out, double *d_pwld) {
  if (n == 0) return VX_OK;
  VX_KERNEL(ctx, "k_pvec_update", k_pvec_update<<<blocks_for(n, FE_THREADS), FE_THREADS, 0, ctx->stream>>>(d_pv_in, n, pose12(x), rowmajor(rot_var9),
                                                                                                          rowmajor(tsl_var9), d_pv_out, d_pwld));
  return VX_OK;
}

}  // namespace vx

extern "C" {

int vx_var_init(vx_ctx *ctx, const float *xyz, int64_t n, const vx_pose *ext, double dept_err, double beam_err, vx_point_var *out) {
  VX_ARG(ctx && ext && n >= 0 && (n == 0 || (xyz && out)));
  VX_TRY(bind(ctx));
  if (n == 0) return VX_OK;
  VX_TRY(ctx->scratch.reserve(sizeof(float) * 3 * n + 64));
  VX_TRY(ctx->scratch2.reserve(sizeof(vx_point_var) * n));
  VX_CUDA(cudaMemcpyAsync(ctx->scratch.p, xyz, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, ctx->stream));
  VX_TRY(var_init_device(ctx, ctx->scratch.as<float>(), n, *ext, dept_err, beam_err, ctx->scratch2.as<vx_point_var>()));
  VX_CUDA(cudaMemcpyAsync(out, ctx->scratch2.p, sizeof(vx_point_var) * n, cudaMemcpyDeviceToHost, ctx->stream));
  VX_CUDA(cudaStreamSynchronize(ctx->stream));
  return VX_OK;
}

int vx_pvec_update(vx_ctx *ctx, const vx_point_var *pv_in, int64_t n, const vx_pose *x, const double *rot_var9, const double *tsl_var9,
                   vx_point_var *pv_out, double *pwld) {
  VX_ARG(ctx && x && rot_var9 && tsl_var9 && n >= 0 && (n == 0 || (pv_in && pv_out)));
  VX_TRY(bind(ctx));
  if (n == 0) return VX_OK;
  VX_TRY(ctx->scratch.reserve(sizeof(vx_point_var) * n));
  VX_TRY(ctx->scratch2.reserve(sizeof(double) * 3 * n));
  VX_CUDA(cudaMemcpyAsync(ctx->scratch.p, pv_in, sizeof(vx_point_var) * n, cudaMemcpyHostToDevice, ctx->stream));
  VX_TRY(pvec_update_device(ctx, ctx->scratch.as<vx_point_var>(), n, *x, rot_var9, tsl_var9, ctx->scratch.as<vx_point_var>(), ctx->scratch2.as<double>()));
  VX_CUDA(cudaMemcpyAsync(pv_out, ctx->scratch.p, sizeof(vx_point_var) * n, cudaMemcpyDeviceToHost, ctx->stream));
  if (pwld) VX_CUDA(cudaMemcpyAsync(pwld, ctx->scratch2.p, sizeof(double) * 3 * n, cudaMemcpyDeviceToHost, ctx->stream));
  VX_CUDA(cudaStreamSynchronize(ctx->stream));
  return VX_OK;
}

}  // extern "C"

// ---- IMU de-skew: the per-point half of IMUEKF::motion_blur (ekf_imu.hpp:138-163) -----------------------------------------------------
#include <cub/cub.cuh>
#include <thrust/iterator/reverse_iterator.h>
namespace {
struct MinOp { __device__ __forceinline__ float operator()(float a, float b) const { return a < b ? a : b; } };
__global__ void k_deskew_stamp(const vx_point48 *__restrict__ pts, int64_t n, float *__restrict__ stamp) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) stamp[i] = pts[i].curvature;
}
// row-by-row products in the order of a plain triple loop (s = 0; s += a(r,k) * b(k)), no contraction: the arithmetic of a sequential host loop
__device__ __forceinline__ void mv(const double *M, const double *v, double *o) {      // M column-major
#pragma unroll
  for (int r = 0; r < 3; r++) o[r] = __dadd_rn(__dadd_rn(__dadd_rn(0.0, __dmul_rn(M[r], v[0])), __dmul_rn(M[3 + r], v[1])), __dmul_rn(M[6 + r], v[2]));
}
__device__ __forceinline__ void mtv(const double *M, const double *v, double *o) {     // M^T v
#pragma unroll
  for (int r = 0; r < 3; r++) o[r] = __dadd_rn(__dadd_rn(__dadd_rn(0.0, __dmul_rn(M[3 * r], v[0])), __dmul_rn(M[3 * r + 1], v[1])), __dmul_rn(M[3 * r + 2], v[2]));
}
// Exp(ang_vel, dt), tools.hpp:68-84: I + sin(a) K + (1 - cos(a)) K K, K = hat(axis), a = |w| dt; identity when |w| <= 1e-7
__device__ __forceinline__ void exp_rate(const double *w, double dt, double *E) {
  const double n = sqrt(__dadd_rn(__dadd_rn(__dadd_rn(0.0, __dmul_rn(w[0], w[0])), __dmul_rn(w[1], w[1])), __dmul_rn(w[2], w[2])));
  E[0] = E[4] = E[8] = 1.0; E[1] = E[2] = E[3] = E[5] = E[6] = E[7] = 0.0;
  if (!(n > 1e-7)) return;
  const double x = w[0] / n, y = w[1] / n, z = w[2] / n;
  const double K[9] = {0, z, -y, -z, 0, x, y, -x, 0};     // column-major hat(axis)
  const double a = __dmul_rn(n, dt), s = sin(a), c1 = __dadd_rn(1.0, -cos(a));
#pragma unroll
  for (int c = 0; c < 3; c++)
#pragma unroll
    for (int r = 0; r < 3; r++) {
      const double kk = __dadd_rn(__dadd_rn(__dadd_rn(0.0, __dmul_rn(K[r], K[3 * c])), __dmul_rn(K[3 + r], K[3 * c + 1])), __dmul_rn(K[6 + r], K[3 * c + 2]));
      E[3 * c + r] = __dadd_rn(__dadd_rn(E[3 * c + r], __dmul_rn(s, K[3 * c + r])), __dmul_rn(c1, kk));
    }
}
struct DeskewFrame { double xR[9], xp[3], lR[9], lp[3]; };
__device__ __forceinline__ void deskew_one(const vx_imu_pose &h, const DeskewFrame &F, float stamp, float &x, float &y, float &z) {
  const double dt = __dadd_rn((double)stamp, -h.t);
  double E[9], Ri[9];
  exp_rate(h.angvel, dt, E);
#pragma unroll
  for (int c = 0; c < 3; c++) mv(h.R, E + 3 * c, Ri + 3 * c);             // R_i = R_imu * Exp(angvel_avr, dt)
  double T[3];
#pragma unroll
  for (int k = 0; k < 3; k++)                                             // pos + vel dt + 0.5 acc dt dt - xc.p, left to right
    T[k] = __dadd_rn(__dadd_rn(__dadd_rn(h.p[k], __dmul_rn(h.v[k], dt)), __dmul_rn(__dmul_rn(__dmul_rn(0.5, h.acc[k]), dt), dt)), -F.xp[k]);
  const double P[3] = {(double)x, (double)y, (double)z};
  double a[3], b[3], c[3], d[3];
  mv(F.lR, P, a);
#pragma unroll
  for (int k = 0; k < 3; k++) a[k] = __dadd_rn(a[k], F.lp[k]);            // Lid_rot_to_IMU * P_i + Lid_offset_to_IMU
  mv(Ri, a, b);
#pragma unroll
  for (int k = 0; k < 3; k++) b[k] = __dadd_rn(b[k], T[k]);
  mtv(F.xR, b, c);
#pragma unroll
  for (int k = 0; k < 3; k++) c[k] = __dadd_rn(c[k], -F.lp[k]);
  mtv(F.lR, c, d);
  x = (float)d[0]; y = (float)d[1]; z = (float)d[2];
}
__global__ void k_deskew(vx_point48 *__restrict__ pts, const float *__restrict__ sufmin, int64_t n, const vx_imu_pose *__restrict__ poses, int np, DeskewFrame F,
                         int32_t *__restrict__ assign) {
  const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (j >= n) return;
  const double m = (double)sufmin[j];
  int lo = -1, hi = np;                       // largest i with poses[i].t < m (t ascending)
  while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (poses[mid].t < m) lo = mid; else hi = mid; }
  if (assign) assign[j] = lo;
  if (lo < 0) return;
  float x = pts[j].x, y = pts[j].y, z = pts[j].z;
  const float stamp = pts[j].curvature;
  deskew_one(poses[lo], F, stamp, x, y, z);
  if (j == 0) {                               // pcl_in.begin() is compensated again by every earlier pose (the walk never moves past it)
    for (int i = lo - 1; i >= 0; i--)
      if ((double)stamp > poses[i].t) { deskew_one(poses[i], F, stamp, x, y, z); if (assign) assign[0] = i; }
  }
  pts[j].x = x; pts[j].y = y; pts[j].z = z;
}
}  // namespace

extern "C" int vx_deskew(vx_ctx *ctx, vx_point48 *pts, int64_t n, const vx_imu_pose *poses, int n_poses, const vx_pose *xc, const vx_pose *lidar_to_imu,
                         int32_t *assign) {
  VX_ARG(ctx && n >= 0 && n < ((int64_t)1 << 31) && n_poses >= 0 && xc && lidar_to_imu && (n == 0 || pts) && (n_poses == 0 || poses));
  VX_TRY(bind(ctx));
  if (n == 0) return VX_OK;
  if (n_poses == 0) { if (assign) for (int64_t k = 0; k < n; k++) assign[k] = -1; return VX_OK; }
  for (int i = 1; i < n_poses; i++) VX_ARG(poses[i].t >= poses[i - 1].t);
  cudaStream_t st = ctx->stream;
  DevBuf d_pts, d_st, d_min, d_poses, d_as;
  DeskewFrame F;
  for (int k = 0; k < 9; k++) { F.xR[k] = xc->R[k]; F.lR[k] = lidar_to_imu->R[k]; }
  for (int k = 0; k < 3; k++) { F.xp[k] = xc->p[k]; F.lp[k] = lidar_to_imu->p[k]; }
  auto body = [&]() -> int {
    VX_TRY(d_pts.reserve(sizeof(vx_point48) * n)); VX_TRY(d_st.reserve(4 * n)); VX_TRY(d_min.reserve(4 * n));
    VX_TRY(d_poses.reserve(sizeof(vx_imu_pose) * n_poses)); VX_TRY(d_as.reserve(4 * n));
    VX_CUDA(cudaMemcpyAsync(d_pts.p, pts, sizeof(vx_point48) * n, cudaMemcpyHostToDevice, st));
    VX_CUDA(cudaMemcpyAsync(d_poses.p, poses, sizeof(vx_imu_pose) * n_poses, cudaMemcpyHostToDevice, st));
    VX_KERNEL(ctx, "k_deskew_stamp", k_deskew_stamp<<<blocks_for(n, 256), 256, 0, st>>>(d_pts.as<vx_point48>(), n, d_st.as<float>()));
    // minimum stamp from each point to the end of the sweep: an inclusive min-scan over the reversed array
    auto rin = thrust::make_reverse_iterator(d_st.as<float>() + n);
    auto rout = thrust::make_reverse_iterator(d_min.as<float>() + n);
    size_t tb = 0;
    VX_CUDA(cub::DeviceScan::InclusiveScan(nullptr, tb, rin, rout, MinOp(), (int)n, st));
    VX_TRY(ctx->scratch.reserve(tb));
    VX_CUDA(cub::DeviceScan::InclusiveScan(ctx->scratch.p, tb, rin, rout, MinOp(), (int)n, st));
    ctx->launches += 2;
    VX_KERNEL(ctx, "k_deskew", k_deskew<<<blocks_for(n, 128), 128, 0, st>>>(d_pts.as<vx_point48>(), d_min.as<float>(), n, d_poses.as<vx_imu_pose>(), n_poses, F,
                                                                         assign ? d_as.as<int32_t>() : nullptr));
    VX_CUDA(cudaMemcpyAsync(pts, d_pts.p, sizeof(vx_point48) * n, cudaMemcpyDeviceToHost, st));
    if (assign) VX_CUDA(cudaMemcpyAsync(assign, d_as.p, 4 * n, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    return VX_OK;
  };
  const int rc = body();
  cudaStreamSynchronize(st);
  for (DevBuf *b : {&d_pts, &d_st, &d_min, &d_poses, &d_as}) b->release();
  return rc;
}
