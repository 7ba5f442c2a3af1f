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
