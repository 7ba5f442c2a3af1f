// vx_ba.cu — Lidar_BA_Optimizer::damping_iter (voxel_map.hpp:422-497) over the device LidarFactor.
//
// The LM control flow, the gauge fix, the (6W x 6W) LDL^T solve and the pose retraction run on the
// host (W <= 20: 60..120 unknowns, microseconds); acc_evaluate2 / evaluate_only_residual run on the
// device.  A factor created with n_windows > 1 is a BATCH of independent problems (bottom-level GBA
// windows, voxelslam.cpp:3096): every window keeps its own damping state (u, v), accept/reject
// sequence and stop test, and each device pass evaluates all windows at once.
#include "vx_factor.cuh"
#include "vx_hostmat.hpp"
#include <cmath>
#include <limits>
#include <algorithm>
#include <cstdlib>

using namespace vx;

namespace {

// R <- R * Exp(w)   (tools.hpp:51-66: Rodrigues, identity below 1e-11)
void retract_rot(const double *R, const double *w, double *out) {
  const double n = std::sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
  double E[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};   // column-major
  if (n >= 1e-11) {
    const double x = w[0] / n, y = w[1] / n, z = w[2] / n;
    const double K[9] = {0, z, -y, -z, 0, x, y, -x, 0};   // column-major hat(axis)
    double KK[9];
    for (int c = 0; c < 3; c++)
      for (int r = 0; r < 3; r++) KK[c * 3 + r] = K[0 * 3 + r] * K[c * 3 + 0] + K[1 * 3 + r] * K[c * 3 + 1] + K[2 * 3 + r] * K[c * 3 + 2];
    const double s = std::sin(n), c1 = 1.0 - std::cos(n);
    for (int k = 0; k < 9; k++) E[k] += s * K[k] + c1 * KK[k];
  }
  for (int c = 0; c < 3; c++)
    for (int r = 0; r < 3; r++) out[c * 3 + r] = R[0 * 3 + r] * E[c * 3 + 0] + R[1 * 3 + r] * E[c * 3 + 1] + R[2 * 3 + r] * E[c * 3 + 2];
}

}  // namespace

namespace vx {
int ldlt_lower_solve_device(vx_ctx *ctx, const double *A, int n, const double *b, double *x);   // vx_ldlt.cu
// win_status (optional, n_windows): when given, a window with fewer voxels than thd_num is flagged
// VX_ERR_TOO_FEW and left untouched instead of failing the whole batch.
int ba_lidar_impl(vx_ctx *ctx, vx_factor *f, vx_pose *poses, double *hess_out, double *resis, int max_iter, int thd_num,
                  int *converged, int *accept_trace, int *n_iters, int *win_status) {
  VX_ARG(f != nullptr && poses != nullptr && max_iter >= 0);
  VX_TRY(bind(ctx));
  const int W = f->W, NW = f->n_win, jl = 6 * W, F = W * NW;
  std::vector<char> few;
  // Lidar_BA_Optimizer::only_residual exits the process when a problem has fewer voxels than threads
  // (voxel_map.hpp:399-403); the drop-in reports it instead.
  {
    std::vector<int32_t> wv(NW + 1);
    VX_CUDA(cudaMemcpyAsync(wv.data(), f->wv_off.p, sizeof(int32_t) * (NW + 1), cudaMemcpyDeviceToHost, ctx->stream));
    VX_CUDA(cudaStreamSynchronize(ctx->stream));
    few.assign(NW, 0);
    for (int w = 0; w < NW; w++) {
      if (win_status) win_status[w] = VX_OK;
      if (wv[w + 1] - wv[w] < thd_num) {
        if (!win_status) { set_error("Too Less Voxel: window %d has %d voxels < thd_num %d", w, wv[w + 1] - wv[w], thd_num); return VX_ERR_TOO_FEW; }
        win_status[w] = VX_ERR_TOO_FEW; few[w] = 1;
      }
    }
  }
  static const int ldlt_device_min = [] { const char *e = std::getenv("VX_LDLT_DEVICE_MIN"); return e ? std::atoi(e) : 384; }();
  const size_t hsz = (size_t)jl * jl;
  std::vector<double> Hess(hsz * NW), JacT((size_t)jl * NW), Hnew(hsz * NW), Jnew((size_t)jl * NW), r_new(NW), r1(NW, 0.0), r2(NW, 0.0);
  std::vector<double> u(NW, 0.01), v(NW, 2.0), A(hsz), dxi(jl), D(jl), nb(jl), tmp, q1s(NW, 0.0);
  std::vector<char> calc(NW, 1), active(NW, 1), conv(NW, 1);
  std::vector<int> iters(NW, 0), tr;
  std::vector<vx_pose> x(poses, poses + F), xt(poses, poses + F);
  if (hess_out) std::fill(hess_out, hess_out + hsz * NW, 0.0);
  if (resis) std::fill(resis, resis + 2 * (size_t)NW, 0.0);
  for (int w = 0; w < NW; w++) if (few[w]) { active[w] = 0; conv[w] = 0; }

  for (int it = 0; it < max_iter; it++) {
    bool any_active = false, any_calc = false;
    for (int w = 0; w < NW; w++) { any_active |= active[w] != 0; any_calc |= (active[w] && calc[w]); }
    if (!any_active) break;
    if (any_calc) {
      VX_TRY(vx_factor_hessian(ctx, f, x.data(), 0, -1, Hnew.data(), Jnew.data(), r_new.data()));
      for (int w = 0; w < NW; w++)
        if (active[w] && calc[w]) {
          std::copy(Hnew.begin() + hsz * w, Hnew.begin() + hsz * (w + 1), Hess.begin() + hsz * w);
          std::copy(Jnew.begin() + (size_t)jl * w, Jnew.begin() + (size_t)jl * (w + 1), JacT.begin() + (size_t)jl * w);
          r1[w] = r_new[w];
          if (hess_out) std::copy(Hnew.begin() + hsz * w, Hnew.begin() + hsz * (w + 1), hess_out + hsz * w);   // *hess = Hess, before the gauge fix (:446)
        }
    }
    for (int w = 0; w < NW; w++) {
      if (!active[w]) continue;
      if (it == 0 && resis) resis[2 * w] = r1[w];
      double *H = Hess.data() + hsz * w, *J = JacT.data() + (size_t)jl * w;
      // gauge: pose 0 fixed (voxel_map.hpp:452-455)
      for (int c = 0; c < jl; c++) for (int r = 0; r < 6; r++) H[(size_t)c * jl + r] = 0.0;
      for (int c = 0; c < 6; c++) for (int r = 0; r < jl; r++) H[(size_t)c * jl + r] = 0.0;
      for (int r = 0; r < 6; r++) { H[(size_t)r * jl + r] = 1.0; J[r] = 0.0; }
      for (int r = 0; r < jl; r++) D[r] = H[(size_t)r * jl + r];
      std::copy(H, H + hsz, A.begin());
      for (int r = 0; r < jl; r++) { A[(size_t)r * jl + r] += u[w] * D[r]; nb[r] = -J[r]; }
      // small windows: the host sweep (microseconds); the top-level GBA window (W ~ 399, 2394 unknowns, 4.6 GFLOP) goes to
      // the blocked device factorisation (vx_ldlt.cu)
      if (jl >= ldlt_device_min) VX_TRY(ldlt_lower_solve_device(ctx, A.data(), jl, nb.data(), dxi.data()));
      else hm::ldlt_lower_solve(A.data(), jl, nb.data(), dxi.data(), tr, tmp);
      for (int j = 0; j < W; j++) {
        const vx_pose &s = x[(size_t)w * W + j];
        vx_pose &d = xt[(size_t)w * W + j];
        retract_rot(s.R, &dxi[6 * j], d.R);
        for (int k = 0; k < 3; k++) d.p[k] = s.p[k] + dxi[6 * j + 3 + k];
      }
      double q1 = 0.0;
      for (int r = 0; r < jl; r++) q1 += dxi[r] * (u[w] * D[r] * dxi[r] - J[r]);
      q1s[w] = 0.5 * q1;
    }
    VX_TRY(vx_factor_residual(ctx, f, xt.data(), 0, -1, r_new.data()));
    for (int w = 0; w < NW; w++) {
      if (!active[w]) continue;
      r2[w] = r_new[w];
      double q = r1[w] - r2[w];
      const double q1 = q1s[w];
      iters[w]++;
      if (q > 0) {
        for (int j = 0; j < W; j++) x[(size_t)w * W + j] = xt[(size_t)w * W + j];
        q = q / q1;
        v[w] = 2;
        q = 1 - std::pow(2 * q - 1, 3);
        u[w] *= (q < 1.0 / 3 ? 1.0 / 3 : q);
        calc[w] = 1;
        if (accept_trace) accept_trace[(size_t)w * max_iter + it] = 1;
      } else {
        u[w] = u[w] * v[w];
        v[w] = 2 * v[w];
        calc[w] = 0;
        conv[w] = 0;
        if (accept_trace) accept_trace[(size_t)w * max_iter + it] = 0;
        // keep feeding this window's rejected trial to later residual passes until a new one replaces it
      }
      if (std::fabs((r1[w] - r2[w]) / r1[w]) < 1e-6) {
        active[w] = 0;
      }
    }
    // windows that stopped keep the cache of their LAST trial (voxel_map.hpp:317-319): pin xt to it
  }
  for (int w = 0; w < NW; w++) {
    if (resis) resis[2 * w + 1] = r2[w];
    if (converged) converged[w] = conv[w];
    if (n_iters) n_iters[w] = iters[w];
  }
  std::copy(x.begin(), x.end(), poses);
  return VX_OK;
}
}  // namespace vx

extern "C" int vx_ba_lidar(vx_ctx *ctx, vx_factor *f, vx_pose *poses, double *hess_out, double *resis, int max_iter, int thd_num,
                           int *converged, int *accept_trace, int *n_iters) {
  return vx::ba_lidar_impl(ctx, f, poses, hess_out, resis, max_iter, thd_num, converged, accept_trace, n_iters, nullptr);
}
