// vx_ba.cu — Lidar_BA_Optimizer::damping_iter (voxel_map.hpp:422-497) over the device LidarFactor.
//
// acc_evaluate2 / evaluate_only_residual run on the device; so do the gauge fix and the (6W x 6W) LDL^T solve of every
// window with up to 128 unknowns (k_lm_solve below: W <= 21, i.e. every local BA and every bottom-level GBA window).  The
// LM control flow and the pose retraction (6W numbers per window) stay on the host.  Larger systems (top-level GBA) take
// the host sweep or the blocked device factorisation of vx_ldlt.cu.  A factor created with n_windows > 1 is a BATCH of independent problems (bottom-level GBA
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


// ---- batched LM step on the device (6W <= LM_MAXN unknowns per window) ------------------------------------------
// One CTA per window, the system in shared memory, one thread per row.  The arithmetic is the host sweep's
// (hm::ldlt_lower_solve, vx_hostmat.hpp: Eigen's unblocked diagonal-pivot LDL^T on the lower triangle) operation for
// operation — explicit round-to-nearest multiplies and adds, no FMA contraction, sums in the same order — so dx, and with
// it every accept / reject decision of the LM loop, is bitwise what the host solve gives.  What it removes is the traffic:
// the 6W x 6W Hessians stay in HBM (11.7 MB per LM iteration at 399 windows), only dx and q1 (6W + 1 doubles per window)
// cross PCIe, and 399 solves run side by side instead of one after another on a host core.
constexpr int LM_MAXN = 128;

// windows with mask[w] != 0 take the new evaluation; the others keep their old Hess / JacT (voxel_map.hpp:486-489)
__global__ void k_lm_keep(const double *__restrict__ H, const double *__restrict__ J, double *__restrict__ Hk, double *__restrict__ Jk,
                          const int32_t *__restrict__ mask, int n) {
  const int w = blockIdx.x;
  if (!mask[w]) return;
  const size_t nn = (size_t)n * n;
  const size_t t0 = blockIdx.y * (size_t)blockDim.x + threadIdx.x, step = (size_t)gridDim.y * blockDim.x;
  if (nn % 2 == 0) {
    const double2 *src = reinterpret_cast<const double2 *>(H + nn * w);
    double2 *dst = reinterpret_cast<double2 *>(Hk + nn * w);
    for (size_t k = t0; k < nn / 2; k += step) dst[k] = src[k];
  } else {
    for (size_t k = t0; k < nn; k += step) Hk[nn * w + k] = H[nn * w + k];
  }
  for (size_t k = t0; k < (size_t)n; k += step) Jk[(size_t)n * w + k] = J[(size_t)n * w + k];
}

template <int NT>
__global__ void __launch_bounds__(NT)
k_lm_solve(const double *__restrict__ Hk, const double *__restrict__ Jk, const double *__restrict__ u_w, const int32_t *__restrict__ act,
           int n, double *__restrict__ dx, double *__restrict__ q1) {
  extern __shared__ double lm_sm[];
  double *A = lm_sm;                 // [n][n] column-major, lower triangle used
  double *tmp = A + (size_t)n * n;   // [n]
  double *x = tmp + n;               // [n]
  double *D = x + n;                 // [n]
  double *Jg = D + n;                // [n]
  __shared__ int s_tr[LM_MAXN];
  __shared__ double s_d[LM_MAXN];
  const int w = blockIdx.x, t = threadIdx.x, lane = t & 31, wp = t >> 5;
  if (!act[w]) return;
  const double u = u_w[w];
#define LM_AT(r, c) A[(size_t)(c) * n + (r)]
  // gauge: pose 0 fixed (voxel_map.hpp:452-455)
  for (int idx = t; idx < n * n; idx += NT) {
    const int c = idx / n, r = idx - c * n;
    double v = Hk[(size_t)n * n * w + idx];
    if (r < 6 || c < 6) v = r == c ? 1.0 : 0.0;
    A[idx] = v;
  }
  __syncthreads();
  if (t < n) {
    D[t] = LM_AT(t, t);
    Jg[t] = t < 6 ? 0.0 : Jk[(size_t)n * w + t];
  }
  __syncthreads();
  if (t < n) {
    LM_AT(t, t) = __dadd_rn(LM_AT(t, t), __dmul_rn(u, D[t]));
    x[t] = -Jg[t];
  }
  __syncthreads();
  // Pivot sequence up front.  The sweep is left-looking: when step k picks its pivot the diagonal entries at positions >= k
  // still hold their input values (moved only by the earlier interchanges), so the whole sequence is a function of the input
  // diagonal — one warp runs the host's selection (first index of the largest |diagonal| in (k, n) if it beats |A(k,k)|;
  // comparisons with NaN are false) on a copy of the diagonal instead of two CTA barriers inside every column step.
  if (t < n) s_d[t] = fabs(LM_AT(t, t));
  __syncthreads();
  if (wp == 0) {
    for (int k = 0; k < n; k++) {
      double bv = -1.0; int bi = n;
      for (int i = k + 1 + lane; i < n; i += 32) { const double v = s_d[i]; if (v > bv) { bv = v; bi = i; } }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
      }
      const double dk = s_d[k];
      const int big = bv > dk ? bi : k;
      __syncwarp();
      if (lane == 0) {
        s_tr[k] = big;
        if (big != k) { s_d[k] = s_d[big]; s_d[big] = dk; }
      }
      __syncwarp();
    }
  }
  __syncthreads();
  for (int k = 0; k < n; k++) {
    const int big = s_tr[k];
    if (big != k) {   // symmetric interchange restricted to the lower triangle
      if (t < k) { const double a = LM_AT(k, t); LM_AT(k, t) = LM_AT(big, t); LM_AT(big, t) = a; }
      else if (t > big && t < n) { const double a = LM_AT(t, k); LM_AT(t, k) = LM_AT(t, big); LM_AT(t, big) = a; }
      else if (t > k && t < big) { const double a = LM_AT(t, k); LM_AT(t, k) = LM_AT(big, t); LM_AT(big, t) = a; }
      else if (t == k) { const double a = LM_AT(k, k); LM_AT(k, k) = LM_AT(big, big); LM_AT(big, big) = a; }
      __syncthreads();
    }
    if (k > 0) {
      if (t < k) tmp[t] = __dmul_rn(LM_AT(t, t), LM_AT(k, t));
      __syncthreads();
      // only the running sum is a dependent chain: loads and products are independent, so the loops are unrolled and the host's
      // "skip a zero tmp[j]" is a select, not a branch (a - 0 would turn a = -0 into +0)
      if (t == k) {
        double acc = 0.0;
#pragma unroll 4
        for (int j = 0; j < k; j++) acc = __dadd_rn(acc, __dmul_rn(LM_AT(k, j), tmp[j]));
        LM_AT(k, k) = __dadd_rn(LM_AT(k, k), -acc);
      } else if (t > k && t < n) {
        double a = LM_AT(t, k);
#pragma unroll 4
        for (int j = 0; j < k; j++) {
          const double tj = tmp[j];
          const double na = __dadd_rn(a, -__dmul_rn(LM_AT(t, j), tj));
          a = tj == 0.0 ? a : na;
        }
        LM_AT(t, k) = a;
      }
      __syncthreads();
    }
    const double akk = LM_AT(k, k);
    if (fabs(akk) > 0.0 && t > k && t < n) LM_AT(t, k) = LM_AT(t, k) / akk;
    __syncthreads();
  }
  // solve: P, L, D, L^T, P^T
  if (t == 0) for (int k = 0; k < n; k++) if (s_tr[k] != k) { const double a = x[k]; x[k] = x[s_tr[k]]; x[s_tr[k]] = a; }
  __syncthreads();
  for (int j = 0; j < n; j++) {
    const double xj = x[j];
    if (xj != 0.0 && t > j && t < n) x[t] = __dadd_rn(x[t], -__dmul_rn(LM_AT(t, j), xj));
    __syncthreads();
  }
  if (t < n) {
    const double tol = 1.0 / 1.7976931348623157e308;
    x[t] = fabs(LM_AT(t, t)) > tol ? x[t] / LM_AT(t, t) : 0.0;
  }
  __syncthreads();
  if (t == 0) {
    for (int j = n - 1; j >= 0; j--) {
      double s = x[j];
#pragma unroll 4
      for (int i = j + 1; i < n; i++) s = __dadd_rn(s, -__dmul_rn(LM_AT(i, j), x[i]));
      x[j] = s;
    }
    for (int k = n - 1; k >= 0; k--) if (s_tr[k] != k) { const double a = x[k]; x[k] = x[s_tr[k]]; x[s_tr[k]] = a; }
    // q1 = 0.5 dx^T (u D dx - JacT)   (voxel_map.hpp:470)
    double q = 0.0;
    for (int r = 0; r < n; r++) q = __dadd_rn(q, __dmul_rn(x[r], __dadd_rn(__dmul_rn(__dmul_rn(u, D[r]), x[r]), -Jg[r])));
    q1[w] = 0.5 * q;
  }
  __syncthreads();
  if (t < n) dx[(size_t)n * w + t] = x[t];
#undef LM_AT
}

}  // namespace

namespace vx {
int ldlt_lower_solve_device(vx_ctx *ctx, const double *A, int n, const double *b, double *x);   // vx_ldlt.cu
int lm_solve_large_device(vx_ctx *ctx, const double *dH, const double *dJ, double u, int n, double *dx, double *q1);   // vx_ldlt.cu
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
  // device LM step: k_lm_keep + k_lm_solve for windows of up to LM_MAXN unknowns, the blocked factorisation of vx_ldlt.cu from
  // ldlt_device_min unknowns on (top-level GBA); Hess and JacT never leave the device.  In between (129...383 unknowns) and
  // under VX_LM_HOST the host sweep.
  const bool small = jl <= LM_MAXN;
  const bool dev = (small || jl >= ldlt_device_min) && f->V > 0 && f->nnz > 0 && getenv("VX_LM_HOST") == nullptr;
  const bool host_mats = !dev;
  std::vector<double> Hess(host_mats ? hsz * NW : 0), JacT(host_mats ? (size_t)jl * NW : 0), Hnew(Hess.size()), Jnew(JacT.size()), r_new(NW), r1(NW, 0.0), r2(NW, 0.0);
  std::vector<double> u(NW, 0.01), v(NW, 2.0), A(host_mats ? hsz : 0), dxi(jl), D(jl), nb(jl), tmp, q1s(NW, 0.0);
  std::vector<char> calc(NW, 1), active(NW, 1), conv(NW, 1);
  std::vector<int> iters(NW, 0), tr;
  std::vector<vx_pose> x(poses, poses + F), xt(poses, poses + F);
  if (hess_out && (!dev || max_iter == 0)) std::fill(hess_out, hess_out + hsz * NW, 0.0);   // the device path writes it whole at the end
  if (resis) std::fill(resis, resis + 2 * (size_t)NW, 0.0);
  for (int w = 0; w < NW; w++) if (few[w]) { active[w] = 0; conv[w] = 0; }

  std::vector<int32_t> mask(NW);
  std::vector<double> dx_all;
  cudaStream_t st = ctx->stream;
  if (dev) {
    VX_TRY(f->lm_H.reserve(sizeof(double) * hsz * NW)); VX_TRY(f->lm_J.reserve(sizeof(double) * jl * NW));
    VX_TRY(f->lm_dx.reserve(sizeof(double) * jl * NW)); VX_TRY(f->lm_q1.reserve(sizeof(double) * NW));
    VX_TRY(f->lm_u.reserve(sizeof(double) * NW)); VX_TRY(f->lm_mask.reserve(sizeof(int32_t) * NW));
    dx_all.resize((size_t)jl * NW);
    const size_t smem = sizeof(double) * ((size_t)jl * jl + 4 * jl);
    if (!small) {}
    else if (jl <= 64) VX_CUDA(cudaFuncSetAttribute(k_lm_solve<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    else VX_CUDA(cudaFuncSetAttribute(k_lm_solve<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  bool evaluated = false;

  for (int it = 0; it < max_iter; it++) {
    bool any_active = false, any_calc = false;
    for (int w = 0; w < NW; w++) { any_active |= active[w] != 0; any_calc |= (active[w] && calc[w]); }
    if (!any_active) break;
    if (any_calc && dev) {
      VX_TRY(vx_factor_set_poses(ctx, f, x.data()));
      VX_TRY(vx_factor_hessian_async(ctx, f, 0, -1));
      for (int w = 0; w < NW; w++) mask[w] = active[w] && calc[w];
      VX_CUDA(cudaMemcpyAsync(f->lm_mask.p, mask.data(), sizeof(int32_t) * NW, cudaMemcpyHostToDevice, st));
      VX_KERNEL(ctx, "k_lm_keep", k_lm_keep<<<dim3(NW, (unsigned)std::max<size_t>(1, std::min<size_t>(hsz / 4096, 1024))), 256, 0, st>>>(f->H.as<double>(), f->J.as<double>(), f->lm_H.as<double>(), f->lm_J.as<double>(),
                                                              f->lm_mask.as<int32_t>(), jl));
      VX_CUDA(cudaMemcpyAsync(r_new.data(), f->cost_h.p, sizeof(double) * NW, cudaMemcpyDeviceToHost, st));
      VX_CUDA(cudaStreamSynchronize(st));
      for (int w = 0; w < NW; w++) if (mask[w]) r1[w] = r_new[w];
      evaluated = true;
    } else if (any_calc) {
      VX_TRY(vx_factor_hessian(ctx, f, x.data(), 0, -1, Hnew.data(), Jnew.data(), r_new.data()));
      for (int w = 0; w < NW; w++)
        if (active[w] && calc[w]) {
          std::copy(Hnew.begin() + hsz * w, Hnew.begin() + hsz * (w + 1), Hess.begin() + hsz * w);
          std::copy(Jnew.begin() + (size_t)jl * w, Jnew.begin() + (size_t)jl * (w + 1), JacT.begin() + (size_t)jl * w);
          r1[w] = r_new[w];
          if (hess_out) std::copy(Hnew.begin() + hsz * w, Hnew.begin() + hsz * (w + 1), hess_out + hsz * w);   // *hess = Hess, before the gauge fix (:446)
        }
    }
    if (dev && !small) {
      for (int w = 0; w < NW; w++)
        if (active[w])
          VX_TRY(lm_solve_large_device(ctx, f->lm_H.as<double>() + hsz * w, f->lm_J.as<double>() + (size_t)jl * w, u[w], jl,
                                       dx_all.data() + (size_t)jl * w, &q1s[w]));
    } else if (dev) {
      for (int w = 0; w < NW; w++) mask[w] = active[w];
      VX_CUDA(cudaMemcpyAsync(f->lm_mask.p, mask.data(), sizeof(int32_t) * NW, cudaMemcpyHostToDevice, st));
      VX_CUDA(cudaMemcpyAsync(f->lm_u.p, u.data(), sizeof(double) * NW, cudaMemcpyHostToDevice, st));
      const size_t smem = sizeof(double) * ((size_t)jl * jl + 4 * jl);
      if (jl <= 64) {
        VX_KERNEL(ctx, "k_lm_solve", k_lm_solve<64><<<NW, 64, smem, st>>>(f->lm_H.as<double>(), f->lm_J.as<double>(), f->lm_u.as<double>(),
                                                                        f->lm_mask.as<int32_t>(), jl, f->lm_dx.as<double>(), f->lm_q1.as<double>()));
      } else {
        VX_KERNEL(ctx, "k_lm_solve", k_lm_solve<128><<<NW, 128, smem, st>>>(f->lm_H.as<double>(), f->lm_J.as<double>(), f->lm_u.as<double>(),
                                                                          f->lm_mask.as<int32_t>(), jl, f->lm_dx.as<double>(), f->lm_q1.as<double>()));
      }
      VX_CUDA(cudaMemcpyAsync(dx_all.data(), f->lm_dx.p, sizeof(double) * jl * NW, cudaMemcpyDeviceToHost, st));
      VX_CUDA(cudaMemcpyAsync(q1s.data(), f->lm_q1.p, sizeof(double) * NW, cudaMemcpyDeviceToHost, st));
      VX_CUDA(cudaStreamSynchronize(st));
    }
    for (int w = 0; w < NW; w++) {
      if (!active[w]) continue;
      if (it == 0 && resis) resis[2 * w] = r1[w];
      const double *dxw = dxi.data();
      if (dev) {
        dxw = dx_all.data() + (size_t)jl * w;
      } else {
        double *H = Hess.data() + hsz * w, *J = JacT.data() + (size_t)jl * w;
        // gauge: pose 0 fixed (voxel_map.hpp:452-455)
        for (int c = 0; c < jl; c++) for (int r = 0; r < 6; r++) H[(size_t)c * jl + r] = 0.0;
        for (int c = 0; c < 6; c++) for (int r = 0; r < jl; r++) H[(size_t)c * jl + r] = 0.0;
        for (int r = 0; r < 6; r++) { H[(size_t)r * jl + r] = 1.0; J[r] = 0.0; }
        for (int r = 0; r < jl; r++) D[r] = H[(size_t)r * jl + r];
        std::copy(H, H + hsz, A.begin());
        for (int r = 0; r < jl; r++) { A[(size_t)r * jl + r] += u[w] * D[r]; nb[r] = -J[r]; }
        // the host sweep, or for the top-level GBA window (W ~ 399, 2394 unknowns, 4.6 GFLOP) the blocked device
        // factorisation (vx_ldlt.cu)
        if (jl >= ldlt_device_min) VX_TRY(ldlt_lower_solve_device(ctx, A.data(), jl, nb.data(), dxi.data()));
        else hm::ldlt_lower_solve(A.data(), jl, nb.data(), dxi.data(), tr, tmp);
        double q1 = 0.0;
        for (int r = 0; r < jl; r++) q1 += dxi[r] * (u[w] * D[r] * dxi[r] - J[r]);
        q1s[w] = 0.5 * q1;
      }
      for (int j = 0; j < W; j++) {
        const vx_pose &s = x[(size_t)w * W + j];
        vx_pose &d = xt[(size_t)w * W + j];
        retract_rot(s.R, &dxw[6 * j], d.R);
        for (int k = 0; k < 3; k++) d.p[k] = s.p[k] + dxw[6 * j + 3 + k];
      }
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
  if (dev && hess_out && evaluated) {   // *hess = Hess of the last accepted evaluation, before the gauge fix (:446)
    VX_CUDA(cudaMemcpyAsync(hess_out, f->lm_H.p, sizeof(double) * hsz * NW, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    for (int w = 0; w < NW; w++) if (few[w]) std::fill(hess_out + hsz * w, hess_out + hsz * (w + 1), 0.0);
  } else if (dev && hess_out && max_iter > 0) {
    std::fill(hess_out, hess_out + hsz * NW, 0.0);
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
