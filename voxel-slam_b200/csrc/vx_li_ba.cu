// vx_li_ba.cu — the LiDAR-inertial LM optimizers over the device LidarFactor:
//   B6  IMU_PRE                        preintegration.hpp:11-303   (host, a POD the caller owns)
//   B4  LI_BA_Optimizer::damping_iter  voxel_map.hpp:624-713      (15W unknowns)
//   B5  LI_BA_OptimizerGravity::damping_iter  voxel_map.hpp:878-975 (15W+3 unknowns)
// Everything between two LM decisions runs on the device: acc_evaluate2 / evaluate_only_residual, the W-1 IMU
// chain factors (give_evaluate / give_evaluate_g, one thread each: 15x15 inverse + 15x33 Jacobian), the assembly (IMU blocks in
// chain order, * imu_coef, LiDAR blocks scattered stride 6 -> 15: the reference's own order of additions), the gauge fix and the
// blocked (15W[+3])^2 LDL^T of vx_ldlt.cu.  Per LM iteration the states go up and dx, q1 and W costs come back; the host keeps the
// LM control flow, the retraction and the IMU_PRE bias bookkeeping.  VX_LI_HOST=1 selects the round-1 host assembly + host sweep.
#include "vx_factor.cuh"
#include "vx_hostmat.hpp"
#include <algorithm>
#include <cstdlib>

using namespace vx;
using namespace vx::hm;

namespace {

constexpr int DIM = 15;

struct Pre {   // unpacked vx_imu_pre
  M3 R_delta, R_bg, p_bg, p_ba, v_bg, v_ba;
  V3 p_delta, v_delta, bg, ba, dbg, dba;
  double dtime;
  M15 cov;
};
VX_HD Pre unpack(const vx_imu_pre *p) {
  Pre q;
  q.R_delta = M3::from(p->R_delta); q.R_bg = M3::from(p->R_bg);
  q.p_bg = M3::from(p->p_bg); q.p_ba = M3::from(p->p_ba); q.v_bg = M3::from(p->v_bg); q.v_ba = M3::from(p->v_ba);
  q.p_delta = V3::from(p->p_delta); q.v_delta = V3::from(p->v_delta);
  q.bg = V3::from(p->bg); q.ba = V3::from(p->ba); q.dbg = V3::from(p->dbg); q.dba = V3::from(p->dba);
  q.dtime = p->dtime;
  q.cov = M15::from(p->cov);
  return q;
}
struct State { M3 R; V3 p, v, bg, ba, g; };
VX_HD State unpack(const vx_imu_state *s) {
  State x;
  x.R = M3::from(s->R); x.p = V3::from(s->p); x.v = V3::from(s->v); x.bg = V3::from(s->bg); x.ba = V3::from(s->ba); x.g = V3::from(s->g);
  return x;
}

// preintegration.hpp:75-135
void add_imu(vx_imu_pre *out, const vx_imu_noise *nz, const V3 &gyr, const V3 &acc, double dt) {
  Pre q = unpack(out);
  q.dtime += dt;
  const M3 R_inc = so3_exp(gyr, dt);
  const M3 R_jr = so3_jr(gyr * dt);
  const M3 R_dt = dt * q.R_delta;
  const M3 R_h = (0.5 * dt * dt) * q.R_delta;
  const M3 ax = hat(acc);

  q.p_ba = q.p_ba + q.v_ba * dt - R_h;
  q.p_bg = q.p_bg + q.v_bg * dt - R_h * ax * q.R_bg;
  q.v_ba = q.v_ba - R_dt;
  q.v_bg = q.v_bg - R_dt * ax * q.R_bg;
  q.R_bg = R_inc.T() * q.R_bg - R_jr * dt;

  M<9, 9> A = M<9, 9>::eye();
  M<9, 6> B;
  A.put(0, 0, R_inc.T());
  A.put(3, 0, -(R_h * ax));
  A.put(3, 6, M3::eye() * dt);
  A.put(6, 0, -(R_dt * ax));
  B.put(0, 0, R_jr * dt);
  B.put(3, 3, R_h);
  B.put(6, 3, R_dt);
  const M<6, 6> nm = M<6, 6>::from(nz->noise_meas), nw = M<6, 6>::from(nz->noise_walk);
  const M<9, 9> c9 = A * q.cov.blk<9, 9>(0, 0) * A.T() + B * nm * B.T();
  q.cov.put(0, 0, c9);
  q.cov.put(9, 9, q.cov.blk<6, 6>(9, 9) + nw * dt);

  q.p_delta += q.v_delta * dt + R_h * acc;
  q.v_delta += R_dt * acc;
  q.R_delta = q.R_delta * R_inc;

  q.R_delta.to(out->R_delta); q.R_bg.to(out->R_bg); q.p_bg.to(out->p_bg); q.p_ba.to(out->p_ba); q.v_bg.to(out->v_bg); q.v_ba.to(out->v_ba);
  q.p_delta.to(out->p_delta); q.v_delta.to(out->v_delta);
  out->dtime = q.dtime;
  q.cov.to(out->cov);
}

// give_evaluate (preintegration.hpp:137-212) for NC = 30, give_evaluate_g (:214-294) for NC = 33: residual rr, information
// cov^-1 and (when J != nullptr) the 15 x NC Jacobian.  Shared by the host evaluation and the device kernel.
template <int NC>
VX_HD void imu_parts(const vx_imu_pre *pre, const State &s1, const State &s2, V15 &rr, M15 &info, M<DIM, NC> *Jp) {
  const Pre q = unpack(pre);
  const M3 R_corr = q.R_delta * so3_exp(q.R_bg * q.dbg);
  const V3 t_corr = q.p_delta + q.p_bg * q.dbg + q.p_ba * q.dba;
  const V3 v_corr = q.v_delta + q.v_bg * q.dbg + q.v_ba * q.dba;
  const M3 R1t = s1.R.T();
  const double dt = q.dtime;

  const M3 res_r = R_corr.T() * R1t * s2.R;
  const V3 exp_v = R1t * (s2.v - s1.v - dt * s1.g);
  const V3 exp_t = R1t * (s2.p - s1.p - s1.v * dt - (0.5 * dt * dt) * s1.g);
  rr.put(0, 0, so3_log(res_r));
  rr.put(3, 0, exp_t - t_corr);
  rr.put(6, 0, exp_v - v_corr);
  rr.put(9, 0, s2.bg - s1.bg);
  rr.put(12, 0, s2.ba - s1.ba);
  info = inverse(q.cov);   // a dense inverse per call, as the reference (:166 / :244)

  if (Jp) {
    M<DIM, NC> &J = *Jp;
    const M3 I = M3::eye();
    const M3 Jri = so3_jr_inv(res_r);
    // d/d(state 1)
    J.put(0, 0, -(Jri * s2.R.T() * s1.R));
    J.put(0, 9, -(Jri * res_r.T() * so3_jr(q.R_bg * q.dbg) * q.R_bg));
    J.put(3, 0, hat(exp_t));
    J.put(3, 3, -R1t);
    J.put(3, 6, -(R1t * dt));
    J.put(3, 9, -q.p_bg);
    J.put(3, 12, -q.p_ba);
    J.put(6, 0, hat(exp_v));
    J.put(6, 6, -R1t);
    J.put(6, 9, -q.v_bg);
    J.put(6, 12, -q.v_ba);
    J.put(9, 9, -I);
    J.put(12, 12, -I);
    // d/d(state 2)
    J.put(0, DIM + 0, Jri);
    J.put(3, DIM + 3, R1t);
    J.put(6, DIM + 6, R1t);
    J.put(9, DIM + 9, I);
    J.put(12, DIM + 12, I);
    if (NC == 2 * DIM + 3) {   // d/d(gravity), :277-283
      J.put(3, 2 * DIM, R1t * (-0.5 * dt * dt));
      J.put(6, 2 * DIM, R1t * (-dt));
    }
  }
}
template <int NC>
double imu_evaluate(const vx_imu_pre *pre, const State &s1, const State &s2, M<NC, NC> *jtj, M<NC, 1> *gg) {
  V15 rr; M15 info;
  if (jtj && gg) {
    M<DIM, NC> J;
    imu_parts<NC>(pre, s1, s2, rr, info, &J);
    const M<NC, DIM> JtI = J.T() * info;
    *jtj = JtI * J;
    *gg = JtI * rr;
  } else {
    imu_parts<NC>(pre, s1, s2, rr, info, (M<DIM, NC> *)nullptr);
  }
  return rr.dot(info * rr);
}

// R <- R Exp(w) on the POD
void retract(const vx_imu_state &s, const double *d, vx_imu_state &o) {
  const M3 R = M3::from(s.R) * so3_exp(v3(d[0], d[1], d[2]));
  R.to(o.R);
  for (int k = 0; k < 3; k++) {
    o.p[k] = s.p[k] + d[3 + k]; o.v[k] = s.v[k] + d[6 + k]; o.bg[k] = s.bg[k] + d[9 + k]; o.ba[k] = s.ba[k] + d[12 + k];
  }
}

// ---- device path -----------------------------------------------------------------------------------------------------------------
// One WARP per IMU factor i (states i, i+1): lane 0 forms the residual, cov^-1 and the 15 x NC Jacobian (3 x 3 algebra and one
// 15 x 15 inverse) in shared memory; the warp then takes J^T cov^-1 (NC x 15), J^T cov^-1 J (NC x NC) and J^T cov^-1 r entry by
// entry, each a sum over k in ascending order from 0.0 — the host's sequence of products and sums.  jtj column-major, cost always.
template <int NC>
__global__ void __launch_bounds__(32) k_li_imu(const vx_imu_pre *__restrict__ imus, const vx_imu_state *__restrict__ xs, int nf, int jac,
                                              double *__restrict__ jtj, double *__restrict__ gg, double *__restrict__ cost) {
  __shared__ double sJ[DIM * NC], sI[DIM * DIM], sT[NC * DIM], sr[DIM], sIr[DIM];     // J [k][c] col-major (k + 15 c), info col-major, JtI (r + NC k)
  const int i = blockIdx.x, lane = threadIdx.x;
  if (i >= nf) return;
  if (lane == 0) {
    const State s1 = unpack(&xs[i]), s2 = unpack(&xs[i + 1]);
    V15 rr; M15 info;
    if (jac) {
      M<DIM, NC> J;
      imu_parts<NC>(&imus[i], s1, s2, rr, info, &J);
      for (int e = 0; e < DIM * NC; e++) sJ[e] = J[e];
    } else {
      imu_parts<NC>(&imus[i], s1, s2, rr, info, (M<DIM, NC> *)nullptr);
    }
    for (int e = 0; e < DIM * DIM; e++) sI[e] = info[e];
    for (int e = 0; e < DIM; e++) sr[e] = rr[e];
  }
  __syncwarp();
  if (lane < DIM) {                          // info * rr, then rr . (info rr)
    double s = 0;
    for (int k = 0; k < DIM; k++) s += sI[lane + DIM * k] * sr[k];
    sIr[lane] = s;
  }
  __syncwarp();
  if (lane == 0) {
    double s = 0;
    for (int k = 0; k < DIM; k++) s += sr[k] * sIr[k];
    cost[i] = s;
  }
  if (!jac) return;
  for (int e = lane; e < NC * DIM; e += 32) {           // JtI(r, c) = sum_k J(k, r) info(k, c)
    const int r = e % NC, c = e / NC;
    double s = 0;
    for (int k = 0; k < DIM; k++) s += sJ[k + DIM * r] * sI[k + DIM * c];
    sT[r + NC * c] = s;
  }
  __syncwarp();
  double *out = jtj + (size_t)i * NC * NC;
  for (int e = lane; e < NC * NC; e += 32) {            // jtj(r, c) = sum_k JtI(r, k) J(k, c)
    const int r = e % NC, c = e / NC;
    double s = 0;
    for (int k = 0; k < DIM; k++) s += sT[r + NC * k] * sJ[k + DIM * c];
    out[e] = s;
  }
  for (int r = lane; r < NC; r += 32) {                 // gg(r) = sum_k JtI(r, k) rr(k)
    double s = 0;
    for (int k = 0; k < DIM; k++) s += sT[r + NC * k] * sr[k];
    gg[(size_t)i * NC + r] = s;
  }
}
// Hess / JacT of divide_thread (voxel_map.hpp:519-584 / :746-825) + hess_plus (:509-517 / :726-738), one thread per entry: the IMU
// contributions of an entry are added in chain order from 0.0, scaled by imu_coef, then the LiDAR block entry is added — the
// sequence of roundings of the reference's loops.  NC = 30 (no gravity) or 33 (gravity: last 3 unknowns).
template <int NC>
__global__ void k_li_assemble(const double *__restrict__ jtj, const double *__restrict__ gg, int W, double coef, const double *__restrict__ Hl,
                              const double *__restrict__ Jl, int n, double *__restrict__ H, double *__restrict__ J) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x, c = blockIdx.y;
  if (r >= n) return;
  const int nf = W - 1, nx = DIM * W, jl = 6 * W;
  auto factor_index = [&](int i, int a) -> int {   // index of unknown a inside factor i's NC-vector, or -1
    if (a >= nx) return NC == 33 ? 30 + (a - nx) : -1;
    const int b = a / DIM;
    if (b == i) return a % DIM;
    if (b == i + 1) return DIM + a % DIM;
    return -1;
  };
  int lo = 0, hi = nf - 1;                          // factors that can touch (r, c)
  if (r < nx) { lo = max(lo, r / DIM - 1); hi = min(hi, r / DIM); }
  if (c < nx) { lo = max(lo, c / DIM - 1); hi = min(hi, c / DIM); }
  double s = 0.0;
  for (int i = lo; i <= hi; i++) {
    const int a = factor_index(i, r), b = factor_index(i, c);
    if (a >= 0 && b >= 0) s = __dadd_rn(s, jtj[(size_t)i * NC * NC + (size_t)b * NC + a]);
  }
  s = __dmul_rn(s, coef);
  if (r < nx && c < nx && r % DIM < 6 && c % DIM < 6) s = __dadd_rn(s, Hl[(size_t)((c / DIM) * 6 + c % DIM) * jl + (r / DIM) * 6 + r % DIM]);
  H[(size_t)c * n + r] = s;
  if (c == 0) {
    double g = 0.0;
    lo = 0; hi = nf - 1;
    if (r < nx) { lo = max(lo, r / DIM - 1); hi = min(hi, r / DIM); }
    for (int i = lo; i <= hi; i++) { const int a = factor_index(i, r); if (a >= 0) g = __dadd_rn(g, gg[(size_t)i * NC + a]); }
    g = __dmul_rn(g, coef);
    if (r < nx && r % DIM < 6) g = __dadd_rn(g, Jl[(r / DIM) * 6 + r % DIM]);
    J[r] = g;
  }
}
}  // namespace
namespace vx {
int lm_solve_large_device(vx_ctx *ctx, const double *dH, const double *dJ, double u, int n, double *dx, double *q1, int nfix);   // vx_ldlt.cu
}
namespace {

// Common body of the two damping_iter()s (they differ in unknown count, factor width, gauge, iteration cap, g update).
int li_damping(vx_ctx *ctx, vx_factor *f, vx_imu_state *xs, vx_imu_pre *imus, double imu_coef, bool gravity, int max_iter, double *resis2,
               double *hess_out, int *accept_trace, int *n_iters) {
  VX_ARG(ctx && f && xs && max_iter >= 0);
  VX_ARG(f->n_win == 1);
  const int W = f->W;
  VX_ARG(W == 1 || imus != nullptr);
  VX_TRY(bind(ctx));
  const int jl = 6 * W, n = DIM * W + (gravity ? 3 : 0), nfix = gravity ? 6 : DIM;
  const size_t nn = (size_t)n * n;
  std::vector<double> Hess(nn), JacT(n), A(nn), dxi(n), D(n), nb(n), tmp, hl((size_t)jl * jl), jlv(jl);
  std::vector<int> tr;
  std::vector<vx_imu_state> x(xs, xs + W), xt(xs, xs + W);
  std::vector<vx_pose> lp(W);
  auto lidar_poses = [&](const std::vector<vx_imu_state> &s) {
    for (int j = 0; j < W; j++) { std::copy(s[j].R, s[j].R + 9, lp[j].R); std::copy(s[j].p, s[j].p + 3, lp[j].p); }
  };
#define H_(r, c) Hess[(size_t)(c) * n + (r)]
  double u = 0.01, v = 2, r1 = 0, r2 = 0;
  bool calc = true;
  int iters = 0;
  if (resis2) resis2[0] = resis2[1] = 0;
  // device path: see the file header
  const bool devp = std::getenv("VX_LI_HOST") == nullptr && W >= 2 && f->V > 0 && f->nnz > 0;
  const int nf = W - 1, NCf = gravity ? 33 : 30;
  cudaStream_t st = ctx->stream;
  DevBuf d_imus, d_x, d_jtj, d_gg, d_cost, d_H, d_J;
  struct Rel { std::vector<DevBuf *> b; ~Rel() { for (DevBuf *x : b) x->release(); } } rel{{&d_imus, &d_x, &d_jtj, &d_gg, &d_cost, &d_H, &d_J}};
  std::vector<double> costs(std::max(nf, 1));
  if (devp) {
    VX_TRY(d_imus.reserve(sizeof(vx_imu_pre) * nf)); VX_TRY(d_x.reserve(sizeof(vx_imu_state) * W));
    VX_TRY(d_jtj.reserve(sizeof(double) * NCf * NCf * nf)); VX_TRY(d_gg.reserve(sizeof(double) * NCf * nf)); VX_TRY(d_cost.reserve(sizeof(double) * nf));
    VX_TRY(d_H.reserve(sizeof(double) * nn)); VX_TRY(d_J.reserve(sizeof(double) * n));
  }
  // IMU chain at the states s (device): costs[] <- per-factor rr^T cov^-1 rr; with jac also jtj / gg
  auto imu_device = [&](const std::vector<vx_imu_state> &s, bool jac) -> int {
    VX_CUDA(cudaMemcpyAsync(d_x.p, s.data(), sizeof(vx_imu_state) * W, cudaMemcpyHostToDevice, st));
    VX_CUDA(cudaMemcpyAsync(d_imus.p, imus, sizeof(vx_imu_pre) * nf, cudaMemcpyHostToDevice, st));
    if (gravity) VX_KERNEL(ctx, "k_li_imu", k_li_imu<33><<<nf, 32, 0, st>>>(d_imus.as<vx_imu_pre>(), d_x.as<vx_imu_state>(), nf, jac ? 1 : 0, d_jtj.as<double>(),
                                                                                     d_gg.as<double>(), d_cost.as<double>()));
    else VX_KERNEL(ctx, "k_li_imu", k_li_imu<30><<<nf, 32, 0, st>>>(d_imus.as<vx_imu_pre>(), d_x.as<vx_imu_state>(), nf, jac ? 1 : 0, d_jtj.as<double>(),
                                                                              d_gg.as<double>(), d_cost.as<double>()));
    VX_CUDA(cudaMemcpyAsync(costs.data(), d_cost.p, sizeof(double) * nf, cudaMemcpyDeviceToHost, st));
    return VX_OK;
  };

  for (int it = 0; it < max_iter; it++) {
    if (calc && devp) {
      VX_TRY(imu_device(x, true));
      lidar_poses(x);
      VX_TRY(vx_factor_set_poses(ctx, f, lp.data()));
      VX_TRY(vx_factor_hessian_async(ctx, f, 0, -1));
      if (gravity) VX_KERNEL(ctx, "k_li_assemble", k_li_assemble<33><<<dim3(blocks_for(n, 128), n), 128, 0, st>>>(d_jtj.as<double>(), d_gg.as<double>(), W, imu_coef, f->H.as<double>(),
                                                                                                      f->J.as<double>(), n, d_H.as<double>(), d_J.as<double>()));
      else VX_KERNEL(ctx, "k_li_assemble", k_li_assemble<30><<<dim3(blocks_for(n, 128), n), 128, 0, st>>>(d_jtj.as<double>(), d_gg.as<double>(), W, imu_coef, f->H.as<double>(),
                                                                                               f->J.as<double>(), n, d_H.as<double>(), d_J.as<double>()));
      double lc = 0;
      VX_CUDA(cudaMemcpyAsync(&lc, f->cost_h.p, sizeof(double), cudaMemcpyDeviceToHost, st));
      if (hess_out) VX_CUDA(cudaMemcpyAsync(hess_out, d_H.p, sizeof(double) * nn, cudaMemcpyDeviceToHost, st));   // *hess = Hess, before the gauge fix
      VX_CUDA(cudaStreamSynchronize(st));
      double cost = 0;
      for (int i = 0; i < nf; i++) cost += costs[i];
      cost *= imu_coef * 0.5;
      r1 = cost + lc;
    } else if (calc) {
      // divide_thread (voxel_map.hpp:519-584 / :746-825): IMU chain first, scaled by imu_coef, then the LiDAR blocks on top
      std::fill(Hess.begin(), Hess.end(), 0.0);
      std::fill(JacT.begin(), JacT.end(), 0.0);
      double cost = 0;
      for (int i = 0; i + 1 < W; i++) {
        const State s1 = unpack(&x[i]), s2 = unpack(&x[i + 1]);
        if (!gravity) {
          M<30, 30> jtj; M<30, 1> gg;
          cost += imu_evaluate<30>(&imus[i], s1, s2, &jtj, &gg);
          for (int c = 0; c < 30; c++) for (int r = 0; r < 30; r++) H_(i * DIM + r, i * DIM + c) += jtj(r, c);
          for (int r = 0; r < 30; r++) JacT[i * DIM + r] += gg[r];
        } else {
          M<33, 33> jtj; M<33, 1> gg;
          cost += imu_evaluate<33>(&imus[i], s1, s2, &jtj, &gg);
          const int g0 = n - 3;
          for (int c = 0; c < 30; c++) for (int r = 0; r < 30; r++) H_(i * DIM + r, i * DIM + c) += jtj(r, c);
          for (int c = 0; c < 3; c++) for (int r = 0; r < 30; r++) { H_(i * DIM + r, g0 + c) += jtj(r, 30 + c); H_(g0 + c, i * DIM + r) += jtj(30 + c, r); }
          for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) H_(g0 + r, g0 + c) += jtj(30 + r, 30 + c);
          for (int r = 0; r < 30; r++) JacT[i * DIM + r] += gg[r];
          for (int r = 0; r < 3; r++) JacT[g0 + r] += gg[30 + r];
        }
      }
      for (auto &h : Hess) h *= imu_coef;
      for (auto &j : JacT) j *= imu_coef;
      cost *= imu_coef * 0.5;
      double lc = 0;
      lidar_poses(x);
      VX_TRY(vx_factor_hessian(ctx, f, lp.data(), 0, -1, hl.data(), jlv.data(), &lc));
      for (int i = 0; i < W; i++) {   // hess_plus (voxel_map.hpp:509-517 / :726-738)
        for (int r = 0; r < 6; r++) JacT[i * DIM + r] += jlv[i * 6 + r];
        for (int j = 0; j < W; j++)
          for (int c = 0; c < 6; c++) for (int r = 0; r < 6; r++) H_(i * DIM + r, j * DIM + c) += hl[(size_t)(j * 6 + c) * jl + i * 6 + r];
      }
      r1 = cost + lc;
      if (hess_out) std::copy(Hess.begin(), Hess.end(), hess_out);   // *hess = Hess, before the gauge fix
    }
    if (it == 0 && resis2) resis2[0] = r1;

    double q1 = 0;
    if (devp) {
      VX_TRY(vx::lm_solve_large_device(ctx, d_H.as<double>(), d_J.as<double>(), u, n, dxi.data(), &q1, nfix));
    } else {
      for (int c = 0; c < n; c++) for (int r = 0; r < nfix; r++) H_(r, c) = 0.0;
      for (int c = 0; c < nfix; c++) for (int r = 0; r < n; r++) H_(r, c) = 0.0;
      for (int r = 0; r < nfix; r++) { H_(r, r) = 1.0; JacT[r] = 0.0; }
      for (int r = 0; r < n; r++) D[r] = H_(r, r);
      A = Hess;
      for (int r = 0; r < n; r++) { A[(size_t)r * n + r] += u * D[r]; nb[r] = -JacT[r]; }
      ldlt_lower_solve(A.data(), n, nb.data(), dxi.data(), tr, tmp);
      for (int r = 0; r < n; r++) q1 += dxi[r] * (u * D[r] * dxi[r] - JacT[r]);
      q1 *= 0.5;
    }

    for (int j = 0; j < W; j++) retract(x[j], &dxi[DIM * j], xt[j]);
    if (gravity) {   // voxel_map.hpp:921-930: the trial g accumulates on x_stats_temp[0] even across rejected steps
      for (int k = 0; k < 3; k++) xt[0].g[k] += dxi[n - 3 + k];
      for (int j = 1; j < W; j++) std::copy(xt[0].g, xt[0].g + 3, xt[j].g);
    }
    for (int j = 0; j + 1 < W; j++) {   // IMU_PRE::update_state (:296-303)
      for (int k = 0; k < 3; k++) {
        imus[j].dbg_buf[k] = imus[j].dbg[k]; imus[j].dba_buf[k] = imus[j].dba[k];
        imus[j].dbg[k] += dxi[DIM * j + 9 + k]; imus[j].dba[k] += dxi[DIM * j + 12 + k];
      }
    }

    // only_residual (voxel_map.hpp:586-622 / :831-870) at the trial states
    double ci = 0;
    if (devp) {
      VX_TRY(imu_device(xt, false));
      VX_CUDA(cudaStreamSynchronize(st));
      for (int i = 0; i < nf; i++) ci += costs[i];
    } else {
      for (int i = 0; i + 1 < W; i++) {
        const State s1 = unpack(&xt[i]), s2 = unpack(&xt[i + 1]);
        ci += gravity ? imu_evaluate<33>(&imus[i], s1, s2, nullptr, nullptr) : imu_evaluate<30>(&imus[i], s1, s2, nullptr, nullptr);
      }
    }
    ci *= imu_coef * 0.5;
    double lc = 0;
    lidar_poses(xt);
    VX_TRY(vx_factor_residual(ctx, f, lp.data(), 0, -1, &lc));
    r2 = ci + lc;
    double q = r1 - r2;
    iters++;
    if (q > 0) {
      x = xt;
      q = q / q1;
      v = 2;
      q = 1 - std::pow(2 * q - 1, 3);
      u *= (q < 1.0 / 3 ? 1.0 / 3 : q);
      calc = true;
      if (accept_trace) accept_trace[it] = 1;
    } else {
      u = u * v;
      v = 2 * v;
      calc = false;
      for (int j = 0; j + 1 < W; j++)
        for (int k = 0; k < 3; k++) { imus[j].dbg[k] = imus[j].dbg_buf[k]; imus[j].dba[k] = imus[j].dba_buf[k]; }
      if (accept_trace) accept_trace[it] = 0;
    }
    if (std::fabs((r1 - r2) / r1) < 1e-6) break;
  }
#undef H_
  if (resis2) resis2[1] = r2;
  if (n_iters) *n_iters = iters;
  std::copy(x.begin(), x.end(), xs);
  return VX_OK;
}

}  // namespace

extern "C" {

int vx_imu_pre_init(vx_imu_pre *pre, const double *bg3, const double *ba3) {
  VX_ARG(pre != nullptr);
  std::memset(pre, 0, sizeof(*pre));
  pre->R_delta[0] = pre->R_delta[4] = pre->R_delta[8] = 1.0;
  for (int k = 0; k < 3; k++) { pre->bg[k] = bg3 ? bg3[k] : 0.0; pre->ba[k] = ba3 ? ba3[k] : 0.0; }
  return VX_OK;
}

int vx_imu_pre_add(vx_imu_pre *pre, const vx_imu_noise *nz, const double *gyr3, const double *acc3, double dt) {
  VX_ARG(pre && nz && gyr3 && acc3);
  add_imu(pre, nz, V3::from(gyr3), V3::from(acc3), dt);
  return VX_OK;
}

int vx_imu_pre_push(vx_imu_pre *pre, const vx_imu_noise *nz, const vx_imu_sample *s, int n) {
  VX_ARG(pre && nz && (n <= 0 || s));
  const V3 bg = V3::from(pre->bg), ba = V3::from(pre->ba);
  for (int k = 1; k < n; k++) {   // mid-point rule, preintegration.hpp:52-71
    const double dt = s[k].t - s[k - 1].t;
    const V3 gyr = 0.5 * (V3::from(s[k - 1].gyr) + V3::from(s[k].gyr)) - bg;
    const V3 acc = (0.5 * (V3::from(s[k - 1].acc) + V3::from(s[k].acc))) * nz->scale_gravity - ba;
    add_imu(pre, nz, gyr, acc, dt);
  }
  return VX_OK;
}

int vx_imu_pre_evaluate(const vx_imu_pre *pre, const vx_imu_state *st1, const vx_imu_state *st2, int with_gravity, double *jtj, double *gg,
                        double *cost) {
  VX_ARG(pre && st1 && st2 && cost);
  const State s1 = unpack(st1), s2 = unpack(st2);
  const bool jac = jtj && gg;
  if (with_gravity) {
    M<33, 33> h; M<33, 1> g;
    *cost = imu_evaluate<33>(pre, s1, s2, jac ? &h : nullptr, jac ? &g : nullptr);
    if (jac) { h.to(jtj); g.to(gg); }
  } else {
    M<30, 30> h; M<30, 1> g;
    *cost = imu_evaluate<30>(pre, s1, s2, jac ? &h : nullptr, jac ? &g : nullptr);
    if (jac) { h.to(jtj); g.to(gg); }
  }
  return VX_OK;
}

// give_evaluate / give_evaluate_g of n_factors chained factors on the device (factor i between states i and i+1), the kernel the
// LI optimizers use (k_li_imu): jtj n_factors x (nc x nc) column-major, gg n_factors x nc, cost n_factors; nc = 30 or 33.
int vx_imu_pre_evaluate_device(vx_ctx *ctx, const vx_imu_pre *pres, const vx_imu_state *states, int n_factors, int with_gravity, double *jtj,
                               double *gg, double *cost) {
  VX_ARG(ctx && pres && states && cost && n_factors >= 1);
  VX_TRY(bind(ctx));
  cudaStream_t st = ctx->stream;
  const int nc = with_gravity ? 33 : 30, jac = (jtj && gg) ? 1 : 0;
  DevBuf d_p, d_s, d_h, d_g, d_c;
  auto body = [&]() -> int {
    VX_TRY(d_p.reserve(sizeof(vx_imu_pre) * n_factors)); VX_TRY(d_s.reserve(sizeof(vx_imu_state) * (n_factors + 1)));
    VX_TRY(d_h.reserve(sizeof(double) * nc * nc * n_factors)); VX_TRY(d_g.reserve(sizeof(double) * nc * n_factors)); VX_TRY(d_c.reserve(sizeof(double) * n_factors));
    VX_CUDA(cudaMemcpyAsync(d_p.p, pres, sizeof(vx_imu_pre) * n_factors, cudaMemcpyHostToDevice, st));
    VX_CUDA(cudaMemcpyAsync(d_s.p, states, sizeof(vx_imu_state) * (n_factors + 1), cudaMemcpyHostToDevice, st));
    if (with_gravity) VX_KERNEL(ctx, "k_li_imu", k_li_imu<33><<<n_factors, 32, 0, st>>>(d_p.as<vx_imu_pre>(), d_s.as<vx_imu_state>(), n_factors, jac, d_h.as<double>(),
                                                                                                 d_g.as<double>(), d_c.as<double>()));
    else VX_KERNEL(ctx, "k_li_imu", k_li_imu<30><<<n_factors, 32, 0, st>>>(d_p.as<vx_imu_pre>(), d_s.as<vx_imu_state>(), n_factors, jac, d_h.as<double>(),
                                                                                          d_g.as<double>(), d_c.as<double>()));
    if (jac) {
      VX_CUDA(cudaMemcpyAsync(jtj, d_h.p, sizeof(double) * nc * nc * n_factors, cudaMemcpyDeviceToHost, st));
      VX_CUDA(cudaMemcpyAsync(gg, d_g.p, sizeof(double) * nc * n_factors, cudaMemcpyDeviceToHost, st));
    }
    VX_CUDA(cudaMemcpyAsync(cost, d_c.p, sizeof(double) * n_factors, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    return VX_OK;
  };
  const int rc = body();
  cudaStreamSynchronize(st);
  for (DevBuf *b : {&d_p, &d_s, &d_h, &d_g, &d_c}) b->release();
  return rc;
}

int vx_imu_pre_update_state(vx_imu_pre *pre, const double *dxi15) {
  VX_ARG(pre && dxi15);
  for (int k = 0; k < 3; k++) {
    pre->dbg_buf[k] = pre->dbg[k]; pre->dba_buf[k] = pre->dba[k];
    pre->dbg[k] += dxi15[9 + k]; pre->dba[k] += dxi15[12 + k];
  }
  return VX_OK;
}

int vx_ba_li(vx_ctx *ctx, vx_factor *f, vx_imu_state *x, vx_imu_pre *imus, double imu_coef, double *hess_out, int *accept_trace, int *n_iters) {
  return li_damping(ctx, f, x, imus, imu_coef, false, 3, nullptr, hess_out, accept_trace, n_iters);
}

int vx_ba_li_gravity(vx_ctx *ctx, vx_factor *f, vx_imu_state *x, vx_imu_pre *imus, double imu_coef, double *resis2, double *hess_out,
                     int max_iter, int *accept_trace, int *n_iters) {
  return li_damping(ctx, f, x, imus, imu_coef, true, max_iter, resis2, hess_out, accept_trace, n_iters);
}

}  // extern "C"
