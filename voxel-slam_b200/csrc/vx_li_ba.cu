// vx_li_ba.cu — the LiDAR-inertial LM optimizers over the device LidarFactor:
//   B6  IMU_PRE                        preintegration.hpp:11-303   (host, a POD the caller owns)
//   B4  LI_BA_Optimizer::damping_iter  voxel_map.hpp:624-713      (15W unknowns)
//   B5  LI_BA_OptimizerGravity::damping_iter  voxel_map.hpp:878-975 (15W+3 unknowns)
// The LiDAR part (acc_evaluate2 / evaluate_only_residual) runs on the device; the W-1 IMU chain factors (15x15
// inverse + 15x33 Jacobian each), the scatter stride 6 -> 15, the gauge fix and the (15W[+3])^2 LDL^T stay on the
// host: W <= 20 means <= 303 unknowns, microseconds of work that sits between two device passes either way.
#include "vx_factor.cuh"
#include "vx_hostmat.hpp"
#include <algorithm>

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
Pre unpack(const vx_imu_pre *p) {
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
State unpack(const vx_imu_state *s) {
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

// give_evaluate (preintegration.hpp:137-212) for NC = 30, give_evaluate_g (:214-294) for NC = 33.
template <int NC>
double imu_evaluate(const vx_imu_pre *pre, const State &s1, const State &s2, M<NC, NC> *jtj, M<NC, 1> *gg) {
  const Pre q = unpack(pre);
  const M3 R_corr = q.R_delta * so3_exp(q.R_bg * q.dbg);
  const V3 t_corr = q.p_delta + q.p_bg * q.dbg + q.p_ba * q.dba;
  const V3 v_corr = q.v_delta + q.v_bg * q.dbg + q.v_ba * q.dba;
  const M3 R1t = s1.R.T();
  const double dt = q.dtime;

  const M3 res_r = R_corr.T() * R1t * s2.R;
  const V3 exp_v = R1t * (s2.v - s1.v - dt * s1.g);
  const V3 exp_t = R1t * (s2.p - s1.p - s1.v * dt - (0.5 * dt * dt) * s1.g);
  V15 rr;
  rr.put(0, 0, so3_log(res_r));
  rr.put(3, 0, exp_t - t_corr);
  rr.put(6, 0, exp_v - v_corr);
  rr.put(9, 0, s2.bg - s1.bg);
  rr.put(12, 0, s2.ba - s1.ba);
  const M15 info = inverse(q.cov);   // a dense inverse per call, as the reference (:166 / :244)

  if (jtj && gg) {
    const M3 I = M3::eye();
    const M3 Jri = so3_jr_inv(res_r);
    M<DIM, NC> J;
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
    const M<NC, DIM> JtI = J.T() * info;
    *jtj = JtI * J;
    *gg = JtI * rr;
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

  for (int it = 0; it < max_iter; it++) {
    if (calc) {
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

    for (int c = 0; c < n; c++) for (int r = 0; r < nfix; r++) H_(r, c) = 0.0;
    for (int c = 0; c < nfix; c++) for (int r = 0; r < n; r++) H_(r, c) = 0.0;
    for (int r = 0; r < nfix; r++) { H_(r, r) = 1.0; JacT[r] = 0.0; }
    for (int r = 0; r < n; r++) D[r] = H_(r, r);
    A = Hess;
    for (int r = 0; r < n; r++) { A[(size_t)r * n + r] += u * D[r]; nb[r] = -JacT[r]; }
    ldlt_lower_solve(A.data(), n, nb.data(), dxi.data(), tr, tmp);

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
    double q1 = 0;
    for (int r = 0; r < n; r++) q1 += dxi[r] * (u * D[r] * dxi[r] - JacT[r]);
    q1 *= 0.5;

    // only_residual (voxel_map.hpp:586-622 / :831-870) at the trial states
    double ci = 0;
    for (int i = 0; i + 1 < W; i++) {
      const State s1 = unpack(&xt[i]), s2 = unpack(&xt[i + 1]);
      ci += gravity ? imu_evaluate<33>(&imus[i], s1, s2, nullptr, nullptr) : imu_evaluate<30>(&imus[i], s1, s2, nullptr, nullptr);
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
