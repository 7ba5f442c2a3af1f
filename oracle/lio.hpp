// ORACLE — TEST INFRASTRUCTURE ONLY (see vmath.hpp header). PARITY UNPINNED.
//
// lio.hpp — restatement of the LIO ESIKF point-to-plane update and the per-point covariance
// front-end that feeds it:
//   /root/reference/VoxelSLAM/src/voxelslam.hpp:180-200   calcBodyVar
//   /root/reference/VoxelSLAM/src/voxelslam.hpp:210-234   var_init
//   /root/reference/VoxelSLAM/src/voxelslam.hpp:242-265   pvec_update
//   /root/reference/VoxelSLAM/src/voxelslam.cpp:962-1098  VOXEL_SLAM::lio_state_estimation
#pragma once
#include "octree.hpp"

namespace orc {

// DEG2RAD comes from PCL 1.10 pcl/pcl_macros.h: #define DEG2RAD(x) ((x)*0.017453293) (third-party,
// absent from /root/reference; README.md:26 pins PCL 1.10 in prose).
inline double deg2rad(double d) { return d * 0.017453293; }

// voxelslam.hpp:180-200 (note: mutates pb[2] when it is exactly zero; float range / range_var)
inline void calcBodyVar(V3 &pb, const float range_inc, const float degree_inc, M3 &var) {
  if (pb[2] == 0) pb[2] = 0.0001;
  float range = (float)std::sqrt(pb[0] * pb[0] + pb[1] * pb[1] + pb[2] * pb[2]);
  float range_var = range_inc * range_inc;
  double dv = std::pow(std::sin(deg2rad(degree_inc)), 2);
  Mat<2, 2> direction_var;
  direction_var(0, 0) = dv; direction_var(1, 1) = dv;
  V3 direction = pb;
  direction /= direction.norm();
  M3 direction_hat = hat(direction);
  V3 base_vector1 = vec3(1, 1, -(direction[0] + direction[1]) / direction[2]);
  base_vector1 /= base_vector1.norm();
  V3 base_vector2 = cross(base_vector1, direction);
  base_vector2 /= base_vector2.norm();
  Mat<3, 2> N;
  for (int r = 0; r < 3; r++) { N(r, 0) = base_vector1[r]; N(r, 1) = base_vector2[r]; }
  Mat<3, 2> A = (double)range * (direction_hat * N);
  var = direction * (double)range_var * direction.transpose() + A * direction_var * A.transpose();
}

// voxelslam.hpp:210-234
inline void var_init(const IMUST &ext, const std::vector<PointXYZIN> &pl_cur, PVec &pvec, double dept_err, double beam_err) {
  int plsize = (int)pl_cur.size();
  pvec.clear();
  pvec.resize(plsize);
  for (int i = 0; i < plsize; i++) {
    const PointXYZIN &ap = pl_cur[i];
    pointVar &pv = pvec[i];
    pv.pnt = vec3(ap.x, ap.y, ap.z);
    calcBodyVar(pv.pnt, (float)dept_err, (float)beam_err, pv.var);
    pv.pnt = ext.R * pv.pnt + ext.p;
    pv.var = ext.R * pv.var * ext.R.transpose();
  }
}

// voxelslam.hpp:242-265
inline void pvec_update(PVec &pvec, const IMUST &x_curr, std::vector<V3> &pwld) {
  M3 rot_var = x_curr.cov.block<3, 3>(0, 0);
  M3 tsl_var = x_curr.cov.block<3, 3>(3, 3);
  for (pointVar &pv : pvec) {
    M3 phat = hat(pv.pnt);
    pv.var = x_curr.R * pv.var * x_curr.R.transpose() + phat * rot_var * phat.transpose() + tsl_var;
    pwld.push_back(x_curr.R * pv.pnt + x_curr.p);
  }
}

struct LioPassOut {   // what one pass over the points produces (voxelslam.cpp:1005-1050)
  M6 HTH; V6 HTz; M3 nnt; int match_num = 0;
};
struct LioTrace {     // test introspection
  std::vector<LioPassOut> passes;
  std::vector<std::vector<int>> flags;   // per pass, per point association flag
  int iters = 0;
};

// One pass of voxelslam.cpp:1005-1050 with the leaf cache `octos` (in/out).
inline void lio_pass(SurfMap &surf_map, const PVec &pvec, const IMUST &x_curr, std::vector<OctoTree *> &octos,
                     MapCtx *ctx, LioPassOut &o, std::vector<int> *flags_out = nullptr) {
  o.HTH.setZero(); o.HTz.setZero(); o.nnt.setZero(); o.match_num = 0;
  M3 rot_var = x_curr.cov.block<3, 3>(0, 0);
  M3 tsl_var = x_curr.cov.block<3, 3>(3, 3);
  int psize = (int)pvec.size();
  if (flags_out) flags_out->assign(psize, 0);
  for (int i = 0; i < psize; i++) {
    const pointVar &pv = pvec[i];
    M3 phat = hat(pv.pnt);
    M3 var_world = x_curr.R * pv.var * x_curr.R.transpose() + phat * rot_var * phat.transpose() + tsl_var;
    V3 wld = x_curr.R * pv.pnt + x_curr.p;
    double sigma_d = 0;
    Plane *pla = nullptr;
    int flag = 0;
    if (octos[i] != nullptr && octos[i]->inside(wld)) {
      double max_prob = 0;
      flag = octos[i]->match(wld, pla, max_prob, var_world, sigma_d, octos[i]);
    } else {
      flag = match(surf_map, wld, pla, var_world, sigma_d, octos[i], ctx);
    }
    if (flag) {
      Plane &pp = *pla;
      double R_inv = 1.0 / (0.0005 + sigma_d);
      double resi = pp.normal.dot(wld - pp.center);
      V6 jac;
      jac.setBlock(0, 0, phat * x_curr.R.transpose() * pp.normal);
      jac.setBlock(3, 0, pp.normal);
      o.HTH += R_inv * (jac * jac.transpose());
      o.HTz -= R_inv * jac * resi;
      o.nnt += pp.normal * pp.normal.transpose();
      o.match_num++;
      if (flags_out) (*flags_out)[i] = 1;
    }
  }
}

// voxelslam.cpp:962-1098.  Returns the degeneracy verdict (true = healthy).
inline bool lio_state_estimation(SurfMap &surf_map, const PVec &pvec, IMUST &x_curr, MapCtx *ctx, LioTrace *trace = nullptr) {
  IMUST x_prop = x_curr;
  const int num_max_iter = 4;
  bool EKF_stop_flg = 0, flg_EKF_converged = 0;
  M15 G, H_T_H, I_STATE;
  I_STATE.setIdentity();
  int rematch_num = 0;
  int psize = (int)pvec.size();
  std::vector<OctoTree *> octos(psize, nullptr);
  M3 nnt;
  M15 cov_inv = inverse_lu(x_curr.cov);

  for (int iterCount = 0; iterCount < num_max_iter; iterCount++) {
    LioPassOut o;
    std::vector<int> flags;
    lio_pass(surf_map, pvec, x_curr, octos, ctx, o, trace ? &flags : nullptr);
    nnt = o.nnt;
    if (trace) { trace->passes.push_back(o); trace->flags.push_back(flags); trace->iters++; }

    H_T_H.setBlock(0, 0, o.HTH);
    M15 K_1 = inverse_lu(H_T_H + cov_inv);
    Mat<15, 6> K6 = K_1.block<15, 6>(0, 0);
    G.setBlock(0, 0, K6 * o.HTH);
    V15 vec = x_prop - x_curr;
    V15 solution = K6 * o.HTz + vec - G.block<15, 6>(0, 0) * vec.block<6, 1>(0, 0);

    x_curr += solution;
    V3 rot_add = solution.block<3, 1>(0, 0);
    V3 tra_add = solution.block<3, 1>(3, 0);

    EKF_stop_flg = false;
    flg_EKF_converged = false;
    if ((rot_add.norm() * 57.3 < 0.01) && (tra_add.norm() * 100 < 0.015)) flg_EKF_converged = true;
    if (flg_EKF_converged || ((rematch_num == 0) && (iterCount == num_max_iter - 2))) rematch_num++;
    if (rematch_num >= 2 || (iterCount == num_max_iter - 1)) {
      x_curr.cov = (I_STATE - G) * x_curr.cov;
      EKF_stop_flg = true;
    }
    if (EKF_stop_flg) break;
  }
  V3 evalue; M3 evec;
  eig3_sym(nnt, evalue, evec);
  return !(evalue[0] < 14);
}

}  // namespace orc
