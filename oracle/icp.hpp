// ORACLE — TEST INFRASTRUCTURE ONLY (see vmath.hpp header). PARITY UNPINNED.
//
// icp.hpp — restatement of icp_normal (/root/reference/VoxelSLAM/src/loop_refine.hpp:47-145): the point-to-plane ICP with
// normal-compatibility gates that verifies a loop-closure candidate.
//
// Third-party piece: the reference finds the nearest target point with pcl::KdTreeFLANN<pcl::PointXYZ> (PCL, version not
// pinned in the reference tree; FLANN's KDTreeSingleIndex with L2_Simple<float>, exact search, eps = 0).  Its published
// contract is the exact 1-nearest neighbour under the FLOAT squared distance  ((dx*dx) + dy*dy) + dz*dz  accumulated in
// x, y, z order on float coordinates (L2_Simple::operator(): `result += diff * diff` per dimension).  Which of several
// points at exactly the same float distance is returned is an implementation detail of the tree; this restatement (and the
// device kernel) return the lowest target index among them.  Here the search is brute force.
#pragma once
#include "vmath.hpp"
#include "linalg.hpp"
#include <vector>

namespace orc {

struct IcpTrace {
  int iters = 0;
  std::vector<int> match_num;     // per iteration
  std::vector<double> resi;       // per iteration
  double eig[3] = {0, 0, 0};      // eigenvalues of sum tni tni^T of the last iteration
  int is_converge = 0;
};

// src / tar: n x 3 float points and n x 3 float normals (PointXYZINormal's x,y,z and normal_x,y,z).
// pose = (p, R) as in the reference's pair<Vector3d, Matrix3d>; updated in place.  Returns the reference's bool.
inline bool icp_normal(const float *src_xyz, const float *src_nrm, int ns, const float *tar_xyz, const float *tar_nrm, int nt, V3 &p, M3 &R,
                       double icp_eigval, IcpTrace *trace) {
  double paras[4] = {0.2, 0.2, 0.5, 3};                       // loop_refine.hpp:63
  int is_converge = 0;
  M3 mat_norm;
  IcpTrace tr;
  for (int iterCount = 0; iterCount < 20; iterCount++) {      // :70
    Mat<6, 6> Hess; Mat<6, 1> JacT;
    double resi = 0;
    int match_num = 0;
    mat_norm.setZero();
    for (int i = 0; i < ns; i++) {
      V3 plocal; plocal[0] = src_xyz[3 * i]; plocal[1] = src_xyz[3 * i + 1]; plocal[2] = src_xyz[3 * i + 2];
      V3 pi = R * plocal + p;                                  // :82
      const float qx = (float)pi[0], qy = (float)pi[1], qz = (float)pi[2];   // use_search_point (:84-86)
      V3 ni; ni[0] = src_nrm[3 * i]; ni[1] = src_nrm[3 * i + 1]; ni[2] = src_nrm[3 * i + 2];
      ni = R * ni;                                             // :88
      if (nt <= 0) continue;                                   // nearestKSearch(...) > 0 (:89)
      int best = 0; float bd = 0;
      for (int j = 0; j < nt; j++) {
        const float dx = qx - tar_xyz[3 * j], dy = qy - tar_xyz[3 * j + 1], dz = qz - tar_xyz[3 * j + 2];
        const float d2 = (dx * dx + dy * dy) + dz * dz;         // built with -ffp-contract=off
        if (j == 0 || d2 < bd) { bd = d2; best = j; }
      }
      V3 tpi, tni;
      for (int k = 0; k < 3; k++) { tpi[k] = tar_xyz[3 * best + k]; tni[k] = tar_nrm[3 * best + k]; }
      const V3 normal_inc = ni - tni, normal_add = ni + tni;
      const V3 dlt = pi - tpi;
      const double point_to_point_dis = dlt.norm();
      const double point_to_plane = std::fabs(tni.dot(dlt));
      if ((normal_inc.norm() < paras[0] || normal_add.norm() < paras[1]) && point_to_plane < paras[2] && point_to_point_dis < paras[3]) {   // :98
        const double rr = tni.dot(dlt);
        const V3 jh = hat(plocal) * (R.transpose() * tni);     // :103
        Mat<6, 1> jac;
        for (int k = 0; k < 3; k++) { jac[k] = jh[k]; jac[3 + k] = tni[k]; }
        for (int c = 0; c < 6; c++) for (int r = 0; r < 6; r++) Hess(r, c) += jac[r] * jac[c];
        for (int r = 0; r < 6; r++) JacT[r] += jac[r] * rr;
        resi += 0.5 * rr * rr;
        match_num++;
        for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) mat_norm(r, c) += tni[r] * tni[c];
      }
    }
    MatX A(6, 6); VecX b(6), dxi(6);
    for (int c = 0; c < 6; c++) for (int r = 0; r < 6; r++) A(r, c) = Hess(r, c);
    for (int r = 0; r < 6; r++) b[r] = -JacT[r];
    ldlt_solve(A, b, dxi);                                     // :116
    V3 w, t;
    for (int k = 0; k < 3; k++) { w[k] = dxi[k]; t[k] = dxi[3 + k]; }
    R = R * Exp(w);                                            // :117
    p = p + t;
    tr.iters = iterCount + 1; tr.match_num.push_back(match_num); tr.resi.push_back(resi);
    if (w.norm() < 1e-3 && t.norm() < 1e-3) {                  // :122
      if (is_converge) break;
      paras[0] = 0.1; paras[1] = 0.1; paras[2] = 0.1; paras[3] = 1;
      is_converge = 1;
    }
  }
  V3 ev; M3 U;
  eig3_sym(mat_norm, ev, U);                                   // :139
  for (int k = 0; k < 3; k++) tr.eig[k] = ev[k];
  tr.is_converge = is_converge;
  if (trace) *trace = tr;
  return ev[0] > icp_eigval && is_converge == 1;               // :142
}

}  // namespace orc
