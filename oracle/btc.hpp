// ORACLE — TEST INFRASTRUCTURE ONLY (see vmath.hpp header). PARITY UNPINNED.
//
// btc.hpp — restatement of the plane-voxel extraction of the BTC loop detector:
//   /root/reference/VoxelSLAM/src/BTC.cpp:279-320   STDescManager::init_voxel_map (its own key rule: double quotient, truncated)
//   /root/reference/VoxelSLAM/src/BTC.cpp:90-143    BTCOctoTree::init_octo_tree / init_plane
//   /root/reference/VoxelSLAM/src/BTC.cpp:322-334   STDescManager::get_plane
// Third-party piece: init_plane calls Eigen::EigenSolver<Matrix3d> (Eigen 3.3.7, real Schur) on a matrix that is symmetric by
// construction; its real eigenpairs are those of the self-adjoint solver restated in linalg.hpp, to rounding, and the SIGN of
// an eigenvector is unspecified in both.  The output order of the reference is the iteration order of its unordered_map; here
// voxels are visited in ascending (x, y, z) key order.
#pragma once
#include "vmath.hpp"
#include "linalg.hpp"
#include <map>
#include <array>
#include <vector>

namespace orc {

struct BtcPlane { double center[3], normal[3], min_eigen_value, radius, d, points_size; };

inline std::vector<BtcPlane> btc_plane_cloud(const float *xyz, int64_t n, double voxel_size, double plane_thre, int init_num) {
  std::map<std::array<int64_t, 3>, std::vector<int64_t>> vox;
  for (int64_t i = 0; i < n; i++) {
    std::array<int64_t, 3> key;
    for (int j = 0; j < 3; j++) {
      double loc = (double)xyz[3 * i + j] / voxel_size;          // :289
      if (loc < 0) loc -= 1.0;
      key[j] = (int64_t)loc;
    }
    vox[key].push_back(i);
  }
  std::vector<BtcPlane> out;
  for (auto &kv : vox) {
    const std::vector<int64_t> &ids = kv.second;
    if (!((int)ids.size() > init_num)) continue;                  // :91
    M3 cov; V3 c;
    for (int64_t i : ids) {
      V3 pi = vec3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
      for (int cc = 0; cc < 3; cc++) for (int r = 0; r < 3; r++) cov(r, cc) += pi[r] * pi[cc];   // :103
      c += pi;
    }
    const double N = (double)ids.size();
    c = c / N;
    for (int cc = 0; cc < 3; cc++) for (int r = 0; r < 3; r++) cov(r, cc) = cov(r, cc) / N - c[r] * c[cc];   // :107-109
    V3 w; M3 U;
    eig3_sym(cov, w, U);
    if (w[0] < plane_thre) {                                      // :120
      BtcPlane p;
      for (int k = 0; k < 3; k++) { p.center[k] = c[k]; p.normal[k] = U(k, 0); }
      p.min_eigen_value = w[0]; p.radius = std::sqrt(w[2]);
      p.d = -(p.normal[0] * c[0] + p.normal[1] * c[1] + p.normal[2] * c[2]);
      p.points_size = N;
      out.push_back(p);
    }
  }
  return out;
}

}  // namespace orc
