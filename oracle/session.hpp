// ORACLE — TEST INFRASTRUCTURE ONLY (see vmath.hpp header). PARITY UNPINNED.
//
// session.hpp — restatement of the stored-session formats and the keyframe assembly of the reference:
//   /root/reference/VoxelSLAM/src/voxelslam.cpp:166-179   save_pcd        <i>.pcd, binary, fields x y z intensity
//   /root/reference/VoxelSLAM/src/voxelslam.cpp:181-204   save_pose       alidarState.txt
//   /root/reference/VoxelSLAM/src/voxelslam.hpp:268-306   read_lidarstate
//   /root/reference/VoxelSLAM/src/voxelslam.cpp:207-279   pgo_edges_io    edge.txt
//   /root/reference/VoxelSLAM/src/voxelslam.cpp:345-378   keyframe assembly (win scans -> frame of the last, voxel_size / 10)
// Written with iostreams exactly as the reference does (the product side uses snprintf / strtod: two independent codings).
// Third-party pieces restated from their published algorithms: Eigen 3.3.7 Quaterniond(Matrix3d) and Quaterniond::matrix();
// the PCD v0.7 header PCL's savePCDFileBinary writes for pcl::PointXYZI.
#pragma once
#include "types.hpp"
#include "octree.hpp"
#include <cstring>
#include <fstream>
#include <sstream>
#include <iomanip>
#include <string>

namespace orc {

struct ScanState { double t, p[3], R[9], v[3], bg[3], ba[3], g[3], v6[6]; };   // IMUST fields + ScanPose::v6, R column-major

// Eigen::Quaterniond(w, x, y, z).matrix()  (QuaternionBase::toRotationMatrix)
inline void quat_to_R(double w, double x, double y, double z, double *R) {
  const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  const double twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y, tyz = tz * y, tzz = tz * z;
  R[0] = 1 - (tyy + tzz); R[3] = txy - twz;       R[6] = txz + twy;
  R[1] = txy + twz;       R[4] = 1 - (txx + tzz); R[7] = tyz - twx;
  R[2] = txz - twy;       R[5] = tyz + twx;       R[8] = 1 - (txx + tyy);
}
// Eigen::Quaterniond(Matrix3d)  (quaternionbase_assign_impl<Other,3,3>); q = x y z w
inline void R_to_quat(const double *R, double *q) {
  auto m = [&](int r, int c) { return R[c * 3 + r]; };
  double t = m(0, 0) + m(1, 1) + m(2, 2);
  if (t > 0) {
    t = std::sqrt(t + 1.0);
    q[3] = 0.5 * t;
    t = 0.5 / t;
    q[0] = (m(2, 1) - m(1, 2)) * t; q[1] = (m(0, 2) - m(2, 0)) * t; q[2] = (m(1, 0) - m(0, 1)) * t;
  } else {
    int i = 0;
    if (m(1, 1) > m(0, 0)) i = 1;
    if (m(2, 2) > m(i, i)) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    t = std::sqrt(m(i, i) - m(j, j) - m(k, k) + 1.0);
    q[i] = 0.5 * t;
    t = 0.5 / t;
    q[3] = (m(k, j) - m(j, k)) * t;
    q[j] = (m(j, i) + m(i, j)) * t;
    q[k] = (m(k, i) + m(i, k)) * t;
  }
}

inline bool write_lidarstate(const std::string &path, const std::vector<ScanState> &xs) {      // voxelslam.cpp:187-203
  std::ofstream posfile(path);
  if (!posfile.is_open()) return false;
  for (const ScanState &xx : xs) {
    double q[4];
    R_to_quat(xx.R, q);
    posfile << std::fixed << std::setprecision(6) << xx.t << " ";
    posfile << std::setprecision(7) << xx.p[0] << " " << xx.p[1] << " " << xx.p[2] << " ";
    posfile << q[0] << " " << q[1] << " " << q[2] << " " << q[3];
    posfile << " " << xx.v[0] << " " << xx.v[1] << " " << xx.v[2];
    posfile << " " << xx.bg[0] << " " << xx.bg[1] << " " << xx.bg[2];
    posfile << " " << xx.ba[0] << " " << xx.ba[1] << " " << xx.ba[2];
    posfile << " " << xx.g[0] << " " << xx.g[1] << " " << xx.g[2];
    for (int j = 0; j < 6; j++) posfile << " " << xx.v6[j];
    posfile << std::endl;
  }
  return true;
}

inline bool read_lidarstate(const std::string &path, std::vector<ScanState> &out) {            // voxelslam.hpp:268-306
  std::ifstream file(path);
  if (!file.is_open()) return false;
  std::string lineStr, str;
  std::vector<double> nums;
  while (std::getline(file, lineStr)) {
    nums.clear();
    std::stringstream ss(lineStr);
    while (std::getline(ss, str, ' ')) nums.push_back(std::stod(str));
    ScanState xx;
    std::memset(&xx, 0, sizeof(xx));
    xx.R[0] = xx.R[4] = xx.R[8] = 1.0;
    xx.g[2] = -9.8;                                            // IMUST::setZero (tools.hpp:188-197)
    for (int j = 0; j < 6; j++) xx.v6[j] = 0.0;
    xx.t = nums[0];
    xx.p[0] = nums[1]; xx.p[1] = nums[2]; xx.p[2] = nums[3];
    quat_to_R(nums[7], nums[4], nums[5], nums[6], xx.R);
    if (nums.size() >= 20)
      for (int k = 0; k < 3; k++) { xx.v[k] = nums[8 + k]; xx.bg[k] = nums[11 + k]; xx.ba[k] = nums[14 + k]; xx.g[k] = nums[17 + k]; }
    if (nums.size() >= 26)
      for (int i = 0; i < 6; i++) xx.v6[i] = nums[i + 20];
    out.push_back(xx);
  }
  return true;
}

// pcl::io::savePCDFileBinary of a PointCloud<PointXYZI>: PCD v0.7 header, then x y z intensity packed (16 B per point)
inline bool write_pcd_xyzi(const std::string &path, const float *xyzi, int64_t n) {
  std::ofstream f(path, std::ios::binary);
  if (!f.is_open()) return false;
  f << "# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS x y z intensity\nSIZE 4 4 4 4\nTYPE F F F F\nCOUNT 1 1 1 1\n"
    << "WIDTH " << n << "\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS " << n << "\nDATA binary\n";
  f.write(reinterpret_cast<const char *>(xyzi), (std::streamsize)(16 * n));
  return true;
}

struct EdgeLine { std::string n1, n2; int id1, id2; double tra[3], q[4]; };   // q = x y z w
inline bool write_edges(const std::string &path, const std::vector<EdgeLine> &es) {             // voxelslam.cpp:258-276
  std::ofstream outfile(path);
  if (!outfile.is_open()) return false;
  for (const EdgeLine &e : es) {
    outfile << e.n1 << " " << e.n2 << " " << e.id1 << " " << e.id2 << " ";
    outfile << std::setprecision(7) << e.tra[0] << " " << e.tra[1] << " " << e.tra[2] << " ";
    outfile << e.q[0] << " " << e.q[1] << " " << e.q[2] << " " << e.q[3] << std::endl;
  }
  return true;
}

// voxelslam.cpp:345-378: scans into the frame of the last one, down_sampling_voxel(voxel_size / 10)
inline std::vector<PointXYZIN> keyframe_assemble(const std::vector<const float *> &xyz, const std::vector<int64_t> &n, const std::vector<IMUST> &xs,
                                                 double voxel_size) {
  const int win = (int)xs.size();
  const IMUST &xc = xs[win - 1];
  std::vector<PointXYZIN> pl;
  for (int j = 0; j < win; j++) {
    V3 dp = xc.R.transpose() * (xs[j].p - xc.p);
    M3 dR = xc.R.transpose() * xs[j].R;
    for (int64_t k = 0; k < n[j]; k++) {
      PointXYZIN pt;
      std::memset(&pt, 0, sizeof(pt));
      V3 v3 = vec3(xyz[j][3 * k], xyz[j][3 * k + 1], xyz[j][3 * k + 2]);
      v3 = dR * v3 + dp;
      pt.x = (float)v3[0]; pt.y = (float)v3[1]; pt.z = (float)v3[2];
      pl.push_back(pt);
    }
  }
  down_sampling_voxel(pl, voxel_size / 10);
  return pl;
}

}  // namespace orc
