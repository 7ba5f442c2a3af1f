// ORACLE — TEST INFRASTRUCTURE ONLY (see vmath.hpp header). PARITY UNPINNED.
//
// types.hpp — value types of the hot path, restated from
//   /root/reference/VoxelSLAM/src/tools.hpp:24-49   VOXEL_LOC + hash
//   /root/reference/VoxelSLAM/src/tools.hpp:135-199 IMUST
//   /root/reference/VoxelSLAM/src/tools.hpp:304-365 PointCluster
//   /root/reference/VoxelSLAM/src/voxel_map.hpp:18-37 pointVar / PVec
//   /root/reference/VoxelSLAM/src/voxel_map.hpp:83-104 Plane + the global tunables
#pragma once
#include "linalg.hpp"
#include <unordered_map>
#include <memory>

namespace orc {

constexpr int DIM = 15;   // tools.hpp:16
constexpr int DVEL = 6;   // voxel_map.hpp:502

// ---- voxel key (A1) -----------------------------------------------------------------
// voxel_map.hpp:1907-1918 (and 6 more inlined copies): double/double -> float, float "-1"
// for negatives, truncation toward zero.
inline int64_t key_axis_d(double coord, double voxel_size) {
  float loc = (float)(coord / voxel_size);
  if (loc < 0) loc -= 1.0f;
  return (int64_t)loc;
}
// tools.hpp:211 — the coordinate is a float promoted to double before the divide.
inline int64_t key_axis_f(float coord, double voxel_size) {
  float loc = (float)((double)coord / voxel_size);
  if (loc < 0) loc -= 1.0;   // "-= 1.0" is a double literal: float->double, subtract, ->float
  return (int64_t)loc;
}

struct VOXEL_LOC {
  int64_t x, y, z;
  VOXEL_LOC(int64_t vx = 0, int64_t vy = 0, int64_t vz = 0) : x(vx), y(vy), z(vz) {}
  bool operator==(const VOXEL_LOC &o) const { return x == o.x && y == o.y && z == o.z; }
};
struct VoxelHash {  // tools.hpp:37-49 with HASH_P=116101, MAX_N=1e10 (tools.hpp:9-10)
  size_t operator()(const VOXEL_LOC &s) const {
    const size_t HASH_P = 116101;
    const size_t MAX_N = 10000000000ULL;
    return ((((size_t)std::hash<int64_t>()(s.z) * HASH_P) % MAX_N + std::hash<int64_t>()(s.y)) * HASH_P) % MAX_N + std::hash<int64_t>()(s.x);
  }
};
template <class T> using VoxelMap = std::unordered_map<VOXEL_LOC, T, VoxelHash>;

inline VOXEL_LOC voxel_of(const V3 &p, double voxel_size) {
  return VOXEL_LOC(key_axis_d(p[0], voxel_size), key_axis_d(p[1], voxel_size), key_axis_d(p[2], voxel_size));
}

// ---- IMUST --------------------------------------------------------------------------
struct IMUST {
  double t;
  M3 R; V3 p, v, bg, ba, g;
  M15 cov;
  IMUST() { setZero(); }
  void setZero() {  // tools.hpp:188-197 (g left untouched there; zero here as Eigen default-inits garbage)
    t = 0; R.setIdentity();
    p.setZero(); v.setZero(); bg.setZero(); ba.setZero();
    cov.setIdentity(); cov *= 0.0001;
    for (int i = 9; i < 15; i++) for (int j = 9; j < 15; j++) cov(i, j) = (i == j) ? 0.00001 : 0.0;
  }
  IMUST &operator+=(const V15 &ist) {  // tools.hpp:154-162
    R = R * Exp(ist.block<3, 1>(0, 0));
    p += ist.block<3, 1>(3, 0);
    v += ist.block<3, 1>(6, 0);
    bg += ist.block<3, 1>(9, 0);
    ba += ist.block<3, 1>(12, 0);
    return *this;
  }
  V15 operator-(const IMUST &b) const {  // tools.hpp:164-173
    V15 a;
    a.setBlock(0, 0, Log(b.R.transpose() * R));
    a.setBlock(3, 0, p - b.p);
    a.setBlock(6, 0, v - b.v);
    a.setBlock(9, 0, bg - b.bg);
    a.setBlock(12, 0, ba - b.ba);
    return a;
  }
};

// ---- PointCluster -------------------------------------------------------------------
struct PointCluster {
  M3 P; V3 v; int N = 0;
  void clear() { P.setZero(); v.setZero(); N = 0; }
  void push(const V3 &vec) { N++; P += vec * vec.transpose(); v += vec; }   // tools.hpp:326-331
  M3 cov() const { V3 c = v / (double)N; return P / (double)N - c * c.transpose(); }  // tools.hpp:333-337
  PointCluster &operator+=(const PointCluster &s) { P += s.P; v += s.v; N += s.N; return *this; }
  PointCluster &operator-=(const PointCluster &s) { P -= s.P; v -= s.v; N -= s.N; return *this; }
  void transform(const PointCluster &s, const IMUST &st) {  // tools.hpp:357-363
    N = s.N;
    v = st.R * s.v + (double)N * st.p;
    M3 rp = st.R * s.v * st.p.transpose();
    P = st.R * s.P * st.R.transpose() + rp + rp.transpose() + (double)N * (st.p * st.p.transpose());
  }
};

struct pointVar { V3 pnt; M3 var; };       // voxel_map.hpp:18-34
using PVec = std::vector<pointVar>;
using PVecPtr = std::shared_ptr<PVec>;

struct Plane {                              // voxel_map.hpp:83-96
  V3 center, normal;
  M6 plane_var;
  float radius = 0;
  bool is_plane = false;
};

// The reference's global tunables (voxel_map.hpp:98-104, :500; loop_refine.hpp:269-271)
// gathered into one struct that every oracle entry point receives explicitly.
struct Params {
  double voxel_size = 1.0;
  double min_eigen_value = 0.0025;
  int max_layer = 2;
  int max_points = 100;
  double min_point[4] = {5, 5, 5, 5};                      // voxelslam.cpp:917
  double plane_eigen_value_thre[4] = {1, 1, 1, 1};         // stored inverted, voxelslam.cpp:930-931
  double imu_coef = 1e-4;
  double gba_voxel_size = 1.0;
  double gba_min_eigen_value = 0.01;
  double gba_eigen_value_array[4] = {0.5, 0.5, 0.5, 0.5};  // stored inverted, voxelslam.cpp:3023-3024
};

}  // namespace orc
