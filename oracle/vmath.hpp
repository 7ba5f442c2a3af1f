// ORACLE — TEST INFRASTRUCTURE ONLY. Nothing under oracle/ may be linked, imported
// or executed by the product path (voxel-slam_b200/). Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs use it, and only as the checker.
//
// PARITY UNPINNED: the reference ships no tests / golden vectors and
// cannot be built here (needs Eigen 3.3.7 + PCL + ROS + GTSAM, none installed). This is a
// dependency-free C++17 restatement, validated by self-consistency KATs in tests/.
//
// vmath.hpp — tiny fixed-size column-major matrix type (Eigen's default storage order,
// which is what the reference's `Matrix3d`, `Matrix<double,6,6>` ... use) plus the SO(3)
// helpers of /root/reference/VoxelSLAM/src/tools.hpp:51-133.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#include <algorithm>
#include <limits>

namespace orc {

template <int R, int C>
struct Mat {
  double d[R * C];  // column-major
  Mat() { for (int i = 0; i < R * C; i++) d[i] = 0.0; }
  double &operator()(int r, int c) { return d[c * R + r]; }
  const double &operator()(int r, int c) const { return d[c * R + r]; }
  double &operator[](int i) { return d[i]; }
  const double &operator[](int i) const { return d[i]; }
  void setZero() { for (int i = 0; i < R * C; i++) d[i] = 0.0; }
  void setIdentity() { setZero(); for (int i = 0; i < (R < C ? R : C); i++) (*this)(i, i) = 1.0; }
  static Mat Identity() { Mat m; m.setIdentity(); return m; }
  static Mat Zero() { return Mat(); }
  Mat<C, R> transpose() const {
    Mat<C, R> t;
    for (int r = 0; r < R; r++) for (int c = 0; c < C; c++) t(c, r) = (*this)(r, c);
    return t;
  }
  Mat &operator+=(const Mat &o) { for (int i = 0; i < R * C; i++) d[i] += o.d[i]; return *this; }
  Mat &operator-=(const Mat &o) { for (int i = 0; i < R * C; i++) d[i] -= o.d[i]; return *this; }
  Mat &operator*=(double s) { for (int i = 0; i < R * C; i++) d[i] *= s; return *this; }
  Mat &operator/=(double s) { for (int i = 0; i < R * C; i++) d[i] /= s; return *this; }
  double squaredNorm() const { double s = 0; for (int i = 0; i < R * C; i++) s += d[i] * d[i]; return s; }
  double norm() const { return std::sqrt(squaredNorm()); }
  double dot(const Mat &o) const { double s = 0; for (int i = 0; i < R * C; i++) s += d[i] * o.d[i]; return s; }
  double trace() const { double s = 0; for (int i = 0; i < (R < C ? R : C); i++) s += (*this)(i, i); return s; }
  template <int BR, int BC>
  Mat<BR, BC> block(int r0, int c0) const {
    Mat<BR, BC> b;
    for (int r = 0; r < BR; r++) for (int c = 0; c < BC; c++) b(r, c) = (*this)(r0 + r, c0 + c);
    return b;
  }
  template <int BR, int BC>
  void setBlock(int r0, int c0, const Mat<BR, BC> &b) {
    for (int r = 0; r < BR; r++) for (int c = 0; c < BC; c++) (*this)(r0 + r, c0 + c) = b(r, c);
  }
  template <int BR, int BC>
  void addBlock(int r0, int c0, const Mat<BR, BC> &b) {
    for (int r = 0; r < BR; r++) for (int c = 0; c < BC; c++) (*this)(r0 + r, c0 + c) += b(r, c);
  }
  Mat<R, 1> col(int c) const { Mat<R, 1> v; for (int r = 0; r < R; r++) v[r] = (*this)(r, c); return v; }
};

template <int R, int C> Mat<R, C> operator+(Mat<R, C> a, const Mat<R, C> &b) { a += b; return a; }
template <int R, int C> Mat<R, C> operator-(Mat<R, C> a, const Mat<R, C> &b) { a -= b; return a; }
template <int R, int C> Mat<R, C> operator-(Mat<R, C> a) { for (int i = 0; i < R * C; i++) a.d[i] = -a.d[i]; return a; }
template <int R, int C> Mat<R, C> operator*(Mat<R, C> a, double s) { a *= s; return a; }
template <int R, int C> Mat<R, C> operator*(double s, Mat<R, C> a) { a *= s; return a; }
template <int R, int C> Mat<R, C> operator/(Mat<R, C> a, double s) { a /= s; return a; }
template <int R, int K, int C>
Mat<R, C> operator*(const Mat<R, K> &a, const Mat<K, C> &b) {
  Mat<R, C> m;
  for (int c = 0; c < C; c++)
    for (int r = 0; r < R; r++) {
      double s = 0;
      for (int k = 0; k < K; k++) s += a(r, k) * b(k, c);
      m(r, c) = s;
    }
  return m;
}

using V3 = Mat<3, 1>;
using M3 = Mat<3, 3>;
using V6 = Mat<6, 1>;
using M6 = Mat<6, 6>;
using M36 = Mat<3, 6>;
using V15 = Mat<15, 1>;
using M15 = Mat<15, 15>;

inline V3 vec3(double x, double y, double z) { V3 v; v[0] = x; v[1] = y; v[2] = z; return v; }
inline V3 cross(const V3 &a, const V3 &b) {
  return vec3(a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]);
}

// tools.hpp:93-100 (and the SKEW_SYM_MATRX macro, tools.hpp:11)
inline M3 hat(const V3 &v) {
  M3 m;
  m(0, 1) = -v[2]; m(0, 2) = v[1];
  m(1, 0) = v[2];  m(1, 2) = -v[0];
  m(2, 0) = -v[1]; m(2, 1) = v[0];
  return m;
}

// tools.hpp:51-66
inline M3 Exp(const V3 &ang) {
  double n = ang.norm();
  if (n >= 1e-11) {
    V3 ax = ang / n;
    M3 K = hat(ax);
    return M3::Identity() + std::sin(n) * K + (1.0 - std::cos(n)) * (K * K);
  }
  return M3::Identity();
}

// tools.hpp:68-84
inline M3 Exp(const V3 &ang_vel, double dt) {
  double n = ang_vel.norm();
  if (n > 1e-7) {
    V3 ax = ang_vel / n;
    M3 K = hat(ax);
    double a = n * dt;
    return M3::Identity() + std::sin(a) * K + (1.0 - std::cos(a)) * (K * K);
  }
  return M3::Identity();
}

// tools.hpp:86-91
inline V3 Log(const M3 &R) {
  double theta = (R.trace() > 3.0 - 1e-6) ? 0.0 : std::acos(0.5 * (R.trace() - 1));
  V3 K = vec3(R(2, 1) - R(1, 2), R(0, 2) - R(2, 0), R(1, 0) - R(0, 1));
  return (std::fabs(theta) < 0.001) ? (0.5 * K) : (0.5 * theta / std::sin(theta) * K);
}

// tools.hpp:102-116
inline M3 jr(V3 vec) {
  double ang = vec.norm();
  if (ang < 1e-9) return M3::Identity();
  vec /= ang;
  double ra = std::sin(ang) / ang;
  return ra * M3::Identity() + (1 - ra) * (vec * vec.transpose()) - (1 - std::cos(ang)) / ang * hat(vec);
}

// Eigen::AngleAxisd(const Matrix3d&) = Quaternion(mat) -> angle/axis (Eigen 3.3.7
// Geometry/Quaternion.h quaternionbase_assign_impl<Other,3,3>, Geometry/AngleAxis.h operator=).
inline void angle_axis_from_matrix(const M3 &m, double &angle, V3 &axis) {
  double q[4];  // x y z w
  double t = m.trace();
  if (t > 0.0) {
    t = std::sqrt(t + 1.0);
    q[3] = 0.5 * t;
    t = 0.5 / t;
    q[0] = (m(2, 1) - m(1, 2)) * t;
    q[1] = (m(0, 2) - m(2, 0)) * t;
    q[2] = (m(1, 0) - m(0, 1)) * t;
  } else {
    int i = 0;
    if (m(1, 1) > m(0, 0)) i = 1;
    if (m(2, 2) > m(i, i)) i = 2;
    int j = (i + 1) % 3, k = (j + 1) % 3;
    t = std::sqrt(m(i, i) - m(j, j) - m(k, k) + 1.0);
    q[i] = 0.5 * t;
    t = 0.5 / t;
    q[3] = (m(k, j) - m(j, k)) * t;
    q[j] = (m(j, i) + m(i, j)) * t;
    q[k] = (m(k, i) + m(i, k)) * t;
  }
  double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
  if (n != 0.0) {
    angle = 2.0 * std::atan2(n, std::fabs(q[3]));
    if (q[3] < 0) n = -n;
    axis = vec3(q[0] / n, q[1] / n, q[2] / n);
  } else {
    angle = 0.0;
    axis = vec3(1, 0, 0);
  }
}

// tools.hpp:118-133
inline M3 jr_inv(const M3 &rotR) {
  double ang; V3 axi;
  angle_axis_from_matrix(rotR, ang, axi);
  if (ang < 1e-9) return M3::Identity();
  double ctt = ang / 2 / std::tan(ang / 2);
  return ctt * M3::Identity() + (1 - ctt) * (axi * axi.transpose()) + ang / 2 * hat(axi);
}

// ---------------------------------------------------------------------------------------
// Dense dynamic matrix / vector, column-major (Eigen::MatrixXd / VectorXd stand-ins).
// ---------------------------------------------------------------------------------------
struct MatX {
  int rows = 0, cols = 0;
  std::vector<double> d;
  MatX() {}
  MatX(int r, int c) : rows(r), cols(c), d((size_t)r * c, 0.0) {}
  void resize(int r, int c) { rows = r; cols = c; d.assign((size_t)r * c, 0.0); }
  void setZero() { std::fill(d.begin(), d.end(), 0.0); }
  double &operator()(int r, int c) { return d[(size_t)c * rows + r]; }
  const double &operator()(int r, int c) const { return d[(size_t)c * rows + r]; }
  MatX &operator+=(const MatX &o) { for (size_t i = 0; i < d.size(); i++) d[i] += o.d[i]; return *this; }
  MatX &operator*=(double s) { for (auto &x : d) x *= s; return *this; }
  template <int BR, int BC>
  void addBlock(int r0, int c0, const Mat<BR, BC> &b) {
    for (int c = 0; c < BC; c++) for (int r = 0; r < BR; r++) (*this)(r0 + r, c0 + c) += b(r, c);
  }
};
struct VecX {
  std::vector<double> d;
  VecX() {}
  explicit VecX(int n) : d(n, 0.0) {}
  void resize(int n) { d.assign(n, 0.0); }
  void setZero() { std::fill(d.begin(), d.end(), 0.0); }
  int size() const { return (int)d.size(); }
  double &operator[](int i) { return d[i]; }
  const double &operator[](int i) const { return d[i]; }
  VecX &operator+=(const VecX &o) { for (size_t i = 0; i < d.size(); i++) d[i] += o.d[i]; return *this; }
  VecX &operator*=(double s) { for (auto &x : d) x *= s; return *this; }
  double dot(const VecX &o) const { double s = 0; for (size_t i = 0; i < d.size(); i++) s += d[i] * o.d[i]; return s; }
  template <int N> Mat<N, 1> seg(int i0) const { Mat<N, 1> v; for (int i = 0; i < N; i++) v[i] = d[i0 + i]; return v; }
};

}  // namespace orc
