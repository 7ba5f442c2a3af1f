// vx_hostmat.hpp — small fixed-size host algebra for the parts of the optimizers that stay on the CPU
// (IMU preintegration factor, LM bookkeeping, 15x15 gains).  Column-major like the boundary structs
// (Eigen's default), so POD arrays map without transposition.
#pragma once
#include <cmath>
#include <cstddef>
#include <cstring>
#include <limits>
#include <utility>
#include <vector>

#ifdef __CUDACC__
#define VX_HD __host__ __device__
#else
#define VX_HD
#endif

namespace vx {
namespace hm {

template <int R, int C>
struct M {
  double a[R * C];
  VX_HD M() { for (int i = 0; i < R * C; i++) a[i] = 0.0; }
  VX_HD double &operator()(int r, int c) { return a[c * R + r]; }
  VX_HD double operator()(int r, int c) const { return a[c * R + r]; }
  VX_HD double &operator[](int i) { return a[i]; }
  VX_HD double operator[](int i) const { return a[i]; }
  VX_HD static M eye() { M m; for (int i = 0; i < (R < C ? R : C); i++) m(i, i) = 1.0; return m; }
  VX_HD static M from(const double *p) { M m; for (int i = 0; i < R * C; i++) m.a[i] = p[i]; return m; }
  VX_HD void to(double *p) const { for (int i = 0; i < R * C; i++) p[i] = a[i]; }
  VX_HD M<C, R> T() const { M<C, R> t; for (int r = 0; r < R; r++) for (int c = 0; c < C; c++) t(c, r) = (*this)(r, c); return t; }
  template <int BR, int BC> VX_HD M<BR, BC> blk(int r0, int c0) const {
    M<BR, BC> b; for (int c = 0; c < BC; c++) for (int r = 0; r < BR; r++) b(r, c) = (*this)(r0 + r, c0 + c); return b;
  }
  template <int BR, int BC> VX_HD void put(int r0, int c0, const M<BR, BC> &b) {
    for (int c = 0; c < BC; c++) for (int r = 0; r < BR; r++) (*this)(r0 + r, c0 + c) = b(r, c);
  }
  VX_HD M &operator+=(const M &o) { for (int i = 0; i < R * C; i++) a[i] += o.a[i]; return *this; }
  VX_HD M &operator-=(const M &o) { for (int i = 0; i < R * C; i++) a[i] -= o.a[i]; return *this; }
  VX_HD M &operator*=(double s) { for (int i = 0; i < R * C; i++) a[i] *= s; return *this; }
  VX_HD double dot(const M &o) const { double s = 0; for (int i = 0; i < R * C; i++) s += a[i] * o.a[i]; return s; }
  VX_HD double norm() const { return std::sqrt(dot(*this)); }
  VX_HD double trace() const { double s = 0; for (int i = 0; i < (R < C ? R : C); i++) s += (*this)(i, i); return s; }
};
template <int R, int C> VX_HD M<R, C> operator+(M<R, C> x, const M<R, C> &y) { x += y; return x; }
template <int R, int C> VX_HD M<R, C> operator-(M<R, C> x, const M<R, C> &y) { x -= y; return x; }
template <int R, int C> VX_HD M<R, C> operator-(M<R, C> x) { for (int i = 0; i < R * C; i++) x.a[i] = -x.a[i]; return x; }
template <int R, int C> VX_HD M<R, C> operator*(M<R, C> x, double s) { x *= s; return x; }
template <int R, int C> VX_HD M<R, C> operator*(double s, M<R, C> x) { x *= s; return x; }
template <int R, int C> VX_HD M<R, C> operator/(M<R, C> x, double s) { for (int i = 0; i < R * C; i++) x.a[i] /= s; return x; }
template <int R, int K, int C> VX_HD M<R, C> operator*(const M<R, K> &x, const M<K, C> &y) {
  M<R, C> m;
  for (int c = 0; c < C; c++)
    for (int r = 0; r < R; r++) { double s = 0; for (int k = 0; k < K; k++) s += x(r, k) * y(k, c); m(r, c) = s; }
  return m;
}
// x^T y without forming the transpose (same sums, same order as x.T() * y)
template <int K, int R, int C> VX_HD M<R, C> mul_tn(const M<K, R> &x, const M<K, C> &y) {
  M<R, C> m;
  for (int c = 0; c < C; c++)
    for (int r = 0; r < R; r++) { double s = 0; for (int k = 0; k < K; k++) s += x(k, r) * y(k, c); m(r, c) = s; }
  return m;
}
using V3 = M<3, 1>;
using M3 = M<3, 3>;
using V15 = M<15, 1>;
using M15 = M<15, 15>;

VX_HD inline V3 v3(double x, double y, double z) { V3 v; v[0] = x; v[1] = y; v[2] = z; return v; }
VX_HD inline M3 hat(const V3 &v) {   // tools.hpp:93-100
  M3 m;
  m(0, 1) = -v[2]; m(0, 2) = v[1];
  m(1, 0) = v[2];  m(1, 2) = -v[0];
  m(2, 0) = -v[1]; m(2, 1) = v[0];
  return m;
}
VX_HD inline M3 so3_exp(const V3 &w) {   // tools.hpp:51-66
  const double n = w.norm();
  if (n < 1e-11) return M3::eye();
  const M3 K = hat(w / n);
  return M3::eye() + std::sin(n) * K + (1.0 - std::cos(n)) * (K * K);
}
VX_HD inline M3 so3_exp(const V3 &rate, double dt) {   // tools.hpp:68-84
  const double n = rate.norm();
  if (!(n > 1e-7)) return M3::eye();
  const M3 K = hat(rate / n);
  const double ang = n * dt;
  return M3::eye() + std::sin(ang) * K + (1.0 - std::cos(ang)) * (K * K);
}
VX_HD inline V3 so3_log(const M3 &R) {   // tools.hpp:86-91
  const double tr = R.trace();
  const double th = (tr > 3.0 - 1e-6) ? 0.0 : std::acos(0.5 * (tr - 1));
  const V3 k = v3(R(2, 1) - R(1, 2), R(0, 2) - R(2, 0), R(1, 0) - R(0, 1));
  return (std::fabs(th) < 0.001) ? 0.5 * k : (0.5 * th / std::sin(th)) * k;
}
VX_HD inline M3 so3_jr(V3 w) {   // tools.hpp:102-116
  const double ang = w.norm();
  if (ang < 1e-9) return M3::eye();
  w = w / ang;
  const double ra = std::sin(ang) / ang;
  return ra * M3::eye() + (1 - ra) * (w * w.T()) - ((1 - std::cos(ang)) / ang) * hat(w);
}
// tools.hpp:118-133 builds an Eigen::AngleAxisd from the matrix: matrix -> quaternion (largest-diagonal branch
// selection) -> angle = 2 atan2(|xyz|, |w|), axis = xyz / (+-|xyz|).
VX_HD inline M3 so3_jr_inv(const M3 &m) {
  double q[4], t = m.trace();
  if (t > 0.0) {
    t = std::sqrt(t + 1.0); q[3] = 0.5 * t; t = 0.5 / t;
    q[0] = (m(2, 1) - m(1, 2)) * t; q[1] = (m(0, 2) - m(2, 0)) * t; q[2] = (m(1, 0) - m(0, 1)) * t;
  } else {
    int i = 0;
    if (m(1, 1) > m(0, 0)) i = 1;
    if (m(2, 2) > m(i, i)) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    t = std::sqrt(m(i, i) - m(j, j) - m(k, k) + 1.0); q[i] = 0.5 * t; t = 0.5 / t;
    q[3] = (m(k, j) - m(j, k)) * t; q[j] = (m(j, i) + m(i, j)) * t; q[k] = (m(k, i) + m(i, k)) * t;
  }
  double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
  if (n == 0.0) return M3::eye();
  const double ang = 2.0 * std::atan2(n, std::fabs(q[3]));
  if (q[3] < 0) n = -n;
  const V3 ax = v3(q[0] / n, q[1] / n, q[2] / n);
  if (ang < 1e-9) return M3::eye();
  const double ctt = ang / 2 / std::tan(ang / 2);
  return ctt * M3::eye() + (1 - ctt) * (ax * ax.T()) + (ang / 2) * hat(ax);
}

// inverse by partial-pivot LU — what Eigen's fixed-size inverse() does above 4x4 (preintegration.hpp:166, voxelslam.cpp:1054)
template <int N> VX_HD M<N, N> inverse(const M<N, N> &in) {
  M<N, N> A = in, inv;
  int perm[N];
  for (int i = 0; i < N; i++) perm[i] = i;
  for (int k = 0; k < N; k++) {
    int p = k; double best = std::fabs(A(k, k));
    for (int i = k + 1; i < N; i++) if (std::fabs(A(i, k)) > best) { best = std::fabs(A(i, k)); p = i; }
    if (p != k) { for (int c = 0; c < N; c++) { const double t_ = A(k, c); A(k, c) = A(p, c); A(p, c) = t_; } const int tp = perm[k]; perm[k] = perm[p]; perm[p] = tp; }
    const double piv = A(k, k);
    for (int i = k + 1; i < N; i++) A(i, k) /= piv;
    for (int c = k + 1; c < N; c++) { const double akc = A(k, c); for (int i = k + 1; i < N; i++) A(i, c) -= A(i, k) * akc; }
  }
  for (int c = 0; c < N; c++) {
    double y[N];
    for (int i = 0; i < N; i++) y[i] = (perm[i] == c) ? 1.0 : 0.0;
    for (int i = 0; i < N; i++) for (int j = 0; j < i; j++) y[i] -= A(i, j) * y[j];
    for (int i = N - 1; i >= 0; i--) { for (int j = i + 1; j < N; j++) y[i] -= A(i, j) * y[j]; y[i] /= A(i, i); }
    for (int i = 0; i < N; i++) inv(i, c) = y[i];
  }
  return inv;
}

// LDL^T with diagonal pivoting on the LOWER triangle (what Eigen's MatrixXd::ldlt() reads, voxel_map.hpp:458/:659/:918;
// Hess is not exactly symmetric — SURVEY appendix A.7), then solve A x = b.  A is n x n column-major, destroyed.
inline void ldlt_lower_solve(double *A, int n, const double *b, double *x, std::vector<int> &tr, std::vector<double> &tmp) {
  tr.resize(n); tmp.resize(n);
#define VX_AT(r, c) A[(size_t)(c) * n + (r)]
  for (int k = 0; k < n; k++) {
    int big = k; double best = std::fabs(VX_AT(k, k));
    for (int i = k + 1; i < n; i++) { const double v = std::fabs(VX_AT(i, i)); if (v > best) { best = v; big = i; } }
    tr[k] = big;
    if (big != k) {   // symmetric row/column interchange restricted to the lower triangle
      for (int j = 0; j < k; j++) std::swap(VX_AT(k, j), VX_AT(big, j));
      for (int i = big + 1; i < n; i++) std::swap(VX_AT(i, k), VX_AT(i, big));
      std::swap(VX_AT(k, k), VX_AT(big, big));
      for (int i = k + 1; i < big; i++) std::swap(VX_AT(i, k), VX_AT(big, i));
    }
    if (k > 0) {
      double acc = 0.0;
      for (int j = 0; j < k; j++) { tmp[j] = VX_AT(j, j) * VX_AT(k, j); acc += VX_AT(k, j) * tmp[j]; }
      VX_AT(k, k) -= acc;
      for (int j = 0; j < k; j++) {
        const double tj = tmp[j];
        if (tj == 0.0) continue;
        for (int i = k + 1; i < n; i++) VX_AT(i, k) -= VX_AT(i, j) * tj;
      }
    }
    const double akk = VX_AT(k, k);
    if (std::fabs(akk) > 0.0)
      for (int i = k + 1; i < n; i++) VX_AT(i, k) /= akk;
  }
  for (int i = 0; i < n; i++) x[i] = b[i];
  for (int k = 0; k < n; k++) if (tr[k] != k) std::swap(x[k], x[tr[k]]);
  for (int j = 0; j < n; j++) { const double xj = x[j]; if (xj != 0.0) for (int i = j + 1; i < n; i++) x[i] -= VX_AT(i, j) * xj; }
  const double tol = 1.0 / std::numeric_limits<double>::max();
  for (int i = 0; i < n; i++) x[i] = std::fabs(VX_AT(i, i)) > tol ? x[i] / VX_AT(i, i) : 0.0;
  for (int j = n - 1; j >= 0; j--) { double s = x[j]; for (int i = j + 1; i < n; i++) s -= VX_AT(i, j) * x[i]; x[j] = s; }
  for (int k = n - 1; k >= 0; k--) if (tr[k] != k) std::swap(x[k], x[tr[k]]);
#undef VX_AT
}

}  // namespace hm
}  // namespace vx
