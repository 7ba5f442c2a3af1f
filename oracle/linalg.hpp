// ORACLE — TEST INFRASTRUCTURE ONLY (see vmath.hpp header). PARITY UNPINNED.
//
// linalg.hpp — restatement of the three Eigen 3.3.7 decompositions whose numerics sit on
// the hot path (the dependency is pinned only in prose, /root/reference/README.md:26):
//   * SelfAdjointEigenSolver<Matrix3d>  (voxel_map.hpp:312,1416,1525; loop_refine.hpp:363;
//     voxelslam.cpp:1090): scale -> 3x3 Householder tridiagonalisation -> implicit symmetric
//     QR with Wilkinson shift -> ascending sort.  (Eigen/src/Eigenvalues/SelfAdjointEigenSolver.h,
//     Tridiagonalization.h, Jacobi/Jacobi.h)
//   * MatrixXd::ldlt().solve()  (voxel_map.hpp:458,659,918): diagonally pivoted LDL^T on the
//     LOWER triangle (Eigen/src/Cholesky/LDLT.h).
//   * fixed 15x15 inverse()  (voxelslam.cpp:987,1054; preintegration.hpp:166,244): partial
//     pivot LU (Eigen/src/LU/PartialPivLU.h via InverseImpl.h).
#pragma once
#include "vmath.hpp"

namespace orc {

struct Givens { double c, s; };
inline Givens make_givens(double p, double q) {
  Givens g;
  if (q == 0.0) { g.c = p < 0 ? -1.0 : 1.0; g.s = 0.0; }
  else if (p == 0.0) { g.c = 0.0; g.s = q < 0 ? 1.0 : -1.0; }
  else if (std::fabs(p) > std::fabs(q)) {
    double t = q / p, u = std::sqrt(1.0 + t * t);
    if (p < 0) u = -u;
    g.c = 1.0 / u; g.s = -t * g.c;
  } else {
    double t = p / q, u = std::sqrt(1.0 + t * t);
    if (q < 0) u = -u;
    g.s = -1.0 / u; g.c = -t * g.s;
  }
  return g;
}

// Returns ascending eigenvalues `w` and unit eigenvectors in the columns of `U`.
// Only the lower triangle of `A` is read, as in Eigen.
inline void eig3_sym(const M3 &A, V3 &w, M3 &U) {
  double a00 = A(0, 0), a10 = A(1, 0), a20 = A(2, 0), a11 = A(1, 1), a21 = A(2, 1), a22 = A(2, 2);
  double scale = std::max({std::fabs(a00), std::fabs(a10), std::fabs(a20), std::fabs(a11), std::fabs(a21), std::fabs(a22)});
  if (scale == 0.0) scale = 1.0;
  a00 /= scale; a10 /= scale; a20 /= scale; a11 /= scale; a21 /= scale; a22 /= scale;

  double diag[3], sub[2];
  M3 Q;
  const double tiny = std::numeric_limits<double>::min();
  diag[0] = a00;
  double v1norm2 = a20 * a20;
  if (v1norm2 <= tiny) {
    diag[1] = a11; diag[2] = a22; sub[0] = a10; sub[1] = a21;
    Q.setIdentity();
  } else {
    double beta = std::sqrt(a10 * a10 + v1norm2);
    double invBeta = 1.0 / beta;
    double m01 = a10 * invBeta, m02 = a20 * invBeta;
    double q = 2.0 * m01 * a21 + m02 * (a22 - a11);
    diag[1] = a11 + m02 * q;
    diag[2] = a22 - m02 * q;
    sub[0] = beta;
    sub[1] = a21 - m01 * q;
    Q.setZero();
    Q(0, 0) = 1; Q(1, 1) = m01; Q(1, 2) = m02; Q(2, 1) = m02; Q(2, 2) = -m01;
  }

  const int n = 3;
  int end = n - 1, start = 0, iter = 0;
  const int maxIter = 30;
  const double eps = std::numeric_limits<double>::epsilon();
  while (end > 0) {
    for (int i = start; i < end; ++i) {
      // Eigen 3.3.7: isMuchSmallerThan(|sub|, |d_i|+|d_{i+1}|, precision=2*eps) || |sub| <= considerAsZero
      if (std::fabs(sub[i]) <= (std::fabs(diag[i]) + std::fabs(diag[i + 1])) * (2 * eps) || std::fabs(sub[i]) <= tiny)
        sub[i] = 0.0;
    }
    while (end > 0 && sub[end - 1] == 0.0) end--;
    if (end <= 0) break;
    iter++;
    if (iter > maxIter * n) break;
    start = end - 1;
    while (start > 0 && sub[start - 1] != 0.0) start--;

    // tridiagonal_qr_step
    double td = (diag[end - 1] - diag[end]) * 0.5;
    double e = sub[end - 1];
    double mu = diag[end];
    if (td == 0.0) {
      mu -= std::fabs(e);
    } else {
      double e2 = e * e;
      double h = std::hypot(td, e);
      if (e2 == 0.0) mu -= (e / (td + (td > 0 ? 1 : -1))) * (e / h);
      else mu -= e2 / (td + (td > 0 ? h : -h));
    }
    double x = diag[start] - mu;
    double z = sub[start];
    for (int k = start; k < end; ++k) {
      Givens r = make_givens(x, z);
      double sdk = r.s * diag[k] + r.c * sub[k];
      double dkp1 = r.s * sub[k] + r.c * diag[k + 1];
      diag[k] = r.c * (r.c * diag[k] - r.s * sub[k]) - r.s * (r.c * sub[k] - r.s * diag[k + 1]);
      diag[k + 1] = r.s * sdk + r.c * dkp1;
      sub[k] = r.c * sdk - r.s * dkp1;
      if (k > start) sub[k - 1] = r.c * sub[k - 1] - r.s * z;
      x = sub[k];
      if (k < end - 1) {
        z = -r.s * sub[k + 1];
        sub[k + 1] = r.c * sub[k + 1];
      }
      // Q = Q * G  (applyOnTheRight(k, k+1, rot))
      for (int i = 0; i < n; i++) {
        double xi = Q(i, k), yi = Q(i, k + 1);
        Q(i, k) = r.c * xi - r.s * yi;
        Q(i, k + 1) = r.s * xi + r.c * yi;
      }
    }
  }
  // ascending selection sort with column swaps
  for (int i = 0; i < n - 1; ++i) {
    int k = 0;
    double mn = diag[i];
    for (int j = 1; j < n - i; j++) if (diag[i + j] < mn) { mn = diag[i + j]; k = j; }
    if (k > 0) {
      std::swap(diag[i], diag[k + i]);
      for (int r = 0; r < n; r++) std::swap(Q(r, i), Q(r, k + i));
    }
  }
  for (int i = 0; i < 3; i++) w[i] = diag[i] * scale;
  U = Q;
}

// Pivoted LDL^T on the lower triangle of A (n x n, column-major), then solve A x = b.
// Mirrors Eigen::LDLT<MatrixXd, Lower>::compute + solve (unblocked kernel).
inline void ldlt_solve(MatX A, const VecX &b, VecX &x) {
  const int n = A.rows;
  std::vector<int> tr(n);
  std::vector<double> temp(n);
  for (int k = 0; k < n; ++k) {
    int big = k; double best = std::fabs(A(k, k));
    for (int i = k + 1; i < n; i++) { double v = std::fabs(A(i, i)); if (v > best) { best = v; big = i; } }
    tr[k] = big;
    if (k != big) {
      int s = n - big - 1;
      for (int j = 0; j < k; j++) std::swap(A(k, j), A(big, j));
      for (int i = 0; i < s; i++) std::swap(A(big + 1 + i, k), A(big + 1 + i, big));
      std::swap(A(k, k), A(big, big));
      for (int i = k + 1; i < big; ++i) { double t = A(i, k); A(i, k) = A(big, i); A(big, i) = t; }
    }
    int rs = n - k - 1;
    if (k > 0) {
      for (int j = 0; j < k; j++) temp[j] = A(j, j) * A(k, j);
      double acc = 0; for (int j = 0; j < k; j++) acc += A(k, j) * temp[j];
      A(k, k) -= acc;
      for (int j = 0; j < k; j++) {
        double tj = temp[j];
        if (tj == 0.0) continue;
        for (int i = 0; i < rs; i++) A(k + 1 + i, k) -= A(k + 1 + i, j) * tj;
      }
    }
    double akk = A(k, k);
    if (rs > 0 && std::fabs(akk) > 0.0)
      for (int i = 0; i < rs; i++) A(k + 1 + i, k) /= akk;
  }
  // solve: x = P^T L^-T D^-1 L^-1 P b
  x = b;
  for (int k = 0; k < n; k++) if (tr[k] != k) std::swap(x[k], x[tr[k]]);
  for (int j = 0; j < n; j++) { double xj = x[j]; if (xj != 0.0) for (int i = j + 1; i < n; i++) x[i] -= A(i, j) * xj; }
  const double tol = 1.0 / std::numeric_limits<double>::max();
  for (int i = 0; i < n; i++) { if (std::fabs(A(i, i)) > tol) x[i] /= A(i, i); else x[i] = 0.0; }
  for (int j = n - 1; j >= 0; j--) { double s = x[j]; for (int i = j + 1; i < n; i++) s -= A(i, j) * x[i]; x[j] = s; }
  for (int k = n - 1; k >= 0; k--) if (tr[k] != k) std::swap(x[k], x[tr[k]]);
}

// General inverse by partial-pivot LU (Eigen's path for fixed sizes > 4).
template <int N>
Mat<N, N> inverse_lu(const Mat<N, N> &Ain) {
  Mat<N, N> A = Ain;
  int perm[N];
  for (int i = 0; i < N; i++) perm[i] = i;
  for (int k = 0; k < N; k++) {
    int p = k; double best = std::fabs(A(k, k));
    for (int i = k + 1; i < N; i++) if (std::fabs(A(i, k)) > best) { best = std::fabs(A(i, k)); p = i; }
    if (p != k) { for (int c = 0; c < N; c++) std::swap(A(k, c), A(p, c)); std::swap(perm[k], perm[p]); }
    double piv = A(k, k);
    for (int i = k + 1; i < N; i++) A(i, k) /= piv;
    for (int c = k + 1; c < N; c++) { double akc = A(k, c); for (int i = k + 1; i < N; i++) A(i, c) -= A(i, k) * akc; }
  }
  Mat<N, N> inv;
  for (int c = 0; c < N; c++) {
    double y[N];
    for (int i = 0; i < N; i++) y[i] = (perm[i] == c) ? 1.0 : 0.0;
    for (int i = 0; i < N; i++) for (int j = 0; j < i; j++) y[i] -= A(i, j) * y[j];
    for (int i = N - 1; i >= 0; i--) { for (int j = i + 1; j < N; j++) y[i] -= A(i, j) * y[j]; y[i] /= A(i, i); }
    for (int i = 0; i < N; i++) inv(i, c) = y[i];
  }
  return inv;
}

}  // namespace orc
