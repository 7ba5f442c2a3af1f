// vx_math.cuh — FP64 3-vector / 3x3 device helpers and the batched symmetric 3x3 eigen-solve.
// Everything is register-resident (no dynamically indexed local arrays).
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace vx {

struct d3 { double x, y, z; };
// symmetric 3x3: xx xy xz yy yz zz
struct s33 { double xx, xy, xz, yy, yz, zz; };
// general 3x3 row-major by name: m<row><col>
struct m33 { double m00, m01, m02, m10, m11, m12, m20, m21, m22; };

__host__ __device__ __forceinline__ d3 mk3(double x, double y, double z) { d3 r; r.x = x; r.y = y; r.z = z; return r; }
__host__ __device__ __forceinline__ d3 operator+(d3 a, d3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__host__ __device__ __forceinline__ d3 operator-(d3 a, d3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__host__ __device__ __forceinline__ d3 operator*(double s, d3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
__host__ __device__ __forceinline__ double dot(d3 a, d3 b) { return fma(a.x, b.x, fma(a.y, b.y, a.z * b.z)); }
__host__ __device__ __forceinline__ d3 cross(d3 a, d3 b) {
  return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
// a + s*b
__host__ __device__ __forceinline__ d3 axpy(double s, d3 b, d3 a) { return mk3(fma(s, b.x, a.x), fma(s, b.y, a.y), fma(s, b.z, a.z)); }

// R stored column-major in 9 doubles (Eigen): R(r,c) = R[c*3+r]
struct rot { double r00, r10, r20, r01, r11, r21, r02, r12, r22; };
__device__ __forceinline__ rot load_rot(const double *__restrict__ p) {
  rot R;
  R.r00 = p[0]; R.r10 = p[1]; R.r20 = p[2];
  R.r01 = p[3]; R.r11 = p[4]; R.r21 = p[5];
  R.r02 = p[6]; R.r12 = p[7]; R.r22 = p[8];
  return R;
}
__host__ __device__ __forceinline__ d3 rmul(const rot &R, d3 v) {   // R v
  return mk3(fma(R.r00, v.x, fma(R.r01, v.y, R.r02 * v.z)),
             fma(R.r10, v.x, fma(R.r11, v.y, R.r12 * v.z)),
             fma(R.r20, v.x, fma(R.r21, v.y, R.r22 * v.z)));
}
__host__ __device__ __forceinline__ d3 rtmul(const rot &R, d3 v) {  // R^T v
  return mk3(fma(R.r00, v.x, fma(R.r10, v.y, R.r20 * v.z)),
             fma(R.r01, v.x, fma(R.r11, v.y, R.r21 * v.z)),
             fma(R.r02, v.x, fma(R.r12, v.y, R.r22 * v.z)));
}
__host__ __device__ __forceinline__ d3 smul(const s33 &P, d3 v) {   // P v (P symmetric)
  return mk3(fma(P.xx, v.x, fma(P.xy, v.y, P.xz * v.z)),
             fma(P.xy, v.x, fma(P.yy, v.y, P.yz * v.z)),
             fma(P.xz, v.x, fma(P.yz, v.y, P.zz * v.z)));
}

// R P R^T for symmetric P (result symmetric)
__host__ __device__ __forceinline__ s33 rprt(const rot &R, const s33 &P) {
  // T = R P  (rows of T)
  d3 t0 = mk3(fma(R.r00, P.xx, fma(R.r01, P.xy, R.r02 * P.xz)), fma(R.r00, P.xy, fma(R.r01, P.yy, R.r02 * P.yz)), fma(R.r00, P.xz, fma(R.r01, P.yz, R.r02 * P.zz)));
  d3 t1 = mk3(fma(R.r10, P.xx, fma(R.r11, P.xy, R.r12 * P.xz)), fma(R.r10, P.xy, fma(R.r11, P.yy, R.r12 * P.yz)), fma(R.r10, P.xz, fma(R.r11, P.yz, R.r12 * P.zz)));
  d3 t2 = mk3(fma(R.r20, P.xx, fma(R.r21, P.xy, R.r22 * P.xz)), fma(R.r20, P.xy, fma(R.r21, P.yy, R.r22 * P.yz)), fma(R.r20, P.xz, fma(R.r21, P.yz, R.r22 * P.zz)));
  s33 o;
  o.xx = fma(t0.x, R.r00, fma(t0.y, R.r01, t0.z * R.r02));
  o.xy = fma(t0.x, R.r10, fma(t0.y, R.r11, t0.z * R.r12));
  o.xz = fma(t0.x, R.r20, fma(t0.y, R.r21, t0.z * R.r22));
  o.yy = fma(t1.x, R.r10, fma(t1.y, R.r11, t1.z * R.r12));
  o.yz = fma(t1.x, R.r20, fma(t1.y, R.r21, t1.z * R.r22));
  o.zz = fma(t2.x, R.r20, fma(t2.y, R.r21, t2.z * R.r22));
  return o;
}

// 1 / sqrt(x) for normal, positive x (the two call sites below guarantee it): MUFU.RSQ64H (2^-22.9) and one third-order
// step y (1 + e/2 + 3 e^2 / 8), e = 1 - x y^2, which leaves ~2^-68 before rounding.  The library rsqrt() wraps the same seed
// in special-case handling (zero, denormal, infinity, NaN: a branch and a slow-path call per use) that this solve never needs.
__host__ __device__ __forceinline__ double rsqrt_pos(double x) {
#ifdef __CUDA_ARCH__
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double e = fma(-(x * y), y, 1.0);
  return fma(y, e * fma(0.375, e, 0.5), y);
#else
  return 1.0 / sqrt(x);
#endif
}

// One Jacobi rotation annihilating a_pq of a symmetric 3x3 held in scalars; V accumulates columns.
// app,aqq,apq: the 2x2 pivot block; arp,arq: the remaining row's couplings; vXp,vXq: columns p,q of V.
__host__ __device__ __forceinline__ void jacobi_rot(double &app, double &aqq, double &apq, double &arp, double &arq,
                                           double &v0p, double &v0q, double &v1p, double &v1q, double &v2p, double &v2q) {
  if (apq == 0.0) return;
  // Rotation angle phi with tan(2 phi) = 2 apq / (aqq - app) =: b / d, |phi| <= pi/4 (the smaller root t = tan phi of
  // t^2 + 2 t d / b - 1 = 0), from the double angle: cos(2 phi) = |d| / r, r = sqrt(d^2 + b^2), c^2 = (1 + cos 2phi) / 2,
  // s = sin(2 phi) / (2 c), t = s / c.  Two rsqrt and no division or square root per rotation: the batched solve is bound by
  // FP64 issue slots, and rsqrt is the cheapest of the three (rsqrt_pos above).
  const double d = aqq - app, b2 = 2.0 * apq;
  const double h = fma(d, d, b2 * b2);
  if (h < 1e-290) { apq = 0.0; return; }   // |apq| and |d| below 1e-145 of the matrix scale
  const double rs = rsqrt_pos(h);
  const double cc = fma(0.5 * fabs(d), rs, 0.5);      // c^2 in [1/2, 1]
  const double q = rsqrt_pos(cc);                      // 1 / c
  const double c = cc * q;
  const double s = (d < 0.0 ? -b2 : b2) * rs * (0.5 * q);
  const double t = s * q;
  app = fma(-t, apq, app);
  aqq = fma(t, apq, aqq);
  apq = 0.0;
  double rp = arp, rq = arq;
  arp = fma(c, rp, -s * rq);
  arq = fma(s, rp, c * rq);
  double a, b;
  a = v0p; b = v0q; v0p = fma(c, a, -s * b); v0q = fma(s, a, c * b);
  a = v1p; b = v1q; v1p = fma(c, a, -s * b); v1q = fma(s, a, c * b);
  a = v2p; b = v2q; v2p = fma(c, a, -s * b); v2q = fma(s, a, c * b);
}

// Symmetric 3x3 eigen-decomposition, ascending eigenvalues, unit eigenvectors in the columns of U
// (column-major 9 doubles), i.e. the contract of Eigen::SelfAdjointEigenSolver<Matrix3d> used at
// voxel_map.hpp:312 / :1416 / loop_refine.hpp:363.  Cyclic Jacobi on the max-scaled matrix:
// relative accuracy on every eigenvalue is O(eps), which is what the BA cost (lambda_0) needs.
__host__ __device__ inline void eig3_jacobi(const s33 &A, double lam[3], double U[9]) {
  double scale = fmax(fmax(fmax(fabs(A.xx), fabs(A.yy)), fmax(fabs(A.zz), fabs(A.xy))), fmax(fabs(A.xz), fabs(A.yz)));
  if (scale == 0.0) scale = 1.0;
  double inv = 1.0 / scale;
  double a00 = A.xx * inv, a11 = A.yy * inv, a22 = A.zz * inv, a01 = A.xy * inv, a02 = A.xz * inv, a12 = A.yz * inv;
  double v00 = 1, v01 = 0, v02 = 0, v10 = 0, v11 = 1, v12 = 0, v20 = 0, v21 = 0, v22 = 1;
  // Plane covariances take 3-4 sweeps (never fewer than 2): the first three run without the convergence test, which costs
  // nearly as much as a rotation; a rotation of an already diagonal pair returns at once.
#pragma unroll 1
  for (int sweep = 0; sweep < 12; sweep++) {
    if (sweep >= 3) {
      double off = fabs(a01) + fabs(a02) + fabs(a12);
      if (off <= 1e-300) break;
      // converged when every off-diagonal is negligible against its two diagonal neighbours
      const double eps = 1.1102230246251565e-16;
      bool c01 = fabs(a01) <= eps * 0.125 * fmin(fabs(a00), fabs(a11)) || a01 == 0.0;
      bool c02 = fabs(a02) <= eps * 0.125 * fmin(fabs(a00), fabs(a22)) || a02 == 0.0;
      bool c12 = fabs(a12) <= eps * 0.125 * fmin(fabs(a11), fabs(a22)) || a12 == 0.0;
      if (c01 && c02 && c12) break;
    }
    // (p,q) = (0,1): remaining index r = 2 couples through a02 (r,p) and a12 (r,q)
    jacobi_rot(a00, a11, a01, a02, a12, v00, v01, v10, v11, v20, v21);
    // (0,2): r = 1 couples through a01 (r,p) and a12 (r,q)
    jacobi_rot(a00, a22, a02, a01, a12, v00, v02, v10, v12, v20, v22);
    // (1,2): r = 0 couples through a01 (r,p) and a02 (r,q)
    jacobi_rot(a11, a22, a12, a01, a02, v01, v02, v11, v12, v21, v22);
  }
  double l0 = a00, l1 = a11, l2 = a22;
  // ascending sort with column swaps
#define VX_SWAPCOL(la, lb, xa, xb, ya, yb, za, zb) \
  if (lb < la) { double _t = la; la = lb; lb = _t; _t = xa; xa = xb; xb = _t; _t = ya; ya = yb; yb = _t; _t = za; za = zb; zb = _t; }
  VX_SWAPCOL(l0, l1, v00, v01, v10, v11, v20, v21)
  VX_SWAPCOL(l1, l2, v01, v02, v11, v12, v21, v22)
  VX_SWAPCOL(l0, l1, v00, v01, v10, v11, v20, v21)
#undef VX_SWAPCOL
  lam[0] = l0 * scale; lam[1] = l1 * scale; lam[2] = l2 * scale;
  U[0] = v00; U[1] = v10; U[2] = v20;
  U[3] = v01; U[4] = v11; U[5] = v21;
  U[6] = v02; U[7] = v12; U[8] = v22;
}

}  // namespace vx
