// vx_terms.cuh — closed-form per-voxel-frame quantities of LidarFactor::acc_evaluate2
// (voxel_map.hpp:203-252), shared by the pair-list path (vx_factor.cu) and the fused
// small-window path (vx_hess_small.cu).
#pragma once
#include "vx_math.cuh"

namespace vx {

struct VfTerms {
  double g[18];   // g1[6] g2[6] w[6]
  double jj[6];   // Auk^T uk
  double E[36];   // diagonal-block corrections, column-major 6x6
};

// All per-voxel-frame quantities of voxel_map.hpp:203-252 in closed form (see file header).
__device__ __forceinline__ void vf_terms(const rot &R, d3 p, const s33 &P, d3 v, double n, d3 uk, d3 u1, d3 u2, d3 vbar,
                                         double iNN /* 1 / NN */, VfTerms &o) {
  d3 a = rtmul(R, uk), b1 = rtmul(R, u1), b2 = rtmul(R, u2);
  d3 Pa = smul(P, a);
  d3 t = p - vbar;
  double c = dot(uk, t);
  d3 combo2 = axpy(n, t, rmul(R, v));
  d3 va = cross(v, a);                       // viRiTuk
  double c2uk = dot(combo2, uk);
  // gradient
  d3 jr = (2.0 * iNN) * axpy(c, va, cross(Pa, a));
  d3 jt = (2.0 * iNN * c2uk) * uk;
  o.jj[0] = jr.x; o.jj[1] = jr.y; o.jj[2] = jr.z; o.jj[3] = jt.x; o.jj[4] = jt.y; o.jj[5] = jt.z;
  // g_m
  {
    d3 q = axpy(dot(t, u1), v, smul(P, b1));
    d3 gr = iNN * (cross(Pa, b1) - cross(a, q) + c * cross(v, b1));
    d3 gt = iNN * axpy(c2uk, u1, dot(combo2, u1) * uk);
    o.g[0] = gr.x; o.g[1] = gr.y; o.g[2] = gr.z; o.g[3] = gt.x; o.g[4] = gt.y; o.g[5] = gt.z;
  }
  {
    d3 q = axpy(dot(t, u2), v, smul(P, b2));
    d3 gr = iNN * (cross(Pa, b2) - cross(a, q) + c * cross(v, b2));
    d3 gt = iNN * axpy(c2uk, u2, dot(combo2, u2) * uk);
    o.g[6] = gr.x; o.g[7] = gr.y; o.g[8] = gr.z; o.g[9] = gt.x; o.g[10] = gt.y; o.g[11] = gt.z;
  }
  o.g[12] = va.x; o.g[13] = va.y; o.g[14] = va.z; o.g[15] = n * uk.x; o.g[16] = n * uk.y; o.g[17] = n * uk.z;
  // E00 = (2/NN) (combo1 - [a]x P) [a]x - 0.5 [jj_rot]x,  combo1 = [Pa + c v]x =: [q]x
  //   [q]x [a]x = a q^T - (q.a) I ;   ([a]x P) [a]x : rows m_r of [a]x P give rows (m_r x a)
  d3 q = axpy(c, v, Pa);
  double qa = dot(q, a);
  d3 p0 = mk3(P.xx, P.xy, P.xz), p1 = mk3(P.xy, P.yy, P.yz), p2 = mk3(P.xz, P.yz, P.zz);
  d3 c0 = cross(a, p0), c1 = cross(a, p1), c2 = cross(a, p2);   // columns of [a]x P
  d3 m0 = mk3(c0.x, c1.x, c2.x), m1 = mk3(c0.y, c1.y, c2.y), m2 = mk3(c0.z, c1.z, c2.z);  // rows
  d3 r0 = cross(m0, a), r1 = cross(m1, a), r2 = cross(m2, a);
  const double s2 = 2.0 * iNN;
#pragma unroll
  for (int k = 0; k < 36; k++) o.E[k] = 0.0;
  // rot-rot (rows 0..2, cols 0..2), column-major index c*6+r
  o.E[0 * 6 + 0] = s2 * (a.x * q.x - qa - r0.x);
  o.E[1 * 6 + 0] = s2 * (a.x * q.y - r0.y) + 0.5 * jr.z;
  o.E[2 * 6 + 0] = s2 * (a.x * q.z - r0.z) - 0.5 * jr.y;
  o.E[0 * 6 + 1] = s2 * (a.y * q.x - r1.x) - 0.5 * jr.z;
  o.E[1 * 6 + 1] = s2 * (a.y * q.y - qa - r1.y);
  o.E[2 * 6 + 1] = s2 * (a.y * q.z - r1.z) + 0.5 * jr.x;
  o.E[0 * 6 + 2] = s2 * (a.z * q.x - r2.x) + 0.5 * jr.y;
  o.E[1 * 6 + 2] = s2 * (a.z * q.y - r2.y) - 0.5 * jr.x;
  o.E[2 * 6 + 2] = s2 * (a.z * q.z - qa - r2.z);
  // rot-tra: (2/NN) va uk^T and its transpose
  const double vaa[3] = {va.x, va.y, va.z}, uka[3] = {uk.x, uk.y, uk.z};
#pragma unroll
  for (int r = 0; r < 3; r++)
#pragma unroll
    for (int cc = 0; cc < 3; cc++) {
      double x = s2 * vaa[r] * uka[cc];
      o.E[(3 + cc) * 6 + r] = x;          // (r, 3+cc)
      o.E[r * 6 + (3 + cc)] = x;          // (3+cc, r)
      o.E[(3 + cc) * 6 + (3 + r)] = s2 * n * uka[r] * uka[cc];
    }
}


}  // namespace vx
