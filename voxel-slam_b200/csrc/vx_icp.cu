// vx_icp.cu — icp_normal (loop_refine.hpp:47-145): point-to-plane ICP with normal-compatibility gates, the geometric check of
// a loop-closure candidate.
//
// The reference asks a pcl::KdTreeFLANN for the nearest target point of every transformed source point, 20 times.  Here the
// target cloud is binned once into a uniform grid (cell = 1 m, stable radix sort by cell, so a cell's points stay in index
// order), and one thread per source point scans the cells that intersect the gate ball |pi - tpi| < paras[3] (3 m, then 1 m):
// a nearest neighbour outside that ball is rejected by the reference's own gate (:98), so the accepted set is the same.
// Everything that decides an association is evaluated in the reference's operation order with explicit round-to-nearest
// multiplies and adds — the double transform R*p + t, its narrowing to float, the float squared distance
// ((dx*dx) + dy*dy) + dz*dz of FLANN's L2_Simple, lowest target index among equal distances — so the association is exact.
// Hess (21 unique) | JacT 6 | resi | match_num | sum tni tni^T (6 unique) are reduced by warp shuffles and a fixed second
// stage; the 6x6 LDL^T solve, the retraction and the convergence logic (:116-134) run on the host.
#include "vx_common.cuh"
#include "vx_math.cuh"
#include "vx_hostmat.hpp"
#include <cub/cub.cuh>
#include <algorithm>
#include <cmath>
#include <vector>

using namespace vx;

namespace {

constexpr int ICP_NT = 128;
constexpr int ICP_NACC = 35;     // Hess upper 21 | JacT 6 | resi 1 | match 1 | mat_norm upper 6

struct IcpGrid { float ox, oy, oz, inv_c; int nx, ny, nz; };
struct IcpPose { double R[9]; double p[3]; double paras[4]; };

__device__ __forceinline__ int icp_cell(float v, float o, float inv_c, int n) {
  int i = (int)floorf((v - o) * inv_c);
  return i < 0 ? 0 : (i >= n ? n - 1 : i);
}

__global__ void k_icp_cells(const float *__restrict__ xyz, int64_t n, IcpGrid g, int32_t *__restrict__ cell, int32_t *__restrict__ idx) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int ix = icp_cell(xyz[3 * i], g.ox, g.inv_c, g.nx), iy = icp_cell(xyz[3 * i + 1], g.oy, g.inv_c, g.ny), iz = icp_cell(xyz[3 * i + 2], g.oz, g.inv_c, g.nz);
  cell[i] = ix + g.nx * (iy + g.ny * iz);
  idx[i] = (int32_t)i;
}

// start[c] = first sorted position with cell >= c (c = 0..ncell)
__global__ void k_icp_starts(const int32_t *__restrict__ cell_sorted, int64_t n, int64_t ncell, int32_t *__restrict__ start) {
  const int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (c > ncell) return;
  int64_t lo = 0, hi = n;
  while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (cell_sorted[mid] < c) lo = mid + 1; else hi = mid; }
  start[c] = (int32_t)lo;
}

// sorted target records: xyz + original index in one 16-byte load
__global__ void k_icp_gather(const float *__restrict__ xyz, const int32_t *__restrict__ idx_sorted, int64_t n, float4 *__restrict__ rec) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int32_t j = idx_sorted[i];
  rec[i] = make_float4(xyz[3 * (int64_t)j], xyz[3 * (int64_t)j + 1], xyz[3 * (int64_t)j + 2], __int_as_float(j));
}

// s = ((a0*b0 + a1*b1) + a2*b2), no contraction: the order of the reference's 3-term products
__device__ __forceinline__ double dot3_rn(double a0, double a1, double a2, double b0, double b1, double b2) {
  return __dadd_rn(__dadd_rn(__dmul_rn(a0, b0), __dmul_rn(a1, b1)), __dmul_rn(a2, b2));
}

__global__ void __launch_bounds__(ICP_NT)
k_icp_pass(const float *__restrict__ src_xyz, const float *__restrict__ src_nrm, int64_t ns, IcpPose X, IcpGrid g,
           const int32_t *__restrict__ start, const float4 *__restrict__ rec, const float *__restrict__ tar_nrm, int64_t nt,
           double *__restrict__ partial) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  double a[ICP_NACC];
#pragma unroll
  for (int k = 0; k < ICP_NACC; k++) a[k] = 0.0;
  if (i < ns && nt > 0) {
    const double lx = (double)src_xyz[3 * i], ly = (double)src_xyz[3 * i + 1], lz = (double)src_xyz[3 * i + 2];
    // pi = R * plocal + p, ni = R * ni   (column-major R; loop_refine.hpp:82, :88)
    const double px = __dadd_rn(dot3_rn(X.R[0], X.R[3], X.R[6], lx, ly, lz), X.p[0]);
    const double py = __dadd_rn(dot3_rn(X.R[1], X.R[4], X.R[7], lx, ly, lz), X.p[1]);
    const double pz = __dadd_rn(dot3_rn(X.R[2], X.R[5], X.R[8], lx, ly, lz), X.p[2]);
    const double sx = (double)src_nrm[3 * i], sy = (double)src_nrm[3 * i + 1], sz = (double)src_nrm[3 * i + 2];
    const double nx = dot3_rn(X.R[0], X.R[3], X.R[6], sx, sy, sz);
    const double ny = dot3_rn(X.R[1], X.R[4], X.R[7], sx, sy, sz);
    const double nz = dot3_rn(X.R[2], X.R[5], X.R[8], sx, sy, sz);
    const float qx = (float)px, qy = (float)py, qz = (float)pz;          // use_search_point (:84-86)
    // cells that intersect the gate ball (+ a margin for the narrowing of the query)
    const float r = (float)X.paras[3] + 1e-3f;
    int bi = -1; float bd = 0.f, bx = 0.f, by = 0.f, bz = 0.f;
    const float fx0 = (qx - r - g.ox) * g.inv_c, fx1 = (qx + r - g.ox) * g.inv_c;
    const float fy0 = (qy - r - g.oy) * g.inv_c, fy1 = (qy + r - g.oy) * g.inv_c;
    const float fz0 = (qz - r - g.oz) * g.inv_c, fz1 = (qz + r - g.oz) * g.inv_c;
    if (fx1 >= 0.f && fy1 >= 0.f && fz1 >= 0.f && fx0 < (float)g.nx && fy0 < (float)g.ny && fz0 < (float)g.nz) {
      const int x0 = max(0, (int)floorf(fx0)), x1 = min(g.nx - 1, (int)floorf(fx1));
      const int y0 = max(0, (int)floorf(fy0)), y1 = min(g.ny - 1, (int)floorf(fy1));
      const int z0 = max(0, (int)floorf(fz0)), z1 = min(g.nz - 1, (int)floorf(fz1));
      for (int iz = z0; iz <= z1; iz++)
        for (int iy = y0; iy <= y1; iy++) {
          // cells x0..x1 of one row are contiguous in the sorted order
          const int c0 = x0 + g.nx * (iy + g.ny * iz);
          const int32_t jb = start[c0], je = start[c0 + (x1 - x0) + 1];
          for (int32_t j = jb; j < je; j++) {
            const float4 t = rec[j];
            const float dx = __fsub_rn(qx, t.x), dy = __fsub_rn(qy, t.y), dz = __fsub_rn(qz, t.z);
            const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
            const int id = __float_as_int(t.w);
            if (bi < 0 || d2 < bd || (d2 == bd && id < bi)) { bd = d2; bi = id; bx = t.x; by = t.y; bz = t.z; }
          }
        }
    }
    if (bi >= 0) {
      const float *tn = tar_nrm + 3 * (int64_t)bi;
      const double tnx = (double)tn[0], tny = (double)tn[1], tnz = (double)tn[2];
      const double dx = __dadd_rn(px, -(double)bx), dy = __dadd_rn(py, -(double)by), dz = __dadd_rn(pz, -(double)bz);   // pi - tpi
      // gates (:93-98)
      const double ix_ = __dadd_rn(nx, -tnx), iy_ = __dadd_rn(ny, -tny), iz_ = __dadd_rn(nz, -tnz);
      const double ax_ = __dadd_rn(nx, tnx), ay_ = __dadd_rn(ny, tny), az_ = __dadd_rn(nz, tnz);
      const double inc = sqrt(dot3_rn(ix_, iy_, iz_, ix_, iy_, iz_)), add = sqrt(dot3_rn(ax_, ay_, az_, ax_, ay_, az_));
      const double p2p = sqrt(dot3_rn(dx, dy, dz, dx, dy, dz));
      const double rr = dot3_rn(tnx, tny, tnz, dx, dy, dz);
      if ((inc < X.paras[0] || add < X.paras[1]) && fabs(rr) < X.paras[2] && p2p < X.paras[3]) {
        // jac = [ hat(plocal) R^T tni ; tni ]   (:102-104)
        const double v0 = dot3_rn(X.R[0], X.R[1], X.R[2], tnx, tny, tnz);
        const double v1 = dot3_rn(X.R[3], X.R[4], X.R[5], tnx, tny, tnz);
        const double v2 = dot3_rn(X.R[6], X.R[7], X.R[8], tnx, tny, tnz);
        double j[6];
        j[0] = __dadd_rn(__dmul_rn(-lz, v1), __dmul_rn(ly, v2));
        j[1] = __dadd_rn(__dmul_rn(lz, v0), __dmul_rn(-lx, v2));
        j[2] = __dadd_rn(__dmul_rn(-ly, v0), __dmul_rn(lx, v1));
        j[3] = tnx; j[4] = tny; j[5] = tnz;
        int k = 0;
#pragma unroll
        for (int c = 0; c < 6; c++)
#pragma unroll
          for (int rI = 0; rI <= c; rI++) a[k++] = __dmul_rn(j[rI], j[c]);       // Hess upper, column by column
#pragma unroll
        for (int rI = 0; rI < 6; rI++) a[21 + rI] = __dmul_rn(j[rI], rr);
        a[27] = __dmul_rn(__dmul_rn(0.5, rr), rr);
        a[28] = 1.0;
        a[29] = __dmul_rn(tnx, tnx); a[30] = __dmul_rn(tnx, tny); a[31] = __dmul_rn(tnx, tnz);
        a[32] = __dmul_rn(tny, tny); a[33] = __dmul_rn(tny, tnz); a[34] = __dmul_rn(tnz, tnz);
      }
    }
  }
  // fixed-shape reduction: warp shuffles, then the block's warps in order
  __shared__ double sh[ICP_NT / 32][ICP_NACC];
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < ICP_NACC; k++) {
    double x = a[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
    if (lane == 0) sh[wp][k] = x;
  }
  __syncthreads();
  if (threadIdx.x < ICP_NACC) {
    double sacc = 0.0;
#pragma unroll
    for (int w = 0; w < ICP_NT / 32; w++) sacc += sh[w][threadIdx.x];
    partial[(size_t)blockIdx.x * ICP_NACC + threadIdx.x] = sacc;
  }
}

// out[k] = sum over blocks, ascending
__global__ void k_icp_fin(const double *__restrict__ partial, int nblocks, double *__restrict__ out) {
  const int k = threadIdx.x;
  if (k >= ICP_NACC) return;
  double s = 0.0;
  for (int b = 0; b < nblocks; b++) s += partial[(size_t)b * ICP_NACC + k];
  out[k] = s;
}

}  // namespace

extern "C" int vx_icp_normal(vx_ctx *ctx, const float *src_xyz, const float *src_nrm, int64_t ns, const float *tar_xyz, const float *tar_nrm,
                             int64_t nt, vx_pose *pose, double icp_eigval, int32_t *accepted, int32_t *iters, int32_t *match_nums20,
                             double *resis20, double *eig3, int32_t *is_converge_out) {
  VX_ARG(pose != nullptr && ns >= 0 && nt >= 0 && (ns == 0 || (src_xyz && src_nrm)) && (nt == 0 || (tar_xyz && tar_nrm)));
  VX_ARG(ns < ((int64_t)1 << 31) && nt < ((int64_t)1 << 31));
  VX_TRY(bind(ctx));
  cudaStream_t st = ctx->stream;
  // grid over the target's bounding box: 1 m cells, coarser when the box is huge
  IcpGrid g{0.f, 0.f, 0.f, 1.f, 1, 1, 1};
  if (nt > 0) {
    float lo[3] = {tar_xyz[0], tar_xyz[1], tar_xyz[2]}, hi[3] = {tar_xyz[0], tar_xyz[1], tar_xyz[2]};
    for (int64_t i = 0; i < nt; i++)
      for (int k = 0; k < 3; k++) { const float v = tar_xyz[3 * i + k]; VX_ARG(std::isfinite(v)); lo[k] = std::min(lo[k], v); hi[k] = std::max(hi[k], v); }
    double c = 1.0;
    for (;;) {
      const double nx = std::floor((hi[0] - lo[0]) / c) + 1, ny = std::floor((hi[1] - lo[1]) / c) + 1, nz = std::floor((hi[2] - lo[2]) / c) + 1;
      if (nx * ny * nz <= 8e6) { g.nx = (int)nx; g.ny = (int)ny; g.nz = (int)nz; break; }
      c *= 2.0;
    }
    g.ox = lo[0]; g.oy = lo[1]; g.oz = lo[2]; g.inv_c = (float)(1.0 / c);
  }
  const int64_t ncell = (int64_t)g.nx * g.ny * g.nz;
  DevBuf d_sx, d_sn, d_tx, d_tn, d_c0, d_c1, d_i0, d_i1, d_start, d_rec, d_part;
  hm::M3 R = hm::M3::from(pose->R);
  hm::V3 p = hm::v3(pose->p[0], pose->p[1], pose->p[2]);
  double paras[4] = {0.2, 0.2, 0.5, 3};                       // loop_refine.hpp:63
  int is_converge = 0, n_it = 0;
  double mat_norm[6] = {0, 0, 0, 0, 0, 0};
  auto body = [&]() -> int {
    const int nblocks = std::max(1, blocks_for(ns, ICP_NT));
    VX_TRY(d_sx.reserve(12 * std::max<int64_t>(ns, 1))); VX_TRY(d_sn.reserve(12 * std::max<int64_t>(ns, 1)));
    VX_TRY(d_tx.reserve(12 * std::max<int64_t>(nt, 1))); VX_TRY(d_tn.reserve(12 * std::max<int64_t>(nt, 1)));
    VX_TRY(d_c0.reserve(4 * std::max<int64_t>(nt, 1))); VX_TRY(d_c1.reserve(4 * std::max<int64_t>(nt, 1)));
    VX_TRY(d_i0.reserve(4 * std::max<int64_t>(nt, 1))); VX_TRY(d_i1.reserve(4 * std::max<int64_t>(nt, 1)));
    VX_TRY(d_start.reserve(4 * (ncell + 2))); VX_TRY(d_rec.reserve(16 * std::max<int64_t>(nt, 1)));
    VX_TRY(d_part.reserve(sizeof(double) * ICP_NACC * ((size_t)nblocks + 1)));
    if (ns) {
      VX_CUDA(cudaMemcpyAsync(d_sx.p, src_xyz, 12 * ns, cudaMemcpyHostToDevice, st));
      VX_CUDA(cudaMemcpyAsync(d_sn.p, src_nrm, 12 * ns, cudaMemcpyHostToDevice, st));
    }
    if (nt) {
      VX_CUDA(cudaMemcpyAsync(d_tx.p, tar_xyz, 12 * nt, cudaMemcpyHostToDevice, st));
      VX_CUDA(cudaMemcpyAsync(d_tn.p, tar_nrm, 12 * nt, cudaMemcpyHostToDevice, st));
      VX_KERNEL(ctx, "k_icp_cells", k_icp_cells<<<blocks_for(nt, 256), 256, 0, st>>>(d_tx.as<float>(), nt, g, d_c0.as<int32_t>(), d_i0.as<int32_t>()));
      int bits = 1; while (((int64_t)1 << bits) < ncell) bits++;
      size_t tb = 0;
      VX_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, d_c0.as<int32_t>(), d_c1.as<int32_t>(), d_i0.as<int32_t>(), d_i1.as<int32_t>(), (int)nt, 0, bits, st));
      VX_TRY(ctx->scratch.reserve(tb));
      VX_CUDA(cub::DeviceRadixSort::SortPairs(ctx->scratch.p, tb, d_c0.as<int32_t>(), d_c1.as<int32_t>(), d_i0.as<int32_t>(), d_i1.as<int32_t>(), (int)nt, 0, bits, st));
      ctx->launches += 1;
      VX_KERNEL(ctx, "k_icp_starts", k_icp_starts<<<blocks_for(ncell + 1, 256), 256, 0, st>>>(d_c1.as<int32_t>(), nt, ncell, d_start.as<int32_t>()));
      VX_KERNEL(ctx, "k_icp_gather", k_icp_gather<<<blocks_for(nt, 256), 256, 0, st>>>(d_tx.as<float>(), d_i1.as<int32_t>(), nt, d_rec.as<float4>()));
    }
    double *fin = d_part.as<double>() + (size_t)nblocks * ICP_NACC;
    std::vector<int> tr; std::vector<double> tmp;
    for (int iterCount = 0; iterCount < 20; iterCount++) {    // :70
      IcpPose X;
      R.to(X.R); for (int k = 0; k < 3; k++) X.p[k] = p[k];
      for (int k = 0; k < 4; k++) X.paras[k] = paras[k];
      double acc[ICP_NACC] = {0};
      if (ns > 0) {
        VX_KERNEL(ctx, "k_icp_pass", k_icp_pass<<<nblocks, ICP_NT, 0, st>>>(d_sx.as<float>(), d_sn.as<float>(), ns, X, g, d_start.as<int32_t>(),
                                                                            d_rec.as<float4>(), d_tn.as<float>(), nt, d_part.as<double>()));
        VX_KERNEL(ctx, "k_icp_fin", k_icp_fin<<<1, 64, 0, st>>>(d_part.as<double>(), nblocks, fin));
        VX_CUDA(cudaMemcpyAsync(ctx->pinned, fin, sizeof(double) * ICP_NACC, cudaMemcpyDeviceToHost, st));
        VX_CUDA(cudaStreamSynchronize(st));
        memcpy(acc, ctx->pinned, sizeof(double) * ICP_NACC);
      }
      double A[36], b[6], dxi[6];
      for (int c = 0, k = 0; c < 6; c++)
        for (int r = 0; r <= c; r++, k++) { A[c * 6 + r] = acc[k]; A[r * 6 + c] = acc[k]; }
      for (int r = 0; r < 6; r++) b[r] = -acc[21 + r];
      hm::ldlt_lower_solve(A, 6, b, dxi, tr, tmp);             // Hess.ldlt().solve(-JacT) (:116)
      R = R * hm::so3_exp(hm::v3(dxi[0], dxi[1], dxi[2]));    // :117
      for (int k = 0; k < 3; k++) p[k] += dxi[3 + k];
      n_it = iterCount + 1;
      if (match_nums20) match_nums20[iterCount] = (int32_t)std::llround(acc[28]);
      if (resis20) resis20[iterCount] = acc[27];
      for (int k = 0; k < 6; k++) mat_norm[k] = acc[29 + k];
      const double rn = std::sqrt(dxi[0] * dxi[0] + dxi[1] * dxi[1] + dxi[2] * dxi[2]);
      const double tn = std::sqrt(dxi[3] * dxi[3] + dxi[4] * dxi[4] + dxi[5] * dxi[5]);
      if (rn < 1e-3 && tn < 1e-3) {                            // :122-131
        if (is_converge) break;
        paras[0] = 0.1; paras[1] = 0.1; paras[2] = 0.1; paras[3] = 1;
        is_converge = 1;
      }
    }
    return VX_OK;
  };
  const int rc = body();
  cudaStreamSynchronize(st);
  for (DevBuf *b : {&d_sx, &d_sn, &d_tx, &d_tn, &d_c0, &d_c1, &d_i0, &d_i1, &d_start, &d_rec, &d_part}) b->release();
  VX_TRY(rc);
  R.to(pose->R);
  for (int k = 0; k < 3; k++) pose->p[k] = p[k];
  // smallest eigenvalue of sum tni tni^T of the last pass (:139-142)
  s33 M; M.xx = mat_norm[0]; M.xy = mat_norm[1]; M.xz = mat_norm[2]; M.yy = mat_norm[3]; M.yz = mat_norm[4]; M.zz = mat_norm[5];
  double lam[3], U[9];
  eig3_jacobi(M, lam, U);
  if (eig3) for (int k = 0; k < 3; k++) eig3[k] = lam[k];
  if (iters) *iters = n_it;
  if (is_converge_out) *is_converge_out = is_converge;
  if (accepted) *accepted = (lam[0] > icp_eigval && is_converge == 1) ? 1 : 0;
  return VX_OK;
}
