// vx_voxelize.cu — A1 voxel keys and the three down-sampling operators (A2, A3, A4) as sort/segment passes.
//   vx_keys               the key rule of voxel_map.hpp:1907-1918 / tools.hpp:206-216, bit-exact
//   vx_downsample_voxel   tools.hpp:201-238: per-cell sequential float running mean.  A STABLE radix
//                         sort by cell key keeps the input order inside every cell, so one thread per
//                         cell replays exactly the reference's float operations: the centroids are
//                         bit-identical to the reference's, only the output ORDER differs (ascending
//                         packed key instead of unordered_map iteration order).
//   vx_downsample_close   tools.hpp:240-302 (A4): the input point nearest the float cell centroid.
//   vx_downsample_pvec    voxel_map.hpp:39-81 (A3): running double mean of point and covariance, diag(var) out.
#include "vx_voxelize.cuh"
#include <cub/cub.cuh>

using namespace vx;

namespace {

__global__ void k_keys(const double *__restrict__ xd, const float *__restrict__ xf, int64_t n3, double vs, int64_t *__restrict__ out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n3) return;
  out[i] = xd ? key_axis_d(xd[i], vs) : key_axis_f(xf[i], vs);
}

__global__ void k_ds_key(const vx_point48 *__restrict__ pts, int64_t n, double vs, uint64_t *__restrict__ keys, int32_t *__restrict__ idx,
                         int *__restrict__ bad) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const vx_point48 &p = pts[i];
  uint64_t k;
  if (!pack_key(key_axis_f(p.x, vs), key_axis_f(p.y, vs), key_axis_f(p.z, vs), k)) *bad = 1;
  keys[i] = k;
  idx[i] = (int32_t)i;
}
__global__ void k_head_flags(const uint64_t *__restrict__ keys, int64_t n, int32_t *__restrict__ flag) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  flag[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1 : 0;
}
// start[c] = sorted position of the first point of cell c; flag scanned inclusively gives cell id + 1
__global__ void k_cell_starts(const int32_t *__restrict__ flag, const int32_t *__restrict__ incl, int64_t n, int32_t *__restrict__ start) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (flag[i]) start[incl[i] - 1] = (int32_t)i;
  if (i == n - 1) start[incl[i]] = (int32_t)n;
}
// One thread per cell: x = (x*c + p.x) / (c + 1) in float, c kept in `curvature` (tools.hpp:227-230).
__global__ void k_ds_mean(const vx_point48 *__restrict__ pts, const int32_t *__restrict__ idx, const int32_t *__restrict__ start, int64_t ncell,
                          vx_point48 *__restrict__ out) {
  int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (c >= ncell) return;
  const int32_t b = start[c], e = start[c + 1];
  vx_point48 pp = pts[idx[b]];
  pp.curvature = 1.0f;
  for (int32_t k = b + 1; k < e; k++) {
    const vx_point48 &q = pts[idx[k]];
    const float c1 = __fadd_rn(pp.curvature, 1.0f);
    pp.x = __fdiv_rn(__fadd_rn(__fmul_rn(pp.x, pp.curvature), q.x), c1);
    pp.y = __fdiv_rn(__fadd_rn(__fmul_rn(pp.y, pp.curvature), q.y), c1);
    pp.z = __fdiv_rn(__fadd_rn(__fmul_rn(pp.z, pp.curvature), q.z), c1);
    pp.curvature = c1;
  }
  out[c] = pp;
}

// tools.hpp:240-302: float centroid by sequential sums, then the input point nearest to it (first minimum, strict <,
// starting from ndis = 100).  The differences are formed in float and widened (double xx = pb.x - pp.x).
__global__ void k_ds_close(const vx_point48 *__restrict__ pts, const int32_t *__restrict__ idx, const int32_t *__restrict__ start, int64_t ncell,
                           vx_point48 *__restrict__ out) {
  int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (c >= ncell) return;
  const int32_t b = start[c], e = start[c + 1];
  float sx = pts[idx[b]].x, sy = pts[idx[b]].y, sz = pts[idx[b]].z;
  for (int32_t k = b + 1; k < e; k++) {
    const vx_point48 &q = pts[idx[k]];
    sx = __fadd_rn(sx, q.x); sy = __fadd_rn(sy, q.y); sz = __fadd_rn(sz, q.z);
  }
  const float cnt = (float)(e - b);
  sx = __fdiv_rn(sx, cnt); sy = __fdiv_rn(sy, cnt); sz = __fdiv_rn(sz, cnt);
  double ndis = 100;
  int32_t best = b;
  for (int32_t k = b; k < e; k++) {
    const vx_point48 &q = pts[idx[k]];
    const double xx = (double)__fsub_rn(sx, q.x), yy = (double)__fsub_rn(sy, q.y), zz = (double)__fsub_rn(sz, q.z);
    const double dis = __dadd_rn(__dadd_rn(__dmul_rn(xx, xx), __dmul_rn(yy, yy)), __dmul_rn(zz, zz));
    if (dis < ndis) { best = k; ndis = dis; }
  }
  out[c] = pts[idx[best]];
}

__global__ void k_ds_key_pvec(const vx_point_var *__restrict__ pts, int64_t n, double vs, uint64_t *__restrict__ keys, int32_t *__restrict__ idx,
                              int *__restrict__ bad) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const vx_point_var &p = pts[i];
  uint64_t k;
  if (!pack_key(key_axis_d(p.pnt[0], vs), key_axis_d(p.pnt[1], vs), key_axis_d(p.pnt[2], vs), k)) *bad = 1;
  keys[i] = k;
  idx[i] = (int32_t)i;
}
// voxel_map.hpp:39-81: running DOUBLE mean (x*c + p)/(c+1) of the point and of the covariance; only diag(var) leaves the
// function (:73-78), and every matrix element runs its own recurrence, so the three diagonal ones suffice.
__global__ void k_ds_pvec(const vx_point_var *__restrict__ pts, const int32_t *__restrict__ idx, const int32_t *__restrict__ start, int64_t ncell,
                          vx_point48 *__restrict__ out) {
  int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (c >= ncell) return;
  const int32_t b = start[c], e = start[c + 1];
  const vx_point_var &f = pts[idx[b]];
  double m[6] = {f.pnt[0], f.pnt[1], f.pnt[2], f.var[0], f.var[4], f.var[8]};
  int cnt = 1;
  for (int32_t k = b + 1; k < e; k++) {
    const vx_point_var &q = pts[idx[k]];
    const double in[6] = {q.pnt[0], q.pnt[1], q.pnt[2], q.var[0], q.var[4], q.var[8]};
    const double cd = (double)cnt, c1 = (double)(cnt + 1);
    for (int j = 0; j < 6; j++) m[j] = __ddiv_rn(__dadd_rn(__dmul_rn(m[j], cd), in[j]), c1);
    cnt++;
  }
  vx_point48 ap;
  ap.x = (float)m[0]; ap.y = (float)m[1]; ap.z = (float)m[2]; ap._p0 = 1.0f;     // pcl::PointXYZINormal(): data[3] = 1
  ap.normal_x = (float)m[3]; ap.normal_y = (float)m[4]; ap.normal_z = (float)m[5]; ap._p1 = 0.0f;
  ap.intensity = 0.0f; ap.curvature = 0.0f; ap._p2 = ap._p3 = 0.0f;
  out[c] = ap;
}

}  // namespace

namespace vx {
// Shared driver of the three down-sampling operators on DEVICE-resident input: key -> stable sort -> cell starts -> one
// thread per cell replaying the reference's sequential arithmetic in input order.  Output order = ascending packed cell key.
int downsample_device(vx_ctx *ctx, const void *d_pts, int64_t n, double voxel_size, DsMode mode, vx_point48 *d_out, int32_t *ncell_out) {
  VX_ARG(n < ((int64_t)1 << 31));
  cudaStream_t st = ctx->stream;
  *ncell_out = 0;
  if (n == 0) return VX_OK;
  DevBuf d_k0, d_k1, d_i0, d_i1, d_flag, d_incl, d_start, d_bad;
  auto body = [&]() -> int {
    VX_TRY(d_k0.reserve(8 * n)); VX_TRY(d_k1.reserve(8 * n)); VX_TRY(d_i0.reserve(4 * n)); VX_TRY(d_i1.reserve(4 * n));
    VX_TRY(d_flag.reserve(4 * n)); VX_TRY(d_incl.reserve(4 * n)); VX_TRY(d_start.reserve(4 * (n + 1))); VX_TRY(d_bad.reserve(4));
    VX_CUDA(cudaMemsetAsync(d_bad.p, 0, 4, st));
    if (mode == DS_PVEC)
      VX_KERNEL(ctx, "k_ds_key_pvec", k_ds_key_pvec<<<blocks_for(n, 256), 256, 0, st>>>((const vx_point_var *)d_pts, n, voxel_size, d_k0.as<uint64_t>(), d_i0.as<int32_t>(), d_bad.as<int>()));
    else
      VX_KERNEL(ctx, "k_ds_key", k_ds_key<<<blocks_for(n, 256), 256, 0, st>>>((const vx_point48 *)d_pts, n, voxel_size, d_k0.as<uint64_t>(), d_i0.as<int32_t>(), d_bad.as<int>()));
    size_t tb = 0;
    VX_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, d_k0.as<uint64_t>(), d_k1.as<uint64_t>(), d_i0.as<int32_t>(), d_i1.as<int32_t>(), (int)n, 0, 63, st));
    VX_TRY(ctx->scratch.reserve(tb));
    VX_CUDA(cub::DeviceRadixSort::SortPairs(ctx->scratch.p, tb, d_k0.as<uint64_t>(), d_k1.as<uint64_t>(), d_i0.as<int32_t>(), d_i1.as<int32_t>(), (int)n, 0, 63, st));
    ctx->launches += 4;
    VX_KERNEL(ctx, "k_head_flags", k_head_flags<<<blocks_for(n, 256), 256, 0, st>>>(d_k1.as<uint64_t>(), n, d_flag.as<int32_t>()));
    VX_CUDA(cub::DeviceScan::InclusiveSum(nullptr, tb, d_flag.as<int32_t>(), d_incl.as<int32_t>(), (int)n, st));
    VX_TRY(ctx->scratch.reserve(tb));
    VX_CUDA(cub::DeviceScan::InclusiveSum(ctx->scratch.p, tb, d_flag.as<int32_t>(), d_incl.as<int32_t>(), (int)n, st));
    ctx->launches += 2;
    VX_KERNEL(ctx, "k_cell_starts", k_cell_starts<<<blocks_for(n, 256), 256, 0, st>>>(d_flag.as<int32_t>(), d_incl.as<int32_t>(), n, d_start.as<int32_t>()));
    int32_t ncell = 0; int bad = 0;
    VX_CUDA(cudaMemcpyAsync(&ncell, d_incl.as<int32_t>() + (n - 1), 4, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaMemcpyAsync(&bad, d_bad.p, 4, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    if (bad) { set_error("down-sampling: a cell index does not fit the 21-bit packed key"); return VX_ERR_ARG; }
    const int nb = blocks_for(ncell, 128);
    if (mode == DS_MEAN)
      VX_KERNEL(ctx, "k_ds_mean", k_ds_mean<<<nb, 128, 0, st>>>((const vx_point48 *)d_pts, d_i1.as<int32_t>(), d_start.as<int32_t>(), ncell, d_out));
    else if (mode == DS_CLOSE)
      VX_KERNEL(ctx, "k_ds_close", k_ds_close<<<nb, 128, 0, st>>>((const vx_point48 *)d_pts, d_i1.as<int32_t>(), d_start.as<int32_t>(), ncell, d_out));
    else
      VX_KERNEL(ctx, "k_ds_pvec", k_ds_pvec<<<nb, 128, 0, st>>>((const vx_point_var *)d_pts, d_i1.as<int32_t>(), d_start.as<int32_t>(), ncell, d_out));
    *ncell_out = ncell;
    return VX_OK;
  };
  const int status = body();
  cudaStreamSynchronize(st);
  for (DevBuf *b : {&d_k0, &d_k1, &d_i0, &d_i1, &d_flag, &d_incl, &d_start, &d_bad}) b->release();
  return status;
}
}  // namespace vx

namespace {
int downsample_cells(vx_ctx *ctx, const void *pts, size_t rec, int64_t n, double voxel_size, DsMode mode, vx_point48 *out, int64_t *out_n) {
  VX_TRY(bind(ctx));
  VX_ARG(n < ((int64_t)1 << 31));
  cudaStream_t st = ctx->stream;
  if (n == 0) { *out_n = 0; return VX_OK; }
  DevBuf d_pts, d_out;
  auto body = [&]() -> int {
    VX_TRY(d_pts.reserve(rec * n)); VX_TRY(d_out.reserve(sizeof(vx_point48) * n));
    VX_CUDA(cudaMemcpyAsync(d_pts.p, pts, rec * n, cudaMemcpyHostToDevice, st));
    int32_t ncell = 0;
    VX_TRY(downsample_device(ctx, d_pts.p, n, voxel_size, mode, d_out.as<vx_point48>(), &ncell));
    VX_CUDA(cudaMemcpyAsync(out, d_out.p, sizeof(vx_point48) * ncell, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    *out_n = ncell;
    return VX_OK;
  };
  const int status = body();
  cudaStreamSynchronize(st);
  d_pts.release(); d_out.release();
  return status;
}
}  // namespace

extern "C" {

int vx_keys(vx_ctx *ctx, const void *xyz, int64_t n, double voxel_size, int is_float_input, int64_t *out_keys) {
  VX_TRY(bind(ctx));
  VX_ARG(n >= 0 && (n == 0 || (xyz && out_keys)) && voxel_size > 0);
  if (n == 0) return VX_OK;
  const size_t esz = is_float_input ? sizeof(float) : sizeof(double);
  DevBuf &in = ctx->scratch, &out = ctx->scratch2;
  VX_TRY(in.reserve(esz * 3 * n));
  VX_TRY(out.reserve(sizeof(int64_t) * 3 * n));
  VX_CUDA(cudaMemcpyAsync(in.p, xyz, esz * 3 * n, cudaMemcpyHostToDevice, ctx->stream));
  VX_KERNEL(ctx, "k_keys", k_keys<<<blocks_for(3 * n, 256), 256, 0, ctx->stream>>>(is_float_input ? nullptr : in.as<double>(), is_float_input ? in.as<float>() : nullptr,
                                                                               3 * n, voxel_size, out.as<int64_t>()));
  VX_CUDA(cudaMemcpyAsync(out_keys, out.p, sizeof(int64_t) * 3 * n, cudaMemcpyDeviceToHost, ctx->stream));
  VX_CUDA(cudaStreamSynchronize(ctx->stream));
  return VX_OK;
}

int vx_downsample_voxel(vx_ctx *ctx, const vx_point48 *pts, int64_t n, double voxel_size, vx_point48 *out, int64_t *out_n) {
  VX_ARG(n >= 0 && out_n != nullptr && (n == 0 || (pts && out)));
  if (voxel_size < 0.001) {   // tools.hpp:203: nothing happens
    if (n) memcpy(out, pts, sizeof(vx_point48) * n);
    *out_n = n;
    return VX_OK;
  }
  return downsample_cells(ctx, pts, sizeof(vx_point48), n, voxel_size, DS_MEAN, out, out_n);
}

int vx_downsample_close(vx_ctx *ctx, const vx_point48 *pts, int64_t n, double voxel_size, vx_point48 *out, int64_t *out_n) {
  VX_ARG(n >= 0 && out_n != nullptr && (n == 0 || (pts && out)));
  if (voxel_size < 0.001) {   // tools.hpp:242
    if (n) memcpy(out, pts, sizeof(vx_point48) * n);
    *out_n = n;
    return VX_OK;
  }
  return downsample_cells(ctx, pts, sizeof(vx_point48), n, voxel_size, DS_CLOSE, out, out_n);
}

int vx_downsample_pvec(vx_ctx *ctx, const vx_point_var *pvec, int64_t n, double voxel_size, vx_point48 *out, int64_t *out_n) {
  VX_ARG(n >= 0 && out_n != nullptr && voxel_size > 0 && (n == 0 || (pvec && out)));
  return downsample_cells(ctx, pvec, sizeof(vx_point_var), n, voxel_size, DS_PVEC, out, out_n);
}

}  // extern "C"
