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
#include "vx_math.cuh"
#include <cub/cub.cuh>
#include <mutex>

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

// ---- BTC plane voxels (BTC.cpp:279-334, :96-143) -----------------------------------------------
// STDescManager::init_voxel_map's key rule is NOT the map's: the quotient stays a double (p / voxel_size, minus 1.0 when
// negative) and is truncated to int64 (BTC.cpp:288-296).
__global__ void k_btc_key(const float *__restrict__ xyz, int64_t n, double vs, uint64_t *__restrict__ keys, int32_t *__restrict__ idx, int *__restrict__ bad) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int64_t c[3];
#pragma unroll
  for (int j = 0; j < 3; j++) {
    double loc = __ddiv_rn((double)xyz[3 * i + j], vs);
    if (loc < 0) loc = __dadd_rn(loc, -1.0);
    c[j] = (int64_t)loc;
  }
  uint64_t k;
  if (!pack_key(c[0], c[1], c[2], k)) *bad = 1;
  keys[i] = k;
  idx[i] = (int32_t)i;
}
// One thread per voxel: BTCOctoTree::init_octo_tree / init_plane (BTC.cpp:90-143) with the sums in insertion order; a plane
// voxel becomes the PointXYZINormal of STDescManager::get_plane (:322-334): center, normal of the smallest eigenvalue.
__global__ void k_btc_plane(const float *__restrict__ xyz, const int32_t *__restrict__ idx, const int32_t *__restrict__ start, int64_t ncell,
                            int init_num, double thre, vx_point48 *__restrict__ out, double *__restrict__ aux, int32_t *__restrict__ flag) {
  int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (c >= ncell) return;
  const int32_t b = start[c], e = start[c + 1];
  flag[c] = 0;
  if (e - b <= init_num) return;                       // voxel_points_.size() > voxel_init_num_ (:91)
  double S[6] = {0, 0, 0, 0, 0, 0}, m[3] = {0, 0, 0};
  for (int32_t k = b; k < e; k++) {
    const float *q = xyz + 3 * (int64_t)idx[k];
    const double x = (double)q[0], y = (double)q[1], z = (double)q[2];
    S[0] = __dadd_rn(S[0], __dmul_rn(x, x)); S[1] = __dadd_rn(S[1], __dmul_rn(x, y)); S[2] = __dadd_rn(S[2], __dmul_rn(x, z));
    S[3] = __dadd_rn(S[3], __dmul_rn(y, y)); S[4] = __dadd_rn(S[4], __dmul_rn(y, z)); S[5] = __dadd_rn(S[5], __dmul_rn(z, z));
    m[0] = __dadd_rn(m[0], x); m[1] = __dadd_rn(m[1], y); m[2] = __dadd_rn(m[2], z);
  }
  const double N = (double)(e - b);
  const double cx = __ddiv_rn(m[0], N), cy = __ddiv_rn(m[1], N), cz = __ddiv_rn(m[2], N);
  s33 C;
  C.xx = __dadd_rn(__ddiv_rn(S[0], N), -__dmul_rn(cx, cx)); C.xy = __dadd_rn(__ddiv_rn(S[1], N), -__dmul_rn(cx, cy));
  C.xz = __dadd_rn(__ddiv_rn(S[2], N), -__dmul_rn(cx, cz)); C.yy = __dadd_rn(__ddiv_rn(S[3], N), -__dmul_rn(cy, cy));
  C.yz = __dadd_rn(__ddiv_rn(S[4], N), -__dmul_rn(cy, cz)); C.zz = __dadd_rn(__ddiv_rn(S[5], N), -__dmul_rn(cz, cz));
  double lam[3], U[9];
  eig3_jacobi(C, lam, U);                              // ascending: lam[0] = evalsReal(evalsMin), lam[2] = evalsReal(evalsMax)
  if (!(lam[0] < thre)) return;
  vx_point48 pi;
  pi.x = (float)cx; pi.y = (float)cy; pi.z = (float)cz; pi._p0 = 1.0f;
  pi.normal_x = (float)U[0]; pi.normal_y = (float)U[1]; pi.normal_z = (float)U[2]; pi._p1 = 0.0f;
  pi.intensity = 0.0f; pi.curvature = 0.0f; pi._p2 = pi._p3 = 0.0f;
  out[c] = pi;
  // BTCPlane's doubles for callers that want them: center 3 | normal 3 | min_eigen_value | radius | d | points_size
  double *a = aux + 10 * c;
  a[0] = cx; a[1] = cy; a[2] = cz; a[3] = U[0]; a[4] = U[1]; a[5] = U[2]; a[6] = lam[0]; a[7] = sqrt(lam[2]);
  a[8] = -__dadd_rn(__dadd_rn(__dmul_rn(U[0], cx), __dmul_rn(U[1], cy)), __dmul_rn(U[2], cz)); a[9] = N;
  flag[c] = 1;
}
__global__ void k_btc_compact(const vx_point48 *__restrict__ in, const double *__restrict__ aux_in, const int32_t *__restrict__ flag,
                              const int32_t *__restrict__ incl, int64_t ncell, vx_point48 *__restrict__ out, double *__restrict__ aux_out) {
  int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (c >= ncell || !flag[c]) return;
  const int32_t o = incl[c] - 1;
  out[o] = in[c];
  for (int k = 0; k < 10; k++) aux_out[10 * (int64_t)o + k] = aux_in[10 * c + k];
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

// ---- vx_keys without any allocation (no counterpart in the reference) ----------
// vx_keys stages the caller's host coordinates and the keys in the context's grow-on-demand device scratch (cached_malloc ->
// cudaMalloc).  Callers that must not allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc) get
// the same kernel on a FIXED scratch: a file-scope device array, [coordinates | keys], sized for VX_KEYS_STATIC_MAX_POINTS
// double-precision points (6 MiB + 6 MiB; every GPU that loads the library holds it until exit).  In host code the array's name
// is only a stand-in: the device address comes from cudaGetSymbolAddress, once per device with that device current.
// All calls on a device share the array and it keeps the last call's contents; nothing in it has to start at zero (the kernel
// writes every key that is read back), but two calls that overlap on one GPU race on it.
namespace {
constexpr int64_t KEYS_STATIC_MAX = VX_KEYS_STATIC_MAX_POINTS;
constexpr size_t KEYS_STATIC_IN_BYTES = sizeof(double) * 3 * KEYS_STATIC_MAX;
constexpr size_t KEYS_STATIC_BYTES = KEYS_STATIC_IN_BYTES + sizeof(int64_t) * 3 * KEYS_STATIC_MAX;
static __device__ __align__(256) unsigned char g_keys_static_scratch[KEYS_STATIC_BYTES];
constexpr int KEYS_MAX_DEVICES = 64;
std::mutex g_keys_static_mu;
unsigned char *g_keys_static_ptr[KEYS_MAX_DEVICES] = {};
}  // namespace

int vx_keys_static(vx_ctx *ctx, const void *xyz, int64_t n, double voxel_size, int is_float_input, int64_t *out_keys) {
  VX_TRY(bind(ctx));                       // cudaSetDevice(ctx->device): the symbol is looked up with its device current
  VX_ARG(n >= 0 && (n == 0 || (xyz && out_keys)) && voxel_size > 0);
  if (n > KEYS_STATIC_MAX) {
    set_error("vx_keys_static: %lld points exceed the static scratch of %lld points (VX_KEYS_STATIC_MAX_POINTS); use vx_keys",
              (long long)n, (long long)KEYS_STATIC_MAX);
    return VX_ERR_ARG;
  }
  VX_ARG(ctx->device >= 0 && ctx->device < KEYS_MAX_DEVICES);
  if (n == 0) return VX_OK;
  unsigned char *base = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_keys_static_mu);
    if (!g_keys_static_ptr[ctx->device]) {
      void *p = nullptr;
      VX_CUDA(cudaGetSymbolAddress(&p, g_keys_static_scratch));
      g_keys_static_ptr[ctx->device] = (unsigned char *)p;
    }
    base = g_keys_static_ptr[ctx->device];
  }
  const size_t esz = is_float_input ? sizeof(float) : sizeof(double);
  void *d_in = base;
  int64_t *d_out = reinterpret_cast<int64_t *>(base + KEYS_STATIC_IN_BYTES);
  cudaStream_t st = ctx->stream;
  VX_CUDA(cudaMemcpyAsync(d_in, xyz, esz * 3 * n, cudaMemcpyHostToDevice, st));
  k_keys<<<blocks_for(3 * n, 256), 256, 0, st>>>(is_float_input ? nullptr : (const double *)d_in, is_float_input ? (const float *)d_in : nullptr, 3 * n, voxel_size, d_out);
  VX_LAUNCH_CHECK(ctx);
  VX_CUDA(cudaMemcpyAsync(out_keys, d_out, sizeof(int64_t) * 3 * n, cudaMemcpyDeviceToHost, st));
  VX_CUDA(cudaStreamSynchronize(st));
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

// STDescManager::init_voxel_map + get_plane (BTC.cpp:279-334): the plane cloud icp_normal works on.  xyz: n x 3 floats
// (PointXYZI x,y,z).  out / aux (10 doubles per plane, may be NULL): planes in ascending voxel-key order.
int vx_btc_plane_cloud(vx_ctx *ctx, const float *xyz, int64_t n, double voxel_size, double plane_thre, int init_num, vx_point48 *out,
                       double *aux, int64_t out_cap, int64_t *n_out) {
  VX_ARG(ctx && n_out && n >= 0 && voxel_size > 0 && out_cap >= 0 && (n == 0 || xyz) && n < ((int64_t)1 << 31));
  VX_TRY(bind(ctx));
  *n_out = 0;
  if (n == 0) return VX_OK;
  cudaStream_t st = ctx->stream;
  DevBuf d_xyz, d_k0, d_k1, d_i0, d_i1, d_flag, d_incl, d_start, d_bad, d_pl, d_aux, d_pf, d_pi, d_out, d_aout;
  auto body = [&]() -> int {
    VX_TRY(d_xyz.reserve(12 * n)); VX_TRY(d_k0.reserve(8 * n)); VX_TRY(d_k1.reserve(8 * n)); VX_TRY(d_i0.reserve(4 * n)); VX_TRY(d_i1.reserve(4 * n));
    VX_TRY(d_flag.reserve(4 * n)); VX_TRY(d_incl.reserve(4 * n)); VX_TRY(d_start.reserve(4 * (n + 1))); VX_TRY(d_bad.reserve(4));
    VX_CUDA(cudaMemcpyAsync(d_xyz.p, xyz, 12 * n, cudaMemcpyHostToDevice, st));
    VX_CUDA(cudaMemsetAsync(d_bad.p, 0, 4, st));
    VX_KERNEL(ctx, "k_btc_key", k_btc_key<<<blocks_for(n, 256), 256, 0, st>>>(d_xyz.as<float>(), n, voxel_size, d_k0.as<uint64_t>(), d_i0.as<int32_t>(), d_bad.as<int>()));
    size_t tb = 0;
    VX_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, d_k0.as<uint64_t>(), d_k1.as<uint64_t>(), d_i0.as<int32_t>(), d_i1.as<int32_t>(), (int)n, 0, 63, st));
    VX_TRY(ctx->scratch.reserve(tb));
    VX_CUDA(cub::DeviceRadixSort::SortPairs(ctx->scratch.p, tb, d_k0.as<uint64_t>(), d_k1.as<uint64_t>(), d_i0.as<int32_t>(), d_i1.as<int32_t>(), (int)n, 0, 63, st));
    VX_KERNEL(ctx, "k_head_flags", k_head_flags<<<blocks_for(n, 256), 256, 0, st>>>(d_k1.as<uint64_t>(), n, d_flag.as<int32_t>()));
    VX_CUDA(cub::DeviceScan::InclusiveSum(nullptr, tb, d_flag.as<int32_t>(), d_incl.as<int32_t>(), (int)n, st));
    VX_TRY(ctx->scratch.reserve(tb));
    VX_CUDA(cub::DeviceScan::InclusiveSum(ctx->scratch.p, tb, d_flag.as<int32_t>(), d_incl.as<int32_t>(), (int)n, st));
    ctx->launches += 6;
    VX_KERNEL(ctx, "k_cell_starts", k_cell_starts<<<blocks_for(n, 256), 256, 0, st>>>(d_flag.as<int32_t>(), d_incl.as<int32_t>(), n, d_start.as<int32_t>()));
    int32_t ncell = 0; int bad = 0;
    VX_CUDA(cudaMemcpyAsync(&ncell, d_incl.as<int32_t>() + (n - 1), 4, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaMemcpyAsync(&bad, d_bad.p, 4, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    if (bad) { set_error("vx_btc_plane_cloud: a cell index does not fit the 21-bit packed key"); return VX_ERR_ARG; }
    VX_TRY(d_pl.reserve(sizeof(vx_point48) * ncell)); VX_TRY(d_aux.reserve(80 * (size_t)ncell)); VX_TRY(d_pf.reserve(4 * (size_t)ncell)); VX_TRY(d_pi.reserve(4 * (size_t)ncell));
    VX_KERNEL(ctx, "k_btc_plane", k_btc_plane<<<blocks_for(ncell, 128), 128, 0, st>>>(d_xyz.as<float>(), d_i1.as<int32_t>(), d_start.as<int32_t>(), ncell, init_num,
                                                                                    plane_thre, d_pl.as<vx_point48>(), d_aux.as<double>(), d_pf.as<int32_t>()));
    VX_CUDA(cub::DeviceScan::InclusiveSum(nullptr, tb, d_pf.as<int32_t>(), d_pi.as<int32_t>(), ncell, st));
    VX_TRY(ctx->scratch.reserve(tb));
    VX_CUDA(cub::DeviceScan::InclusiveSum(ctx->scratch.p, tb, d_pf.as<int32_t>(), d_pi.as<int32_t>(), ncell, st));
    ctx->launches += 1;
    int32_t np = 0;
    VX_CUDA(cudaMemcpyAsync(&np, d_pi.as<int32_t>() + (ncell - 1), 4, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    *n_out = np;
    if (np > out_cap) { set_error("vx_btc_plane_cloud: %d planes, capacity %lld", np, (long long)out_cap); return VX_ERR_ARG; }
    if (np == 0) return VX_OK;
    VX_TRY(d_out.reserve(sizeof(vx_point48) * np)); VX_TRY(d_aout.reserve(80 * (size_t)np));
    VX_KERNEL(ctx, "k_btc_compact", k_btc_compact<<<blocks_for(ncell, 256), 256, 0, st>>>(d_pl.as<vx_point48>(), d_aux.as<double>(), d_pf.as<int32_t>(), d_pi.as<int32_t>(),
                                                                                        ncell, d_out.as<vx_point48>(), d_aout.as<double>()));
    if (out) VX_CUDA(cudaMemcpyAsync(out, d_out.p, sizeof(vx_point48) * np, cudaMemcpyDeviceToHost, st));
    if (aux) VX_CUDA(cudaMemcpyAsync(aux, d_aout.p, 80 * (size_t)np, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    return VX_OK;
  };
  const int status = body();
  cudaStreamSynchronize(st);
  for (DevBuf *b : {&d_xyz, &d_k0, &d_k1, &d_i0, &d_i1, &d_flag, &d_incl, &d_start, &d_bad, &d_pl, &d_aux, &d_pf, &d_pi, &d_out, &d_aout}) b->release();
  return status;
}

}  // extern "C"
